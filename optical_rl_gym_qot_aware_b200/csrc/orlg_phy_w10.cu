// PhyRMSA-v0 env_kernel instantiations for W = 10 words per link row (see engine_launch.cuh)
#define ORLG_INSTANTIATE_PHY 10
#include "engine_launch.cuh"
