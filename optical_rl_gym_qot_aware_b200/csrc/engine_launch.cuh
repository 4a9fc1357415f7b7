// engine_launch.cuh — host-side launch of env_kernel<W, BD, MT, PHY>.  Each word count W is
// instantiated in its own translation unit (orlg_w2.cu / orlg_w4.cu / orlg_w10.cu) so the
// three compile in parallel; orlg.cu only sees the declaration of launch_w.
#pragma once
#include <cuda_runtime.h>

#include "engine_kernel.cuh"

namespace orlg {

struct LaunchGeom {
  int BD, NP, device;  // BD = threads per CTA (= envs per staged tile)
  size_t smem_bytes;
  bool info;    // engine created with info_stats: INFO instantiations (staged tiles of 128 envs, or direct)
  bool direct;  // env-major occupancy, no staging (DIRECT instantiations, CTAs of DIRECT_CTA threads)
};
constexpr int DIRECT_CTA = 64;

template <int W>
cudaError_t launch_w(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st);
// PhyRMSA-v0 instantiations (tile of 32 envs), compiled in orlg_phy.cu
template <int W>
cudaError_t launch_phy(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st);

#if defined(ORLG_INSTANTIATE_W) || defined(ORLG_INSTANTIATE_PHY)
template <int W, int BD, int MT, bool PHY, bool DIRECT, bool INFO = false>
static cudaError_t launch_t(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
  static size_t configured[64] = {};  // largest dynamic smem opted into, per device
  if (g.smem_bytes > 48 * 1024 && configured[g.device & 63] < g.smem_bytes) {
    cudaError_t rc = cudaFuncSetAttribute(env_kernel<W, BD, MT, PHY, DIRECT, INFO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)g.smem_bytes);
    if (rc != cudaSuccess) return rc;
    configured[g.device & 63] = g.smem_bytes;
  }
  env_kernel<W, BD, MT, PHY, DIRECT, INFO><<<(g.NP + BD - 1) / BD, BD, g.smem_bytes, st>>>(p, io);
  return cudaGetLastError();
}

template <int W, int BD, bool PHY, bool DIRECT = false>
static cudaError_t launch_m(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
  switch (io.mode) {
    case MODE_STEP: return launch_t<W, BD, MODE_STEP, PHY, DIRECT>(g, p, io, st);
    case MODE_HEURISTIC: return launch_t<W, BD, MODE_HEURISTIC, PHY, DIRECT>(g, p, io, st);
    case MODE_ROLLOUT: return launch_t<W, BD, MODE_ROLLOUT, PHY, DIRECT>(g, p, io, st);
    default: return launch_t<W, BD, -1, PHY, DIRECT>(g, p, io, st);
  }
}

// engines with info_stats: only the modes that change state carry the info code
template <int W, int BD, bool DIRECT>
static cudaError_t launch_info(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
  switch (io.mode) {
    case MODE_STEP: return launch_t<W, BD, MODE_STEP, false, DIRECT, true>(g, p, io, st);
    case MODE_HEURISTIC: return launch_t<W, BD, MODE_HEURISTIC, false, DIRECT>(g, p, io, st);
    case MODE_ROLLOUT: return launch_t<W, BD, MODE_ROLLOUT, false, DIRECT, true>(g, p, io, st);
    default: return launch_t<W, BD, -1, false, DIRECT, true>(g, p, io, st);
  }
}

#endif
#ifdef ORLG_INSTANTIATE_W
template <int W>
cudaError_t launch_w(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
#ifdef ORLG_FAST_AB  // A/B builds (tools/build_variant_w4.sh): only the staged 128-env tile without info_stats
  if (g.info || g.direct || g.BD != 128) return cudaErrorNotSupported;
  return launch_m<W, 128, false>(g, p, io, st);
#endif
  if (g.info) return g.direct ? launch_info<W, DIRECT_CTA, true>(g, p, io, st) : launch_info<W, 128, false>(g, p, io, st);
  if (g.direct) return launch_m<W, DIRECT_CTA, false, true>(g, p, io, st);
  switch (g.BD) {
    case 32: return launch_m<W, 32, false>(g, p, io, st);
    case 64: return launch_m<W, 64, false>(g, p, io, st);
    default: return launch_m<W, 128, false>(g, p, io, st);
  }
}
template cudaError_t launch_w<ORLG_INSTANTIATE_W>(const LaunchGeom&, const Params&, const StepIO&, cudaStream_t);
#endif
#ifdef ORLG_INSTANTIATE_PHY
// PhyRMSA engines with periodic defragmentation: the modes that step carry it (INFO flag of the PHY kernels)
template <int W, int BD, bool DIRECT>
static cudaError_t launch_phy_defrag(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
  switch (io.mode) {
    case MODE_STEP: return launch_t<W, BD, MODE_STEP, true, DIRECT, true>(g, p, io, st);
    case MODE_ROLLOUT: return launch_t<W, BD, MODE_ROLLOUT, true, DIRECT, true>(g, p, io, st);
    case MODE_HEURISTIC: return launch_t<W, BD, MODE_HEURISTIC, true, DIRECT>(g, p, io, st);
    default: return launch_t<W, BD, -1, true, DIRECT>(g, p, io, st);
  }
}
template <int W>
cudaError_t launch_phy(const LaunchGeom& g, const Params& p, const StepIO& io, cudaStream_t st) {
  if (p.phy_defrag_period > 0)
    return g.direct ? launch_phy_defrag<W, DIRECT_CTA, true>(g, p, io, st) : launch_phy_defrag<W, 32, false>(g, p, io, st);
  if (g.direct) return launch_m<W, DIRECT_CTA, true, true>(g, p, io, st);
  return launch_m<W, 32, true>(g, p, io, st);
}
template cudaError_t launch_phy<ORLG_INSTANTIATE_PHY>(const LaunchGeom&, const Params&, const StepIO&, cudaStream_t);
#endif

}  // namespace orlg
