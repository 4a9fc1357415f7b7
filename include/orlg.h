/* orlg.h — C-ABI of the B200-native batched stepping engine for the QoT-aware
 * Optical RL-Gym environments (RMSA-v0, DeepRMSA-v0, RMCSA-v0, PhyRMSA-v0).
 *
 * The reference (pure Python) has no FFI; the boundary this ABI replaces is the
 * gym.Env protocol of its envs:
 *   orlg_create      <- RMSAEnv.__init__       optical_rl_gym/envs/rmsa_env.py:29-220
 *                       DeepRMSAEnv.__init__   optical_rl_gym/envs/deeprmsa_env.py:10-46
 *                       RMCSAEnv.__init__      optical_rl_gym/envs/rmcsa_env.py:29-207
 *   orlg_reset       <- RMSAEnv.reset          rmsa_env.py:343-457 (full / counters-only)
 *   orlg_step        <- RMSAEnv.step           rmsa_env.py:222-341
 *                       DeepRMSAEnv.step       deeprmsa_env.py:48-58
 *                       RMCSAEnv.step          rmcsa_env.py:209-339
 *   orlg_observation <- DeepRMSAEnv.observation deeprmsa_env.py:60-121,
 *                       SimpleMatrixObservation rmsa_env.py:940-971
 *   orlg_heuristic   <- shortest_path_first_fit rmsa_env.py:854-871,
 *                       shortest_available_path_first_fit rmsa_env.py:901-913,
 *                       least_loaded_path_first_fit rmsa_env.py:916-937,
 *                       deeprmsa_env.py:135-155, rmcsa_env.py:882-911
 *   orlg_rollout     <- utils.evaluate_heuristic utils.py:124-162 (heuristic+step loop)
 *   orlg_load_trace  <- the five rng draws of _next_service rmsa_env.py:646-659
 *   orlg_get_counters<- the counters behind the info dict rmsa_env.py:293-307
 *   orlg_export_state<- topology.graph["available_slots"] rmsa_env.py:396-398
 *   orlg_gsnr        <- calculate_osnr examples/calculate_osnr.py:9-56
 *   PhyRMSA-v0 (env_kind 3): orlg_step <- PhyRMSAEnv.step phy_rmsa_env.py:272-351,
 *                       orlg_heuristic <- phy_rmsa_env.py:1254-1438,1676-1737,
 *                       orlg_phy_get_stats <- the info dict phy_rmsa_env.py:319-348
 *
 * All pointers named dev_* are CUDA device pointers owned by the caller; the
 * engine owns its internal state.  Functions return 0 or a negative ORLG_E*.
 * Work is asynchronous on the given stream (a cudaStream_t passed as void*).
 * No C++ exceptions and no torch types cross this boundary.
 */
#ifndef ORLG_H
#define ORLG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORLG_ABI_VERSION 2

enum {
  ORLG_OK = 0,
  ORLG_EINVAL = -1,   /* bad config / argument */
  ORLG_ENOMEM = -2,   /* device or host allocation failed */
  ORLG_ECUDA = -3,    /* CUDA runtime error (see orlg_last_error) */
  ORLG_ESTATE = -4,   /* call not valid in this state (e.g. replay without trace) */
  ORLG_EFAULT = -5    /* device-side sticky fault (heap overflow ...) */
};

enum { ORLG_ENV_RMSA = 0, ORLG_ENV_DEEPRMSA = 1, ORLG_ENV_RMCSA = 2, ORLG_ENV_PHYRMSA = 3 };

enum { ORLG_SRC_REPLAY = 0, ORLG_SRC_PHILOX = 1 };

/* heuristic ids (orlg_heuristic / orlg_rollout) */
enum {
  ORLG_H_SP_FF = 0,    /* rmsa_env.py:854-871 ; DeepRMSA: deeprmsa_env.py:135-143 */
  ORLG_H_SAP_FF = 1,   /* rmsa_env.py:901-913 ; DeepRMSA: deeprmsa_env.py:146-155 */
  ORLG_H_LLP_FF = 2,   /* rmsa_env.py:916-937 */
  ORLG_H_SAP_BM_FC_FF = 3, /* rmcsa_env.py:882-911 */
  ORLG_H_RANDOM = 4,   /* uniform action from the env's Philox stream */
  /* PhyRMSA-v0, physical layer (grooming off): */
  ORLG_H_PHY_SAP_FF = 5,     /* sapff_rmsa           phy_rmsa_env.py:1676-1737 */
  ORLG_H_PHY_SAP_BM = 6,     /* phy_aware_sapbm_rmsa phy_rmsa_env.py:1254-1314 */
  ORLG_H_PHY_BM_FF = 7,      /* phy_aware_bmff_rmsa  phy_rmsa_env.py:1317-1372 */
  ORLG_H_PHY_BM_FA_CUT = 8,  /* phy_aware_bmfa_rmsa  phy_rmsa_env.py:1375-1438 */
  ORLG_H_PATH_FF = 9,        /* PathOnlyFirstFitAction rmsa_env.py:974-1008 (orlg_path_first_fit only) */
  ORLG_H_PHY_FA_FF = 10,     /* phy_aware_faff_rmsa  phy_rmsa_env.py:1508-1569 */
  ORLG_H_PHY_BM_FA_RSS = 11, /* phy_aware_bmfa_rss_rmsa phy_rmsa_env.py:1441-1505 (calculate_r_spatial :1085-1108; <= 64 links) */
  ORLG_H_PHY_FA_FF_RSS = 12, /* phy_aware_faff_rss_rmsa phy_rmsa_env.py:1572-1633 */
  /* engines with the GN-model QoT gate (orlg_config.qot_gate): the walk of shortest_available_path_first_fit
   * (rmsa_env.py:901-913) / ..._best_modulation_first_core_first_fit (rmcsa_env.py:882-911) that returns the
   * first candidate passing the gate; with qot_adaptive every modulation is tried per path, most efficient first */
  ORLG_H_SAP_GN_FF = 13
};

enum {
  ORLG_OBS_NONE = 0, ORLG_OBS_F64 = 1, ORLG_OBS_F32 = 2, ORLG_OBS_MATRIX_U8 = 3,
  /* PhyRMSA-v0: f64 (number_cuts_total, rss_total_metric) of the step's info dict
   * (phy_rmsa_env.py:313-314,337-338; what UseInfoReward wrappers.py:4-16 feeds back) */
  ORLG_OBS_PHY_INFO = 4,
  ORLG_OBS_PHY_CUTS = 5 /* number_cuts_total only (the RSS pass is ~5x the cost of a step) */
};

#define ORLG_PHY_ACTION_DIM 9 /* PhyRMSA action: (path, 8 channel numbers), -1 terminated; path -2 = block */

/* One request as drawn by _next_service (rmsa_env.py:646-659).  `arrival` is the
 * absolute clock value after the draw (current_time + expovariate), `holding`
 * the holding time; both raw f64 so that replay is bit-exact. */
typedef struct orlg_request {
  double arrival;
  double holding;
  int32_t bit_rate;
  int16_t src; /* node ids in graph["node_indices"] order */
  int16_t dst;
} orlg_request; /* 24 bytes */

/* ---- GN-model GSNR: calculate_osnr, examples/calculate_osnr.py:9-56 --------------------
 * Physical inputs the reference leaves undefined are explicit here.  Every
 * link is split into span_count[l] equal spans of span_length_km[l]
 * (examples/create_topology_gn.py:122-134: int(length // 80) + 1 spans).  A service that
 * occupies slots [first, first+n) has bandwidth n*slot_hz and centre frequency
 * f0_hz + (first + n/2)*slot_hz. */
typedef struct orlg_gsnr_params {
  double launch_power_w;    /* per-channel launch power (1e-3 = 0 dBm) */
  double attenuation_per_m; /* field attenuation alpha [1/m]: power falls as exp(-2*alpha*L) */
  double noise_figure_lin;  /* amplifier noise figure, linear */
  double f0_hz;             /* lower edge of slot 0 */
  double slot_hz;           /* slot width in Hz (channel_width * 1e9) */
  const int32_t* span_count;    /* host, [n_edges] */
  const double* span_length_km; /* host, [n_edges] */
} orlg_gsnr_params;


typedef struct orlg_config {
  int32_t abi_version; /* = ORLG_ABI_VERSION */
  int32_t env_kind;    /* ORLG_ENV_* */
  int32_t device;      /* CUDA device ordinal */
  int32_t num_envs;    /* envs held by this engine (this GPU's shard) */
  int64_t global_env_offset; /* global id of local env 0 (Philox key, trace rows) */

  /* topology tables, host pointers, copied by orlg_create */
  int32_t n_nodes, n_edges, k_paths, max_hops;
  const int16_t* path_links;       /* [n_nodes*n_nodes][k_paths][max_hops], -1 padded, by edge "index" */
  const int16_t* path_links_by_id; /* same through edge "id" (rmsa_env.py:750, F9) or NULL = path_links */
  const int32_t* path_hops;        /* [n_nodes*n_nodes][k_paths], 0 = absent */
  const double* path_length;       /* [n_nodes*n_nodes][k_paths] km */
  const int32_t* path_se;          /* spectral efficiency of best_modulation */
  int32_t n_mods;
  const int32_t* mod_se;           /* [n_mods] */
  const double* mod_max_length;    /* [n_mods] */
  const double* mod_reach;         /* RMCSA: [n_mods] lmax_xt (rmcsa_env.py:377-379), host-evaluated */
  const double* mod_snr_coeff;     /* RMCSA: [n_mods] SNR_min*h*f*G*NF prefix product (rmcsa_env.py:366-374) */

  /* env parameters (constructor kwargs of the reference) */
  int32_t num_slots;     /* num_spectrum_resources */
  int32_t num_cores;     /* num_spatial_resources (RMCSA), else 1 */
  int32_t j;             /* DeepRMSA blocks per path */
  int32_t episode_length;
  int32_t allow_rejection;
  int32_t auto_reset;    /* VecEnv semantics: counters-only reset in-kernel when done */
  double channel_width;  /* GHz per slot (12.5) */
  double mean_iat;       /* mean_service_inter_arrival_time (optical_network_env.py:127-129) */
  double mean_holding;   /* mean_service_holding_time */

  /* request source */
  int32_t request_source; /* ORLG_SRC_* */
  int32_t n_bit_rates;    /* discrete table size, 0 = continuous randint(lo, hi) */
  uint64_t seed;          /* Philox key */
  const int32_t* bit_rates;     /* [n_bit_rates] */
  const double* bit_rate_cum;   /* [n_bit_rates] cumulative weights (random.choices) */
  int32_t bit_rate_lo, bit_rate_hi;
  const double* src_cum;        /* [n_nodes] cumulative node weights */
  const double* dst_cum;        /* [n_nodes][n_nodes] cumulative weights given src (optical_network_env.py:201-206) */

  int32_t heap_capacity;  /* release-queue entries per env; 0 = provable bound */
  int32_t obs_kind;       /* ORLG_OBS_* written by orlg_step */
  /* PhyRMSA virtual layer (grooming, phy_rmsa_env.py:625-659,1650-1673): 0 = channel_state not kept, physical
   * layer only; 1 = env.grooming True; 2 = env.grooming False with channel_state kept, which is the reference's
   * behaviour: sapff/sapbm/bmff/faff/faff_rss call use_existing_channels unconditionally, bmfa/bmfa_rss only
   * under env.grooming.  A virtual action is (idp + 20, channels); shares are taken as the reference's
   * use_existing_channels takes them (all the free capacity of each channel, the last one the remainder). */
  int32_t phy_grooming;
  /* RMSA-v0 / DeepRMSA-v0: 1 = keep what the info dict of RMSAEnv.step needs (rmsa_env.py:276-332): the
   * time-weighted link utilization / compactness of _update_link_stats :562-641, the per-bit-rate histograms
   * and the running services' slots x hops; orlg_set_info_output then makes orlg_step write the dict */
  int32_t info_stats;
  int32_t reserved[4];

  /* PhyRMSA-v0 (phy_rmsa_env.py:51-53,93-95): num_slots = number of channels, num_cores = 1 */
  int32_t phy_n_tables;   /* rows of connections_detail */
  int32_t phy_max_terms;  /* row length of phy_cut_links (<= 63 used per path) */
  const int32_t* phy_table_id;   /* [n_nodes*n_nodes] first row matching the unordered pair, -1 = none */
  const uint8_t* phy_mod_level;  /* [phy_n_tables][num_slots][k_paths], levels 1..8 */
  const double* phy_gsnr;        /* [phy_n_tables][num_slots][k_paths] dB */
  const uint8_t* phy_cut_links;  /* [n_nodes*n_nodes*k_paths][phy_max_terms] neighbour link per r_cut term */
  const int32_t* phy_cut_count;  /* [n_nodes*n_nodes*k_paths] */

  /* ---- QoT gate: GN-model GSNR feeding the QoT feasibility and modulation-format check (RMSA-v0, RMCSA-v0) ----
   * The reference ships the GN model (examples/calculate_osnr.py:9-56) and the SD-FEC thresholds it is meant to be
   * compared with (minimum_osnr = target_SNR_dB_vec, examples/create_topology_gn.py:65-110) but no env that calls
   * them.  With qot_gate = 1 a step accepts a lightpath iff, beyond the env's own checks
   * (rmsa_env.py:233-252; rmcsa_env.py:222-258 incl. _crosstalk_is_acceptable), (a) its GSNR given the running
   * services on its links (and core) is >= qot_thr_db[its spectral efficiency] and (b) the GSNR of every
   * neighbour - a running service sharing a link and core with it - re-evaluated with the candidate present is
   * >= qot_thr_db[the neighbour's spectral efficiency].
   * RMSA-v0: the modulation is path.best_modulation as in get_number_slots (rmsa_env.py:708-719), or with
   * qot_adaptive = 1 the env assigns the most efficient modulation that passes at (path, initial_slot): spectral
   * efficiencies qot_se[0..qot_n_se) in that (descending) order; one whose slot count does not fit ends the
   * search (wider ones fit even less); one whose interference-free GSNR (ASE + self-channel NLI only) misses its
   * threshold is skipped.  RMCSA-v0: the action names the modulation; qot_adaptive only changes ORLG_H_SAP_GN_FF.
   * The engine keeps per running service the cross-channel sums of calculate_osnr.py:31-46 up to date (one term
   * per shared link enters / leaves when a lightpath is provisioned / released), so a decision costs
   * O(running services) instead of O(neighbours x interferers); the occupancy of such engines is env-major. */
  int32_t qot_gate, qot_adaptive, qot_n_se, qot_pad;
  int32_t qot_se[8];
  double qot_thr_db[16];               /* by spectral efficiency; +inf = never acceptable */
  const orlg_gsnr_params* qot_gsnr;    /* physical inputs of the GN model */

  /* PhyRMSA-v0 periodic defragmentation (phy_rmsa_env.py:54-56, 355-417): every phy_defrag_period processed
   * services the step also runs _groom_defragmentation (:703-733) and then reallocates up to phy_number_moves whole
   * channels, candidates ranked by the gain in the fragmentation metric (phy_metric 0 = "cut" calculate_r_cut
   * :1123-1193, 1 = "rss" calculate_r_spatial :1085-1108; "rss" needs <= 64 links) and by age (_move :662-697).
   * 0 = off.  Needs phy_grooming != 0 (the grooming moves read channel_state). */
  int32_t phy_defrag_period, phy_number_moves, phy_metric, phy_pad;
} orlg_config;

typedef struct orlg_engine orlg_engine;

/* counters exported per env by orlg_get_counters, int64 each */
enum {
  ORLG_C_SERVICES_PROCESSED = 0,
  ORLG_C_SERVICES_ACCEPTED = 1,
  ORLG_C_EP_SERVICES_PROCESSED = 2,
  ORLG_C_EP_SERVICES_ACCEPTED = 3,
  ORLG_C_BIT_RATE_REQUESTED = 4,
  ORLG_C_BIT_RATE_PROVISIONED = 5,
  ORLG_C_EP_BIT_RATE_REQUESTED = 6,
  ORLG_C_EP_BIT_RATE_PROVISIONED = 7,
  ORLG_C_REQUESTS_DRAWN = 8,   /* number of _next_service draws so far */
  ORLG_C_QUEUE_LEN = 9,        /* live entries of the release heap */
  ORLG_C_FAULT = 10,           /* sticky fault bits */
  ORLG_C_EPISODES = 11,        /* completed episodes */
  ORLG_NUM_COUNTERS = 12
};

int orlg_abi_version(void);
const char* orlg_build_info(void);

int orlg_create(const orlg_config* cfg, orlg_engine** out);
int orlg_destroy(orlg_engine* e);
const char* orlg_last_error(const orlg_engine* e);

/* Geometry helpers so callers can size their buffers. */
int orlg_action_dim(const orlg_engine* e);
int orlg_obs_dim(const orlg_engine* e);        /* elements per env of the configured obs */
int64_t orlg_state_words(const orlg_engine* e); /* uint32 words per env of the packed occupancy */
int64_t orlg_hbm_bytes(const orlg_engine* e);   /* device bytes owned by the engine */

/* Replay source: dev_trace is [n_requests][num_envs] orlg_request (request r of
 * local env i at dev_trace[r*num_envs + i]); the engine keeps the pointer (caller
 * keeps the buffer alive).  Request 0 is consumed by orlg_reset(full=1). */
int orlg_load_trace(orlg_engine* e, const orlg_request* dev_trace, int64_t n_requests);

/* full=1: constructor-time reset (rmsa_env.py:391-457) incl. drawing the first
 * request; full=0: per-episode counters only (rmsa_env.py:344-389). */
int orlg_reset(orlg_engine* e, int full, void* stream);

/* One env step for every env.  dev_actions: int32 [num_envs][action_dim].
 * Outputs (any may be NULL): dev_reward int32[num_envs], dev_done uint8[num_envs],
 * dev_obs [num_envs][obs_dim] of the configured obs type, dev_alloc int32
 * [num_envs][4] = (accepted, path, initial_slot, n_slots) of the decision taken
 * (core in bits 16.. of path for RMCSA). */
int orlg_step(orlg_engine* e, const int32_t* dev_actions, int32_t* dev_reward,
              uint8_t* dev_done, void* dev_obs, int32_t* dev_alloc, void* stream);

/* Same, with HOST buffers: copies actions H2D, steps, copies outputs D2H and
 * synchronises the stream (the end-to-end path measured by bench.py "e2e"). */
int orlg_step_host(orlg_engine* e, const int32_t* host_actions, int32_t* host_reward,
                   uint8_t* host_done, void* host_obs, int32_t* host_alloc, void* stream);

/* The two halves of orlg_step_host, the VecEnv step_async / step_wait pair: _async enqueues the H2D
 * copy, the step and the D2H copies on `stream` and returns; the host buffers (pinned memory for real
 * overlap) are valid after orlg_step_host_wait(stream).  Engines driven on different streams
 * overlap their copies, kernels and the caller's host work. */
int orlg_step_host_async(orlg_engine* e, const int32_t* host_actions, int32_t* host_reward,
                         uint8_t* host_done, void* host_obs, int32_t* host_alloc, void* stream);
int orlg_step_host_wait(orlg_engine* e, void* stream);
/* How orlg_step_host_wait waits: 0 (default) = cudaStreamSynchronize (the driver spins: lowest latency when
 * every engine's host thread has a core to itself), 1 = sleep on a blocking-sync event recorded by the _async
 * call (for hosts with fewer cores than busy threads, e.g. eight per-GPU processes on a 16-thread box, where
 * spinning waiters take the cores the policy needs). */
int orlg_set_host_wait(orlg_engine* e, int blocking);

/* Layout of dev_obs for orlg_step / orlg_observation, in elements: feature f of
 * env i is written at dev_obs[i*env_stride + f*feat_stride].  Default (obs_dim, 1)
 * = row-major [num_envs][obs_dim]; (1, num_envs) = feature-major, which makes the
 * kernel's stores coalesced (the Python wrapper exposes it as a transposed view). */
int orlg_set_obs_layout(orlg_engine* e, int64_t env_stride, int64_t feat_stride);

/* Observation of the pending request without stepping. */
int orlg_observation(orlg_engine* e, int obs_kind, void* dev_obs, void* stream);

/* Device-side heuristic: writes the action the reference heuristic would return.
 * QoT-gated engines, ORLG_H_SAP_GN_FF: the launch also records its verdict (lightpath, modulation, running sums) as a
 * proposal.  If the next state-touching call on the engine is orlg_step (or orlg_step_host*), every env whose action
 * equals the proposal takes that verdict instead of evaluating the gate a second time - the reference's agent loop
 * `action = heuristic(env); env.step(action)` costs one gate evaluation, not two.  Any other action is evaluated as
 * usual; orlg_reset, orlg_load_trace or another step in between end the proposal.  Results are identical either way. */
int orlg_heuristic(orlg_engine* e, int heuristic_id, int32_t* dev_actions, void* stream);

/* Fused next-action hint.  With a non-NULL dev_next_actions (int32 [num_envs][action_dim], kept by
 * the engine until replaced or cleared with NULL) every following orlg_step / orlg_step_host also
 * writes, for the NEW pending request of every env, the action the reference heuristic
 * `heuristic_id` takes on the updated network: exactly what an orlg_heuristic launch after the step
 * would write, without a second launch and a second read of the occupancy.  The buffer may be the
 * dev_actions buffer of the same orlg_step call (an env reads its action before it writes its hint). */
int orlg_set_next_action_hint(orlg_engine* e, int heuristic_id, int32_t* dev_next_actions);

/* PathOnlyFirstFitAction.action (rmsa_env.py:982-1005): dev_paths int32[num_envs] (path index or k_paths =
 * reject) -> RMSA-v0: dev_actions int32[num_envs][2] = (path, first-fit slot) or (k_paths, num_slots).
 * RMCSA-v0 (rmcsa_env.py:950-982; as shipped that wrapper calls get_number_slots / is_path_free with RMSA's
 * signatures and raises, so its intent is built): int32[num_envs][4] = (path, best modulation by reach, first core
 * with room, first-fit slot) or the reject action. */
int orlg_path_first_fit(orlg_engine* e, const int32_t* dev_paths, int32_t* dev_actions, void* stream);

/* nsteps x (heuristic -> step) in one launch, state staged on chip between steps.
 * dev_reward_sum (optional) int64[num_envs] accumulates rewards. */
int orlg_rollout(orlg_engine* e, int heuristic_id, int nsteps, int64_t* dev_reward_sum,
                 void* stream);

int orlg_get_counters(orlg_engine* e, int64_t* dev_out /* [num_envs][ORLG_NUM_COUNTERS] */,
                      void* stream);

/* With a non-NULL dev_snapshot (int64 [num_envs][8], kept by the engine until replaced or cleared with NULL) every
 * following orlg_step also writes the counters ORLG_C_SERVICES_PROCESSED .. ORLG_C_EP_BIT_RATE_PROVISIONED as they are
 * where the reference builds its info dict: after the decision, before _next_service counts the next request and
 * before an episode reset (rmsa_env.py:293-307 precede :334-335) - the values behind service_blocking_rate,
 * episode_service_blocking_rate, bit_rate_blocking_rate and episode_bit_rate_blocking_rate. */
int orlg_set_counter_snapshot(orlg_engine* e, int64_t* dev_snapshot);

/* PhyRMSA-v0 per-env statistics, f64 [num_envs][8]: total_path_length_episode, total_gsnr_episode,
 * total_path_index_episode, total_modulation_level_episode, channels_accepted_episode,
 * physical_services_accepted_episode (phy_rmsa_env.py:431-439), then number_cuts_total and
 * rss_total_metric (:1195-1203, :1110-1121) of the CURRENT channel map. */
int orlg_phy_get_stats(orlg_engine* e, double* dev_out, void* stream);

/* The info dict of RMSAEnv.step (rmsa_env.py:293-332) for every following orlg_step, f64
 * [num_envs][orlg_info_dim]: service_blocking_rate, episode_service_blocking_rate, bit_rate_blocking_rate,
 * episode_bit_rate_blocking_rate, network_compactness (:806-851), network_compactness_difference,
 * avg_link_compactness, avg_link_utilization (np.mean over the edges, numpy's summation order), fairness, then
 * bit_rate_blocking_<b> for every configured bit rate — evaluated where the reference evaluates them: after the
 * provisioning, before the next request is drawn and the releases are processed.  NULL switches it off.
 * Needs info_stats = 1 (RMSA-v0 / DeepRMSA-v0). */
int orlg_info_dim(const orlg_engine* e);
int orlg_set_info_output(orlg_engine* e, double* dev_info);

/* PhyRMSA-v0 defragmentation counters of the current episode, int64 [num_envs][3]: counted_moves,
 * counted_moves_groom, counted_defrag_cycles (phy_rmsa_env.py:110-112; info keys num_moves = [0] / 2 + [1],
 * num_moves_groom, num_defrag_cycle :340-342). */
int orlg_phy_get_moves(orlg_engine* e, int64_t* dev_out, void* stream);

/* PhyRMSA-v0 channel_state (phy_rmsa_env.py:158-166,599-601): the partially used channels of every env in
 * append order, int32 [num_envs][max_entries][6] = (src_id * n_nodes + dst_id, idp, channel, used, free,
 * capacity) in units of 100 Gb/s, and their number per env in dev_count [num_envs] (entries past
 * max_entries are counted, not written).  Needs phy_grooming != 0. */
int orlg_phy_get_channel_state(orlg_engine* e, int32_t* dev_out, int32_t max_entries, int32_t* dev_count,
                               void* stream);

/* Unpacked occupancy: uint8 [num_envs][num_cores][n_edges][num_slots], 1 = free,
 * i.e. topology.graph["available_slots"] of every env. */
int orlg_export_state(orlg_engine* e, uint8_t* dev_out, void* stream);

/* Pending request of every env (current_service). */
int orlg_get_requests(orlg_engine* e, orlg_request* dev_out, void* stream);

/* ---- GN-model GSNR calls (parameter block: orlg_gsnr_params above) ---- */
int orlg_gsnr_configure(orlg_engine* e, const orlg_gsnr_params* params);

/* For every env: dev_query int32[num_envs][4] = (path | core << 16, first_slot, n_slots,
 * spectral efficiency of the candidate's modulation or 0 = the path's best modulation) of a
 * candidate lightpath for the pending request (not provisioned).  dev_out f64[num_envs][2]:
 *   [0] GSNR in dB of the candidate given the env's running services;
 *   [1] the lowest GSNR in dB among its neighbours - running services that share a link (and
 *       core) with the candidate - re-evaluated with the candidate present; +inf if none.
 * A query with path < 0 or path >= k_paths yields NaN in both. */
int orlg_gsnr(orlg_engine* e, const int32_t* dev_query, double* dev_out, void* stream);

/* The same two values for the candidate a step action describes: dev_actions is an action tensor in the
 * env's own format for the pending request - RMSA-v0 int32[num_envs][2] (path, first slot) with the path's
 * best modulation, RMCSA-v0 int32[num_envs][4] (path, modulation, core, first slot) - e.g. what
 * orlg_heuristic or the fused next-action hint wrote; the slot count is get_number_slots of the pending
 * bit rate (rmsa_env.py:708-719, rmcsa_env.py:753-765).  A rejection action yields NaN.  This is how the
 * QoT-aware RMCSA configuration evaluates the GN model for every decision without a host round trip. */
int orlg_gsnr_for_actions(orlg_engine* e, const int32_t* dev_actions, double* dev_out, void* stream);

/* One step with the device heuristic as the policy - one iteration of utils.evaluate_heuristic's loop
 * (utils.py:124-162): the action `heuristic_id` takes for the pending request is written to dev_actions_out
 * (optional) and stepped; outputs as orlg_step.  Engines with the QoT gate and ORLG_H_SAP_GN_FF evaluate the gate
 * once for both (2 launches); everything else is orlg_heuristic + orlg_step. */
int orlg_step_auto(orlg_engine* e, int heuristic_id, int32_t* dev_actions_out, int32_t* dev_reward, uint8_t* dev_done,
                   void* dev_obs, int32_t* dev_alloc, void* stream);

/* orlg_step_auto for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same step as orlg_step_auto(e, heuristic_id, NULL, ...), but the action tensor that travels from the heuristic
 * launch to the step launch lives in a static device array inside liborlg.so instead of a lazily cudaMalloc'ed
 * buffer; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: num_envs * action_dim <= ORLG_NOALLOC_MAX_INTS int32 (1 MiB of device memory, held by every GPU that loads
 * the library until the process exits); larger engines get ORLG_EINVAL with a message naming the limit.
 * The array is shared by every engine and every call on a device and keeps the last call's contents (each call
 * overwrites all the entries it reads, nothing needs clearing): calls that overlap on the GPU - two streams, two
 * engines, two host threads - race on it; serialise them or use orlg_step_auto with buffers of your own. */
#define ORLG_NOALLOC_MAX_INTS 262144
int orlg_step_auto_noalloc(orlg_engine* e, int heuristic_id, int32_t* dev_reward, uint8_t* dev_done, void* dev_obs,
                           int32_t* dev_alloc, void* stream);

/* QoT gate: f64 [num_envs][2] = (GSNR in dB of the last candidate the gate evaluated for the env, lowest GSNR among
 * its neighbours with the candidate present; +inf without neighbours); NaN where nothing was evaluated (invalid
 * action, no free spectrum) and, for [1], where the candidate itself missed its threshold.  After orlg_step_auto
 * with ORLG_H_SAP_GN_FF (gate evaluated once for heuristic and step) a blocked request reports the last candidate
 * the heuristic looked at. */
int orlg_qot_get(orlg_engine* e, double* dev_out, void* stream);
/* QoT gate self-check (tests): dev_out int64 [num_envs] = number of (running service, hop) rows whose running sums
 * differ from a from-scratch evaluation over the release queue; 0 everywhere unless the engine is broken. */
int orlg_qot_check(orlg_engine* e, int64_t* dev_out, void* stream);

/* The running services of every env (topology.graph["running_services"], rmsa_env.py:484-487; the owner map
 * spectrum_slots_allocation :481-483 follows from it): int32 [num_envs][max_rows][6] = (src_id * n_nodes + dst_id,
 * path, core, first slot, slots, spectral efficiency) in release-queue order, and their number per env in
 * dev_count (rows past max_rows are counted, not written).  PhyRMSA: one row per (service, channel):
 * (pair, path, share, channel, position in the action, level). */
int orlg_export_services(orlg_engine* e, int32_t* dev_out, int32_t max_rows, int32_t* dev_count, void* stream);

/* Measurement support (bench.py): nsteps steps - mode 0: orlg_step(dev_actions, ...) as configured (fused hint
 * included); mode 1: orlg_step_auto(heuristic_id, dev_actions as its output or NULL, ...) - each bracketed by its own
 * CUDA event pair, with an optional L2-evicting memset of dev_flush[flush_bytes] before every pair, issued from one C
 * loop (no interpreter between an event record and the launch it times) or, with use_graph, captured into one CUDA
 * graph and launched once.  ms_step[nsteps] = duration of every step; ms_first[nsteps] (optional) = duration of
 * the step's first launch when it has two (gate / heuristic kernel), else 0.  Synchronous. */
int orlg_timed_steps(orlg_engine* e, int mode, int heuristic_id, int32_t* dev_actions, int32_t* dev_reward,
                     uint8_t* dev_done, void* dev_obs, int nsteps, void* dev_flush, size_t flush_bytes, int use_graph,
                     float* ms_step, float* ms_first, void* stream);

/* kernels launched by this engine so far (bench.py "gpu_launches") */
int64_t orlg_launch_count(const orlg_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* ORLG_H */
