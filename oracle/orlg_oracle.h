/* orlg_oracle.h — CPU restatement of the reference's env stepping path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this.  The
 * product (optical_rl_gym_qot_aware_b200/, include/orlg.h) never links it.
 *
 * Parity status: PINNED for RMSA-v0 / DeepRMSA-v0 / RMCSA-v0 against golden
 * traces exported from the unmodified reference run in the build container
 * (oracle/refharness/gen_golden.py -> tests/golden/ (npz files), checked by
 * tests/test_oracle_golden.py).  The reference itself holds no golden vectors
 * (SURVEY.md F13).  PhyRMSA-v0: PINNED the same way (gen_phy_golden.py, tests/test_phy_oracle.py).
 * GN-model GSNR (examples/calculate_osnr.py): the formula is PINNED against numbers produced by the
 * reference's own function body (gen_gsnr_golden.py injects the missing Span/Link classes, SURVEY.md
 * F2); the physical inputs the reference never defines are ours (spec.gsnr_params).
 *
 * Deliberately literal: one byte per (core, link, slot) like the reference's
 * numpy int arrays, run-length encoding by scanning, heap as in heapq.  It
 * shares no code with the CUDA engine.
 */
#ifndef ORLG_ORACLE_H
#define ORLG_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_request { /* same layout as orlg_request */
  double arrival;
  double holding;
  int32_t bit_rate;
  int16_t src;
  int16_t dst;
} oracle_request;

/* GN-model GSNR (examples/calculate_osnr.py:9-56); same parameter block as orlg_gsnr_params */
typedef struct oracle_gsnr_params {
  double launch_power_w, attenuation_per_m, noise_figure_lin, f0_hz, slot_hz;
  const int32_t* span_count;
  const double* span_length_km;
} oracle_gsnr_params;

typedef struct oracle_config {
  int32_t env_kind; /* 0 RMSA, 1 DeepRMSA, 2 RMCSA, 3 PhyRMSA */
  int32_t n_nodes, n_edges, k_paths, max_hops;
  const int16_t* path_links;
  const int16_t* path_links_by_id; /* may be NULL */
  const int32_t* path_hops;
  const double* path_length;
  const int32_t* path_se;
  int32_t n_mods;
  const int32_t* mod_se;
  const double* mod_max_length;
  const double* mod_reach;
  const double* mod_snr_coeff;
  int32_t num_slots, num_cores, j, episode_length, allow_rejection, auto_reset;
  double channel_width, mean_iat, mean_holding;
  int32_t request_source; /* 0 replay, 1 philox */
  int32_t n_bit_rates;
  uint64_t seed;
  const int32_t* bit_rates;
  const double* bit_rate_cum;
  int32_t bit_rate_lo, bit_rate_hi;
  const double* src_cum;
  const double* dst_cum;
  /* PhyRMSA-v0 (env_kind 3, phy_rmsa_env.py): channels take the place of slots */
  int32_t phy_n_tables;
  /* virtual layer (phy_rmsa_env.py:625-659, 1650-1673): 0 = channel_state not kept (physical layer only, the
   * heuristics never groom), 1 = env.grooming True, 2 = env.grooming False with channel_state kept, which is what
   * the reference does: sapff/sapbm/bmff/faff/faff_rss call use_existing_channels unconditionally (:1256,1321,1510,
   * 1574,1678), bmfa/bmfa_rss only if env.grooming (:1377,1443) */
  int32_t phy_grooming;
  const int32_t* phy_table_id;   /* [n_nodes*n_nodes] first connections_detail row of the unordered pair */
  const uint8_t* phy_mod_level;  /* [n_tables][num_slots][k_paths]  modulation_level[table][channel][idp] */
  const double* phy_gsnr;        /* [n_tables][num_slots][k_paths]  gsnr[table][channel][idp] */
  const int16_t* path_nodes;     /* [n_nodes*n_nodes][k_paths][max_hops+1] node ids, -1 padded */
  const int32_t* edge_nodes;     /* [n_edges][2] */
  /* 1: keep what the info dict of RMSAEnv.step needs (rmsa_env.py:276-332): time-weighted link statistics
   * (_update_link_stats :562-641), per-bit-rate histograms; RMSA-v0 / DeepRMSA-v0 */
  int32_t info_stats, info_pad;
  /* GN-model QoT gate (RMSA-v0 / RMCSA-v0).  The reference has no env that calls calculate_osnr;
   * the gate is what BASELINE.json's north_star describes: the GN-model GSNR of the candidate lightpath and of
   * its neighbours (running services sharing a link and core with it, re-evaluated with the candidate present)
   * feeds the QoT feasibility check - accept iff every one of them meets the SD-FEC threshold of its own
   * modulation (minimum_osnr = target_SNR_dB_vec, examples/create_topology_gn.py:65-110) - and, with
   * qot_adaptive, the modulation-format choice (RMSA-v0: the env assigns the most efficient modulation that
   * passes, spectral efficiencies qot_se[] in descending order; a modulation is tried only if the channel's
   * interference-free GSNR meets its threshold).  Every GSNR is one literal call of calculate_osnr
   * (examples/calculate_osnr.py:9-56). */
  int32_t qot_gate, qot_adaptive, qot_n_se, qot_pad;
  int32_t qot_se[8];
  double qot_thr_db[16];             /* by spectral efficiency */
  const oracle_gsnr_params* qot_gsnr;
  /* PhyRMSA-v0 periodic defragmentation (phy_rmsa_env.py:355-417, 662-764): every phy_defrag_period processed
   * services, grooming moves (_groom_defragmentation) then up to phy_number_moves channel reallocations ranked by the
   * fragmentation gain (phy_metric 0 = "cut", 1 = "rss"); 0 = off */
  int32_t phy_defrag_period, phy_number_moves, phy_metric, phy_pad;
} oracle_config;

typedef struct oracle_env oracle_env;

oracle_env* oracle_create(const oracle_config* cfg, int64_t global_env_id);
void oracle_destroy(oracle_env* e);
/* per-env contiguous trace; stride in requests between consecutive requests */
void oracle_set_trace(oracle_env* e, const oracle_request* trace, int64_t n, int64_t stride);
void oracle_reset(oracle_env* e, int full);
int oracle_action_dim(const oracle_env* e);
int oracle_obs_dim(const oracle_env* e);
/* alloc[4] = accepted, path (| core<<16), initial_slot, n_slots */
void oracle_step(oracle_env* e, const int32_t* action, int32_t* reward, uint8_t* done,
                 double* obs, int32_t* alloc);
void oracle_observation(oracle_env* e, double* obs);
void oracle_matrix_observation(oracle_env* e, uint8_t* obs);
void oracle_heuristic(oracle_env* e, int heuristic_id, int32_t* action);
void oracle_counters(const oracle_env* e, int64_t* out12);
void oracle_state(const oracle_env* e, uint8_t* out); /* [cores][E][S] */
void oracle_request_now(const oracle_env* e, oracle_request* out);

/* n_envs independent envs, nsteps of heuristic+step each, on `threads` host
 * threads; returns total steps.  Used as the timed CPU baseline. */
int64_t oracle_rollout_many(const oracle_config* cfg, int64_t env0, int64_t n_envs,
                            int nsteps, int heuristic_id, int threads,
                            int64_t* counters_out /* [n_envs][12] or NULL */,
                            uint8_t* state_out /* [n_envs][cores*E*S] or NULL */,
                            int64_t* reward_sum_out /* [n_envs] or NULL */);

/* PhyRMSA: actions are int32[ORACLE_PHY_ADIM] = (path, channel_0 .. channel_7), -1 terminated;
 * heuristic ids 5 SAP-FF, 6 SAP-BM, 7 BM-FF, 8 BM-FA-Cut, 10 FA-FF (physical layer, no grooming).
 * stats out8 = total_path_length_episode, total_gsnr_episode, total_path_index_episode,
 * total_modulation_level_episode, channels_accepted_episode, physical_services_accepted_episode,
 * number_cuts_total, rss_total_metric (phy_rmsa_env.py:313-348). */
#define ORACLE_PHY_ADIM 9
void oracle_phy_stats(const oracle_env* e, double* out8);
/* out3 = counted_moves, counted_moves_groom, counted_defrag_cycles (phy_rmsa_env.py:110-112; per episode :433-435) */
void oracle_phy_moves(const oracle_env* e, int64_t* out3);
/* info dict of the LAST step (rmsa_env.py:293-332): service_blocking_rate, episode_service_blocking_rate,
 * bit_rate_blocking_rate, episode_bit_rate_blocking_rate, network_compactness, network_compactness_difference,
 * avg_link_compactness, avg_link_utilization, fairness, then bit_rate_blocking_<b> for every bit rate.
 * Returns the number of values (9 + n_bit_rates).  Needs info_stats = 1. */
int oracle_info(const oracle_env* e, double* out, int max_values);
/* channel_state (phy_rmsa_env.py:158-166) in append order, rows (src*N+dst, idp, channel, used, free, capacity) */
int64_t oracle_phy_channel_state(const oracle_env* e, int32_t* out, int64_t max_entries);

/* query[4] = (path | core << 16, first_slot, n_slots, se or 0); out2 = (candidate dB, min neighbour dB) */
void oracle_gsnr(const oracle_env* e, const oracle_gsnr_params* g, const int32_t* query, double* out2);
int64_t oracle_services(const oracle_env* e, int32_t* out6, int64_t max_rows);
/* QoT gate of the last step / heuristic call: out2 = (candidate GSNR dB, lowest neighbour GSNR dB with the
 * candidate present) of the last candidate evaluated; NaN where nothing was evaluated */
void oracle_qot_stats(long long* out4); /* process-wide: gate calls, candidate misses, neighbours, neighbour hops */
void oracle_qot_last(const oracle_env* e, double* out2);

/* persistent pool of envs stepped with a heuristic on host threads (the envs keep
 * their state between calls); the timed CPU arm of bench.py */
typedef struct oracle_pool oracle_pool;
oracle_pool* oracle_pool_create(const oracle_config* cfg, int64_t env0, int64_t n_envs);
void oracle_pool_destroy(oracle_pool* p);
int64_t oracle_pool_rollout(oracle_pool* p, int nsteps, int heuristic_id, int threads);
oracle_env* oracle_pool_env(oracle_pool* p, int64_t i);
int64_t oracle_pool_reward_sum(const oracle_pool* p);
/* pool rollouts that also evaluate the GN model for every decision's candidate (RMSA-v0 / RMCSA-v0 actions) */
void oracle_pool_set_gsnr(oracle_pool* p, const oracle_gsnr_params* g);
double oracle_pool_gsnr_sum(const oracle_pool* p);

/* deterministic RNG pieces, exposed for tests */
void oracle_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                       uint32_t k1, uint32_t out[4]);
double oracle_dlog(double x);
void oracle_philox_request(const oracle_config* cfg, int64_t global_env_id, int64_t req_index,
                           double now, oracle_request* out);

#ifdef __cplusplus
}
#endif
#endif
