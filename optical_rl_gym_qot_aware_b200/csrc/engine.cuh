// engine.cuh — device-side env stepping for RMSA-v0 / DeepRMSA-v0 / RMCSA-v0.
//
// Execution model: one thread owns one env for the whole launch (a warp owns 32
// envs).  A CTA of BD threads owns the tile of BD consecutive envs; the tile's
// packed occupancy lives in HBM as occ[tile][plane][BD] (plane = (core*E+link)*W
// + word), i.e. exactly the shared-memory image, so one TMA bulk copy
// (cp.async.bulk, mbarrier completion) stages it and one bulk copy writes it
// back.  In shared memory word (plane, tid) sits at plane*BD+tid: lanes hit
// consecutive banks whatever plane each lane addresses, so the per-env gathers
// over path links are conflict free.
//
// Reference functions restated here (paths under optical_rl_gym/envs/):
//   is_path_free            rmsa_env.py:721-734, rmcsa_env.py:767-794
//   get_number_slots        rmsa_env.py:708-719, rmcsa_env.py:753-765
//   _provision/_release     rmsa_env.py:462-535, rmcsa_env.py:488-558
//   get_available_blocks    rmsa_env.py:774-804
//   _next_service           rmsa_env.py:643-695, rmcsa_env.py:690-739
//   step                    rmsa_env.py:222-341, deeprmsa_env.py:48-58, rmcsa_env.py:209-339
//   observation             deeprmsa_env.py:60-121
//   heuristics              rmsa_env.py:854-937, deeprmsa_env.py:135-155, rmcsa_env.py:882-911
#pragma once
#include <cstdint>

#include "../../include/orlg.h"
#include "bitrow.cuh"
#include "rng.cuh"

namespace orlg {

enum Mode {
  MODE_STEP = 0,        // actions from a device tensor (gym / VecEnv step)
  MODE_HEURISTIC = 1,   // only compute the heuristic's action
  MODE_ROLLOUT = 2,     // nsteps x (heuristic -> step) with state kept on chip
  MODE_OBS = 3,         // only write the observation
  MODE_RESET_FULL = 4,  // constructor-time reset
  MODE_RESET_EPISODE = 5
};

// internal per-env counters, SoA [CNT_N][NP]
enum { CNT_PROC = 0, CNT_ACC, CNT_EP_PROC, CNT_EP_ACC, CNT_BR_REQ, CNT_BR_PROV, CNT_EP_BR_REQ,
       CNT_EP_BR_PROV, CNT_DRAWN, CNT_EPISODES, CNT_FAULT, CNT_N };

struct Params {
  int kind, num_envs, NP, n_nodes, E, k, S, cores, j, n_mods;
  int planes, rec_u4, episode_length, allow_rejection, auto_reset, request_source;
  int n_bit_rates, br_lo, br_hi, heap_cap, heap_stride, action_dim;
  double cw, mean_iat, mean_holding;
  int cw_x2;                 // 2*channel_width when that is a small integer (12.5 GHz -> 25), else 0
  unsigned slot_magic[16];   // floor(2^32 / (cw_x2*se)) + 1: exact division by multiplication for se < 16
  const double* obs_pos;     // [S+1]  2*(x - 0.5*S)/S   (deeprmsa_env.py:85-87,101)
  const double* obs_slots;   // [S+2]  (n - 5.5)/3.5     (deeprmsa_env.py:97)
  unsigned long long seed;
  long long genv0;
  const unsigned char* hot_tab;  // path records + request distributions in one allocation of hot_tab_lines * 128 B
  int hot_tab_lines;
  const uint4* path_rec;  // [N*N*k][rec_u4]: byte0 hops, 1 se, 2 best modulation, 3.. link index per hop
  const double* path_length;
  const int* mod_se;
  const double* mod_reach;
  const double* mod_coeff;
  const int* bit_rates;
  const double* br_cum;
  const double* src_cum;
  const double* dst_cum;
  const unsigned char* guide;  // [2 + N][64] bisect guide tables: src, bit rate, dst given src (rng.cuh)
  uint32_t* occ;
  double* now;
  double* holding;
  uint32_t* meta;  // bit_rate | src << 16 | dst << 24
  double* next_rel;
  int* heap_len;
  ulonglong2* heap;  // [NP][heap_stride] {key bits, payload}
  long long* counters;
  const orlg_request* trace;
  long long trace_len;
  // PhyRMSA-v0 (phy_rmsa_env.py): channels take the place of slots, one core
  const int* phy_table_id;              // [N*N] row of the QoT tables for the unordered pair
  const unsigned char* phy_mod;         // [T][S][k] modulation_level
  const double* phy_gsnr;               // [T][S][k] gsnr
  const uint32_t* phy_level_mask;       // [T][k][PHY_LEVELS][W] channels whose level == m
  const unsigned char* phy_cut_links;   // [N*N*k][phy_max_terms] neighbour link of every r_cut term
  const int* phy_cut_count;             // [N*N*k]
  int phy_max_terms, phy_mask_w;
  // info dict of RMSAEnv.step (rmsa_env.py:276-332), RMSA-v0 / DeepRMSA-v0 with info_stats
  int info_stats;
  double* lnk_stat;       // [3][E][NP] topology[n1][n2]["last_update" | "utilization" | "compactness"] (:562-641)
  int* br_hist;           // [2][n_bit_rates][NP] bit_rate_requested_histogram / bit_rate_provisioned_histogram
  long long* slot_hops;   // [NP] sum of number_slots * hops over the running services (:809-812)
  // virtual layer (phy_rmsa_env.py:625-659, 1650-1673): channel_state of every env as one list in append order,
  // entry i of env e at phy_cs[i * NP + e] = {pair | idp << 16 | channel << 20, used | free << 8 | capacity << 16}
  int phy_grooming;                     // 0 off, 1 env.grooming, 2 channel_state kept with env.grooming False
  int phy_cs_cap;
  uint2* phy_cs;
  uint32_t* phy_cs_len;                 // [NP] entries in use | tombstones << 16
  double* phy_f;                        // [2][NP] episode sums: path length, gsnr
  long long* phy_i;                     // [4][NP] episode sums: path index, modulation level, channels, physical services
  // PhyRMSA service table (engine_env.cuh PhySvc): one 48-byte record per running service
  uint4* phy_svc;                       // [NP][phy_svc_cap][3]
  unsigned short* phy_free;             // [NP][phy_svc_cap] stack of free rows
  int* phy_free_n;                      // [NP]
  int phy_svc_cap;
  // periodic defragmentation (phy_rmsa_env.py:355-417): order list of the running services, move counters, candidates
  int phy_defrag_period, phy_number_moves, phy_metric, phy_run_cap, phy_cand_cap;
  unsigned short* phy_run;              // [NP][phy_run_cap] running_services in list order, 0xFFFF = removed
  uint32_t* phy_run_len;                // [NP] slots in use | tombstones << 16
  long long* phy_moves;                 // [3][NP] counted_moves, counted_moves_groom, counted_defrag_cycles
  double2* phy_cand;                    // [NP][phy_cand_cap] (gain, age) of the defragmentation candidates
  uint32_t* phy_cand_id;                // [NP][phy_cand_cap] service row | channel << 16
  // GN-model QoT gate (qot.cuh): RMSA-v0 / RMCSA-v0, env-major occupancy
  int qot, qot_adaptive, qot_n_se, qot_hops, qot_slot_cap;
  int qot_se[8];                        // adaptive RMSA-v0: spectral efficiencies tried, descending
  int qot_mod_order[16];                // adaptive RMCSA-v0: modulation rows by descending spectral efficiency (stable)
  double qot_thr[16];                   // GSNR threshold in dB by spectral efficiency
  double qot_thr_lin[16];               // the same as the largest acceptable 1 / GSNR (linear): 10^(-thr / 10)
  longlong2* qot_sums;                  // [NP][qot_slot_cap][qot_hops] running (S_A, S_X) of every service row, 2^-44 fixed point
  unsigned short* qot_free;             // [NP][qot_slot_cap] stack of free rows
  int* qot_free_n;                      // [NP]
  unsigned long long* qot_rel;          // [NP][QOT_REL_CAP] payloads of the services the last step released
  int* qot_rel_n;                       // [NP]
  double* qot_gp;                       // [NP][2] GSNR pair of a heuristic launch's proposal (qot.cuh)
  unsigned long long* qot_add;          // [NP] payload of the lightpath the last qot_kernel launch accepted (0: none); it enters its
                                        // neighbours' sums at the start of the next launch
  int4* qot_verdict;                    // [NP][2] (accept, path | core << 16, first slot, slots), (se, row, modulation, 0)
  double* qot_g;                        // [NP][2] 1 / GSNR (linear) of the last candidate evaluated and of its worst neighbour with it present; < 0 = none
  const double2* qot_chan;              // [S + 1] per slot count: (B h NF / P, (P / B)^3 8 / (27 pi |b2|) g^2 B / P)
};
constexpr int PHY_LEVELS = 8;   // modulation levels 1..8 (100 Gb/s per level per channel, phy_rmsa_env.py:1424)
constexpr int PHY_ADIM = 9;     // action = (path, 8 channel numbers, -1 terminated)

struct StepIO {
  int mode, heuristic, nsteps, obs_kind;
  const int* actions;
  int* actions_out;
  int* reward;
  unsigned char* done;
  void* obs;
  long long obs_es, obs_fs;  // element strides: env, feature
  int* alloc;
  long long* reward_sum;
  double* info;        // optional f64 [num_envs][9 + n_bit_rates]: the step's info dict (needs info_stats)
  long long* snap;     // optional int64 [num_envs][8]: the first eight counters at the info dict's point in the step
  int* hint_actions;   // optional: the heuristic's action for the NEW pending request (fused next-action hint)
  int hint_heuristic;
};

// ---- path records ----------------------------------------------------------
// 16 B (hops <= 12) or 32 B records: word 0 = hops | se << 8 | best modulation << 16, then one
// byte per hop (edge "index"), four hops per word.
#ifndef ORLG_LINK_UNROLL
#define ORLG_LINK_UNROLL 1
#endif

struct PathRec {
  uint32_t w[8];
  __device__ __forceinline__ int hops() const { return (int)(w[0] & 0xFF); }
  __device__ __forceinline__ int se() const { return (int)((w[0] >> 8) & 0xFF); }
  __device__ __forceinline__ int best_mod() const { return (int)((w[0] >> 16) & 0xFF); }
  __device__ __forceinline__ int link(int h) const {  // random access (cold paths only)
    const int g = (h >> 2) + 1;
    uint32_t v = w[1];
#pragma unroll
    for (int i = 2; i < 8; i++) v = (g == i) ? w[i] : v;
    return (int)((v >> ((h & 3) * 8)) & 0xFFu);
  }
};

// f(link) for every hop in order; hops are unrolled with static byte positions
template <typename F>
__device__ __forceinline__ void for_each_link(const PathRec& r, F f) {
  const int h = r.hops();
#if ORLG_LINK_UNROLL
#pragma unroll
  for (int g = 0; g < 3; g++) {
    if (4 * g < h) {
      const uint32_t wd = r.w[g + 1];
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (4 * g + u < h) f((int)((wd >> (8 * u)) & 0xFFu));
    }
  }
  for (int i = 12; i < h; i++) f(r.link(i));  // long paths (32-byte records)
#else
  // one copy of the four-hop body; the group's word is picked by selects (no indexed registers)
#pragma unroll 1
  for (int g = 0; 4 * g < h; g++) {
    uint32_t wd = r.w[1];
#pragma unroll
    for (int i = 2; i < 8; i++) wd = (g + 1 == i) ? r.w[i] : wd;
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (4 * g + u < h) f((int)((wd >> (8 * u)) & 0xFFu));
  }
#endif
}

__device__ __forceinline__ PathRec load_path(const Params& p, int pair, int path) {
  const uint4* a = p.path_rec + ((size_t)pair * p.k + path) * p.rec_u4;
  PathRec r;
  uint4 v = __ldg(a);
  r.w[0] = v.x;
  r.w[1] = v.y;
  r.w[2] = v.z;
  r.w[3] = v.w;
  r.w[4] = r.w[5] = r.w[6] = r.w[7] = 0u;
  if (p.rec_u4 > 1) {
    uint4 u = __ldg(a + 1);
    r.w[4] = u.x;
    r.w[5] = u.y;
    r.w[6] = u.z;
    r.w[7] = u.w;
  }
  return r;
}

// ---- occupancy tile in shared memory ----------------------------------------
template <int W, int BD>
struct Tile {
  uint32_t* s;  // [planes][BD]: shared-memory tile, or (BD == 1) this env's rows in HBM
  int tid, E;
  __device__ __forceinline__ uint32_t& word(int core, int link, int wi) {
    return s[(((core * E) + link) * W + wi) * BD + tid];
  }
  // AND of the rows of every link of the path (core plane `core`).  Not inlined: one copy of the
  // unrolled hop loop serves every call site.
  __device__ __forceinline__ Row<W> path_avail(const PathRec& r, int core) {
    Row<W> a = row_ones<W>();
    const uint32_t* base = s + (size_t)core * E * W * BD + tid;
    for_each_link(r, [&](int l) {
      const uint32_t* q = base + l * (W * BD);
      if constexpr (BD == 1 && W % 4 == 0) {  // env-major rows in HBM: 16-byte aligned
#pragma unroll
        for (int wi = 0; wi < W; wi += 4) {
          uint4 v = *reinterpret_cast<const uint4*>(q + wi);
          a.w[wi] &= v.x;
          a.w[wi + 1] &= v.y;
          a.w[wi + 2] &= v.z;
          a.w[wi + 3] &= v.w;
        }
      } else if constexpr (BD == 1 && W % 2 == 0) {  // 8-byte aligned (W = 2, 10)
#pragma unroll
        for (int wi = 0; wi < W; wi += 2) {
          uint2 v = *reinterpret_cast<const uint2*>(q + wi);
          a.w[wi] &= v.x;
          a.w[wi + 1] &= v.y;
        }
      } else {
#pragma unroll
        for (int wi = 0; wi < W; wi++) a.w[wi] &= q[wi * BD];
      }
    });
    return a;
  }
  // clear (provision) or set (release) slots [start, start+n) on every link of the path
  __device__ __forceinline__ void apply_range(const PathRec& r, int core, int start, int n, bool set) {
    uint32_t m[W];
#pragma unroll
    for (int wi = 0; wi < W; wi++) m[wi] = mask_word(wi, start, start + n);
    uint32_t* base = s + (size_t)core * E * W * BD + tid;
    for_each_link(r, [&](int l) {
      uint32_t* q = base + l * (W * BD);
#pragma unroll
      for (int wi = 0; wi < W; wi++)
        if (m[wi]) {
          if constexpr (BD == 1) {  // rows in HBM: a reduction that returns nothing, so no hop waits for its row
            if (set) atomicOr(q + wi, m[wi]);
            else atomicAnd(q + wi, ~m[wi]);
          } else {
            q[wi * BD] = set ? (q[wi * BD] | m[wi]) : (q[wi * BD] & ~m[wi]);
          }
        }
    });
  }
  // one slot / channel on every link of the path (PhyRMSA allocates whole channels): one word per link
  __device__ __forceinline__ void apply_bit(const PathRec& r, int core, int pos, bool set) {
    const uint32_t m = 1u << (pos & 31);
    uint32_t* base = s + (size_t)core * E * W * BD + tid + (pos >> 5) * BD;
    for_each_link(r, [&](int l) {
      uint32_t* q = base + l * (W * BD);
      if constexpr (BD == 1) {
        if (set) atomicOr(q, m);
        else atomicAnd(q, ~m);
      } else {
        *q = set ? (*q | m) : (*q & ~m);
      }
    });
  }
  __device__ __forceinline__ void clear_range(const PathRec& r, int core, int start, int n) {
    apply_range(r, core, start, n, false);
  }
  __device__ __forceinline__ void set_range(const PathRec& r, int core, int start, int n) {
    apply_range(r, core, start, n, true);
  }
};

// get_number_slots: ceil(bit_rate / (SE * channel_width)) + 1.  Python evaluates it in f64; when
// 2*channel_width is a small integer the quotient's distance to the next integer is >= 1/(2cw*SE),
// far above an ulp, so the integer form ceil(2*br / (2cw*SE)) is exactly the same number.
__device__ __forceinline__ int number_slots(int bit_rate, int se, const Params& p) {
  if (p.cw_x2 > 0 && se < 16) {
    // floor(x / d) == umulhi(x, floor(2^32/d) + 1) for x * d < 2^32 (x <= 2*65535 + d, d <= 2*cw*15)
    const unsigned d = (unsigned)(p.cw_x2 * se);
    return (int)__umulhi(2u * (unsigned)bit_rate + d - 1u, p.slot_magic[se]) + 1;
  }
  return (int)ceil((double)bit_rate / ((double)se * p.cw)) + 1;
}

// first start in [0, S - n) with n free slots (exclusive bound: SURVEY.md F8), -1 if none
template <int W>
__device__ __forceinline__ int first_fit(const Row<W>& avail, int n, int S) {
  if (n >= S) return -1;
  Row<W> b = window_starts<W>(avail, n);
#pragma unroll
  for (int i = 0; i < W; i++) b.w[i] &= mask_word(i, 0, S - n);
  return lowest_set<W>(b);
}

}  // namespace orlg
