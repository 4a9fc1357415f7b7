// engine_util.cuh — small non-template kernels behind the read-only ABI calls
// (orlg_get_counters, orlg_export_state, orlg_get_requests).  Included by orlg.cu only.
#pragma once
#include "engine_env.cuh"

namespace orlg {

// ---- small utility kernels ----------------------------------------------------
// counters [CNT_N][NP] (+ heap_len) -> [num_envs][ORLG_NUM_COUNTERS]
__global__ void counters_kernel(const Params p, long long* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  long long* o = out + (size_t)e * ORLG_NUM_COUNTERS;
  auto c = [&](int i) { return p.counters[(size_t)i * p.NP + e]; };
  o[ORLG_C_SERVICES_PROCESSED] = c(CNT_PROC);
  o[ORLG_C_SERVICES_ACCEPTED] = c(CNT_ACC);
  o[ORLG_C_EP_SERVICES_PROCESSED] = c(CNT_EP_PROC);
  o[ORLG_C_EP_SERVICES_ACCEPTED] = c(CNT_EP_ACC);
  o[ORLG_C_BIT_RATE_REQUESTED] = c(CNT_BR_REQ);
  o[ORLG_C_BIT_RATE_PROVISIONED] = c(CNT_BR_PROV);
  o[ORLG_C_EP_BIT_RATE_REQUESTED] = c(CNT_EP_BR_REQ);
  o[ORLG_C_EP_BIT_RATE_PROVISIONED] = c(CNT_EP_BR_PROV);
  o[ORLG_C_REQUESTS_DRAWN] = c(CNT_DRAWN);
  o[ORLG_C_QUEUE_LEN] = p.heap_len[e];
  o[ORLG_C_FAULT] = c(CNT_FAULT);
  o[ORLG_C_EPISODES] = c(CNT_EPISODES);
}

// packed tiles -> uint8 [num_envs][cores][E][S]
__global__ void export_state_kernel(const Params p, int W, int BD, unsigned char* out) {
  long long cells = (long long)p.cores * p.E * p.S;
  long long total = cells * p.num_envs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long e = i / cells, c = i % cells;
    int row = (int)(c / p.S), s = (int)(c % p.S);
    size_t w = ((size_t)(e / BD) * p.planes + (size_t)row * W + (s >> 5)) * BD + (e % BD);
    out[i] = (unsigned char)((p.occ[w] >> (s & 31)) & 1u);
  }
}

__global__ void requests_kernel(const Params p, orlg_request* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  uint32_t m = p.meta[e];
  orlg_request r;
  r.arrival = p.now[e];
  r.holding = p.holding[e];
  r.bit_rate = (int)(m & 0xFFFFu);
  r.src = (int16_t)((m >> 16) & 0xFFu);
  r.dst = (int16_t)(m >> 24);
  out[e] = r;
}


// running services of every env = the entries of its release queue, in queue order: rows of
// (src_id * n_nodes + dst_id, path, core, first slot, slots, spectral efficiency) - what the reference keeps as
// topology.graph["running_services"] (rmsa_env.py:484-487); the owner map spectrum_slots_allocation (:481-483)
// follows from it.  PhyRMSA: one row per (service, channel): (pair, path, share, channel, position, level).
__global__ void services_kernel(const Params p, int* out, int max_rows, int* count) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  const ulonglong2* heap = p.heap + (size_t)e * p.heap_stride + HEAP_PAD;
  const int n = p.heap_len[e];
  if (p.kind == ORLG_ENV_PHYRMSA) {  // one queue entry per service: its channels are in the service's record
    int rows = 0;
    for (int i = 0; i < n; i++) {
      const uint4* r = p.phy_svc + ((size_t)e * p.phy_svc_cap + (size_t)heap[i].y) * 3;
      const uint4 a = r[0], b = r[1], c = r[2];
      const int nch = (int)(a.z >> 24);
      const uint32_t cw[4] = {b.x, b.y, b.z, b.w}, sw[2] = {c.x, c.y}, pw[2] = {c.z, c.w};
      for (int q = 0; q < nch; q++, rows++) {
        if (rows >= max_rows) continue;
        int* o = out + ((size_t)e * max_rows + rows) * 6;
        o[0] = (int)(a.z & 0xFFFFu);
        o[1] = (int)((a.z >> 16) & 0xFFu);
        o[2] = (int)((sw[q >> 2] >> (8 * (q & 3))) & 0xFFu);
        o[3] = (int)((cw[q >> 1] >> (16 * (q & 1))) & 0xFFFFu);
        o[4] = q;
        o[5] = (int)((pw[q >> 2] >> (8 * (q & 3))) & 0xFFu);
      }
    }
    count[e] = rows;
    return;
  }
  for (int i = 0; i < n && i < max_rows; i++) {
    const unsigned long long s = heap[i].y;
    int* o = out + ((size_t)e * max_rows + i) * 6;
    o[0] = svc_pair(s);
    o[1] = svc_path(s);
    o[2] = svc_core(s);
    o[3] = svc_start(s);
    o[4] = svc_slots(s);
    o[5] = svc_se(s);
  }
  count[e] = n;
}

__global__ void phy_moves_kernel(const Params p, long long* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  for (int i = 0; i < 3; i++) out[(size_t)e * 3 + i] = p.phy_moves ? p.phy_moves[(size_t)i * p.NP + e] : 0;
}

// PhyRMSA-v0 channel_state without the tombstones, thread per env
__global__ void phy_channel_state_kernel(const Params p, int* out, int max_entries, int* count) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  const int len = (int)(p.phy_cs_len[e] & 0xFFFFu);
  int n = 0;
  for (int i = 0; i < len; i++) {
    const uint2 v = p.phy_cs[(size_t)i * p.NP + e];
    if (v.x == 0xFFFFFFFFu) continue;
    if (n < max_entries) {
      int* o = out + ((size_t)e * max_entries + n) * 6;
      o[0] = (int)(v.x & 0xFFFFu);
      o[1] = (int)((v.x >> 16) & 15u);
      o[2] = (int)(v.x >> 20);
      o[3] = (int)(v.y & 0xFFu);
      o[4] = (int)((v.y >> 8) & 0xFFu);
      o[5] = (int)(v.y >> 16);
    }
    n++;
  }
  count[e] = n;
}

// PhyRMSA-v0 statistics, one warp per env: episode sums + fragmentation metrics of the current
// channel map (phy_rmsa_env.py:1195-1203, 1110-1121).  Lanes take channels round-robin; the
// per-channel terms are reduced with warp shuffles.  (The sums over channels are therefore
// reassociated: number_cuts_total is an exact integer ratio, rss_total_metric agrees with the
// reference's sequential f64 sum to ~1e-15 relative.)
__global__ void __launch_bounds__(128) phy_stats_kernel(const Params p, int W, int BD, double* out) {
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (e >= p.num_envs) return;
  const uint32_t* tile = p.occ + (size_t)(e / BD) * p.planes * BD + (e % BD);
  long long cuts = 0;
  double rss = 0.0;
  for (int ch = lane; ch < p.S; ch += 32) {
    const int wi = ch >> 5, b = ch & 31;
    long long sum_sq = 0;
    int sum = 0, run = 0, prev = 0;
    for (int l = 0; l < p.E; l++) {
      int bit = (tile[(size_t)(l * W + wi) * BD] >> b) & 1u;
      cuts += bit & ~prev;
      prev = bit;
      if (bit) {
        run++;
      } else {
        sum_sq += (long long)run * run;
        sum += run;
        run = 0;
      }
    }
    sum_sq += (long long)run * run;
    sum += run;
    rss += sqrt((double)sum_sq) / (double)(sum + 1);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cuts += __shfl_xor_sync(0xFFFFFFFFu, cuts, o);
    rss += __shfl_xor_sync(0xFFFFFFFFu, rss, o);
  }
  if (lane == 0) {
    double* o = out + (size_t)e * 8;
    o[0] = p.phy_f[e];
    o[1] = p.phy_f[(size_t)p.NP + e];
    for (int i = 0; i < 4; i++) o[2 + i] = (double)p.phy_i[(size_t)i * p.NP + e];
    o[6] = (double)cuts / (double)p.S;
    o[7] = rss / (double)p.S;
  }
}

}  // namespace orlg
