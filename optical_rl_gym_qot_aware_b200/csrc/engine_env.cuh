// engine_env.cuh — per-thread env context: release heap, request draw, decision.
#pragma once
#include "engine.cuh"

namespace orlg {

constexpr int HEAP_D = 4;    // 4-ary implicit min-heap keyed by release time
constexpr int HEAP_PAD = 3;  // logical index i lives in slot i+3 -> child groups are 64B aligned

// payload of a heap entry: what _release_path needs (rmsa_env.py:515-535)
__device__ __forceinline__ unsigned long long pack_service(int pair, int path, int core, int start, int n) {
  return (unsigned long long)pair | ((unsigned long long)path << 16) | ((unsigned long long)core << 24) |
         ((unsigned long long)start << 32) | ((unsigned long long)n << 48);
}

template <int W, int BD>
struct Env {
  const Params& p;
  Tile<W, BD> tile;
  int e;  // local env index
  ulonglong2* heap;
  double now, holding, next_rel;
  int bit_rate, src, dst, heap_len;
  long long cnt[CNT_N];

  __device__ __forceinline__ Env(const Params& p_, uint32_t* smem, int tid, int e_)
      : p(p_), tile{smem, tid, p_.E}, e(e_) {
    heap = p.heap + (size_t)e * p.heap_stride + HEAP_PAD;
  }
  __device__ __forceinline__ int pair() const { return src * p.n_nodes + dst; }

  // ---- scalars <-> HBM (SoA, coalesced) ----
  __device__ __forceinline__ void load_scalars() {
    now = p.now[e];
    holding = p.holding[e];
    uint32_t m = p.meta[e];
    bit_rate = (int)(m & 0xFFFFu);
    src = (int)((m >> 16) & 0xFFu);
    dst = (int)(m >> 24);
    next_rel = p.next_rel[e];
    heap_len = p.heap_len[e];
#pragma unroll
    for (int i = 0; i < CNT_N; i++) cnt[i] = p.counters[(size_t)i * p.NP + e];
  }
  __device__ __forceinline__ void store_scalars() {
    p.now[e] = now;
    p.holding[e] = holding;
    p.meta[e] = (uint32_t)bit_rate | ((uint32_t)src << 16) | ((uint32_t)dst << 24);
    p.next_rel[e] = next_rel;
    p.heap_len[e] = heap_len;
#pragma unroll
    for (int i = 0; i < CNT_N; i++) p.counters[(size_t)i * p.NP + e] = cnt[i];
  }

  // ---- release queue: heapq of (arrival+holding, service), optical_network_env.py:178-189 ----
  __device__ __forceinline__ double key_of(const ulonglong2& v) const {
    return __longlong_as_double((long long)v.x);
  }
  __device__ __forceinline__ void heap_push(double key, unsigned long long payload) {
    if (heap_len >= p.heap_cap) {  // cannot happen with the provable bound; sticky fault
      cnt[CNT_FAULT] |= 1;
      return;
    }
    int i = heap_len++;
    while (i > 0) {
      int parent = (i - 1) / HEAP_D;
      ulonglong2 pv = heap[parent];
      if (key < key_of(pv)) {
        heap[i] = pv;
        i = parent;
      } else {
        break;
      }
    }
    heap[i] = make_ulonglong2((unsigned long long)__double_as_longlong(key), payload);
    if (i == 0) next_rel = key;
  }
  // removes the minimum and returns its payload
  __device__ __forceinline__ unsigned long long heap_pop() {
    ulonglong2 root = heap[0];
    ulonglong2 last = heap[--heap_len];
    int n = heap_len;
    if (n > 0) {
      double lk = key_of(last);
      int i = 0;
      for (;;) {
        int c = HEAP_D * i + 1;
        if (c >= n) break;
        ulonglong2 best = heap[c];
        int bi = c;
#pragma unroll
        for (int d = 1; d < HEAP_D; d++) {
          if (c + d < n) {
            ulonglong2 v = heap[c + d];
            if (key_of(v) < key_of(best)) {
              best = v;
              bi = c + d;
            }
          }
        }
        if (key_of(best) < lk) {
          heap[i] = best;
          i = bi;
        } else {
          break;
        }
      }
      heap[i] = last;
      next_rel = (i == 0) ? lk : key_of(heap[0]);
    } else {
      next_rel = __longlong_as_double(0x7FF0000000000000ll);  // +inf
    }
    return root.y;
  }

  // ---- _next_service rmsa_env.py:643-695 / rmcsa_env.py:690-739 ----
  __device__ __forceinline__ void next_service() {
    long long r = cnt[CNT_DRAWN];
    if (p.request_source == ORLG_SRC_REPLAY) {
      long long rr = r < p.trace_len ? r : p.trace_len - 1;
      if (r >= p.trace_len) cnt[CNT_FAULT] |= 2;  // trace exhausted
      const orlg_request* q = p.trace + (size_t)rr * p.num_envs + e;
      now = q->arrival;
      holding = q->holding;
      bit_rate = q->bit_rate;
      src = q->src;
      dst = q->dst;
    } else {
      unsigned long long g = (unsigned long long)(p.genv0 + e);
      uint2 key = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
      uint4 R0 = philox4x32_10(make_uint4((uint32_t)r, (uint32_t)((unsigned long long)r >> 32), (uint32_t)g, 0u), key);
      uint4 R1 = philox4x32_10(make_uint4((uint32_t)r, (uint32_t)((unsigned long long)r >> 32), (uint32_t)g, 1u), key);
      double lam_a = 1.0 / p.mean_iat, lam_h = 1.0 / p.mean_holding;
      now = now + (-det_log(1.0 - res53(R0.x, R0.y)) / lam_a);
      holding = -det_log(1.0 - res53(R0.z, R0.w)) / lam_h;
      const double inv32 = 1.0 / 4294967296.0;
      int n = p.n_nodes;
      src = bisect_right(p.src_cum, ((double)R1.x * inv32) * __ldg(p.src_cum + n - 1), n - 1);
      const double* dc = p.dst_cum + (size_t)src * n;
      dst = bisect_right(dc, ((double)R1.y * inv32) * __ldg(dc + n - 1), n - 1);
      double ub = (double)R1.z * inv32;
      if (p.n_bit_rates > 0) {
        int nb = p.n_bit_rates;
        bit_rate = __ldg(p.bit_rates + bisect_right(p.br_cum, ub * __ldg(p.br_cum + nb - 1), nb - 1));
      } else {
        bit_rate = p.br_lo + (int)floor(ub * (double)(p.br_hi - p.br_lo + 1));
      }
    }
    cnt[CNT_DRAWN] = r + 1;
    if (p.kind != ORLG_ENV_RMCSA) {  // rmsa_env.py:673-674
      cnt[CNT_PROC]++;
      cnt[CNT_EP_PROC]++;
    }
    cnt[CNT_BR_REQ] += bit_rate;  // rmsa_env.py:677-678, rmcsa_env.py:730-731
    cnt[CNT_EP_BR_REQ] += bit_rate;
    while (next_rel <= now) {  // rmsa_env.py:689-695
      unsigned long long s = heap_pop();
      PathRec rec = load_path(p, (int)(s & 0xFFFF), (int)((s >> 16) & 0xFF));
      tile.set_range(rec, (int)((s >> 24) & 0xFF), (int)((s >> 32) & 0xFFFF), (int)(s >> 48));
    }
  }

  // counters-only reset, rmsa_env.py:344-389 (the pending service is re-counted)
  __device__ __forceinline__ void reset_episode() {
    cnt[CNT_EP_ACC] = 0;
    cnt[CNT_EP_BR_PROV] = 0;
    cnt[CNT_EP_PROC] = 1;
    cnt[CNT_EP_BR_REQ] = bit_rate;
  }

  // _provision_path + _add_release (state and counters only)
  __device__ __forceinline__ void provision(const PathRec& rec, int path, int core, int start, int n) {
    tile.clear_range(rec, core, start, n);
    cnt[CNT_ACC]++;
    cnt[CNT_EP_ACC]++;
    cnt[CNT_BR_PROV] += bit_rate;
    cnt[CNT_EP_BR_PROV] += bit_rate;
    heap_push(now + holding, pack_service(pair(), path, core, start, n));
  }

  // _crosstalk_is_acceptable rmcsa_env.py:341-384 (modulation-only factors precomputed on host)
  __device__ __forceinline__ bool crosstalk_ok(int mod, double path_length) const {
    double denom = __ldg(p.mod_coeff + mod) * ((double)bit_rate / (double)__ldg(p.mod_se + mod));
    denom = denom * 1e9;
    double lmax_snr = 100.0 / denom;
    lmax_snr = lmax_snr / 1000.0;
    return path_length < __ldg(p.mod_reach + mod) && path_length < lmax_snr;
  }
};

}  // namespace orlg
