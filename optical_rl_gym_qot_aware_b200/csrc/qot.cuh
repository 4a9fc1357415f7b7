// qot.cuh — GN-model QoT gate of RMSA-v0 / RMCSA-v0 (include/orlg.h "QoT gate"), one warp per env.
//
// The gate: a candidate lightpath is accepted iff its GSNR (calculate_osnr, examples/calculate_osnr.py:9-56)
// meets the SD-FEC threshold of its modulation (minimum_osnr of examples/create_topology_gn.py:65-110) AND the
// GSNR of every neighbour — a running service that shares a link and core with it — re-evaluated with the
// candidate present still meets its own.  With qot_adaptive the modulation is chosen by the same check.
//
// Cost model.  Literally, calculate_osnr re-sums every interferer of every neighbour: ~neighbours x interferers
// asinh pairs per decision (what gsnr_kernel_fast does, 15 ms per 65 536 decisions on the RMCSA workload).  Adding
// or removing ONE lightpath changes the cross-channel sum of a neighbour by exactly one term per shared link, so
// the engine keeps, for every running service and every hop of its path, the running sums
//     S_A = sum_i [asinh(..(df + B_i/2)) - asinh(..(df - B_i/2))]      (calculate_osnr.py:36-41)
//     S_X = sum_i Phi(SE_i) (B_i / |df|) 5/3                            (:43-45, without the per-link l_eff / L)
// in a per-env service table next to the release queue (the queue payload carries the service's row).  A decision
// is then one pass over the env's queue: O(services) membership tests, O(neighbours) table lookups.
//
// The sums are 64-bit fixed point (2^-44): integer addition is associative, so provision / release in any order
// leave exactly the sums a from-scratch evaluation would give (qot_check_kernel verifies that bit for bit) and
// there is no drift however long an env runs.  Quantisation is ~6e-14 per term against O(1) sums: ~1e-12 dB.
#pragma once
#include "gsnr.cuh"

namespace orlg {

enum { QOT_VERDICT = 0, QOT_HEUR_VERDICT = 1, QOT_HEURISTIC = 2 };
#ifndef QOT_WARPS_N  // envs (warps) per CTA (A/B knob): a one-warp CTA gives its SM slot back the moment its env is done
#define QOT_WARPS_N 1
#endif
constexpr int QOT_WARPS = QOT_WARPS_N;
constexpr int QOT_REL_CAP = 32;  // services released by one step (overflow: sticky fault bit 16)
static_assert(QOT_REL_CAP == 32, "qot_kernel keeps one released payload per lane");
constexpr double QOT_SCALE = 17592186044416.0;  // 2^44
constexpr double QOT_INV_SCALE = 1.0 / 17592186044416.0;

struct QotIO {
  int mode;
  const int* actions;  // QOT_VERDICT: the step's action tensor
  int* actions_out;    // heuristic modes: the action chosen (may be null in QOT_HEUR_VERDICT)
  int use_proposal;    // QOT_VERDICT: the last launch was QOT_HEURISTIC on this very state - an action equal to its
                       // proposal takes that launch's verdict (and the sums it left in the next free row)
};

#ifndef QOT_MIN_CTAS  // resident CTAs per SM the register allocation must allow (A/B knob)
#define QOT_MIN_CTAS 32
#endif
// QOT_CAP = running services staged per window (an env with more is walked window by window); the kernel is
// instantiated for 64 / 128 / 256 and the launch picks the smallest that covers the offered load + 1 sigma, because
// shared memory per CTA is what bounds the resident warps here
template <int QOT_CAP, bool NARROW>
struct QotWarpSmem {
  unsigned long long pay[QOT_CAP];      // release-queue payloads of the staged window
  // their paths as link sets, staged when the topology has at most 64 links (NARROW) and the window is small; wider
  // sets and the 256-entry window (where the shared memory is worth more as L1) take them from the path table
  static constexpr bool STAGED = NARROW && QOT_CAP <= 128;
  unsigned long long mask[STAGED ? QOT_CAP : 1];
  // the candidate's own sums, per hop, as (low, high) 32-bit halves: shared memory has native 32-bit adds only (a
  // 64-bit atomicAdd is a compare-and-swap loop); the carry out of the low half is seen by the add that causes it
  unsigned cSA_lo[32], cSA_hi[32], cSX_lo[32], cSX_hi[32];
  __device__ __forceinline__ static void add64(unsigned* lo, unsigned* hi, unsigned long long x) {
    const unsigned xl = (unsigned)x, xh = (unsigned)(x >> 32);
    const unsigned old = atomicAdd(lo, xl);
    const unsigned up = xh + ((unsigned)(old + xl) < xl ? 1u : 0u);
    if (up) atomicAdd(hi, up);
  }
  __device__ __forceinline__ long long SA(int h) const { return (long long)(((unsigned long long)cSA_hi[h] << 32) | cSA_lo[h]); }
  __device__ __forceinline__ long long SX(int h) const { return (long long)(((unsigned long long)cSX_hi[h] << 32) | cSX_lo[h]); }
  unsigned short nb[QOT_CAP];           // compacted neighbour list of the current pass
  unsigned char hop_of[128];            // link -> hop of the candidate
};

__device__ __forceinline__ double phi_mod(int se) {  // phi_modulation_format[se - 1], calculate_osnr.py:20
  return se <= 2 ? 1.0 : se == 3 ? 2.0 / 3.0 : se == 4 ? 17.0 / 25.0 : se == 5 ? 69.0 / 100.0 : 13.0 / 21.0;
}
__device__ __forceinline__ long long qot_q(double x) { return __double2ll_rn(x * QOT_SCALE); }

// A lightpath of n slots centred at f, on one link: every span adds (P_ASE + P_NLI) / P (calculate_osnr.py:45-52).
// The factors that do not depend on the link are hoisted: v = (ase * lc.z + nli * lc.x * sum_phi) / P.  The
// reference adds span by span and divides twice (1 / (P / x)); here spans are a multiplication - the GSNR agrees to
// ~1e-14 dB (tolerance of the path: 1e-4 dB).
struct QotChan {
  double ase, nli;  // B h f NF / P ;  (P / B)^3 8 / (27 pi |b2|) g^2 B / P
};
__device__ __forceinline__ QotChan qot_chan(const Params& p, const GsnrParams& g, int start, int n) {
  const double2 t = __ldg(p.qot_chan + n);  // per slot count: (B h NF / P, (P / B)^3 8 / (27 pi |b2|) g^2 B / P), built at create
  QotChan c;
  c.ase = t.x * (g.f0 + ((double)start + 0.5 * (double)n) * g.slot_hz);
  c.nli = t.y;
  return c;
}
__device__ __forceinline__ double qot_sum_phi(const GsnrParams& g, int n, long long SA, long long SX, double lol) {
  return __ldg(g.sci_tab + n) + ((double)SA * QOT_INV_SCALE - ((double)SX * QOT_INV_SCALE) * lol);
}

template <int QOT_CAP, bool NARROW, bool MULTICORE>
struct QotCtx {
  const Params& p;
  const GsnrParams& g;
  QotWarpSmem<QOT_CAP, NARROW>& sm;
  const double4* link_c;  // per-link factors (l_eff, l_eff / L, exp(2 a L) - 1, spans), staged in shared memory
  int e, lane;
  const ulonglong2* heap;
  int heap_len;
  longlong2* sums;  // this env's service table [slot][qot_hops]
  int bit_rate, pair;
  double invP;
  int win;  // staged window of the release queue

  __device__ __forceinline__ longlong2* row(int slot) const { return sums + (size_t)slot * p.qot_hops; }
  __device__ __forceinline__ const unsigned char* path_bytes(int pr, int path) const {
    return reinterpret_cast<const unsigned char*>(p.path_rec + ((size_t)pr * p.k + path) * p.rec_u4);
  }

  // link set of a staged service: staged (NARROW) or from the path table (a few KB, L1 resident)
  __device__ __forceinline__ ulonglong2 svc_mask(unsigned long long s, int i) const {
    if constexpr (QotWarpSmem<QOT_CAP, NARROW>::STAGED) return make_ulonglong2(sm.mask[i], 0ull);
    else return __ldg(g.path_mask + (size_t)svc_pair(s) * p.k + svc_path(s));
  }
  __device__ __forceinline__ void stage(int w) {
    __syncwarp();
    const int base = w * QOT_CAP, n = min(QOT_CAP, heap_len - base);
    // every round's payload is asked for before the first one is used: one memory round trip per window, not one per
    // 32 services
    unsigned long long s[QOT_CAP / 32];
#pragma unroll
    for (int u = 0; u < QOT_CAP / 32; u++) s[u] = (lane + 32 * u < n) ? heap[base + lane + 32 * u].y : 0ull;
#pragma unroll
    for (int u = 0; u < QOT_CAP / 32; u++) {
      const int i = lane + 32 * u;
      if (i < n) {
        sm.pay[i] = s[u];
        if constexpr (QotWarpSmem<QOT_CAP, NARROW>::STAGED)
          sm.mask[i] = __ldg(reinterpret_cast<const unsigned long long*>(g.path_mask + (size_t)svc_pair(s[u]) * p.k + svc_path(s[u])));
      }
    }
    win = w;
    __syncwarp();
  }

  // `f(payload, common links)` for every running service on `core` that shares a link with `mask`: the window's
  // services are filtered 32 at a time from shared memory, the hits compacted (ballot), then visited with one lane
  // each, so the per-neighbour work runs with full lanes.
  template <typename F>
  __device__ __forceinline__ void for_neighbours(int core, ulonglong2 mask, F f) {
    const int nwin = (heap_len + QOT_CAP - 1) / QOT_CAP;
    for (int w = 0; w < nwin; w++) {
      if (win != w) stage(w);
      const int n = min(QOT_CAP, heap_len - w * QOT_CAP);
      int cnt = 0;
      for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        bool hit = false;
        if (i < n && (!MULTICORE || svc_core(sm.pay[i]) == core)) {  // RMSA-v0: one core, nothing to compare
          const ulonglong2 im = svc_mask(sm.pay[i], i);
          hit = ((im.x & mask.x) | (im.y & mask.y)) != 0ull;
        }
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, hit);
        if (hit) sm.nb[cnt + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)i;
        cnt += __popc(bal);
      }
      __syncwarp();
      for (int j = lane; j < cnt; j += 32) {
        const int i = sm.nb[j];
        const ulonglong2 im = svc_mask(sm.pay[i], i);
        f(sm.pay[i], make_ulonglong2(im.x & mask.x, im.y & mask.y));
      }
      __syncwarp();
    }
  }

  // sum over the hops of (spans x per-span term); lane h < hops holds link h's term
  __device__ __forceinline__ double hop_sum(double v, int hops) const {
    double acc = 0.0;
    for (int h = 0; h < hops; h++) acc += __shfl_sync(0xFFFFFFFFu, v, h);
    return acc;
  }

  // the channel alone on its path: ASE and self-channel NLI only.  1 / GSNR (linear).
  __device__ __forceinline__ double alone_acc(const PathRec& rec, int start, int n) const {
    const QotChan c = qot_chan(p, g, start, n);
    const int hops = rec.hops();
    double v = 0.0;
    if (lane < hops) {
      const double4 lc = link_c[rec.link(lane)];
      v = lc.w * (c.ase * lc.z + c.nli * lc.x * __ldg(g.sci_tab + n));
    }
    return hop_sum(v, hops);
  }
  __device__ __forceinline__ bool alone_ok(const PathRec& rec, int start, int n, int se) const {
    return alone_acc(rec, start, n) <= p.qot_thr_lin[se];
  }

  // The gate for one candidate; warp-collective.  Returns pass / fail; acc_c = 1 / GSNR of the candidate,
  // acc_w = the largest 1 / GSNR among its neighbours with the candidate present (0 without neighbours).
  // Leaves the candidate's sums in sm.cSA_* / sm.cSX_*.
  __device__ __forceinline__ bool gate(const PathRec& rec, int path, int core, int start, int n, int se, double& acc_c, double& acc_w) {
    const int cpos2 = 2 * start + n;
    const ulonglong2 cmask = __ldg(g.path_mask + (size_t)pair * p.k + path);
    const int hops = rec.hops();
    __syncwarp();
    sm.cSA_lo[lane] = sm.cSA_hi[lane] = sm.cSX_lo[lane] = sm.cSX_hi[lane] = 0u;
    if (lane < hops) sm.hop_of[rec.link(lane)] = (unsigned char)lane;
    __syncwarp();
    const int cols = 4 * g.tab_S + 1;
    const double pm_c = phi_mod(se);
    bool fail = false;
    double worst = 0.0;
    for_neighbours(core, cmask, [&](unsigned long long s, ulonglong2 common) {
      const int ni = svc_slots(s), sei = svc_se(s);
      const int d = (2 * svc_start(s) + ni) - cpos2;  // centre distance in half slots, interferer - candidate
      const double2 t_ci = __ldg(g.xci_tab + (size_t)ni * cols + (d + 2 * g.tab_S));  // service i seen by the candidate
      const double2 t_ic = __ldg(g.xci_tab + (size_t)n * cols + (-d + 2 * g.tab_S));  // the candidate seen by service i
      const unsigned char* ri = path_bytes(svc_pair(s), svc_path(s));  // byte 0 hops, bytes 4.. link per hop
      const longlong2* rw = row(svc_slot(s));
      const unsigned long long qa_ci = (unsigned long long)qot_q(t_ci.x),
                               qx_ci = (unsigned long long)qot_q(phi_mod(sei) * t_ci.y);
      const long long qa_ic = qot_q(t_ic.x), qx_ic = qot_q(pm_c * t_ic.y);
      unsigned long long c0 = common.x, c1 = common.y;
      while (c0) {
        const int h = sm.hop_of[__ffsll((long long)c0) - 1];
        sm.add64(&sm.cSA_lo[h], &sm.cSA_hi[h], qa_ci);
        sm.add64(&sm.cSX_lo[h], &sm.cSX_hi[h], qx_ci);
        c0 &= c0 - 1;
      }
      while (c1) {
        const int h = sm.hop_of[64 + __ffsll((long long)c1) - 1];
        sm.add64(&sm.cSA_lo[h], &sm.cSA_hi[h], qa_ci);
        sm.add64(&sm.cSX_lo[h], &sm.cSX_hi[h], qx_ci);
        c1 &= c1 - 1;
      }
      // the neighbour with the candidate present: its stored sums + the candidate's term on the shared links
      const QotChan ci = qot_chan(p, g, svc_start(s), ni);
      const double sci = __ldg(g.sci_tab + ni);
      double acc = 0.0;
      const int hi = __ldg(ri);
      for (int h0 = 0; h0 < hi; h0 += 4) {  // four independent row loads in flight; no branch per hop (a hop past
        longlong2 sv[4];                     // the end reads link 0 of the record's padding and adds +0.0)
#pragma unroll
        for (int u = 0; u < 4; u++) sv[u] = (h0 + u < hi) ? rw[h0 + u] : make_longlong2(0, 0);
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int l = __ldg(ri + 4 + h0 + u);
          const double4 lc = link_c[l];
          const bool shared = mask_has(common, l);
          sv[u].x += shared ? qa_ic : 0ll;
          sv[u].y += shared ? qx_ic : 0ll;
          const double sum_phi = sci + ((double)sv[u].x * QOT_INV_SCALE - ((double)sv[u].y * QOT_INV_SCALE) * lc.y);
          const double t = lc.w * (ci.ase * lc.z + ci.nli * lc.x * sum_phi);
          acc += (h0 + u < hi) ? t : 0.0;
        }
      }
      worst = fmax(worst, acc);
      if (!(acc <= p.qot_thr_lin[sei])) fail = true;
    });
    fail = __any_sync(0xFFFFFFFFu, fail);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xFFFFFFFFu, worst, o));
    const QotChan cc = qot_chan(p, g, start, n);
    double v = 0.0;
    if (lane < hops) {
      const double4 lc = link_c[rec.link(lane)];
      v = lc.w * (cc.ase * lc.z + cc.nli * lc.x * qot_sum_phi(g, n, sm.SA(lane), sm.SX(lane), lc.y));
    }
    acc_c = hop_sum(v, hops);
    acc_w = worst;
    return acc_c <= p.qot_thr_lin[se] && !fail;
  }

  // add (sign = +1) or remove (-1) lightpath (path mask, core, start, n, se) from the sums of its neighbours; the
  // service in row `skip_row` is the lightpath itself
  __device__ __forceinline__ void apply_to_neighbours(ulonglong2 mask, int core, int start, int n, int se, int sign, int skip_row) {
    const int cols = 4 * g.tab_S + 1;
    const int pos2 = 2 * start + n;
    const double pm = phi_mod(se);
    for_neighbours(core, mask, [&](unsigned long long s, ulonglong2 common) {
      if (svc_slot(s) == skip_row) return;
      const int d = (2 * svc_start(s) + svc_slots(s)) - pos2;
      const double2 t = __ldg(g.xci_tab + (size_t)n * cols + (-d + 2 * g.tab_S));  // this lightpath seen by service i
      const long long qa = sign * qot_q(t.x), qx = sign * qot_q(pm * t.y);
      const unsigned char* ri = path_bytes(svc_pair(s), svc_path(s));
      longlong2* rw = row(svc_slot(s));
      const int hi = __ldg(ri);
      // reductions without a return value: nothing waits for the rows.  This env's rows are read (by the gate, below
      // in the same warp, after a __syncwarp) only after every event has been issued, and nobody else touches them.
      for (int h = 0; h < hi; h++)
        if (mask_has(common, __ldg(ri + 4 + h))) {
          atomicAdd(reinterpret_cast<unsigned long long*>(&rw[h].x), (unsigned long long)qa);
          atomicAdd(reinterpret_cast<unsigned long long*>(&rw[h].y), (unsigned long long)qx);
        }
    });
  }
};

__device__ __forceinline__ bool qot_crosstalk_ok(const Params& p, int bit_rate, int mod, double path_length) {
  // _crosstalk_is_acceptable rmcsa_env.py:341-384 (as Env::crosstalk_ok)
  double denom = __ldg(p.mod_coeff + mod) * ((double)bit_rate / (double)__ldg(p.mod_se + mod));
  denom = denom * 1e9;
  double lmax_snr = 100.0 / denom;
  lmax_snr = lmax_snr / 1000.0;
  return path_length < __ldg(p.mod_reach + mod) && path_length < lmax_snr;
}

// One lane's candidate lightpath for the pending request
struct QotCand {
  int path, core, start, n, se, mod;
};
__device__ __forceinline__ QotCand qot_cand_from(const QotCand& c, int src_lane) {
  QotCand r;
  r.path = __shfl_sync(0xFFFFFFFFu, c.path, src_lane);
  r.core = __shfl_sync(0xFFFFFFFFu, c.core, src_lane);
  r.start = __shfl_sync(0xFFFFFFFFu, c.start, src_lane);
  r.n = __shfl_sync(0xFFFFFFFFu, c.n, src_lane);
  r.se = __shfl_sync(0xFFFFFFFFu, c.se, src_lane);
  r.mod = __shfl_sync(0xFFFFFFFFu, c.mod, src_lane);
  return r;
}

// the channel alone on its path, evaluated by ONE lane (hops walked serially): lanes hold different candidates
__device__ __forceinline__ bool qot_alone_ok_lane(const Params& p, const GsnrParams& g, const double4* link_c,
                                                  const PathRec& rec, int start, int n, int se, double invP) {
  const QotChan c = qot_chan(p, g, start, n);
  const double sci = __ldg(g.sci_tab + n);
  double acc = 0.0;
  // lanes hold paths of different lengths: the hops of a group of four are computed without branches (a hop past the
  // end reads link 0 and adds +0.0, which leaves the sum as it is), only the groups are skipped
  const int h = rec.hops();
#pragma unroll
  for (int gq = 0; gq < 3; gq++) {
    if (4 * gq < h) {
      const uint32_t wd = rec.w[gq + 1];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const double4 lc = link_c[(wd >> (8 * u)) & 0xFFu];
        const double t = lc.w * (c.ase * lc.z + c.nli * lc.x * sci);
        acc += (4 * gq + u < h) ? t : 0.0;
      }
    }
  }
  for (int i = 12; i < h; i++) {  // long paths (32-byte records)
    const double4 lc = link_c[rec.link(i)];
    acc += lc.w * (c.ase * lc.z + c.nli * lc.x * sci);
  }
  return acc <= p.qot_thr_lin[se];
}

// One warp per env: (1) the services the last step released leave their neighbours' sums, (2) the candidates of the
// action / of heuristic 13 are generated one per lane - (path, modulation) pairs: slot count, first fit, RMCSA's
// closed-form gate, interference-free GSNR - and the viable ones put through the gate in the order the reference's
// first-fit heuristics walk them, (3) an accepted candidate gets a row of the service table and enters its
// neighbours' sums; the verdict tells the step kernel what to provision.
template <int W, int QOT_CAP, bool RMCSA, int MODE, bool NARROW>
__global__ void __launch_bounds__(QOT_WARPS * 32, QOT_MIN_CTAS) qot_kernel(const __grid_constant__ Params p, const __grid_constant__ GsnrParams g, const QotIO io) {
  __shared__ QotWarpSmem<QOT_CAP, NARROW> smem[QOT_WARPS];
  extern __shared__ double4 s_link[];  // per-link factors, p.E entries (sized at launch)
  for (int l = threadIdx.x; l < p.E; l += blockDim.x) s_link[l] = g.link_c[l];
  __syncthreads();
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (e >= p.num_envs) return;
  const uint32_t m = p.meta[e];
  const int bit_rate = (int)(m & 0xFFFFu), src = (int)((m >> 16) & 0xFFu), dst = (int)(m >> 24);
  QotCtx<QOT_CAP, NARROW, RMCSA> cx{p, g, smem[threadIdx.x >> 5], s_link, e, lane, p.heap + (size_t)e * p.heap_stride + HEAP_PAD, p.heap_len[e],
            p.qot_sums + (size_t)e * p.qot_slot_cap * p.qot_hops, bit_rate, src * p.n_nodes + dst, 1.0 / g.P, -1};
  unsigned short* free_stack = p.qot_free + (size_t)e * p.qot_slot_cap;
  int free_n = p.qot_free_n[e];
  if (cx.heap_len > 0) cx.stage(0);

  // (1) the lightpath the last launch accepted enters its neighbours' sums (it is in the queue by now - unless the
  // step released it at once), the services the last step released leave theirs.  Integer sums: any order gives the
  // same table.
  const unsigned long long added = p.qot_add[e];
  const int nrel = p.qot_rel_n[e], nadd = added != 0ull ? 1 : 0;
  // lane r holds released payload r (QOT_REL_CAP = 32 = one per lane), fetched before the loop needs them; the row an
  // accept will take is on top of the free stack now unless a release pushes one
  const unsigned long long my_rel = lane < nrel ? p.qot_rel[(size_t)e * QOT_REL_CAP + lane] : 0ull;
  int top_row = (nrel == 0 && free_n > 0) ? (int)free_stack[free_n - 1] : 0;
  for (int ev = 0; ev < nadd + nrel; ev++) {
    const bool is_add = ev < nadd;
    const unsigned long long rs = __shfl_sync(0xFFFFFFFFu, my_rel, (ev - nadd) & 31);
    const unsigned long long s = is_add ? added : rs;
    const ulonglong2 em = __ldg(g.path_mask + (size_t)svc_pair(s) * p.k + svc_path(s));
    cx.apply_to_neighbours(em, svc_core(s), svc_start(s), svc_slots(s), svc_se(s), is_add ? 1 : -1, is_add ? svc_slot(s) : -1);
    if (!is_add) {
      if (lane == 0) free_stack[free_n] = (unsigned short)svc_slot(s);
      free_n++;
      top_row = svc_slot(s);
    }
  }
  __syncwarp();

  // (2) candidates
  Tile<W, 1> tile{p.occ + (size_t)e * p.planes, 0, p.E};
  constexpr bool rmcsa = RMCSA;  // p.kind == ORLG_ENV_RMCSA (else RMSA-v0): each instantiation carries one env's code only
  constexpr bool verdict = MODE == QOT_VERDICT;  // MODE == io.mode: each instantiation carries one mode's code only
  // modulations tried per path, most efficient first: RMSA-v0 adaptive qot_se[], RMCSA-v0 adaptive qot_mod_order[],
  // else the path's best modulation; a verdict on an RMCSA action names its modulation
  const int nt = (rmcsa && verdict) ? 1 : !p.qot_adaptive ? 1 : rmcsa ? p.n_mods : p.qot_n_se;
  int act[4] = {0, 0, 0, 0};
  bool act_ok = true;
  if (verdict) {
    if (!rmcsa) {
      const int2 a = reinterpret_cast<const int2*>(io.actions)[e];
      act[0] = a.x, act[3] = a.y;
      act_ok = a.x >= 0 && a.y >= 0 && a.x < p.k && a.y < p.S;  // rmsa_env.py:233-235
    } else {
      const int4 a = reinterpret_cast<const int4*>(io.actions)[e];
      act[0] = a.x, act[1] = a.y, act[2] = a.z, act[3] = a.w;
      act_ok = a.x >= 0 && a.y >= 0 && a.z >= 0 && a.w >= 0 && a.x < p.k && a.y < p.n_mods && a.z < p.cores && a.w < p.S;
    }
  }
  // An agent loop asks the heuristic and steps with its answer: the verdict launch would repeat the heuristic launch's
  // gate.  When the host says nothing changed in between (io.use_proposal) and the action is the proposal, its verdict
  // stands: the lightpath, its modulation, and its sums, which that launch left in the next free row.
  bool cached = false;
  QotCand A{-1, 0, -1, 0, 0, 0};  // the last candidate evaluated
  if (verdict && io.use_proposal && act_ok) {
    const int4 v0 = p.qot_verdict[2 * (size_t)e], v1 = p.qot_verdict[2 * (size_t)e + 1];
    const int vp = v0.y & 0xFFFF, vc = v0.y >> 16;
    cached = v0.x != 0 && act[0] == vp && act[3] == v0.z && (!rmcsa || (act[1] == v1.z && act[2] == vc));
    if (cached) A = QotCand{vp, vc, v0.z, v0.w, v1.x, v1.z};
  }
  const int n_paths = (verdict ? 1 : p.k) * ((act_ok && !cached) ? 1 : 0);
  const int ppr = max(1, 32 / nt);  // whole paths per round of 32 lanes
  double gc = -1.0, gmin = 0.0;  // 1 / GSNR (linear) of the last candidate evaluated (< 0: none) and of its worst neighbour
  bool accept = cached;
  // One loop, one gate evaluation per trip (a single inlined copy of the gate).  phase 0: the walk of the heuristic,
  // candidates generated one per lane, a round of whole paths at a time, the viable ones gated in order.  phase 1
  // (RMSA-v0, QOT_HEUR_VERDICT): the action (path, slot) does not name the modulation - the step assigns the most
  // efficient one that passes AT THIS SLOT, which can be better than the one that passed at its own first-fit slot;
  // walk the step's search over the more efficient ones.  phase 2: none passed, the chosen one is evaluated again.
  QotCand c{-1, 0, -1, 0, 0, 0}, chosen{-1, 0, -1, 0, 0, 0};
  unsigned viable = 0u;
  int p0 = 0, phase = 0, cur_t = 0;
  double gc0 = -1.0, gmin0 = 0.0;
  while (true) {
    if (phase == 0 && !viable) {
      if (p0 >= n_paths) break;
      // ---- generation, one candidate per lane ----
      c = QotCand{-1, 0, -1, 0, 0, 0};
      bool ok = false;
      const int pl = lane / nt, t = lane - pl * nt;
      if (pl < ppr && p0 + pl < n_paths) {
        c.path = verdict ? act[0] : p0 + pl;
        const PathRec rec = load_path(p, cx.pair, c.path);
        if (rec.hops() > 0) {
          if (!rmcsa) {
            c.se = p.qot_adaptive ? p.qot_se[t] : rec.se();
            c.n = number_slots(bit_rate, c.se, p);
            const Row<W> av = tile.path_avail(rec, 0);
            if (verdict) {  // RMSAEnv.step: the slot is given
              c.start = act[3];
              ok = c.start + c.n <= p.S && range_is_free<W>(av, c.start, c.n);
            } else {  // walk of rmsa_env.py:901-913
              c.start = first_fit<W>(av, c.n, p.S);
              ok = c.start >= 0;
            }
          } else {
            c.mod = verdict ? act[1] : p.qot_adaptive ? p.qot_mod_order[t] : rec.best_mod();
            if (c.mod < p.n_mods) {  // else: no modulation reaches (get_best_modulation_format raises)
              c.se = __ldg(p.mod_se + c.mod);
              c.n = number_slots(bit_rate, c.se, p);
              // RMCSAEnv.step: core and slot are given; else the walk of rmcsa_env.py:882-911: the first core with room
              for (int core = verdict ? act[2] : 0; core < (verdict ? act[2] + 1 : p.cores) && !ok; core++) {
                const Row<W> av = tile.path_avail(rec, core);
                c.core = core;
                if (verdict) {
                  c.start = act[3];
                  ok = c.start + c.n <= p.S && range_is_free<W>(av, c.start, c.n);
                } else {
                  c.start = first_fit<W>(av, c.n, p.S);
                  ok = c.start >= 0;
                }
              }
              ok = ok && qot_crosstalk_ok(p, bit_rate, c.mod, __ldg(p.path_length + (size_t)cx.pair * p.k + c.path));
            }
          }
          // a slot count that does not fit ends a path's search in the reference walk; wider channels (later t) fit even
          // less, so dropping every candidate without room is the same thing
          if (ok && p.qot_adaptive) ok = qot_alone_ok_lane(p, g, s_link, rec, c.start, c.n, c.se, cx.invP);
        }
      }
      viable = __ballot_sync(0xFFFFFFFFu, ok);
      p0 += ppr;
      continue;
    }
    if (phase == 1 && !viable) {
      phase = 2;
      A = chosen;
    } else if (phase < 2) {
      const int L = __ffs((int)viable) - 1;
      viable &= viable - 1;
      A = qot_cand_from(c, L);
      cur_t = L % nt;
    }
    // ---- the gate ----
    const PathRec rec = load_path(p, cx.pair, A.path);
    accept = cx.gate(rec, A.path, A.core, A.start, A.n, A.se, gc, gmin);
    if (phase == 2) break;
    if (phase == 0 && accept && !rmcsa && MODE != QOT_VERDICT && cur_t > 0) {
      chosen = A;
      gc0 = gc, gmin0 = gmin;  // what the heuristic's own walk evaluated last (QOT_HEURISTIC reports that)
      c = chosen;
      bool ok2 = false;
      if (lane < cur_t) {
        c.se = p.qot_se[lane];
        c.n = number_slots(bit_rate, c.se, p);  // narrower than the chosen channel: it fits where that one fits
        ok2 = qot_alone_ok_lane(p, g, s_link, rec, c.start, c.n, c.se, cx.invP);
      }
      viable = __ballot_sync(0xFFFFFFFFu, ok2);
      phase = 1;
      accept = false;
      continue;
    }
    if (accept) break;
  }

  if (!verdict && io.actions_out != nullptr && lane == 0) {
    if (!rmcsa)
      reinterpret_cast<int2*>(io.actions_out)[e] = accept ? make_int2(A.path, A.start) : make_int2(p.k, p.S);
    else
      reinterpret_cast<int4*>(io.actions_out)[e] =
          accept ? make_int4(A.path, A.mod, A.core, A.start) : make_int4(p.k, p.n_mods, p.cores, p.S);
  }
  if (lane == 0) {  // 1 / GSNR (linear); orlg_qot_get converts to dB on demand (no log10 on the step path)
    double* q = p.qot_g + 2 * (size_t)e;   // the pair orlg_qot_get reports
    double* qp = p.qot_gp + 2 * (size_t)e; // the pair of a QOT_HEURISTIC launch's proposal, for the verdict launch that takes it
    if (cached) {
      q[0] = qp[0], q[1] = qp[1];
    } else {
      const double fc = gc;  // < 0: nothing evaluated
      const double fw = (gc >= 0.0 && gc <= p.qot_thr_lin[A.se]) ? gmin : -1.0;  // not reported when the candidate failed
      if (MODE == QOT_HEURISTIC) {
        qp[0] = fc, qp[1] = fw;
        // the heuristic's own walk ended at its first-fit candidate; the modulation search at that slot came after
        q[0] = gc0 >= 0.0 ? gc0 : fc, q[1] = gc0 >= 0.0 ? gmin0 : fw;
      } else {
        q[0] = fc, q[1] = fw;
      }
    }
  }

  // (3) the accepted candidate enters the service table (QOT_HEURISTIC: its sums are left in the row the next pop
  // would return and the verdict is recorded as the proposal; nothing is taken)
  int slot = 0;
  if (accept && free_n == 0) {  // more running services than rows (the release queue is as full): refuse
    accept = false;
    if (MODE != QOT_HEURISTIC && lane == 0) atomicOr((unsigned long long*)&p.counters[(size_t)CNT_FAULT * p.NP + e], 32ull);
  }
  if (accept) {
    slot = top_row;  // == free_stack[free_n - 1]
    if (MODE != QOT_HEURISTIC) free_n--;
    if (!cached) {
      const PathRec rec = load_path(p, cx.pair, A.path);
      if (lane < rec.hops())
        cx.row(slot)[lane] = make_longlong2(cx.sm.SA(lane), cx.sm.SX(lane));
    }
  }
  if (lane == 0) {
    p.qot_verdict[2 * (size_t)e] = make_int4(accept ? 1 : 0, A.path | (A.core << 16), A.start, A.n);
    p.qot_verdict[2 * (size_t)e + 1] = make_int4(A.se, slot, A.mod, 0);
  }
  if (lane == 0) {
    p.qot_free_n[e] = free_n;
    p.qot_rel_n[e] = 0;
    // the accepted lightpath enters its neighbours' sums at the start of the next launch (one copy of that code)
    p.qot_add[e] = (accept && MODE != QOT_HEURISTIC) ? pack_service(cx.pair, A.path, A.core, A.start, A.n, A.se, slot) : 0ull;
  }
}

// 1 / GSNR (linear, < 0 = none) -> dB (NaN = none) for orlg_qot_get
__global__ void qot_db_kernel(const double* in, double* out, long long n) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i] >= 0.0 ? 10.0 * log10(1.0 / in[i]) : __longlong_as_double(0x7FF8000000000000ll);
}

// Test / debug: recompute every running service's sums from scratch and compare with the table (exact).
// out[e] = number of (service, hop) rows that differ.  One thread per env.
__global__ void qot_check_kernel(const Params p, const GsnrParams g, long long* out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.num_envs) return;
  const ulonglong2* heap = p.heap + (size_t)e * p.heap_stride + HEAP_PAD;
  const int n_act = p.heap_len[e];
  const longlong2* sums = p.qot_sums + (size_t)e * p.qot_slot_cap * p.qot_hops;
  const int cols = 4 * g.tab_S + 1;
  long long bad = 0;
  for (int a = 0; a < n_act; a++) {
    const unsigned long long t = heap[a].y;
    const PathRec rt = load_path(p, svc_pair(t), svc_path(t));
    const int pos2 = 2 * svc_start(t) + svc_slots(t);
    for (int h = 0; h < rt.hops(); h++) {
      const int l = rt.link(h);
      long long SA = 0, SX = 0;
      for (int i = 0; i < n_act; i++) {
        if (i == a) continue;
        const unsigned long long s = heap[i].y;
        if (svc_core(s) != svc_core(t)) continue;
        const ulonglong2 im = __ldg(g.path_mask + (size_t)svc_pair(s) * p.k + svc_path(s));
        if (!mask_has(im, l)) continue;
        const int d = (2 * svc_start(s) + svc_slots(s)) - pos2;
        const double2 x = __ldg(g.xci_tab + (size_t)svc_slots(s) * cols + (d + 2 * g.tab_S));
        SA += qot_q(x.x);
        SX += qot_q(phi_mod(svc_se(s)) * x.y);
      }
      const longlong2 st = sums[(size_t)svc_slot(t) * p.qot_hops + h];
      if (st.x != SA || st.y != SX) bad++;
    }
  }
  out[e] = bad;
}
}  // namespace orlg
