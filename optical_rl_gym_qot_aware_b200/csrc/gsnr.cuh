// gsnr.cuh — GN-model GSNR of candidate lightpaths (calculate_osnr, examples/calculate_osnr.py:9-56).
//
// One warp per env.  The env's running services are the entries of its release heap
// (payload = pair, path, core, first slot, slots, spectral efficiency).  Work items are
// "targets": item 0 is the candidate, item j >= 1 the running service j-1 if it shares a link
// and core with the candidate (a neighbour, re-evaluated with the candidate as one more
// interferer).  Lanes take items round-robin; the neighbours' minimum is a warp shuffle
// reduction.  All arithmetic is f64 (asinh / exp / log10 of the CUDA math library).
//
// Per span of every link of the target (calculate_osnr.py:22-52):
//   l_eff_a = 1/(2a), l_eff = (1 - exp(-2 a L))/(2a)
//   sum_phi = asinh(pi^2 |b2| B^2 / (4a))                                      (SCI)
//           + sum_i [asinh(pi^2|b2| l_eff_a B_i (df_i + B_i/2)) - asinh(... (df_i - B_i/2))
//                    - Phi(SE_i) (B_i/|df_i|) 5/3 (l_eff/L)]                     (XCI)
//   P_NLI = (P/B)^3 8/(27 pi |b2|) g^2 l_eff sum_phi B,  P_ASE = B h f (exp(2aL) - 1) NF
//   acc  += 1 / (P / (P_ASE + P_NLI));   GSNR_dB = 10 log10(1/acc)
// `sum_phi += phi` is applied only for services other than the target (the reference adds the
// previous interferer's phi once more when it meets the target itself in running_services,
// SURVEY.md 7.4-7; a candidate is never in that list, so the two coincide for output [0]).
#pragma once
#include "engine_env.cuh"

namespace orlg {

struct GsnrParams {
  double P, alpha, nf, f0, slot_hz;
  const int* span_count;
  const double* span_km;
  const ulonglong2* path_mask;  // [N*N*k] links of every path as a bit set (fast kernel; null when E > 128)
  const double4* link_c;        // [E] (l_eff, l_eff / L, exp(2 a L) - 1, span count)
  // [S + 1][4 S + 1]: the asinh pair of calculate_osnr.py:36-44 for an interferer of n slots whose centre lies
  // d half-slots from the target's (column d + 2 S).  Channels live on the slot grid, so the pair only depends on
  // (n, d); null when the table would exceed 16 MB.
  const double2* xci_tab;  // (asinh pair, (B / |df|) * 5 / 3)
  const double* sci_tab;   // [S + 1] self-channel term asinh(pi^2 |b2| B^2 / (4 a)) of an n-slot channel (:30)
  int tab_S;
};

__device__ __forceinline__ bool path_has_link(const PathRec& r, int link) {
  bool f = false;
  int h = r.hops();
  for (int i = 0; i < h; i++) f = f || (r.link(i) == link);
  return f;
}

struct Chan {  // a lightpath as the GN model sees it
  double B, f;
  int se;
  int n, pos2;  // slots, and twice the centre position in slots (2 * first + n): the integer grid behind B and f
};
__device__ __forceinline__ Chan chan_of(const GsnrParams& g, int start, int n, int se) {
  Chan c;
  c.B = (double)n * g.slot_hz;
  c.f = g.f0 + ((double)start + 0.5 * (double)n) * g.slot_hz;
  c.se = se;
  c.n = n;
  c.pos2 = 2 * start + n;
  return c;
}

__device__ __forceinline__ double xci_phi(const Chan& t, const Chan& i, double l_eff_a, double leff_over_L) {
  const double pi = 3.141592653589793, beta2 = 21.3e-27;
  double df = i.f - t.f;
  double c = pi * pi * beta2 * l_eff_a * i.B;
  // phi_modulation_format[spectral_efficiency - 1], calculate_osnr.py:20
  const double pm = i.se <= 2 ? 1.0 : i.se == 3 ? 2.0 / 3.0 : i.se == 4 ? 17.0 / 25.0 : i.se == 5 ? 69.0 / 100.0 : 13.0 / 21.0;
  return (asinh(c * (df + i.B / 2)) - asinh(c * (df - i.B / 2))) - pm * (i.B / fabs(df)) * 5.0 / 3.0 * leff_over_L;
}

// GSNR (dB) of target (rec, core, ch) given the heap's services except index `skip`, plus the
// optional extra interferer (xrec, xcore, xch).
__device__ double gsnr_target(const Params& p, const GsnrParams& g, const ulonglong2* heap, int n_act,
                              const PathRec& rec, int core, const Chan& ch, int skip, bool has_extra,
                              const PathRec& xrec, int xcore, const Chan& xch) {
  const double pi = 3.141592653589793, beta2 = 21.3e-27, gamma = 1.3e-3, h_planck = 6.626e-34;
  const double a = g.alpha;
  const double l_eff_a = 1.0 / (2.0 * a);
  const double sci = asinh(pi * pi * beta2 * ch.B * ch.B / (4.0 * a));
  double acc = 0.0;
  const int hops = rec.hops();
  for (int hi = 0; hi < hops; hi++) {
    const int l = rec.link(hi);
    const double L = __ldg(g.span_km + l) * 1e3;
    const int ns = __ldg(g.span_count + l);
    const double l_eff = (1.0 - exp(-2.0 * a * L)) / (2.0 * a);
    double sum_phi = sci;
    for (int i = 0; i < n_act; i++) {
      if (i == skip) continue;
      unsigned long long s = heap[i].y;
      if (svc_core(s) != core) continue;
      PathRec r = load_path(p, svc_pair(s), svc_path(s));
      if (!path_has_link(r, l)) continue;
      sum_phi += xci_phi(ch, chan_of(g, svc_start(s), svc_slots(s), svc_se(s)), l_eff_a, l_eff / L);
    }
    if (has_extra && xcore == core && path_has_link(xrec, l)) sum_phi += xci_phi(ch, xch, l_eff_a, l_eff / L);
    double pr = g.P / ch.B;
    double p_nli = pr * pr * pr * (8.0 / (27.0 * pi * beta2)) * gamma * gamma * l_eff * sum_phi * ch.B;
    double p_ase = ch.B * h_planck * ch.f * (exp(2.0 * a * L) - 1.0) * g.nf;
    double v = 1.0 / (g.P / (p_ase + p_nli));
    for (int k = 0; k < ns; k++) acc += v;  // identical spans, summed span by span like the reference
  }
  return 10.0 * log10(1.0 / acc);
}

// Candidate of env e: explicit query (path | core << 16, first slot, slots, se or 0), or — FROM_ACTION — the env's
// own step action for the pending request: RMSA-v0 (path, first slot) with the path's best modulation, RMCSA-v0
// (path, modulation, core, first slot); the slot count then comes from the pending bit rate through
// get_number_slots, so a heuristic's (or the fused hint's) action tensor is evaluated as it is.
template <bool FROM_ACTION>
__device__ __forceinline__ bool resolve_candidate(const Params& p, const int* __restrict__ query, int e, int4& q,
                                                  PathRec& crec, int& pair) {
  uint32_t m = p.meta[e];
  pair = (int)((m >> 16) & 0xFFu) * p.n_nodes + (int)(m >> 24);
  int mod = -1;
  if (FROM_ACTION) {
    if (p.action_dim == 2) {
      int2 a = reinterpret_cast<const int2*>(query)[e];
      q = make_int4(a.x, a.y, 0, 0);
    } else {
      int4 a = reinterpret_cast<const int4*>(query)[e];
      mod = a.y;
      q = make_int4((a.x & 0xFFFF) | (a.z << 16), a.w, 0, 0);
      if (a.x < 0 || a.x >= p.k || a.z < 0 || a.z >= p.cores || mod < 0 || mod >= p.n_mods) q.x = -1;
    }
  } else {
    q = reinterpret_cast<const int4*>(query)[e];
  }
  const int path = q.x & 0xFFFF, core = q.x >> 16;
  bool ok = q.x >= 0 && path < p.k && core < p.cores;
  if (ok) {
    crec = load_path(p, pair, path);
    ok = crec.hops() > 0;
  }
  if (ok && FROM_ACTION) {
    q.w = mod >= 0 ? __ldg(p.mod_se + mod) : crec.se();
    q.z = number_slots((int)(m & 0xFFFFu), q.w, p);
  }
  return ok && q.y >= 0 && q.z >= 1 && q.y + q.z <= p.S;
}

// Generic kernel (any number of links): lanes take targets round-robin, every target rescans the env's services
// per hop.  Kept for topologies with more than 128 links; everything else runs gsnr_kernel_fast below.
template <bool FROM_ACTION>
__global__ void __launch_bounds__(128) gsnr_kernel_generic(const Params p, const GsnrParams g, const int* __restrict__ query,
                                                           double* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= p.num_envs) return;
  const int e = warp;
  const double nan = __longlong_as_double(0x7FF8000000000000ll), inf = __longlong_as_double(0x7FF0000000000000ll);
  int4 q;
  PathRec crec{};
  int pair;
  if (!resolve_candidate<FROM_ACTION>(p, query, e, q, crec, pair)) {
    if (lane == 0) out[2 * (size_t)e] = out[2 * (size_t)e + 1] = nan;
    return;
  }
  const int core = q.x >> 16;
  const Chan cch = chan_of(g, q.y, q.z, q.w > 0 ? q.w : crec.se());
  const ulonglong2* heap = p.heap + (size_t)e * p.heap_stride + HEAP_PAD;
  const int n_act = p.heap_len[e];
  double mine = inf;
  for (int item = lane; item <= n_act; item += 32) {
    if (item == 0) {
      double v = gsnr_target(p, g, heap, n_act, crec, core, cch, -1, false, crec, core, cch);
      out[2 * (size_t)e] = v;
    } else {
      unsigned long long s = heap[item - 1].y;
      if (svc_core(s) != core) continue;
      PathRec r = load_path(p, svc_pair(s), svc_path(s));
      bool shares = false;
      for (int hi = 0; hi < crec.hops(); hi++) shares = shares || path_has_link(r, crec.link(hi));
      if (!shares) continue;
      double v = gsnr_target(p, g, heap, n_act, r, core, chan_of(g, svc_start(s), svc_slots(s), svc_se(s)),
                             item - 1, true, crec, core, cch);
      mine = fmin(mine, v);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine = fmin(mine, __shfl_xor_sync(0xFFFFFFFFu, mine, o));
  if (lane == 0) out[2 * (size_t)e + 1] = mine;
}

// ---- fast kernel -------------------------------------------------------------------------------------------
// Static tables built once by gsnr_init_tables (orlg_gsnr_configure):
//   path_mask[pair*k + path]  the links of a path as a 128-bit set, so "does service i use link l" and "do two
//                             paths share a link" are bit tests instead of loops over path records;
//   link_c[l] = (l_eff, l_eff / L, exp(2 a L) - 1, span count), the per-link factors of calculate_osnr.py:25,51.
// The warp of an env walks its targets one at a time (the candidate, then every running service that shares a
// link and core with it).  For one target the 32 lanes (1) scan the env's services 32 at a time and queue the
// ones that interfere (same core, at least one common link) by ballot compaction, (2) whenever 32 are queued
// evaluate their cross-channel terms with one lane per interferer — the two asinh of calculate_osnr.py:36-44 do
// not depend on the link, only the last term scales with l_eff / L — and (3) for every hop of the target the
// terms of the interferers present on that link are summed over the warp (fixed butterfly order) into lane h.
// Per link the terms are thus added batch by batch in heap order and tree-wise inside a batch: deterministic,
// and within ~1e-15 relative of the generic kernel's strictly sequential sum (tolerance of the path: 1e-4 dB).
constexpr int GS_WARPS = 4;
struct GsnrWarpSmem {
  unsigned queue[64];
  int tlink[32];  // links of the current target, hop order
  unsigned char hop_of[128];  // link -> hop of the current target (valid for the target's links only)
};

__global__ void gsnr_init_tables(const Params p, const GsnrParams g, int npaths, ulonglong2* path_mask, double4* link_c) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < npaths) {
    PathRec r = load_path(p, i / p.k, i % p.k);
    unsigned long long lo = 0ull, hi = 0ull;
    for (int h = 0; h < r.hops(); h++) {
      int l = r.link(h);
      if (l < 64) lo |= 1ull << l;
      else hi |= 1ull << (l & 63);
    }
    path_mask[i] = make_ulonglong2(lo, hi);
  }
  if (i < p.E) {
    const double a = g.alpha, L = g.span_km[i] * 1e3;
    const double l_eff = (1.0 - exp(-2.0 * a * L)) / (2.0 * a);
    link_c[i] = make_double4(l_eff, l_eff / L, exp(2.0 * a * L) - 1.0, (double)g.span_count[i]);
  }
}

__global__ void gsnr_init_xci(const GsnrParams g, int S, double2* tab, double* sci) {
  const int cols = 4 * S + 1;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= S) {
    const double Bn = (double)i * g.slot_hz;
    sci[i] = asinh(3.141592653589793 * 3.141592653589793 * 21.3e-27 * Bn * Bn / (4.0 * g.alpha));
  }
  if (i >= (long long)(S + 1) * cols) return;
  const int n = (int)(i / cols), d = (int)(i % cols) - 2 * S;
  const double pi = 3.141592653589793, beta2 = 21.3e-27;
  const double B = (double)n * g.slot_hz, df = (double)d * (0.5 * g.slot_hz);
  const double c = pi * pi * beta2 * (1.0 / (2.0 * g.alpha)) * B;
  tab[i] = make_double2(asinh(c * (df + B / 2)) - asinh(c * (df - B / 2)), (B / fabs(df)) * 5.0 / 3.0);
}

__device__ __forceinline__ bool mask_has(ulonglong2 m, int l) { return (((l < 64) ? m.x : m.y) >> (l & 63)) & 1ull; }

#ifndef GS_MIN_CTAS  // resident CTAs per SM the register allocation must allow (A/B knob)
#define GS_MIN_CTAS 6
#endif
template <bool FROM_ACTION>
__global__ void __launch_bounds__(GS_WARPS * 32, GS_MIN_CTAS) gsnr_kernel_fast(const Params p, const GsnrParams g,
                                                                  const int* __restrict__ query, double* __restrict__ out) {
  __shared__ GsnrWarpSmem smem[GS_WARPS];
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= p.num_envs) return;
  GsnrWarpSmem& sm = smem[threadIdx.x >> 5];
  const int e = warp;
  const double nan = __longlong_as_double(0x7FF8000000000000ll), inf = __longlong_as_double(0x7FF0000000000000ll);
  const double pi = 3.141592653589793, beta2 = 21.3e-27, gamma = 1.3e-3, h_planck = 6.626e-34;
  int4 q;
  PathRec crec{};
  int pair;
  if (!resolve_candidate<FROM_ACTION>(p, query, e, q, crec, pair)) {  // uniform per warp
    if (lane == 0) out[2 * (size_t)e] = out[2 * (size_t)e + 1] = nan;
    return;
  }
  const int ccore = q.x >> 16;
  const Chan cch = chan_of(g, q.y, q.z, q.w > 0 ? q.w : crec.se());
  const ulonglong2 cmask = __ldg(g.path_mask + (size_t)pair * p.k + (q.x & 0xFFFF));
  const ulonglong2* heap = p.heap + (size_t)e * p.heap_stride + HEAP_PAD;
  const int n_act = p.heap_len[e];
  const double a = g.alpha, l_eff_a = 1.0 / (2.0 * a);
  const double cxci = pi * pi * beta2 * l_eff_a;
  double worst = inf;

  for (int item = 0; item <= n_act; item++) {
    // ---- the target (uniform over the warp) ----
    PathRec trec;
    Chan tch;
    ulonglong2 tmask;
    if (item == 0) {
      trec = crec;
      tch = cch;
      tmask = cmask;
    } else {
      const unsigned long long s = heap[item - 1].y;
      if (svc_core(s) != ccore) continue;
      const size_t pi_ = (size_t)svc_pair(s) * p.k + svc_path(s);
      tmask = __ldg(g.path_mask + pi_);
      if (((tmask.x & cmask.x) | (tmask.y & cmask.y)) == 0ull) continue;
      trec = load_path(p, svc_pair(s), svc_path(s));
      tch = chan_of(g, svc_start(s), svc_slots(s), svc_se(s));
    }
    const int hops = trec.hops();
    const int my_link = lane < hops ? trec.link(lane) : 0;
    double4 lc = make_double4(0.0, 0.0, 0.0, 0.0);
    if (lane < hops) lc = g.link_c[my_link];
    sm.tlink[lane] = my_link;
    if (lane < hops) sm.hop_of[my_link] = (unsigned char)lane;
    __syncwarp();
    double sum_phi = g.sci_tab ? __ldg(g.sci_tab + tch.n)  // SCI, calculate_osnr.py:30
                               : asinh(pi * pi * beta2 * tch.B * tch.B / (4.0 * a));

    // ---- interferers: queue by ballot compaction, evaluate 32 at a time ----
    int queued = 0;
    const int n_src = n_act + (item > 0 ? 1 : 0);  // neighbours also see the candidate, last (index n_act)
    auto drain = [&](int count) {  // evaluate queue[0 .. count), count <= 32, one interferer per lane
      double A = 0.0, X = 0.0;
      unsigned hm = 0u;
      if (lane < count) {
        const int i = (int)sm.queue[lane];
        Chan ic;
        ulonglong2 im;
        if (i == n_act) {
          ic = cch;
          im = cmask;
        } else {
          const unsigned long long s = heap[i].y;
          ic = chan_of(g, svc_start(s), svc_slots(s), svc_se(s));
          im = __ldg(g.path_mask + (size_t)svc_pair(s) * p.k + svc_path(s));
        }
        const double pm = ic.se <= 2 ? 1.0 : ic.se == 3 ? 2.0 / 3.0 : ic.se == 4 ? 17.0 / 25.0 : ic.se == 5 ? 69.0 / 100.0 : 13.0 / 21.0;
        // asinh(u) - asinh(v), calculate_osnr.py:36-44.  Channels that share a link and core do not overlap, so
        // u and v have one sign and the difference is log((|u'| + sqrt(u'^2 + 1)) / (|v'| + sqrt(v'^2 + 1)))
        // with u' the larger magnitude: one log instead of two asinh (agreement ~1e-15 relative).
        if (g.xci_tab != nullptr) {  // both channels sit on the slot grid: the pair and B/|df| are a table entry
          const double2 t = __ldg(g.xci_tab + (size_t)ic.n * (4 * g.tab_S + 1) + (ic.pos2 - tch.pos2 + 2 * g.tab_S));
          A = t.x;
          X = pm * t.y;
        } else {
          const double df = ic.f - tch.f;
          const double c = cxci * ic.B;
          double u = c * (df + ic.B / 2), v = c * (df - ic.B / 2);
          if (u > 0.0 && v < 0.0) {
            A = asinh(u) - asinh(v);  // overlapping spectra (a candidate that is not free): literal form
          } else {
            if (u <= 0.0) {
              const double t = -u;
              u = -v;
              v = t;
            }
            A = log((u + sqrt(u * u + 1.0)) / (v + sqrt(v * v + 1.0)));
          }
          X = pm * (ic.B / fabs(df)) * 5.0 / 3.0;
        }
        // hops of the target the interferer is on: walk the common links (usually one or two)
        unsigned long long c0 = im.x & tmask.x, c1 = im.y & tmask.y;
        while (c0) {
          hm |= 1u << sm.hop_of[__ffsll((long long)c0) - 1];
          c0 &= c0 - 1;
        }
        while (c1) {
          hm |= 1u << sm.hop_of[64 + __ffsll((long long)c1) - 1];
          c1 &= c1 - 1;
        }
      }
      // per hop of the target: sum over the 32 lanes (fixed butterfly order), lane h keeps the link's total
      for (int h = 0; h < hops; h++) {
        const double lol = __shfl_sync(0xFFFFFFFFu, lc.y, h);
        double t = ((hm >> h) & 1u) ? A - X * lol : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xFFFFFFFFu, t, o);
        if (lane == h) sum_phi += t;
      }
      __syncwarp();
    };
    for (int base = 0; base < n_src; base += 32) {
      const int i = base + lane;
      bool hit = false;
      if (i < n_act) {
        if (i != item - 1) {
          const unsigned long long s = heap[i].y;
          if (svc_core(s) == ccore) {
            const ulonglong2 im = __ldg(g.path_mask + (size_t)svc_pair(s) * p.k + svc_path(s));
            hit = ((im.x & tmask.x) | (im.y & tmask.y)) != 0ull;
          }
        }
      } else if (i == n_act && item > 0) {
        hit = true;  // the candidate shares a link and the core with every neighbour by construction
      }
      const unsigned bal = __ballot_sync(0xFFFFFFFFu, hit);
      if (hit) sm.queue[queued + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
      queued += __popc(bal);
      __syncwarp();
      if (queued >= 32) {
        drain(32);
        if (lane < queued - 32) sm.queue[lane] = sm.queue[32 + lane];  // keep the overflow, in order
        queued -= 32;
        __syncwarp();
      }
    }
    if (queued > 0) drain(queued);

    // ---- per link: P_NLI, P_ASE, one term per span (calculate_osnr.py:45-54), links in path order ----
    double v = 0.0;
    if (lane < hops) {
      const double pr = g.P / tch.B;
      const double p_nli = pr * pr * pr * (8.0 / (27.0 * pi * beta2)) * gamma * gamma * lc.x * sum_phi * tch.B;
      const double p_ase = tch.B * h_planck * tch.f * lc.z * g.nf;
      v = 1.0 / (g.P / (p_ase + p_nli));
    }
    double acc = 0.0;
    for (int h = 0; h < hops; h++) {
      const double vh = __shfl_sync(0xFFFFFFFFu, v, h);
      const int ns = (int)__shfl_sync(0xFFFFFFFFu, lc.w, h);
      for (int k = 0; k < ns; k++) acc += vh;
    }
    const double gs = 10.0 * log10(1.0 / acc);
    if (item == 0) {
      if (lane == 0) out[2 * (size_t)e] = gs;
    } else {
      worst = fmin(worst, gs);
    }
  }
  if (lane == 0) out[2 * (size_t)e + 1] = worst;
}

}  // namespace orlg
