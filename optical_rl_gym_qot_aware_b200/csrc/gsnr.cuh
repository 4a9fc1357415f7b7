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
};
__device__ __forceinline__ Chan chan_of(const GsnrParams& g, int start, int n, int se) {
  Chan c;
  c.B = (double)n * g.slot_hz;
  c.f = g.f0 + ((double)start + 0.5 * (double)n) * g.slot_hz;
  c.se = se;
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

// FROM_ACTION = false: `query` holds explicit candidates (path | core << 16, first slot, slots, se or 0).
// FROM_ACTION = true : `query` holds the env's own step actions for the pending request — RMSA-v0 (path, first slot)
// with the path's best modulation, RMCSA-v0 (path, modulation, core, first slot) — and the slot count comes from
// the pending bit rate through get_number_slots, so a heuristic's (or the fused hint's) action tensor can be
// evaluated as it is, without a host round trip.
template <bool FROM_ACTION>
__global__ void __launch_bounds__(128) gsnr_kernel(const Params p, const GsnrParams g, const int* __restrict__ query,
                                                   double* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= p.num_envs) return;
  const int e = warp;
  const double nan = __longlong_as_double(0x7FF8000000000000ll), inf = __longlong_as_double(0x7FF0000000000000ll);
  uint32_t m = p.meta[e];
  const int pair = (int)((m >> 16) & 0xFFu) * p.n_nodes + (int)(m >> 24);
  int4 q;
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
  PathRec crec{};
  bool ok = q.x >= 0 && path < p.k && core < p.cores;
  if (ok) {
    crec = load_path(p, pair, path);
    ok = crec.hops() > 0;
  }
  if (ok && FROM_ACTION) {
    q.w = mod >= 0 ? __ldg(p.mod_se + mod) : crec.se();
    q.z = number_slots((int)(m & 0xFFFFu), q.w, p);
  }
  ok = ok && q.y >= 0 && q.z >= 1 && q.y + q.z <= p.S;
  if (!ok) {
    if (lane == 0) out[2 * (size_t)e] = out[2 * (size_t)e + 1] = nan;
    return;
  }
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

}  // namespace orlg
