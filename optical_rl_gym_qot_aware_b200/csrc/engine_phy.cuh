// engine_phy.cuh — PhyRMSA-v0 on the stepping engine (optical_rl_gym/envs/phy_rmsa_env.py,
// physical layer: grooming=False, no periodic defragmentation).
//
// Channels take the place of slots: the tile holds available_channels[E][C] as packed rows
// (bit = 1 = free, phy_rmsa_env.py:185-189).  A service owns a *set* of whole channels on its
// path; it enters the release heap as one entry per channel (same key), so release is the
// generic "set one bit on every link of the path".
//
//   step (physical path)      phy_rmsa_env.py:272-309, _provision_path :544-623,
//                             _service_acceptance :767-778, _release_path :781-861
//   is_channel_free           :1029-1035  -> AND of the path's link rows
//   calculate_r_cut(modified) :1140-1193  -> T - 2 * (number of term links whose bit is set), per
//                             channel, counted for all channels at once with bit-sliced adders
//   _calculate_total_cuts     :1195-1203, calculate_total_r_spatial :1110-1121 (info metrics)
//   heuristics                sapff_rmsa :1676-1737, phy_aware_sapbm_rmsa :1254-1314,
//                             phy_aware_bmff_rmsa :1317-1372, phy_aware_bmfa_rmsa :1375-1438,
//                             phy_aware_faff_rmsa :1508-1569, and the RSS-ranked pair
//                             phy_aware_bmfa_rss_rmsa :1441-1505, phy_aware_faff_rss_rmsa :1572-1633
//
// Python's `sorted(row, key=(-mod, -frag))` + "take channels until the bit rate is covered" is
// evaluated without sorting: modulation levels are visited from high to low through per-level
// channel masks, inside a level the channels with the lowest neighbour count (highest r_cut) come
// first, ties in ascending channel order (stability of sorted()).
#pragma once
#include "engine_step.cuh"

namespace orlg {

constexpr int PHY_PLANES = 6;  // bit-sliced neighbour counters: up to 63 r_cut terms per path

template <int W>
__device__ __forceinline__ bool row_any(const Row<W>& r) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < W; i++) o |= r.w[i];
  return o != 0;
}
template <int W>
__device__ __forceinline__ bool row_bit(const Row<W>& r, int pos) {
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < W; i++) v = (pos >> 5) == i ? r.w[i] : v;
  return (v >> (pos & 31)) & 1u;
}
template <int W>
__device__ __forceinline__ void row_clear_bit(Row<W>& r, int pos) {
#pragma unroll
  for (int i = 0; i < W; i++)
    if ((pos >> 5) == i) r.w[i] &= ~(1u << (pos & 31));
}

template <int W>
__device__ __forceinline__ Row<W> level_mask(const Params& p, int table, int idp, int m) {
  const uint32_t* a = p.phy_level_mask + (((size_t)table * p.k + idp) * PHY_LEVELS + (m - 1)) * W;
  Row<W> r;
  if constexpr (W % 2 == 0) {  // rows are W words apart: 8-byte aligned for even W
#pragma unroll
    for (int i = 0; i < W; i += 2) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(a + i));
      r.w[i] = v.x;
      r.w[i + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < W; i++) r.w[i] = __ldg(a + i);
  }
  return r;
}

template <int W>
struct CutPlanes {
  Row<W> b[PHY_PLANES];
};

// carry[] enters the counters at plane q0; before it every counter is < 2^np, so the ripple stops at plane np - 1
template <int W>
__device__ __forceinline__ void ripple_add(CutPlanes<W>& c, uint32_t (&carry)[W], int q0, int np) {
#pragma unroll
  for (int q = 0; q < PHY_PLANES; q++) {
    if (q >= q0 && q < np) {
#pragma unroll
      for (int i = 0; i < W; i++) {
        uint32_t t2 = c.b[q].w[i] & carry[i];
        c.b[q].w[i] ^= carry[i];
        carry[i] = t2;
      }
    }
  }
}

// per-channel number of r_cut terms whose neighbour link has the channel free.  Bit-sliced counters; the rows of
// four terms at a time go through carry-save adders (two 3:2 steps into the ones, one into the twos), so only one
// carry row per four terms ripples through the planes above - about half the word operations of adding row by row.
// np bounds the planes that can hold a bit after t terms: it depends on the loop counter only.
template <int W, int BD, typename ENV>
__device__ __forceinline__ void count_terms(ENV& env, int pathg, CutPlanes<W>& c) {
  const Params& p = env.p;
#pragma unroll
  for (int q = 0; q < PHY_PLANES; q++)
#pragma unroll
    for (int i = 0; i < W; i++) c.b[q].w[i] = 0u;
  const int T = __ldg(p.phy_cut_count + pathg);
  const unsigned char* links = p.phy_cut_links + (size_t)pathg * p.phy_max_terms;
  int t = 0;
  for (; t + 4 <= T; t += 4) {
    const int l0 = __ldg(links + t), l1 = __ldg(links + t + 1), l2 = __ldg(links + t + 2), l3 = __ldg(links + t + 3);
    uint32_t carry[W];
#pragma unroll
    for (int i = 0; i < W; i++) {
      const uint32_t x0 = env.tile.word(0, l0, i), x1 = env.tile.word(0, l1, i), x2 = env.tile.word(0, l2, i),
                     x3 = env.tile.word(0, l3, i);
      uint32_t ones = c.b[0].w[i], twos = c.b[1].w[i];
      uint32_t u = ones ^ x0;
      const uint32_t ha = (ones & x0) | (u & x1);
      ones = u ^ x1;
      u = ones ^ x2;
      const uint32_t hb = (ones & x2) | (u & x3);
      ones = u ^ x3;
      u = twos ^ ha;
      carry[i] = (twos & ha) | (u & hb);
      twos = u ^ hb;
      c.b[0].w[i] = ones;
      c.b[1].w[i] = twos;
    }
    ripple_add<W>(c, carry, 2, 32 - __clz(t + 4));
  }
  if (t + 2 <= T) {
    const int l0 = __ldg(links + t), l1 = __ldg(links + t + 1);
    uint32_t carry[W];
#pragma unroll
    for (int i = 0; i < W; i++) {
      const uint32_t x0 = env.tile.word(0, l0, i), x1 = env.tile.word(0, l1, i);
      const uint32_t ones = c.b[0].w[i], u = ones ^ x0;
      carry[i] = (ones & x0) | (u & x1);
      c.b[0].w[i] = u ^ x1;
    }
    t += 2;
    ripple_add<W>(c, carry, 1, 32 - __clz(t));
  }
  if (t < T) {
    const int l = __ldg(links + t);
    uint32_t carry[W];
#pragma unroll
    for (int i = 0; i < W; i++) carry[i] = env.tile.word(0, l, i);
    ripple_add<W>(c, carry, 0, 32 - __clz(t + 1));
  }
}

// subset of `cand` with the smallest counter value
template <int W>
__device__ __forceinline__ Row<W> min_count_subset(const Row<W>& cand, const CutPlanes<W>& c) {
  Row<W> s = cand;
#pragma unroll
  for (int q = PHY_PLANES - 1; q >= 0; q--) {
    Row<W> t;
#pragma unroll
    for (int i = 0; i < W; i++) t.w[i] = s.w[i] & ~c.b[q].w[i];
    if (row_any<W>(t)) s = t;
  }
  return s;
}
template <int W>
__device__ __forceinline__ int count_at(const CutPlanes<W>& c, int ch) {
  int v = 0;
#pragma unroll
  for (int q = 0; q < PHY_PLANES; q++) v |= (int)row_bit<W>(c.b[q], ch) << q;
  return v;
}

// ---- RSS-ranked heuristics (phy_aware_bmfa_rss_rmsa :1441-1505, phy_aware_faff_rss_rmsa :1572-1633) ----
// The fragmentation key is calculate_r_spatial (:1085-1108), an f64: rss(column with the path's links
// taken) - rss(column), rss = sqrt(sum len^2) / (sum len + 1) over the runs of free links of the
// channel's column in link-index order.  A column is gathered into one CT word (E <= 8*sizeof(CT)
// links), its runs are peeled with x & (x + lowbit); sqrt and the division are IEEE, as in numpy.
template <typename CT>
__device__ __forceinline__ double column_rss(CT x) {
  int sum = 0, sum_sq = 0;
  while (x) {
    const CT y = x + (x & (CT)(0 - x));  // the carry runs through the lowest run of ones
    const CT run = x & ~y;
    const int len = sizeof(CT) == 8 ? __popcll((unsigned long long)run) : __popc((unsigned)run);
    sum += len;
    sum_sq += len * len;
    x &= y;
  }
  return sqrt((double)sum_sq) / (double)(sum + 1);
}

template <int W, int BD, typename CT, typename ENV>
__device__ __forceinline__ double r_spatial(ENV& env, int ch, CT path_links) {
  CT col = 0;
  const int wi = ch >> 5, b = ch & 31;
  for (int l = 0; l < env.p.E; l++) col |= (CT)((env.tile.word(0, l, wi) >> b) & 1u) << l;
  return column_rss<CT>(col & ~path_links) - column_rss<CT>(col);
}

template <int W, int BD, typename CT, typename ENV>
__device__ void phy_heuristic_rss(ENV& env, int hid, int* a) {
  const Params& p = env.p;
  const int pr = env.pair();
  const int table = __ldg(p.phy_table_id + pr);
  if (table < 0) return;
  const bool by_mod = hid == ORLG_H_PHY_BM_FA_RSS;  // key (-mod, -frag) :1469, else (-frag) :1601
  const double ninf = __longlong_as_double(0xFFF0000000000000ll);
  int best = -1, best_mod = -1;
  double best_frag = ninf;
  for (int idp = 0; idp < p.k; idp++) {
    PathRec rec = load_path(p, pr, idp);
    if (rec.hops() <= 0) continue;
    Row<W> A = env.tile.path_avail(rec, 0);
    if (!row_any<W>(A)) continue;
    CT links = 0;
    for_each_link(rec, [&](int l) { links |= (CT)1 << l; });
    // head of the sorted row (strict '>' in ascending channel order = stability of sorted()) and
    // what the whole row can carry
    int cap = 0, hmod = -1;
    double hfrag = ninf;
#pragma unroll
    for (int i = 0; i < W; i++) {
      uint32_t m = A.w[i];
      while (m) {
        const int ch = 32 * i + (__ffs((int)m) - 1);
        m &= m - 1;
        const int mod = __ldg(p.phy_mod + ((size_t)table * p.S + ch) * p.k + idp);
        cap += mod;
        if (by_mod && mod < hmod) continue;
        const double f = r_spatial<W, BD, CT>(env, ch, links);
        if (by_mod ? (mod > hmod || f > hfrag) : (f > hfrag)) {
          hmod = mod;
          hfrag = f;
        }
      }
    }
    if (cap * 100 < env.bit_rate) continue;  // the row would be tried, found short and popped
    if (by_mod ? (hmod > best_mod || (hmod == best_mod && hfrag > best_frag)) : (hfrag > best_frag)) {
      best = idp;
      best_mod = hmod;
      best_frag = hfrag;
    }
  }
  if (best < 0) return;
  // the first channels of the chosen row in sorted order: insertion into a sorted list of 8 held in
  // registers (strict '>' keeps the earlier channel in front of an equal key)
  PathRec rec = load_path(p, pr, best);
  Row<W> A = env.tile.path_avail(rec, 0);
  CT links = 0;
  for_each_link(rec, [&](int l) { links |= (CT)1 << l; });
  double fk[PHY_ADIM - 1];
  int fm[PHY_ADIM - 1], fc[PHY_ADIM - 1];
#pragma unroll
  for (int q = 0; q < PHY_ADIM - 1; q++) {
    fk[q] = ninf;
    fm[q] = -1;
    fc[q] = -1;
  }
#pragma unroll
  for (int i = 0; i < W; i++) {
    uint32_t m = A.w[i];
    while (m) {
      int ch = 32 * i + (__ffs((int)m) - 1);
      m &= m - 1;
      int mod = __ldg(p.phy_mod + ((size_t)table * p.S + ch) * p.k + best);
      if (by_mod && mod < fm[PHY_ADIM - 2]) continue;
      double f = r_spatial<W, BD, CT>(env, ch, links);
      bool shifting = false;  // once placed, the displaced entries move down unconditionally
#pragma unroll
      for (int q = 0; q < PHY_ADIM - 1; q++) {
        const bool before = shifting || fc[q] < 0 ||
                            (by_mod ? (mod > fm[q] || (mod == fm[q] && f > fk[q])) : (f > fk[q]));
        if (before) {
          shifting = true;
          const double tk = fk[q];
          const int tm = fm[q], tc = fc[q];
          fk[q] = f;
          fm[q] = mod;
          fc[q] = ch;
          f = tk;
          mod = tm;
          ch = tc;
          if (ch < 0) break;
        }
      }
    }
  }
  int need = env.bit_rate, n = 0;
#pragma unroll
  for (int q = 0; q < PHY_ADIM - 1; q++)
    if (need > 0 && fc[q] >= 0) {
      a[1 + n++] = fc[q];
      need -= fm[q] * 100;
    }
  if (need <= 0) {
    a[0] = best;
  } else {  // more than 8 channels would be needed: treated as blocked (documented cap)
#pragma unroll
    for (int i = 1; i < PHY_ADIM; i++) a[i] = -1;
  }
}

// use_existing_channels (phy_rmsa_env.py:1650-1673): the first path whose partially used channels can carry
// the request; their channels in list order, each giving all its free capacity.  a[0] = idp + 20 (:1380).
template <typename ENV>
__device__ bool phy_use_existing_channels(ENV& env, int* a) {
  const Params& p = env.p;
  const uint32_t pr = (uint32_t)env.pair();
  int sum[8];
#pragma unroll
  for (int q = 0; q < 8; q++) sum[q] = 0;
  for (int i = 0; i < env.cs_len; i++) {
    const uint2 v = env.cs_at(i);
    if ((v.x & 0xFFFFu) != pr || v.x == ENV::CS_DEAD) continue;
    const int idp = (int)((v.x >> 16) & 15u), fr = (int)((v.y >> 8) & 0xFFu);
#pragma unroll
    for (int q = 0; q < 8; q++) sum[q] += (idp == q) ? fr : 0;
  }
  int idp = -1;
#pragma unroll
  for (int q = 7; q >= 0; q--)
    if (q < p.k && sum[q] * 100 >= env.bit_rate) idp = q;
  if (idp < 0) return false;
  int need = env.bit_rate, n = 0;
  const uint32_t want = pr | ((uint32_t)idp << 16);
  for (int i = 0; i < env.cs_len && need > 0; i++) {
    const uint2 v = env.cs_at(i);
    const int fr = (int)((v.y >> 8) & 0xFFu);
    if ((v.x & 0xFFFFFu) != want || v.x == ENV::CS_DEAD || fr <= 0) continue;
    if (n == PHY_ADIM - 1) {  // documented cap of 8 channels per action: fall back to the physical layer
#pragma unroll
      for (int q = 1; q < PHY_ADIM; q++) a[q] = -1;
      return false;
    }
    a[1 + n++] = (int)(v.x >> 20);
    need -= fr * 100;
  }
  a[0] = idp + 20;
  return true;
}

// The reference heuristic's action for the pending request: a[0] = path or -2, a[1..8] channels.
template <int W, int BD, typename ENV>
__device__ void phy_heuristic(ENV& env, int hid, int* a) {
  const Params& p = env.p;
#pragma unroll
  for (int i = 0; i < PHY_ADIM; i++) a[i] = -1;
  a[0] = -2;
  // the virtual layer first: unconditionally in sapff/sapbm/bmff/faff/faff_rss (:1256,1321,1510,1574,1678),
  // under env.grooming in bmfa/bmfa_rss (:1377,1443)
  if (p.phy_grooming == 1 ||
      (p.phy_grooming == 2 && hid != ORLG_H_PHY_BM_FA_CUT && hid != ORLG_H_PHY_BM_FA_RSS))
    if (phy_use_existing_channels(env, a)) return;
  if (hid == ORLG_H_PHY_BM_FA_RSS || hid == ORLG_H_PHY_FA_FF_RSS) {
    if (p.E <= 32) phy_heuristic_rss<W, BD, uint32_t>(env, hid, a);
    else phy_heuristic_rss<W, BD, unsigned long long>(env, hid, a);  // E <= 64, checked by the host
    return;
  }
  const int pr = env.pair();
  const int table = __ldg(p.phy_table_id + pr);
  if (table < 0) return;
  int best = -1, best_mod = -1, best_frag = -(1 << 30);
  CutPlanes<W> cp_best;  // the counters of the best row so far (BM-FA-Cut, FA-FF): the channel walk below needs them again
#pragma unroll
  for (int q = 0; q < PHY_PLANES; q++)
#pragma unroll
    for (int i = 0; i < W; i++) cp_best.b[q].w[i] = 0u;
  for (int idp = 0; idp < p.k; idp++) {
    PathRec rec = load_path(p, pr, idp);
    if (rec.hops() <= 0) continue;
    Row<W> A = env.tile.path_avail(rec, 0);
    if (!row_any<W>(A)) continue;
    int cap = 0, head = 0;
    static_assert(PHY_LEVELS % 4 == 0, "the level masks are fetched four at a time");
#pragma unroll 1
    for (int m = PHY_LEVELS; m >= 1; m -= 4) {  // four masks in flight: one trip to L2 per four levels
      Row<W> L[4];
#pragma unroll
      for (int u = 0; u < 4; u++) L[u] = level_mask<W>(p, table, idp, m - u);
#pragma unroll
      for (int u = 0; u < 4; u++) {
        and_into(L[u], A);
        const int c = popcount<W>(L[u]);
        cap += (m - u) * c;
        if (head == 0 && c > 0) head = m - u;
      }
    }
    if (cap * 100 < env.bit_rate) continue;  // the row would be tried, found short and popped
    if (hid == ORLG_H_PHY_SAP_FF || hid == ORLG_H_PHY_SAP_BM) {  // first row that can carry the request
      best = idp;
      break;
    }
    if (hid == ORLG_H_PHY_BM_FF) {  // :1350-1355, strict '>' keeps the earliest row
      if (head > best_mod) {
        best = idp;
        best_mod = head;
      }
      continue;
    }
    // BM-FA-Cut :1414-1421: head of the row = (highest level, highest r_cut inside it);
    // FA-FF :1546-1550: head = highest r_cut over the whole row, levels ignored
    CutPlanes<W> cp;
    const int pathg = pr * p.k + idp;
    count_terms<W, BD>(env, pathg, cp);
    Row<W> cand = A;
    if (hid == ORLG_H_PHY_FA_FF) {
      head = 0;
    } else {
      cand = level_mask<W>(p, table, idp, head);
      and_into(cand, A);
    }
    Row<W> s = min_count_subset<W>(cand, cp);
    int frag = __ldg(p.phy_cut_count + pathg) - 2 * count_at<W>(cp, lowest_set<W>(s));
    if (head > best_mod || (head == best_mod && frag > best_frag)) {
      best = idp;
      best_mod = head;
      best_frag = frag;
      cp_best = cp;
    }
  }
  if (best < 0) return;
  // channels of the chosen row in the heuristic's order until the bit rate is covered
  PathRec rec = load_path(p, pr, best);
  Row<W> A = env.tile.path_avail(rec, 0);
  int need = env.bit_rate, n = 0;
  if (hid == ORLG_H_PHY_SAP_FF || hid == ORLG_H_PHY_FA_FF) {
    // SAP-FF: ascending channel number, whatever the level (:1704); FA-FF: descending r_cut, then
    // ascending channel (:1537); each channel counts with its own level
    while (need > 0 && n < PHY_ADIM - 1) {
      int ch = lowest_set<W>(hid == ORLG_H_PHY_FA_FF ? min_count_subset<W>(A, cp_best) : A);
      if (ch < 0) break;
      row_clear_bit<W>(A, ch);
      a[1 + n++] = ch;
      need -= (int)__ldg(p.phy_mod + ((size_t)table * p.S + ch) * p.k + best) * 100;
    }
  } else {
    // the scan above knows the highest level the row offers (BM-FF, BM-FA-Cut): the levels above it are empty
#pragma unroll 1
    for (int m = best_mod > 0 ? best_mod : PHY_LEVELS; m >= 1 && need > 0; m--) {
      Row<W> cand = level_mask<W>(p, table, best, m);
      and_into(cand, A);
      while (need > 0 && n < PHY_ADIM - 1 && row_any<W>(cand)) {
        Row<W> s = cand;
        if (hid == ORLG_H_PHY_BM_FA_CUT) s = min_count_subset<W>(cand, cp_best);
        int ch = lowest_set<W>(s);
        row_clear_bit<W>(cand, ch);
        a[1 + n++] = ch;
        need -= m * 100;
      }
    }
  }
  if (need <= 0) {
    a[0] = best;
  } else {  // more than 8 channels would be needed: treated as blocked (documented cap)
#pragma unroll
    for (int i = 1; i < PHY_ADIM; i++) a[i] = -1;
  }
}

// ---- periodic defragmentation: phy_rmsa_env.py:355-417 (step), _groom_defragmentation :703-733, _move_virtual
// :735-764, _move :662-697 --------------------------------------------------------------------------------------
// Runs inside the step kernel after _next_service, every phy_defrag_period processed services (all envs of a batch
// reach that count in the same step, so the warp does not diverge on it).  running_services is the order list
// with tombstones kept next to the service table; Python's index-based list iteration over a list whose current
// element is removed and re-appended (the next element is skipped, the moved one is met again at the end) is
// reproduced on it.

// number of r_cut terms of the path whose neighbour link has the channel free
template <int W, int BD, typename ENV>
__device__ int count_terms_at(ENV& env, int pathg, int ch) {
  const Params& p = env.p;
  const int T = __ldg(p.phy_cut_count + pathg), wi = ch >> 5, b = ch & 31;
  const unsigned char* links = p.phy_cut_links + (size_t)pathg * p.phy_max_terms;
  int c = 0;
  for (int t = 0; t < T; t++) c += (int)((env.tile.word(0, __ldg(links + t), wi) >> b) & 1u);
  return c;
}

template <int W, int BD, typename CT, typename ENV>
__device__ double defrag_gain_rss(ENV& env, int ch, CT path_links) {  // calculate_r_spatial(ch, links, True) :1085-1108
  CT col = 0;
  const int wi = ch >> 5, b = ch & 31;
  for (int l = 0; l < env.p.E; l++) col |= (CT)((env.tile.word(0, l, wi) >> b) & 1u) << l;
  return column_rss<CT>(col | path_links) - column_rss<CT>(col);
}

template <int W, int BD, typename CT, typename ENV>
__device__ __noinline__ void phy_defragmentation_t(ENV& env) {
  const Params& p = env.p;
  typedef typename ENV::PhySvc Svc;
  if (p.phy_run_cap - env.run_len < 2 * p.phy_number_moves + 8) env.run_squeeze();  // room for this cycle's re-appends
  // ---- _groom_defragmentation :703-733 ----
  int gm = 0;
  bool full = false;
  for (int pos = 0; pos < env.run_len && !full; pos++) {  // for service in running_services (live list)
    const int slot = env.run_at(pos);
    if (slot == 0xFFFF) continue;
    Svc sv = env.svc_load(slot);
    bool moved = false;
    for (int ci = 0; ci < sv.nch && !full; ci++) {  // for channel in service.channels (live list)
      const int c_ch = sv.ch[ci], c_share = sv.share[ci], c_cap = sv.cap[ci];
      if (c_share != c_cap) {
        const int at = env.cs_find(ENV::cs_key(sv.pair, sv.path, c_ch));  // corresponding_channel
        if (at >= 0 && (int)(env.cs_at(at).y & 0xFFu) == c_share) {      // nobody else uses it: it can move
          const uint32_t want = (uint32_t)sv.pair | ((uint32_t)sv.path << 16);
          for (int j = 0; j < env.cs_len; j++) {  // for target_channel in channel_state[key]
            const uint2 v = env.cs_at(j);
            if (v.x == ENV::CS_DEAD || (v.x & 0xFFFFFu) != want || j == at) continue;
            const int used = (int)(v.y & 0xFFu), fr = (int)((v.y >> 8) & 0xFFu), cap = (int)(v.y >> 16), tch = (int)(v.x >> 20);
            if (fr >= c_share) {
              env.cs_remove(j);  // two tuples removed, one appended :717-724
              env.cs_remove(at);
              env.cs_append(ENV::cs_key(sv.pair, sv.path, tch), used + c_share, fr - c_share, cap);
              PathRec rec = load_path(p, sv.pair, sv.path);  // _move_virtual :735-764
              env.tile.apply_bit(rec, 0, c_ch, true);
              env.run_remove(sv.run_pos);
              sv.run_pos = env.run_append(slot);
              sv.remove_channel(ci);
              sv.append_channel(tch, c_share, cap);
              moved = true;
              gm++;
              break;
            }
          }
        }
      }
      if (gm == p.phy_number_moves) full = true;  // :731
    }
    if (moved) {
      env.svc_store(slot, sv);
      if (!full) {  // the element that slid into this index is skipped by Python's list iterator
        int q = pos + 1;
        while (q < env.run_len && env.run_at(q) == 0xFFFFu) q++;
        pos = q;
      }
    }
  }
  env.mv[1] = gm;  // counted_moves_groom :357
  if (gm > p.phy_number_moves) return;
  // ---- candidates :360-380: whole channels whose release gains fragmentation metric, running_services order ----
  double2* cd = p.phy_cand + (size_t)env.e * p.phy_cand_cap;
  uint32_t* cid = p.phy_cand_id + (size_t)env.e * p.phy_cand_cap;
  int ncand = 0;
  for (int pos = 0; pos < env.run_len; pos++) {
    const int slot = env.run_at(pos);
    if (slot == 0xFFFF) continue;
    const Svc sv = env.svc_load(slot);
    const int pathg = sv.pair * p.k + sv.path;
    CT links = 0;
    if (p.phy_metric != 0) {
      PathRec rec = load_path(p, sv.pair, sv.path);
      for_each_link(rec, [&](int l) { links |= (CT)1 << l; });
    }
    for (int ci = 0; ci < sv.nch; ci++) {
      if (sv.share[ci] != sv.cap[ci]) continue;
      const double diff = p.phy_metric == 0
                              ? (double)(2 * count_terms_at<W, BD>(env, pathg, sv.ch[ci]) - __ldg(p.phy_cut_count + pathg))
                              : defrag_gain_rss<W, BD, CT>(env, sv.ch[ci], links);
      if (diff > 0) {
        if (ncand < p.phy_cand_cap) {
          cd[ncand] = make_double2(diff, env.now - sv.arrival);
          cid[ncand] = (uint32_t)slot | ((uint32_t)sv.ch[ci] << 16);
          ncand++;
        } else {
          env.cnt[CNT_FAULT] |= 128;
        }
      }
    }
  }
  // ---- reallocation in sorted(key=(-gain, -age)) order (stable: the earliest of equal keys first) :381-415 ----
  int num_moves = 0;
  for (;;) {
    int best = -1;
    double bx = 0.0, by = 0.0;
    for (int i = 0; i < ncand; i++) {
      const double2 d = cd[i];
      if (d.x > 0.0 && (best < 0 || d.x > bx || (d.x == bx && d.y > by))) {
        best = i;
        bx = d.x;
        by = d.y;
      }
    }
    if (best < 0) break;
    cd[best].x = -1.0;  // taken
    const int slot = (int)(cid[best] & 0xFFFFu), cch = (int)(cid[best] >> 16);
    Svc sv = env.svc_load(slot);
    int ci = -1;
    for (int q = 0; q < sv.nch; q++)
      if (ci < 0 && sv.ch[q] == cch && sv.share[q] == sv.cap[q]) ci = q;
    if (ci >= 0) {
      // idp: the candidate's path looked up among the CURRENT request's k paths (dataclass equality): its own index
      // when both connect the same two nodes, else the loop variable's last value k - 1 (:386-391)
      const int sa = sv.pair / p.n_nodes, sb = sv.pair % p.n_nodes;
      const bool same = (sa == env.src && sb == env.dst) || (sa == env.dst && sb == env.src);
      const int idp = same ? sv.path : p.k - 1;
      const int table = __ldg(p.phy_table_id + sv.pair);
      const int level = (int)__ldg(p.phy_mod + ((size_t)table * p.S + cch) * p.k + idp);
      const int pathg = sv.pair * p.k + sv.path;
      PathRec rec = load_path(p, sv.pair, sv.path);
      Row<W> opt = level_mask<W>(p, table, idp, level);
      and_into(opt, env.tile.path_avail(rec, 0));  // free on the path and of the same modulation level
      if (row_any<W>(opt)) {
        int nch_ = -1;
        double metric = 0.0;
        if (p.phy_metric == 0) {  // best of sorted(key=(-r_cut, channel)): fewest free neighbours, lowest channel
          CutPlanes<W> cp;
          count_terms<W, BD>(env, pathg, cp);
          nch_ = lowest_set<W>(min_count_subset<W>(opt, cp));
          metric = (double)(__ldg(p.phy_cut_count + pathg) - 2 * count_at<W>(cp, nch_));
        } else {
          CT links = 0;
          for_each_link(rec, [&](int l) { links |= (CT)1 << l; });
#pragma unroll 1
          for (int i = 0; i < W; i++) {
            uint32_t mword = opt.w[i];
            while (mword) {
              const int ch = 32 * i + (__ffs((int)mword) - 1);
              mword &= mword - 1;
              const double f = r_spatial<W, BD, CT>(env, ch, links);
              if (nch_ < 0 || f > metric) {
                nch_ = ch;
                metric = f;
              }
            }
          }
        }
        if (-1.0 * metric < bx) {  // :401-405 it pays: _move :662-697
          env.tile.apply_bit(rec, 0, nch_, false);
          env.tile.apply_bit(rec, 0, cch, true);
          env.run_remove(sv.run_pos);
          sv.run_pos = env.run_append(slot);
          const int sh = sv.share[ci], cpv = sv.cap[ci];
          sv.remove_channel(ci);
          sv.append_channel(nch_, sh, cpv);
          env.svc_store(slot, sv);
          num_moves++;
          env.mv[0]++;
        }
      }
    }
    if ((long long)num_moves + env.mv[1] > (long long)p.phy_number_moves) break;  // :414
  }
  if (num_moves != 0) env.mv[2]++;
}

template <int W, int BD, typename ENV>
__device__ __forceinline__ void phy_defragmentation(ENV& env) {
  if (env.p.E <= 32) phy_defragmentation_t<W, BD, uint32_t>(env);
  else phy_defragmentation_t<W, BD, unsigned long long>(env);
}

// PhyRMSAEnv.step (phy_rmsa_env.py:272-309).  The action carries channel numbers; the (used, free) fields of
// the reference's channel tuples are what its heuristics put there: every channel is used up to its capacity
// (physical layer) or its free part (virtual layer) until the bit rate is covered, the last one takes the rest.
template <int W, int BD, typename ENV>
__device__ Decision phy_decide(ENV& env, const int* a) {
  const Params& p = env.p;
  Decision d{0, -1, 0, -1, 0};
  const int path = a[0], pr = env.pair();
  int nch = 0;
  while (nch < PHY_ADIM - 1 && a[1 + nch] >= 0) nch++;
  const double key = env.now + env.holding;
  if (p.phy_grooming && path >= 20 && path < 20 + p.k && nch > 0) {
    // virtual layer :280-288, _provision_virtual_path :625-659
    const int idp = path - 20;
    int at[PHY_ADIM - 1], share[PHY_ADIM - 1];
    uint32_t yv[PHY_ADIM - 1];
#pragma unroll
    for (int i = 0; i < PHY_ADIM - 1; i++) at[i] = -1;
    for (int j = 0; j < env.cs_len; j++) {  // one pass finds every listed channel
      const uint2 v = env.cs_at(j);
      if ((v.x & 0xFFFFFu) != ((uint32_t)pr | ((uint32_t)idp << 16)) || v.x == ENV::CS_DEAD) continue;
      const int ch = (int)(v.x >> 20);
#pragma unroll
      for (int i = 0; i < PHY_ADIM - 1; i++)
        if (i < nch && a[1 + i] == ch && at[i] < 0) {
          at[i] = j;
          yv[i] = v.y;
        }
    }
    bool ok = true;
    int need = env.bit_rate / 100;
#pragma unroll
    for (int i = 0; i < PHY_ADIM - 1; i++)
      if (i < nch) {
#pragma unroll
        for (int q = 0; q < i; q++) ok = ok && a[1 + q] != a[1 + i];
        const int fr = at[i] >= 0 ? (int)((yv[i] >> 8) & 0xFFu) : 0;
        ok = ok && fr > 0 && need > 0;
        share[i] = fr < need ? fr : need;
        need -= share[i];
      }
    if (!ok) return d;
    if (env.sv_free_n == 0 || env.heap_len >= p.heap_cap) {  // no row for the service: refuse, state untouched
      env.cnt[CNT_FAULT] |= 1;
      return d;
    }
    typename ENV::PhySvc sv;
    sv.arrival = env.now;
    sv.pair = pr;
    sv.path = idp;
    sv.nch = nch;
    sv.run_pos = 0;
#pragma unroll
    for (int i = 0; i < PHY_ADIM - 1; i++) sv.ch[i] = sv.share[i] = sv.cap[i] = 0;
#pragma unroll
    for (int i = 0; i < PHY_ADIM - 1; i++)
      if (i < nch) env.cs_remove(at[i]);
#pragma unroll
    for (int i = 0; i < PHY_ADIM - 1; i++)
      if (i < nch) {  // :643-645 the updated tuple goes to the end of the list
        const int used = (int)(yv[i] & 0xFFu), fr = (int)((yv[i] >> 8) & 0xFFu), cap = (int)(yv[i] >> 16);
        env.cs_append(ENV::cs_key(pr, idp, a[1 + i]), used + share[i], fr - share[i], cap);
        sv.ch[i] = a[1 + i];
        sv.share[i] = share[i];
        sv.cap[i] = cap;
      }
    env.svc_provision(sv, key);
    env.cnt[CNT_ACC]++;  // _service_acceptance(True) :767-778
    env.cnt[CNT_EP_ACC]++;
    env.cnt[CNT_BR_PROV] += env.bit_rate;
    env.cnt[CNT_EP_BR_PROV] += env.bit_rate;
    return Decision{1, path, 0, a[1], nch};
  }
  if (path < 0 || path >= p.k || nch == 0) return d;
  const int table = __ldg(p.phy_table_id + pr);
  PathRec rec = load_path(p, pr, path);
  if (table < 0 || rec.hops() <= 0) return d;
  Row<W> A = env.tile.path_avail(rec, 0);
  bool ok = true;
  int need = env.bit_rate / 100;
  for (int i = 0; i < nch; i++) {  // is_path_free_on_channels :1019-1027 (+ range / duplicate checks)
    int ch = a[1 + i];
    ok = ok && ch < p.S && row_bit<W>(A, ch < p.S ? ch : 0);
    if (ok) row_clear_bit<W>(A, ch);
    if (ok && p.phy_grooming) {  // with channel_state kept, every listed channel must carry part of the request
      ok = need > 0;
      need -= (int)__ldg(p.phy_mod + ((size_t)table * p.S + ch) * p.k + path);
    }
  }
  if (!ok) return d;
  if (env.sv_free_n == 0 || env.heap_len >= p.heap_cap) {  // no row for the service: refuse, state untouched
    env.cnt[CNT_FAULT] |= 1;
    return d;
  }
  typename ENV::PhySvc sv;
  sv.arrival = env.now;
  sv.pair = pr;
  sv.path = path;
  sv.nch = nch;
  sv.run_pos = 0;
#pragma unroll
  for (int i = 0; i < PHY_ADIM - 1; i++) sv.ch[i] = sv.share[i] = sv.cap[i] = 0;
  need = env.bit_rate / 100;
  for (int i = 0; i < nch; i++) {  // _provision_path :576-601
    const int ch = a[1 + i];
    env.tile.apply_bit(rec, 0, ch, false);
    const size_t tix = ((size_t)table * p.S + ch) * p.k + path;
    const int level = __ldg(p.phy_mod + tix);
    env.fsum[1] += __ldg(p.phy_gsnr + tix);
    env.isum[1] += level;
    env.isum[2]++;
    int share = 0;
    if (p.phy_grooming) {
      share = level < need ? level : need;
      need -= share;
      if (share != level) env.cs_append(ENV::cs_key(pr, path, ch), share, level - share, level);  // :599-601
    }
#pragma unroll
    for (int q = 0; q < PHY_ADIM - 1; q++)
      if (q == i) {
        sv.ch[q] = ch;
        sv.share[q] = share;
        sv.cap[q] = level;
      }
  }
  env.svc_provision(sv, key);
  env.cnt[CNT_ACC]++;  // _service_acceptance(False) :767-778
  env.cnt[CNT_EP_ACC]++;
  env.fsum[0] += __ldg(p.path_length + (size_t)pr * p.k + path);
  env.isum[0] += path + 1;
  env.isum[3]++;
  env.cnt[CNT_BR_PROV] += env.bit_rate;
  env.cnt[CNT_EP_BR_PROV] += env.bit_rate;
  d = Decision{1, path, 0, a[1], nch};
  return d;
}

// number_cuts_total and rss_total_metric of the step's info dict (phy_rmsa_env.py:313-314),
// evaluated where the reference evaluates them: after the provisioning, before the releases.
template <int W, int BD, typename ENV>
__device__ void phy_info(ENV& env, double* o, long long fs, bool want_rss) {
  const Params& p = env.p;
  // _calculate_total_cuts :1195-1203: runs of free links (in link-index order) per channel
  long long cuts = 0;
  uint32_t prev[W];
#pragma unroll
  for (int i = 0; i < W; i++) prev[i] = 0u;
  for (int l = 0; l < p.E; l++) {
#pragma unroll
    for (int i = 0; i < W; i++) {
      const uint32_t cur = env.tile.word(0, l, i);
      cuts += __popc(cur & ~prev[i]);
      prev[i] = cur;
    }
  }
  o[0] = (double)cuts / (double)p.S;
  if (!want_rss) return;
  // calculate_total_r_spatial :1110-1121: sqrt(sum len^2) / (sum len + 1) over the free runs
  double r_spatial = 0.0;
  for (int i = 0; i < W; i++) {
    for (int b = 0; b < 32; b++) {
      const int ch = 32 * i + b;
      if (ch >= p.S) break;
      long long sum_sq = 0;
      int sum = 0, run = 0;
      for (int l = 0; l < p.E; l++) {
        if ((env.tile.word(0, l, i) >> b) & 1u) {
          run++;
        } else {
          sum_sq += (long long)run * run;
          sum += run;
          run = 0;
        }
      }
      sum_sq += (long long)run * run;
      sum += run;
      r_spatial += sqrt((double)sum_sq) / (double)(sum + 1);
    }
  }
  o[fs] = r_spatial / (double)p.S;
}

}  // namespace orlg
