// engine_env.cuh — per-thread env context: release heap, request draw, decision.
#pragma once
#include "engine.cuh"

namespace orlg {

constexpr int HEAP_D = 8;    // 8-ary implicit min-heap keyed by release time: one 128 B line per child group
constexpr int HEAP_PAD = 7;  // logical index i lives in slot i+7 -> the children of i are slots 8i+8..8i+15 (line aligned)
constexpr int QOT_REL_CAP_ = 32;  // = QOT_REL_CAP (qot.cuh)
constexpr int HEAP_TAIL = 8; // slack after the last entry so that a whole child group can be read unconditionally

// payload of a heap entry: what _release_path needs (rmsa_env.py:515-535) plus the spectral efficiency of
// the service's modulation (the GN model's per-interferer correction needs it) and, for engines with the
// GN-model QoT gate, the service's row in the per-env table of running cross-channel sums (qot.cuh).
// bits: pair 0-15 | path 16-19 | core 20-24 | first slot 25-33 | slots 34-42 | se 43-46 | row 47-60
__device__ __forceinline__ unsigned long long pack_service(int pair, int path, int core, int start, int n, int se,
                                                           int row = 0) {
  return (unsigned long long)pair | ((unsigned long long)(path & 15) << 16) | ((unsigned long long)(core & 31) << 20) |
         ((unsigned long long)(start & 511) << 25) | ((unsigned long long)(n & 511) << 34) |
         ((unsigned long long)(se & 15) << 43) | ((unsigned long long)(row & 0x3FFF) << 47);
}
__device__ __forceinline__ int svc_pair(unsigned long long s) { return (int)(s & 0xFFFF); }
__device__ __forceinline__ int svc_path(unsigned long long s) { return (int)((s >> 16) & 15); }
__device__ __forceinline__ int svc_core(unsigned long long s) { return (int)((s >> 20) & 31); }
__device__ __forceinline__ int svc_start(unsigned long long s) { return (int)((s >> 25) & 511); }
__device__ __forceinline__ int svc_slots(unsigned long long s) { return (int)((s >> 34) & 511); }
__device__ __forceinline__ int svc_se(unsigned long long s) { return (int)((s >> 43) & 15); }
__device__ __forceinline__ int svc_slot(unsigned long long s) { return (int)((s >> 47) & 0x3FFF); }

// INFO: the instantiation keeps the info-dict state of RMSAEnv.step (only the kernels launched for engines
// created with info_stats carry that code)
template <int W, int BD, bool PHY = false, bool INFO = false>
struct Env {
  // the INFO instantiations keep RMSAEnv's info-dict state; for PhyRMSA the same flag selects the kernels that carry
  // the periodic defragmentation (engine_phy.cuh), so engines without it run kernels whose env context never
  // leaves the registers
  static constexpr bool INFO_STATS = INFO && !PHY;
  static constexpr bool PHY_DEFRAG = INFO && PHY;
  const Params& p;
  Tile<W, BD> tile;
  int e;  // local env index
  ulonglong2* heap;
  double now, holding, next_rel;
  int bit_rate, src, dst, heap_len;
  long long cnt[CNT_N];
  double fsum[2];      // PhyRMSA episode sums (phy_rmsa_env.py:431-439)
  long long isum[4];
  int cs_len, cs_dead; // PhyRMSA channel_state list: slots in use (tombstones included), tombstones
  long long slot_hops; // info_stats: sum of number_slots * hops over the running services
  int rel_n = 0;       // QoT gate: services released by this step, handed to the next qot_kernel launch
  // PhyRMSA service table (one queue entry per service; its channels live in the service's record)
  int sv_free_n = 0;            // free rows of the table
  int run_len = 0, run_dead = 0;  // running_services order list (defragmentation only): slots in use, tombstones
  long long mv[3] = {0, 0, 0};  // counted_moves, counted_moves_groom, counted_defrag_cycles (phy_rmsa_env.py:110-112)

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
    slot_hops = 0;
    rel_n = p.qot ? p.qot_rel_n[e] : 0;  // releases not yet seen by qot_kernel survive an episode reset
    if constexpr (INFO_STATS) slot_hops = p.slot_hops[e];
    if constexpr (PHY) {
      fsum[0] = p.phy_f[e];
      fsum[1] = p.phy_f[(size_t)p.NP + e];
#pragma unroll
      for (int i = 0; i < 4; i++) isum[i] = p.phy_i[(size_t)i * p.NP + e];
      load_cs_len();
      sv_free_n = p.phy_free_n[e];
      if (p.phy_defrag_period > 0) {
        const uint32_t v = p.phy_run_len[e];
        run_len = (int)(v & 0xFFFFu);
        run_dead = (int)(v >> 16);
#pragma unroll
        for (int i = 0; i < 3; i++) mv[i] = p.phy_moves[(size_t)i * p.NP + e];
      }
    }
  }
  // ---- info dict of RMSAEnv.step (rmsa_env.py:276-332), kept only with p.info_stats ----
  __device__ __forceinline__ double& lnk(int f, int l) const { return p.lnk_stat[((size_t)f * p.E + l) * p.NP + e]; }
  // used slots of a link row: bits [0, S) that are not free
  __device__ __forceinline__ Row<W> used_row(int l) {
    Row<W> u;
#pragma unroll
    for (int i = 0; i < W; i++) u.w[i] = ~tile.word(0, l, i) & mask_word(i, 0, p.S);
    return u;
  }
  // _update_link_stats :562-641: time-weighted utilization and link compactness (the used blocks of a row are
  // the runs of `used`; between lambda_min and lambda_max there are as many used runs as the row has)
  __device__ __forceinline__ void info_update_link(int l) {
    const double last_update = lnk(0, l);
    const double time_diff = now - last_update;
    if (now > 0) {
      const Row<W> u = used_row(l);
      const int used = popcount<W>(u), blocks = count_runs<W>(u);
      const double cur_util = (double)used / (double)p.S;
      lnk(1, l) = ((lnk(1, l) * last_update) + (cur_util * time_diff)) / now;
      double cur = 0.0;
      if (used < p.S) {  // np.sum(slot_allocation) > 0
        cur = 1.0;
        if (blocks > 1) {
          const int span = highest_set<W>(u) + 1 - lowest_set<W>(u);
          cur = ((double)span / (double)used) * (1.0 / (double)blocks);
        }
      }
      lnk(2, l) = ((lnk(2, l) * last_update) + (cur * time_diff)) / now;
    }
    lnk(0, l) = now;
  }
  // _get_network_compactness :806-851
  __device__ __forceinline__ double info_network_compactness() {
    long long sum_occupied = 0, sum_unused_blocks = 0;
    for (int l = 0; l < p.E; l++) {
      const Row<W> u = used_row(l);
      const int blocks = count_runs<W>(u);
      if (blocks > 1) {
        sum_occupied += highest_set<W>(u) + 1 - lowest_set<W>(u);
        sum_unused_blocks += blocks - 1;  // one free run between two consecutive used blocks
      }
    }
    if (sum_unused_blocks > 0)
      return ((double)sum_occupied / (double)slot_hops) * ((double)p.E / (double)sum_unused_blocks);
    return 1.0;
  }
  __device__ __forceinline__ int bit_rate_index(int br) const {
    int idx = -1;
    for (int i = 0; i < p.n_bit_rates; i++)
      if (__ldg(p.bit_rates + i) == br) idx = idx < 0 ? i : idx;
    return idx;
  }
  __device__ __forceinline__ void info_count(int which, int br) {
    const int bi = bit_rate_index(br);
    if (bi >= 0) p.br_hist[((size_t)which * p.n_bit_rates + bi) * p.NP + e]++;
  }
  // np.mean of the per-link values in edge order: numpy's pairwise summation (8 accumulators, n <= 128)
  __device__ __forceinline__ double info_link_mean(int f) const {
    const int n = p.E;
    double res;
    if (n < 8) {
      res = 0.0;
      for (int i = 0; i < n; i++) res += lnk(f, i);
    } else {
      double r[8];
#pragma unroll
      for (int j = 0; j < 8; j++) r[j] = lnk(f, j);
      int i = 8;
      for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int j = 0; j < 8; j++) r[j] += lnk(f, i + j);
      }
      res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
      for (; i < n; i++) res += lnk(f, i);
    }
    return res / (double)n;
  }
  // the info dict, taken after the provisioning and before _next_service (:276-332)
  __device__ __forceinline__ void info_write(double* o, double previous_compactness) {
    const double cur = info_network_compactness();
    o[0] = (double)(cnt[CNT_PROC] - cnt[CNT_ACC]) / (double)cnt[CNT_PROC];
    o[1] = (double)(cnt[CNT_EP_PROC] - cnt[CNT_EP_ACC]) / (double)cnt[CNT_EP_PROC];
    o[2] = (double)(cnt[CNT_BR_REQ] - cnt[CNT_BR_PROV]) / (double)cnt[CNT_BR_REQ];
    o[3] = (double)(cnt[CNT_EP_BR_REQ] - cnt[CNT_EP_BR_PROV]) / (double)cnt[CNT_EP_BR_REQ];
    o[4] = cur;
    o[5] = previous_compactness - cur;
    o[6] = info_link_mean(2);
    o[7] = info_link_mean(1);
    double mx = 0.0, mn = 0.0;
    for (int i = 0; i < p.n_bit_rates; i++) {  // :277-289, fairness :326-328
      const int rq = p.br_hist[((size_t)i) * p.NP + e], pv = p.br_hist[((size_t)p.n_bit_rates + i) * p.NP + e];
      const double b = rq > 0 ? (double)(rq - pv) / (double)rq : 0.0;
      o[9 + i] = b;
      if (i == 0 || b > mx) mx = b;
      if (i == 0 || b < mn) mn = b;
    }
    o[8] = mx - mn;
  }

  __device__ __forceinline__ void load_cs_len() {
    cs_len = cs_dead = 0;
    if (PHY && p.phy_grooming) {
      uint32_t v = p.phy_cs_len[e];
      cs_len = (int)(v & 0xFFFFu);
      cs_dead = (int)(v >> 16);
    }
  }
  // only the pending request (enough for the heuristics and the observation)
  __device__ __forceinline__ void load_request() {
    uint32_t m = p.meta[e];
    bit_rate = (int)(m & 0xFFFFu);
    src = (int)((m >> 16) & 0xFFu);
    dst = (int)(m >> 24);
    if constexpr (PHY) load_cs_len();
  }
  __device__ __forceinline__ void store_scalars() {
    p.now[e] = now;
    p.holding[e] = holding;
    p.meta[e] = (uint32_t)bit_rate | ((uint32_t)src << 16) | ((uint32_t)dst << 24);
    p.next_rel[e] = next_rel;
    p.heap_len[e] = heap_len;
#pragma unroll
    for (int i = 0; i < CNT_N; i++) p.counters[(size_t)i * p.NP + e] = cnt[i];
    if constexpr (INFO_STATS) p.slot_hops[e] = slot_hops;
    if (p.qot) p.qot_rel_n[e] = rel_n;
    if constexpr (PHY) {
      p.phy_f[e] = fsum[0];
      p.phy_f[(size_t)p.NP + e] = fsum[1];
#pragma unroll
      for (int i = 0; i < 4; i++) p.phy_i[(size_t)i * p.NP + e] = isum[i];
      if (p.phy_grooming) p.phy_cs_len[e] = (uint32_t)cs_len | ((uint32_t)cs_dead << 16);
      p.phy_free_n[e] = sv_free_n;
      if (p.phy_defrag_period > 0) {
        p.phy_run_len[e] = (uint32_t)run_len | ((uint32_t)run_dead << 16);
#pragma unroll
        for (int i = 0; i < 3; i++) p.phy_moves[(size_t)i * p.NP + e] = mv[i];
      }
    }
  }

  // ---- PhyRMSA channel_state (phy_rmsa_env.py:599-601, 625-659, 801-846) ----
  // The reference keeps one Python list of (channel, used, free, capacity) per (src_id, dst_id, idp) and
  // updates a tuple by remove + append.  Here every env has one list for all keys in append order; a removed
  // entry becomes a tombstone and the list is squeezed when tombstones pile up, which keeps the order.
  static constexpr uint32_t CS_DEAD = 0xFFFFFFFFu;
  __device__ __forceinline__ static uint32_t cs_key(int pr, int idp, int ch) {
    return (uint32_t)pr | ((uint32_t)idp << 16) | ((uint32_t)ch << 20);
  }
  __device__ __forceinline__ uint2& cs_at(int i) const { return p.phy_cs[(size_t)i * p.NP + e]; }
  __device__ int cs_find(uint32_t key) const {
    for (int i = 0; i < cs_len; i++)
      if (cs_at(i).x == key) return i;
    return -1;
  }
  __device__ __forceinline__ void cs_remove(int i) {
    cs_at(i).x = CS_DEAD;
    cs_dead++;
  }
  __device__ void cs_squeeze() {
    int j = 0;
    for (int i = 0; i < cs_len; i++) {
      uint2 v = cs_at(i);
      if (v.x != CS_DEAD) {
        if (j != i) cs_at(j) = v;
        j++;
      }
    }
    cs_len = j;
    cs_dead = 0;
  }
  __device__ void cs_append(uint32_t key, int used, int free_, int cap) {
    if (cs_len == p.phy_cs_cap || cs_dead >= 32 + ((cs_len - cs_dead) >> 1)) cs_squeeze();
    if (cs_len == p.phy_cs_cap) {
      cnt[CNT_FAULT] |= 4;  // more partially used channels than live lightpaths can have: capacity misconfigured
      return;
    }
    cs_at(cs_len++) = make_uint2(key, (uint32_t)used | ((uint32_t)free_ << 8) | ((uint32_t)cap << 16));
  }
  // ---- PhyRMSA service table: Service objects with their channel tuples (phy_rmsa_env.py:610-621) ----
  // 48-byte records [slot][3 x uint4]: arrival f64 | pair u16, path u8, nch u8 | position in the order list;
  // channel number u16 x 8; share u8 x 8 (tuple field [1]) and capacity u8 x 8 (field [3]), units of 100 Gb/s.
  struct PhySvc {
    double arrival;
    int pair, path, nch, run_pos;
    int ch[8], share[8], cap[8];
    __device__ __forceinline__ void remove_channel(int ci) {  // service.channels.remove(old_channel)
#pragma unroll
      for (int i = 0; i < 7; i++)
        if (i >= ci) {
          ch[i] = ch[i + 1];
          share[i] = share[i + 1];
          cap[i] = cap[i + 1];
        }
      nch--;
    }
    __device__ __forceinline__ void append_channel(int c, int sh, int cp) {  // service.channels.append(new_channel)
#pragma unroll
      for (int i = 0; i < 8; i++)
        if (i == nch) {
          ch[i] = c;
          share[i] = sh;
          cap[i] = cp;
        }
      nch++;
    }
  };
  __device__ __forceinline__ uint4* svc_rec(int slot) const { return p.phy_svc + ((size_t)e * p.phy_svc_cap + slot) * 3; }
  __device__ PhySvc svc_load(int slot) const {
    const uint4* r = svc_rec(slot);
    const uint4 a = r[0], b = r[1], c = r[2];
    PhySvc v;
    v.arrival = __hiloint2double((int)a.y, (int)a.x);
    v.pair = (int)(a.z & 0xFFFFu);
    v.path = (int)((a.z >> 16) & 0xFFu);
    v.nch = (int)(a.z >> 24);
    v.run_pos = (int)a.w;
    const uint32_t cw[4] = {b.x, b.y, b.z, b.w}, sw[2] = {c.x, c.y}, pw[2] = {c.z, c.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
      v.ch[i] = (int)((cw[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
      v.share[i] = (int)((sw[i >> 2] >> (8 * (i & 3))) & 0xFFu);
      v.cap[i] = (int)((pw[i >> 2] >> (8 * (i & 3))) & 0xFFu);
    }
    return v;
  }
  __device__ void svc_store(int slot, const PhySvc& v) const {
    uint32_t cw[4] = {0, 0, 0, 0}, sw[2] = {0, 0}, pw[2] = {0, 0};
#pragma unroll
    for (int i = 0; i < 8; i++) {
      cw[i >> 1] |= (uint32_t)(v.ch[i] & 0xFFFF) << (16 * (i & 1));
      sw[i >> 2] |= (uint32_t)(v.share[i] & 0xFF) << (8 * (i & 3));
      pw[i >> 2] |= (uint32_t)(v.cap[i] & 0xFF) << (8 * (i & 3));
    }
    uint4* r = svc_rec(slot);
    r[0] = make_uint4((uint32_t)__double2loint(v.arrival), (uint32_t)__double2hiint(v.arrival),
                      (uint32_t)v.pair | ((uint32_t)v.path << 16) | ((uint32_t)v.nch << 24), (uint32_t)v.run_pos);
    r[1] = make_uint4(cw[0], cw[1], cw[2], cw[3]);
    r[2] = make_uint4(sw[0], sw[1], pw[0], pw[1]);
  }
  __device__ __forceinline__ unsigned short& run_at(int i) const { return p.phy_run[(size_t)e * p.phy_run_cap + i]; }
  // topology.graph["running_services"] as an append-only list with tombstones (kept for defragmentation only)
  __device__ void run_squeeze() {
    int j = 0;
    for (int i = 0; i < run_len; i++) {
      const unsigned short sl = run_at(i);
      if (sl != 0xFFFFu) {
        if (j != i) {
          run_at(j) = sl;
          svc_rec(sl)[0].w = (uint32_t)j;
        }
        j++;
      }
    }
    run_len = j;
    run_dead = 0;
  }
  __device__ int run_append(int slot) {  // returns the position
    if (run_len == p.phy_run_cap) run_squeeze();
    if (run_len == p.phy_run_cap) {
      cnt[CNT_FAULT] |= 64;
      return 0;
    }
    run_at(run_len) = (unsigned short)slot;
    return run_len++;
  }
  __device__ __forceinline__ void run_remove(int pos) {
    run_at(pos) = 0xFFFFu;
    run_dead++;
  }
  // a new Service enters the table, the release queue and (defragmentation) the order list; false = table full
  __device__ bool svc_provision(PhySvc& v, double key) {
    if (sv_free_n == 0 || heap_len >= p.heap_cap) {
      cnt[CNT_FAULT] |= 1;
      return false;
    }
    const int slot = p.phy_free[(size_t)e * p.phy_svc_cap + --sv_free_n];
    v.run_pos = p.phy_defrag_period > 0 ? run_append(slot) : 0;
    svc_store(slot, v);
    heap_push(key, (unsigned long long)slot);
    return true;
  }
  // _release_path :781-861: the channels of the service in the order of service.channels
  __device__ void phy_release(int slot) {
    const PhySvc v = svc_load(slot);
    PathRec rec = load_path(p, v.pair, v.path);
    // the record's fields are only ever indexed statically (the arrays stay in registers); the loop body is
    // selected per iteration instead
#pragma unroll 1
    for (int i = 0; i < v.nch; i++) {
      int share = v.share[0], cap = v.cap[0], ch = v.ch[0];
#pragma unroll
      for (int q = 1; q < 8; q++)
        if (q == i) {
          share = v.share[q];
          cap = v.cap[q];
          ch = v.ch[q];
        }
      bool free_it = !p.phy_grooming || share == cap;  // :784 channel[1] == channel[3]
      if (!free_it) {
        const uint32_t key = cs_key(v.pair, v.path, ch);
        const int at = cs_find(key);
        if (at < 0) {  // unreachable from step (the reference would raise)
          cnt[CNT_FAULT] |= 8;
          continue;
        }
        const uint32_t y = cs_at(at).y;
        const int used = (int)(y & 0xFFu), fr = (int)((y >> 8) & 0xFFu);
        cs_remove(at);
        if (used == share) free_it = true;                              // :821 nobody else is on the channel
        else cs_append(key, used - share, fr + share, (int)(y >> 16));  // :841-844
      }
      if (free_it) tile.apply_bit(rec, 0, ch, true);
    }
    if (p.phy_defrag_period > 0) {
      run_remove(v.run_pos);  // running_services.remove(service) :861
      if (run_dead >= 64 + ((run_len - run_dead) >> 1)) run_squeeze();
    }
    p.phy_free[(size_t)e * p.phy_svc_cap + sv_free_n++] = (unsigned short)slot;
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
  // removes the minimum and returns its payload.  The HEAP_D children of a node are read with
  // independent loads (slots past the end are masked by index), so one level = one memory round trip.
  __device__ __forceinline__ unsigned long long heap_pop() {
    ulonglong2 root = heap[0];
    ulonglong2 last = heap[--heap_len];
    const int n = heap_len;
    const double inf = __longlong_as_double(0x7FF0000000000000ll);
    if (n > 0) {
      const double lk = key_of(last);
      double rk = lk;  // key that ends up at the root
      int i = 0;
      for (;;) {
        const int c = HEAP_D * i + 1;
        if (c >= n) break;
        ulonglong2 v[HEAP_D];
#pragma unroll
        for (int d = 0; d < HEAP_D; d++) v[d] = heap[c + d];
        double bk = inf;
        int bd = 0;
#pragma unroll
        for (int d = 0; d < HEAP_D; d++) {
          double k = (c + d < n) ? key_of(v[d]) : inf;
          if (k < bk) {
            bk = k;
            bd = d;
          }
        }
        if (bk < lk) {
          unsigned long long pay = v[0].y;
#pragma unroll
          for (int d = 1; d < HEAP_D; d++) pay = (bd == d) ? v[d].y : pay;
          heap[i] = make_ulonglong2((unsigned long long)__double_as_longlong(bk), pay);
          if (i == 0) rk = bk;
          i = c + bd;
        } else {
          break;
        }
      }
      heap[i] = last;
      next_rel = rk;
    } else {
      next_rel = inf;
    }
    return root.y;
  }

  // ---- _next_service rmsa_env.py:643-695 / rmcsa_env.py:690-739 ----
  // The five draws of the next request depend only on the request index and the clock, not on
  // the decision being taken, so they can be evaluated ahead of it (draw_request) and applied
  // after it (commit_request); the reference's order of state changes is kept by commit_request.
  struct Req {
    double now, holding;
    int bit_rate, src, dst;
  };
  __device__ __forceinline__ Req draw_request() {
    Req q;
    const long long r = cnt[CNT_DRAWN];
    if (p.request_source == ORLG_SRC_REPLAY) {
      long long rr = r < p.trace_len ? r : p.trace_len - 1;
      if (r >= p.trace_len) cnt[CNT_FAULT] |= 2;  // trace exhausted
      const orlg_request* t = p.trace + (size_t)rr * p.num_envs + e;
      q.now = t->arrival;
      q.holding = t->holding;
      q.bit_rate = t->bit_rate;
      q.src = t->src;
      q.dst = t->dst;
    } else {
      unsigned long long g = (unsigned long long)(p.genv0 + e);
      uint2 key = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
      uint4 R0 = philox4x32_10(make_uint4((uint32_t)r, (uint32_t)((unsigned long long)r >> 32), (uint32_t)g, 0u), key);
      uint4 R1 = philox4x32_10(make_uint4((uint32_t)r, (uint32_t)((unsigned long long)r >> 32), (uint32_t)g, 1u), key);
      double lam_a = 1.0 / p.mean_iat, lam_h = 1.0 / p.mean_holding;
      q.now = now + (-det_log(1.0 - res53(R0.x, R0.y)) / lam_a);
      q.holding = -det_log(1.0 - res53(R0.z, R0.w)) / lam_h;
      const double inv32 = 1.0 / 4294967296.0;
      int n = p.n_nodes;
      q.src = bisect_right_guided(p.src_cum, p.guide, R1.x, ((double)R1.x * inv32) * __ldg(p.src_cum + n - 1), n - 1);
      const double* dc = p.dst_cum + (size_t)q.src * n;
      q.dst = bisect_right_guided(dc, p.guide + 64 * (2 + q.src), R1.y, ((double)R1.y * inv32) * __ldg(dc + n - 1), n - 1);
      double ub = (double)R1.z * inv32;
      if (p.n_bit_rates > 0) {
        int nb = p.n_bit_rates;
        q.bit_rate = __ldg(p.bit_rates + bisect_right_guided(p.br_cum, p.guide + 64, R1.z, ub * __ldg(p.br_cum + nb - 1), nb - 1));
      } else {
        q.bit_rate = p.br_lo + (int)floor(ub * (double)(p.br_hi - p.br_lo + 1));
      }
    }
    return q;
  }
  __device__ __forceinline__ void commit_request(const Req& q) {
    now = q.now;
    holding = q.holding;
    bit_rate = q.bit_rate;
    src = q.src;
    dst = q.dst;
    cnt[CNT_DRAWN]++;
    if (p.kind != ORLG_ENV_RMCSA) {  // rmsa_env.py:673-674
      cnt[CNT_PROC]++;
      cnt[CNT_EP_PROC]++;
    }
    cnt[CNT_BR_REQ] += bit_rate;  // rmsa_env.py:677-678, rmcsa_env.py:730-731
    cnt[CNT_EP_BR_REQ] += bit_rate;
    if constexpr (INFO_STATS) info_count(0, bit_rate);  // bit_rate_requested_histogram :680
    while (next_rel <= now) {  // rmsa_env.py:689-695
      unsigned long long s = heap_pop();
      if constexpr (PHY) {
        phy_release((int)s);  // the payload is the service's row
      } else {
        PathRec rec = load_path(p, svc_pair(s), svc_path(s));
        tile.set_range(rec, svc_core(s), svc_start(s), svc_slots(s));
        if (p.qot) {  // its neighbours' running sums are corrected by the next qot_kernel launch (qot.cuh)
          if (rel_n < QOT_REL_CAP_) p.qot_rel[(size_t)e * QOT_REL_CAP_ + rel_n++] = s;
          else cnt[CNT_FAULT] |= 16;
        }
        if constexpr (INFO_STATS) {  // _release_path :531-535 (current_time is already the new arrival)
          for_each_link(rec, [&](int l) { info_update_link(l); });
          slot_hops -= (long long)svc_slots(s) * rec.hops();
        }
      }
    }
  }
  __device__ __forceinline__ void next_service() { commit_request(draw_request()); }

  // Pull the heap lines the coming push / pops will touch towards L2 while the tile is in flight.
  __device__ __forceinline__ void prefetch_queue(double upcoming_now) const {
    if (heap_len > 0) {
      const ulonglong2* parent = heap + (heap_len - 1) / HEAP_D;  // parent of the insertion slot
      asm volatile("prefetch.global.L2 [%0];" ::"l"(parent));
      if (next_rel <= upcoming_now) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(heap));       // root (slot 7, line 0)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(heap + 1));   // its child group (line 1)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(heap + heap_len - 1));  // the entry that refills the root
      }
    }
  }

  // counters-only reset, rmsa_env.py:344-389 (the pending service is re-counted)
  __device__ __forceinline__ void reset_episode() {
    cnt[CNT_EP_ACC] = 0;
    cnt[CNT_EP_BR_PROV] = 0;
    cnt[CNT_EP_PROC] = 1;
    cnt[CNT_EP_BR_REQ] = bit_rate;
    if constexpr (PHY) {
      fsum[0] = fsum[1] = 0.0;
      isum[0] = isum[1] = isum[2] = isum[3] = 0;
      mv[0] = mv[1] = mv[2] = 0;  // phy_rmsa_env.py:433-435
    }
  }

  // _provision_path + _add_release (state and counters only)
  __device__ __forceinline__ void provision(const PathRec& rec, int path, int core, int start, int n, int se, int row = 0) {
    tile.clear_range(rec, core, start, n);
    if constexpr (INFO_STATS) {  // _provision_path :494, :510
      for_each_link(rec, [&](int l) { info_update_link(l); });
      slot_hops += (long long)n * rec.hops();
      info_count(1, bit_rate);
    }
    cnt[CNT_ACC]++;
    cnt[CNT_EP_ACC]++;
    cnt[CNT_BR_PROV] += bit_rate;
    cnt[CNT_EP_BR_PROV] += bit_rate;
    heap_push(now + holding, pack_service(pair(), path, core, start, n, se, row));
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
