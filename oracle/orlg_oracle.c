/* orlg_oracle.c — CPU restatement of the reference env stepping path.
 * TEST INFRASTRUCTURE ONLY (see orlg_oracle.h for the rules and parity status).
 *
 * Every function cites the reference lines it follows; paths are relative to
 * /root/reference/optical_rl_gym/envs/.
 */
#include "orlg_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ------------------------------------------------------------------ RNG -- */
/* Philox4x32-10 (Salmon et al., SC'11).  Used only by the "philox" request
 * source, which replaces CPython's MT19937 stream. */
void oracle_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                       uint32_t k1, uint32_t out[4]) {
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Natural log built from IEEE add/mul/div only (fdlibm e_log.c scheme, no FMA)
 * so that the CUDA engine and this oracle agree bit for bit in philox mode. */
double oracle_dlog(double x) {
  static const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10,
                      Lg1 = 6.666666666666735130e-01, Lg2 = 3.999999999940941908e-01,
                      Lg3 = 2.857142874366239149e-01, Lg4 = 2.222219843214978396e-01,
                      Lg5 = 1.818357216161805012e-01, Lg6 = 1.531383769920937332e-01,
                      Lg7 = 1.479819860511658591e-01;
  uint64_t bits;
  memcpy(&bits, &x, 8);
  int k = 0;
  if ((bits >> 52) == 0) { /* subnormal or zero: scale up (never hit for 1-u, u<1) */
    x *= 18014398509481984.0; /* 2^54 */
    memcpy(&bits, &x, 8);
    k = -54;
  }
  k += (int)(bits >> 52) - 1023;
  uint64_t mant = bits & 0x000FFFFFFFFFFFFFull;
  /* bring m into [sqrt(1/2), sqrt(2)) */
  if (mant >= 0x6A09E667F3BCDull) {
    k += 1;
    bits = mant | 0x3FE0000000000000ull;
  } else {
    bits = mant | 0x3FF0000000000000ull;
  }
  double m;
  memcpy(&m, &bits, 8);
  double f = m - 1.0;
  double s = f / (2.0 + f);
  double z = s * s;
  double w = z * z;
  double t1 = w * (Lg2 + w * (Lg4 + w * Lg6));
  double t2 = z * (Lg1 + w * (Lg3 + w * (Lg5 + w * Lg7)));
  double R = t2 + t1;
  double hfsq = 0.5 * f * f;
  double dk = (double)k;
  return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
}

static double res53(uint32_t a, uint32_t b) { /* MT genrand_res53 form */
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

/* bisect.bisect_right(cum, x, 0, hi) as used by random.choices */
static int bisect_right(const double* cum, double x, int hi) {
  int lo = 0;
  while (lo < hi) {
    int mid = (lo + hi) / 2;
    if (x < cum[mid]) hi = mid; else lo = mid + 1;
  }
  return lo;
}

/* The five draws of _next_service (rmsa_env.py:646-659; _get_node_pair
 * optical_network_env.py:191-208) on the Philox stream of one env. */
void oracle_philox_request(const oracle_config* c, int64_t genv, int64_t r, double now,
                           oracle_request* out) {
  uint32_t R0[4], R1[4];
  uint32_t k0 = (uint32_t)c->seed, k1 = (uint32_t)(c->seed >> 32);
  oracle_philox4x32((uint32_t)r, (uint32_t)((uint64_t)r >> 32), (uint32_t)genv, 0u, k0, k1, R0);
  oracle_philox4x32((uint32_t)r, (uint32_t)((uint64_t)r >> 32), (uint32_t)genv, 1u, k0, k1, R1);
  double u_at = res53(R0[0], R0[1]), u_ht = res53(R0[2], R0[3]);
  double lam_a = 1.0 / c->mean_iat, lam_h = 1.0 / c->mean_holding;
  out->arrival = now + (-oracle_dlog(1.0 - u_at) / lam_a);
  out->holding = -oracle_dlog(1.0 - u_ht) / lam_h;
  double us = (double)R1[0] * (1.0 / 4294967296.0);
  double ud = (double)R1[1] * (1.0 / 4294967296.0);
  double ub = (double)R1[2] * (1.0 / 4294967296.0);
  int n = c->n_nodes;
  int src = bisect_right(c->src_cum, us * c->src_cum[n - 1], n - 1);
  const double* dc = c->dst_cum + (size_t)src * n;
  int dst = bisect_right(dc, ud * dc[n - 1], n - 1);
  out->src = (int16_t)src;
  out->dst = (int16_t)dst;
  if (c->n_bit_rates > 0) {
    int nb = c->n_bit_rates;
    out->bit_rate = c->bit_rates[bisect_right(c->bit_rate_cum, ub * c->bit_rate_cum[nb - 1], nb - 1)];
  } else {
    out->bit_rate = c->bit_rate_lo + (int32_t)floor(ub * (double)(c->bit_rate_hi - c->bit_rate_lo + 1));
  }
}

/* ------------------------------------------------------------ env state -- */
typedef struct svc { /* a provisioned Service (utils.py:56-80), what release needs */
  double release_time;
  int32_t pair, path, core, initial_slot, number_slots;
  int32_t se; /* current_modulation.spectral_efficiency */
  int32_t nch;      /* PhyRMSA: service.channels */
  int16_t ch[8];
  int8_t share[8], cap[8]; /* channel tuple fields [1] (capacity used by this service) and [3] (level), units of 100 Gb/s */
  int64_t uid;    /* identity of the Service object (service_id repeats across episodes, F7) */
  double arrival; /* service.arrival_time (age of a defragmentation candidate, phy_rmsa_env.py:372) */
} svc;

/* one tuple of channel_state[src_id, dst_id, idp] (phy_rmsa_env.py:599-601): (channel, used, free, capacity);
 * the per-key Python lists are kept as one list in append order (remove + append moves an entry to the end) */
typedef struct chan_state { int32_t pair, idp, ch, used, free_, cap; } chan_state;

struct oracle_env {
  oracle_config c;
  int64_t genv;
  uint8_t* avail; /* [cores][E][S], 1 = free: topology.graph["available_slots"] */
  svc* heap;      /* heapq list of (release_time, service) */
  /* info dict state (info_stats): topology[n1][n2]["last_update"|"utilization"|"compactness"], histograms */
  double *lnk_last_update, *lnk_utilization, *lnk_compactness;
  int64_t *hist_requested, *hist_provisioned; /* bit_rate_requested_histogram / _provisioned_ by bit-rate index */
  double info[64];
  int info_n;
  chan_state* cs; /* channel_state, all keys, append order */
  int64_t cs_len, cs_cap;
  int64_t heap_len, heap_cap;
  double now;     /* current_time */
  oracle_request cur;
  int new_service; /* _new_service */
  int64_t services_processed, services_accepted, ep_processed, ep_accepted;
  int64_t br_requested, br_provisioned, ep_br_requested, ep_br_provisioned;
  int64_t requests_drawn, episodes;
  /* PhyRMSA episode sums (phy_rmsa_env.py:431-439) */
  double total_path_length_ep, total_gsnr_ep;
  int64_t total_path_index_ep, total_mod_level_ep, channels_accepted_ep, physical_accepted_ep;
  double info_cuts, info_rss; /* number_cuts_total / rss_total_metric of the last step's info (:313-314) */
  int64_t* run; /* topology.graph["running_services"] in list order (uids): PhyRMSA defragmentation walks it */
  int64_t run_len, run_cap, next_uid;
  int64_t counted_moves, counted_moves_groom, counted_defrag_cycles; /* phy_rmsa_env.py:110-112 */
  double qot_last[2];         /* QoT gate: GSNR of the last candidate evaluated, lowest neighbour GSNR with it present */
  int16_t* link_of; /* [N][N] edge index or -1 (topology[a][b]["index"]) */
  const oracle_request* trace;
  int64_t trace_len, trace_stride;
};

#define AV(e, core, link) ((e)->avail + ((size_t)(core) * (e)->c.n_edges + (link)) * (e)->c.num_slots)

oracle_env* oracle_create(const oracle_config* cfg, int64_t genv) {
  oracle_env* e = (oracle_env*)calloc(1, sizeof(*e));
  e->c = *cfg;
  if (e->c.num_cores < 1) e->c.num_cores = 1;
  if (e->c.j < 1) e->c.j = 1;
  e->genv = genv;
  e->avail = (uint8_t*)malloc((size_t)e->c.num_cores * e->c.n_edges * e->c.num_slots);
  e->heap_cap = 64;
  e->heap = (svc*)malloc(sizeof(svc) * e->heap_cap);
  if (e->c.info_stats) {
    e->lnk_last_update = (double*)calloc((size_t)e->c.n_edges, sizeof(double));
    e->lnk_utilization = (double*)calloc((size_t)e->c.n_edges, sizeof(double));
    e->lnk_compactness = (double*)calloc((size_t)e->c.n_edges, sizeof(double));
    e->hist_requested = (int64_t*)calloc((size_t)(e->c.n_bit_rates > 0 ? e->c.n_bit_rates : 1), sizeof(int64_t));
    e->hist_provisioned = (int64_t*)calloc((size_t)(e->c.n_bit_rates > 0 ? e->c.n_bit_rates : 1), sizeof(int64_t));
  }
  if (e->c.env_kind == 3 && e->c.edge_nodes) {
    int N = e->c.n_nodes;
    e->link_of = (int16_t*)malloc(sizeof(int16_t) * N * N);
    for (int i = 0; i < N * N; i++) e->link_of[i] = -1;
    for (int l = 0; l < e->c.n_edges; l++) {
      int a = e->c.edge_nodes[2 * l], b = e->c.edge_nodes[2 * l + 1];
      e->link_of[a * N + b] = e->link_of[b * N + a] = (int16_t)l;
    }
  }
  return e;
}

void oracle_destroy(oracle_env* e) {
  if (!e) return;
  free(e->avail);
  free(e->heap);
  free(e->cs);
  free(e->lnk_last_update); free(e->lnk_utilization); free(e->lnk_compactness);
  free(e->hist_requested); free(e->hist_provisioned);
  free(e->link_of);
  free(e->run);
  free(e);
}

void oracle_set_trace(oracle_env* e, const oracle_request* t, int64_t n, int64_t stride) {
  e->trace = t;
  e->trace_len = n;
  e->trace_stride = stride;
}

int oracle_action_dim(const oracle_env* e) {
  return e->c.env_kind == 0 ? 2 : e->c.env_kind == 1 ? 1 : e->c.env_kind == 2 ? 4 : ORACLE_PHY_ADIM;
}
int oracle_obs_dim(const oracle_env* e) { /* deeprmsa_env.py:35-37 */
  return 1 + 2 * e->c.n_nodes + (2 * e->c.j + 3) * e->c.k_paths;
}

/* --------------------------------------------------------------- heapq -- */
/* heapq.heappush / heappop on (release_time, service): _add_release
 * optical_network_env.py:178-189, consumer rmsa_env.py:689-695. */
static void heap_push(oracle_env* e, svc item) {
  if (e->heap_len == e->heap_cap) {
    e->heap_cap *= 2;
    e->heap = (svc*)realloc(e->heap, sizeof(svc) * e->heap_cap);
  }
  int64_t pos = e->heap_len++;
  while (pos > 0) { /* _siftdown */
    int64_t parent = (pos - 1) >> 1;
    if (item.release_time < e->heap[parent].release_time) {
      e->heap[pos] = e->heap[parent];
      pos = parent;
    } else break;
  }
  e->heap[pos] = item;
}

static svc heap_pop(oracle_env* e) {
  svc last = e->heap[--e->heap_len];
  if (e->heap_len == 0) return last;
  svc ret = e->heap[0];
  /* _siftup: bubble the smaller child up to a leaf, then sift `last` down */
  int64_t pos = 0, end = e->heap_len, child = 1;
  while (child < end) {
    int64_t right = child + 1;
    if (right < end && !(e->heap[child].release_time < e->heap[right].release_time)) child = right;
    e->heap[pos] = e->heap[child];
    pos = child;
    child = 2 * pos + 1;
  }
  while (pos > 0) {
    int64_t parent = (pos - 1) >> 1;
    if (last.release_time < e->heap[parent].release_time) {
      e->heap[pos] = e->heap[parent];
      pos = parent;
    } else break;
  }
  e->heap[pos] = last;
  return ret;
}

/* ------------------------------------------------------------ core ops -- */
static const int16_t* links_of(const oracle_env* e, int pair, int path, int by_id) {
  const int16_t* t = (by_id && e->c.path_links_by_id) ? e->c.path_links_by_id : e->c.path_links;
  return t + ((size_t)pair * e->c.k_paths + path) * e->c.max_hops;
}
static int hops_of(const oracle_env* e, int pair, int path) {
  return e->c.path_hops[(size_t)pair * e->c.k_paths + path];
}
static int cur_pair(const oracle_env* e) { return e->cur.src * e->c.n_nodes + e->cur.dst; }

/* get_number_slots rmsa_env.py:708-719 / rmcsa_env.py:753-765 */
static int number_slots(const oracle_env* e, int se) {
  return (int)ceil((double)e->cur.bit_rate / ((double)se * e->c.channel_width)) + 1;
}

/* is_path_free rmsa_env.py:721-734 / rmcsa_env.py:767-794 */
static int is_path_free(const oracle_env* e, int pair, int path, int core, int start, int n) {
  if (start + n > e->c.num_slots) return 0;
  const int16_t* L = links_of(e, pair, path, 0);
  int h = hops_of(e, pair, path);
  for (int i = 0; i < h; i++) {
    const uint8_t* row = AV(e, core, L[i]);
    for (int s = start; s < start + n; s++)
      if (row[s] == 0) return 0;
  }
  return 1;
}

/* _provision_path rmsa_env.py:462-513 / rmcsa_env.py:488-534 (state + counters;
 * the info-only link/network statistics are out of scope) */
static int rle(const uint8_t* a, int n, int* pos, uint8_t* val, int* len);

/* _get_network_compactness rmsa_env.py:806-851 */
static double network_compactness(const oracle_env* e) {
  int E = e->c.n_edges, S = e->c.num_slots;
  int64_t sum_slots_paths = 0, sum_occupied = 0, sum_unused_spectrum_blocks = 0;
  for (int64_t i = 0; i < e->heap_len; i++) /* running services = the release queue */
    sum_slots_paths += (int64_t)e->heap[i].number_slots * hops_of(e, e->heap[i].pair, e->heap[i].path);
  int pos[1024], len[1024], ipos[1024], ilen[1024];
  uint8_t val[1024], ival[1024];
  for (int l = 0; l < E; l++) {
    const uint8_t* row = AV(e, 0, l);
    int nr = rle(row, S, pos, val, len), first = -1, last = -1, used = 0;
    for (int r = 0; r < nr; r++)
      if (val[r] == 0) { if (first < 0) first = r; last = r; used++; }
    if (used > 1) {
      int lambda_min = pos[first], lambda_max = pos[last] + len[last];
      sum_occupied += lambda_max - lambda_min;
      int ni = rle(row + lambda_min, lambda_max - lambda_min, ipos, ival, ilen);
      for (int r = 0; r < ni; r++) sum_unused_spectrum_blocks += ival[r]; /* np.sum(internal_values) */
    }
  }
  if (sum_unused_spectrum_blocks > 0)
    return ((double)sum_occupied / (double)sum_slots_paths) * ((double)E / (double)sum_unused_spectrum_blocks);
  return 1.0;
}

/* _update_link_stats rmsa_env.py:562-641 (utilization and compactness; external_fragmentation is not in the info dict) */
static void update_link_stats(oracle_env* e, int l) {
  if (!e->c.info_stats) return;
  int S = e->c.num_slots;
  double last_update = e->lnk_last_update[l];
  double time_diff = e->now - last_update;
  if (e->now > 0) {
    const uint8_t* row = AV(e, 0, l);
    int64_t free_slots = 0;
    for (int s = 0; s < S; s++) free_slots += row[s];
    double last_util = e->lnk_utilization[l];
    double cur_util = (double)(S - free_slots) / (double)S;
    e->lnk_utilization[l] = ((last_util * last_update) + (cur_util * time_diff)) / e->now;
    double last_compactness = e->lnk_compactness[l], cur_link_compactness = 0.0;
    if (free_slots > 0) {
      int pos[1024], len[1024], ipos[1024], ilen[1024];
      uint8_t val[1024], ival[1024];
      int nr = rle(row, S, pos, val, len), first = -1, last = -1, used = 0;
      for (int r = 0; r < nr; r++)
        if (val[r] == 0) { if (first < 0) first = r; last = r; used++; }
      if (used > 1) {
        int lambda_min = pos[first], lambda_max = pos[last] + len[last];
        int ni = rle(row + lambda_min, lambda_max - lambda_min, ipos, ival, ilen);
        int64_t unused_spectrum_slots = 0;
        for (int r = 0; r < ni; r++) unused_spectrum_slots += 1 - ival[r]; /* np.sum(1 - internal_values) */
        if (unused_spectrum_slots > 0)
          cur_link_compactness = ((double)(lambda_max - lambda_min) / (double)(S - free_slots)) *
                                 (1.0 / (double)unused_spectrum_slots);
        else cur_link_compactness = 1.0;
      } else cur_link_compactness = 1.0;
    }
    e->lnk_compactness[l] = ((last_compactness * last_update) + (cur_link_compactness * time_diff)) / e->now;
  }
  e->lnk_last_update[l] = e->now;
}

/* numpy's pairwise summation of a contiguous f64 array (what np.mean of a list runs), n < 128 */
static double np_sum(const double* a, int n) {
  if (n < 8) {
    double res = 0.;
    for (int i = 0; i < n; i++) res += a[i];
    return res;
  }
  double r[8];
  int i;
  for (i = 0; i < 8; i++) r[i] = a[i];
  for (i = 8; i < n - (n % 8); i += 8)
    for (int j = 0; j < 8; j++) r[j] += a[i + j];
  double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  for (; i < n; i++) res += a[i];
  return res;
}

static int bit_rate_index(const oracle_env* e, int bit_rate) {
  for (int i = 0; i < e->c.n_bit_rates; i++)
    if (e->c.bit_rates[i] == bit_rate) return i;
  return -1;
}

/* the info dict of RMSAEnv.step rmsa_env.py:276-332, taken after the provisioning and before _next_service */
static void make_info(oracle_env* e, double previous_network_compactness) {
  double* o = e->info;
  int E = e->c.n_edges, nb = e->c.n_bit_rates;
  double cur = network_compactness(e);
  o[0] = (double)(e->services_processed - e->services_accepted) / (double)e->services_processed;
  o[1] = (double)(e->ep_processed - e->ep_accepted) / (double)e->ep_processed;
  o[2] = (double)(e->br_requested - e->br_provisioned) / (double)e->br_requested;
  o[3] = (double)(e->ep_br_requested - e->ep_br_provisioned) / (double)e->ep_br_requested;
  o[4] = cur;
  o[5] = previous_network_compactness - cur;
  o[6] = np_sum(e->lnk_compactness, E) / (double)E;
  o[7] = np_sum(e->lnk_utilization, E) / (double)E;
  double mx = 0, mn = 0;
  for (int i = 0; i < nb && i < 50; i++) { /* :277-289 */
    double b = e->hist_requested[i] > 0
                   ? (double)(e->hist_requested[i] - e->hist_provisioned[i]) / (double)e->hist_requested[i] : 0.0;
    o[9 + i] = b;
    if (i == 0 || b > mx) mx = b;
    if (i == 0 || b < mn) mn = b;
  }
  o[8] = mx - mn; /* fairness :326-328 */
  e->info_n = 9 + (nb < 50 ? nb : 50);
}

int oracle_info(const oracle_env* e, double* out, int max_values) {
  for (int i = 0; i < e->info_n && i < max_values; i++) out[i] = e->info[i];
  return e->info_n;
}

static void provision_path(oracle_env* e, int pair, int path, int core, int start, int n) {
  const int16_t* L = links_of(e, pair, path, 0);
  int h = hops_of(e, pair, path);
  for (int i = 0; i < h; i++) {
    memset(AV(e, core, L[i]) + start, 0, (size_t)n);
    if (e->c.env_kind <= 1) update_link_stats(e, L[i]); /* rmsa_env.py:494 */
  }
  if (e->c.info_stats && e->c.env_kind <= 1) {
    int bi = bit_rate_index(e, e->cur.bit_rate);
    if (bi >= 0) e->hist_provisioned[bi]++; /* :510 */
  }
  e->services_accepted++;
  e->ep_accepted++;
  e->br_provisioned += e->cur.bit_rate;
  e->ep_br_provisioned += e->cur.bit_rate;
}

static int64_t cs_find(const oracle_env* e, int pair, int idp, int ch) {
  for (int64_t i = 0; i < e->cs_len; i++)
    if (e->cs[i].pair == pair && e->cs[i].idp == idp && e->cs[i].ch == ch) return i;
  return -1;
}
static void cs_remove(oracle_env* e, int64_t at) {
  memmove(e->cs + at, e->cs + at + 1, sizeof(chan_state) * (size_t)(e->cs_len - at - 1));
  e->cs_len--;
}
static void cs_append(oracle_env* e, int pair, int idp, int ch, int used, int free_, int cap) {
  if (e->cs_len == e->cs_cap) {
    e->cs_cap = e->cs_cap ? 2 * e->cs_cap : 64;
    e->cs = (chan_state*)realloc(e->cs, sizeof(chan_state) * (size_t)e->cs_cap);
  }
  chan_state r = {pair, idp, ch, used, free_, cap};
  e->cs[e->cs_len++] = r;
}

/* topology.graph["running_services"]: list.append / list.remove */
static void run_append(oracle_env* e, int64_t uid) {
  if (e->run_len == e->run_cap) {
    e->run_cap = e->run_cap ? 2 * e->run_cap : 256;
    e->run = (int64_t*)realloc(e->run, sizeof(int64_t) * (size_t)e->run_cap);
  }
  e->run[e->run_len++] = uid;
}
static void run_remove(oracle_env* e, int64_t uid) {
  for (int64_t i = 0; i < e->run_len; i++)
    if (e->run[i] == uid) {
      memmove(e->run + i, e->run + i + 1, sizeof(int64_t) * (size_t)(e->run_len - i - 1));
      e->run_len--;
      return;
    }
}

/* _release_path rmsa_env.py:515-535 / rmcsa_env.py:536-558 */
static void release_path(oracle_env* e, const svc* s) {
  if (e->c.env_kind == 3) run_remove(e, s->uid); /* phy_rmsa_env.py:861 (order of this line does not matter) */
  const int16_t* L = links_of(e, s->pair, s->path, 0);
  int h = hops_of(e, s->pair, s->path);
  if (e->c.env_kind == 3 && !e->c.phy_grooming) { /* no channel_state: every channel of the service is freed */
    for (int i = 0; i < h; i++)
      for (int c = 0; c < s->nch; c++) AV(e, 0, L[i])[s->ch[c]] = 1;
    return;
  }
  if (e->c.env_kind == 3) { /* phy_rmsa_env.py:781-861 */
    for (int c = 0; c < s->nch; c++) {
      int free_it = s->share[c] == s->cap[c]; /* :784 channel[1] == channel[3] */
      if (!free_it) {                        /* :801-846 */
        int64_t at = cs_find(e, s->pair, s->path, s->ch[c]);
        if (at < 0) continue; /* the reference would raise IndexError; unreachable from step */
        chan_state r = e->cs[at];
        cs_remove(e, at);
        if (r.used == s->share[c]) free_it = 1; /* :821 nobody else is on the channel */
        else cs_append(e, r.pair, r.idp, r.ch, r.used - s->share[c], r.free_ + s->share[c], r.cap); /* :841-844 */
      }
      if (free_it)
        for (int i = 0; i < h; i++) AV(e, 0, L[i])[s->ch[c]] = 1;
    }
    return;
  }
  for (int i = 0; i < h; i++) {
    memset(AV(e, s->core, L[i]) + s->initial_slot, 1, (size_t)s->number_slots);
    if (e->c.env_kind <= 1) update_link_stats(e, L[i]); /* rmsa_env.py:531-533 */
  }
}

/* get_available_slots rmsa_env.py:745-756 (rows by edge "id", F9); core 0 plane
 * as in rmcsa_env.py:796-807 is not used by any in-scope caller. */
static void available_slots(const oracle_env* e, int pair, int path, uint8_t* out) {
  const int16_t* L = links_of(e, pair, path, 1);
  int h = hops_of(e, pair, path), S = e->c.num_slots;
  memset(out, 1, (size_t)S);
  for (int i = 0; i < h; i++) {
    const uint8_t* row = AV(e, 0, L[i]);
    for (int s = 0; s < S; s++) out[s] = (uint8_t)(out[s] * row[s]);
  }
}

/* rle rmsa_env.py:758-772: returns number of runs */
static int rle(const uint8_t* a, int n, int* pos, uint8_t* val, int* len) {
  int r = 0, start = 0;
  for (int i = 1; i <= n; i++) {
    if (i == n || a[i] != a[i - 1]) {
      pos[r] = start; val[r] = a[start]; len[r] = i - start;
      r++; start = i;
    }
  }
  return r;
}

/* get_available_blocks rmsa_env.py:774-804: first j runs of ones with length>=slots */
static int available_blocks(const oracle_env* e, int pair, int path, int slots, int* starts, int* lens) {
  int S = e->c.num_slots;
  uint8_t av[1024], val[1024];
  int pos[1024], len[1024];
  available_slots(e, pair, path, av);
  int r = rle(av, S, pos, val, len), nb = 0;
  for (int i = 0; i < r && nb < e->c.j; i++)
    if (val[i] == 1 && len[i] >= slots) { starts[nb] = pos[i]; lens[nb] = len[i]; nb++; }
  return nb;
}

/* _next_service rmsa_env.py:643-695 ; rmcsa_env.py:690-739 */
static void next_service(oracle_env* e) {
  if (e->new_service) return;
  oracle_request rq;
  if (e->c.request_source == 0) {
    int64_t r = e->requests_drawn < e->trace_len ? e->requests_drawn : e->trace_len - 1;
    rq = e->trace[r * e->trace_stride];
  } else {
    oracle_philox_request(&e->c, e->genv, e->requests_drawn, e->now, &rq);
  }
  e->requests_drawn++;
  e->now = rq.arrival; /* at = current_time + expovariate(...); current_time = at */
  e->cur = rq;
  e->new_service = 1;
  if (e->c.env_kind != 2) { /* rmsa_env.py:673-678 */
    e->services_processed++;
    e->ep_processed++;
  }
  e->br_requested += rq.bit_rate; /* rmsa_env.py:677-678 ; rmcsa_env.py:730-731 */
  e->ep_br_requested += rq.bit_rate;
  if (e->c.info_stats && e->c.env_kind <= 1) {
    int bi = bit_rate_index(e, rq.bit_rate);
    if (bi >= 0) e->hist_requested[bi]++; /* :680 */
  }
  while (e->heap_len > 0) { /* rmsa_env.py:689-695 */
    if (e->heap[0].release_time <= e->now) {
      svc s = heap_pop(e);
      release_path(e, &s);
    } else break; /* pop + push back of the same item leaves the heap's content unchanged */
  }
}

/* reset rmsa_env.py:343-457 / rmcsa_env.py:386-483 */
void oracle_reset(oracle_env* e, int full) {
  e->ep_br_requested = e->ep_br_provisioned = 0;
  e->ep_processed = e->ep_accepted = 0;
  e->total_path_length_ep = e->total_gsnr_ep = 0.0; /* phy_rmsa_env.py:431-439 */
  e->total_path_index_ep = e->total_mod_level_ep = e->channels_accepted_ep = e->physical_accepted_ep = 0;
  e->counted_moves = e->counted_moves_groom = e->counted_defrag_cycles = 0; /* phy_rmsa_env.py:433-435 */
  if (!full) {
    if (e->new_service) { /* pending service is re-counted, rmsa_env.py:370-374 */
      e->ep_processed++;
      e->ep_br_requested += e->cur.bit_rate;
    }
    return;
  }
  e->cs_len = 0;   /* channel_state phy_rmsa_env.py:455-461 */
  e->run_len = 0;
  if (e->c.info_stats) { /* optical_network_env.py:257-260, rmsa_env.py:444-452 */
    for (int l = 0; l < e->c.n_edges; l++) e->lnk_last_update[l] = e->lnk_utilization[l] = e->lnk_compactness[l] = 0.0;
    for (int i = 0; i < e->c.n_bit_rates; i++) e->hist_requested[i] = e->hist_provisioned[i] = 0;
  }
  e->heap_len = 0; /* OpticalNetworkEnv.reset optical_network_env.py:216-222 */
  e->now = 0.0;
  e->services_processed = e->services_accepted = 0;
  e->br_requested = e->br_provisioned = 0;
  e->requests_drawn = 0;
  e->episodes = 0;
  memset(e->avail, 1, (size_t)e->c.num_cores * e->c.n_edges * e->c.num_slots);
  e->new_service = 0;
  next_service(e);
}

/* ------------------------------------------------------------- stepping -- */
/* GN-model QoT gate (defined with the GN model at the end of this file) */
static int qot_gate_eval(oracle_env* e, int pair, int path, int core, int start, int n, int se);
static int qot_alone_ok(const oracle_env* e, int pair, int path, int start, int n, int se);
static int mods_by_se_desc(const oracle_env* e, int* order);
/* _crosstalk_is_acceptable rmcsa_env.py:341-384.  The modulation-only factors
 * (10**x terms) are evaluated by the caller in Python with the reference's own
 * expression order and arrive as mod_snr_coeff / mod_reach. */
static int crosstalk_ok(const oracle_env* e, int mod, double path_length) {
  double denom = e->c.mod_snr_coeff[mod] * ((double)e->cur.bit_rate / (double)e->c.mod_se[mod]);
  denom = denom * 1e9;
  double lmax_snr = 100.0 / denom;
  lmax_snr = lmax_snr / 1000.0;
  double lmax_xt = e->c.mod_reach[mod];
  return path_length < lmax_xt && path_length < lmax_snr;
}

static void phy_defragmentation(oracle_env* e);
static void finish_step(oracle_env* e, int accepted, int32_t* reward, uint8_t* done) {
  if (e->c.env_kind == 2) { /* rmcsa_env.py:292-295 (bit rate counted a second time) */
    e->services_processed++;
    e->ep_processed++;
    e->br_requested += e->cur.bit_rate;
    e->ep_br_requested += e->cur.bit_rate;
  }
  /* reward: optical_network_env.py:213-214 ; deeprmsa_env.py:123-124 */
  int r = e->c.env_kind == 1 ? (accepted ? 1 : -1) : (accepted ? 1 : 0);
  e->new_service = 0;
  next_service(e);
  if (e->c.env_kind == 3 && e->c.phy_defrag_period > 0 && e->services_processed % e->c.phy_defrag_period == 0)
    phy_defragmentation(e); /* phy_rmsa_env.py:355-417 */
  int d = e->ep_processed == e->c.episode_length; /* rmsa_env.py:339 */
  if (d) {
    e->episodes++;
    if (e->c.auto_reset) oracle_reset(e, 0);
  }
  if (reward) *reward = r;
  if (done) *done = (uint8_t)d;
}

/* ------------------------------------------------------- PhyRMSA-v0 -- */
/* phy_rmsa_env.py, physical layer only: grooming=False, defrag_period=None
 * (SURVEY.md 8f N2 is a later row).  Channels take the place of slots:
 * available_channels[E][C] lives in e->avail with one core. */
static int phy_level(const oracle_env* e, int table, int ch, int idp) {
  return e->c.phy_mod_level[((size_t)table * e->c.num_slots + ch) * e->c.k_paths + idp];
}
static const int16_t* nodes_of(const oracle_env* e, int pair, int path) {
  return e->c.path_nodes + ((size_t)pair * e->c.k_paths + path) * (e->c.max_hops + 1);
}

/* is_channel_free phy_rmsa_env.py:1029-1035 */
static int is_channel_free(const oracle_env* e, int pair, int path, int ch) {
  const int16_t* L = links_of(e, pair, path, 0);
  for (int i = 0; i < hops_of(e, pair, path); i++)
    if (AV(e, 0, L[i])[ch] == 0) return 0;
  return 1;
}

/* calculate_r_cut(channel, links, defrag_flag=False, path, modified=True) phy_rmsa_env.py:1140-1193 */
static int calculate_r_cut_flag(const oracle_env* e, int pair, int path, int ch, int defrag_flag) {
  int N = e->c.n_nodes, E = e->c.n_edges, h = hops_of(e, pair, path);
  const int16_t* nl = nodes_of(e, pair, path);
  const int16_t* L = links_of(e, pair, path, 0);
  int nn = h + 1;
  uint8_t tmp[256];
  int result[2];
  for (int pass = 0; pass < 2; pass++) {
    for (int l = 0; l < E; l++) tmp[l] = AV(e, 0, l)[ch];
    if (pass == 1) for (int i = 0; i < h; i++) tmp[L[i]] = defrag_flag ? 1 : 0; /* temporary_channels2[i] = 1 / 0 (:1167-1172) */
    int cut = 0;
    for (int i = 0; i < nn; i++) {
      for (int key = 0; key < N; key++) { /* for node_key in self.topology[path.node_list[i]] */
        int lk = e->link_of[nl[i] * N + key];
        if (lk < 0) continue;
        int on_path = 0;
        for (int q = 0; q < nn; q++) on_path |= (nl[q] == key);
        if (on_path) continue;
        if (i == nn - 1) cut += abs(tmp[e->link_of[nl[i - 1] * N + nl[i]]] - tmp[lk]);
        else if (i == 0) cut += abs(tmp[e->link_of[nl[i] * N + nl[i + 1]]] - tmp[lk]);
        else cut += abs(tmp[e->link_of[nl[i] * N + nl[i + 1]]] - tmp[lk]) +
                    abs(tmp[e->link_of[nl[i - 1] * N + nl[i]]] - tmp[lk]);
      }
    }
    result[pass] = cut;
  }
  return result[0] - result[1];
}

static int calculate_r_cut(const oracle_env* e, int pair, int path, int ch) { return calculate_r_cut_flag(e, pair, path, ch, 0); }

/* sqrt(sum len^2) / (sum len + 1) over the runs of free links of a channel column, link-index order
 * (the summand of calculate_r_spatial / calculate_total_r_spatial phy_rmsa_env.py:1085-1121) */
static double column_rss(const uint8_t* col, int E) {
  long long sum_sq = 0, sum = 0, run = 0;
  for (int l = 0; l < E; l++) {
    if (col[l]) run++;
    else { sum_sq += run * run; sum += run; run = 0; }
  }
  sum_sq += run * run; sum += run;
  return sqrt((double)sum_sq) / (double)(sum + 1);
}

/* calculate_r_spatial(channel, links, defrag_flag=False) phy_rmsa_env.py:1085-1108 */
static double calculate_r_spatial_flag(const oracle_env* e, int pair, int path, int ch, int defrag_flag) {
  int E = e->c.n_edges, h = hops_of(e, pair, path);
  const int16_t* L = links_of(e, pair, path, 0);
  uint8_t tmp[256];
  for (int l = 0; l < E; l++) tmp[l] = AV(e, 0, l)[ch];
  double before = column_rss(tmp, E);
  for (int i = 0; i < h; i++) tmp[L[i]] = defrag_flag ? 1 : 0; /* :1095-1100 */
  double after = column_rss(tmp, E);
  return after - before;
}
static double calculate_r_spatial(const oracle_env* e, int pair, int path, int ch) {
  return calculate_r_spatial_flag(e, pair, path, ch, 0);
}

static int use_existing_channels(const oracle_env* e, int32_t* a);

typedef struct phy_cand { int mod; double frag; int ch, idp; } phy_cand;

/* sorted(row, key=...) is stable: insertion sort by the heuristic's key */
static int cand_before(const phy_cand* a, const phy_cand* b, int hid) {
  if (hid == 5) return a->ch < b->ch;                                  /* sapff_rmsa :1704 key x[1] */
  if (hid == 6 || hid == 7) return a->mod != b->mod ? a->mod > b->mod : a->ch < b->ch; /* (-mod, ch) :1282,1343 */
  if (hid == 10 || hid == 12) return a->frag > b->frag;               /* phy_aware_faff(_rss)_rmsa key (-frag) :1537,1601 */
  return a->mod != b->mod ? a->mod > b->mod : a->frag > b->frag;      /* (-mod, -frag) :1405 */
}

/* phy_aware_bmfa_rmsa :1375-1438 (grooming False), phy_aware_bmff_rmsa :1317-1372,
 * phy_aware_sapbm_rmsa :1254-1314, sapff_rmsa :1676-1737 with use_existing_channels -> (-3, []) */
static void phy_heuristic(oracle_env* e, int hid, int32_t* a) {
  int k = e->c.k_paths, C = e->c.num_slots, pair = cur_pair(e);
  for (int i = 0; i < ORACLE_PHY_ADIM; i++) a[i] = -1;
  a[0] = -2; /* (-2, []) : blocked */
  /* the virtual layer first: unconditionally in sapff/sapbm/bmff/faff/faff_rss, under env.grooming in bmfa/bmfa_rss */
  if (e->c.phy_grooming == 1 || (e->c.phy_grooming == 2 && hid != 8 && hid != 11)) {
    if (use_existing_channels(e, a)) return;
    for (int i = 1; i < ORACLE_PHY_ADIM; i++) a[i] = -1;
  }
  int table = e->c.phy_table_id[pair];
  if (table < 0) return;
  phy_cand* rows = (phy_cand*)malloc(sizeof(phy_cand) * (size_t)k * C);
  int* len = (int*)calloc(k, sizeof(int));
  int* alive = (int*)calloc(k, sizeof(int));
  for (int idp = 0; idp < k; idp++) {
    phy_cand* row = rows + (size_t)idp * C;
    if (hops_of(e, pair, idp) <= 0) continue;
    for (int ch = 0; ch < C; ch++) {
      if (!is_channel_free(e, pair, idp, ch)) continue;
      double frag = 0.0;
      if (hid == 8 || hid == 10) frag = (double)calculate_r_cut(e, pair, idp, ch);
      else if (hid == 11 || hid == 12) frag = calculate_r_spatial(e, pair, idp, ch); /* _bmfa_rss_ :1467, _faff_rss_ :1597 */
      phy_cand c = {phy_level(e, table, ch, idp), frag, ch, idp};
      int pos = len[idp]++;
      while (pos > 0 && cand_before(&c, &row[pos - 1], hid)) { row[pos] = row[pos - 1]; pos--; }
      row[pos] = c;
    }
    alive[idp] = 1;
  }
  for (;;) {
    int best = -1;
    for (int i = 0; i < k; i++) {
      if (!alive[i] || len[i] == 0) continue;
      const phy_cand* hd = rows + (size_t)i * C;
      if (hid == 5 || hid == 6) { best = i; break; }          /* first non-empty row :1289-1297 */
      if (best < 0) { best = i; continue; }
      const phy_cand* bh = rows + (size_t)best * C;
      if (hid == 7) { if (hd->mod > bh->mod) best = i; }      /* :1350-1355 */
      else if (hid == 10 || hid == 12) { if (hd->frag > bh->frag) best = i; } /* :1546-1550, :1610-1614 */
      else if (hd->mod > bh->mod || (hd->mod == bh->mod && hd->frag > bh->frag)) best = i; /* :1414-1421 */
    }
    if (best < 0) break;
    int unassigned = e->cur.bit_rate, n = 0;
    const phy_cand* row = rows + (size_t)best * C;
    for (int i = 0; i < len[best]; i++) {
      unassigned -= row[i].mod * 100;
      if (n < ORACLE_PHY_ADIM - 1) a[1 + n] = row[i].ch;
      n++;
      if (unassigned <= 0) {
        a[0] = n <= ORACLE_PHY_ADIM - 1 ? row[i].idp : -2;
        if (a[0] == -2) for (int q = 1; q < ORACLE_PHY_ADIM; q++) a[q] = -1;
        free(rows); free(len); free(alive);
        return;
      }
    }
    for (int q = 1; q < ORACLE_PHY_ADIM; q++) a[q] = -1;
    alive[best] = 0; /* sorted_free_channels.pop(row) */
  }
  free(rows); free(len); free(alive);
}

/* use_existing_channels phy_rmsa_env.py:1650-1673: first path whose partially used channels can carry the
 * request; channels in list order, each gives all its free capacity, the last one what is still missing */
static int use_existing_channels(const oracle_env* e, int32_t* a) {
  int k = e->c.k_paths, pair = cur_pair(e), need = e->cur.bit_rate;
  for (int idp = 0; idp < k; idp++) {
    int sum = 0;
    for (int64_t i = 0; i < e->cs_len; i++)
      if (e->cs[i].pair == pair && e->cs[i].idp == idp) sum += e->cs[i].free_;
    if (sum * 100 < need) continue; /* sum(free) >= unassigned_bitrate / 100 */
    int n = 0;
    for (int64_t i = 0; i < e->cs_len; i++) {
      if (e->cs[i].pair != pair || e->cs[i].idp != idp || e->cs[i].free_ <= 0) continue;
      need -= e->cs[i].free_ * 100;
      if (n < ORACLE_PHY_ADIM - 1) a[1 + n] = e->cs[i].ch;
      n++;
      if (need <= 0) {
        if (n > ORACLE_PHY_ADIM - 1) return 0; /* documented cap of 8 channels per action */
        a[0] = idp + 20;
        return 1;
      }
    }
  }
  return 0;
}

/* PhyRMSAEnv.step phy_rmsa_env.py:272-309.  The action carries channel numbers; the (used, free) fields of
 * the reference's channel tuples are what its heuristics put there: every channel is used up to its capacity
 * (physical) or its free part (virtual) until the bit rate is covered, the last one takes the remainder. */
static int phy_decide(oracle_env* e, const int32_t* a, int32_t* alloc) {
  int k = e->c.k_paths, C = e->c.num_slots, pair = cur_pair(e);
  int path = a[0], nch = 0, accepted = 0;
  const int groom = e->c.phy_grooming != 0;
  while (nch < ORACLE_PHY_ADIM - 1 && a[1 + nch] >= 0) nch++;
  if (groom && path >= 20 && path < 20 + k && nch > 0) { /* virtual layer :280-288, _provision_virtual_path :625-659 */
    int idp = path - 20, need = e->cur.bit_rate / 100, ok = 1, share[8];
    for (int i = 0; ok && i < nch; i++) {
      for (int q = 0; ok && q < i; q++) ok = a[1 + q] != a[1 + i];
      int64_t at = ok ? cs_find(e, pair, idp, a[1 + i]) : -1;
      ok = ok && at >= 0 && e->cs[at].free_ > 0 && need > 0;
      if (ok) {
        share[i] = e->cs[at].free_ < need ? e->cs[at].free_ : need;
        need -= share[i];
      }
    }
    if (ok) {
      svc s = {e->cur.arrival + e->cur.holding, pair, idp, 0, a[1], nch, 0, nch, {0}, {0}, {0}};
      for (int i = 0; i < nch; i++) {
        int64_t at = cs_find(e, pair, idp, a[1 + i]);
        chan_state r = e->cs[at];
        cs_remove(e, at); /* :643-645 remove, then append the updated tuple */
        cs_append(e, r.pair, r.idp, r.ch, r.used + share[i], r.free_ - share[i], r.cap);
        s.ch[i] = (int16_t)a[1 + i];
        s.share[i] = (int8_t)share[i];
        s.cap[i] = (int8_t)r.cap;
      }
      e->services_accepted++; /* _service_acceptance(True) :767-778 */
      e->ep_accepted++;
      e->br_provisioned += e->cur.bit_rate;
      e->ep_br_provisioned += e->cur.bit_rate;
      s.uid = e->next_uid++;
      s.arrival = e->cur.arrival;
      run_append(e, s.uid); /* :656 */
      heap_push(e, s);
      accepted = 1;
    }
    if (alloc) {
      alloc[0] = accepted;
      alloc[1] = accepted ? path : -1;
      alloc[2] = accepted ? a[1] : -1;
      alloc[3] = accepted ? nch : 0;
    }
    return accepted;
  }
  int ok = path >= 0 && path < k && nch > 0 && hops_of(e, pair, path) > 0 && e->c.phy_table_id[pair] >= 0;
  for (int i = 0; ok && i < nch; i++) {
    ok = a[1 + i] < C;
    for (int q = 0; ok && q < i; q++) ok = a[1 + q] != a[1 + i];
  }
  for (int i = 0; ok && i < nch; i++) ok = is_channel_free(e, pair, path, a[1 + i]); /* is_path_free_on_channels */
  if (ok && groom) { /* with channel_state kept, every listed channel must carry part of the request */
    int need = e->cur.bit_rate / 100, table = e->c.phy_table_id[pair];
    for (int i = 0; ok && i < nch; i++) {
      ok = need > 0;
      need -= phy_level(e, table, a[1 + i], path);
    }
  }
  if (ok) {
    int table = e->c.phy_table_id[pair];
    const int16_t* L = links_of(e, pair, path, 0);
    svc s = {e->cur.arrival + e->cur.holding, pair, path, 0, a[1], nch, 0, nch, {0}, {0}, {0}};
    for (int i = 0; i < hops_of(e, pair, path); i++)
      for (int c = 0; c < nch; c++) AV(e, 0, L[i])[a[1 + c]] = 0; /* _provision_path :576-590 */
    int need = e->cur.bit_rate / 100;
    for (int c = 0; c < nch; c++) { /* :592-598 */
      int level = phy_level(e, table, a[1 + c], path);
      s.ch[c] = (int16_t)a[1 + c];
      s.cap[c] = (int8_t)level;
      s.share[c] = (int8_t)(level < need ? level : need);
      need -= s.share[c];
      if (groom && s.share[c] != level) /* :599-601 channel[2] != 0 */
        cs_append(e, pair, path, a[1 + c], s.share[c], level - s.share[c], level);
      e->total_gsnr_ep += e->c.phy_gsnr[((size_t)table * C + a[1 + c]) * k + path];
      e->total_mod_level_ep += level;
      e->channels_accepted_ep++;
    }
    e->services_accepted++; /* _service_acceptance(False) :767-778 */
    e->ep_accepted++;
    e->total_path_length_ep += e->c.path_length[(size_t)pair * k + path];
    e->total_path_index_ep += path + 1;
    e->physical_accepted_ep++;
    e->br_provisioned += e->cur.bit_rate;
    e->ep_br_provisioned += e->cur.bit_rate;
    s.uid = e->next_uid++;
    s.arrival = e->cur.arrival;
    run_append(e, s.uid); /* :620 */
    heap_push(e, s);
    accepted = 1;
  }
  if (alloc) {
    alloc[0] = accepted;
    alloc[1] = accepted ? path : -1;
    alloc[2] = accepted ? a[1] : -1;
    alloc[3] = accepted ? nch : 0;
  }
  return accepted;
}

/* ---- PhyRMSA periodic defragmentation: phy_rmsa_env.py:355-417 (step), _groom_defragmentation :703-733,
 * _move_virtual :735-764, _move :662-697.  Python iterates its lists by index while the loop bodies remove and
 * re-append elements; the loops below do the same on arrays, so the elements skipped / visited twice are the same. */
static svc* svc_by_uid(oracle_env* e, int64_t uid) {
  for (int64_t i = 0; i < e->heap_len; i++)
    if (e->heap[i].uid == uid) return &e->heap[i];
  return NULL;
}
static void svc_channel_remove(svc* s, int ci) { /* service.channels.remove(old_channel) */
  for (int i = ci; i + 1 < s->nch; i++) { s->ch[i] = s->ch[i + 1]; s->share[i] = s->share[i + 1]; s->cap[i] = s->cap[i + 1]; }
  s->nch--;
}
static void svc_channel_append(svc* s, int ch, int share, int cap) {
  s->ch[s->nch] = (int16_t)ch; s->share[s->nch] = (int8_t)share; s->cap[s->nch] = (int8_t)cap;
  s->nch++;
}
static void set_channel_on_path(oracle_env* e, const svc* s, int ch, int value) {
  const int16_t* L = links_of(e, s->pair, s->path, 0);
  for (int i = 0; i < hops_of(e, s->pair, s->path); i++) AV(e, 0, L[i])[ch] = (uint8_t)value;
}

/* _groom_defragmentation :703-733 */
static int groom_defragmentation(oracle_env* e) {
  int grooming_moves = 0;
  for (int64_t i = 0; i < e->run_len; i++) {       /* for service in running_services (live list) */
    const int64_t uid = e->run[i];
    svc* s = svc_by_uid(e, uid);
    if (!s) continue;
    for (int ci = 0; ci < s->nch; ci++) {          /* for channel in service.channels (live list) */
      const int c_ch = s->ch[ci], c_share = s->share[ci], c_cap = s->cap[ci];
      if (c_share != c_cap) {
        int64_t at = cs_find(e, s->pair, s->path, c_ch); /* corresponding_channel = [t for t in ... if t[0] == channel[0]][0] */
        if (at >= 0 && e->cs[at].used == c_share) {      /* nobody else uses the channel: it can move */
          for (int64_t j = 0; j < e->cs_len; j++) {      /* for target_channel in channel_state[key] */
            chan_state t = e->cs[j];
            if (t.pair != s->pair || t.idp != s->path) continue;
            if (j == at) continue;                       /* target_channel != corresponding_channel */
            if (t.free_ >= c_share) {
              /* two tuples removed, one appended (:717-724) */
              if (j > at) { cs_remove(e, j); cs_remove(e, at); } else { cs_remove(e, at); cs_remove(e, j); }
              cs_append(e, t.pair, t.idp, t.ch, t.used + c_share, t.free_ - c_share, t.cap);
              /* _move_virtual(service, (target[0], channel[1], updated[2], target[3], True), channel) */
              set_channel_on_path(e, s, c_ch, 1);
              run_remove(e, uid);
              run_append(e, uid);
              svc_channel_remove(s, ci);
              svc_channel_append(s, t.ch, c_share, t.cap);
              grooming_moves++;
              break;
            }
          }
        }
      }
      if (grooming_moves == e->c.phy_number_moves) return grooming_moves;
    }
  }
  return grooming_moves;
}

typedef struct defrag_cand { double diff, age; int ch; int64_t uid; int share, cap; } defrag_cand;

static void phy_defragmentation(oracle_env* e) {
  const int C = e->c.num_slots, k = e->c.k_paths, N = e->c.n_nodes;
  e->counted_moves_groom = groom_defragmentation(e); /* :357 */
  if (e->counted_moves_groom > e->c.phy_number_moves) return;
  /* candidates :360-380, in running_services order, then sorted by (-gain, -age), stable */
  int64_t cap = 64, n = 0;
  defrag_cand* cand = (defrag_cand*)malloc(sizeof(defrag_cand) * (size_t)cap);
  for (int64_t i = 0; i < e->run_len; i++) {
    const svc* s = svc_by_uid(e, e->run[i]);
    if (!s) continue;
    for (int ci = 0; ci < s->nch; ci++) {
      if (s->share[ci] != s->cap[ci]) continue; /* only whole channels are reallocated */
      double diff = e->c.phy_metric == 0 ? (double)calculate_r_cut_flag(e, s->pair, s->path, s->ch[ci], 1)
                                         : calculate_r_spatial_flag(e, s->pair, s->path, s->ch[ci], 1);
      if (diff > 0) {
        if (n == cap) { cap *= 2; cand = (defrag_cand*)realloc(cand, sizeof(defrag_cand) * (size_t)cap); }
        defrag_cand c = {diff, e->now - s->arrival, s->ch[ci], s->uid, s->share[ci], s->cap[ci]};
        int64_t pos = n++;
        while (pos > 0 && (c.diff > cand[pos - 1].diff || (c.diff == cand[pos - 1].diff && c.age > cand[pos - 1].age))) {
          cand[pos] = cand[pos - 1];
          pos--;
        }
        cand[pos] = c;
      }
    }
  }
  int num_moves = 0;
  const int cur_a = e->cur.src, cur_b = e->cur.dst;
  for (int64_t i = 0; i < n; i++) {
    svc* s = svc_by_uid(e, cand[i].uid);
    if (!s) continue;
    const int sa = s->pair / N, sb = s->pair % N;
    /* idp: the index of the candidate's path among the CURRENT request's k paths (dataclass equality), which is
     * the path's own index when both are between the same two nodes and otherwise the last value of the loop
     * variable, k - 1 (:386-391) */
    const int same_pair = (sa == cur_a && sb == cur_b) || (sa == cur_b && sb == cur_a);
    const int idp = same_pair ? s->path : k - 1;
    const int table = e->c.phy_table_id[s->pair];
    int have = 0, best_ch = -1;
    double best_metric = 0;
    for (int ch = 0; ch < C; ch++) { /* reallocation options :382-399, best of sorted(key=(-metric, channel)) */
      if (!is_channel_free(e, s->pair, s->path, ch)) continue;
      if (phy_level(e, table, ch, idp) != phy_level(e, table, cand[i].ch, idp)) continue;
      double metric = e->c.phy_metric == 0 ? (double)calculate_r_cut(e, s->pair, s->path, ch)
                                           : calculate_r_spatial(e, s->pair, s->path, ch);
      if (!have || metric > best_metric) { have = 1; best_metric = metric; best_ch = ch; }
    }
    if (have && -1 * best_metric < cand[i].diff) { /* :401-405 */
      /* _move(service, new_channel_number, old_channel) :662-697 */
      int ci = -1;
      for (int q = 0; q < s->nch; q++)
        if (s->ch[q] == cand[i].ch && s->share[q] == cand[i].share && s->cap[q] == cand[i].cap) { ci = q; break; }
      if (ci >= 0) {
        set_channel_on_path(e, s, best_ch, 0);
        set_channel_on_path(e, s, cand[i].ch, 1);
        run_remove(e, s->uid);
        run_append(e, s->uid);
        svc_channel_remove(s, ci);
        svc_channel_append(s, best_ch, cand[i].share, cand[i].cap);
        num_moves++;
        e->counted_moves++;
      }
    }
    if (num_moves + e->counted_moves_groom > e->c.phy_number_moves) break; /* :414 */
  }
  if (num_moves != 0) e->counted_defrag_cycles++;
  free(cand);
}

void oracle_phy_moves(const oracle_env* e, int64_t* out3) {
  out3[0] = e->counted_moves; out3[1] = e->counted_moves_groom; out3[2] = e->counted_defrag_cycles;
}

/* channel_state in append order: rows (pair, idp, channel, used, free, capacity); returns the number of entries */
int64_t oracle_phy_channel_state(const oracle_env* e, int32_t* out, int64_t max_entries) {
  for (int64_t i = 0; i < e->cs_len && i < max_entries; i++) {
    const chan_state* r = &e->cs[i];
    int32_t* o = out + i * 6;
    o[0] = r->pair; o[1] = r->idp; o[2] = r->ch; o[3] = r->used; o[4] = r->free_; o[5] = r->cap;
  }
  return e->cs_len;
}

/* _calculate_total_cuts :1195-1203, calculate_total_r_spatial :1110-1121 */
static void phy_fragmentation(const oracle_env* e, double* cuts, double* rss) {
  int E = e->c.n_edges, C = e->c.num_slots;
  int64_t number_cut = 0;
  double r_spatial = 0;
  uint8_t col[256];
  int pos[256], len[256];
  uint8_t val[256];
  for (int ch = 0; ch < C; ch++) {
    for (int l = 0; l < E; l++) col[l] = AV(e, 0, l)[ch];
    int nr = rle(col, E, pos, val, len);
    int64_t sum_sq = 0, sum = 0;
    for (int r = 0; r < nr; r++) {
      number_cut += val[r]; /* np.sum(values) */
      if (val[r] == 1) { sum_sq += (int64_t)len[r] * len[r]; sum += len[r]; }
    }
    r_spatial += sqrt((double)sum_sq) / (double)(sum + 1);
  }
  *cuts = (double)number_cut / (double)C;
  *rss = r_spatial / (double)C;
}

/* out[6], out[7]: as computed by the last step() for its info dict, i.e. after the provisioning
 * and BEFORE the releases of _next_service (phy_rmsa_env.py:313-314 precede :351) */
void oracle_phy_stats(const oracle_env* e, double* out) {
  out[0] = e->total_path_length_ep;
  out[1] = e->total_gsnr_ep;
  out[2] = (double)e->total_path_index_ep;
  out[3] = (double)e->total_mod_level_ep;
  out[4] = (double)e->channels_accepted_ep;
  out[5] = (double)e->physical_accepted_ep;
  out[6] = e->info_cuts;
  out[7] = e->info_rss;
}

/* RMSAEnv.step rmsa_env.py:222-341 (decision part) */
static int rmsa_decide(oracle_env* e, int path, int slot, int32_t* alloc) {
  int pair = cur_pair(e), accepted = 0, n = 0;
  if (e->c.qot_gate && e->c.env_kind == 0) {
    /* RMSA-v0 with the GN-model QoT gate: the modulation is path.best_modulation (rmsa_env.py:708-719) or,
     * with qot_adaptive, the most efficient one whose candidate passes the gate at (path, slot) */
    e->qot_last[0] = e->qot_last[1] = NAN;
    if (path >= 0 && slot >= 0 && path < e->c.k_paths && slot < e->c.num_slots && hops_of(e, pair, path) > 0) {
      int nse = e->c.qot_adaptive ? e->c.qot_n_se : 1;
      for (int t = 0; t < nse && !accepted; t++) {
        int se = e->c.qot_adaptive ? e->c.qot_se[t] : e->c.path_se[(size_t)pair * e->c.k_paths + path];
        n = number_slots(e, se);
        if (!is_path_free(e, pair, path, 0, slot, n)) break; /* a wider channel does not fit either */
        if (e->c.qot_adaptive && !qot_alone_ok(e, pair, path, slot, n, se)) continue;
        if (qot_gate_eval(e, pair, path, 0, slot, n, se)) {
          provision_path(e, pair, path, 0, slot, n);
          accepted = 1;
          svc s = {e->cur.arrival + e->cur.holding, pair, path, 0, slot, n, se};
          heap_push(e, s);
        }
      }
    }
    if (alloc) {
      alloc[0] = accepted;
      alloc[1] = accepted ? path : -1;
      alloc[2] = accepted ? slot : -1;
      alloc[3] = accepted ? n : 0;
    }
    return accepted;
  }
  if (path >= 0 && slot >= 0 && path < e->c.k_paths && slot < e->c.num_slots &&
      hops_of(e, pair, path) > 0) {
    n = number_slots(e, e->c.path_se[(size_t)pair * e->c.k_paths + path]);
    if (is_path_free(e, pair, path, 0, slot, n)) {
      provision_path(e, pair, path, 0, slot, n);
      accepted = 1;
      svc s = {e->cur.arrival + e->cur.holding, pair, path, 0, slot, n,
               e->c.path_se[(size_t)pair * e->c.k_paths + path]};
      heap_push(e, s); /* _add_release */
    }
  }
  if (alloc) {
    alloc[0] = accepted;
    alloc[1] = accepted ? path : -1;
    alloc[2] = accepted ? slot : -1;
    alloc[3] = accepted ? n : 0;
  }
  return accepted;
}

void oracle_step(oracle_env* e, const int32_t* a, int32_t* reward, uint8_t* done, double* obs,
                 int32_t* alloc) {
  int accepted = 0;
  int k = e->c.k_paths, S = e->c.num_slots;
  const int want_info = e->c.info_stats && e->c.env_kind <= 1;
  double previous_network_compactness = want_info ? network_compactness(e) : 0.0; /* rmsa_env.py:227-229 */
  if (e->c.env_kind == 0) {
    accepted = rmsa_decide(e, a[0], a[1], alloc);
  } else if (e->c.env_kind == 1) { /* DeepRMSAEnv.step deeprmsa_env.py:48-58 */
    int act = a[0], path = k, slot = S;
    int pair = cur_pair(e);
    if (act >= 0 && act < k * e->c.j) {
      int route = act / e->c.j, block = act % e->c.j;
      if (hops_of(e, pair, route) > 0) {
        int starts[64], lens[64];
        int n = number_slots(e, e->c.path_se[(size_t)pair * k + route]);
        int nb = available_blocks(e, pair, route, n, starts, lens);
        if (block < nb) { path = route; slot = starts[block]; }
      }
    }
    accepted = rmsa_decide(e, path, slot, alloc);
  } else if (e->c.env_kind == 3) {
    accepted = phy_decide(e, a, alloc);
    phy_fragmentation(e, &e->info_cuts, &e->info_rss);
  } else { /* RMCSAEnv.step rmcsa_env.py:209-339 */
    int path = a[0], mod = a[1], core = a[2], slot = a[3], n = 0;
    int pair = cur_pair(e);
    if (e->c.qot_gate) e->qot_last[0] = e->qot_last[1] = NAN;
    if (path >= 0 && mod >= 0 && core >= 0 && slot >= 0 && path < k && mod < e->c.n_mods &&
        core < e->c.num_cores && slot < S && hops_of(e, pair, path) > 0) {
      n = number_slots(e, e->c.mod_se[mod]);
      if (is_path_free(e, pair, path, core, slot, n) &&
          crosstalk_ok(e, mod, e->c.path_length[(size_t)pair * k + path]) &&
          (!e->c.qot_gate || qot_gate_eval(e, pair, path, core, slot, n, e->c.mod_se[mod]))) {
        provision_path(e, pair, path, core, slot, n);
        accepted = 1;
        svc s = {e->cur.arrival + e->cur.holding, pair, path, core, slot, n, e->c.mod_se[mod]};
        heap_push(e, s);
      }
    }
    if (alloc) {
      alloc[0] = accepted;
      alloc[1] = accepted ? (path | (core << 16)) : -1;
      alloc[2] = accepted ? slot : -1;
      alloc[3] = accepted ? n : 0;
    }
  }
  if (want_info) make_info(e, previous_network_compactness);
  finish_step(e, accepted, reward, done);
  if (obs && e->c.env_kind == 1) oracle_observation(e, obs);
}

/* DeepRMSAEnv.observation deeprmsa_env.py:60-121 */
void oracle_observation(oracle_env* e, double* obs) {
  int N = e->c.n_nodes, k = e->c.k_paths, j = e->c.j, S = e->c.num_slots;
  int dim = oracle_obs_dim(e), pair = cur_pair(e);
  double* tau = obs + 1;
  double* sp = obs + 1 + 2 * N;
  for (int i = 0; i < dim; i++) obs[i] = 0.0;
  obs[0] = (double)e->cur.bit_rate / 100.0;
  int mn = e->cur.src < e->cur.dst ? e->cur.src : e->cur.dst;
  int mx = e->cur.src < e->cur.dst ? e->cur.dst : e->cur.src;
  tau[mn] = 1.0;
  tau[N + mx] = 1.0;
  for (int i = 0; i < k * (2 * j + 3); i++) sp[i] = -1.0;
  uint8_t av[1024], val[1024];
  int pos[1024], len[1024], starts[64], lens[64];
  for (int p = 0; p < k; p++) {
    if (hops_of(e, pair, p) <= 0) continue;
    double* row = sp + (size_t)p * (2 * j + 3);
    available_slots(e, pair, p, av);
    int n = number_slots(e, e->c.path_se[(size_t)pair * k + p]);
    int nb = available_blocks(e, pair, p, n, starts, lens);
    for (int b = 0; b < nb; b++) {
      row[2 * b] = 2.0 * ((double)starts[b] - 0.5 * (double)S) / (double)S;
      row[2 * b + 1] = ((double)lens[b] - 8.0) / 8.0;
    }
    row[2 * j] = ((double)n - 5.5) / 3.5;
    int r = rle(av, S, pos, val, len);
    long tot = 0, cnt = 0, sumlen = 0;
    for (int s = 0; s < S; s++) tot += av[s];
    row[2 * j + 1] = 2.0 * ((double)tot - 0.5 * (double)S) / (double)S;
    for (int i = 0; i < r; i++)
      if (val[i] == 1) { cnt++; sumlen += len[i]; }
    if (cnt > 0) row[2 * j + 2] = ((double)sumlen / (double)cnt - 4.0) / 4.0; /* np.mean */
  }
}

/* SimpleMatrixObservation rmsa_env.py:940-971 / rmcsa_env.py:914-947 */
void oracle_matrix_observation(oracle_env* e, uint8_t* obs) {
  int N = e->c.n_nodes;
  size_t cells = (size_t)e->c.num_cores * e->c.n_edges * e->c.num_slots;
  memset(obs, 0, 2 * (size_t)N);
  int mn = e->cur.src < e->cur.dst ? e->cur.src : e->cur.dst;
  int mx = e->cur.src < e->cur.dst ? e->cur.dst : e->cur.src;
  obs[mn] = 1;
  obs[N + mx] = 1;
  memcpy(obs + 2 * N, e->avail, cells);
}

/* first fit with the reference's exclusive upper bound range(0, S - n) (F8) */
static int first_fit(const oracle_env* e, int pair, int path, int core, int n) {
  for (int s = 0; s < e->c.num_slots - n; s++)
    if (is_path_free(e, pair, path, core, s, n)) return s;
  return -1;
}

/* Shortest available path, first fit, GN-model QoT (heuristic id 13): the first candidate, in the order the
 * reference's first-fit heuristics walk them (rmsa_env.py:901-913, rmcsa_env.py:882-911), that passes the QoT
 * gate.  RMSA-v0: per path the modulation is path.best_modulation or, adaptive, every spectral efficiency in
 * descending order; RMCSA-v0: per path the best modulation by reach (utils.py:105-117) or, adaptive, every
 * modulation by descending spectral efficiency, on the first core that has room. */
static void qot_heuristic(oracle_env* e, int32_t* a) {
  int k = e->c.k_paths, S = e->c.num_slots, pair = cur_pair(e);
  e->qot_last[0] = e->qot_last[1] = NAN;
  if (e->c.env_kind == 0) {
    a[0] = k; a[1] = S;
    int nse = e->c.qot_adaptive ? e->c.qot_n_se : 1;
    for (int p = 0; p < k; p++) {
      if (hops_of(e, pair, p) <= 0) continue;
      for (int t = 0; t < nse; t++) {
        int se = e->c.qot_adaptive ? e->c.qot_se[t] : e->c.path_se[(size_t)pair * k + p];
        int n = number_slots(e, se);
        int s = first_fit(e, pair, p, 0, n);
        if (s < 0) break;
        if (e->c.qot_adaptive && !qot_alone_ok(e, pair, p, s, n, se)) continue;
        if (qot_gate_eval(e, pair, p, 0, s, n, se)) { a[0] = p; a[1] = s; return; }
      }
    }
    return;
  }
  a[0] = k; a[1] = e->c.n_mods; a[2] = e->c.num_cores; a[3] = S;
  int order[64];
  int nm = mods_by_se_desc(e, order);
  for (int p = 0; p < k; p++) {
    if (hops_of(e, pair, p) <= 0) continue;
    double len = e->c.path_length[(size_t)pair * k + p];
    int first = -1; /* get_best_modulation_format: the first in SE-descending order that reaches */
    for (int i = 0; i < nm && first < 0; i++)
      if (len <= e->c.mod_max_length[order[i]]) first = i;
    int lo = e->c.qot_adaptive ? 0 : first, hi = e->c.qot_adaptive ? nm : first + 1;
    if (lo < 0) continue;
    for (int i = lo; i < hi; i++) {
      int m = order[i], se = e->c.mod_se[m];
      int n = number_slots(e, se), core = -1, s = -1;
      for (int c = 0; c < e->c.num_cores && s < 0; c++) {
        s = first_fit(e, pair, p, c, n);
        if (s >= 0) core = c;
      }
      if (s < 0) break; /* no core has room; a wider channel does not fit either */
      if (!crosstalk_ok(e, m, len)) continue;
      if (e->c.qot_adaptive && !qot_alone_ok(e, pair, p, s, n, se)) continue;
      if (qot_gate_eval(e, pair, p, core, s, n, se)) { a[0] = p; a[1] = m; a[2] = core; a[3] = s; return; }
    }
  }
}

void oracle_heuristic(oracle_env* e, int hid, int32_t* a) {
  if (e->c.env_kind == 3) { phy_heuristic(e, hid, a); return; }
  if (hid == 13 && e->c.qot_gate && (e->c.env_kind == 0 || e->c.env_kind == 2)) { qot_heuristic(e, a); return; }
  int k = e->c.k_paths, S = e->c.num_slots, pair = cur_pair(e);
  if (e->c.env_kind == 0) {
    a[0] = k; a[1] = S;
    if (hid == 0) { /* shortest_path_first_fit rmsa_env.py:854-871 */
      if (hops_of(e, pair, 0) > 0) {
        int s = first_fit(e, pair, 0, 0, number_slots(e, e->c.path_se[(size_t)pair * k]));
        if (s >= 0) { a[0] = 0; a[1] = s; }
      }
    } else if (hid == 1) { /* shortest_available_path_first_fit rmsa_env.py:901-913 */
      for (int p = 0; p < k; p++) {
        if (hops_of(e, pair, p) <= 0) continue;
        int s = first_fit(e, pair, p, 0, number_slots(e, e->c.path_se[(size_t)pair * k + p]));
        if (s >= 0) { a[0] = p; a[1] = s; return; }
      }
    } else if (hid == 2) { /* least_loaded_path_first_fit rmsa_env.py:916-937 */
      long max_free = 0;
      uint8_t av[1024];
      for (int p = 0; p < k; p++) {
        if (hops_of(e, pair, p) <= 0) continue;
        int s = first_fit(e, pair, p, 0, number_slots(e, e->c.path_se[(size_t)pair * k + p]));
        if (s >= 0) {
          available_slots(e, pair, p, av);
          long fr = 0;
          for (int i = 0; i < S; i++) fr += av[i];
          if (fr > max_free) { a[0] = p; a[1] = s; max_free = fr; }
        }
      }
    }
  } else if (e->c.env_kind == 1) {
    int starts[64], lens[64];
    if (hid == 0) { /* deeprmsa_env.py:135-143 */
      a[0] = 0;
      if (e->c.allow_rejection) {
        int nb = hops_of(e, pair, 0) > 0
                     ? available_blocks(e, pair, 0, number_slots(e, e->c.path_se[(size_t)pair * k]), starts, lens)
                     : 0;
        a[0] = nb > 0 ? 0 : k * e->c.j;
      }
    } else { /* deeprmsa_env.py:146-155 */
      a[0] = k * e->c.j;
      for (int p = 0; p < k; p++) {
        if (hops_of(e, pair, p) <= 0) continue;
        int nb = available_blocks(e, pair, p, number_slots(e, e->c.path_se[(size_t)pair * k + p]), starts, lens);
        if (nb > 0) { a[0] = p * e->c.j; return; }
      }
    }
  } else { /* shortest_available_path_best_modulation_first_core_first_fit rmcsa_env.py:882-911 */
    a[0] = k; a[1] = e->c.n_mods; a[2] = e->c.num_cores; a[3] = S;
    for (int p = 0; p < k; p++) {
      if (hops_of(e, pair, p) <= 0) continue;
      double len = e->c.path_length[(size_t)pair * k + p];
      /* get_best_modulation_format utils.py:105-117: highest SE first (stable), first that reaches */
      int best = -1;
      for (int se_rank = 0; se_rank < e->c.n_mods && best < 0; se_rank++) {
        /* pick the se_rank-th modulation in (SE desc, original order) */
        int chosen = -1, seen = 0;
        for (int pass_se = 1 << 30; chosen < 0;) {
          int top = -1;
          for (int m = 0; m < e->c.n_mods; m++)
            if (e->c.mod_se[m] < pass_se && (top < 0 || e->c.mod_se[m] > e->c.mod_se[top])) top = m;
          if (top < 0) break;
          int se = e->c.mod_se[top];
          for (int m = 0; m < e->c.n_mods; m++)
            if (e->c.mod_se[m] == se) { if (seen == se_rank) { chosen = m; break; } seen++; }
          pass_se = se;
        }
        if (chosen >= 0 && len <= e->c.mod_max_length[chosen]) best = chosen;
      }
      if (best < 0) continue;
      int n = number_slots(e, e->c.mod_se[best]);
      for (int core = 0; core < e->c.num_cores; core++) {
        int s = first_fit(e, pair, p, core, n);
        if (s >= 0) { a[0] = p; a[1] = best; a[2] = core; a[3] = s; return; }
      }
    }
  }
}

void oracle_counters(const oracle_env* e, int64_t* o) {
  o[0] = e->services_processed; o[1] = e->services_accepted;
  o[2] = e->ep_processed; o[3] = e->ep_accepted;
  o[4] = e->br_requested; o[5] = e->br_provisioned;
  o[6] = e->ep_br_requested; o[7] = e->ep_br_provisioned;
  o[8] = e->requests_drawn; o[9] = e->heap_len; o[10] = 0; o[11] = e->episodes;
}

void oracle_state(const oracle_env* e, uint8_t* out) {
  memcpy(out, e->avail, (size_t)e->c.num_cores * e->c.n_edges * e->c.num_slots);
}

void oracle_request_now(const oracle_env* e, oracle_request* out) { *out = e->cur; }

/* ---- timed CPU baseline: independent envs spread over host threads ---- */
typedef struct rollout_job {
  const oracle_config* cfg;
  int64_t env0, n_envs;
  int nsteps, hid;
  int64_t* counters_out;
  uint8_t* state_out;
  int64_t* reward_sum_out;
  int64_t next; /* shared work counter (atomic) */
} rollout_job;

static void* rollout_worker(void* arg) {
  rollout_job* jb = (rollout_job*)arg;
  const oracle_config* cfg = jb->cfg;
  size_t cells = (size_t)(cfg->num_cores < 1 ? 1 : cfg->num_cores) * cfg->n_edges * cfg->num_slots;
  for (;;) {
    int64_t i = __atomic_fetch_add(&jb->next, 1, __ATOMIC_RELAXED);
    if (i >= jb->n_envs) break;
    oracle_env* e = oracle_create(cfg, jb->env0 + i);
    oracle_reset(e, 1);
    int32_t a[16], r;
    uint8_t d;
    int64_t rs = 0;
    double obs[1024];
    for (int t = 0; t < jb->nsteps; t++) {
      oracle_heuristic(e, jb->hid, a);
      oracle_step(e, a, &r, &d, cfg->env_kind == 1 ? obs : NULL, NULL);
      rs += r;
    }
    if (jb->counters_out) oracle_counters(e, jb->counters_out + i * 12);
    if (jb->state_out) oracle_state(e, jb->state_out + (size_t)i * cells);
    if (jb->reward_sum_out) jb->reward_sum_out[i] = rs;
    oracle_destroy(e);
  }
  return NULL;
}

int64_t oracle_rollout_many(const oracle_config* cfg, int64_t env0, int64_t n_envs, int nsteps,
                            int hid, int threads, int64_t* counters_out, uint8_t* state_out,
                            int64_t* reward_sum_out) {
  rollout_job jb = {cfg, env0, n_envs, nsteps, hid, counters_out, state_out, reward_sum_out, 0};
  if (threads < 1) threads = 1;
  if (threads > 256) threads = 256;
  pthread_t th[256];
  for (int t = 1; t < threads; t++) pthread_create(&th[t], NULL, rollout_worker, &jb);
  rollout_worker(&jb);
  for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
  return n_envs * (int64_t)nsteps;
}

/* ---- persistent pool of envs (bench.py --impl reference / cpu_baseline): the
 * envs keep their network state between calls, like a VecEnv of reference envs ---- */
struct oracle_pool {
  oracle_env** envs;
  int64_t n;
  int64_t next;
  int nsteps, hid;
  int64_t reward_sum;
  const oracle_gsnr_params* gsnr; /* non-NULL: evaluate calculate_osnr for every decision (bench C3 arm) */
  double gsnr_sum;
};

oracle_pool* oracle_pool_create(const oracle_config* cfg, int64_t env0, int64_t n_envs) {
  oracle_pool* p = (oracle_pool*)calloc(1, sizeof(oracle_pool));
  p->envs = (oracle_env**)calloc((size_t)n_envs, sizeof(oracle_env*));
  p->n = n_envs;
  for (int64_t i = 0; i < n_envs; i++) {
    p->envs[i] = oracle_create(cfg, env0 + i);
    oracle_reset(p->envs[i], 1);
  }
  return p;
}

void oracle_pool_destroy(oracle_pool* p) {
  if (!p) return;
  for (int64_t i = 0; i < p->n; i++) oracle_destroy(p->envs[i]);
  free(p->envs);
  free(p);
}

static pthread_mutex_t pool_sum_lock = PTHREAD_MUTEX_INITIALIZER;
static void* pool_worker(void* arg) {
  oracle_pool* p = (oracle_pool*)arg;
  int64_t rs = 0;
  double gs = 0.0;
  double obs[1024];
  for (;;) {
    int64_t i = __atomic_fetch_add(&p->next, 1, __ATOMIC_RELAXED);
    if (i >= p->n) break;
    oracle_env* e = p->envs[i];
    int32_t a[16], r;
    uint8_t d;
    for (int t = 0; t < p->nsteps; t++) {
      oracle_heuristic(e, p->hid, a);
      if (p->gsnr && (e->c.env_kind == 0 || e->c.env_kind == 2)) {
        /* the candidate the action describes, slot count from the pending bit rate (get_number_slots) */
        int path = a[0], k = e->c.k_paths;
        if (path >= 0 && path < k) {
          int pair = e->cur.src * e->c.n_nodes + e->cur.dst, q[4];
          double o2[2];
          if (e->c.env_kind == 0) {
            int se = e->c.path_se[(size_t)pair * k + path];
            q[0] = path; q[1] = a[1]; q[2] = number_slots(e, se); q[3] = se;
          } else {
            int se = e->c.mod_se[a[1]];
            q[0] = path | (a[2] << 16); q[1] = a[3]; q[2] = number_slots(e, se); q[3] = se;
          }
          oracle_gsnr(e, p->gsnr, q, o2);
          if (o2[0] == o2[0]) gs += o2[0];
        }
      }
      oracle_step(e, a, &r, &d, e->c.env_kind == 1 ? obs : NULL, NULL);
      rs += r;
    }
  }
  __atomic_fetch_add(&p->reward_sum, rs, __ATOMIC_RELAXED);
  if (p->gsnr) {
    pthread_mutex_lock(&pool_sum_lock);
    p->gsnr_sum += gs;
    pthread_mutex_unlock(&pool_sum_lock);
  }
  return NULL;
}

/* evaluate the GN model for every decision of the pool's rollouts (NULL switches it off); g must outlive them */
void oracle_pool_set_gsnr(oracle_pool* p, const oracle_gsnr_params* g) { p->gsnr = g; }
double oracle_pool_gsnr_sum(const oracle_pool* p) { return p->gsnr_sum; }

int64_t oracle_pool_rollout(oracle_pool* p, int nsteps, int hid, int threads) {
  p->next = 0;
  p->nsteps = nsteps;
  p->hid = hid;
  if (threads < 1) threads = 1;
  if (threads > 256) threads = 256;
  pthread_t th[256];
  for (int t = 1; t < threads; t++) pthread_create(&th[t], NULL, pool_worker, p);
  pool_worker(p);
  for (int t = 1; t < threads; t++) pthread_join(th[t], NULL);
  return p->n * (int64_t)nsteps;
}

oracle_env* oracle_pool_env(oracle_pool* p, int64_t i) { return (i >= 0 && i < p->n) ? p->envs[i] : NULL; }
int64_t oracle_pool_reward_sum(const oracle_pool* p) { return p->reward_sum; }


/* ---- GN-model GSNR: calculate_osnr, examples/calculate_osnr.py:9-56 ------------------
 * Literal restatement: explicit Span objects per link (the commented-out sketch at
 * optical_rl_gym/utils.py:38-54, built as in examples/create_topology_gn.py:122-134), a
 * running_services list per link, one loop iteration per span.  Parity: the reference
 * function is not importable as shipped; tests/golden/gsnr_*.npz were made by
 * calling the UNMODIFIED function body through oracle/refharness/gen_gsnr_golden.py, which
 * injects the missing Span/Link classes into optical_rl_gym.utils before the import. */
typedef struct gn_span { double length, attenuation_normalized, noise_figure_normalized; } gn_span;
typedef struct gn_service { int id; double bandwidth, center_frequency, launch_power; int se; } gn_service;

static double gn_calculate_osnr(const gn_service* cur, int n_links, const gn_span* const* spans,
                                const int* n_spans, const gn_service* const* running, const int* n_running) {
  const double pi = 3.141592653589793;
  double beta_2 = -21.3e-27, gamma = 1.3e-3, h_plank = 6.626e-34; /* :10-12 */
  double acc_gsnr = 0, l_eff_a, l_eff, phi, sum_phi, power_ase, power_nli_span;
  const double phi_modulation_format[6] = {1, 1, 2.0 / 3, 17.0 / 25, 69.0 / 100, 13.0 / 21}; /* :20 */
  for (int li = 0; li < n_links; li++) {           /* for link in current_service.path.links */
    for (int si = 0; si < n_spans[li]; si++) {     /* for span in ...["link"].spans */
      const gn_span* span = &spans[li][si];
      l_eff_a = 1 / (2 * span->attenuation_normalized);
      l_eff = (1 - exp(-2 * span->attenuation_normalized * span->length * 1e3)) / (2 * span->attenuation_normalized);
      sum_phi = asinh(pi * pi * fabs(beta_2) * (cur->bandwidth * cur->bandwidth) / (4 * span->attenuation_normalized));
      for (int ri = 0; ri < n_running[li]; ri++) { /* for service in ...["running_services"] */
        const gn_service* sv = &running[li][ri];
        if (sv->id != cur->id) {
          phi = (asinh(pi * pi * fabs(beta_2) * l_eff_a * sv->bandwidth *
                       (sv->center_frequency - cur->center_frequency + (sv->bandwidth / 2))) -
                 asinh(pi * pi * fabs(beta_2) * l_eff_a * sv->bandwidth *
                       (sv->center_frequency - cur->center_frequency - (sv->bandwidth / 2)))) -
                (phi_modulation_format[sv->se - 1] *
                 (sv->bandwidth / fabs(sv->center_frequency - cur->center_frequency)) * 5 / 3 *
                 (l_eff / (span->length * 1e3)));
          sum_phi += phi; /* applied to other services only, see the header of csrc/gsnr.cuh */
        }
      }
      power_nli_span = pow(cur->launch_power / cur->bandwidth, 3) * (8 / (27 * pi * fabs(beta_2))) *
                       (gamma * gamma) * l_eff * sum_phi * cur->bandwidth;
      power_ase = cur->bandwidth * h_plank * cur->center_frequency *
                  (exp(2 * span->attenuation_normalized * span->length * 1e3) - 1) * span->noise_figure_normalized;
      acc_gsnr += 1 / (cur->launch_power / (power_ase + power_nli_span));
    }
  }
  return 10 * log10(1 / acc_gsnr);
}

static gn_service gn_service_of(const oracle_gsnr_params* g, int id, int start, int n, int se) {
  gn_service s;
  s.id = id;
  s.bandwidth = (double)n * g->slot_hz;
  s.center_frequency = g->f0_hz + ((double)start + 0.5 * (double)n) * g->slot_hz;
  s.launch_power = g->launch_power_w;
  s.se = se;
  return s;
}

static int path_uses(const oracle_env* e, int pair, int path, int link) {
  const int16_t* L = links_of(e, pair, path, 0);
  for (int h = 0; h < hops_of(e, pair, path); h++) if (L[h] == link) return 1;
  return 0;
}

/* GSNR of `target` (id tid; tid = -1: the candidate) on its path, with running_services per
 * link = the env's live services on that link and core (+ the candidate when add_cand). */
static double gsnr_of(const oracle_env* e, const oracle_gsnr_params* g, int tpair, int tpath, int tcore,
                      const gn_service* target, int add_cand, int cpair, int cpath, int ccore,
                      const gn_service* cand) {
  int hops = hops_of(e, tpair, tpath);
  const int16_t* L = links_of(e, tpair, tpath, 0);
  gn_span** spans = (gn_span**)calloc(hops, sizeof(gn_span*));
  gn_service** running = (gn_service**)calloc(hops, sizeof(gn_service*));
  int* n_spans = (int*)calloc(hops, sizeof(int));
  int* n_running = (int*)calloc(hops, sizeof(int));
  for (int h = 0; h < hops; h++) {
    int l = L[h];
    n_spans[h] = g->span_count[l];
    spans[h] = (gn_span*)malloc(sizeof(gn_span) * n_spans[h]);
    for (int s = 0; s < n_spans[h]; s++) {
      spans[h][s].length = g->span_length_km[l];
      spans[h][s].attenuation_normalized = g->attenuation_per_m;
      spans[h][s].noise_figure_normalized = g->noise_figure_lin;
    }
    running[h] = (gn_service*)malloc(sizeof(gn_service) * (e->heap_len + 1));
    for (int64_t i = 0; i < e->heap_len; i++) {
      const svc* s = &e->heap[i];
      if (s->core == tcore && path_uses(e, s->pair, s->path, l))
        running[h][n_running[h]++] = gn_service_of(g, (int)i, s->initial_slot, s->number_slots, s->se);
    }
    if (add_cand && ccore == tcore && path_uses(e, cpair, cpath, l)) running[h][n_running[h]++] = *cand;
  }
  double v = gn_calculate_osnr(target, hops, (const gn_span* const*)spans, n_spans,
                               (const gn_service* const*)running, n_running);
  for (int h = 0; h < hops; h++) { free(spans[h]); free(running[h]); }
  free(spans); free(running); free(n_spans); free(n_running);
  return v;
}

void oracle_gsnr(const oracle_env* e, const oracle_gsnr_params* g, const int32_t* query, double* out2) {
  int path = query[0] & 0xFFFF, core = query[0] >> 16, start = query[1], n = query[2];
  int pair = cur_pair(e);
  out2[0] = out2[1] = NAN;
  if (query[0] < 0 || path >= e->c.k_paths || core >= e->c.num_cores || start < 0 || n < 1 ||
      start + n > e->c.num_slots || hops_of(e, pair, path) <= 0)
    return;
  int se = query[3] > 0 ? query[3] : e->c.path_se[(size_t)pair * e->c.k_paths + path];
  gn_service cand = gn_service_of(g, -1, start, n, se);
  out2[0] = gsnr_of(e, g, pair, path, core, &cand, 0, pair, path, core, &cand);
  double mn = INFINITY;
  const int16_t* CL = links_of(e, pair, path, 0);
  for (int64_t i = 0; i < e->heap_len; i++) {
    const svc* s = &e->heap[i];
    if (s->core != core) continue;
    int shares = 0;
    for (int h = 0; h < hops_of(e, pair, path); h++) shares |= path_uses(e, s->pair, s->path, CL[h]);
    if (!shares) continue;
    gn_service t = gn_service_of(g, (int)i, s->initial_slot, s->number_slots, s->se);
    double v = gsnr_of(e, g, s->pair, s->path, s->core, &t, 1, pair, path, core, &cand);
    if (v < mn) mn = v;
  }
  out2[1] = mn;
}

/* live services of the env: rows of (pair, path, core, initial_slot, number_slots, se) */
int64_t oracle_services(const oracle_env* e, int32_t* out6, int64_t max_rows) {
  int64_t n = e->heap_len < max_rows ? e->heap_len : max_rows;
  for (int64_t i = 0; i < n; i++) {
    const svc* s = &e->heap[i];
    int32_t* o = out6 + 6 * i;
    o[0] = s->pair; o[1] = s->path; o[2] = s->core; o[3] = s->initial_slot; o[4] = s->number_slots; o[5] = s->se;
  }
  return e->heap_len;
}

/* ---- GN-model QoT gate (oracle_config.qot_gate) ---------------------------------------------------
 * accept iff the candidate's GSNR meets the threshold of its modulation and every neighbour's GSNR,
 * re-evaluated with the candidate present, still meets its own.  Each GSNR is one literal call of
 * calculate_osnr; nothing is cached.  qot_last = (candidate, lowest neighbour; NaN when the candidate
 * itself failed and the neighbours were not looked at). */
static long long g_qot_stat[4]; /* workload statistics for DESIGN.md: gate calls, candidate misses, neighbours, neighbour hops */
void oracle_qot_stats(long long* out4) { for (int i = 0; i < 4; i++) out4[i] = __atomic_load_n(&g_qot_stat[i], __ATOMIC_RELAXED); }
static int qot_gate_eval(oracle_env* e, int pair, int path, int core, int start, int n, int se) {
  const oracle_gsnr_params* g = e->c.qot_gsnr;
  gn_service cand = gn_service_of(g, -1, start, n, se);
  double gc = gsnr_of(e, g, pair, path, core, &cand, 0, pair, path, core, &cand);
  e->qot_last[0] = gc;
  e->qot_last[1] = NAN;
  __atomic_fetch_add(&g_qot_stat[0], 1, __ATOMIC_RELAXED);
  if (!(gc >= e->c.qot_thr_db[se])) { __atomic_fetch_add(&g_qot_stat[1], 1, __ATOMIC_RELAXED); return 0; }
  int ok = 1;
  double mn = INFINITY;
  const int16_t* CL = links_of(e, pair, path, 0);
  for (int64_t i = 0; i < e->heap_len; i++) {
    const svc* s = &e->heap[i];
    if (s->core != core) continue;
    int shares = 0;
    for (int h = 0; h < hops_of(e, pair, path); h++) shares |= path_uses(e, s->pair, s->path, CL[h]);
    if (!shares) continue;
    gn_service t = gn_service_of(g, (int)i, s->initial_slot, s->number_slots, s->se);
    double v = gsnr_of(e, g, s->pair, s->path, s->core, &t, 1, pair, path, core, &cand);
    __atomic_fetch_add(&g_qot_stat[2], 1, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_qot_stat[3], hops_of(e, s->pair, s->path), __ATOMIC_RELAXED);
    if (v < mn) mn = v;
    if (!(v >= e->c.qot_thr_db[s->se])) ok = 0;
  }
  e->qot_last[1] = mn;
  return ok;
}

/* the channel alone on its path (no interferer): ASE and self-channel NLI only */
static int qot_alone_ok(const oracle_env* e, int pair, int path, int start, int n, int se) {
  const oracle_gsnr_params* g = e->c.qot_gsnr;
  gn_service cand = gn_service_of(g, -1, start, n, se);
  int hops = hops_of(e, pair, path);
  const int16_t* L = links_of(e, pair, path, 0);
  gn_span* spans[64] = {0};
  const gn_service* running[64] = {0};
  int n_spans[64] = {0}, n_running[64] = {0};
  for (int h = 0; h < hops; h++) {
    int l = L[h];
    n_spans[h] = g->span_count[l];
    spans[h] = (gn_span*)malloc(sizeof(gn_span) * n_spans[h]);
    for (int s = 0; s < n_spans[h]; s++) {
      spans[h][s].length = g->span_length_km[l];
      spans[h][s].attenuation_normalized = g->attenuation_per_m;
      spans[h][s].noise_figure_normalized = g->noise_figure_lin;
    }
    running[h] = NULL;
    n_running[h] = 0;
  }
  double v = gn_calculate_osnr(&cand, hops, (const gn_span* const*)spans, n_spans, running, n_running);
  for (int h = 0; h < hops; h++) free(spans[h]);
  return v >= e->c.qot_thr_db[se];
}

/* modulation rows by descending spectral efficiency, stable (the order get_best_modulation_format walks, utils.py:105-117) */
static int mods_by_se_desc(const oracle_env* e, int* order) {
  int nm = e->c.n_mods < 64 ? e->c.n_mods : 64;
  for (int m = 0; m < nm; m++) order[m] = m;
  for (int a = 1; a < nm; a++)
    for (int b = a; b > 0 && e->c.mod_se[order[b]] > e->c.mod_se[order[b - 1]]; b--) {
      int t = order[b]; order[b] = order[b - 1]; order[b - 1] = t;
    }
  return nm;
}

void oracle_qot_last(const oracle_env* e, double* out2) { out2[0] = e->qot_last[0]; out2[1] = e->qot_last[1]; }
