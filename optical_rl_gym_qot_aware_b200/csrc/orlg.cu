// orlg.cu — C-ABI (include/orlg.h) over the sm_100a stepping kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -shared -Xcompiler -fPIC
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "engine_launch.cuh"
#include "engine_util.cuh"
#include "gsnr.cuh"
#include "qot.cuh"

using namespace orlg;

struct orlg_engine {
  Params p{};
  orlg_config cfg{};
  int W = 0, BD = 0, device = 0;  // BD = threads per CTA
  double* info_out = nullptr;     // orlg_set_info_output
  int layout_bd = 0;              // envs per occupancy tile in HBM (1 = env-major, direct kernels)
  bool direct = false;
  size_t smem_bytes = 0;
  int obs_kind = 0, obs_dim = 0;
  std::vector<void*> allocs;
  size_t hbm_bytes = 0;
  long long launches = 0;
  std::string err;
  // pinned/device staging for orlg_step_host
  int* d_actions = nullptr;
  int* d_reward = nullptr;
  unsigned char* d_done = nullptr;
  void* d_obs = nullptr;
  int* d_alloc = nullptr;
  size_t obs_elem = 0;
  long long obs_es = 0, obs_fs = 1;  // 0 = row-major default
  GsnrParams gp{};
  bool gsnr_ready = false;
  cudaEvent_t host_ev = nullptr;  // end of the last orlg_step_host_async (blocking-sync event)
  bool blocking_wait = false;
  int* hint_actions = nullptr;
  int hint_heuristic = 0;
  int* auto_actions = nullptr;          // orlg_step_auto without a caller buffer
  long long* auto_reward_sum = nullptr; // set by orlg_rollout around its orlg_step_auto loop
  long long* snap = nullptr;            // orlg_set_counter_snapshot
  bool qot_proposal_valid = false;      // the last state-touching call was a heuristic-13 launch (launch_qot)
  int qot_cap = 64;                     // services staged per window by qot_kernel (64 / 128 / 256)
  cudaEvent_t mid_ev = nullptr;         // orlg_timed_steps: recorded between the launches of a multi-launch step
  bool mid_external = false;
};

static thread_local std::string g_create_error;

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess) {                                                              \
      e->err = std::string(#call) + ": " + cudaGetErrorString(_e);                        \
      return ORLG_ECUDA;                                                                  \
    }                                                                                     \
  } while (0)

template <typename T>
static int dev_alloc(orlg_engine* e, T** out, size_t count, bool zero = true) {
  void* ptr = nullptr;
  size_t bytes = count * sizeof(T);
  if (bytes == 0) bytes = 16;
  if (cudaMalloc(&ptr, bytes) != cudaSuccess) {
    e->err = "cudaMalloc failed for " + std::to_string(bytes) + " bytes";
    return ORLG_ENOMEM;
  }
  if (zero) cudaMemset(ptr, 0, bytes);
  e->allocs.push_back(ptr);
  e->hbm_bytes += bytes;
  *out = (T*)ptr;
  return ORLG_OK;
}

template <typename T>
static int dev_upload(orlg_engine* e, const T** out, const T* host, size_t count) {
  T* d = nullptr;
  int rc = dev_alloc(e, &d, count, false);
  if (rc) return rc;
  if (count && cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) {
    e->err = "cudaMemcpy H2D failed";
    return ORLG_ECUDA;
  }
  *out = d;
  return ORLG_OK;
}

extern "C" int orlg_abi_version(void) { return ORLG_ABI_VERSION; }
extern "C" const char* orlg_build_info(void) {
  return "orlg sm_100a engine: thread-per-env, TMA-staged tiles (BD in {32,64,128}) or env-major direct rows, W in {2,4,10}, 8-ary release heap";
}
extern "C" const char* orlg_last_error(const orlg_engine* e) {
  return e ? e->err.c_str() : g_create_error.c_str();
}

static int launch(orlg_engine* e, const StepIO& io, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  LaunchGeom g{e->BD, e->p.NP, e->device, e->smem_bytes, e->p.info_stats != 0, e->direct};
  cudaError_t rc;
  if (e->p.kind == ORLG_ENV_PHYRMSA) {
    switch (e->W) {
      case 2: rc = launch_phy<2>(g, e->p, io, st); break;
      case 4: rc = launch_phy<4>(g, e->p, io, st); break;
      default: rc = launch_phy<10>(g, e->p, io, st); break;
    }
  } else {
    switch (e->W) {
      case 2: rc = launch_w<2>(g, e->p, io, st); break;
      case 4: rc = launch_w<4>(g, e->p, io, st); break;
      default: rc = launch_w<10>(g, e->p, io, st); break;
    }
  }
  if (rc != cudaSuccess) {
    e->err = std::string("kernel launch: ") + cudaGetErrorString(rc);
    return ORLG_ECUDA;
  }
  e->launches++;
  return ORLG_OK;
}

static int fail_create(orlg_engine* e, int code, const std::string& msg) {
  g_create_error = msg;
  if (e) {
    for (void* a : e->allocs) cudaFree(a);
    delete e;
  }
  return code;
}

static int configure_gsnr(orlg_engine* e, const orlg_gsnr_params* g);

extern "C" int orlg_create(const orlg_config* c, orlg_engine** out) {
  if (!c || !out) return fail_create(nullptr, ORLG_EINVAL, "null argument");
  *out = nullptr;
  if (c->abi_version != ORLG_ABI_VERSION) return fail_create(nullptr, ORLG_EINVAL, "abi_version mismatch");
  if (c->env_kind < ORLG_ENV_RMSA || c->env_kind > ORLG_ENV_PHYRMSA)
    return fail_create(nullptr, ORLG_EINVAL, "env_kind not supported by this build");
  const bool phy = c->env_kind == ORLG_ENV_PHYRMSA;
  if (phy && (!c->phy_table_id || !c->phy_mod_level || !c->phy_gsnr || !c->phy_cut_links || !c->phy_cut_count ||
              c->phy_n_tables < 1 || c->phy_max_terms < 1))
    return fail_create(nullptr, ORLG_EINVAL, "PhyRMSA needs the QoT tables (table_id, mod_level, gsnr, cut terms)");
  if (c->num_envs < 1 || c->n_nodes < 2 || c->n_nodes > 255 || c->n_edges < 1 || c->n_edges > 255 ||
      c->k_paths < 1 || c->k_paths > 15 || c->max_hops < 1 || c->max_hops > 28 || c->num_slots < 2 ||
      c->num_slots > 320 || c->episode_length < 1)
    return fail_create(nullptr, ORLG_EINVAL, "config out of range (nodes/edges<=255, k<=15, hops<=28, slots<=320)");
  if (!c->path_links || !c->path_hops || !c->path_se || !c->path_length)
    return fail_create(nullptr, ORLG_EINVAL, "missing topology tables");
  if (!(c->mean_iat > 0) || !(c->mean_holding > 0) || !(c->channel_width > 0))
    return fail_create(nullptr, ORLG_EINVAL, "mean_iat, mean_holding, channel_width must be > 0");
  const int cores = c->env_kind == ORLG_ENV_RMCSA ? c->num_cores : 1;
  if (cores < 1 || cores > 32) return fail_create(nullptr, ORLG_EINVAL, "num_cores out of range");
  if (c->env_kind == ORLG_ENV_RMCSA && (!c->mod_reach || !c->mod_snr_coeff || !c->mod_se || c->n_mods < 1))
    return fail_create(nullptr, ORLG_EINVAL, "RMCSA needs modulation tables");
  if (c->request_source == ORLG_SRC_PHILOX &&
      (!c->src_cum || !c->dst_cum || (c->n_bit_rates > 0 && (!c->bit_rates || !c->bit_rate_cum))))
    return fail_create(nullptr, ORLG_EINVAL, "philox source needs src_cum/dst_cum/bit-rate tables");
  const size_t npairs = (size_t)c->n_nodes * c->n_nodes, npaths = npairs * c->k_paths;
  if (c->path_links_by_id &&
      memcmp(c->path_links_by_id, c->path_links, npaths * c->max_hops * sizeof(int16_t)) != 0)
    return fail_create(nullptr, ORLG_EINVAL, "edge 'id' != 'index' (SURVEY F9) is not supported");

  if (cudaSetDevice(c->device) != cudaSuccess) return fail_create(nullptr, ORLG_ECUDA, "cudaSetDevice failed");
  orlg_engine* e = new orlg_engine();
  e->cfg = *c;
  e->device = c->device;
  Params& p = e->p;
  e->W = c->num_slots <= 64 ? 2 : c->num_slots <= 128 ? 4 : 10;
  p.kind = c->env_kind;
  p.num_envs = c->num_envs;
  p.n_nodes = c->n_nodes;
  p.E = c->n_edges;
  p.k = c->k_paths;
  p.S = c->num_slots;
  p.cores = cores;
  p.j = c->env_kind == ORLG_ENV_DEEPRMSA ? (c->j < 1 ? 1 : c->j) : 1;
  p.n_mods = c->n_mods;
  p.planes = cores * p.E * e->W;
  p.episode_length = c->episode_length;
  p.allow_rejection = c->allow_rejection;
  p.auto_reset = c->auto_reset;
  p.request_source = c->request_source;
  p.n_bit_rates = c->n_bit_rates;
  p.br_lo = c->bit_rate_lo;
  p.br_hi = c->bit_rate_hi;
  p.action_dim = c->env_kind == ORLG_ENV_RMSA ? 2 : c->env_kind == ORLG_ENV_DEEPRMSA ? 1 : phy ? PHY_ADIM : 4;
  p.cw = c->channel_width;
  {
    double x2 = 2.0 * c->channel_width;
    p.cw_x2 = (x2 >= 1.0 && x2 <= 1024.0 && x2 == std::floor(x2)) ? (int)x2 : 0;
    for (int se = 0; se < 16; se++)
      p.slot_magic[se] = (p.cw_x2 > 0 && se > 0) ? (unsigned)(4294967296ull / (unsigned long long)(p.cw_x2 * se)) + 1u : 0u;
  }
  p.mean_iat = c->mean_iat;
  p.mean_holding = c->mean_holding;
  p.seed = c->seed;
  p.genv0 = c->global_env_offset;

  // tile geometry: as many envs per CTA as keep >= 3 CTAs resident per SM.  Envs whose occupancy
  // is too large for that (fewer than ~8 warps of tiles per SM) are not staged at all: env-major
  // occupancy, rows read in place (DIRECT kernels).  ORLG_DIRECT=0/1 overrides the choice.
  const size_t bytes_per_env = (size_t)p.planes * 4;
  int bd = 128;
  while (bd > 32 && bytes_per_env * bd > 72 * 1024) bd >>= 1;
  if (phy) bd = 32;  // PhyRMSA kernels are instantiated for 32-env tiles only
  if (const char* s = getenv("ORLG_BD"); s && !phy) {
    int v = atoi(s);
    if (v == 32 || v == 64 || v == 128) bd = v;
  }
  // measured on B200 (65 536 envs): staged wins at 352 B/env (C1/C2) and for PhyRMSA, whose metrics read
  // every row anyway; direct wins from 896 B/env (C5 +28 %/+64 %) to 4 760 B/env (C3, 4.2x)
  bool direct = !phy && bytes_per_env * 256 > 200 * 1024;
  if (const char* s = getenv("ORLG_DIRECT"); s && *s) direct = atoi(s) != 0;
  if (c->info_stats) {  // the info-dict kernels exist for staged tiles of 128 envs and for direct rows
    bd = 128;
    if (bytes_per_env * bd > 100 * 1024) direct = true;
  }
  if (!direct && bytes_per_env * bd > 220 * 1024) direct = true;
  if (c->qot_gate) direct = true;  // qot_kernel (one warp per env) reads the env's rows in place
  e->direct = direct;
  e->BD = direct ? DIRECT_CTA : bd;
  e->layout_bd = direct ? 1 : bd;
  e->smem_bytes = direct ? 0 : bytes_per_env * bd;
  p.NP = (c->num_envs + e->BD - 1) / e->BD * e->BD;

  // path records
  p.rec_u4 = c->max_hops <= 12 ? 1 : 2;
  std::vector<unsigned char> rec(npaths * 16 * p.rec_u4, 0);
  int max_se = 1;
  for (size_t i = 0; i < npaths; i++) {
    unsigned char* r = rec.data() + i * 16 * p.rec_u4;
    int h = c->path_hops[i];
    if (h < 0 || h > c->max_hops) return fail_create(e, ORLG_EINVAL, "path_hops out of range");
    r[0] = (unsigned char)h;
    r[1] = (unsigned char)c->path_se[i];
    r[2] = 255;
    if (h > 0 && c->path_se[i] < 1) return fail_create(e, ORLG_EINVAL, "path_se must be >= 1");
    if (h > 0 && c->path_se[i] > max_se) max_se = c->path_se[i];
    if (c->n_mods > 0 && c->mod_se && c->mod_max_length && h > 0) {
      // get_best_modulation_format utils.py:105-117: highest SE first (stable), first that reaches
      int best = -1;
      std::vector<int> order(c->n_mods);
      for (int m = 0; m < c->n_mods; m++) order[m] = m;
      for (int a = 1; a < c->n_mods; a++)  // stable insertion sort by SE descending
        for (int b = a; b > 0 && c->mod_se[order[b]] > c->mod_se[order[b - 1]]; b--) std::swap(order[b], order[b - 1]);
      for (int m = 0; m < c->n_mods && best < 0; m++)
        if (c->path_length[i] <= c->mod_max_length[order[m]]) best = order[m];
      if (best >= 0) r[2] = (unsigned char)best;
    }
    for (int k = 0; k < h; k++) {
      int l = c->path_links[i * c->max_hops + k];
      if (l < 0 || l >= c->n_edges) return fail_create(e, ORLG_EINVAL, "path_links out of range");
      r[4 + k] = (unsigned char)l;
    }
  }
  if (c->mod_se)
    for (int m = 0; m < c->n_mods; m++) max_se = c->mod_se[m] > max_se ? c->mod_se[m] : max_se;

  // release-queue capacity: every live service holds >= min_slots slots on >= 1 link
  int min_br = c->bit_rate_lo;
  if (c->n_bit_rates > 0 && c->bit_rates) {
    min_br = c->bit_rates[0];
    for (int i = 1; i < c->n_bit_rates; i++) min_br = c->bit_rates[i] < min_br ? c->bit_rates[i] : min_br;
  }
  if (c->request_source == ORLG_SRC_REPLAY || min_br < 1) min_br = 1;
  int min_slots = (int)std::ceil((double)min_br / ((double)max_se * c->channel_width)) + 1;
  if (phy) min_slots = 1;  // a service holds >= 1 channel on >= 1 link
  long long bound = (long long)cores * p.E * (p.S / min_slots);
  const long long lightpaths = bound;  // PhyRMSA: occupied (link, channel) cells
  if (phy && c->phy_grooming) bound *= PHY_LEVELS;  // a level-m channel can carry m groomed 100G services, one entry each
  // Default capacity: the offered load is Poisson (M/G/inf): mean A = holding / inter-arrival services in the
  // system if nothing were blocked, so A + 8 sqrt(A) + 32 entries are never exceeded in practice (8 sigma), and
  // never more than the provable bound.  Overflow raises the sticky fault bit (ORLG_C_FAULT) and the request is
  // refused; it never corrupts state silently.
  {
    const double A = c->mean_holding / c->mean_iat;
    double est = A + 8.0 * std::sqrt(A) + 32.0;
    if (c->request_source == ORLG_SRC_PHILOX && est < (double)bound) bound = (long long)std::ceil(est);
    // PhyRMSA addresses a service's row with 16 bits
    if (phy && bound > 0xFFFE) bound = est < 65534.0 ? (long long)std::ceil(est) : 0xFFFE;
    // engines with the QoT gate address a service's row with 14 bits: sized by load for every request source
    if (c->qot_gate && bound > 0x3FFF) bound = est < 16383.0 ? (long long)std::ceil(est) : 0x3FFF;
  }
  p.heap_cap = c->heap_capacity > 0 ? c->heap_capacity : (int)(bound < 16 ? 16 : bound);
  p.heap_stride = (p.heap_cap + HEAP_PAD + HEAP_TAIL + 7) & ~7;  // 128 B multiples: child groups stay line aligned

  int rc = 0;
  {  // the tables every step reads (path records, request distributions) share one contiguous
     // allocation (a few L2 lines instead of five scattered ones).  Warming L1 with them at kernel
     // start was measured and changed nothing: the early stalls are the cold per-env scalar loads.
    auto pad = [](size_t b) { return (b + 127) & ~(size_t)127; };
    const size_t o_rec = 0, b_rec = npaths * 16 * p.rec_u4;
    const bool have_br = c->bit_rates && c->n_bit_rates > 0;
    const size_t o_src = pad(o_rec + b_rec), b_src = c->src_cum ? (size_t)c->n_nodes * 8 : 0;
    const size_t o_brc = o_src + b_src, b_brc = have_br ? (size_t)c->n_bit_rates * 8 : 0;
    const size_t o_brv = o_brc + b_brc, b_brv = have_br ? (size_t)c->n_bit_rates * 4 : 0;
    const size_t o_dst = pad(o_brv + b_brv), b_dst = c->dst_cum ? npairs * 8 : 0;
    const size_t o_gd = pad(o_dst + b_dst), b_gd = (b_src && b_dst) ? (size_t)(2 + c->n_nodes) * 64 : 0;
    std::vector<unsigned char> blob(pad(o_gd + b_gd), 0);
    if (b_gd) {  // guide tables of bisect_right_guided: insertion point of the smallest value of each of 64 buckets
      auto fill = [&](unsigned char* g, const double* cum, int n) {
        if (n > 255) n = 255;
        for (int b = 0; b < 64; b++) {
          const double x = ((double)((unsigned long long)b << 26) * (1.0 / 4294967296.0)) * cum[n - 1];
          int cnt = 0;
          for (int i = 0; i < n - 1; i++) cnt += (x >= cum[i]) ? 1 : 0;
          g[b] = (unsigned char)cnt;
        }
      };
      unsigned char* g = blob.data() + o_gd;
      fill(g, c->src_cum, c->n_nodes);
      if (have_br) fill(g + 64, c->bit_rate_cum, c->n_bit_rates);
      for (int s0 = 0; s0 < c->n_nodes; s0++) fill(g + 64 * (2 + s0), c->dst_cum + (size_t)s0 * c->n_nodes, c->n_nodes);
    }
    memcpy(blob.data() + o_rec, rec.data(), b_rec);
    if (b_src) memcpy(blob.data() + o_src, c->src_cum, b_src);
    if (b_brc) memcpy(blob.data() + o_brc, c->bit_rate_cum, b_brc);
    if (b_brv) memcpy(blob.data() + o_brv, c->bit_rates, b_brv);
    if (b_dst) memcpy(blob.data() + o_dst, c->dst_cum, b_dst);
    const unsigned char* d = nullptr;
    if ((rc = dev_upload(e, &d, blob.data(), blob.size()))) return fail_create(e, rc, e->err);
    p.hot_tab = d;
    p.hot_tab_lines = (int)(blob.size() / 128);
    p.path_rec = reinterpret_cast<const uint4*>(d + o_rec);
    if (b_src) p.src_cum = reinterpret_cast<const double*>(d + o_src);
    if (b_brc) p.br_cum = reinterpret_cast<const double*>(d + o_brc);
    if (b_brv) p.bit_rates = reinterpret_cast<const int*>(d + o_brv);
    if (b_dst) p.dst_cum = reinterpret_cast<const double*>(d + o_dst);
    if (b_gd) p.guide = d + o_gd;
  }
  if ((rc = dev_upload(e, &p.path_length, c->path_length, npaths))) return fail_create(e, rc, e->err);
  if (c->mod_se && (rc = dev_upload(e, &p.mod_se, c->mod_se, c->n_mods))) return fail_create(e, rc, e->err);
  if (c->mod_reach && (rc = dev_upload(e, &p.mod_reach, c->mod_reach, c->n_mods))) return fail_create(e, rc, e->err);
  if (c->mod_snr_coeff && (rc = dev_upload(e, &p.mod_coeff, c->mod_snr_coeff, c->n_mods))) return fail_create(e, rc, e->err);

  {  // observation normalisation tables, f64 in the reference's operation order
    std::vector<double> pos(p.S + 1), sl(p.S + 2);
    const double dS = (double)p.S;
    for (int x = 0; x <= p.S; x++) pos[x] = 2.0 * ((double)x - 0.5 * dS) / dS;
    for (int n = 0; n <= p.S + 1; n++) sl[n] = ((double)n - 5.5) / 3.5;
    if ((rc = dev_upload(e, &p.obs_pos, pos.data(), pos.size()))) return fail_create(e, rc, e->err);
    if ((rc = dev_upload(e, &p.obs_slots, sl.data(), sl.size()))) return fail_create(e, rc, e->err);
  }
  if (phy) {  // QoT tables, per-level channel masks, r_cut term lists
    const size_t T = c->phy_n_tables, S = p.S, K = p.k, Wd = e->W;
    for (size_t i = 0; i < npaths; i++)
      if (c->phy_cut_count[i] < 0 || c->phy_cut_count[i] > 63 || c->phy_cut_count[i] > c->phy_max_terms)
        return fail_create(e, ORLG_EINVAL, "more than 63 r_cut terms on a path");
    std::vector<uint32_t> masks(T * K * PHY_LEVELS * Wd, 0u);
    for (size_t t = 0; t < T; t++)
      for (size_t ch = 0; ch < S; ch++)
        for (size_t ip = 0; ip < K; ip++) {
          int m = c->phy_mod_level[(t * S + ch) * K + ip];
          if (m < 1 || m > PHY_LEVELS) return fail_create(e, ORLG_EINVAL, "modulation levels must be in 1..8");
          masks[((t * K + ip) * PHY_LEVELS + (m - 1)) * Wd + (ch >> 5)] |= 1u << (ch & 31);
        }
    for (size_t i = 0; i < npairs; i++)
      if (c->phy_table_id[i] >= c->phy_n_tables) return fail_create(e, ORLG_EINVAL, "phy_table_id out of range");
    {  // an action carries at most 8 channels; the reference's greedy take has no cap (phy_rmsa_env.py:1425-1432), so a
       // bit-rate table that could need more is refused here instead of diverging silently
      int min_level = PHY_LEVELS, max_br = c->bit_rate_hi;
      for (size_t i = 0; i < T * S * K; i++) min_level = c->phy_mod_level[i] < min_level ? c->phy_mod_level[i] : min_level;
      if (c->n_bit_rates > 0 && c->bit_rates) {
        max_br = 0;
        for (int i = 0; i < c->n_bit_rates; i++) max_br = c->bit_rates[i] > max_br ? c->bit_rates[i] : max_br;
      }
      if (max_br > (PHY_ADIM - 1) * 100 * min_level)
        return fail_create(e, ORLG_EINVAL, "PhyRMSA: the largest bit rate could need more than 8 channels of the lowest "
                                           "modulation level in the table (action format: 8 channels)");
    }
    if ((rc = dev_upload(e, &p.phy_table_id, c->phy_table_id, npairs)) ||
        (rc = dev_upload(e, &p.phy_mod, c->phy_mod_level, T * S * K)) ||
        (rc = dev_upload(e, &p.phy_gsnr, c->phy_gsnr, T * S * K)) ||
        (rc = dev_upload(e, &p.phy_level_mask, masks.data(), masks.size())) ||
        (rc = dev_upload(e, &p.phy_cut_links, c->phy_cut_links, npaths * (size_t)c->phy_max_terms)) ||
        (rc = dev_upload(e, &p.phy_cut_count, c->phy_cut_count, npaths)))
      return fail_create(e, rc, e->err);
    p.phy_max_terms = c->phy_max_terms;
    p.phy_mask_w = e->W;
  }
  const size_t NP = p.NP;
  if (phy && ((rc = dev_alloc(e, &p.phy_f, 2 * NP)) || (rc = dev_alloc(e, &p.phy_i, 4 * NP))))
    return fail_create(e, rc, e->err);
  if (phy) {  // service table: one row per running service (= release-queue capacity)
    if (p.heap_cap > 0xFFFE) return fail_create(e, ORLG_EINVAL, "PhyRMSA: release-queue capacity above 65534 services");
    p.phy_svc_cap = p.heap_cap;
    if ((rc = dev_alloc(e, &p.phy_svc, NP * (size_t)p.phy_svc_cap * 3)) ||
        (rc = dev_alloc(e, &p.phy_free, NP * (size_t)p.phy_svc_cap)) || (rc = dev_alloc(e, &p.phy_free_n, NP)))
      return fail_create(e, rc, e->err);
    if (c->phy_defrag_period > 0) {
      if (!c->phy_grooming) return fail_create(e, ORLG_EINVAL, "phy_defrag_period needs phy_grooming != 0 (channel_state)");
      if (c->phy_number_moves < 0 || c->phy_metric < 0 || c->phy_metric > 1 || (c->phy_metric == 1 && p.E > 64))
        return fail_create(e, ORLG_EINVAL, "defragmentation: number_moves >= 0, metric 0 (cut) or 1 (rss, <= 64 links)");
      p.phy_defrag_period = c->phy_defrag_period;
      p.phy_number_moves = c->phy_number_moves;
      p.phy_metric = c->phy_metric;
      long long rcap = 2ll * p.phy_svc_cap + 4ll * c->phy_number_moves + 64;
      p.phy_run_cap = (int)(rcap > 65535 ? 65535 : rcap);
      if (p.phy_run_cap < p.phy_svc_cap + 2 * c->phy_number_moves + 16)
        return fail_create(e, ORLG_EINVAL, "defragmentation: too many services for the order list");
      // candidates: every whole channel of every running service, at most the occupied (link, channel) cells
      long long ccap = (long long)p.phy_svc_cap * (PHY_ADIM - 1);
      if (ccap > lightpaths) ccap = lightpaths;
      p.phy_cand_cap = (int)ccap;
      if ((rc = dev_alloc(e, &p.phy_run, NP * (size_t)p.phy_run_cap)) || (rc = dev_alloc(e, &p.phy_run_len, NP)) ||
          (rc = dev_alloc(e, &p.phy_moves, 3 * NP)) || (rc = dev_alloc(e, &p.phy_cand, NP * (size_t)p.phy_cand_cap)) ||
          (rc = dev_alloc(e, &p.phy_cand_id, NP * (size_t)p.phy_cand_cap)))
        return fail_create(e, rc, e->err);
    }
  }
  if (c->info_stats) {
    if (c->env_kind != ORLG_ENV_RMSA && c->env_kind != ORLG_ENV_DEEPRMSA)
      return fail_create(e, ORLG_EINVAL, "info_stats is built for RMSA-v0 / DeepRMSA-v0");
    if (p.n_bit_rates <= 0 || p.n_bit_rates > 55) return fail_create(e, ORLG_EINVAL, "info_stats needs 1..55 discrete bit rates");
    p.info_stats = 1;
    if ((rc = dev_alloc(e, &p.lnk_stat, (size_t)3 * p.E * NP)) ||
        (rc = dev_alloc(e, &p.br_hist, (size_t)2 * p.n_bit_rates * NP)) || (rc = dev_alloc(e, &p.slot_hops, NP)))
      return fail_create(e, rc, e->err);
  }
  if (phy && c->phy_grooming) {
    if (c->phy_grooming < 0 || c->phy_grooming > 2) return fail_create(e, ORLG_EINVAL, "phy_grooming must be 0, 1 or 2");
    if (npairs > 65536 || p.k > 8 || p.S > 4096) return fail_create(e, ORLG_EINVAL, "channel_state keys: pairs <= 65536, k <= 8, channels <= 4096");
    p.phy_grooming = c->phy_grooming;
    // a partially used channel belongs to one live lightpath; + slack for tombstones between squeezes
    long long cap = lightpaths + 64 + lightpaths / 2;
    p.phy_cs_cap = (int)(cap > 65535 ? 65535 : cap);
    if ((rc = dev_alloc(e, &p.phy_cs, (size_t)p.phy_cs_cap * NP)) || (rc = dev_alloc(e, &p.phy_cs_len, NP)))
      return fail_create(e, rc, e->err);
  }
  if ((rc = dev_alloc(e, &p.occ, (size_t)p.planes * NP)) || (rc = dev_alloc(e, &p.now, NP)) ||
      (rc = dev_alloc(e, &p.holding, NP)) || (rc = dev_alloc(e, &p.meta, NP)) ||
      (rc = dev_alloc(e, &p.next_rel, NP)) || (rc = dev_alloc(e, &p.heap_len, NP)) ||
      (rc = dev_alloc(e, &p.heap, NP * p.heap_stride)) ||
      (rc = dev_alloc(e, &p.counters, (size_t)CNT_N * NP)))
    return fail_create(e, rc, e->err);

  if (c->qot_gate) {
    if (c->env_kind != ORLG_ENV_RMSA && c->env_kind != ORLG_ENV_RMCSA)
      return fail_create(e, ORLG_EINVAL, "qot_gate is defined for RMSA-v0 and RMCSA-v0");
    if (!c->qot_gsnr || p.E > 128 || c->n_mods > 16 || c->qot_n_se < 0 || c->qot_n_se > 8 || c->info_stats)
      return fail_create(e, ORLG_EINVAL, "qot_gate needs qot_gsnr, <= 128 links, <= 16 modulations, no info_stats");
    if (p.heap_cap > 0x3FFF) return fail_create(e, ORLG_EINVAL, "qot_gate: release-queue capacity above 16383 rows");
    if ((rc = configure_gsnr(e, c->qot_gsnr))) return fail_create(e, rc, e->err);
    if (!e->gp.xci_tab) return fail_create(e, ORLG_EINVAL, "qot_gate: the cross-channel table does not fit (slots too many)");
    p.qot = 1;
    p.qot_adaptive = c->qot_adaptive ? 1 : 0;
    p.qot_n_se = c->qot_n_se;
    for (int i = 0; i < 8; i++) p.qot_se[i] = c->qot_se[i];
    for (int i = 0; i < 16; i++) {
      p.qot_thr[i] = c->qot_thr_db[i];
      p.qot_thr_lin[i] = std::pow(10.0, -c->qot_thr_db[i] / 10.0);  // +inf dB -> 0 (never), -inf dB -> inf (always)
    }
    for (int i = 0; i < c->qot_n_se; i++)
      if (c->qot_se[i] < 1 || c->qot_se[i] > 15) return fail_create(e, ORLG_EINVAL, "qot_se out of range");
    if (p.qot_adaptive && c->env_kind == ORLG_ENV_RMSA && c->qot_n_se < 1)
      return fail_create(e, ORLG_EINVAL, "qot_adaptive needs qot_se");
    for (int m = 0; m < 16; m++) p.qot_mod_order[m] = m;
    for (int a = 1; a < c->n_mods; a++)  // stable, by descending spectral efficiency (utils.py:105-117)
      for (int b = a; b > 0 && c->mod_se[p.qot_mod_order[b]] > c->mod_se[p.qot_mod_order[b - 1]]; b--)
        std::swap(p.qot_mod_order[b], p.qot_mod_order[b - 1]);
    {  // window of the release queue qot_kernel stages in shared memory: the offered load + 1 sigma covers the
       // queue nearly always (blocking keeps it below the offered load); a longer queue costs a second window
      const double A = c->mean_holding / c->mean_iat;
      int cap = (int)std::ceil(A + std::sqrt(A));
      if (const char* s = getenv("ORLG_QOT_CAP"); s && *s) cap = atoi(s);
      e->qot_cap = cap <= 64 ? 64 : cap <= 128 ? 128 : 256;
    }
    {  // the link-independent factors of P_ASE / P and P_NLI / P (calculate_osnr.py:45-52) per slot count
      const double pi = 3.141592653589793, beta2 = 21.3e-27, gamma = 1.3e-3, h_planck = 6.626e-34;
      const orlg_gsnr_params* gq = c->qot_gsnr;
      std::vector<double> tab(2 * ((size_t)p.S + 1), 0.0);
      for (int n = 1; n <= p.S; n++) {
        const double B = (double)n * gq->slot_hz, pr = gq->launch_power_w / B;
        tab[2 * n] = B * h_planck * gq->noise_figure_lin / gq->launch_power_w;
        tab[2 * n + 1] = pr * pr * pr * (8.0 / (27.0 * pi * beta2)) * gamma * gamma * B / gq->launch_power_w;
      }
      const double* d = nullptr;
      if ((rc = dev_upload(e, &d, tab.data(), tab.size()))) return fail_create(e, rc, e->err);
      p.qot_chan = reinterpret_cast<const double2*>(d);
    }
    p.qot_hops = c->max_hops;
    p.qot_slot_cap = p.heap_cap;
    if ((rc = dev_alloc(e, &p.qot_sums, NP * (size_t)p.qot_slot_cap * p.qot_hops)) ||
        (rc = dev_alloc(e, &p.qot_free, NP * (size_t)p.qot_slot_cap)) || (rc = dev_alloc(e, &p.qot_free_n, NP)) ||
        (rc = dev_alloc(e, &p.qot_rel, NP * (size_t)QOT_REL_CAP)) || (rc = dev_alloc(e, &p.qot_rel_n, NP)) || (rc = dev_alloc(e, &p.qot_add, NP)) || (rc = dev_alloc(e, &p.qot_gp, NP * 2)) ||
        (rc = dev_alloc(e, &p.qot_verdict, NP * 2)) || (rc = dev_alloc(e, &p.qot_g, NP * 2)))
      return fail_create(e, rc, e->err);
  }

  e->obs_kind = c->obs_kind;
  if (c->obs_kind == ORLG_OBS_F64 || c->obs_kind == ORLG_OBS_F32) {
    e->obs_dim = 1 + 2 * p.n_nodes + (2 * p.j + 3) * p.k;
    e->obs_elem = c->obs_kind == ORLG_OBS_F64 ? 8 : 4;
  } else if (c->obs_kind == ORLG_OBS_MATRIX_U8) {
    e->obs_dim = 2 * p.n_nodes + cores * p.E * p.S;
    e->obs_elem = 1;
  } else if (c->obs_kind == ORLG_OBS_PHY_INFO || c->obs_kind == ORLG_OBS_PHY_CUTS) {
    if (!phy) return fail_create(e, ORLG_EINVAL, "ORLG_OBS_PHY_* needs env_kind PhyRMSA");
    e->obs_dim = c->obs_kind == ORLG_OBS_PHY_INFO ? 2 : 1;
    e->obs_elem = 8;
  }
  if (phy && (c->obs_kind == ORLG_OBS_F64 || c->obs_kind == ORLG_OBS_F32))
    return fail_create(e, ORLG_EINVAL, "the DeepRMSA observation is not defined for PhyRMSA");
  *out = e;
  return ORLG_OK;
}

extern "C" int orlg_destroy(orlg_engine* e) {
  if (!e) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  for (void* a : e->allocs) cudaFree(a);
  if (e->host_ev) cudaEventDestroy(e->host_ev);
  delete e;
  return ORLG_OK;
}

extern "C" int orlg_action_dim(const orlg_engine* e) { return e ? e->p.action_dim : ORLG_EINVAL; }
extern "C" int orlg_obs_dim(const orlg_engine* e) { return e ? e->obs_dim : ORLG_EINVAL; }
extern "C" int64_t orlg_state_words(const orlg_engine* e) { return e ? e->p.planes : ORLG_EINVAL; }
extern "C" int64_t orlg_hbm_bytes(const orlg_engine* e) { return e ? (int64_t)e->hbm_bytes : ORLG_EINVAL; }
extern "C" int64_t orlg_launch_count(const orlg_engine* e) { return e ? e->launches : ORLG_EINVAL; }

extern "C" int orlg_set_obs_layout(orlg_engine* e, int64_t env_stride, int64_t feat_stride) {
  if (!e || env_stride < 1 || feat_stride < 1) return ORLG_EINVAL;
  e->obs_es = env_stride;
  e->obs_fs = feat_stride;
  return ORLG_OK;
}

extern "C" int orlg_load_trace(orlg_engine* e, const orlg_request* dev_trace, int64_t n) {
  if (!e || !dev_trace || n < 1) return ORLG_EINVAL;
  e->p.trace = dev_trace;
  e->p.trace_len = n;
  e->qot_proposal_valid = false;
  return ORLG_OK;
}

static int check_source(orlg_engine* e) {
  if (e->p.request_source == ORLG_SRC_REPLAY && !e->p.trace) {
    e->err = "replay source without a trace (call orlg_load_trace)";
    return ORLG_ESTATE;
  }
  return ORLG_OK;
}

static bool heuristic_ok(const orlg_engine* e, int hid) {
  if (!e) return false;
  if (e->p.kind == ORLG_ENV_PHYRMSA)
    return (hid >= ORLG_H_PHY_SAP_FF && hid <= ORLG_H_PHY_BM_FA_CUT) || hid == ORLG_H_PHY_FA_FF ||
           ((hid == ORLG_H_PHY_BM_FA_RSS || hid == ORLG_H_PHY_FA_FF_RSS) && e->p.E <= 64);  // columns are gathered into 64 bits
  if (hid == ORLG_H_SAP_GN_FF) return e->p.qot != 0;
  return hid >= ORLG_H_SP_FF && hid <= ORLG_H_SAP_BM_FC_FF;
}

static void record_mid(orlg_engine* e, void* stream) {
  if (e->mid_ev) cudaEventRecordWithFlags(e->mid_ev, (cudaStream_t)stream, e->mid_external ? cudaEventRecordExternal : cudaEventRecordDefault);
}

// qot_kernel launch (engines with the QoT gate): one warp per env
static int launch_qot(orlg_engine* e, int mode, const int32_t* actions, int32_t* actions_out, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  // the verdict launch that follows a heuristic-13 launch on the same state may take its verdict (qot.cuh); any
  // other launch of this function, a reset or a new trace ends that
  QotIO q{mode, actions, actions_out, (mode == QOT_VERDICT && e->qot_proposal_valid) ? 1 : 0};
  e->qot_proposal_valid = (mode == QOT_HEURISTIC);
  const int grid = (e->p.num_envs + QOT_WARPS - 1) / QOT_WARPS;
  const size_t dsm = (size_t)e->p.E * sizeof(double4);  // the per-link factors staged by every CTA
#define ORLG_QOT_LAUNCH_N(WW, RM, MD, NR)                                                                          \
  do {                                                                                                             \
    if (e->qot_cap <= 64) qot_kernel<WW, 64, RM, MD, NR><<<grid, QOT_WARPS * 32, dsm, st>>>(e->p, e->gp, q);       \
    else if (e->qot_cap <= 128) qot_kernel<WW, 128, RM, MD, NR><<<grid, QOT_WARPS * 32, dsm, st>>>(e->p, e->gp, q); \
    else qot_kernel<WW, 256, RM, MD, NR><<<grid, QOT_WARPS * 32, dsm, st>>>(e->p, e->gp, q);                        \
  } while (0)
#define ORLG_QOT_LAUNCH_M(WW, RM, MD)                         \
  do {                                                        \
    if (e->p.E <= 64) ORLG_QOT_LAUNCH_N(WW, RM, MD, true);    \
    else ORLG_QOT_LAUNCH_N(WW, RM, MD, false);                \
  } while (0)
#define ORLG_QOT_LAUNCH_K(WW, RM)                                          \
  do {                                                                     \
    if (mode == QOT_VERDICT) ORLG_QOT_LAUNCH_M(WW, RM, QOT_VERDICT);       \
    else if (mode == QOT_HEUR_VERDICT) ORLG_QOT_LAUNCH_M(WW, RM, QOT_HEUR_VERDICT); \
    else ORLG_QOT_LAUNCH_M(WW, RM, QOT_HEURISTIC);                         \
  } while (0)
#define ORLG_QOT_LAUNCH(WW)                                   \
  do {                                                        \
    if (e->p.kind == ORLG_ENV_RMCSA) ORLG_QOT_LAUNCH_K(WW, true); \
    else ORLG_QOT_LAUNCH_K(WW, false);                        \
  } while (0)
  switch (e->W) {
    case 2: ORLG_QOT_LAUNCH(2); break;
    case 4: ORLG_QOT_LAUNCH(4); break;
    default: ORLG_QOT_LAUNCH(10); break;
  }
#undef ORLG_QOT_LAUNCH
#undef ORLG_QOT_LAUNCH_K
#undef ORLG_QOT_LAUNCH_M
#undef ORLG_QOT_LAUNCH_N
  cudaError_t rc = cudaGetLastError();
  if (rc != cudaSuccess) {
    e->err = std::string("qot_kernel launch: ") + cudaGetErrorString(rc);
    return ORLG_ECUDA;
  }
  e->launches++;
  return ORLG_OK;
}

static StepIO make_io(orlg_engine* e, int mode) {
  StepIO io{};
  io.mode = mode;
  io.obs_kind = e->obs_kind;
  io.obs_es = e->obs_es ? e->obs_es : e->obs_dim;
  io.obs_fs = e->obs_es ? e->obs_fs : 1;
  io.nsteps = 1;
  io.snap = e->snap;
  return io;
}

extern "C" int orlg_reset(orlg_engine* e, int full, void* stream) {
  if (!e) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  cudaSetDevice(e->device);
  StepIO io = make_io(e, full ? MODE_RESET_FULL : MODE_RESET_EPISODE);
  io.obs = nullptr;
  e->qot_proposal_valid = false;
  return launch(e, io, stream);
}

extern "C" int orlg_step(orlg_engine* e, const int32_t* dev_actions, int32_t* dev_reward, uint8_t* dev_done,
                         void* dev_obs, int32_t* dev_alloc, void* stream) {
  if (!e || !dev_actions) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  cudaSetDevice(e->device);
  if (e->p.qot) {  // the gate decides (verdict), the step kernel applies
    if (int rc = launch_qot(e, QOT_VERDICT, dev_actions, nullptr, stream)) return rc;
    record_mid(e, stream);
  }
  StepIO io = make_io(e, MODE_STEP);
  io.actions = dev_actions;
  io.reward = dev_reward;
  io.done = dev_done;
  io.obs = e->obs_kind ? dev_obs : nullptr;
  io.alloc = dev_alloc;
  io.hint_actions = e->hint_actions;
  io.hint_heuristic = e->hint_heuristic;
  io.info = e->info_out;
  return launch(e, io, stream);
}

extern "C" int orlg_step_auto(orlg_engine* e, int hid, int32_t* dev_actions_out, int32_t* dev_reward, uint8_t* dev_done,
                              void* dev_obs, int32_t* dev_alloc, void* stream) {
  if (!e || !heuristic_ok(e, hid)) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  cudaSetDevice(e->device);
  int rc;
  if (!dev_actions_out) {  // the action still has to live somewhere between the two launches
    if (!e->auto_actions && (rc = ::dev_alloc(e, &e->auto_actions, (size_t)e->p.num_envs * e->p.action_dim))) return rc;
    dev_actions_out = e->auto_actions;
  }
  if (e->p.qot && hid == ORLG_H_SAP_GN_FF) {
    if ((rc = launch_qot(e, QOT_HEUR_VERDICT, nullptr, dev_actions_out, stream))) return rc;
  } else {
    if ((rc = orlg_heuristic(e, hid, dev_actions_out, stream))) return rc;
    if (e->p.qot && (rc = launch_qot(e, QOT_VERDICT, dev_actions_out, nullptr, stream))) return rc;
  }
  record_mid(e, stream);
  StepIO io = make_io(e, MODE_STEP);
  io.actions = dev_actions_out;
  io.reward = dev_reward;
  io.done = dev_done;
  io.obs = e->obs_kind ? dev_obs : nullptr;
  io.alloc = dev_alloc;
  io.info = e->info_out;
  io.reward_sum = e->auto_reward_sum;
  return launch(e, io, stream);
}

extern "C" int orlg_observation(orlg_engine* e, int obs_kind, void* dev_obs, void* stream) {
  if (!e || !dev_obs || obs_kind < ORLG_OBS_F64 || obs_kind > ORLG_OBS_PHY_CUTS) return ORLG_EINVAL;
  const bool deep = obs_kind == ORLG_OBS_F64 || obs_kind == ORLG_OBS_F32;
  const bool phyobs = obs_kind == ORLG_OBS_PHY_INFO || obs_kind == ORLG_OBS_PHY_CUTS;
  if (deep && (e->p.kind == ORLG_ENV_RMCSA || e->p.kind == ORLG_ENV_PHYRMSA)) return ORLG_EINVAL;
  if (phyobs && e->p.kind != ORLG_ENV_PHYRMSA) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  StepIO io = make_io(e, MODE_OBS);
  io.obs_kind = obs_kind;
  io.obs = dev_obs;
  if (!e->obs_es)
    io.obs_es = obs_kind == ORLG_OBS_MATRIX_U8 ? 2 * e->p.n_nodes + e->p.cores * e->p.E * e->p.S
                : obs_kind == ORLG_OBS_PHY_INFO ? 2 : obs_kind == ORLG_OBS_PHY_CUTS ? 1
                : 1 + 2 * e->p.n_nodes + (2 * e->p.j + 3) * e->p.k;
  return launch(e, io, stream);
}

extern "C" int orlg_heuristic(orlg_engine* e, int hid, int32_t* dev_actions, void* stream) {
  if (!e || !dev_actions || !heuristic_ok(e, hid)) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  if (hid == ORLG_H_SAP_GN_FF) return launch_qot(e, QOT_HEURISTIC, nullptr, dev_actions, stream);
  StepIO io = make_io(e, MODE_HEURISTIC);
  io.heuristic = hid;
  io.actions_out = dev_actions;
  io.obs = nullptr;
  return launch(e, io, stream);
}

extern "C" int orlg_path_first_fit(orlg_engine* e, const int32_t* dev_paths, int32_t* dev_actions, void* stream) {
  if (!e || !dev_paths || !dev_actions) return ORLG_EINVAL;
  if (e->p.kind != ORLG_ENV_RMSA && e->p.kind != ORLG_ENV_RMCSA) {
    e->err = "orlg_path_first_fit is defined for RMSA-v0 (rmsa_env.py:974-1008) and RMCSA-v0 (rmcsa_env.py:950-982)";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  StepIO io = make_io(e, MODE_HEURISTIC);
  io.heuristic = ORLG_H_PATH_FF;
  io.actions = dev_paths;
  io.actions_out = dev_actions;
  io.obs = nullptr;
  return launch(e, io, stream);
}

extern "C" int orlg_set_next_action_hint(orlg_engine* e, int hid, int32_t* dev_next_actions) {
  if (!e) return ORLG_EINVAL;
  if (dev_next_actions && (!heuristic_ok(e, hid) || hid == ORLG_H_PATH_FF || hid == ORLG_H_SAP_GN_FF)) return ORLG_EINVAL;
  e->hint_actions = dev_next_actions;
  e->hint_heuristic = hid;
  return ORLG_OK;
}

extern "C" int orlg_rollout(orlg_engine* e, int hid, int nsteps, int64_t* dev_reward_sum, void* stream) {
  if (!e || nsteps < 1 || !heuristic_ok(e, hid)) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  cudaSetDevice(e->device);
  if (e->p.qot) {  // the gate is a launch of its own: nsteps x orlg_step_auto
    e->auto_reward_sum = (long long*)dev_reward_sum;
    int rc = ORLG_OK;
    for (int t = 0; t < nsteps && rc == ORLG_OK; t++)
      rc = orlg_step_auto(e, hid, nullptr, nullptr, nullptr, nullptr, nullptr, stream);
    e->auto_reward_sum = nullptr;
    return rc;
  }
  StepIO io = make_io(e, MODE_ROLLOUT);
  io.heuristic = hid;
  io.nsteps = nsteps;
  io.reward_sum = (long long*)dev_reward_sum;
  io.obs = nullptr;
  return launch(e, io, stream);
}

extern "C" int orlg_get_counters(orlg_engine* e, int64_t* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  counters_kernel<<<(e->p.num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(e->p, (long long*)dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_export_state(orlg_engine* e, uint8_t* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  export_state_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(e->p, e->W, e->layout_bd, dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_get_requests(orlg_engine* e, orlg_request* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  requests_kernel<<<(e->p.num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(e->p, dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

// Host-buffer step: H2D actions, step, D2H outputs, stream sync (bench.py "e2e").
extern "C" int orlg_step_host_async(orlg_engine* e, const int32_t* h_actions, int32_t* h_reward, uint8_t* h_done,
                                    void* h_obs, int32_t* h_alloc, void* stream) {
  if (!e || !h_actions) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = e->p.num_envs;
  int rc = 0;
  if (!e->d_actions) {
    if ((rc = dev_alloc(e, &e->d_actions, n * e->p.action_dim)) || (rc = dev_alloc(e, &e->d_reward, n)) ||
        (rc = dev_alloc(e, &e->d_done, n)) || (rc = dev_alloc(e, &e->d_alloc, n * 4)))
      return rc;
    if (e->obs_dim) {
      unsigned char* o = nullptr;
      if ((rc = dev_alloc(e, &o, n * e->obs_dim * e->obs_elem))) return rc;
      e->d_obs = o;
    }
  }
  CK(cudaMemcpyAsync(e->d_actions, h_actions, n * e->p.action_dim * sizeof(int), cudaMemcpyHostToDevice, st));
  if ((rc = orlg_step(e, e->d_actions, e->d_reward, e->d_done, e->d_obs, e->d_alloc, stream))) return rc;
  if (h_reward) CK(cudaMemcpyAsync(h_reward, e->d_reward, n * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (h_done) CK(cudaMemcpyAsync(h_done, e->d_done, n, cudaMemcpyDeviceToHost, st));
  if (h_alloc) CK(cudaMemcpyAsync(h_alloc, e->d_alloc, n * 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (h_obs && e->d_obs)
    CK(cudaMemcpyAsync(h_obs, e->d_obs, n * e->obs_dim * e->obs_elem, cudaMemcpyDeviceToHost, st));
  if (e->blocking_wait) {
    if (!e->host_ev) CK(cudaEventCreateWithFlags(&e->host_ev, cudaEventBlockingSync | cudaEventDisableTiming));
    CK(cudaEventRecord(e->host_ev, st));
  }
  return ORLG_OK;
}

extern "C" int orlg_step_host_wait(orlg_engine* e, void* stream) {
  if (!e) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  if (e->blocking_wait && e->host_ev)
    CK(cudaEventSynchronize(e->host_ev));  // the thread sleeps instead of spinning on the stream
  else
    CK(cudaStreamSynchronize((cudaStream_t)stream));
  return ORLG_OK;
}

extern "C" int orlg_set_host_wait(orlg_engine* e, int blocking) {
  if (!e) return ORLG_EINVAL;
  e->blocking_wait = blocking != 0;
  return ORLG_OK;
}

extern "C" int orlg_step_host(orlg_engine* e, const int32_t* h_actions, int32_t* h_reward, uint8_t* h_done,
                              void* h_obs, int32_t* h_alloc, void* stream) {
  if (int rc = orlg_step_host_async(e, h_actions, h_reward, h_done, h_obs, h_alloc, stream)) return rc;
  return orlg_step_host_wait(e, stream);
}

extern "C" int orlg_gsnr_configure(orlg_engine* e, const orlg_gsnr_params* g) {
  if (e && e->p.qot) {
    e->err = "the GN model of an engine with the QoT gate is configured by orlg_create (orlg_config.qot_gsnr)";
    return ORLG_ESTATE;
  }
  return configure_gsnr(e, g);
}
static int configure_gsnr(orlg_engine* e, const orlg_gsnr_params* g) {
  if (!e || !g || !g->span_count || !g->span_length_km) return ORLG_EINVAL;
  if (!(g->launch_power_w > 0) || !(g->attenuation_per_m > 0) || !(g->noise_figure_lin > 0) || !(g->slot_hz > 0) ||
      !(g->f0_hz > 0)) {
    e->err = "gsnr parameters must be positive";
    return ORLG_EINVAL;
  }
  for (int l = 0; l < e->p.E; l++)
    if (g->span_count[l] < 1 || !(g->span_length_km[l] > 0)) {
      e->err = "span_count >= 1 and span_length_km > 0 required for every link";
      return ORLG_EINVAL;
    }
  cudaSetDevice(e->device);
  int rc;
  if ((rc = dev_upload(e, &e->gp.span_count, g->span_count, (size_t)e->p.E))) return rc;
  if ((rc = dev_upload(e, &e->gp.span_km, g->span_length_km, (size_t)e->p.E))) return rc;
  e->gp.P = g->launch_power_w;
  e->gp.alpha = g->attenuation_per_m;
  e->gp.nf = g->noise_figure_lin;
  e->gp.f0 = g->f0_hz;
  e->gp.slot_hz = g->slot_hz;
  e->gp.path_mask = nullptr;
  e->gp.link_c = nullptr;
  e->gp.xci_tab = nullptr;
  e->gp.sci_tab = nullptr;
  if (e->p.E <= 128) {  // tables of the fast kernel: link sets of the paths, per-link span factors
    const int npaths = e->p.n_nodes * e->p.n_nodes * e->p.k;
    ulonglong2* pm = nullptr;
    double4* lc = nullptr;
    if ((rc = dev_alloc(e, &pm, (size_t)npaths, false))) return rc;
    if ((rc = dev_alloc(e, &lc, (size_t)e->p.E, false))) return rc;
    const int n = npaths > e->p.E ? npaths : e->p.E;
    gsnr_init_tables<<<(n + 127) / 128, 128>>>(e->p, e->gp, npaths, pm, lc);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    e->gp.path_mask = pm;
    e->gp.link_c = lc;
    e->gp.xci_tab = nullptr;
    e->gp.tab_S = e->p.S;
    const long long entries = (long long)(e->p.S + 1) * (4 * e->p.S + 1);
    const char* notab = getenv("ORLG_GSNR_NO_TABLE");  // A/B: evaluate the asinh pair instead of reading it
    if (entries * 16 <= (16ll << 20) && !(notab && *notab == '1')) {
      double2* tab = nullptr;
      if ((rc = dev_alloc(e, &tab, (size_t)entries, false))) return rc;
      double* sci = nullptr;
      if ((rc = dev_alloc(e, &sci, (size_t)e->p.S + 1, false))) return rc;
      gsnr_init_xci<<<(unsigned)((entries + 255) / 256), 256>>>(e->gp, e->p.S, tab, sci);
      CK(cudaGetLastError());
      CK(cudaDeviceSynchronize());
      e->gp.xci_tab = tab;
      e->gp.sci_tab = sci;
    }
  }
  e->gsnr_ready = true;
  return ORLG_OK;
}

template <bool FROM_ACTION>
static cudaError_t launch_gsnr(orlg_engine* e, const int32_t* dev_in, double* dev_out, cudaStream_t st) {
  const int grid = (e->p.num_envs + GS_WARPS - 1) / GS_WARPS;
  const char* force = getenv("ORLG_GSNR_GENERIC");  // A/B: the previous kernel
  if (e->gp.path_mask && !(force && *force == '1'))
    gsnr_kernel_fast<FROM_ACTION><<<grid, GS_WARPS * 32, 0, st>>>(e->p, e->gp, dev_in, dev_out);
  else
    gsnr_kernel_generic<FROM_ACTION><<<grid, GS_WARPS * 32, 0, st>>>(e->p, e->gp, dev_in, dev_out);
  return cudaGetLastError();
}

extern "C" int orlg_gsnr(orlg_engine* e, const int32_t* dev_query, double* dev_out, void* stream) {
  if (!e || !dev_query || !dev_out) return ORLG_EINVAL;
  if (!e->gsnr_ready) {
    e->err = "orlg_gsnr before orlg_gsnr_configure";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  CK(launch_gsnr<false>(e, dev_query, dev_out, (cudaStream_t)stream));
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_gsnr_for_actions(orlg_engine* e, const int32_t* dev_actions, double* dev_out, void* stream) {
  if (!e || !dev_actions || !dev_out) return ORLG_EINVAL;
  if (!e->gsnr_ready) {
    e->err = "orlg_gsnr_for_actions before orlg_gsnr_configure";
    return ORLG_ESTATE;
  }
  if (e->p.kind != ORLG_ENV_RMSA && e->p.kind != ORLG_ENV_RMCSA) {
    e->err = "orlg_gsnr_for_actions takes RMSA-v0 (path, slot) or RMCSA-v0 (path, modulation, core, slot) actions";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  CK(launch_gsnr<true>(e, dev_actions, dev_out, (cudaStream_t)stream));
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_info_dim(const orlg_engine* e) { return (e && e->p.info_stats) ? 9 + e->p.n_bit_rates : 0; }
extern "C" int orlg_set_info_output(orlg_engine* e, double* dev_info) {
  if (!e) return ORLG_EINVAL;
  if (dev_info && !e->p.info_stats) {
    e->err = "orlg_set_info_output needs an engine created with info_stats = 1";
    return ORLG_ESTATE;
  }
  e->info_out = dev_info;
  return ORLG_OK;
}

extern "C" int orlg_phy_get_channel_state(orlg_engine* e, int32_t* dev_out, int32_t max_entries, int32_t* dev_count,
                                          void* stream) {
  if (!e || !dev_count || max_entries < 0 || (max_entries > 0 && !dev_out)) return ORLG_EINVAL;
  if (e->p.kind != ORLG_ENV_PHYRMSA || !e->p.phy_grooming) {
    e->err = "orlg_phy_get_channel_state needs a PhyRMSA engine with phy_grooming != 0";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  phy_channel_state_kernel<<<(e->p.num_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(e->p, dev_out, max_entries,
                                                                                          dev_count);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_phy_get_stats(orlg_engine* e, double* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  if (e->p.kind != ORLG_ENV_PHYRMSA) {
    e->err = "orlg_phy_get_stats needs env_kind PhyRMSA";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  const int warps_per_cta = 4;
  phy_stats_kernel<<<(e->p.num_envs + warps_per_cta - 1) / warps_per_cta, warps_per_cta * 32, 0, (cudaStream_t)stream>>>(
      e->p, e->W, e->layout_bd, dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_qot_get(orlg_engine* e, double* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  if (!e->p.qot) {
    e->err = "orlg_qot_get needs an engine created with qot_gate = 1";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  const long long n = 2ll * e->p.num_envs;
  qot_db_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(e->p.qot_g, dev_out, n);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_qot_check(orlg_engine* e, int64_t* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  if (!e->p.qot) {
    e->err = "orlg_qot_check needs an engine created with qot_gate = 1";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  // the services released by the last step leave the sums first (heuristic mode changes nothing else)
  if (!e->auto_actions) {
    int rc = dev_alloc(e, &e->auto_actions, (size_t)e->p.num_envs * e->p.action_dim);
    if (rc) return rc;
  }
  if (int rc = launch_qot(e, QOT_HEURISTIC, nullptr, e->auto_actions, stream)) return rc;
  qot_check_kernel<<<(e->p.num_envs + 63) / 64, 64, 0, (cudaStream_t)stream>>>(e->p, e->gp, (long long*)dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

// ---- orlg_step_auto_noalloc: orlg_step_auto without any allocation -------------------------------------------
// orlg_step_auto needs scratch for the action tensor that travels from the heuristic launch to the step launch; when
// the caller passes no buffer it allocates one (cudaMalloc) on first use.  Callers that must not allocate (CUDA graph
// capture, serving loops, harnesses that forbid it) use this variant: the scratch is a file-scope device array of
// ORLG_NOALLOC_MAX_INTS int32 (1 MiB; every GPU that loads the library holds it until exit).  Host code must not
// pass the symbol itself to a kernel (it is a host-side stand-in, not a device address): the address is taken once
// per device with cudaGetSymbolAddress, with that device current.
static __device__ __align__(256) int32_t g_noalloc_actions[ORLG_NOALLOC_MAX_INTS];
static int32_t* g_noalloc_ptr[64] = {};  // per device, resolved on first use

extern "C" int orlg_step_auto_noalloc(orlg_engine* e, int hid, int32_t* dev_reward, uint8_t* dev_done, void* dev_obs,
                                      int32_t* dev_alloc, void* stream) {
  if (!e || !heuristic_ok(e, hid)) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  const long long need = (long long)e->p.num_envs * e->p.action_dim;
  if (need > (long long)ORLG_NOALLOC_MAX_INTS) {
    e->err = "orlg_step_auto_noalloc: num_envs * action_dim = " + std::to_string(need) + " int32 exceed the static scratch of " +
             std::to_string((long long)ORLG_NOALLOC_MAX_INTS) + " (ORLG_NOALLOC_MAX_INTS); use orlg_step_auto with a caller buffer";
    return ORLG_EINVAL;
  }
  if (e->device < 0 || e->device >= 64) {
    e->err = "orlg_step_auto_noalloc: device ordinal out of range";
    return ORLG_EINVAL;
  }
  CK(cudaSetDevice(e->device));
  if (!g_noalloc_ptr[e->device]) {  // set-up, once per device, that device current
    void* ptr = nullptr;
    CK(cudaGetSymbolAddress(&ptr, g_noalloc_actions));
    g_noalloc_ptr[e->device] = (int32_t*)ptr;
  }
  // every env's action is written by the heuristic launch before the step launch reads it, so what an earlier
  // call left in the array is never read: nothing has to be cleared
  return orlg_step_auto(e, hid, g_noalloc_ptr[e->device], dev_reward, dev_done, dev_obs, dev_alloc, stream);
}

extern "C" int orlg_export_services(orlg_engine* e, int32_t* dev_out, int32_t max_rows, int32_t* dev_count, void* stream) {
  if (!e || !dev_count || max_rows < 0 || (max_rows > 0 && !dev_out)) return ORLG_EINVAL;
  cudaSetDevice(e->device);
  services_kernel<<<(e->p.num_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(e->p, dev_out, max_rows, dev_count);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

// ---- orlg_timed_steps: K steps timed on the device with nothing but C between the event records ----------------
extern "C" int orlg_timed_steps(orlg_engine* e, int mode, int hid, int32_t* dev_actions, int32_t* dev_reward,
                                uint8_t* dev_done, void* dev_obs, int nsteps, void* dev_flush, size_t flush_bytes,
                                int use_graph, float* ms_step, float* ms_first, void* stream) {
  if (!e || nsteps < 1 || nsteps > 100000 || !ms_step || (mode != 0 && mode != 1) || (mode == 0 && !dev_actions))
    return ORLG_EINVAL;
  if (mode == 1 && !heuristic_ok(e, hid)) return ORLG_EINVAL;
  if (int rc = check_source(e)) return rc;
  CK(cudaSetDevice(e->device));
  cudaStream_t caller = (cudaStream_t)stream, st = caller;
  std::vector<cudaEvent_t> ev(3 * (size_t)nsteps, nullptr);
  for (auto& x : ev) CK(cudaEventCreate(&x));
  int rc = ORLG_OK;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaStream_t own = nullptr;
  if (mode == 1 && !dev_actions) {  // allocate the action scratch outside the timed / captured region
    if (!e->auto_actions && (rc = ::dev_alloc(e, &e->auto_actions, (size_t)e->p.num_envs * e->p.action_dim))) return rc;
    dev_actions = e->auto_actions;
  }
  if (use_graph) {  // the whole K-step sequence as one graph launch: no host work between the nodes
    CK(cudaStreamSynchronize(caller));
    CK(cudaStreamCreateWithFlags(&own, cudaStreamNonBlocking));
    st = own;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  }
  const unsigned flag = use_graph ? cudaEventRecordExternal : cudaEventRecordDefault;
  e->mid_external = use_graph != 0;
  for (int t = 0; t < nsteps && rc == ORLG_OK; t++) {
    if (dev_flush && flush_bytes) cudaMemsetAsync(dev_flush, t & 1, flush_bytes, st);  // evicts L2, outside the event pair
    cudaEventRecordWithFlags(ev[3 * t], st, flag);
    e->mid_ev = ev[3 * t + 1];
    rc = mode == 0 ? orlg_step(e, dev_actions, dev_reward, dev_done, dev_obs, nullptr, st)
                   : orlg_step_auto(e, hid, dev_actions, dev_reward, dev_done, dev_obs, nullptr, st);
    cudaEventRecordWithFlags(ev[3 * t + 2], st, flag);
  }
  e->mid_ev = nullptr;
  e->mid_external = false;
  cudaError_t ce = cudaSuccess;
  if (use_graph) {
    ce = cudaStreamEndCapture(st, &graph);
    if (ce == cudaSuccess && rc == ORLG_OK) ce = cudaGraphInstantiate(&exec, graph, 0);
    if (ce == cudaSuccess && rc == ORLG_OK) ce = cudaGraphLaunch(exec, st);
  }
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
  if (ce == cudaSuccess && rc == ORLG_OK) {
    const bool two = mode == 1 || e->p.qot;
    for (int t = 0; t < nsteps && ce == cudaSuccess; t++) {
      ce = cudaEventElapsedTime(&ms_step[t], ev[3 * t], ev[3 * t + 2]);
      if (ms_first) {
        ms_first[t] = 0.f;
        if (two && ce == cudaSuccess) ce = cudaEventElapsedTime(&ms_first[t], ev[3 * t], ev[3 * t + 1]);
      }
    }
  }
  if (exec) cudaGraphExecDestroy(exec);
  if (graph) cudaGraphDestroy(graph);
  if (own) cudaStreamDestroy(own);
  for (auto& x : ev) cudaEventDestroy(x);
  if (rc != ORLG_OK) return rc;
  if (ce != cudaSuccess) {
    e->err = std::string("orlg_timed_steps: ") + cudaGetErrorString(ce);
    cudaGetLastError();
    return ORLG_ECUDA;
  }
  return ORLG_OK;
}

extern "C" int orlg_phy_get_moves(orlg_engine* e, int64_t* dev_out, void* stream) {
  if (!e || !dev_out) return ORLG_EINVAL;
  if (e->p.kind != ORLG_ENV_PHYRMSA) {
    e->err = "orlg_phy_get_moves needs env_kind PhyRMSA";
    return ORLG_ESTATE;
  }
  cudaSetDevice(e->device);
  phy_moves_kernel<<<(e->p.num_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(e->p, (long long*)dev_out);
  CK(cudaGetLastError());
  e->launches++;
  return ORLG_OK;
}

extern "C" int orlg_set_counter_snapshot(orlg_engine* e, int64_t* dev_snapshot) {
  if (!e) return ORLG_EINVAL;
  e->snap = (long long*)dev_snapshot;
  return ORLG_OK;
}
