// engine_kernel.cuh — the stepping kernel: TMA-staged tile in, step(s), tile out.
#pragma once
#include "engine_step.cuh"

namespace orlg {

// ---- sm_100a async-copy plumbing (TMA 1-D bulk copies + mbarrier) ------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// One launch = one pass of the hot path over every env (or nsteps passes in
// rollout mode).  Grid = NP / BD tiles, one thread per env.
template <int W, int BD>
__global__ void __launch_bounds__(BD) env_kernel(const Params p, const StepIO io) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  uint32_t* tile_s = reinterpret_cast<uint32_t*>(smem_raw);
  const int tid = threadIdx.x;
  const int e = blockIdx.x * BD + tid;
  const bool active = e < p.num_envs;
  const int mode = io.mode;
  const uint32_t tile_bytes = (uint32_t)p.planes * BD * 4u;
  uint32_t* tile_g = p.occ + (size_t)blockIdx.x * p.planes * BD;
  const bool need_tile = mode != MODE_RESET_FULL && mode != MODE_RESET_EPISODE;
  const bool dirty_tile = mode == MODE_STEP || mode == MODE_ROLLOUT || mode == MODE_RESET_FULL;

  if (need_tile) {
    if (tid == 0) mbar_init(&bar, 1);
    __syncthreads();
    if (tid == 0) {
      mbar_expect_tx(&bar, tile_bytes);
      bulk_g2s(tile_s, tile_g, tile_bytes, &bar);
    }
  }

  Env<W, BD> env(p, tile_s, tid, active ? e : 0);
  if (mode == MODE_RESET_FULL) {
    // rmsa_env.py:391-457: empty network, clock 0, counters 0, then the first request
    for (int pl = 0; pl < p.planes; pl++) tile_s[pl * BD + tid] = mask_word(pl % W, 0, p.S);
    env.now = 0.0;
    env.holding = 0.0;
    env.bit_rate = env.src = env.dst = 0;
    env.heap_len = 0;
    env.next_rel = __longlong_as_double(0x7FF0000000000000ll);
#pragma unroll
    for (int i = 0; i < CNT_N; i++) env.cnt[i] = 0;
    if (active) env.next_service();
  } else if (active) {
    env.load_scalars();
  }
  if (need_tile) mbar_wait(&bar, 0);

  if (active) {
    if (mode == MODE_RESET_EPISODE) {
      env.reset_episode();
    } else if (mode == MODE_HEURISTIC) {
      int a[4];
      heuristic_action<W, BD>(env, io.heuristic, a);
      for (int i = 0; i < p.action_dim; i++) io.actions_out[(size_t)e * p.action_dim + i] = a[i];
    } else if (mode == MODE_STEP || mode == MODE_ROLLOUT) {
      const int nsteps = (mode == MODE_ROLLOUT) ? io.nsteps : 1;
      long long rsum = 0;
      for (int t = 0; t < nsteps; t++) {
        int a[4];
        if (mode == MODE_ROLLOUT) {
          heuristic_action<W, BD>(env, io.heuristic, a);
        } else {
          for (int i = 0; i < p.action_dim; i++) a[i] = io.actions[(size_t)e * p.action_dim + i];
        }
        Decision d;
        if (p.kind == ORLG_ENV_RMSA) d = rmsa_decide<W, BD>(env, a[0], a[1]);
        else if (p.kind == ORLG_ENV_DEEPRMSA) d = deeprmsa_decide<W, BD>(env, a[0]);
        else d = rmcsa_decide<W, BD>(env, a[0], a[1], a[2], a[3]);
        if (p.kind == ORLG_ENV_RMCSA) {  // rmcsa_env.py:292-295
          env.cnt[CNT_PROC]++;
          env.cnt[CNT_EP_PROC]++;
          env.cnt[CNT_BR_REQ] += env.bit_rate;
          env.cnt[CNT_EP_BR_REQ] += env.bit_rate;
        }
        // reward: optical_network_env.py:213-214 ; deeprmsa_env.py:123-124
        int r = (p.kind == ORLG_ENV_DEEPRMSA) ? (d.accepted ? 1 : -1) : (d.accepted ? 1 : 0);
        env.next_service();
        int dn = env.cnt[CNT_EP_PROC] == (long long)p.episode_length;  // rmsa_env.py:339
        if (dn) {
          env.cnt[CNT_EPISODES]++;
          if (p.auto_reset) env.reset_episode();
        }
        rsum += r;
        if (mode == MODE_STEP) {
          if (io.reward) io.reward[e] = r;
          if (io.done) io.done[e] = (unsigned char)dn;
          if (io.alloc) {
            int4 v = make_int4(d.accepted, d.accepted ? (d.path | (d.core << 16)) : -1,
                               d.accepted ? d.start : -1, d.n);
            reinterpret_cast<int4*>(io.alloc)[e] = v;
          }
        }
      }
      if (io.reward_sum) io.reward_sum[e] += rsum;
    }
    if ((mode == MODE_STEP || mode == MODE_OBS || mode == MODE_RESET_FULL) && io.obs != nullptr) {
      if (io.obs_kind == ORLG_OBS_F64)
        write_deeprmsa_obs<W, BD, double>(env, reinterpret_cast<double*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
      else if (io.obs_kind == ORLG_OBS_F32)
        write_deeprmsa_obs<W, BD, float>(env, reinterpret_cast<float*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
      else if (io.obs_kind == ORLG_OBS_MATRIX_U8)
        write_matrix_obs<W, BD>(env, reinterpret_cast<unsigned char*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
    }
    if (mode != MODE_HEURISTIC && mode != MODE_OBS) env.store_scalars();
  }

  if (dirty_tile) {
    fence_async_smem();  // make this thread's smem writes visible to the async proxy
    __syncthreads();
    if (tid == 0) bulk_s2g(tile_g, tile_s, tile_bytes);
  }
}

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

}  // namespace orlg
