// engine_kernel.cuh — the stepping kernel: TMA-staged tile in, step(s), tile out.
#pragma once
#include "engine_phy.cuh"

namespace orlg {

// ---- sm_100a async-copy plumbing (TMA 1-D bulk copies + mbarrier) ------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  // single-CTA use: make the init visible to the async proxy (no cluster-scope fence -> no L1 flush)
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
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
  // the kernel may exit once the source tile has been read; the global writes complete on their own
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// One launch = one pass of the hot path over every env (or nsteps passes in
// rollout mode).  Grid = NP / BD tiles, one thread per env.  MT is the mode the
// instantiation is specialised for (MODE_STEP / MODE_HEURISTIC / MODE_ROLLOUT), or
// -1 for the rarely used modes (observation, resets), which read io.mode.
//
// DIRECT = false: the tile of BD envs is staged through shared memory (occ[tile][plane][BD]).
// DIRECT = true : nothing is staged; the occupancy is env-major in HBM (occ[env][plane], i.e. the
// same layout with a tile of one env) and every thread reads and writes its own rows in place
// with vector loads.  This is the variant for envs whose occupancy is so large (multi-core, wide
// spectra, many links) that only a warp or two of staged tiles would fit on an SM.
#ifdef ORLG_MIN_CTAS  // A/B knob: resident CTAs per SM the register allocation must allow
#define ORLG_LAUNCH_BOUNDS(BD) __launch_bounds__(BD, ORLG_MIN_CTAS)
#else
#define ORLG_LAUNCH_BOUNDS(BD) __launch_bounds__(BD)
#endif
template <int W, int BD, int MT, bool PHY, bool DIRECT = false, bool INFO = false>
__global__ void ORLG_LAUNCH_BOUNDS(BD) env_kernel(const Params p, const StepIO io) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  constexpr int PS = DIRECT ? 1 : BD;  // plane stride of the occupancy this thread sees
  const int tid = threadIdx.x;
  const int e = blockIdx.x * BD + tid;
  const bool active = e < p.num_envs;
  uint32_t* tile_s = DIRECT ? p.occ + (size_t)(active ? e : 0) * p.planes : reinterpret_cast<uint32_t*>(smem_raw);
  const int mode = MT >= 0 ? MT : io.mode;
  const uint32_t tile_bytes = (uint32_t)p.planes * BD * 4u;
  uint32_t* tile_g = p.occ + (size_t)blockIdx.x * p.planes * BD;
  const bool need_tile = !DIRECT && mode != MODE_RESET_FULL && mode != MODE_RESET_EPISODE;
  const bool dirty_tile = !DIRECT && (mode == MODE_STEP || mode == MODE_ROLLOUT || mode == MODE_RESET_FULL);
  const bool stepping = mode == MODE_STEP || mode == MODE_ROLLOUT;

  if (need_tile) {
    if (tid == 0) {
      mbar_init(&bar, 1);
      mbar_expect_tx(&bar, tile_bytes);
      bulk_g2s(tile_s, tile_g, tile_bytes, &bar);
    }
  }

  Env<W, PS, PHY, INFO> env(p, tile_s, DIRECT ? 0 : tid, active ? e : 0);
  typename Env<W, PS, PHY, INFO>::Req nx{};
  constexpr int ADIM = PHY ? PHY_ADIM : 4;
  int a[ADIM];
#pragma unroll
  for (int i = 0; i < ADIM; i++) a[i] = 0;
  if (mode == MODE_RESET_FULL) {
    // rmsa_env.py:391-457: empty network, clock 0, counters 0, then the first request
    if (!DIRECT || active)
      for (int pl = 0; pl < p.planes; pl++) tile_s[pl * PS + (DIRECT ? 0 : tid)] = mask_word(pl % W, 0, p.S);
    env.now = 0.0;
    env.holding = 0.0;
    env.bit_rate = env.src = env.dst = 0;
    env.heap_len = 0;
    env.cs_len = env.cs_dead = 0;  // channel_state phy_rmsa_env.py:455-461
    env.slot_hops = 0;
    if (INFO && !PHY && active) {  // optical_network_env.py:257-260, rmsa_env.py:444-452
      for (int l = 0; l < p.E; l++) env.lnk(0, l) = env.lnk(1, l) = env.lnk(2, l) = 0.0;
      for (int i = 0; i < 2 * p.n_bit_rates; i++) p.br_hist[(size_t)i * p.NP + env.e] = 0;
    }
    env.next_rel = __longlong_as_double(0x7FF0000000000000ll);
    if (PHY && active) {  // every row of the service table is free; empty order list
      for (int i = 0; i < p.phy_svc_cap; i++)
        p.phy_free[(size_t)env.e * p.phy_svc_cap + i] = (unsigned short)(p.phy_svc_cap - 1 - i);
      env.sv_free_n = p.phy_svc_cap;
      env.run_len = env.run_dead = 0;
      env.mv[0] = env.mv[1] = env.mv[2] = 0;
    }
    if (p.qot && active) {  // every row of the service table is free, nothing waits to leave the sums
      for (int i = 0; i < p.qot_slot_cap; i++)
        p.qot_free[(size_t)env.e * p.qot_slot_cap + i] = (unsigned short)(p.qot_slot_cap - 1 - i);
      p.qot_free_n[env.e] = p.qot_slot_cap;
      p.qot_add[env.e] = 0ull;
    }
#pragma unroll
    for (int i = 0; i < CNT_N; i++) env.cnt[i] = 0;
    env.fsum[0] = env.fsum[1] = 0.0;
    env.isum[0] = env.isum[1] = env.isum[2] = env.isum[3] = 0;
    if (active) env.next_service();
  } else if (active) {
    // everything that does not need the tile happens while the bulk copy is in flight
    if (mode == MODE_HEURISTIC || mode == MODE_OBS) {
      env.load_request();
    } else {
      env.load_scalars();
      if (mode == MODE_STEP && PHY) {
#pragma unroll
        for (int i = 0; i < ADIM; i++) a[i] = io.actions[(size_t)e * ADIM + i];
      } else if (mode == MODE_STEP) {
        if (p.action_dim == 1) {
          a[0] = io.actions[e];
        } else if (p.action_dim == 2) {
          int2 v = reinterpret_cast<const int2*>(io.actions)[e];
          a[0] = v.x;
          a[1] = v.y;
        } else {
          int4 v = reinterpret_cast<const int4*>(io.actions)[e];
          a[0] = v.x;
          a[1] = v.y;
          a[2] = v.z;
          a[3] = v.w;
        }
      }
      if (stepping) {
        nx = env.draw_request();
        env.prefetch_queue(nx.now);
      }
    }
  }
  if (need_tile) {
    __syncthreads();  // the barrier init by thread 0 is visible to every waiter
    mbar_wait(&bar, 0);
  }

  if (active) {
    if (mode == MODE_RESET_EPISODE) {
      env.reset_episode();
    } else if (mode == MODE_HEURISTIC && PHY) {
      if constexpr (PHY) phy_heuristic<W, PS>(env, io.heuristic, a);
#pragma unroll
      for (int i = 0; i < ADIM; i++) io.actions_out[(size_t)e * ADIM + i] = a[i];
    } else if (mode == MODE_HEURISTIC) {
      if constexpr (!PHY)
        heuristic_action<W, PS>(env, io.heuristic, a,
                                (io.heuristic == ORLG_H_PATH_FF && io.actions) ? io.actions[e] : 0);
      if (p.action_dim == 1) {
        io.actions_out[e] = a[0];
      } else if (p.action_dim == 2) {
        reinterpret_cast<int2*>(io.actions_out)[e] = make_int2(a[0], a[1]);
      } else {
        reinterpret_cast<int4*>(io.actions_out)[e] = make_int4(a[0], a[1], a[2], a[3]);
      }
    } else if (stepping) {
      const int nsteps = (mode == MODE_ROLLOUT) ? io.nsteps : 1;
      long long rsum = 0;
      for (int t = 0; t < nsteps; t++) {
        if (mode == MODE_ROLLOUT) {
          if constexpr (PHY) phy_heuristic<W, PS>(env, io.heuristic, a);
          else heuristic_action<W, PS>(env, io.heuristic, a);
          if (t > 0) nx = env.draw_request();
        }
        Decision d;
        double prev_compactness = 0.0;
        const bool want_info = INFO && !PHY && mode == MODE_STEP && io.info != nullptr;
        if constexpr (INFO && !PHY)
          if (want_info) prev_compactness = env.info_network_compactness();  // rmsa_env.py:227-229
        if constexpr (PHY) {
          d = phy_decide<W, PS>(env, a);
          // the info metrics are taken after the provisioning and before the releases (:313-314 vs :351)
          if (mode == MODE_STEP && io.obs != nullptr &&
              (io.obs_kind == ORLG_OBS_PHY_INFO || io.obs_kind == ORLG_OBS_PHY_CUTS))
            phy_info<W, PS>(env, reinterpret_cast<double*>(io.obs) + (long long)e * io.obs_es, io.obs_fs,
                            io.obs_kind == ORLG_OBS_PHY_INFO);
        } else if (p.qot) {
          // GN-model QoT gate: qot_kernel has taken the decision (validity, free spectrum, RMCSA's closed-form
          // gate, GSNR of the candidate and its neighbours); apply its verdict
          const int4 v0 = p.qot_verdict[2 * (size_t)e], v1 = p.qot_verdict[2 * (size_t)e + 1];
          d = Decision{0, -1, 0, -1, 0};
          if (v0.x) {
            const int path = v0.y & 0xFFFF, core = v0.y >> 16;
            PathRec rec = load_path(p, env.pair(), path);
            env.provision(rec, path, core, v0.z, v0.w, v1.x, v1.y);
            d = Decision{1, path, core, v0.z, v0.w};
          }
        } else {
          if (p.kind == ORLG_ENV_RMSA) d = rmsa_decide<W, PS>(env, a[0], a[1]);
          else if (p.kind == ORLG_ENV_DEEPRMSA) d = deeprmsa_decide<W, PS>(env, a[0]);
          else d = rmcsa_decide<W, PS>(env, a[0], a[1], a[2], a[3]);
        }
        if constexpr (INFO && !PHY)
          if (want_info) env.info_write(io.info + (size_t)e * (9 + p.n_bit_rates), prev_compactness);
        if (p.kind == ORLG_ENV_RMCSA) {  // rmcsa_env.py:292-295
          env.cnt[CNT_PROC]++;
          env.cnt[CNT_EP_PROC]++;
          env.cnt[CNT_BR_REQ] += env.bit_rate;
          env.cnt[CNT_EP_BR_REQ] += env.bit_rate;
        }
        if (mode == MODE_STEP && io.snap != nullptr) {
          // the counters as the reference's info dict reads them: after the decision, before _next_service counts
          // the next request and before an episode reset (rmsa_env.py:293-307 precede :334-335)
          long long* o = io.snap + (size_t)e * 8;
#pragma unroll
          for (int i = 0; i < 8; i++) o[i] = env.cnt[i];
        }
        // reward: optical_network_env.py:213-214 ; deeprmsa_env.py:123-124
        int r = (p.kind == ORLG_ENV_DEEPRMSA) ? (d.accepted ? 1 : -1) : (d.accepted ? 1 : 0);
        env.commit_request(nx);
        if constexpr (PHY && INFO)  // periodic defragmentation, after _next_service (phy_rmsa_env.py:355-417)
          if (p.phy_defrag_period > 0 && env.cnt[CNT_PROC] % p.phy_defrag_period == 0) phy_defragmentation<W, PS>(env);
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
    unsigned fit_mask = 0u;
    bool have_fit_mask = false;
    if ((mode == MODE_STEP || mode == MODE_OBS || mode == MODE_RESET_FULL) && io.obs != nullptr) {
      if constexpr (!PHY) {
        if (io.obs_kind == ORLG_OBS_F64) {
          fit_mask = write_deeprmsa_obs<W, PS, double>(env, reinterpret_cast<double*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
          have_fit_mask = true;
        } else if (io.obs_kind == ORLG_OBS_F32) {
          fit_mask = write_deeprmsa_obs<W, PS, float>(env, reinterpret_cast<float*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
          have_fit_mask = true;
        }
      } else if (mode != MODE_STEP && (io.obs_kind == ORLG_OBS_PHY_INFO || io.obs_kind == ORLG_OBS_PHY_CUTS)) {
        phy_info<W, PS>(env, reinterpret_cast<double*>(io.obs) + (long long)e * io.obs_es, io.obs_fs,
                        io.obs_kind == ORLG_OBS_PHY_INFO);
      }
      if (io.obs_kind == ORLG_OBS_MATRIX_U8)
        write_matrix_obs<W, PS>(env, reinterpret_cast<unsigned char*>(io.obs) + (long long)e * io.obs_es, io.obs_fs);
    }
    if (mode == MODE_STEP && io.hint_actions != nullptr) {
      // fused next-action hint: what orlg_heuristic would write if it were launched now
      if constexpr (PHY) {
        phy_heuristic<W, PS>(env, io.hint_heuristic, a);
#pragma unroll
        for (int i = 0; i < ADIM; i++) io.hint_actions[(size_t)e * ADIM + i] = a[i];
      } else {
        if (p.kind == ORLG_ENV_DEEPRMSA && have_fit_mask &&
            (io.hint_heuristic == ORLG_H_SAP_FF || io.hint_heuristic == ORLG_H_SP_FF)) {
          // deeprmsa_env.py:135-155 decided from the blocks the observation just enumerated
          if (io.hint_heuristic == ORLG_H_SP_FF)
            a[0] = (!p.allow_rejection || (fit_mask & 1u)) ? 0 : p.k * p.j;
          else
            a[0] = fit_mask ? (__ffs((int)fit_mask) - 1) * p.j : p.k * p.j;
        } else {
          heuristic_action<W, PS>(env, io.hint_heuristic, a);
        }
        if (p.action_dim == 1) {
          io.hint_actions[e] = a[0];
        } else if (p.action_dim == 2) {
          reinterpret_cast<int2*>(io.hint_actions)[e] = make_int2(a[0], a[1]);
        } else {
          reinterpret_cast<int4*>(io.hint_actions)[e] = make_int4(a[0], a[1], a[2], a[3]);
        }
      }
    }
    if (mode != MODE_HEURISTIC && mode != MODE_OBS) env.store_scalars();
  }

  if (dirty_tile) {
    fence_async_smem();  // make this thread's smem writes visible to the async proxy
    __syncthreads();
    if (tid == 0) bulk_s2g(tile_g, tile_s, tile_bytes);
  }
}

}  // namespace orlg
