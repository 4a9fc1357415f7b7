// engine_step.cuh — decisions, heuristics and the DeepRMSA observation, per thread.
#pragma once
#include "engine_env.cuh"

namespace orlg {

struct Decision {
  int accepted, path, core, start, n;
};

// first j maximal runs of ones with length >= n, ascending start
// (get_available_blocks rmsa_env.py:774-804).  Calls f(block_index, start, length).
template <int W, typename F>
__device__ __forceinline__ int for_each_block(const Row<W>& avail, int n, int j, F f) {
  Row<W> b = window_starts<W>(avail, n);
  int nb = 0;
  while (nb < j) {
    int s = lowest_set<W>(b);
    if (s < 0) break;
    int len = run_length_from<W>(avail, s);
    f(nb, s, len);
    nb++;
    clear_below<W>(b, s + len);
  }
  return nb;
}

// RMSAEnv.step decision (rmsa_env.py:233-271)
template <int W, int BD, typename ENV>
__device__ __forceinline__ Decision rmsa_decide(ENV& env, int path, int slot) {
  const Params& p = env.p;
  Decision d{0, -1, 0, -1, 0};
  if (path >= 0 && slot >= 0 && path < p.k && slot < p.S) {
    PathRec rec = load_path(p, env.pair(), path);
    if (rec.hops() > 0) {
      int n = number_slots(env.bit_rate, rec.se(), p);
      if (slot + n <= p.S) {
        Row<W> a = env.tile.path_avail(rec, 0);
        if (range_is_free<W>(a, slot, n)) {
          env.provision(rec, path, 0, slot, n, rec.se());
          d = Decision{1, path, 0, slot, n};
        }
      }
    }
  }
  return d;
}

// DeepRMSAEnv.step (deeprmsa_env.py:48-58): action -> (route, block) -> start of that block
template <int W, int BD, typename ENV>
__device__ __forceinline__ Decision deeprmsa_decide(ENV& env, int act) {
  const Params& p = env.p;
  Decision d{0, -1, 0, -1, 0};
  if (act >= 0 && act < p.k * p.j) {
    int route = act / p.j, block = act % p.j;
    PathRec rec = load_path(p, env.pair(), route);
    if (rec.hops() > 0) {
      int n = number_slots(env.bit_rate, rec.se(), p);
      Row<W> a = env.tile.path_avail(rec, 0);
      int start = -1;
      for_each_block<W>(a, n, block + 1, [&](int b, int s, int) { if (b == block) start = s; });
      if (start >= 0) {  // a block of >= n free slots: is_path_free holds by construction
        env.provision(rec, route, 0, start, n, rec.se());
        d = Decision{1, route, 0, start, n};
      }
    }
  }
  return d;
}

// RMCSAEnv.step decision (rmcsa_env.py:222-289)
template <int W, int BD, typename ENV>
__device__ __forceinline__ Decision rmcsa_decide(ENV& env, int path, int mod, int core, int slot) {
  const Params& p = env.p;
  Decision d{0, -1, 0, -1, 0};
  if (path >= 0 && mod >= 0 && core >= 0 && slot >= 0 && path < p.k && mod < p.n_mods &&
      core < p.cores && slot < p.S) {
    int pr = env.pair();
    PathRec rec = load_path(p, pr, path);
    if (rec.hops() > 0) {
      int n = number_slots(env.bit_rate, __ldg(p.mod_se + mod), p);
      if (slot + n <= p.S) {
        Row<W> a = env.tile.path_avail(rec, core);
        if (range_is_free<W>(a, slot, n) &&
            env.crosstalk_ok(mod, __ldg(p.path_length + (size_t)pr * p.k + path))) {
          env.provision(rec, path, core, slot, n, __ldg(p.mod_se + mod));
          d = Decision{1, path, core, slot, n};
        }
      }
    }
  }
  return d;
}

// Reference heuristics; `a` receives action_dim ints.
template <int W, int BD, typename ENV>
__device__ __forceinline__ void heuristic_action(ENV& env, int hid, int* a, int in_path = 0) {
  const Params& p = env.p;
  int pr = env.pair();
  if (p.kind == ORLG_ENV_RMSA && hid == ORLG_H_PATH_FF) {  // PathOnlyFirstFitAction rmsa_env.py:982-1005
    a[0] = p.k;
    a[1] = p.S;
    if (in_path >= 0 && in_path < p.k) {
      PathRec rec = load_path(p, pr, in_path);
      if (rec.hops() > 0) {
        int s = first_fit<W>(env.tile.path_avail(rec, 0), number_slots(env.bit_rate, rec.se(), p), p.S);
        if (s >= 0) {
          a[0] = in_path;
          a[1] = s;
        }
      }
    }
  } else if (p.kind == ORLG_ENV_RMSA) {
    a[0] = p.k;
    a[1] = p.S;
    long long best_free = 0;
    int npaths = (hid == ORLG_H_SP_FF) ? 1 : p.k;
    for (int path = 0; path < npaths; path++) {
      PathRec rec = load_path(p, pr, path);
      if (rec.hops() <= 0) continue;
      int n = number_slots(env.bit_rate, rec.se(), p);
      Row<W> av = env.tile.path_avail(rec, 0);
      int s = first_fit<W>(av, n, p.S);
      if (s < 0) continue;
      if (hid == ORLG_H_LLP_FF) {  // rmsa_env.py:916-937, strict '>' keeps the first best
        int fr = popcount<W>(av);
        if (fr > best_free) {
          a[0] = path;
          a[1] = s;
          best_free = fr;
        }
      } else {  // SP-FF rmsa_env.py:854-871 / SAP-FF rmsa_env.py:901-913
        a[0] = path;
        a[1] = s;
        break;
      }
    }
  } else if (p.kind == ORLG_ENV_DEEPRMSA) {
    if (hid == ORLG_H_SP_FF && !p.allow_rejection) {  // deeprmsa_env.py:135-137
      a[0] = 0;
      return;
    }
    a[0] = p.k * p.j;
    int npaths = (hid == ORLG_H_SP_FF) ? 1 : p.k;
    for (int path = 0; path < npaths; path++) {  // deeprmsa_env.py:138-155
      PathRec rec = load_path(p, pr, path);
      if (rec.hops() <= 0) continue;
      int n = number_slots(env.bit_rate, rec.se(), p);
      Row<W> av = env.tile.path_avail(rec, 0);
      Row<W> b = window_starts<W>(av, n);
      if (lowest_set<W>(b) >= 0) {
        a[0] = path * p.j;
        break;
      }
    }
  } else {  // rmcsa_env.py:882-911
    a[0] = p.k;
    a[1] = p.n_mods;
    a[2] = p.cores;
    a[3] = p.S;
    // ORLG_H_PATH_FF: the walk restricted to the given path (the device form of rmcsa_env.py:950-982, which as
    // shipped calls get_number_slots / is_path_free with RMSA's signatures and cannot run)
    const bool one = hid == ORLG_H_PATH_FF;
    if (one && (in_path < 0 || in_path >= p.k)) return;
    for (int path = one ? in_path : 0; path < (one ? in_path + 1 : p.k); path++) {
      PathRec rec = load_path(p, pr, path);
      int mod = rec.best_mod();
      if (rec.hops() <= 0 || mod >= p.n_mods) continue;
      int n = number_slots(env.bit_rate, __ldg(p.mod_se + mod), p);
      for (int core = 0; core < p.cores; core++) {
        Row<W> av = env.tile.path_avail(rec, core);
        int s = first_fit<W>(av, n, p.S);
        if (s >= 0) {
          a[0] = path;
          a[1] = mod;
          a[2] = core;
          a[3] = s;
          return;
        }
      }
    }
  }
}

// DeepRMSAEnv.observation (deeprmsa_env.py:60-121), f64 arithmetic in the reference's
// operation order; OUT is double or float (one rounding at the store).
// Returns the bitmask of paths that have at least one fitting block (what the K-SP-FF / SP-FF
// heuristics of deeprmsa_env.py:135-155 decide on), so a fused next-action hint costs nothing extra.
template <int W, int BD, typename OUT, typename ENV>
__device__ __forceinline__ unsigned write_deeprmsa_obs(ENV& env, OUT* o, long long fs) {
  unsigned fit_mask = 0u;
  const Params& p = env.p;
  const int N = p.n_nodes, S = p.S, j = p.j;
  o[0] = (OUT)((double)env.bit_rate / 100.0);
  int mn = min(env.src, env.dst), mx = max(env.src, env.dst);
  for (int i = 0; i < N; i++) {
    o[(1 + i) * fs] = (OUT)(i == mn ? 1.0 : 0.0);
    o[(1 + N + i) * fs] = (OUT)(i == mx ? 1.0 : 0.0);
  }
  int pr = env.pair();
  const int rowlen = 2 * j + 3;
  for (int path = 0; path < p.k; path++) {
    OUT* row = o + (long long)(1 + 2 * N + path * rowlen) * fs;
    PathRec rec = load_path(p, pr, path);
    if (rec.hops() <= 0) {
      for (int i = 0; i < rowlen; i++) row[i * fs] = (OUT)-1.0;
      continue;
    }
    int n = number_slots(env.bit_rate, rec.se(), p);
    Row<W> av = env.tile.path_avail(rec, 0);
    int nb = for_each_block<W>(av, n, j, [&](int b, int s, int len) {
      row[(2 * b) * fs] = (OUT)__ldg(p.obs_pos + s);
      row[(2 * b + 1) * fs] = (OUT)(((double)len - 8.0) * 0.125);  // == /8 exactly
    });
    if (nb > 0) fit_mask |= 1u << path;
    for (int b = nb; b < j; b++) {
      row[(2 * b) * fs] = (OUT)-1.0;
      row[(2 * b + 1) * fs] = (OUT)-1.0;
    }
    row[(2 * j) * fs] = (OUT)(n <= S + 1 ? __ldg(p.obs_slots + n) : ((double)n - 5.5) / 3.5);
    int tot = popcount<W>(av), runs = count_runs<W>(av);
    row[(2 * j + 1) * fs] = (OUT)__ldg(p.obs_pos + tot);
    row[(2 * j + 2) * fs] = (OUT)(runs > 0 ? ((double)tot / (double)runs - 4.0) * 0.25 : -1.0);  // == /4 exactly
  }
  return fit_mask;
}

// SimpleMatrixObservation (rmsa_env.py:940-971, rmcsa_env.py:914-947): one-hot pair + unpacked slots
template <int W, int BD, typename ENV>
__device__ __forceinline__ void write_matrix_obs(ENV& env, unsigned char* o, long long fs) {
  const Params& p = env.p;
  int mn = min(env.src, env.dst), mx = max(env.src, env.dst);
  for (int i = 0; i < p.n_nodes; i++) {
    o[i * fs] = (unsigned char)(i == mn);
    o[(p.n_nodes + i) * fs] = (unsigned char)(i == mx);
  }
  long long base = 2 * p.n_nodes;
  for (int c = 0; c < p.cores; c++)
    for (int l = 0; l < p.E; l++)
      for (int s = 0; s < p.S; s++)
        o[(base + ((long long)(c * p.E + l) * p.S + s)) * fs] =
            (unsigned char)((env.tile.word(c, l, s >> 5) >> (s & 31)) & 1u);
}

}  // namespace orlg
