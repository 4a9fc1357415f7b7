// bitrow.cuh — W-word spectrum rows held in registers, one env per thread.
//
// A row is the packed form of one line of the reference's
// topology.graph["available_slots"] (rmsa_env.py:396-398): bit s of word s/32
// is 1 iff slot s is free.  Bits >= num_slots are always 0, so they can never
// be part of a free block.  Everything here is written with static word
// indices (no dynamically indexed register arrays -> no local memory).
#pragma once
#include <cstdint>

namespace orlg {

template <int W>
struct Row {
  uint32_t w[W];
};

template <int W>
__device__ __forceinline__ Row<W> row_ones() {
  Row<W> r;
#pragma unroll
  for (int i = 0; i < W; i++) r.w[i] = 0xFFFFFFFFu;
  return r;
}

// bits [0, len) that fall into word `wi`: the funnel shift clamps its amount to 32, so a prefix that covers
// the whole word (or more) needs no special case
__device__ __forceinline__ uint32_t prefix_word(int wi, int len) {
  return ~__funnelshift_lc(0u, 0xFFFFFFFFu, (uint32_t)max(len - 32 * wi, 0));
}
// bits of [lo, hi) that fall into word `wi` (empty when hi <= lo)
__device__ __forceinline__ uint32_t mask_word(int wi, int lo, int hi) {
  return prefix_word(wi, hi) & ~prefix_word(wi, lo);
}

// logical right shift of the whole row by 2^B bits (zeros enter at the top)
template <int W, int B>
__device__ __forceinline__ Row<W> shr_pow2(const Row<W>& x) {
  Row<W> r;
  if constexpr (B < 5) {
    constexpr int s = 1 << B;
#pragma unroll
    for (int i = 0; i < W; i++) {
      uint32_t hi = (i + 1 < W) ? x.w[i + 1] : 0u;
      r.w[i] = __funnelshift_r(x.w[i], hi, s);
    }
  } else {
    constexpr int ws = 1 << (B - 5);
#pragma unroll
    for (int i = 0; i < W; i++) r.w[i] = (i + ws < W) ? x.w[i + ws] : 0u;
  }
  return r;
}

template <int W>
__device__ __forceinline__ void and_into(Row<W>& a, const Row<W>& b) {
#pragma unroll
  for (int i = 0; i < W; i++) a.w[i] &= b.w[i];
}

// number of shift-amount bits that matter for a W-word row: amounts < 32*W
template <int W>
struct ShiftBits {
  static constexpr int value = W <= 1 ? 5 : W <= 2 ? 6 : W <= 4 ? 7 : W <= 8 ? 8 : W <= 16 ? 9 : 10;
};

template <int W, int B>
__device__ __forceinline__ void shr_words(Row<W>& x, int amount) {
  if ((amount >> B) & 1) x = shr_pow2<W, B>(x);
  if constexpr (B + 1 < ShiftBits<W>::value) shr_words<W, B + 1>(x, amount);
}

// right shift by an arbitrary amount in [0, 32*W): one funnel shift per word for the sub-word part,
// then conditional whole-word moves
template <int W>
__device__ __forceinline__ Row<W> shr_var(Row<W> x, int amount) {
  Row<W> r;
  const int s = amount & 31;
#pragma unroll
  for (int i = 0; i < W; i++) r.w[i] = __funnelshift_r(x.w[i], (i + 1 < W) ? x.w[i + 1] : 0u, s);
  shr_words<W, 5>(r, amount);
  return r;
}

template <int W, int B>
__device__ __forceinline__ void window_step(Row<W>& b, int& len, int n) {
  if (2 * len <= n) {
    Row<W> s = shr_pow2<W, B>(b);
    and_into(b, s);
    len *= 2;
  }
  if constexpr (B + 1 < ShiftBits<W>::value) window_step<W, B + 1>(b, len, n);
}

// bit p of the result is 1 iff bits p .. p+n-1 of `a` are all 1 (n >= 1)
template <int W>
__device__ __forceinline__ Row<W> window_starts(const Row<W>& a, int n) {
  Row<W> b = a;
  if (n > 32 * W) {  // longer than the row: no window (also keeps every shift amount below 32*W)
#pragma unroll
    for (int i = 0; i < W; i++) b.w[i] = 0u;
    return b;
  }
  int len = 1;
  window_step<W, 0>(b, len, n);
  int r = n - len;  // 0 <= r < len
  if (r > 0) {
    Row<W> s = shr_var<W>(b, r);
    and_into(b, s);
  }
  return b;
}

// index of the lowest set bit, -1 if none
template <int W>
__device__ __forceinline__ int lowest_set(const Row<W>& x) {
  int r = -1;
#pragma unroll
  for (int i = W - 1; i >= 0; i--)
    if (x.w[i]) r = 32 * i + (__ffs((int)x.w[i]) - 1);
  return r;
}

// index of the highest set bit, -1 if none
template <int W>
__device__ __forceinline__ int highest_set(const Row<W>& x) {
  int r = -1;
#pragma unroll
  for (int i = 0; i < W; i++)
    if (x.w[i]) r = 32 * i + 31 - __clz((int)x.w[i]);
  return r;
}

// clear every bit below `pos`
template <int W>
__device__ __forceinline__ void clear_below(Row<W>& x, int pos) {
#pragma unroll
  for (int i = 0; i < W; i++) x.w[i] &= ~mask_word(i, 0, pos);
}

template <int W>
__device__ __forceinline__ int popcount(const Row<W>& x) {
  int c = 0;
#pragma unroll
  for (int i = 0; i < W; i++) c += __popc(x.w[i]);
  return c;
}

// number of maximal runs of ones (the value==1 entries of rle(), rmsa_env.py:758-772)
template <int W>
__device__ __forceinline__ int count_runs(const Row<W>& x) {
  int c = 0;
#pragma unroll
  for (int i = 0; i < W; i++) {
    uint32_t prev = (x.w[i] << 1) | (i > 0 ? (x.w[i - 1] >> 31) : 0u);
    c += __popc(x.w[i] & ~prev);
  }
  return c;
}

// are bits [start, start+n) all ones?
template <int W>
__device__ __forceinline__ bool range_is_free(const Row<W>& a, int start, int n) {
  bool ok = true;
#pragma unroll
  for (int i = 0; i < W; i++) {
    uint32_t m = mask_word(i, start, start + n);
    ok = ok && ((a.w[i] & m) == m);
  }
  return ok;
}

// length of the run of ones that starts at `pos` (bit pos must be 1; bits >= S are 0)
template <int W>
__device__ __forceinline__ int run_length_from(const Row<W>& a, int pos) {
  Row<W> z;
#pragma unroll
  for (int i = 0; i < W; i++) z.w[i] = ~a.w[i];
  clear_below<W>(z, pos);
  int e = lowest_set<W>(z);
  return (e < 0 ? 32 * W : e) - pos;
}

}  // namespace orlg
