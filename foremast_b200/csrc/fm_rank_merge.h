// fm_rank_merge.h -- per-lane rank bookkeeping of the rank-test kernel (fm_rank.cu).
//
// After the warp has sorted the baseline and current samples separately, 32 lanes split the
// pooled (merged) order into equal spans with a merge-path search and walk the tie groups that
// START in their span, accumulating exact integer statistics:
//     2*R1 += cnt_base(group) * (2*s + t + 1)       (twice the average rank of positions [s, s+t))
//     T    += t^3 - t
// Long tie groups (error counters that sit at 0 all day) are crossed by galloping, not stepping.
// The functions are plain C++ on arrays in the kernel's padded shared-memory layout so the same
// code is unit-tested on the host (tests/test_rank_merge_host.py compiles it with g++).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define FM_HD __host__ __device__ __forceinline__
#define FM_HD_COLD __host__ __device__ __noinline__  // tie-group slow path: keep it out of the hot footprint
#else
#define FM_HD inline
#define FM_HD_COLD inline
#endif

// shared-memory index of sorted element e: one pad word per 32 keeps the warp's post-sort
// stores (lane stride R words) and the analysis reads spread over banks
FM_HD int fm_pidx(int e) { return e + (e >> 5); }

// a[start..n) is sorted; returns the first index >= start whose element differs from v
// (start itself if a[start] != v).  Gallop + binary search: O(log run length).
template <typename T>
FM_HD_COLD int fm_run_end(const T *a, int n, int start, T v) {
  int e = start;
  if (e < n && a[fm_pidx(e)] == v) {
    int step = 1;
    while (e + step < n && a[fm_pidx(e + step)] == v) { e += step; step <<= 1; }
    int lo = e + 1, hi = (e + step < n) ? e + step : n;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (a[fm_pidx(mid)] == v) lo = mid + 1; else hi = mid;
    }
    e = lo;
  }
  return e;
}

// Mann-Whitney / Kruskal-Wallis share this: A = sorted baseline (na), B = sorted current (nb),
// both as order-preserving u32 keys of the fp32 samples.
FM_HD_COLD void fm_mw_lane(const uint32_t *A, int na, const uint32_t *B, int nb, int lane, long long &r2, long long &ties) {
  const int n = na + nb, c = (n + 31) >> 5;
  int p0 = lane * c;
  if (p0 > n) p0 = n;
  int p1 = p0 + c;
  if (p1 > n) p1 = n;
  r2 = 0;
  ties = 0;
  if (p0 >= p1) return;
  // merge path: i = how many baseline samples are among the first p0 pooled (baseline first on ties)
  int lo = p0 > nb ? p0 - nb : 0, hi = p0 < na ? p0 : na;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (A[fm_pidx(mid)] <= B[fm_pidx(p0 - 1 - mid)]) lo = mid + 1; else hi = mid;
  }
  int i = lo, j = p0 - lo;
  if (p0 > 0) {  // a group that started in an earlier span belongs to that lane: skip its tail
    uint32_t prev;
    if (i > 0 && j > 0) {
      uint32_t x = A[fm_pidx(i - 1)], y = B[fm_pidx(j - 1)];
      prev = x > y ? x : y;
    } else if (i > 0) prev = A[fm_pidx(i - 1)];
    else prev = B[fm_pidx(j - 1)];
    i = fm_run_end(A, na, i, prev);
    j = fm_run_end(B, nb, j, prev);
  }
  while (i + j < p1) {
    uint32_t v;
    if (i < na && j < nb) {
      uint32_t x = A[fm_pidx(i)], y = B[fm_pidx(j)];
      v = x < y ? x : y;
    } else if (i < na) v = A[fm_pidx(i)];
    else v = B[fm_pidx(j)];
    const int i2 = fm_run_end(A, na, i, v), j2 = fm_run_end(B, nb, j, v);
    const long long ca = i2 - i, t = (long long)(i2 - i) + (j2 - j), s = (long long)i + j;
    r2 += ca * (2 * s + t + 1);
    ties += t * t * t - t;
    i = i2;
    j = j2;
  }
}

// Fast path of fm_mw_lane for tie-free data (continuous metrics): one balanced merge walk with a
// uniform trip count and no group bookkeeping.  plt accumulates, over the baseline samples taken
// in this lane's span, how many current samples are smaller; without ties
// 2*R1 = na*(na+1) + 2*sum(plt) and the tie term is 0.  Returns true if a baseline and a current
// sample compare equal anywhere in the span (then the caller must use fm_mw_lane instead).
// Both runs must be followed by the 0xFFFFFFFF sentinel (index na / nb readable).
FM_HD bool fm_mw_lane_fast(const uint32_t *A, int na, const uint32_t *B, int nb, int lane, int &plt) {
  const int n = na + nb, c = (n + 31) >> 5;
  int p0 = lane * c;
  if (p0 > n) p0 = n;
  int p1 = p0 + c;
  if (p1 > n) p1 = n;
  plt = 0;
  if (p0 >= p1) return false;
  int lo = p0 > nb ? p0 - nb : 0, hi = p0 < na ? p0 : na;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (A[fm_pidx(mid)] <= B[fm_pidx(p0 - 1 - mid)]) lo = mid + 1; else hi = mid;
  }
  int i = lo, j = p0 - lo;
  uint32_t ha = A[fm_pidx(i)], hb = B[fm_pidx(j)];
  bool tie = false;
  for (int p = p0; p < p1; ++p) {
    tie = tie || (ha == hb);  // both heads real here: at most one run is exhausted before p1
    if (ha <= hb) {           // baseline first on ties
      plt += j;
      ++i;
      ha = A[fm_pidx(i)];
    } else {
      ++j;
      hb = B[fm_pidx(j)];
    }
  }
  return tie;
}

// Wilcoxon signed-rank: K = sorted keys (|d| bits << 1 | (d < 0)), n = number of non-zero pairs.
FM_HD_COLD void fm_w_lane(const uint32_t *K, int n, int lane, long long &wp2, long long &ties) {
  const int c = (n + 31) >> 5;
  int p0 = lane * c;
  if (p0 > n) p0 = n;
  int p1 = p0 + c;
  if (p1 > n) p1 = n;
  wp2 = 0;
  ties = 0;
  if (p0 >= p1) return;
  int s = p0;
  if (p0 > 0) {
    const uint32_t pm = K[fm_pidx(p0 - 1)] >> 1;
    s = fm_run_end<uint32_t>(K, n, s, pm << 1);
    s = fm_run_end<uint32_t>(K, n, s, (pm << 1) | 1u);
  }
  while (s < p1) {
    const uint32_t m = K[fm_pidx(s)] >> 1;
    const int e0 = fm_run_end<uint32_t>(K, n, s, m << 1);          // positive differences first
    const int e = fm_run_end<uint32_t>(K, n, e0, (m << 1) | 1u);
    const long long t = e - s, cp = e0 - s;
    wp2 += cp * (2 * (long long)s + t + 1);
    ties += t * t * t - t;
    s = e;
  }
}
