// fm_rank_merge.h -- merge-path rank bookkeeping of the rank-test kernel (fm_rank.cu).
//
// After the warp has sorted the baseline (A) and current (B) samples separately, the 32 lanes
// split the pooled (merged) order into equal spans with a merge-path search and walk their span
// with a UNIFORM trip count (no per-group branching, no data-dependent loops):
//     plt += #{b < a}   for every baseline sample a taken in the span (baseline first on ties)
// and, only at the first sample of a baseline run that meets an equal current sample,
//     peq += cA*cB,  x3 += cA*cB*(cA+cB)
// with the run lengths cA, cB looked up in small tables the sort pass left beside the keys.
// The exact integer statistics follow from
//     2*R1 = nA(nA+1) + 2*sum(plt) + sum(peq)
//     T    = T_A + T_B + 3*sum(x3)          (T_A, T_B: tie terms inside each sample)
// because (cA+cB)^3-(cA+cB) = (cA^3-cA) + (cB^3-cB) + 3*cA*cB*(cA+cB).
// Plain C++ on arrays in the kernel's padded shared-memory layout, so the same code is
// unit-tested on the host (tests/test_rank_merge_host.py compiles it with g++).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define FM_HD __host__ __device__ __forceinline__
#else
#define FM_HD inline
#endif

// shared-memory index of sorted element e: one pad word per 32 keeps the warp's post-sort
// stores (lane stride R words) and the walk's reads spread over banks
FM_HD int fm_pidx(int e) { return e + (e >> 5); }

// T: one tile holding both sorted runs of order-preserving u32 keys in a single element space --
// baseline element e at T[fm_pidx(e)], current element e at T[fm_pidx(K + e)], K a multiple of 32
// (so fm_pidx(K + e) = fm_pidx(K) + fm_pidx(e)) -- each run followed by a 0xFFFFFFFF sentinel.
// rlA / rlB: run length stored at the index of each run's first element (read only there), or
// nullptr when that sample has no equal values at all (every run has length 1).
FM_HD void fm_mw_walk(const uint32_t *T, int K, int na, int nb, const uint16_t *rlA, const uint16_t *rlB, int lane,
                      int &plt, long long &peq, long long &x3) {
  const int n = na + nb, c = (n + 31) >> 5;
  int p0 = lane * c;
  if (p0 > n) p0 = n;
  int p1 = p0 + c;
  if (p1 > n) p1 = n;
  plt = 0;
  peq = 0;
  x3 = 0;
  if (p0 >= p1) return;
  // merge path: i = how many baseline samples are among the first p0 pooled (baseline first on ties)
  int lo = p0 > nb ? p0 - nb : 0, hi = p0 < na ? p0 : na;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (T[fm_pidx(mid)] <= T[fm_pidx(K + p0 - 1 - mid)]) lo = mid + 1; else hi = mid;
  }
  const int i0 = lo;
  int i = lo, jk = K + p0 - lo;  // heads: baseline element i, current element jk - K
  uint32_t ha = T[fm_pidx(i)], hb = T[fm_pidx(jk)];
  uint32_t pa = i > 0 ? T[fm_pidx(i - 1)] : 0xFFFFFFFFu;  // previous baseline key (sentinel: none)
  // Branch-free step (lanes disagree on which side to take at nearly every step, so an if/else
  // body ran both sides each time): one load at a selected element index of the shared space.
  int acc = 0;  // sum of jk over the baseline samples taken; K * (their count) comes off at the end
  for (int p = p0; p < p1; ++p) {
    const bool take = ha <= hb;  // take a baseline sample; all current samples below it are already taken
    if (take && ha == hb && ha != pa) {  // first of its run and it meets an equal current run (which starts at jk)
      const long long ca = rlA ? rlA[i] : 1, cb = rlB ? rlB[jk - K] : 1;
      peq += ca * cb;
      x3 += ca * cb * (ca + cb);
    }
    acc += take ? jk : 0;
    pa = take ? ha : pa;
    i += take ? 1 : 0;
    jk += take ? 0 : 1;
    const uint32_t nxt = T[fm_pidx(take ? i : jk)];
    ha = take ? nxt : ha;
    hb = take ? hb : nxt;
  }
  plt = acc - K * (i - i0);
}
