// fm_rank_search.h -- Mann-Whitney / Kruskal-Wallis rank bookkeeping WITHOUT sorting the current
// sample (fm_rank.cu, k_mw).
//
// The warp sorts only the baseline sample A (n1 keys) into a linear shared-memory tile.  Every
// current sample b is then located in A by a branch-free lower bound:
//     lt(b) = #{a < b},   eq(b) = #{a == b}
// and the only thing still needed from the current sample B itself is cB(b), how often the value b
// occurs in B -- found by the rounds of fm_dup_resolve_host below (exact: keys are compared, not
// fingerprints).  With S_lt = sum lt(b), S_eq = sum eq(b):
//     2*U1 = 2*n1*n2 - 2*S_lt - S_eq            (pairs b < a count 2, pairs a == b count 1)
//     2*R1 = n1*(n1+1) + 2*U1
//     T    = T_A + sum_b (cB(b)^2 - 1) + 3 * sum_b eq(b) * (eq(b) + cB(b))
// because a pooled tie group with cA baseline and cB current members contributes
// (cA+cB)^3 - (cA+cB) = (cA^3-cA) + (cB^3-cB) + 3*cA*cB*(cA+cB), and summing over the cB members of
// the group gives cB*(cB^2-1) and cB*3*cA*(cA+cB).
// Plain C++ on arrays, so the same code is unit-tested on the host (tests/test_rank_search_host.py
// compiles it with g++).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define FM_HD __host__ __device__ __forceinline__
#else
#ifndef FM_HD
#define FM_HD inline
#endif
#endif

#define FM_KEY_EMPTY 0xFFFFFFFFu

// lower bound over exactly N sorted keys (slots past the real ones hold 0xFFFFFFFF), fixed trip
// count, one load + one compare + one conditional add per step
template <int N>
FM_HD uint32_t fm_lower_bound(const uint32_t *a, uint32_t key) {
  uint32_t pos = 0;
  int len = N;
#pragma unroll
  while (len > 1) {
    const int half = len >> 1;
    pos += a[pos + half] < key ? (uint32_t)half : 0u;
    len -= half;
  }
  return pos + (a[pos] < key ? 1u : 0u);
}

// Slot of `key` in the id table (m slots, any m) of round `round` of the duplicate search below.
FM_HD uint32_t fm_dup_slot(uint32_t key, uint32_t round, uint32_t m) {
  const uint32_t h = (key ^ (round * 0x85EBCA6Bu)) * 0x9E3779B1u;
#ifdef __CUDA_ARCH__
  return __umulhi(h, m);
#else
  return (uint32_t)(((uint64_t)h * m) >> 32);
#endif
}

// Which current samples are equal to each other?  Exact, without sorting and without atomics:
// every key has an id (its index in keys[]).  In a round, all unresolved keys WRITE their id to
// table[slot(key)] (last writer wins), then all READ the slot back:
//   - it holds my id                      -> I represent my value; resolved;
//   - it holds the id of an EQUAL key     -> that key represents me; resolved, rep[me] = its id;
//   - it holds the id of a different key  -> a hash collision; I stay for the next round (new hash).
// Equal keys share the slot in every round, so either one of them wins it (and all the others see
// it at once) or they all lose to a stranger together: a value never ends up with two
// representatives.  The slot winner always resolves, so the rounds terminate.
// This host form (one call = the whole warp's n keys) is what tests/test_rank_search_host.py checks;
// k_mw (fm_rank.cu) runs the same rounds with the keys spread over the lanes' registers.
// rep[i] = i for representatives, else the representative's index.  Returns the number of rounds.
inline int fm_dup_resolve_host(const uint32_t *keys, int n, uint16_t *table, uint32_t m, uint16_t *rep) {
  int pending = n, round = 0;
  for (int i = 0; i < n; ++i) rep[i] = 0xFFFF;
  while (pending > 0) {
    for (int i = 0; i < n; ++i)
      if (rep[i] == 0xFFFF) table[fm_dup_slot(keys[i], (uint32_t)round, m)] = (uint16_t)i;
    for (int i = 0; i < n; ++i) {
      if (rep[i] != 0xFFFF) continue;
      const uint16_t v = table[fm_dup_slot(keys[i], (uint32_t)round, m)];
      if (v == i || keys[v] == keys[i]) {
        rep[i] = v;
        --pending;
      }
    }
    ++round;
  }
  return round;
}

// Tie term of a sorted sample from each element's index i inside its run of equal values:
// t^3 - t = sum_{i<t} 3*i*(i+1), so T = 3 * sum over elements of i*(i+1) -- no run lengths needed.
FM_HD uint32_t fm_tie_step(uint32_t i) { return i * (i + 1u); }

// the statistics of DECISIONS.md "Rank statistics" from the per-sample sums
FM_HD void fm_mw_finish(long long n1, long long n2, long long s_lt, long long s_eq, long long t_a, long long t_b,
                        long long x3, long long *r1x2, long long *u1x2, long long *ties) {
  *u1x2 = 2 * n1 * n2 - 2 * s_lt - s_eq;
  *r1x2 = n1 * (n1 + 1) + *u1x2;
  *ties = t_a + t_b + 3 * x3;
}
