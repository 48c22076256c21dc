// Host harness for foremast_b200/csrc/fm_rank_search.h: the statistics k_mw derives from a sorted
// baseline tile, lower bounds of the current samples and the duplicate rounds over the current keys.
// Input (stdin): cases, then per case   na nb  A(sorted u32 keys)...  B(u32 keys, any order)...
// Output per case: 2R1 2U1 T
#include <cstdio>
#include <vector>
#include "../../foremast_b200/csrc/fm_rank_search.h"

template <int N>
static void one_case(const std::vector<uint32_t> &A, const std::vector<uint32_t> &B) {
  const int na = (int)A.size(), nb = (int)B.size();
  std::vector<uint32_t> tile(N + 4, 0xFFFFFFFFu);
  for (int i = 0; i < na; ++i) tile[i] = A[i];
  // run lengths of A at run heads + T_A (what run_pass leaves)
  std::vector<uint32_t> rl(na + 2, 0);
  long long t_a = 0;
  for (int s = 0; s < na;) {
    int e = s;
    while (e < na && A[e] == A[s]) ++e;
    rl[s] = (uint32_t)(e - s);
    for (int i = 0; i < e - s; ++i) t_a += 3ll * fm_tie_step((uint32_t)i);  // the kernel's per-element form of t^3 - t
    s = e;
  }
  // equal values inside B: representatives + multiplicities by the kernel's rounds
  const uint32_t m = (uint32_t)(N + N / 4 + 37);  // any size >= the key count works; the kernel's is set by its shared-memory budget
  std::vector<uint16_t> table(m, 0), rep(nb + 1, 0);
  fm_dup_resolve_host(B.data(), nb, table.data(), m, rep.data());
  std::vector<uint32_t> cnt(nb + 1, 0);
  for (int i = 0; i < nb; ++i) cnt[rep[i]] += 1;
  long long s_lt = 0, s_eq = 0, x3 = 0, t_b = 0;
  for (int i = 0; i < nb; ++i) {
    const uint32_t k = B[i];
    const uint32_t pos = fm_lower_bound<N>(tile.data(), k);
    const long long eq = tile[pos] == k ? rl[pos] : 0u;
    s_lt += pos;
    s_eq += eq;
    if (rep[i] == i) {  // one contribution per distinct value, as in the kernel
      const long long c = cnt[i];
      x3 += c * eq * (eq + c);
      t_b += c * c * c - c;
    }
  }
  long long r1, u1, t;
  fm_mw_finish(na, nb, s_lt, s_eq, t_a, t_b, x3, &r1, &u1, &t);
  printf("%lld %lld %lld\n", r1, u1, t);
}

int main() {
  int cases;
  if (scanf("%d", &cases) != 1) return 1;
  while (cases--) {
    int na, nb;
    if (scanf("%d %d", &na, &nb) != 2) return 1;
    std::vector<uint32_t> A(na), B(nb);
    for (auto &x : A) if (scanf("%u", &x) != 1) return 1;
    for (auto &x : B) if (scanf("%u", &x) != 1) return 1;
    const int m = na > nb ? na : nb;
    // the tile sizes of the kernel's instantiations that matter for the halving sequence: 2^q and 3*2^q
    if (m <= 32) one_case<32>(A, B);
    else if (m <= 96) one_case<96>(A, B);
    else if (m <= 768) one_case<768>(A, B);
    else if (m <= 1024) one_case<1024>(A, B);
    else one_case<2048>(A, B);
  }
  return 0;
}
