// Host harness for foremast_b200/csrc/fm_rank_merge.h: emulates the 32 lanes sequentially and
// prints the reduced statistics.  Input (stdin): cases, then per case
//   na nb  A... B...   (sorted u32 keys)
// Output per case: sum(plt) sum(peq) sum(x3)
#include <cstdio>
#include <vector>
#include "../../foremast_b200/csrc/fm_rank_merge.h"

// both runs in one element space, the second starting at element K (a multiple of 32)
static std::vector<uint32_t> tile(const std::vector<uint32_t> &a, const std::vector<uint32_t> &b, int K) {
  std::vector<uint32_t> o(fm_pidx(2 * K) + 4, 0);
  for (size_t i = 0; i < a.size(); ++i) o[fm_pidx((int)i)] = a[i];
  o[fm_pidx((int)a.size())] = 0xFFFFFFFFu;  // the kernel's runs end with a sentinel
  for (size_t i = 0; i < b.size(); ++i) o[fm_pidx(K + (int)i)] = b[i];
  o[fm_pidx(K + (int)b.size())] = 0xFFFFFFFFu;
  return o;
}
static std::vector<uint16_t> runlen(const std::vector<uint32_t> &a) {  // length at each run head, 0xFFFF elsewhere
  std::vector<uint16_t> r(a.size() + 2, 0xFFFF);
  for (size_t s = 0; s < a.size();) {
    size_t e = s;
    while (e < a.size() && a[e] == a[s]) ++e;
    r[s] = (uint16_t)(e - s);
    s = e;
  }
  return r;
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
    const int K = ((na > nb ? na : nb) + 1 + 31) / 32 * 32;
    auto t = tile(A, B, K);
    auto ra = runlen(A), rb = runlen(B);
    long long plt = 0, peq = 0, x3 = 0;
    for (int lane = 0; lane < 32; ++lane) {
      int a;
      long long b, c;
      fm_mw_walk(t.data(), K, na, nb, ra.data(), rb.data(), lane, a, b, c);
      plt += a; peq += b; x3 += c;
    }
    printf("%lld %lld %lld\n", plt, peq, x3);
  }
  return 0;
}
