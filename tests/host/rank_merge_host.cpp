// Host harness for foremast_b200/csrc/fm_rank_merge.h: emulates the 32 lanes sequentially and
// prints the reduced statistics.  Input (stdin): "m" na nb  A...  B...  (sorted u32 keys)  or  "w" n K...
#include <cstdio>
#include <vector>
#include "../../foremast_b200/csrc/fm_rank_merge.h"

static std::vector<uint32_t> padu(const std::vector<uint32_t> &a) {
  std::vector<uint32_t> o(a.size() + a.size() / 32 + 4);
  for (size_t i = 0; i < a.size(); ++i) o[fm_pidx((int)i)] = a[i];
  return o;
}
int main() {
  int cases;
  if (scanf("%d", &cases) != 1) return 1;
  while (cases--) {
    char kind[4];
    if (scanf("%3s", kind) != 1) return 1;
    if (kind[0] == 'm') {
      int na, nb;
      if (scanf("%d %d", &na, &nb) != 2) return 1;
      std::vector<uint32_t> A(na), B(nb);
      for (auto &x : A) if (scanf("%u", &x) != 1) return 1;
      for (auto &x : B) if (scanf("%u", &x) != 1) return 1;
      auto pa = padu(A), pb = padu(B);
      pa[fm_pidx(na)] = 0xFFFFFFFFu;  // the kernel's tiles end with a sentinel
      pb[fm_pidx(nb)] = 0xFFFFFFFFu;
      long long r2 = 0, t = 0, fast = 0;
      bool cross = false;
      for (int lane = 0; lane < 32; ++lane) {
        long long a, b;
        fm_mw_lane(pa.data(), na, pb.data(), nb, lane, a, b);
        r2 += a; t += b;
        int plt;
        cross = fm_mw_lane_fast(pa.data(), na, pb.data(), nb, lane, plt) || cross;
        fast += plt;
      }
      // third column: the fast path's 2*R1 (valid only when the data has no ties), or -1 if it saw a cross tie
      printf("%lld %lld %lld\n", r2, t, cross ? -1ll : (long long)na * (na + 1) + 2 * fast);
    } else {
      int n;
      if (scanf("%d", &n) != 1) return 1;
      std::vector<uint32_t> K(n);
      for (auto &x : K) if (scanf("%u", &x) != 1) return 1;
      auto pk = padu(K);
      long long w = 0, t = 0;
      for (int lane = 0; lane < 32; ++lane) {
        long long a, b;
        fm_w_lane(pk.data(), n, lane, a, b);
        w += a; t += b;
      }
      printf("%lld %lld\n", w, t);
    }
  }
  return 0;
}
