// fm_rank.cu -- pairwise rank tests baseline vs current: Mann-Whitney U,
// Wilcoxon signed-rank, Kruskal-Wallis, and the ALL/ANY combiner that decides whether the
// threshold is lowered.
//
// Replaces the brain's "perform pairwise algorithm to check if both have same distribution
// pattern" step (docs/guides/design.md:35; ML_PAIRWISE_ALGORITHM foremast-brain/README.md:34-38;
// MIN_*_DATA_POINTS deploy/foremast/3_brain/foremast-brain.yaml:74-79).
//
// Mapping: ONE WARP PER SERIES-WINDOW.  Each lane keeps R keys in registers (32*R >= segment
// length); the warp sorts them with a register-resident bitonic network -- compare-exchanges
// inside a lane are two VIMNMX, the 15 lane-crossing stages use SHFL.BFLY -- so the sort never
// touches memory.  Two launches: k_wilcoxon sorts one u32 key per pair (|d| with the sign in bit 0)
// and ranks it; k_mw sorts the baseline only and locates every current sample in it by a lower
// bound over a shared-memory tile (fm_rank_search.h), so group membership never rides through a
// sort.  Exact int64 rank sums and tie terms come out of both; k_rank_finalize turns them into
// p-values.  Reads are plain coalesced 4-byte loads / cp.async: segment starts are not 16 B aligned
// in general, and these kernels are bound by instruction issue, not by HBM.
// Algorithmic bytes per unit: 4*(n_base+n_cur) + 8 + 12 + 64.
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include <utility>

#include "fm_common.cuh"
#include "fm_rank_search.h"

namespace {

constexpr int kWarpsPerBlock = 8;


constexpr uint32_t kKeySentinel = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t sel_minmax(bool lower, uint32_t a, uint32_t b) { return lower ? min(a, b) : max(a, b); }
__device__ __forceinline__ void cex(uint32_t &a, uint32_t &b) { uint32_t lo = min(a, b), hi = max(a, b); a = lo; b = hi; }

// order-preserving map of a finite float onto uint32 (x + 0.0f folds -0.0 into +0.0)
__device__ __forceinline__ uint32_t float_key(float x) {
  const uint32_t u = __float_as_uint(x + 0.0f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// ---- register-resident sort of 32*R u32 keys, R = 2^q or 3*2^q ---------------------------------
// Lane `lane` holds v[r] = element lane*R + r.  Three pieces:
//  1. lane_sort: the R keys of each lane, by Batcher's merge-exchange network generated at compile
//     time for exactly R inputs (132 comparators for R = 24, 191 for R = 32);
//  2. five cross-lane merges (2, 4, .. 32 lanes) as a ROLLED loop: a "mirror" step (partner element
//     2L-1-e, so every compare-exchange points the same way and no direction flags are needed)
//     followed by half-cleaners over lane bits -- lane masks are runtime values, register indices
//     stay static, so the whole sort is ~1k instructions and stays in the instruction cache (the
//     fully unrolled network, 12k instructions per kernel, stalled 7 of 8 issue slots on fetch);
//  3. lane_bitonic_merge: each lane's R registers are then a bitonic sequence; half-cleaners while
//     the length is even (valid for any even length), and a 3-sorter per group when R = 3*2^q.
// R = 3*2^q exists because the sort is the kernel's cost and it is paid per SLOT, not per sample:
// 720-sample segments in a 1024-slot network waste 30 % of it, in a 768-slot one 6 %.
template <int R>
struct LaneNet {
  int n = 0;
  unsigned char a[R * 9 + 1] = {}, b[R * 9 + 1] = {};
};

template <int R>
__host__ __device__ constexpr LaneNet<R> make_lane_net() {
  LaneNet<R> t{};
  for (int p = 1; p < R; p *= 2)
    for (int k = p; k >= 1; k /= 2)
      for (int j = k % p; j <= R - 1 - k; j += 2 * k)
        for (int i = 0; i <= (k - 1 < R - j - k - 1 ? k - 1 : R - j - k - 1); ++i)
          if ((i + j) / (2 * p) == (i + j + k) / (2 * p)) {
            t.a[t.n] = (unsigned char)(i + j);
            t.b[t.n] = (unsigned char)(i + j + k);
            ++t.n;
          }
  return t;
}

template <int R, int... I>
__device__ __forceinline__ void lane_sort_impl(uint32_t (&v)[R], std::integer_sequence<int, I...>) {
  constexpr LaneNet<R> net = make_lane_net<R>();
  (cex(v[net.a[I]], v[net.b[I]]), ...);
}

template <int R>
__device__ __forceinline__ void lane_sort(uint32_t (&v)[R]) {
  constexpr int n = make_lane_net<R>().n;
  if constexpr (n > 0) lane_sort_impl<R>(v, std::make_integer_sequence<int, n>{});
}

__host__ __device__ constexpr int odd_part(int r) { return (r & 1) ? r : odd_part(r >> 1); }
__host__ __device__ constexpr int halvings(int r) { return (r & 1) ? 0 : 1 + halvings(r >> 1); }

template <int R>
__device__ __forceinline__ void lane_bitonic_merge(uint32_t (&v)[R]) {
  constexpr int kHalvings = halvings(R);
#pragma unroll
  for (int s = 0; s < kHalvings; ++s) {
    const int len = R >> s, half = len >> 1;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (r % len < half) cex(v[r], v[r + half]);
  }
  if constexpr (odd_part(R) == 3) {
#pragma unroll
    for (int g = 0; g < R; g += 3) {  // 3-sorter in four ALU ops: min3, max3, and the middle by subtraction (mod 2^32)
      const uint32_t lo = __vimin3_u32(v[g], v[g + 1], v[g + 2]), hi = __vimax3_u32(v[g], v[g + 1], v[g + 2]);
      const uint32_t mid = v[g] + v[g + 1] + v[g + 2] - lo - hi;
      v[g] = lo;
      v[g + 1] = mid;
      v[g + 2] = hi;
    }
  }
}

template <int R>
__device__ __forceinline__ void warp_bitonic_sort(uint32_t (&v)[R], const unsigned lane) {
  static_assert(odd_part(R) == 1 || odd_part(R) == 3, "R must be 2^q or 3*2^q");
  lane_sort<R>(v);
  // rolled for the long networks (instruction cache); the short ones (up to 256 slots) are small enough to be laid
  // out flat, which saves the loop control and the run-time lane masks of their fifteen cross-lane steps
  constexpr int kFlat = R <= 8 ? 5 : 1;
#pragma unroll kFlat
  for (unsigned kk = 1; kk <= 16; kk <<= 1) {  // merges 2*kk lanes, L = R*kk elements a side
    {
      const unsigned lm = 2 * kk - 1;
      const bool lower = (lane & kk) == 0;
#pragma unroll
      for (int r = 0; r < R / 2; ++r) {
        const uint32_t a = __shfl_xor_sync(0xffffffffu, v[R - 1 - r], lm);
        const uint32_t c = __shfl_xor_sync(0xffffffffu, v[r], lm);
        v[r] = sel_minmax(lower, v[r], a);
        v[R - 1 - r] = sel_minmax(lower, v[R - 1 - r], c);
      }
      if constexpr (R & 1) {  // the middle register mirrors onto itself
        const uint32_t a = __shfl_xor_sync(0xffffffffu, v[R / 2], lm);
        v[R / 2] = sel_minmax(lower, v[R / 2], a);
      }
    }
#pragma unroll kFlat
    for (unsigned jj = kk >> 1; jj >= 1; jj >>= 1) {
      const bool lower = (lane & jj) == 0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t a = __shfl_xor_sync(0xffffffffu, v[r], jj);
        v[r] = sel_minmax(lower, v[r], a);
      }
    }
    lane_bitonic_merge<R>(v);
  }
}

// Tie bookkeeping of a sorted run that is still in registers (lane l holds elements l*R .. l*R+R-1, the
// first n are real; keys compare equal after dropping their low `SH` bits).  Every element learns the first
// and the last index of its run of equal keys -- a forward sweep with a max-scan of run heads over the lanes,
// a backward sweep with a min-scan of run tails -- and contributes
//   * fm_tie_step(i) = i*(i+1) to the tie sum, i = its index inside the run (3 * the sum = sum of t^3 - t);
//   * if it is `counted` (Wilcoxon: a positive difference, sign bit 0 clear): its twice-average-rank
//     2s + t + 1 = first + last + 2.
// Only units that have ties come here.
template <int R, int SH, bool RANKS>
__device__ __forceinline__ void tie_sums(const uint32_t (&key)[R], const unsigned lane, const uint32_t n, uint32_t &tie_acc,
                                         long long &rank2_acc) {
  const int base = (int)(lane * R), rem = (int)n - base;  // real keys in this lane's slots (may be <= 0 or >= R)
  const uint32_t nxt = __shfl_down_sync(0xffffffffu, key[0], 1), prv = __shfl_up_sync(0xffffffffu, key[R - 1], 1);
  // forward: last run head inside this lane, carried over the lanes
  int last_head = -1;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const bool head = r < rem && (r == 0 ? (lane == 0 || (key[0] >> SH) != (prv >> SH)) : (key[r] >> SH) != (key[r - 1] >> SH));
    if (head) last_head = base + r;
  }
  int carry = last_head;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, carry, o);
    if ((int)lane >= o) carry = max(carry, t);
  }
  int start = __shfl_up_sync(0xffffffffu, carry, 1);  // lane 0 never reads it: its first element is a head
  // backward: first run tail inside this lane, carried back over the lanes
  int first_tail = 0x7FFFFFFF;
  if (RANKS) {
#pragma unroll
    for (int r = R - 1; r >= 0; --r) {
      const bool tail = r < rem && (base + r + 1 == (int)n || (r + 1 < R ? (key[r + 1] >> SH) != (key[r] >> SH) : (nxt >> SH) != (key[R - 1] >> SH)));
      if (tail) first_tail = base + r;
    }
    int back = first_tail;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_down_sync(0xffffffffu, back, o);
      if ((int)lane + o < 32) back = min(back, t);
    }
    first_tail = __shfl_down_sync(0xffffffffu, back, 1);  // the nearest tail in the lanes after this one
  }
  // per element: run start so far (forward), run end (from the element's own lane backwards, else the carry)
  int ends[R];
  if (RANKS) {
    int end = first_tail;
#pragma unroll
    for (int r = R - 1; r >= 0; --r) {
      const bool tail = r < rem && (base + r + 1 == (int)n || (r + 1 < R ? (key[r + 1] >> SH) != (key[r] >> SH) : (nxt >> SH) != (key[R - 1] >> SH)));
      if (tail) end = base + r;
      ends[r] = end;
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (r < rem) {
      const int e = base + r;
      const bool head = r == 0 ? (lane == 0 || (key[0] >> SH) != (prv >> SH)) : (key[r] >> SH) != (key[r - 1] >> SH);
      if (head) start = e;
      tie_acc += fm_tie_step((uint32_t)(e - start));
      if (RANKS && !(key[r] & 1u)) rank2_acc += (long long)(start + ends[r] + 2);
    }
  }
}

// p-values: same expressions, same order as oracle/DECISIONS.md "p-values".
__device__ __noinline__ double p_mann_whitney(long long u1x2, uint32_t n1, uint32_t n2, long long ties) {
  double nn = (double)n1 * (double)n2, N = (double)n1 + (double)n2;
  double var = nn / 12.0 * ((N + 1.0) - (double)ties / (N * (N - 1.0)));
  if (!(var > 0.0)) return 1.0;
  long long u2x2 = 2ll * n1 * n2 - u1x2;
  long long umax = u1x2 > u2x2 ? u1x2 : u2x2;
  double z = ((double)(umax - (long long)n1 * n2) * 0.5 - 0.5) / sqrt(var);
  double p = erfc(z * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

__device__ __noinline__ double p_kruskal(long long r1x2, uint32_t n1, uint32_t n2, long long ties, double *h_out) {
  double N = (double)n1 + (double)n2;
  long long Ni = (long long)n1 + n2;
  long long r2x2 = Ni * (Ni + 1) - r1x2;
  double r1 = 0.5 * (double)r1x2, r2 = 0.5 * (double)r2x2;
  double c = 1.0 - (double)ties / (N * N * N - N);
  if (!(c > 0.0)) { *h_out = 0.0; return 1.0; }
  double h = 12.0 / (N * (N + 1.0)) * (r1 * r1 / (double)n1 + r2 * r2 / (double)n2) - 3.0 * (N + 1.0);
  h = h / c;
  if (!(h > 0.0)) h = 0.0;
  *h_out = h;
  return erfc(sqrt(h * 0.5));
}

__device__ __noinline__ double p_wilcoxon(long long wp2, long long wm2, uint32_t n, long long ties) {
  if (n == 0) return 1.0;
  double nd = (double)n;
  double se2 = nd * (nd + 1.0) * (2.0 * nd + 1.0) / 24.0 - (double)ties / 48.0;
  if (!(se2 > 0.0)) return 1.0;
  long long tmin = wp2 < wm2 ? wp2 : wm2;
  double z = ((double)tmin * 0.5 - nd * (nd + 1.0) / 4.0) / sqrt(se2);
  double p = erfc(fabs(z) * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

// Wilcoxon signed-rank pass: one warp per unit, one sort of |d| keys (sign in bit 0).  The NEXT unit's
// paired samples are staged in shared memory with 4-byte cp.async while the current unit is being
// sorted in registers; ranks and the tie term are read off the sorted registers (tie_sums when there are ties).  A separate launch from the Mann-Whitney
// pass: each keeps its instruction footprint inside the instruction cache (one fused kernel lost 4-5
// issue slots per issued instruction to fetch stalls), and both are bound by instruction issue, so the
// second read of the samples (from L2/HBM) costs nothing measurable.
// The units a launch covers: all of the batch (list == nullptr), or the units of ONE sort-size class of a batch that has
// several -- list holds the unit indices grouped by class, hist[c] the class sizes (fm_rank_classes below), so a batch
// of short windows with a few long ones among them does not sort every unit in the long ones' network.
struct UnitSet {
  const uint32_t *list;
  const uint32_t *hist;
  int cls;
};
template <bool LISTED>  // a template parameter of the kernels: the all-units form must not pay for the other (it did: + 8 %)
__device__ __forceinline__ uint64_t unit_set_begin(const UnitSet &us, const DevBatch &b, const uint32_t *&lst) {
  if (!LISTED) { lst = nullptr; return b.n_units; }
  uint32_t off = 0;
  for (int c = 0; c < us.cls; ++c) off += us.hist[c];
  lst = us.list + off;
  return us.hist[us.cls];
}

template <int R, bool LISTED>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, R > 32 ? 1 : 3) k_wilcoxon(DevBatch b, const DevParams *__restrict__ dp,
                                                                                  fm_series_score *__restrict__ scores,
                                                                                  const int raw_cap, const UnitSet us) {
  extern __shared__ uint32_t s_tiles[];  // [warps][2][raw_cap] staged samples (the sorted keys never leave the registers)
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  const uint32_t pairwise = dp->pairwise;
  const bool want_w = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_WILCOXON;
  float *const rawA = reinterpret_cast<float *>(s_tiles) + (size_t)warp * 2 * raw_cap;
  float *const rawB = rawA + raw_cap;
  const uint32_t *lst;
  const uint64_t n_set = unit_set_begin<LISTED>(us, b, lst);  // positions 0 .. n_set-1 of this launch; unit = lst ? lst[pos] : pos
  auto stage_pairs = [&](uint64_t pos) {
    if (pos < n_set) {
      const uint64_t uu = LISTED ? lst[pos] : pos;
      const int64_t o = b.offsets[uu];
      const uint32_t h0 = b.seg_lens[3 * uu], n1 = b.seg_lens[3 * uu + 1], n2 = b.seg_lens[3 * uu + 2];
      const uint32_t mm = min(min(n1, n2), (uint32_t)raw_cap);
      const float *pa = b.values + o + h0, *pb = pa + n1;
      for (uint32_t e = lane; e < mm; e += 32) {
        cp_async4(rawA + e, pa + e);
        cp_async4(rawB + e, pb + e);
      }
    }
    cp_async_commit();
  };
  stage_pairs((uint64_t)blockIdx.x * kWarpsPerBlock + warp);

  // Rounds: in round i warp w of block c owns unit (i*gridDim + c)*kWarpsPerBlock + w.
  const uint64_t n_rounds = (n_set + (uint64_t)gridDim.x * kWarpsPerBlock - 1) / ((uint64_t)gridDim.x * kWarpsPerBlock);
  for (uint64_t round = 0; round < n_rounds; ++round) {
    const uint64_t pos = (round * gridDim.x + blockIdx.x) * kWarpsPerBlock + warp;
    const bool have = pos < n_set;
    const uint64_t u = LISTED && have ? lst[pos] : pos;
    uint32_t nb = 0, nc = 0;
    if (have) { nb = b.seg_lens[3 * u + 1]; nc = b.seg_lens[3 * u + 2]; }
    const uint32_t m = min(nb, nc);  // the pairs are the first m samples of each segment (D8)
    // D17: at most FM_RANK_MAX_SEGMENT pairs are ranked; the launch picks R so that every ranked unit fits
    const bool run_w = have && want_w && m >= dp->min_wilcoxon && m > 0 && m <= 32u * R && m <= (uint32_t)FM_RANK_MAX_SEGMENT;
    long long w_2wp = 0, w_ties = 0;
    uint32_t w_n = 0;
    cp_async_wait<0>();
    __syncwarp();
    if (!run_w) {  // nothing to rank here, but the staging pipeline must keep moving
      stage_pairs(pos + (uint64_t)gridDim.x * kWarpsPerBlock);
    } else {
      uint32_t key[R];
      uint32_t nz = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t e = r * 32 + lane;
        uint32_t k = kKeySentinel;
        if (e < m) {
          // |d| bits << 1 with the sign in bit 0 is a rotate; +-0 rotate to 0 / 1
          const uint32_t du = __float_as_uint(rawB[e] - rawA[e]), rot = __funnelshift_l(du, du, 1);
          if (rot > 1u) {
            k = rot;
            ++nz;
          }
        }
        key[r] = k;
      }
      const uint32_t n_valid = w_n = __reduce_add_sync(0xffffffffu, nz);
      __syncwarp();  // every lane has its keys in registers: the staging buffers are free again
      stage_pairs(pos + (uint64_t)gridDim.x * kWarpsPerBlock);
      warp_bitonic_sort<R>(key, lane);

      // ties between neighbours of the sorted run (the lane holds elements lane*R .. lane*R+R-1); the keys
      // carry the sign in bit 0.  Slots of this lane that hold real samples are compared against literals
      // below: a bound written as lane*R + r < n_valid gets hoisted into R loop-invariant registers, and
      // those spilled
      const int rem = (int)n_valid - (int)(lane * R);
      bool tie = false;
#pragma unroll
      for (int r = 0; r + 1 < R; ++r) tie = tie || (((key[r] >> 1) == (key[r + 1] >> 1)) && (r + 1 < rem));
      const uint32_t nxt = __shfl_down_sync(0xffffffffu, key[0], 1);
      tie = tie || (lane < 31 && ((key[R - 1] >> 1) == (nxt >> 1)) && (R < rem));
      const bool any_tie = __any_sync(0xffffffffu, tie);
      long long acc = 0;
      uint32_t tie_acc = 0;
      if (!any_tie) {
        // tie-free |d|: the rank of sorted position e is e+1
        int a32 = 0, npos = 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (r < rem && !(key[r] & 1u)) { a32 += 2 * (r + 1); ++npos; }
        acc = a32 + 2 * (int)(lane * R) * npos;
      } else {
        tie_sums<R, 1, true>(key, lane, n_valid, tie_acc, acc);
      }
      w_2wp = warp_sum_i64(acc);
      w_ties = any_tie ? 3 * warp_sum_i64((long long)tie_acc) : 0;
    }
    // exact integer statistics out; p-values and the combiner run one thread per unit in k_rank_finalize
    if (lane == 0 && have) {
      fm_series_score *o = scores + u;
      o->status = run_w ? FM_S_WILCOXON_RUN : 0u;  // first writer (scores start zeroed)
      o->w_n = w_n;
      o->w_2wplus = w_2wp;
      o->w_2wminus = run_w ? (long long)w_n * ((long long)w_n + 1) - w_2wp : 0;
      o->w_ties = w_ties;
    }
  }
}


// R interleaved lower bounds over a linear tile of N sorted keys in shared memory (the algorithm of
// fm_lower_bound<N>, fm_rank_search.h).  adr[r] is the running lower bound of key[r] as a byte address in
// the shared window; every step is one LDS at a literal offset, one compare and one predicated add per
// key.  The loads are volatile asm so that the R independent chains stay interleaved as written (left to
// itself the compiler ran the searches one after the other, 11 exposed shared-memory round trips each).
template <int R, int LEN>
struct LowerBound {
  static __device__ __forceinline__ void run(uint32_t (&adr)[R], const uint32_t (&key)[R]) {
    constexpr int half = LEN > 1 ? LEN / 2 : 0;  // LEN == 1: the closing comparison, offset 0, step 1
    uint32_t v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v[r]) : "r"(adr[r]), "n"(half * 4));
#pragma unroll
    for (int r = 0; r < R; ++r)
      asm("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %1, %2;\n\t@p add.u32 %0, %0, %3;\n\t}"
          : "+r"(adr[r]) : "r"(v[r]), "r"(key[r]), "n"(LEN > 1 ? half * 4 : 4));
    if constexpr (LEN > 1) LowerBound<R, LEN - half>::run(adr, key);
  }
};

// ---- Mann-Whitney / Kruskal-Wallis pass: sort the baseline, SEARCH the current samples ---------------
// One warp per unit.  The current samples are staged raw by cp.async, turned into keys in place (id =
// element index, lane l owns ids l*R .. l*R+R-1), and equal values among them are found exactly by the
// write-then-read rounds of fm_rank_search.h (no sort, no atomics; the baseline's loads are in flight
// meanwhile).  The baseline's 32*R key slots (real keys, then 0xFFFFFFFF) are then sorted in registers;
// if it has ties, each element's index inside its run gives the tie term and the run length is left at
// the run's first index; the sorted keys go to a LINEAR shared-memory tile with 128-bit stores.  Finally
// every current key is located in the tile by a fixed-trip lower bound: lt = #{a < b}, eq = #{a == b}
// (the run length; 1 when the baseline is tie-free).  A value's representative carries its whole group
// with weight c = its multiplicity among the current samples, the other members weigh 0:
//     S_lt += c*lt, S_eq += c*eq, X3 += c*eq*(eq+c), T_B += c^3 - c
// Everything outside the sort runs in rolled loops over chunks of CH keys, so the kernel stays inside the
// instruction cache and needs few registers.  Compared with sorting both samples and walking the merge:
// no second sort (1504 min/max per lane), no merge-path walk, and the searches run on the load/store
// pipe while other warps sort on the ALU pipe.
// Shared memory per warp: tile u32[N+4] | current keys u32[key_words] (raw floats first; the
// representative's id for keys that turn out to have one) | id table u16[n_slots] of the rounds, whose
// bytes afterwards hold the per-id counters u16[N] and the baseline's run lengths u16[rl_entries].
template <int R, int W, bool LISTED>
__global__ void __launch_bounds__(W * 32, R <= 24 ? 3 : 1) k_mw(DevBatch b, const DevParams *__restrict__ dp,
                                                                fm_series_score *__restrict__ scores, const int rl_entries,
                                                                const int key_words, const int n_slots, const UnitSet us) {
  constexpr int N = 32 * R;
  constexpr int kTileWords = N + 4;
  constexpr int CH = R % 8 == 0 ? 8 : (R <= 6 ? R : 4);  // keys per chunk of the rolled loops
  constexpr int NCH = R / CH;
  static_assert(R % CH == 0, "chunks tile the lane's keys");
  extern __shared__ __align__(16) uint32_t s_mw[];
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  const int tab_words = (n_slots + 7) / 8 * 4;  // u16 slots, rounded to 16 bytes
  uint32_t *const tile = s_mw + (size_t)warp * (kTileWords + key_words + tab_words);
  uint32_t *const keysB = tile + kTileWords;  // key_words = the batch's longest segment rounded up to a chunk
  uint16_t *const tabH = reinterpret_cast<uint16_t *>(keysB + key_words);
  uint32_t *const cntW = keysB + key_words;  // once the rounds are over the table's bytes hold u16 counters per id
  const uint16_t *const cntB = tabH;         // (incremented two to a word) ...
  uint16_t *const rlA = tabH + N;            // ... and the run lengths of the baseline (host: n_slots >= N + rl_entries)
  (void)rl_entries;
  const uint32_t pairwise = dp->pairwise;
  const bool want_mw = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_MANN_WHITE;
  const bool want_k = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_KRUSKAL;
  const uint32_t min_mw = dp->min_mann_white, min_k = dp->min_kruskal;
  const uint32_t id0 = lane * R, m_slots = (uint32_t)n_slots;
  using Mask = typename std::conditional<(R > 32), unsigned long long, uint32_t>::type;
  auto mask_ffs = [](Mask m) { return sizeof(Mask) == 8 ? __ffsll((long long)m) : __ffs((int)m); };
  auto load_chunk = [&](const uint32_t *src, uint32_t (&k)[CH]) {
    if constexpr (CH % 4 == 0) {
#pragma unroll
      for (int i = 0; i < CH; i += 4) {
        const uint4 q = *reinterpret_cast<const uint4 *>(src + i);
        k[i] = q.x; k[i + 1] = q.y; k[i + 2] = q.z; k[i + 3] = q.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < CH; ++i) k[i] = src[i];
    }
  };
  auto store_chunk = [&](uint32_t *dst, const uint32_t (&k)[CH]) {
    if constexpr (CH % 4 == 0) {
#pragma unroll
      for (int i = 0; i < CH; i += 4) *reinterpret_cast<uint4 *>(dst + i) = make_uint4(k[i], k[i + 1], k[i + 2], k[i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < CH; ++i) dst[i] = k[i];
    }
  };

  const uint64_t stride_units = (uint64_t)gridDim.x * W;
  const uint32_t *lst;
  const uint64_t n_set = unit_set_begin<LISTED>(us, b, lst);
  for (uint64_t pos = (uint64_t)blockIdx.x * W + warp; pos < n_set; pos += stride_units) {
    const uint64_t u = LISTED ? lst[pos] : pos;
    const int64_t off = b.offsets[u];
    const uint32_t nh = b.seg_lens[3 * u], nb = b.seg_lens[3 * u + 1], nc = b.seg_lens[3 * u + 2];
    {  // pull the next round's samples towards L2 while this round works
      const uint64_t pn = pos + stride_units;
      if (pn < n_set) {
        const uint64_t un = LISTED ? lst[pn] : pn;
        const int64_t offn = b.offsets[un];
        const uint32_t nhn = b.seg_lens[3 * un], len = b.seg_lens[3 * un + 1] + b.seg_lens[3 * un + 2];
        const char *p = reinterpret_cast<const char *>(b.values + offn + nhn);
        for (uint32_t byte = lane * 128; byte < len * 4 + 128; byte += 32 * 128)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p + byte));
      }
    }
    const float *base = b.values + off + nh, *cur = base + nb;
    // D17: segments longer than FM_RANK_MAX_SEGMENT are not ranked; the launch picks R so that all others fit
    const bool fits = nb <= (uint32_t)N && nc <= (uint32_t)N && nb > 0 && nc > 0 && nb <= (uint32_t)FM_RANK_MAX_SEGMENT &&
                      nc <= (uint32_t)FM_RANK_MAX_SEGMENT;
    const bool run_mw = want_mw && nb >= min_mw && nc >= min_mw && fits;
    const bool run_k = want_k && nb >= min_k && nc >= min_k && fits;
    long long r1x2 = 0, u1x2 = 0, ties = 0;
    if (run_mw || run_k) {  // warp-uniform
      // ---- both samples requested: the current one raw into shared memory, the baseline into registers
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t e = (uint32_t)(r * 32) + lane;
        if (e < nc) cp_async4(keysB + e, cur + e);
      }
      cp_async_commit();
      float x[R];
      {
        const uint32_t last = nb - 1;
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = __ldg(base + min((uint32_t)(r * 32) + lane, last));
      }
      cp_async_wait<0>();
      __syncwarp();
      // ---- raw samples -> keys in place; round 0 of the duplicate search: every key writes its id
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const uint32_t idc = id0 + (uint32_t)(c * CH);
        if (idc >= nc) break;  // ids from here on hold no sample (and may lie past the key region)
        uint32_t k[CH];
        load_chunk(keysB + idc, k);
#pragma unroll
        for (int i = 0; i < CH; ++i) {
          const bool valid = idc + i < nc;
          k[i] = valid ? float_key(__uint_as_float(k[i])) : kKeySentinel;
          if (valid) tabH[fm_dup_slot(k[i], 0u, m_slots)] = (uint16_t)(idc + i);
        }
        store_chunk(keysB + idc, k);
      }
      __syncwarp();
      // ---- round 0: read.  pend / nonrep: one bit per key slot r of this lane
      Mask pend = 0, nonrep = 0;
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const uint32_t idc = id0 + (uint32_t)(c * CH);
        if (idc >= nc) break;
        uint32_t k[CH];
        load_chunk(keysB + idc, k);
        Mask p = 0, q = 0;
#pragma unroll
        for (int i = 0; i < CH; ++i) {
          if (idc + i < nc) {
            const uint32_t v = tabH[fm_dup_slot(k[i], 0u, m_slots)];
            if (v != idc + i) {
              if (keysB[v] == k[i]) { q |= (Mask)1 << i; keysB[idc + i] = v; }
              else p |= (Mask)1 << i;
            }
          }
        }
        pend |= p << (c * CH);
        nonrep |= q << (c * CH);
      }
      __syncwarp();
      // ---- further rounds for the few keys that lost their slot to a different value
      for (uint32_t round = 1; __any_sync(0xffffffffu, pend != 0); ++round) {
        for (Mask m = pend; m; m &= m - 1) {
          const uint32_t id = id0 + (uint32_t)(mask_ffs(m) - 1);
          tabH[fm_dup_slot(keysB[id], round, m_slots)] = (uint16_t)id;
        }
        __syncwarp();
        for (Mask m = pend; m; m &= m - 1) {
          const int r = mask_ffs(m) - 1;
          const uint32_t id = id0 + (uint32_t)r, k = keysB[id];
          const uint32_t v = tabH[fm_dup_slot(k, round, m_slots)];
          if (v == id) pend &= ~((Mask)1 << r);
          else if (keysB[v] == k) { pend &= ~((Mask)1 << r); nonrep |= (Mask)1 << r; keysB[id] = v; }
        }
        __syncwarp();
      }
      const bool dup_b = __any_sync(0xffffffffu, nonrep != 0);
      if (dup_b) {  // multiplicity of every represented value: counters over the (now dead) id table
        for (int i = lane; i < N / 2; i += 32) cntW[i] = 0u;
        __syncwarp();
        for (Mask m = nonrep; m; m &= m - 1) {
          const uint32_t rep = keysB[id0 + (uint32_t)(mask_ffs(m) - 1)];
          atomicAdd(cntW + (rep >> 1), 1u << ((rep & 1u) * 16u));
        }
      }
      // ---- baseline: keys, sort, ties, linear tile
      bool tie_a;
      uint32_t tie_acc = 0;  // sum over this lane's elements of i*(i+1), i = index inside the run of equal keys
      {
        uint32_t key[R];
#pragma unroll
        for (int r = 0; r < R; ++r) key[r] = (uint32_t)(r * 32) + lane < nb ? float_key(x[r]) : kKeySentinel;
        warp_bitonic_sort<R>(key, lane);
        const int rem = (int)nb - (int)(lane * R);  // real keys in this lane's slots (may be <= 0 or >= R)
        const uint32_t nxt = __shfl_down_sync(0xffffffffu, key[0], 1), prv = __shfl_up_sync(0xffffffffu, key[R - 1], 1);
        bool tie = false;
#pragma unroll
        for (int r = 0; r + 1 < R; ++r) tie = tie || ((key[r] == key[r + 1]) && (r + 1 < rem));
        tie = tie || (lane < 31 && key[R - 1] == nxt && R < rem);
        tie_a = __any_sync(0xffffffffu, tie);
        if (tie_a) {
          // run starts: the last one inside this lane, then the nearest one in the lanes before (max-scan)
          int last_head = -1;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const bool head = r < rem && (r == 0 ? (lane == 0 || key[0] != prv) : key[r] != key[r - 1]);
            if (head) last_head = (int)id0 + r;
          }
          int carry = last_head;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, carry, o);
            if ((int)lane >= o) carry = max(carry, t);
          }
          int start = __shfl_up_sync(0xffffffffu, carry, 1);  // lane 0 never reads it: its first element is a head
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if (r < rem) {
              const int e = (int)id0 + r;
              const bool head = r == 0 ? (lane == 0 || key[0] != prv) : key[r] != key[r - 1];
              if (head) start = e;
              const uint32_t i = (uint32_t)(e - start);
              tie_acc += fm_tie_step(i);
              const bool last = e + 1 == (int)nb || (r + 1 < R ? key[r + 1] != key[r] : nxt != key[R - 1]);
              if (last) rlA[start] = (uint16_t)(i + 1u);
            }
          }
        }
        if constexpr (R % 4 == 0) {
#pragma unroll
          for (int r = 0; r < R; r += 4)
            *reinterpret_cast<uint4 *>(tile + id0 + r) = make_uint4(key[r], key[r + 1], key[r + 2], key[r + 3]);
        } else {
#pragma unroll
          for (int r = 0; r < R; ++r) tile[id0 + r] = key[r];
        }
        if (lane == 0) tile[N] = kKeySentinel;
      }
      __syncwarp();
      // ---- lower bounds of CH keys at a time, weights, sums
      const uint32_t tile_s = (uint32_t)__cvta_generic_to_shared(tile);
      uint32_t s_lt = 0, s_eq = 0;
      long long x3 = 0, t_b = 0;
      const bool plain = !tie_a && !dup_b;  // eq is 0 or 1 and every value of the current sample occurs once
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const uint32_t idc = id0 + (uint32_t)(c * CH);
        if (idc >= nc) break;
        uint32_t k[CH], adr[CH];
        load_chunk(keysB + idc, k);
#pragma unroll
        for (int i = 0; i < CH; ++i) adr[i] = tile_s;
        LowerBound<CH, N>::run(adr, k);
        const uint32_t skip = (uint32_t)(nonrep >> (c * CH));  // members represented by another key weigh 0
        if (plain) {
#pragma unroll
          for (int i = 0; i < CH; ++i) {
            const bool w = idc + i < nc;
            uint32_t a;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(adr[i]));
            s_lt += w ? (adr[i] - tile_s) >> 2 : 0u;
            s_eq += (w && a == k[i]) ? 1u : 0u;
          }
        } else {
#pragma unroll
          for (int i = 0; i < CH; ++i) {
            const uint32_t pos = (adr[i] - tile_s) >> 2;
            const bool w = idc + i < nc && !((skip >> i) & 1u);
            const uint32_t cw = w ? (dup_b ? (uint32_t)cntB[idc + i] + 1u : 1u) : 0u;
            const uint32_t eq = tile[pos] == k[i] ? (tie_a ? (uint32_t)rlA[pos] : 1u) : 0u;
            s_lt += cw * pos;
            s_eq += cw * eq;
            x3 += (long long)((unsigned long long)(eq * (eq + cw)) * cw);
            t_b += (long long)((unsigned long long)(cw * cw) * cw) - (long long)cw;
          }
        }
      }
      if (plain) x3 = 2ll * s_eq;
      const long long lt_sum = warp_sum_i64((long long)s_lt), eq_sum = warp_sum_i64((long long)s_eq);
      const long long x3_sum = warp_sum_i64(x3), tb_sum = dup_b ? warp_sum_i64(t_b) : 0;
      const long long ta_sum = tie_a ? 3 * warp_sum_i64((long long)tie_acc) : 0;
      fm_mw_finish((long long)nb, (long long)nc, lt_sum, eq_sum, ta_sum, tb_sum, x3_sum, &r1x2, &u1x2, &ties);
      __syncwarp();  // tile, keys and table are rewritten by the next unit
    }
    if (lane == 0) {
      fm_series_score *o = scores + u;
      // fire-and-forget OR (the Wilcoxon launch, or the zero fill, wrote the word earlier in the stream)
      if (run_mw || run_k) atomicOr(&o->status, (run_mw ? FM_S_MW_RUN : 0u) | (run_k ? FM_S_KRUSKAL_RUN : 0u));
      o->mw_2u1 = u1x2;
      o->mw_2r1 = r1x2;
      o->mw_ties = ties;
    }
  }
}

// One thread per unit: normal-approximation p-values and the ALL/ANY combiner
// (ML_PAIRWISE_ALGORITHM / ML_PAIRWISE_THRESHOLD, foremast-brain/README.md:34-35).
__global__ void __launch_bounds__(256) k_rank_finalize(DevBatch b, const DevParams *__restrict__ dp,
                                                        fm_series_score *__restrict__ scores) {
  const uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= b.n_units) return;
  fm_series_score *o = scores + u;
  const uint32_t st = o->status, nb = b.seg_lens[3 * u + 1], nc = b.seg_lens[3 * u + 2];
  const bool run_mw = st & FM_S_MW_RUN, run_w = st & FM_S_WILCOXON_RUN, run_k = st & FM_S_KRUSKAL_RUN;
  double p_mw = 1.0, p_w = 1.0, p_k = 1.0, kw_h = 0.0;
  if (run_mw) p_mw = p_mann_whitney(o->mw_2u1, nb, nc, o->mw_ties);
  if (run_w) p_w = p_wilcoxon(o->w_2wplus, o->w_2wminus, o->w_n, o->w_ties);
  if (run_k) p_k = p_kruskal(o->mw_2r1, nb, nc, o->mw_ties, &kw_h);
  const double alpha = dp->pairwise_threshold;
  const uint32_t pairwise = dp->pairwise;
  const int n_run = (int)run_mw + (int)run_w + (int)run_k;
  const int n_rej = (int)(run_mw && p_mw < alpha) + (int)(run_w && p_w < alpha) + (int)(run_k && p_k < alpha);
  bool differs = false;
  if (pairwise == FM_PW_ALL) differs = n_run > 0 && n_rej == n_run;
  else if (pairwise != FM_PW_NONE) differs = n_rej > 0;
  if (differs) o->status = st | FM_S_DIFFERS;
  o->p_mw = p_mw;
  o->p_w = p_w;
  o->p_k = p_k;
  o->kw_h = kw_h;
}

constexpr uint64_t kBlocksPerSm = 64;  // many more blocks than are resident: tie-heavy units take longer, the block
                                       // scheduler evens it out (a persistent 5-blocks-per-SM grid measured 19 % slower)

// ---- grouping the units of a mixed batch by sort-size class -----------------------------------------
// hist[0..11]: class sizes by pair count (Wilcoxon), hist[12..23]: by longer segment (Mann-Whitney / Kruskal);
// cursor = hist + 24.  One atomic per class and warp (match.any), not per unit.
__device__ __forceinline__ uint32_t warp_class_slot(uint32_t cls, uint32_t *counter, unsigned lane) {
  const unsigned peers = __match_any_sync(0xffffffffu, cls);
  const int leader = __ffs(peers) - 1;
  uint32_t base = 0;
  if ((int)lane == leader) base = atomicAdd(counter + cls, (uint32_t)__popc(peers));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
}
__global__ void __launch_bounds__(256) k_rank_class_hist(DevBatch b, uint32_t *__restrict__ hist) {
  const uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t nb = 0, nc = 0;
  if (u < b.n_units) { nb = b.seg_lens[3 * u + 1]; nc = b.seg_lens[3 * u + 2]; }
  // lanes past the end count into class 0 of a third, unused histogram
  const unsigned lane = fm_lane();
  warp_class_slot(u < b.n_units ? fm_rank_class(min(nb, nc)) : 36u, hist, lane);
  warp_class_slot(u < b.n_units ? 12u + fm_rank_class(max(nb, nc)) : 37u, hist, lane);
}
__global__ void __launch_bounds__(256) k_rank_class_scatter(DevBatch b, uint32_t *__restrict__ hist, uint32_t *__restrict__ list_w,
                                                            uint32_t *__restrict__ list_m) {
  const uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool have = u < b.n_units;
  uint32_t nb = 0, nc = 0;
  if (have) { nb = b.seg_lens[3 * u + 1]; nc = b.seg_lens[3 * u + 2]; }
  const unsigned lane = fm_lane();
  uint32_t *const cursor = hist + 40;
  const uint32_t cw = have ? fm_rank_class(min(nb, nc)) : 36u, cm = have ? 12u + fm_rank_class(max(nb, nc)) : 37u;
  const uint32_t pw = warp_class_slot(cw, cursor, lane), pm = warp_class_slot(cm, cursor, lane);
  if (have) {
    uint32_t ow = 0, om = 0;
    for (uint32_t c = 0; c < cw; ++c) ow += hist[c];
    for (uint32_t c = 12; c < cm; ++c) om += hist[c];
    list_w[ow + pw] = (uint32_t)u;
    list_m[om + pm] = (uint32_t)u;
  }
}

template <int R>
cudaError_t launch_wilcoxon(const DevBatch &b, const DevParams *dp, uint32_t max_pairs, fm_series_score *scores, int sm_count,
                            cudaStream_t st, UnitSet us = UnitSet{nullptr, nullptr, 0}) {
  const uint64_t blocks = std::min<uint64_t>((b.n_units + kWarpsPerBlock - 1) / kWarpsPerBlock, (uint64_t)sm_count * kBlocksPerSm);
  if (blocks == 0) return cudaSuccess;
  const int raw_cap = ((int)std::min<uint32_t>(max_pairs, 32u * R) + 3) & ~3;
  const int smem = kWarpsPerBlock * 2 * raw_cap * 4;
  cudaError_t e = us.list ? cudaFuncSetAttribute(k_wilcoxon<R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                          : cudaFuncSetAttribute(k_wilcoxon<R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  if (us.list) k_wilcoxon<R, true><<<(unsigned)blocks, kWarpsPerBlock * 32, smem, st>>>(b, dp, scores, raw_cap, us);
  else k_wilcoxon<R, false><<<(unsigned)blocks, kWarpsPerBlock * 32, smem, st>>>(b, dp, scores, raw_cap, us);
  return cudaGetLastError();
}

template <int R>
cudaError_t launch_mw(const DevBatch &b, const DevParams *dp, uint32_t max_seg, fm_series_score *scores, int sm_count,
                      cudaStream_t st, UnitSet us = UnitSet{nullptr, nullptr, 0}) {
  constexpr int W = R >= 48 ? 4 : (R == 32 ? 6 : 8);
  constexpr int N = 32 * R;
  const int seg_cap = (int)std::min<uint32_t>(max_seg, 32u * R);
  const int rl_cap = seg_cap + 2, key_words = ((seg_cap + 7) & ~7) + 8;  // a chunk may start just below seg_cap
  // id table of the duplicate search: as many u16 slots as fit while three blocks stay resident per SM
  // (fewer slots -> more keys need a second round), never fewer than its later tenants need (counters +
  // run lengths) nor than 1.25 keys' worth, and no more than 4 per key
  const int budget = (227 * 1024 - 3 * 1024) / (3 * W) / 16 * 16 - ((N + 4) + key_words) * 4;
  int n_slots = std::max(std::max(N + rl_cap, seg_cap + seg_cap / 4), std::min(budget / 2, 4 * seg_cap));
  n_slots = (n_slots + 7) & ~7;
  const int smem = W * ((N + 4) + key_words + n_slots / 2) * 4;
  cudaError_t e = us.list ? cudaFuncSetAttribute(k_mw<R, W, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                          : cudaFuncSetAttribute(k_mw<R, W, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const uint64_t blocks = std::min<uint64_t>((b.n_units + W - 1) / W, (uint64_t)sm_count * kBlocksPerSm);
  if (blocks == 0) return cudaSuccess;
  if (us.list) k_mw<R, W, true><<<(unsigned)blocks, W * 32, smem, st>>>(b, dp, scores, rl_cap, key_words, n_slots, us);
  else k_mw<R, W, false><<<(unsigned)blocks, W * 32, smem, st>>>(b, dp, scores, rl_cap, key_words, n_slots, us);
  return cudaGetLastError();
}

// the sort sizes: R keys per lane, R = 2^q or 3*2^q
#define FM_FOR_SORT_SIZE(n, CALL)                                                                    \
  ((n) <= 32 ? CALL(1) : (n) <= 64 ? CALL(2) : (n) <= 96 ? CALL(3) : (n) <= 128 ? CALL(4) : (n) <= 192 ? CALL(6)      \
   : (n) <= 256 ? CALL(8) : (n) <= 384 ? CALL(12) : (n) <= 512 ? CALL(16) : (n) <= 768 ? CALL(24)   \
   : (n) <= 1024 ? CALL(32) : (n) <= 1536 ? CALL(48) : CALL(64))

}  // namespace

// max_base / max_cur = longest baseline / current segment in the batch.  Units whose segments exceed
// FM_RANK_MAX_SEGMENT samples are not ranked (D17: Mann-Whitney / Kruskal need both segments within the
// limit, Wilcoxon its min(n_base, n_cur) pairs); the sort size is chosen for the rest.
cudaError_t fm_launch_rank(const DevBatch &b, const DevParams *dp, uint32_t pairwise, uint32_t max_base, uint32_t max_cur,
                           fm_series_score *scores, int sm_count, cudaStream_t st, cudaEvent_t after_wilcoxon,
                           uint32_t mask_w, uint32_t mask_m, uint32_t *class_scratch, uint32_t *class_lists) {
  const bool all = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY;
  const bool w = all || pairwise == FM_PW_WILCOXON;
  const bool pooled = all || pairwise == FM_PW_MANN_WHITE || pairwise == FM_PW_KRUSKAL;
  const uint32_t cap = FM_RANK_MAX_SEGMENT;
  const uint32_t max_pairs = std::min(std::min(max_base, max_cur), cap);  // bounds min(n_base, n_cur) of every unit
  const uint32_t max_seg = std::min(std::max(max_base, max_cur), cap);
  cudaError_t e = cudaSuccess;
  // several sort sizes in the batch: group the units by class, one launch per class that occurs
  const bool split_w = w && class_scratch && class_lists && (mask_w & (mask_w - 1u)) != 0;
  const bool split_m = pooled && class_scratch && class_lists && (mask_m & (mask_m - 1u)) != 0;
  uint32_t *const hist = class_scratch, *const list_w = class_lists, *const list_m = class_lists + b.n_units;
  if ((split_w || split_m) && b.n_units) {
    e = cudaMemsetAsync(hist, 0, 64 * sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((b.n_units + 255) / 256);
    k_rank_class_hist<<<blocks, 256, 0, st>>>(b, hist);
    k_rank_class_scatter<<<blocks, 256, 0, st>>>(b, hist, list_w, list_m);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
#define FM_W(R) launch_wilcoxon<R>(b, dp, pairs_c, scores, sm_count, st, us)
#define FM_M(R) launch_mw<R>(b, dp, seg_c, scores, sm_count, st, us)
  if (w) {
    for (int c = 0; c < 12 && e == cudaSuccess; ++c) {
      if (split_w && !((mask_w >> c) & 1u)) continue;
      if (!split_w && c) break;
      const uint32_t pairs_c = split_w ? std::min(kRankClassSize[c], max_pairs) : max_pairs;
      const UnitSet us = split_w ? UnitSet{list_w, hist, c} : UnitSet{nullptr, nullptr, 0};
      e = FM_FOR_SORT_SIZE(pairs_c, FM_W);
    }
  }
  if (e != cudaSuccess) return e;
  if (after_wilcoxon) cudaEventRecord(after_wilcoxon, st);
  if (pooled) {
    for (int c = 0; c < 12 && e == cudaSuccess; ++c) {
      if (split_m && !((mask_m >> c) & 1u)) continue;
      if (!split_m && c) break;
      const uint32_t seg_c = split_m ? std::min(kRankClassSize[c], max_seg) : max_seg;
      const UnitSet us = split_m ? UnitSet{list_m, hist + 12, c} : UnitSet{nullptr, nullptr, 0};
      e = FM_FOR_SORT_SIZE(seg_c, FM_M);
    }
  }
  if (e != cudaSuccess) return e;
#undef FM_W
#undef FM_M
  if (b.n_units) k_rank_finalize<<<(unsigned)((b.n_units + 255) / 256), 256, 0, st>>>(b, dp, scores);
  return cudaGetLastError();
}
