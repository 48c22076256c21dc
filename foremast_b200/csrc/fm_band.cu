// fm_band.cu -- band models + anomaly flagging: moving-average band,
// EWMA and double-exponential smoothing, fused with the flagging of the current window.
//
// Replaces the brain's "compute the historical model ... use threshold to detect current anomaly
// data points" step (docs/guides/design.md:35, models list design.md:68-70, thresholds
// deploy/foremast/3_brain/foremast-brain.yaml:26-73).
//
// Mapping: ONE THREAD PER SERIES-WINDOW, a warp owns 32 consecutive units.  The recurrences are
// evaluated sequentially in sample order -- exactly the order of oracle/DECISIONS.md -- so sums,
// bands and therefore flags are bit-identical to the CPU oracle; the parallelism comes from the
// 10^5..10^7 independent series, not from scanning inside one.  HBM access stays coalesced by
// staging: each warp pulls, per stage, a 128 B row (32 samples) of each of its 32 units with
// 16 B cp.async copies (every 128 B line fully used by one instruction) into a 2-deep shared
// memory ring, and each lane then walks its own row with conflict-free 128-bit LDS (row stride
// 36 words).  Algorithmic bytes per unit: 4*(n_hist+n_base+n_cur) + 8 (offset) + 12 (lengths)
// + results.
#include <algorithm>

#include "fm_common.cuh"

namespace {

constexpr int kChunk = 32;        // samples per row per stage
constexpr int kRowWords = 36;     // 32 + 4 pad words: 16 B aligned rows, conflict-free LDS.128
constexpr int kStages = 2;        // current chunk + one in flight (deeper rings cost occupancy: 3 -> 2 gained 12 %)
constexpr int kWarpsPerBlock = 8;  // launch bound; small batches run 6-warp blocks (launch_mode)
// Variants with the rolling band keep one more slot: the chunk BEFORE the current one stays intact,
// so the lagging sample (W <= 32 back) comes from shared memory instead of a scattered global load.
__host__ __device__ constexpr int ring_slots(uint32_t algs) { return (algs & FM_ALG_MA_ROLLING) ? kStages + 1 : kStages; }
__host__ __device__ constexpr int ring_bytes(uint32_t algs, int wpb) { return wpb * ring_slots(algs) * 32 * kRowWords * 4; }
__host__ __device__ constexpr int lo_bytes(int wpb) { return wpb * 32 * kRowWords * 4; }  // MODE 1: staging of the lower band, one chunk per warp

struct Rule {
  double k, mlb;
  uint32_t bound;
};

template <int N>
struct IntC { static constexpr int value = N; };

template <uint32_t ALGS>
struct State {
  // MA_ALL
  double x0, s1, s2, ma_up, ma_lo, ma_mean, ma_sd;
  // EWMA
  double es, esse, esig;
  // DES
  double dl, db, dsse, dsig;
  // MA_ROLLING: running prefix sums of d and d*d (rp*) and the same sums W samples behind (rq*)
  double rp1, rp2, rq1, rq2, rsd;
  uint32_t cnt[4], first[4];  // MA_ALL, EWMA, DES, MA_ROLLING
};

// One sample of the stream W positions back, for the lagging prefix sums of the rolling band.
// If it lies in the current or the previous 32-sample chunk it is still in this lane's row of the
// shared-memory ring (4-way bank conflict at worst); otherwise it is read from global memory, where
// the main stream fetched the line moments ago (L1/L2 hit, no HBM traffic, but 32 different lines
// per warp instruction -- 6x slower, measured, when every sample went that way).
struct Lag {
  const float *unit;   // values + offset of this unit
  const float *srow;   // this lane's row in ring slot 0
  uint32_t a0, ns, c0, W, last_j;  // last_j: index of the last current point
  uint32_t lo_pos, cur_chunk, cur_off, prev_off;  // smem holds unit positions >= lo_pos; slot offsets in floats
  __device__ __forceinline__ bool at(uint32_t q, double &x) const {
    if (q < W) return false;
    const uint32_t sp = q - W;
    const uint32_t p = sp < ns ? a0 + sp : c0 + (sp - ns);
    if (p >= lo_pos) x = (double)srow[((p >> 5) == cur_chunk ? cur_off : prev_off) + (p & 31u)];
    else x = (double)__ldg(unit + p);
    return true;
  }
};

__device__ __forceinline__ bool outside(double x, double up, double lo, uint32_t bound) {
  return ((bound & 1u) && x > up) || ((bound & 2u) && x < lo);
}

// Lagging sample of the rolling band for chunks that lie at least one whole chunk inside their window, with
// W <= 32: sample I of the chunk lags onto index I - W of the same row, or onto 32 + I - W of the previous
// chunk's row, which the ring still holds.  cur_w / prev_w are those rows shifted by -W / 32 - W, so the
// access is one warp-uniform select and one LDS at a literal offset.
struct DeepLag {
  const float *cur_w, *prev_w;
  uint32_t W;
  double Wd;  // (double)W: the window is full on these paths
  // i = index of the sample in its chunk = K + 4*v4, with cur_w / prev_w already advanced by 4*v4
  template <int K>
  __device__ __forceinline__ double at(uint32_t i) const { return (double)(i >= W ? cur_w : prev_w)[K]; }
};

template <uint32_t ALGS, int K>
__device__ __forceinline__ void roll_advance_deep(State<ALGS> &s, double xd, const DeepLag &dl, uint32_t i) {
  const double d = xd - s.x0;
  s.rp1 += d;
  s.rp2 = fma(d, d, s.rp2);
  const double dq = dl.at<K>(i) - s.x0;
  s.rq1 += dq;
  s.rq2 = fma(dq, dq, s.rq2);
}

template <uint32_t ALGS>
__device__ __forceinline__ void roll_advance(State<ALGS> &s, double xd, uint32_t q, const Lag &lag, bool valid) {
  const double d = xd - s.x0;
  s.rp1 += d;
  s.rp2 = fma(d, d, s.rp2);
  double xl;
  if (valid && lag.at(q, xl)) {  // lanes riding along in a warp-uniform fast path must not touch memory
    const double dq = xl - s.x0;
    s.rq1 += dq;
    s.rq2 = fma(dq, dq, s.rq2);
  }
}

template <uint32_t ALGS, bool DEEP = false, int K = 0>
__device__ __forceinline__ void src_update(State<ALGS> &s, double xd, double alpha, double ab, uint32_t q, const Lag &lag,
                                           bool valid, const DeepLag *dl = nullptr, uint32_t i = 0) {
  if (ALGS & FM_ALG_MA_ROLLING) {
    if (DEEP) roll_advance_deep<ALGS, K>(s, xd, *dl, i);
    else roll_advance<ALGS>(s, xd, q, lag, valid);
  }
  if (ALGS & FM_ALG_MA_ALL) {
    double d = xd - s.x0;
    s.s1 += d;
    s.s2 = fma(d, d, s.s2);
  }
  if (ALGS & FM_ALG_EWMA) {
    double e = xd - s.es;
    s.esse = fma(e, e, s.esse);
    s.es = fma(alpha, e, s.es);
  }
  if (ALGS & FM_ALG_DES) {
    double f = s.dl + s.db, e = xd - f;
    s.dsse = fma(e, e, s.dsse);
    s.dl = fma(alpha, e, f);
    s.db = fma(ab, e, s.db);
  }
}

// ---- forecast bands (EWMA, DES) without the band arithmetic per point -----------------------------------
// The oracle flags x against up = fma(k, sig, f) and lo = fma(-k, sig, f).  With e = fl(x - f) and K = k sig,
// x > up needs x - f > K - u (K + |f|) (u = 2^-53, the rounding of up), and symmetrically below.  So
//     |e| <= fl(K) (1 - 2^-30) - 2^-40 (|f| + |x|)
// proves the point inside: the two margins exceed the roundings of e, of fl(K), of up / lo and of this test itself
// (each at most a few u of K + |f| + |x|) by factors of 2^20 and 2^10.  Points that fail it -- anomalies, points
// within 1e-9 of an edge, NaNs (every comparison false) -- take the oracle's arithmetic, so flags stay bit-identical;
// x < min_lower_bound is a separate, exact comparison.  A threshold k that is negative or not finite disables the
// filter (-1: nothing is <= it).
constexpr double kBandRel = 1.0 - 9.313225746154785e-10;  // 1 - 2^-30
constexpr double kBandAbs = 9.094947017729282e-13;        // 2^-40
__device__ __forceinline__ double band_clear(double k, double sig) {
  return (k >= 0.0 && k < 1.0e150) ? (k * sig) * kBandRel : -1.0;
}

template <uint32_t ALGS>
__device__ __forceinline__ void finalize(State<ALGS> &s, const Rule &r, uint32_t ns) {
  if (ALGS & FM_ALG_MA_ALL) {
    double m1 = s.s1 / (double)ns, mean = s.x0 + m1;
    double var = fma(-m1, m1, s.s2 / (double)ns);
    if (!(var > 0.0)) var = 0.0;
    double sd = sqrt(var);
    s.ma_mean = mean;
    s.ma_sd = sd;
    s.ma_up = fma(r.k, sd, mean);
    double lo = fma(-r.k, sd, mean);
    s.ma_lo = lo < r.mlb ? r.mlb : lo;
    // the prefix sums are done with: their registers carry the bounds cur_fast compares with, a side that the rule
    // does not have pushed out to infinity (two comparisons per point, no rule bits)
    s.s1 = (r.bound & 1u) ? s.ma_up : __longlong_as_double(0x7FF0000000000000ll);
    s.s2 = (r.bound & 2u) ? s.ma_lo : __longlong_as_double(0xFFF0000000000000ll);
  }
  if (ALGS & FM_ALG_EWMA) s.esig = ns >= 2 ? sqrt(s.esse / (double)(ns - 1)) : 0.0;
  if (ALGS & FM_ALG_DES) s.dsig = ns >= 2 ? sqrt(s.dsse / (double)(ns - 1)) : 0.0;
  // the residual sums are done with: their registers carry the inside-the-band thresholds of cur_fast from here on
  if (ALGS & FM_ALG_EWMA) s.esse = band_clear(r.k, s.esig);
  if (ALGS & FM_ALG_DES) s.dsse = band_clear(r.k, s.dsig);
}

// Per-thread form of the rule for the fast path: x < max(mlb, lo) is evaluated as
// (x < lo) || (x < mlb), which needs no fp64 max, and a disabled side is a constant-false predicate.
struct FastRule {
  double k, nk, mlb;
  double k2e;  // k*k*(1 - kRollEps), 0 when k is not a usable (finite, >= 0) threshold: every point then takes the exact path
  bool up_on, lo_on;
};

// ---- rolling band without a division or a square root per point ------------------------------------
// The oracle flags a current point x against up = fma(k, sd, mean), mean = x0 + a1/c, sd = sqrt(var),
// var = max(0, fma(-m1, m1, a2/c)) with a1, a2 the window sums of d = x - x0 and d*d: two fp64 divisions
// and a square root per point, and the band kernel spent its time in them.  Nearly every point lies well
// inside its band, and that can be seen from the sums alone: with D = c*(x - x0) - a1 = c*(x - mean) and
// V = c*a2 - a1^2 = c^2*var (both exact up to rounding), x > up would need D > k*sqrt(V).  The filter
// below proves "not flagged" whenever (|D| + eD)^2 <= k^2*(V - eV)*(1 - eps) (or D is on the harmless
// side of a one-sided band), where eD and eV bound -- with a factor > 1000 to spare, eps = 2^-40 against
// unit roundoff 2^-53 -- every rounding the oracle's own evaluation and this one can have made:
//   |c*mean_o - (c*x0 + a1)| <= 2^-52 (c|x0| + |a1|),  |c^2 var_o - V| <= 2^-51 (c|a2| + a1^2),
//   sd_o and up_o within 2^-53 relative of sqrt(var_o) and k*sd_o + mean_o.
// Points the filter cannot clear (real anomalies, points within 1e-12 of the band edge, degenerate
// windows) are decided by roll_exact, the oracle's arithmetic verbatim, so flags stay bit-identical.
constexpr double kRollEps = 9.094947017729282e-13;  // 2^-40

struct RollExact {
  bool flag;
  double sd;
};
__device__ __noinline__ RollExact roll_exact(double a1, double a2, uint32_t cnt, double x0, double k, double mlb,
                                             uint32_t bound, double xd) {
  const double m1 = a1 / (double)cnt, mean = x0 + m1;
  double var = fma(-m1, m1, a2 / (double)cnt);
  if (!(var > 0.0)) var = 0.0;
  const double sd = sqrt(var), up = fma(k, sd, mean);
  double lo = fma(-k, sd, mean);
  lo = lo < mlb ? mlb : lo;
  RollExact r;
  r.flag = ((bound & 1u) && xd > up) || ((bound & 2u) && xd < lo);
  r.sd = sd;
  return r;
}

// true when the point certainly is NOT outside the rolling band (see above); c = window count as a double
__device__ __forceinline__ bool roll_clear(double a1, double a2, double c, double d, double xd, double x0,
                                           const FastRule &fr) {
  const double D = fma(c, d, -a1);
  const double a1sq = a1 * a1;
  const double V = fma(c, a2, -a1sq);
  const double eD = kRollEps * fma(c, fabs(xd) + fabs(x0), fabs(a1));
  const double eV = kRollEps * fma(c, fabs(a2), a1sq);
  const double kin = fr.k2e * (V - eV);
  const double dp = D + eD, dm = eD - D;  // upper side: D + eD, lower side: -D + eD
  const bool in_up = !fr.up_on | (dp <= 0.0) | (dp * dp <= kin);
  const bool in_lo = !fr.lo_on | (dm <= 0.0) | (dm * dm <= kin);
  return in_up & in_lo;
}

__device__ __forceinline__ bool outside_fast(double x, double up, double lo, const FastRule &fr) {
  return (fr.up_on && x > up) || (fr.lo_on && (x < lo || x < fr.mlb));
}

// Fast-path twin of cur_update for chunks that lie entirely inside the current window: identical
// arithmetic, but flags are gathered into one bit mask per algorithm per 32-sample chunk (bit I)
// and counted once per chunk instead of being counted point by point.
// K = position of the sample inside its group of four, i = its index in the chunk (the bit it owns in mask).
template <uint32_t ALGS, int K, bool DEEP = false>
__device__ __forceinline__ void cur_fast(State<ALGS> &s, const FastRule &fr, double xd, uint32_t j, bool valid,
                                         double alpha, double ab, uint32_t (&mask)[4], const Lag &lag, const uint32_t i,
                                         const DeepLag *dl = nullptr) {
  if (ALGS & FM_ALG_MA_ROLLING) {
    // DEEP: at least 32 current points lie behind this one, so the window is full (W <= 32)
    const uint32_t q = lag.ns + j, cnt = DEEP ? dl->W : (lag.W < q ? lag.W : q);
    const double c = DEEP ? dl->Wd : (double)cnt;
    const double a1 = s.rp1 - s.rq1, a2 = s.rp2 - s.rq2;
    const double d = xd - s.x0;
    // the last current point also reports the sd it was judged with (sigma[1]): exact path.  Bitwise logic:
    // the filter is evaluated by every lane, only the exact path is a branch
    const bool need = valid & (!roll_clear(a1, a2, c, d, xd, s.x0, fr) | (j == lag.last_j));
    bool f = fr.lo_on & (xd < fr.mlb);  // the clamp of the lower band is a plain comparison
    if (need) {
      // A flat window with the point on it (a gauge that does not move, an error count that stays 0: common) is the
      // one case the filter can never clear, since V = 0.  The oracle's arithmetic on it is exact: a1 = a2 = 0 and
      // x = x0 give mean = x0, sd = 0, up = lo = x0, so only the lower clamp (f already holds it) can flag the point.
      if ((a1 == 0.0) & (a2 == 0.0) & (d == 0.0)) {
        s.rsd = 0.0;
      } else {
        const RollExact e = roll_exact(a1, a2, cnt, s.x0, fr.k, fr.mlb, (fr.up_on ? 1u : 0u) | (fr.lo_on ? 2u : 0u), xd);
        f = e.flag;
        s.rsd = e.sd;
      }
    }
    if (f) mask[3] |= 1u << i;
    if (DEEP) roll_advance_deep<ALGS, K>(s, xd, *dl, i);
    else roll_advance<ALGS>(s, xd, q, lag, valid);
  }
  if (ALGS & FM_ALG_MA_ALL) {
    // s.s1 / s.s2 = ma_up / ma_lo (which carries the clamp) or +-infinity for a side the rule does not have (finalize)
    if ((xd > s.s1) | (xd < s.s2)) mask[0] |= 1u << i;
  }
  const bool below = fr.lo_on & (xd < fr.mlb);  // the clamp of a lower band, shared by the forecast bands
  if (ALGS & FM_ALG_EWMA) {
    const double f0 = s.es, e = xd - f0;
    if (!(fabs(e) <= fma(-kBandAbs, fabs(f0) + fabs(xd), s.esse)) | below) {  // not provably inside (band_clear)
      const double up = fma(fr.k, s.esig, f0), lo = fma(fr.nk, s.esig, f0);
      if (outside_fast(xd, up, lo, fr)) mask[1] |= 1u << i;
    }
    s.es = fma(alpha, e, s.es);
  }
  if (ALGS & FM_ALG_DES) {
    const double f0 = s.dl + s.db, e = xd - f0;
    if (!(fabs(e) <= fma(-kBandAbs, fabs(f0) + fabs(xd), s.dsse)) | below) {
      const double up = fma(fr.k, s.dsig, f0), lo = fma(fr.nk, s.dsig, f0);
      if (outside_fast(xd, up, lo, fr)) mask[2] |= 1u << i;
    }
    s.dl = fma(alpha, e, f0);
    s.db = fma(ab, e, s.db);
  }
}

// One current point: flag against each band, then keep the smoothers running (D4).
template <uint32_t ALGS>
__device__ __forceinline__ void cur_update(State<ALGS> &s, const Rule &r, double xd, uint32_t j, bool valid,
                                           double alpha, double ab, float &bup, float &blo, uint32_t band_alg,
                                           const Lag &lag) {
  if (ALGS & FM_ALG_MA_ROLLING) {
    const uint32_t q = lag.ns + j, cnt = lag.W < q ? lag.W : q;
    const double a1 = s.rp1 - s.rq1, a2 = s.rp2 - s.rq2;
    const double m1 = a1 / (double)cnt, mean = s.x0 + m1;
    double var = fma(-m1, m1, a2 / (double)cnt);
    if (!(var > 0.0)) var = 0.0;
    const double sd = sqrt(var), up = fma(r.k, sd, mean);
    double lo = fma(-r.k, sd, mean);
    lo = lo < r.mlb ? r.mlb : lo;
    if (valid) s.rsd = sd;
    const bool f = valid && outside(xd, up, lo, r.bound);
    if (f && s.cnt[3]++ == 0) s.first[3] = j;
    if (band_alg == FM_ALG_MA_ROLLING) { bup = (float)up; blo = (float)lo; }
    roll_advance<ALGS>(s, xd, q, lag, valid);
  }
  if (ALGS & FM_ALG_MA_ALL) {
    bool f = valid && outside(xd, s.ma_up, s.ma_lo, r.bound);
    if (f && s.cnt[0]++ == 0) s.first[0] = j;
    if (band_alg == FM_ALG_MA_ALL) { bup = (float)s.ma_up; blo = (float)s.ma_lo; }
  }
  if (ALGS & FM_ALG_EWMA) {
    double f0 = s.es, up = fma(r.k, s.esig, f0), lo = fma(-r.k, s.esig, f0);
    lo = lo < r.mlb ? r.mlb : lo;
    bool f = valid && outside(xd, up, lo, r.bound);
    if (f && s.cnt[1]++ == 0) s.first[1] = j;
    if (band_alg == FM_ALG_EWMA) { bup = (float)up; blo = (float)lo; }
    double e = xd - f0;
    s.es = fma(alpha, e, s.es);
  }
  if (ALGS & FM_ALG_DES) {
    double f0 = s.dl + s.db, up = fma(r.k, s.dsig, f0), lo = fma(-r.k, s.dsig, f0);
    lo = lo < r.mlb ? r.mlb : lo;
    bool f = valid && outside(xd, up, lo, r.bound);
    if (f && s.cnt[2]++ == 0) s.first[2] = j;
    if (band_alg == FM_ALG_DES) { bup = (float)up; blo = (float)lo; }
    double e = xd - f0;
    s.dl = fma(alpha, e, f0);
    s.db = fma(ab, e, s.db);
  }
}

// MODE 0: score.  MODE 1: score and materialise the per-point band of dp->band_algorithm.
// MODE 2: emit pass of the anomaly list -- ALGS is one algorithm, only units with anomalies are
// read again, and every flagged point is written to (emit_idx, emit_val) at emit_off[u] + k.
// Two 8-warp blocks per SM need at most 128 registers; the kernels with three or four scorers where ptxas would take a
// few more (and halve the occupancy) are held to that.  The others are left alone: told "two blocks", ptxas spends the
// allowance (91 -> 126 registers for EWMA + DES), which costs the 6-warp blocks of small batches their third block.
constexpr int band_min_blocks(uint32_t algs) { return (algs == 7u || algs == 11u || algs == 14u || algs == 15u) ? 2 : 0; }  // 0: no statement

template <uint32_t ALGS, int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, band_min_blocks(ALGS)) k_band(DevBatch b, const DevParams *__restrict__ dp,
                                                              bool rank_ran, fm_series_score *__restrict__ scores,
                                                              uint32_t *__restrict__ unit_status,
                                                              uint32_t *__restrict__ unit_points,
                                                              float *__restrict__ band_up, float *__restrict__ band_lo,
                                                              const unsigned long long *__restrict__ emit_off,
                                                              uint32_t *__restrict__ emit_idx, float *__restrict__ emit_val) {
  constexpr bool BANDS = MODE == 1, EMIT = MODE == 2;
  // chunk bodies: eight groups of four samples.  Kernels that carry the rolling band keep them rolled (its
  // per-point code times 32 samples times five variants was 11 k instructions and waited on instruction fetch);
  // the others fit the instruction cache unrolled and are 3-8 % faster that way
  constexpr int kChunkUnroll = (ALGS & FM_ALG_MA_ROLLING) ? 1 : 8;
  constexpr int kSlot = (ALGS & FM_ALG_MA_ALL) ? 0 : (ALGS & FM_ALG_EWMA) ? 1 : (ALGS & FM_ALG_DES) ? 2 : 3;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  constexpr int kSlots = ring_slots(ALGS);
  const int wpb = (int)(blockDim.x >> 5), kRingB = ring_bytes(ALGS, wpb);  // warps per block is a launch-time choice
  float *const s_ring = reinterpret_cast<float *>(smem_raw) + warp * (kSlots * 32 * kRowWords);
  // per row: the address of the unit's first sample (as a number: the staging adds the chunk's byte offset to it) ...
  unsigned long long *const s_src = reinterpret_cast<unsigned long long *>(smem_raw + kRingB) + warp * 32;
  uint32_t *const s_len = reinterpret_cast<uint32_t *>(smem_raw + kRingB + wpb * 32 * 8) + warp * 32;
  float *const s_lo = reinterpret_cast<float *>(smem_raw + kRingB + wpb * 32 * (8 + 4)) + warp * (32 * kRowWords);  // MODE 1 only
  const uint64_t gid = ((uint64_t)blockIdx.x * wpb + warp) * 32 + lane;
  const bool have = gid < b.n_units;
  const uint64_t u = have && b.order ? (uint64_t)b.order[gid] : gid;  // see fm_launch_band_order

  // ---- per-lane unit description
  int64_t off = 0;
  uint32_t nh = 0, nb = 0, nc = 0, mt = 0, status = 0;
  if (have) {
    off = b.offsets[u];
    nh = b.seg_lens[3 * u]; nb = b.seg_lens[3 * u + 1]; nc = b.seg_lens[3 * u + 2];
    mt = b.metric_type[u];
    if (rank_ran || EMIT) status = scores[u].status;
  }
  unsigned long long emit_base = 0;
  bool emit_any = true;
  if (EMIT && have) {
    emit_base = emit_off[u];
    emit_any = emit_off[u + 1] > emit_base;
  }
  const uint32_t ntot = nh + nb + nc;
  // model window (D15)
  uint32_t a0 = 0, ns = 0;
  const uint32_t min_hist = dp->min_historical;
  if (nh >= min_hist && nh > 0) { a0 = 0; ns = nh; }
  else if (nb >= min_hist && nb > 0) { a0 = nh; ns = nb; status |= FM_S_MODEL_BASE; }
  const bool active = have && ns > 0 && nc > 0 && emit_any;
  if (have && !active) status |= FM_S_UNKNOWN;
  // ranges of this lane in unit coordinates; empty when inactive
  const uint32_t a1 = active ? a0 + ns : 0, c0 = active ? nh + nb : 0, c1 = active ? ntot : 0;
  if (!active) a0 = 0;

  Rule r;
  {
    const DevRule &dr = dp->rules[mt < FM_MAX_METRIC_TYPES ? mt : 0];
    r.k = (status & FM_S_DIFFERS) ? dr.threshold * dp->differ_scale : dr.threshold;
    r.mlb = dr.min_lower_bound;
    r.bound = dr.bound;
  }
  FastRule fr;
  fr.k = r.k; fr.nk = -r.k; fr.mlb = r.mlb;
  fr.k2e = (r.k >= 0.0 && r.k < 1.0e150) ? r.k * r.k * (1.0 - kRollEps) : 0.0;
  fr.up_on = (r.bound & 1u) != 0; fr.lo_on = (r.bound & 2u) != 0;
  const double alpha = dp->alpha, ab = dp->alpha * dp->beta;
  const uint32_t band_alg = BANDS ? dp->band_algorithm : 0u;

  State<ALGS> s;
  s.x0 = s.s1 = s.s2 = s.ma_up = s.ma_lo = s.ma_mean = s.ma_sd = 0.0;
  s.es = s.esse = s.esig = 0.0;
  s.dl = s.db = s.dsse = s.dsig = 0.0;
  s.rp1 = s.rp2 = s.rq1 = s.rq2 = s.rsd = 0.0;
#pragma unroll
  for (int a = 0; a < 4; ++a) { s.cnt[a] = 0; s.first[a] = FM_NONE_IDX; }
  Lag lag;
  lag.unit = b.values + off;
  lag.srow = s_ring + lane * kRowWords;
  lag.lo_pos = 0xFFFFFFFFu; lag.cur_chunk = 0; lag.cur_off = 0; lag.prev_off = 0;
  lag.a0 = a0; lag.ns = ns; lag.c0 = c0; lag.last_j = active ? nc - 1 : 0xFFFFFFFFu;
  lag.W = active ? dp->ma_window : 0xFFFFFFFFu;  // lanes without a unit never touch memory
  const uint32_t w_uniform = dp->ma_window;

  // ---- staging set-up
  s_src[lane] = (unsigned long long)reinterpret_cast<uintptr_t>(b.values + off);
  s_len[lane] = (active && ALGS) ? ((ntot + 3u) & ~3u) : 0u;   // inactive units are not even read
  __syncwarp();
  // with no band algorithm selected the kernel only resolves statuses: nothing is staged
  const uint32_t nchunks = ALGS ? (warp_max_u32(active ? ntot : 0u) + kChunk - 1) / kChunk : 0u;
  const unsigned row_sub = lane >> 3, col4 = lane & 7;

  auto issue = [&](uint32_t ch) {
    if (ch < nchunks) {
      float *dst = s_ring + (ch % kSlots) * (32 * kRowWords);
      const uint32_t pos = ch * kChunk + col4 * 4;
      uint32_t len[8];
      unsigned long long src[8];
#pragma unroll
      for (int rr = 0; rr < 8; ++rr) { len[rr] = s_len[rr * 4 + row_sub]; src[rr] = s_src[rr * 4 + row_sub]; }
#pragma unroll
      for (int rr = 0; rr < 8; ++rr) {
        const unsigned row = rr * 4 + row_sub;
        if (pos < len[rr])
          cp_async16(dst + row * kRowWords + col4 * 4, reinterpret_cast<const float *>((uintptr_t)(src[rr] + 4ull * pos)));
      }
    }
    cp_async_commit();
  };

#pragma unroll
  for (int st = 0; st < kStages - 1; ++st) issue(st);

  for (uint32_t ch = 0; ch < nchunks; ++ch) {
    cp_async_wait<kStages - 2>();
    __syncwarp();
    issue(ch + kStages - 1);

    const float *row = s_ring + (ch % kSlots) * (32 * kRowWords) + lane * kRowWords;
    if (ALGS & FM_ALG_MA_ROLLING) {
      lag.cur_chunk = ch;
      lag.cur_off = (ch % kSlots) * (32 * kRowWords);
      lag.prev_off = ((ch + kSlots - 1) % kSlots) * (32 * kRowWords);
      lag.lo_pos = ch ? (ch - 1) * kChunk : 0u;
    }
    const uint32_t cs = ch * kChunk, ce = cs + kChunk;
    const bool done = cs >= c1, not_started = ce <= a0;
    const bool cls_src = cs >= a0 + 2 && ce <= a1;
    const bool cls_cur = cs > c0 && ce <= c1;
    // rolling band, W <= 32: the chunk before this one lies in the same window, so every lagging sample is
    // in this lane's row of the current or the previous ring slot
    const bool roll_deep = (ALGS & FM_ALG_MA_ROLLING) && w_uniform <= 32u;
    DeepLag dlag;
    if (ALGS & FM_ALG_MA_ROLLING) {
      dlag.W = w_uniform;
      dlag.Wd = (double)w_uniform;
      dlag.cur_w = row - w_uniform;
      dlag.prev_w = s_ring + ((ch + kSlots - 1) % kSlots) * (32 * kRowWords) + lane * kRowWords + 32 - w_uniform;
    }
    const bool cls_skip = not_started || done || (cs >= a1 && ce <= c0);
    const bool dontcare = !active || done || not_started;

    if (roll_deep && __all_sync(0xffffffffu, (cls_src && cs >= a0 + 32u) || dontcare)) {
      // rolled over the eight groups of four samples: the unrolled chunk bodies of all variants together
      // were 11 k instructions in the four-scorer kernel and it waited on instruction fetch
#pragma unroll kChunkUnroll
      for (int v4 = 0; v4 < 8; ++v4) {
        const float4 x = *reinterpret_cast<const float4 *>(row + v4 * 4);
        const uint32_t i0 = v4 * 4;
        src_update<ALGS, true, 0>(s, (double)x.x, alpha, ab, 0u, lag, true, &dlag, i0 + 0);
        src_update<ALGS, true, 1>(s, (double)x.y, alpha, ab, 0u, lag, true, &dlag, i0 + 1);
        src_update<ALGS, true, 2>(s, (double)x.z, alpha, ab, 0u, lag, true, &dlag, i0 + 2);
        src_update<ALGS, true, 3>(s, (double)x.w, alpha, ab, 0u, lag, true, &dlag, i0 + 3);
        dlag.cur_w += 4; dlag.prev_w += 4;
      }
    } else if (__all_sync(0xffffffffu, cls_src || dontcare)) {
      // whole chunk inside the model window for every lane: no range checks
#pragma unroll kChunkUnroll
      for (int v4 = 0; v4 < 8; ++v4) {
        const float4 x = *reinterpret_cast<const float4 *>(row + v4 * 4);
        const uint32_t q = cs + v4 * 4 - a0;
        src_update<ALGS>(s, (double)x.x, alpha, ab, q + 0, lag, cls_src);
        src_update<ALGS>(s, (double)x.y, alpha, ab, q + 1, lag, cls_src);
        src_update<ALGS>(s, (double)x.z, alpha, ab, q + 2, lag, cls_src);
        src_update<ALGS>(s, (double)x.w, alpha, ab, q + 3, lag, cls_src);
      }
    } else if (MODE == 1 && __all_sync(0xffffffffu, cls_cur || dontcare)) {
      // Band materialisation: each lane writes its 32 upper values over the samples it has just
      // consumed (same smem row) and its 32 lower values into a second tile; then the warp stores
      // both tiles with the mirror image of the staging pattern -- 8 lanes x 16 B per unit row --
      // so the band arrays are written in full 128 B lines instead of 4 B at a 5.7 KB stride.
      float *uprow = s_ring + (ch % kSlots) * (32 * kRowWords) + lane * kRowWords;
      float *lorow = s_lo + lane * kRowWords;
#pragma unroll kChunkUnroll
      for (int v4 = 0; v4 < 8; ++v4) {
        const float4 x = *reinterpret_cast<const float4 *>(uprow + v4 * 4);
        const uint32_t j = cs + v4 * 4 - c0;
        float4 bu, bl;
        cur_update<ALGS>(s, r, (double)x.x, j + 0, cls_cur, alpha, ab, bu.x, bl.x, band_alg, lag);
        cur_update<ALGS>(s, r, (double)x.y, j + 1, cls_cur, alpha, ab, bu.y, bl.y, band_alg, lag);
        cur_update<ALGS>(s, r, (double)x.z, j + 2, cls_cur, alpha, ab, bu.z, bl.z, band_alg, lag);
        cur_update<ALGS>(s, r, (double)x.w, j + 3, cls_cur, alpha, ab, bu.w, bl.w, band_alg, lag);
        *reinterpret_cast<float4 *>(uprow + v4 * 4) = bu;
        *reinterpret_cast<float4 *>(lorow + v4 * 4) = bl;
      }
      const unsigned rows_ok = __ballot_sync(0xffffffffu, cls_cur);
      __syncwarp();
      const float *ups = s_ring + (ch % kSlots) * (32 * kRowWords);
#pragma unroll
      for (int rr = 0; rr < 8; ++rr) {
        const unsigned rw = rr * 4 + row_sub;
        if ((rows_ok >> rw) & 1u) {
          const int64_t at = (int64_t)((s_src[rw] - (unsigned long long)reinterpret_cast<uintptr_t>(b.values)) >> 2) + cs + col4 * 4;  // unit start and chunk start are multiples of 4: 16 B aligned
          *reinterpret_cast<float4 *>(band_up + at) = *reinterpret_cast<const float4 *>(ups + rw * kRowWords + col4 * 4);
          *reinterpret_cast<float4 *>(band_lo + at) = *reinterpret_cast<const float4 *>(s_lo + rw * kRowWords + col4 * 4);
        }
      }
    } else if (MODE == 0 && roll_deep && __all_sync(0xffffffffu, (cls_cur && cs >= c0 + 32u) || dontcare)) {
      uint32_t mask[4] = {0u, 0u, 0u, 0u};
      const uint32_t j0 = cs - c0;
#pragma unroll kChunkUnroll
      for (int v4 = 0; v4 < 8; ++v4) {
        const float4 x = *reinterpret_cast<const float4 *>(row + v4 * 4);
        const uint32_t i0 = v4 * 4;
        cur_fast<ALGS, 0, true>(s, fr, (double)x.x, j0 + i0 + 0, cls_cur, alpha, ab, mask, lag, i0 + 0, &dlag);
        cur_fast<ALGS, 1, true>(s, fr, (double)x.y, j0 + i0 + 1, cls_cur, alpha, ab, mask, lag, i0 + 1, &dlag);
        cur_fast<ALGS, 2, true>(s, fr, (double)x.z, j0 + i0 + 2, cls_cur, alpha, ab, mask, lag, i0 + 2, &dlag);
        cur_fast<ALGS, 3, true>(s, fr, (double)x.w, j0 + i0 + 3, cls_cur, alpha, ab, mask, lag, i0 + 3, &dlag);
        dlag.cur_w += 4; dlag.prev_w += 4;
      }
      if (cls_cur) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
          if (mask[a]) {
            if (s.cnt[a] == 0) s.first[a] = j0 + (uint32_t)(__ffs(mask[a]) - 1);
            s.cnt[a] += __popc(mask[a]);
          }
      }
    } else if (MODE == 0 && __all_sync(0xffffffffu, cls_cur || dontcare)) {
      uint32_t mask[4] = {0u, 0u, 0u, 0u};
      const uint32_t j0 = cs - c0;
#pragma unroll kChunkUnroll
      for (int v4 = 0; v4 < 8; ++v4) {
        const float4 x = *reinterpret_cast<const float4 *>(row + v4 * 4);
        const uint32_t i0 = v4 * 4;
        cur_fast<ALGS, 0>(s, fr, (double)x.x, j0 + i0 + 0, cls_cur, alpha, ab, mask, lag, i0 + 0);
        cur_fast<ALGS, 1>(s, fr, (double)x.y, j0 + i0 + 1, cls_cur, alpha, ab, mask, lag, i0 + 1);
        cur_fast<ALGS, 2>(s, fr, (double)x.z, j0 + i0 + 2, cls_cur, alpha, ab, mask, lag, i0 + 2);
        cur_fast<ALGS, 3>(s, fr, (double)x.w, j0 + i0 + 3, cls_cur, alpha, ab, mask, lag, i0 + 3);
      }
      if (cls_cur) {  // lanes that only ride along (done / not started) keep their counters
#pragma unroll
        for (int a = 0; a < 4; ++a)
          if (mask[a]) {
            if (s.cnt[a] == 0) s.first[a] = j0 + (uint32_t)(__ffs(mask[a]) - 1);
            s.cnt[a] += __popc(mask[a]);
          }
      }
    } else if (!__all_sync(0xffffffffu, cls_skip)) {
      // generic path: chunk straddles a segment edge for at least one lane
      if (active && !done && !not_started) {
        for (int i = 0; i < kChunk; ++i) {
          const uint32_t t = cs + i;
          const double xd = (double)row[i];
          if (t >= a0 && t < a1) {
            if (t == a0) {
              s.x0 = xd; s.s1 = 0.0; s.s2 = 0.0;
              s.es = xd; s.esse = 0.0;
              s.dl = xd; s.db = 0.0; s.dsse = 0.0;
              s.rp1 = 0.0; s.rp2 = 0.0; s.rq1 = 0.0; s.rq2 = 0.0;
              if (ALGS & FM_ALG_MA_ROLLING) roll_advance<ALGS>(s, xd, 0u, lag, true);  // d = 0 like the oracle's first step
            } else {
              if (t == a0 + 1) s.db = xd - s.dl;
              src_update<ALGS>(s, xd, alpha, ab, t - a0, lag, true);
            }
          } else if (t >= c0 && t < c1) {
            if (t == c0) finalize<ALGS>(s, r, ns);
            float bu = 0.f, bl = 0.f;
            const uint32_t before = s.cnt[kSlot];
            cur_update<ALGS>(s, r, xd, t - c0, true, alpha, ab, bu, bl, band_alg, lag);
            if (BANDS && band_alg) { band_up[off + t] = bu; band_lo[off + t] = bl; }
            if (EMIT && s.cnt[kSlot] != before) {
              emit_idx[emit_base + before] = t - c0;
              emit_val[emit_base + before] = row[i];
            }
          }
        }
      }
    }
  }
  cp_async_wait<0>();

  // ---- results
  if (have && !EMIT) {
    fm_series_score *o = scores + u;
    uint32_t pts = 0;
#pragma unroll
    for (int a = 0; a < FM_N_ALG; ++a) o->first_idx[a] = FM_NONE_IDX;
    if (active) {
      const uint32_t min_pts = dp->min_anomaly_points;
      if (ALGS & FM_ALG_MA_ALL) {
        status |= 1u << (FM_S_ALG_SHIFT + 0);
        o->n_anom[0] = s.cnt[0]; o->first_idx[0] = s.first[0];
        o->ma_mean = s.ma_mean; o->ma_std = s.ma_sd; o->sigma[0] = s.ma_sd;
        if (s.cnt[0] > 0 && s.cnt[0] >= min_pts) status |= FM_S_ANOMALY;
        pts = max(pts, s.cnt[0]);
      }
      if (ALGS & FM_ALG_EWMA) {
        status |= 1u << (FM_S_ALG_SHIFT + 2);
        o->n_anom[2] = s.cnt[1]; o->first_idx[2] = s.first[1]; o->sigma[2] = s.esig;
        if (s.cnt[1] > 0 && s.cnt[1] >= min_pts) status |= FM_S_ANOMALY;
        pts = max(pts, s.cnt[1]);
      }
      if (ALGS & FM_ALG_MA_ROLLING) {
        status |= 1u << (FM_S_ALG_SHIFT + 1);
        o->n_anom[1] = s.cnt[3]; o->first_idx[1] = s.first[3]; o->sigma[1] = s.rsd;
        if (s.cnt[3] > 0 && s.cnt[3] >= min_pts) status |= FM_S_ANOMALY;
        pts = max(pts, s.cnt[3]);
      }
      if ((ALGS & FM_ALG_DES) && ns >= 2) {
        status |= 1u << (FM_S_ALG_SHIFT + 3);
        o->n_anom[3] = s.cnt[2]; o->first_idx[3] = s.first[2]; o->sigma[3] = s.dsig;
        if (s.cnt[2] > 0 && s.cnt[2] >= min_pts) status |= FM_S_ANOMALY;
        pts = max(pts, s.cnt[2]);
      }
    }
    o->status = status;
    o->k_eff = r.k;
    if (!rank_ran) { o->p_mw = 1.0; o->p_w = 1.0; o->p_k = 1.0; }
    unit_status[u] = status;
    unit_points[u] = pts;
  }
}

template <uint32_t ALGS, int MODE>
cudaError_t launch_mode(const DevBatch &b, const DevParams *dp, bool rank_ran, fm_series_score *scores,
                        uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                        const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, cudaStream_t st) {
  if (b.n_units == 0) return cudaSuccess;
  // Block size against the tail effect: 8-warp blocks sit two to an SM, 6-warp blocks three (18
  // resident warps instead of 16), so a batch that needs a second, mostly empty wave with the one
  // may fit the first wave of the other (80 k series: 0.143 -> 0.119 ms).  Large batches keep 8
  // warps, which measured 4-7 % faster on the HBM-bound single-algorithm passes.
  int sm_count = 148, dev = 0;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
  const uint64_t n_warps = (b.n_units + 31) / 32;
  auto smem_of = [](int wpb) { return ring_bytes(ALGS, wpb) + wpb * 32 * (8 + 4) + (MODE == 1 ? lo_bytes(wpb) : 0); };
  auto cost = [&](int wpb, uint64_t *waves_out) {  // waves x warps resident per SM (~96 registers per thread)
    const uint64_t per_sm = std::max<uint64_t>(1, std::min<uint64_t>((227u * 1024u) / (smem_of(wpb) + 1024), 65536u / (96u * 32u * wpb)));
    const uint64_t blocks = (n_warps + wpb - 1) / wpb, waves = (blocks + sm_count * per_sm - 1) / (sm_count * per_sm);
    if (waves_out) *waves_out = waves;
    return waves * per_sm * wpb;
  };
  uint64_t w8 = 0;
  const uint64_t c8 = cost(8, &w8), c6 = cost(6, nullptr);
  const int wpb = (w8 <= 4 && c6 < c8) ? 6 : 8;
  const unsigned grid = (unsigned)((n_warps + wpb - 1) / wpb);
  const int smem = smem_of(wpb);
  cudaError_t e = cudaFuncSetAttribute(k_band<ALGS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       smem_of(kWarpsPerBlock));
  if (e != cudaSuccess) return e;
  k_band<ALGS, MODE><<<grid, wpb * 32, smem, st>>>(b, dp, rank_ran, scores, unit_status, unit_points, band_up, band_lo,
                                                   emit_off, emit_idx, emit_val);
  return cudaGetLastError();
}

template <uint32_t ALGS>
cudaError_t launch(const DevBatch &b, const DevParams *dp, bool rank_ran, fm_series_score *scores,
                   uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo, cudaStream_t st) {
  if (band_up && band_lo)
    return launch_mode<ALGS, 1>(b, dp, rank_ran, scores, unit_status, unit_points, band_up, band_lo, nullptr, nullptr, nullptr, st);
  return launch_mode<ALGS, 0>(b, dp, rank_ran, scores, unit_status, unit_points, nullptr, nullptr, nullptr, nullptr, nullptr, st);
}

// ---- unit order for ragged batches -------------------------------------------------------------------
// k_band gives a thread a unit and a warp the longest of its 32 units' time; chunks in which the lanes are in
// different segments take the guarded path.  A batch of jobs with different window lengths therefore ran at
// half the speed of a uniform batch of the same size (scripts/ragged_time.py).  This kernel groups the units of
// every tile of 2048 by length -- one block, a counting sort in shared memory over 32 x 32 classes of (total
// length, current length) in chunks, the classes spread over the range the tile actually has -- and k_band takes
// unit order[t] instead of unit t.  A tile of equally long units (the usual batch) gets the identity; units
// stay inside their tile, so the samples a block reads stay close together.  Within a class the order is the
// order of arrival: which units share a warp may differ from run to run, what is computed for a unit does not.
constexpr int kOrderTile = 2048;

__global__ void __launch_bounds__(kOrderTile / 2) k_band_order(DevBatch b, uint32_t *__restrict__ order) {
  __shared__ uint32_t s_bin[1024];   // class sizes, then running positions
  __shared__ uint32_t s_red[4][32];  // per-warp min / max of the two lengths
  const uint64_t base = (uint64_t)blockIdx.x * kOrderTile;
  const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
  uint32_t tc[2], cc[2];  // total and current length in chunks of this thread's two units (t, t + 1024)
  uint32_t lo_t = 0xFFFFFFFFu, hi_t = 0, lo_c = 0xFFFFFFFFu, hi_c = 0;
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const uint64_t u = base + t + e * (kOrderTile / 2);
    tc[e] = cc[e] = 0xFFFFFFFFu;
    if (u < b.n_units) {
      const uint32_t nc = b.seg_lens[3 * u + 2], tot = b.seg_lens[3 * u] + b.seg_lens[3 * u + 1] + nc;
      tc[e] = tot / kChunk; cc[e] = nc / kChunk;
      lo_t = min(lo_t, tc[e]); hi_t = max(hi_t, tc[e]); lo_c = min(lo_c, cc[e]); hi_c = max(hi_c, cc[e]);
    }
  }
  s_bin[t] = 0;
  lo_t = __reduce_min_sync(0xffffffffu, lo_t); hi_t = __reduce_max_sync(0xffffffffu, hi_t);
  lo_c = __reduce_min_sync(0xffffffffu, lo_c); hi_c = __reduce_max_sync(0xffffffffu, hi_c);
  if (lane == 0) { s_red[0][warp] = lo_t; s_red[1][warp] = hi_t; s_red[2][warp] = lo_c; s_red[3][warp] = hi_c; }
  __syncthreads();
  lo_t = __reduce_min_sync(0xffffffffu, s_red[0][lane]); hi_t = __reduce_max_sync(0xffffffffu, s_red[1][lane]);
  lo_c = __reduce_min_sync(0xffffffffu, s_red[2][lane]); hi_c = __reduce_max_sync(0xffffffffu, s_red[3][lane]);
  if (lo_t >= hi_t && lo_c >= hi_c) {  // equally long units (or none): identity
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const uint64_t u = base + t + e * (kOrderTile / 2);
      if (u < b.n_units) order[u] = (uint32_t)u;
    }
    return;
  }
  // 32 classes over [lo, hi] of each length (exact when the range has at most 32 values)
  const uint32_t st = (hi_t - lo_t) / 32u + 1u, sc = (hi_c - lo_c) / 32u + 1u;
  uint32_t bin[2];
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    bin[e] = 0xFFFFFFFFu;
    if (tc[e] != 0xFFFFFFFFu) {
      bin[e] = ((tc[e] - lo_t) / st) * 32u + (cc[e] - lo_c) / sc;
      atomicAdd(&s_bin[bin[e]], 1u);
    }
  }
  __syncthreads();
  {  // exclusive scan of the 1024 class sizes: warp scans, then the warp totals
    const uint32_t v = s_bin[t];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
      if ((int)lane >= o) inc += x;
    }
    __syncthreads();
    if (lane == 31) s_red[0][warp] = inc;
    __syncthreads();
    uint32_t w = s_red[0][lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, winc, o);
      if ((int)lane >= o) winc += x;
    }
    const uint32_t before = __shfl_sync(0xffffffffu, winc - w, warp);  // units in the warps' classes before this one's
    s_bin[t] = before + inc - v;
  }
  __syncthreads();
#pragma unroll
  for (int e = 0; e < 2; ++e)
    if (bin[e] != 0xFFFFFFFFu) order[base + atomicAdd(&s_bin[bin[e]], 1u)] = (uint32_t)(base + t + e * (kOrderTile / 2));
}

}  // namespace

cudaError_t fm_launch_band_order(const DevBatch &b, uint32_t *order, cudaStream_t st) {
  if (!b.n_units) return cudaSuccess;
  k_band_order<<<(unsigned)((b.n_units + kOrderTile - 1) / kOrderTile), kOrderTile / 2, 0, st>>>(b, order);
  return cudaGetLastError();
}

// algorithms: subset of {MA_ALL, MA_ROLLING, EWMA, DES}; 0 still runs (it resolves unknown units and statuses).
cudaError_t fm_launch_band(const DevBatch &b, const DevParams *dp, uint32_t algorithms, bool rank_ran,
                           fm_series_score *scores, uint32_t *unit_status, uint32_t *unit_points,
                           float *band_up, float *band_lo, cudaStream_t st) {
  const uint32_t a = algorithms & (FM_ALG_MA_ALL | FM_ALG_MA_ROLLING | FM_ALG_EWMA | FM_ALG_DES);
#define FM_BAND_CASE(M) \
  case M: return launch<M>(b, dp, rank_ran, scores, unit_status, unit_points, band_up, band_lo, st);
  switch (a) {
    FM_BAND_CASE(0u) FM_BAND_CASE(1u) FM_BAND_CASE(2u) FM_BAND_CASE(3u) FM_BAND_CASE(4u) FM_BAND_CASE(5u)
    FM_BAND_CASE(6u) FM_BAND_CASE(7u) FM_BAND_CASE(8u) FM_BAND_CASE(9u) FM_BAND_CASE(10u) FM_BAND_CASE(11u)
    FM_BAND_CASE(12u) FM_BAND_CASE(13u) FM_BAND_CASE(14u) FM_BAND_CASE(15u)
  }
#undef FM_BAND_CASE
  return cudaErrorInvalidValue;
}

// Emit pass of the anomaly list for ONE algorithm of {MA_ALL, MA_ROLLING, EWMA, DES}: re-runs the
// same deterministic recurrences over the units that have anomalies and writes each flagged
// (current index, sample) at emit_off[u] + k.
cudaError_t fm_launch_band_emit(const DevBatch &b, const DevParams *dp, uint32_t algorithm, fm_series_score *scores,
                                const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, cudaStream_t st) {
  switch (algorithm) {
    case FM_ALG_MA_ALL: return launch_mode<FM_ALG_MA_ALL, 2>(b, dp, true, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, st);
    case FM_ALG_MA_ROLLING: return launch_mode<FM_ALG_MA_ROLLING, 2>(b, dp, true, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, st);
    case FM_ALG_EWMA: return launch_mode<FM_ALG_EWMA, 2>(b, dp, true, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, st);
    case FM_ALG_DES: return launch_mode<FM_ALG_DES, 2>(b, dp, true, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, st);
  }
  return cudaErrorInvalidValue;
}
