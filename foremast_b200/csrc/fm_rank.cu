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
// inside a lane are two FMNMX/VIMNMX, the 15 lane-crossing stages use SHFL.BFLY -- so the sort
// never touches memory.  Baseline and current are sorted separately and payload-free (group
// membership is recovered by merging, not carried through the sort); Wilcoxon sorts one u32 key
// per pair with the sign in bit 0.  The sorted runs go to a padded shared-memory tile and the 32
// lanes split the merged order by merge path (fm_rank_merge.h) to accumulate exact int64 rank
// sums and tie terms.  Reads are plain coalesced 4-byte loads: segment starts are not 16 B
// aligned in general, and this kernel is issue-bound on the sort, not on HBM.
// Algorithmic bytes per unit: 4*(n_base+n_cur) + 8 + 12 + 64.
#include <math_constants.h>

#include "fm_common.cuh"
#include "fm_rank_merge.h"

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr uint32_t kKeySentinel = 0xFFFFFFFFu;

__device__ __forceinline__ float sel_minmax(bool lower, float a, float b) { return lower ? fminf(a, b) : fmaxf(a, b); }
__device__ __forceinline__ uint32_t sel_minmax(bool lower, uint32_t a, uint32_t b) { return lower ? min(a, b) : max(a, b); }
__device__ __forceinline__ void cex(float &a, float &b) { float lo = fminf(a, b), hi = fmaxf(a, b); a = lo; b = hi; }
__device__ __forceinline__ void cex(uint32_t &a, uint32_t &b) { uint32_t lo = min(a, b), hi = max(a, b); a = lo; b = hi; }

// Ascending bitonic sort of 32*R keys held as v[r] = element lane*R + r.  Every merge starts
// with a "mirror" step (partner = e ^ (k-1)) so all compare-exchanges point the same way and no
// direction flags are needed; steps whose stride stays below R are pure register work.
template <int R, typename T>
__device__ __forceinline__ void warp_bitonic_sort(T (&v)[R], const unsigned lane) {
#pragma unroll
  for (int k = 2; k <= 32 * R; k <<= 1) {
    if (k <= R) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int pr = r ^ (k - 1);
        if (pr > r) cex(v[r], v[pr]);
      }
    } else {
      const unsigned lm = k / R - 1;
      const bool lower = (lane & (k / R / 2)) == 0;
      if (R == 1) {
        T a = __shfl_xor_sync(0xffffffffu, v[0], lm);
        v[0] = sel_minmax(lower, v[0], a);
      } else {
#pragma unroll
        for (int r = 0; r < R / 2; ++r) {
          T a = __shfl_xor_sync(0xffffffffu, v[R - 1 - r], lm);
          T c = __shfl_xor_sync(0xffffffffu, v[r], lm);
          v[r] = sel_minmax(lower, v[r], a);
          v[R - 1 - r] = sel_minmax(lower, v[R - 1 - r], c);
        }
      }
    }
#pragma unroll
    for (int j = k >> 2; j >= 1; j >>= 1) {
      if (j < R) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int pr = r ^ j;
          if (pr > r) cex(v[r], v[pr]);
        }
      } else {
        const unsigned lm = j / R;
        const bool lower = (lane & lm) == 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          T a = __shfl_xor_sync(0xffffffffu, v[r], lm);
          v[r] = sel_minmax(lower, v[r], a);
        }
      }
    }
  }
}

// p-values: same expressions, same order as oracle/DECISIONS.md "p-values".
__device__ double p_mann_whitney(long long u1x2, uint32_t n1, uint32_t n2, long long ties) {
  double nn = (double)n1 * (double)n2, N = (double)n1 + (double)n2;
  double var = nn / 12.0 * ((N + 1.0) - (double)ties / (N * (N - 1.0)));
  if (!(var > 0.0)) return 1.0;
  long long u2x2 = 2ll * n1 * n2 - u1x2;
  long long umax = u1x2 > u2x2 ? u1x2 : u2x2;
  double z = ((double)(umax - (long long)n1 * n2) * 0.5 - 0.5) / sqrt(var);
  double p = erfc(z * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

__device__ double p_kruskal(long long r1x2, uint32_t n1, uint32_t n2, long long ties, double *h_out) {
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

__device__ double p_wilcoxon(long long wp2, long long wm2, uint32_t n, long long ties) {
  if (n == 0) return 1.0;
  double nd = (double)n;
  double se2 = nd * (nd + 1.0) * (2.0 * nd + 1.0) / 24.0 - (double)ties / 48.0;
  if (!(se2 > 0.0)) return 1.0;
  long long tmin = wp2 < wm2 ? wp2 : wm2;
  double z = ((double)tmin * 0.5 - nd * (nd + 1.0) / 4.0) / sqrt(se2);
  double p = erfc(fabs(z) * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

template <int R>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_rank(DevBatch b, const DevParams *__restrict__ dp,
                                                              fm_series_score *__restrict__ scores) {
  constexpr int kTile = 33 * R + 1;  // padded words per sorted run
  __shared__ float s_tile[kWarpsPerBlock][2][kTile];

  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  const uint32_t pairwise = dp->pairwise;
  const bool want_mw = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_MANN_WHITE;
  const bool want_w = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_WILCOXON;
  const bool want_k = pairwise == FM_PW_ALL || pairwise == FM_PW_ANY || pairwise == FM_PW_KRUSKAL;
  float *const sA = s_tile[warp][0], *const sB = s_tile[warp][1];
  uint32_t *const sK = reinterpret_cast<uint32_t *>(sA);

  for (uint64_t u = (uint64_t)blockIdx.x * kWarpsPerBlock + warp; u < b.n_units;
       u += (uint64_t)gridDim.x * kWarpsPerBlock) {
    const int64_t off = b.offsets[u];
    const uint32_t nh = b.seg_lens[3 * u], nb = b.seg_lens[3 * u + 1], nc = b.seg_lens[3 * u + 2];
    const float *base = b.values + off + nh, *cur = base + nb;
    const uint32_t m = min(nb, nc);
    const bool fits = nb <= 32u * R && nc <= 32u * R;
    const bool run_mw = want_mw && nb >= dp->min_mann_white && nc >= dp->min_mann_white && m > 0 && fits;
    const bool run_w = want_w && m >= dp->min_wilcoxon && m > 0 && fits;
    const bool run_k = want_k && nb >= dp->min_kruskal && nc >= dp->min_kruskal && m > 0 && fits;

    long long w_2wp = 0, w_ties = 0, mw_2r1 = 0, mw_ties = 0;
    uint32_t w_n = 0;

    if (run_w) {
      uint32_t key[R];
      uint32_t nz = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t e = r * 32 + lane;
        uint32_t k = kKeySentinel;
        if (e < m) {
          const float d = cur[e] - base[e];
          if (d != 0.0f) {
            k = (__float_as_uint(fabsf(d)) << 1) | (d < 0.0f ? 1u : 0u);
            ++nz;
          }
        }
        key[r] = k;
      }
      w_n = __reduce_add_sync(0xffffffffu, nz);
      warp_bitonic_sort<R, uint32_t>(key, lane);
      __syncwarp();
#pragma unroll
      for (int r = 0; r < R; ++r) sK[fm_pidx(lane * R + r)] = key[r];
      __syncwarp();
      fm_w_lane(sK, (int)w_n, (int)lane, w_2wp, w_ties);
      w_2wp = warp_sum_i64(w_2wp);
      w_ties = warp_sum_i64(w_ties);
      __syncwarp();
    }
    if (run_mw || run_k) {
#pragma unroll 1
      for (int seg = 0; seg < 2; ++seg) {
        const float *src = seg ? cur : base;
        const uint32_t n = seg ? nc : nb;
        float *dst = seg ? sB : sA;
        float v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint32_t e = r * 32 + lane;
          v[r] = e < n ? src[e] : CUDART_INF_F;
        }
        warp_bitonic_sort<R, float>(v, lane);
#pragma unroll
        for (int r = 0; r < R; ++r) dst[fm_pidx(lane * R + r)] = v[r];
      }
      __syncwarp();
      fm_mw_lane(sA, (int)nb, sB, (int)nc, (int)lane, mw_2r1, mw_ties);
      mw_2r1 = warp_sum_i64(mw_2r1);
      mw_ties = warp_sum_i64(mw_ties);
      __syncwarp();
    }

    // ---- p-values on lanes 0..2, combiner on lane 0
    const long long mw_2u1 = mw_2r1 - (long long)nb * ((long long)nb + 1);
    const long long w_2wm = (long long)w_n * ((long long)w_n + 1) - w_2wp;
    double p = 1.0, kw_h = 0.0;
    if (lane == 0 && run_mw) p = p_mann_whitney(mw_2u1, nb, nc, mw_ties);
    if (lane == 1 && run_w) p = p_wilcoxon(w_2wp, w_2wm, w_n, w_ties);
    if (lane == 2 && run_k) p = p_kruskal(mw_2r1, nb, nc, mw_ties, &kw_h);
    const double p_mw = __shfl_sync(0xffffffffu, p, 0), p_w = __shfl_sync(0xffffffffu, p, 1);
    const double p_k = __shfl_sync(0xffffffffu, p, 2);
    kw_h = __shfl_sync(0xffffffffu, kw_h, 2);
    if (lane == 0) {
      const double alpha = dp->pairwise_threshold;
      const int n_run = (int)run_mw + (int)run_w + (int)run_k;
      const int n_rej = (int)(run_mw && p_mw < alpha) + (int)(run_w && p_w < alpha) + (int)(run_k && p_k < alpha);
      bool differs = false;
      if (pairwise == FM_PW_ALL) differs = n_run > 0 && n_rej == n_run;
      else if (pairwise != FM_PW_NONE) differs = n_rej > 0;
      uint32_t status = (differs ? FM_S_DIFFERS : 0u) | (run_mw ? FM_S_MW_RUN : 0u) |
                        (run_w ? FM_S_WILCOXON_RUN : 0u) | (run_k ? FM_S_KRUSKAL_RUN : 0u);
      fm_series_score *o = scores + u;
      o->status = status;
      o->w_n = w_n;
      const bool pooled = run_mw || run_k;
      o->mw_2u1 = pooled ? mw_2u1 : 0;
      o->mw_2r1 = pooled ? mw_2r1 : 0;
      o->mw_ties = pooled ? mw_ties : 0;
      o->w_2wplus = w_2wp;
      o->w_2wminus = run_w ? w_2wm : 0;
      o->w_ties = w_ties;
      o->p_mw = p_mw;
      o->p_w = p_w;
      o->p_k = p_k;
      o->kw_h = kw_h;
    }
  }
}

template <int R>
cudaError_t launch(const DevBatch &b, const DevParams *dp, fm_series_score *scores, int sm_count, cudaStream_t st) {
  uint64_t blocks = (b.n_units + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const uint64_t cap = (uint64_t)sm_count * 64;  // grid-stride beyond a few waves
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  k_rank<R><<<(unsigned)blocks, kWarpsPerBlock * 32, 0, st>>>(b, dp, scores);
  return cudaGetLastError();
}

}  // namespace

// max_seg = longest baseline / current segment in the batch; segments above 1024 samples are
// not ranked by this kernel (the caller reports FM_E_UNSUPPORTED).
cudaError_t fm_launch_rank(const DevBatch &b, const DevParams *dp, uint32_t max_seg, fm_series_score *scores,
                           int sm_count, cudaStream_t st) {
  if (max_seg <= 32) return launch<1>(b, dp, scores, sm_count, st);
  if (max_seg <= 64) return launch<2>(b, dp, scores, sm_count, st);
  if (max_seg <= 128) return launch<4>(b, dp, scores, sm_count, st);
  if (max_seg <= 256) return launch<8>(b, dp, scores, sm_count, st);
  if (max_seg <= 512) return launch<16>(b, dp, scores, sm_count, st);
  if (max_seg <= 1024) return launch<32>(b, dp, scores, sm_count, st);
  return cudaErrorInvalidValue;
}
