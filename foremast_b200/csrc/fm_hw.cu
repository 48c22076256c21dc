// fm_hw.cu -- Holt-Winters additive seasonal scoring: "Holt-Winters" in the
// brain's model list (docs/guides/design.md:71) over the 7-day history that barrelman requests
// (foremast-barrelman/pkg/client/metrics/metricsquery.go:95-97), period 1440 = one day at 60 s.
//
// Mapping: ONE WARP PER SERIES-WINDOW.  The seasonal array (period fp64 values, 11.25 KB at 1440)
// lives in shared memory, so a thread-per-series layout is impossible.  Given last season's S,
// level and trend obey a 2-state linear recurrence with constant matrix
//     [l'; b'] = A [l; b] + g * (x - S),  A = [[1-a, 1-a], [-ab, 1-ab]],  g = [a; ab]
// The stream is cut into super-tiles of 32*k samples (k = period/32, so a super-tile never
// touches a seasonal slot twice) and each lane owns k CONSECUTIVE samples of it:
//   pass 1  every lane runs the recurrence over its k samples from a zero state (sequential,
//           5 fp64 ops per sample, no shuffles) -> its chunk's affine offset u;
//   scan    one inclusive warp scan of the 32 offsets with the precomputed A^(k*2^i), plus the
//           carry of the previous super-tile through A^(k*lane) -> the true state entering each chunk;
//   pass 2  every lane re-runs its k samples from that state: one-step error, season update,
//           residual sum / flags.
// This reduce-then-scan form does ~12 fp64 ops per sample; the first version scanned every
// 32-sample tile across lanes (five SHFL+2x2-matvec rounds per tile) and was 5x slower.
// fp64 throughout; the scan re-associates the sums, so floats match the sequential oracle to
// ~1e-15 relative (tested at 1e-5), not bit for bit.
// Algorithmic bytes per unit: 4*(n_model + n_cur) + 8 + 12 + 32.
// This kernel serves the periods that cannot be cut into 16..32 equal chunks of at most 64 samples
// (fm_hw_tm_lanes; and FOREMAST_GPU_HW_SMEM=1, for tests); the others go to fm_hw_tm.cu, which keeps the
// seasonals in tensor memory.  The per-unit
// gate, the scan and the result bookkeeping are shared (fm_hw_common.cuh).
#include <cstdlib>

#include "fm_hw_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 4;

using fm_hw::V2;
using fm_hw::matvec;

// MODE: see fm_hw_common.cuh.
template <int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_hw(DevBatch b, const DevParams *__restrict__ dp,
                                                            fm_series_score *__restrict__ scores,
                                                            uint32_t *__restrict__ unit_status,
                                                            uint32_t *__restrict__ unit_points,
                                                            float *__restrict__ band_up, float *__restrict__ band_lo,
                                                            const unsigned long long *__restrict__ emit_off,
                                                            uint32_t *__restrict__ emit_idx, float *__restrict__ emit_val,
                                                            double *__restrict__ state) {
  constexpr bool BANDS = MODE == 1, EMIT = MODE == 2, FIT = MODE == 3, RESUME = MODE == 4;
  extern __shared__ double s_dyn[];  // [warps][period] seasonals, then [warps][32*k] fp32 samples
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  const uint32_t m = dp->period, k = m / 32, lt = 32 * k;
  double *const S = s_dyn + (size_t)warp * m;
  float *const X = reinterpret_cast<float *>(s_dyn + (size_t)kWarpsPerBlock * m) + (size_t)warp * lt;
  const double alpha = dp->alpha, ab = dp->alpha * dp->beta, g1 = dp->gamma * (1.0 - dp->alpha);
  // A^2 and A*g for the two-samples-per-iteration form of the recurrence
  const double a1 = 1.0 - alpha, d1 = 1.0 - ab;
  const double a2[4] = {a1 * a1 - a1 * ab, a1 * a1 + a1 * d1, -ab * a1 - ab * d1, -ab * a1 + d1 * d1};
  const double agl = a1 * alpha + a1 * ab, agb = -ab * alpha + d1 * ab;
  const double *const Apow = dp->hw_pow;               // A^(k*2^i), i = 0..4
  const double *const Alane = dp->hw_lane + 4 * lane;  // A^(k*lane)

  for (uint64_t u = (uint64_t)blockIdx.x * kWarpsPerBlock + warp; u < b.n_units;
       u += (uint64_t)gridDim.x * kWarpsPerBlock) {
    fm_hw::Unit w;
    if (!fm_hw::unit_begin<MODE>(b, dp, u, lane, m, scores, unit_status, unit_points, emit_off, state, w)) continue;
    const int64_t off = w.off;
    const uint32_t a0 = w.a0, ns = w.ns, c0 = w.c0, n_stream = w.n_stream, bound = w.bound;
    const float *const unit = w.unit;
    double *const ust = w.ust;
    const double kk = w.kk, mlb = w.mlb;
    auto sample = [&](uint32_t q) -> float { return *w.at(q); };

    V2 carry;
    double sse = 0.0, sigma = 0.0;
    bool sigma_ready = false;
    uint32_t phase = 0;  // seasonal slot of stream position 0
    if (RESUME) {
      carry.l = ust[0];
      carry.b = ust[1];
      sigma = ust[2];
      sigma_ready = true;
      phase = (uint32_t)ust[3];
      __syncwarp();
      for (uint32_t i = lane; i < m; i += 32) S[i] = ust[4 + i];
    } else {
      // ---- initialisation from the first two seasons (D5)
      double s0 = 0.0, s1 = 0.0;
      for (uint32_t i = lane; i < m; i += 32) { s0 += (double)sample(i); s1 += (double)sample(m + i); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      }
      const double L0 = s0 / (double)m, M2 = s1 / (double)m;
      carry.l = L0;
      carry.b = (M2 - L0) / (double)m;
      __syncwarp();
      for (uint32_t i = lane; i < m; i += 32) S[i] = (double)sample(i) - L0;
    }
    uint32_t n_anom = 0, first = FM_NONE_IDX;

    for (uint32_t q0 = RESUME ? 0u : m; q0 < n_stream; q0 += lt) {
      // stage the super-tile (coalesced) and find this lane's chunk
      __syncwarp();
      {  // 4-byte async copies, all in flight together (segments are only 4 B aligned in general)
        const uint32_t nv = min(lt, n_stream - q0);
        if (q0 + nv <= ns || q0 >= ns) {  // the super-tile lies in one segment: one contiguous source
          const float *src = unit + (q0 < ns ? a0 + q0 : c0 + (q0 - ns));
#pragma unroll 9
          for (uint32_t i = lane; i < nv; i += 32) cp_async4(X + i, src + i);
        } else {
          for (uint32_t i = lane; i < nv; i += 32) {
            const uint32_t q = q0 + i;
            cp_async4(X + i, unit + (q < ns ? a0 + q : c0 + (q - ns)));
          }
        }
        cp_async_commit();
        cp_async_wait<0>();
      }
      __syncwarp();
      const uint32_t p0 = lane * k, qa = q0 + p0;           // first local / stream position of the chunk
      const uint32_t idx0 = (phase + qa) % m;                // its seasonal slot
      const uint32_t kv = qa < n_stream ? min(k, n_stream - qa) : 0u;  // real samples in this lane's chunk
      // seasonal slot of local sample i; chunks that cross the end of the season take the slow loops
      const bool wrap = __any_sync(0xffffffffu, idx0 + kv > m);
      auto sidx = [&](uint32_t i) { const uint32_t j = idx0 + i; return j >= m ? j - m : j; };
      // pass 1: chunk offset from a zero state
      V2 uo;
      uo.l = 0.0;
      uo.b = 0.0;
      {
        // (chunks past the end of the stream only feed lanes that are past the end too)
        // Two samples per iteration: v'' = A^2 v + (A g y1 + g y2).  The bracket does not depend on
        // the state, so the dependent chain is one 2x2 mat-vec per TWO samples instead of four
        // dependent fp64 ops per sample -- this kernel is fp64-latency bound.
        uint32_t i = 0;
        if (!wrap) {
          const float *xp = X + p0;
          const double *sp = S + idx0;
#pragma unroll 4
          for (; i + 1 < kv; i += 2) {
            const double y1 = (double)xp[i] - sp[i], y2 = (double)xp[i + 1] - sp[i + 1];
            const double cl = fma(agl, y1, alpha * y2), cb = fma(agb, y1, ab * y2);
            const double nl = fma(a2[0], uo.l, fma(a2[1], uo.b, cl));
            const double nb2 = fma(a2[2], uo.l, fma(a2[3], uo.b, cb));
            uo.l = nl;
            uo.b = nb2;
          }
        } else {
          for (; i + 1 < kv; i += 2) {
            const double y1 = (double)X[p0 + i] - S[sidx(i)], y2 = (double)X[p0 + i + 1] - S[sidx(i + 1)];
            const double cl = fma(agl, y1, alpha * y2), cb = fma(agb, y1, ab * y2);
            const double nl = fma(a2[0], uo.l, fma(a2[1], uo.b, cl));
            const double nb2 = fma(a2[2], uo.l, fma(a2[3], uo.b, cb));
            uo.l = nl;
            uo.b = nb2;
          }
        }
        if (i < kv) {
          const double y = (double)X[p0 + i] - S[sidx(i)];
          const double t = uo.l + uo.b, w = y - t;
          uo.l = fma(alpha, w, t);
          uo.b = fma(ab, w, uo.b);
        }
      }
      // the state entering this lane's chunk: warp scan of the chunk maps v -> A^k v + u
      const V2 vin = fm_hw::scan_chunks(uo, carry, Apow, Alane, lane);
      // pass 2: the real recurrence over the chunk; flags collected in a 64-bit mask (k <= 64)
      double l = vin.l, bb = vin.b;
      unsigned long long fmask = 0ull;
      uint32_t i = 0;
      const bool tile_has_cur = q0 + lt > ns;
      for (int half = 0; half < 2; ++half) {
        // half 0: samples of the model window; half 1: current samples (sigma must be final first)
        const uint32_t i_end = half == 0 ? (qa < ns ? min(kv, ns - qa) : 0u) : kv;
        // pairs: the state two samples ahead comes straight from A^2 (short dependent chain); the
        // state one sample ahead, the two one-step errors and the season updates hang off it
        auto pair_step = [&](uint32_t ii, const float *xq, double *sq1, double *sq2) {
          const double x1 = (double)xq[0], x2 = (double)xq[1], s1 = *sq1, s2 = *sq2;
          const double y1 = x1 - s1, y2 = x2 - s2;
          const double lb1 = l + bb, e1 = y1 - lb1;
          const double l1 = fma(alpha, e1, lb1), b1 = fma(ab, e1, bb);      // state after the first sample
          const double cl = fma(agl, y1, alpha * y2), cb = fma(agb, y1, ab * y2);
          const double l2 = fma(a2[0], l, fma(a2[1], bb, cl)), b2 = fma(a2[2], l, fma(a2[3], bb, cb));
          const double lb2 = l1 + b1, e2 = y2 - lb2;
          *sq1 = fma(g1, e1, s1);
          *sq2 = fma(g1, e2, s2);
          if (half == 0) {
            sse = fma(e1, e1, sse);
            sse = fma(e2, e2, sse);
          } else {
            const double f1 = lb1 + s1, f2 = lb2 + s2;
            const double up1 = fma(kk, sigma, f1), up2 = fma(kk, sigma, f2);
            double lo1 = fma(-kk, sigma, f1), lo2 = fma(-kk, sigma, f2);
            lo1 = lo1 < mlb ? mlb : lo1;
            lo2 = lo2 < mlb ? mlb : lo2;
            // guarded comparison (DECISIONS "HW"): excursions below 1e-9 relative do not count
            const double g1x = 1e-9 * (fabs(f1) + fabs(x1)), g2x = 1e-9 * (fabs(f2) + fabs(x2));
            if (((bound & 1u) && x1 - up1 > g1x) || ((bound & 2u) && lo1 - x1 > g1x)) fmask |= 1ull << ii;
            if (((bound & 1u) && x2 - up2 > g2x) || ((bound & 2u) && lo2 - x2 > g2x)) fmask |= 2ull << ii;
            if (BANDS) {
              const int64_t at = off + c0 + (qa + ii - ns);
              band_up[at] = (float)up1; band_lo[at] = (float)lo1;
              band_up[at + 1] = (float)up2; band_lo[at + 1] = (float)lo2;
            }
          }
          l = l2;
          bb = b2;
        };
        if (!wrap) {
          const float *xp = X + p0;
          double *sp = S + idx0;
#pragma unroll 2
          for (; i + 1 < i_end; i += 2) pair_step(i, xp + i, sp + i, sp + i + 1);
        } else {
          for (; i + 1 < i_end; i += 2) pair_step(i, X + p0 + i, S + sidx(i), S + sidx(i + 1));
        }
        for (; i < i_end; ++i) {
          const uint32_t idx = sidx(i);
          const double x = (double)X[p0 + i], seas = S[idx];
          const double lb = l + bb, f = lb + seas, e = x - f;
          l = fma(alpha, e, lb);
          bb = fma(ab, e, bb);
          S[idx] = fma(g1, e, seas);
          if (half == 0) {
            sse = fma(e, e, sse);
          } else {
            const double up = fma(kk, sigma, f);
            double lo = fma(-kk, sigma, f);
            lo = lo < mlb ? mlb : lo;
            const double gx = 1e-9 * (fabs(f) + fabs(x));
            if (((bound & 1u) && x - up > gx) || ((bound & 2u) && lo - x > gx)) fmask |= 1ull << i;
            if (BANDS) {
              band_up[off + c0 + (qa + i - ns)] = (float)up;
              band_lo[off + c0 + (qa + i - ns)] = (float)lo;
            }
          }
        }
        if (half == 0) {
          if (!tile_has_cur) break;
          if (!sigma_ready) {  // the model window ends in this super-tile: close the residual sum first
            double t = sse;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            sigma = sqrt(t / (double)(ns - m));
            sigma_ready = true;
          }
        }
      }
      // next super-tile's carry (and the state the model cache keeps): the state after the last real
      // sample of this super-tile, which lives in lane (nv-1)/k
      {
        const uint32_t nv = min(lt, n_stream - q0);
        const int src = (int)((nv - 1) / k);
        carry.l = __shfl_sync(0xffffffffu, l, src);
        carry.b = __shfl_sync(0xffffffffu, bb, src);
      }
      if (tile_has_cur) fm_hw::tile_flags<EMIT>(fmask, lane, qa, ns, w.emit_base, emit_idx, emit_val, X + p0, n_anom, first);
    }
    if (FIT || RESUME) {  // save the model: state after the last sample, residual sigma, phase, seasonals
      if (FIT && !sigma_ready) {
        double t = sse;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        sigma = sqrt(t / (double)(ns - m));
      }
      __syncwarp();
      for (uint32_t i = lane; i < m; i += 32) ust[4 + i] = S[i];
      if (lane == 0) fm_hw::save_state_scalars(ust, carry, sigma, phase, n_stream, m);
      __syncwarp();
    }
    if (lane == 0 && !EMIT && !FIT) fm_hw::unit_finish<MODE>(dp, u, w, n_anom, first, sigma, scores, unit_status, unit_points);
  }
}

}  // namespace

template <int MODE>
static cudaError_t launch_hw(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                             uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                             const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, int sm_count,
                             cudaStream_t st, double *state = nullptr) {
  if (fm_hw_tm_supported(period) && !getenv("FOREMAST_GPU_HW_SMEM"))  // seasonals in tensor memory: 20 units per SM instead of 12
    return fm_launch_hw_tm(MODE, b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx,
                           emit_val, state, sm_count, st);
  uint64_t blocks = (b.n_units + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const uint64_t cap = (uint64_t)sm_count * 32;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  const int smem = kWarpsPerBlock * ((int)period * (int)sizeof(double) + 32 * ((int)period / 32) * (int)sizeof(float));
  cudaError_t e = cudaFuncSetAttribute(k_hw<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  k_hw<MODE><<<(unsigned)blocks, kWarpsPerBlock * 32, smem, st>>>(b, dp, scores, unit_status, unit_points, band_up, band_lo,
                                                               emit_off, emit_idx, emit_val, state);
  return cudaGetLastError();
}

// Runs after fm_launch_band (it merges into the statuses that kernel wrote).  period must be in
// [32, 2048]; the caller checks.
cudaError_t fm_launch_hw(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                         uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                         int sm_count, cudaStream_t st) {
  if (band_up && band_lo)
    return launch_hw<1>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, nullptr, nullptr, nullptr, sm_count, st);
  return launch_hw<0>(b, dp, period, scores, unit_status, unit_points, nullptr, nullptr, nullptr, nullptr, nullptr, sm_count, st);
}

cudaError_t fm_launch_hw_emit(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                              const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, int sm_count,
                              cudaStream_t st) {
  return launch_hw<2>(b, dp, period, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, sm_count, st);
}

// Continuous-mode model cache: fit = model window only, state saved; resume = score the current
// segment from the saved state and advance it.  state holds (period + 4) doubles per unit.
cudaError_t fm_launch_hw_fit(const DevBatch &b, const DevParams *dp, uint32_t period, double *state, int sm_count,
                             cudaStream_t st) {
  return launch_hw<3>(b, dp, period, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, sm_count, st, state);
}

cudaError_t fm_launch_hw_resume(const DevBatch &b, const DevParams *dp, uint32_t period, double *state,
                                fm_series_score *scores, uint32_t *unit_status, uint32_t *unit_points, int sm_count,
                                cudaStream_t st) {
  return launch_hw<4>(b, dp, period, scores, unit_status, unit_points, nullptr, nullptr, nullptr, nullptr, nullptr, sm_count, st, state);
}
