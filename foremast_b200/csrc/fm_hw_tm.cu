// fm_hw_tm.cu -- Holt-Winters scoring with the seasonal state in TENSOR MEMORY.
//
// Same model, same reduce-then-scan formulation and the same modes as fm_hw.cu (see there for the
// reference citations and the maths); what changes is where the period-long seasonal array lives.
// k_hw keeps it in shared memory, 8 B x period per unit (11.25 KB at period 1440), and that is what
// caps it at 12 warps per SM -- the kernel is latency-bound, so occupancy is its speed.  Here a
// super-tile is exactly one season (32*k samples, k = period/32), so lane `l` meets the SAME k
// seasonal slots in every super-tile: the array is lane-private, and Blackwell's tensor memory
// (256 KB per SM: 128 lanes x 512 32-bit columns, each thread of a warp reaching the columns of its
// own lane with tcgen05.ld/st.32x32b) holds it as plain scratch -- no tensor-core instruction is
// involved.  Shared memory then carries only the fp32 samples of the season (5.6 KB per unit) and
// one CTA of 20 warps per SM runs 20 units at a time instead of 12.  Seasonals move between tensor
// memory and registers eight at a time (16 columns), so the inner loops also lose their per-sample
// LDS/STS and all of k_hw's wrap-around index handling (storage is indexed by the position inside
// the lane's chunk, never by the absolute slot).  Seasons arrive by TMA: one cp.async.bulk per season
// into one of two buffers with an mbarrier each (4-byte cp.async when source or size is not 16 B
// aligned), the next season in flight while this one is scored.
// Used when period % 32 == 0; other periods take k_hw.
#include <algorithm>

#include "fm_hw_common.cuh"

namespace {

constexpr int kMaxWarps = 18;  // measured: 16 -> 4.51, 18 -> 4.36, 19 -> 4.53, 20 (single sample buffer) -> 5.07 ms per 200k x 10080

using fm_hw::V2;
using fm_hw::matvec;

// eight seasonals = 16 columns of this thread's tensor-memory lane
__device__ __forceinline__ void tm_ld8(uint32_t taddr, double (&d)[8]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
  for (int j = 0; j < 8; ++j) d[j] = __hiloint2double((int)r[2 * j + 1], (int)r[2 * j]);
}

__device__ __forceinline__ void tm_st8(uint32_t taddr, const double (&d)[8]) {
  uint32_t r[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    r[2 * j] = (uint32_t)__double2loint(d[j]);
    r[2 * j + 1] = (uint32_t)__double2hiint(d[j]);
  }
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// The same 16 columns as raw words, load and arrival apart: the next block's seasonals are requested
// before this block is worked on and awaited after it, so the tensor-memory round trip is hidden.
// The arrival names the registers as in/out operands: nothing that reads them can move above it.
__device__ __forceinline__ void tm_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tm_ld_arrive(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ void tm_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// ---- shared-memory reads at a register address + literal offset.  The sample and weight addresses of a
//      lane are computed once and passed through keep(), otherwise the compiler re-derives them from
//      threadIdx and the kernel parameters at every use (15 instructions per block of eight samples).
__device__ __forceinline__ uint32_t keep(uint32_t v) {
  asm volatile("" : "+r"(v));
  return v;
}
#define FM_LDS_F32(dst, addr, imm) asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(dst) : "r"(addr), "n"(imm))
#define FM_LDS_V2F64(dst, addr, imm) \
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2+%3];" : "=d"((dst).x), "=d"((dst).y) : "r"(addr), "n"(imm))
__device__ __forceinline__ void lds8(uint32_t a, float (&x)[8]) {
  FM_LDS_F32(x[0], a, 0); FM_LDS_F32(x[1], a, 4); FM_LDS_F32(x[2], a, 8); FM_LDS_F32(x[3], a, 12);
  FM_LDS_F32(x[4], a, 16); FM_LDS_F32(x[5], a, 20); FM_LDS_F32(x[6], a, 24); FM_LDS_F32(x[7], a, 28);
}
__device__ __forceinline__ void ldw4(uint32_t a, double2 (&w)[4]) {
  FM_LDS_V2F64(w[0], a, 0); FM_LDS_V2F64(w[1], a, 16); FM_LDS_V2F64(w[2], a, 32); FM_LDS_V2F64(w[3], a, 48);
}
__device__ __forceinline__ double tm_get(const uint32_t (&r)[16], int j) {
  return __hiloint2double((int)r[2 * j + 1], (int)r[2 * j]);
}
__device__ __forceinline__ void tm_put(uint32_t (&r)[16], int j, double v) {
  r[2 * j] = (uint32_t)__double2loint(v);
  r[2 * j + 1] = (uint32_t)__double2hiint(v);
}

// ---- TMA: one 1-D bulk copy (cp.async.bulk, UBLKCP) brings a whole season into shared memory and
//      signals an mbarrier, instead of 45 four-byte cp.async per lane; needs 16 B aligned source and size
__device__ __forceinline__ void mbar_init(uint64_t *bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void bulk_load(float *dst, const float *src, uint32_t bytes, uint64_t *bar) {
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar), d = (uint32_t)__cvta_generic_to_shared(dst);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // earlier generic reads of dst come first
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(d),
               "l"(src), "r"(bytes), "r"(b)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(b),
      "r"(parity)
      : "memory");
}

// MODE as in fm_hw.cu: 0 score, 1 score + bands, 2 emit pass of the anomaly list, 3 fit (model
// window only, state saved), 4 resume (score the current segment from the saved state, advance it).
template <int MODE>
__global__ void __launch_bounds__(kMaxWarps * 32, 1) k_hw_tm(DevBatch b, const DevParams *__restrict__ dp,
                                                             fm_series_score *__restrict__ scores,
                                                             uint32_t *__restrict__ unit_status,
                                                             uint32_t *__restrict__ unit_points,
                                                             float *__restrict__ band_up, float *__restrict__ band_lo,
                                                             const unsigned long long *__restrict__ emit_off,
                                                             uint32_t *__restrict__ emit_idx, float *__restrict__ emit_val,
                                                             double *__restrict__ state) {
  constexpr bool BANDS = MODE == 1, EMIT = MODE == 2, FIT = MODE == 3, RESUME = MODE == 4;
  extern __shared__ float s_x[];  // [warps][2][period] fp32 samples: the season being scored and the next one in flight
  __shared__ uint32_t s_tm_base;
  __shared__ double s_apow[5 * 4];
  __shared__ double2 s_w[64];  // chunk weights (DevParams::hw_w)
  __shared__ __align__(8) uint64_t s_bar[kMaxWarps][2];  // one mbarrier per warp per sample buffer
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
  const uint32_t m = dp->period, k = m / 32, kb = (k + 7) / 8;  // kb blocks of eight seasonals per lane
  float *const Xb = s_x + (size_t)warp * 2 * m;

  // ---- tensor memory: the CTA takes all 512 columns; warp w works in lane quarter w % 4 (the only
  //      lanes its tcgen05.ld/st can reach) and column range (w / 4) * 16 * kb
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&s_tm_base)),
                 "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (threadIdx.x < 20) s_apow[threadIdx.x] = dp->hw_pow[threadIdx.x];
  if (threadIdx.x < 64) s_w[threadIdx.x] = make_double2(dp->hw_w[2 * threadIdx.x], dp->hw_w[2 * threadIdx.x + 1]);
  if (lane == 0) {
    mbar_init(&s_bar[warp][0]);
    mbar_init(&s_bar[warp][1]);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tm_base = s_tm_base;
  const uint32_t tS = tm_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * 16u * kb;

  const double alpha = dp->alpha, ab = dp->alpha * dp->beta, g1 = dp->gamma * (1.0 - dp->alpha);
  const double *const Apow = s_apow;  // A^(k*2^i), i = 0..4
  const double Alane[4] = {dp->hw_lane[4 * lane], dp->hw_lane[4 * lane + 1], dp->hw_lane[4 * lane + 2],
                           dp->hw_lane[4 * lane + 3]};  // A^(k*lane)
  const uint32_t p0 = lane * k;  // this lane's chunk inside a season-long super-tile
  const uint32_t xs0 = keep((uint32_t)__cvta_generic_to_shared(Xb + p0));  // ... as a shared-window address, buffer 0
  const uint32_t ws0 = keep((uint32_t)__cvta_generic_to_shared(s_w));
  const uint32_t buf_bytes = m * 4u;
  uint32_t par = 0, bulk = 0;    // per sample buffer: mbarrier phase to wait for next / filled by a bulk copy

  for (uint64_t u = (uint64_t)blockIdx.x * n_warps + warp; u < b.n_units; u += (uint64_t)gridDim.x * n_warps) {
    fm_hw::Unit w;
    if (!fm_hw::unit_begin<MODE>(b, dp, u, lane, m, scores, unit_status, unit_points, emit_off, state, w)) continue;
    const int64_t off = w.off;
    const uint32_t a0 = w.a0, ns = w.ns, c0 = w.c0, n_stream = w.n_stream, bound = w.bound;
    const float *const unit = w.unit;
    double *const ust = w.ust;
    const double kk = w.kk, mlb = w.mlb;

    // Stream positions [q0s, q0s + m) into sample buffer `buf`: one TMA bulk copy when source and
    // size are 16 B aligned (the usual case: model window = history, whole seasons), else 4-byte
    // async copies, all in flight together.  Completion is awaited later (wait_buf).
    auto stage = [&](uint32_t q0s, uint32_t buf) {
      float *const dst = Xb + buf * m;
      const uint32_t nv = min(m, n_stream - q0s);
      if (q0s + nv <= ns || q0s >= ns) {  // one segment: one contiguous source
        const float *src = unit + (q0s < ns ? a0 + q0s : c0 + (q0s - ns));
        if (((reinterpret_cast<uintptr_t>(src) & 15u) | (nv & 3u)) == 0) {
          if (lane == 0) bulk_load(dst, src, nv * 4u, &s_bar[warp][buf]);
          bulk |= 1u << buf;
          return;
        }
#pragma unroll 3
        for (uint32_t i = lane; i < nv; i += 32) cp_async4(dst + i, src + i);
      } else {
        for (uint32_t i = lane; i < nv; i += 32) {
          const uint32_t q = q0s + i;
          cp_async4(dst + i, unit + (q < ns ? a0 + q : c0 + (q - ns)));
        }
      }
      cp_async_commit();
      bulk &= ~(1u << buf);
    };
    auto wait_buf = [&](uint32_t buf) {
      if (bulk & (1u << buf)) {
        mbar_wait(&s_bar[warp][buf], (par >> buf) & 1u);
        par ^= 1u << buf;
      } else {
        cp_async_wait<0>();
      }
      __syncwarp();
    };

    V2 carry;
    double sse = 0.0, sigma = 0.0;
    bool sigma_ready = false;
    uint32_t phase = 0;  // seasonal slot of stream position 0
    uint32_t cur = 0;  // buffer that holds (or is receiving) the super-tile at q0
    bool landed = false;
    __syncwarp();      // the previous unit is done with both buffers
    // sum of this lane's k samples of the season at shared address xa (two accumulators)
    auto chunk_sum = [&](uint32_t xa) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll 1
      for (uint32_t blk = 0; blk < kb; ++blk, xa += 32u) {
        float x[8];
        lds8(xa, x);
        const uint32_t rem = k - 8 * blk;  // warp-uniform
        if (rem >= 8) {
#pragma unroll
          for (int jj = 0; jj < 8; jj += 2) { s0 += (double)x[jj]; s1 += (double)x[jj + 1]; }
        } else {
#pragma unroll
          for (int jj = 0; jj < 7; ++jj)
            if ((uint32_t)jj < rem) s0 += (double)x[jj];
        }
      }
      s0 += s1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      return s0;
    };
    if (RESUME) {
      stage(0, 0);
      carry.l = ust[0];
      carry.b = ust[1];
      sigma = ust[2];
      sigma_ready = true;
      phase = (uint32_t)ust[3];
      uint32_t slot = phase + p0;  // < 2m: phase < m, p0 + i < m
#pragma unroll 1
      for (uint32_t blk = 0; blk < kb; ++blk) {
        double s[8];
#pragma unroll
        for (int jj = 0; jj < 8; ++jj, ++slot) s[jj] = 8 * blk + jj < k ? ust[4 + (slot >= m ? slot - m : slot)] : 0.0;
        tm_st8(tS + 16 * blk, s);
      }
    } else {
      // ---- initialisation from the first two seasons (D5); both lie inside the model window.  The
      //      second one is also the first super-tile scored below, so it stays where it lands.
      stage(0, 0);
      stage(m, 1);
      wait_buf(0);
      const double L0 = chunk_sum(xs0) / (double)m;
      {
        uint32_t xa = xs0;
#pragma unroll 1
        for (uint32_t blk = 0; blk < kb; ++blk, xa += 32u) {
          float x[8];
          lds8(xa, x);
          double s[8];
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) s[jj] = 8 * blk + jj < k ? (double)x[jj] - L0 : 0.0;
          tm_st8(tS + 16 * blk, s);
        }
      }
      wait_buf(1);
      const double M2 = chunk_sum(xs0 + buf_bytes) / (double)m;
      carry.l = L0;
      carry.b = (M2 - L0) / (double)m;
      cur = 1;
      landed = true;  // the loop's first super-tile is already in its buffer
    }
    tm_wait_st();
    uint32_t n_anom = 0, first = FM_NONE_IDX;
    for (uint32_t q0 = RESUME ? 0u : m; q0 < n_stream; q0 += m, cur ^= 1u) {
      // wait for this super-tile; every lane is then done with the other buffer: start the next one
      if (landed) { landed = false; __syncwarp(); }
      else wait_buf(cur);
      if (q0 + m < n_stream) stage(q0 + m, cur ^ 1u);
      const uint32_t qa = q0 + p0;                                      // stream position of the chunk
      const uint32_t kv = qa < n_stream ? min(k, n_stream - qa) : 0u;  // real samples in this lane's chunk
      const float *const xp = Xb + cur * m + p0;
      const uint32_t xs = keep(xs0 + (cur ? buf_bytes : 0u));  // the chunk's samples as a shared-window address
      // A whole season inside the model window (all but the last one or two super-tiles of a unit) is the
      // fast case: every lane has its k samples and only the residual sum is wanted, so nothing is guarded
      // per lane; seasonals come eight at a time, the next block requested before this one is worked on.
      // Every other super-tile (the one the model window ends in, the current samples, a ragged end) takes
      // the guarded form.
      const bool whole = !RESUME && q0 + m <= ns;  // warp-uniform
      uint32_t ra[16], rb[16];

      // ---- pass 1: chunk offset from a zero state.  Fast case: a dot product with the precomputed
      //      chunk weights (two accumulators per component, no dependent chain).  Chunks past the end of
      //      the stream only feed lanes that are past the end too.
      V2 uo;
      if (whole) {
        double ul0 = 0.0, ub0 = 0.0, ul1 = 0.0, ub1 = 0.0;
        auto p1 = [&](uint32_t blk, uint32_t (&c)[16], uint32_t (&n)[16]) {
          const uint32_t xa = xs + 32u * blk, wa = ws0 + 128u * blk, rem = k - 8u * blk;
          tm_ld_arrive(c);
          if (blk + 1 < kb) tm_ld16_issue(tS + 16 * (blk + 1), n);
          float x[8];
          lds8(xa, x);
          double2 wA[4], wB[4];
          ldw4(wa, wA);
          if (rem >= 8) {
            ldw4(wa + 64u, wB);
#pragma unroll
            for (int jj = 0; jj < 4; jj += 2) {
              const double y0 = (double)x[jj] - tm_get(c, jj), y1 = (double)x[jj + 1] - tm_get(c, jj + 1);
              ul0 = fma(wA[jj].x, y0, ul0); ub0 = fma(wA[jj].y, y0, ub0);
              ul1 = fma(wA[jj + 1].x, y1, ul1); ub1 = fma(wA[jj + 1].y, y1, ub1);
            }
#pragma unroll
            for (int jj = 0; jj < 4; jj += 2) {
              const double y0 = (double)x[4 + jj] - tm_get(c, 4 + jj), y1 = (double)x[5 + jj] - tm_get(c, 5 + jj);
              ul0 = fma(wB[jj].x, y0, ul0); ub0 = fma(wB[jj].y, y0, ub0);
              ul1 = fma(wB[jj + 1].x, y1, ul1); ub1 = fma(wB[jj + 1].y, y1, ub1);
            }
          } else {  // the last block of a chunk whose length is not a multiple of eight; rem is warp-uniform
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              if ((uint32_t)jj < rem) {
                const double y0 = (double)x[jj] - tm_get(c, jj);
                ul0 = fma(wA[jj].x, y0, ul0); ub0 = fma(wA[jj].y, y0, ub0);
              }
            }
            if (rem > 4) {
              ldw4(wa + 64u, wB);
#pragma unroll
              for (int jj = 0; jj < 3; ++jj) {
                if ((uint32_t)(4 + jj) < rem) {
                  const double y0 = (double)x[4 + jj] - tm_get(c, 4 + jj);
                  ul0 = fma(wB[jj].x, y0, ul0); ub0 = fma(wB[jj].y, y0, ub0);
                }
              }
            }
          }
        };
        tm_ld16_issue(tS, ra);
#pragma unroll 1
        for (uint32_t blk = 0; blk < kb; blk += 2) {
          p1(blk, ra, rb);
          if (blk + 1 < kb) p1(blk + 1, rb, ra);
        }
        uo.l = ul0 + ul1;
        uo.b = ub0 + ub1;
      } else {
        uo.l = 0.0;
        uo.b = 0.0;
#pragma unroll 1
        for (uint32_t blk = 0; blk < kb; ++blk) {
          if (__all_sync(0xffffffffu, 8 * blk >= kv)) break;  // no lane has samples from here on
          double s[8];
          float x[8];
          lds8(xs + 32u * blk, x);
          tm_ld8(tS + 16 * blk, s);
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            if (8 * blk + jj < kv) {
              const double y = (double)x[jj] - s[jj];
              const double t = uo.l + uo.b, d = y - t;
              uo.l = fma(alpha, d, t);
              uo.b = fma(ab, d, uo.b);
            }
          }
        }
      }
      // ---- the state entering this lane's chunk: warp scan of the chunk maps v -> A^k v + u
      const V2 vin = fm_hw::scan_chunks(uo, carry, Apow, Alane, lane);

      // ---- pass 2: the real recurrence over the chunk
      double l = vin.l, bb = vin.b;
      unsigned long long fmask = 0ull;  // flags of the chunk's current samples (k <= 64)
      const bool tile_has_cur = q0 + m > ns;
      if (whole) {
        auto one = [&](float xf, uint32_t (&c)[16], int jj) {
          const double sv = tm_get(c, jj);
          const double y = (double)xf - sv;
          const double lb = l + bb, e = y - lb;
          l = fma(alpha, e, lb);
          bb = fma(ab, e, bb);
          sse = fma(e, e, sse);
          tm_put(c, jj, fma(g1, e, sv));
        };
        auto p2 = [&](uint32_t blk, uint32_t (&c)[16], uint32_t (&n)[16]) {
          const uint32_t rem = k - 8u * blk;
          tm_ld_arrive(c);
          if (blk + 1 < kb) tm_ld16_issue(tS + 16 * (blk + 1), n);
          float x[8];
          lds8(xs + 32u * blk, x);
          if (rem >= 8) {
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) one(x[jj], c, jj);
          } else {
#pragma unroll
            for (int jj = 0; jj < 7; ++jj)
              if ((uint32_t)jj < rem) one(x[jj], c, jj);
          }
          tm_st16(tS + 16 * blk, c);
        };
        tm_ld16_issue(tS, ra);
#pragma unroll 1
        for (uint32_t blk = 0; blk < kb; blk += 2) {
          p2(blk, ra, rb);
          if (blk + 1 < kb) p2(blk + 1, rb, ra);
        }
        tm_wait_st();
      } else {
        // Sweep 0 takes the samples of the model window, sweep 1 the current ones (sigma must be final
        // first); each sweep walks the seasonals block by block through registers.
        const uint32_t i_m = qa < ns ? min(kv, ns - qa) : 0u;  // model samples at the head of the chunk
        double clear = 0.0;  // |x - forecast| up to this is inside the band whatever the rounding (see below)
#pragma unroll 1
        for (int sweep = 0; sweep < 2; ++sweep) {
          if (sweep == 0 && q0 >= ns) continue;  // nothing of the model window in this super-tile
          if (sweep == 1) {
            if (!tile_has_cur) break;
            if (!sigma_ready) {  // the model window ends in this super-tile: close the residual sum first
              double t = sse;
#pragma unroll
              for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
              sigma = sqrt(t / (double)(ns - m));
              sigma_ready = true;
            }
            clear = kk * sigma * 0.999999;
          }
          const uint32_t lo = sweep ? i_m : 0u, hi = sweep ? kv : i_m;
#pragma unroll 1
          for (uint32_t blk = 0; blk < kb; ++blk) {
            if (__all_sync(0xffffffffu, 8 * blk >= hi || 8 * blk + 8 <= lo)) continue;  // no lane has work in this block
            double s[8];
            float xf[8];
            lds8(xs + 32u * blk, xf);
            tm_ld8(tS + 16 * blk, s);
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
              const uint32_t i = 8 * blk + jj;
              if (i >= lo && i < hi) {  // one sample: one-step error, state and season update, residual sum or flag
                const double x = (double)xf[jj];
                const double lb = l + bb, f = lb + s[jj], e = x - f;
                l = fma(alpha, e, lb);
                bb = fma(ab, e, bb);
                if (sweep == 0) {
                  sse = fma(e, e, sse);
                } else if (BANDS || !(fabs(e) <= clear && x >= mlb)) {
                  // Not provably inside: the oracle's arithmetic.  (Inside means |x - f| <= 0.999999 k sigma
                  // and x not below the lower clamp: x - up and lw - x are then negative by 1e-6 k sigma,
                  // their roundings are 2^-52 of |f| + k sigma + |x|, and the comparison has 1e-9 (|f| + |x|)
                  // of slack on top -- neither test can fire.)
                  const double up = fma(kk, sigma, f);
                  double lw = fma(-kk, sigma, f);
                  lw = lw < mlb ? mlb : lw;
                  // guarded comparison (DECISIONS "HW"): excursions below 1e-9 relative do not count
                  const double gx = 1e-9 * (fabs(f) + fabs(x));
                  if (((bound & 1u) && x - up > gx) || ((bound & 2u) && lw - x > gx)) fmask |= 1ull << i;
                  if (BANDS) {
                    band_up[off + c0 + (qa + i - ns)] = (float)up;
                    band_lo[off + c0 + (qa + i - ns)] = (float)lw;
                  }
                }
                s[jj] = fma(g1, e, s[jj]);
              }
            }
            tm_st8(tS + 16 * blk, s);
          }
          tm_wait_st();
        }
      }
      // next super-tile's carry (and the state the model cache keeps): the state after the last real
      // sample of this super-tile, which lives in lane (nv-1)/k
      {
        const uint32_t nv = min(m, n_stream - q0);
        const int src = whole ? 31 : (int)((nv - 1) / k);
        carry.l = __shfl_sync(0xffffffffu, l, src);
        carry.b = __shfl_sync(0xffffffffu, bb, src);
      }
      if (tile_has_cur) fm_hw::tile_flags<EMIT>(fmask, lane, qa, ns, w.emit_base, emit_idx, emit_val, xp, n_anom, first);
    }
    if (FIT || RESUME) {  // save the model: state after the last sample, residual sigma, phase, seasonals
      if (FIT && !sigma_ready) {
        double t = sse;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        sigma = sqrt(t / (double)(ns - m));
      }
      for (uint32_t blk = 0; blk < kb; ++blk) {
        double s[8];
        tm_ld8(tS + 16 * blk, s);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t i = 8 * blk + j;
          if (i < k) ust[4 + (phase + p0 + i) % m] = s[j];
        }
      }
      if (lane == 0) fm_hw::save_state_scalars(ust, carry, sigma, phase, n_stream, m);
      __syncwarp();
    }
    if (lane == 0 && !EMIT && !FIT) fm_hw::unit_finish<MODE>(dp, u, w, n_anom, first, sigma, scores, unit_status, unit_points);
  }

  // ---- give the tensor memory back
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tm_base), "n"(512));
}

template <int MODE>
cudaError_t launch_tm(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                      uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                      const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, double *state, int sm_count,
                      cudaStream_t st) {
  const uint32_t k = period / 32, kb = (k + 7) / 8;
  // warps per CTA: four lane quarters x as many 16*kb-column ranges as fit in 512 columns, capped by
  // the launch bound and by the fp32 season each warp keeps in shared memory
  int warps = (int)std::min<uint32_t>(4 * (512 / (16 * kb)), kMaxWarps);
  warps = std::min(warps, (220 * 1024 - 32) / ((int)period * 8));
  if (warps < 1) return cudaErrorInvalidValue;
  uint64_t blocks = (b.n_units + warps - 1) / warps;
  const uint64_t cap = (uint64_t)sm_count;  // persistent: one CTA is resident per SM (it owns the tensor memory)
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  const int smem = warps * 2 * (int)period * (int)sizeof(float) + 32;  // lds8 may read up to 7 samples past a chunk
  cudaError_t e = cudaFuncSetAttribute(k_hw_tm<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  k_hw_tm<MODE><<<(unsigned)blocks, warps * 32, smem, st>>>(b, dp, scores, unit_status, unit_points, band_up, band_lo,
                                                           emit_off, emit_idx, emit_val, state);
  return cudaGetLastError();
}

}  // namespace

bool fm_hw_tm_supported(uint32_t period) { return period >= 32 && period % 32 == 0 && period <= 2048; }

cudaError_t fm_launch_hw_tm(int mode, const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                            uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                            const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, double *state,
                            int sm_count, cudaStream_t st) {
  switch (mode) {
    case 0: return launch_tm<0>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx, emit_val, state, sm_count, st);
    case 1: return launch_tm<1>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx, emit_val, state, sm_count, st);
    case 2: return launch_tm<2>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx, emit_val, state, sm_count, st);
    case 3: return launch_tm<3>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx, emit_val, state, sm_count, st);
    case 4: return launch_tm<4>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx, emit_val, state, sm_count, st);
  }
  return cudaErrorInvalidValue;
}
