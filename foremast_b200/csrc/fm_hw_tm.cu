// fm_hw_tm.cu -- Holt-Winters scoring with the seasonal state in TENSOR MEMORY.
//
// Same model, same reduce-then-scan formulation and the same modes as fm_hw.cu (see there for the
// reference citations and the maths); what changes is where the period-long seasonal array lives.
// k_hw keeps it in shared memory, 8 B x period per unit (11.25 KB at period 1440), and that is what
// caps it at 12 warps per SM -- the kernel is latency-bound, so occupancy is its speed.  Here a
// super-tile is exactly one season (L chunks of k samples, one per lane), so lane `l` meets the SAME k
// seasonal slots in every super-tile: the array is lane-private, and Blackwell's tensor memory
// (256 KB per SM: 128 lanes x 512 32-bit columns, each thread of a warp reaching the columns of its
// own lane with tcgen05.ld/st.32x32b) holds it as plain scratch -- no tensor-core instruction is
// involved.  Shared memory then carries only the fp32 samples of the season (5.6 KB per unit) and
// one CTA of 20 warps per SM runs 20 units at a time instead of 12.  Seasonals move between tensor
// memory and registers four at a time (8 columns), the next four requested before these are worked
// on.  Seasons arrive by TMA: one cp.async.bulk per season into one of two buffers with an mbarrier
// each (4-byte cp.async when source or size is not 16 B aligned), the next season in flight while
// this one is scored, and the next unit's first two seasons pulled into L2 meanwhile.
//
// A season that lies wholly inside the model window -- all but the last one or two super-tiles of a
// unit -- takes an unguarded path: the chunk offsets of pass 1 are a dot product with precomputed
// weights (DevParams::hw_w), pass 2 is the plain one-step recurrence.  The super-tile the model window
// ends in, the current samples and ragged ends take the guarded form, where a current sample whose
// error is provably inside the band skips the band arithmetic.
// A season is cut into L chunks of k samples, L the largest divisor of the period up to 32 (fm_hw_tm_lanes: 1440 = 32 x 45,
// 1000 = 25 x 40, 168 = 28 x 6); lanes past L ride along on lane 0's samples and nothing they compute is kept.
// Periods without such a divisor (k > 64 or fewer than 16 lanes) take k_hw.
#include <algorithm>

#include "fm_hw_common.cuh"

namespace {

constexpr int kMaxWarps = 20;  // per 200k x 10080: 18 -> 3.30, 19 -> 3.15, 20 -> 3.01 ms (tensor memory holds 20 x 96 columns, shared memory 20 x 2 seasons)

using fm_hw::V2;
using fm_hw::matvec;

// ---- tensor memory: four seasonals = 8 columns of this thread's lane.  Load and arrival are apart:
//      the next block is requested before this one is worked on and awaited after it.  The arrival
//      names the registers as in/out operands, so nothing that reads them can move above it.
__device__ __forceinline__ void tm_ld_issue(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tm_ld_arrive(uint32_t (&r)[8]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
               :
               : "memory");
}
__device__ __forceinline__ void tm_st(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n" ::"r"(taddr), "r"(r[0]),
               "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ double tm_get(const uint32_t (&r)[8], int j) {
  return __hiloint2double((int)r[2 * j + 1], (int)r[2 * j]);
}
__device__ __forceinline__ void tm_put(uint32_t (&r)[8], int j, double v) {
  r[2 * j] = (uint32_t)__double2loint(v);
  r[2 * j + 1] = (uint32_t)__double2hiint(v);
}
__device__ __forceinline__ void tm_ld4(uint32_t taddr, double (&d)[4]) {  // load and wait
  uint32_t r[8];
  tm_ld_issue(taddr, r);
  tm_ld_arrive(r);
#pragma unroll
  for (int j = 0; j < 4; ++j) d[j] = tm_get(r, j);
}
__device__ __forceinline__ void tm_st4(uint32_t taddr, const double (&d)[4]) {
  uint32_t r[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) tm_put(r, j, d[j]);
  tm_st(taddr, r);
}

// ---- shared-memory reads at a register address + literal offset.  The addresses a lane needs are
//      computed once and passed through keep(): left alone, the compiler re-derives them from threadIdx
//      and the kernel parameters at every use (15 instructions per block of samples).
__device__ __forceinline__ uint32_t keep(uint32_t v) {
  asm volatile("" : "+r"(v));
  return v;
}
#define FM_LDS_F32(dst, addr, imm) asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(dst) : "r"(addr), "n"(imm))
#define FM_LDS_V2F64(dst, addr, imm) \
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2+%3];" : "=d"((dst).x), "=d"((dst).y) : "r"(addr), "n"(imm))
template <bool VEC>  // VEC: the address is 16 B aligned (chunk lengths that are multiples of four)
__device__ __forceinline__ void lds4(uint32_t a, float (&x)[4]) {
  if constexpr (VEC) {
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]) : "r"(a));
  } else {
    FM_LDS_F32(x[0], a, 0); FM_LDS_F32(x[1], a, 4); FM_LDS_F32(x[2], a, 8); FM_LDS_F32(x[3], a, 12);
  }
}
__device__ __forceinline__ void ldw4(uint32_t a, double2 (&w)[4]) {
  FM_LDS_V2F64(w[0], a, 0); FM_LDS_V2F64(w[1], a, 16); FM_LDS_V2F64(w[2], a, 32); FM_LDS_V2F64(w[3], a, 48);
}

// ---- TMA: one 1-D bulk copy (cp.async.bulk, UBLKCP) brings a whole season into shared memory and
//      signals an mbarrier, instead of 45 four-byte cp.async per lane; needs 16 B aligned source and size.
//      All shared operands are shared-window addresses.
__device__ __forceinline__ void mbar_init(uint32_t bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_expect(uint32_t bar, uint32_t bytes) {  // one arrival, `bytes` to come
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy(uint32_t dst, const float *src, uint32_t bytes, uint32_t bar) {
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // earlier generic reads of dst come first
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// shared -> global bulk copy (the model cache's seasonals go out as one 8 * period byte store)
__device__ __forceinline__ void bulk_store(void *dst, uint32_t src, uint32_t bytes) {
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // the generic writes of src come first
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");  // src may be overwritten from here on
}
__device__ __forceinline__ double lds_f64(uint32_t a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void cp_async4_s(uint32_t dst, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(gmem) : "memory");
}

// MODE as in fm_hw.cu: 0 score, 1 score + bands, 2 emit pass of the anomaly list, 3 fit (model
// window only, state saved), 4 resume (score the current segment from the saved state, advance it).
// VEC: the chunk length is a multiple of four.  A lane's chunk then starts 16 B aligned and is read with 128-bit
// loads; when it is a multiple of eight as well, the chunks are laid out four words apart (stride k + 4), because at a
// stride of k words with k % 8 == 0 the lanes of a quarter warp would all read the same banks (period 1024: one bank
// for the whole warp).  Each lane then brings its own chunk in with its own bulk copy.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(kMaxWarps * 32, 1) k_hw_tm(DevBatch b, const DevParams *__restrict__ dp,
                                                             fm_series_score *__restrict__ scores,
                                                             uint32_t *__restrict__ unit_status,
                                                             uint32_t *__restrict__ unit_points,
                                                             float *__restrict__ band_up, float *__restrict__ band_lo,
                                                             const unsigned long long *__restrict__ emit_off,
                                                             uint32_t *__restrict__ emit_idx, float *__restrict__ emit_val,
                                                             double *__restrict__ state) {
  constexpr bool BANDS = MODE == 1, EMIT = MODE == 2, FIT = MODE == 3, RESUME = MODE == 4;
  extern __shared__ __align__(16) float s_x[];  // [warps][2][period] fp32 samples: the season being scored and the next one in flight
  __shared__ uint32_t s_tm_base;
  __shared__ double s_apow[5 * 4];    // A^(k*2^i), i = 0..4
  __shared__ double2 s_w[64];         // chunk weights (DevParams::hw_w)
  __shared__ __align__(8) uint64_t s_bar[kMaxWarps][2];  // one mbarrier per warp per sample buffer
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
  const uint32_t m = dp->period, k = dp->hw_tk, nl = dp->hw_tl;  // a season = nl chunks of k samples, one per lane
  const bool busy = lane < nl;         // lanes past the season's chunks ride along on lane 0's samples; nothing they compute is kept
  const uint32_t cs = VEC && (k & 7u) == 0 ? k + 4u : k;      // chunk stride in shared memory, words
  const uint32_t bw = ((cs == k ? m : nl * cs) + 3u) & ~3u;  // words per sample buffer (keeps the second buffer 16 B aligned)
  const uint32_t kq = (k + 3) / 4;  // blocks of four seasonals per lane
  const uint32_t npair = min(k / 8, (kq - 1) / 2);  // leading pairs of whole blocks that have a block after them

  // ---- tensor memory: the CTA takes all 512 columns; warp w works in lane quarter w % 4 (the only
  //      lanes its tcgen05.ld/st can reach) and column range (w / 4) * 16 * ceil(k / 8)
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&s_tm_base)),
                 "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (threadIdx.x < 20) s_apow[threadIdx.x] = dp->hw_tpow[threadIdx.x];
  if (threadIdx.x < 64) s_w[threadIdx.x] = make_double2(dp->hw_w[2 * threadIdx.x], dp->hw_w[2 * threadIdx.x + 1]);
  const uint32_t bar_s = keep((uint32_t)__cvta_generic_to_shared(&s_bar[warp][0]));  // buffer 1's barrier: + 8
  if (lane == 0) {
    mbar_init(bar_s);
    mbar_init(bar_s + 8u);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const uint32_t tm_base = s_tm_base;
  const uint32_t tS = keep(tm_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * 16u * ((k + 7) / 8));

  const double alpha = dp->alpha, ab = dp->alpha * dp->beta, g1 = dp->gamma * (1.0 - dp->alpha);
  const uint32_t p0 = busy ? lane * k : 0u;  // this lane's chunk inside a season-long super-tile
  const uint32_t buf_bytes = bw * 4u;
  const uint32_t xb_s = keep((uint32_t)__cvta_generic_to_shared(s_x + (size_t)warp * 2 * bw));  // this warp's two buffers
  const uint32_t pw = busy ? lane * cs : 0u;  // ... and where it lies in a sample buffer
  const uint32_t xs0 = keep(xb_s + 4u * pw);  // this lane's chunk in buffer 0
  const uint32_t ws0 = keep((uint32_t)__cvta_generic_to_shared(s_w));
  uint32_t pb = 0;  // bits 0-1: mbarrier phase to wait for next, per sample buffer; bits 2-3: buffer filled by a bulk copy

  const uint64_t stride_units = (uint64_t)gridDim.x * n_warps;
  for (uint64_t u = (uint64_t)blockIdx.x * n_warps + warp; u < b.n_units; u += stride_units) {
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
      const uint32_t dst = xb_s + (buf ? buf_bytes : 0u), bar = bar_s + 8u * buf;
      const uint32_t nv = min(m, n_stream - q0s);
      const bool one_seg = q0s + nv <= ns || q0s >= ns;  // one contiguous source
      const float *const src = unit + (q0s < ns ? a0 + q0s : c0 + (q0s - ns));
      pb &= ~(4u << buf);
      if (cs == k) {  // chunks back to back: the season lies in shared memory as it does in the stream
        if (one_seg && ((reinterpret_cast<uintptr_t>(src) & 15u) | (nv & 3u)) == 0) {
          if (lane == 0) {
            bulk_expect(bar, nv * 4u);
            bulk_copy(dst, src, nv * 4u, bar);
          }
          pb |= 4u << buf;
          return;
        }
        if (one_seg) {
#pragma unroll 3
          for (uint32_t i = lane; i < nv; i += 32) cp_async4_s(dst + 4u * i, src + i);
        } else {
          for (uint32_t i = lane; i < nv; i += 32) {
            const uint32_t q = q0s + i;
            cp_async4_s(dst + 4u * i, unit + (q < ns ? a0 + q : c0 + (q - ns)));
          }
        }
      } else {  // padded chunks: every lane fetches its own (k * 4 bytes is a multiple of 32)
        const uint32_t first = lane * k, dl = dst + 4u * pw;
        const uint32_t cnt = busy && first < nv ? min(k, nv - first) : 0u;  // this lane's samples in the super-tile
        if (one_seg && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
          const uint32_t nfull = nv / k;  // lanes whose chunk is whole: bulk copies; the ragged one after them: 4-byte copies
          if (nfull) {
            if (lane == 0) bulk_expect(bar, nfull * k * 4u);
            __syncwarp();
            if (lane < nfull) bulk_copy(dl, src + first, k * 4u, bar);
            pb |= 4u << buf;
          }
          if (lane == nfull)
            for (uint32_t i = 0; i < cnt; ++i) cp_async4_s(dl + 4u * i, src + first + i);
        } else {
          for (uint32_t i = 0; i < cnt; ++i) {
            const uint32_t q = q0s + first + i;
            cp_async4_s(dl + 4u * i, unit + (q < ns ? a0 + q : c0 + (q - ns)));
          }
        }
      }
      cp_async_commit();
    };
    auto wait_buf = [&](uint32_t buf) {
      if (pb & (4u << buf)) {
        mbar_wait(bar_s + 8u * buf, (pb >> buf) & 1u);
        pb ^= 1u << buf;
      }
      cp_async_wait<0>();
      __syncwarp();
    };
    // sum over the season of the lane's k samples at shared address xa (two accumulators per lane)
    auto season_sum = [&](uint32_t xa) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll 1
      for (uint32_t blk = 0; blk < kq; ++blk, xa += 16u) {
        float x[4];
        lds4<VEC>(xa, x);
        const uint32_t rem = k - 4 * blk;  // warp-uniform
        if (rem >= 4) {
          s0 += (double)x[0]; s1 += (double)x[1]; s0 += (double)x[2]; s1 += (double)x[3];
        } else {
#pragma unroll
          for (int j = 0; j < 3; ++j)
            if ((uint32_t)j < rem) s0 += (double)x[j];
        }
      }
      s0 = busy ? s0 + s1 : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      return s0;
    };

    V2 carry;
    double sse = 0.0, sigma = 0.0;
    bool sigma_ready = false;
    uint32_t phase = 0;  // seasonal slot of stream position 0
    uint32_t cur = 0;    // buffer that holds (or is receiving) the super-tile at q0
    bool landed = false;
    __syncwarp();        // the previous unit is done with both buffers
    if (RESUME) {
      // The cached seasonals (slot order, 8 * period bytes -- as much as the two sample buffers hold) come in as one
      // bulk copy (coalesced 8-byte loads when the record is not 16 B aligned); each lane then picks the slots
      // of its chunk.  Read straight from global, a lane's 8-byte loads are a period / 32 apart from its
      // neighbours': four times the traffic, and slower than scoring the whole history again.
      if ((reinterpret_cast<uintptr_t>(ust) & 15u) == 0 && (m & 1u) == 0) {
        if (lane == 0) {
          bulk_expect(bar_s, m * 8u);
          bulk_copy(xb_s, reinterpret_cast<const float *>(ust + 4), m * 8u, bar_s);
        }
        pb |= 4u;
      } else {
        for (uint32_t i = lane; i < m; i += 32) sts_f64(xb_s + 8u * i, ust[4 + i]);
        pb &= ~4u;
      }
      carry.l = ust[0];
      carry.b = ust[1];
      sigma = ust[2];
      sigma_ready = true;
      phase = (uint32_t)ust[3];
      wait_buf(0);
      uint32_t slot = phase + p0;  // < 2m: phase < m and p0 + i < m
#pragma unroll 1
      for (uint32_t blk = 0; blk < kq; ++blk) {
        double s[4];
#pragma unroll
        for (int j = 0; j < 4; ++j, ++slot) s[j] = 4 * blk + j < k ? lds_f64(xb_s + 8u * (slot >= m ? slot - m : slot)) : 0.0;
        tm_st4(tS + 8 * blk, s);
      }
      __syncwarp();  // every lane has its seasonals: the buffers can take samples
      stage(0, 0);
    } else {
      // ---- initialisation from the first two seasons (D5); both lie inside the model window.  The
      //      second one is also the first super-tile scored below, so it stays where it lands.
      stage(0, 0);
      stage(m, 1);
      wait_buf(0);
      const double L0 = season_sum(xs0) / (double)m;
      {
        uint32_t xa = xs0;
#pragma unroll 1
        for (uint32_t blk = 0; blk < kq; ++blk, xa += 16u) {
          float x[4];
          lds4<VEC>(xa, x);
          double s[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) s[j] = 4 * blk + j < k ? (double)x[j] - L0 : 0.0;
          tm_st4(tS + 8 * blk, s);
        }
      }
      wait_buf(1);
      const double M2 = season_sum(xs0 + buf_bytes) / (double)m;
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
      if (q0 + m >= n_stream) {  // last super-tile: pull the head of this warp's next unit (its first two seasons) towards L2
        const uint64_t un = u + stride_units;
        if (un < b.n_units) {
          const int64_t on = b.offsets[un], len = b.offsets[un + 1] - on;
          const uint32_t bytes = (uint32_t)min((int64_t)(2u * m), len) * 4u;
          const char *p = reinterpret_cast<const char *>(b.values + on);
          for (uint32_t byte = lane * 128u; byte < bytes; byte += 32u * 128u) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + byte));
        }
      }
      const uint32_t qa = q0 + p0;                                      // stream position of the chunk
      const uint32_t kv = busy && qa < n_stream ? min(k, n_stream - qa) : 0u;  // real samples in this lane's chunk
      const uint32_t xs = keep(xs0 + (cur ? buf_bytes : 0u));          // the chunk's samples (shared-window address)
      const bool whole = !RESUME && q0 + m <= ns;  // a whole season of the model window (warp-uniform)
      uint32_t ra[8], rb[8];

      // ---- pass 1: chunk offset from a zero state, as a dot product with the chunk weights (two
      //      accumulators per component, no dependent chain).  The weights are those of a FULL chunk, and that
      //      is enough: an offset is only ever used by the lanes after it, and the one chunk of a super-tile
      //      that can be short is the last one with samples (whatever this computes there, and in the empty
      //      chunks behind it, only reaches lanes that have no samples).  Skipped when one lane holds it all.
      //      The offsets are those of a DEVIATION: with (r_l, r_b) the carried state, the state is written as
      //      (r_l, 0) + r_b A^i e_2 + d_i -- the first term is a fixed point of the recurrence under the constant
      //      input r_l, the second the free motion of the carried trend, and d obeys the same recurrence with
      //      inputs y_i - r_l from d_0 = 0.  So the scan needs no carry term, what it adds up is small next to the
      //      level, and a constant series (deviation exactly zero) comes out exactly as in the sequential form
      //      (with the carry pushed through powers of A it came out with 1e-16 of noise per super-tile, which a
      //      marginally stable parameter set amplified past the 1e-9 guard of the flags).
      V2 uo;
      uo.l = 0.0;
      uo.b = 0.0;
      const double r_l = carry.l, r_b = carry.b;
      if (min(m, n_stream - q0) > k) {
        double ul0 = 0.0, ub0 = 0.0, ul1 = 0.0, ub1 = 0.0;
        auto p1 = [&](uint32_t blk, uint32_t (&c)[8], uint32_t (&n)[8]) {
          float x[4];
          double2 wv[4];
          lds4<VEC>(xs + 16u * blk, x);
          ldw4(ws0 + 64u * blk, wv);
          tm_ld_arrive(c);
          if (blk + 1 < kq) tm_ld_issue(tS + 8 * (blk + 1), n);
          const uint32_t rem = k - 4u * blk;  // warp-uniform
          if (rem >= 4) {
            const double y0 = ((double)x[0] - tm_get(c, 0)) - r_l, y1 = ((double)x[1] - tm_get(c, 1)) - r_l;
            const double y2 = ((double)x[2] - tm_get(c, 2)) - r_l, y3 = ((double)x[3] - tm_get(c, 3)) - r_l;
            ul0 = fma(wv[0].x, y0, ul0); ub0 = fma(wv[0].y, y0, ub0);
            ul1 = fma(wv[1].x, y1, ul1); ub1 = fma(wv[1].y, y1, ub1);
            ul0 = fma(wv[2].x, y2, ul0); ub0 = fma(wv[2].y, y2, ub0);
            ul1 = fma(wv[3].x, y3, ul1); ub1 = fma(wv[3].y, y3, ub1);
          } else {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
              if ((uint32_t)j < rem) {
                const double y = ((double)x[j] - tm_get(c, j)) - r_l;
                ul0 = fma(wv[j].x, y, ul0);
                ub0 = fma(wv[j].y, y, ub0);
              }
            }
          }
        };
        // the leading pairs of blocks are whole and have a block after them: no guards at all
        auto p1w = [&](uint32_t blk, uint32_t (&c)[8], uint32_t (&n)[8]) {
          float x[4];
          double2 wv[4];
          lds4<VEC>(xs + 16u * blk, x);
          ldw4(ws0 + 64u * blk, wv);
          tm_ld_arrive(c);
          tm_ld_issue(tS + 8 * (blk + 1), n);
          const double y0 = ((double)x[0] - tm_get(c, 0)) - r_l, y1 = ((double)x[1] - tm_get(c, 1)) - r_l;
          const double y2 = ((double)x[2] - tm_get(c, 2)) - r_l, y3 = ((double)x[3] - tm_get(c, 3)) - r_l;
          ul0 = fma(wv[0].x, y0, ul0); ub0 = fma(wv[0].y, y0, ub0);
          ul1 = fma(wv[1].x, y1, ul1); ub1 = fma(wv[1].y, y1, ub1);
          ul0 = fma(wv[2].x, y2, ul0); ub0 = fma(wv[2].y, y2, ub0);
          ul1 = fma(wv[3].x, y3, ul1); ub1 = fma(wv[3].y, y3, ub1);
        };
        tm_ld_issue(tS, ra);
#pragma unroll 1
        for (uint32_t blk = 0; blk < 2 * npair; blk += 2) {
          p1w(blk, ra, rb);
          p1w(blk + 1, rb, ra);
        }
#pragma unroll 1
        for (uint32_t blk = 2 * npair; blk < kq; blk += 2) {
          p1(blk, ra, rb);
          if (blk + 1 < kq) p1(blk + 1, rb, ra);
        }
        uo.l = ul0 + ul1;
        uo.b = ub0 + ub1;
      }
      // ---- the state entering this lane's chunk: warp scan of the chunk maps v -> A^k v + u
      const V2 vin = fm_hw::scan_chunks<false>(uo, carry, s_apow, nullptr, lane);

      // ---- pass 2: the real recurrence over the chunk, from (r_l, 0) + r_b A^p0 e_2 + the deviation
      double l = r_l + fma(r_b, dp->hw_tlane[4 * lane + 1], vin.l), bb = fma(r_b, dp->hw_tlane[4 * lane + 3], vin.b);
      unsigned long long fmask = 0ull;  // flags of the chunk's current samples (k <= 64)
      const bool tile_has_cur = q0 + m > ns;
      if (!sigma_ready && q0 >= ns) {  // the model window ended with the previous super-tile: close the residual sum
        double t = busy ? sse : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        sigma = sqrt(t / (double)(ns - m));
        sigma_ready = true;
      }
      if (whole) {
        auto one = [&](float xf, uint32_t (&c)[8], int j) {
          const double sv = tm_get(c, j);
          const double y = (double)xf - sv;
          const double lb = l + bb, e = y - lb;
          l = fma(alpha, e, lb);
          bb = fma(ab, e, bb);
          sse = fma(e, e, sse);
          tm_put(c, j, fma(g1, e, sv));
        };
        auto p2 = [&](uint32_t blk, uint32_t (&c)[8], uint32_t (&n)[8]) {
          float x[4];
          lds4<VEC>(xs + 16u * blk, x);
          tm_ld_arrive(c);
          if (blk + 1 < kq) tm_ld_issue(tS + 8 * (blk + 1), n);
          const uint32_t rem = k - 4u * blk;
          if (rem >= 4) {
#pragma unroll
            for (int j = 0; j < 4; ++j) one(x[j], c, j);
          } else {
#pragma unroll
            for (int j = 0; j < 3; ++j)
              if ((uint32_t)j < rem) one(x[j], c, j);
          }
          tm_st(tS + 8 * blk, c);
        };
        auto p2w = [&](uint32_t blk, uint32_t (&c)[8], uint32_t (&n)[8]) {
          float x[4];
          lds4<VEC>(xs + 16u * blk, x);
          tm_ld_arrive(c);
          tm_ld_issue(tS + 8 * (blk + 1), n);
#pragma unroll
          for (int j = 0; j < 4; ++j) one(x[j], c, j);
          tm_st(tS + 8 * blk, c);
        };
        tm_ld_issue(tS, ra);
#pragma unroll 1
        for (uint32_t blk = 0; blk < 2 * npair; blk += 2) {
          p2w(blk, ra, rb);
          p2w(blk + 1, rb, ra);
        }
#pragma unroll 1
        for (uint32_t blk = 2 * npair; blk < kq; blk += 2) {
          p2(blk, ra, rb);
          if (blk + 1 < kq) p2(blk + 1, rb, ra);
        }
        tm_wait_st();
      } else if (!FIT && !RESUME && sigma_ready && q0 >= ns && q0 + m >= n_stream) {
        // The unit's last super-tile, current samples only, and no state is kept afterwards: the seasonals
        // are dead after this read, and so are level and trend after a lane's last real sample.  So nothing
        // is guarded per lane (what a lane computes past its samples is masked out of the flags) and nothing
        // is written back.
        const double clear = kk * sigma * 0.999999;  // see the guarded form below
        const uint32_t kmax = min(k, n_stream - q0);  // the longest chunk of the super-tile (lane 0's)
#pragma unroll 1
        for (uint32_t blk = 0; 4 * blk < kmax; ++blk) {
          float xf[4];
          double s[4];
          lds4<VEC>(xs + 16u * blk, xf);
          tm_ld4(tS + 8 * blk, s);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t i = 4 * blk + j;
            const double x = (double)xf[j];
            const double lb = l + bb, f = lb + s[j], e = x - f;
            l = fma(alpha, e, lb);
            bb = fma(ab, e, bb);
            if (BANDS || !(fabs(e) <= clear && x >= mlb)) {
              const double up = fma(kk, sigma, f);
              double lw = fma(-kk, sigma, f);
              lw = lw < mlb ? mlb : lw;
              const double gx = 1e-9 * (fabs(f) + fabs(x));
              if (((bound & 1u) && x - up > gx) || ((bound & 2u) && lw - x > gx)) fmask |= 1ull << i;
              if (BANDS && i < kv) {
                band_up[off + c0 + (qa + i - ns)] = (float)up;
                band_lo[off + c0 + (qa + i - ns)] = (float)lw;
              }
            }
          }
        }
        fmask &= kv >= 64 ? ~0ull : (1ull << kv) - 1ull;
      } else {
        // Sweep 0 takes the samples of the model window, sweep 1 the current ones (sigma must be final
        // first); each sweep walks the seasonals block by block through registers.
        const uint32_t i_m = qa < ns ? min(kv, ns - qa) : 0u;  // model samples at the head of the chunk
        double clear = 0.0;  // a current sample with |x - forecast| up to this is inside the band (see below)
#pragma unroll 1
        for (int sweep = 0; sweep < 2; ++sweep) {
          if (sweep == 0 && q0 >= ns) continue;  // nothing of the model window in this super-tile
          if (sweep == 1) {
            if (!tile_has_cur) break;
            if (!sigma_ready) {  // the model window ends in this super-tile: close the residual sum first
              double t = busy ? sse : 0.0;
#pragma unroll
              for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
              sigma = sqrt(t / (double)(ns - m));
              sigma_ready = true;
            }
            clear = kk * sigma * 0.999999;
          }
          const uint32_t lo = sweep ? i_m : 0u, hi = sweep ? kv : i_m;
#pragma unroll 1
          for (uint32_t blk = 0; blk < kq; ++blk) {
            if (__all_sync(0xffffffffu, 4 * blk >= hi || 4 * blk + 4 <= lo)) continue;  // no lane has work in this block
            float xf[4];
            double s[4];
            lds4<VEC>(xs + 16u * blk, xf);
            tm_ld4(tS + 8 * blk, s);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t i = 4 * blk + j;
              if (i >= lo && i < hi) {  // one sample: one-step error, state and season update, residual sum or flag
                const double x = (double)xf[j];
                const double lb = l + bb, f = lb + s[j], e = x - f;
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
                s[j] = fma(g1, e, s[j]);
              }
            }
            tm_st4(tS + 8 * blk, s);
          }
          tm_wait_st();
        }
      }
      // next super-tile's carry (and the state the model cache keeps): the state after the last real
      // sample of this super-tile, which lives in lane (nv-1)/k
      {
        const uint32_t nv = min(m, n_stream - q0);
        const int src = whole ? (int)nl - 1 : (int)((nv - 1) / k);
        carry.l = __shfl_sync(0xffffffffu, l, src);
        carry.b = __shfl_sync(0xffffffffu, bb, src);
      }
      if (tile_has_cur) {
        const float *const xp = s_x + (size_t)warp * 2 * bw + cur * bw + pw;
        fm_hw::tile_flags<EMIT>(fmask, lane, qa, ns, w.emit_base, emit_idx, emit_val, xp, n_anom, first);
      }
    }
    if (FIT || RESUME) {  // save the model: state after the last sample, residual sigma, phase, seasonals
      if (FIT && !sigma_ready) {
        double t = busy ? sse : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        sigma = sqrt(t / (double)(ns - m));
      }
      // seasonals: into the (now idle) sample buffers in slot order, then out as one bulk store
      __syncwarp();
      uint32_t slot = phase + p0;
#pragma unroll 1
      for (uint32_t blk = 0; blk < kq; ++blk) {
        double s[4];
        tm_ld4(tS + 8 * blk, s);
#pragma unroll
        for (int j = 0; j < 4; ++j, ++slot)
          if (4 * blk + j < k && busy) sts_f64(xb_s + 8u * (slot >= m ? slot - m : slot), s[j]);
      }
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // every lane's writes before the bulk store reads them
      __syncwarp();
      if ((reinterpret_cast<uintptr_t>(ust) & 15u) == 0 && (m & 1u) == 0) {
        if (lane == 0) bulk_store(ust + 4, xb_s, m * 8u);
      } else {
        for (uint32_t i = lane; i < m; i += 32) ust[4 + i] = lds_f64(xb_s + 8u * i);
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

template <int MODE, bool VEC>
cudaError_t launch_tm2(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                       uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                       const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, double *state, int sm_count,
                       cudaStream_t st) {
  const uint32_t nl = fm_hw_tm_lanes(period);
  if (!nl) return cudaErrorInvalidValue;
  const uint32_t k = period / nl, kb = (k + 7) / 8;
  const uint32_t cs = VEC && (k & 7u) == 0 ? k + 4u : k, bw = ((cs == k ? period : nl * cs) + 3u) & ~3u;  // as in the kernel
  // warps per CTA: four lane quarters x as many 16*kb-column ranges as fit in 512 columns, capped by
  // the launch bound and by the two fp32 seasons each warp keeps in shared memory
  int warps = (int)std::min<uint32_t>(4 * (512 / (16 * kb)), kMaxWarps);
  warps = std::min(warps, (227 * 1024 - 1536 - 32) / ((int)bw * 8));  // 1.5 KB: the kernel's static shared memory
  if (warps < 1) return cudaErrorInvalidValue;
  uint64_t blocks = (b.n_units + warps - 1) / warps;
  const uint64_t cap = (uint64_t)sm_count;  // persistent: one CTA is resident per SM (it owns the tensor memory)
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  const int smem = warps * 2 * (int)bw * (int)sizeof(float) + 32;  // scalar lds4 may read up to 3 samples past a chunk
  cudaError_t e = cudaFuncSetAttribute(k_hw_tm<MODE, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  k_hw_tm<MODE, VEC><<<(unsigned)blocks, warps * 32, smem, st>>>(b, dp, scores, unit_status, unit_points, band_up, band_lo,
                                                                emit_off, emit_idx, emit_val, state);
  return cudaGetLastError();
}

template <int MODE>
cudaError_t launch_tm(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                      uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                      const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, double *state, int sm_count,
                      cudaStream_t st) {
  const uint32_t nl = fm_hw_tm_lanes(period);
  if (nl && (period / nl) % 4 == 0)
    return launch_tm2<MODE, true>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx,
                                  emit_val, state, sm_count, st);
  return launch_tm2<MODE, false>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, emit_off, emit_idx,
                                 emit_val, state, sm_count, st);
}

}  // namespace

bool fm_hw_tm_supported(uint32_t period) { return fm_hw_tm_lanes(period) != 0; }

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
