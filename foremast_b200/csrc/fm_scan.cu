// fm_scan.cu -- exclusive prefix sum of per-unit anomaly counts: turns
// "unit u has c[u] anomalous current points" into offsets of a compact (index, value) list, the
// device-side form of the anomaly pairs barrelman decodes as [t0, v0, t1, v1, ...]
// (foremast-barrelman/pkg/controller/DeploymentController.go:443-452).
//
// Single pass, decoupled look-back: tiles of 2048 counts take a ticket from an atomic counter,
// publish their aggregate, and resolve their exclusive prefix by looking back over the published
// tile states (status in the top 2 bits of a 64-bit word, value below).  The tile states and the
// ticket counter are SCRATCH: global memory that is neither input nor output and must start at
// zero.  Algorithmic bytes: 4 in + 8 out per unit.
#include <mutex>

#include "fm_common.cuh"

namespace {

constexpr int kThreads = 256, kItems = 8, kTile = kThreads * kItems;
constexpr unsigned long long kAgg = 1ull << 62, kPrefix = 2ull << 62, kValMask = (1ull << 62) - 1;

// in[i] is read at byte offset i*stride (so the counts can be gathered straight out of the
// fm_series_score array); out has n+1 entries, out[n] = total.
__global__ void __launch_bounds__(kThreads) k_scan_counts(const unsigned char *__restrict__ in, size_t stride, uint64_t n,
                                                          unsigned long long *__restrict__ out,
                                                          unsigned long long *__restrict__ scratch) {
  __shared__ unsigned long long s_warp[kThreads / 32];
  __shared__ unsigned long long s_base;
  __shared__ unsigned int s_tile;
  unsigned long long *ticket = scratch;     // scratch[0]: next tile ticket
  unsigned long long *state = scratch + 1;  // scratch[1 + tile]
  if (threadIdx.x == 0) s_tile = (unsigned int)atomicAdd(ticket, 1ull);
  __syncthreads();
  const uint64_t tile = s_tile;
  const uint64_t first = tile * kTile + (uint64_t)threadIdx.x * kItems;
  uint32_t v[kItems];
  unsigned long long local = 0;
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const uint64_t e = first + i;
    v[i] = e < n ? *reinterpret_cast<const uint32_t *>(in + e * stride) : 0u;
    local += v[i];
  }
  // block-wide exclusive scan of the per-thread sums
  unsigned long long incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((threadIdx.x & 31) >= o) incl += t;
  }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();
  unsigned long long warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < kThreads / 32; ++w) {
    if (w < (int)(threadIdx.x >> 5)) warp_off += s_warp[w];
    block_total += s_warp[w];
  }
  // publish, then look back (one warp; each lane inspects one predecessor per round)
  if (threadIdx.x < 32) {
    if (threadIdx.x == 0) {
      const unsigned long long word = (tile == 0 ? kPrefix : kAgg) | block_total;
      atomicExch(state + tile, word);
    }
    unsigned long long base = 0;
    if (tile > 0) {
      int64_t look = (int64_t)tile - 1 - (int64_t)threadIdx.x;
      while (true) {
        unsigned long long w = 0;
        if (look >= 0) {
          do { w = *reinterpret_cast<volatile unsigned long long *>(state + look); } while ((w >> 62) == 0);
        } else {
          w = kPrefix;  // before tile 0: an inclusive prefix of zero
        }
        const unsigned has_prefix = __ballot_sync(0xffffffffu, (w >> 62) == 2);
        // add aggregates of the lanes up to and including the nearest lane that holds a full prefix
        const int stop = has_prefix ? __ffs(has_prefix) - 1 : 32;
        unsigned long long contrib = ((int)threadIdx.x <= stop) ? (w & kValMask) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
        base += contrib;
        if (has_prefix) break;
        look -= 32;
      }
      if (threadIdx.x == 0) atomicExch(state + tile, kPrefix | (base + block_total));
    }
    if (threadIdx.x == 0) s_base = base;
  }
  __syncthreads();
  unsigned long long run = s_base + warp_off + (incl - local);
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const uint64_t e = first + i;
    if (e < n) out[e] = run;
    run += v[i];
    if (e + 1 == n) out[n] = run;
  }
  if (n == 0 && tile == 0 && threadIdx.x == 0) out[0] = 0;
}

}  // namespace

uint64_t fm_scan_tiles(uint64_t n) { return n ? (n + kTile - 1) / kTile : 1; }
uint64_t fm_scan_scratch_bytes(uint64_t n) { return (fm_scan_tiles(n) + 1) * sizeof(unsigned long long); }

// scratch must hold fm_scan_scratch_bytes(n) bytes; it is cleared here (stream-ordered).
cudaError_t fm_launch_scan_counts(const void *in, size_t stride, uint64_t n, unsigned long long *out,
                                  unsigned long long *scratch, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(scratch, 0, fm_scan_scratch_bytes(n), st);
  if (e != cudaSuccess) return e;
  k_scan_counts<<<(unsigned)fm_scan_tiles(n), kThreads, 0, st>>>(static_cast<const unsigned char *>(in), stride, n, out, scratch);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ allocation-free variant
//
// fm_scan_counts_static: the same exclusive scan for callers that are not allowed to allocate
// device memory (see include/foremast_gpu.h).  Its tile states and ticket live in a fixed array
// inside this module instead of a cudaMallocAsync'ed buffer, so the call is a memset plus a kernel
// launch and nothing else.
//
// g_static_scratch is a __device__ symbol.  In host code that name is only a stand-in for the
// symbol, NOT a device address -- handing it to a kernel would fault (or, where the GPU can reach
// host memory, silently scribble on the wrong memory).  The address the kernel needs comes from
// cudaGetSymbolAddress, asked once per device with that device current, and is kept in
// g_static_ptr[device].
namespace {

constexpr uint64_t kStaticMaxCounts = FM_SCAN_STATIC_MAX_COUNTS;
constexpr uint64_t kStaticMaxTiles = (kStaticMaxCounts + kTile - 1) / kTile;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned long long g_static_scratch[kStaticMaxTiles + 1];  // 64 KiB + 8 B per GPU

std::mutex g_static_mu;
unsigned long long *g_static_ptr[kMaxDevices] = {nullptr};

}  // namespace

extern "C" int fm_scan_counts_static(const uint32_t *d_counts, uint64_t n, uint64_t *d_offsets, void *cuda_stream) {
  if (!d_offsets || (n && !d_counts)) return FM_E_INVALID;
  if (n > kStaticMaxCounts) return FM_E_TOO_LARGE;  // the fixed scratch cannot hold more tile states
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    return FM_E_NO_DEVICE;
  }
  unsigned long long *scratch;
  {
    std::lock_guard<std::mutex> g(g_static_mu);
    if (!g_static_ptr[dev]) {  // first call on this device: resolve the symbol to a device address
      void *p = nullptr;
      if (cudaGetSymbolAddress(&p, g_static_scratch) != cudaSuccess || !p) {
        cudaGetLastError();
        return FM_E_CUDA;
      }
      g_static_ptr[dev] = static_cast<unsigned long long *>(p);
    }
    scratch = g_static_ptr[dev];
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  // the array keeps whatever the previous call left in it: ticket and tile states must start at zero
  if (cudaMemsetAsync(scratch, 0, fm_scan_scratch_bytes(n), st) != cudaSuccess) {
    cudaGetLastError();
    return FM_E_CUDA;
  }
  k_scan_counts<<<(unsigned)fm_scan_tiles(n), kThreads, 0, st>>>(reinterpret_cast<const unsigned char *>(d_counts),
                                                               sizeof(uint32_t), n,
                                                               reinterpret_cast<unsigned long long *>(d_offsets), scratch);
  if (cudaGetLastError() != cudaSuccess) return FM_E_CUDA;
  return FM_OK;
}
