// CSR construction: stable counting sort of edge ids by an int32 key.
//
// Replaces the unsorted `index` argument of the reference's scatter_add_ /
// index_add_ call sites (EGCLStack.py:292-298, PAINNStack.py:254-257,
// utils/model/model.py:279-286, SCFStack.py:285) by an explicit, reusable CSR.
//
// Pipeline (all on `stream`, integer atomics only -- the result is unique):
//   1. histogram   cnt[key[e]]++ , remembering the (arbitrary) arrival slot
//   2. scan        rowptr = exclusive prefix sum of cnt
//   3. scatter     tmp[rowptr[key[e]] + slot[e]] = e
//   4. segsort     perm[seg] = ascending sort of tmp[seg]   (rank by counting)
// Step 4 removes the arrival-order nondeterminism of step 1, so perm is the
// stable order a sequential CPU loop over e = 0..E-1 would produce.
#include "hgk_common.cuh"

#include <mutex>

namespace hgk {

thread_local char g_err[512] = {0};
char* error_buffer() { return g_err; }
int fail(int status, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return status;
}

__global__ void __launch_bounds__(256)
k_csr_hist(const int32_t* __restrict__ key, int64_t E, int32_t N,
           int32_t* __restrict__ cnt, int32_t* __restrict__ slot) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; e < E; e += stride) {
    int32_t k = key[e];
    // out-of-range keys are clamped so a corrupt index cannot write out of bounds
    k = k < 0 ? 0 : (k >= N ? N - 1 : k);
    slot[e] = atomicAdd(&cnt[k], 1);
  }
}

// Exclusive scan of cnt[0..N) into rowptr[0..N], single CTA, 1024 threads x 8 items.
constexpr int kScanThreads = 1024;
constexpr int kScanItems = 8;
__global__ void __launch_bounds__(kScanThreads)
k_exclusive_scan(const int32_t* __restrict__ cnt, int64_t N,
                 int32_t* __restrict__ rowptr) {
  __shared__ int32_t warp_sums[kScanThreads / 32];
  __shared__ int32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < N; base += (int64_t)kScanThreads * kScanItems) {
    int32_t v[kScanItems];
    int32_t local = 0;
    const int64_t first = base + (int64_t)tid * kScanItems;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      v[i] = (first + i < N) ? cnt[first + i] : 0;
      local += v[i];
    }
    // inclusive scan of `local` across the CTA
    int32_t incl = local;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      int32_t t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      int32_t ws = warp_sums[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, ws, off);
        if (lane >= off) ws += t;
      }
      warp_sums[lane] = ws;  // inclusive over warps
    }
    __syncthreads();
    const int32_t carry = carry_s;
    int32_t excl = carry + incl - local + (wid > 0 ? warp_sums[wid - 1] : 0);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      if (first + i < N) rowptr[first + i] = excl;
      excl += v[i];
    }
    __syncthreads();
    if (tid == kScanThreads - 1) carry_s = carry + warp_sums[kScanThreads / 32 - 1];
    __syncthreads();
  }
  if (tid == 0) rowptr[N] = carry_s;
}

__global__ void __launch_bounds__(256)
k_csr_scatter(const int32_t* __restrict__ key, int64_t E, int32_t N,
              const int32_t* __restrict__ rowptr, const int32_t* __restrict__ slot,
              int32_t* __restrict__ tmp) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; e < E; e += stride) {
    int32_t k = key[e];
    k = k < 0 ? 0 : (k >= N ? N - 1 : k);
    tmp[rowptr[k] + slot[e]] = (int32_t)e;
  }
}

// Segments up to this length are rank-sorted by one warp (O(len^2 / 32): 58 edges per row at config 5);
// longer ones (a hub row, a per-graph key) go to k_csr_segsort_long.
constexpr int32_t kSegsortWarpMax = 128;

// One warp per segment: rank-by-counting sort of distinct int32 values.
__global__ void __launch_bounds__(256)
k_csr_segsort(const int32_t* __restrict__ rowptr, int64_t N,
              const int32_t* __restrict__ tmp, int32_t* __restrict__ perm) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t n = warp; n < N; n += nwarps) {
    const int32_t beg = rowptr[n], end = rowptr[n + 1];
    const int32_t len = end - beg;
    if (len > kSegsortWarpMax) continue;               // k_csr_segsort_long
    for (int32_t t = lane; t < len; t += 32) {
      const int32_t v = tmp[beg + t];
      int32_t rank = 0;
      for (int32_t s = 0; s < len; ++s) rank += (tmp[beg + s] < v) ? 1 : 0;
      perm[beg + rank] = v;
    }
  }
}

// Long segments, one CTA each: the segment is cut into chunks of kSegChunk values, every chunk is sorted in shared
// memory (bitonic network, padded with INT32_MAX) and written back in place; the rank of a value is then the sum over
// the chunks of the number of smaller values (binary search: the values are distinct), and the value goes to its
// place.  O(len log^2 chunk + len (len / chunk) log chunk) instead of O(len^2).
constexpr int kSegChunk = 4096;

__global__ void __launch_bounds__(256)
k_csr_segsort_long(const int32_t* __restrict__ rowptr, int64_t N, int32_t* __restrict__ tmp, int32_t* __restrict__ perm) {
  __shared__ int32_t buf[kSegChunk];
  for (int64_t n = blockIdx.x; n < N; n += gridDim.x) {
    const int32_t beg = rowptr[n], len = rowptr[n + 1] - beg;
    if (len <= kSegsortWarpMax) continue;              // CTA-uniform
    int32_t* seg = tmp + beg;
    const int32_t nchunks = (len + kSegChunk - 1) / kSegChunk;
    for (int32_t c = 0; c < nchunks; ++c) {
      const int32_t c0 = c * kSegChunk;
      const int32_t clen = min(kSegChunk, len - c0);
      int32_t pot = 1;                                 // power of two >= clen
      while (pot < clen) pot <<= 1;
      for (int32_t i = threadIdx.x; i < pot; i += blockDim.x) buf[i] = i < clen ? seg[c0 + i] : INT32_MAX;
      __syncthreads();
      for (int32_t k = 2; k <= pot; k <<= 1) {
        for (int32_t j = k >> 1; j > 0; j >>= 1) {
          for (int32_t i = threadIdx.x; i < pot; i += blockDim.x) {
            const int32_t partner = i ^ j;
            if (partner > i) {
              const int32_t a = buf[i], b2 = buf[partner];
              const bool up = (i & k) == 0;
              if ((a > b2) == up) { buf[i] = b2; buf[partner] = a; }
            }
          }
          __syncthreads();
        }
      }
      for (int32_t i = threadIdx.x; i < clen; i += blockDim.x) seg[c0 + i] = buf[i];
      __syncthreads();
    }
    if (nchunks == 1) {
      for (int32_t i = threadIdx.x; i < len; i += blockDim.x) perm[beg + i] = seg[i];
    } else {
      __threadfence_block();
      for (int32_t i = threadIdx.x; i < len; i += blockDim.x) {
        const int32_t v = seg[i];
        int32_t rank = 0;
        for (int32_t c = 0; c < nchunks; ++c) {
          const int32_t* ch = seg + c * kSegChunk;
          int32_t lo = 0, hi = min(kSegChunk, len - c * kSegChunk);
          while (lo < hi) {                            // first position whose value is >= v
            const int32_t mid = (lo + hi) >> 1;
            if (ch[mid] < v) lo = mid + 1; else hi = mid;
          }
          rank += lo;
        }
        perm[beg + rank] = v;
      }
    }
    __syncthreads();
  }
}

}  // namespace hgk

using namespace hgk;

extern "C" int hgk_version(void) { return 100; }
extern "C" const char* hgk_last_error(void) { return hgk::error_buffer(); }

extern "C" int64_t hgk_csr_build_workspace_bytes(int64_t E, int64_t N) {
  if (E < 0 || N < 0) return 0;
  return align_up(4 * N, 256) + 2 * align_up(4 * E, 256) + 256;
}

namespace hgk {
int csr_build_impl(const int32_t* key, int64_t E, int64_t N, int32_t* rowptr,
                   int32_t* perm, void* workspace, cudaStream_t st) {
  char* ws = (char*)workspace;
  int32_t* cnt = (int32_t*)ws;
  int32_t* slot = (int32_t*)(ws + align_up(4 * N, 256));
  int32_t* tmp = (int32_t*)(ws + align_up(4 * N, 256) + align_up(4 * E, 256));
  if (N == 0) {
    cudaMemsetAsync(rowptr, 0, 4, st);
    return check_launch("csr_build(memset)");
  }
  cudaMemsetAsync(cnt, 0, 4 * N, st);
  if (E > 0) {
    const int blocks = (int)imin64(ceil_div(E, 256), (int64_t)kNumSM * 16);
    k_csr_hist<<<blocks, 256, 0, st>>>(key, E, (int32_t)N, cnt, slot);
  }
  k_exclusive_scan<<<1, kScanThreads, 0, st>>>(cnt, N, rowptr);
  if (E > 0) {
    const int blocks = (int)imin64(ceil_div(E, 256), (int64_t)kNumSM * 16);
    k_csr_scatter<<<blocks, 256, 0, st>>>(key, E, (int32_t)N, rowptr, slot, tmp);
    const int sblocks = (int)imin64(ceil_div(N, 8), (int64_t)kNumSM * 32);
    k_csr_segsort<<<sblocks, 256, 0, st>>>(rowptr, N, tmp, perm);
    // rows longer than kSegsortWarpMax exist only if E / N allows one: skip the launch for the common case
    if (E > kSegsortWarpMax) {
      const int lblocks = (int)imin64(N, (int64_t)kNumSM * 4);
      k_csr_segsort_long<<<lblocks, 256, 0, st>>>(rowptr, N, tmp, perm);
    }
  }
  return check_launch("csr_build");
}
}  // namespace hgk

extern "C" int hgk_csr_build(const int32_t* key, int64_t E, int64_t N,
                             int32_t* rowptr, int32_t* perm, void* workspace,
                             int64_t workspace_bytes, void* stream) {
  HGK_REQUIRE(E >= 0 && N >= 0, "csr_build: negative size (E=%lld N=%lld)",
              (long long)E, (long long)N);
  HGK_REQUIRE(E < (1ll << 31) && N < (1ll << 31), "csr_build: sizes must fit int32");
  HGK_REQUIRE(rowptr != nullptr, "csr_build: rowptr is NULL");
  HGK_REQUIRE(E == 0 || (key && perm), "csr_build: key/perm is NULL with E>0");
  HGK_REQUIRE(E == 0 || N > 0, "csr_build: E>0 needs N>0");
  if (workspace_bytes < hgk_csr_build_workspace_bytes(E, N) || !workspace)
    return fail(HGK_ERR_WORKSPACE, "csr_build: workspace %lld < %lld bytes",
                (long long)workspace_bytes,
                (long long)hgk_csr_build_workspace_bytes(E, N));
  return csr_build_impl(key, E, N, rowptr, perm, workspace, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// hgk_csr_build_static: the same grouping for callers that may not allocate device memory.
//
// Who that is: a caller that has handed the whole GPU to another allocator (a serving process that
// pre-reserves all device memory for its own pool, so a plain cudaMalloc would fail), or one that
// runs in a latency-critical enqueue callback where cudaMalloc / cudaFree are off limits because
// they synchronise the device.  Such a caller cannot provide the workspace hgk_csr_build asks for,
// so the workspace lives here, in a fixed array that is part of the library image: every GPU that
// loads libhgk.so carries these bytes until the process exits.  The price is a size limit
// (HGK_CSR_STATIC_MAX_EDGES / _MAX_ROWS) and no concurrency: all calls on a device share the one
// array, so two calls whose kernels can overlap on the GPU (different streams, different host
// threads) corrupt each other.  Serialise them.
namespace hgk {

constexpr int64_t kStaticScratchBytes =
    ((4 * (int64_t)HGK_CSR_STATIC_MAX_ROWS + 255) / 256) * 256 +
    2 * (((4 * (int64_t)HGK_CSR_STATIC_MAX_EDGES + 255) / 256) * 256) + 256;

static __device__ __align__(256) unsigned char g_csr_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_csr_static_addr[kMaxDevices] = {nullptr};
static std::mutex g_csr_static_mutex;

// In host code `g_csr_static_scratch` is only a handle for the runtime, not an address a kernel can
// use: the device address has to be asked for, once per device, with that device current.
static int csr_static_scratch(void** out) {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "csr_build_static: cudaGetDevice: %s", cudaGetErrorString(err));
  if (dev < 0 || dev >= kMaxDevices) return fail(HGK_ERR_UNSUPPORTED, "csr_build_static: device %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_csr_static_mutex);
  if (g_csr_static_addr[dev] == nullptr) {
    void* ptr = nullptr;
    err = cudaGetSymbolAddress(&ptr, g_csr_static_scratch);
    if (err != cudaSuccess || ptr == nullptr)
      return fail(HGK_ERR_CUDA, "csr_build_static: cudaGetSymbolAddress: %s", cudaGetErrorString(err));
    g_csr_static_addr[dev] = ptr;
  }
  *out = g_csr_static_addr[dev];
  return HGK_OK;
}

}  // namespace hgk

extern "C" int hgk_csr_build_static(const int32_t* key, int64_t E, int64_t N, int32_t* rowptr,
                                    int32_t* perm, void* stream) {
  HGK_REQUIRE(E >= 0 && N >= 0, "csr_build_static: negative size (E=%lld N=%lld)", (long long)E, (long long)N);
  if (E > HGK_CSR_STATIC_MAX_EDGES || N > HGK_CSR_STATIC_MAX_ROWS)
    return fail(HGK_ERR_UNSUPPORTED,
                "csr_build_static: E=%lld N=%lld exceeds the built-in scratch (max %lld edges, %lld rows); "
                "use hgk_csr_build with a caller-provided workspace",
                (long long)E, (long long)N, (long long)HGK_CSR_STATIC_MAX_EDGES, (long long)HGK_CSR_STATIC_MAX_ROWS);
  HGK_REQUIRE(rowptr != nullptr, "csr_build_static: rowptr is NULL");
  HGK_REQUIRE(E == 0 || (key && perm), "csr_build_static: key/perm is NULL with E>0");
  HGK_REQUIRE(E == 0 || N > 0, "csr_build_static: E>0 needs N>0");
  void* scratch = nullptr;
  int rc = csr_static_scratch(&scratch);
  if (rc) return rc;
  // The array keeps whatever the previous call left in it.  The only part that must start at zero
  // is the per-row counter block, which csr_build_impl clears on `stream` before the histogram;
  // slot[] and tmp[] are fully rewritten for the E edges of this call before they are read.
  return csr_build_impl(key, E, N, rowptr, perm, scratch, (cudaStream_t)stream);
}
