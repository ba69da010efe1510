// noalloc.cu — irf_depth_scan_static: the depth scan (difference counters -> per-base depth) for callers that must not
// allocate device memory.
//
// Who cannot allocate?  Two kinds of hosts we have met:
//   * a service whose GPU memory is carved up once at start-up (everything handed to its own pool, nothing left for the
//     CUDA allocator), where a late cudaMalloc simply fails or eats into another tenant's share;
//   * a latency-sensitive loop that keeps several streams busy: cudaMalloc and cudaFree synchronise the whole device, so
//     one allocation inside the loop stalls every stream, not only the one doing the scan.
// The regular path (irf_finalize) keeps its tile-sum scratch in a cudaMalloc'ed buffer owned by the context.  This
// variant keeps the scratch in a fixed array inside the library image instead, so a call touches only the caller's input
// and output buffers plus that array.
//
// Cost and limits, stated plainly:
//   * NA_MAX_ELEMS = 2^27 counters per call (512 MB of int32); the scratch is NA_MAX_TILES * 4 B = 128 KB.  Every GPU that
//     loads this library carries those 128 KB until the process exits, whether or not the entry point is ever used.
//   * All calls on one GPU share the one array and it keeps whatever the last call left there.  The call zeroes the part it
//     is about to use before launching, so stale contents never leak into a result — but two calls that overlap in time on
//     the same GPU (different streams, different threads) write the same array and corrupt each other.  Serialise them.
//   * The address of the array is looked up once per GPU, by the first call made while that GPU is current.  That first
//     call performs a symbol look-up in the CUDA runtime, so make it during set-up, before any stream capture begins.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <mutex>

#include "../../include/irfinder_b200.h"

namespace {

constexpr int NA_THREADS = 256, NA_ITEMS = 16, NA_TILE = NA_THREADS * NA_ITEMS;   // 4096 counters per tile
constexpr unsigned long long NA_MAX_ELEMS = 1ull << 27;
constexpr unsigned int NA_MAX_TILES = (unsigned int)(NA_MAX_ELEMS / NA_TILE);      // 32768

// The scratch itself.  In host code this name is only a handle for the runtime, not a device address: a kernel handed
// `na_tile_sums` directly from the host would read the wrong memory.  Host code below therefore asks the runtime for the
// real address (cudaGetSymbolAddress) and passes that pointer to the kernels.
static __device__ __align__(256) int32_t na_tile_sums[NA_MAX_TILES];

__global__ void __launch_bounds__(NA_THREADS) na_tile_sums_kernel(const int32_t* __restrict__ in, unsigned long long n, int32_t* __restrict__ sums) {
  unsigned long long base = (unsigned long long)blockIdx.x * NA_TILE; int acc = 0;
  for (int k = 0; k < NA_ITEMS; k++) { unsigned long long i = base + (unsigned long long)k * NA_THREADS + threadIdx.x; if (i < n) acc += in[i]; }
  __shared__ int sh[NA_THREADS / 32];
  for (int o = 16; o; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < NA_THREADS / 32; w++) t += sh[w]; sums[blockIdx.x] = t; }
}
// exclusive scan of the tile sums, one CTA (at most 32768 values)
__global__ void __launch_bounds__(1024) na_scan_sums_kernel(int32_t* sums, unsigned int nt) {
  __shared__ int sh[1024]; __shared__ int carry; if (threadIdx.x == 0) carry = 0; __syncthreads();
  for (unsigned int base = 0; base < nt; base += 1024) {
    unsigned int i = base + threadIdx.x; int v = i < nt ? sums[i] : 0; sh[threadIdx.x] = v; __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0; __syncthreads(); sh[threadIdx.x] += t; __syncthreads(); }
    int incl = sh[threadIdx.x], c = carry; if (i < nt) sums[i] = c + incl - v; __syncthreads();
    if (threadIdx.x == 1023) carry = c + incl; __syncthreads();
  }
}
// out[i] = in[0] + ... + in[i]; `out` may be the same buffer as `in`
__global__ void __launch_bounds__(NA_THREADS) na_apply_kernel(const int32_t* in, int32_t* out, unsigned long long n, const int32_t* __restrict__ sums) {
  unsigned long long base = (unsigned long long)blockIdx.x * NA_TILE + (unsigned long long)threadIdx.x * NA_ITEMS; int v[NA_ITEMS]; int tot = 0;
  for (int k = 0; k < NA_ITEMS; k++) v[k] = base + k < n ? in[base + k] : 0;
#pragma unroll
  for (int k = 0; k < NA_ITEMS; k++) { tot += v[k]; v[k] = tot; }
  int incl = tot; const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  __shared__ int wsum[NA_THREADS / 32];
  if (lane == 31) wsum[warp] = incl; __syncthreads();
  int off = sums[blockIdx.x] + incl - tot; for (int w = 0; w < warp; w++) off += wsum[w];
  for (int k = 0; k < NA_ITEMS; k++) if (base + k < n) out[base + k] = v[k] + off;
}

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_lock;
int32_t* g_scratch[NA_MAX_DEVICES] = {nullptr};   // device address of na_tile_sums on each GPU, filled on first use

int fail(int code, char* errbuf, size_t len, const char* fmt, const char* detail, unsigned long long a, unsigned long long b) {
  if (errbuf && len) snprintf(errbuf, len, fmt, detail, a, b);
  return code;
}

}  // namespace

extern "C" int irf_depth_scan_static(const int32_t* diff_dev, int32_t* depth_dev, size_t n, void* cuda_stream, char* errbuf, size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (n == 0) return IRF_OK;
  if (!diff_dev || !depth_dev) return fail(IRF_EINVAL, errbuf, errbuf_len, "irf_depth_scan_static: %s (n = %llu)%.0llu", "null device pointer", (unsigned long long)n, 0);
  if ((unsigned long long)n > NA_MAX_ELEMS)
    return fail(IRF_ECAPACITY, errbuf, errbuf_len, "irf_depth_scan_static: %s: %llu counters requested, the built-in scratch covers at most %llu per call; split the array or use irf_finalize",
                "size over the fixed limit", (unsigned long long)n, NA_MAX_ELEMS);
  int dev = -1; cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) return fail(IRF_ECUDA, errbuf, errbuf_len, "irf_depth_scan_static: %s (device %llu)%.0llu", e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range", (unsigned long long)(dev < 0 ? 0 : dev), 0);
  int32_t* sums = nullptr;
  {
    std::lock_guard<std::mutex> g(g_lock);
    if (!g_scratch[dev]) {   // first call with this GPU current: resolve the array's device address once
      void* p = nullptr; e = cudaGetSymbolAddress(&p, na_tile_sums);
      if (e != cudaSuccess || !p) return fail(IRF_ECUDA, errbuf, errbuf_len, "irf_depth_scan_static: cudaGetSymbolAddress failed: %s%.0llu%.0llu", cudaGetErrorString(e), 0, 0);
      g_scratch[dev] = (int32_t*)p;
    }
    sums = g_scratch[dev];
  }
  cudaStream_t s = (cudaStream_t)cuda_stream;
  const unsigned int nt = (unsigned int)((n + NA_TILE - 1) / NA_TILE);
  // the array still holds the previous call's tile sums: start this call's part from zero
  e = cudaMemsetAsync(sums, 0, (size_t)nt * sizeof(int32_t), s);
  if (e != cudaSuccess) return fail(IRF_ECUDA, errbuf, errbuf_len, "irf_depth_scan_static: clearing the scratch failed: %s%.0llu%.0llu", cudaGetErrorString(e), 0, 0);
  na_tile_sums_kernel<<<nt, NA_THREADS, 0, s>>>(diff_dev, (unsigned long long)n, sums);
  na_scan_sums_kernel<<<1, 1024, 0, s>>>(sums, nt);
  na_apply_kernel<<<nt, NA_THREADS, 0, s>>>(diff_dev, depth_dev, (unsigned long long)n, sums);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(IRF_ECUDA, errbuf, errbuf_len, "irf_depth_scan_static: kernel launch failed: %s%.0llu%.0llu", cudaGetErrorString(e), 0, 0);
  return IRF_OK;
}
