// Summary statistics for the chief's monitoring path (reference main.py:95-109 via monitor_cb: SummaryHist,
// SummaryNorm2, SummaryGradient, SummaryGWRatio): min / max / sum / sum-of-squares of a tensor in one pass, and
// the TensorBoard histogram (TensorFlow's exponential bucket boundaries) — computed on the device so that only a
// few hundred bytes cross PCIe every `test_interval` steps instead of the 13 MB of weights and activations.
//
// tensor_stats is a two-level reduction: every CTA writes one partial (min, max, sum, sumsq as doubles) into a
// scratch buffer, the last CTA to finish (ticket counter at the head of the scratch) folds the partials in index
// order, so the result does not depend on scheduling.  The ticket is reset by that CTA: the scratch only has to be
// zero before its first use.
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include <float.h>

namespace dtm {

constexpr int ST_THREADS = 256;
constexpr int ST_MAX_BLOCKS = 1024;
constexpr size_t ST_HEADER = 64;  // bytes: ticket + padding

struct StatPartial {
  double mn, mx, sum, sumsq;
};

template <typename T>
__device__ __forceinline__ float load_as_float(const T* p, size_t i);
template <>
__device__ __forceinline__ float load_as_float<float>(const float* p, size_t i) { return p[i]; }
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) { return __bfloat162float(p[i]); }

template <typename T>
__global__ void __launch_bounds__(ST_THREADS)
tensor_stats_kernel(const T* __restrict__ x, size_t n, float* __restrict__ out, uint8_t* __restrict__ scratch) {
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  StatPartial* partials = reinterpret_cast<StatPartial*>(scratch + ST_HEADER);
  double mn = DBL_MAX, mx = -DBL_MAX, sum = 0.0, sumsq = 0.0;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const double v = load_as_float<T>(x, i);
    mn = fmin(mn, v);
    mx = fmax(mx, v);
    sum += v;
    sumsq += v * v;
  }
  __shared__ StatPartial s_part[ST_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
  }
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = StatPartial{mn, mx, sum, sumsq};
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) {
    StatPartial p = s_part[0];
    for (int w = 1; w < ST_THREADS / 32; ++w) {
      p.mn = fmin(p.mn, s_part[w].mn);
      p.mx = fmax(p.mx, s_part[w].mx);
      p.sum += s_part[w].sum;
      p.sumsq += s_part[w].sumsq;
    }
    partials[blockIdx.x] = p;
    __threadfence();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    StatPartial t = partials[0];
    for (unsigned int b = 1; b < gridDim.x; ++b) {
      const StatPartial p = partials[b];
      t.mn = fmin(t.mn, p.mn);
      t.mx = fmax(t.mx, p.mx);
      t.sum += p.sum;
      t.sumsq += p.sumsq;
    }
    out[0] = static_cast<float>(t.mn);
    out[1] = static_cast<float>(t.mx);
    out[2] = static_cast<float>(t.sum);
    out[3] = static_cast<float>(t.sumsq);
    out[4] = static_cast<float>(n);
    out[5] = static_cast<float>(sqrt(t.sumsq));  // L2 norm
    *ticket = 0;                                  // ready for the next call
  }
}

// counts[b] += #elements whose value falls in bucket b, bucket b = (limits[b-1], limits[b]] boundaries as in
// tensorflow::histogram: index = upper_bound(limits, v), clamped to the last bucket.
template <typename T>
__global__ void __launch_bounds__(256)
histogram_kernel(const T* __restrict__ x, size_t n, const double* __restrict__ limits, int nbins, float* __restrict__ counts) {
  extern __shared__ unsigned int s_cnt[];
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) s_cnt[i] = 0;
  __syncthreads();
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const double v = load_as_float<T>(x, i);
    int lo = 0, hi = nbins;  // first index with limits[idx] > v
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (limits[mid] > v) hi = mid; else lo = mid + 1;
    }
    atomicAdd(&s_cnt[min(lo, nbins - 1)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    if (s_cnt[i]) atomicAdd(&counts[i], static_cast<float>(s_cnt[i]));
}

static int stats_grid(size_t n) {
  size_t b = (n + ST_THREADS * 8 - 1) / (ST_THREADS * 8);
  if (b < 1) b = 1;
  if (b > ST_MAX_BLOCKS) b = ST_MAX_BLOCKS;
  return static_cast<int>(b);
}

}  // namespace dtm

extern "C" size_t dtm_tensor_stats_scratch_bytes(size_t n) {
  return dtm::ST_HEADER + static_cast<size_t>(dtm::stats_grid(n)) * sizeof(dtm::StatPartial);
}

extern "C" int dtm_tensor_stats(const void* x, size_t n, int is_bf16, float* out6, void* scratch, size_t scratch_bytes,
                                void* stream) {
  using namespace dtm;
  if (n == 0) return dtm_fail("tensor_stats: empty tensor");
  if (!scratch || scratch_bytes < dtm_tensor_stats_scratch_bytes(n)) return dtm_fail("tensor_stats: scratch buffer too small");
  if (reinterpret_cast<uintptr_t>(scratch) % 8 != 0) return dtm_fail("tensor_stats: scratch must be 8-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = stats_grid(n);
  if (is_bf16)
    tensor_stats_kernel<__nv_bfloat16><<<grid, ST_THREADS, 0, s>>>(static_cast<const __nv_bfloat16*>(x), n, out6, static_cast<uint8_t*>(scratch));
  else
    tensor_stats_kernel<float><<<grid, ST_THREADS, 0, s>>>(static_cast<const float*>(x), n, out6, static_cast<uint8_t*>(scratch));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_histogram(const void* x, size_t n, int is_bf16, const double* limits, int nbins, float* counts, void* stream) {
  using namespace dtm;
  if (n == 0 || nbins <= 0) return dtm_fail("histogram: empty input");
  if (nbins > 8192) return dtm_fail("histogram: at most 8192 buckets");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = stats_grid(n);
  const size_t smem = static_cast<size_t>(nbins) * sizeof(unsigned int);
  if (is_bf16)
    histogram_kernel<__nv_bfloat16><<<grid, 256, smem, s>>>(static_cast<const __nv_bfloat16*>(x), n, limits, nbins, counts);
  else
    histogram_kernel<float><<<grid, 256, smem, s>>>(static_cast<const float*>(x), n, limits, nbins, counts);
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// dtm_tensor_stats_static: the same reduction for callers that have no device memory to hand over as scratch.
//
// Who that is: a monitoring or diagnostics hook that fires while every byte of the GPU is already spoken for (the
// trainer sized its buffers to the card, or an out-of-memory / NaN post-mortem is being written), a watcher thread
// that must not call cudaMalloc because that call can synchronise the whole device and stall the very training
// streams it observes, or a plain-C host (this ABI works without PyTorch) that has no device allocator of its own.
//
// Instead of a caller-provided workspace the per-CTA partials live in the file-scope array below.  It is part of the
// library image, so every GPU that loads libdtm_b200 carries these 16 KiB until the process exits — hence the cap.
// Size limit: the partials of dtm_tensor_stats for n elements must fit, i.e. n <= DTM_STATS_STATIC_MAX_ELEMS;
// larger tensors are refused (use dtm_tensor_stats with your own scratch).
//
// On the host, `g_stats_static_scratch` is only a stand-in symbol, not a device address; the address a kernel can use is
// obtained once per device with cudaGetSymbolAddress while that device is current.  That look-up happens inside the
// first call made on each device, so make that first call before starting any stream capture.
//
// All calls on a device share the one array: two calls whose kernels may run at the same time (different streams,
// different threads) race on the partials and on the ticket and return garbage — serialise them.  The array keeps
// whatever the previous call left behind; the only part that has to start at zero is the 64-byte ticket header, and
// the entry point clears it on the caller's stream before every launch, so a call that died half-way cannot poison
// the next one.  Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

#define DTM_STATS_STATIC_SCRATCH_BYTES 16384

static __device__ __align__(256) unsigned char g_stats_static_scratch[DTM_STATS_STATIC_SCRATCH_BYTES];

namespace dtm {

constexpr size_t ST_STATIC_MAX_BLOCKS = (DTM_STATS_STATIC_SCRATCH_BYTES - ST_HEADER) / sizeof(StatPartial);   // 510
constexpr size_t ST_STATIC_MAX_ELEMS = ST_STATIC_MAX_BLOCKS * ST_THREADS * 8;                               // 1 044 480

static_assert(DTM_STATS_STATIC_MAX_ELEMS == ST_STATIC_MAX_ELEMS, "header constant out of sync with the kernel geometry");

// Device address of the static scratch for the *current* device, resolved on first use per device.
static int static_scratch_address(void** out) {
  static std::mutex mu;
  static void* addr[64] = {};
  int dev = -1;
  DTM_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return dtm_fail("tensor_stats_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    DTM_CUDA_OK(cudaGetSymbolAddress(&p, g_stats_static_scratch));
    if (!p) return dtm_fail("tensor_stats_static: cudaGetSymbolAddress returned a null address");
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}

}  // namespace dtm

extern "C" int dtm_tensor_stats_static(const void* x, size_t n, int is_bf16, float* out6, void* stream) {
  using namespace dtm;
  if (n == 0) return dtm_fail("tensor_stats_static: empty tensor");
  if (!x || !out6) return dtm_fail("tensor_stats_static: null input or output pointer");
  if (n > ST_STATIC_MAX_ELEMS || dtm_tensor_stats_scratch_bytes(n) > DTM_STATS_STATIC_SCRATCH_BYTES)
    return dtm_fail("tensor_stats_static: tensor too large for the built-in 16 KiB scratch (limit 1044480 elements); "
                    "call dtm_tensor_stats with a caller-provided scratch buffer instead");
  void* scratch = nullptr;
  if (int rc = static_scratch_address(&scratch)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DTM_CUDA_OK(cudaMemsetAsync(scratch, 0, ST_HEADER, s));   // the ticket must start at zero whatever ran before
  const int grid = stats_grid(n);
  if (is_bf16)
    tensor_stats_kernel<__nv_bfloat16><<<grid, ST_THREADS, 0, s>>>(static_cast<const __nv_bfloat16*>(x), n, out6, static_cast<uint8_t*>(scratch));
  else
    tensor_stats_kernel<float><<<grid, ST_THREADS, 0, s>>>(static_cast<const float*>(x), n, out6, static_cast<uint8_t*>(scratch));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
