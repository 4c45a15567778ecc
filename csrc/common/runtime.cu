// Runtime glue of libdtm_b200: error strings, device queries, TMA tensor-map construction.
#include <stdlib.h>
#include "host_util.h"
#include "../../include/dtm_b200.h"
#include <mutex>
#include <stdio.h>
#include <string.h>

namespace dtm {

static thread_local char g_last_error[512] = "";

int dtm_fail(const char* msg) {
  snprintf(g_last_error, sizeof(g_last_error), "%s", msg);
  return 1;
}

int dtm_fail_cuda(cudaError_t e, const char* expr, const char* file, int line) {
  snprintf(g_last_error, sizeof(g_last_error), "CUDA error %d (%s) in `%s` at %s:%d", (int)e, cudaGetErrorString(e),
           expr, file, line);
  return 2;
}

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_outer, uint32_t swizzle_bytes) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return dtm_fail("cuTensorMapEncodeTiled is not available (no CUDA driver?)");
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return dtm_fail("TMA base pointer must be 16-byte aligned");
  if (row_stride_bytes % 16 != 0) return dtm_fail("TMA row stride must be a multiple of 16 bytes");
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 32) sw = CU_TENSOR_MAP_SWIZZLE_32B;
  else if (swizzle_bytes == 64) sw = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 128) sw = CU_TENSOR_MAP_SWIZZLE_128B;
  else if (swizzle_bytes != 0) return dtm_fail("unsupported TMA swizzle");
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu box=%ux%u)", (int)r,
             (unsigned long long)inner, (unsigned long long)outer, box_inner, box_outer);
    return dtm_fail(buf);
  }
  return 0;
}

int device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return 0;
  return dev < 64 ? dev : 63;
}

int device_sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// 0 = off, 1 = strict (a kernel reads nothing from global memory before its dependency wait), 2 = early weights (kernels
// may stage WEIGHTS before the wait; only valid when the caller knows that no weight is written by the kernel launched
// immediately before - true for the training engine's schedule, which opts in around its own launches).
static thread_local int g_pdl_mode = -1;   // per host thread: the towers of a single-process job set it around their own launches
int pdl_mode() {
  if (g_pdl_mode >= 0) return g_pdl_mode;
  const char* e = getenv("DTM_PDL");   // read per launch: a graph capture bakes the choice in, tools can A/B in one process
  if (!e) return 1;
  return e[0] == '0' ? 0 : (e[0] == '2' ? 2 : 1);
}
bool pdl_enabled() { return pdl_mode() != 0; }

}  // namespace dtm

extern "C" int dtm_set_pdl_mode(int mode) {
  const int prev = dtm::g_pdl_mode;
  dtm::g_pdl_mode = mode < 0 ? -1 : (mode > 2 ? 2 : mode);
  return prev;
}

// ---- persisting-L2 window of the optimizer state --------------------------------------------------------------------
// B200's L2 (126 MB) is larger than the whole optimizer state of the MNIST models (fp32 weights, gradient, two Adam slots,
// EMA shadows, bf16 shadow: 72 MB for the baseline preset).  The update kernels are launched with an access-policy window
// over that state: their lines are kept in the persisting carve-out from one step to the next, so the read-modify-write of
// the optimizer is served by the L2 instead of HBM.
namespace dtm {
static L2Window g_l2_window[64];
const L2Window* update_l2_window() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  return g_l2_window[dev].bytes ? &g_l2_window[dev] : nullptr;
}
}  // namespace dtm

extern "C" int dtm_set_update_l2_window(const void* base, size_t bytes, size_t* carveout_bytes, size_t* max_window_bytes) {
  using namespace dtm;
  int dev = 0;
  DTM_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return dtm_fail("l2 window: device index out of range");
  int max_persist = 0, max_window = 0;
  DTM_CUDA_OK(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
  DTM_CUDA_OK(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev));
  if (carveout_bytes) *carveout_bytes = static_cast<size_t>(max_persist);
  if (max_window_bytes) *max_window_bytes = static_cast<size_t>(max_window);
  if (!base || bytes == 0 || bytes > static_cast<size_t>(max_persist) || bytes > static_cast<size_t>(max_window)) {
    // off - also when the state does not fit the carve-out: a partial window makes the update kernels thrash the lines they
    // pinned, and a carve-out left behind by an earlier engine takes 60 % of the L2 away from everybody else (measured on the
    // wide preset: 497 -> 709 us per step)
    g_l2_window[dev] = L2Window{};
    if (max_persist > 0) {
      DTM_CUDA_OK(cudaDeviceSynchronize());
      DTM_CUDA_OK(cudaCtxResetPersistingL2Cache());
      DTM_CUDA_OK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0));
    }
    return 0;
  }
  if (max_persist <= 0 || max_window <= 0) return dtm_fail("l2 window: this device has no persisting L2 cache");
  size_t want = bytes;
  size_t have = 0;
  DTM_CUDA_OK(cudaDeviceGetLimit(&have, cudaLimitPersistingL2CacheSize));
  if (have < want) DTM_CUDA_OK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
  size_t got = 0;
  DTM_CUDA_OK(cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize));
  L2Window w;
  w.base = const_cast<void*>(base);
  w.bytes = bytes < static_cast<size_t>(max_window) ? bytes : static_cast<size_t>(max_window);
  w.hit_ratio = got >= w.bytes ? 1.0f : static_cast<float>(static_cast<double>(got) / static_cast<double>(w.bytes));
  g_l2_window[dev] = w;
  if (carveout_bytes) *carveout_bytes = got;
  return 0;
}

extern "C" const char* dtm_last_error(void) { return dtm::g_last_error; }

extern "C" int dtm_abi_version(void) { return DTM_ABI_VERSION; }

extern "C" int dtm_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin) {
  int dev = 0;
  DTM_CUDA_OK(cudaGetDevice(&dev));
  cudaDeviceProp p;
  DTM_CUDA_OK(cudaGetDeviceProperties(&p, dev));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (smem_optin) *smem_optin = p.sharedMemPerBlockOptin;
  return 0;
}
