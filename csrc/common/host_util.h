// Host-side helpers shared by every translation unit of libdtm_b200: error reporting through the
// C ABI (integer status + thread-local message) and the lazily resolved driver entry points.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dtm {

// Records `msg` as the calling thread's last error and returns a non-zero status.
int dtm_fail(const char* msg);
int dtm_fail_cuda(cudaError_t e, const char* expr, const char* file, int line);

#define DTM_CUDA_OK(expr)                                                        \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess) return ::dtm::dtm_fail_cuda(_e, #expr, __FILE__, __LINE__); \
  } while (0)

// cuTensorMapEncodeTiled resolved through cudaGetDriverEntryPoint, so the library carries no link-time
// dependency on libcuda.so and can be dlopen'ed on a machine without a driver (build checks, CPU CI).
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PFN_encodeTiled get_encode_tiled();

// 2-D bf16 row-major tensor [outer][inner] (inner contiguous), box {box_inner, box_outer}.
// swizzle_bytes in {0, 32, 64, 128}. Out-of-bounds elements read as zero.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_outer, uint32_t swizzle_bytes);

int device_sm_count();
int device_slot();   // current device index clamped to [0, 64): index of per-device one-time set-up flags

// Access-policy window the optimizer update kernels are launched with (dtm_set_update_l2_window); null when off.
struct L2Window {
  void* base = nullptr;
  size_t bytes = 0;
  float hit_ratio = 0.0f;
};
const L2Window* update_l2_window();

// Programmatic dependent launch mode: 0 off (every kernel starts after its predecessor has drained), 1 strict, 2 early
// weights (see runtime.cu).  DTM_PDL=0|1|2 in the environment, or dtm_set_pdl_mode().
int pdl_mode();
bool pdl_enabled();

// Launch with the programmatic-stream-serialization attribute: the kernel may begin its prologue while the previous
// kernel of the stream is finishing; the kernel itself calls pdl_wait() (ptx.cuh) before it touches that kernel's output.
// cluster_z > 1 additionally groups cluster_z consecutive CTAs along grid z into a thread-block cluster.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster_z,
                                      Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 1;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = static_cast<unsigned>(cluster_z);
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  return launch_pdl_cluster(kern, grid, block, smem, s, 1, static_cast<Args&&>(args)...);
}

}  // namespace dtm
