// umma_probe: a hardware-truth harness for tcgen05 shared-memory descriptors and TMEM layouts.
//
// The host hands over (a) a raw byte image that is copied verbatim into shared memory, (b) a list of
// tcgen05.mma operations given as raw 64-bit A/B descriptors (start addresses relative to the image
// base), the 32-bit instruction descriptor, the destination TMEM column and the accumulate flag.  The
// kernel zero-fills TMEM, issues the MMAs from one thread, and dumps all 128 TMEM lanes x ncols back
// to global memory.  Python builds the layouts/descriptors (tools/umma_probe.py) and compares the dump
// with a float32 matmul, so every layout hypothesis used by the conv/GEMM kernels is checked against
// the silicon rather than against documentation.
#include "../common/ptx.cuh"
#include "../../include/dtm_b200.h"
#include "../common/host_util.h"

namespace dtm {

struct ProbeOp {
  uint64_t desc_a;
  uint64_t desc_b;
  uint32_t idesc;
  uint32_t tmem_col;
  uint32_t accumulate;
  uint32_t pad;
};

DTM_DEVICE void tmem_st_zero_x16(uint32_t taddr) {
  uint32_t z = 0;
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr),
      "r"(z)
      : "memory");
}

__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(const uint8_t* __restrict__ image, uint32_t image_bytes, const ProbeOp* __restrict__ ops,
                  uint32_t n_ops, float* __restrict__ dump, uint32_t ncols) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_base_slot;

  const uint32_t warp = threadIdx.x >> 5;
  if (warp == 0) {
    tmem_alloc(&tmem_base_slot, 512);
    tmem_relinquish();
  }
  if (threadIdx.x == 32) {
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  for (uint32_t i = threadIdx.x * 16; i < image_bytes; i += blockDim.x * 16)
    *reinterpret_cast<uint4*>(smem + i) = *reinterpret_cast<const uint4*>(image + i);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  // zero the accumulator window
  for (uint32_t c = 0; c < ncols; c += 16) tmem_st_zero_x16(tmem_base + ((warp * 32u) << 16) + c);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (threadIdx.x == 0) {
    const uint32_t base16 = smem_u32(smem) >> 4;
    for (uint32_t i = 0; i < n_ops; ++i) {
      ProbeOp op = ops[i];
      uint64_t da = (op.desc_a & ~0x3FFFull) | ((op.desc_a + base16) & 0x3FFFull);
      uint64_t db = (op.desc_b & ~0x3FFFull) | ((op.desc_b + base16) & 0x3FFFull);
      umma_f16(tmem_base + op.tmem_col, da, db, op.idesc, op.accumulate);
    }
    umma_commit(&done_bar);
  }
  mbar_wait(&done_bar, 0, 99);
  tc_fence_after();

  for (uint32_t c = 0; c < ncols; c += 16) {
    uint32_t r[16];
    tmem_ld_x16(tmem_base + ((warp * 32u) << 16) + c, r);
    tmem_ld_wait();
    float* out = dump + static_cast<size_t>(threadIdx.x) * ncols + c;
#pragma unroll
    for (int j = 0; j < 16; ++j) out[j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

}  // namespace dtm

extern "C" int dtm_umma_probe(const void* image, uint32_t image_bytes, const void* ops, uint32_t n_ops, float* dump,
                              uint32_t ncols, void* stream) {
  using namespace dtm;
  if (image_bytes % 16 != 0 || image_bytes > 200 * 1024) return dtm_fail("umma_probe: image must be 16B-multiple <= 200 KiB");
  if (ncols % 16 != 0 || ncols > 512) return dtm_fail("umma_probe: ncols must be a multiple of 16, <= 512");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DTM_CUDA_OK(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  umma_probe_kernel<<<1, 128, 200 * 1024, s>>>(static_cast<const uint8_t*>(image), image_bytes,
                                               static_cast<const ProbeOp*>(ops), n_ops, dump, ncols);
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
