// Re-packs the conv2 filter (bf16, TensorFlow layout [tap][ci][co]) into the shared-memory operand images of the two
// tensor-core kernels, once per step, so that every CTA obtains its 100 KiB of weights with one cp.async.bulk instead of
// 51 200 two-byte gathers (ncu: that prologue was 27 % of conv2_fwd's samples).
//   forward image: [k/8][co = 64][k%8], k = tap*32 + ci            value W2[tap][ci][co]
//   dgrad image  : [k/8][ci = 32][k%8], k = tap'*64 + co           value W2[24 - tap'][ci][co]
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"

namespace dtm {

__global__ void __launch_bounds__(256)
conv2_pack_kernel(const __nv_bfloat16* __restrict__ w2, __nv_bfloat16* __restrict__ fwd, __nv_bfloat16* __restrict__ dgrad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;   // flat index into W2: (tap*32 + ci)*64 + co
  if (i >= 25 * 32 * 64) return;
  const int co = i & 63, ci = (i >> 6) & 31, tap = i >> 11;
  const __nv_bfloat16 v = w2[i];
  const int k = tap * 32 + ci;
  fwd[((k >> 3) * 64 + co) * 8 + (k & 7)] = v;
  const int kd = (24 - tap) * 64 + co;
  dgrad[((kd >> 3) * 32 + ci) * 8 + (kd & 7)] = v;
}

}  // namespace dtm

extern "C" int dtm_conv2_pack_weights(const void* w2_bf16, void* packed_fwd, void* packed_dgrad, void* stream) {
  using namespace dtm;
  if (!w2_bf16 || !packed_fwd || !packed_dgrad) return dtm_fail("conv2_pack_weights: null argument");
  conv2_pack_kernel<<<(25 * 32 * 64 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(w2_bf16), static_cast<__nv_bfloat16*>(packed_fwd), static_cast<__nv_bfloat16*>(packed_dgrad));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
