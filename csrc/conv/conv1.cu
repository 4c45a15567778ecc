// conv1 weight gradient on CUDA cores: the round-1 kernel, kept as the `DTM_WGRAD_FUSED=0` arm (the default is the fused
// tensor-core kernel in conv_wgrad_fused.cu; conv1_wgrad_tc.cu is a second opt-in variant; the forward pass lives in
// conv1_sw.cu / conv1_tc.cu).
//
// Weight gradient (autodiff of the same lines; conv1 has no input gradient): the pooled gradient is routed
// through the saved arg-max to one of four positions, so each (pooled pixel, channel) contributes exactly one
// 5x5 input patch; lanes keep the 25*C_in partial sums of their channel in registers.
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int C1_OUT = 32;
constexpr int IN_HW = 28;
constexpr int IN_PITCH = 34;  // 28 + 2*2 halo, +2 so the four pooling positions hit distinct banks
constexpr int IN_ROWS = 32;

template <int CIN>
DTM_DEVICE void load_image_to_smem(float* s_in, const void* x, int in_is_float, int img) {
  // s_in[ci][row][col] with a 2-pixel zero halo; row pitch IN_PITCH
  for (int i = threadIdx.x; i < CIN * IN_ROWS * IN_PITCH; i += blockDim.x) s_in[i] = 0.0f;
  __syncthreads();
  const int n = IN_HW * IN_HW * CIN;
  if (in_is_float) {
    const float* src = static_cast<const float*>(x) + static_cast<size_t>(img) * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int ci = i % CIN, p = i / CIN;
      s_in[(ci * IN_ROWS + p / IN_HW + 2) * IN_PITCH + p % IN_HW + 2] = src[i];
    }
  } else {
    const uint8_t* src = static_cast<const uint8_t*>(x) + static_cast<size_t>(img) * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int ci = i % CIN, p = i / CIN;
      s_in[(ci * IN_ROWS + p / IN_HW + 2) * IN_PITCH + p % IN_HW + 2] = static_cast<float>(src[i]) * (1.0f / 255.0f) - 0.5f;
    }
  }
  __syncthreads();
}

// dW1[kh][kw][ci][co] += sum over images / pooled pixels; db1[co] likewise.  grad_pooled is the gradient with
// respect to the pooled conv1 activation, already multiplied by the ReLU mask ([B][196][32] bf16).
template <int CIN>
__global__ void __launch_bounds__(256)
conv1_wgrad_kernel(const void* __restrict__ x, int in_is_float, const __nv_bfloat16* __restrict__ grad_pooled,
                   const uint8_t* __restrict__ argmax, int batch, float* __restrict__ dw, float* __restrict__ db) {
  __shared__ float s_in[CIN * IN_ROWS * IN_PITCH];
  __shared__ float s_red[25 * CIN + 1][C1_OUT];
  __shared__ __align__(16) __nv_bfloat16 s_g[IMG * IMG * C1_OUT];   // 12.25 KiB: this image's pooled gradient
  __shared__ __align__(16) uint8_t s_a[IMG * IMG * C1_OUT];         // 6.1 KiB: its arg-max
  const int c = threadIdx.x & 31;
  const int grp = threadIdx.x >> 5;
  float acc[25 * CIN];
#pragma unroll
  for (int t = 0; t < 25 * CIN; ++t) acc[t] = 0.0f;
  float accb = 0.0f;
  pdl_wait();
  pdl_launch_dependents();

  for (int img = blockIdx.x; img < batch; img += gridDim.x) {
    __syncthreads();
    // one coalesced 16-byte sweep brings gradient and arg-max on chip; the per-pixel loop below then never waits on
    // a global load (it used to: ~25 dependent round trips per warp and image)
    {
      const uint4* g_src = reinterpret_cast<const uint4*>(grad_pooled + static_cast<size_t>(img) * (IMG * IMG) * C1_OUT);
      const uint4* a_src = reinterpret_cast<const uint4*>(argmax + static_cast<size_t>(img) * (IMG * IMG) * C1_OUT);
      for (int i = threadIdx.x; i < IMG * IMG * C1_OUT * 2 / 16; i += blockDim.x) reinterpret_cast<uint4*>(s_g)[i] = g_src[i];
      for (int i = threadIdx.x; i < IMG * IMG * C1_OUT / 16; i += blockDim.x) reinterpret_cast<uint4*>(s_a)[i] = a_src[i];
    }
    load_image_to_smem<CIN>(s_in, x, in_is_float, img);
    for (int pp = grp; pp < IMG * IMG; pp += 8) {
      const float g = __bfloat162float(s_g[pp * C1_OUT + c]);
      if (g == 0.0f) continue;
      const int a = s_a[pp * C1_OUT + c];
      const int ph = pp / IMG, pw = pp % IMG;
      accb += g;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        const float* src = s_in + (ci * IN_ROWS + 2 * ph + (a >> 1)) * IN_PITCH + 2 * pw + (a & 1);
#pragma unroll
        for (int kh = 0; kh < 5; ++kh)
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) acc[(kh * 5 + kw) * CIN + ci] = fmaf(src[kh * IN_PITCH + kw], g, acc[(kh * 5 + kw) * CIN + ci]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (25 * CIN + 1) * C1_OUT; i += blockDim.x) (&s_red[0][0])[i] = 0.0f;
  __syncthreads();
#pragma unroll
  for (int t = 0; t < 25 * CIN; ++t) atomicAdd(&s_red[t][c], acc[t]);
  atomicAdd(&s_red[25 * CIN][c], accb);
  __syncthreads();
  for (int i = threadIdx.x; i < (25 * CIN + 1) * C1_OUT; i += blockDim.x) {
    const int t = i / C1_OUT, cc = i % C1_OUT;
    const float s = s_red[t][cc];
    if (t < 25 * CIN) red_add_f32(dw + t * C1_OUT + cc, s);
    else red_add_f32(db + cc, s);
  }
}

}  // namespace dtm

extern "C" int dtm_conv1_wgrad_tc(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax,
                                  int batch, float* dw, float* db, void* stream);

extern "C" int dtm_conv1_wgrad(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax,
                               int batch, float* dw, float* db, void* stream) {
  using namespace dtm;
  // DTM_C1W_TC=1 selects the tensor-core kernel (conv1_wgrad_tc.cu).  It is correct and needs 1.7x fewer instructions, but
  // as measured in round 1 it is slower than this kernel (41 vs 29 us alone, 60 vs 44 us beside conv2_wgrad: its producers
  // wait on per-row global loads, and conv2_wgrad's 512-column TMEM allocation keeps it off the SM) - so it is opt-in.
  {
    const char* e = getenv("DTM_C1W_TC");
    if (e && e[0] == '1') return dtm_conv1_wgrad_tc(x, in_is_float, channels, grad_pooled, argmax, batch, dw, db, stream);
  }
  if (channels != 1 && channels != 3) return dtm_fail("conv1_wgrad: channels must be 1 or 3");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Every CTA ends with 26*C_in*32 red.add into the same few cache lines, so the CTA count sets the length of that serialised
  // tail (512 single-image CTAs: ~16 K atomics per line, most of the kernel's time); a persistent grid of a couple of CTAs
  // per SM keeps enough warps for the latency-bound main loop and cuts the atomics by the number of images per CTA.
  const char* e = getenv("DTM_C1W_CTAS_PER_SM");
  const int per_sm = e ? max(1, atoi(e)) : 2;
  const int grid = min(batch, per_sm * device_sm_count());
  const __nv_bfloat16* g = static_cast<const __nv_bfloat16*>(grad_pooled);
  const uint8_t* a = static_cast<const uint8_t*>(argmax);
  if (channels == 1) DTM_CUDA_OK(launch_pdl(conv1_wgrad_kernel<1>, dim3(grid), dim3(256), 0, s, x, in_is_float, g, a, batch, dw, db));
  else               DTM_CUDA_OK(launch_pdl(conv1_wgrad_kernel<3>, dim3(grid), dim3(256), 0, s, x, in_is_float, g, a, batch, dw, db));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
