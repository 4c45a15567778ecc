// conv1 of the MNIST CNN, forward and weight-gradient, on CUDA cores.
//
// Forward (reference mnist_input.py:39, :134-153): uint8 pixel -> x/255 - 0.5 -> 5x5 SAME conv (C_in 1 or 3 -> 32)
// -> +bias -> ReLU -> 2x2/2 max-pool [-> LRN(radius 4) for the `reference` preset], all in one kernel.  With
// K = 25*C_in <= 75 the layer is tensor-core hostile (and is 4 % of the model's FLOPs), so it is a direct
// convolution: one CTA per image, lane = output channel (so the LRN window is a handful of warp shuffles), the
// 6x6 input patch of a pooling window is broadcast from shared memory.  The output is written in the plane
// layout (plane_layout.cuh) that conv2's tcgen05 kernel consumes without any im2col.
//
// Weight gradient (autodiff of the same lines; conv1 has no input gradient): the pooled gradient is routed
// through the saved arg-max to one of four positions, so each (pooled pixel, channel) contributes exactly one
// 5x5 input patch; lanes keep the 25*C_in partial sums of their channel in registers.
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int C1_OUT = 32;
constexpr int IN_HW = 28;
constexpr int IN_PITCH = 34;  // 28 + 2*2 halo, +2 so the four pooling positions hit distinct banks
constexpr int IN_ROWS = 32;

constexpr float LRN_BIAS = 1.0f;
constexpr float LRN_ALPHA = 0.001f / 9.0f;
constexpr float LRN_BETA = 0.75f;
constexpr int LRN_RADIUS = 4;

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

template <int CIN, bool LRN>
__global__ void __launch_bounds__(256)
conv1_fwd_kernel(const void* __restrict__ x, int in_is_float, const float* __restrict__ w, const float* __restrict__ bias,
                 uint8_t* __restrict__ p1_planes, uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ pool_raw) {
  __shared__ float s_in[CIN * IN_ROWS * IN_PITCH];
  __shared__ float s_w[25 * CIN * C1_OUT];
  const int img = blockIdx.x;
  for (int i = threadIdx.x; i < 25 * CIN * C1_OUT; i += blockDim.x) s_w[i] = w[i];
  load_image_to_smem<CIN>(s_in, x, in_is_float, img);

  const int c = threadIdx.x & 31;
  const int grp = threadIdx.x >> 5;
  float wr[25 * CIN];
#pragma unroll
  for (int t = 0; t < 25 * CIN; ++t) wr[t] = s_w[t * C1_OUT + c];  // TF layout [kh][kw][ci][co]
  const float b = bias[c];

  uint8_t* out_img = p1_planes + static_cast<size_t>(img) * plane_image_bytes(C1_OUT);
  for (int pp = grp; pp < IMG * IMG; pp += 8) {
    const int ph = pp / IMG, pw = pp % IMG;
    float acc[2][2] = {{b, b}, {b, b}};
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci) {
      float patch[6][6];
      const float* src = s_in + (ci * IN_ROWS + 2 * ph) * IN_PITCH + 2 * pw;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int q = 0; q < 6; ++q) patch[r][q] = src[r * IN_PITCH + q];
#pragma unroll
      for (int kh = 0; kh < 5; ++kh)
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) {
          const float wv = wr[(kh * 5 + kw) * CIN + ci];
          acc[0][0] = fmaf(patch[kh][kw], wv, acc[0][0]);
          acc[0][1] = fmaf(patch[kh][kw + 1], wv, acc[0][1]);
          acc[1][0] = fmaf(patch[kh + 1][kw], wv, acc[1][0]);
          acc[1][1] = fmaf(patch[kh + 1][kw + 1], wv, acc[1][1]);
        }
    }
    // ReLU + 2x2 max-pool, first maximum wins (scan order (0,0),(0,1),(1,0),(1,1))
    float best = fmaxf(acc[0][0], 0.0f);
    int arg = 0;
    float v = fmaxf(acc[0][1], 0.0f); if (v > best) { best = v; arg = 1; }
    v = fmaxf(acc[1][0], 0.0f);       if (v > best) { best = v; arg = 2; }
    v = fmaxf(acc[1][1], 0.0f);       if (v > best) { best = v; arg = 3; }
    argmax[(static_cast<size_t>(img) * (IMG * IMG) + pp) * C1_OUT + c] = static_cast<uint8_t>(arg);
    float y = best;
    if (LRN) {
      pool_raw[(static_cast<size_t>(img) * (IMG * IMG) + pp) * C1_OUT + c] = __float2bfloat16(best);
      // the backward pass sees the bf16-rounded activation; normalise the same value
      const float a = __bfloat162float(__float2bfloat16(best));
      const float sq = a * a;
      float s = sq;
#pragma unroll
      for (int d = 1; d <= LRN_RADIUS; ++d) {
        const float up = __shfl_up_sync(0xffffffffu, sq, d);
        const float dn = __shfl_down_sync(0xffffffffu, sq, d);
        if (c - d >= 0) s += up;
        if (c + d < C1_OUT) s += dn;
      }
      y = a * __powf(LRN_BIAS + LRN_ALPHA * s, -LRN_BETA);
    }
    *reinterpret_cast<__nv_bfloat16*>(out_img + plane_offset(ph, pw, c)) = __float2bfloat16(y);
  }
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

extern "C" int dtm_conv1_fwd(const void* x, int in_is_float, int channels, const float* w, const float* bias, int batch,
                             int lrn, void* p1_planes, void* argmax, void* pool_raw, void* stream) {
  using namespace dtm;
  if (channels != 1 && channels != 3) return dtm_fail("conv1_fwd: channels must be 1 or 3");
  if (batch <= 0) return dtm_fail("conv1_fwd: batch must be positive");
  if (lrn && !pool_raw) return dtm_fail("conv1_fwd: LRN needs the pool_raw output buffer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* p = static_cast<uint8_t*>(p1_planes);
  uint8_t* a = static_cast<uint8_t*>(argmax);
  __nv_bfloat16* r = static_cast<__nv_bfloat16*>(pool_raw);
  if (channels == 1) {
    if (lrn) conv1_fwd_kernel<1, true><<<batch, 256, 0, s>>>(x, in_is_float, w, bias, p, a, r);
    else     conv1_fwd_kernel<1, false><<<batch, 256, 0, s>>>(x, in_is_float, w, bias, p, a, r);
  } else {
    if (lrn) conv1_fwd_kernel<3, true><<<batch, 256, 0, s>>>(x, in_is_float, w, bias, p, a, r);
    else     conv1_fwd_kernel<3, false><<<batch, 256, 0, s>>>(x, in_is_float, w, bias, p, a, r);
  }
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_conv1_wgrad(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax,
                               int batch, float* dw, float* db, void* stream) {
  using namespace dtm;
  if (channels != 1 && channels != 3) return dtm_fail("conv1_wgrad: channels must be 1 or 3");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = min(batch, 4 * device_sm_count());   // 3-4 CTAs per SM: the loop is LDS/FMA latency bound
  const __nv_bfloat16* g = static_cast<const __nv_bfloat16*>(grad_pooled);
  const uint8_t* a = static_cast<const uint8_t*>(argmax);
  if (channels == 1) conv1_wgrad_kernel<1><<<grid, 256, 0, s>>>(x, in_is_float, g, a, batch, dw, db);
  else               conv1_wgrad_kernel<3><<<grid, 256, 0, s>>>(x, in_is_float, g, a, batch, dw, db);
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
