// Backward of conv2's pooling stage: routes the gradient of the pooled activation [B][7][7][64] back to the
// 14x14 grid and writes it in the plane layout consumed by the tcgen05 dgrad / wgrad kernels, plus the conv2
// bias gradient.  Two variants:
//   baseline : conv -> ReLU -> pool.            dZ2 = route(dP2)   (dP2 already carries the ReLU mask: the fc GEMM
//              epilogue multiplied it by [P2 > 0], and pooled > 0  <=>  the arg-max pixel passed the ReLU)
//   reference: conv -> ReLU -> LRN -> pool      (mnist_input.py:163-172) dZ2 = relu'(a) * lrn_bwd(route(dP2))
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "../conv/plane_layout.cuh"

namespace dtm {

constexpr int UP_C = 64;

__global__ void __launch_bounds__(256)
unpool2_kernel(const __nv_bfloat16* __restrict__ dp2, const uint8_t* __restrict__ argmax, int batch,
               uint8_t* __restrict__ dz2_planes, float* __restrict__ db2) {
  __shared__ float s_db[32][UP_C + 1];
  pdl_wait();
  pdl_launch_dependents();
  const int c8 = threadIdx.x & 7;
  const int r = threadIdx.x >> 3;
  float part[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) part[i] = 0.0f;
  const int rows = batch * 49;
  for (int row = blockIdx.x * 32 + r; row < rows; row += gridDim.x * 32) {
    const int img = row / 49, pp = row % 49;
    const int ph = pp / 7, pw = pp % 7;
    const uint4 gv = *reinterpret_cast<const uint4*>(dp2 + static_cast<size_t>(row) * UP_C + c8 * 8);
    const uint2 av = *reinterpret_cast<const uint2*>(argmax + static_cast<size_t>(row) * UP_C + c8 * 8);
    const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
    uint32_t arg[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) { arg[i] = (av.x >> (8 * i)) & 0xFF; arg[4 + i] = (av.y >> (8 * i)) & 0xFF; }
#pragma unroll
    for (int i = 0; i < 4; ++i) { part[2 * i] += bf16_lo(gw[i]); part[2 * i + 1] += bf16_hi(gw[i]); }
    uint8_t* base = dz2_planes + static_cast<size_t>(img) * plane_image_bytes(UP_C);
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t lo = (arg[2 * i] == d) ? (gw[i] & 0xFFFFu) : 0u;
        const uint32_t hi = (arg[2 * i + 1] == d) ? (gw[i] & 0xFFFF0000u) : 0u;
        o[i] = lo | hi;
      }
      *reinterpret_cast<uint4*>(base + plane_offset(2 * ph + (d >> 1), 2 * pw + (d & 1), c8 * 8)) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) s_db[r][c8 * 8 + i] = part[i];
  __syncthreads();
  if (threadIdx.x < UP_C) {
    float s = 0.0f;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) s += s_db[k][threadIdx.x];
    red_add_f32(db2 + threadIdx.x, s);
  }
}

// reference preset.  Eight lanes per pixel, eight channels per lane: the +-4-channel LRN windows reach into the two
// neighbouring lanes only, so the window sums need eight shuffles instead of 64 channels in one thread's registers (round 1:
// one thread per pixel, 255 registers with spills, 88 us for 100 K pixels).
__global__ void __launch_bounds__(256)
unpool2_lrn_kernel(const __nv_bfloat16* __restrict__ dp2, const uint8_t* __restrict__ argmax,
                   const __nv_bfloat16* __restrict__ act2, int batch, uint8_t* __restrict__ dz2_planes,
                   float* __restrict__ db2) {
  __shared__ float s_db[UP_C];
  if (threadIdx.x < UP_C) s_db[threadIdx.x] = 0.0f;
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  const int c8 = threadIdx.x & 7;
  const int total = batch * IMG * IMG;
  const float alpha = 0.001f / 9.0f, beta = 0.75f;
  float dbp[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) dbp[i] = 0.0f;
  // every lane of a warp runs the same number of rounds (the shuffles need all 32 lanes)
  for (int base = blockIdx.x * 32; base < total; base += gridDim.x * 32) {
    const int pix = base + (threadIdx.x >> 3);
    const bool active = pix < total;
    float g[8], a[8];
    int img = 0, h = 0, w = 0;
    if (active) {
      img = pix / (IMG * IMG);
      const int p = pix - img * (IMG * IMG);
      h = p / IMG;
      w = p - h * IMG;
      const uint32_t mypos = (h & 1) * 2 + (w & 1);
      const size_t prow = (static_cast<size_t>(img) * 49 + (h >> 1) * 7 + (w >> 1)) * UP_C + c8 * 8;
      const uint4 gv = *reinterpret_cast<const uint4*>(dp2 + prow);
      const uint2 av = *reinterpret_cast<const uint2*>(argmax + prow);
      const uint4 xv = *reinterpret_cast<const uint4*>(act2 + static_cast<size_t>(pix) * UP_C + c8 * 8);
      const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w};
      const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t a0 = ((i < 2 ? av.x : av.y) >> (16 * (i & 1))) & 0xFF;
        const uint32_t a1 = ((i < 2 ? av.x : av.y) >> (16 * (i & 1) + 8)) & 0xFF;
        g[2 * i] = a0 == mypos ? bf16_lo(gw[i]) : 0.0f;
        g[2 * i + 1] = a1 == mypos ? bf16_hi(gw[i]) : 0.0f;
        a[2 * i] = bf16_lo(xw[i]);
        a[2 * i + 1] = bf16_hi(xw[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) { g[i] = 0.0f; a[i] = 0.0f; }
    }
    // ext[0..3] = left neighbour's channels 4..7, ext[4..11] = own, ext[12..15] = right neighbour's channels 0..3
    auto widen = [&](const float (&v)[8], float (&ext)[16]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float l = __shfl_up_sync(0xffffffffu, v[4 + i], 1);
        const float r = __shfl_down_sync(0xffffffffu, v[i], 1);
        ext[i] = c8 > 0 ? l : 0.0f;
        ext[12 + i] = c8 < 7 ? r : 0.0f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ext[4 + i] = v[i];
    };
    float sq[8], ext[16], t1[8], sp[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) sq[i] = a[i] * a[i];
    widen(sq, ext);
    float win[8];
    window9_sums<8>(ext, win);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float s = 1.0f + alpha * win[i];
      float inv_s;
      const float pw = lrn_pow_m075(s, &inv_s);
      sp[i] = pw;
      t1[i] = g[i] * a[i] * pw * inv_s;
    }
    widen(t1, ext);
    window9_sums<8>(ext, win);
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float da = g[i] * sp[i] - 2.0f * alpha * beta * a[i] * win[i];
      o[i] = a[i] > 0.0f ? da : 0.0f;
    }
    if (active) {
      uint8_t* dst = dz2_planes + static_cast<size_t>(img) * plane_image_bytes(UP_C) + plane_offset(h, w, c8 * 8);
      const uint4 pk = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
      *reinterpret_cast<uint4*>(dst) = pk;
      // bias gradient: the bf16-rounded values the wgrad kernel will see
      const uint32_t pw4[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) { dbp[2 * i] += bf16_lo(pw4[i]); dbp[2 * i + 1] += bf16_hi(pw4[i]); }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float v = dbp[i];
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if (lane_id() < 8) atomicAdd(&s_db[c8 * 8 + i], v);
  }
  __syncthreads();
  if (threadIdx.x < UP_C) red_add_f32(db2 + threadIdx.x, s_db[threadIdx.x]);
}

// out[n] += sum_m X[m][n]  (bias gradients of layers whose dY comes out of a GEMM epilogue)
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, int M, int N, float* __restrict__ out) {
  const int n = blockIdx.x * 32 + (threadIdx.x & 31);
  const int r = threadIdx.x >> 5;
  __shared__ float s[8][33];
  pdl_wait();
  pdl_launch_dependents();
  float acc = 0.0f;
  if (n < N) {
    // four independent loads per pass: with one the loop ran at a single L2 round trip per element (48 us for 3 MB)
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const int step = gridDim.y * 8;
    int m = blockIdx.y * 8 + r;
    for (; m + 3 * step < M; m += 4 * step) {
      a0 += __bfloat162float(x[static_cast<size_t>(m) * N + n]);
      a1 += __bfloat162float(x[static_cast<size_t>(m + step) * N + n]);
      a2 += __bfloat162float(x[static_cast<size_t>(m + 2 * step) * N + n]);
      a3 += __bfloat162float(x[static_cast<size_t>(m + 3 * step) * N + n]);
    }
    for (; m < M; m += step) a0 += __bfloat162float(x[static_cast<size_t>(m) * N + n]);
    acc = (a0 + a1) + (a2 + a3);
  }
  s[r][threadIdx.x & 31] = acc;
  __syncthreads();
  if (r == 0 && n < N) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s[k][threadIdx.x];
    red_add_f32(out + n, t);
  }
}

}  // namespace dtm

extern "C" int dtm_unpool2(const void* dp2, const void* argmax, const void* act2, int batch, int lrn, void* dz2_planes,
                           float* db2, void* stream) {
  using namespace dtm;
  if (batch <= 0) return dtm_fail("unpool2: batch must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (lrn) {
    if (!act2) return dtm_fail("unpool2: LRN backward needs act2");
    const int pixels = batch * IMG * IMG;
    const int grid = min((pixels + 31) / 32, 8 * device_sm_count());
    DTM_CUDA_OK(launch_pdl(unpool2_lrn_kernel, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dp2),
                           static_cast<const uint8_t*>(argmax), static_cast<const __nv_bfloat16*>(act2), batch,
                           static_cast<uint8_t*>(dz2_planes), db2));
  } else {
    const int rows = batch * 49;
    const int grid = min((rows + 31) / 32, 4 * device_sm_count());
    DTM_CUDA_OK(launch_pdl(unpool2_kernel, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dp2),
                           static_cast<const uint8_t*>(argmax), batch, static_cast<uint8_t*>(dz2_planes), db2));
  }
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_colsum_bf16(const void* x, int M, int N, float* out, void* stream) {
  using namespace dtm;
  if (M <= 0 || N <= 0) return dtm_fail("colsum: M and N must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  dim3 grid((N + 31) / 32, min(2 * device_sm_count(), (M + 7) / 8));
  DTM_CUDA_OK(launch_pdl(colsum_bf16_kernel, grid, dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), M, N, out));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
