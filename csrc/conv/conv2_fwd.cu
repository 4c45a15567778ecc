// conv2 forward on tcgen05: 5x5 SAME conv (32 -> 64) + bias + ReLU [+ LRN] + 2x2 max-pool, one persistent kernel.
// Reference graph: mnist_input.py:156-172 (65 % of the model's forward FLOPs).
//
// Implicit GEMM without im2col.  The pooled conv1 activation arrives in the plane layout (plane_layout.cuh):
// one cp.async.bulk drops a whole 16 KiB image into shared memory, and each of the 25 filter taps is the *same*
// bytes seen through a UMMA descriptor whose start address is moved by (kh*16 + kw - 2) pixels.  Per image:
//   2 M-tiles (image rows 0-7 / 8-15 of the 16-wide grid) x 25 taps x 2 K-chunks = 100 tcgen05.mma (128x64x16).
// A work unit is ONE M-tile (finer load balance than whole images); its 64-column accumulator is one of a ring of four
// in TMEM, so the MMA issuer runs up to three units ahead of the epilogue.  The 100 KiB of
// filter weights are re-packed once per CTA into the canonical K-major operand layout and stay resident.
//
// Warp roles: 0 = bulk-copy producer, 1 = MMA issuer, 2 = TMEM allocator, 4-7 = epilogue (TMEM -> registers ->
// bias/ReLU/LRN -> pooling by warp shuffles: the four pixels of a pooling window sit in lanes l, l^1, l^16, l^17).
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int C2_IN = 32;
constexpr int C2_OUT = 64;
constexpr int C2_K = 25 * C2_IN;                       // 800
constexpr int C2_W_BYTES = C2_K * C2_OUT * 2;          // 102400
constexpr int C2_IMG_BYTES = plane_image_bytes(C2_IN); // 16384
constexpr int C2_BUFS = 3;
constexpr int C2_ACC = 4;                               // TMEM accumulator ring: 4 x 64 columns
constexpr int C2_SMEM = C2_W_BYTES + PL_GUARD_BYTES + C2_BUFS * C2_IMG_BYTES + PL_GUARD_BYTES + 128;

constexpr float LRN2_BIAS = 1.0f;
constexpr float LRN2_ALPHA = 0.001f / 9.0f;
constexpr float LRN2_BETA = 0.75f;
static_assert(LRN2_BETA == 0.75f, "the LRN epilogue computes s^-0.75 with lrn_pow_m075");

// Epilogue width: the pooling/arg-max code is a long dependent chain (3 shuffles + compares per channel), and with one
// epilogue warp per scheduler it, not the tensor core, set the kernel time (ncu: tensor pipe 24 % active).  Without
// LRN the 64 channels split cleanly, so two warps share each TMEM lane quadrant (32 channels each, 8 epilogue warps);
// With LRN a warp still owns 32 output channels: it also reads the four channels beyond its half from TMEM (the +-4-channel
// window), so the 8-warp layout holds there too (round 1 kept LRN on four warps with 64 channels per thread: 54.6 us against
// 21.3 us without LRN).
template <bool LRN> struct C2Epi { static constexpr int WARPS = 8; static constexpr int CH = 32; };

template <bool LRN>
__global__ void __launch_bounds__(128 + 32 * C2Epi<LRN>::WARPS, 1)
conv2_fwd_kernel(const uint8_t* __restrict__ p1_planes, const __nv_bfloat16* __restrict__ w2,
                 const uint8_t* __restrict__ w_packed, const float* __restrict__ bias, int batch, __nv_bfloat16* __restrict__ p2, uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ act2,
                 int early_weights) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t img_full[C2_BUFS], img_empty[C2_BUFS], acc_full[C2_ACC], acc_empty[C2_ACC], w_full;
  __shared__ uint32_t tmem_slot;
  __shared__ float s_bias[C2_OUT];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* s_w = smem;                                   // [k8 = 100][n = 64][8] bf16
  uint8_t* s_guard = smem + C2_W_BYTES;                  // zero guard, image buffers, zero guard
  uint8_t* s_img = s_guard + PL_GUARD_BYTES;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();

  // ---- one-time setup -------------------------------------------------------------------------
  if (!early_weights) pdl_wait();   // strict mode: nothing is read from global memory before the dependency wait
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < C2_BUFS; ++i) { mbar_init(&img_full[i], 1); mbar_init(&img_empty[i], 1); }
    for (int i = 0; i < C2_ACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 32 * C2Epi<LRN>::WARPS); }
    mbar_init(&w_full, 1);
    fence_mbar_init();
    // early-weights mode: the packed filter image was written by the previous step's update kernel, so this 100 KiB
    // copy (and the bias / self-pack reads below) may overlap the tail of the preceding kernel
    if (w_packed != nullptr) {
      mbar_arrive_expect_tx(&w_full, C2_W_BYTES);
      bulk_load_1d(s_w, w_packed, C2_W_BYTES, &w_full);
    }
  }
  if (warp == 2) { tmem_alloc(&tmem_slot, 256); tmem_relinquish(); }
  if (threadIdx.x < C2_OUT) s_bias[threadIdx.x] = bias[threadIdx.x];
  {  // zero guards + image buffers (bottom halo of the last plane is the next buffer's first rows)
    uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x * 16; i < 2 * PL_GUARD_BYTES + C2_BUFS * C2_IMG_BYTES; i += blockDim.x * 16)
      *reinterpret_cast<uint4*>(s_guard + i) = z;
  }
  // weights: TF layout [k = tap*32 + ci][n = co] (n contiguous) -> [k/8][n][k%8]; skipped when the caller supplies
  // the packed image (dtm_conv2_pack_weights), which the producer then fetches with one 100 KiB bulk copy
  for (int chunk = threadIdx.x; w_packed == nullptr && chunk < (C2_K / 8) * C2_OUT; chunk += blockDim.x) {
    const int k8 = chunk / C2_OUT, n = chunk % C2_OUT;
    uint32_t v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint16_t lo = *reinterpret_cast<const uint16_t*>(w2 + (k8 * 8 + 2 * i) * C2_OUT + n);
      const uint16_t hi = *reinterpret_cast<const uint16_t*>(w2 + (k8 * 8 + 2 * i + 1) * C2_OUT + n);
      v[i] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
    }
    *reinterpret_cast<uint4*>(s_w + chunk * 16) = make_uint4(v[0], v[1], v[2], v[3]);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  pdl_wait();                 // conv1's planes are complete and visible from here on
  pdl_launch_dependents();

  // work unit = one M-tile (half an image: grid rows 0-7 or 8-15).  1024 units over 148 CTAs balance to 98.8 %; whole
  // images (3.46 per CTA) left the last of four rounds two-thirds empty.
  const int first = blockIdx.x;
  const int stride = gridDim.x;
  const int units = 2 * batch;

  if (warp == 0) {
    // ---- producer: one 16 KiB bulk copy per image ------------------------------------------------
    if (elect_one()) {
      int it = 0;
      for (int u = first; u < units; u += stride, ++it) {
        const int b = it % C2_BUFS;
        mbar_wait(&img_empty[b], ((it / C2_BUFS) & 1) ^ 1, 10);
        mbar_arrive_expect_tx(&img_full[b], C2_IMG_BYTES);
        bulk_load_1d(s_img + b * C2_IMG_BYTES, p1_planes + static_cast<size_t>(u >> 1) * C2_IMG_BYTES, C2_IMG_BYTES, &img_full[b]);
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer ----------------------------------------------------------------------------
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, C2_OUT, false, false);
      const UmmaDescBase w_desc = umma_desc_base(smem_u32(s_w), C2_OUT * 16, 128, UMMA_SWIZZLE_NONE);
      if (w_packed != nullptr) mbar_wait(&w_full, 0, 14);
      int it = 0;
      for (int u = first; u < units; u += stride, ++it) {
        const int b = it % C2_BUFS;
        const int a = it % C2_ACC;
        mbar_wait(&acc_empty[a], ((it / C2_ACC) & 1) ^ 1, 11);
        mbar_wait(&img_full[b], (it / C2_BUFS) & 1, 12);
        tc_fence_after();
        const UmmaDescBase a_base = umma_desc_base(smem_u32(s_img + b * C2_IMG_BYTES), PL_PLANE_BYTES, 128, UMMA_SWIZZLE_NONE);
        {
          const int t = u & 1;
          const uint32_t d_tmem = tmem_base + a * 64;
          const int tile_off = t * 128 * 16;
#pragma unroll
          for (int tap = 0; tap < 25; ++tap) {
            const int a_off = ((tap / 5) * PL_GRID + (tap % 5) - 2) * 16;   // compile-time after unrolling
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint64_t da = umma_desc_at(a_base, tile_off + a_off + j * 2 * PL_PLANE_BYTES);
              const uint64_t db = umma_desc_at(w_desc, (tap * 4 + j * 2) * (C2_OUT * 16));
              if (tap == 0 && j == 0) umma_f16_c<false>(d_tmem, da, db, idesc);
              else                    umma_f16_c<true>(d_tmem, da, db, idesc);
            }
          }
        }
        umma_commit(&img_empty[b]);
        umma_commit(&acc_full[a]);
      }
    }
  } else if (warp >= 4) {
    // ---- epilogue ------------------------------------------------------------------------------
    constexpr int CH = C2Epi<LRN>::CH;     // channels handled by this warp
    const uint32_t q = warp & 3;           // TMEM lane quadrant
    const int c0 = static_cast<int>((warp - 4) >> 2) * 32;   // first channel of this warp
    const int w_px = lane & 15;            // column inside the 16-wide grid
    const int h_sub = lane >> 4;           // 0/1: row inside the warp's row pair
    int it = 0;
    for (int u = first; u < units; u += stride, ++it) {
      const int a = it % C2_ACC;
      const int img = u >> 1;
      mbar_wait(&acc_full[a], (it / C2_ACC) & 1, 13);
      tc_fence_after();
      {
        const int t = u & 1;
        const int h = t * 8 + q * 2 + h_sub;         // image row of this thread's pixel
        const int ph = t * 4 + q, pw = w_px >> 1;    // pooled coordinates of its window
        const bool writer = (lane < 16) && ((lane & 1) == 0) && pw < 7 && ph < 7;
        const bool pixel_valid = h < IMG && w_px < IMG;
        float v[CH];
#pragma unroll
        for (int cb = 0; cb < CH; cb += 32) {
          uint32_t r[32];
          tmem_ld_x32(tmem_base + ((q * 32u) << 16) + a * 64 + c0 + cb, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[cb + j] = fmaxf(__uint_as_float(r[j]) + s_bias[c0 + cb + j], 0.0f);
        }
        if (LRN) {
          // act2 keeps the post-ReLU activation (bf16) for the backward pass; LRN runs on the rounded values
#pragma unroll
          for (int j = 0; j < CH; ++j) v[j] = __bfloat162float(__float2bfloat16(v[j]));
          if (pixel_valid) {
            uint4* dst = reinterpret_cast<uint4*>(act2 + (static_cast<size_t>(img) * (IMG * IMG) + h * IMG + w_px) * C2_OUT + c0);
#pragma unroll
            for (int j = 0; j < CH; j += 8)
              dst[j >> 3] = make_uint4(pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                                       pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
          }
          // squares of this warp's 32 channels with the 4 channels on either side (zero outside 0..63): the lower half reads
          // channels 32..35 from the accumulator as well, the upper half 28..31
          float ext[CH + 8];
          {
            uint32_t hr[16];
            tmem_ld_x16(tmem_base + ((q * 32u) << 16) + a * 64 + (c0 == 0 ? 32 : 16), hr);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float lo = fmaxf(__uint_as_float(hr[k]) + s_bias[32 + k], 0.0f);         // channels 32..35
              const float hi = fmaxf(__uint_as_float(hr[12 + k]) + s_bias[28 + k], 0.0f);    // channels 28..31
              const float lo_r = __bfloat162float(__float2bfloat16(lo)), hi_r = __bfloat162float(__float2bfloat16(hi));
              ext[k] = c0 == 0 ? 0.0f : hi_r * hi_r;
              ext[CH + 4 + k] = c0 == 0 ? lo_r * lo_r : 0.0f;
            }
          }
#pragma unroll
          for (int j = 0; j < CH; ++j) ext[4 + j] = v[j] * v[j];
          float win[CH];
          window9_sums<CH>(ext, win);
#pragma unroll
          for (int j = 0; j < CH; ++j) v[j] *= lrn_pow_m075(LRN2_BIAS + LRN2_ALPHA * win[j]);
        }
        // 2x2 max-pool across lanes l, l^1 (w+1), l^16 (h+1), l^17
        __nv_bfloat16* p2_dst = p2 + (static_cast<size_t>(img) * 49 + ph * 7 + pw) * C2_OUT + c0;
        uint8_t* am_dst = argmax + (static_cast<size_t>(img) * 49 + ph * 7 + pw) * C2_OUT + c0;
#pragma unroll
        for (int j0 = 0; j0 < CH; j0 += 8) {
          float best[8];
          uint32_t arg_lo = 0, arg_hi = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float v00 = v[j0 + j];
            const float v01 = __shfl_xor_sync(0xffffffffu, v00, 1);
            const float v10 = __shfl_xor_sync(0xffffffffu, v00, 16);
            const float v11 = __shfl_xor_sync(0xffffffffu, v00, 17);
            float m = v00;
            uint32_t arg = 0;
            if (v01 > m) { m = v01; arg = 1; }
            if (v10 > m) { m = v10; arg = 2; }
            if (v11 > m) { m = v11; arg = 3; }
            best[j] = m;
            if (j < 4) arg_lo |= arg << (8 * j);
            else       arg_hi |= arg << (8 * (j - 4));
          }
          if (writer) {
            *reinterpret_cast<uint4*>(p2_dst + j0) = make_uint4(pack_bf16x2(best[0], best[1]), pack_bf16x2(best[2], best[3]),
                                                                pack_bf16x2(best[4], best[5]), pack_bf16x2(best[6], best[7]));
            *reinterpret_cast<uint2*>(am_dst + j0) = make_uint2(arg_lo, arg_hi);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[a]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 256);
}

}  // namespace dtm

extern "C" int dtm_conv2_fwd(const void* p1_planes, const void* w2_bf16, const void* packed, const float* bias, int batch, int lrn,
                             void* p2, void* argmax, void* act2, void* stream) {
  using namespace dtm;
  if (batch <= 0) return dtm_fail("conv2_fwd: batch must be positive");
  if (lrn && !act2) return dtm_fail("conv2_fwd: LRN needs the act2 buffer");
  if (!w2_bf16 && !packed) return dtm_fail("conv2_fwd: needs weights (w2_bf16 or packed)");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv2_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, C2_SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv2_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, C2_SMEM));
    configured[dev_slot] = true;
  }
  const int grid = min(2 * batch, device_sm_count());
  auto args = [&](auto kern) -> int {
    DTM_CUDA_OK(launch_pdl(kern, dim3(grid), dim3(384), C2_SMEM, s, static_cast<const uint8_t*>(p1_planes), static_cast<const __nv_bfloat16*>(w2_bf16),
                                    static_cast<const uint8_t*>(packed), bias, batch, static_cast<__nv_bfloat16*>(p2), static_cast<uint8_t*>(argmax),
                                    static_cast<__nv_bfloat16*>(act2), pdl_mode() == 2 ? 1 : 0));
    return 0;
  };
  if (int rc = lrn ? args(conv2_fwd_kernel<true>) : args(conv2_fwd_kernel<false>)) return rc;
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
