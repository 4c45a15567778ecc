// conv2 input gradient on tcgen05 (autodiff of mnist_input.py:161): dP1 = dZ2 (*) flip(W2), fused with the backward
// of what precedes conv2 in the graph — ReLU mask of conv1's pooled output, or LRN backward + ReLU mask for the
// `reference` preset (mnist_input.py:149-153).
//
// Same zero-copy-tap scheme as the forward kernel: dZ2 arrives in the plane layout (64 channels = 8 planes,
// 32 KiB per image, one bulk copy), every tap is a re-based descriptor.  Per image 2 M-tiles x 25 taps x 4
// K-chunks = 200 tcgen05.mma of shape 128 x 32 x 16.  The B operand is W2 with both spatial axes flipped and
// (ci, co) swapped: B[k = tap'*64 + co][n = ci] = W2[24 - tap'][ci][co]; eight consecutive co are contiguous in
// the TensorFlow layout, so the re-pack into [k/8][n][8] is one 16-byte load per chunk.
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int DG_CI = 32;                                // output channels of this GEMM (conv2's input channels)
constexpr int DG_CO = 64;                                // reduction channels per tap
constexpr int DG_K = 25 * DG_CO;                         // 1600
constexpr int DG_W_BYTES = DG_K * DG_CI * 2;             // 102400
constexpr int DG_IMG_BYTES = plane_image_bytes(DG_CO);   // 32768
constexpr int DG_BUFS = 3;
constexpr int DG_ACC = 4;                                // TMEM accumulator ring: 4 x 32 columns
constexpr int DG_SMEM = DG_W_BYTES + PL_GUARD_BYTES + DG_BUFS * DG_IMG_BYTES + PL_GUARD_BYTES + 128;

// Like the forward kernel, the epilogue (TMEM read, activation fetch, mask, store) is a latency chain: without LRN two
// warps share each TMEM lane quadrant (16 of the 32 channels each).  LRN backward keeps that split: a warp reads all 32
// gradient / activation channels (its +-8-channel context) but evaluates the window terms only where its 16 outputs need
// them (round 1: four warps with 32 channels per thread, 69.9 us against 27.9 us without LRN).
template <bool LRN> struct DgEpi { static constexpr int WARPS = 8; static constexpr int CH = 16; };

// LRN backward (TF semantics) for output channels [16 HALF, 16 HALF + 16) of a 32-channel pixel:
//   s_c = 1 + alpha sum_{|d|<=4} a_{c+d}^2 ;  t_c = dy_c a_c s_c^(-beta-1) ;  da_j = dy_j s_j^-beta - 2 alpha beta a_j sum_{|d|<=4} t_{j+d}
template <int HALF>
DTM_DEVICE void dg_lrn_backward_half(const float (&dy)[32], const float (&a)[32], float (&out)[16]) {
  constexpr int J0 = HALF * 16;
  constexpr int T0 = J0 - 4 < 0 ? 0 : J0 - 4, T1 = J0 + 20 > 32 ? 32 : J0 + 20;    // channels whose t_c is needed
  constexpr int NT = T1 - T0;                                                      // 20
  const float alpha = 0.001f / 9.0f, beta = 0.75f;
  float ext[NT + 8], ss[NT], sp[NT];
#pragma unroll
  for (int k = 0; k < NT + 8; ++k) {                       // squares of channels T0 - 4 .. T1 + 3, zero outside 0..31
    const int c = T0 - 4 + k;
    ext[k] = (c >= 0 && c < 32) ? a[c < 0 ? 0 : (c > 31 ? 31 : c)] * a[c < 0 ? 0 : (c > 31 ? 31 : c)] : 0.0f;
  }
  window9_sums<NT>(ext, ss);
  float text[16 + 8];                                      // t of channels J0 - 4 .. J0 + 19, zero outside T0..T1-1
#pragma unroll
  for (int m = 0; m < 24; ++m) text[m] = 0.0f;
#pragma unroll
  for (int i = 0; i < NT; ++i) {
    const int c = T0 + i;
    float inv_s;
    const float s = 1.0f + alpha * ss[i];
    const float pw = lrn_pow_m075(s, &inv_s);
    sp[i] = pw;
    text[c - (J0 - 4)] = dy[c] * a[c] * pw * inv_s;
  }
  float acc[16];
  window9_sums<16>(text, acc);
#pragma unroll
  for (int j = 0; j < 16; ++j) out[j] = dy[J0 + j] * sp[J0 + j - T0] - 2.0f * alpha * beta * a[J0 + j] * acc[j];
}

// The non-LRN kernel declares 512 threads (it is launched with 384) to cap itself at 128 registers per thread: 384 x 128 =
// 48 K registers leave room for a CTA of the optimizer-update kernel that runs beside the conv backward - at 165 registers
// the two kernels could not share an SM and simply took turns.
template <bool LRN>
__global__ void __launch_bounds__(LRN ? 384 : 512, 1)
conv2_dgrad_kernel(const uint8_t* __restrict__ dz2_planes, const __nv_bfloat16* __restrict__ w2,
                   const uint8_t* __restrict__ w_packed, const uint8_t* __restrict__ p1_planes, const __nv_bfloat16* __restrict__ pool_raw, int batch,
                   __nv_bfloat16* __restrict__ dpool1, int early_weights) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t img_full[DG_BUFS], img_empty[DG_BUFS], acc_full[DG_ACC], acc_empty[DG_ACC], w_full;
  __shared__ uint32_t tmem_slot;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* s_w = smem;                                   // [k8 = 200][n = 32][8] bf16
  uint8_t* s_guard = smem + DG_W_BYTES;
  uint8_t* s_img = s_guard + PL_GUARD_BYTES;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();

  if (!(early_weights & 1)) pdl_wait();   // strict mode: nothing is read from global memory before the dependency wait
  if (warp == 1 && elect_one()) {
    for (int i = 0; i < DG_BUFS; ++i) { mbar_init(&img_full[i], 1); mbar_init(&img_empty[i], 1); }
    for (int i = 0; i < DG_ACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 32 * DgEpi<LRN>::WARPS); }
    mbar_init(&w_full, 1);
    fence_mbar_init();
    if (w_packed != nullptr) {   // written by the previous step's update kernel: fetch before the dependency wait
      mbar_arrive_expect_tx(&w_full, DG_W_BYTES);
      bulk_load_1d(s_w, w_packed, DG_W_BYTES, &w_full);
    }
  }
  if (warp == 2) { tmem_alloc(&tmem_slot, 128); tmem_relinquish(); }
  {
    uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x * 16; i < 2 * PL_GUARD_BYTES + DG_BUFS * DG_IMG_BYTES; i += blockDim.x * 16)
      *reinterpret_cast<uint4*>(s_guard + i) = z;
  }
  for (int chunk = threadIdx.x; w_packed == nullptr && chunk < (DG_K / 8) * DG_CI; chunk += blockDim.x) {
    const int k8 = chunk / DG_CI, ci = chunk % DG_CI;
    const int tap_f = k8 >> 3, co8 = k8 & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(w2 + ((24 - tap_f) * DG_CI + ci) * DG_CO + co8 * 8);
    *reinterpret_cast<uint4*>(s_w + chunk * 16) = v;
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // work unit = one M-tile (half an image), as in the forward kernel: 1024 units balance over 148 CTAs to 98.8 %
  const int first = blockIdx.x, stride = gridDim.x;
  const int units = 2 * batch;
  pdl_wait();                 // the unpooled gradient planes (and everything upstream) are complete from here on
  pdl_launch_dependents();

  if (warp == 0) {
    if (elect_one()) {
      int it = 0;
      for (int u = first; u < units; u += stride, ++it) {
        const int b = it % DG_BUFS;
        mbar_wait(&img_empty[b], ((it / DG_BUFS) & 1) ^ 1, 20);
        mbar_arrive_expect_tx(&img_full[b], DG_IMG_BYTES);
        bulk_load_1d(s_img + b * DG_IMG_BYTES, dz2_planes + static_cast<size_t>(u >> 1) * DG_IMG_BYTES, DG_IMG_BYTES, &img_full[b]);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, DG_CI, false, false);
      const UmmaDescBase w_desc = umma_desc_base(smem_u32(s_w), DG_CI * 16, 128, UMMA_SWIZZLE_NONE);
      if (w_packed != nullptr) mbar_wait(&w_full, 0, 24);
      int it = 0;
      for (int u = first; u < units; u += stride, ++it) {
        const int b = it % DG_BUFS;
        const int a = it % DG_ACC;
        mbar_wait(&acc_empty[a], ((it / DG_ACC) & 1) ^ 1, 21);
        mbar_wait(&img_full[b], (it / DG_BUFS) & 1, 22);
        tc_fence_after();
        const UmmaDescBase a_base = umma_desc_base(smem_u32(s_img + b * DG_IMG_BYTES), PL_PLANE_BYTES, 128, UMMA_SWIZZLE_NONE);
        {
          const int t = u & 1;
          const uint32_t d_tmem = tmem_base + a * 32;
          const int tile_off = t * 128 * 16;
#pragma unroll
          for (int tap = 0; tap < 25; ++tap) {
            const int a_off = ((tap / 5) * PL_GRID + (tap % 5) - 2) * 16;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint64_t da = umma_desc_at(a_base, tile_off + a_off + j * 2 * PL_PLANE_BYTES);
              const uint64_t db = umma_desc_at(w_desc, (tap * 8 + j * 2) * (DG_CI * 16));
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
    constexpr int CH = DgEpi<LRN>::CH;
    const uint32_t q = warp & 3;
    const int c0 = static_cast<int>((warp - 4) >> 2) * 16;
    const int w_px = lane & 15, h_sub = lane >> 4;
    int it = 0;
    for (int u = first; u < units; u += stride, ++it) {
      const int a = it % DG_ACC;
      const int img = u >> 1;
      {  // the activation that gates this pixel's gradient: start fetching it before the accumulator wait
        const int hp = (u & 1) * 8 + q * 2 + h_sub;
        if (!(early_weights & 2) && hp < IMG && w_px < IMG) {
          if (LRN) prefetch_l1(pool_raw + (static_cast<size_t>(img) * (IMG * IMG) + hp * IMG + w_px) * DG_CI);
          else     prefetch_l1(p1_planes + static_cast<size_t>(img) * plane_image_bytes(DG_CI) + plane_offset(hp, w_px, c0));
        }
      }
      mbar_wait(&acc_full[a], (it / DG_ACC) & 1, 23);
      tc_fence_after();
      do {   // single pass; `continue` below leaves it for threads whose grid pixel is padding
        const int t = u & 1;
        const int h = t * 8 + q * 2 + h_sub;
        const bool valid = h < IMG && w_px < IMG;
        float g[CH];
        float act[CH];  // the activation whose sign gates the gradient (pooled conv1 output, pre-LRN)
        if (LRN) {
          float dy[32], a32[32];
          {
            uint32_t r[32];
            tmem_ld_x32(tmem_base + ((q * 32u) << 16) + a * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) dy[j] = __uint_as_float(r[j]);
          }
          if (!valid) continue;
          const uint4* src = reinterpret_cast<const uint4*>(pool_raw + (static_cast<size_t>(img) * (IMG * IMG) + h * IMG + w_px) * DG_CI);
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {
            const uint4 u4 = src[c8];
            const uint32_t wds[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) { a32[c8 * 8 + 2 * i] = bf16_lo(wds[i]); a32[c8 * 8 + 2 * i + 1] = bf16_hi(wds[i]); }
          }
          if (c0 == 0) {
            dg_lrn_backward_half<0>(dy, a32, g);
#pragma unroll
            for (int j = 0; j < CH; ++j) act[j] = a32[j];
          } else {
            dg_lrn_backward_half<1>(dy, a32, g);
#pragma unroll
            for (int j = 0; j < CH; ++j) act[j] = a32[16 + j];
          }
        } else {
          {
            uint32_t r[16];
            tmem_ld_x16(tmem_base + ((q * 32u) << 16) + a * 32 + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < CH; ++j) g[j] = __uint_as_float(r[j]);
          }
          if (!valid) continue;
          const uint8_t* src = p1_planes + static_cast<size_t>(img) * plane_image_bytes(DG_CI);
#pragma unroll
          for (int c8 = 0; c8 < CH / 8; ++c8) {
            const uint4 u = *reinterpret_cast<const uint4*>(src + plane_offset(h, w_px, c0 + c8 * 8));
            const uint32_t wds[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) { act[c8 * 8 + 2 * i] = bf16_lo(wds[i]); act[c8 * 8 + 2 * i + 1] = bf16_hi(wds[i]); }
          }
        }
        uint4* dst = reinterpret_cast<uint4*>(dpool1 + (static_cast<size_t>(img) * (IMG * IMG) + h * IMG + w_px) * DG_CI + c0);
#pragma unroll
        for (int c8 = 0; c8 < CH / 8; ++c8) {
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) o[i] = act[c8 * 8 + i] > 0.0f ? g[c8 * 8 + i] : 0.0f;
          dst[c8] = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
        }
      } while (0);
      tc_fence_before();
      mbar_arrive(&acc_empty[a]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 128);
}

}  // namespace dtm

extern "C" int dtm_conv2_dgrad(const void* dz2_planes, const void* w2_bf16, const void* packed, const void* p1_planes,
                               const void* pool_raw, int batch, int lrn, void* dpool1, void* stream) {
  using namespace dtm;
  if (batch <= 0) return dtm_fail("conv2_dgrad: batch must be positive");
  if (lrn && !pool_raw) return dtm_fail("conv2_dgrad: LRN backward needs pool_raw");
  if (!lrn && !p1_planes) return dtm_fail("conv2_dgrad: needs the conv1 activation for the ReLU mask");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv2_dgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv2_dgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SMEM));
    configured[dev_slot] = true;
  }
  const int grid = min(2 * batch, device_sm_count());
  const char* pe = getenv("DTM_EPI_PREFETCH");          // bit 1 of the mode word: epilogue prefetch off (A/B switch)
  const int no_prefetch = (pe && pe[0] == '0') ? 2 : 0;
  auto launch = [&](auto kern) -> int {
    DTM_CUDA_OK(launch_pdl(kern, dim3(grid), dim3(384), DG_SMEM, s, static_cast<const uint8_t*>(dz2_planes),
                           static_cast<const __nv_bfloat16*>(w2_bf16), static_cast<const uint8_t*>(packed),
                           static_cast<const uint8_t*>(p1_planes), static_cast<const __nv_bfloat16*>(pool_raw), batch,
                           static_cast<__nv_bfloat16*>(dpool1), (pdl_mode() == 2 ? 1 : 0) | no_prefetch));
    return 0;
  };
  if (int rc = lrn ? launch(conv2_dgrad_kernel<true>) : launch(conv2_dgrad_kernel<false>)) return rc;
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
