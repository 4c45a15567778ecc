// conv1 forward on tcgen05 WITHOUT im2col ("shifted window" operand, round 2): same contract as conv1_tc.cu (reference
// mnist_input.py:39, :134-153: uint8 pixel -> x/255 - 0.5 -> 5x5 SAME conv (C_in 1|3 -> 32) -> +bias -> ReLU -> 2x2/2
// max-pool [-> LRN] -> plane layout), an eighth of its producer work.
//
// conv1_tc.cu builds an explicit im2col tile per (pooled-pixel group, window position): 4096 16-byte stores and 32 K
// 2-byte gathers per image, and that - not the tensor core - set its time.  Here the producers only write, per image channel
// and per column parity dx, ONE 16-byte entry per (image row y, pooled column pw):
//        F[dx][y][pw][s] = Xp[y][2 pw + dx + s],   s = 0..7          (Xp = normalised image with a 2-pixel zero halo)
// i.e. 896 stores per image and channel.  Read as a K-major operand, a tile's row m = (ph, pw) for window position
// d = (dy, dx) and K index k = kh * 8 + kw is
//        A[m][k] = F[dx][2 ph + dy + kh][pw][kw] = Xp[2 ph + dy + kh][2 pw + dx + kw]
// - the pixel that tap (kh, kw) pairs with conv pixel (2 ph + dy, 2 pw + dx) - and the descriptor alone expresses it:
// the 8 kw of a K chunk are the 8 bf16 of an entry, the kh chunks are one F row (16 entries, 256 B) apart (LBO), and
// 8-row groups (pw = 8 g .. 8 g + 7 of one pooled row ph) are two image rows apart (SBO = 512 B).  The pooled columns split
// into g = 0 (pw 0-7) and g = 1 (pw 8-13), so a tile is 14 valid 8-row groups (112 of 128 rows).  Filter rows kw = 5..7 and
// kh = 5 are zero, K = 48 per channel: three tcgen05.mma (128 x 32 x 16) per tile and channel.
//
// Tile order, TMEM layout and epilogue are conv1_tc.cu's: the four window positions of a group accumulate into four
// neighbouring 32-column blocks, an epilogue thread owns ONE pooled pixel, max-pool / arg-max are register compares, LRN
// stays in the thread.
//
// Warp roles: 0-3 epilogue (TMEM lane quadrant = warp index), 4-7 producers, 8 MMA issuer, 9 TMEM allocator.
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int S1_OUT = 32;
constexpr int S1_HW = 28;
constexpr int S1_XP_PITCH = 40;                 // bf16 halo image: 32 rows x 40 columns (column 2 pw + dx + 7 <= 35)
constexpr int S1_F_ROWS = 38;                   // F rows reachable: 2 * 15 + 1 + 5 = 36 (garbage row groups included) + slack
constexpr int S1_F_ROW_BYTES = 16 * 16;         // 16 pooled columns x 16 B
constexpr int S1_F_BYTES = S1_F_ROWS * S1_F_ROW_BYTES;          // one (channel, dx) array: 9728 B
constexpr int S1_ACC = 2;                       // TMEM accumulator ring: 2 groups x (4 positions x 32 columns)
constexpr int S1_KCH = 6;                       // K chunks per channel (kh 0..4 + one zero chunk) -> 3 MMAs of K = 16
constexpr float S1_LRN_BIAS = 1.0f, S1_LRN_ALPHA = 0.001f / 9.0f, S1_LRN_BETA = 0.75f;
static_assert(S1_LRN_BETA == 0.75f, "the LRN epilogue computes s^-0.75 with lrn_pow_m075");

template <int CIN> struct S1Geom {
  static constexpr int F_SET = 2 * CIN * S1_F_BYTES;                       // both parities, every channel: one image
  static constexpr int B_BYTES = CIN * S1_KCH * S1_OUT * 16;               // [ci][k/8][n][8] bf16
  static constexpr int XP_BYTES = CIN * 32 * S1_XP_PITCH * 2;
  static constexpr int SMEM = 2 * F_SET + B_BYTES + XP_BYTES + 256;
};

template <int CIN, bool LRN>
__global__ void __launch_bounds__(320, CIN == 1 ? 2 : 1)
conv1_sw_kernel(const void* __restrict__ x, int in_is_float, const float* __restrict__ w, const float* __restrict__ bias, int batch,
                uint8_t* __restrict__ p1_planes, uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ pool_raw) {
  using G = S1Geom<CIN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[2], empty_bar[2], acc_full[S1_ACC], acc_empty[S1_ACC];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_bias[S1_OUT];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* s_f = smem;                                                   // [buf][ci][dx][row][pw][8] bf16
  uint8_t* s_b = s_f + 2 * G::F_SET;                                     // [ci][k/8][32][8] bf16
  __nv_bfloat16* s_xp = reinterpret_cast<__nv_bfloat16*>(s_b + G::B_BYTES);   // [ci][32][40] bf16, halo = 0

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int first = blockIdx.x, stride = gridDim.x;

  // ---- one-time setup ---------------------------------------------------------------------------------------------
  if (warp == 8 && elect_one()) {
    for (int i = 0; i < 2; ++i) { mbar_init(&full_bar[i], 4); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < S1_ACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
    fence_mbar_init();
  }
  if (warp == 9) { tmem_alloc(&tmem_slot, S1_ACC * 4 * S1_OUT); tmem_relinquish(); }
  // F arrays (rows that pair with the halo stay zero for good) and the halo image
  {
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x * 16; i < 2 * G::F_SET; i += blockDim.x * 16) *reinterpret_cast<uint4*>(s_f + i) = z;
    for (int i = threadIdx.x * 16; i < G::XP_BYTES; i += blockDim.x * 16) *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(s_xp) + i) = z;
  }
  // conv1 opens the step: its predecessor in the stream is the previous step's update kernel, which writes these very
  // weights - so the dependency wait comes before the first global read
  pdl_wait();
  pdl_launch_dependents();
  if (threadIdx.x < S1_OUT) s_bias[threadIdx.x] = bias[threadIdx.x];
  // filter: TF layout [kh][kw][ci][co] fp32 -> per channel B[k = kh*8 + kw][n = co] as [k/8][n][k%8] bf16, zero for kw > 4, kh = 5
  for (int i = threadIdx.x; i < CIN * S1_KCH * S1_OUT * 8; i += blockDim.x) {
    const int kw = i & 7, n = (i >> 3) % S1_OUT, kh = (i / (8 * S1_OUT)) % S1_KCH, ci = i / (8 * S1_OUT * S1_KCH);
    const float v = (kw < 5 && kh < 5) ? w[((kh * 5 + kw) * CIN + ci) * S1_OUT + n] : 0.0f;
    reinterpret_cast<__nv_bfloat16*>(s_b)[i] = __float2bfloat16(v);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp >= 4 && warp < 8) {
    // ================================ producers (128 threads) ================================
    const int t = threadIdx.x - 128;
    int img_it = 0;
    for (int img = first; img < batch; img += stride, ++img_it) {
      const int buf = img_it & 1;
      // 1. halo image: normalise, convert (interior only - the border stays zero).  Safe to overwrite: the previous image's
      //    entries were finished before the bar.sync at the end of the last round.
      const int npx = S1_HW * S1_HW * CIN;
      if (in_is_float) {
        const float* src = static_cast<const float*>(x) + static_cast<size_t>(img) * npx;
        for (int i = t; i < npx; i += 128) {
          const int ci = i % CIN, p = i / CIN;
          s_xp[(ci * 32 + p / S1_HW + 2) * S1_XP_PITCH + p % S1_HW + 2] = __float2bfloat16(src[i]);
        }
      } else {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(x) + static_cast<size_t>(img) * npx);
        for (int wd = t; wd < npx / 4; wd += 128) {      // 784 * C_in is a multiple of 4: whole words
          const uint32_t v = src[wd];
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const int i = wd * 4 + b, ci = i % CIN, p = i / CIN;
            s_xp[(ci * 32 + p / S1_HW + 2) * S1_XP_PITCH + p % S1_HW + 2] =
                __float2bfloat16(static_cast<float>((v >> (8 * b)) & 0xFF) * (1.0f / 255.0f) - 0.5f);
          }
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");   // producers only: the halo image is complete
      mbar_wait_warp(&empty_bar[buf], ((img_it >> 1) & 1) ^ 1, 40);      // the MMAs of two images ago have released this set
      // 2. entries of rows 2..29 (rows 0, 1, 30.. pair with the zero halo and were cleared once):
      //    F[ci][dx][y][pw][s] = Xp[ci][y][2 pw + dx + s]
      uint8_t* fset = s_f + buf * G::F_SET;
      for (int e = t; e < CIN * 2 * 28 * 16; e += 128) {
        const int pw = e & 15, y = 2 + (e >> 4) % 28, dx = (e / (16 * 28)) & 1, ci = e / (16 * 28 * 2);
        const uint32_t* row = reinterpret_cast<const uint32_t*>(s_xp + (ci * 32 + y) * S1_XP_PITCH) + pw;    // column 2 pw
        uint32_t r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];
        if (dx) {
          const uint32_t r4 = row[4];
          r0 = __funnelshift_r(r0, r1, 16); r1 = __funnelshift_r(r1, r2, 16);
          r2 = __funnelshift_r(r2, r3, 16); r3 = __funnelshift_r(r3, r4, 16);
        }
        *reinterpret_cast<uint4*>(fset + (ci * 2 + dx) * S1_F_BYTES + (y * 16 + pw) * 16) = make_uint4(r0, r1, r2, r3);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_bar[buf]);
      asm volatile("bar.sync 1, 128;" ::: "memory");   // nobody still reads the halo image when the next one is staged
    }
  } else if (warp == 8) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, S1_OUT, false, false);
      int img_it = 0, group_it = 0;
      for (int img = first; img < batch; img += stride, ++img_it) {
        const int buf = img_it & 1;
        mbar_wait(&full_bar[buf], (img_it >> 1) & 1, 42);
        tc_fence_after();
        const uint32_t fset = smem_u32(s_f + buf * G::F_SET);
#pragma unroll 1
        for (int g = 0; g < 2; ++g, ++group_it) {
          const int a = group_it % S1_ACC;
          mbar_wait(&acc_empty[a], ((group_it / S1_ACC) & 1) ^ 1, 41);
          tc_fence_after();
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const int dy = d >> 1, dx = d & 1;
            const uint32_t d_tmem = tmem_base + a * (4 * S1_OUT) + d * S1_OUT;
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) {
              // K-major, no swizzle: K chunks (kh) one F row apart (LBO = 256 B), 8-row groups two image rows apart (SBO = 512 B)
              const UmmaDescBase a_desc = umma_desc_base(fset + (ci * 2 + dx) * S1_F_BYTES + (dy * 16 + g * 8) * 16, S1_F_ROW_BYTES,
                                                         2 * S1_F_ROW_BYTES, UMMA_SWIZZLE_NONE);
              const UmmaDescBase b_desc = umma_desc_base(smem_u32(s_b + ci * S1_KCH * S1_OUT * 16), S1_OUT * 16, 128, UMMA_SWIZZLE_NONE);
#pragma unroll
              for (int j = 0; j < S1_KCH / 2; ++j) {
                const uint64_t da = umma_desc_at(a_desc, j * 2 * S1_F_ROW_BYTES);
                const uint64_t db = umma_desc_at(b_desc, j * 2 * (S1_OUT * 16));
                if (ci == 0 && j == 0) umma_f16_c<false>(d_tmem, da, db, idesc);
                else                   umma_f16_c<true>(d_tmem, da, db, idesc);
              }
            }
          }
          umma_commit(&acc_full[a]);
        }
        umma_commit(&empty_bar[buf]);
      }
    }
  } else if (warp < 4) {
    // ================================ epilogue (conv1_tc.cu's, with this kernel's row -> pooled pixel map) ================================
    const uint32_t q = warp;
    int group_it = 0;
    for (int img = first; img < batch; img += stride) {
      uint8_t* out_img = p1_planes + static_cast<size_t>(img) * plane_image_bytes(S1_OUT);
#pragma unroll 1
      for (int g = 0; g < 2; ++g, ++group_it) {
        const int a = group_it % S1_ACC;
        mbar_wait(&acc_full[a], (group_it / S1_ACC) & 1, 43);
        tc_fence_after();
        const int m = q * 32 + lane;                       // tile row: pooled row m / 8, pooled column 8 g + m % 8
        const int ph = m >> 3, pw = g * 8 + (m & 7);
        const uint32_t t_addr = tmem_base + ((q * 32u) << 16) + a * (4 * S1_OUT);
        float best[S1_OUT];
        uint32_t arg[S1_OUT / 4];
#pragma unroll
        for (int j = 0; j < S1_OUT / 4; ++j) arg[j] = 0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r0[16], r1[16], r2[16], r3[16];         // the four window positions, 16 channels
          tmem_ld_x16(t_addr + 0 * S1_OUT + half * 16, r0);
          tmem_ld_x16(t_addr + 1 * S1_OUT + half * 16, r1);
          tmem_ld_x16(t_addr + 2 * S1_OUT + half * 16, r2);
          tmem_ld_x16(t_addr + 3 * S1_OUT + half * 16, r3);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int c = half * 16 + j;
            float mx = __uint_as_float(r0[j]);
            uint32_t ax = 0;                                // first maximum in position order wins
            const float v1 = __uint_as_float(r1[j]), v2 = __uint_as_float(r2[j]), v3 = __uint_as_float(r3[j]);
            if (v1 > mx) { mx = v1; ax = 1; }
            if (v2 > mx) { mx = v2; ax = 2; }
            if (v3 > mx) { mx = v3; ax = 3; }
            best[c] = fmaxf(mx + s_bias[c], 0.0f);          // bias and ReLU commute with the max
            arg[c >> 2] |= ax << (8 * (c & 3));
          }
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[a]);                         // accumulators are in registers: hand the TMEM group back
        if (ph >= IMG || pw >= IMG) continue;
        const int pp = ph * IMG + pw;
        uint4* am = reinterpret_cast<uint4*>(argmax + (static_cast<size_t>(img) * (IMG * IMG) + pp) * S1_OUT);
        am[0] = make_uint4(arg[0], arg[1], arg[2], arg[3]);
        am[1] = make_uint4(arg[4], arg[5], arg[6], arg[7]);
        if (LRN) {
          uint4* raw = reinterpret_cast<uint4*>(pool_raw + (static_cast<size_t>(img) * (IMG * IMG) + pp) * S1_OUT);
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8)
            raw[c8] = make_uint4(pack_bf16x2(best[c8 * 8], best[c8 * 8 + 1]), pack_bf16x2(best[c8 * 8 + 2], best[c8 * 8 + 3]),
                                 pack_bf16x2(best[c8 * 8 + 4], best[c8 * 8 + 5]), pack_bf16x2(best[c8 * 8 + 6], best[c8 * 8 + 7]));
          float act[S1_OUT], yv[S1_OUT];
#pragma unroll
          for (int c = 0; c < S1_OUT; ++c) act[c] = __bfloat162float(__float2bfloat16(best[c]));   // what the backward pass sees
          float ext[S1_OUT + 8], win[S1_OUT];
#pragma unroll
          for (int k = 0; k < 4; ++k) ext[k] = ext[S1_OUT + 4 + k] = 0.0f;
#pragma unroll
          for (int c = 0; c < S1_OUT; ++c) ext[4 + c] = act[c] * act[c];
          window9_sums<S1_OUT>(ext, win);
#pragma unroll
          for (int c = 0; c < S1_OUT; ++c) yv[c] = act[c] * lrn_pow_m075(S1_LRN_BIAS + S1_LRN_ALPHA * win[c]);
#pragma unroll
          for (int c = 0; c < S1_OUT; ++c) best[c] = yv[c];
        }
#pragma unroll
        for (int c8 = 0; c8 < 4; ++c8)
          *reinterpret_cast<uint4*>(out_img + plane_offset(ph, pw, c8 * 8)) =
              make_uint4(pack_bf16x2(best[c8 * 8], best[c8 * 8 + 1]), pack_bf16x2(best[c8 * 8 + 2], best[c8 * 8 + 3]),
                         pack_bf16x2(best[c8 * 8 + 4], best[c8 * 8 + 5]), pack_bf16x2(best[c8 * 8 + 6], best[c8 * 8 + 7]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, S1_ACC * 4 * S1_OUT);
}

}  // namespace dtm

// Called by dtm_conv1_fwd (conv1_tc.cu) unless DTM_CONV1_IM2COL=1 asks for the round-1 kernel.
int dtm_conv1_fwd_shifted_window(const void* x, int in_is_float, int channels, const float* w, const float* bias, int batch, int lrn,
                                 void* p1_planes, void* argmax, void* pool_raw, void* stream) {
  using namespace dtm;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* p = static_cast<uint8_t*>(p1_planes);
  uint8_t* a = static_cast<uint8_t*>(argmax);
  __nv_bfloat16* r = static_cast<__nv_bfloat16*>(pool_raw);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_sw_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, S1Geom<1>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_sw_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, S1Geom<1>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_sw_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, S1Geom<3>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_sw_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, S1Geom<3>::SMEM));
    configured[dev_slot] = true;
  }
  const int per_sm = channels == 1 ? 2 : 1;
  const int grid = min(batch, per_sm * device_sm_count());
  if (channels == 1) {
    if (lrn) DTM_CUDA_OK(launch_pdl(conv1_sw_kernel<1, true>, dim3(grid), dim3(320), S1Geom<1>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
    else     DTM_CUDA_OK(launch_pdl(conv1_sw_kernel<1, false>, dim3(grid), dim3(320), S1Geom<1>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
  } else {
    if (lrn) DTM_CUDA_OK(launch_pdl(conv1_sw_kernel<3, true>, dim3(grid), dim3(320), S1Geom<3>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
    else     DTM_CUDA_OK(launch_pdl(conv1_sw_kernel<3, false>, dim3(grid), dim3(320), S1Geom<3>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
  }
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
