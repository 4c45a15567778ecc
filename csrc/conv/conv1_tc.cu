// conv1 forward on tcgen05 (reference mnist_input.py:39, :134-153): uint8 pixel -> x/255 - 0.5 -> 5x5 SAME conv
// (C_in 1|3 -> 32) -> +bias -> ReLU -> 2x2/2 max-pool [-> LRN] -> plane layout, as an implicit GEMM.
//
// K = 25*C_in (25 or 75) looks tensor-core hostile, but the CUDA-core version needed ~630 k FMAs + 230 k shared loads
// per image and was issue bound (27 us for a batch of 512, ncu: profiles/).  Here producer warps build an explicit
// im2col tile in shared memory — 16-byte chunks of 8 taps, which is exactly one row of a UMMA core matrix — and the
// tensor core does the arithmetic:
//     A  [128 conv pixels x K_pad]  K-major, no swizzle  ([k/8][row][8], LBO = 2048 B, SBO = 128 B), bf16
//     B  [32 channels x K_pad]      K-major, no swizzle  ([k/8][n][8],   LBO =  512 B, SBO = 128 B), bf16, resident
//     D  [128 x 32] fp32 in TMEM per tile, 8 tiles per image (2 groups of 128 pooled pixels x 4 window positions), the
//        accumulators of a group double buffered against the epilogue.
// Because the im2col tile is built by hand its row order is free, and it is chosen for the epilogue: tile (g, d) holds
// position d = dy*2 + dx of the pooling windows of pooled pixels g*128 .. g*128+127, and the four tiles of a group
// accumulate into four neighbouring 32-column TMEM blocks.  An epilogue thread therefore owns ONE pooled pixel and reads
// its four candidates from four column blocks of its own TMEM lane: max-pool and arg-max are plain register compares (no
// shuffles, no idle lanes), bias and ReLU are applied once after the max (both commute with it), and all 32 channels of
// the pooled pixel sit in one thread, so LRN (window +-4 channels) stays in registers too.
//
// Warp roles: 0-3 epilogue (TMEM lane quadrant = warp index), 4-7 im2col producers, 8 MMA issuer, 9 TMEM allocator.
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int T1_OUT = 32;
constexpr int T1_HW = 28;
constexpr int T1_PITCH = 32;            // halo image: 32 x 32 per input channel (2-pixel zero border)
constexpr int T1_TILE_M = 128;
constexpr int T1_GROUPS = 2;            // ceil(196 pooled pixels / 128)
constexpr int T1_TILES = 4 * T1_GROUPS;   // (group, window position)
constexpr int T1_SLOTS = 4;             // A-tile ring
constexpr int T1_ACC = 2;               // TMEM accumulator ring: 2 groups x (4 positions x 32 columns)

constexpr float T1_LRN_BIAS = 1.0f, T1_LRN_ALPHA = 0.001f / 9.0f, T1_LRN_BETA = 0.75f;
static_assert(T1_LRN_BETA == 0.75f, "the LRN epilogue computes s^-0.75 with lrn_pow_m075");

template <int CIN> struct T1Geom {
  static constexpr int K = 25 * CIN;
  static constexpr int K16 = (K + 15) / 16;          // MMA K steps
  static constexpr int K8 = 2 * K16;                 // 16-byte chunks per row
  static constexpr int A_TILE_BYTES = K8 * T1_TILE_M * 16;
  static constexpr int B_BYTES = K8 * T1_OUT * 16;
  static constexpr int IMG_BYTES = CIN * T1_PITCH * T1_PITCH * 2;
  static constexpr int SMEM = T1_SLOTS * A_TILE_BYTES + B_BYTES + 2 * IMG_BYTES + 256;
};

template <int CIN, bool LRN>
__global__ void __launch_bounds__(320, 2)
conv1_tc_kernel(const void* __restrict__ x, int in_is_float, const float* __restrict__ w, const float* __restrict__ bias, int batch,
                uint8_t* __restrict__ p1_planes, uint8_t* __restrict__ argmax, __nv_bfloat16* __restrict__ pool_raw) {
  using G = T1Geom<CIN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[T1_SLOTS], empty_bar[T1_SLOTS], acc_full[T1_ACC], acc_empty[T1_ACC];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_bias[T1_OUT];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* s_a = smem;                                                   // [slot][k8][128 rows][8] bf16
  uint8_t* s_b = s_a + T1_SLOTS * G::A_TILE_BYTES;                       // [k8][32][8] bf16
  __nv_bfloat16* s_img = reinterpret_cast<__nv_bfloat16*>(s_b + G::B_BYTES);   // [2][CIN][32][32] bf16, halo = 0

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int first = blockIdx.x, stride = gridDim.x;

  // ---- one-time setup ---------------------------------------------------------------------------------------------
  if (warp == 8 && elect_one()) {
    for (int i = 0; i < T1_SLOTS; ++i) { mbar_init(&full_bar[i], 128); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < T1_ACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
    fence_mbar_init();
  }
  if (warp == 9) { tmem_alloc(&tmem_slot, T1_ACC * 4 * T1_OUT); tmem_relinquish(); }
  for (int i = threadIdx.x; i < 2 * CIN * T1_PITCH * T1_PITCH; i += blockDim.x) s_img[i] = __float2bfloat16(0.0f);
  // conv1 opens the step: its predecessor in the stream is the previous step's update kernel, which writes these very
  // weights - so the dependency wait comes before the first global read
  pdl_wait();
  pdl_launch_dependents();
  if (threadIdx.x < T1_OUT) s_bias[threadIdx.x] = bias[threadIdx.x];
  // filter: TF layout [tap][ci][co] fp32 -> B[k = tap*CIN + ci][n = co] as [k/8][n][k%8] bf16, zero for k >= K
  for (int i = threadIdx.x; i < G::K8 * T1_OUT * 8; i += blockDim.x) {
    const int kk = i & 7, n = (i >> 3) % T1_OUT, k8 = i / (8 * T1_OUT);
    const int k = k8 * 8 + kk;
    const float v = k < G::K ? w[k * T1_OUT + n] : 0.0f;
    reinterpret_cast<__nv_bfloat16*>(s_b)[i] = __float2bfloat16(v);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp >= 4 && warp < 8) {
    // ================================ im2col producers (128 threads, one GEMM row each) ================================
    const int t = threadIdx.x - 128;
    int tile_it = 0, img_it = 0;
    for (int img = first; img < batch; img += stride, ++img_it) {
      __nv_bfloat16* im = s_img + (img_it & 1) * (CIN * T1_PITCH * T1_PITCH);
      // stage the image: normalise, convert, place inside the halo (interior only - the border stays zero)
      const int npx = T1_HW * T1_HW * CIN;
      if (in_is_float) {
        const float* src = static_cast<const float*>(x) + static_cast<size_t>(img) * npx;
        for (int i = t; i < npx; i += 128) {
          const int ci = i % CIN, p = i / CIN;
          im[(ci * T1_PITCH + p / T1_HW + 2) * T1_PITCH + p % T1_HW + 2] = __float2bfloat16(src[i]);
        }
      } else {
        const uint8_t* src = static_cast<const uint8_t*>(x) + static_cast<size_t>(img) * npx;
        for (int i = t; i < npx; i += 128) {
          const int ci = i % CIN, p = i / CIN;
          im[(ci * T1_PITCH + p / T1_HW + 2) * T1_PITCH + p % T1_HW + 2] =
              __float2bfloat16(static_cast<float>(src[i]) * (1.0f / 255.0f) - 0.5f);
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");   // producers only: the image is complete
      const uint16_t* imu = reinterpret_cast<const uint16_t*>(im);
#pragma unroll 1
      for (int tile = 0; tile < T1_TILES; ++tile, ++tile_it) {
        const int slot = tile_it % T1_SLOTS;
        mbar_wait(&empty_bar[slot], ((tile_it / T1_SLOTS) & 1) ^ 1, 40);
        const int pp = (tile >> 2) * T1_TILE_M + t, d = tile & 3;   // this row: pooled pixel pp, window position d
        const bool valid = pp < IMG * IMG;
        const int y = 2 * (pp / IMG) + (d >> 1), xx = 2 * (pp % IMG) + (d & 1);   // conv output pixel; halo coords == (y, xx)
        uint8_t* dst = s_a + slot * G::A_TILE_BYTES + t * 16;
#pragma unroll
        for (int k8 = 0; k8 < G::K8; ++k8) {
          uint32_t pk[4] = {0, 0, 0, 0};
          if (valid) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              const int k = k8 * 8 + kk;
              if (k < G::K) {
                const int tap = k / CIN, ci = k % CIN;
                const uint32_t v = imu[(ci * T1_PITCH + y + tap / 5) * T1_PITCH + xx + tap % 5];
                pk[kk >> 1] |= v << (16 * (kk & 1));
              }
            }
          }
          *reinterpret_cast<uint4*>(dst + k8 * (T1_TILE_M * 16)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[slot]);
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");   // nobody still reads this image buffer when it is re-staged (2 images later)
    }
  } else if (warp == 8) {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(T1_TILE_M, T1_OUT, false, false);
      const UmmaDescBase b_desc = umma_desc_base(smem_u32(s_b), T1_OUT * 16, 128, UMMA_SWIZZLE_NONE);
      int tile_it = 0, group_it = 0;
      for (int img = first; img < batch; img += stride) {
#pragma unroll 1
        for (int g = 0; g < T1_GROUPS; ++g, ++group_it) {
          const int a = group_it % T1_ACC;
          mbar_wait(&acc_empty[a], ((group_it / T1_ACC) & 1) ^ 1, 41);
#pragma unroll 1
          for (int d = 0; d < 4; ++d, ++tile_it) {
            const int slot = tile_it % T1_SLOTS;
            mbar_wait(&full_bar[slot], (tile_it / T1_SLOTS) & 1, 42);
            tc_fence_after();
            const UmmaDescBase a_desc = umma_desc_base(smem_u32(s_a + slot * G::A_TILE_BYTES), T1_TILE_M * 16, 128, UMMA_SWIZZLE_NONE);
            const uint32_t d_tmem = tmem_base + a * (4 * T1_OUT) + d * T1_OUT;
#pragma unroll
            for (int j = 0; j < G::K16; ++j) {
              const uint64_t da = umma_desc_at(a_desc, j * 2 * (T1_TILE_M * 16));
              const uint64_t db = umma_desc_at(b_desc, j * 2 * (T1_OUT * 16));
              if (j == 0) umma_f16_c<false>(d_tmem, da, db, idesc);
              else        umma_f16_c<true>(d_tmem, da, db, idesc);
            }
            umma_commit(&empty_bar[slot]);
          }
          umma_commit(&acc_full[a]);
        }
      }
    }
  } else if (warp < 4) {
    // ================================ epilogue ================================
    const uint32_t q = warp;
    int group_it = 0;
    for (int img = first; img < batch; img += stride) {
      uint8_t* out_img = p1_planes + static_cast<size_t>(img) * plane_image_bytes(T1_OUT);
#pragma unroll 1
      for (int g = 0; g < T1_GROUPS; ++g, ++group_it) {
        const int a = group_it % T1_ACC;
        mbar_wait(&acc_full[a], (group_it / T1_ACC) & 1, 43);
        tc_fence_after();
        const int pp = g * T1_TILE_M + q * 32 + lane;     // the pooled pixel this thread owns
        const uint32_t t_addr = tmem_base + ((q * 32u) << 16) + a * (4 * T1_OUT);
        float best[T1_OUT];
        uint32_t arg[T1_OUT / 4];
#pragma unroll
        for (int j = 0; j < T1_OUT / 4; ++j) arg[j] = 0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r0[16], r1[16], r2[16], r3[16];         // the four window positions, 16 channels
          tmem_ld_x16(t_addr + 0 * T1_OUT + half * 16, r0);
          tmem_ld_x16(t_addr + 1 * T1_OUT + half * 16, r1);
          tmem_ld_x16(t_addr + 2 * T1_OUT + half * 16, r2);
          tmem_ld_x16(t_addr + 3 * T1_OUT + half * 16, r3);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int c = half * 16 + j;
            float m = __uint_as_float(r0[j]);
            uint32_t ax = 0;                                // first maximum in position order wins
            const float v1 = __uint_as_float(r1[j]), v2 = __uint_as_float(r2[j]), v3 = __uint_as_float(r3[j]);
            if (v1 > m) { m = v1; ax = 1; }
            if (v2 > m) { m = v2; ax = 2; }
            if (v3 > m) { m = v3; ax = 3; }
            best[c] = fmaxf(m + s_bias[c], 0.0f);           // bias and ReLU commute with the max
            arg[c >> 2] |= ax << (8 * (c & 3));
          }
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[a]);                         // accumulators are in registers: hand the TMEM group back
        if (pp >= IMG * IMG) continue;
        const int ph = pp / IMG, pw = pp % IMG;
        uint4* am = reinterpret_cast<uint4*>(argmax + (static_cast<size_t>(img) * (IMG * IMG) + pp) * T1_OUT);
        am[0] = make_uint4(arg[0], arg[1], arg[2], arg[3]);
        am[1] = make_uint4(arg[4], arg[5], arg[6], arg[7]);
        if (LRN) {
          uint4* raw = reinterpret_cast<uint4*>(pool_raw + (static_cast<size_t>(img) * (IMG * IMG) + pp) * T1_OUT);
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8)
            raw[c8] = make_uint4(pack_bf16x2(best[c8 * 8], best[c8 * 8 + 1]), pack_bf16x2(best[c8 * 8 + 2], best[c8 * 8 + 3]),
                                 pack_bf16x2(best[c8 * 8 + 4], best[c8 * 8 + 5]), pack_bf16x2(best[c8 * 8 + 6], best[c8 * 8 + 7]));
          float act[T1_OUT], y[T1_OUT];
#pragma unroll
          for (int c = 0; c < T1_OUT; ++c) act[c] = __bfloat162float(__float2bfloat16(best[c]));   // what the backward pass sees
          float ext[T1_OUT + 8], win[T1_OUT];
#pragma unroll
          for (int k = 0; k < 4; ++k) ext[k] = ext[T1_OUT + 4 + k] = 0.0f;
#pragma unroll
          for (int c = 0; c < T1_OUT; ++c) ext[4 + c] = act[c] * act[c];
          window9_sums<T1_OUT>(ext, win);
#pragma unroll
          for (int c = 0; c < T1_OUT; ++c) y[c] = act[c] * lrn_pow_m075(T1_LRN_BIAS + T1_LRN_ALPHA * win[c]);
#pragma unroll
          for (int c = 0; c < T1_OUT; ++c) best[c] = y[c];
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
  if (warp == 9) tmem_dealloc(tmem_base, T1_ACC * 4 * T1_OUT);
}

}  // namespace dtm

// conv1_sw.cu: the same contract without im2col (default since round 2)
int dtm_conv1_fwd_shifted_window(const void* x, int in_is_float, int channels, const float* w, const float* bias, int batch, int lrn,
                                 void* p1_planes, void* argmax, void* pool_raw, void* stream);

extern "C" int dtm_conv1_fwd(const void* x, int in_is_float, int channels, const float* w, const float* bias, int batch,
                             int lrn, void* p1_planes, void* argmax, void* pool_raw, void* stream) {
  using namespace dtm;
  if (channels != 1 && channels != 3) return dtm_fail("conv1_fwd: channels must be 1 or 3");
  if (batch <= 0) return dtm_fail("conv1_fwd: batch must be positive");
  if (lrn && !pool_raw) return dtm_fail("conv1_fwd: LRN needs the pool_raw output buffer");
  {
    const char* e = getenv("DTM_CONV1_IM2COL");      // 1 = the round-1 kernel below (explicit im2col tiles), kept as the A/B arm
    if (!(e && e[0] == '1'))
      return dtm_conv1_fwd_shifted_window(x, in_is_float, channels, w, bias, batch, lrn, p1_planes, argmax, pool_raw, stream);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* p = static_cast<uint8_t*>(p1_planes);
  uint8_t* a = static_cast<uint8_t*>(argmax);
  __nv_bfloat16* r = static_cast<__nv_bfloat16*>(pool_raw);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, T1Geom<1>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, T1Geom<1>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_tc_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, T1Geom<3>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_tc_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, T1Geom<3>::SMEM));
    configured[dev_slot] = true;
  }
  // persistent CTAs; the 1-channel geometry needs ~43 KiB of shared memory, so two CTAs share an SM and overlap their phases
  const int per_sm = channels == 1 ? 2 : 1;
  const int grid = min(batch, per_sm * device_sm_count());
  if (channels == 1) {
    if (lrn) DTM_CUDA_OK(launch_pdl(conv1_tc_kernel<1, true>, dim3(grid), dim3(320), T1Geom<1>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
    else     DTM_CUDA_OK(launch_pdl(conv1_tc_kernel<1, false>, dim3(grid), dim3(320), T1Geom<1>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
  } else {
    if (lrn) DTM_CUDA_OK(launch_pdl(conv1_tc_kernel<3, true>, dim3(grid), dim3(320), T1Geom<3>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
    else     DTM_CUDA_OK(launch_pdl(conv1_tc_kernel<3, false>, dim3(grid), dim3(320), T1Geom<3>::SMEM, s, x, in_is_float, w, bias, batch, p, a, r));
  }
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
