// conv1 weight + bias gradient on tcgen05 (autodiff of mnist_input.py:142-147; conv1 has no input gradient).
// OPT-IN (DTM_C1W_TC=1): numerically verified for C_in = 1 and 3, but in round 1 still slower than the CUDA-core kernel in
// conv1.cu (see the note there).  Next steps if it is picked up again: bulk-prefetch dpool1/arg-max per image into shared
// memory instead of per-row global loads, keep `imu` in the shared address space (the im2col loads compile to generic
// LD.E.U16), and shrink conv2_wgrad's TMEM allocation so that both kernels can hold accumulators on one SM.
//
//   dW1[k = tap*C_in + ci][co] = sum over images and conv pixels of  patch[pixel][k] * dZ1[pixel][co]
//
// where dZ1 is the pooled gradient routed through the saved arg-max (one of the four pixels of a pooling window gets the
// value, the others zero).  The CUDA-core version spent 23 K warp instructions per image on one LDS + one FFMA per
// multiply-add and was issue bound (ncu: 11.8 M instructions, the most of any kernel of the step, 29 us; 44 us when it
// shares the SMs with conv2_wgrad and the optimizer update).  Here the reduction runs on the tensor core with K = pixels:
//
//     A  [M = taps (26 -> 64 | 76 -> 128)] x [K = 112 pixels]   MN-major, no swizzle: [k-group of 8 taps][pixel][8 taps]
//     B  [N = 32 channels]                 x [K = 112 pixels]   MN-major, no swizzle: [8 channels][pixel][8 channels]
//     D  [M x 32] fp32 in TMEM, accumulated over EVERY work unit of the CTA - one epilogue per CTA.
//
// A is the same im2col row conv1's forward kernel builds (16-byte chunks of 8 taps); row k = 25*C_in is the constant 1,
// so D[25*C_in][co] is the bias gradient for free.  B is the un-pooled gradient, built on the fly from dpool1 + arg-max.
// The pixel order inside a work unit is arbitrary as long as both operands agree: unit = two pooled rows, pixel index
// = pooled pixel * 4 + window position, 112 pixels = 7 MMA K steps.  Taps beyond the real ones (M padding) read whatever
// follows in shared memory: rows of D are independent, and those rows are never stored.
//
// Warp roles: 0-7 producers in two groups of four warps that take alternate work units (one K row = one conv pixel per
// thread; the kernel is bound by the producers' instruction latency, so two units are built concurrently and the global
// loads of a unit are issued before the wait for its shared-memory slot); warps 0-3 also run the epilogue; warp 8 is the
// MMA issuer and TMEM allocator.
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int W1_OUT = 32;
constexpr int W1_PITCH = 32;                  // halo image: 32 x 32 per input channel (2-pixel zero border)
constexpr int W1_ROWS = 112;                  // K rows per work unit: 28 pooled pixels x 4 window positions
constexpr int W1_UNITS = 7;                   // units per image (14 pooled rows / 2)
constexpr int W1_GROUP_BYTES = W1_ROWS * 16;  // one group of 8 taps (or 8 channels) for every row of the unit
constexpr int W1_STAGES = 3;

template <int CIN> struct W1Geom {
  static constexpr int KBIAS = 25 * CIN;                       // index of the all-ones row
  static constexpr int TG = (KBIAS + 1 + 7) / 8;               // tap groups actually built (4 | 10)
  static constexpr int UM = TG * 8 <= 64 ? 64 : 128;           // UMMA M
  static constexpr int A_BYTES = TG * W1_GROUP_BYTES;
  static constexpr int B_BYTES = (W1_OUT / 8) * W1_GROUP_BYTES;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // the M padding of the last stage reads up to (UM/8) groups from its A base
  static constexpr int TAIL = (UM / 8) * W1_GROUP_BYTES > STAGE_BYTES ? (UM / 8) * W1_GROUP_BYTES - STAGE_BYTES : 0;
  static constexpr int IMG_BYTES = CIN * W1_PITCH * W1_PITCH * 2;
  static constexpr int SMEM = W1_STAGES * STAGE_BYTES + TAIL + 2 * IMG_BYTES + 256;
};

constexpr int W1_PRODUCERS = 256;

template <int CIN>
__global__ void __launch_bounds__(W1_PRODUCERS + 32, 2)
conv1_wgrad_tc_kernel(const void* __restrict__ x, int in_is_float, const __nv_bfloat16* __restrict__ grad_pooled,
                      const uint8_t* __restrict__ argmax, int batch, float* __restrict__ dw, float* __restrict__ db) {
  using G = W1Geom<CIN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[W1_STAGES], empty_bar[W1_STAGES], done_bar;
  __shared__ uint32_t tmem_slot;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* s_stage = smem;
  __nv_bfloat16* s_img = reinterpret_cast<__nv_bfloat16*>(smem + W1_STAGES * G::STAGE_BYTES + G::TAIL);   // [2][CIN][32][32]

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int first = blockIdx.x, stride = gridDim.x;

  if (warp == 8) {
    if (elect_one()) {
      for (int i = 0; i < W1_STAGES; ++i) { mbar_init(&full_bar[i], 128); mbar_init(&empty_bar[i], 1); }
      mbar_init(&done_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(&tmem_slot, 32);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < 2 * CIN * W1_PITCH * W1_PITCH; i += blockDim.x) s_img[i] = __float2bfloat16(0.0f);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  pdl_wait();                 // grad_pooled is the output of the kernel in front of this one
  pdl_launch_dependents();

  if (warp < 8) {
    // ================================ producers: one K row (conv pixel) per thread ================================
    const int grp = warp >> 2;              // producer group: takes the units with (unit counter & 1) == grp
    const int t = threadIdx.x & 127;        // row inside the unit, rows 112..127 idle
    const int pl = t >> 2, d = t & 3;       // pooled pixel inside the unit, window position
    int it = 0, img_it = 0;
    for (int img = first; img < batch; img += stride, ++img_it) {
      __nv_bfloat16* im = s_img + (img_it & 1) * (CIN * W1_PITCH * W1_PITCH);
      const int npx = 28 * 28 * CIN;
      if (in_is_float) {
        const float* src = static_cast<const float*>(x) + static_cast<size_t>(img) * npx;
        for (int i = threadIdx.x; i < npx; i += W1_PRODUCERS) {
          const int ci = i % CIN, p = i / CIN;
          im[(ci * W1_PITCH + p / 28 + 2) * W1_PITCH + p % 28 + 2] = __float2bfloat16(src[i]);
        }
      } else {
        const uint8_t* src = static_cast<const uint8_t*>(x) + static_cast<size_t>(img) * npx;
        for (int i = threadIdx.x; i < npx; i += W1_PRODUCERS) {
          const int ci = i % CIN, p = i / CIN;
          im[(ci * W1_PITCH + p / 28 + 2) * W1_PITCH + p % 28 + 2] =
              __float2bfloat16(static_cast<float>(src[i]) * (1.0f / 255.0f) - 0.5f);
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");   // producers only: the image is complete
      const uint16_t* imu = reinterpret_cast<const uint16_t*>(im);
#pragma unroll 1
      for (int u = 0; u < W1_UNITS; ++u, ++it) {
        if ((it & 1) != grp) continue;
        const int slot = it % W1_STAGES;
        const int pp = u * 28 + pl;
        const size_t prow = (static_cast<size_t>(img) * (IMG * IMG) + (t < W1_ROWS ? pp : 0)) * W1_OUT;
        // the global loads of this row first: they are in flight while the slot is awaited and the im2col row is built
        uint4 gv[W1_OUT / 8];
        uint2 av[W1_OUT / 8];
#pragma unroll
        for (int cg = 0; cg < W1_OUT / 8; ++cg) {
          gv[cg] = *reinterpret_cast<const uint4*>(grad_pooled + prow + cg * 8);
          av[cg] = *reinterpret_cast<const uint2*>(argmax + prow + cg * 8);
        }
        mbar_wait(&empty_bar[slot], ((it / W1_STAGES) & 1) ^ 1, 50);
        if (t < W1_ROWS) {
          const int ph = pp / IMG, pw = pp - ph * IMG;
          const int y = 2 * ph + (d >> 1), xx = 2 * pw + (d & 1);   // conv output pixel == top-left of its patch in halo coords
          uint8_t* a_dst = s_stage + slot * G::STAGE_BYTES + t * 16;
#pragma unroll
          for (int k8 = 0; k8 < G::TG; ++k8) {
            uint32_t pk[4] = {0, 0, 0, 0};
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              const int k = k8 * 8 + kk;
              uint32_t v = 0;
              if (k < G::KBIAS) {
                const int tap = k / CIN, ci = k % CIN;
                v = imu[(ci * W1_PITCH + y + tap / 5) * W1_PITCH + xx + tap % 5];
              } else if (k == G::KBIAS) {
                v = 0x3F80u;   // bf16 1.0: this row of D is the bias gradient
              }
              pk[kk >> 1] |= v << (16 * (kk & 1));
            }
            *reinterpret_cast<uint4*>(a_dst + k8 * W1_GROUP_BYTES) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          }
          // un-pooled gradient row: channel c gets the pooled value iff its arg-max names this window position
          uint8_t* b_dst = a_dst + G::A_BYTES;
#pragma unroll
          for (int cg = 0; cg < W1_OUT / 8; ++cg) {
            const uint32_t gw[4] = {gv[cg].x, gv[cg].y, gv[cg].z, gv[cg].w};
            uint32_t z[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a0 = ((j < 2 ? av[cg].x : av[cg].y) >> (16 * (j & 1))) & 0xFF;
              const uint32_t a1 = ((j < 2 ? av[cg].x : av[cg].y) >> (16 * (j & 1) + 8)) & 0xFF;
              z[j] = (a0 == static_cast<uint32_t>(d) ? (gw[j] & 0x0000FFFFu) : 0u) | (a1 == static_cast<uint32_t>(d) ? (gw[j] & 0xFFFF0000u) : 0u);
            }
            *reinterpret_cast<uint4*>(b_dst + cg * W1_GROUP_BYTES) = make_uint4(z[0], z[1], z[2], z[3]);
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[slot]);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");   // nobody still reads this image buffer when it is re-staged
    }
    if (warp >= 4) goto done;   // the epilogue needs one warp per TMEM lane quadrant
    // ================================ epilogue (once per CTA) ================================
    mbar_wait(&done_bar, 0, 52);
    tc_fence_after();
    const uint32_t q = warp;
    uint32_t r[32];
    tmem_ld_x32(tmem_base + ((q * 32u) << 16), r);
    tmem_ld_wait();
    // M = 64: rows 16q..16q+15 sit in lanes 0-15 of quadrant q; M = 128: row = TMEM lane
    const bool lane_ok = G::UM == 128 || lane < 16;
    const int k = G::UM == 128 ? static_cast<int>(q * 32 + lane) : static_cast<int>(q * 16 + lane);
    if (lane_ok && first < batch) {
      if (k < G::KBIAS) {
#pragma unroll
        for (int co = 0; co < W1_OUT; ++co) red_add_f32(dw + k * W1_OUT + co, __uint_as_float(r[co]));
      } else if (k == G::KBIAS) {
#pragma unroll
        for (int co = 0; co < W1_OUT; ++co) red_add_f32(db + co, __uint_as_float(r[co]));
      }
    }
  } else {
    // ================================ MMA issuer ================================
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(G::UM, W1_OUT, true, true);
      int it = 0;
      for (int img = first; img < batch; img += stride) {
#pragma unroll 1
        for (int u = 0; u < W1_UNITS; ++u, ++it) {
          const int slot = it % W1_STAGES;
          mbar_wait(&full_bar[slot], (it / W1_STAGES) & 1, 51);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(s_stage + slot * G::STAGE_BYTES);
          // MN-major, no swizzle: 8-pixel groups 128 B apart (LBO), 8-tap / 8-channel groups one W1_GROUP_BYTES apart (SBO)
          const UmmaDescBase a_desc = umma_desc_base(a_addr, 128, W1_GROUP_BYTES, UMMA_SWIZZLE_NONE);
          const UmmaDescBase b_desc = umma_desc_base(a_addr + G::A_BYTES, 128, W1_GROUP_BYTES, UMMA_SWIZZLE_NONE);
#pragma unroll
          for (int j = 0; j < W1_ROWS / 16; ++j) {
            const uint64_t da = umma_desc_at(a_desc, j * 256);
            const uint64_t dbb = umma_desc_at(b_desc, j * 256);
            umma_f16(tmem_base, da, dbb, idesc, (it > 0 || j > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[slot]);
        }
      }
      umma_commit(&done_bar);
    }
  }

done:
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_base, 32);
}

}  // namespace dtm

extern "C" int dtm_conv1_wgrad_tc(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax,
                                  int batch, float* dw, float* db, void* stream) {
  using namespace dtm;
  if (channels != 1 && channels != 3) return dtm_fail("conv1_wgrad: channels must be 1 or 3");
  if (batch <= 0) return dtm_fail("conv1_wgrad: batch must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, W1Geom<1>::SMEM));
    DTM_CUDA_OK(cudaFuncSetAttribute(conv1_wgrad_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, W1Geom<3>::SMEM));
    configured[dev_slot] = true;
  }
  // persistent CTAs (one accumulator and one epilogue each); two per SM overlap staging with MMA issue
  const int grid = min(batch, 2 * device_sm_count());
  const __nv_bfloat16* g = static_cast<const __nv_bfloat16*>(grad_pooled);
  const uint8_t* a = static_cast<const uint8_t*>(argmax);
  if (channels == 1)
    DTM_CUDA_OK(launch_pdl(conv1_wgrad_tc_kernel<1>, dim3(grid), dim3(W1_PRODUCERS + 32), W1Geom<1>::SMEM, s, x, in_is_float, g, a, batch, dw, db));
  else
    DTM_CUDA_OK(launch_pdl(conv1_wgrad_tc_kernel<3>, dim3(grid), dim3(W1_PRODUCERS + 32), W1Geom<3>::SMEM, s, x, in_is_float, g, a, batch, dw, db));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
