// conv2 weight gradient on tcgen05 (autodiff of mnist_input.py:161):
//   dW2[kh][kw][ci][co] = sum_{image, pixel} P1[pixel + tap][ci] * dZ2[pixel][co]
//
// Here the reduction (K) dimension is *pixels*.  Both operands are read straight from their plane-layout images
// as MN-major UMMA operands: a plane's 8 pixels x 8 channels x 2 B block is a canonical MN-major core matrix
// (channels contiguous, pixels 16 B apart), planes are the M/N groups (SBO = 4096 B), and the filter tap is once
// more a re-based start address on the P1 side.  Per image and tap: 14 tcgen05.mma of shape 64(co) x 32(ci) x 16
// (one per image row — the zero rows 0-1 of the dZ2 plane are skipped), accumulated over all images of the CTA
// in TMEM.  25 taps x 32 columns exceed the 512 TMEM columns, so blockIdx.y splits the taps 13 / 12.
// The M = 64 accumulator occupies TMEM lanes {0-15, 32-47, 64-79, 96-111} (measured: profiles/umma_probe_r1.json).
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "plane_layout.cuh"

namespace dtm {

constexpr int WG_CI = 32, WG_CO = 64;
constexpr int WG_P1_BYTES = plane_image_bytes(WG_CI);    // 16384
constexpr int WG_DZ_BYTES = plane_image_bytes(WG_CO);    // 32768
constexpr int WG_STAGE_BYTES = PL_GUARD_BYTES + WG_P1_BYTES + PL_GUARD_BYTES + WG_DZ_BYTES;  // 53248
constexpr int WG_STAGES = 3;
constexpr int WG_SMEM = WG_STAGES * WG_STAGE_BYTES + 128;
constexpr int WG_TAPS_Y0 = 13;

// All MMAs of one image for this CTA's tap group.  Taps are unrolled up to the larger group (13) with a uniform
// predicate for the 12-tap group, K steps (image rows) stay a loop: every descriptor is base + compile-time offset
// + ks * 256 B.
template <bool FIRST_IMAGE>
DTM_DEVICE void issue_image(uint32_t tmem_base, const UmmaDescBase& dz_base, const UmmaDescBase& p1_base, int tap0, int ntaps,
                            uint32_t idesc) {
#pragma unroll 1
  for (int ks = PL_ROW_OFF; ks < PL_GRID; ++ks) {   // one K step = one image row of 16 grid pixels
    const uint64_t da = umma_desc_at(dz_base, ks * 256);
    const bool first = FIRST_IMAGE && ks == PL_ROW_OFF;
    // tap = tap0 + tl, tap0 in {0, 13}: (kh, kw) offsets differ between the two groups, so both tables are unrolled
#pragma unroll
    for (int tl = 0; tl < WG_TAPS_Y0; ++tl) {
      if (tl < ntaps) {
        const int tap_a = tl, tap_b = WG_TAPS_Y0 + tl;
        const int off_a = (((tap_a / 5) - 2) * PL_GRID + (tap_a % 5) - 2) * 16;
        const int off_b = (((tap_b / 5) - 2) * PL_GRID + (tap_b % 5) - 2) * 16;
        const uint64_t db = umma_desc_at(p1_base, ks * 256 + (tap0 == 0 ? off_a : off_b));
        if (first) umma_f16_c<false>(tmem_base + tl * WG_CI, da, db, idesc);
        else       umma_f16_c<true>(tmem_base + tl * WG_CI, da, db, idesc);
      }
    }
  }
}

__global__ void __launch_bounds__(256, 1)
conv2_wgrad_kernel(const uint8_t* __restrict__ p1_planes, const uint8_t* __restrict__ dz2_planes, int batch,
                   float* __restrict__ dw2) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[WG_STAGES], empty_bar[WG_STAGES], done_bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int tap0 = blockIdx.y == 0 ? 0 : WG_TAPS_Y0;
  const int ntaps = blockIdx.y == 0 ? WG_TAPS_Y0 : 25 - WG_TAPS_Y0;

  if (warp == 1 && elect_one()) {
    for (int i = 0; i < WG_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) { tmem_alloc(&tmem_slot, 512); tmem_relinquish(); }
  {  // zero the guards around every P1 buffer (taps over-read into them; they must hold finite zeros)
    uint4 z = make_uint4(0, 0, 0, 0);
    for (int s = 0; s < WG_STAGES; ++s) {
      uint8_t* st = smem + s * WG_STAGE_BYTES;
      for (int i = threadIdx.x * 16; i < PL_GUARD_BYTES; i += blockDim.x * 16) {
        *reinterpret_cast<uint4*>(st + i) = z;
        *reinterpret_cast<uint4*>(st + PL_GUARD_BYTES + WG_P1_BYTES + i) = z;
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int first = blockIdx.x, stride = gridDim.x;
  pdl_wait();                 // both operand plane sets come from preceding kernels
  pdl_launch_dependents();

  if (warp == 0) {
    if (elect_one()) {
      int it = 0;
      for (int img = first; img < batch; img += stride, ++it) {
        const int s = it % WG_STAGES;
        mbar_wait(&empty_bar[s], ((it / WG_STAGES) & 1) ^ 1, 30);
        mbar_arrive_expect_tx(&full_bar[s], WG_P1_BYTES + WG_DZ_BYTES);
        uint8_t* st = smem + s * WG_STAGE_BYTES;
        bulk_load_1d(st + PL_GUARD_BYTES, p1_planes + static_cast<size_t>(img) * WG_P1_BYTES, WG_P1_BYTES, &full_bar[s]);
        bulk_load_1d(st + 2 * PL_GUARD_BYTES + WG_P1_BYTES, dz2_planes + static_cast<size_t>(img) * WG_DZ_BYTES, WG_DZ_BYTES,
                     &full_bar[s]);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(64, WG_CI, true, true);
      int it = 0;
      for (int img = first; img < batch; img += stride, ++it) {
        const int s = it % WG_STAGES;
        mbar_wait(&full_bar[s], (it / WG_STAGES) & 1, 31);
        tc_fence_after();
        const UmmaDescBase p1_base = umma_desc_base(smem_u32(smem + s * WG_STAGE_BYTES + PL_GUARD_BYTES), 128, PL_PLANE_BYTES, UMMA_SWIZZLE_NONE);
        const UmmaDescBase dz_base = umma_desc_base(smem_u32(smem + s * WG_STAGE_BYTES + 2 * PL_GUARD_BYTES + WG_P1_BYTES), 128,
                                                    PL_PLANE_BYTES, UMMA_SWIZZLE_NONE);
        if (it == 0) issue_image<true>(tmem_base, dz_base, p1_base, tap0, ntaps, idesc);
        else         issue_image<false>(tmem_base, dz_base, p1_base, tap0, ntaps, idesc);
        umma_commit(&empty_bar[s]);
      }
      umma_commit(&done_bar);
    }
  } else if (warp >= 4) {
    const uint32_t q = warp & 3;
    mbar_wait_warp(&done_bar, 0, 32);
    tc_fence_after();
    const int co = q * 16 + lane;  // valid for lane < 16
#pragma unroll 1
    for (int tl = 0; tl < ntaps; ++tl) {
      uint32_t r[32];
      tmem_ld_x32(tmem_base + ((q * 32u) << 16) + tl * WG_CI, r);
      tmem_ld_wait();
      if (lane < 16) {
        float* dst = dw2 + static_cast<size_t>(tap0 + tl) * WG_CI * WG_CO + co;
#pragma unroll
        for (int ci = 0; ci < WG_CI; ++ci) red_add_f32(dst + ci * WG_CO, __uint_as_float(r[ci]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace dtm

extern "C" int dtm_conv2_wgrad(const void* p1_planes, const void* dz2_planes, int batch, float* dw2, void* stream) {
  using namespace dtm;
  if (batch <= 0) return dtm_fail("conv2_wgrad: batch must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    configured[dev_slot] = true;
  }
  const int gx = min(batch, max(1, device_sm_count() / 2));
  DTM_CUDA_OK(launch_pdl(conv2_wgrad_kernel, dim3(gx, 2), dim3(256), WG_SMEM, s, static_cast<const uint8_t*>(p1_planes),
                         static_cast<const uint8_t*>(dz2_planes), batch, dw2));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
