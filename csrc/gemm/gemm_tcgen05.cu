// Fully-connected-layer GEMM on 5th-gen tensor cores:  D[M,N] = A[M,K] . B[K,N]  (bf16 in, fp32 TMEM accumulate)
//
//   * operands staged by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) through a 4-deep mbarrier ring,
//   * one elected thread issues tcgen05.mma (UMMA 128 x 64 x 16), accumulator lives in TMEM,
//   * four epilogue warps read TMEM with tcgen05.ld and apply the layer epilogue in registers:
//       bias + ReLU (+ Philox dropout) -> bf16       (forward of local3/local4, mnist_input.py:184,192)
//       ReLU-mask of a saved activation -> bf16      (dX of the same layers)
//       plain fp32 store / split-K reduction         (dW written straight into the flat gradient buffer)
//
// Either operand may be K-major (K contiguous in global memory) or MN-major (M/N contiguous); the
// TensorFlow [in, out] weight layout therefore serves forward (B is MN-major), dX (B is K-major) and
// dW (both MN-major) without a transposed copy.
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/philox.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"
#include "../conv/plane_layout.cuh"

namespace dtm {

constexpr int BM = 128;
constexpr int BN = 64;
constexpr int BK = 64;
constexpr int UK = 16;  // UMMA K for 16-bit inputs
constexpr uint32_t A_STAGE_BYTES = BM * BK * 2;
constexpr uint32_t B_STAGE_BYTES = BN * BK * 2;
constexpr uint32_t gemm_smem_bytes(int stages) { return stages * (A_STAGE_BYTES + B_STAGE_BYTES) + 1024; }  // + alignment slack

struct GemmArgs {
  int M, N, K;
  int k_blocks_per_split;
  int epilogue;  // DTM_EPI_*
  int ldd;       // leading dimension (elements) of D and of `ref`
  void* d;
  const float* bias;            // [N] or null
  const __nv_bfloat16* ref;     // [M, ldd] activation whose sign gates the output (EPI_MASK)
  float alpha;                  // scale applied to the accumulator (fp32 epilogues)
  float keep_prob;              // dropout keep probability; >= 1 disables
  uint64_t seed;
  const long long* step_ptr;    // device global_step, makes the dropout mask step dependent
  // DTM_EPI_MASK_UNPOOL_BF16: D is the gradient of the 7x7x64 pooled conv2 activation ([M][49*64], one N tile = one pooled
  // pixel); besides D the epilogue routes it through the pooling arg-max into the 14x14 plane-layout tensor conv2's backward
  // kernels consume - the separate un-pooling kernel and its launch gap disappear from the step's critical path
  const uint8_t* argmax;        // [M][49][64]
  uint8_t* planes;              // [M] plane-layout images, 64 channels
  // Cluster split-K: the two CTAs of a (1,1,2) cluster each accumulate half of K for the SAME output tile; CTA 1 ships its
  // fp32 partial tile into CTA 0's shared memory (DSMEM), CTA 0 adds it and runs the fused epilogue.  Doubles the number
  // of SMs streaming operands when the tile count is below half the machine (the fc forward: 64 tiles, K = 3136).
  int cluster_reduce;
  int prefetch;                 // epilogue operand prefetch before the accumulator wait (DTM_EPI_PREFETCH=0 switches it off)
};

// STAGES = 4: two CTAs per SM (grids larger than the machine, short K).  STAGES = 8: one CTA per SM with twice the
// bytes in flight — when the grid does not fill the GPU (the fc forward: 64 tiles, K = 3136) the kernel is bound by
// TMA round-trip latency per K block, and depth is the only lever.
template <bool A_MN, bool B_MN, int STAGES>
__global__ void __launch_bounds__(256, STAGES <= 4 ? 2 : 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[STAGES];
  __shared__ uint64_t empty_bar[STAGES];
  __shared__ uint64_t accum_bar;
  __shared__ uint32_t tmem_slot;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;

  const uint32_t warp = threadIdx.x >> 5;
  const int m0 = blockIdx.y * BM;
  const int n0 = blockIdx.x * BN;
  const int total_kb = (g.K + BK - 1) / BK;
  const int kb_begin = blockIdx.z * g.k_blocks_per_split;
  const int kb_end = min(total_kb, kb_begin + g.k_blocks_per_split);
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(&tmem_slot, BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_slot;
  pdl_wait();                 // operands produced by the preceding kernel are complete from here on
  pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      for (int i = 0; i < num_kb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1, 1);
        mbar_arrive_expect_tx(&full_bar[s], A_STAGE_BYTES + B_STAGE_BYTES);
        const int k0 = (kb_begin + i) * BK;
        uint8_t* a_dst = smem_a + s * A_STAGE_BYTES;
        uint8_t* b_dst = smem_b + s * B_STAGE_BYTES;
        if (A_MN) {  // global [K][M]: two 64-wide M atoms, each [BK k-rows][128 B]
          tma_load_2d(a_dst, &tmap_a, &full_bar[s], m0, k0);
          tma_load_2d(a_dst + BK * 128, &tmap_a, &full_bar[s], m0 + 64, k0);
        } else {     // global [M][K]: one box [BM rows][128 B]
          tma_load_2d(a_dst, &tmap_a, &full_bar[s], k0, m0);
        }
        if (B_MN) {  // global [K][N]
          tma_load_2d(b_dst, &tmap_b, &full_bar[s], n0, k0);
        } else {     // global [N][K]
          tma_load_2d(b_dst, &tmap_b, &full_bar[s], k0, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    if (elect_one()) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, A_MN, B_MN);
      for (int i = 0; i < num_kb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&full_bar[s], ph, 2);
        tc_fence_after();
        // K-major SW128: 8-row groups 1024 B apart, a K step of 16 elements is +32 B inside the row.
        // MN-major SW128: swizzle atoms (64 elements of M/N) BK*128 B apart, 8-k groups 1024 B apart,
        //                 a K step of 16 rows is +2048 B.
        const UmmaDescBase a_base = umma_desc_base(smem_u32(smem_a + s * A_STAGE_BYTES), A_MN ? BK * 128 : 0, 1024, UMMA_SWIZZLE_128B);
        const UmmaDescBase b_base = umma_desc_base(smem_u32(smem_b + s * B_STAGE_BYTES), B_MN ? BK * 128 : 0, 1024, UMMA_SWIZZLE_128B);
#pragma unroll
        for (int j = 0; j < BK / UK; ++j) {
          const uint64_t da = umma_desc_at(a_base, j * (A_MN ? UK * 128 : UK * 2));
          const uint64_t db = umma_desc_at(b_base, j * (B_MN ? UK * 128 : UK * 2));
          if (j == 0) umma_f16(tmem_d, da, db, idesc, i > 0 ? 1u : 0u);
          else        umma_f16_c<true>(tmem_d, da, db, idesc);
        }
        umma_commit(&empty_bar[s]);  // frees this smem stage once the MMAs above have read it
      }
      umma_commit(&accum_bar);       // accumulator complete
    }
  } else if (warp >= 4) {
    // ------------------------------ epilogue ------------------------------
    const uint32_t q = warp & 3;  // TMEM lane quadrant this warp may access
    const int m = m0 + q * 32 + lane_id();
    // epilogue operands on their way while the main loop runs: this row's reference activation (128 B), its arg-max bytes,
    // the bias slice
    if (g.prefetch && m < g.M && n0 < g.N) {
      if (g.ref) prefetch_l1(g.ref + static_cast<size_t>(m) * g.ldd + n0);
      if (g.epilogue == DTM_EPI_MASK_UNPOOL_BF16) prefetch_l1(g.argmax + (static_cast<size_t>(m) * 49 + (n0 >> 6)) * 64);
      if (g.bias && lane_id() < 2) prefetch_l1(g.bias + n0 + lane_id() * 32);
    }
    if (num_kb > 0) {
      mbar_wait(&accum_bar, 0, 3);
      tc_fence_after();
    }
    // partial tile in CTA 0's (idle) stage memory: float4 (c/4, row) at ((c/4) * 128 + row) * 16 - consecutive rows are
    // consecutive 16-byte words, so neither the remote stores nor the loads below conflict
    const uint32_t part_row = smem_u32(smem_a) + (q * 32 + lane_id()) * 16;
    if (g.cluster_reduce) {
      cluster_sync();   // #1: both CTAs have finished their MMAs - CTA 0's stage buffers are free
      if (blockIdx.z == 1) {
        const uint32_t dst = cluster_map_shared(part_row, 0);
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t r[32];
          if (num_kb > 0) {
            tmem_ld_x32(tmem_d + ((q * 32u) << 16) + c, r);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0;
          }
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            st_shared_cluster_v4(dst + ((c + j) >> 2) * (BM * 16), __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                 __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
      }
      cluster_sync();   // #2: the partial tile has landed in CTA 0
    }
    if (!g.cluster_reduce || blockIdx.z == 0) {
    long long step = 0;
    if (g.epilogue == DTM_EPI_BIAS_RELU_BF16 && g.keep_prob < 1.0f && g.step_ptr) step = *g.step_ptr;
    const uint32_t keep_thresh = static_cast<uint32_t>(g.keep_prob * 16777216.0f);
    const float inv_keep = g.keep_prob < 1.0f ? 1.0f / g.keep_prob : 1.0f;

#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      if (num_kb > 0) {
        tmem_ld_x32(tmem_d + ((q * 32u) << 16) + c, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0;
      }
      if (g.cluster_reduce) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float4 p;
          asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(p.x), "=f"(p.y), "=f"(p.z), "=f"(p.w)
                       : "r"(part_row + ((c + j) >> 2) * (BM * 16)));
          r[j] = __float_as_uint(__uint_as_float(r[j]) + p.x);
          r[j + 1] = __float_as_uint(__uint_as_float(r[j + 1]) + p.y);
          r[j + 2] = __float_as_uint(__uint_as_float(r[j + 2]) + p.z);
          r[j + 3] = __float_as_uint(__uint_as_float(r[j + 3]) + p.w);
        }
      }
      const int n = n0 + c;
      if (m >= g.M || n >= g.N) continue;
      const int valid = min(32, g.N - n);  // N % 8 == 0 is enforced by the host wrapper
      if (g.epilogue == DTM_EPI_STORE_F32) {
        float* out = static_cast<float*>(g.d) + static_cast<size_t>(m) * g.ldd + n;
        if (gridDim.z > 1 && !g.cluster_reduce) {
          for (int j = 0; j < valid; j += 4)
            red_add_v4_f32(out + j, g.alpha * __uint_as_float(r[j]), g.alpha * __uint_as_float(r[j + 1]),
                           g.alpha * __uint_as_float(r[j + 2]), g.alpha * __uint_as_float(r[j + 3]));
        } else {
          for (int j = 0; j < valid; j += 4) {
            float4 v = make_float4(g.alpha * __uint_as_float(r[j]), g.alpha * __uint_as_float(r[j + 1]),
                                   g.alpha * __uint_as_float(r[j + 2]), g.alpha * __uint_as_float(r[j + 3]));
            *reinterpret_cast<float4*>(out + j) = v;
          }
        }
      } else {
        __nv_bfloat16* out = static_cast<__nv_bfloat16*>(g.d) + static_cast<size_t>(m) * g.ldd + n;
        const __nv_bfloat16* ref = g.ref ? g.ref + static_cast<size_t>(m) * g.ldd + n : nullptr;
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          if (j >= valid) break;
          float v[8];
#pragma unroll
          for (int t = 0; t < 8; ++t) v[t] = __uint_as_float(r[j + t]);
          if (g.epilogue == DTM_EPI_BIAS_RELU_BF16 || g.epilogue == DTM_EPI_BIAS_BF16) {
            if (g.bias) {
              const float4 b0 = *reinterpret_cast<const float4*>(g.bias + n + j);
              const float4 b1 = *reinterpret_cast<const float4*>(g.bias + n + j + 4);
              v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
              v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
            }
            if (g.epilogue == DTM_EPI_BIAS_RELU_BF16) {
#pragma unroll
              for (int t = 0; t < 8; ++t) v[t] = fmaxf(v[t], 0.0f);
              if (g.keep_prob < 1.0f) {
                const uint64_t e = static_cast<uint64_t>(m) * g.N + (n + j);  // multiple of 4
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                  const Philox4 bits = dropout_bits((e >> 2) + h, static_cast<uint64_t>(step), g.seed);
                  v[h * 4 + 0] = ((bits.x >> 8) < keep_thresh) ? v[h * 4 + 0] * inv_keep : 0.0f;
                  v[h * 4 + 1] = ((bits.y >> 8) < keep_thresh) ? v[h * 4 + 1] * inv_keep : 0.0f;
                  v[h * 4 + 2] = ((bits.z >> 8) < keep_thresh) ? v[h * 4 + 2] * inv_keep : 0.0f;
                  v[h * 4 + 3] = ((bits.w >> 8) < keep_thresh) ? v[h * 4 + 3] * inv_keep : 0.0f;
                }
              }
            }
          } else if (g.epilogue == DTM_EPI_MASK_BF16 || g.epilogue == DTM_EPI_MASK_UNPOOL_BF16) {
            const uint4 rv = *reinterpret_cast<const uint4*>(ref + j);
            const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              v[2 * t] = bf16_lo(rw[t]) > 0.0f ? v[2 * t] : 0.0f;
              v[2 * t + 1] = bf16_hi(rw[t]) > 0.0f ? v[2 * t + 1] : 0.0f;
            }
          }
          uint4 o;
          o.x = pack_bf16x2(v[0], v[1]);
          o.y = pack_bf16x2(v[2], v[3]);
          o.z = pack_bf16x2(v[4], v[5]);
          o.w = pack_bf16x2(v[6], v[7]);
          *reinterpret_cast<uint4*>(out + j) = o;
          if (g.epilogue == DTM_EPI_MASK_UNPOOL_BF16) {
            // this thread holds 8 channels of pooled pixel pp = n / 64 of image m: write the four positions of its 2x2
            // window (the arg-max one gets the value, the others zero) as 16-byte plane pixels
            const int pp = n >> 6, ch = (n & 63) + j;
            const int ph = pp / 7, pw = pp - ph * 7;
            const uint2 av = *reinterpret_cast<const uint2*>(g.argmax + (static_cast<size_t>(m) * 49 + pp) * 64 + ch);
            const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
            uint8_t* img = g.planes + static_cast<size_t>(m) * plane_image_bytes(64);
#pragma unroll
            for (int d = 0; d < 4; ++d) {
              uint32_t z[4];
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const uint32_t a0 = ((t < 2 ? av.x : av.y) >> (16 * (t & 1))) & 0xFF;
                const uint32_t a1 = ((t < 2 ? av.x : av.y) >> (16 * (t & 1) + 8)) & 0xFF;
                z[t] = (a0 == static_cast<uint32_t>(d) ? (ow[t] & 0x0000FFFFu) : 0u) | (a1 == static_cast<uint32_t>(d) ? (ow[t] & 0xFFFF0000u) : 0u);
              }
              *reinterpret_cast<uint4*>(img + plane_offset(2 * ph + (d >> 1), 2 * pw + (d & 1), ch)) = make_uint4(z[0], z[1], z[2], z[3]);
            }
          }
        }
      }
    }
    }   // CTA 0 of a cluster / every CTA without cluster split
  }
  if (warp < 4 && g.cluster_reduce) {   // the cluster barrier counts every thread of both CTAs
    __syncwarp();
    cluster_sync();
    cluster_sync();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_d, BN);
}

template <bool A_MN, bool B_MN, int STAGES>
static int launch_gemm_s(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, dim3 grid, cudaStream_t s) {
  auto kern = gemm_bf16_kernel<A_MN, B_MN, STAGES>;
  static bool configured[64] = {};     // per device: function attributes belong to a device's context
  const int dev_slot = device_slot();
  if (!configured[dev_slot]) {
    DTM_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_bytes(STAGES)));
    configured[dev_slot] = true;
  }
  DTM_CUDA_OK(launch_pdl_cluster(kern, grid, dim3(256), gemm_smem_bytes(STAGES), s, g.cluster_reduce ? 2 : 1, ta, tb, g));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

template <bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, int splits, cudaStream_t s) {
  dim3 grid((g.N + BN - 1) / BN, (g.M + BM - 1) / BM, splits);
  const int tiles = grid.x * grid.y * grid.z;
  if (tiles <= device_sm_count() && g.k_blocks_per_split > 8) return launch_gemm_s<A_MN, B_MN, 8>(ta, tb, g, grid, s);
  return launch_gemm_s<A_MN, B_MN, 4>(ta, tb, g, grid, s);
}

}  // namespace dtm

namespace dtm {
static int gemm_run(const void* a, int a_mn_major, const void* b, int b_mn_major, void* d, int M, int N, int K, int ldd,
                    int epilogue, const float* bias, const void* ref, float alpha, float keep_prob, uint64_t seed,
                    const long long* step_ptr, int split_k, const void* argmax, void* planes, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return dtm_fail("gemm: M, N, K must be positive");
  if (N % 8 != 0 || ldd % 8 != 0) return dtm_fail("gemm: N and ldd must be multiples of 8");
  if (K % 8 != 0) return dtm_fail("gemm: K must be a multiple of 8 (16-byte TMA rows)");
  if (M % 8 != 0 && a_mn_major) return dtm_fail("gemm: MN-major A needs M % 8 == 0");
  if (epilogue < 0 || epilogue > DTM_EPI_MASK_UNPOOL_BF16) return dtm_fail("gemm: unknown epilogue");
  if ((epilogue == DTM_EPI_MASK_BF16 || epilogue == DTM_EPI_MASK_UNPOOL_BF16) && !ref)
    return dtm_fail("gemm: EPI_MASK needs a reference activation");
  if (epilogue == DTM_EPI_MASK_UNPOOL_BF16 && (!argmax || !planes || N != 49 * 64 || ldd != N))
    return dtm_fail("gemm: EPI_MASK_UNPOOL needs argmax, planes and N = ldd = 3136");
  const int total_kb = (K + BK - 1) / BK;
  int splits = split_k < 1 ? 1 : split_k;
  if (splits > total_kb) splits = total_kb;
  if (splits > 1 && epilogue != DTM_EPI_STORE_F32) return dtm_fail("gemm: split-K requires the fp32 store epilogue");
  // bf16 epilogues with few tiles and a deep K: split K over the two CTAs of a cluster (reduction through DSMEM).
  // Opt-in (DTM_GEMM_CLUSTER=1): measured on the fc forward (64 tiles, K = 3136) the kernel alone gains 17.0 -> 15.7 us,
  // but inside the training step's programmatic-launch chain the cluster launch costs more than that (170.6 -> 180.3 us
  // per step), so the engine leaves it off.
  const int tiles_mn = ((N + BN - 1) / BN) * ((M + BM - 1) / BM);
  const bool cluster_reduce = splits == 1 && epilogue != DTM_EPI_STORE_F32 && 2 * tiles_mn <= device_sm_count() && total_kb >= 16 &&
                              (getenv("DTM_GEMM_CLUSTER") && getenv("DTM_GEMM_CLUSTER")[0] == '1');
  if (cluster_reduce) splits = 2;
  const int per = (total_kb + splits - 1) / splits;
  splits = (total_kb + per - 1) / per;

  CUtensorMap ta, tb;
  int rc;
  if (a_mn_major) rc = make_tmap_2d_bf16(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)M * 2, 64, BK, 128);
  else            rc = make_tmap_2d_bf16(&ta, a, (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, BK, BM, 128);
  if (rc) return rc;
  if (b_mn_major) rc = make_tmap_2d_bf16(&tb, b, (uint64_t)N, (uint64_t)K, (uint64_t)N * 2, 64, BK, 128);
  else            rc = make_tmap_2d_bf16(&tb, b, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, BK, BN, 128);
  if (rc) return rc;

  GemmArgs g;
  g.M = M; g.N = N; g.K = K;
  g.k_blocks_per_split = per;
  g.epilogue = epilogue;
  g.ldd = ldd;
  g.d = d;
  g.bias = bias;
  g.ref = static_cast<const __nv_bfloat16*>(ref);
  g.alpha = alpha;
  g.keep_prob = keep_prob;
  g.seed = seed;
  g.step_ptr = step_ptr;
  g.cluster_reduce = (cluster_reduce && splits == 2) ? 1 : 0;
  {
    const char* pe = getenv("DTM_EPI_PREFETCH");
    g.prefetch = (pe && pe[0] == '0') ? 0 : 1;
  }
  g.argmax = static_cast<const uint8_t*>(argmax);
  g.planes = static_cast<uint8_t*>(planes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a_mn_major && b_mn_major) return launch_gemm<true, true>(ta, tb, g, splits, s);
  if (a_mn_major && !b_mn_major) return launch_gemm<true, false>(ta, tb, g, splits, s);
  if (!a_mn_major && b_mn_major) return launch_gemm<false, true>(ta, tb, g, splits, s);
  return launch_gemm<false, false>(ta, tb, g, splits, s);
}
}  // namespace dtm

extern "C" int dtm_gemm_bf16(const void* a, int a_mn_major, const void* b, int b_mn_major, void* d, int M, int N,
                             int K, int ldd, int epilogue, const float* bias, const void* ref, float alpha,
                             float keep_prob, uint64_t seed, const long long* step_ptr, int split_k, void* stream) {
  if (epilogue == DTM_EPI_MASK_UNPOOL_BF16) return dtm::dtm_fail("gemm: use dtm_gemm_mask_unpool for the un-pooling epilogue");
  return dtm::gemm_run(a, a_mn_major, b, b_mn_major, d, M, N, K, ldd, epilogue, bias, ref, alpha, keep_prob, seed, step_ptr,
                       split_k, nullptr, nullptr, stream);
}

extern "C" int dtm_gemm_mask_unpool(const void* dy, const void* w_bf16, const void* p2, const void* argmax, void* dp2,
                                    void* dz2_planes, int batch, int hidden, void* stream) {
  return dtm::gemm_run(dy, 0, w_bf16, 0, dp2, batch, 49 * 64, hidden, 49 * 64, DTM_EPI_MASK_UNPOOL_BF16, nullptr, p2, 1.0f,
                       1.0f, 0, nullptr, 1, argmax, dz2_planes, stream);
}
