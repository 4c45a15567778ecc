// Thin inline-PTX layer for sm_100a: mbarrier, TMA (cp.async.bulk[.tensor]), tcgen05 (UMMA / TMEM).
// Everything here is hand-written PTX; no CUTLASS/CuTe types are used on any path of this library.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

namespace dtm {

#define DTM_DEVICE __device__ __forceinline__

// ---------------------------------------------------------------------------------------------
// Generic helpers
// ---------------------------------------------------------------------------------------------
DTM_DEVICE uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

DTM_DEVICE uint32_t lane_id() {
  uint32_t l;
  asm volatile("mov.u32 %0, %%laneid;" : "=r"(l));
  return l;
}

// ---- thread-block clusters / distributed shared memory ------------------------------------------------------------
DTM_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// Barrier over every thread of the cluster (non-aligned form: warps whose lanes diverged may call it lane by lane).
// Release/acquire: shared-memory writes made before it - including writes into a peer CTA's shared memory - are visible
// to every thread of the cluster after it.
DTM_DEVICE void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
// Address of `local_smem_addr` (a shared::cta address of this CTA's window) in CTA `rank` of the cluster.
DTM_DEVICE uint32_t cluster_map_shared(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
DTM_DEVICE void st_shared_cluster_v4(uint32_t cluster_addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(cluster_addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Programmatic dependent launch (PDL).  A kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still running; everything it does before pdl_wait() must be independent of the
// predecessor's output (barrier init, TMEM allocation, staging of weights that were written at least two kernels
// upstream).  pdl_wait() returns once the predecessor grid has completed and its memory is visible; without the launch
// attribute it returns immediately.  EVERY thread of EVERY kernel in a PDL chain calls it, so that "grid complete"
// stays transitive along the chain.  pdl_launch_dependents() lets the next kernel's CTAs be scheduled as soon as SM
// resources free up.
DTM_DEVICE void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
DTM_DEVICE void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

DTM_DEVICE bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

// Watchdog budget for every spin-wait in this library (SM clock cycles, ~2 s at 2 GHz). A wait that
// exceeds it prints a tag and traps: a wedged pipeline becomes a launch failure, never a hung GPU.
#ifndef DTM_WATCHDOG_CYCLES
#define DTM_WATCHDOG_CYCLES (4000000000ll)
#endif

DTM_DEVICE void watchdog_trap(const char* what, int a, int b) {
  printf("[dtm watchdog] %s timed out (block %d,%d thread %d; info %d,%d)\n", what, blockIdx.x, blockIdx.y,
         threadIdx.x, a, b);
  __trap();
}

// ---------------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------------
DTM_DEVICE void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
DTM_DEVICE void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy writes to smem -> visible to the async proxy (TMA / UMMA operand reads)
DTM_DEVICE void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

DTM_DEVICE void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
DTM_DEVICE void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
DTM_DEVICE bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
DTM_DEVICE void mbar_wait(uint64_t* bar, uint32_t parity, int tag = 0) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > DTM_WATCHDOG_CYCLES) watchdog_trap("mbarrier wait", tag, (int)parity);
  }
}

// Warp-wide wait for LONG waits (an epilogue warp waiting for the whole main loop, a builder warp waiting for the tensor
// core to release a buffer): lane 0 polls with a sleep between polls, the other 31 lanes park at the warp barrier, so
// that a waiting warp does not take issue slots from the MMA-issuing warp of the same scheduler.  Every lane must call
// it (converged warp).
DTM_DEVICE void mbar_wait_warp(uint64_t* bar, uint32_t parity, int tag = 0) {
  if (lane_id() == 0 && !mbar_try_wait(bar, parity)) {
    const long long t0 = clock64();
    do {
      __nanosleep(100);
      if (clock64() - t0 > DTM_WATCHDOG_CYCLES) watchdog_trap("mbarrier wait (warp)", tag, (int)parity);
    } while (!mbar_try_wait(bar, parity));
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// TMA
// ---------------------------------------------------------------------------------------------
DTM_DEVICE void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}

// 2-D tiled tensor load: box lands densely (inner dim first, swizzled per the tensor map) at `dst`.
DTM_DEVICE void tma_load_2d(void* dst, const void* tmap, uint64_t* bar, int32_t c_inner, int32_t c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
      : "memory");
}

// 1-D bulk copy global -> shared (UBLKCP), completion signalled on an mbarrier. bytes % 16 == 0.
DTM_DEVICE void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// 1-D bulk copy shared -> global (bulk-group completion).
DTM_DEVICE void bulk_store_1d(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(reinterpret_cast<uint64_t>(dst)),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
DTM_DEVICE void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
DTM_DEVICE void bulk_wait_group_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
DTM_DEVICE void bulk_wait_group() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05: TMEM allocation, MMA, commit, loads
// ---------------------------------------------------------------------------------------------
DTM_DEVICE void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {  // whole warp, ncols pow2 >= 32
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
}
DTM_DEVICE void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
DTM_DEVICE void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp (the allocating one)
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
DTM_DEVICE void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DTM_DEVICE void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16/fp16 inputs with fp32 accumulate.
DTM_DEVICE void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Same with the accumulate flag known at compile time (no predicate set-up in the issue loop).
template <bool ACCUMULATE>
DTM_DEVICE void umma_f16_c(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc) {
  if (ACCUMULATE) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.eq.b32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc)
        : "memory");
  }
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed.
DTM_DEVICE void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// TMEM -> registers, shape 32x32b: thread i of the warp reads TMEM lane (lane_base + i), N consecutive
// 32-bit columns. taddr = (lane << 16) | column. A warp may only touch lanes [32*(warp%4), +32).
DTM_DEVICE void tmem_ld_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
DTM_DEVICE void tmem_ld_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
DTM_DEVICE void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// UMMA descriptors (bit layouts per the PTX ISA "tcgen05 matrix / instruction descriptor" tables)
// ---------------------------------------------------------------------------------------------
enum : uint32_t { UMMA_SWIZZLE_NONE = 0, UMMA_SWIZZLE_128B = 2, UMMA_SWIZZLE_64B = 4, UMMA_SWIZZLE_32B = 6 };

// 64-bit shared-memory matrix descriptor. Offsets are bytes (multiples of 16).
//   K-major,  no swizzle : rows of a core matrix 16 B apart; LBO = K-direction core-matrix stride,
//                          SBO = M/N-direction 8-row-group stride.
//   K-major,  swizzle    : SBO = stride between 8-row groups (8 * swizzle bytes); LBO unused.
//   MN-major, no swizzle : SBO = stride between 8-element MN groups, LBO = stride between 8-k groups.
//   MN-major, swizzle    : LBO = stride between swizzle atoms along MN, SBO = stride between 8-k groups.
DTM_DEVICE uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t swizzle) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;  // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(swizzle & 7u) << 61;
  return d;
}

// Descriptor arithmetic for issue loops.  Building a 64-bit descriptor from scratch costs ~20 dependent ALU ops, and the
// single MMA-issuing thread executes them back to back: that, not the tensor core, bounds the issue rate.  The start
// address occupies the low 14 bits of the low word, so moving an operand by a (compile-time) byte offset is one
// 32-bit add on a pre-built base; the high word never changes.
struct UmmaDescBase {
  uint32_t lo, hi;
};
DTM_DEVICE UmmaDescBase umma_desc_base(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t swizzle) {
  const uint64_t d = umma_smem_desc(smem_addr, lbo_bytes, sbo_bytes, swizzle);
  return UmmaDescBase{static_cast<uint32_t>(d), static_cast<uint32_t>(d >> 32)};
}
DTM_DEVICE uint64_t umma_desc_at(const UmmaDescBase& b, int byte_offset) {
  return (static_cast<uint64_t>(b.hi) << 32) | static_cast<uint64_t>(b.lo + static_cast<uint32_t>(byte_offset >> 4));
}

// 32-bit instruction descriptor for kind::f16 with bf16 A/B and fp32 D.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(uint32_t M, uint32_t N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4)                          // D format: F32
         | (1u << 7)                        // A format: BF16
         | (1u << 10)                       // B format: BF16
         | ((a_mn_major ? 1u : 0u) << 15)   // A major-ness
         | ((b_mn_major ? 1u : 0u) << 16)   // B major-ness
         | ((N >> 3) << 17)                 // N / 8
         | ((M >> 4) << 24);                // M / 16
}

// ---------------------------------------------------------------------------------------------
// Small numeric helpers
// ---------------------------------------------------------------------------------------------
DTM_DEVICE uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
DTM_DEVICE float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
DTM_DEVICE float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xFFFF0000u); }

// Cache hint without a destination register: an epilogue issues it for the operands it will need BEFORE it waits for the
// accumulator, so that the later loads hit L1/L2 instead of starting their DRAM/L2 round trip after the wait
// (profiles/stall_report_r1.md: 10-22 % of the GEMM / dgrad stall samples sat on exactly those loads).
DTM_DEVICE void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// LRN's s^(-3/4) for s >= 1 (tf.nn.lrn with beta = 0.75, bias 1): r = s^(-1/2) from MUFU.RSQ, then r * sqrt(r) with
// MUFU.SQRT - two special-function instructions and one multiply where __powf costs lg2 + mul + ex2 and range fix-ups;
// r * r is 1 / s for the backward pass, which would otherwise divide.
DTM_DEVICE float lrn_pow_m075(float s, float* inv_s = nullptr) {
  float r, q;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(q) : "f"(r));
  if (inv_s) *inv_s = r * r;
  return r * q;
}

// LRN's channel window: out[j] = ext[j] + ... + ext[j + 8] for N consecutive channels whose values (squares, or the
// backward pass's t terms) sit in ext[4 + j] with four neighbours (zero outside the tensor) on either side.  Running prefix
// sums make that one addition and one subtraction per channel instead of eight additions; the prefix holds at most N + 8
// terms of one sign class of magnitude, so the difference is exact to a few ulp of the largest prefix.
template <int N>
DTM_DEVICE void window9_sums(const float (&ext)[N + 8], float (&out)[N]) {
  float p[N + 9];
  p[0] = 0.0f;
#pragma unroll
  for (int k = 0; k < N + 8; ++k) p[k + 1] = p[k] + ext[k];
#pragma unroll
  for (int j = 0; j < N; ++j) out[j] = p[j + 9] - p[j];
}

DTM_DEVICE float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

DTM_DEVICE void red_add_f32(float* addr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
DTM_DEVICE void red_add_v4_f32(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

}  // namespace dtm
