// Peer-memory communication over NVLink 5 / NVSwitch — the replacement for the reference's gRPC parameter-server
// traffic (variable pulls main.py:80-93, gradient pushes + Apply ops mnist_input.py:262-264, global_step sync,
// SURVEY §2f X1-X5).  No NCCL on this path: kernels load and store peer GPUs' memory directly.
//
//   symmetric memory   every rank cudaMalloc's the same buffers and exchanges CUDA IPC handles once; a kernel gets
//                      the N peer base pointers by value (dtm_peer_ctx).
//   cross-GPU barrier  flags[N] per rank in symmetric memory; rank r publishes a monotonically increasing epoch into
//                      slot r of every peer (st.release.sys) and waits for all its own slots (ld.acquire.sys).
//                      Inside a kernel CTA 0 does the NVLink part, the other CTAs wait on a local flag; kernels are
//                      launched cooperatively so that all CTAs are co-resident.  Every spin has a watchdog.
//   fused step         [barrier] -> reduce-scatter (each rank sums its 1/N shard of the flat gradient from all peers,
//                      16-byte P2P loads, fixed rank order => bit-identical everywhere) -> 1/N scale + weight decay +
//                      SGD/Momentum/Adam + EMA on the shard *in registers* (optimizer state is sharded) ->
//                      all-gather of the updated fp32 weights by P2P stores into every peer -> [barrier] ->
//                      local pass: bf16 shadow weights, clear accumulated gradients, step bookkeeping.
//                      Link bytes per GPU and direction: (N-1)/N * S in (gradients) and out (weights).
//   all-reduce         stand-alone fp32 sum on a symmetric buffer: one-shot (every rank reads everything; latency
//                      optimal for small messages) or two-shot (reduce-scatter + all-gather; bandwidth optimal).
//   broadcast          every rank pulls the root's buffer (initial parameters, restored checkpoints, global_step).
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../optim/update_math.cuh"
#include "../../include/dtm_b200.h"

namespace dtm {

struct CommCtrl {          // rank-local control block (device memory, zero-initialised once)
  unsigned int epoch;      // last barrier epoch used
  unsigned int arrive;     // intra-kernel CTA arrival counter
  unsigned int go;         // intra-kernel release flag (epoch value)
  unsigned int ticket;     // last-CTA election
};

DTM_DEVICE unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DTM_DEVICE void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DTM_DEVICE unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DTM_DEVICE void st_release_gpu(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DTM_DEVICE void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peer loads/stores: system scope so that they are never served from a stale L1 line
DTM_DEVICE float4 ld_peer_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
DTM_DEVICE void st_peer_v4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVLS (NVLink SHARP) through a multicast address: one load that the NVSwitch answers with the SUM of every rank's copy,
// one store that the switch replicates into every rank's copy.
DTM_DEVICE float4 mc_ld_reduce_v4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
DTM_DEVICE void mc_st_v4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
DTM_DEVICE void mc_st_bf16x4(__nv_bfloat16* p, uint32_t a, uint32_t b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
DTM_DEVICE void mc_st_bf16x8(__nv_bfloat16* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Ranks may reach a collective seconds apart (start-up, graph capture), so cross-GPU waits get a longer leash than
// the intra-kernel pipelines: ~2 minutes of SM clocks.
__device__ long long g_comm_watchdog_cycles = 240000000000ll;   // dtm_set_comm_watchdog() overrides it (tests: seconds, not minutes)
#define COMM_WATCHDOG_CYCLES g_comm_watchdog_cycles

DTM_DEVICE void spin_until_ge_sys(const unsigned int* p, unsigned int target, int tag) {
  const long long t0 = clock64();
  while (static_cast<int>(ld_acquire_sys(p) - target) < 0)
    if (clock64() - t0 > COMM_WATCHDOG_CYCLES) watchdog_trap("peer barrier (a rank did not arrive)", tag, (int)target);
}
DTM_DEVICE void spin_until_ge_gpu(const unsigned int* p, unsigned int target, int tag) {
  const long long t0 = clock64();
  while (static_cast<int>(ld_acquire_gpu(p) - target) < 0)
    if (clock64() - t0 > COMM_WATCHDOG_CYCLES) watchdog_trap("grid barrier", tag, (int)target);
}

// Two rendezvous flavours over the same epoch flags (`k` strictly increasing; slot r of every rank is written by rank r).
//
// barrier_ready  - "everything the kernels in front of this one wrote on rank r is visible": nothing of THIS kernel has to
//                  be waited for, so CTA 0 publishes at once and every CTA polls its own rank's flags (local memory) by
//                  itself - one NVLink store latency, no grid-wide arrive, no release broadcast.
// barrier_done   - "every CTA of this kernel on every rank has finished its peer loads and stores": CTAs count in, the
//                  last one publishes to the peers, and again every CTA polls the local flags itself.
// Kernels are launched cooperatively so that all CTAs are co-resident while they spin.
DTM_DEVICE void barrier_ready_publish(const dtm_peer_ctx& ctx, unsigned int k) {
  if (blockIdx.x == 0 && threadIdx.x < ctx.world) st_release_sys(ctx.flags[threadIdx.x] + ctx.rank, k);
}
DTM_DEVICE void barrier_ready_wait(const dtm_peer_ctx& ctx, unsigned int k) {
  if (threadIdx.x < ctx.world) spin_until_ge_sys(ctx.flags[ctx.rank] + threadIdx.x, k, threadIdx.x);
  __syncthreads();
}
DTM_DEVICE void barrier_ready(const dtm_peer_ctx& ctx, unsigned int k) {
  barrier_ready_publish(ctx, k);
  barrier_ready_wait(ctx, k);
}
// The last CTA of the grid to get here publishes epoch `k` to every peer (all CTAs call it; no waiting).
DTM_DEVICE void publish_when_grid_arrived(const dtm_peer_ctx& ctx, CommCtrl* ctrl, unsigned int k) {
  __shared__ unsigned int s_publish;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    unsigned int prev;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(prev) : "l"(&ctrl->arrive) : "memory");
    s_publish = prev == gridDim.x - 1 ? 1u : 0u;
    if (s_publish) ctrl->arrive = 0;   // nobody counts in again before the peers have answered this epoch
  }
  __syncthreads();
  if (s_publish && threadIdx.x < ctx.world) st_release_sys(ctx.flags[threadIdx.x] + ctx.rank, k);
}
DTM_DEVICE void barrier_done(const dtm_peer_ctx& ctx, CommCtrl* ctrl, unsigned int k) {
  publish_when_grid_arrived(ctx, ctrl, k);
  if (threadIdx.x < ctx.world) spin_until_ge_sys(ctx.flags[ctx.rank] + threadIdx.x, k, threadIdx.x);
  __syncthreads();
}

// Sum of one float4 over all ranks, fixed rank order (bit-identical whichever rank computes it).  The compile-time
// rank count lets every peer load be issued before the first add: N-1 NVLink round trips overlap instead of chaining.
template <int NR>
DTM_DEVICE float4 reduce_peers_fixed(float* const* bufs, int i) {
  float4 v[NR];
#pragma unroll
  for (int j = 0; j < NR; ++j) v[j] = ld_peer_v4(bufs[j] + i);
  float4 a = v[0];
#pragma unroll
  for (int j = 1; j < NR; ++j) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
  return a;
}
DTM_DEVICE float4 reduce_peers(float* const* bufs, int world, int i) {
  switch (world) {
    case 2: return reduce_peers_fixed<2>(bufs, i);
    case 4: return reduce_peers_fixed<4>(bufs, i);
    case 8: return reduce_peers_fixed<8>(bufs, i);
    default: break;
  }
  float4 a = ld_peer_v4(bufs[0] + i);
#pragma unroll 1
  for (int j = 1; j < world; ++j) {
    const float4 v = ld_peer_v4(bufs[j] + i);
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  return a;
}

DTM_DEVICE void shard_range(int lo, int hi, int rank, int world, int& s_lo, int& s_hi) {
  const int n = hi - lo;
  int per = (n + world - 1) / world;
  per = (per + 63) / 64 * 64;
  s_lo = min(hi, lo + rank * per);
  s_hi = min(hi, s_lo + per);
}

// ---------------------------------------------------------------------------------------------------------
// Two instantiations.  NVLS: 256 threads and at most 64 registers (launch bounds 256 x 4), one CTA per SM - small enough to
// sit on an SM NEXT TO a conv-backward CTA (conv2_dgrad: 384 threads x 128 registers), so the gradient exchange of the fc
// bucket really runs under the conv backward instead of taking SMs away from it; the in-switch reduction needs few loads in
// flight (8-GPU sweep: 32 CTAs reach the same time as 148).  P2P: 512 threads, N-1 peer loads in flight per thread.
// (NVLS with LIGHT = false: 512 threads and four reductions in flight per thread - for shards of millions of elements,
// where the rank-local optimizer traffic needs the memory parallelism: wide preset, 2 GPUs: 741 -> ~600 us per step.)
template <bool NVLS, bool LIGHT>
__global__ void __launch_bounds__(LIGHT ? 256 : 512, LIGHT ? 4 : 1)
fused_allreduce_update_kernel(const dtm_peer_ctx ctx, const dtm_update_config cfg, dtm_train_state* __restrict__ st,
                              int lo, int hi, float* __restrict__ slot0, float* __restrict__ slot1,
                              float* __restrict__ ema, __nv_bfloat16* __restrict__ w_bf16, CommCtrl* __restrict__ ctrl,
                              int do_finish, int depth) {
  const int N = ctx.world, r = ctx.rank;
  pdl_wait();                 // launched with a programmatic edge: the launch latency hides under the preceding kernel's tail
  pdl_launch_dependents();
  const unsigned int e0 = ld_acquire_gpu(&ctrl->epoch);
  // "My gradients are complete" may be told to the peers at once - unless this launch first sums squares of weights
  // (weight-decay loss terms): a peer that has seen the flag may already multicast NEW weights into the very copies that
  // sum is reading, so then the flag goes out only after every CTA has finished the sum.
  bool prepass = false;
  for (int s = 0; s < cfg.n_segments; ++s)
    prepass = prepass || (cfg.seg_wd[s] != 0.0f && max(lo, cfg.seg_offset[s]) < min(hi, cfg.seg_offset[s + 1]));
  if (!prepass) barrier_ready_publish(ctx, e0 + 1);
  const UpdateScalars u = make_update_scalars(cfg, st, 1.0f / static_cast<float>(N));
  float* w_local = ctx.w[r];
  float* g_local = ctx.grad[r];
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;

  // ---- while peers are still finishing their backward pass: sum of squares of the decayed variables (the
  //      l2_loss terms of the loss value use the weights of the step in flight, mnist_input.py:113)
  {
    __shared__ float s_wl[DTM_MAX_SEGMENTS];
    if (threadIdx.x < DTM_MAX_SEGMENTS) s_wl[threadIdx.x] = 0.0f;
    __syncthreads();
    bool any = false;
    for (int s = 0; s < cfg.n_segments; ++s) {
      if (cfg.seg_wd[s] == 0.0f) continue;
      const int a = max(lo, cfg.seg_offset[s]), b = min(hi, cfg.seg_offset[s + 1]);
      if (a >= b) continue;
      any = true;
      float sq = 0.0f;
      if (cfg.seg_gather_bf16[s]) {
        // fp32 masters of these variables are only current on their owners; the bf16 shadow is current everywhere
        // (this sum only feeds the *reported* loss value - the decay gradient itself uses the owner's fp32 weights)
        for (int i = a + tid * 4; i < b; i += nthreads * 4) {
          const uint2 v = *reinterpret_cast<const uint2*>(w_bf16 + i);
          const float x0 = bf16_lo(v.x), x1 = bf16_hi(v.x), x2 = bf16_lo(v.y), x3 = bf16_hi(v.y);
          sq += x0 * x0 + x1 * x1 + x2 * x2 + x3 * x3;
        }
      } else {
        for (int i = a + tid * 4; i < b; i += nthreads * 4) {
          const float4 v = *reinterpret_cast<const float4*>(w_local + i);
          sq += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
      }
      sq = warp_sum(sq);
      if (lane_id() == 0 && sq != 0.0f) atomicAdd(&s_wl[s], sq);
    }
    __syncthreads();
    if (any && threadIdx.x < cfg.n_segments && s_wl[threadIdx.x] != 0.0f) atomicAdd(&st->wl_sum[threadIdx.x], s_wl[threadIdx.x]);
  }

  // ---- barrier A: every rank's gradients are complete
  if (prepass) publish_when_grid_arrived(ctx, ctrl, e0 + 1);
  barrier_ready_wait(ctx, e0 + 1);

  // ---- reduce-scatter + optimizer on my shard + all-gather of the new weights
  int s_lo, s_hi;
  shard_range(lo, hi, r, N, s_lo, s_hi);
  constexpr bool nvls = NVLS;
  int seg = 0;
  // optimizer + EMA on one float4 of my shard, then publish it: multicast store (NVLS) or N-1 peer stores (P2P).  With
  // `defer_bf16` the packed bf16 value of a tensor-core-only variable is handed back instead of stored (the NVLS loop pairs
  // two neighbouring float4 into one 16-byte multicast store).
  auto apply = [&](int i, const float4 g, bool defer_bf16, uint2& pk_out) -> bool {
    seg = segment_advance(cfg, seg, i);
    const float wd = cfg.seg_wd[seg];
    const float4 wv = *reinterpret_cast<const float4*>(w_local + i);
    const float4 o = update_vec4(cfg, u, wv, g, wd, slot0, slot1, ema, i);
    *reinterpret_cast<float4*>(w_local + i) = o;
    if (cfg.seg_gather_bf16[seg]) {
      // tensor-core-only variable: peers need the bf16 value, nothing else (half the link bytes, no conversion pass)
      const uint2 pk = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
      if (nvls) {
        if (defer_bf16) { pk_out = pk; return true; }
        mc_st_bf16x4(static_cast<__nv_bfloat16*>(ctx.mc_w_bf16) + i, pk.x, pk.y);
      } else {
#pragma unroll 1
        for (int j = 0; j < N; ++j) {   // start with this rank itself, then walk the ring
          const int p = (r + j) % N;
          asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1,%2};" ::"l"(static_cast<__nv_bfloat16*>(ctx.w_bf16[p]) + i), "r"(pk.x), "r"(pk.y) : "memory");
        }
      }
    } else if (nvls) {
      mc_st_v4(ctx.mc_w + i, o);
    } else {
#pragma unroll 1
      for (int j = 1; j < N; ++j) {   // start with the next rank so that the N ranks do not all hit the same peer
        const int p = (r + j) % N;
        st_peer_v4(ctx.w[p] + i, o);
      }
    }
    return false;
  };
  const int step = nthreads * 4;
  if (nvls) {
    // `depth` (<= U) in-switch reductions in flight per thread before the first one is consumed.  Consecutive threads take
    // consecutive float4: a warp's multimem.ld_reduce covers 512 contiguous bytes.  (Measured in round 2: giving each thread
    // 8 consecutive elements - two 16-byte reductions 32 bytes apart per lane, for the sake of one 16-byte bf16 multicast
    // store - nearly halves the switch throughput: wide exchange 207 -> 342 us at 8 GPUs.  Reverted.)
    constexpr int U = LIGHT ? 2 : 4;
    depth = LIGHT ? (depth > U ? U : depth) : U;
    for (int base = s_lo + tid * 4; base < s_hi; base += step * depth) {
      float4 g[U];
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (k < depth && base + k * step < s_hi) g[k] = mc_ld_reduce_v4(ctx.mc_grad + base + k * step);
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (k < depth && base + k * step < s_hi) { uint2 unused; apply(base + k * step, g[k], false, unused); }
    }
  } else if (!LIGHT && (N == 8 || N == 4 || N == 2)) {
    // two float4 per thread and iteration: 2(N-1) NVLink round trips overlap, and the second one hides the optimizer
    // math of the first
    for (int base = s_lo + tid * 4; base < s_hi; base += step * 2) {
      const bool two = base + step < s_hi;
      const float4 g0 = reduce_peers(ctx.grad, N, base);
      float4 g1 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (two) g1 = reduce_peers(ctx.grad, N, base + step);
      uint2 unused;
      apply(base, g0, false, unused);
      if (two) apply(base + step, g1, false, unused);
    }
  } else {
    // software-pipelined: the peer loads of the next visit are in flight while this one is updated (a visit is one NVLink
    // round trip + the optimizer's local loads; with few visits per thread - 5 at two ranks - neither hides the other)
    int i = s_lo + tid * 4;
    float4 g_next = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < s_hi) g_next = reduce_peers(ctx.grad, N, i);
    for (; i < s_hi; i += step) {
      const float4 g = g_next;
      if (i + step < s_hi) g_next = reduce_peers(ctx.grad, N, i + step);
      uint2 unused;
      apply(i, g, false, unused);
    }
  }

  // ---- barrier B: all shards of the new weights have landed everywhere, nobody reads my gradients any more
  barrier_done(ctx, ctrl, e0 + 2);

  // ---- local pass, only over the variables that were gathered in fp32 (the small ones): bf16 shadow, operand images,
  //      and clearing of gradients that are accumulated with atomics
  for (int s = 0; s < cfg.n_segments; ++s) {
    const int a = max(lo, cfg.seg_offset[s]), b = min(hi, cfg.seg_offset[s + 1]);
    if (a >= b) continue;
    const bool convert = !cfg.seg_gather_bf16[s];
    const bool zero = cfg.seg_zero_grad[s] != 0;
    if (!convert && !zero) continue;
    for (int i = a + tid * 4; i < b; i += nthreads * 4) {
      if (convert) {
        const float4 v = *reinterpret_cast<const float4*>(w_local + i);
        *reinterpret_cast<uint2*>(w_bf16 + i) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
        if (s == cfg.pack_seg) pack_conv2_vec4(cfg, i - cfg.seg_offset[s], v);
      }
      if (zero) *reinterpret_cast<float4*>(g_local + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }

  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(&ctrl->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {   // the whole CTA: finish_step spreads the bookkeeping over its threads
    __threadfence();
    if (threadIdx.x == 0) {
      ctrl->ticket = 0;
      st_release_gpu(&ctrl->epoch, e0 + 2);
    }
    if (do_finish) finish_step(cfg, st, u.lr_plain);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Replicated one-shot update for SMALL buckets (the conv variables: 52 K elements closing every step).  The fused kernel
// above pays a chain of dependent latencies at the end of the step - reduce-scatter, all-gather, a second cross-GPU barrier
// that waits for every remote store - about 30 us for 200 KB.  Here every rank reads ALL peers' gradients itself (N-1 x
// 200 KB over NVLink, fixed rank order: bit-identical sums everywhere) and applies the whole update to its own copy; the
// optimizer state of the bucket is replicated.  No all-gather, nothing remote is written, ONE barrier ("the peers' gradients
// are complete"), and no CTA waits for another CTA, so the launch is an ordinary one.
//
// Gradients of this bucket are double-buffered by step parity: a rank may still be reading my buffer of step k when I
// finish, so step k+1 accumulates into the other buffer, and the buffer of step k is cleared during step k+1 - after that
// step's barrier, which no peer can pass before it has left step k's kernel.
__global__ void __launch_bounds__(256, 4)
replicated_update_kernel(const dtm_peer_ctx ctx, const dtm_update_config cfg, dtm_train_state* __restrict__ st, int lo, int hi,
                         float* const* __restrict__ grad_now, float* __restrict__ grad_prev_local, float* __restrict__ slot0,
                         float* __restrict__ slot1, float* __restrict__ ema, __nv_bfloat16* __restrict__ w_bf16,
                         CommCtrl* __restrict__ ctrl, int do_finish) {
  const int N = ctx.world, r = ctx.rank;
  pdl_wait();
  pdl_launch_dependents();
  const unsigned int e0 = ld_acquire_gpu(&ctrl->epoch);
  barrier_ready_publish(ctx, e0 + 1);
  const UpdateScalars u = make_update_scalars(cfg, st, 1.0f / static_cast<float>(N));
  float* w_local = ctx.w[r];
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  __shared__ float s_wl[DTM_MAX_SEGMENTS];
  __shared__ float* s_src[DTM_MAX_RANKS];
  if (threadIdx.x < DTM_MAX_SEGMENTS) s_wl[threadIdx.x] = 0.0f;
  if (threadIdx.x < N) s_src[threadIdx.x] = grad_now[threadIdx.x];
  barrier_ready_wait(ctx, e0 + 1);      // ends with __syncthreads
  int seg = 0, cur_seg = -1;
  float cur_sq = 0.0f;
  for (int i = lo + tid * 4; i < hi; i += nthreads * 4) {
    const float4 g = reduce_peers(s_src, N, i);
    seg = segment_advance(cfg, seg, i);
    const float wd = cfg.seg_wd[seg];
    const float4 wv = *reinterpret_cast<const float4*>(w_local + i);
    if (wd != 0.0f) {
      if (seg != cur_seg) {
        if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
        cur_seg = seg;
        cur_sq = 0.0f;
      }
      cur_sq += wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    }
    const float4 o = update_vec4(cfg, u, wv, g, wd, slot0, slot1, ema, i);
    *reinterpret_cast<float4*>(w_local + i) = o;
    *reinterpret_cast<uint2*>(w_bf16 + i) = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
    if (seg == cfg.pack_seg) pack_conv2_vec4(cfg, i - cfg.seg_offset[seg], o);
    // last step's buffer: every peer has finished with it (see above); the next step accumulates into it again
    if (grad_prev_local) *reinterpret_cast<float4*>(grad_prev_local + i) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
  __syncthreads();
  if (threadIdx.x < cfg.n_segments && s_wl[threadIdx.x] != 0.0f) atomicAdd(&st->wl_sum[threadIdx.x], s_wl[threadIdx.x]);
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(&ctrl->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {
    __threadfence();
    if (threadIdx.x == 0) {
      ctrl->ticket = 0;
      st_release_gpu(&ctrl->epoch, e0 + 2);
    }
    if (do_finish) finish_step(cfg, st, u.lr_plain);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Stand-alone all-reduce (sum) of n floats at element offset `off` of the symmetric buffer `buf`.
//   mode 0 (one-shot): out[i] = sum_j buf_j[i], every rank reads all peers.
//   mode 1 (two-shot): shard owner reduces and writes the sum back into every rank's buf (in place); out = buf.
__global__ void __launch_bounds__(512, 1)
allreduce_kernel(const dtm_peer_ctx ctx, float* const* __restrict__ bufs, float* __restrict__ mc, int off, int n,
                 float* __restrict__ out, int mode, CommCtrl* __restrict__ ctrl) {
  const int N = ctx.world, r = ctx.rank;
  const unsigned int e0 = ld_acquire_gpu(&ctrl->epoch);
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  barrier_ready(ctx, e0 + 1);
  const int step = nthreads * 4;
  if (mode == 0) {
    if (mc) {
      // NVLS one-shot: every rank asks the switch for the sum of every element
      for (int base = tid * 4; base < n; base += step * 4) {
        float4 a[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) if (base + k * step < n) a[k] = mc_ld_reduce_v4(mc + off + base + k * step);
#pragma unroll
        for (int k = 0; k < 4; ++k) if (base + k * step < n) *reinterpret_cast<float4*>(out + base + k * step) = a[k];
      }
    } else {
      for (int i = tid * 4; i < n; i += step) {
        const float4 a = reduce_peers(bufs, N, off + i);
        *reinterpret_cast<float4*>(out + i) = a;
      }
    }
  } else {
    int s_lo, s_hi;
    shard_range(0, n, r, N, s_lo, s_hi);
    if (mc) {
      // NVLS two-shot: in-switch reduction of my shard, multicast store of the sums into every rank's buffer
      for (int base = s_lo + tid * 4; base < s_hi; base += step * 4) {
        float4 a[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) if (base + k * step < s_hi) a[k] = mc_ld_reduce_v4(mc + off + base + k * step);
#pragma unroll
        for (int k = 0; k < 4; ++k) if (base + k * step < s_hi) mc_st_v4(mc + off + base + k * step, a[k]);
      }
    } else {
      for (int i = s_lo + tid * 4; i < s_hi; i += step) {
        const float4 a = reduce_peers(bufs, N, off + i);
#pragma unroll 1
        for (int j = 0; j < N; ++j) st_peer_v4(bufs[(r + j) % N] + off + i, a);
      }
    }
  }
  barrier_done(ctx, ctrl, e0 + 2);
  if (mode == 1 && out != bufs[r] + off)
    for (int i = tid * 4; i < n; i += nthreads * 4)
      *reinterpret_cast<float4*>(out + i) = *reinterpret_cast<const float4*>(bufs[r] + off + i);
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) { __threadfence(); s_last = (atomicAdd(&ctrl->ticket, 1u) == gridDim.x - 1) ? 1u : 0u; }
  __syncthreads();
  if (s_last && threadIdx.x == 0) { ctrl->ticket = 0; st_release_gpu(&ctrl->epoch, e0 + 2); }
}

// Every rank copies n floats (and optionally the training state) from the root's symmetric buffer.
__global__ void __launch_bounds__(512, 1)
broadcast_kernel(const dtm_peer_ctx ctx, float* const* __restrict__ bufs, int off, int n, int root, CommCtrl* __restrict__ ctrl) {
  const int r = ctx.rank;
  const unsigned int e0 = ld_acquire_gpu(&ctrl->epoch);
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.x * blockDim.x;
  barrier_ready(ctx, e0 + 1);
  if (r != root)
    for (int i = tid * 4; i < n; i += nthreads * 4)
      *reinterpret_cast<float4*>(bufs[r] + off + i) = ld_peer_v4(bufs[root] + off + i);
  barrier_done(ctx, ctrl, e0 + 2);
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) { __threadfence(); s_last = (atomicAdd(&ctrl->ticket, 1u) == gridDim.x - 1) ? 1u : 0u; }
  __syncthreads();
  if (s_last && threadIdx.x == 0) { ctrl->ticket = 0; st_release_gpu(&ctrl->epoch, e0 + 2); }
}

// Cooperative launch (all CTAs co-resident: they spin on each other).  With `pdl` the launch also carries the programmatic
// stream-serialization attribute, so that the kernel is scheduled while its predecessor drains; should the driver refuse the
// combination, the first refusal switches it off for good.
template <typename Kernel, typename... Args>
static int launch_cooperative(Kernel kern, int grid, int block, cudaStream_t s, bool pdl, Args... args) {
  static bool pdl_ok = true;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (pdl && pdl_ok && pdl_enabled()) {
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.numAttrs = 2;
    if (cudaLaunchKernelEx(&cfg, kern, args...) == cudaSuccess) return 0;
    (void)cudaGetLastError();
    pdl_ok = false;
    cfg.numAttrs = 1;
  }
  DTM_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, args...));
  return 0;
}

static int comm_grid(size_t elems, int max_ctas = 0, int threads = 512) {
  // enough CTAs to keep 18 NVLinks busy, never more than one per SM (co-residency for the spin barriers)
  size_t want = (elems / 4 + threads - 1) / threads;
  int cap = device_sm_count();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (want < 1) want = 1;
  return static_cast<int>(want > static_cast<size_t>(cap) ? cap : want);
}

// in-flight multimem.ld_reduce per thread in the LIGHT NVLS instantiation: DTM_COMM_DEPTH = 1 | 2 (clamped; the heavy
// instantiation always keeps 4 in flight - round-1 8-GPU sweep: profiles/fused_exchange_8gpu_sweep_r1.json)
static int comm_depth() {
  const char* e = getenv("DTM_COMM_DEPTH");   // read per launch (cheap) so that tools can sweep it inside one process
  const int v = e ? atoi(e) : 2;
  return v < 1 ? 1 : (v > 2 ? 2 : v);
}

static int check_ctx(const dtm_peer_ctx* ctx) {
  if (!ctx) return dtm_fail("comm: null context");
  if (ctx->world < 1 || ctx->world > DTM_MAX_RANKS) return dtm_fail("comm: world size out of range");
  if (ctx->rank < 0 || ctx->rank >= ctx->world) return dtm_fail("comm: bad rank");
  return 0;
}

}  // namespace dtm

// ---- symmetric memory (CUDA IPC) ----------------------------------------------------------------------------
extern "C" int dtm_symm_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  using namespace dtm;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (!ptr || !handle64 || bytes == 0) return dtm_fail("symm_alloc: bad arguments");
  void* p = nullptr;
  DTM_CUDA_OK(cudaMalloc(&p, bytes));
  DTM_CUDA_OK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  DTM_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  *ptr = p;
  return 0;
}
extern "C" int dtm_symm_open(const unsigned char* handle64, void** peer_ptr) {
  using namespace dtm;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DTM_CUDA_OK(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int dtm_symm_close(void* peer_ptr) {
  using namespace dtm;
  DTM_CUDA_OK(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}
extern "C" int dtm_symm_free(void* ptr) {
  using namespace dtm;
  DTM_CUDA_OK(cudaFree(ptr));
  return 0;
}

extern "C" int dtm_set_comm_watchdog(long long cycles) {
  using namespace dtm;
  if (cycles <= 0) return dtm_fail("comm watchdog: cycles must be positive");
  DTM_CUDA_OK(cudaMemcpyToSymbol(g_comm_watchdog_cycles, &cycles, sizeof(cycles)));
  return 0;
}

// ---- collectives ------------------------------------------------------------------------------------------
extern "C" int dtm_fused_allreduce_update(const dtm_peer_ctx* ctx, const dtm_update_config* cfg, dtm_train_state* state,
                                          int lo, int hi, float* slot0, float* slot1, float* ema, void* w_bf16, void* ctrl,
                                          int do_finish, int max_ctas, void* stream) {
  using namespace dtm;
  if (int rc = check_ctx(ctx)) return rc;
  if (!cfg || !state || !ctrl || !w_bf16) return dtm_fail("fused_allreduce_update: null argument");
  if (lo < 0 || hi > cfg->total || lo >= hi || lo % 64 || hi % 64) return dtm_fail("fused_allreduce_update: bad element range");
  if (cfg->optimizer == DTM_OPT_ADAM && (!slot0 || !slot1)) return dtm_fail("fused_allreduce_update: Adam needs both slots");
  if (cfg->optimizer == DTM_OPT_MOMENTUM && !slot0) return dtm_fail("fused_allreduce_update: Momentum needs slot0");
  const bool nvls = ctx->mc_grad != nullptr && ctx->mc_w != nullptr && ctx->mc_w_bf16 != nullptr;
  // light variant while a rank's shard stays below 2 M elements (every preset but `wide`)
  const bool light = static_cast<size_t>(hi - lo) / ctx->world <= (2u << 20);
  if (nvls && light)
    return launch_cooperative(fused_allreduce_update_kernel<true, true>, comm_grid(static_cast<size_t>(hi - lo), max_ctas, 256), 256,
                              static_cast<cudaStream_t>(stream), true, *ctx, *cfg, state, lo, hi, slot0, slot1, ema,
                              static_cast<__nv_bfloat16*>(w_bf16), static_cast<CommCtrl*>(ctrl), do_finish, comm_depth());
  if (!nvls && light && ctx->world <= 4)      // small-footprint P2P form (co-resident with the conv backward, like NVLS light)
    return launch_cooperative(fused_allreduce_update_kernel<false, true>, comm_grid(static_cast<size_t>(hi - lo), max_ctas, 256), 256,
                              static_cast<cudaStream_t>(stream), true, *ctx, *cfg, state, lo, hi, slot0, slot1, ema,
                              static_cast<__nv_bfloat16*>(w_bf16), static_cast<CommCtrl*>(ctrl), do_finish, comm_depth());
  if (nvls)
    return launch_cooperative(fused_allreduce_update_kernel<true, false>, comm_grid(static_cast<size_t>(hi - lo), max_ctas), 512,
                              static_cast<cudaStream_t>(stream), true, *ctx, *cfg, state, lo, hi, slot0, slot1, ema,
                              static_cast<__nv_bfloat16*>(w_bf16), static_cast<CommCtrl*>(ctrl), do_finish, comm_depth());
  return launch_cooperative(fused_allreduce_update_kernel<false, false>, comm_grid(static_cast<size_t>(hi - lo), max_ctas), 512,
                            static_cast<cudaStream_t>(stream), true, *ctx, *cfg, state, lo, hi, slot0, slot1, ema,
                            static_cast<__nv_bfloat16*>(w_bf16), static_cast<CommCtrl*>(ctrl), do_finish, comm_depth());
}

extern "C" int dtm_replicated_allreduce_update(const dtm_peer_ctx* ctx, const dtm_update_config* cfg, dtm_train_state* state, int lo,
                                               int hi, float* const* grad_now_dev, float* grad_prev_local, float* slot0,
                                               float* slot1, float* ema, void* w_bf16, void* ctrl, int do_finish, void* stream) {
  using namespace dtm;
  if (int rc = check_ctx(ctx)) return rc;
  if (!cfg || !state || !ctrl || !w_bf16 || !grad_now_dev) return dtm_fail("replicated_update: null argument");
  if (lo < 0 || hi > cfg->total || lo >= hi || lo % 64 || hi % 64) return dtm_fail("replicated_update: bad element range");
  if (cfg->optimizer == DTM_OPT_ADAM && (!slot0 || !slot1)) return dtm_fail("replicated_update: Adam needs both slots");
  if (cfg->optimizer == DTM_OPT_MOMENTUM && !slot0) return dtm_fail("replicated_update: Momentum needs slot0");
  for (int sg = 0; sg < cfg->n_segments; ++sg)
    if (cfg->seg_gather_bf16[sg] && cfg->seg_offset[sg] < hi && cfg->seg_offset[sg + 1] > lo)
      return dtm_fail("replicated_update: the range holds a variable whose fp32 master is sharded");
  const int n4 = (hi - lo) / 4;
  const int grid = max(1, min((n4 + 255) / 256, device_sm_count()));
  DTM_CUDA_OK(launch_pdl(replicated_update_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), *ctx, *cfg, state, lo, hi,
                         grad_now_dev, grad_prev_local, slot0, slot1, ema, static_cast<__nv_bfloat16*>(w_bf16),
                         static_cast<CommCtrl*>(ctrl), do_finish));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_allreduce_f32(const dtm_peer_ctx* ctx, float* const* bufs_dev, float* mc_buf, int off, int n, float* out, int mode,
                                 void* ctrl, void* stream) {
  using namespace dtm;
  if (int rc = check_ctx(ctx)) return rc;
  if (n <= 0 || n % 4 || off % 4) return dtm_fail("allreduce: n and offset must be positive multiples of 4");
  if (mode != 0 && mode != 1) return dtm_fail("allreduce: mode 0 (one-shot) or 1 (two-shot)");
  const size_t work = mode == 0 ? static_cast<size_t>(n) : static_cast<size_t>(n) / ctx->world + 1;
  return launch_cooperative(allreduce_kernel, comm_grid(work), 512, static_cast<cudaStream_t>(stream), false, *ctx, bufs_dev, mc_buf, off, n,
                            out, mode, static_cast<CommCtrl*>(ctrl));
}

extern "C" int dtm_broadcast_f32(const dtm_peer_ctx* ctx, float* const* bufs_dev, int off, int n, int root, void* ctrl, void* stream) {
  using namespace dtm;
  if (int rc = check_ctx(ctx)) return rc;
  if (n <= 0 || n % 4 || off % 4) return dtm_fail("broadcast: n and offset must be positive multiples of 4");
  if (root < 0 || root >= ctx->world) return dtm_fail("broadcast: bad root");
  return launch_cooperative(broadcast_kernel, comm_grid(static_cast<size_t>(n)), 512, static_cast<cudaStream_t>(stream), false, *ctx,
                            bufs_dev, off, n, root, static_cast<CommCtrl*>(ctrl));
}
