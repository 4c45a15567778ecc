// Single-GPU optimizer step: one kernel replaces the reference's per-variable Apply<Opt> ops, the l2 weight-decay
// gradient, the staircase learning-rate schedule, the EMA(0.9999) shadow update, the loss moving averages and the
// global_step increment (mnist_input.py:246-270, :288-290), and refreshes the bf16 shadow weights that the
// tensor-core kernels read.  All state is device resident, so the step needs no host involvement.
// (With more than one GPU the same math runs inside the fused all-reduce kernel: csrc/comm/.)
#include <stdlib.h>
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "update_math.cuh"

namespace dtm {

// Grid-stride form (every CTA owns a fixed 1/grid share).  Default: measured faster than the chunk dispenser below both
// stand-alone (18.7 vs 20.8 us uncapped, 26.9 vs 29.8 us at one CTA per SM) and inside the step.
__global__ void __launch_bounds__(512)
fused_update_stride_kernel(const dtm_update_config cfg, dtm_train_state* __restrict__ st, float* __restrict__ w,
                    float* __restrict__ grad, float* __restrict__ slot0, float* __restrict__ slot1,
                    float* __restrict__ ema, __nv_bfloat16* __restrict__ w_bf16, float grad_scale, int lo, int hi,
                    int do_finish) {
  const UpdateScalars u = make_update_scalars(cfg, st, grad_scale);
  const int lo4 = lo >> 2, n4 = hi >> 2;
  __shared__ float s_wl[DTM_MAX_SEGMENTS];
  if (threadIdx.x < DTM_MAX_SEGMENTS) s_wl[threadIdx.x] = 0.0f;
  __syncthreads();
  int cur_seg = -1;
  float cur_sq = 0.0f;
  int seg = 0;
  for (int i4 = lo4 + blockIdx.x * blockDim.x + threadIdx.x; i4 < n4; i4 += gridDim.x * blockDim.x) {
    const int i = i4 << 2;
    seg = segment_advance(cfg, seg, i);         // segment offsets are multiples of 64: a float4 never straddles
    const float wd = cfg.seg_wd[seg];
    const float4 wv = *reinterpret_cast<const float4*>(w + i);
    const float4 gv = *reinterpret_cast<const float4*>(grad + i);
    if (wd != 0.0f) {
      if (seg != cur_seg) {
        if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
        cur_seg = seg;
        cur_sq = 0.0f;
      }
      cur_sq += wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    }
    const float4 o = update_vec4(cfg, u, wv, gv, wd, slot0, slot1, ema, i);
    *reinterpret_cast<float4*>(w + i) = o;
    if (cfg.seg_zero_grad[seg]) *reinterpret_cast<float4*>(grad + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (w_bf16) *reinterpret_cast<uint2*>(w_bf16 + i) = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
    if (seg == cfg.pack_seg) pack_conv2_vec4(cfg, i - cfg.seg_offset[seg], o);
  }
  if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
  __syncthreads();
  if (threadIdx.x < cfg.n_segments && s_wl[threadIdx.x] != 0.0f) atomicAdd(&st->wl_sum[threadIdx.x], s_wl[threadIdx.x]);
  // last block to finish closes the step.  A launch that does not close it (the early update of the fc bucket) takes no
  // ticket: with the full grid that was 1184 returning atomics on one address, ~14 ns apiece in the L2, during which the
  // finished CTAs kept their SM slots and the conv backward kernel queued behind them could not start (reference preset:
  // the update spanned 48 us on the timeline for 6 us of memory traffic).
  if (!do_finish) return;
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {   // the whole CTA: finish_step spreads the bookkeeping over its threads
    __threadfence();
    if (threadIdx.x == 0) st->ticket = 0;
    finish_step(cfg, st, u.lr_plain);
  }
}

// (DTM_UPDATE_DYNAMIC=1)  Work is handed out in chunks of 128 float4 per warp from a counter in the training state instead of a fixed grid
// stride: the early update of the fc bucket runs BESIDE the conv backward kernels with however many CTAs fit next to them
// (one per SM beside conv2_dgrad, more once it has drained), and a CTA that becomes resident late simply takes what is
// left - no CTA owns a fixed share that would make it the straggler.  Each lane keeps two float4 elements (10 16-byte
// loads) in flight; 64 registers, so that a 256-thread CTA still fits next to a 384 x 128-register conv2_dgrad CTA.
constexpr int UPD_CHUNK4 = 128;   // float4 per warp chunk: 32 lanes x 2 x 2

__global__ void __launch_bounds__(256, 4)
fused_update_kernel(const dtm_update_config cfg, dtm_train_state* __restrict__ st, float* __restrict__ w,
                    float* __restrict__ grad, float* __restrict__ slot0, float* __restrict__ slot1,
                    float* __restrict__ ema, __nv_bfloat16* __restrict__ w_bf16, float grad_scale, int lo, int hi,
                    int do_finish) {
  const UpdateScalars u = make_update_scalars(cfg, st, grad_scale);
  const int lo4 = lo >> 2, n4 = hi >> 2;
  __shared__ float s_wl[DTM_MAX_SEGMENTS];
  if (threadIdx.x < DTM_MAX_SEGMENTS) s_wl[threadIdx.x] = 0.0f;
  __syncthreads();
  int cur_seg = -1;
  float cur_sq = 0.0f;
  int seg = 0;
  const uint32_t lane = lane_id();
  // one element: segment look-up, weight-decay sum of squares, optimizer + EMA, bf16 shadow, operand images
  auto apply = [&](int i, const float4 wv, const float4 gv) {
    seg = segment_advance(cfg, seg, i);         // segment offsets are multiples of 64: a float4 never straddles
    const float wd = cfg.seg_wd[seg];
    if (wd != 0.0f) {
      if (seg != cur_seg) {
        if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
        cur_seg = seg;
        cur_sq = 0.0f;
      }
      cur_sq += wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    }
    const float4 o = update_vec4(cfg, u, wv, gv, wd, slot0, slot1, ema, i);
    *reinterpret_cast<float4*>(w + i) = o;
    if (cfg.seg_zero_grad[seg]) *reinterpret_cast<float4*>(grad + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (w_bf16) *reinterpret_cast<uint2*>(w_bf16 + i) = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
    if (seg == cfg.pack_seg) pack_conv2_vec4(cfg, i - cfg.seg_offset[seg], o);
  };
  for (;;) {
    unsigned int c = 0;
    if (lane == 0) c = atomicAdd(&st->work, 1u);
    c = __shfl_sync(0xffffffffu, c, 0);
    const long long base = static_cast<long long>(lo4) + static_cast<long long>(c) * UPD_CHUNK4;
    if (base >= n4) break;                       // chunk indices only grow: every later claim is past the end too
    const int b4 = static_cast<int>(base);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int i4a = b4 + h * 64 + lane, i4b = i4a + 32;
      const bool oka = i4a < n4, okb = i4b < n4;
      float4 wa, ga, wb, gb;
      if (oka) { wa = *reinterpret_cast<const float4*>(w + (i4a << 2)); ga = *reinterpret_cast<const float4*>(grad + (i4a << 2)); }
      if (okb) { wb = *reinterpret_cast<const float4*>(w + (i4b << 2)); gb = *reinterpret_cast<const float4*>(grad + (i4b << 2)); }
      if (oka) apply(i4a << 2, wa, ga);
      if (okb) apply(i4b << 2, wb, gb);
    }
  }
  if (cur_seg >= 0) atomicAdd(&s_wl[cur_seg], cur_sq);
  __syncthreads();
  if (threadIdx.x < cfg.n_segments && s_wl[threadIdx.x] != 0.0f) atomicAdd(&st->wl_sum[threadIdx.x], s_wl[threadIdx.x]);
  // last block to finish closes the step
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {   // the whole CTA: finish_step spreads the bookkeeping over its threads
    __threadfence();
    if (threadIdx.x == 0) { st->ticket = 0; st->work = 0; }
    if (do_finish) finish_step(cfg, st, u.lr_plain);
  }
}

__global__ void state_init_kernel(dtm_train_state* st, long long step, float b1, float b2) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    dtm_train_state z = {};
    z.global_step = step;
    // TF keeps beta^t as variables initialised to beta and multiplied once per step: after `step` steps beta^(step+1)
    z.beta1_power = powf(b1, static_cast<float>(step + 1));
    z.beta2_power = powf(b2, static_cast<float>(step + 1));
    *st = z;
  }
}

}  // namespace dtm

extern "C" int dtm_fused_update_range(const dtm_update_config* cfg, dtm_train_state* state, float* w, float* grad, float* slot0,
                                      float* slot1, float* ema, void* w_bf16, float grad_scale, int lo, int hi, int do_finish,
                                      int max_ctas, void* stream) {
  using namespace dtm;
  if (!cfg || !state || !w || !grad) return dtm_fail("fused_update: null argument");
  if (cfg->n_segments < 1 || cfg->n_segments > DTM_MAX_SEGMENTS) return dtm_fail("fused_update: bad segment count");
  if (cfg->total % 64 != 0) return dtm_fail("fused_update: total must be a multiple of 64");
  for (int s = 0; s <= cfg->n_segments; ++s)
    if (cfg->seg_offset[s] % 64 != 0) return dtm_fail("fused_update: segment offsets must be multiples of 64");
  if (lo < 0 || hi > cfg->total || lo >= hi || lo % 64 || hi % 64) return dtm_fail("fused_update: bad element range");
  if (cfg->optimizer == DTM_OPT_ADAM && (!slot0 || !slot1)) return dtm_fail("fused_update: Adam needs both slots");
  if (cfg->optimizer == DTM_OPT_MOMENTUM && !slot0) return dtm_fail("fused_update: Momentum needs slot0");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int n4 = (hi - lo) / 4;
  const char* dyn = getenv("DTM_UPDATE_DYNAMIC");
  const bool dynamic = dyn && dyn[0] == '1';
  int grid = dynamic ? min((n4 + 2 * UPD_CHUNK4 - 1) / (2 * UPD_CHUNK4), 4 * device_sm_count()) : min((n4 + 255) / 256, 8 * device_sm_count());
  if (max_ctas > 0) grid = min(grid, max_ctas);
  if (grid < 1) grid = 1;
  if (cfg->pack_seg >= cfg->n_segments || (cfg->pack_seg >= 0 && (!cfg->pack_fwd || !cfg->pack_dgrad)))
    return dtm_fail("fused_update: pack_seg set without both pack buffers");
  // threads per CTA of the capped (max_ctas) grid-stride form: more warps = more loads in flight next to the conv backward
  // kernels; 384 x 40 registers still fits beside a 384 x 128-register conv2_dgrad CTA (DTM_UPDATE_THREADS=256|384|512)
  int threads = 256;
  if (!dynamic && max_ctas > 0) {
    const char* te = getenv("DTM_UPDATE_THREADS");
    const int tv = te ? atoi(te) : 384;   // same box: 256 -> 153.0, 384 -> 147.8, 512 -> 147.9 us per step
    threads = (tv == 384 || tv == 512) ? tv : 256;
  }
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(grid);
  lc.blockDim = dim3(threads);
  lc.stream = s;
  cudaLaunchAttribute attr[1];
  if (const L2Window* win = update_l2_window()) {   // optimizer state stays in the persisting L2 between steps
    attr[0].id = cudaLaunchAttributeAccessPolicyWindow;
    attr[0].val.accessPolicyWindow.base_ptr = win->base;
    attr[0].val.accessPolicyWindow.num_bytes = win->bytes;
    attr[0].val.accessPolicyWindow.hitRatio = win->hit_ratio;
    attr[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    lc.attrs = attr;
    lc.numAttrs = 1;
  }
  __nv_bfloat16* wb = static_cast<__nv_bfloat16*>(w_bf16);
  if (dynamic) DTM_CUDA_OK(cudaLaunchKernelEx(&lc, fused_update_kernel, *cfg, state, w, grad, slot0, slot1, ema, wb, grad_scale, lo, hi, do_finish));
  else         DTM_CUDA_OK(cudaLaunchKernelEx(&lc, fused_update_stride_kernel, *cfg, state, w, grad, slot0, slot1, ema, wb, grad_scale, lo, hi, do_finish));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_fused_update(const dtm_update_config* cfg, dtm_train_state* state, float* w, float* grad, float* slot0,
                                float* slot1, float* ema, void* w_bf16, float grad_scale, void* stream) {
  if (!cfg) return dtm::dtm_fail("fused_update: null argument");
  return dtm_fused_update_range(cfg, state, w, grad, slot0, slot1, ema, w_bf16, grad_scale, 0, cfg->total, 1, 0, stream);
}

extern "C" int dtm_state_init(dtm_train_state* state, long long global_step, float beta1, float beta2, void* stream) {
  using namespace dtm;
  state_init_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(state, global_step, beta1, beta2);
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
