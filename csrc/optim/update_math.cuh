// Per-element optimizer math shared by the single-GPU update kernel and the fused all-reduce+update kernel.
// Semantics follow the TF-1.x ops the reference builds (mnist_input.py:246-270):
//   grad      = mean gradient + wd * w          (the l2_loss(w)*wd terms of the 'losses' collection, :113)
//   lr        = base_lr * decay_rate ^ floor(step / decay_steps)            (exponential_decay, staircase, :252-256)
//   SGD       : w -= lr * grad
//   Momentum  : a = mu * a + grad ; w -= lr * a
//   Adam      : m = b1 m + (1-b1) grad ; v = b2 v + (1-b2) grad^2 ; w -= lr * sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps)
//   EMA       : s -= (1 - d) (s - w_new),  d = min(decay, (1 + step) / (10 + step))     (:265-267, num_updates=step)
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/dtm_b200.h"

namespace dtm {

struct UpdateScalars {
  float lr;          // decayed learning rate (Adam: already multiplied by the bias-correction factor)
  float ema_keep;    // d
  float grad_scale;  // 1 / world_size
  float lr_plain;    // decayed learning rate without the Adam factor (what the summaries report)
};

__device__ __forceinline__ UpdateScalars make_update_scalars(const dtm_update_config& cfg, const dtm_train_state* st,
                                                             float grad_scale) {
  UpdateScalars u;
  const long long step = st->global_step;
  const float exponent = cfg.decay_steps > 0 ? floorf(static_cast<float>(step / cfg.decay_steps)) : 0.0f;
  float lr = cfg.base_lr * powf(cfg.decay_rate, exponent);
  u.lr_plain = lr;
  if (cfg.optimizer == DTM_OPT_ADAM) lr *= sqrtf(1.0f - st->beta2_power) / (1.0f - st->beta1_power);
  u.lr = lr;
  const float t = static_cast<float>(step);
  u.ema_keep = fminf(cfg.ema_decay, (1.0f + t) / (10.0f + t));
  u.grad_scale = grad_scale;
  return u;
}

// Grid-stride loops visit increasing indices, so the segment only ever moves forward: amortised O(1) instead of a
// scan per element.
__device__ __forceinline__ int segment_advance(const dtm_update_config& cfg, int seg, int i) {
#pragma unroll 1
  while (seg + 1 < cfg.n_segments && i >= cfg.seg_offset[seg + 1]) ++seg;
  return seg;
}

__device__ __forceinline__ int segment_of(const dtm_update_config& cfg, int i) {
  int s = 0;
#pragma unroll 1
  while (s + 1 < cfg.n_segments && i >= cfg.seg_offset[s + 1]) ++s;
  return s;
}

// Updates one element; returns the new weight. slot0/slot1 may be null depending on the optimizer.
__device__ __forceinline__ float update_element(const dtm_update_config& cfg, const UpdateScalars& u, float w, float g_sum,
                                                float wd, float* slot0, float* slot1, float* ema, int i) {
  const float g = g_sum * u.grad_scale + wd * w;
  float w_new;
  if (cfg.optimizer == DTM_OPT_ADAM) {
    const float m = cfg.beta1 * slot0[i] + (1.0f - cfg.beta1) * g;
    const float v = cfg.beta2 * slot1[i] + (1.0f - cfg.beta2) * g * g;
    slot0[i] = m;
    slot1[i] = v;
    w_new = w - u.lr * m / (sqrtf(v) + cfg.epsilon);
  } else if (cfg.optimizer == DTM_OPT_MOMENTUM) {
    const float a = cfg.momentum * slot0[i] + g;
    slot0[i] = a;
    w_new = w - u.lr * a;
  } else {
    w_new = w - u.lr * g;
  }
  if (ema) {
    const float s = ema[i];
    ema[i] = s - (1.0f - u.ema_keep) * (s - w_new);
  }
  return w_new;
}

// Four consecutive elements at once: every state array is touched with one 16-byte load and one 16-byte store.
__device__ __forceinline__ float4 update_vec4(const dtm_update_config& cfg, const UpdateScalars& u, float4 w, float4 gs, float wd,
                                              float* __restrict__ slot0, float* __restrict__ slot1, float* __restrict__ ema, int i) {
  float wv[4] = {w.x, w.y, w.z, w.w};
  float g[4] = {gs.x * u.grad_scale + wd * w.x, gs.y * u.grad_scale + wd * w.y, gs.z * u.grad_scale + wd * w.z,
                gs.w * u.grad_scale + wd * w.w};
  float o[4];
  if (cfg.optimizer == DTM_OPT_ADAM) {
    const float4 m4 = *reinterpret_cast<const float4*>(slot0 + i);
    const float4 v4 = *reinterpret_cast<const float4*>(slot1 + i);
    float m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      m[k] = cfg.beta1 * m[k] + (1.0f - cfg.beta1) * g[k];
      v[k] = cfg.beta2 * v[k] + (1.0f - cfg.beta2) * g[k] * g[k];
      // sqrt via one MUFU.RSQ (v * rsqrt(v)) and an approximate divide: ~1e-7 relative error, a fifth of the
      // instructions of the IEEE sequences - this kernel was issue-bound, not memory-bound, with them
      const float sq = v[k] > 0.0f ? v[k] * rsqrtf(v[k]) : 0.0f;
      o[k] = wv[k] - u.lr * __fdividef(m[k], sq + cfg.epsilon);
    }
    *reinterpret_cast<float4*>(slot0 + i) = make_float4(m[0], m[1], m[2], m[3]);
    *reinterpret_cast<float4*>(slot1 + i) = make_float4(v[0], v[1], v[2], v[3]);
  } else if (cfg.optimizer == DTM_OPT_MOMENTUM) {
    const float4 a4 = *reinterpret_cast<const float4*>(slot0 + i);
    float a[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      a[k] = cfg.momentum * a[k] + g[k];
      o[k] = wv[k] - u.lr * a[k];
    }
    *reinterpret_cast<float4*>(slot0 + i) = make_float4(a[0], a[1], a[2], a[3]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = wv[k] - u.lr * g[k];
  }
  if (ema) {
    const float4 s4 = *reinterpret_cast<const float4*>(ema + i);
    const float k1 = 1.0f - u.ema_keep;
    *reinterpret_cast<float4*>(ema + i) = make_float4(s4.x - k1 * (s4.x - o[0]), s4.y - k1 * (s4.y - o[1]), s4.z - k1 * (s4.z - o[2]),
                                                      s4.w - k1 * (s4.w - o[3]));
  }
  return make_float4(o[0], o[1], o[2], o[3]);
}

// Emits the conv2 operand images for 4 consecutive filter elements starting at flat index `rel` inside the conv2
// segment ((tap*32 + ci)*64 + co, co % 4 == 0).  Same mapping as csrc/conv/conv2_pack.cu.
__device__ __forceinline__ void pack_conv2_vec4(const dtm_update_config& cfg, int rel, float4 o) {
  const int co = rel & 63, ci = (rel >> 6) & 31, tap = rel >> 11;
  if (tap >= 25) return;
  __nv_bfloat16* fwd = static_cast<__nv_bfloat16*>(cfg.pack_fwd);
  __nv_bfloat16* dg = static_cast<__nv_bfloat16*>(cfg.pack_dgrad);
  const int k = tap * 32 + ci;
  const float v[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
  for (int t = 0; t < 4; ++t) fwd[((k >> 3) * 64 + co + t) * 8 + (k & 7)] = __float2bfloat16(v[t]);
  const int kd = (24 - tap) * 64 + co;          // co..co+3 stay inside one group of 8
  uint2 pk;
  __nv_bfloat162 lo2 = __floats2bfloat162_rn(o.x, o.y), hi2 = __floats2bfloat162_rn(o.z, o.w);
  pk.x = *reinterpret_cast<uint32_t*>(&lo2);
  pk.y = *reinterpret_cast<uint32_t*>(&hi2);
  *reinterpret_cast<uint2*>(dg + ((kd >> 3) * 32 + ci) * 8 + (kd & 7)) = pk;
}

// End-of-step bookkeeping, executed by exactly one thread after every element has been updated:
// publishes the step's metrics, advances the loss moving averages (mnist_input.py:288-290, decay 0.9),
// the Adam beta powers and global_step, and clears the per-step accumulators.
// Closes the step: loss bookkeeping, beta powers, global_step.  Called by EVERY thread of the (single) CTA that finishes
// last; blockDim.x >= DTM_MAX_SEGMENTS.  The per-variable work runs one thread per variable and thread 0 loads all of its
// scalars before its first store, so the step ends after ~2 dependent memory round trips instead of ~25 (this tail is
// on the critical path of every training step).
__device__ __forceinline__ void finish_step(const dtm_update_config& cfg, dtm_train_state* st, float lr_plain) {
  __shared__ float s_fin_wl[DTM_MAX_SEGMENTS];
  const int t = threadIdx.x;
  if (t < DTM_MAX_SEGMENTS) {
    float wl = 0.0f;
    if (t < cfg.n_segments) {
      const float sum = st->wl_sum[t];
      const float avg = st->loss_avg[t];
      wl = 0.5f * cfg.seg_wd[t] * sum;
      st->last_wl[t] = wl;
      st->loss_avg[t] = avg - 0.1f * (avg - wl);
      st->wl_sum[t] = 0.0f;
    }
    s_fin_wl[t] = wl;
  }
  __syncthreads();
  if (t == 0) {
    const float xent = st->loss_sum;
    const float acc = st->correct;
    const float avg_x = st->loss_avg[DTM_MAX_SEGMENTS];
    const float avg_t = st->loss_avg[DTM_MAX_SEGMENTS + 1];
    const float b1 = st->beta1_power, b2 = st->beta2_power;
    const long long step = st->global_step;
    float total = xent;
    for (int s = 0; s < cfg.n_segments; ++s) total += s_fin_wl[s];   // same summation order as ever
    st->loss_avg[DTM_MAX_SEGMENTS] = avg_x - 0.1f * (avg_x - xent);
    st->loss_avg[DTM_MAX_SEGMENTS + 1] = avg_t - 0.1f * (avg_t - total);
    st->last_xent = xent;
    st->last_total = total;
    st->last_acc = acc;
    st->last_lr = lr_plain;
    st->loss_sum = 0.0f;
    st->correct = 0.0f;
    if (cfg.optimizer == DTM_OPT_ADAM) {
      st->beta1_power = b1 * cfg.beta1;
      st->beta2_power = b2 * cfg.beta2;
    }
    st->global_step = step + 1;
  }
}

}  // namespace dtm
