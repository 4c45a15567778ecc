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
};

__device__ __forceinline__ UpdateScalars make_update_scalars(const dtm_update_config& cfg, const dtm_train_state* st,
                                                             float grad_scale) {
  UpdateScalars u;
  const long long step = st->global_step;
  const float exponent = cfg.decay_steps > 0 ? floorf(static_cast<float>(step / cfg.decay_steps)) : 0.0f;
  float lr = cfg.base_lr * powf(cfg.decay_rate, exponent);
  if (cfg.optimizer == DTM_OPT_ADAM) lr *= sqrtf(1.0f - st->beta2_power) / (1.0f - st->beta1_power);
  u.lr = lr;
  const float t = static_cast<float>(step);
  u.ema_keep = fminf(cfg.ema_decay, (1.0f + t) / (10.0f + t));
  u.grad_scale = grad_scale;
  return u;
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

// End-of-step bookkeeping, executed by exactly one thread after every element has been updated:
// publishes the step's metrics, advances the loss moving averages (mnist_input.py:288-290, decay 0.9),
// the Adam beta powers and global_step, and clears the per-step accumulators.
__device__ __forceinline__ void finish_step(const dtm_update_config& cfg, dtm_train_state* st, float lr_plain) {
  const float xent = st->loss_sum;
  float total = xent;
  for (int s = 0; s < cfg.n_segments; ++s) {
    const float wl = 0.5f * cfg.seg_wd[s] * st->wl_sum[s];
    st->last_wl[s] = wl;
    total += wl;
    st->loss_avg[s] -= 0.1f * (st->loss_avg[s] - wl);
    st->wl_sum[s] = 0.0f;
  }
  st->loss_avg[DTM_MAX_SEGMENTS] -= 0.1f * (st->loss_avg[DTM_MAX_SEGMENTS] - xent);
  st->loss_avg[DTM_MAX_SEGMENTS + 1] -= 0.1f * (st->loss_avg[DTM_MAX_SEGMENTS + 1] - total);
  st->last_xent = xent;
  st->last_total = total;
  st->last_acc = st->correct;
  st->last_lr = lr_plain;
  st->loss_sum = 0.0f;
  st->correct = 0.0f;
  if (cfg.optimizer == DTM_OPT_ADAM) {
    st->beta1_power *= cfg.beta1;
    st->beta2_power *= cfg.beta2;
  }
  st->global_step += 1;
}

}  // namespace dtm
