// The 10-class head of the MNIST CNN: softmax_linear (mnist_input.py:200-205), sparse softmax cross-entropy +
// batch mean (:224-226), top-1 accuracy (:238-240), the NaN check that NanTensorHook performs on the host
// (main.py:138) and the whole backward of the head, in two small CUDA-core kernels (N = 10 is GEMV-shaped; a
// 128-row tensor-core tile would be 92 % padding).
//
//   head_forward : one warp per example.  logits = h . W + b ; loss/accuracy accumulated into the device-side
//                  training state ; dlogits = (softmax - onehot) * grad_scale.
//   head_backward: one CTA per 32 hidden units.  dH = dlogits . W^T, gated by the sign of the forward activation
//                  (ReLU backward; for a dropped-out activation the stored h is 0, so the same test also applies
//                  the dropout mask and only the 1/keep scale remains), dW = h^T . dlogits, both bias gradients.
#include "../common/ptx.cuh"
#include "../common/host_util.h"
#include "../../include/dtm_b200.h"

namespace dtm {

constexpr int NCLS = 10;

constexpr int HF_CHUNK = 1024;     // hidden units staged per pass
constexpr int HF_PITCH = NCLS + 1; // 11 floats per weight row: lane stride 11 words is a permutation of the 32 banks

__global__ void __launch_bounds__(256)
head_forward_kernel(const __nv_bfloat16* __restrict__ h, int batch, int hidden, const float* __restrict__ w,
                    const float* __restrict__ bias, const int* __restrict__ labels, float grad_scale, float loss_scale,
                    float* __restrict__ logits_out, float* __restrict__ probs_out, float* __restrict__ dlogits,
                    __nv_bfloat16* __restrict__ dh, float inv_keep, dtm_train_state* __restrict__ state, int early_weights) {
  __shared__ float s_w[HF_CHUNK * HF_PITCH];   // 44 KiB: the weight rows of the current chunk, shared by the CTA's 8 rows
  __shared__ float s_loss[8], s_corr[8];
  __shared__ unsigned int s_nan;
  const int wid = threadIdx.x >> 5;
  const int row = blockIdx.x * 8 + wid;
  const uint32_t lane = lane_id();
  const bool row_ok = row < batch;
  if (threadIdx.x == 0) s_nan = 0;
  if (!early_weights) pdl_wait();   // strict mode: nothing is read from global memory before the dependency wait
  float acc[NCLS];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) acc[c] = 0.0f;
  const __nv_bfloat16* hr = h + static_cast<size_t>(row_ok ? row : 0) * hidden;
  // stage one chunk of weight rows with 16-byte loads, several in flight per thread (a serial 4-byte loop cost 20 us)
  auto stage_chunk = [&](int k0, int kc) {
    const float4* w4 = reinterpret_cast<const float4*>(w + static_cast<size_t>(k0) * NCLS);
    const int n4 = (kc * NCLS) >> 2;
#pragma unroll 5
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
      const float4 v = w4[i];
      const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int idx = 4 * i + t;
        s_w[(idx / NCLS) * HF_PITCH + idx % NCLS] = e[t];
      }
    }
    for (int idx = (n4 << 2) + threadIdx.x; idx < kc * NCLS; idx += blockDim.x)
      s_w[(idx / NCLS) * HF_PITCH + idx % NCLS] = w[static_cast<size_t>(k0) * NCLS + idx];
  };
  for (int k0 = 0; k0 < hidden; k0 += HF_CHUNK) {
    const int kc = min(HF_CHUNK, hidden - k0);
    __syncthreads();
    stage_chunk(k0, kc);
    __syncthreads();
    if (k0 == 0) { pdl_wait(); pdl_launch_dependents(); }   // the weights above are older than the preceding kernel; h is not
    if (row_ok) {
#pragma unroll 8
      for (int k = lane; k < kc; k += 32) {
        const float hv = __bfloat162float(hr[k0 + k]);
        const float* wk = s_w + k * HF_PITCH;
#pragma unroll
        for (int c = 0; c < NCLS; ++c) acc[c] = fmaf(hv, wk[c], acc[c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NCLS; ++c) acc[c] = warp_sum(acc[c]) + bias[c];

  float my_loss = 0.0f, my_corr = 0.0f;
  float d[NCLS];   // dlogits of this warp's row (lane 0 computes, everyone gets a copy below)
#pragma unroll
  for (int c = 0; c < NCLS; ++c) d[c] = 0.0f;
  if (row_ok && lane == 0) {
    float mx = acc[0];
#pragma unroll
    for (int c = 1; c < NCLS; ++c) mx = fmaxf(mx, acc[c]);
    float e[NCLS], sum = 0.0f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) { e[c] = __expf(acc[c] - mx); sum += e[c]; }
    const float inv = 1.0f / sum;
    if (logits_out) {
#pragma unroll
      for (int c = 0; c < NCLS; ++c) logits_out[static_cast<size_t>(row) * NCLS + c] = acc[c];
    }
    if (probs_out) {
#pragma unroll
      for (int c = 0; c < NCLS; ++c) probs_out[static_cast<size_t>(row) * NCLS + c] = e[c] * inv;
    }
    if (labels) {
      const int y = labels[row];
      float target = 0.0f;
      int greater = 0;
#pragma unroll
      for (int c = 0; c < NCLS; ++c) if (c == y) target = acc[c];
#pragma unroll
      for (int c = 0; c < NCLS; ++c) greater += (acc[c] > target) ? 1 : 0;
      const float loss = (mx - target) + __logf(sum);
      if (dlogits) {
#pragma unroll
        for (int c = 0; c < NCLS; ++c) {
          d[c] = (e[c] * inv - (c == y ? 1.0f : 0.0f)) * grad_scale;
          dlogits[static_cast<size_t>(row) * NCLS + c] = d[c];
        }
      }
      my_loss = loss * loss_scale;
      my_corr = greater == 0 ? loss_scale : 0.0f;
      if (!(fabsf(loss) <= 3.0e38f)) atomicOr(&s_nan, 1u);  // catches NaN and Inf
    }
  }
  // gradient w.r.t. the hidden activation, dH = (dlogits . W^T) * inv_keep * [h > 0], while the weight rows are still in
  // shared memory: the fc gradient GEMMs that follow need nothing else from the head, so its weight/bias gradients
  // (head_backward) leave the critical path
  if (dh != nullptr && dlogits != nullptr && labels != nullptr) {
#pragma unroll
    for (int c = 0; c < NCLS; ++c) d[c] = __shfl_sync(0xffffffffu, d[c], 0);
    for (int k0 = 0; k0 < hidden; k0 += HF_CHUNK) {
      const int kc = min(HF_CHUNK, hidden - k0);
      if (hidden > HF_CHUNK) {   // more than one chunk: the buffer holds the last one, stage again
        __syncthreads();
        stage_chunk(k0, kc);
        __syncthreads();
      }
      if (row_ok) {
#pragma unroll 4
        for (int k = lane; k < kc; k += 32) {
          const float hv = __bfloat162float(hr[k0 + k]);
          const float* wk = s_w + k * HF_PITCH;
          float g = 0.0f;
#pragma unroll
          for (int c = 0; c < NCLS; ++c) g = fmaf(d[c], wk[c], g);
          dh[static_cast<size_t>(row) * hidden + k0 + k] = __float2bfloat16(hv > 0.0f ? g * inv_keep : 0.0f);
        }
      }
    }
  }
  if (lane == 0) { s_loss[wid] = my_loss; s_corr[wid] = my_corr; }
  __syncthreads();
  if (threadIdx.x == 0 && state && labels) {   // one atomic per CTA instead of one per example
    float l = 0.0f, cnt = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { l += s_loss[i]; cnt += s_corr[i]; }
    atomicAdd(&state->loss_sum, l);
    if (cnt != 0.0f) atomicAdd(&state->correct, cnt);
    if (s_nan) atomicOr(&state->nan_flag, 1u);
  }
}

constexpr int HB_COLS = 32;

__global__ void __launch_bounds__(256)
head_backward_kernel(const __nv_bfloat16* __restrict__ h, int batch, int hidden, const float* __restrict__ w,
                     const float* __restrict__ dlogits, float inv_keep, __nv_bfloat16* __restrict__ dh,
                     float* __restrict__ dw, float* __restrict__ db_out, float* __restrict__ db_hidden, int dh_ready) {
  __shared__ float s_red[8][NCLS + 1][HB_COLS + 1];
  const int kk = threadIdx.x & 31;
  const int r = threadIdx.x >> 5;
  const int k = blockIdx.x * HB_COLS + kk;
  const bool col_ok = k < hidden;
  pdl_wait();                 // dlogits / h come from the preceding kernels
  pdl_launch_dependents();
  float wk[NCLS];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) wk[c] = col_ok ? w[static_cast<size_t>(k) * NCLS + c] : 0.0f;
  float adw[NCLS];
#pragma unroll
  for (int c = 0; c < NCLS; ++c) adw[c] = 0.0f;
  float adb = 0.0f;
  for (int b = blockIdx.y * 8 + r; b < batch; b += gridDim.y * 8) {
    const float* dl = dlogits + static_cast<size_t>(b) * NCLS;   // same address across the warp: broadcast load
    float d[NCLS];
#pragma unroll
    for (int c = 0; c < NCLS; ++c) d[c] = dl[c];
    if (col_ok) {
      const float hv = __bfloat162float(h[static_cast<size_t>(b) * hidden + k]);
      float g = 0.0f;
#pragma unroll
      for (int c = 0; c < NCLS; ++c) { g = fmaf(d[c], wk[c], g); adw[c] = fmaf(hv, d[c], adw[c]); }
      __nv_bfloat16 gb;
      if (dh_ready) {   // head_forward already produced dH: only its column sums (hidden-bias gradient) are needed
        gb = dh[static_cast<size_t>(b) * hidden + k];
      } else {
        gb = __float2bfloat16(hv > 0.0f ? g * inv_keep : 0.0f);
        dh[static_cast<size_t>(b) * hidden + k] = gb;
      }
      adb += __bfloat162float(gb);
    }
  }
#pragma unroll
  for (int c = 0; c < NCLS; ++c) s_red[r][c][kk] = adw[c];
  s_red[r][NCLS][kk] = adb;
  __syncthreads();
  for (int i = threadIdx.x; i < (NCLS + 1) * HB_COLS; i += blockDim.x) {
    const int c = i / HB_COLS, col = i % HB_COLS;
    const int kq = blockIdx.x * HB_COLS + col;
    if (kq >= hidden) continue;
    float s = 0.0f;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += s_red[q][c][col];
    if (c < NCLS) red_add_f32(dw + static_cast<size_t>(kq) * NCLS + c, s);
    else if (db_hidden) red_add_f32(db_hidden + kq, s);
  }
  // output-bias gradient: column sums of dlogits, done once (first column block), split over blockIdx.y
  if (blockIdx.x == 0 && threadIdx.x < NCLS) {
    float s = 0.0f;
    for (int b = blockIdx.y; b < batch; b += gridDim.y) s += dlogits[static_cast<size_t>(b) * NCLS + threadIdx.x];
    red_add_f32(db_out + threadIdx.x, s);
  }
}

}  // namespace dtm

extern "C" int dtm_head_forward(const void* h, int batch, int hidden, const float* w, const float* bias, const int* labels,
                                float grad_scale, float loss_scale, float* logits, float* probs, float* dlogits, void* dh,
                                float inv_keep, dtm_train_state* state, void* stream) {
  using namespace dtm;
  if (batch <= 0 || hidden <= 0) return dtm_fail("head_forward: batch and hidden must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DTM_CUDA_OK(launch_pdl(head_forward_kernel, dim3((batch + 7) / 8), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(h), batch, hidden,
                         w, bias, labels, grad_scale, loss_scale, logits, probs, dlogits, static_cast<__nv_bfloat16*>(dh), inv_keep, state,
                         pdl_mode() == 2 ? 1 : 0));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dtm_head_backward(const void* h, int batch, int hidden, const float* w, const float* dlogits, float inv_keep,
                                 void* dh, float* dw, float* db_out, float* db_hidden, int dh_ready, void* stream) {
  using namespace dtm;
  if (batch <= 0 || hidden <= 0) return dtm_fail("head_backward: batch and hidden must be positive");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int col_blocks = (hidden + HB_COLS - 1) / HB_COLS;
  int split = 1;
  while (col_blocks * split < 2 * device_sm_count() && split * 16 <= batch && split < 16) split *= 2;
  DTM_CUDA_OK(launch_pdl(head_backward_kernel, dim3(col_blocks, split), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(h), batch,
                         hidden, w, dlogits, inv_keep, static_cast<__nv_bfloat16*>(dh), dw, db_out, db_hidden, dh_ready));
  DTM_CUDA_OK(cudaGetLastError());
  return 0;
}
