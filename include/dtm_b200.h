/* libdtm_b200 — C ABI of the sm_100a kernel library behind distributed_tensorflow_ibm_mnist_b200.
 *
 * Conventions
 *   - every entry point returns 0 on success, non-zero on failure; dtm_last_error() then holds a
 *     human-readable message for the calling thread.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued asynchronously on it.
 *   - pointers are device pointers unless a parameter is documented as host memory.
 *   - "plane layout" (activations that feed a tensor-core convolution): per image, C/8 planes of
 *     16 x 16 pixels x 8 channels (bf16, 4096 B per plane); plane row r holds image row r-2, rows 0-1 and
 *     columns 14-15 are zero.  A 5x5 SAME convolution tap is then a constant byte offset, so the UMMA
 *     shared-memory descriptor is simply re-based per tap and no im2col copy is ever made.
 */
#ifndef DTM_B200_H_
#define DTM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DTM_ABI_VERSION 1

/* GEMM epilogues */
#define DTM_EPI_STORE_F32 0      /* D(fp32)  = alpha * acc          (split-K: red.add into pre-zeroed D) */
#define DTM_EPI_BIAS_BF16 1      /* D(bf16)  = acc + bias                                            */
#define DTM_EPI_BIAS_RELU_BF16 2 /* D(bf16)  = dropout(relu(acc + bias))                              */
#define DTM_EPI_MASK_BF16 3      /* D(bf16)  = acc * (ref > 0)      (ReLU backward)                   */

/* ---- runtime ---------------------------------------------------------------------------------- */
const char* dtm_last_error(void);
int dtm_abi_version(void);
int dtm_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);

/* ---- tcgen05 probe (tools/umma_probe.py) -------------------------------------------------------- */
/* Copies `image` into shared memory, issues the listed tcgen05.mma ops, dumps 128 TMEM lanes x ncols.
 * No counterpart in the reference; used to pin descriptor/layout semantics on real silicon. */
int dtm_umma_probe(const void* image, uint32_t image_bytes, const void* ops, uint32_t n_ops, float* dump,
                   uint32_t ncols, void* stream);

/* ---- fully connected layers (reference: mnist_input.py:184, :192 — tf.matmul + bias + relu) ------- */
/* D[M,N] = A[M,K] . B[K,N], bf16 operands, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A stored [M][K] (K contiguous); 1: A stored [K][M] (M contiguous).
 *   b_mn_major = 0: B stored [N][K] (K contiguous); 1: B stored [K][N] (N contiguous, TensorFlow layout).
 * N, K, ldd multiples of 8.  `ref` (bf16, [M][ldd]) gates DTM_EPI_MASK_BF16.  keep_prob >= 1 disables
 * dropout; otherwise the mask is Philox(seed, *step_ptr, element).  split_k > 1 needs DTM_EPI_STORE_F32
 * and a zero-initialised D. */
int dtm_gemm_bf16(const void* a, int a_mn_major, const void* b, int b_mn_major, void* d, int M, int N, int K,
                  int ldd, int epilogue, const float* bias, const void* ref, float alpha, float keep_prob,
                  uint64_t seed, const long long* step_ptr, int split_k, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTM_B200_H_ */
