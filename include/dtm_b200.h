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

#define DTM_ABI_VERSION 11

/* GEMM epilogues */
#define DTM_EPI_STORE_F32 0      /* D(fp32)  = alpha * acc          (split-K: red.add into pre-zeroed D) */
#define DTM_EPI_BIAS_BF16 1      /* D(bf16)  = acc + bias                                            */
#define DTM_EPI_BIAS_RELU_BF16 2 /* D(bf16)  = dropout(relu(acc + bias))                              */
#define DTM_EPI_MASK_BF16 3      /* D(bf16)  = acc * (ref > 0)      (ReLU backward)                   */
#define DTM_EPI_MASK_UNPOOL_BF16 4 /* as 3, plus un-pooling into conv2's plane layout (dtm_gemm_mask_unpool only) */

/* Optimizers (reference: tf_parameter_mgr.getOptimizer(lr), mnist_input.py:261) */
#define DTM_OPT_SGD 0
#define DTM_OPT_MOMENTUM 1
#define DTM_OPT_ADAM 2

#define DTM_MAX_SEGMENTS 16

/* Device-resident training state: everything the hot loop needs lives on the GPU so that a training step is a
 * pure kernel sequence (CUDA-graph replayable, no host synchronisation).  Mirrors the scalar variables the
 * reference keeps in its graph: global_step (main.py:83), beta powers of Adam, the loss moving averages
 * (mnist_input.py:288-290), plus the loss/accuracy/NaN accumulators read by the hooks (main.py:104-107, :138). */
typedef struct dtm_train_state {
  long long global_step;
  float beta1_power, beta2_power;
  float loss_sum;          /* cross-entropy mean of the step in flight (accumulated by dtm_head_forward)   */
  float correct;           /* accuracy of the step in flight                                              */
  unsigned int nan_flag;   /* sticky: set when a non-finite loss is seen                                  */
  unsigned int ticket;     /* internal: last-block election of the update kernels                         */
  float last_xent, last_total, last_acc, last_lr; /* metrics of the last completed step                  */
  float wl_sum[DTM_MAX_SEGMENTS];   /* sum of squares per variable (weights used by the step in flight)   */
  float last_wl[DTM_MAX_SEGMENTS];  /* wd * l2_loss per variable of the last completed step               */
  float loss_avg[DTM_MAX_SEGMENTS + 2]; /* EMA(0.9): per-variable weight_loss, [16] cross_entropy, [17] total_loss */
  unsigned int work;       /* internal: chunk dispenser of dtm_fused_update_range (zero between launches)         */
  unsigned int reserved0;
} dtm_train_state;

/* Flat parameter space: variables ("segments") are laid out back to back, each offset a multiple of 64 elements. */
typedef struct dtm_update_config {
  int optimizer;
  int n_segments;
  int total;                /* number of fp32 elements incl. padding (multiple of 64) */
  int reserved;
  long long decay_steps;
  float base_lr, decay_rate, momentum, beta1, beta2, epsilon, ema_decay;
  int seg_offset[DTM_MAX_SEGMENTS + 1];
  float seg_wd[DTM_MAX_SEGMENTS];
  int seg_zero_grad[DTM_MAX_SEGMENTS]; /* 1: gradient is accumulated with atomics and must be cleared after use */
  /* Optional: while writing the bf16 copy of segment `pack_seg` (the conv2 filter, [5][5][32][64]) also emit the two
   * tensor-core operand images of dtm_conv2_pack_weights, so that no separate packing pass is needed.  -1 = off. */
  /* Multi-GPU only: 1 = after the update the owner sends just the bf16 value of this variable to its peers (the
   * variable is consumed by tensor-core kernels only); the fp32 master then lives on the owning rank alone and is
   * gathered on demand (checkpoints).  0 = the fp32 value is all-gathered and every rank derives bf16 locally. */
  int seg_gather_bf16[DTM_MAX_SEGMENTS];
  int pack_seg;
  int reserved2;
  void* pack_fwd;
  void* pack_dgrad;
} dtm_update_config;

/* Peer-memory communication context, passed to kernels by value: base pointers of every rank's symmetric buffers
 * (flat gradient, flat fp32 weights, barrier flags), obtained through CUDA IPC (dtm_symm_*). */
#define DTM_MAX_RANKS 8
typedef struct dtm_peer_ctx {
  int rank, world;
  float* grad[DTM_MAX_RANKS];
  float* w[DTM_MAX_RANKS];
  unsigned int* flags[DTM_MAX_RANKS]; /* DTM_MAX_RANKS epochs per rank; slot r of rank p is written by rank r */
  void* w_bf16[DTM_MAX_RANKS];        /* bf16 shadow weights (symmetric): target of the bf16-only all-gather */
  /* NVLink-SHARP (NVLS) multicast addresses of the same three buffers, or NULL: one address that the NVSwitch maps onto
   * every rank's copy.  When all three are set the collectives reduce inside the switch (multimem.ld_reduce) and
   * all-gather with one multicast store (multimem.st) instead of N-1 peer loads / stores per element. */
  float* mc_grad;
  float* mc_w;
  void* mc_w_bf16;
} dtm_peer_ctx;

/* ---- runtime ---------------------------------------------------------------------------------- */
/* Keep [base, base + bytes) - the flat optimizer state - in the persisting part of the L2 for the update kernels of the
 * current device (access-policy window on their launches; the carve-out is raised to min(bytes, device maximum)).
 * base = NULL switches it off.  The reference has no counterpart (its Apply ops run inside TensorFlow's runtime,
 * mnist_input.py:262-264); this is a B200 memory-hierarchy decision. */
int dtm_set_update_l2_window(const void* base, size_t bytes, size_t* carveout_bytes, size_t* max_window_bytes);
const char* dtm_last_error(void);
int dtm_abi_version(void);
int dtm_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);
/* Programmatic dependent launch (kernel prologues overlap the tail of the preceding kernel of the stream).
 * 0 = off; 1 = strict: no kernel reads global memory before its dependency wait (default; always safe);
 * 2 = early weights: conv2 / head kernels stage their WEIGHT operands before the wait - only valid if no weight buffer is
 * written by the kernel launched immediately before them (the training engine's schedule; it switches to 2 around its own
 * launches).  -1 = back to the DTM_PDL environment variable.  Process-wide, host-side; returns the previous setting. */
int dtm_set_pdl_mode(int mode);

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

/* Gradient GEMM of the first fully connected layer fused with the max-pool backward of conv2 (non-LRN presets):
 *   dp2[B][3136] (bf16) = (dy[B][hidden] . w_bf16[3136][hidden]^T) * [p2 > 0]         (what DTM_EPI_MASK_BF16 computes)
 *   dz2_planes (plane layout, 64 ch): every pooled value routed to its arg-max position, the other three positions of the
 *   window zeroed - exactly what dtm_unpool2 writes.  One N tile of the GEMM is one pooled pixel (64 channels), so the
 *   epilogue thread that owns an (image, pooled pixel) writes its 2x2 window directly.  The conv2 bias gradient (column sums
 *   of dp2) is left to dtm_colsum_bf16, off the critical path. */
int dtm_gemm_mask_unpool(const void* dy, const void* w_bf16, const void* p2, const void* argmax, void* dp2, void* dz2_planes,
                         int batch, int hidden, void* stream);

/* ---- conv1 (reference: mnist_input.py:39, :134-153) ----------------------------------------------------------- */
/* x: uint8 [B][28*28*C] (in_is_float = 0, normalised x/255-0.5 in-kernel) or float32 already normalised.
 * w: fp32 TF layout [5][5][C][32].  Outputs: p1_planes (plane layout, 32 ch), argmax uint8 [B][196][32],
 * pool_raw bf16 [B][196][32] (only with lrn: the pooled activation before LRN, needed by the backward pass). */
int dtm_conv1_fwd(const void* x, int in_is_float, int channels, const float* w, const float* bias, int batch, int lrn,
                  void* p1_planes, void* argmax, void* pool_raw, void* stream);
/* grad_pooled bf16 [B][196][32]: gradient w.r.t. the pooled activation, ReLU mask applied.  dw/db are accumulated
 * (red.add) and must be zero on entry. */
int dtm_conv1_wgrad(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax, int batch,
                    float* dw, float* db, void* stream);
/* Same contract on tcgen05 (K = pixels; bf16 operands, so pixels are rounded to bf16 after normalisation).  Verified, but
 * slower than the CUDA-core kernel in round 1: dtm_conv1_wgrad only dispatches to it when DTM_C1W_TC=1. */
int dtm_conv1_wgrad_tc(const void* x, int in_is_float, int channels, const void* grad_pooled, const void* argmax, int batch,
                       float* dw, float* db, void* stream);

/* ---- conv2 on tcgen05 (reference: mnist_input.py:156-172) ------------------------------------------------------ */
/* w2_bf16: bf16 copy of the TF-layout weights [5][5][32][64].  p2: bf16 [B][7][7][64] (== [B][3136] in TF flatten
 * order), argmax uint8 same shape, act2 bf16 [B][196][64] (only with lrn: post-ReLU, pre-LRN activation). */
int dtm_conv2_fwd(const void* p1_planes, const void* w2_bf16, const void* packed, const float* bias, int batch, int lrn,
                  void* p2, void* argmax, void* act2, void* stream);
/* Re-packs the bf16 conv2 weights into the two shared-memory operand images the tensor-core kernels use
 * (forward: [k/8][co][8], k = tap*32+ci; dgrad: [k/8][ci][8], k = (24-tap)*64+co), 102400 bytes each, so that each CTA
 * fetches its weights with one bulk copy instead of re-packing them itself.  Pass the results as `packed` below
 * (NULL = let the kernel pack from w2_bf16). */
int dtm_conv2_pack_weights(const void* w2_bf16, void* packed_fwd, void* packed_dgrad, void* stream);
/* dz2_planes: plane layout, 64 ch (from dtm_unpool2).  dpool1: bf16 [B][196][32], gradient w.r.t. conv1's pooled
 * activation with the ReLU mask (and LRN backward when lrn) applied. */
int dtm_conv2_dgrad(const void* dz2_planes, const void* w2_bf16, const void* packed, const void* p1_planes,
                    const void* pool_raw, int batch, int lrn, void* dpool1, void* stream);
/* dw2 fp32 [5][5][32][64], accumulated (red.add): zero on entry. */
int dtm_conv2_wgrad(const void* p1_planes, const void* dz2_planes, int batch, float* dw2, void* stream);
/* dtm_conv2_wgrad and dtm_conv1_wgrad (uint8 pixels, channels = 1 or 3) in ONE tcgen05 kernel: the warps that idle in conv2_wgrad
 * build conv1's operands (16-byte shifted-window entries of the image, un-pooled gradient planes) and the MMA warp
 * accumulates dW1 / db1 in spare TMEM columns (autodiff of mnist_input.py:142-147, :161).  Same contracts: dw2, dw1, db1
 * are accumulated with red.add and must be zero on entry; pixels are rounded to bf16 after normalisation. */
int dtm_conv_wgrad_fused(const void* p1_planes, const void* dz2_planes, const void* x_u8, int channels, const void* grad_pooled,
                         const void* argmax, int batch, float* dw2, float* dw1, float* db1, void* stream);
/* dp2 bf16 [B][49][64] -> dz2_planes (every byte of the valid region is written) ; db2[64] accumulated. */
int dtm_unpool2(const void* dp2, const void* argmax, const void* act2, int batch, int lrn, void* dz2_planes, float* db2,
                void* stream);
int dtm_colsum_bf16(const void* x, int M, int N, float* out, void* stream);

/* ---- classifier head + loss (reference: mnist_input.py:200-205, :224-226, :238-240; main.py:138) ---------------- */
/* h bf16 [B][hidden]; w fp32 [hidden][10]; labels int32 (may be NULL for inference).  Any of logits / probs /
 * dlogits ([B][10] fp32) / dh / state may be NULL.  loss_scale is 1/B so that state->loss_sum is the batch mean.
 * dh (needs labels and dlogits): bf16 [B][hidden] = (dlogits . w^T) * inv_keep * [h > 0], computed while the weight rows are
 * still in shared memory, so that the gradient GEMMs of the layer below can start without waiting for dtm_head_backward. */
int dtm_head_forward(const void* h, int batch, int hidden, const float* w, const float* bias, const int* labels,
                     float grad_scale, float loss_scale, float* logits, float* probs, float* dlogits, void* dh, float inv_keep,
                     dtm_train_state* state, void* stream);
/* dw [hidden][10], db_out [10], db_hidden [hidden] (may be NULL) are accumulated (red.add).  dh_ready = 0: also computes and
 * stores dh as above; dh_ready = 1: dh was produced by dtm_head_forward and is only read (for db_hidden). */
int dtm_head_backward(const void* h, int batch, int hidden, const float* w, const float* dlogits, float inv_keep, void* dh,
                      float* dw, float* db_out, float* db_hidden, int dh_ready, void* stream);

/* ---- optimizer (reference: mnist_input.py:246-270) ------------------------------------------------------------ */
/* One kernel: lr schedule, weight decay, SGD/Momentum/Adam, EMA(0.9999), bf16 shadow weights, loss averages,
 * global_step++.  grad_scale multiplies the summed gradient (1/world_size).  slot0/slot1/ema/w_bf16 may be NULL. */
int dtm_fused_update(const dtm_update_config* cfg, dtm_train_state* state, float* w, float* grad, float* slot0,
                     float* slot1, float* ema, void* w_bf16, float grad_scale, void* stream);
/* Same for the flat elements [lo, hi) only (multiples of 64).  Lets a caller update a bucket of variables as soon as
 * its gradients are complete, beside the rest of the backward pass; only the call with do_finish = 1 closes the step.
 * max_ctas > 0 caps the grid so that the kernel leaves room for the kernels it is meant to overlap with. */
int dtm_fused_update_range(const dtm_update_config* cfg, dtm_train_state* state, float* w, float* grad, float* slot0,
                           float* slot1, float* ema, void* w_bf16, float grad_scale, int lo, int hi, int do_finish,
                           int max_ctas, void* stream);
int dtm_state_init(dtm_train_state* state, long long global_step, float beta1, float beta2, void* stream);

/* ---- summary statistics (reference: main.py:95-109 through monitor_cb) ----------------------------------------- */
/* out6 = {min, max, sum, sum of squares, element count, L2 norm}.  `scratch` is device memory of at least
 * dtm_tensor_stats_scratch_bytes(n) bytes, 8-byte aligned, whose first 64 bytes are zero before the first call
 * (the kernel leaves them zero again); calls sharing one scratch buffer must not overlap on the GPU. */
size_t dtm_tensor_stats_scratch_bytes(size_t n);
int dtm_tensor_stats(const void* x, size_t n, int is_bf16, float* out6, void* scratch, size_t scratch_bytes, void* stream);
/* Allocation-free variant of dtm_tensor_stats (no counterpart in the reference).  For callers that cannot hand over
 * device scratch: hooks that run when the GPU's memory is fully committed, watcher threads that must not call
 * cudaMalloc (it can synchronise the device and stall the streams being watched), plain-C hosts without a device
 * allocator.  The partials live in a 16 KiB array inside the library image (held by every GPU that loads the library
 * until process exit), which limits n to DTM_STATS_STATIC_MAX_ELEMS elements; larger inputs are rejected with an error.
 * The first call on each device resolves the array's device address (make that call before any stream capture).
 * Calls share the array: do not let two of them overlap on one GPU. */
#define DTM_STATS_STATIC_MAX_ELEMS 1044480
int dtm_tensor_stats_static(const void* x, size_t n, int is_bf16, float* out6, void* stream);
/* counts[nbins] (float, accumulated) over the sorted bucket limits (device doubles), TensorFlow histogram rule. */
int dtm_histogram(const void* x, size_t n, int is_bf16, const double* limits, int nbins, float* counts, void* stream);
/* host helper: CRC-32C of a host buffer, continuing from `crc` (TFRecord / checkpoint / event-file codecs). */
uint32_t dtm_crc32c(const char* data, size_t n, uint32_t crc);

/* ---- native input pipeline (host code; reference: the queue runners of mnist_input.py:44-73) --------------------- */
/* Reader threads parse TFRecord-framed tf.train.Example records (`image_raw` bytes of exactly image_bytes, `label` int64)
 * from `files` (shuffled file order, infinite epochs) into a RandomShuffleQueue of capacity min_after_dequeue + 3*batch;
 * dtm_loader_next blocks until more than min_after_dequeue records are queued, then copies `batch` uniformly drawn records
 * into pixels_out (uint8 [batch][image_bytes]) and labels_out (int32 [batch]) - typically two regions of one pinned buffer.
 * shuffle = 0: one reader, file order, FIFO.  A missing file fails dtm_loader_create ("Failed to find file: ..."); corrupt
 * or mis-sized records fail dtm_loader_next.  Both return their message through dtm_last_error(). */
void* dtm_loader_create(const char* const* files, int n_files, int batch, int image_bytes, int num_threads, uint64_t seed,
                        int shuffle, int min_after_dequeue);
int dtm_loader_next(void* loader, uint8_t* pixels_out, int32_t* labels_out);
/* Batch ring: after dtm_loader_ring_start a thread of the loader fills the caller's n_buffers (>= 2) buffers - each
 * label_offset + 4*batch bytes, pixels at 0 and int32 labels at label_offset, typically pinned - with whole batches ahead
 * of the caller.  dtm_loader_ring_acquire blocks for the next filled buffer and returns its index (-1: error, see
 * dtm_last_error); the buffer belongs to the caller until dtm_loader_ring_release(index) (e.g. once its H2D copy has
 * finished).  dtm_loader_next is not available while a ring runs. */
int dtm_loader_ring_start(void* loader, uint8_t* const* buffers, int n_buffers, uint64_t label_offset);
int dtm_loader_ring_acquire(void* loader);
int dtm_loader_ring_release(void* loader, int index);
uint64_t dtm_loader_records_read(void* loader);
void dtm_loader_destroy(void* loader);

/* ---- peer-memory communication (replaces the reference's gRPC parameter-server traffic: main.py:58-60, :80-93;
 *      mnist_input.py:262-267) ------------------------------------------------------------------------------------ */
/* Cycle budget of the cross-GPU spin-waits on the current device (default 2.4e11, about two minutes: ranks may reach a
 * collective seconds apart).  A wait that exceeds it prints "[dtm watchdog] peer barrier (a rank did not arrive)" and
 * traps - the kernel-level counterpart of the reference's session failing when a worker disappears (main.py:140-146). */
int dtm_set_comm_watchdog(long long cycles);
/* Symmetric memory: cudaMalloc + zero-fill + CUDA IPC handle (64 bytes, host memory) to be exchanged between the
 * ranks of one node; dtm_symm_open maps a peer's allocation into this process. */
int dtm_symm_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int dtm_symm_open(const unsigned char* handle64, void** peer_ptr);
int dtm_symm_close(void* peer_ptr);
int dtm_symm_free(void* ptr);
/* Symmetric memory on the virtual-memory-management API, with NVLink-SHARP multicast binding (csrc/comm/symm_vmm.cu) - the
 * communicator's own allocator (SURVEY X-a); dtm_symm_* above is the cudaMalloc + CUDA-IPC fallback, which cannot be bound
 * to a multicast object.  All sizes are multiples of the granularity dtm_vmm_supported reports.
 *   create -> export_fd -> (fd travels to the peer process) -> import_fd -> map      peer access between processes
 *   create -> grant(other device)                                                   peer access inside one process
 *   mc_create (once, N devices) -> [export/import] -> mc_add_device (every device) -> barrier -> mc_bind (every device, its
 *   own allocation) -> barrier -> map(mc handle)                                    one address that reaches all N copies */
int dtm_vmm_supported(int device, int* vmm, int* multicast, size_t* granularity);
int dtm_vmm_create(size_t bytes, int device, unsigned long long* handle, void** va);
int dtm_vmm_export_fd(unsigned long long handle, int* fd);
int dtm_vmm_import_fd(int fd, unsigned long long* handle);
int dtm_vmm_map(unsigned long long handle, size_t bytes, int device, void** va);
int dtm_vmm_grant(void* va, size_t bytes, int device);
int dtm_vmm_unmap(void* va, size_t bytes);
int dtm_vmm_release(unsigned long long handle);
int dtm_mc_create(int num_devices, size_t bytes, unsigned long long* mc_handle);
int dtm_mc_add_device(unsigned long long mc_handle, int device);
int dtm_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes);
/* One kernel: cross-GPU barrier, reduce-scatter of grad[lo:hi) over peer loads, scale 1/N + weight decay + optimizer
 * + EMA on this rank's shard, all-gather of the new weights by peer stores (bf16 only for seg_gather_bf16 variables,
 * fp32 otherwise), barrier, local bf16 shadow of the fp32-gathered variables + gradient clearing; with do_finish the
 * step bookkeeping of dtm_fused_update.  `ctrl` = 16 bytes of zero-initialised device
 * memory per communicator.  max_ctas > 0 caps the grid (use when the kernel overlaps with compute kernels).
 * All ranks must issue the same sequence of collectives. */
int dtm_fused_allreduce_update(const dtm_peer_ctx* ctx, const dtm_update_config* cfg, dtm_train_state* state, int lo, int hi,
                               float* slot0, float* slot1, float* ema, void* w_bf16, void* ctrl, int do_finish, int max_ctas,
                               void* stream);
/* fp32 sum over ranks of n floats at element offset off of a symmetric buffer (bufs_dev: device array of the N base
 * pointers).  mode 0 one-shot -> out (local); mode 1 two-shot, in place in every rank's buffer (+ copy to out).
 * mc_buf: NVLS multicast address of the same buffer or NULL; when set, mode 0 reads every element with one in-switch
 * reduction and mode 1 reduces its shard in the switch and multicasts the sums back. */
/* Small buckets (the conv variables): every rank sums ALL ranks' gradients of grad[lo:hi) itself (fixed rank order) and
 * updates its own replica - one cross-GPU barrier, no all-gather, ordinary launch.  grad_now_dev: device array of the N
 * ranks' gradient buffers of THIS step; grad_prev_local: this rank's buffer of the previous step, cleared here (gradients
 * of the bucket are double-buffered by step parity), or NULL.  Replaces the Apply ops of mnist_input.py:262-264. */
int dtm_replicated_allreduce_update(const dtm_peer_ctx* ctx, const dtm_update_config* cfg, dtm_train_state* state, int lo, int hi,
                                    float* const* grad_now_dev, float* grad_prev_local, float* slot0, float* slot1, float* ema,
                                    void* w_bf16, void* ctrl, int do_finish, void* stream);
int dtm_allreduce_f32(const dtm_peer_ctx* ctx, float* const* bufs_dev, float* mc_buf, int off, int n, float* out, int mode,
                      void* ctrl, void* stream);
int dtm_broadcast_f32(const dtm_peer_ctx* ctx, float* const* bufs_dev, int off, int n, int root, void* ctrl, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTM_B200_H_ */
