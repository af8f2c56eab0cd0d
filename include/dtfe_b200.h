/* C ABI of libdtfe_cuda.so - the sm_100a kernels of the B200 parameter-server framework.
 *
 * Conventions
 *   - every function returns 0 on success; otherwise dtfe_last_error() (thread-local) explains;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - weight matrices use the reference's tf.layers.dense layout [in, out], row-major;
 *   - "shard" pointers may be peer-mapped HBM of another GPU (see dtfe_ipc_*): kernels read them
 *     with L1-bypassing loads and update them with red.add at .sys scope;
 *   - file:line citations refer to /root/reference (rubythonode/DistributedTensorFlowExample).
 */
#ifndef DTFE_B200_H_
#define DTFE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------ */
const char* dtfe_last_error(void);
int dtfe_abi_version(void);
int dtfe_struct_sizes(size_t* out, int n);   /* sizeof {dtfe_mlp2_launch, dtfe_sync_exchange, dtfe_gemm_tc}; returns 3 */
int dtfe_device_count(int* n);
/* out[6] = {cc major, cc minor, SM count, max opt-in smem/CTA, L2 bytes, cooperative launch} */
int dtfe_device_info(int dev, int* out);

/* Parameter shard = what a ps task owns (reference main.py:67-68: tf.train.Server for job 'ps';
 * TF1 keeps the variables in that process's host RAM).  Here: zeroed cudaMalloc memory on the
 * current device, exportable to other processes on the box as a 64-byte CUDA IPC handle. */
int dtfe_shard_alloc(size_t bytes, void** out);
int dtfe_shard_free(void* p);
int dtfe_ipc_export(void* p, void* handle64);
int dtfe_ipc_open(const void* handle64, void** out);
int dtfe_ipc_close(void* p);
int dtfe_enable_peer(int dev, int peer);

/* Benchmark hygiene: overwrite `bytes` (> L2 size) so the next kernel starts with a cold L2. */
int dtfe_l2_flush(void* buf, size_t bytes, void* stream);

/* global_step += inc with a system-scope atomic; *observed = pre-increment value
 * (reference worker.py:29-32 counter_op, :136 fetch of global_step, :141 sess.run(counter_op)). */
int dtfe_global_step_add(int* global_step, int inc, int* observed_or_null, void* stream);

/* ---- dense layers, fp32 SIMT path ---------------------------------------------------------- */
/* C[M,N] = act(A[M,K] . W[K,N] + bias[N]); act 0 = identity, 1 = sigmoid
 * (reference worker.py:50-53 tf.layers.dense; shard_reads=1 when W/bias are shard memory = fused pull,
 *  worker.py:81-85). */
int dtfe_dense_fwd_f32(const float* A, long lda, const float* W, long ldw, const float* bias, float* C,
                       long ldc, int M, int N, int K, int act, int shard_reads, void* stream);
/* dX[M,K] = (dY[M,N] . W[K,N]^T) * Hprev*(1-Hprev)   (MatMul grad + SigmoidGrad of tf.gradients,
 * reference worker.py:72-73) */
int dtfe_dense_bwd_dx_f32(const float* dY, long lddy, const float* W, long ldw, const float* Hprev, long ldh,
                          float* dX, long lddx, int M, int N, int K, int act, int shard_reads, void* stream);
/* W[K,N] += alpha * X[M,K]^T . dY[M,N]  - gradient push fused with the SGD apply
 * (reference worker.py:75-79 apply_gradients on the global variables).  atomic=1: red.add (lossless
 * Hogwild); atomic=0: unlocked read-modify-write like TF1 use_locking=False.  W may be NULL and/or
 * G non-NULL to get the raw gradient (sync mode). */
int dtfe_dense_bwd_dw_push_f32(const float* X, long ldx, const float* dY, long lddy, float* W, long ldw,
                               float* G, long ldg, int M, int N, int K, float alpha, int atomic, void* stream);
/* b[N] += alpha * sum_m dY[m,:]   (TF1 BiasAddGrad + ApplyGradientDescent) */
int dtfe_bias_grad_push_f32(const float* dY, long lddy, int M, int N, float* b, float* gout, float alpha,
                            int atomic, void* stream);

/* bf16 gradient in; bias may be a bf16 or an fp32 shard (TF1 BiasAddGrad + apply on the wide bf16 path). */
int dtfe_bias_grad_push_bf16(const void* dY_bf16, long lddy, int M, int N, void* b, int b_is_bf16, float* gout,
                             float alpha, void* stream);

/* ---- loss / metrics ------------------------------------------------------------------------ */
/* Softmax cross-entropy forward+backward with per-row loss and argmax-correct flags
 * (reference worker.py:59-68 loss, :87-90 accuracy).  Exactly one of label_ids (int32 [M]) and
 * label_dense (float [M,C], one-hot or soft) is non-NULL.  dZ = (softmax - y) * scale. */
int dtfe_softmax_xent_f32(const float* logits, long ldz, const int* label_ids, const float* label_dense,
                          long ldl, int M, int C, float scale, float* dZ, long lddz, float* rowloss,
                          int* rowcorrect, void* stream);
/* Same with class ids only, additionally writing a bf16 copy of dZ for the tcgen05 backward GEMMs. */
int dtfe_softmax_xent_f32_bf16grad(const float* logits, long ldz, const int* label_ids, int M, int C, float scale,
                                   float* dZ, long lddz, void* dZ_bf16, long lddzb, float* rowloss,
                                   int* rowcorrect, void* stream);
/* Deterministic means: *out_loss = mean(rowloss), *out_acc = mean(rowcorrect), *out_count = sum. */
int dtfe_mean_loss_acc(const float* rowloss, const int* rowcorrect, int M, float* out_loss, float* out_acc,
                       int* out_count, void* stream);

/* ---- flat push / pull / optimizers over peer memory ------------------------------------------ */
/* shard[i] += alpha*grad[i]: the asynchronous push+apply as one kernel (reference worker.py:79). */
int dtfe_push_sgd_f32(float* shard, const float* grad, size_t n, float alpha, int atomic, void* stream);
int dtfe_push_sgd_bf16(void* shard_bf16, const float* grad, size_t n, float alpha, void* stream);
/* dst[i] = shard[i] with L1-bypassing loads: the stand-alone pull (reference worker.py:81-85 sync_op). */
int dtfe_pull_f32(float* dst, const float* shard, size_t n, void* stream);
/* The same pull on a copy engine (cudaMemcpyAsync, source may be a peer-mapped shard): for bulk pulls that run
 * next to GEMMs and must not take SMs from them (reference sync_op, worker.py:81-85). */
int dtfe_memcpy_async(void* dst, const void* src, size_t bytes, void* stream);
/* Adam at the owner: consumes and clears the accumulation slot workers pushed into
 * (no counterpart in the reference, which only has GradientDescentOptimizer, worker.py:71). */
int dtfe_adam_apply_f32(float* param, float* m, float* v, float* slot, size_t n, float lr_t, float beta1,
                        float beta2, float eps, float grad_scale, void* stream);
/* Synchronous SGD: every rank reduces the flat-gradient ranges [seg_begin,seg_end) of the variables it
 * owns out of all ranks' gradient buffers in peer memory and applies them in place (seg_param); flag barriers in peer memory on entry and exit.  One rank per
 * GPU (the kernel spins on peers).  epoch = 0 takes the epoch from local_sync[3] (device side), which lets
 * the launch be replayed from a CUDA graph.  A wait is abandoned - parameters untouched, local_sync[2] != 0 - when
 * *abort_flag becomes non-zero (a rank is leaving) or after timeout_s seconds, so a lost rank cannot hang
 * the others' GPUs.  No counterpart in the reference (async only); semantic model is
 * TF1 SyncReplicasOptimizer. */
int dtfe_sync_rs_apply_f32(const float* const* grads, unsigned* const* flags, int rank, int world, int nseg,
                           float* const* seg_param, const size_t* seg_begin, const size_t* seg_end, float alpha,
                           unsigned epoch, unsigned* local_sync, int* global_step_or_null,
                           const unsigned* abort_flag_or_null, double timeout_s, void* stream);
int dtfe_sync_max_ranks(void);
/* On-device batch sampling (TF1 DataSet.next_batch on resident data): gathers the rows
 * idx[*cursor * B .. +B) of X/labels into a dense batch; the cursor is device memory. */
int dtfe_gather_rows_f32(const float* X, long ldx, const int* labels, const int* idx, const int* cursor, int B,
                         int D, float* Xout, int* Yout, void* stream);

/* Arguments of the in-kernel synchronous exchange (phase D of dtfe_mlp2_step): same protocol and same
 * buffers as dtfe_sync_rs_apply_f32 (exchange buffers + flag arrays from dtfe_shard_alloc, opened by all
 * ranks), but run at the end of every step inside the step kernel, so a launch can cover many synchronous
 * steps.  One rank per GPU. */
typedef struct dtfe_sync_exchange {
  const float* grads[16];     /* every rank's gradient buffer (this rank's own one is where the step writes) */
  unsigned* flags[16];        /* every rank's flag array, 32 u32: [0..15] arrive, [16..31] depart */
  int rank, world;
  int nseg;                   /* variables this rank owns: flat-gradient range and parameter base */
  float* seg_param[8];
  size_t seg_begin[8], seg_end[8];
  float alpha;                /* -lr / world */
  unsigned* epoch;            /* device word: number of completed exchanges (shared with dtfe_sync_rs_apply_f32's local_sync[3]) */
  unsigned* status;           /* device word: set non-zero when a wait was abandoned (1 arrive, 2 depart) */
  const unsigned* abort_flag; /* optional peer-visible stop word */
  unsigned long long timeout_ns;
  int* global_step;           /* non-NULL on the rank that owns global_step */
  /* mode 1 - all-reduce fused with the apply (dtfe_mlp2_step only).  Every rank keeps its own copy of the
   * parameters and of global_step ("home": the shard itself on the rank that owns the variable, a private
   * replica elsewhere) and passes THOSE as W1/b1/W2/b2/global_step.  After its backward pass every CTA
   * publishes the gradients it produced, waits for the same CTA of every other rank (per-CTA flags, no grid
   * barrier), sums all ranks' gradients for exactly the elements it produced - in rank order, so every
   * replica stays bit-identical - and applies alpha * sum to its home copy (the W1 slice stays in shared
   * memory across steps).  Nothing is pulled; the only peer traffic is the flag stores and the gradient
   * loads.  The exchange buffers hold two copies of the flat gradient, grad_stride floats apart, used by
   * alternate steps, which makes a second barrier per step unnecessary.  seg_* and global_step above are
   * ignored. */
  /* mode 2 - the same all-reduce with the apply, low-latency transport: a gradient element travels as the
   * 8-byte pair {value, step number}, stored by the producing thread straight into EVERY rank's receive buffer
   * (ll_recv[p], peer memory) as soon as it is computed.  The pair is the message and its own arrival flag:
   * no fence, no flag store, no round trip - a consumer spins on its LOCAL buffer until the step number of
   * every rank's pair matches.  Receive buffer of a rank: [2 alternating copies][world sources][grad_stride]
   * pairs = 2 * world * grad_stride * 8 bytes, zeroed once, from dtfe_shard_alloc.  cta_flags is unused. */
  int mode;                   /* 0: reduce-scatter by the owners (default), 1: all-reduce with apply, 2: the same, low latency */
  size_t grad_stride;         /* modes 1, 2: floats between the two copies of the flat gradient */
  unsigned* cta_flags[16];    /* mode 1: every rank's per-CTA arrival words, [16 ranks][160 CTAs] u32 */
  float* ll_recv[16];         /* mode 2: every rank's receive buffer */
} dtfe_sync_exchange;

/* ---- the whole reference training step in one cooperative kernel ------------------------------- */
/* pull -> fwd -> xent/accuracy -> bwd -> push+apply -> global_step for nsteps consecutive steps
 * (reference worker.py:129-141, model worker.py:46-57).  Limits: B<=128, H<=128, C<=16, D/ks<=148. */
typedef struct dtfe_mlp2_launch {
  int B, D, H, C, ks;                 /* batch, inputs, hidden, classes, K-slice per CTA (8|16) */
  float *W1, *b1, *W2, *b2;           /* shard memory: [D,H], [H], [H,C], [C] */
  int* global_step;                   /* shard memory, int32 */
  const float* X; long ldx;           /* dataset / batch rows */
  const int* labels;                  /* class ids, indexed like X rows */
  const int* batch_idx;               /* [nsteps*B] row indices or NULL */
  long rows_per_step;                 /* batch_idx==NULL: row = step*rows_per_step + r */
  float* workspace;                   /* dtfe_mlp2_step_workspace_words() 4-byte words, zeroed once */
  float* loss_out; float* acc_out; int* step_out;   /* [nsteps] each */
  float lr; int nsteps; int atomic; int train; int remote;
  long long* dbg;                     /* NULL, or [nsteps][dtfe_mlp2_step_ctas()][16] globaltimer stamps */
  float *gW1, *gb1, *gW2, *gb2;       /* all NULL (async: push), or sync mode: raw gradients stored here, no push,
                                         no global_step increment (dtfe_sync_rs_apply_f32 does both) */
  int* cursor;                        /* NULL, or device int: batches start at batch_idx[*cursor * B]; advanced by
                                         nsteps at kernel end (lets a captured graph walk through the data) */
  const dtfe_sync_exchange* sync_exchange;  /* NULL, or: synchronous SGD with the exchange inside the kernel */
} dtfe_mlp2_launch;
int dtfe_mlp2_step_grid(int D, int ks);          /* W1 slices = split-K factor of layer 1 */
int dtfe_mlp2_step_ctas(int B, int D, int H, int ks);  /* CTAs per launch on the current device (slices + row/helper
                                                          CTAs, at most one per SM); -1 on error */
size_t dtfe_mlp2_step_workspace_words(int B, int D, int H, int C, int ks);
int dtfe_mlp2_step(const dtfe_mlp2_launch* launch, void* stream);

/* ---- thin bf16 layer (output width C <= 16, e.g. 8192 -> 10 logits) ------------------------------------------
 * Three streaming kernels for a layer whose products would waste >= 87 % of a 128-wide tensor-core tile:
 * forward logits = act.W + b (reference worker.py:53), input gradient dX = (dZ.W^T) * H(1-H) (worker.py:72-73),
 * and the weight gradient pushed into the shard, W += alpha * act^T.dZ (worker.py:75-79; partial sums of the row
 * chunks meet in `scratch`, the last CTA per column block pushes once per element).  bf16 matrices with 16-byte
 * aligned rows, K % 8 == 0; W is [K][ldw] with ldw >= 16 when C > 8. */
int dtfe_thin_fwd_bf16(const void* act, long lda, const void* W, long ldw, const void* bias_or_null, float* out, long ldo,
                       int M, int K, int C, void* stream);
int dtfe_thin_dx_bf16(const void* dZ, long lddz, const void* W, long ldw, const void* Hact, long ldh, void* dX, long lddx,
                      int M, int K, int C, void* stream);
size_t dtfe_thin_dw_scratch_words(int M, int K);      /* 4-byte words, zeroed once */
int dtfe_thin_dw_push_bf16(const void* act, long lda, const void* dZ, long lddz, void* Wshard, long ldw, float alpha, int M,
                           int K, int C, float* scratch, void* stream);

/* dtfe_mlp2_step on the tensor cores (same launch struct, workspace and results): X.W1 and X^T.dA run as
 * tcgen05.mma kind::tf32 with fp32 accumulators in TMEM on one cluster of 8 CTAs; CTA i owns a slice of the
 * features, pulls its slice of W1 from the shard with TMA (one step ahead, own update added in shared memory)
 * and pushes -lr*dW1 from the TMEM epilogue with red.add.v4.  tf32 rounds X, W1 and dA to 10 mantissa bits;
 * everything else is fp32.  Asynchronous lossless training only (train = 1, atomic = 1, no gradient buffers, no
 * sync_exchange); `ks` is ignored.  Limits: B <= 104, D <= 832, H <= 128 with H % 4 == 0, C <= 16, 16-byte
 * aligned parameters.  Reference: the same step, worker.py:129-141. */
int dtfe_mlp2_step_tc_supported(int B, int D, int H, int C);
int dtfe_mlp2_step_tc(const dtfe_mlp2_launch* launch, void* stream);

/* Host-fed loop (the reference's feed_dict path, worker.py:133-137): `nsteps` steps, step i fed from the
 * pinned host buffers x_host[i] ([B,D] fp32) / y_host[i] ([B] int32) through double-buffered staging
 * and private copy streams; the loss of step i is copied back to loss_host[i] (pinned).  Up to
 * steps_per_launch (<= 16) consecutive steps share one kernel launch; every batch still gets its own H2D
 * copy and every loss its D2H read.  The feed handle owns the staging buffers, events and streams. */
int dtfe_hostfeed_create(int B, int D, int steps_per_launch, void** out);
int dtfe_hostfeed_destroy(void* handle);
int dtfe_mlp2_train_host(void* handle, const dtfe_mlp2_launch* launch, const float* const* x_host,
                         const int* const* y_host, int nsteps, float* loss_host, void* stream);

/* dtfe_mlp2_step for callers that cannot allocate device memory (no counterpart in the reference).
 * Some users of this library have no way to get a workspace: a worker that lives inside a fixed
 * fraction of a shared GPU, a host program whose own memory pool already owns the device, or code
 * at a point where cudaMalloc (which synchronises the whole device) must not be called.  For them
 * the scratch is a 4 MiB array built into the library image (dtfe_mlp2_step_noalloc_max_words()
 * 4-byte words, held by every GPU that loads the library until the process exits);
 * `launch->workspace` is ignored, inputs and outputs are passed as usual.
 *   - Shapes whose scratch does not fit are refused (non-zero return, dtfe_last_error() names the
 *     size needed and the limit).  The reference model (B=100, D=784, H=100, C=10, ks=8) fits.
 *   - The first call on each device looks up the array's device address and configures the kernel;
 *     make that first call outside any stream capture.  Later calls only enqueue a 16-byte memset
 *     and the launch.  Nothing is ever allocated or freed.
 *   - All calls on a device share the one array: two calls whose kernels could overlap on the GPU
 *     (different streams, or different host threads) corrupt each other.  Serialise them. */
size_t dtfe_mlp2_step_noalloc_max_words(void);
int dtfe_mlp2_step_noalloc(const dtfe_mlp2_launch* launch, void* stream);

/* ---- tcgen05 GEMM with fused epilogues (bf16 operands, fp32 accumulation in TMEM) ---------------- */
/* D[M,N] (op)= A . B through TMA -> swizzled smem -> tcgen05.mma -> TMEM -> tcgen05.ld.
 *   a_mn_major = 0: A is [M][lda], K contiguous;   1: A is [K][lda], M contiguous
 *   b_mn_major = 0: B is [N][ldb], K contiguous;   1: B is [K][ldb], N contiguous
 * so with the reference's [in,out] kernels (worker.py:50-53):
 *   forward  X.W        : A=X (0), B=W (1)      epilogue 1: D = act(acc + bias)      worker.py:50-53
 *   dX       dY.W^T     : A=dY (0), B=W (0)     epilogue 2: D = acc * H*(1-H)        worker.py:72-73
 *   dW push  X^T.dY     : A=X (1), B=dY (1)     epilogue 3: D += alpha*acc (red.add) worker.py:75-79
 * Operands may be peer-mapped shard memory: B = shard weights makes the forward GEMM the pull
 * (worker.py:81-85).  Row pitches must be multiples of 8 elements, bases 16-byte aligned.
 * epilogue: 0 store, 1 bias+act (act: 0 none, 1 sigmoid), 2 sigmoid-grad with H, 3 push into D
 * (D may be NULL when only the raw fp32 gradient G is wanted).  tile_n: 0 auto, 128, 256. */
typedef struct dtfe_gemm_tc {
  int M, N, K;
  const void* A; long lda; int a_mn_major;
  const void* B; long ldb; int b_mn_major;
  void* D; long ldd; int out_bf16;
  int epilogue;
  const void* bias; int bias_bf16; int act;
  const void* H; long ldh;
  float alpha; float* G; long ldg;
  int tile_n;
  int nparts; int rows_per_part;      /* push epilogue into a row-partitioned variable: row m of the result goes to */
  void* D_parts[8];                   /* D_parts[m / rows_per_part] (one block per ps shard); D is then ignored        */
  int tma_push;                       /* push epilogue of the 2-CTA kernel through TMA reduce-adds (bf16 targets whose   */
                                      /* row blocks are multiples of 128 rows): 16 KB bulk adds, meant for peer shards   */
} dtfe_gemm_tc;
int dtfe_gemm_tc_bf16(const dtfe_gemm_tc* p, void* stream);
/* Same struct, fp32 operands consumed by the tensor cores as tf32 (kind::tf32), fp32 H/bias/outputs: fp32
 * parameter shards - the reference's dtype (worker.py:47-53) - go from TMA straight into the MMA.  Row pitches
 * must be multiples of 4 elements; tile_n 0, 128 or 512. */
int dtfe_gemm_tc_tf32(const dtfe_gemm_tc* p, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTFE_B200_H_ */
