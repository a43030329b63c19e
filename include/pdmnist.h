/*
 * pdmnist.h -- C ABI of libpdmnist_b200.so, the sm_100a native core of
 * pytorch_distributed_mnist_b200.
 *
 * Everything below is `extern "C"`, takes raw device/host pointers and a CUDA
 * stream, and returns 0 on success or a negative pdm_status.  After a failure
 * pdm_last_error() (thread-local) describes it.  Python binds these through
 * ctypes (pytorch_distributed_mnist_b200/ops/_lib.py); nothing here links
 * against libtorch, so the library also serves C/C++ callers.
 *
 * "ref:N" cites /root/reference/multi_proc_single_gpu.py line N: the reference
 * behaviour each entry point re-implements.  The reference itself ships no
 * native code; each of these replaces a PyTorch/NCCL/cuBLAS
 * library call on its path.
 */
#ifndef PDMNIST_H_
#define PDMNIST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* pdm_stream_t; /* cudaStream_t */

enum pdm_status {
  PDM_OK = 0,
  PDM_ERR_ARG = -1,     /* bad argument */
  PDM_ERR_CUDA = -2,    /* a CUDA runtime call failed */
  PDM_ERR_STATE = -3,   /* object used before it was set up */
  PDM_ERR_IO = -4,      /* socket / file error */
  PDM_ERR_TIMEOUT = -5, /* a peer did not show up in time */
  PDM_ERR_LIMIT = -6    /* request exceeds a documented maximum */
};

enum pdm_dtype { PDM_U8 = 0, PDM_F32 = 1, PDM_I64 = 2, PDM_I32 = 3 };
enum pdm_opt { PDM_OPT_ADAM = 0, PDM_OPT_SGD = 1, PDM_OPT_NONE = 2 };

const char* pdm_last_error(void);
int pdm_version(void);
/* sizeof() of the structs below as this build sees them, for bindings to check their own declarations:
 * 0 pdm_hyper, 1 pdm_peers, 2 pdm_net_step_args, 3 pdm_net_tc_args, 4 pdm_gemm_args, 5 pdm_cnn_params,
 * 6 pdm_cnn_plan, 7 pdm_reducer_config; 0 for an unknown index. */
size_t pdm_abi_sizeof(int which);
/* 1 when the calling process can see a CUDA device of compute capability 10.x. */
int pdm_device_ok(int device);

/* ------------------------------------------------------------------------
 * Geometry of the reference model `Net` = Linear(784, 10)  (ref:119-126).
 * Flat parameter / optimizer-state buffers are PDM_NET_FLAT floats:
 *   [0, 7840)     fc.weight, row-major (10, 784)
 *   [7840, 7850)  fc.bias
 *   [7850, 7852)  padding (keeps every float4 aligned); never updated
 * ---------------------------------------------------------------------- */
#define PDM_NET_IN 784
#define PDM_NET_OUT 10
#define PDM_NET_PARAMS 7850
#define PDM_NET_FLAT 7852
/* Per-CTA stride of the partial-gradient scratch, in floats. */
#define PDM_NET_SCRATCH_STRIDE 7856
#define PDM_NET_MAX_CTAS 148
/* Floats of scratch pdm_net_train_step needs for a launch of `ctas` CTAs. */
#define PDM_NET_SCRATCH_FLOATS(ctas) ((size_t)(ctas) * PDM_NET_SCRATCH_STRIDE)

/* Optimizer hyper-parameters, by value per launch.
 * Adam matches torch.optim.Adam (ref:191): bias-corrected, eps outside the
 * sqrt, L2 weight decay added to the gradient.  SGD matches torch.optim.SGD
 * with momentum/dampening 0/nesterov off (the commented-out ref:192-194). */
typedef struct pdm_hyper {
  int32_t opt;        /* pdm_opt */
  float lr;
  float beta1;        /* Adam beta1, or SGD momentum */
  float beta2;
  float eps;
  float weight_decay;
  int64_t step;       /* 1-based index of THIS update (Adam bias correction;
                         SGD uses step==1 to seed the momentum buffer) */
} pdm_hyper;

/* Peer links for the in-kernel all-reduce (filled by pdm_comm_*; world==1 or
 * a zeroed struct means single GPU). */
#define PDM_MAX_RANKS 8
typedef struct pdm_peers {
  int32_t world;
  int32_t rank;
  uint32_t seq;                      /* flag value of this collective; never 0 */
  uint32_t pad_;
  uint64_t* inbox[PDM_MAX_RANKS];    /* inbox[p]: rank p's LL inbox, mapped here */
  int32_t* status;                   /* device word: nonzero after a peer timeout */
  uint64_t timeout_ns;               /* 0 = wait forever */
} pdm_peers;

/* One fused training step of `Net` (ref:83-95 in ONE launch):
 *   logits = x.W^T + b; loss = cross_entropy(logits, y) (mean over B);
 *   dW, db; all-reduce over `peers` (sum in rank order, then * 1/world);
 *   optimizer update of params/m/v in place.
 *  x        [B,784] device, PDM_U8 (raw MNIST bytes, normalised in-kernel as
 *           (x/255-0.1307)/0.3081, ref:134) or PDM_F32 (already normalised)
 *  y        [B] device labels, PDM_U8 / PDM_I64 / PDM_I32
 *  params,m,v  PDM_NET_FLAT floats each (m,v may be NULL for PDM_OPT_NONE;
 *           v may be NULL for SGD)
 *  grad_out optional PDM_NET_FLAT floats: receives the reduced mean gradient
 *  scratch  PDM_NET_SCRATCH_FLOATS(ctas) floats of device scratch
 *  sync     3 x uint32 device words, zero-initialised once (grid barrier + step-to-step flag)
 *  metrics  optional 4 doubles, device: += {loss*B, correct, B, 0}
 *  step_out optional 2 floats, device or mapped host: {mean loss, #correct}
 *  ctas     0 = choose from B; else 1..PDM_NET_MAX_CTAS
 * All CTAs of the launch must fit on the GPU at once (they meet at a grid barrier): the launch is
 * a programmatic-dependent one when the grid fits one CTA per SM, else (or with
 * PDM_NET_LAUNCH=coop) a cooperative one, which the driver refuses if it cannot be co-resident. */
typedef struct pdm_net_step_args {
  const void* x; int32_t x_dtype;
  const void* y; int32_t y_dtype;
  int32_t batch;
  int32_t ctas;
  float* params; float* m; float* v;
  float* grad_out;
  float* scratch;
  uint32_t* sync;
  double* metrics;
  float* step_out;
  pdm_hyper hyper;
  pdm_peers peers;
  const int32_t* row_index;   /* optional, device: batch row r is row row_index[r] of x and y.  Lets the
                                 kernel gather a shuffled batch itself from a dataset that is resident in
                                 device memory -- or in pinned, mapped host memory (zero-copy over PCIe). */
  int32_t flag_dependency;    /* 1: the caller can order this step after the previous one through sync[2] (a
                                 count of finished CTAs) instead of grid completion; 2: this launch is counted in
                                 sync[2] but still waits for its predecessor by grid completion (use for the first
                                 step of a run, when uncounted work may precede it on the stream).  Used by default
                                 in multi-GPU jobs, where it is faster (12.6 vs 17.8 us/step at 2 GPUs together
                                 with the pulled LL exchange), not on one GPU (10.3 vs 8.7); PDM_NET_DEP=grid|flag
                                 forces one.  Needs `sync` to be at least 3 words and done_before to be exact. */
  uint32_t done_before;       /* sum of pdm_net_planned_ctas() over all earlier steps issued on these sync words */
  int32_t inputs_stable;      /* 1: x, y and row_index were NOT written by the kernel that precedes this call on the
                                 stream (they come from a memcpy, or are a resident dataset).  The launch is a
                                 programmatic dependent one, and with this flag the kernel loads its first input
                                 tile before it waits for its predecessor (saves ~1 us/step).  Leave 0 when another
                                 kernel (a gather, a normalise, .contiguous()) produces the inputs right before:
                                 only the wait makes that kernel's writes visible. */
} pdm_net_step_args;

int pdm_net_train_step(const pdm_net_step_args* args, pdm_stream_t stream);
/* CTAs pdm_net_train_step will launch for a batch when args->ctas == 0. */
int pdm_net_default_ctas(int batch);
/* CTAs it will launch for (batch, args->ctas). */
int pdm_net_planned_ctas(int batch, int want_ctas);

/* The same fused step for callers that must not allocate.  (This entry point has no
 * counterpart in the reference, which ships no native code and never manages device memory.)
 *
 * Who needs it: a host that gave its whole device-memory budget to its own pool at start-up
 * and forbids allocation afterwards (cudaMalloc/cudaFree also synchronise the device, which
 * stalls other streams of a serving or real-time loop); code called from inside another
 * allocator or under its lock; a recovery path after an out-of-memory error.  None of them can
 * provide pdm_net_train_step's `scratch`/`sync`, so here those live in a fixed __device__ array
 * inside the library: args->scratch and args->sync are ignored (may be NULL); x, y, params,
 * m, v, grad_out, metrics, step_out are still the caller's.
 *
 * Limit: the array holds the partial gradients of at most PDM_NET_STATIC_MAX_CTAS CTAs
 * (PDM_NET_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until exit).
 * args->ctas above that returns PDM_ERR_LIMIT; ctas == 0 picks the usual count clipped to the
 * limit.  Every batch size works.
 *
 * First call on a device (with that device current) looks the array's device address up and
 * caches it: make that call before any stream capture.  Nothing here allocates or frees; the
 * PDM_NET_TRACE clock tracing (whose buffer is allocated on first use) is ignored on this path.
 * With peers the exchange form (PDM_NET_LL, default pull) is chosen exactly as in
 * pdm_net_train_step, so ranks may mix the two entry points; consecutive static steps are always
 * ordered by grid completion (flag_dependency / done_before are ignored).
 * All calls on a device share the array and it keeps the previous call's contents; what must
 * be zero is cleared in stream order on every call, but calls that overlap on one GPU race on
 * the array -- serialise them. */
#define PDM_NET_STATIC_MAX_CTAS 64
#define PDM_NET_STATIC_SCRATCH_BYTES (PDM_NET_STATIC_MAX_CTAS * PDM_NET_SCRATCH_STRIDE * 4)
int pdm_net_train_step_static(const pdm_net_step_args* args, pdm_stream_t stream);

/* The training loop of `Net` as ONE persistent kernel on the tensor cores (net_tc.cu): a cluster
 * of PDM_TC_CLUSTER CTAs keeps the model (weights, optimizer moments) on chip and runs `nsteps`
 * steps of ref:83-95 -- tcgen05.mma forward and weight-gradient GEMMs with TMEM accumulators,
 * cross-CTA logit reduction through distributed shared memory, in-kernel LL all-reduce with the
 * peers, Adam/SGD -- in a single launch.  Two input modes:
 *   pool  (row_index == NULL) step i reads batch ((first_batch + i) mod (pool_rows / batch)) of x / y
 *   plan  (row_index != NULL) step i reads rows row_index[i*batch ...] (last step ragged if count is
 *         not a multiple of batch); x / y may be device memory or pinned, mapped host memory
 *  x [rows,784] PDM_U8 raw pixels (normalised in-kernel), 16-byte aligned; y [rows] PDM_U8
 *  params,m,v PDM_NET_FLAT floats, read at the start and written back after the last step
 *  grad_out   optional PDM_NET_FLAT floats: the LAST step's reduced mean gradient
 *  logits_out optional [batch,10]: the LAST step's logits (before the update)
 *  step_out   optional [nsteps][2] {mean loss, #correct}, device or mapped host memory
 *  metrics    optional 4 doubles, device: += {loss*B, correct, B, 0} per step
 *  hyper.step 1-based index of the first update; peers.seq the LL sequence number of the first
 *  step (the run consumes nsteps consecutive numbers).
 * Math: raw pixels are exact in bf16; weights and dlogits enter the MMAs as three bf16 terms
 * (hi + mid + lo), so products are exact and accumulation is fp32 -- fp32-faithful, not bf16.
 * Softmax uses ex2.approx, the Adam update sqrt.approx / div.approx (<= 2 ulp each). */
#define PDM_TC_CLUSTER 13
/* Optional feeder ring (plan mode): device scratch of PDM_TC_RING_BYTES.  With it the launch carries a
 * second cluster that copies WHOLE rows of the coming blocks (784 contiguous bytes each: full-size PCIe
 * reads when x / y are pinned host memory) into the ring, and the training cluster reads its 64-byte
 * column slices from there through L2 instead of issuing 64-byte reads over PCIe. */
#define PDM_TC_RING_SLOTS 4
#define PDM_TC_RING_SLOT_BYTES (512 * 784 + 512)
#define PDM_TC_RING_BYTES ((size_t)PDM_TC_RING_SLOTS * PDM_TC_RING_SLOT_BYTES + 256)
#define PDM_TC_MAX_LAUNCH_STEPS 8192   /* longer runs are cut into launches of at most this many steps */
typedef struct pdm_net_tc_args {
  const uint8_t* x;
  const uint8_t* y;
  const int32_t* row_index;
  int64_t pool_rows;
  int64_t first_batch;
  int64_t count;
  int32_t batch;
  int64_t nsteps;
  float* params; float* m; float* v;
  float* grad_out;
  float* logits_out;
  double* metrics;
  float* step_out;
  pdm_hyper hyper;
  pdm_peers peers;
  void* ring;            /* optional feeder ring, see PDM_TC_RING_BYTES (NULL: every CTA reads its slices directly) */
  size_t ring_bytes;
} pdm_net_tc_args;
/* 1 when `device` can launch the cluster kernel (compute capability 10.x, 13-CTA clusters). */
int pdm_net_tc_supported(int device);
int pdm_net_tc_run(const pdm_net_tc_args* args, pdm_stream_t stream);
/* PDM_TC_TRACE=1 only: per-CTA clock stamps of the last run, [13][64][32] int64 (benchmarks/tc_trace.py). */
int pdm_net_tc_trace_read(long long* out);

/* Forward + loss + top-1 only (ref:106-114).  metrics (4 doubles, device) is
 * accumulated atomically: += {loss*B, correct, B, 0}.  Any grid size; no scratch. */
int pdm_net_eval_step(const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                      const float* params, float* logits_out /* optional [B,10] */,
                      double* metrics, pdm_stream_t stream);

/* logits[B,10] = x.W^T + b  (Net.forward, ref:126). */
int pdm_net_forward(const void* x, int x_dtype, int batch, const float* params,
                    float* logits, pdm_stream_t stream);

/* ------------------------------------------------------------------------
 * Flat fused optimizers (one launch for the whole model; replaces the eight
 * _foreach launches of torch.optim.Adam).
 * grad may alias a communication bucket.  grad_scale multiplies the gradient
 * first (1/world for a summed bucket).
 * ---------------------------------------------------------------------- */
int pdm_optim_step(float* params, const float* grad, float* m, float* v, size_t n,
                   const pdm_hyper* hyper, float grad_scale, pdm_stream_t stream);

/* ------------------------------------------------------------------------
 * Fused cross-entropy (ref:88): loss, top-1 count and dlogits in one pass.
 *  logits [B,C] fp32, C <= 32.  dlogits may be NULL (evaluation).
 * ---------------------------------------------------------------------- */
int pdm_cross_entropy(const float* logits, const void* y, int y_dtype, int batch, int classes,
                      float* loss_mean /* device scalar */, float* dlogits, double* metrics,
                      pdm_stream_t stream);

/* ------------------------------------------------------------------------
 * tcgen05 / TMA GEMM at tf32 with fp32 accumulation (gemm_tcgen05.cu):
 *     D[M,N] (+)= A[M,K] . B[N,K]^T
 * The GEMM-shaped layers of the CNN extension: fc1 and, through
 * the shifted-row implicit-GEMM modes, the 3x3 convolutions.  Replaces the
 * cuBLAS / cuDNN calls the stock stack would make.
 *  a, b      row-major fp32 matrices as stored: [a_rows][a_cols], pitch lda
 *            (elements; base and pitch 16-byte aligned)
 *  *_mn_major 0: stored [MN][K] (K contiguous); 1: stored [K][MN]
 *  mode      0 plain GEMM
 *            1 conv fwd/dgrad: A = NHWC activations as [pixels][C] (K-major);
 *              k-block kb -> tap = kb / cblocks, A rows shifted by
 *              shift_sign * ((tap/3)*img_w + tap%3); B = [N][taps*C] K-major
 *            2 conv wgrad: A = dY [pixels][Cout] and B = X [pixels][Cin], both
 *              MN-major; N = taps*Cin, column block j reads X shifted by its tap
 *  epilogue  + bias[n], ReLU, zero where mask[m][n] <= 0; `accumulate` or
 *            splits > 1 switches the store to atomic adds (d must be initialised)
 * ---------------------------------------------------------------------- */
typedef struct pdm_gemm_args {
  const float* a; int64_t a_rows, a_cols, lda; int32_t a_mn_major;
  const float* b; int64_t b_rows, b_cols, ldb; int32_t b_mn_major;
  float* d; int64_t ldd;
  int32_t M, N, K;
  const float* bias; int32_t relu;
  const float* mask; int64_t ld_mask;
  int32_t block_n;      /* 0 = auto; multiple of 32, <= 256 */
  int32_t splits;       /* split-K factor (0/1 = none) */
  int32_t accumulate;
  int32_t mode, img_w, cblocks, shift_sign;
} pdm_gemm_args;
int pdm_gemm_tf32(const pdm_gemm_args* args, pdm_stream_t stream);

/* ------------------------------------------------------------------------
 * Non-GEMM CNN layers (cnn_kernels.cu), each fused with its neighbours.  The
 * reference has no CNN; these serve the BASELINE.json extension
 * and replace what cuDNN / ATen would run for it.  Activations are NHWC viewed
 * as [B*26*26, C] (26-wide rows; conv2's valid outputs are the 24x24 corner).
 * Dropout masks are Philox4x32-10(seed, element index) and are recomputed in
 * the backward kernels, never stored.
 * ---------------------------------------------------------------------- */
/* conv1 1->c1 3x3 + bias + ReLU.  x [B,784] PDM_U8 (normalised in-kernel) or PDM_F32; w [c1][9]. */
int pdm_conv1_fwd(const void* x, int x_dtype, const float* w, const float* bias, float* out,
                  int batch, int c1, pdm_stream_t stream);
/* dw [c1][9] += sum dz*x, db [c1] += sum dz   (dz already ReLU-masked). */
int pdm_conv1_bwd(const void* x, int x_dtype, const float* dz, float* dw, float* db, int batch,
                  int c1, pdm_stream_t stream);
/* max-pool 2x2 + bias + ReLU + dropout(p) over raw conv accumulators z [B*676, C];
 * out [B, C*144] in NCHW-flatten order; code [B*C*144] replay bytes for the backward. */
int pdm_pool_fwd(const float* z, const float* bias, float* out, uint8_t* code, int batch, int C,
                 float p, uint64_t seed, pdm_stream_t stream);
/* dy [B*676, C] (fully written, zero outside the arg-max positions); dbias [C] += (may be NULL). */
int pdm_pool_bwd(const float* dp, const uint8_t* code, float* dy, float* dbias, int batch, int C,
                 float p, uint64_t seed, pdm_stream_t stream);
/* logits [B,10] = dropout(h, p) . w^T + bias,  h [B,H], w [10][H]. */
int pdm_fcout_fwd(const float* h, const float* w, const float* bias, float* logits, int batch,
                  int H, float p, uint64_t seed, pdm_stream_t stream);
/* dz1 [B,H] = (dl . w) * dropout * (h > 0); dw [10][H] +=, db [10] +=, db1 [H] += colsum(dz1). */
int pdm_fcout_bwd(const float* dl, const float* h, const float* w, float* dz1, float* dw, float* db,
                  float* db1, int batch, int H, float p, uint64_t seed, pdm_stream_t stream);
/* z [rows][cols] = relu(z + bias[col]) in place (epilogue of a split-K GEMM). */
int pdm_bias_relu(float* z, const float* bias, int64_t rows, int cols, pdm_stream_t stream);
/* dir 0: w [Co][Ci][9] -> a [Co][9][Ci], b [Ci][9][Co] (b may be NULL);  dir 1: a [Co][Ci][9] += src [Co][9][Ci]. */
int pdm_conv_relayout(const float* src, float* a, float* b, int cout, int cin, int dir, pdm_stream_t stream);

/* ------------------------------------------------------------------------
 * Native executor for one CNN step (cnn_step.cu): every kernel of the forward, loss and backward
 * pass issued from one call, gradients written straight into the flat gradient buffer, and the
 * reducer's per-bucket fused all-reduce + optimizer launched on a side stream as soon as a bucket
 * is final (static schedule; the role of the c10d Reducer's hooks for this model).
 * Parameter pointers follow the module's layout (conv weights [Co][Ci][3][3], fc weights [out][in]).
 * ---------------------------------------------------------------------- */
typedef struct pdm_cnn_ctx pdm_cnn_ctx;
typedef struct pdm_cnn_params {
  float *conv1_w, *conv1_b, *conv2_w, *conv2_b, *fc1_w, *fc1_b, *fc2_w, *fc2_b;
} pdm_cnn_params;
struct pdm_comm;
struct pdm_nvls;
typedef struct pdm_cnn_plan {
  int32_t n_buckets;            /* 1..4, in launch order (bucket 0 = the LAST parameters) */
  int32_t max_ctas;             /* CTAs for each reducer kernel (0 = default) */
  int64_t lo[4], hi[4];         /* float ranges of the buckets inside the flat buffers */
  struct pdm_comm* comm;        /* NULL: no peers, optimizer only */
  size_t params_off, grads_off, m_off, v_off;   /* comm != NULL: byte offsets of the flat buffers in its data area */
  float *params, *grads, *m, *v;                /* comm == NULL: the flat buffers themselves */
  pdm_hyper hyper;
  float grad_scale;             /* 1/world */
  struct pdm_nvls* nvls;        /* non-NULL: the flat buffers live in this NVLS region (offsets as above, into ITS data
                                   area) and every bucket is reduced in the switch (pdm_nvls_allreduce_optim) */
} pdm_cnn_plan;
int pdm_cnn_create(pdm_cnn_ctx** out, int device, int width, int max_batch);
int pdm_cnn_eval_step(pdm_cnn_ctx* c, const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                      const pdm_cnn_params* params, float* logits_out, double* metrics, pdm_stream_t stream);
int pdm_cnn_train_step(pdm_cnn_ctx* c, const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                       const pdm_cnn_params* params, const pdm_cnn_params* grads, float p1, float p2,
                       uint64_t seed, double* metrics, float* loss_out, const pdm_cnn_plan* plan,
                       pdm_stream_t stream);
int pdm_cnn_destroy(pdm_cnn_ctx* c);

/* Reducer bucket assignment (c10d _compute_bucket_assignment_by_size): reverse parameter
 * order, a bucket closes once it reaches its cap, the first bucket uses first_cap_bytes.
 * out_bucket[i] = bucket of parameter i; returns the number of buckets. */
int pdm_assign_buckets(const int64_t* numels, int n, int64_t first_cap_bytes, int64_t cap_bytes,
                       int elem_size, int32_t* out_bucket);

/* ------------------------------------------------------------------------
 * Peer-memory communicator (comm.cu).  Replaces ProcessGroupNCCL + NCCL +
 * the c10d Reducer on the gradient path (ref:167-168, ref:188-189, ref:91;
 * SURVEY N2-N4, N7, K9, K14-K16).  One process per GPU; every rank allocates
 * a symmetric region, exports a 64-byte CUDA IPC handle, and maps all peers.
 * The handle exchange itself goes through whatever key-value store the caller
 * has (the tcp:// rendezvous store).  All ranks must issue the same sequence
 * of collectives.  Offsets are bytes into the region's data area.
 * ---------------------------------------------------------------------- */
typedef struct pdm_comm pdm_comm;
int pdm_comm_create(pdm_comm** out, int rank, int world, int device, size_t data_bytes);
int pdm_comm_handle(pdm_comm* c, void* out64);                 /* this rank's IPC handle */
int pdm_comm_connect(pdm_comm* c, const void* handles);        /* world x 64 bytes, rank order */
void* pdm_comm_data(pdm_comm* c);                              /* local data area (device ptr) */
size_t pdm_comm_data_bytes(pdm_comm* c);
int pdm_comm_set_timeout(pdm_comm* c, double seconds);         /* 0 = wait forever; default 30 s */
int pdm_comm_peers(pdm_comm* c, pdm_peers* out, int advance);  /* links for pdm_net_train_step */
int pdm_comm_peers_block(pdm_comm* c, pdm_peers* out, int64_t nsteps); /* links + nsteps sequence numbers (pdm_net_tc_run) */
int pdm_comm_status(pdm_comm* c, int* status);                 /* 0, or 1 + rank that timed out */
int pdm_comm_barrier(pdm_comm* c, pdm_stream_t stream);
/* Every rank ends up with root's bytes [offset, offset+bytes) (DDP's initial
 * parameter broadcast, K15): non-root ranks pull over NVLink. */
int pdm_comm_broadcast(pdm_comm* c, size_t offset, size_t bytes, int root, pdm_stream_t stream);
/* Fused all-reduce (rank-ordered sum) * grad_scale + optimizer update.
 *  algo 0 auto / 1 one-shot (every rank reduces and updates everything) /
 *       2 two-shot (rank r reduces and updates slice r, then stores the new
 *         parameters into every replica; m/v of other slices stay stale here)
 *  hyper->opt == PDM_OPT_NONE: plain all-reduce, result at out_off (must
 *  differ from grad_off in one-shot mode).
 *  max_ctas 0 = up to one CTA per SM (the kernel is launched cooperatively). */
int pdm_comm_allreduce_optim(pdm_comm* c, size_t grad_off, size_t out_off, size_t m_off, size_t v_off,
                             size_t n_floats, const pdm_hyper* hyper, float grad_scale, int algo,
                             int max_ctas, pdm_stream_t stream);
int pdm_comm_destroy(pdm_comm* c);

/* ------------------------------------------------------------------------
 * Gradient reducer for arbitrary models (reducer.cu; the role of c10d's Reducer).
 * Gradients are slices of a flat buffer inside the communicator's (or NVLS region's) data area;
 * a bucket is a 4-float-aligned range of it.  pdm_reducer_mark_ready is called from autograd's
 * post-accumulate hook; when a bucket's last gradient has landed -- and every earlier bucket has
 * been launched -- the reducer orders the communication stream after the compute stream and
 * launches ONE fused all-reduce + 1/W + optimizer kernel over the range (hyper->opt NONE / NULL:
 * the averaged gradient in place), and after the last bucket hands the compute stream the result.
 * comm == nvls == NULL: bookkeeping only; the caller reduces the buckets reported in first/count.
 * ---------------------------------------------------------------------- */
struct pdm_nvls;
typedef struct pdm_reducer pdm_reducer;
typedef struct pdm_reducer_config {
  int32_t n_params, n_buckets;
  const int32_t* bucket_of;        /* [n_params] */
  const int64_t* bucket_lo;        /* [n_buckets] float offsets into the flat buffers */
  const int64_t* bucket_hi;
  size_t grads_off, params_off, m_off, v_off;   /* byte offsets of the flat buffers in the data area */
  pdm_comm* comm;                  /* peer-memory communicator, or NULL */
  struct pdm_nvls* nvls;           /* NVLS region (takes precedence), or NULL */
  int32_t world;
  int32_t device;                  /* CUDA device of the streams; -1 without native communication */
} pdm_reducer_config;
int pdm_reducer_create(pdm_reducer** out, const pdm_reducer_config* cfg);
int pdm_reducer_arm(pdm_reducer* r);   /* start of a training forward; PDM_ERR_STATE if the last pass is unfinished */
int pdm_reducer_mark_ready(pdm_reducer* r, int32_t param, const pdm_hyper* hyper, pdm_stream_t compute,
                           pdm_stream_t comm_stream, int32_t* first, int32_t* count);
/* find_unused_parameters: list the parameters without a gradient in this pass (zero their slices), then
 * finalize marks them ready and launches the remaining buckets. */
int pdm_reducer_unmarked(pdm_reducer* r, int32_t* out, int32_t* count);
int pdm_reducer_finalize(pdm_reducer* r, const pdm_hyper* hyper, pdm_stream_t compute, pdm_stream_t comm_stream,
                         int32_t* first, int32_t* count);
/* Order in which the parameters became ready in this pass, and whether a completed bucket had to wait for
 * an earlier one (input of the bucket rebuild after the first iteration). */
int pdm_reducer_ready_order(pdm_reducer* r, int32_t* out, int32_t* n, int32_t* held_back);
int pdm_reducer_state(pdm_reducer* r, int32_t* next, int64_t* launches, int32_t* pending, int32_t* order,
                      int32_t* n_order);
int pdm_reducer_destroy(pdm_reducer* r);

/* ------------------------------------------------------------------------
 * NVLS (multicast / multimem) region (nvls.cu): the in-switch path NCCL takes for large
 * all-reduces.  The reducer uses it for models of >= 8 MB at 4 or more GPUs
 * (parallel/ddp.py; PDM_NVLS=0 / 1 forces it off / on); it moves 1 + 1/W of the payload per GPU
 * and direction where the peer-memory two-shot kernel moves 2 (W-1)/W.
 * A symmetric region of VMM memory bound to a multicast object, so that
 * `multimem.ld_reduce` sums an address over all replicas inside the NVSwitch and
 * `multimem.st` writes all replicas with one store.  Bootstrap, in order:
 *   every rank : pdm_nvls_create
 *   rank 0     : pdm_nvls_export_fd, descriptor sent to the others (SCM_RIGHTS)
 *   others     : pdm_nvls_import_fd
 *   every rank : pdm_nvls_add_device, host barrier, pdm_nvls_bind, host barrier
 * pdm_nvls_allreduce_optim has the contract of the two-shot
 * pdm_comm_allreduce_optim; offsets are bytes into pdm_nvls_data(). */
typedef struct pdm_nvls pdm_nvls;
int pdm_nvls_supported(int device);                  /* 1 yes, 0 no, <0 error */
int pdm_nvls_create(pdm_nvls** out, int rank, int world, int device, size_t data_bytes);
int pdm_nvls_export_fd(pdm_nvls* nv, int* fd);
int pdm_nvls_import_fd(pdm_nvls* nv, int fd);
int pdm_nvls_add_device(pdm_nvls* nv);
int pdm_nvls_bind(pdm_nvls* nv);
void* pdm_nvls_data(pdm_nvls* nv);
size_t pdm_nvls_data_bytes(pdm_nvls* nv);
void pdm_nvls_set_timeout(pdm_nvls* nv, double seconds);
int pdm_nvls_status(pdm_nvls* nv, int* status);
int pdm_nvls_allreduce_optim(pdm_nvls* nv, size_t grad_off, size_t out_off, size_t m_off, size_t v_off,
                             size_t n_floats, const pdm_hyper* hyper, float grad_scale, pdm_stream_t stream);
int pdm_nvls_destroy(pdm_nvls* nv);

/* ------------------------------------------------------------------------
 * TCP key-value rendezvous store (store.cu; host code only).  The role of the
 * c10d TCPStore behind the reference's tcp:// init method (ref:167-168,
 * ref:323-327) as far as this framework needs it: publishing the
 * communicator's IPC handles and counting arrivals.  The rank with is_server
 * hosts the store on `port`; every rank connects as a client.
 * ---------------------------------------------------------------------- */
typedef struct pdm_store pdm_store;
int pdm_store_create(pdm_store** out, const char* host, int port, int is_server, double timeout_s);
int pdm_store_set(pdm_store* s, const char* key, const void* val, size_t len);
/* Blocks until `key` exists (or timeout).  *len = value size; at most cap bytes copied. */
int pdm_store_get(pdm_store* s, const char* key, void* buf, size_t cap, size_t* len);
int pdm_store_add(pdm_store* s, const char* key, int64_t delta, int64_t* result);
int pdm_store_destroy(pdm_store* s);

/* ------------------------------------------------------------------------
 * Native epoch executor for the fused Net path (engine.cu).  Replaces the
 * DataLoader worker processes, the per-step pageable H2D copies and the two
 * per-step host syncs of Trainer.train()/evaluate() (ref:77-116; SURVEY N13,
 * N14, K1, K2, K12, K13).
 * ---------------------------------------------------------------------- */
typedef struct pdm_engine pdm_engine;
int pdm_engine_create(pdm_engine** out, int device, int max_batch, int slots, int workers);
int pdm_engine_bind(pdm_engine* e, float* params, float* m, float* v, pdm_comm* comm /* may be NULL */);
/* Host-resident dataset: images [n,784] uint8, labels [n] uint8.  Not copied; must outlive the runs. */
int pdm_engine_set_dataset(pdm_engine* e, const uint8_t* images, const uint8_t* labels, int64_t n);
/* ceil(count/batch) steps in `indices` order.  train: hyper->step = index of the first update;
 * step_results (host, [nsteps][2]) gets {mean loss, correct} per step.  Blocks until done. */
int pdm_engine_run(pdm_engine* e, const int64_t* indices, int64_t count, int batch, int train,
                   const pdm_hyper* hyper, int ctas, float* step_results, float* device_ms);
/* nsteps training steps over a device-resident pool of samples; no host traffic.  With raw uint8
 * inputs and ctas == 0 the pass is ONE launch of the persistent tensor-core kernel (pdm_net_tc_run)
 * on devices that support it (PDM_NET_TC=0 disables); the same holds for pdm_engine_run_gather. */
int pdm_engine_run_device(pdm_engine* e, const uint8_t* dev_images, const uint8_t* dev_labels,
                          int64_t pool_samples, int batch, int64_t nsteps, const pdm_hyper* hyper,
                          int ctas, float* device_ms);
/* The same with typed inputs: x PDM_U8 or PDM_F32 (already normalised), y PDM_U8 / PDM_I64 / PDM_I32. */
int pdm_engine_run_device_typed(pdm_engine* e, const void* dev_images, int x_dtype, const void* dev_labels,
                                int y_dtype, int64_t pool_samples, int batch, int64_t nsteps,
                                const pdm_hyper* hyper, int ctas, float* device_ms);
/* In-kernel gather path (B200-first: the dataset fits HBM many times over).  place: mode 0 copies
 * the dataset to device memory once; mode 1 page-locks + maps the caller's host arrays so the step
 * kernel reads its rows straight from host memory over PCIe (zero-copy).  The last four placements
 * are cached (key: host pointers, length, mode); placing one of them again only selects it.
 * run_gather uploads the index plan (4 B/sample) once, then issues one launch per step. */
int pdm_engine_place_dataset(pdm_engine* e, const uint8_t* images, const uint8_t* labels, int64_t n, int mode);
int pdm_engine_run_gather(pdm_engine* e, const int64_t* indices, int64_t count, int batch,
                          const pdm_hyper* hyper, int ctas, float* step_results, float* device_ms);
int pdm_engine_metrics(pdm_engine* e, double out[4], int reset);
/* Kernels launched by the last pdm_engine_run* call (1 when the persistent tensor-core kernel took the pass). */
int64_t pdm_engine_last_launches(pdm_engine* e);
void* pdm_engine_scratch(pdm_engine* e);
void* pdm_engine_sync_words(pdm_engine* e);
void* pdm_engine_metrics_ptr(pdm_engine* e);
int pdm_engine_destroy(pdm_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* PDMNIST_H_ */
