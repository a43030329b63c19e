// cnn_step.cu -- native executor for one training / evaluation step of the CNN extension.
//
// The autograd route (ops/cnn_ops.py) costs ~0.7 ms of Python per step for ~0.8 ms of GPU work
// at width 1 (profiles/README.md section 3).  This executor issues the same kernels -- conv1,
// tcgen05 implicit-GEMM conv2, fused pool, tcgen05 fc1, fused fc2, cross-entropy and their
// backward counterparts -- from one C call, writes every gradient straight into the model's
// flat gradient buffer (no autograd accumulate pass), and launches the reducer's fused
// all-reduce + optimizer kernel for each bucket on a side stream as soon as the bucket's last
// gradient exists, so bucket 0 (fc2, fc1) overlaps the convolution backward.
// It is the CNN analogue of engine.cu for Net and replaces, for this model, the c10d Reducer's
// hook/bucket/launch logic with a static schedule known at build time.
#include "common.cuh"

#include <new>

extern "C" int pdm_comm_allreduce_optim(struct pdm_comm* c, size_t grad_off, size_t out_off, size_t m_off, size_t v_off,
                                        size_t n_floats, const pdm_hyper* h, float grad_scale, int algo, int max_ctas,
                                        pdm_stream_t stream);
extern "C" void* pdm_comm_data(struct pdm_comm* c);

namespace {
constexpr int kPix = 676;   // 26 x 26 conv1 outputs per image
}

struct pdm_cnn_ctx {
  int device = 0, width = 1, max_batch = 0;
  int c1 = 32, c2 = 64, H = 128;
  float *a1 = nullptr, *z2 = nullptr, *pooled = nullptr, *h1 = nullptr, *logits = nullptr, *dlogits = nullptr;
  float *dz1 = nullptr, *dpool = nullptr, *wk = nullptr, *wd = nullptr, *gw = nullptr, *loss = nullptr;
  uint8_t* code = nullptr;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_bucket[2] = {nullptr, nullptr};
  cudaEvent_t ev_comm_done = nullptr;
};

namespace {

int gemm(const float* a, int64_t ar, int64_t ac, int64_t lda, int amn, const float* b, int64_t br, int64_t bc, int64_t ldb,
         int bmn, float* d, int64_t ldd, int M, int N, int K, const float* bias, int relu, const float* mask, int64_t ldm,
         int block_n, int splits, int mode, int cblocks, int shift_sign, cudaStream_t s) {
  pdm_gemm_args g;
  memset(&g, 0, sizeof(g));
  g.a = a; g.a_rows = ar; g.a_cols = ac; g.lda = lda; g.a_mn_major = amn;
  g.b = b; g.b_rows = br; g.b_cols = bc; g.ldb = ldb; g.b_mn_major = bmn;
  g.d = d; g.ldd = ldd; g.M = M; g.N = N; g.K = K;
  g.bias = bias; g.relu = relu; g.mask = mask; g.ld_mask = ldm;
  g.block_n = block_n; g.splits = splits; g.mode = mode; g.img_w = 26; g.cblocks = cblocks; g.shift_sign = shift_sign;
  return pdm_gemm_tf32(&g, s);
}

int splits_for(int tiles, int kblocks) {
  int want = (2 * 148 + tiles - 1) / tiles;
  if (want < 1) want = 1;
  if (want > kblocks) want = kblocks;
  return want;
}

}  // namespace

extern "C" {

int pdm_cnn_destroy(pdm_cnn_ctx* c);

int pdm_cnn_create(pdm_cnn_ctx** out, int device, int width, int max_batch) {
  PDM_REQUIRE(out && width >= 1 && width <= 8 && max_batch >= 1, PDM_ERR_ARG, "pdm_cnn_create: bad arguments (width 1..8)");
  PDM_CHECK_CUDA(cudaSetDevice(device));
  pdm_cnn_ctx* c = new (std::nothrow) pdm_cnn_ctx();
  PDM_REQUIRE(c != nullptr, PDM_ERR_STATE, "pdm_cnn_create: out of host memory");
  c->device = device; c->width = width; c->max_batch = max_batch;
  c->c1 = 32 * width; c->c2 = 64 * width; c->H = 128 * width;
  const size_t B = (size_t)max_batch;
#define CNN_ALLOC(ptr, count)                                                                         \
  do {                                                                                                \
    cudaError_t _e = cudaMalloc((void**)&(ptr), (size_t)(count));                                     \
    if (_e != cudaSuccess) {                                                                          \
      pdm_set_error("pdm_cnn_create: cudaMalloc(%zu) -> %s", (size_t)(count), cudaGetErrorString(_e)); \
      pdm_cnn_destroy(c);                                                                             \
      return PDM_ERR_CUDA;                                                                            \
    }                                                                                                 \
  } while (0)
  CNN_ALLOC(c->a1, B * kPix * c->c1 * 4);
  CNN_ALLOC(c->z2, B * kPix * c->c2 * 4);
  CNN_ALLOC(c->pooled, B * c->c2 * 144 * 4);
  CNN_ALLOC(c->dpool, B * c->c2 * 144 * 4);
  CNN_ALLOC(c->code, B * c->c2 * 144);
  CNN_ALLOC(c->h1, B * c->H * 4);
  CNN_ALLOC(c->dz1, B * c->H * 4);
  CNN_ALLOC(c->logits, B * 10 * 4);
  CNN_ALLOC(c->dlogits, B * 10 * 4);
  CNN_ALLOC(c->wk, (size_t)c->c2 * 9 * c->c1 * 4);
  CNN_ALLOC(c->wd, (size_t)c->c1 * 9 * c->c2 * 4);
  CNN_ALLOC(c->gw, (size_t)c->c2 * 9 * c->c1 * 4);
  CNN_ALLOC(c->loss, 256);
#undef CNN_ALLOC
  PDM_CHECK_CUDA(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) PDM_CHECK_CUDA(cudaEventCreateWithFlags(&c->ev_bucket[i], cudaEventDisableTiming));
  PDM_CHECK_CUDA(cudaEventCreateWithFlags(&c->ev_comm_done, cudaEventDisableTiming));
  *out = c;
  return PDM_OK;
}

// Forward pass into ctx->logits.  training != 0 applies dropout with the given seed.
static int cnn_forward(pdm_cnn_ctx* c, const void* x, int x_dtype, int B, const pdm_cnn_params* p, int training, float p1,
                       float p2, uint64_t seed, int need_bwd, cudaStream_t s) {
  int rc = pdm_conv1_fwd(x, x_dtype, p->conv1_w, p->conv1_b, c->a1, B, c->c1, s);
  if (rc) return rc;
  rc = pdm_conv_relayout(p->conv2_w, c->wk, need_bwd ? c->wd : nullptr, c->c2, c->c1, 0, s);
  if (rc) return rc;
  const int M = B * kPix;
  rc = gemm(c->a1, M, c->c1, c->c1, 0, c->wk, c->c2, 9 * c->c1, 9 * c->c1, 0, c->z2, c->c2, M, c->c2, 9 * c->c1, nullptr, 0,
            nullptr, 0, 0, 1, 1, c->c1 / 32, 1, s);
  if (rc) return rc;
  rc = pdm_pool_fwd(c->z2, p->conv2_b, c->pooled, c->code, B, c->c2, training ? p1 : 0.f, seed, s);
  if (rc) return rc;
  const int K1 = c->c2 * 144;
  const int bn = c->H >= 128 ? 128 : (c->H + 31) / 32 * 32;
  const int tiles = ((B + 127) / 128) * ((c->H + bn - 1) / bn);
  const int splits = splits_for(tiles, (K1 + 31) / 32);
  if (splits > 1) {
    PDM_CHECK_CUDA(cudaMemsetAsync(c->h1, 0, (size_t)B * c->H * 4, s));
    rc = gemm(c->pooled, B, K1, K1, 0, p->fc1_w, c->H, K1, K1, 0, c->h1, c->H, B, c->H, K1, nullptr, 0, nullptr, 0, bn, splits, 0,
              1, 1, s);
    if (rc) return rc;
    rc = pdm_bias_relu(c->h1, p->fc1_b, B, c->H, s);
  } else {
    rc = gemm(c->pooled, B, K1, K1, 0, p->fc1_w, c->H, K1, K1, 0, c->h1, c->H, B, c->H, K1, p->fc1_b, 1, nullptr, 0, bn, 1, 0, 1,
              1, s);
  }
  if (rc) return rc;
  return pdm_fcout_fwd(c->h1, p->fc2_w, p->fc2_b, c->logits, B, c->H, training ? p2 : 0.f, seed ^ 0x9E3779B97F4A7C15ull, s);
}

// Evaluation: logits (optional copy-out) + loss/top-1 accumulated into metrics.
int pdm_cnn_eval_step(pdm_cnn_ctx* c, const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                      const pdm_cnn_params* params, float* logits_out, double* metrics, pdm_stream_t stream) {
  PDM_REQUIRE(c && x && params, PDM_ERR_ARG, "pdm_cnn_eval_step: NULL argument");
  PDM_REQUIRE(batch >= 1 && batch <= c->max_batch, PDM_ERR_LIMIT, "pdm_cnn_eval_step: batch %d outside [1, %d]", batch, c->max_batch);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = cnn_forward(c, x, x_dtype, batch, params, 0, 0.f, 0.f, 0, 0, s);
  if (rc) return rc;
  if (y != nullptr) {
    rc = pdm_cross_entropy(c->logits, y, y_dtype, batch, 10, c->loss, nullptr, metrics, s);
    if (rc) return rc;
  }
  if (logits_out) PDM_CHECK_CUDA(cudaMemcpyAsync(logits_out, c->logits, (size_t)batch * 10 * 4, cudaMemcpyDeviceToDevice, s));
  return PDM_OK;
}

// One training step.  grads receives the (un-reduced) gradients, overwritten.  When `plan` is
// given the reducer runs too: bucket b covers flat floats [lo[b], hi[b]) of params/grads/m/v
// (offsets from the bases in `plan`); each is reduced over `plan->comm` (or just updated when
// comm is NULL) on the executor's side stream as soon as it is final, and `stream` waits for the
// last one before the call's work is considered complete.
int pdm_cnn_train_step(pdm_cnn_ctx* c, const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                       const pdm_cnn_params* p, const pdm_cnn_params* g, float p1, float p2, uint64_t seed,
                       double* metrics, float* loss_out, const pdm_cnn_plan* plan, pdm_stream_t stream) {
  PDM_REQUIRE(c && x && y && p && g, PDM_ERR_ARG, "pdm_cnn_train_step: NULL argument");
  PDM_REQUIRE(batch >= 1 && batch <= c->max_batch, PDM_ERR_LIMIT, "pdm_cnn_train_step: batch %d outside [1, %d]", batch, c->max_batch);
  PDM_REQUIRE(p1 >= 0.f && p1 < 1.f && p2 >= 0.f && p2 < 1.f, PDM_ERR_ARG, "pdm_cnn_train_step: dropout probabilities must be in [0, 1)");
  cudaStream_t s = (cudaStream_t)stream;
  const int B = batch, c1 = c->c1, c2 = c->c2, H = c->H, M = B * kPix, K1 = c2 * 144;
  const uint64_t seed2 = seed ^ 0x9E3779B97F4A7C15ull;
  int rc = cnn_forward(c, x, x_dtype, B, p, 1, p1, p2, seed, 1, s);
  if (rc) return rc;
  rc = pdm_cross_entropy(c->logits, y, y_dtype, B, 10, loss_out ? loss_out : c->loss, c->dlogits, metrics, s);
  if (rc) return rc;
  // small gradients are accumulated atomically by their kernels: clear them first
  PDM_CHECK_CUDA(cudaMemsetAsync(g->fc2_w, 0, (size_t)10 * H * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->fc2_b, 0, 10 * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->fc1_b, 0, (size_t)H * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->conv2_b, 0, (size_t)c2 * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->conv2_w, 0, (size_t)c2 * c1 * 9 * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->conv1_w, 0, (size_t)c1 * 9 * 4, s));
  PDM_CHECK_CUDA(cudaMemsetAsync(g->conv1_b, 0, (size_t)c1 * 4, s));
  // ---- fc2 backward: dz1 (dropout + ReLU masks folded in), dW2, db2, db1 ----
  rc = pdm_fcout_bwd(c->dlogits, c->h1, p->fc2_w, c->dz1, g->fc2_w, g->fc2_b, g->fc1_b, B, H, p2, seed2, s);
  if (rc) return rc;
  // ---- fc1 wgrad: dW1[H, K1] = dz1^T . pooled (both MN-major, straight into the gradient buffer) ----
  rc = gemm(c->dz1, B, H, H, 1, c->pooled, B, K1, K1, 1, g->fc1_w, K1, H, K1, B, nullptr, 0, nullptr, 0, 64, 1, 0, 1, 1, s);
  if (rc) return rc;
  int next_bucket = 0;
  auto launch_bucket = [&](int b) -> int {
    if (plan == nullptr || b >= plan->n_buckets) return PDM_OK;
    PDM_CHECK_CUDA(cudaEventRecord(c->ev_bucket[b & 1], s));
    PDM_CHECK_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_bucket[b & 1], 0));
    const size_t lo = (size_t)plan->lo[b], n = (size_t)(plan->hi[b] - plan->lo[b]);
    if (plan->nvls != nullptr) {
      return pdm_nvls_allreduce_optim(plan->nvls, plan->grads_off + lo * 4, plan->params_off + lo * 4, plan->m_off + lo * 4,
                                      plan->v_off + lo * 4, n, &plan->hyper, plan->grad_scale, c->comm_stream);
    }
    if (plan->comm != nullptr) {
      return pdm_comm_allreduce_optim(plan->comm, plan->grads_off + lo * 4, plan->params_off + lo * 4, plan->m_off + lo * 4,
                                      plan->v_off + lo * 4, n, &plan->hyper, plan->grad_scale, 0, plan->max_ctas, c->comm_stream);
    }
    return pdm_optim_step(plan->params + lo, plan->grads + lo, plan->m + lo, plan->v + lo, n, &plan->hyper, 1.0f, c->comm_stream);
  };
  // ---- fc1 dgrad: dPooled[B, K1] = dz1 . W1 (W1 stored [H][K1] is the MN-major B operand) ----
  rc = gemm(c->dz1, B, H, H, 0, p->fc1_w, H, K1, K1, 1, c->dpool, K1, B, K1, H, nullptr, 0, nullptr, 0, K1 >= 256 ? 256 : 64, 1, 0, 1,
            1, s);
  if (rc) return rc;
  // bucket 0 = {fc2.b, fc2.w, fc1.b, fc1.w}: its gradients are final and nothing below reads those
  // parameters again (the dgrad above was the last reader of fc1.weight), so reduce + update it
  // now, under the convolution backward
  if (plan != nullptr && plan->n_buckets >= 2) {
    rc = launch_bucket(next_bucket++);
    if (rc) return rc;
  }
  // ---- pool backward: dY2 over the raw conv2 outputs (reuses z2's buffer), db2 ----
  rc = pdm_pool_bwd(c->dpool, c->code, c->z2, g->conv2_b, B, c2, p1, seed, s);
  if (rc) return rc;
  // ---- conv2 wgrad (shifted implicit GEMM, split-K) -> [c2][tap][c1] -> parameter layout ----
  {
    const int tiles = ((c2 + 127) / 128) * ((9 * c1 + 255) / 256);
    const int splits = splits_for(tiles, (M + 31) / 32);
    PDM_CHECK_CUDA(cudaMemsetAsync(c->gw, 0, (size_t)c2 * 9 * c1 * 4, s));
    rc = gemm(c->z2, M, c2, c2, 1, c->a1, M, c1, c1, 1, c->gw, 9 * c1, c2, 9 * c1, M, nullptr, 0, nullptr, 0, 0, splits, 2, c1 / 32, 1, s);
    if (rc) return rc;
    rc = pdm_conv_relayout(c->gw, g->conv2_w, nullptr, c2, c1, 1, s);
    if (rc) return rc;
  }
  // ---- conv2 dgrad, in place over a1 (the epilogue reads a1 as the ReLU mask element by element) ----
  rc = gemm(c->z2, M, c2, c2, 0, c->wd, c1, 9 * c2, 9 * c2, 0, c->a1, c1, M, c1, 9 * c2, nullptr, 0, c->a1, c1, 0, 1, 1, c2 / 32, -1, s);
  if (rc) return rc;
  // ---- conv1 backward ----
  rc = pdm_conv1_bwd(x, x_dtype, c->a1, g->conv1_w, g->conv1_b, B, c1, s);
  if (rc) return rc;
  if (plan != nullptr) {
    while (next_bucket < plan->n_buckets) {
      rc = launch_bucket(next_bucket++);
      if (rc) return rc;
    }
    PDM_CHECK_CUDA(cudaEventRecord(c->ev_comm_done, c->comm_stream));
    PDM_CHECK_CUDA(cudaStreamWaitEvent(s, c->ev_comm_done, 0));
  }
  return PDM_OK;
}

int pdm_cnn_destroy(pdm_cnn_ctx* c) {
  if (!c) return PDM_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  float* bufs[] = {c->a1, c->z2, c->pooled, c->dpool, c->h1, c->dz1, c->logits, c->dlogits, c->wk, c->wd, c->gw, c->loss};
  for (float* b : bufs) if (b) cudaFree(b);
  if (c->code) cudaFree(c->code);
  if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
  for (int i = 0; i < 2; ++i) if (c->ev_bucket[i]) cudaEventDestroy(c->ev_bucket[i]);
  if (c->ev_comm_done) cudaEventDestroy(c->ev_comm_done);
  delete c;
  cudaGetLastError();
  return PDM_OK;
}

}  // extern "C"
