// reducer.cu -- the gradient reducer of the data-parallel wrapper for ARBITRARY models
// (the role of c10d's Reducer, SURVEY N4: per-parameter ready marks, bucket-ready counting,
// strictly in-order bucket launches on a communication stream, the stream hand-over back to the
// compute stream after the last bucket).
//
// Gradients already live in the flat bucket memory (parallel/flat.py), so there is no
// copy-in / copy-out: a ready bucket is ONE launch of the fused all-reduce + 1/W + optimizer
// kernel over its range -- the peer-memory two-shot kernel (comm.cu) or the in-switch multimem
// kernel (nvls.cu).  Python only forwards autograd's post-accumulate hook (parallel/ddp.py).
//
// Without a native communicator (gloo / CPU, or GPUs that cannot map each other) the reducer
// still does the bookkeeping and tells the caller which buckets became launchable; the caller
// reduces them with torch.distributed.
#include "common.cuh"

#include <vector>

struct pdm_reducer {
  int n_params = 0, n_buckets = 0, world = 1, device = -1;
  std::vector<int32_t> bucket_of;       // parameter -> bucket
  std::vector<int64_t> lo, hi;          // bucket -> float range in the flat buffers
  std::vector<int32_t> size;            // bucket -> number of parameters
  std::vector<int32_t> pending;         // parameters still missing in this pass
  std::vector<uint8_t> marked;          // parameter already seen in this pass
  std::vector<int32_t> order;           // launch log of this pass
  std::vector<int32_t> ready;           // parameters in the order they became ready in this pass
  bool held_back = false;               // a bucket completed while an earlier one was still open
  int next = 0;                         // next bucket to launch
  bool started = false;                 // a parameter was marked since the last arm
  size_t grads_off = 0, params_off = 0, m_off = 0, v_off = 0;
  pdm_comm* comm = nullptr;
  pdm_nvls* nvls = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  int64_t launches = 0;
};

extern "C" int pdm_reducer_create(pdm_reducer** out, const pdm_reducer_config* cfg) {
  PDM_REQUIRE(out && cfg && cfg->n_params >= 1 && cfg->n_buckets >= 1 && cfg->bucket_of && cfg->bucket_lo &&
              cfg->bucket_hi && cfg->world >= 1, PDM_ERR_ARG, "pdm_reducer_create: bad arguments");
  pdm_reducer* r = new pdm_reducer();
  r->n_params = cfg->n_params;
  r->n_buckets = cfg->n_buckets;
  r->world = cfg->world;
  r->device = cfg->device;
  r->bucket_of.assign(cfg->bucket_of, cfg->bucket_of + cfg->n_params);
  r->lo.assign(cfg->bucket_lo, cfg->bucket_lo + cfg->n_buckets);
  r->hi.assign(cfg->bucket_hi, cfg->bucket_hi + cfg->n_buckets);
  r->size.assign(cfg->n_buckets, 0);
  for (int i = 0; i < r->n_params; ++i) {
    int b = r->bucket_of[i];
    if (b < 0 || b >= r->n_buckets) {
      delete r;
      PDM_REQUIRE(false, PDM_ERR_ARG, "pdm_reducer_create: parameter %d maps to bucket %d of %d", i, b, cfg->n_buckets);
    }
    r->size[b] += 1;
  }
  for (int b = 0; b < r->n_buckets; ++b) {
    if (r->size[b] == 0 || r->lo[b] < 0 || r->hi[b] <= r->lo[b] || (r->lo[b] & 3) || (r->hi[b] & 3)) {
      int sz = r->size[b];
      long long l = (long long)r->lo[b], h = (long long)r->hi[b];
      delete r;
      PDM_REQUIRE(false, PDM_ERR_ARG, "pdm_reducer_create: bucket %d has %d parameters, range [%lld, %lld) "
                  "(ranges are non-empty and 4-float aligned)", b, sz, l, h);
    }
  }
  r->pending = r->size;
  r->marked.assign(r->n_params, 0);
  r->grads_off = cfg->grads_off;
  r->params_off = cfg->params_off;
  r->m_off = cfg->m_off;
  r->v_off = cfg->v_off;
  r->comm = cfg->comm;
  r->nvls = cfg->nvls;
  if ((r->comm || r->nvls) && r->device >= 0) {
    cudaError_t e = cudaSetDevice(r->device);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_ready, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming);
    if (e != cudaSuccess) {
      delete r;
      PDM_REQUIRE(false, PDM_ERR_CUDA, "pdm_reducer_create: %s", cudaGetErrorString(e));
    }
  }
  *out = r;
  return PDM_OK;
}

// Start of a training forward.  Returns PDM_ERR_STATE (and leaves the state untouched for
// pdm_reducer_state) when the previous backward marked some but not all parameters: a bucket --
// and every later one -- was never reduced and the replicas have drifted apart.
extern "C" int pdm_reducer_arm(pdm_reducer* r) {
  PDM_REQUIRE(r, PDM_ERR_ARG, "pdm_reducer_arm: null reducer");
  PDM_REQUIRE(!(r->started && r->next < r->n_buckets), PDM_ERR_STATE,
              "pdm_reducer_arm: the previous backward pass left buckets %d..%d unreduced", r->next, r->n_buckets - 1);
  r->pending = r->size;
  std::fill(r->marked.begin(), r->marked.end(), 0);
  r->order.clear();
  r->ready.clear();
  r->held_back = false;
  r->next = 0;
  r->started = false;
  return PDM_OK;
}

static int launch_bucket(pdm_reducer* r, int b, const pdm_hyper* hyper, pdm_stream_t compute, pdm_stream_t comm_stream) {
  r->launches += 1;
  r->order.push_back(b);
  if (!r->comm && !r->nvls) return PDM_OK;                   // the caller reduces it
  cudaStream_t cs = (cudaStream_t)comm_stream, ms = (cudaStream_t)compute;
  // the bucket's last gradient was produced on the compute stream
  PDM_CHECK_CUDA(cudaEventRecord(r->ev_ready, ms));
  PDM_CHECK_CUDA(cudaStreamWaitEvent(cs, r->ev_ready, 0));
  const size_t lo = (size_t)r->lo[b] * 4, n = (size_t)(r->hi[b] - r->lo[b]);
  pdm_hyper plain = {};
  plain.opt = PDM_OPT_NONE;
  plain.step = 1;
  const bool fused = hyper && hyper->opt != PDM_OPT_NONE;
  const pdm_hyper* h = fused ? hyper : &plain;
  // fused: new parameters land at params; plain: averaged gradient in place
  const size_t out_off = fused ? r->params_off + lo : r->grads_off + lo;
  const float scale = 1.0f / (float)r->world;
  int rc;
  if (r->nvls)
    rc = pdm_nvls_allreduce_optim(r->nvls, r->grads_off + lo, out_off, r->m_off + lo, r->v_off + lo, n, h, scale, comm_stream);
  else if (r->world > 1 || fused)
    rc = pdm_comm_allreduce_optim(r->comm, r->grads_off + lo, out_off, r->m_off + lo, r->v_off + lo, n, h, scale,
                                  fused ? 0 : 2, 0, comm_stream);
  else
    rc = PDM_OK;
  if (rc != PDM_OK) return rc;
  if (b == r->n_buckets - 1) {
    // the caller's stream continues only after the last bucket is applied
    PDM_CHECK_CUDA(cudaEventRecord(r->ev_done, cs));
    PDM_CHECK_CUDA(cudaStreamWaitEvent(ms, r->ev_done, 0));
  }
  return PDM_OK;
}

// Autograd finished accumulating the gradient of `param`.  Launches every bucket that became
// launchable (buckets go out strictly in index order on every rank, so all ranks issue the same
// collective sequence); *first / *count report which.
extern "C" int pdm_reducer_mark_ready(pdm_reducer* r, int32_t param, const pdm_hyper* hyper, pdm_stream_t compute,
                                      pdm_stream_t comm_stream, int32_t* first, int32_t* count) {
  PDM_REQUIRE(r && param >= 0 && param < r->n_params, PDM_ERR_ARG, "pdm_reducer_mark_ready: bad parameter index %d", (int)param);
  PDM_REQUIRE(!r->marked[param], PDM_ERR_STATE, "pdm_reducer_mark_ready: parameter %d was marked twice in one backward pass "
              "(two backward passes after one forward? wrap the extra ones in no_sync())", (int)param);
  r->marked[param] = 1;
  r->started = true;
  r->ready.push_back(param);
  const int b = r->bucket_of[param];
  r->pending[b] -= 1;
  if (r->pending[b] == 0 && b > r->next) r->held_back = true;
  int f = r->next, c = 0;
  while (r->next < r->n_buckets && r->pending[r->next] == 0) {
    int rc = launch_bucket(r, r->next, hyper, compute, comm_stream);
    if (rc != PDM_OK) return rc;
    r->next += 1;
    c += 1;
  }
  if (first) *first = f;
  if (count) *count = c;
  return PDM_OK;
}

// Parameters that have not been marked in this pass (find_unused_parameters: the caller zeroes their
// gradient slices, then calls pdm_reducer_finalize).  out may be NULL to get the count only.
extern "C" int pdm_reducer_unmarked(pdm_reducer* r, int32_t* out, int32_t* count) {
  PDM_REQUIRE(r && count, PDM_ERR_ARG, "pdm_reducer_unmarked: bad arguments");
  int32_t n = 0;
  for (int i = 0; i < r->n_params; ++i) {
    if (r->marked[i]) continue;
    if (out) out[n] = i;
    n += 1;
  }
  *count = n;
  return PDM_OK;
}

// End of a backward pass in find_unused_parameters mode: every parameter that received no gradient counts
// as ready (its slice holds zeros), so the remaining buckets go out -- in index order, on every rank, whatever
// subset of the model each rank used.  A pass that marked nothing launches nothing.
extern "C" int pdm_reducer_finalize(pdm_reducer* r, const pdm_hyper* hyper, pdm_stream_t compute,
                                    pdm_stream_t comm_stream, int32_t* first, int32_t* count) {
  PDM_REQUIRE(r, PDM_ERR_ARG, "pdm_reducer_finalize: null reducer");
  int f = r->next, c = 0;
  if (r->started) {
    for (int i = 0; i < r->n_params; ++i) {
      if (r->marked[i]) continue;
      r->marked[i] = 1;
      r->ready.push_back(i);
      r->pending[r->bucket_of[i]] -= 1;
    }
    while (r->next < r->n_buckets && r->pending[r->next] == 0) {
      int rc = launch_bucket(r, r->next, hyper, compute, comm_stream);
      if (rc != PDM_OK) return rc;
      r->next += 1;
      c += 1;
    }
  }
  if (first) *first = f;
  if (count) *count = c;
  return PDM_OK;
}

// The order in which the parameters became ready in this pass (out: n_params entries, *n = how many so
// far) and whether a bucket had to wait for an earlier one: the input of the bucket rebuild after the
// first iteration (c10d Reducer::rebuild_buckets).
extern "C" int pdm_reducer_ready_order(pdm_reducer* r, int32_t* out, int32_t* n, int32_t* held_back) {
  PDM_REQUIRE(r && n, PDM_ERR_ARG, "pdm_reducer_ready_order: bad arguments");
  if (out) std::copy(r->ready.begin(), r->ready.end(), out);
  *n = (int32_t)r->ready.size();
  if (held_back) *held_back = r->held_back ? 1 : 0;
  return PDM_OK;
}

// next bucket to launch, total launches so far; pending[n_buckets] and order[<= n_buckets] are optional.
extern "C" int pdm_reducer_state(pdm_reducer* r, int32_t* next, int64_t* launches, int32_t* pending, int32_t* order,
                                 int32_t* n_order) {
  PDM_REQUIRE(r, PDM_ERR_ARG, "pdm_reducer_state: null reducer");
  if (next) *next = r->next;
  if (launches) *launches = r->launches;
  if (pending) std::copy(r->pending.begin(), r->pending.end(), pending);
  if (order) std::copy(r->order.begin(), r->order.end(), order);
  if (n_order) *n_order = (int32_t)r->order.size();
  return PDM_OK;
}

extern "C" int pdm_reducer_destroy(pdm_reducer* r) {
  if (!r) return PDM_OK;
  if (r->ev_ready) cudaEventDestroy(r->ev_ready);
  if (r->ev_done) cudaEventDestroy(r->ev_done);
  delete r;
  return PDM_OK;
}
