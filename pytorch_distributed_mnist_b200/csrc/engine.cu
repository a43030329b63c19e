// engine.cu -- native epoch executor for the fused Net path.
//
// Replaces, for one epoch of Trainer.train()/evaluate() (ref:77-116):
//   * the DataLoader worker processes + shared-memory hand-off:
//     C++ worker threads gather the sampled rows of the in-memory uint8
//     dataset into a ring of pinned staging buffers;
//   * the two pageable cudaMemcpy per step (ref:84-85, K1/K2): ONE async copy
//     of a packed [B*784 image bytes | B label bytes] slot -- 785 B/sample
//     instead of the reference's 3,144 B/sample (fp32 + int64);
//   * the ~25 launches of the step: one fused kernel (net_step.cu);
//   * the two host syncs per step (ref:94-95, K12/K13): the per-step
//     {loss, correct} pair is copied back asynchronously and only read after
//     the epoch's last step.
// Two streams: a copy stream that runs ahead issuing the H2D copies (throttled only by
// "slot consumed" events, `slots` steps back) and a compute stream that runs the step
// kernels back to back, each waiting on its slot's "copy done" event.  The kernel writes
// its {loss, correct} pair straight into mapped pinned host memory (an 8-byte device->host
// store per step), so there is no D2H memcpy in the loop.  Six CUDA calls per step.
#include "common.cuh"

#include <emmintrin.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <stdlib.h>
#include <new>
#include <thread>
#include <vector>

extern "C" int pdm_comm_peers(struct pdm_comm* c, pdm_peers* out, int advance);
extern "C" int pdm_comm_peers_block(struct pdm_comm* c, pdm_peers* out, int64_t nsteps);

namespace {

constexpr int kImg = PDM_NET_IN;
constexpr int kSample = PDM_NET_IN + 1;  // image bytes + one label byte

struct Job {              // one run() call, shared with the workers
  const int64_t* indices = nullptr;
  int64_t count = 0;
  int batch = 0;
  int64_t nsteps = 0;
  std::atomic<int64_t> next{0};   // next batch index to stage (per job: a late worker of an
                                  // earlier job can never claim a batch of this one)
};

bool debug_on() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PDM_ENGINE_DEBUG"); v = (e && *e && *e != '0') ? 1 : 0; }
  return v == 1;
}
// PDM_ENGINE_PROFILE=1: host time spent per section of the step loop, printed after each run.
bool profile_on() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("PDM_ENGINE_PROFILE"); v = (e && *e && *e != '0') ? 1 : 0; }
  return v == 1;
}
struct HostProf {
  double t[6] = {0, 0, 0, 0, 0, 0};   // wait-filled, h2d+events, launch, record+d2h, sync-old, total
  double first_wait = 0;              // wait-filled of step 0: worker wake-up + the first gather
  std::chrono::steady_clock::time_point last;
  void start() { last = std::chrono::steady_clock::now(); }
  void lap(int i) {
    const auto now = std::chrono::steady_clock::now();
    t[i] += std::chrono::duration<double, std::micro>(now - last).count();
    last = now;
  }
};
#define DBG(...) do { if (debug_on()) { fprintf(stderr, "[pdm_engine] " __VA_ARGS__); fputc('\n', stderr); fflush(stderr); } } while (0)

}  // namespace

struct pdm_engine {
  int device = 0;
  int max_batch = 0;
  int slots = 0;       // ring depth (even)
  int nworkers = 0;
  // model / optimizer state (not owned)
  float* params = nullptr;
  float* m = nullptr;
  float* v = nullptr;
  pdm_comm* comm = nullptr;
  int world = 1;                          // ranks of `comm`
  // owned device memory
  float* scratch = nullptr;
  uint32_t* sync = nullptr;
  double* metrics = nullptr;
  float* step_out = nullptr;              // [slots][2]
  uint32_t done_total = 0;                // CTAs launched so far on `sync` (step-to-step flag dependency)
  int64_t last_launches = 0;              // kernels launched by the last run_* call
  std::vector<uint8_t*> dev_slot;         // [slots] packed batch
  std::vector<uint8_t*> host_slot;        // [slots] pinned staging
  float* host_results = nullptr;          // pinned + mapped [capacity][2]
  float* host_results_dev = nullptr;      // device alias of host_results
  int64_t host_results_cap = 0;
  cudaStream_t stream[3] = {nullptr, nullptr, nullptr};   // [0] compute, [1], [2] copy (alternating)
  std::vector<cudaEvent_t> h2d_done;      // per slot, recorded on the copy stream
  std::vector<cudaEvent_t> consumed;      // per slot, recorded on the compute stream after its kernel
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  // dataset (host, not owned)
  const uint8_t* images = nullptr;
  const uint8_t* labels = nullptr;
  int64_t n_samples = 0;
  // datasets as the kernels see them on the in-kernel-gather path (pdm_engine_place_dataset): a small
  // cache, so that switching between loaders (train set / another train set) re-uses the page-locking
  // or the device copy instead of redoing it inside a training pass
  struct Placement {
    int mode = -1;                        // 0 device copy (owned), 1 host arrays registered + mapped
    const uint8_t* host_images = nullptr; // key
    const uint8_t* host_labels = nullptr;
    int64_t n = 0;
    uint8_t* images = nullptr;            // device pointer (copy or alias)
    uint8_t* labels = nullptr;
    uint64_t last_use = 0;
  };
  static constexpr int kMaxPlacements = 4;
  Placement placements[kMaxPlacements];
  int placed = -1;                        // index of the current placement, -1 = none
  uint64_t place_clock = 0;
  int32_t* idx_dev = nullptr;             // index plan of the current pass
  int64_t idx_cap = 0;
  uint8_t* tc_ring = nullptr;             // feeder ring of the persistent tensor-core kernel (zero-copy placements)
  int32_t* idx_host = nullptr;            // pinned staging of the plan (so its upload is truly asynchronous)
  int64_t idx_host_cap = 0;
  // worker pool
  std::vector<std::thread> workers;
  std::mutex mu;
  std::condition_variable cv_work, cv_filled, cv_free;
  bool quit = false;
  uint64_t job_id = 0;
  std::shared_ptr<Job> job;
  std::vector<int64_t> filled_batch;      // per slot: batch index staged there (-1 = none)
  std::vector<int64_t> free_for;          // per slot: highest batch index allowed to be staged there
  int64_t workers_active = 0;
};

namespace {

// Copy into a pinned staging slot with non-temporal stores: the slot is read next by the copy
// engine, never by this core, so write-allocating it into the cache only adds a read-for-
// ownership of every line (a third of the host memory traffic of the gather).  dst and n are
// multiples of 16 bytes by construction (784 = 49 * 16, slots are 256-byte aligned).
inline void stream_copy(uint8_t* dst, const uint8_t* src, size_t n) {
  if ((reinterpret_cast<uintptr_t>(dst) & 15) != 0 || (n & 15) != 0) {
    memcpy(dst, src, n);
    return;
  }
  const __m128i* s = reinterpret_cast<const __m128i*>(src);
  __m128i* d = reinterpret_cast<__m128i*>(dst);
  for (size_t k = 0; k < n / 16; ++k) _mm_stream_si128(d + k, _mm_loadu_si128(s + k));
}

void gather_batch(pdm_engine* e, const Job& job, int64_t b, uint8_t* dst) {
  DBG("worker stages batch %lld", (long long)b);
  const int64_t lo = b * job.batch;
  const int64_t n = std::min<int64_t>(job.batch, job.count - lo);
  uint8_t* img = dst;
  uint8_t* lab = dst + (size_t)n * kImg;
  const int64_t* idx = job.indices + lo;
  constexpr int kAhead = 8;   // rows: a shuffled gather is DRAM-latency bound without prefetch
  int64_t i = 0;
  while (i < n) {  // coalesce runs of consecutive indices into one memcpy
    int64_t j = i + 1;
    while (j < n && idx[j] == idx[j - 1] + 1) ++j;
    if (j == i + 1 && i + kAhead < n) {
      const uint8_t* nxt = e->images + (size_t)idx[i + kAhead] * kImg;
      for (int off = 0; off < kImg; off += 64) __builtin_prefetch(nxt + off, 0, 0);
      __builtin_prefetch(e->labels + idx[i + kAhead], 0, 0);
    }
    stream_copy(img + (size_t)i * kImg, e->images + (size_t)idx[i] * kImg, (size_t)(j - i) * kImg);
    memcpy(lab + i, e->labels + idx[i], (size_t)(j - i));
    i = j;
  }
  _mm_sfence();   // streaming stores must be globally visible before the slot is handed to the DMA
}

void worker_main(pdm_engine* e) {
  uint64_t seen = 0;
  for (;;) {
    std::shared_ptr<Job> jp;
    {
      std::unique_lock<std::mutex> lk(e->mu);
      e->cv_work.wait(lk, [&] { return e->quit || e->job_id != seen; });
      if (e->quit) return;
      seen = e->job_id;
      jp = e->job;
      e->workers_active += 1;
    }
    Job& job = *jp;
    for (;;) {
      const int64_t b = job.next.fetch_add(1);
      if (b >= job.nsteps) break;
      const int s = (int)(b % e->slots);
      {
        std::unique_lock<std::mutex> lk(e->mu);
        e->cv_free.wait(lk, [&] { return e->quit || b <= e->free_for[s]; });
        if (e->quit) break;
      }
      gather_batch(e, job, b, e->host_slot[s]);
      {
        std::lock_guard<std::mutex> lk(e->mu);
        e->filled_batch[s] = b;
      }
      e->cv_filled.notify_all();
    }
    {
      std::lock_guard<std::mutex> lk(e->mu);
      e->workers_active -= 1;
    }
    e->cv_filled.notify_all();
  }
}

// The persistent tensor-core kernel (net_tc.cu) takes a whole pass in one launch: used for raw uint8
// inputs whenever the device can launch it and the caller did not ask for a specific CTA count.
// It runs on one 13-SM cluster, so beyond ~1024 rows per step the 148-SM per-step kernel is faster
// (measured: 7.7 vs 9.7 us at batch 512, 57 vs 27 us at batch 4096; benchmarks/tc_sweep.py).  With peers its
// in-kernel exchange wins at every world size of one node (measured us/step at the 20-step pass of bench.py,
// persistent vs per-step kernel: 2 GPUs 10.8 vs 13.8, 4 GPUs 12.5 vs 14.0, 8 GPUs 14.8 vs 15.4;
// profiles/raw/r2/bench_*gpu_policy_k20.log).  PDM_NET_TC_MAX_WORLD (default 8) caps the world size that
// takes it for resident inputs; passes that read pinned host memory always take it, for its feeder cluster
// (8 GPUs: 21 vs 32 us/step).
int tc_max_world() {
  static const int v = [] { const char* e = getenv("PDM_NET_TC_MAX_WORLD"); return e && *e ? atoi(e) : 8; }();
  return v;
}
bool use_tc(const pdm_engine* e, int x_dtype, int y_dtype, int ctas, int batch, bool from_host) {
  if (!(x_dtype == PDM_U8 && y_dtype == PDM_U8 && ctas == 0 && batch <= 1024 && pdm_net_tc_supported(e->device) == 1)) return false;
  if (from_host || e->comm == nullptr) return true;
  return e->world <= tc_max_world();
}

int ensure_results(pdm_engine* e, int64_t nsteps) {
  if (nsteps <= e->host_results_cap) return PDM_OK;
  if (e->host_results) cudaFreeHost(e->host_results);
  e->host_results = nullptr;
  e->host_results_cap = 0;
  const int64_t cap = std::max<int64_t>(nsteps, 1024);
  PDM_CHECK_CUDA(cudaHostAlloc((void**)&e->host_results, (size_t)cap * 2 * sizeof(float), cudaHostAllocMapped));
  PDM_CHECK_CUDA(cudaHostGetDevicePointer((void**)&e->host_results_dev, e->host_results, 0));
  e->host_results_cap = cap;
  return PDM_OK;
}

}  // namespace

extern "C" {

int pdm_engine_destroy(pdm_engine* e);

int pdm_engine_create(pdm_engine** out, int device, int max_batch, int slots, int workers) {
  PDM_REQUIRE(out != nullptr, PDM_ERR_ARG, "pdm_engine_create: out is NULL");
  PDM_REQUIRE(max_batch >= 1, PDM_ERR_ARG, "pdm_engine_create: max_batch must be >= 1");
  PDM_REQUIRE(max_batch <= (1 << 20), PDM_ERR_LIMIT, "pdm_engine_create: max_batch %d > 1048576", max_batch);
  if (slots < 2) slots = 2;
  if (slots & 1) slots += 1;   // step n and n+slots must land on the same stream
  if (slots > 64) slots = 64;
  if (workers < 1) workers = 1;
  if (workers > 64) workers = 64;
  PDM_CHECK_CUDA(cudaSetDevice(device));
  pdm_engine* e = new (std::nothrow) pdm_engine();
  PDM_REQUIRE(e != nullptr, PDM_ERR_STATE, "pdm_engine_create: out of host memory");
  e->device = device;
  e->max_batch = max_batch;
  e->slots = slots;
  e->nworkers = workers;
  const size_t slot_bytes = ((size_t)max_batch * kSample + 255) / 256 * 256;
#define ENGINE_TRY(expr)                                                                   \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      pdm_set_error("pdm_engine_create: %s -> %s", #expr, cudaGetErrorString(_e));         \
      pdm_engine_destroy(e);                                                               \
      return PDM_ERR_CUDA;                                                                 \
    }                                                                                      \
  } while (0)
  ENGINE_TRY(cudaMalloc((void**)&e->scratch, PDM_NET_SCRATCH_FLOATS(PDM_NET_MAX_CTAS) * sizeof(float)));
  ENGINE_TRY(cudaMalloc((void**)&e->sync, 256));
  ENGINE_TRY(cudaMemset(e->sync, 0, 256));
  ENGINE_TRY(cudaMalloc((void**)&e->metrics, 4 * sizeof(double)));
  ENGINE_TRY(cudaMemset(e->metrics, 0, 4 * sizeof(double)));
  ENGINE_TRY(cudaMalloc((void**)&e->step_out, (size_t)slots * 2 * sizeof(float)));
  e->dev_slot.assign(slots, nullptr);
  e->host_slot.assign(slots, nullptr);
  e->h2d_done.assign(slots, nullptr);
  e->consumed.assign(slots, nullptr);
  for (int s = 0; s < slots; ++s) {
    ENGINE_TRY(cudaMalloc((void**)&e->dev_slot[s], slot_bytes));
    ENGINE_TRY(cudaHostAlloc((void**)&e->host_slot[s], slot_bytes, cudaHostAllocDefault));
    ENGINE_TRY(cudaEventCreateWithFlags(&e->h2d_done[s], cudaEventDisableTiming));
    ENGINE_TRY(cudaEventCreateWithFlags(&e->consumed[s], cudaEventDisableTiming));
  }
  for (int i = 0; i < 3; ++i) {
    ENGINE_TRY(cudaStreamCreateWithFlags(&e->stream[i], cudaStreamNonBlocking));
  }
  ENGINE_TRY(cudaEventCreate(&e->ev_begin));
  ENGINE_TRY(cudaEventCreate(&e->ev_end));
#undef ENGINE_TRY
  e->filled_batch.assign(slots, -1);
  e->free_for.assign(slots, 0);
  for (int w = 0; w < workers; ++w) e->workers.emplace_back(worker_main, e);
  *out = e;
  return PDM_OK;
}

int pdm_engine_bind(pdm_engine* e, float* params, float* m, float* v, pdm_comm* comm) {
  PDM_REQUIRE(e && params, PDM_ERR_ARG, "pdm_engine_bind: engine and params are required");
  e->params = params;
  e->m = m;
  e->v = v;
  e->comm = comm;
  e->world = 1;
  if (comm != nullptr) {
    pdm_peers pr;
    if (pdm_comm_peers(comm, &pr, 0) == PDM_OK) e->world = pr.world > 1 ? pr.world : 1;
  }
  return PDM_OK;
}

int pdm_engine_set_dataset(pdm_engine* e, const uint8_t* images, const uint8_t* labels, int64_t n) {
  PDM_REQUIRE(e && images && labels && n >= 1, PDM_ERR_ARG, "pdm_engine_set_dataset: bad arguments");
  e->images = images;
  e->labels = labels;
  e->n_samples = n;
  return PDM_OK;
}

// Run ceil(count / batch) steps over the host dataset in the order given by `indices`.
//  train != 0: fused training steps (hyper->step is the index of the first update);
//  train == 0: evaluation steps (metrics only).
//  step_results: optional host [nsteps][2] = per-step {mean loss, correct} (training only).
//  device_ms: optional, device time of the whole run (events on the engine's streams).
int pdm_engine_run(pdm_engine* e, const int64_t* indices, int64_t count, int batch, int train,
                   const pdm_hyper* hyper, int ctas, float* step_results, float* device_ms) {
  PDM_REQUIRE(e && indices, PDM_ERR_ARG, "pdm_engine_run: engine and indices are required");
  PDM_REQUIRE(e->params != nullptr, PDM_ERR_STATE, "pdm_engine_run: pdm_engine_bind was not called");
  PDM_REQUIRE(e->images != nullptr, PDM_ERR_STATE, "pdm_engine_run: pdm_engine_set_dataset was not called");
  PDM_REQUIRE(batch >= 1 && batch <= e->max_batch, PDM_ERR_LIMIT, "pdm_engine_run: batch %d outside [1, %d]", batch,
              e->max_batch);
  PDM_REQUIRE(!train || hyper, PDM_ERR_ARG, "pdm_engine_run: training needs hyper-parameters");
  if (count <= 0) return PDM_OK;
  {
    int64_t lo = indices[0], hi = indices[0];
    for (int64_t i = 1; i < count; ++i) {
      lo = indices[i] < lo ? indices[i] : lo;
      hi = indices[i] > hi ? indices[i] : hi;
    }
    PDM_REQUIRE(lo >= 0 && hi < e->n_samples, PDM_ERR_ARG, "pdm_engine_run: index range [%lld, %lld] outside [0, %lld)",
                (long long)lo, (long long)hi, (long long)e->n_samples);
  }
  PDM_CHECK_CUDA(cudaSetDevice(e->device));
  const int64_t nsteps = (count + batch - 1) / batch;
  if (train) {
    int rc = ensure_results(e, nsteps);
    if (rc != PDM_OK) return rc;
  }
  std::shared_ptr<Job> job = std::make_shared<Job>();
  job->indices = indices;
  job->count = count;
  job->batch = batch;
  job->nsteps = nsteps;
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->job = job;
    for (int s = 0; s < e->slots; ++s) { e->filled_batch[s] = -1; e->free_for[s] = s; }
    e->job_id += 1;
  }
  e->last_launches = nsteps;
  DBG("run: %lld steps of batch %d (train=%d)", (long long)nsteps, batch, train);
  e->cv_work.notify_all();

  // both streams start after everything already queued on the device
  PDM_CHECK_CUDA(cudaEventRecord(e->ev_begin, e->stream[0]));
  PDM_CHECK_CUDA(cudaStreamWaitEvent(e->stream[1], e->ev_begin, 0));
  PDM_CHECK_CUDA(cudaStreamWaitEvent(e->stream[2], e->ev_begin, 0));

  int rc = PDM_OK;
  pdm_hyper h;
  if (hyper) h = *hyper;
  else memset(&h, 0, sizeof(h));
  const bool prof = profile_on();
  HostProf hp;
  const auto t_loop0 = std::chrono::steady_clock::now();
  cudaStream_t cs = e->stream[0];   // compute
  for (int64_t b = 0; b < nsteps && rc == PDM_OK; ++b) {
    if (prof) hp.start();
    const int s = (int)(b % e->slots);
    const int64_t lo = b * batch;
    const int n = (int)std::min<int64_t>(batch, count - lo);
    {
      std::unique_lock<std::mutex> lk(e->mu);
      e->cv_filled.wait(lk, [&] { return e->filled_batch[s] == b; });
    }
    if (prof) { hp.lap(0); if (b == 0) hp.first_wait = hp.t[0]; }
    DBG("main issues step %lld (slot %d, n=%d)", (long long)b, s, n);
    // ---- copy stream: device slot s is free once the kernel of step b - slots has run ----
    const size_t bytes = (size_t)n * kSample;
    cudaError_t ce = cudaSuccess;
    // two copy streams take turns so one stream's event bookkeeping overlaps the other's DMA
    cudaStream_t xs = e->stream[1 + (b & 1)];
    // the stream-side wait is only needed when the consuming kernel is not known to be done
    if (b >= e->slots && cudaEventQuery(e->consumed[s]) != cudaSuccess) ce = cudaStreamWaitEvent(xs, e->consumed[s], 0);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(e->dev_slot[s], e->host_slot[s], bytes, cudaMemcpyHostToDevice, xs);
    if (ce == cudaSuccess) ce = cudaEventRecord(e->h2d_done[s], xs);
    // ---- compute stream: kernels run back to back, each after its slot's copy ----
    if (ce == cudaSuccess) ce = cudaStreamWaitEvent(cs, e->h2d_done[s], 0);
    if (ce != cudaSuccess) {
      pdm_set_error("pdm_engine_run: step %lld copy/record -> %s", (long long)b, cudaGetErrorString(ce));
      rc = PDM_ERR_CUDA;
      break;
    }
    if (prof) hp.lap(1);
    if (train) {
      pdm_net_step_args a;
      memset(&a, 0, sizeof(a));
      a.x = e->dev_slot[s];
      a.x_dtype = PDM_U8;
      a.y = e->dev_slot[s] + (size_t)n * kImg;
      a.y_dtype = PDM_U8;
      a.batch = n;
      a.ctas = ctas;
      a.params = e->params;
      a.m = e->m;
      a.v = e->v;
      a.scratch = e->scratch;
      a.sync = e->sync;
      a.metrics = e->metrics;
      a.step_out = e->host_results_dev + 2 * b;     // mapped host memory: the step's D2H transfer
      a.hyper = h;
      a.hyper.step = h.step + b;
      if (e->comm != nullptr) {
        rc = pdm_comm_peers(e->comm, &a.peers, 1);
        if (rc != PDM_OK) break;
      }
      a.flag_dependency = b == 0 ? 2 : 1;   // first step of a run: ordered after whatever ran before by the grid dependency
      a.done_before = e->done_total;
      a.inputs_stable = 1;                  // the slot was filled by a memcpy that completed before the launch
      rc = pdm_net_train_step(&a, cs);
      if (rc == PDM_OK) e->done_total += (uint32_t)pdm_net_planned_ctas(n, ctas);
    } else {
      rc = pdm_net_eval_step(e->dev_slot[s], PDM_U8, e->dev_slot[s] + (size_t)n * kImg, PDM_U8, n, e->params, nullptr,
                             e->metrics, cs);
    }
    if (rc != PDM_OK) break;
    if (prof) hp.lap(2);
    ce = cudaEventRecord(e->consumed[s], cs);
    if (ce != cudaSuccess) {
      pdm_set_error("pdm_engine_run: step %lld record -> %s", (long long)b, cudaGetErrorString(ce));
      rc = PDM_ERR_CUDA;
      break;
    }
    // The pinned staging slot may be refilled once its H2D copy has completed.  Look at the
    // slot used slots/2 steps ago so the wait is normally already satisfied.
    if (prof) hp.lap(3);
    const int64_t old = b - e->slots / 2;
    if (old >= 0) {
      const int so = (int)(old % e->slots);
      ce = cudaEventSynchronize(e->h2d_done[so]);
      if (ce != cudaSuccess) {
        pdm_set_error("pdm_engine_run: step %lld wait -> %s", (long long)b, cudaGetErrorString(ce));
        rc = PDM_ERR_CUDA;
        break;
      }
      {
        std::lock_guard<std::mutex> lk(e->mu);
        e->free_for[so] = old + e->slots;
      }
      e->cv_free.notify_all();
    }
    if (prof) hp.lap(4);
  }
  if (rc != PDM_OK) {
    // release the workers before reporting the failure
    {
      std::lock_guard<std::mutex> lk(e->mu);
      job->next.store(nsteps);
      for (int s = 0; s < e->slots; ++s) e->free_for[s] = INT64_MAX;
    }
    e->cv_free.notify_all();
  }
  if (prof) {
    const double tot = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_loop0).count();
    fprintf(stderr, "[pdm_engine] host us/step over %lld steps: wait-staged %.2f | h2d+events %.2f | launch %.2f | "
            "record %.2f | slot-sync %.2f | loop total %.2f | first batch ready after %.1f us\n", (long long)nsteps,
            hp.t[0] / nsteps, hp.t[1] / nsteps, hp.t[2] / nsteps, hp.t[3] / nsteps, hp.t[4] / nsteps, tot / nsteps,
            hp.first_wait);
  }
  DBG("all steps issued, draining streams");
  cudaStreamSynchronize(e->stream[1]);
  cudaStreamSynchronize(e->stream[2]);
  cudaEventRecord(e->ev_end, e->stream[0]);
  cudaError_t se = cudaStreamSynchronize(e->stream[0]);
  {
    std::unique_lock<std::mutex> lk(e->mu);
    for (int s = 0; s < e->slots; ++s) e->free_for[s] = INT64_MAX;
    e->cv_free.notify_all();
    e->cv_filled.wait(lk, [&] { return e->workers_active == 0 && job->next.load() >= nsteps; });
  }
  DBG("run done");
  if (rc != PDM_OK) return rc;
  if (se != cudaSuccess) {
    pdm_set_error("pdm_engine_run: device error after %lld steps -> %s", (long long)nsteps, cudaGetErrorString(se));
    return PDM_ERR_CUDA;
  }
  if (device_ms) {
    float ms = 0.f;
    PDM_CHECK_CUDA(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
    *device_ms = ms;
  }
  if (train && step_results) memcpy(step_results, e->host_results, (size_t)nsteps * 2 * sizeof(float));
  return PDM_OK;
}

// Run `nsteps` training steps over a DEVICE-resident pool of samples: step i reads batch
// (i mod (pool_samples / batch)).  No host traffic; one launch per step.  x_dtype PDM_U8 (raw bytes,
// normalised in-kernel) or PDM_F32 (already normalised, what the reference's loader yields);
// y_dtype PDM_U8 / PDM_I64 / PDM_I32.
int pdm_engine_run_device_typed(pdm_engine* e, const void* dev_images, int x_dtype, const void* dev_labels, int y_dtype,
                                int64_t pool_samples, int batch, int64_t nsteps, const pdm_hyper* hyper, int ctas,
                                float* device_ms) {
  PDM_REQUIRE(e && dev_images && dev_labels && hyper, PDM_ERR_ARG, "pdm_engine_run_device: NULL argument");
  PDM_REQUIRE(e->params != nullptr, PDM_ERR_STATE, "pdm_engine_run_device: pdm_engine_bind was not called");
  PDM_REQUIRE(batch >= 1 && pool_samples >= batch, PDM_ERR_ARG, "pdm_engine_run_device: pool smaller than one batch");
  PDM_REQUIRE(x_dtype == PDM_U8 || x_dtype == PDM_F32, PDM_ERR_ARG, "pdm_engine_run_device: x_dtype must be PDM_U8 or PDM_F32");
  PDM_REQUIRE(y_dtype == PDM_U8 || y_dtype == PDM_I64 || y_dtype == PDM_I32, PDM_ERR_ARG,
              "pdm_engine_run_device: y_dtype must be PDM_U8, PDM_I64 or PDM_I32");
  const size_t x_row = (size_t)kImg * (x_dtype == PDM_U8 ? 1 : 4);
  const size_t y_row = y_dtype == PDM_U8 ? 1 : (y_dtype == PDM_I64 ? 8 : 4);
  PDM_CHECK_CUDA(cudaSetDevice(e->device));
  const int64_t nb = pool_samples / batch;
  cudaStream_t st = e->stream[0];
  if (nsteps >= 1 && use_tc(e, x_dtype, y_dtype, ctas, batch, false)) {
    pdm_net_tc_args t;
    memset(&t, 0, sizeof(t));
    t.x = static_cast<const uint8_t*>(dev_images);
    t.y = static_cast<const uint8_t*>(dev_labels);
    t.pool_rows = pool_samples;
    t.batch = batch;
    t.nsteps = nsteps;
    t.params = e->params; t.m = e->m; t.v = e->v;
    t.metrics = e->metrics;
    t.hyper = *hyper;
    if (e->comm != nullptr) {
      int rc = pdm_comm_peers_block(e->comm, &t.peers, nsteps);
      if (rc != PDM_OK) return rc;
    }
    PDM_CHECK_CUDA(cudaEventRecord(e->ev_begin, st));
    int rc = pdm_net_tc_run(&t, st);
    if (rc != PDM_OK) return rc;
    PDM_CHECK_CUDA(cudaEventRecord(e->ev_end, st));
    PDM_CHECK_CUDA(cudaStreamSynchronize(st));
    e->last_launches = (nsteps + PDM_TC_MAX_LAUNCH_STEPS - 1) / PDM_TC_MAX_LAUNCH_STEPS;
    if (device_ms) {
      float ms = 0.f;
      PDM_CHECK_CUDA(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
      *device_ms = ms;
    }
    return PDM_OK;
  }
  e->last_launches = nsteps;
  PDM_CHECK_CUDA(cudaEventRecord(e->ev_begin, st));
  for (int64_t i = 0; i < nsteps; ++i) {
    const int64_t b = i % nb;
    pdm_net_step_args a;
    memset(&a, 0, sizeof(a));
    a.x = static_cast<const uint8_t*>(dev_images) + (size_t)b * batch * x_row;
    a.x_dtype = x_dtype;
    a.y = static_cast<const uint8_t*>(dev_labels) + (size_t)b * batch * y_row;
    a.y_dtype = y_dtype;
    a.batch = batch;
    a.ctas = ctas;
    a.params = e->params;
    a.m = e->m;
    a.v = e->v;
    a.scratch = e->scratch;
    a.sync = e->sync;
    a.metrics = e->metrics;
    a.hyper = *hyper;
    a.hyper.step = hyper->step + i;
    if (e->comm != nullptr) {
      int rc = pdm_comm_peers(e->comm, &a.peers, 1);
      if (rc != PDM_OK) return rc;
    }
    a.flag_dependency = i == 0 ? 2 : 1;
    a.done_before = e->done_total;
    a.inputs_stable = 1;                    // resident pool
    int rc = pdm_net_train_step(&a, st);
    if (rc != PDM_OK) return rc;
    e->done_total += (uint32_t)pdm_net_planned_ctas(batch, ctas);
  }
  PDM_CHECK_CUDA(cudaEventRecord(e->ev_end, st));
  PDM_CHECK_CUDA(cudaStreamSynchronize(st));
  if (device_ms) {
    float ms = 0.f;
    PDM_CHECK_CUDA(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
    *device_ms = ms;
  }
  return PDM_OK;
}

int pdm_engine_run_device(pdm_engine* e, const uint8_t* dev_images, const uint8_t* dev_labels, int64_t pool_samples,
                          int batch, int64_t nsteps, const pdm_hyper* hyper, int ctas, float* device_ms) {
  return pdm_engine_run_device_typed(e, dev_images, PDM_U8, dev_labels, PDM_U8, pool_samples, batch, nsteps, hyper, ctas,
                                     device_ms);
}

static void drop_placement(pdm_engine::Placement* p) {
  if (p->mode == 0) {
    if (p->images) cudaFree(p->images);
    if (p->labels) cudaFree(p->labels);
  } else if (p->mode == 1) {
    if (p->host_images) cudaHostUnregister(const_cast<uint8_t*>(p->host_images));
    if (p->host_labels) cudaHostUnregister(const_cast<uint8_t*>(p->host_labels));
  }
  *p = pdm_engine::Placement();
  cudaGetLastError();
}

static void unplace_all(pdm_engine* e) {
  for (auto& p : e->placements) drop_placement(&p);
  e->placed = -1;
}

// Make a dataset visible to the in-kernel gather (pdm_engine_run_gather) and select it.
//   mode 0  copy it into device memory once (60,000 MNIST images are 47 MB of a 180 GB HBM)
//   mode 1  page-lock and map the caller's host arrays; the step kernel then reads its rows
//           straight out of host memory over PCIe (zero-copy) -- no staging copy, no H2D memcpy
// Up to four placements are kept (keyed by host pointer, length and mode): selecting one that
// exists costs nothing, so a trainer can switch loaders between passes without paying the
// page-locking again.  The host arrays must stay alive (mode 1: and unmoved) until the engine is
// destroyed or the placement is evicted (least recently used first).
int pdm_engine_place_dataset(pdm_engine* e, const uint8_t* images, const uint8_t* labels, int64_t n, int mode) {
  PDM_REQUIRE(e && images && labels && n >= 1, PDM_ERR_ARG, "pdm_engine_place_dataset: bad arguments");
  PDM_REQUIRE(mode == 0 || mode == 1, PDM_ERR_ARG, "pdm_engine_place_dataset: mode must be 0 (device) or 1 (zero-copy)");
  PDM_REQUIRE(n <= 0x7fffffffLL, PDM_ERR_LIMIT, "pdm_engine_place_dataset: at most 2^31-1 samples");
  e->place_clock += 1;
  int victim = 0;
  for (int i = 0; i < pdm_engine::kMaxPlacements; ++i) {
    pdm_engine::Placement& p = e->placements[i];
    if (p.mode == mode && p.host_images == images && p.host_labels == labels && p.n == n) {
      p.last_use = e->place_clock;
      e->placed = i;
      return PDM_OK;
    }
    if (p.mode < 0 || (e->placements[victim].mode >= 0 && p.last_use < e->placements[victim].last_use)) victim = i;
  }
  PDM_CHECK_CUDA(cudaSetDevice(e->device));
  pdm_engine::Placement& p = e->placements[victim];
  if (p.mode >= 0) {
    PDM_CHECK_CUDA(cudaDeviceSynchronize());    // a queued kernel may still read the evicted placement
    if (e->placed == victim) e->placed = -1;
    drop_placement(&p);
  }
  if (mode == 0) {
    PDM_CHECK_CUDA(cudaMalloc((void**)&p.images, (size_t)n * kImg));
    p.mode = 0;
    PDM_CHECK_CUDA(cudaMalloc((void**)&p.labels, (size_t)n));
    PDM_CHECK_CUDA(cudaMemcpy(p.images, images, (size_t)n * kImg, cudaMemcpyHostToDevice));
    PDM_CHECK_CUDA(cudaMemcpy(p.labels, labels, (size_t)n, cudaMemcpyHostToDevice));
    p.host_images = images;
    p.host_labels = labels;
  } else {
    cudaError_t ce = cudaHostRegister(const_cast<uint8_t*>(images), (size_t)n * kImg, cudaHostRegisterMapped);
    PDM_REQUIRE(ce == cudaSuccess, PDM_ERR_CUDA, "pdm_engine_place_dataset: cudaHostRegister(images) -> %s", cudaGetErrorString(ce));
    p.mode = 1;
    p.host_images = images;
    ce = cudaHostRegister(const_cast<uint8_t*>(labels), (size_t)n, cudaHostRegisterMapped);
    if (ce != cudaSuccess) {
      drop_placement(&p);
      pdm_set_error("pdm_engine_place_dataset: cudaHostRegister(labels) -> %s", cudaGetErrorString(ce));
      return PDM_ERR_CUDA;
    }
    p.host_labels = labels;
    PDM_CHECK_CUDA(cudaHostGetDevicePointer((void**)&p.images, const_cast<uint8_t*>(images), 0));
    PDM_CHECK_CUDA(cudaHostGetDevicePointer((void**)&p.labels, const_cast<uint8_t*>(labels), 0));
    if (e->tc_ring == nullptr && pdm_net_tc_supported(e->device) == 1)      // feeder ring: allocated here, not inside a pass
      PDM_CHECK_CUDA(cudaMalloc((void**)&e->tc_ring, PDM_TC_RING_BYTES));
  }
  p.n = n;
  p.last_use = e->place_clock;
  e->placed = victim;
  // room for the index plan of a whole pass over this set, so that a training pass never allocates
  if (n > e->idx_cap) {
    if (e->idx_dev) { PDM_CHECK_CUDA(cudaDeviceSynchronize()); cudaFree(e->idx_dev); }
    e->idx_dev = nullptr;
    e->idx_cap = 0;
    PDM_CHECK_CUDA(cudaMalloc((void**)&e->idx_dev, (size_t)n * sizeof(int32_t)));
    e->idx_cap = n;
  }
  if (e->idx_host_cap < e->idx_cap) {
    if (e->idx_host) cudaFreeHost(e->idx_host);
    e->idx_host = nullptr;
    e->idx_host_cap = 0;
    PDM_CHECK_CUDA(cudaHostAlloc((void**)&e->idx_host, (size_t)e->idx_cap * sizeof(int32_t), cudaHostAllocDefault));
    e->idx_host_cap = e->idx_cap;
  }
  return PDM_OK;
}

// Training pass with the gather done by the step kernel: the index plan is uploaded once
// (4 bytes per sample), then every step is a single launch that reads its rows from the placed
// dataset (device memory, or host memory over PCIe).  Same results contract as pdm_engine_run.
int pdm_engine_run_gather(pdm_engine* e, const int64_t* indices, int64_t count, int batch, const pdm_hyper* hyper,
                          int ctas, float* step_results, float* device_ms) {
  PDM_REQUIRE(e && indices && hyper, PDM_ERR_ARG, "pdm_engine_run_gather: engine, indices and hyper are required");
  PDM_REQUIRE(e->params != nullptr, PDM_ERR_STATE, "pdm_engine_run_gather: pdm_engine_bind was not called");
  PDM_REQUIRE(e->placed >= 0, PDM_ERR_STATE, "pdm_engine_run_gather: pdm_engine_place_dataset was not called");
  PDM_REQUIRE(batch >= 1, PDM_ERR_ARG, "pdm_engine_run_gather: batch must be >= 1");
  if (count <= 0) return PDM_OK;
  PDM_CHECK_CUDA(cudaSetDevice(e->device));
  const pdm_engine::Placement& pl = e->placements[e->placed];
  if (count > e->idx_cap || count > e->idx_host_cap) {      // a pass longer than the set (several epochs in one plan)
    PDM_CHECK_CUDA(cudaDeviceSynchronize());
    if (e->idx_dev) cudaFree(e->idx_dev);
    if (e->idx_host) cudaFreeHost(e->idx_host);
    e->idx_dev = nullptr; e->idx_host = nullptr;
    e->idx_cap = e->idx_host_cap = 0;
    PDM_CHECK_CUDA(cudaMalloc((void**)&e->idx_dev, (size_t)count * sizeof(int32_t)));
    e->idx_cap = count;
    PDM_CHECK_CUDA(cudaHostAlloc((void**)&e->idx_host, (size_t)count * sizeof(int32_t), cudaHostAllocDefault));
    e->idx_host_cap = count;
  }
  {
    // narrow to 32 bits and range-check in one pass (min/max instead of a branch per element)
    int64_t lo = indices[0], hi = indices[0];
    int32_t* dst = e->idx_host;
    for (int64_t i = 0; i < count; ++i) {
      const int64_t v = indices[i];
      lo = v < lo ? v : lo;
      hi = v > hi ? v : hi;
      dst[i] = (int32_t)v;
    }
    PDM_REQUIRE(lo >= 0 && hi < pl.n, PDM_ERR_ARG, "pdm_engine_run_gather: index range [%lld, %lld] outside [0, %lld)",
                (long long)lo, (long long)hi, (long long)pl.n);
  }
  const int64_t nsteps = (count + batch - 1) / batch;
  int rc = ensure_results(e, nsteps);
  if (rc != PDM_OK) return rc;
  cudaStream_t cs = e->stream[0];
  PDM_CHECK_CUDA(cudaEventRecord(e->ev_begin, cs));
  PDM_CHECK_CUDA(cudaMemcpyAsync(e->idx_dev, e->idx_host, (size_t)count * sizeof(int32_t), cudaMemcpyHostToDevice, cs));
  if (use_tc(e, PDM_U8, PDM_U8, ctas, batch, pl.mode == 1)) {
    // the whole pass is one launch: the kernel's producer warps gather every step's rows themselves
    pdm_net_tc_args t;
    memset(&t, 0, sizeof(t));
    t.x = pl.images;
    t.y = pl.labels;
    t.row_index = e->idx_dev;
    t.pool_rows = pl.n;
    t.count = count;
    t.batch = batch;
    t.params = e->params; t.m = e->m; t.v = e->v;
    t.metrics = e->metrics;
    t.step_out = e->host_results_dev;
    t.hyper = *hyper;
    if (pl.mode == 1) {                     // rows come over PCIe: let the feeder cluster fetch them whole
      if (e->tc_ring == nullptr) PDM_CHECK_CUDA(cudaMalloc((void**)&e->tc_ring, PDM_TC_RING_BYTES));
      t.ring = e->tc_ring;
      t.ring_bytes = PDM_TC_RING_BYTES;
    }
    if (e->comm != nullptr) {
      rc = pdm_comm_peers_block(e->comm, &t.peers, nsteps);
      if (rc != PDM_OK) return rc;
    }
    rc = pdm_net_tc_run(&t, cs);
    if (rc != PDM_OK) return rc;
    e->last_launches = (nsteps + PDM_TC_MAX_LAUNCH_STEPS - 1) / PDM_TC_MAX_LAUNCH_STEPS;
    PDM_CHECK_CUDA(cudaEventRecord(e->ev_end, cs));
    PDM_CHECK_CUDA(cudaStreamSynchronize(cs));
    if (device_ms) {
      float ms = 0.f;
      PDM_CHECK_CUDA(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
      *device_ms = ms;
    }
    if (step_results) memcpy(step_results, e->host_results, (size_t)nsteps * 2 * sizeof(float));
    return PDM_OK;
  }
  e->last_launches = nsteps;
  for (int64_t b = 0; b < nsteps; ++b) {
    const int n = (int)std::min<int64_t>(batch, count - b * batch);
    pdm_net_step_args a;
    memset(&a, 0, sizeof(a));
    a.x = pl.images;
    a.x_dtype = PDM_U8;
    a.y = pl.labels;
    a.y_dtype = PDM_U8;
    a.row_index = e->idx_dev + b * batch;
    a.batch = n;
    a.ctas = ctas;
    a.params = e->params;
    a.m = e->m;
    a.v = e->v;
    a.scratch = e->scratch;
    a.sync = e->sync;
    a.metrics = e->metrics;
    a.step_out = e->host_results_dev + 2 * b;
    a.hyper = *hyper;
    a.hyper.step = hyper->step + b;
    if (e->comm != nullptr) {
      rc = pdm_comm_peers(e->comm, &a.peers, 1);
      if (rc != PDM_OK) return rc;
    }
    a.flag_dependency = b == 0 ? 2 : 1;
    a.done_before = e->done_total;
    a.inputs_stable = 1;                    // resident dataset; the index plan came by memcpy
    rc = pdm_net_train_step(&a, cs);
    if (rc != PDM_OK) return rc;
    e->done_total += (uint32_t)pdm_net_planned_ctas(n, ctas);
  }
  PDM_CHECK_CUDA(cudaEventRecord(e->ev_end, cs));
  PDM_CHECK_CUDA(cudaStreamSynchronize(cs));
  if (device_ms) {
    float ms = 0.f;
    PDM_CHECK_CUDA(cudaEventElapsedTime(&ms, e->ev_begin, e->ev_end));
    *device_ms = ms;
  }
  if (step_results) memcpy(step_results, e->host_results, (size_t)nsteps * 2 * sizeof(float));
  return PDM_OK;
}

// Copy out {sum loss*n, correct, samples, 0}; reset != 0 zeroes the accumulator.  Synchronises.
int pdm_engine_metrics(pdm_engine* e, double out[4], int reset) {
  PDM_REQUIRE(e && out, PDM_ERR_ARG, "pdm_engine_metrics: NULL argument");
  PDM_CHECK_CUDA(cudaSetDevice(e->device));
  PDM_CHECK_CUDA(cudaDeviceSynchronize());
  PDM_CHECK_CUDA(cudaMemcpy(out, e->metrics, 4 * sizeof(double), cudaMemcpyDeviceToHost));
  if (reset) PDM_CHECK_CUDA(cudaMemset(e->metrics, 0, 4 * sizeof(double)));
  return PDM_OK;
}

int64_t pdm_engine_last_launches(pdm_engine* e) { return e ? e->last_launches : 0; }
void* pdm_engine_scratch(pdm_engine* e) { return e ? e->scratch : nullptr; }
void* pdm_engine_sync_words(pdm_engine* e) { return e ? e->sync : nullptr; }
void* pdm_engine_metrics_ptr(pdm_engine* e) { return e ? e->metrics : nullptr; }

int pdm_engine_destroy(pdm_engine* e) {
  if (!e) return PDM_OK;
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->quit = true;
  }
  e->cv_work.notify_all();
  e->cv_free.notify_all();
  for (auto& t : e->workers)
    if (t.joinable()) t.join();
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  for (auto p : e->dev_slot) if (p) cudaFree(p);
  for (auto p : e->host_slot) if (p) cudaFreeHost(p);
  for (auto ev : e->h2d_done) if (ev) cudaEventDestroy(ev);
  for (auto ev : e->consumed) if (ev) cudaEventDestroy(ev);
  for (int i = 0; i < 3; ++i) {
    if (e->stream[i]) cudaStreamDestroy(e->stream[i]);
  }
  if (e->ev_begin) cudaEventDestroy(e->ev_begin);
  if (e->ev_end) cudaEventDestroy(e->ev_end);
  if (e->host_results) cudaFreeHost(e->host_results);
  unplace_all(e);
  if (e->tc_ring) cudaFree(e->tc_ring);
  if (e->idx_dev) cudaFree(e->idx_dev);
  if (e->idx_host) cudaFreeHost(e->idx_host);
  if (e->scratch) cudaFree(e->scratch);
  if (e->sync) cudaFree(e->sync);
  if (e->metrics) cudaFree(e->metrics);
  if (e->step_out) cudaFree(e->step_out);
  delete e;
  cudaGetLastError();
  return PDM_OK;
}

}  // extern "C"
