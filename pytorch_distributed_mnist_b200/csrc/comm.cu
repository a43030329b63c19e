// comm.cu -- peer-memory communicator: the B200-native replacement for the
// ProcessGroupNCCL + NCCL + c10d Reducer gradient path (SURVEY N2-N4, N7,
// K9, K15, 5.8).
//
// One process per GPU.  Every rank cudaMalloc's one *symmetric region* with the
// same layout, exports it as a CUDA IPC handle, and maps every peer's region
// (NVLink P2P through NVSwitch: every peer is one hop at full bandwidth, so no
// ring or tree is built).  Collectives are then plain kernels that load/store
// peer addresses:
//
//   region = [ LL inbox (fused Net step, net_step.cu) | flag words | data ]
//
//   barrier            flag exchange only
//   broadcast          every rank pulls root's bytes into its own data region
//                      (replaces DDP's _broadcast_coalesced, K15)
//   all-reduce(+optim) one-shot  : each rank reads all peers' gradients, sums
//                                  in rank order, applies the optimizer to
//                                  its full replica (small buckets)
//                      two-shot  : rank r reduces slice r (reduce-scatter),
//                                  applies the optimizer to that slice only,
//                                  and stores the updated *parameters* into
//                                  every peer's replica (all-gather) -- the
//                                  reduced gradient is never written back
//                                  (large buckets; optimizer work is 1/W)
//
// Ordering protocol.  Flags are monotonically increasing
// 32-bit epochs, one word per (kind, source rank), written with
// st.release.sys after a __threadfence_system and polled locally with
// ld.acquire.sys.  Every collective starts with "my inputs are ready" and ends
// with "I am done touching your memory", so a rank never overwrites a buffer a
// peer may still read and never reads parameters a peer is still writing.
// Every wait is bounded: on timeout the kernel records the missing rank in a
// status word and returns instead of hanging the GPU.
// Reduction order is rank 0..W-1 on every rank, so replicas stay bit-identical.
#include "common.cuh"

#include <math.h>
#include <new>

namespace {

constexpr size_t kLLBytes = 2u << 20;  // >= 2 * 8 * 1963 * 32 (net_step.cu) and >= 2 * 8 * 13 * 960 * 8 (net_tc.cu)
constexpr size_t kFlagBytes = 4096;
constexpr size_t kHeaderBytes = kLLBytes + kFlagBytes;
constexpr int kFlagStart = 0;   // word offsets inside the flag block (32 words = 128 B apart)
constexpr int kFlagEnd = 32;
constexpr int kFlagBar = 64;
constexpr int kFlagStatus = 96;
constexpr int kFlagGrid = 128;  // 2 words: local grid barrier

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct Links {
  int world, rank;
  uint32_t epoch;
  uint64_t timeout_ns;
  uint32_t skew_ns;              // fault injection (PDM_SKEW_NS): random pause before every flag this rank raises
  uint8_t* base[PDM_MAX_RANKS];  // region base of every rank, as mapped here
};

__device__ __forceinline__ uint32_t* flag_word(const Links& L, int owner, int kind, int src) {
  return reinterpret_cast<uint32_t*>(L.base[owner] + kLLBytes) + kind + src;
}

// Thread p < world of the calling CTA tells peer p "epoch reached" on `kind`.
// Call after a CTA/grid barrier that orders the data writes before it.
__device__ __forceinline__ void signal_peers(const Links& L, int kind) {
  if ((int)threadIdx.x < L.world) {
    pdm_skew(L.skew_ns, L.epoch, (uint32_t)(L.rank * 8 + threadIdx.x), (uint32_t)kind);
    __threadfence_system();
    st_release_sys(flag_word(L, threadIdx.x, kind, L.rank), L.epoch);
  }
}
// Thread p < world waits until peer p has signalled `kind` for this epoch.
__device__ __forceinline__ void wait_peers(const Links& L, int kind) {
  if ((int)threadIdx.x < L.world) {
    const uint32_t* f = flag_word(L, L.rank, kind, threadIdx.x);
    uint64_t t0 = 0;
    uint32_t spins = 0;
    // signed distance handles wrap-around of the epoch counter
    while ((int32_t)(ld_acquire_sys(f) - L.epoch) < 0) {
      if ((++spins & 255u) == 0u && L.timeout_ns != 0) {
        const uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > L.timeout_ns) {
          atomicExch(reinterpret_cast<int*>(reinterpret_cast<uint32_t*>(L.base[L.rank] + kLLBytes) + kFlagStatus),
                     1 + (int)threadIdx.x);
          break;
        }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint32_t* grid_words(const Links& L) {
  return reinterpret_cast<uint32_t*>(L.base[L.rank] + kLLBytes) + kFlagGrid;
}

__global__ void barrier_kernel(const __grid_constant__ Links L) {
  signal_peers(L, kFlagBar);
  wait_peers(L, kFlagBar);
}

// Every rank copies root's [off, off+bytes) into its own region.
__global__ void __launch_bounds__(256) broadcast_kernel(const __grid_constant__ Links L, size_t off, size_t n16, int root) {
  if (blockIdx.x == 0) signal_peers(L, kFlagStart);  // "my region is quiescent, you may proceed"
  wait_peers(L, kFlagStart);
  if (L.rank != root) {
    const uint4* src = reinterpret_cast<const uint4*>(L.base[root] + kHeaderBytes + off);
    uint4* dst = reinterpret_cast<uint4*>(L.base[L.rank] + kHeaderBytes + off);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) dst[i] = src[i];
  }
  if (gridDim.x > 1) grid_barrier(grid_words(L), gridDim.x);
  else __syncthreads();
  if (blockIdx.x == 0) {
    signal_peers(L, kFlagEnd);   // root must not change the bytes until everyone has them
    wait_peers(L, kFlagEnd);
  }
}

struct OptK {
  int opt, first;
  float lr, beta1, beta2, eps, wd, step_size, bc2_sqrt, gscale;
};

__device__ __forceinline__ void upd(const OptK& k, float g, float& p, float& m, float& v) {
  g *= k.gscale;
  if (k.opt == PDM_OPT_NONE) { p = g; return; }   // plain all-reduce: "p" is the output
  if (k.wd != 0.f) g = fmaf(k.wd, p, g);
  if (k.opt == PDM_OPT_ADAM) {
    m = m + (1.f - k.beta1) * (g - m);
    v = v * k.beta2 + (1.f - k.beta2) * g * g;
    const float denom = sqrtf(v) / k.bc2_sqrt + k.eps;
    p = p - k.step_size * (m / denom);
  } else {
    if (k.beta1 != 0.f) {
      m = k.first ? g : (m * k.beta1 + g);
      g = m;
    }
    p = p - k.lr * g;
  }
}

struct ReduceArgs {
  size_t grad_off, out_off, m_off, v_off;  // byte offsets in the data region; out = params (or all-reduce output)
  size_t n4;                               // float4 count
  int two_shot;
  OptK k;
};

__device__ __forceinline__ float4 ld_peer(const uint8_t* base, size_t off, size_t i) {
  // peer lines are never cached in the local L2 (UVA aperture); the relaxed.sys load also skips L1
  float4 r;
  const float4* p = reinterpret_cast<const float4*>(base + kHeaderBytes + off) + i;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}

// Fused all-reduce + scale + optimizer (+ parameter all-gather in two-shot mode).
__global__ void __launch_bounds__(512) reduce_optim_kernel(const __grid_constant__ Links L, const __grid_constant__ ReduceArgs A) {
  const int W = L.world;
  if (blockIdx.x == 0) signal_peers(L, kFlagStart);   // my gradients are final (stream order)
  wait_peers(L, kFlagStart);                          // everyone's gradients are final

  size_t lo = 0, hi = A.n4;
  if (A.two_shot) {                                   // slice owned by this rank
    const size_t per = (A.n4 + W - 1) / W;
    lo = min(A.n4, per * (size_t)L.rank);
    hi = min(A.n4, lo + per);
  }
  uint8_t* self = L.base[L.rank] + kHeaderBytes;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int p0 = 0; p0 < W; p0 += 4) {               // up to 4 peer loads in flight, summed in rank order
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (p0 + u < W) t[u] = ld_peer(L.base[p0 + u], A.grad_off, i);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (p0 + u < W) { g.x += t[u].x; g.y += t[u].y; g.z += t[u].z; g.w += t[u].w; }
    }
    float4 p4 = make_float4(0.f, 0.f, 0.f, 0.f), m4 = p4, v4 = p4;
    const bool has_state = A.k.opt != PDM_OPT_NONE;
    if (has_state) {
      p4 = reinterpret_cast<float4*>(self + A.out_off)[i];
      m4 = reinterpret_cast<float4*>(self + A.m_off)[i];
      if (A.k.opt == PDM_OPT_ADAM) v4 = reinterpret_cast<float4*>(self + A.v_off)[i];
    }
    upd(A.k, g.x, p4.x, m4.x, v4.x);
    upd(A.k, g.y, p4.y, m4.y, v4.y);
    upd(A.k, g.z, p4.z, m4.z, v4.z);
    upd(A.k, g.w, p4.w, m4.w, v4.w);
    if (has_state) {
      reinterpret_cast<float4*>(self + A.m_off)[i] = m4;
      if (A.k.opt == PDM_OPT_ADAM) reinterpret_cast<float4*>(self + A.v_off)[i] = v4;
    }
    if (A.two_shot) {
      for (int p = 0; p < W; ++p)                     // all-gather: push the owner's result everywhere
        reinterpret_cast<float4*>(L.base[p] + kHeaderBytes + A.out_off)[i] = p4;
    } else {
      reinterpret_cast<float4*>(self + A.out_off)[i] = p4;
    }
  }
  if (gridDim.x > 1) grid_barrier(grid_words(L), gridDim.x);
  else __syncthreads();
  if (blockIdx.x == 0) {
    signal_peers(L, kFlagEnd);   // done reading your gradients / writing your parameters
    wait_peers(L, kFlagEnd);
  }
}

}  // namespace

struct pdm_comm {
  int rank = 0, world = 1, device = 0;
  size_t data_bytes = 0, total_bytes = 0;
  uint8_t* local = nullptr;
  uint8_t* base[PDM_MAX_RANKS] = {};
  bool opened[PDM_MAX_RANKS] = {};
  bool connected = false;
  uint32_t epoch = 0;     // collective counter (flags)
  uint32_t ll_seq = 0;    // fused-step counter (LL flags)
  uint64_t timeout_ns = 30ull * 1000000000ull;
  int sm_count = 148;
};

static Links make_links(pdm_comm* c) {
  Links L;
  memset(&L, 0, sizeof(L));
  L.world = c->world;
  L.rank = c->rank;
  c->epoch += 1;
  if (c->epoch == 0) c->epoch = 1;
  L.epoch = c->epoch;
  L.timeout_ns = c->timeout_ns;
  L.skew_ns = pdm_skew_ns();
  for (int p = 0; p < c->world; ++p) L.base[p] = c->base[p];
  return L;
}

extern "C" {

int pdm_comm_create(pdm_comm** out, int rank, int world, int device, size_t data_bytes) {
  PDM_REQUIRE(out != nullptr, PDM_ERR_ARG, "pdm_comm_create: out is NULL");
  PDM_REQUIRE(world >= 1 && world <= PDM_MAX_RANKS, PDM_ERR_LIMIT, "pdm_comm_create: world %d outside [1, %d]", world,
              PDM_MAX_RANKS);
  PDM_REQUIRE(rank >= 0 && rank < world, PDM_ERR_ARG, "pdm_comm_create: rank %d outside [0, %d)", rank, world);
  PDM_CHECK_CUDA(cudaSetDevice(device));
  pdm_comm* c = new (std::nothrow) pdm_comm();
  PDM_REQUIRE(c != nullptr, PDM_ERR_STATE, "pdm_comm_create: out of host memory");
  c->rank = rank;
  c->world = world;
  c->device = device;
  c->data_bytes = (data_bytes + 255) / 256 * 256;
  c->total_bytes = kHeaderBytes + c->data_bytes;
  cudaError_t e = cudaMalloc((void**)&c->local, c->total_bytes);
  if (e != cudaSuccess) {
    pdm_set_error("pdm_comm_create: cudaMalloc(%zu) -> %s", c->total_bytes, cudaGetErrorString(e));
    delete c;
    return PDM_ERR_CUDA;
  }
  e = cudaMemset(c->local, 0, c->total_bytes);
  if (e != cudaSuccess) {
    pdm_set_error("pdm_comm_create: cudaMemset -> %s", cudaGetErrorString(e));
    cudaFree(c->local);
    delete c;
    return PDM_ERR_CUDA;
  }
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  c->base[rank] = c->local;
  c->connected = (world == 1);
  *out = c;
  return PDM_OK;
}

int pdm_comm_handle(pdm_comm* c, void* out64) {
  PDM_REQUIRE(c && out64, PDM_ERR_ARG, "pdm_comm_handle: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  cudaIpcMemHandle_t h;
  PDM_CHECK_CUDA(cudaIpcGetMemHandle(&h, c->local));
  memcpy(out64, &h, sizeof(h));
  return PDM_OK;
}

int pdm_comm_connect(pdm_comm* c, const void* handles) {
  PDM_REQUIRE(c && handles, PDM_ERR_ARG, "pdm_comm_connect: NULL argument");
  PDM_CHECK_CUDA(cudaSetDevice(c->device));
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank || c->opened[p]) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const uint8_t*)handles + 64 * (size_t)p, sizeof(h));
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      pdm_set_error("pdm_comm_connect: rank %d cannot map rank %d's region: %s (is P2P/NVLink available "
                    "between the two devices and are both visible to this process?)",
                    c->rank, p, cudaGetErrorString(e));
      return PDM_ERR_CUDA;
    }
    c->base[p] = (uint8_t*)ptr;
    c->opened[p] = true;
  }
  c->connected = true;
  return PDM_OK;
}

void* pdm_comm_data(pdm_comm* c) { return c ? c->local + kHeaderBytes : nullptr; }
size_t pdm_comm_data_bytes(pdm_comm* c) { return c ? c->data_bytes : 0; }

int pdm_comm_set_timeout(pdm_comm* c, double seconds) {
  PDM_REQUIRE(c, PDM_ERR_ARG, "pdm_comm_set_timeout: NULL comm");
  c->timeout_ns = seconds <= 0 ? 0 : (uint64_t)(seconds * 1e9);
  return PDM_OK;
}

// Peer links for the fused Net step.  advance != 0 consumes one LL sequence number.
int pdm_comm_peers(pdm_comm* c, pdm_peers* out, int advance) {
  PDM_REQUIRE(c && out, PDM_ERR_ARG, "pdm_comm_peers: NULL argument");
  PDM_REQUIRE(c->connected, PDM_ERR_STATE, "pdm_comm_peers: peers not connected yet");
  memset(out, 0, sizeof(*out));
  out->world = c->world;
  out->rank = c->rank;
  if (advance) {
    c->ll_seq += 1;
    if (c->ll_seq == 0) c->ll_seq = 1;
  }
  out->seq = c->ll_seq ? c->ll_seq : 1;
  for (int p = 0; p < c->world; ++p) out->inbox[p] = reinterpret_cast<uint64_t*>(c->base[p]);
  out->status = reinterpret_cast<int32_t*>(reinterpret_cast<uint32_t*>(c->local + kLLBytes) + kFlagStatus);
  out->timeout_ns = c->timeout_ns;
  return PDM_OK;
}

// Peer links for a persistent multi-step kernel (net_tc.cu): reserves `nsteps` consecutive LL sequence
// numbers; out->seq is the first of them.
int pdm_comm_peers_block(pdm_comm* c, pdm_peers* out, int64_t nsteps) {
  PDM_REQUIRE(nsteps >= 1, PDM_ERR_ARG, "pdm_comm_peers_block: nsteps must be >= 1");
  int rc = pdm_comm_peers(c, out, 1);
  if (rc != PDM_OK) return rc;
  c->ll_seq += (uint32_t)(nsteps - 1);
  if (c->ll_seq < out->seq) {            // wrapped past zero inside the block: start the block after the wrap
    c->ll_seq = 0;
    return pdm_comm_peers_block(c, out, nsteps);
  }
  return PDM_OK;
}

// Nonzero (1 + missing rank) once any wait has timed out.  Synchronises the device.
int pdm_comm_status(pdm_comm* c, int* status) {
  PDM_REQUIRE(c && status, PDM_ERR_ARG, "pdm_comm_status: NULL argument");
  PDM_CHECK_CUDA(cudaMemcpy(status, reinterpret_cast<uint32_t*>(c->local + kLLBytes) + kFlagStatus, sizeof(int),
                            cudaMemcpyDeviceToHost));
  return PDM_OK;
}

int pdm_comm_barrier(pdm_comm* c, pdm_stream_t stream) {
  PDM_REQUIRE(c && c->connected, PDM_ERR_STATE, "pdm_comm_barrier: not connected");
  if (c->world == 1) return PDM_OK;
  Links L = make_links(c);
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(L);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_comm_broadcast(pdm_comm* c, size_t offset, size_t bytes, int root, pdm_stream_t stream) {
  PDM_REQUIRE(c && c->connected, PDM_ERR_STATE, "pdm_comm_broadcast: not connected");
  PDM_REQUIRE(root >= 0 && root < c->world, PDM_ERR_ARG, "pdm_comm_broadcast: root %d outside [0, %d)", root, c->world);
  PDM_REQUIRE(offset % 16 == 0 && bytes % 16 == 0, PDM_ERR_ARG, "pdm_comm_broadcast: offset/bytes must be multiples of 16");
  PDM_REQUIRE(offset + bytes <= c->data_bytes, PDM_ERR_LIMIT, "pdm_comm_broadcast: [%zu, %zu) outside the %zu-byte data region",
              offset, offset + bytes, c->data_bytes);
  if (c->world == 1 || bytes == 0) return PDM_OK;
  Links L = make_links(c);
  const size_t n16 = bytes / 16;
  size_t blocks = (n16 + 256 * 8 - 1) / (256 * 8);
  if (blocks < 1) blocks = 1;
  if (blocks > (size_t)c->sm_count) blocks = c->sm_count;
  broadcast_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(L, offset, n16, root);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

// algo: 0 = auto, 1 = one-shot, 2 = two-shot.  hyper->opt == PDM_OPT_NONE makes this a plain
// all-reduce (mean if grad_scale = 1/world) whose result lands at out_off.
int pdm_comm_allreduce_optim(pdm_comm* c, size_t grad_off, size_t out_off, size_t m_off, size_t v_off, size_t n_floats,
                             const pdm_hyper* h, float grad_scale, int algo, int max_ctas, pdm_stream_t stream) {
  PDM_REQUIRE(c && c->connected, PDM_ERR_STATE, "pdm_comm_allreduce_optim: not connected");
  PDM_REQUIRE(h != nullptr, PDM_ERR_ARG, "pdm_comm_allreduce_optim: hyper is NULL");
  PDM_REQUIRE(n_floats % 4 == 0, PDM_ERR_ARG, "pdm_comm_allreduce_optim: n_floats (%zu) must be a multiple of 4", n_floats);
  PDM_REQUIRE((grad_off | out_off | m_off | v_off) % 16 == 0, PDM_ERR_ARG, "pdm_comm_allreduce_optim: offsets must be 16-byte aligned");
  const size_t bytes = n_floats * 4;
  PDM_REQUIRE(grad_off + bytes <= c->data_bytes && out_off + bytes <= c->data_bytes, PDM_ERR_LIMIT,
              "pdm_comm_allreduce_optim: range outside the %zu-byte data region", c->data_bytes);
  if (h->opt != PDM_OPT_NONE)
    PDM_REQUIRE(m_off + bytes <= c->data_bytes && v_off + bytes <= c->data_bytes, PDM_ERR_LIMIT,
                "pdm_comm_allreduce_optim: state range outside the data region");
  if (n_floats == 0) return PDM_OK;
  if (c->world == 1 && h->opt != PDM_OPT_NONE) {
    // No peers: the bucket kernel would be a cooperative grid barrier around the plain optimizer update
    // (ncu: 20 us for a 3-CTA bucket, half of it at the barrier).  Same update function, same result.
    uint8_t* d = c->local + kHeaderBytes;
    return pdm_optim_step(reinterpret_cast<float*>(d + out_off), reinterpret_cast<const float*>(d + grad_off),
                          reinterpret_cast<float*>(d + m_off), reinterpret_cast<float*>(d + v_off), n_floats, h,
                          grad_scale, stream);
  }
  ReduceArgs A;
  memset(&A, 0, sizeof(A));
  A.grad_off = grad_off; A.out_off = out_off; A.m_off = m_off; A.v_off = v_off;
  A.n4 = n_floats / 4;
  if (algo == 0) algo = (bytes >= (size_t)256 * 1024 && c->world > 1) ? 2 : 1;
  A.two_shot = (algo == 2 && c->world > 1) ? 1 : 0;
  PDM_REQUIRE(!(h->opt == PDM_OPT_NONE && !A.two_shot && out_off == grad_off && c->world > 1), PDM_ERR_ARG,
              "pdm_comm_allreduce_optim: one-shot all-reduce cannot run in place (peers still read the input)");
  A.k.opt = h->opt;
  A.k.first = h->step <= 1;
  A.k.lr = h->lr; A.k.beta1 = h->beta1; A.k.beta2 = h->beta2; A.k.eps = h->eps; A.k.wd = h->weight_decay;
  A.k.gscale = grad_scale;
  A.k.step_size = 0.f; A.k.bc2_sqrt = 1.f;
  if (h->opt == PDM_OPT_ADAM) {
    const double step = (double)(h->step < 1 ? 1 : h->step);
    A.k.step_size = (float)((double)h->lr / (1.0 - pow((double)h->beta1, step)));
    A.k.bc2_sqrt = (float)sqrt(1.0 - pow((double)h->beta2, step));
  }
  Links L = make_links(c);
  const size_t work = A.two_shot ? (A.n4 + c->world - 1) / c->world : A.n4;
  size_t blocks = (work + 512 * 4 - 1) / (512 * 4);
  size_t cap = max_ctas > 0 ? (size_t)max_ctas : (size_t)c->sm_count;
  if (cap > (size_t)c->sm_count) cap = c->sm_count;   // grid barrier needs co-residency
  if (blocks < 1) blocks = 1;
  if (blocks > cap) blocks = cap;
  void* params[] = {&L, &A};
  if (blocks > 1)
    PDM_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)reduce_optim_kernel, dim3((unsigned)blocks), dim3(512), params, 0,
                                               (cudaStream_t)stream));
  else
    PDM_CHECK_CUDA(cudaLaunchKernel((const void*)reduce_optim_kernel, dim3(1), dim3(512), params, 0, (cudaStream_t)stream));
  return PDM_OK;
}

int pdm_comm_destroy(pdm_comm* c) {
  if (!c) return PDM_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (int p = 0; p < c->world; ++p)
    if (c->opened[p]) cudaIpcCloseMemHandle(c->base[p]);
  if (c->local) cudaFree(c->local);
  delete c;
  cudaGetLastError();
  return PDM_OK;
}

}  // extern "C"
