// nvls.cu -- in-switch (NVLS / multimem) all-reduce fused with the optimizer.  EXPERIMENTAL.
//
// STATUS: written and compiled for sm_100a in round 1 but NOT yet run on hardware (the round's
// GPU budget was spent before it existed).  Nothing uses it by default: it is reached only through
// parallel/nvls.py (an explicit NvlsRegion: benchmarks/allreduce_sweep.py --nvls), and the P2P two-shot kernel in
// comm.cu remains the large-bucket path.  Treat every statement below as the design, not a result.
//
// What it replaces: for large buckets the reference's ncclAllReduce uses NVLS when the fabric has
// it.  The P2P two-shot kernel moves (W-1)/W of the payload twice per GPU (peer
// reads for the reduce-scatter, peer writes for the all-gather).  With a multicast object the
// switch does both: `multimem.ld_reduce` returns the sum over all replicas of an address with one
// load, `multimem.st` writes all replicas with one store, so each GPU moves ~1/W of the payload
// in each direction.
//
// Memory: multicast needs VMM allocations (cuMemCreate), not cudaMalloc, so this is its own
// symmetric region: every rank creates `bytes` of physical memory with a POSIX-fd shareable
// handle type, rank 0 creates the multicast object and exports it as a file descriptor (sent to
// the other ranks with SCM_RIGHTS by parallel/nvls.py), every rank adds its device, binds its
// memory, and maps two views: `uc` (its own replica) and `mc` (the multicast address).
//
// Cross-GPU synchronisation uses the multicast too: `multimem.red.add` on a flag word bumps that
// word on EVERY replica, so "everyone arrived" is "my own copy of the flag reached epoch*W".
// Block b of every rank pairs with block b of the other ranks (same grid everywhere, fixed for the
// life of the region), which needs no grid-wide barrier inside a rank.
#include "common.cuh"

#include <cuda.h>
#include <math.h>

#include <new>

namespace {

constexpr int kBlocks = 64;             // fixed grid: flag b belongs to block b on every rank
constexpr int kThreadsNv = 512;
constexpr size_t kFlagBytes = 4096;     // flags live at the start of the region, data after

template <typename F>
F drv_sym(const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
  return reinterpret_cast<F>(sym);
}
#define DRV(fn) static const auto p_##fn = drv_sym<decltype(&fn)>(#fn)
#define CU_CALL(fn, ...)                                                                  \
  do {                                                                                    \
    DRV(fn);                                                                              \
    PDM_REQUIRE(p_##fn != nullptr, PDM_ERR_CUDA, #fn " is not exported by this driver");   \
    CUresult _r = p_##fn(__VA_ARGS__);                                                    \
    PDM_REQUIRE(_r == CUDA_SUCCESS, PDM_ERR_CUDA, "%s:%d: " #fn " -> CUresult %d", __FILE__, __LINE__, (int)_r); \
  } while (0)

struct OptN {
  int opt, first;
  float lr, beta1, beta2, eps, wd, step_size, bc2_sqrt, gscale;
};

__device__ __forceinline__ void upd_n(const OptN& k, float g, float& p, float& m, float& v) {
  g *= k.gscale;
  if (k.opt == PDM_OPT_NONE) { p = g; return; }
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

struct NvArgs {
  uint8_t* uc;            // this rank's replica
  uint8_t* mc;            // multicast view of the same offsets
  size_t grad_off, out_off, m_off, v_off;   // byte offsets from the start of the data area
  size_t n4;
  int rank, world;
  uint32_t epoch;         // call counter, identical on every rank
  uint64_t timeout_ns;
  int32_t* status;        // local device word: nonzero after a timeout
  OptN k;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-level barrier across ranks: block b of every rank adds 1 to flag b of every replica.
__device__ __forceinline__ void mc_block_barrier(const NvArgs& A, uint32_t target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* mc_flag = reinterpret_cast<uint32_t*>(A.mc) + blockIdx.x;
    const uint32_t* my_flag = reinterpret_cast<const uint32_t*>(A.uc) + blockIdx.x;
    __threadfence_system();
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_flag), "r"(1u) : "memory");
    uint64_t t0 = 0;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(my_flag) - target) < 0) {
      if ((++spins & 1023u) == 0u && A.timeout_ns != 0) {
        const uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > A.timeout_ns) {
          if (A.status != nullptr) atomicExch(A.status, 1);
          break;
        }
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreadsNv) nvls_reduce_optim_kernel(const __grid_constant__ NvArgs A) {
  const uint32_t W = (uint32_t)A.world;
  // 1) every rank's gradients are final (stream order on each rank) once its block b is here
  mc_block_barrier(A, (2u * A.epoch + 1u) * W);

  const size_t per = (A.n4 + W - 1) / W;                  // this rank reduces and updates slice `rank`
  const size_t lo = min(A.n4, per * (size_t)A.rank), hi = min(A.n4, lo + per);
  uint8_t* data_uc = A.uc + kFlagBytes;
  uint8_t* data_mc = A.mc + kFlagBytes;
  const bool has_state = A.k.opt != PDM_OPT_NONE;
  // four independent in-switch reductions in flight per thread (one ld_reduce is a full NVLink round trip)
  constexpr int kUnroll = 4;
  const size_t chunk = (size_t)blockDim.x * kUnroll;
  for (size_t base = lo + (size_t)blockIdx.x * chunk + threadIdx.x; base < hi; base += (size_t)gridDim.x * chunk) {
    float4 g[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < hi) {
        const float4* src = reinterpret_cast<const float4*>(data_mc + A.grad_off) + i;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(g[u].x), "=f"(g[u].y), "=f"(g[u].z), "=f"(g[u].w) : "l"(src) : "memory");
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i >= hi) continue;
      float4 p4 = make_float4(0.f, 0.f, 0.f, 0.f), m4 = p4, v4 = p4;
      if (has_state) {
        p4 = reinterpret_cast<const float4*>(data_uc + A.out_off)[i];
        m4 = reinterpret_cast<const float4*>(data_uc + A.m_off)[i];
        if (A.k.opt == PDM_OPT_ADAM) v4 = reinterpret_cast<const float4*>(data_uc + A.v_off)[i];
      }
      upd_n(A.k, g[u].x, p4.x, m4.x, v4.x);
      upd_n(A.k, g[u].y, p4.y, m4.y, v4.y);
      upd_n(A.k, g[u].z, p4.z, m4.z, v4.z);
      upd_n(A.k, g[u].w, p4.w, m4.w, v4.w);
      if (has_state) {
        reinterpret_cast<float4*>(data_uc + A.m_off)[i] = m4;
        if (A.k.opt == PDM_OPT_ADAM) reinterpret_cast<float4*>(data_uc + A.v_off)[i] = v4;
      }
      float4* dst = reinterpret_cast<float4*>(data_mc + A.out_off) + i;   // one store, every replica
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                   ::"l"(dst), "f"(p4.x), "f"(p4.y), "f"(p4.z), "f"(p4.w) : "memory");
    }
  }
  // 2) nobody leaves before every rank's block b has issued its stores (release/acquire at system scope)
  mc_block_barrier(A, (2u * A.epoch + 2u) * W);
}

}  // namespace

struct pdm_nvls {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;                       // whole region (flags + data), granularity-aligned
  CUmemGenericAllocationHandle mem = 0, mc = 0;
  CUdeviceptr uc_ptr = 0, mc_ptr = 0;
  bool have_mem = false, have_mc = false, uc_mapped = false, mc_mapped = false, bound = false;
  uint32_t epoch = 0;
  uint64_t timeout_ns = 30ull * 1000000000ull;
  int32_t* status = nullptr;
};

extern "C" {

int pdm_nvls_destroy(pdm_nvls* nv);

// 1 when the device can take part in a multicast team, 0 when not, <0 on error.
int pdm_nvls_supported(int device) {
  DRV(cuDeviceGet);
  DRV(cuDeviceGetAttribute);
  if (!p_cuDeviceGet || !p_cuDeviceGetAttribute) return 0;
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) return -1;
  CUdevice dev;
  int v = 0;
  if (p_cuDeviceGet(&dev, device) != CUDA_SUCCESS) return -1;
  if (p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return 0;
  return v ? 1 : 0;
}

// Step 1 (every rank): local physical memory + its unicast mapping; rank 0 also creates the
// multicast object.  data_bytes excludes the flag page.
int pdm_nvls_create(pdm_nvls** out, int rank, int world, int device, size_t data_bytes) {
  PDM_REQUIRE(out != nullptr && world >= 2 && world <= PDM_MAX_RANKS && rank >= 0 && rank < world, PDM_ERR_ARG,
              "pdm_nvls_create: bad arguments (world 2..%d)", PDM_MAX_RANKS);
  PDM_REQUIRE(pdm_nvls_supported(device) == 1, PDM_ERR_STATE, "pdm_nvls_create: device %d does not support multicast objects", device);
  PDM_CHECK_CUDA(cudaSetDevice(device));
  pdm_nvls* nv = new (std::nothrow) pdm_nvls();
  PDM_REQUIRE(nv != nullptr, PDM_ERR_STATE, "pdm_nvls_create: out of host memory");
  nv->rank = rank; nv->world = world; nv->device = device;
  *out = nv;                                  // the caller destroys it on any later error
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;   // multicast binds shareable memory only
  CUmulticastObjectProp mprop;
  memset(&mprop, 0, sizeof(mprop));
  mprop.numDevices = (unsigned)world;
  mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mprop.size = kFlagBytes + data_bytes;
  size_t g_mc = 0, g_mem = 0;
  CU_CALL(cuMulticastGetGranularity, &g_mc, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  CU_CALL(cuMemGetAllocationGranularity, &g_mem, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  const size_t gran = g_mc > g_mem ? g_mc : g_mem;      // both are powers of two
  nv->bytes = (kFlagBytes + data_bytes + gran - 1) / gran * gran;
  mprop.size = nv->bytes;
  CU_CALL(cuMemCreate, &nv->mem, nv->bytes, &prop, 0);
  nv->have_mem = true;
  CU_CALL(cuMemAddressReserve, &nv->uc_ptr, nv->bytes, gran, 0, 0);
  CU_CALL(cuMemMap, nv->uc_ptr, nv->bytes, 0, nv->mem, 0);
  nv->uc_mapped = true;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_CALL(cuMemSetAccess, nv->uc_ptr, nv->bytes, &acc, 1);
  PDM_CHECK_CUDA(cudaMemset(reinterpret_cast<void*>(nv->uc_ptr), 0, nv->bytes));
  PDM_CHECK_CUDA(cudaMalloc(&nv->status, sizeof(int32_t)));
  PDM_CHECK_CUDA(cudaMemset(nv->status, 0, sizeof(int32_t)));
  if (rank == 0) {
    CU_CALL(cuMulticastCreate, &nv->mc, &mprop);
    nv->have_mc = true;
  }
  return PDM_OK;
}

// Step 2a (rank 0): file descriptor of the multicast object, to be passed with SCM_RIGHTS.
int pdm_nvls_export_fd(pdm_nvls* nv, int* fd) {
  PDM_REQUIRE(nv && fd && nv->have_mc && nv->rank == 0, PDM_ERR_STATE, "pdm_nvls_export_fd: only rank 0 owns the multicast object");
  int out = -1;
  CU_CALL(cuMemExportToShareableHandle, &out, nv->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  *fd = out;
  return PDM_OK;
}

// Step 2b (other ranks): adopt the multicast object behind a received descriptor.
int pdm_nvls_import_fd(pdm_nvls* nv, int fd) {
  PDM_REQUIRE(nv && !nv->have_mc && fd >= 0, PDM_ERR_STATE, "pdm_nvls_import_fd: bad state");
  CU_CALL(cuMemImportFromShareableHandle, &nv->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
          CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  nv->have_mc = true;
  return PDM_OK;
}

// Step 3 (every rank): join the team.  Binding blocks until all `world` devices have been added.
int pdm_nvls_add_device(pdm_nvls* nv) {
  PDM_REQUIRE(nv && nv->have_mc, PDM_ERR_STATE, "pdm_nvls_add_device: no multicast object yet");
  PDM_CHECK_CUDA(cudaSetDevice(nv->device));
  CUdevice dev;
  CU_CALL(cuDeviceGet, &dev, nv->device);
  CU_CALL(cuMulticastAddDevice, nv->mc, dev);
  return PDM_OK;
}

// Step 4 (every rank, after everyone's step 3): bind the local memory and map the multicast view.
int pdm_nvls_bind(pdm_nvls* nv) {
  PDM_REQUIRE(nv && nv->have_mc && nv->have_mem && !nv->bound, PDM_ERR_STATE, "pdm_nvls_bind: bad state");
  PDM_CHECK_CUDA(cudaSetDevice(nv->device));
  CU_CALL(cuMulticastBindMem, nv->mc, 0, nv->mem, 0, nv->bytes, 0);
  nv->bound = true;
  CU_CALL(cuMemAddressReserve, &nv->mc_ptr, nv->bytes, 0, 0, 0);
  CU_CALL(cuMemMap, nv->mc_ptr, nv->bytes, 0, nv->mc, 0);
  nv->mc_mapped = true;
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = nv->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_CALL(cuMemSetAccess, nv->mc_ptr, nv->bytes, &acc, 1);
  return PDM_OK;
}

void* pdm_nvls_data(pdm_nvls* nv) { return nv ? reinterpret_cast<void*>(nv->uc_ptr + kFlagBytes) : nullptr; }
size_t pdm_nvls_data_bytes(pdm_nvls* nv) { return nv ? nv->bytes - kFlagBytes : 0; }
void pdm_nvls_set_timeout(pdm_nvls* nv, double seconds) {
  if (nv) nv->timeout_ns = seconds > 0 ? (uint64_t)(seconds * 1e9) : 0;
}
int pdm_nvls_status(pdm_nvls* nv, int* out) {
  PDM_REQUIRE(nv && out && nv->status, PDM_ERR_ARG, "pdm_nvls_status: bad arguments");
  PDM_CHECK_CUDA(cudaMemcpy(out, nv->status, sizeof(int32_t), cudaMemcpyDeviceToHost));
  return PDM_OK;
}

// Same contract as pdm_comm_allreduce_optim (two-shot form): offsets are bytes into the data area;
// hyper->opt == PDM_OPT_NONE makes it a plain all-reduce whose result lands at out_off on every rank.
int pdm_nvls_allreduce_optim(pdm_nvls* nv, size_t grad_off, size_t out_off, size_t m_off, size_t v_off, size_t n_floats,
                             const pdm_hyper* h, float grad_scale, pdm_stream_t stream) {
  PDM_REQUIRE(nv && nv->mc_mapped, PDM_ERR_STATE, "pdm_nvls_allreduce_optim: region is not bound");
  PDM_REQUIRE(h != nullptr, PDM_ERR_ARG, "pdm_nvls_allreduce_optim: hyper is NULL");
  PDM_REQUIRE(n_floats % 4 == 0 && (grad_off | out_off | m_off | v_off) % 16 == 0, PDM_ERR_ARG,
              "pdm_nvls_allreduce_optim: ranges must be 16-byte aligned multiples of 4 floats");
  const size_t bytes = n_floats * 4, cap = nv->bytes - kFlagBytes;
  PDM_REQUIRE(grad_off + bytes <= cap && out_off + bytes <= cap, PDM_ERR_LIMIT, "pdm_nvls_allreduce_optim: range outside the region");
  if (h->opt != PDM_OPT_NONE)
    PDM_REQUIRE(m_off + bytes <= cap && v_off + bytes <= cap, PDM_ERR_LIMIT, "pdm_nvls_allreduce_optim: state range outside the region");
  NvArgs A;
  memset(&A, 0, sizeof(A));
  A.uc = reinterpret_cast<uint8_t*>(nv->uc_ptr);
  A.mc = reinterpret_cast<uint8_t*>(nv->mc_ptr);
  A.grad_off = grad_off; A.out_off = out_off; A.m_off = m_off; A.v_off = v_off;
  A.n4 = n_floats / 4;
  A.rank = nv->rank; A.world = nv->world;
  A.epoch = nv->epoch++;
  A.timeout_ns = nv->timeout_ns;
  A.status = nv->status;
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
  // fixed grid (flag b <-> block b on every rank); 64 CTAs x 512 threads all fit on the device at once
  nvls_reduce_optim_kernel<<<kBlocks, kThreadsNv, 0, (cudaStream_t)stream>>>(A);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_nvls_destroy(pdm_nvls* nv) {
  if (!nv) return PDM_OK;
  cudaSetDevice(nv->device);
  cudaDeviceSynchronize();
  DRV(cuMemUnmap);
  DRV(cuMemAddressFree);
  DRV(cuMemRelease);
  DRV(cuMulticastUnbind);
  DRV(cuDeviceGet);
  if (p_cuMemUnmap && p_cuMemAddressFree && p_cuMemRelease) {
    if (nv->mc_mapped) { p_cuMemUnmap(nv->mc_ptr, nv->bytes); }
    if (nv->mc_ptr) p_cuMemAddressFree(nv->mc_ptr, nv->bytes);
    if (nv->bound && p_cuMulticastUnbind && p_cuDeviceGet) {
      CUdevice dev;
      if (p_cuDeviceGet(&dev, nv->device) == CUDA_SUCCESS) p_cuMulticastUnbind(nv->mc, dev, 0, nv->bytes);
    }
    if (nv->uc_mapped) p_cuMemUnmap(nv->uc_ptr, nv->bytes);
    if (nv->uc_ptr) p_cuMemAddressFree(nv->uc_ptr, nv->bytes);
    if (nv->have_mc) p_cuMemRelease(nv->mc);
    if (nv->have_mem) p_cuMemRelease(nv->mem);
  }
  if (nv->status) cudaFree(nv->status);
  delete nv;
  cudaGetLastError();
  return PDM_OK;
}

}  // extern "C"
