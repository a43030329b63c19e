// common.cuh -- shared helpers for the sm_100a native core.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>

#include "pdmnist.h"

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
void pdm_set_error(const char* fmt, ...);
unsigned pdm_skew_ns(void);   /* PDM_SKEW_NS, read once */

#define PDM_CHECK_CUDA(expr)                                                          \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) {                                                          \
      pdm_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return PDM_ERR_CUDA;                                                            \
    }                                                                                 \
  } while (0)

#define PDM_REQUIRE(cond, code, ...)   \
  do {                                 \
    if (!(cond)) {                     \
      pdm_set_error(__VA_ARGS__);      \
      return (code);                   \
    }                                  \
  } while (0)

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Grid-wide barrier for a cooperatively launched grid.  sync[0] = arrival
// count, sync[1] = generation; both zero before first use and left
// consistent (count back to 0) after every barrier, so the same two words
// serve every launch.
__device__ __forceinline__ void grid_barrier(uint32_t* sync, uint32_t nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t gen = ld_acquire_gpu(sync + 1);
    __threadfence();
    uint32_t arrived = atomicAdd(sync, 1u);
    if (arrived == nblocks - 1) {
      sync[0] = 0;
      __threadfence();
      st_release_gpu(sync + 1, gen + 1);
    } else {
      while (ld_acquire_gpu(sync + 1) == gen) { }
    }
    __threadfence();
  }
  __syncthreads();
}

// LL ("low latency") peer protocol unit: 8 bytes = {payload bits, flag}.  An
// aligned 8-byte store is never torn on NVLink, so a reader that sees the
// expected flag also sees that unit's payload; no fence is needed.
__device__ __forceinline__ void ll_store2(uint64_t* dst, float a, float b, uint32_t flag) {
  uint64_t u0 = (uint64_t)__float_as_uint(a) | ((uint64_t)flag << 32);
  uint64_t u1 = (uint64_t)__float_as_uint(b) | ((uint64_t)flag << 32);
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(u0), "l"(u1) : "memory");
}
__device__ __forceinline__ bool ll_load2(const uint64_t* src, uint32_t flag, float& a, float& b) {
  uint64_t u0, u1;
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(u0), "=l"(u1) : "l"(src) : "memory");
  a = __uint_as_float((uint32_t)u0);
  b = __uint_as_float((uint32_t)u1);
  return ((uint32_t)(u0 >> 32) == flag) && ((uint32_t)(u1 >> 32) == flag);
}

// Fault-injection aid for the cross-GPU flag protocols (PDM_SKEW_NS=<max ns>, default off): a pseudo-random
// pause that differs per rank, CTA, thread and collective, placed right before a rank publishes data or
// flags, so that soak runs (benchmarks/soak.py) exercise orderings an idle machine never produces.
__device__ __forceinline__ void pdm_skew(uint32_t max_ns, uint32_t a, uint32_t b, uint32_t c) {
  if (max_ns == 0) return;
  uint32_t h = a * 2654435761u ^ (b + 0x9e3779b9u) * 40503u ^ (c * 2246822519u);
  h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
  __nanosleep(h % max_ns);
}

#endif  // __CUDACC__
