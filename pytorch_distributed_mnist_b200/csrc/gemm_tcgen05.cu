// gemm_tcgen05.cu -- TMA-fed, warp-specialised tcgen05 (UMMA) GEMM at tf32 with fp32
// accumulation in TMEM, for the GEMM-shaped layers of the CNN extension (SURVEY 7.2 step 6,
// 7.4): conv2 forward / dgrad / wgrad as implicit ("shifted-row") GEMMs and fc1
// forward / dgrad / wgrad.  The reference reaches these shapes through cuDNN / cuBLAS
// (SURVEY N8); nothing here calls a library.
//
//   D[M,N] (+)= A[M,K] . B[N,K]^T        tf32 x tf32 -> fp32
//
// Persistent: one CTA per SM walks 128 x BLOCK_N output tiles (optionally K splits); the fp32
// accumulator is double-buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1.
// 192 threads:
//   warp 0      TMA producer   cp.async.bulk.tensor.2d -> 128B-swizzled smem ring (4 stages),
//                              completion by mbarrier complete_tx
//   warp 1      MMA issuer     one lane issues tcgen05.mma.cta_group::1.kind::tf32 (M=128,
//                              N=BLOCK_N, K=8 per instruction, 4 per stage); tcgen05.commit
//                              releases the smem stage / publishes the accumulator
//   warps 2-5   epilogue       tcgen05.ld 32 lanes x 16 columns at a time out of TMEM,
//                              fused bias / ReLU / ReLU-mask, store or red.add (split-K)
//
// Operand layouts.  Each operand is a row-major fp32 matrix in global memory, used either
//   K-major  : stored [MN][K]  -> one TMA box  {32 k (=128 B), rows}  per stage
//   MN-major : stored [K][MN]  -> MN/32 boxes  {32 mn (=128 B), 32 k} per stage ("atoms")
// K-major tiles use CU_TENSOR_MAP_SWIZZLE_128B and the UMMA SWIZZLE_128B canonical layout
// (8 rows x 128 B atoms, SBO = 1024 B).  MN-major tf32 has exactly one legal layout on
// sm_100: 128-byte swizzle with 32-byte atomicity -- CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B on
// the TMA side, UMMA layout type SWIZZLE_128B_BASE32B (4 k-rows x 128 B atoms, SBO = 512 B,
// LBO = 4096 B = stride between 32-wide MN atoms).  Either way the operand is consumed in the
// layout it already has: no transposed copy of an activation or a weight is ever made.
//
// Implicit-GEMM convolution (3x3, stride 1, no padding) uses NHWC activations viewed as a 2-D
// matrix [B*H*W, C] and the identity  in_pixel(q, tap) = q + kh*W + kw : the A tile of filter
// tap (kh,kw) is the SAME matrix shifted by kh*W+kw rows, i.e. an ordinary TMA box at a row
// offset (out-of-range rows zero-fill).  Output positions whose window would wrap (w >= W-2 or
// h >= H-2) are computed and discarded by the consumer; that costs (H*W)/((H-2)*(W-2)) - 1
// extra MMA work and removes every gather.
#include "common.cuh"

#include <cuda.h>
#include <stdlib.h>

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 32;            // tf32 elements = 128 bytes = one swizzle atom
constexpr int kAtom = 32;              // elements per 128-byte atom along MN (MN-major)
constexpr int kStages = 4;
constexpr int kThreads = 192;
constexpr int kUmmaK = 8;              // tf32 MMA K
constexpr uint32_t kABytes = kBlockM * 128;     // 16 KB per stage
constexpr uint32_t kAtomBytes = kBlockK * 128;  // 4 KB: one MN-major atom (32 k-rows x 128 B)

enum { MODE_GEMM = 0, MODE_CONV = 1, MODE_WGRAD = 2 };

struct Params {
  int M, N;
  int block_n;
  int a_mn, b_mn;            // 1 = MN-major operand
  int kblocks_per_split;     // k-blocks per (tile, split)
  int kblocks_total;
  int tiles_m, tiles_n, splits;
  int mode, img_w, cblocks, shift_sign;
  float* D;
  long long ldd;
  const float* bias;
  const float* mask;
  long long ld_mask;
  int relu, atomic;
  uint32_t idesc;
  uint32_t tmem_cols;
  uint32_t mn_layout, mn_sbo, mn_lbo;   // UMMA descriptor fields for MN-major operands
};

// ---------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// UMMA shared-memory descriptor, SWIZZLE_128B, sm_100 version field = 1.
//   bits [0,14) start address >> 4, [16,30) LBO >> 4, [32,46) SBO >> 4, [46,48) version, [61,64) layout type (2 = SW128)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout & 7u) << 61;
  return d;
}

__device__ __forceinline__ int tap_shift(int tap, int img_w) { return (tap / 3) * img_w + (tap % 3); }

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Persistent kernel: CTA c walks tiles c, c + gridDim.x, ... (n fastest, then m, then K split).
// The accumulator is double-buffered in TMEM (2 x BLOCK_N columns): while the epilogue warps
// drain tile i, the MMA warp already accumulates tile i+1, and the TMA ring never stops.
__global__ void __launch_bounds__(kThreads, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B atoms need 1024-byte alignment of every stage buffer
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t b_bytes = (uint32_t)p.block_n * 128u;
  const uint32_t stage_bytes = kABytes + b_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * stage_bytes);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;       // [2]
  uint64_t* tmem_empty = tmem_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_mn = p.tiles_m * p.tiles_n;
  const int tiles_total = tiles_mn * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tmem_full + s, 1); mbar_init(tmem_empty + s, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      const uint32_t tx = kABytes + b_bytes;
      uint32_t it = 0;                                  // k-blocks issued so far (ring position)
      for (int t = blockIdx.x; t < tiles_total; t += gridDim.x) {
        const int z = t / tiles_mn, rem = t % tiles_mn;
        const int m0 = (rem / p.tiles_n) * kBlockM, n0 = (rem % p.tiles_n) * p.block_n;
        const int kb_begin = z * p.kblocks_per_split;
        const int nkb = min(p.kblocks_per_split, p.kblocks_total - kb_begin);
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          mbar_wait(empty + s, ph ^ 1u);
          uint8_t* sa = smem + s * stage_bytes;
          uint8_t* sb = sa + kABytes;
          mbar_expect_tx(full + s, tx);
          const int kb = kb_begin + i;
          // ---- A ----
          if (p.mode == MODE_CONV) {
            const int tap = kb / p.cblocks, cb = kb % p.cblocks;
            tma_load_2d(sa, &map_a, full + s, cb * kBlockK, m0 + p.shift_sign * tap_shift(tap, p.img_w));
          } else if (!p.a_mn) {
            tma_load_2d(sa, &map_a, full + s, kb * kBlockK, m0);
          } else {
            for (int a = 0; a < kBlockM / kAtom; ++a)
              tma_load_2d(sa + a * kAtomBytes, &map_a, full + s, m0 + a * kAtom, kb * kBlockK);
          }
          // ---- B ----
          if (p.mode == MODE_WGRAD) {
            for (int j = 0; j < p.block_n / kAtom; ++j) {
              const int n = n0 + j * kAtom;
              const int tap = n / (p.cblocks * kAtom), cb = (n / kAtom) % p.cblocks;
              tma_load_2d(sb + j * kAtomBytes, &map_b, full + s, cb * kAtom, kb * kBlockK + tap_shift(tap, p.img_w));
            }
          } else if (!p.b_mn) {
            tma_load_2d(sb, &map_b, full + s, kb * kBlockK, n0);
          } else {
            for (int j = 0; j < p.block_n / kAtom; ++j)
              tma_load_2d(sb + j * kAtomBytes, &map_b, full + s, n0 + j * kAtom, kb * kBlockK);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer =================================
    if (lane == 0) {
      uint32_t it = 0, tl = 0;
      for (int t = blockIdx.x; t < tiles_total; t += gridDim.x, ++tl) {
        const int z = t / tiles_mn;
        const int kb_begin = z * p.kblocks_per_split;
        const int nkb = min(p.kblocks_per_split, p.kblocks_total - kb_begin);
        const uint32_t as = tl & 1u, aph = (tl >> 1) & 1u;
        mbar_wait(tmem_empty + as, aph ^ 1u);           // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * (uint32_t)p.block_n;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          mbar_wait(full + s, ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * stage_bytes);
          const uint32_t sb = sa + kABytes;
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // K-major: next 8 k = +32 bytes inside the swizzle atom.  MN-major: next 8 k-rows = +1024 bytes.
            const uint64_t da = p.a_mn ? make_desc(sa + k * 1024, p.mn_lbo, p.mn_sbo, p.mn_layout) : make_desc(sa + k * 32, 0, 1024);
            const uint64_t db = p.b_mn ? make_desc(sb + k * 1024, p.mn_lbo, p.mn_sbo, p.mn_layout) : make_desc(sb + k * 32, 0, 1024);
            umma_tf32(tmem_d, da, db, p.idesc, (i | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty + s);        // stage reusable once these MMAs have read it
        }
        umma_commit(tmem_full + as);     // accumulator complete
      }
    }
  } else {
    // =============================== epilogue ====================================
    const int quad = warp & 3;                           // TMEM lane quadrant this warp may read
    uint32_t tl = 0;
    for (int t = blockIdx.x; t < tiles_total; t += gridDim.x, ++tl) {
      const int rem = t % tiles_mn;
      const int m0 = (rem / p.tiles_n) * kBlockM, n0 = (rem % p.tiles_n) * p.block_n;
      const uint32_t as = tl & 1u, aph = (tl >> 1) & 1u;
      mbar_wait(tmem_full + as, aph);
      tc_fence_after();
      const int row = m0 + quad * 32 + lane;
      const bool row_ok = row < p.M;
      float* drow = p.D + (long long)row * p.ldd;
      const float* mrow = p.mask ? p.mask + (long long)row * p.ld_mask : nullptr;
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + as * (uint32_t)p.block_n;
      for (int c = 0; c < p.block_n; c += 16) {
        float v[16];
        tmem_ld16(taddr + (uint32_t)c, v);
        if (!row_ok) continue;
        const int col0 = n0 + c;
        if (col0 >= p.N) continue;
        if (!p.atomic && !p.bias && !mrow && !p.relu && col0 + 16 <= p.N &&
            (reinterpret_cast<uintptr_t>(drow + col0) & 15) == 0) {
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            *reinterpret_cast<float4*>(drow + col0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          continue;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int col = col0 + i;
          if (col >= p.N) break;
          float x = v[i];
          if (p.bias) x += __ldg(p.bias + col);
          if (p.relu) x = fmaxf(x, 0.f);
          if (mrow) x = __ldg(mrow + col) > 0.f ? x : 0.f;
          if (p.atomic) atomicAdd(drow + col, x);
          else drow[col] = x;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty + as);       // 4 epilogue warps release the accumulator
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---------------------------------------------------------------------------- host
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode() {
  static EncodeFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || !sym) return nullptr;
  fn = (EncodeFn)sym;
  return fn;
}

// 2-D fp32 row-major matrix [rows][cols] with leading dimension ld (elements); box = {32 cols, box_rows}.
int make_map(CUtensorMap* map, const float* base, long long rows, long long cols, long long ld, int box_rows, int mn_major) {
  EncodeFn enc = get_encode();
  PDM_REQUIRE(enc != nullptr, PDM_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  PDM_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 4) % 16 == 0, PDM_ERR_ARG,
              "tcgen05 GEMM operands need a 16-byte aligned base and row pitch (ld=%lld)", ld);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  // MN-major tf32: 128 B swizzle with 32 B atomicity (the only layout UMMA accepts for it)
  const CUtensorMapSwizzle swz = mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PDM_REQUIRE(r == CUDA_SUCCESS, PDM_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld box=%d)",
              (int)r, rows, cols, ld, box_rows);
  return PDM_OK;
}

}  // namespace

extern "C" int pdm_gemm_tf32(const pdm_gemm_args* g, pdm_stream_t stream) {
  PDM_REQUIRE(g != nullptr, PDM_ERR_ARG, "pdm_gemm_tf32: args is NULL");
  PDM_REQUIRE(g->a && g->b && g->d, PDM_ERR_ARG, "pdm_gemm_tf32: a, b and d are required");
  PDM_REQUIRE(g->M >= 1 && g->N >= 1 && g->K >= 1, PDM_ERR_ARG, "pdm_gemm_tf32: bad shape M=%d N=%d K=%d", g->M, g->N, g->K);
  PDM_REQUIRE(g->mode >= MODE_GEMM && g->mode <= MODE_WGRAD, PDM_ERR_ARG, "pdm_gemm_tf32: unknown mode %d", g->mode);
  int block_n = g->block_n > 0 ? g->block_n : (g->N >= 256 ? 256 : (g->N + 31) / 32 * 32);
  PDM_REQUIRE(block_n % 32 == 0 && block_n >= 32 && block_n <= 256, PDM_ERR_LIMIT,
              "pdm_gemm_tf32: block_n %d must be a multiple of 32 in [32, 256]", block_n);
  PDM_REQUIRE(g->K % kBlockK == 0 || g->mode != MODE_CONV, PDM_ERR_ARG, "pdm_gemm_tf32: conv mode needs K %% 32 == 0");

  Params p;
  memset(&p, 0, sizeof(p));
  p.M = g->M;
  p.N = g->N;
  p.block_n = block_n;
  p.a_mn = g->a_mn_major ? 1 : 0;
  p.b_mn = g->b_mn_major ? 1 : 0;
  p.mode = g->mode;
  p.img_w = g->img_w;
  p.cblocks = g->cblocks > 0 ? g->cblocks : 1;
  p.shift_sign = g->shift_sign >= 0 ? 1 : -1;
  if (p.mode == MODE_CONV) PDM_REQUIRE(!p.a_mn && !p.b_mn, PDM_ERR_ARG, "pdm_gemm_tf32: conv mode needs K-major operands");
  if (p.mode == MODE_WGRAD) PDM_REQUIRE(p.a_mn && p.b_mn, PDM_ERR_ARG, "pdm_gemm_tf32: wgrad mode needs MN-major operands");
  p.kblocks_total = (g->K + kBlockK - 1) / kBlockK;
  int splits = g->splits > 0 ? g->splits : 1;
  if (splits > p.kblocks_total) splits = p.kblocks_total;
  p.kblocks_per_split = (p.kblocks_total + splits - 1) / splits;
  splits = (p.kblocks_total + p.kblocks_per_split - 1) / p.kblocks_per_split;
  p.D = g->d;
  p.ldd = g->ldd;
  p.bias = g->bias;
  p.mask = g->mask;
  p.ld_mask = g->ld_mask;
  p.relu = g->relu;
  p.atomic = (g->accumulate || splits > 1) ? 1 : 0;
  PDM_REQUIRE(!(splits > 1 && (g->bias || g->relu || g->mask)), PDM_ERR_ARG,
              "pdm_gemm_tf32: split-K cannot be combined with a bias/ReLU/mask epilogue");
  // instruction descriptor: D=f32 (bits 4-5 = 1), A,B = tf32 (bits 7-9, 10-12 = 2), majors (15, 16), N>>3 (17-22), M>>4 (24-28)
  p.idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
            ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(kBlockM >> 4) << 24);
  const int acc_cols = 2 * block_n;        // double-buffered accumulator
  p.tmem_cols = acc_cols <= 32 ? 32 : (acc_cols <= 64 ? 64 : (acc_cols <= 128 ? 128 : (acc_cols <= 256 ? 256 : 512)));
  p.tiles_m = (g->M + kBlockM - 1) / kBlockM;
  p.tiles_n = (g->N + block_n - 1) / block_n;
  p.splits = splits;
  // MN-major descriptor fields: found by experiment on B200, the only combination UMMA accepts for tf32
  p.mn_layout = 1u;                    // SWIZZLE_128B_BASE32B
  p.mn_sbo = 512u;                     // 4 k-rows x 128 B
  p.mn_lbo = (uint32_t)kAtomBytes;

  CUtensorMap ma, mb;
  int rc = make_map(&ma, g->a, g->a_rows, g->a_cols, g->lda, p.a_mn ? kBlockK : kBlockM, p.a_mn);
  if (rc != PDM_OK) return rc;
  rc = make_map(&mb, g->b, g->b_rows, g->b_cols, g->ldb, p.b_mn ? kBlockK : block_n, p.b_mn);
  if (rc != PDM_OK) return rc;

  // at least half of the SM's shared memory so that two CTAs never share an SM (and its TMEM)
  size_t smem = (size_t)kStages * (kABytes + (size_t)block_n * 128) + 256 + 1024;
  if (smem < 120 * 1024) smem = 120 * 1024;
  static bool attr_set = false;
  if (!attr_set) {
    PDM_CHECK_CUDA(cudaFuncSetAttribute(gemm_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  static int sm_count = 0;
  if (sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (sm_count <= 0) sm_count = 148;
  }
  // one persistent CTA per SM (TMEM and the smem ring are sized for one resident CTA)
  const long long tiles_total = (long long)p.tiles_m * p.tiles_n * splits;
  const int grid = (int)(tiles_total < sm_count ? tiles_total : sm_count);
  gemm_tf32_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(ma, mb, p);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}
