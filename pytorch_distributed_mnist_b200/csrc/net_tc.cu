// net_tc.cu -- the reference model's training loop as ONE persistent thread-block-cluster kernel
// on the 5th-generation tensor cores (tcgen05 / TMEM) of sm_100a.
//
// Reference path being replaced (ref:N = /root/reference/multi_proc_single_gpu.py), per step:
//   ref:84-85 H2D copies, ref:87 addmm (K3), ref:88 cross_entropy (K4), ref:91 backward (K5-K7) +
//   DDP reducer / ncclAllReduce (K8-K10), ref:92 Adam (K11), ref:94-95 metrics + 2 host syncs.
// net_step.cu does one step per launch with FMA math, a grid barrier and a pass over per-CTA
// partial gradients in L2.  This file removes the launch, the grid barrier and the partials:
//
//   * ONE launch runs `nsteps` steps.  A cluster of 13 CTAs owns the model: CTA c owns input
//     columns [64c, 64c+64) (13 x 64 = 832 >= 784) -- their weights, Adam moments and gradient
//     stay in registers / shared memory for the whole run and are written back at the end.
//   * forward   P[512 x 48]    = X[512 x 64] . [W_hi; W_mid; W_lo]^T     tcgen05.mma kind::f16, M=128 N=48 K=16
//     wgrad     G[64(+1) x 48] = X^T[64(+1) x 512] . [dl_hi | dl_mid | dl_lo]
//     Both read the SAME shared-memory copy of the x tile: pixels are stored as bf16 (the integers
//     0..255 are exact in bf16) in 8x8 core matrices without swizzle, which is a legal K-major
//     operand (forward) and a legal MN-major operand (wgrad) at the same time.  fp32 operands
//     (weights, dlogits) are split into three bf16 terms (hi + mid + lo = 24 mantissa bits) that
//     sit side by side as three 16-column groups of ONE instruction, so every product is exact,
//     accumulation is fp32 in TMEM, and the x tile is read from shared memory once per GEMM; the
//     epilogues add the three groups.  The input normalisation (x/255 - mean)/std is folded into
//     the epilogues.  A constant "ones" column appended to the tile makes the same wgrad MMA
//     produce the bias gradient.
//   * the K-split of the forward needs a cross-CTA reduction of the partial logits: every CTA owns
//     40 rows of a 512-row block; partials travel as ONE bulk copy (cp.async.bulk shared::cta ->
//     shared::cluster) per owner that completes a transaction mbarrier there -- no cluster barrier,
//     no fence; the owner soft-maxes its rows and bulk-copies its rows of the wgrad operand
//     (dlogits, already split) into every other CTA.
//   * warp-specialised: 7 producer warps stream the NEXT step's rows (device memory, or pinned host
//     memory over PCIe, optionally gathered through an index plan) while the current step computes;
//     1 MMA-issuing warp; 4 epilogue/optimizer warps (TMEM lanes).  Optimizer thread t of the 128
//     owns column t % 64 and classes 5 (t / 64) .. +4: five weights with their moments in registers
//     (threads 0-9 also keep a replica of the bias).
//   * multi-GPU: the gradient of a CTA's columns is exchanged with the same CTA of every peer
//     with an LL protocol (16-byte {value, flag, value, flag} units pushed over NVLink by all 128
//     epilogue threads), summed in rank order (bit-identical replicas), inside the same kernel.
#include "common.cuh"

#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

namespace {

constexpr int kCtas = PDM_TC_CLUSTER;        // 13
constexpr int kCols = 64;                    // input columns per CTA
constexpr int kIn = PDM_NET_IN;              // 784
constexpr int kOut = PDM_NET_OUT;            // 10
constexpr int kHalf = kOut / 2;              // classes per optimizer thread
constexpr int kBlkRows = 512;                // rows per block (4 M-tiles of 128)
constexpr int kOwnRows = 40;                 // rows of a block owned by one CTA (13 * 40 >= 512)
constexpr int kThreads = 384;
// Warp roles, lowest to highest warp id.  The SM's warp scheduler prefers the highest eligible warp id,
// so the latency-critical epilogue/optimizer warps sit on top, then the MMA issuer, then the producers
// (which mostly wait): measured 14.5 -> 9.4 us/step at batch 512 against the roles the other way round.
constexpr int kProdThreads = 224;            // warps 0-6
constexpr int kMmaWarp = 7;
constexpr int kEpiThreads = 128;             // warps 8-11 (warp % 4 = TMEM lane quadrant)
constexpr int kEpiFirst = 256;
constexpr float kScale = 1.0f / (255.0f * 0.3081f);   // d(normalised x)/d(raw pixel)
constexpr float kShift = 0.1307f / 0.3081f;           // mean / std

// x tile: [row group (8 rows)][column block (8 cols)] core matrices of 128 bytes, 9 column blocks
// per row group (8 real + the "ones" block), no swizzle.
constexpr uint32_t kSc = 128;                // stride between column blocks
constexpr uint32_t kColBlks = 9;
constexpr uint32_t kSr = kColBlks * kSc;     // 1152: stride between row groups
constexpr uint32_t kTileBytes = (kBlkRows / 8) * kSr;          // 73,728
constexpr uint32_t kTilePad = 7 * kSc;       // M=128 wgrad reads 16 column blocks from the last row group
constexpr uint32_t kDlSplitBytes = 2 * kBlkRows * 16;          // [n/8][row][n%8] bf16 = 16,384
constexpr uint32_t kWopSplitBytes = 2 * 8 * 128;               // [n/8][k/8][n%8][k%8] bf16 = 2,048

// shared memory map (bytes)
constexpr uint32_t kOffTile = 0;
constexpr uint32_t kOffDlop = kOffTile + kTileBytes + kTilePad;             // wgrad B operand: [split 3][n/8 2][row 512][8] bf16
constexpr uint32_t kOffWop = kOffDlop + 3 * kDlSplitBytes;                  // forward B operand: [split 3][n/8 2][k/8 8][8][8] bf16
constexpr uint32_t kOffXbuf = kOffWop + 3 * kWopSplitBytes;                 // partial logits of my rows from every CTA [13][40][10] f32
constexpr uint32_t kXbufBytes = kCtas * kOwnRows * kOut * 4;                // 20,800
constexpr uint32_t kOffStage = kOffXbuf + kXbufBytes;                       // my partial logits of all rows [512][10] f32 (bulk-copy source)
constexpr uint32_t kStageBytes = kBlkRows * kOut * 4;                       // 20,480
constexpr uint32_t kOffGbuf = kOffStage + kStageBytes;                      // this CTA's gradient as LL units [325] float2 (multi-GPU exchange)
constexpr uint32_t kGbufBytes = 2624;
constexpr uint32_t kOffMisc = kOffGbuf + kGbufBytes;
constexpr uint32_t kMiscBytes = 2048;
constexpr uint32_t kOffTrace = kOffMisc + kMiscBytes;                       // [kTraceSteps][32] clock stamps (PDM_TC_TRACE=1)
constexpr int kTraceSteps = 64;
constexpr uint32_t kSmemBytes = kOffTrace + kTraceSteps * 32 * 8;
constexpr uint32_t kTmemCols = 256;          // forward: 4 M-tiles x 48 columns, wgrad: 48 columns at column 192
constexpr uint32_t kTmemWg = 192;
// feeder ring: one slot = 512 rows of 784 bytes + 512 labels
constexpr int kRingSlots = PDM_TC_RING_SLOTS;
constexpr size_t kRingSlotBytes = PDM_TC_RING_SLOT_BYTES;
static_assert(kRingSlotBytes == (size_t)kBlkRows * kIn + kBlkRows, "ring slot size");

struct Misc {                     // lives at kOffMisc
  uint64_t bar_xfull, bar_fwd, bar_dl, bar_wg, bar_wop, bar_p1, bar_p2, bar_loss;
  uint32_t tmem_slot;
  uint32_t pad0;
  float bias[16];
  float wsum[2][16];              // per warp partial of sum_k W[n][k] over this CTA's columns
  float db[16];
  float lossw[3][2];              // per warp {loss, correct} of this CTA's rows of the current block
  float lossbuf[kCtas][2];        // CTA 0: per-CTA {loss sum, correct}
  uint8_t ys[kBlkRows];           // labels of the current block (written by the producers with the tile)
};
static_assert(sizeof(Misc) <= kMiscBytes, "Misc outgrew its slot");

constexpr int kInlineSteps = 96;       // launches of up to this many steps need no constant upload

struct TcArgs {
  const uint8_t* x;
  const uint8_t* y;
  const int32_t* row_index;
  long long pool_batches;         // pool mode: batches in the pool
  long long first_batch;
  long long count;                // index mode: samples of the run
  int batch;
  long long nsteps;
  float* params;
  float* m;
  float* v;
  float* grad_out;
  float* logits_out;
  double* metrics;
  float* step_out;
  int opt;
  float lr, beta1, beta2, eps, wd;
  long long step0;                // 1-based index of the first update
  uint8_t* ring;                  // optional device ring of kRingSlots blocks (rows + labels): the feeder cluster fills it
  uint32_t* ring_flags;           // [2][kRingSlots]: blocks made ready (x13 per fill) / consumed (x13 per use)
  const float2* adam;             // per step of this launch: {lr / (1 - beta1^t), 1 / sqrt(1 - beta2^t)} (host-computed, double precision)
  float2 adam_inline[kInlineSteps];   // the same for short launches, carried in the kernel parameters (adam == nullptr)
  float inv_world;
  uint32_t skew_ns;               // fault injection: random pause (< skew_ns) before this rank publishes its gradient
  int dbg;                        // bring-up switches (PDM_TC_DBG bit mask); 0 in production
  long long* trace;               // optional [13][kTraceSteps][32] clock64 stamps per CTA (PDM_TC_TRACE=1)
  pdm_peers peers;
};

__device__ __forceinline__ long long clock_now() {
  long long c;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(c) :: "memory");     // the clobber pins the stamp between its neighbours
  return c;
}
// stamps go to shared memory (a global store per stamp perturbs what it measures) and are dumped at the end
#define TC_TRACE(cond, step, slot) do { if (kTrace && (cond) && (step) < kTraceSteps) \
    reinterpret_cast<long long*>(smem + kOffTrace)[(step) * 32 + (slot)] = clock_now(); } while (0)

// ---------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
// The same wait for warps that are far ahead of the critical path (the producers): sleep between polls so
// that their polling does not compete with the epilogue warps for issue slots and the shared-memory pipe.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  while (true) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(128);
  }
}
__device__ __forceinline__ uint32_t mapa(uint32_t local, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(cta));
  return r;
}
// Store into another CTA's shared memory; the store's bytes complete a transaction on that CTA's mbarrier.
__device__ __forceinline__ void st_async2(uint32_t raddr, float a, float b, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
               ::"r"(raddr), "f"(a), "f"(b), "r"(rbar) : "memory");
}
// Bulk copy own shared memory -> another CTA's shared memory; ONE transaction on that CTA's mbarrier.
__device__ __forceinline__ void bulk_s2s(uint32_t rdst, uint32_t src, uint32_t bytes, uint32_t rbar) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(rdst), "r"(src), "r"(bytes), "r"(rbar) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 16 consecutive columns: thread `lane` of warp w gets row 32*(w%4)+lane
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor without swizzle (layout type 0), sm_100 version field = 1.
//   K-major : LBO = stride between 8-element K chunks, SBO = stride between 8-row groups
//   MN-major: SBO = stride between 8-element MN blocks, LBO = stride between 8-deep K groups
// (both verified on hardware against an fp64 evaluation: tests/test_gpu_net_tc.py)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor: D = f32 (bits 4-5), A = B = bf16 (bits 7-9, 10-12), N >> 3 (17-22), M >> 4 (24-28)
constexpr uint32_t kIdescBase = (1u << 4) | (1u << 7) | (1u << 10) | ((48u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t kIdescFwd = kIdescBase;                              // both operands K-major
constexpr uint32_t kIdescWg = kIdescBase | (1u << 15) | (1u << 16);     // both operands MN-major

__device__ __forceinline__ uint32_t bf16_bits(float f) {                // round-to-nearest-even, one cvt instruction
  return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(f));
}
__device__ __forceinline__ float bf16_val(uint32_t b) { return __uint_as_float(b << 16); }
// f = hi + mid + lo with three bf16 terms (24 mantissa bits: exact for every fp32 in range)
__device__ __forceinline__ void split3(float f, uint32_t& hi, uint32_t& mid, uint32_t& lo) {
  hi = bf16_bits(f);
  const float r1 = f - bf16_val(hi);
  mid = bf16_bits(r1);
  const float r2 = r1 - bf16_val(mid);
  lo = bf16_bits(r2);
}

// inv_bc2 = 1 / sqrt(1 - beta2^t).  sqrt.approx / div.approx (<= 2 ulp): every rank runs the same
// instructions on the same values, so replicas stay bit-identical.
__device__ __forceinline__ void opt_update(int opt, float lr, float beta1, float beta2, float eps, float wd, float step_size,
                                           float inv_bc2, bool first, float g, float& p, float& m, float& v) {
  if (opt == PDM_OPT_ADAM) {
    if (wd != 0.f) g = fmaf(wd, p, g);
    m = m + (1.f - beta1) * (g - m);
    v = v * beta2 + (1.f - beta2) * g * g;
    float sq;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(sq) : "f"(v));
    const float denom = fmaf(sq, inv_bc2, eps);
    p = p - step_size * __fdividef(m, denom);
  } else if (opt == PDM_OPT_SGD) {
    if (wd != 0.f) g = fmaf(wd, p, g);
    if (beta1 != 0.f) {
      m = first ? g : (m * beta1 + g);
      g = m;
    }
    p = p - lr * g;
  }
}

// ---------------------------------------------------------------------------- work enumeration
// A run is `nsteps` steps; a step of n rows is ceil(n / 512) blocks; every role walks the same
// (step, block) sequence and numbers the blocks with one running counter (mbarrier parities).
struct Blk {
  int rows;        // valid rows of this block (1..512)
  int tile_rows;   // rows rounded up to a multiple of 128 (rows the MMAs and the exchanges cover)
  int n_step;      // rows of the whole step
  bool first, last;
};
__device__ __forceinline__ int step_rows(const TcArgs& a, long long step) {
  if (a.row_index == nullptr) return a.batch;
  const long long left = a.count - step * (long long)a.batch;
  return (int)(left < a.batch ? left : a.batch);
}
__device__ __forceinline__ Blk make_blk(int n_step, int blk) {
  Blk b;
  b.n_step = n_step;
  const int left = n_step - blk * kBlkRows;
  b.rows = left < kBlkRows ? left : kBlkRows;
  b.tile_rows = (b.rows + 127) / 128 * 128;
  b.first = blk == 0;
  b.last = left <= kBlkRows;
  return b;
}
// dataset rows of a block: pool mode = base + r, plan mode = index[r] (one 64-bit modulo per block, not per row)
struct Rows {
  const int32_t* index;
  long long base;
  __device__ __forceinline__ long long operator()(int r) const { return index != nullptr ? (long long)__ldg(index + r) : base + r; }
};
__device__ __forceinline__ Rows block_rows(const TcArgs& a, long long step, int blk) {
  Rows q;
  q.index = nullptr;
  q.base = 0;
  if (a.row_index != nullptr) q.index = a.row_index + step * (long long)a.batch + blk * kBlkRows;
  else q.base = ((a.first_batch + step) % a.pool_batches) * (long long)a.batch + blk * kBlkRows;
  return q;
}

__device__ __forceinline__ uint4 ld_stream16(const uint8_t* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// ring data was written by another SM during this kernel: read it through L2, never through the
// non-coherent path
__device__ __forceinline__ uint4 ld_l2_16(const uint8_t* p) {
  uint4 r;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void flag_wait_ge(const uint32_t* flag, uint32_t target) {
  while ((int32_t)(ld_acquire_gpu(flag) - target) < 0) __nanosleep(64);
}
__device__ __forceinline__ void flag_add_release(uint32_t* flag) {
  __threadfence();
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(flag) : "memory");
}
// four raw pixels -> four bf16 (exact), packed as two words.  No int->float conversion instruction
// (quarter rate): 0x4B000000 | b is the float 8388608 + b, one full-rate FADD gives float(b), whose
// upper 16 bits are its bf16.  (fp16 pixels would need one byte permute per pair, but kind::f16 with an
// fp16 A operand and a bf16 B operand is an illegal instruction on this part: measured.)
__device__ __forceinline__ void px4_to_bf16(uint32_t w, uint32_t& lo, uint32_t& hi) {
  const float f0 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540)) - 8388608.0f;
  const float f1 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7541)) - 8388608.0f;
  const float f2 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7542)) - 8388608.0f;
  const float f3 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7543)) - 8388608.0f;
  lo = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
  hi = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
}

// ---------------------------------------------------------------------------- producer warps
// 224 threads stream the rows (and labels) of the next block into registers while the current block
// computes, then (once the wgrad MMAs of the current block have finished reading the tile) convert
// them to bf16 and store them as 8x8 core matrices.
template <bool kTrace>
__device__ __forceinline__ void producer_loop(const TcArgs& a, uint8_t* smem, Misc* ms, int pt, int cta) {
  const int col0 = cta * kCols;
  const int valid_cols = min(kCols, kIn - col0);
  uint8_t* tile = smem + kOffTile;
  uint32_t it = 0;
  for (long long step = 0; step < a.nsteps; ++step) {
    const int n_step = step_rows(a, step);
    const int nblk = (n_step + kBlkRows - 1) / kBlkRows;
    for (int blk = 0; blk < nblk; ++blk, ++it) {
      const Blk b = make_blk(n_step, blk);
      const Rows rows = block_rows(a, step, blk);
      constexpr int kItems = (kBlkRows * 4 + kProdThreads - 1) / kProdThreads;   // 10
      uint4 raw[kItems];
      uint32_t lab[3];                                         // labels travel with the tile (three rows per thread)
      const uint8_t* slot = nullptr;
      if (a.ring != nullptr) {
        // fed mode: the feeder cluster has gathered whole rows (large PCIe reads) into a device ring slot;
        // this CTA picks its 64-byte column slices out of it through L2
        if (pt == 0) flag_wait_ge(a.ring_flags + (it % kRingSlots), (uint32_t)kCtas * (it / kRingSlots + 1));
        asm volatile("bar.sync 2, 224;" ::: "memory");
        slot = a.ring + (size_t)(it % kRingSlots) * kRingSlotBytes;
      }
#pragma unroll
      for (int j = 0; j < kItems; ++j) {
        const int item = pt + kProdThreads * j;
        const int r = item >> 2, q = item & 3;
        raw[j] = make_uint4(0u, 0u, 0u, 0u);
        if (r < b.rows && q * 16 < valid_cols)
          raw[j] = slot != nullptr ? ld_l2_16(slot + (size_t)r * kIn + col0 + q * 16)
                                   : ld_stream16(a.x + rows(r) * kIn + col0 + q * 16);
      }
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int r = pt + kProdThreads * j;
        lab[j] = 0u;
        if (r < b.rows) {
          if (slot != nullptr) {
            uint32_t v;
            asm volatile("ld.relaxed.gpu.global.u8 %0, [%1];" : "=r"(v) : "l"(slot + (size_t)kBlkRows * kIn + r) : "memory");
            lab[j] = v;
          } else {
            lab[j] = min((uint32_t)__ldg(a.y + rows(r)), (uint32_t)(kOut - 1));
          }
        }
      }
      TC_TRACE(pt == 32, step, 16);
      if (it > 0) {                                            // previous block's MMAs are done with the tile
        if (a.dbg & 1) mbar_wait(&ms->bar_wg, (it - 1) & 1u);
        else mbar_wait_relaxed(&ms->bar_wg, (it - 1) & 1u);
      }
      TC_TRACE(pt == 32, step, 17);
#pragma unroll
      for (int j = 0; j < kItems; ++j) {
        const int item = pt + kProdThreads * j;
        const int r = item >> 2, q = item & 3;
        if (r >= b.tile_rows) continue;
        uint4 c0, c1;
        px4_to_bf16(raw[j].x, c0.x, c0.y);
        px4_to_bf16(raw[j].y, c0.z, c0.w);
        px4_to_bf16(raw[j].z, c1.x, c1.y);
        px4_to_bf16(raw[j].w, c1.z, c1.w);
        uint8_t* dst = tile + (uint32_t)(r >> 3) * kSr + (uint32_t)(2 * q) * kSc + (uint32_t)(r & 7) * 16u;
        *reinterpret_cast<uint4*>(dst) = c0;
        *reinterpret_cast<uint4*>(dst + kSc) = c1;
      }
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int r = pt + kProdThreads * j;
        if (r < kBlkRows) ms->ys[r] = (uint8_t)lab[j];
      }
      fence_async_smem();                                      // generic-proxy stores -> visible to the MMA (async proxy)
      mbar_arrive(&ms->bar_xfull);
      if (a.ring != nullptr) {                                 // every producer thread has its slice in registers: the slot may be refilled
        asm volatile("bar.sync 2, 224;" ::: "memory");
        if (pt == 0) flag_add_release(a.ring_flags + kRingSlots + (it % kRingSlots));
      }
      TC_TRACE(pt == 32, step, 18);
    }
  }
}

// ---------------------------------------------------------------------------- feeder cluster
// A second cluster of the same launch (13 CTAs x 12 warps) walks the same block sequence a few blocks
// ahead: every warp copies whole rows -- 784 contiguous bytes, i.e. full-size PCIe read requests when the
// dataset is pinned host memory -- into a ring slot in device memory, labels behind the rows.  The slices
// a trainer CTA needs (64 bytes of every row) would otherwise cross PCIe as 64-byte requests.
__device__ __forceinline__ void feeder_loop(const TcArgs& a, int tid, int cta) {
  const int warp = tid >> 5, lane = tid & 31;
  constexpr int kWarps = kThreads / 32;                        // 12
  uint32_t it = 0;
  for (long long step = 0; step < a.nsteps; ++step) {
    const int n_step = step_rows(a, step);
    const int nblk = (n_step + kBlkRows - 1) / kBlkRows;
    for (int blk = 0; blk < nblk; ++blk, ++it) {
      const Blk b = make_blk(n_step, blk);
      const Rows rows = block_rows(a, step, blk);
      const uint32_t s = it % kRingSlots, gen = it / kRingSlots;
      if (gen > 0) {                                           // the slot's previous block has been picked up by all 13 trainer CTAs
        if (tid == 0) flag_wait_ge(a.ring_flags + kRingSlots + s, (uint32_t)kCtas * gen);
        __syncthreads();
      }
      uint8_t* slot = a.ring + (size_t)s * kRingSlotBytes;
      const int per = (b.rows + kCtas - 1) / kCtas;            // rows of this CTA
      const int lo = cta * per, hi = min(b.rows, lo + per);
      constexpr int kRowsPerWarp = 4;                          // rows in flight per warp (8 x 16 bytes per lane)
      for (int r0 = lo + warp * kRowsPerWarp; r0 < hi; r0 += kWarps * kRowsPerWarp) {
        uint4 v0[kRowsPerWarp], v1[kRowsPerWarp];
        uint32_t yv[kRowsPerWarp];
#pragma unroll
        for (int k = 0; k < kRowsPerWarp; ++k) {
          const int r = r0 + k;
          v0[k] = v1[k] = make_uint4(0u, 0u, 0u, 0u);
          yv[k] = 0u;
          if (r < hi) {
            const long long id = rows(r);
            const uint8_t* src = a.x + id * kIn;
            v0[k] = ld_stream16(src + lane * 16);
            if (lane < 17) v1[k] = ld_stream16(src + 512 + lane * 16);
            if (lane == 0) yv[k] = min((uint32_t)__ldg(a.y + id), (uint32_t)(kOut - 1));
          }
        }
#pragma unroll
        for (int k = 0; k < kRowsPerWarp; ++k) {
          const int r = r0 + k;
          if (r < hi) {
            uint8_t* dst = slot + (size_t)r * kIn;
            *reinterpret_cast<uint4*>(dst + lane * 16) = v0[k];
            if (lane < 17) *reinterpret_cast<uint4*>(dst + 512 + lane * 16) = v1[k];
            if (lane == 0) slot[(size_t)kBlkRows * kIn + r] = (uint8_t)yv[k];
          }
        }
      }
      __threadfence();
      __syncthreads();
      if (tid == 0) flag_add_release(a.ring_flags + s);
    }
  }
}

// ---------------------------------------------------------------------------- MMA warp (one lane)
template <bool kTrace>
__device__ __forceinline__ void mma_loop(const TcArgs& a, uint8_t* smem, Misc* ms, uint32_t tmem_base) {
  const uint32_t tile = smem_u32(smem + kOffTile), dlop = smem_u32(smem + kOffDlop), wop = smem_u32(smem + kOffWop);
  uint32_t it = 0;
  for (long long step = 0; step < a.nsteps; ++step) {
    const int n_step = step_rows(a, step);
    const int nblk = (n_step + kBlkRows - 1) / kBlkRows;
    mbar_wait(&ms->bar_wop, (uint32_t)step & 1u);              // this step's weights are in the B operand
    for (int blk = 0; blk < nblk; ++blk, ++it) {
      const Blk b = make_blk(n_step, blk);
      TC_TRACE(true, step, 8);
      mbar_wait(&ms->bar_xfull, it & 1u);
      tc_fence_after();
      TC_TRACE(true, step, 9);
      // ---- forward: [tile_rows x 48] = X[tile_rows x 64] . [W_hi; W_mid; W_lo]^T ----
      const int ntiles = b.tile_rows / 128;
      for (int mt = 0; mt < ntiles; ++mt) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t da = make_desc(tile + (uint32_t)mt * 16u * kSr + (uint32_t)ks * 2u * kSc, kSc, kSr);
          const uint64_t db = make_desc(wop + (uint32_t)ks * 256u, 128u, 1024u);     // 6 row groups of 8: [split][n / 8]
          umma_bf16(tmem_base + (uint32_t)mt * 48u, da, db, kIdescFwd, ks != 0 ? 1u : 0u);
        }
      }
      umma_commit(&ms->bar_fwd);
      TC_TRACE(true, step, 10);
      // ---- wgrad: G[(64 cols + ones + pad) x 48] (+)= X^T . [dl_hi | dl_mid | dl_lo] ----
      mbar_wait(&ms->bar_dl, it & 1u);                        // this CTA's own rows of the operand (and the forward accumulator is drained)
      mbar_wait(&ms->bar_p2, it & 1u);                        // the other CTAs' rows (bulk copies into the operand)
      tc_fence_after();
      TC_TRACE(true, step, 11);
      const int ksteps = b.tile_rows / 16;
      for (int ks = 0; ks < ksteps; ++ks) {
        // MN-major, no swizzle: SBO = stride between 8-wide MN blocks, LBO = stride between 8-deep K groups
        const uint64_t da = make_desc(tile + (uint32_t)ks * 2u * kSr, kSr, kSc);
        const uint64_t db = make_desc(dlop + (uint32_t)ks * 256u, 128u, 8192u);      // 6 column blocks of 8: [split][n / 8]
        umma_bf16(tmem_base + kTmemWg, da, db, kIdescWg, (!b.first || ks != 0) ? 1u : 0u);
      }
      umma_commit(&ms->bar_wg);
      TC_TRACE(true, step, 12);
    }
  }
}

// ---------------------------------------------------------------------------- LL exchange with the peer GPUs
// A CTA's gradient is 65 value sets of 10 floats (64 input columns + the bias) = 325 16-byte LL units
// ({value, flag} x 2).  They are staged in shared memory and ALL 128 epilogue threads push them into
// every peer's inbox and poll the peers' units out of our own inbox -- consecutive threads handle
// consecutive units of one peer (512 contiguous bytes per warp instruction), and the cost does not grow
// with the number of peers the way one-owner-walks-its-peers did (20.6 -> us/step at 8 GPUs).  The
// owners then add the peers' values in rank order, identically on every rank (bit-identical replicas).
// Two buffer halves alternate by sequence parity (a rank can be at most one step ahead of any other).
constexpr int kLLUnits = 65 * kOut / 2;     // 325 units of 16 bytes per (parity, source rank, CTA)
constexpr int kLLSlot = kLLUnits * 2;       // in uint64
template <bool kTrace>
__device__ __forceinline__ void ll_exchange(const TcArgs& a, uint8_t* smem, uint32_t seq, int step, int cta, int t, float (&g)[kHalf], float& gb) {
  const int W = a.peers.world, me = a.peers.rank;
  const size_t par = (size_t)(seq & 1u) * W;
  float2* gbuf = reinterpret_cast<float2*>(smem + kOffGbuf);               // [325] this CTA's units
  float2* rbuf = reinterpret_cast<float2*>(smem + kOffStage);              // [W - 1][325] the peers' units (the staging buffer is only ever written by this CTA and is idle here)
  // float layout of a CTA's message: [64 columns][10 classes], then the ten bias gradients
  {
    float* gf = reinterpret_cast<float*>(gbuf) + (t & (kCols - 1)) * kOut + (t >> 6) * kHalf;
#pragma unroll
    for (int j = 0; j < kHalf; ++j) gf[j] = g[j];
    if (t < kOut) reinterpret_cast<float*>(gbuf)[kCols * kOut + t] = gb;
  }
  epi_sync();
  pdm_skew(a.skew_ns, seq, (uint32_t)(me * 16 + cta), (uint32_t)(t >> 5));
  const int total = (W - 1) * kLLUnits;
  for (int idx = t; idx < total; idx += kEpiThreads) {
    const int k = idx / kLLUnits, u = idx - k * kLLUnits;
    const int p = k + (k >= me ? 1 : 0);
    const float2 v = gbuf[u];
    ll_store2(a.peers.inbox[p] + ((par + me) * kCtas + cta) * kLLSlot + (size_t)u * 2, v.x, v.y, seq);
  }
  TC_TRACE(t == 0, step, 24);
  // poll: up to four units in flight per thread and round (eight spill registers into the hot loop) (one L2 round trip per round)
  constexpr int kInFlight = 4;
  for (int idx0 = t; idx0 < total; idx0 += kInFlight * kEpiThreads) {
    uint32_t pending = 0;
#pragma unroll
    for (int j = 0; j < kInFlight; ++j) if (idx0 + j * kEpiThreads < total) pending |= 1u << j;
    uint32_t spins = 0;
    uint64_t t0 = 0;
    while (pending != 0u) {
      float x[kInFlight], y[kInFlight];
      uint32_t got = 0;
#pragma unroll
      for (int j = 0; j < kInFlight; ++j) {                   // all loads first ...
        if (!((pending >> j) & 1u)) continue;
        const int idx = idx0 + j * kEpiThreads;
        const int k = idx / kLLUnits, u = idx - k * kLLUnits;
        const int p = k + (k >= me ? 1 : 0);
        if (ll_load2(a.peers.inbox[me] + ((par + p) * kCtas + cta) * kLLSlot + (size_t)u * 2, seq, x[j], y[j])) got |= 1u << j;
      }
#pragma unroll
      for (int j = 0; j < kInFlight; ++j) {                   // ... then the stores
        if (!((got >> j) & 1u)) continue;
        const int idx = idx0 + j * kEpiThreads;
        const int k = idx / kLLUnits, u = idx - k * kLLUnits;
        rbuf[k * kLLUnits + u] = make_float2(x[j], y[j]);
      }
      pending &= ~got;
      if (pending != 0u && (++spins & 1023u) == 0u && a.peers.timeout_ns != 0) {
        const uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > a.peers.timeout_ns) {
          const int idx = idx0 + (__ffs(pending) - 1) * kEpiThreads;
          const int k = idx / kLLUnits;
          if (a.peers.status != nullptr) atomicExch(a.peers.status, 1 + k + (k >= me ? 1 : 0));
          break;
        }
      }
    }
  }
  TC_TRACE(t == 0, step, 25);
  epi_sync();
  TC_TRACE(t == 0, step, 26);
  {
    float sum[kHalf], sb = 0.f;
#pragma unroll
    for (int j = 0; j < kHalf; ++j) sum[j] = 0.f;
    const int mine = (t & (kCols - 1)) * kOut + (t >> 6) * kHalf;
    for (int p = 0; p < W; ++p) {                           // rank order on every rank: bit-identical replicas
      if (p == me) {
#pragma unroll
        for (int j = 0; j < kHalf; ++j) sum[j] += g[j];
        sb += gb;
      } else {
        const float* rf = reinterpret_cast<const float*>(rbuf + (p - (p > me ? 1 : 0)) * kLLUnits);
#pragma unroll
        for (int j = 0; j < kHalf; ++j) sum[j] += rf[mine + j];
        if (t < kOut) sb += rf[kCols * kOut + t];
      }
    }
#pragma unroll
    for (int j = 0; j < kHalf; ++j) g[j] = sum[j];
    gb = sb;
  }
  epi_sync();                                 // gbuf / rbuf are the next step's exchange buffers again
}

// Publish this thread's five weights (column c, classes 5 h .. 5 h + 4) as the next forward's B operand
// (three bf16 terms, K-major 8x8 core matrices [split][n/8][k/8][n%8][k%8]) and its share of sum_k W[n][k].
__device__ __forceinline__ void publish_weights(uint8_t* smem, Misc* ms, int c, int h, int lane, int warp, const float (&p)[kHalf]) {
  uint8_t* wop = smem + kOffWop;
  float r[8];
#pragma unroll
  for (int j = 0; j < kHalf; ++j) {
    const int n = h * kHalf + j;
    uint32_t hi, mid, lo;
    split3(p[j], hi, mid, lo);
    const uint32_t off = (uint32_t)(n >> 3) * 1024u + (uint32_t)(c >> 3) * 128u + (uint32_t)(n & 7) * 16u + (uint32_t)(c & 7) * 2u;
    *reinterpret_cast<uint16_t*>(wop + off) = (uint16_t)hi;
    *reinterpret_cast<uint16_t*>(wop + kWopSplitBytes + off) = (uint16_t)mid;
    *reinterpret_cast<uint16_t*>(wop + 2 * kWopSplitBytes + off) = (uint16_t)lo;
    r[j] = p[j];
  }
  r[5] = r[6] = r[7] = 0.f;
  // Transposing warp reduction: 9 shuffles for the five sums instead of 25.  After the three halving
  // levels lane L holds value (L >> 2) & 7 summed over 8 lanes; two more exchanges complete it.
#pragma unroll
  for (int off = 16, w = 4; off >= 4; off >>= 1, w >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < w; ++i) {
      const float keep = upper ? r[i + w] : r[i];
      const float send = upper ? r[i] : r[i + w];
      r[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  r[0] += __shfl_xor_sync(0xffffffffu, r[0], 2);
  r[0] += __shfl_xor_sync(0xffffffffu, r[0], 1);
  const int which = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
  if ((lane & 3) == 0 && which < kHalf) ms->wsum[warp & 1][h * kHalf + which] = r[0];
}

// ---------------------------------------------------------------------------- epilogue / optimizer warps
template <bool kTrace>
__device__ __forceinline__ void epilogue_loop(const TcArgs& a, uint8_t* smem, Misc* ms, uint32_t tmem_base, int t, int cta) {
  const int warp = t >> 5, lane = t & 31;
  // optimizer roles: thread t owns column c = t % 64, classes 5 h .. 5 h + 4 (h = t / 64): five weights per
  // thread on all four warps; threads 0-9 also keep a replica of bias[t]
  const int c = t & (kCols - 1), h = t >> 6;
  const int col = cta * kCols + c;
  const bool owns_col = col < kIn;
  const bool has_bias = t < kOut;
  float* xbuf = reinterpret_cast<float*>(smem + kOffXbuf);
  float* stage = reinterpret_cast<float*>(smem + kOffStage);
  uint8_t* dlop = smem + kOffDlop;
  const uint32_t xbuf_s = smem_u32(xbuf), stage_s = smem_u32(stage), dlop_s = smem_u32(dlop);
  const uint32_t p1_s = smem_u32(&ms->bar_p1), p2_s = smem_u32(&ms->bar_p2), loss_s = smem_u32(&ms->bar_loss);
  const uint32_t lane_taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const int own_cap = min(kOwnRows, kBlkRows - cta * kOwnRows);

  // ---- persistent per-thread state: weights and optimizer moments ----
  float p[kHalf], m[kHalf], v[kHalf], bp = 0.f, bm = 0.f, bv = 0.f;
#pragma unroll
  for (int j = 0; j < kHalf; ++j) {
    p[j] = m[j] = v[j] = 0.f;
    const int idx = (h * kHalf + j) * kIn + col;
    if (owns_col) {
      p[j] = a.params[idx];
      if (a.m != nullptr) m[j] = a.m[idx];
      if (a.v != nullptr) v[j] = a.v[idx];
    }
  }
  if (has_bias) {
    bp = a.params[kIn * kOut + t];
    if (a.m != nullptr) bm = a.m[kIn * kOut + t];
    if (a.v != nullptr) bv = a.v[kIn * kOut + t];
  }

  float step_loss = 0.f, step_correct = 0.f;               // CTA 0, thread 0
  uint32_t it = 0;
  for (long long step = 0; step < a.nsteps; ++step) {
    const int n_step = step_rows(a, step);
    const int nblk = (n_step + kBlkRows - 1) / kBlkRows;
    const float inv_batch = 1.0f / (float)n_step;
    const bool last_step = step + 1 == a.nsteps;
    // ---- 0. this step's weights -> the forward B operand (three bf16 terms) ----
    publish_weights(smem, ms, c, h, lane, warp, p);
    if (has_bias) ms->bias[t] = bp;
    fence_async_smem();
    tc_fence_before();
    mbar_arrive(&ms->bar_wop);
    epi_sync();
    TC_TRACE(t == 0, step, 13);
    const float2 adam_c = a.adam != nullptr ? __ldg(a.adam + step) : a.adam_inline[step];   // consumed at the end of the step
    float wsumn[kOut];
#pragma unroll
    for (int n = 0; n < kOut; ++n) wsumn[n] = kShift * (ms->wsum[0][n] + ms->wsum[1][n]);

    for (int blk = 0; blk < nblk; ++blk, ++it) {
      const Blk b = make_blk(n_step, blk);
      const uint32_t par = it & 1u;
      const int ntiles = b.tile_rows / 128;
      const int my_rows = max(0, min(own_cap, b.tile_rows - cta * kOwnRows));
      if (t == 0) {
        mbar_expect_tx(&ms->bar_p1, (uint32_t)(kCtas * my_rows * kOut * 4));
        mbar_expect_tx(&ms->bar_p2, (uint32_t)((b.tile_rows - my_rows) * 96));   // 6 x 16 bytes per remote row
        if (cta == 0) mbar_expect_tx(&ms->bar_loss, (uint32_t)(kCtas * 8));
      }

      // ---- 1. partial logits of this CTA's 64 columns -> the rows' owners ----
      TC_TRACE(t == 0, step, 0);
      mbar_wait(&ms->bar_fwd, par);
      tc_fence_after();
      TC_TRACE(t == 0, step, 1);
      // staged in shared memory, then ONE bulk copy per owner: rows [40 o, 40 o + 40) are contiguous
      // and land in slot `cta` of owner o's exchange buffer
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        if (mt >= ntiles) continue;
        uint32_t hi[16], mid[16], lo[16];
        tmem_ld16_issue(lane_taddr + (uint32_t)mt * 48u, hi);
        tmem_ld16_issue(lane_taddr + (uint32_t)mt * 48u + 16u, mid);
        tmem_ld16_issue(lane_taddr + (uint32_t)mt * 48u + 32u, lo);
        tmem_ld_wait();
        float2* dst = reinterpret_cast<float2*>(stage + (mt * 128 + t) * kOut);
#pragma unroll
        for (int n = 0; n < kOut; n += 2) {
          const float z0 = (__uint_as_float(hi[n]) + __uint_as_float(mid[n])) + __uint_as_float(lo[n]);
          const float z1 = (__uint_as_float(hi[n + 1]) + __uint_as_float(mid[n + 1])) + __uint_as_float(lo[n + 1]);
          dst[n / 2] = make_float2(fmaf(kScale, z0, -wsumn[n]), fmaf(kScale, z1, -wsumn[n + 1]));
        }
      }
      fence_async_smem();
      epi_sync();
      if (t < kCtas) {
        const int rows_o = max(0, min(min(kOwnRows, kBlkRows - t * kOwnRows), b.tile_rows - t * kOwnRows));
        if (rows_o > 0)
          bulk_s2s(mapa(xbuf_s + (uint32_t)(cta * kOwnRows * kOut) * 4u, (uint32_t)t), stage_s + (uint32_t)(t * kOwnRows * kOut) * 4u,
                   (uint32_t)(rows_o * kOut * 4), mapa(p1_s, (uint32_t)t));
      }
      mbar_wait(&ms->bar_xfull, par);          // (long complete) acquire the producers' label stores

      // ---- 2. owner: two lanes per row (5 classes each): sum the 13 partials, soft-max, loss, top-1,
      //         dlogits as three bf16 terms straight into the wgrad operand ----
      TC_TRACE(t == 0, step, 2);
      mbar_wait(&ms->bar_p1, par);
      TC_TRACE(t == 0, step, 3);
      float my_loss = 0.f, my_hit = 0.f;
      {
        const int lr = t >> 1, half = t & 1;                 // lr < 64: rows beyond my_rows idle along (shuffles need every lane)
        const bool active = lr < my_rows;
        const int r = cta * kOwnRows + lr;
        const bool valid = active && r < b.rows;
        const int c0 = half * 5;
        float z[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) z[k] = ms->bias[c0 + k];
        if (active) {
#pragma unroll
          for (int src = 0; src < kCtas; ++src) {
#pragma unroll
            for (int k = 0; k < 5; ++k) z[k] += xbuf[(src * kOwnRows + lr) * kOut + c0 + k];
          }
        }
        const int y = active ? (int)ms->ys[r] : 0;
        float mx = fmaxf(fmaxf(fmaxf(z[0], z[1]), fmaxf(z[2], z[3])), z[4]);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        float e[5], s = 0.f, zy = -INFINITY;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          e[k] = __expf(z[k] - mx);                          // ex2.approx: <= 2e-6 relative on the probabilities
          s += e[k];
          zy = (c0 + k) == y ? z[k] : zy;
        }
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        zy = fmaxf(zy, __shfl_xor_sync(0xffffffffu, zy, 1));
        bool beaten = false;
#pragma unroll
        for (int k = 0; k < 5; ++k) beaten = beaten || z[k] > zy || (z[k] == zy && (c0 + k) < y);
        const int other_beaten = __shfl_xor_sync(0xffffffffu, beaten ? 1 : 0, 1);   // every lane shuffles (no short-circuit)
        beaten = beaten || other_beaten != 0;
        const float inv_s = __fdividef(1.0f, s);
        if (active) {
          // wgrad B operand, MN-major 8x8 core matrices: [split][n / 8][row][n % 8] bf16 (columns 10-15 stay zero)
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            const int n = c0 + k;
            const float d = valid ? (e[k] * inv_s - (n == y ? 1.f : 0.f)) * inv_batch : 0.f;
            uint32_t hi, mid, lo;
            split3(d, hi, mid, lo);
            uint8_t* dst = dlop + (uint32_t)(n >> 3) * 8192u + (uint32_t)r * 16u + (uint32_t)(n & 7) * 2u;
            *reinterpret_cast<uint16_t*>(dst) = (uint16_t)hi;
            *reinterpret_cast<uint16_t*>(dst + kDlSplitBytes) = (uint16_t)mid;
            *reinterpret_cast<uint16_t*>(dst + 2 * kDlSplitBytes) = (uint16_t)lo;
            if (a.logits_out != nullptr && last_step && valid) a.logits_out[(size_t)(blk * kBlkRows + r) * kOut + n] = z[k];
          }
        }
        if (valid && half == 0) {
          my_loss = (mx + logf(s)) - zy;
          my_hit = beaten ? 0.f : 1.f;
        }
      }
      if (warp < 3) {                                        // rows 0-15, 16-31, 32-39: fixed butterfly, deterministic
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
          my_loss += __shfl_xor_sync(0xffffffffu, my_loss, o);
          my_hit += __shfl_xor_sync(0xffffffffu, my_hit, o);
        }
        if (lane == 0) { ms->lossw[warp][0] = my_loss; ms->lossw[warp][1] = my_hit; }
      }
      fence_async_smem();
      epi_sync();
      // my rows of the operand -> the same rows of every other CTA's operand: 6 chunks (split, n / 8)
      // of my_rows x 16 bytes per destination, one transaction each
      if (t < 6 * (kCtas - 1) && my_rows > 0) {
        const int dsel = t / 6, chunk = t - dsel * 6;
        const uint32_t dest = (uint32_t)(dsel + (dsel >= cta ? 1 : 0));
        const uint32_t off = (uint32_t)chunk * 8192u + (uint32_t)(cta * kOwnRows) * 16u;
        bulk_s2s(mapa(dlop_s + off, dest), dlop_s + off, (uint32_t)my_rows * 16u, mapa(p2_s, dest));
      }
      if (t == 0)
        st_async2(mapa(smem_u32(&ms->lossbuf[cta][0]), 0u), (ms->lossw[0][0] + ms->lossw[1][0]) + ms->lossw[2][0],
                  (ms->lossw[0][1] + ms->lossw[1][1]) + ms->lossw[2][1], mapa(loss_s, 0u));
      tc_fence_before();
      mbar_arrive(&ms->bar_dl);                              // own rows written, forward accumulator drained
      TC_TRACE(t == 0, step, 4);

      if (cta == 0 && t == 0) {                              // the step's metrics
        mbar_wait(&ms->bar_loss, par);
        for (int i = 0; i < kCtas; ++i) { step_loss += ms->lossbuf[i][0]; step_correct += ms->lossbuf[i][1]; }
        if (b.last) {
          if (a.step_out != nullptr) {
            a.step_out[2 * step] = step_loss * inv_batch;
            a.step_out[2 * step + 1] = step_correct;
          }
          if (a.metrics != nullptr) {
            a.metrics[0] += (double)step_loss;
            a.metrics[1] += (double)step_correct;
            a.metrics[2] += (double)n_step;
          }
          step_loss = step_correct = 0.f;
        }
      }

      // ---- 3. gradient of this CTA's columns, all-reduce, optimizer, next step's B operand ----
      TC_TRACE(t == 0, step, 6);
      mbar_wait(&ms->bar_wg, par);
      TC_TRACE(t == 0, step, 7);
      if (!b.last) continue;
      tc_fence_after();
      // accumulator row c holds column c's ten sums (row 64: the bias gradient): warps 0-1 keep classes 0-4 and
      // hand classes 5-9 to warps 2-3 through shared memory (the staging buffer is idle here)
      float g[kHalf], gb = 0.f;
      float* xg = stage;                                     // [5][64]
      {
        uint32_t hi[16], mid[16], lo[16];
        tmem_ld16_issue(lane_taddr + kTmemWg, hi);
        tmem_ld16_issue(lane_taddr + kTmemWg + 16u, mid);
        tmem_ld16_issue(lane_taddr + kTmemWg + 32u, lo);
        tmem_ld_wait();
        float gsum[kOut];
#pragma unroll
        for (int n = 0; n < kOut; ++n) gsum[n] = (__uint_as_float(hi[n]) + __uint_as_float(mid[n])) + __uint_as_float(lo[n]);
        if (h == 0) {
#pragma unroll
          for (int j = 0; j < kHalf; ++j) { g[j] = gsum[j]; xg[j * kCols + c] = gsum[kHalf + j]; }
        } else if (t == kCols) {
#pragma unroll
          for (int n = 0; n < kOut; ++n) ms->db[n] = gsum[n];
        }
      }
      epi_sync();
      TC_TRACE(t == 0, step, 21);
      if (h == 1) {
#pragma unroll
        for (int j = 0; j < kHalf; ++j) g[j] = xg[j * kCols + c];
      }
#pragma unroll
      for (int j = 0; j < kHalf; ++j) g[j] = owns_col ? fmaf(kScale, g[j], -kShift * ms->db[h * kHalf + j]) : 0.f;
      if (has_bias) gb = ms->db[t];
      if (a.peers.world > 1) ll_exchange<kTrace>(a, smem, a.peers.seq + (uint32_t)step, step, cta, t, g, gb);
      TC_TRACE(t == 0, step, 27);
      {
        const float step_size = adam_c.x, inv_bc2 = adam_c.y;
        const bool first = a.step0 + step <= 1;
        if (owns_col) {
#pragma unroll
          for (int j = 0; j < kHalf; ++j) {
            g[j] *= a.inv_world;
            opt_update(a.opt, a.lr, a.beta1, a.beta2, a.eps, a.wd, step_size, inv_bc2, first, g[j], p[j], m[j], v[j]);
            const int idx = (h * kHalf + j) * kIn + col;
            if (a.grad_out != nullptr) a.grad_out[idx] = g[j];
            if (last_step) {
              a.params[idx] = p[j];
              if (a.m != nullptr) a.m[idx] = m[j];
              if (a.v != nullptr) a.v[idx] = v[j];
            }
          }
        }
        if (has_bias) {
          gb *= a.inv_world;
          opt_update(a.opt, a.lr, a.beta1, a.beta2, a.eps, a.wd, step_size, inv_bc2, first, gb, bp, bm, bv);
          if (cta == 0) {
            const int idx = kIn * kOut + t;
            if (a.grad_out != nullptr) a.grad_out[idx] = gb;
            if (last_step) {
              a.params[idx] = bp;
              if (a.m != nullptr) a.m[idx] = bm;
              if (a.v != nullptr) a.v[idx] = bv;
            }
          }
        }
      }
      TC_TRACE(t == 0, step, 22);
      // Reconverge here: the guarded stores above leave the warp split, and without this the next step's
      // publish code runs once per fragment (measured on the same box: 7.76 -> 7.30 us/step at batch 512).
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------- the kernel
template <bool kTrace>
__global__ void __launch_bounds__(kThreads, 1) net_tc_kernel(const __grid_constant__ TcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  Misc* ms = reinterpret_cast<Misc*>(smem + kOffMisc);
  const int tid = threadIdx.x, warp = tid >> 5;
  uint32_t cta;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta));

  if (blockIdx.x >= kCtas) {              // second cluster of the launch: the feeder (only launched with a ring)
    feeder_loop(a, tid, (int)cta);
    return;
  }
  // Nothing but the bookkeeping block (and the trace area) needs defined contents: every operand element that
  // reaches an accumulator row / column the epilogues read is written before it is used; padding (operand rows
  // 10-15, tile column blocks 9-15, the other seven entries of the ones block) only feeds accumulator rows and
  // columns nobody reads, so whatever bits they hold (NaN included) are harmless.  The ones column is set once.
  for (uint32_t i = tid; i < (kSmemBytes - kOffMisc) / 16; i += kThreads)
    reinterpret_cast<uint4*>(smem + kOffMisc)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  for (int r = tid; r < kBlkRows; r += kThreads)
    *reinterpret_cast<uint16_t*>(smem + kOffTile + (uint32_t)(r >> 3) * kSr + 8u * kSc + (uint32_t)(r & 7) * 16u) = 0x3F80;  // bf16 1.0
  if (tid == 0) {
    mbar_init(&ms->bar_xfull, kProdThreads);
    mbar_init(&ms->bar_fwd, 1);
    mbar_init(&ms->bar_dl, kEpiThreads);
    mbar_init(&ms->bar_wg, 1);
    mbar_init(&ms->bar_wop, kEpiThreads);
    mbar_init(&ms->bar_p1, 1);
    mbar_init(&ms->bar_p2, 1);
    mbar_init(&ms->bar_loss, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) tmem_alloc(&ms->tmem_slot, kTmemCols);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();                     // every CTA's mbarriers exist before anyone copies to them
  const uint32_t tmem_base = ms->tmem_slot;

  if (tid >= kEpiFirst) {
    epilogue_loop<kTrace>(a, smem, ms, tmem_base, tid - kEpiFirst, (int)cta);
  } else if (warp == kMmaWarp) {
    if ((tid & 31) == 0) mma_loop<kTrace>(a, smem, ms, tmem_base);
    __syncwarp();
  } else {
    producer_loop<kTrace>(a, smem, ms, tid, (int)cta);
  }

  tc_fence_before();
  __syncthreads();
  if (kTrace && a.trace != nullptr)
    for (int i = tid; i < kTraceSteps * 32; i += kThreads)
      a.trace[(size_t)cta * kTraceSteps * 32 + i] = reinterpret_cast<const long long*>(smem + kOffTrace)[i];
  cluster_sync_all();                     // no CTA leaves while a peer may still address its shared memory
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace

// ---------------------------------------------------------------------------- host side
static void tc_launch_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, cudaStream_t stream, int clusters = 1) {
  *cfg = cudaLaunchConfig_t{};
  cfg->gridDim = dim3(kCtas * clusters);
  cfg->blockDim = dim3(kThreads);
  cfg->dynamicSmemBytes = kSmemBytes;
  cfg->stream = stream;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg->attrs = attr;
  cfg->numAttrs = 1;
}

extern "C" int pdm_net_tc_supported(int device) {
  static int cached[64];
  static bool known[64];
  if (device < 0 || device >= 64) return 0;
  if (known[device]) return cached[device];
  int ok = 0;
  do {
    const char* e = getenv("PDM_NET_TC");
    if (e && *e == '0') break;
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess) break;
    if (cur != device && cudaSetDevice(device) != cudaSuccess) break;
    int major = 0, cluster = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    cudaDeviceGetAttribute(&cluster, cudaDevAttrClusterLaunch, device);
    if (major == 10 && cluster) {
      if (cudaFuncSetAttribute(net_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes) == cudaSuccess &&
          cudaFuncSetAttribute(net_tc_kernel<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
          cudaFuncSetAttribute(net_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes) == cudaSuccess &&
          cudaFuncSetAttribute(net_tc_kernel<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
        cudaLaunchConfig_t cfg;
        cudaLaunchAttribute attr[1];
        tc_launch_config(&cfg, attr, nullptr);
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, net_tc_kernel<false>, &cfg) == cudaSuccess && nclusters >= 1) ok = 1;
      }
    }
    cudaGetLastError();
    if (cur != device && cur >= 0) cudaSetDevice(cur);
  } while (0);
  cached[device] = ok;
  known[device] = true;
  return ok;
}

// PDM_TC_TRACE=1: every CTA stamps clock64 at its synchronisation points for the first 64 steps of
// a run (profiling aid, benchmarks/tc_trace.py); pdm_net_tc_trace_read copies the stamps out.
static long long* g_tc_trace = nullptr;
static long long* tc_trace_buffer() {
  static const bool on = [] { const char* e = getenv("PDM_TC_TRACE"); return e && *e == '1'; }();
  if (!on) return nullptr;
  const size_t bytes = (size_t)kCtas * kTraceSteps * 32 * sizeof(long long);
  if (g_tc_trace == nullptr && cudaMalloc((void**)&g_tc_trace, bytes) != cudaSuccess) return nullptr;
  cudaMemset(g_tc_trace, 0, bytes);
  return g_tc_trace;
}
extern "C" int pdm_net_tc_trace_read(long long* out) {
  PDM_REQUIRE(out != nullptr, PDM_ERR_ARG, "pdm_net_tc_trace_read: out is NULL");
  PDM_REQUIRE(g_tc_trace != nullptr, PDM_ERR_STATE, "pdm_net_tc_trace_read: tracing is off (set PDM_TC_TRACE=1)");
  PDM_CHECK_CUDA(cudaDeviceSynchronize());
  PDM_CHECK_CUDA(cudaMemcpy(out, g_tc_trace, (size_t)kCtas * kTraceSteps * 32 * sizeof(long long), cudaMemcpyDeviceToHost));
  return PDM_OK;
}

// Adam's per-step bias corrections are computed here, on the host, in double precision (the same arithmetic as
// torch.optim.Adam and net_step.cu) and handed to the kernel through a fixed device array: no pow() in the
// kernel's instruction stream.  A launch covers at most kMaxLaunchSteps steps; longer runs are cut into launches.
// The array is shared by all calls on a device: they must be ordered on one stream (the engine's are).
constexpr long long kMaxLaunchSteps = PDM_TC_MAX_LAUNCH_STEPS;
static __device__ float2 g_adam_consts[kMaxLaunchSteps];

static int adam_array(float2** out) {
  static float2* addr[64];
  int dev = -1;
  PDM_CHECK_CUDA(cudaGetDevice(&dev));
  PDM_REQUIRE(dev >= 0 && dev < 64, PDM_ERR_LIMIT, "pdm_net_tc_run: device ordinal %d outside [0, 64)", dev);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    PDM_CHECK_CUDA(cudaGetSymbolAddress(&p, g_adam_consts));
    addr[dev] = static_cast<float2*>(p);
  }
  *out = addr[dev];
  return PDM_OK;
}

extern "C" int pdm_net_tc_run(const pdm_net_tc_args* g, pdm_stream_t stream) {
  PDM_REQUIRE(g != nullptr, PDM_ERR_ARG, "pdm_net_tc_run: args is NULL");
  PDM_REQUIRE(g->x && g->y && g->params, PDM_ERR_ARG, "pdm_net_tc_run: x, y and params are required");
  PDM_REQUIRE(((uintptr_t)g->x & 15) == 0, PDM_ERR_ARG, "pdm_net_tc_run: x must be 16-byte aligned");
  PDM_REQUIRE(g->ring == nullptr || (((uintptr_t)g->ring & 255) == 0 && g->ring_bytes >= PDM_TC_RING_BYTES), PDM_ERR_ARG,
              "pdm_net_tc_run: ring must be 256-byte aligned device memory of at least PDM_TC_RING_BYTES (%zu) bytes",
              (size_t)PDM_TC_RING_BYTES);
  PDM_REQUIRE(g->batch >= 1, PDM_ERR_ARG, "pdm_net_tc_run: batch must be >= 1 (got %d)", g->batch);
  PDM_REQUIRE(g->nsteps >= 0, PDM_ERR_ARG, "pdm_net_tc_run: nsteps must be >= 0");
  const pdm_hyper& h = g->hyper;
  PDM_REQUIRE(h.opt == PDM_OPT_ADAM || h.opt == PDM_OPT_SGD || h.opt == PDM_OPT_NONE, PDM_ERR_ARG,
              "pdm_net_tc_run: unknown optimizer %d", h.opt);
  PDM_REQUIRE(h.opt != PDM_OPT_ADAM || (g->m && g->v), PDM_ERR_ARG, "pdm_net_tc_run: Adam needs m and v");
  PDM_REQUIRE(!(h.opt == PDM_OPT_SGD && h.beta1 != 0.f) || g->m, PDM_ERR_ARG, "pdm_net_tc_run: SGD with momentum needs m");
  long long nsteps = g->nsteps;
  long long pool_batches = 0;
  if (g->row_index != nullptr) {
    PDM_REQUIRE(g->count >= 0, PDM_ERR_ARG, "pdm_net_tc_run: count must be >= 0");
    const long long need = (g->count + g->batch - 1) / g->batch;
    if (nsteps == 0 || nsteps > need) nsteps = need;
  } else {
    PDM_REQUIRE(g->pool_rows >= g->batch, PDM_ERR_ARG, "pdm_net_tc_run: the pool holds fewer rows than one batch");
    pool_batches = g->pool_rows / g->batch;
  }
  if (nsteps == 0) return PDM_OK;
  const int world = g->peers.world > 1 ? g->peers.world : 1;
  PDM_REQUIRE(world <= PDM_MAX_RANKS, PDM_ERR_LIMIT, "pdm_net_tc_run: world %d > %d", world, PDM_MAX_RANKS);
  if (world > 1) {
    PDM_REQUIRE(g->peers.seq != 0, PDM_ERR_ARG, "pdm_net_tc_run: peers.seq must be non-zero");
    PDM_REQUIRE(g->peers.rank >= 0 && g->peers.rank < world, PDM_ERR_ARG, "pdm_net_tc_run: bad rank");
    for (int p = 0; p < world; ++p)
      PDM_REQUIRE(g->peers.inbox[p] != nullptr, PDM_ERR_STATE, "pdm_net_tc_run: inbox[%d] not mapped", p);
  }
  int dev = 0;
  PDM_CHECK_CUDA(cudaGetDevice(&dev));
  PDM_REQUIRE(pdm_net_tc_supported(dev) == 1, PDM_ERR_STATE,
              "pdm_net_tc_run: device %d cannot launch a %d-CTA cluster with %u bytes of shared memory per CTA "
              "(or PDM_NET_TC=0)", dev, kCtas, kSmemBytes);
  float2* adam_dev = nullptr;
  int rc = adam_array(&adam_dev);
  if (rc != PDM_OK) return rc;

  static thread_local float2 adam_host[kMaxLaunchSteps];
  for (long long done = 0; done < nsteps; done += kMaxLaunchSteps) {
    const long long n = nsteps - done < kMaxLaunchSteps ? nsteps - done : kMaxLaunchSteps;
    TcArgs a;
    memset(&a, 0, sizeof(a));
    a.x = g->x;
    a.y = g->y;
    a.batch = g->batch;
    a.nsteps = n;
    if (g->row_index != nullptr) {
      a.row_index = g->row_index + done * g->batch;
      a.count = g->count - done * g->batch;
    } else {
      a.pool_batches = pool_batches;
      a.first_batch = (g->first_batch + done) % pool_batches;
    }
    a.params = g->params;
    a.m = g->m;
    a.v = g->v;
    a.grad_out = g->grad_out;
    a.logits_out = g->logits_out;
    a.metrics = g->metrics;
    a.step_out = g->step_out != nullptr ? g->step_out + 2 * done : nullptr;
    a.opt = h.opt;
    a.lr = h.lr; a.beta1 = h.beta1; a.beta2 = h.beta2; a.eps = h.eps; a.wd = h.weight_decay;
    a.step0 = h.step + done;
    a.inv_world = world > 1 ? 1.0f / (float)world : 1.0f;
    a.adam = adam_dev;
    a.skew_ns = pdm_skew_ns();
    if (g->ring != nullptr && g->row_index != nullptr) {
      a.ring = static_cast<uint8_t*>(g->ring);
      a.ring_flags = reinterpret_cast<uint32_t*>(a.ring + (size_t)kRingSlots * kRingSlotBytes);
      PDM_CHECK_CUDA(cudaMemsetAsync(a.ring_flags, 0, 2 * kRingSlots * sizeof(uint32_t), (cudaStream_t)stream));
    }
    a.trace = tc_trace_buffer();
    a.peers = g->peers;
    a.peers.world = world;
    a.peers.seq = g->peers.seq + (uint32_t)done;
    float2* consts = n <= kInlineSteps ? a.adam_inline : adam_host;
    for (long long i = 0; i < n; ++i) {
      const long long idx = a.step0 + i;
      const double sidx = (double)(idx < 1 ? 1 : idx);
      consts[i] = make_float2(0.f, 1.f);
      if (h.opt == PDM_OPT_ADAM) {
        consts[i].x = (float)((double)h.lr / (1.0 - pow((double)h.beta1, sidx)));
        consts[i].y = (float)(1.0 / sqrt(1.0 - pow((double)h.beta2, sidx)));
      }
    }
    if (n <= kInlineSteps) {
      a.adam = nullptr;              // short launch: the constants travel in the kernel parameters
    } else {
      // pageable source: the runtime stages it before returning, so adam_host may be reused at once
      PDM_CHECK_CUDA(cudaMemcpyAsync(adam_dev, adam_host, (size_t)n * sizeof(float2), cudaMemcpyHostToDevice, (cudaStream_t)stream));
    }
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    tc_launch_config(&cfg, attr, (cudaStream_t)stream, a.ring != nullptr ? 2 : 1);
    void* params[] = {&a};
    const void* fn = a.trace != nullptr ? (const void*)net_tc_kernel<true> : (const void*)net_tc_kernel<false>;
    PDM_CHECK_CUDA(cudaLaunchKernelExC(&cfg, fn, params));
  }
  return PDM_OK;
}
