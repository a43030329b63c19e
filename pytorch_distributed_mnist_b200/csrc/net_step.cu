// net_step.cu -- the reference model's whole training step as ONE sm_100a kernel.
//
// Reference path being replaced (ref:N = /root/reference/multi_proc_single_gpu.py):
//   ref:87  model(data)            -> cuBLAS addmm            
//   ref:88  F.cross_entropy        -> ATen softmax + nll      (K4)
//   ref:91  loss.backward()        -> nll/softmax bwd, cuBLAS mm, column sum (K5-K7)
//           DDP reducer            -> 4 bucket copies + ncclAllReduce      (K8-K10)
//   ref:92  optimizer.step()       -> 8 _foreach launches                  (K11)
//   ref:94-95 loss.item(), argmax/eq/sum/.item() -> 4 kernels + 2 host syncs (K12-K13)
// ~25 launches and two syncs for 16 MFLOP.  Here: one launch.
//
// Shape of the kernel (B = local batch, G = CTAs, RT = rows per tile):
//   phase 1  CTA c owns a contiguous slice of the batch.  Thread t < 196 owns
//            four input columns [4t, 4t+4): it keeps W[:, 4t..4t+3] (40 regs)
//            and its partial dW (40 regs) in registers for the whole kernel and
//            streams the rows through (raw uint8 -> LUT-normalised fp32, or
//            fp32).  Per tile: partial logits are transpose-reduced with warp
//            shuffles (~N shuffles for N values instead of 5N), finished
//            through smem, soft-maxed by 16-lane groups (loss, top-1, dlogits),
//            then dW += dlogits^T x from the same registers.  x is read once.
//   grid barrier
//   phase 2  The 1963 float4 of the flat gradient are spread over all CTAs.
//            Each thread sums the G partials in CTA order (deterministic),
//            exchanges with the peers (LL protocol: 8-byte {value, flag}
//            units over NVLink -- published in this rank's outbox and read
//            by the peers, or pushed into the peers' inboxes; see the knob
//            comment in pdm_net_train_step), sums in rank order so every
//            rank gets bit-identical parameters, scales by 1/world and
//            applies Adam / SGD-momentum in registers.
// Metrics never leave the device unless asked: {loss*B, correct, B} are
// accumulated into a device buffer, and optionally {mean loss, correct} of
// this step is written to step_out (which may be mapped host memory).
//
// tcgen05 note: with N = 10 the forward/wgrad GEMMs are
// 2 MFLOP at B = 64; a UMMA tile would be 94% padding and the step is bound
// by launch + two memory round trips, not math.  The FMA form above is the
// measured-fastest form for this model; the tcgen05/TMA machinery lives in
// gemm_tcgen05.cu and serves the CNN's GEMM-shaped layers.
#include "common.cuh"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>

namespace {

constexpr int kThreads = 256;
constexpr int kIn = PDM_NET_IN;          // 784
constexpr int kOut = PDM_NET_OUT;        // 10
constexpr int kCols4 = kIn / 4;          // 196 float4 columns
constexpr int kFlat4 = PDM_NET_FLAT / 4; // 1963 float4 in the flat parameter vector
constexpr int kStride = PDM_NET_SCRATCH_STRIDE;
constexpr int kFwdWarps = 7;             // warps that own columns (7*32 = 224 >= 196)
constexpr float kMean = 0.1307f;
constexpr float kStd = 0.3081f;

struct OptConsts {
  int opt;
  int first;        // SGD: step == 1 seeds the momentum buffer
  float lr, beta1, beta2, eps, wd;
  float step_size;  // Adam: lr / (1 - beta1^t)
  float bc2_sqrt;   // Adam: sqrt(1 - beta2^t)
  float inv_world;
};

struct KernelArgs {
  const void* x;
  const void* y;
  int y_dtype;
  int batch;
  int rows_per_cta;
  int active;       // CTAs that own rows; the rest only help in phase 2
  float inv_batch;
  float* params;
  float* m;
  float* v;
  float* grad_out;
  float* scratch;
  uint32_t* sync;
  double* metrics;
  float* step_out;
  float* logits_out;
  long long* trace;   // optional [gridDim.x][8] clock64 stamps (bring-up / profiling aid)
  const int* row_index;   // optional [batch]: batch row r is row row_index[r] of x / y (in-kernel gather)
  int flag_dep;           // 1: bump sync[2] when a CTA is done (step-to-step flag dependency in use)
  int flag_wait;          // 1: wait for the previous step through sync[2] instead of griddepcontrol.wait
  uint32_t done_before;   // value sync[2] reaches once every CTA of every earlier launch has finished
  int trigger_early;      // 1: release the next step's launch as soon as this grid is resident
  int poll_ns;            // back-off between polls of sync[2]
  int ll_pull;            // 1: publish into our own outbox and read the peers' outboxes over NVLink
  uint32_t skew_ns;       // fault injection (PDM_SKEW_NS): random pause before the LL publish
  int inputs_stable;      // 1: x / y / row_index were not written by the kernel just before this one on the
                          //    stream, so the first tile may be loaded before the dependency wait
  OptConsts oc;
  pdm_peers peers;
};

__device__ __forceinline__ int load_label(const void* y, int dtype, int row) {
  int v;
  if (dtype == PDM_U8) v = ((const uint8_t*)y)[row];
  else if (dtype == PDM_I64) v = (int)((const long long*)y)[row];
  else v = ((const int*)y)[row];
  return min(max(v, 0), kOut - 1);
}

// Reduce N per-lane values across the 32 lanes of a warp with N-ish shuffles.
// On return lane L holds in v[i] (i < N/32) the warp total of value L*(N/32)+i.
template <int N>
__device__ __forceinline__ void warp_transpose_reduce(float (&v)[N], int lane) {
  static_assert(N % 32 == 0, "N must be a multiple of 32");
#pragma unroll
  for (int off = 16, h = N / 2; off >= 1; off >>= 1, h >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < h; ++i) {
      const float keep = upper ? v[i + h] : v[i];
      const float send = upper ? v[i] : v[i + h];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
}

template <typename XT> struct RawRow;
template <> struct RawRow<uint8_t> {
  using type = uint32_t;
  static __device__ __forceinline__ type load(const void* x, int row, int t) {
    return __ldg(reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(x) + (size_t)row * kIn) + t);
  }
  static __device__ __forceinline__ void decode(type raw, const float* lut, float (&o)[4]) {
    o[0] = lut[raw & 255u];
    o[1] = lut[(raw >> 8) & 255u];
    o[2] = lut[(raw >> 16) & 255u];
    o[3] = lut[raw >> 24];
  }
};
template <> struct RawRow<float> {
  using type = float4;
  static __device__ __forceinline__ type load(const void* x, int row, int t) {
    return __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(x) + (size_t)row * kIn) + t);
  }
  static __device__ __forceinline__ void decode(type raw, const float*, float (&o)[4]) {
    o[0] = raw.x; o[1] = raw.y; o[2] = raw.z; o[3] = raw.w;
  }
};

template <int RT>
struct Smem {
  static constexpr int kVals = RT * kOut;                 // logits per tile
  static constexpr int kPad = ((kVals + 31) / 32) * 32;   // padded for the transpose-reduce
  float lut[256];
  float red[kFwdWarps][kPad];
  float dl[RT][12];
  int ys[RT];
  float bias[12];
  float wloss[8];
  float wcorrect[8];
};

// Apply one optimizer update to a float4 lane group.
__device__ __forceinline__ void opt_update(const OptConsts& oc, float g, float& p, float& m, float& v) {
  if (oc.opt == PDM_OPT_ADAM) {
    if (oc.wd != 0.f) g = fmaf(oc.wd, p, g);
    m = m + (1.f - oc.beta1) * (g - m);                  // exp_avg.lerp_(grad, 1-beta1)
    v = v * oc.beta2 + (1.f - oc.beta2) * g * g;         // mul_(beta2).addcmul_(g, g, 1-beta2)
    const float denom = sqrtf(v) / oc.bc2_sqrt + oc.eps;
    p = p - oc.step_size * (m / denom);                  // addcdiv_(exp_avg, denom, -step_size)
  } else if (oc.opt == PDM_OPT_SGD) {
    if (oc.wd != 0.f) g = fmaf(oc.wd, p, g);
    if (oc.beta1 != 0.f) {
      m = oc.first ? g : (m * oc.beta1 + g);
      g = m;
    }
    p = p - oc.lr * g;
  }
}

// MODE 0: train (phase 1 + barrier + phase 2).  MODE 1: eval/forward only.
template <int RT, typename XT, int MODE>
__global__ void __launch_bounds__(kThreads, 1) net_kernel(const __grid_constant__ KernelArgs a) {
  using Raw = RawRow<XT>;
  using SM = Smem<RT>;
  __shared__ SM sm;

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  const int cta = blockIdx.x;
  const int G = gridDim.x;
  const bool owns_cols = tid < kCols4;
  // PDL secondary launches once EVERY CTA of this grid has executed launch_dependents, i.e. once the
  // whole grid is resident -- exactly the condition under which the successor may take SMs without
  // starving this grid's barrier.  Only with the flag dependency: the successor then orders itself
  // after this step through sync[2], not through grid completion.
  if (MODE == 0 && a.trigger_early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#define PDM_TRACE(slot) do { if (MODE == 0 && a.trace != nullptr && tid == 0) a.trace[cta * 8 + (slot)] = clock64(); } while (0)
  PDM_TRACE(0);

  // ---- one-time per CTA: LUT, bias, this thread's weight columns ----------
  if (sizeof(XT) == 1) {
    // Bit-exact torchvision ToTensor (/255) + Normalize ((x-mean)/std), ref:132-135.
    sm.lut[tid] = __fdiv_rn(__fdiv_rn((float)tid, 255.0f) - kMean, kStd);
  }
  const bool has_rows = MODE == 1 || cta < a.active;
  float db = 0.f;          // threads 196..205: bias gradient of class tid-196
  float my_loss = 0.f;     // softmax threads with c == 0
  float my_correct = 0.f;

  const int row_begin = cta * a.rows_per_cta;
  const int row_end = min(a.batch, row_begin + a.rows_per_cta);

  // Inputs the caller declared stable (a memcpy, a resident dataset: nothing the preceding kernel
  // wrote) do not depend on the previous step: start streaming them before waiting for it.  Otherwise
  // the first tile is loaded after the wait below -- under a programmatic dependent launch only that
  // wait makes the predecessor's writes visible (e.g. a normalise or gather kernel producing x).
  typename Raw::type raw[RT];
  auto load_first_tile = [&]() {
#pragma unroll
    for (int r = 0; r < RT; ++r) {
      raw[r] = typename Raw::type();
      if (owns_cols && row_begin + r < row_end)
        raw[r] = Raw::load(a.x, a.row_index ? __ldg(a.row_index + row_begin + r) : row_begin + r, tid);
    }
  };
  const bool early_inputs = MODE == 1 || a.inputs_stable != 0;
  if (early_inputs) load_first_tile();

  // Programmatic dependent launch: everything above overlaps the tail of the previous step's
  // kernel; parameters, optimizer state, scratch and barrier words are only touched below.
  // (A no-op when the kernel was not launched with the PDL attribute.)
  //
  // Two ways to wait for the previous step.  griddepcontrol.wait waits for that grid to be
  // complete *and flushed*, which with peers includes the acknowledgement of its remote LL
  // stores (measured ~4-5 us per step at 2 GPUs).  The flag form only waits until all of its
  // CTAs have finished phase 2 (each bumps sync[2] after its last local store, release/acquire
  // at gpu scope), so remote traffic of step n drains under step n+1.
  if (MODE == 0) {
    if (a.flag_wait) {
      if (tid == 0) {
        while ((int32_t)(ld_acquire_gpu(a.sync + 2) - a.done_before) < 0) {
          if (a.poll_ns > 0) __nanosleep(a.poll_ns);
        }
      }
      __syncthreads();
    } else {
      asm volatile("griddepcontrol.wait;" ::: "memory");
    }
  }
  if (!early_inputs) load_first_tile();

  uint32_t bar_base = 0;
  if (MODE == 0 && tid == 0 && G > 1) bar_base = ld_acquire_gpu(a.sync + 1);   // latency hidden by phase 1
  if (tid < 12) sm.bias[tid] = tid < kOut ? a.params[kIn * kOut + tid] : 0.f;

  float w[kOut][4];
  float g[kOut][4];
#pragma unroll
  for (int c = 0; c < kOut; ++c) {
#pragma unroll
    for (int i = 0; i < 4; ++i) { w[c][i] = 0.f; g[c][i] = 0.f; }
  }
  if (owns_cols && has_rows) {
#pragma unroll
    for (int c = 0; c < kOut; ++c) {
      const float4 t4 = *reinterpret_cast<const float4*>(a.params + c * kIn + 4 * tid);
      w[c][0] = t4.x; w[c][1] = t4.y; w[c][2] = t4.z; w[c][3] = t4.w;
    }
  }
  __syncthreads();
  PDM_TRACE(1);

  for (int row0 = row_begin; row0 < row_end; row0 += RT) {
    const int nrows = min(RT, row_end - row0);
    if (a.y != nullptr && tid < RT)
      sm.ys[tid] = tid < nrows ? load_label(a.y, a.y_dtype, a.row_index ? __ldg(a.row_index + row0 + tid) : row0 + tid) : 0;

    // decode this tile, prefetch the next one
    float xr[RT][4];
#pragma unroll
    for (int r = 0; r < RT; ++r) Raw::decode(raw[r], sm.lut, xr[r]);
#pragma unroll
    for (int r = 0; r < RT; ++r) {
      const int nr = row0 + RT + r;
      if (owns_cols && nr < row_end) raw[r] = Raw::load(a.x, a.row_index ? __ldg(a.row_index + nr) : nr, tid);
    }

    // ---- forward partials: v[r*10+c] = sum_i x[r][i] * W[c][i] ------------
    if (warp < kFwdWarps) {
      float v[SM::kPad];
#pragma unroll
      for (int r = 0; r < RT; ++r) {
#pragma unroll
        for (int c = 0; c < kOut; ++c) {
          float s = xr[r][0] * w[c][0];
          s = fmaf(xr[r][1], w[c][1], s);
          s = fmaf(xr[r][2], w[c][2], s);
          s = fmaf(xr[r][3], w[c][3], s);
          v[r * kOut + c] = owns_cols ? s : 0.f;
        }
      }
#pragma unroll
      for (int i = SM::kVals; i < SM::kPad; ++i) v[i] = 0.f;
      warp_transpose_reduce<SM::kPad>(v, lane);
      constexpr int per = SM::kPad / 32;
#pragma unroll
      for (int i = 0; i < per; ++i) sm.red[warp][lane * per + i] = v[i];
    }
    __syncthreads();

    // ---- softmax / loss / top-1 / dlogits: 16-lane group per row ----------
    if (tid < RT * 16) {
      const int r = tid >> 4;
      const int c = tid & 15;
      const bool live = c < kOut;
      float z = -INFINITY;
      if (live) {
        z = sm.bias[c];
#pragma unroll
        for (int wv = 0; wv < kFwdWarps; ++wv) z += sm.red[wv][r * kOut + c];
        if (a.logits_out != nullptr && r < nrows) a.logits_out[(size_t)(row0 + r) * kOut + c] = z;
      }
      if (a.y != nullptr) {
        float mx = z;
#pragma unroll
        for (int off = 8; off >= 1; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        const float e = live ? expf(z - mx) : 0.f;
        float s = e;
#pragma unroll
        for (int off = 8; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        const int y = sm.ys[r];
        const float zy = __shfl_sync(0xffffffffu, z, (lane & 16) + y);
        const bool beats = live && (z > zy || (z == zy && c < y));
        const uint32_t ballot = __ballot_sync(0xffffffffu, beats);
        const bool hit = ((ballot >> (lane & 16)) & 0xffffu) == 0u;
        const bool valid = r < nrows;
        if (MODE == 0) {
          float d = 0.f;
          if (valid && live) d = (e / s - (c == y ? 1.f : 0.f)) * a.inv_batch;
          if (c < 12) sm.dl[r][c] = d;
        }
        if (c == 0 && valid) {
          my_loss += (mx + logf(s)) - zy;
          my_correct += hit ? 1.f : 0.f;
        }
      }
    }
    if (MODE == 0) {
      __syncthreads();
      // ---- backward: dW[:, cols] += dlogits^T x, db += colsum(dlogits) -----
      if (owns_cols) {
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          const float4 d0 = *reinterpret_cast<const float4*>(&sm.dl[r][0]);
          const float4 d1 = *reinterpret_cast<const float4*>(&sm.dl[r][4]);
          const float4 d2 = *reinterpret_cast<const float4*>(&sm.dl[r][8]);
          const float d[kOut] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x, d2.y};
#pragma unroll
          for (int c = 0; c < kOut; ++c) {
#pragma unroll
            for (int i = 0; i < 4; ++i) g[c][i] = fmaf(d[c], xr[r][i], g[c][i]);
          }
        }
      } else if (tid < kCols4 + kOut) {
#pragma unroll
        for (int r = 0; r < RT; ++r) db += sm.dl[r][tid - kCols4];
      }
    } else {
      __syncthreads();  // red/ys are rewritten by the next tile
    }
  }

  // ---- CTA totals of loss / correct ----------------------------------------
  if (a.y != nullptr) {
    // only lanes 0 and 16 of warps < RT*16/32 carry values
    float l = my_loss + __shfl_xor_sync(0xffffffffu, my_loss, 16);
    float k = my_correct + __shfl_xor_sync(0xffffffffu, my_correct, 16);
    if (lane == 0) { sm.wloss[warp] = l; sm.wcorrect[warp] = k; }
  }
  __syncthreads();

  if (MODE == 1) {
    if (a.y != nullptr && tid == 0) {
      float l = 0.f, k = 0.f;
      constexpr int nw = (RT * 16 + 31) / 32;
#pragma unroll
      for (int i = 0; i < nw; ++i) { l += sm.wloss[i]; k += sm.wcorrect[i]; }
      const int rows = max(0, row_end - row_begin);
      if (a.metrics != nullptr && rows > 0) {
        atomicAdd(a.metrics + 0, (double)l);
        atomicAdd(a.metrics + 1, (double)k);
        atomicAdd(a.metrics + 2, (double)rows);
      }
    }
    return;
  }

  // ---- publish this CTA's partial gradient ---------------------------------
  float* mine = a.scratch + (size_t)cta * kStride;
  if (!has_rows) {
    // helper CTA: nothing to publish
  } else if (owns_cols) {
#pragma unroll
    for (int c = 0; c < kOut; ++c)
      __stcg(reinterpret_cast<float4*>(mine + c * kIn + 4 * tid), make_float4(g[c][0], g[c][1], g[c][2], g[c][3]));
  } else if (tid < kCols4 + kOut) {
    __stcg(mine + kIn * kOut + (tid - kCols4), db);
  } else if (tid == kCols4 + kOut) {
    float l = 0.f, k = 0.f;
    constexpr int nw = (RT * 16 + 31) / 32;
#pragma unroll
    for (int i = 0; i < nw; ++i) { l += sm.wloss[i]; k += sm.wcorrect[i]; }
    __stcg(mine + PDM_NET_PARAMS, l);
    __stcg(mine + PDM_NET_PARAMS + 1, k);
  }

  PDM_TRACE(2);
  // ---- phase 2 set-up that does not need the barrier: which float4 this lane group owns, and
  //      its parameter / optimizer-state loads (written last by the previous launch) ----------
  const int per_cta = (kFlat4 + G - 1) / G;
  const int j_begin = cta * per_cta;
  const int j_count = max(0, min(kFlat4, j_begin + per_cta) - j_begin);
  const int W = a.peers.world > 1 ? a.peers.world : 1;
  const uint32_t seq = a.peers.seq;
  const int sub = tid & 7;
  const bool has_opt = a.oc.opt != PDM_OPT_NONE;
  float4 p4_pre = make_float4(0.f, 0.f, 0.f, 0.f), m4_pre = p4_pre, v4_pre = p4_pre;
  if (has_opt && sub == 0 && (tid >> 3) < j_count) {
    const int j = j_begin + (tid >> 3);
    p4_pre = *reinterpret_cast<const float4*>(a.params + 4 * j);
    if (a.m != nullptr) m4_pre = *reinterpret_cast<const float4*>(a.m + 4 * j);
    if (a.v != nullptr) v4_pre = *reinterpret_cast<const float4*>(a.v + 4 * j);
  }

  // ---- grid barrier: sync[0] counts arrivals (never reset), sync[1] holds the count at which
  //      this launch started.  Arrival is a fire-and-forget red; only the polls are round trips.
  if (G > 1) {
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(a.sync) : "memory");
      const uint32_t target = bar_base + (uint32_t)G;
      while ((int32_t)(ld_acquire_gpu(a.sync) - target) < 0) { }
      if (cta == 0) a.sync[1] = target;          // base of the next launch on these words
      __threadfence();
    }
    __syncthreads();
  } else {
    __syncthreads();
  }
  // Let the next step's kernel start its input prefetch (PDL); it still waits for this grid to
  // finish before it touches parameters or scratch.
  if (!a.trigger_early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  PDM_TRACE(3);

  // ---- phase 2: reduce partials, all-reduce with peers, optimizer ----------
  // 8 lanes share one float4 of the flat gradient: lane l sums partials l, l+8, ... (loads all in
  // flight: one L2 round trip), then a fixed xor tree combines the 8 lanes, so the summation order
  // depends only on the CTA count -- deterministic run to run.
  for (int base = 0; base < j_count; base += kThreads / 8) {
    const int jj = base + (tid >> 3);
    const bool valid = jj < j_count;
    const int j = j_begin + (valid ? jj : 0);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      const float* src = a.scratch + 4 * j;
      for (int c0 = sub; c0 < a.active; c0 += 64) {
        float4 t[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = c0 + 8 * u;
          t[u] = c < a.active ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)c * kStride))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
      }
    }
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) {
      acc.x += __shfl_xor_sync(0xffffffffu, acc.x, off);
      acc.y += __shfl_xor_sync(0xffffffffu, acc.y, off);
      acc.z += __shfl_xor_sync(0xffffffffu, acc.z, off);
      acc.w += __shfl_xor_sync(0xffffffffu, acc.w, off);
    }
    // after the xor tree every lane of the group holds the same total
    if (valid && j == kFlat4 - 1) {
      // lanes z, w of the last float4 are this rank's {sum of losses, correct}
      if (sub == 0) {
        if (a.step_out != nullptr) {
          a.step_out[0] = acc.z * a.inv_batch;
          a.step_out[1] = acc.w;
        }
        if (a.metrics != nullptr) {
          a.metrics[0] += (double)acc.z;
          a.metrics[1] += (double)acc.w;
          a.metrics[2] += (double)a.batch;
        }
      }
      acc.z = 0.f;
      acc.w = 0.f;
    }

    if (W > 1) {
      // The 8 lanes of a group split the peers: lane `sub` pushes this float4 to rank `sub` and
      // polls rank `sub`'s slot in our inbox, so the per-peer stores and waits run side by side
      // instead of one lane walking all peers.  The sum is then formed in rank order from the
      // lanes' values, identically on every rank.
      const size_t par = (size_t)(seq & 1u) * W;
      const bool remote = valid && sub < W && sub != a.peers.rank;
      pdm_skew(a.skew_ns, seq, (uint32_t)(a.peers.rank * 256 + cta), (uint32_t)(tid >> 5));
      // push (default): write our value into slot `rank` of every peer's inbox, poll our own inbox.
      // pull (PDM_NET_LL=pull): write it once into slot `rank` of OUR inbox and poll slot `sub` of
      // peer `sub`'s inbox over NVLink: a round trip per poll instead of a one-way store, but the
      // kernel retires with no remote stores in flight, so the next step does not wait for their
      // acknowledgements.
      if (a.ll_pull) {
        if (valid && sub == a.peers.rank) {
          uint64_t* dst = a.peers.inbox[a.peers.rank] + ((par + a.peers.rank) * kFlat4 + j) * 4;
          ll_store2(dst, acc.x, acc.y, seq);
          ll_store2(dst + 2, acc.z, acc.w, seq);
        }
      } else if (remote) {
        uint64_t* dst = a.peers.inbox[sub] + ((par + a.peers.rank) * kFlat4 + j) * 4;
        ll_store2(dst, acc.x, acc.y, seq);
        ll_store2(dst + 2, acc.z, acc.w, seq);
      }
      float4 q = (valid && sub == a.peers.rank) ? acc : make_float4(0.f, 0.f, 0.f, 0.f);
      if (remote) {
        const uint64_t* src = a.peers.inbox[a.ll_pull ? sub : a.peers.rank] + ((par + sub) * kFlat4 + j) * 4;
        bool ok0 = false, ok1 = false;
        uint64_t t0 = 0;
        uint32_t spins = 0;
        while (true) {
          if (!ok0) ok0 = ll_load2(src, seq, q.x, q.y);
          if (!ok1) ok1 = ll_load2(src + 2, seq, q.z, q.w);
          if (ok0 && ok1) break;
          if ((++spins & 1023u) == 0u && a.peers.timeout_ns != 0) {
            const uint64_t now = globaltimer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > a.peers.timeout_ns) {
              if (a.peers.status != nullptr) atomicExch(a.peers.status, 1 + sub);
              break;
            }
          }
        }
      }
      float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
      const int group_base = lane & ~7;
#pragma unroll
      for (int p = 0; p < PDM_MAX_RANKS; ++p) {
        const float tx = __shfl_sync(0xffffffffu, q.x, group_base + p);
        const float ty = __shfl_sync(0xffffffffu, q.y, group_base + p);
        const float tz = __shfl_sync(0xffffffffu, q.z, group_base + p);
        const float tw = __shfl_sync(0xffffffffu, q.w, group_base + p);
        if (p < W) { sum.x += tx; sum.y += ty; sum.z += tz; sum.w += tw; }
      }
      acc = sum;
    }
    if (!valid || sub != 0) continue;

    const float iw = a.oc.inv_world;
    acc.x *= iw; acc.y *= iw; acc.z *= iw; acc.w *= iw;
    if (a.grad_out != nullptr) *reinterpret_cast<float4*>(a.grad_out + 4 * j) = acc;

    if (has_opt) {
      float4 p4 = p4_pre, m4 = m4_pre, v4 = v4_pre;
      if (base != 0) {          // only the first pass was prefetched before the barrier
        p4 = *reinterpret_cast<float4*>(a.params + 4 * j);
        if (a.m != nullptr) m4 = *reinterpret_cast<float4*>(a.m + 4 * j);
        if (a.v != nullptr) v4 = *reinterpret_cast<float4*>(a.v + 4 * j);
      }
      opt_update(a.oc, acc.x, p4.x, m4.x, v4.x);
      opt_update(a.oc, acc.y, p4.y, m4.y, v4.y);
      if (j != kFlat4 - 1) {  // the last two floats are padding, not parameters
        opt_update(a.oc, acc.z, p4.z, m4.z, v4.z);
        opt_update(a.oc, acc.w, p4.w, m4.w, v4.w);
      }
      *reinterpret_cast<float4*>(a.params + 4 * j) = p4;
      if (a.m != nullptr) *reinterpret_cast<float4*>(a.m + 4 * j) = m4;
      if (a.v != nullptr) *reinterpret_cast<float4*>(a.v + 4 * j) = v4;
    }
  }
  if (a.flag_dep || a.trace != nullptr) __syncthreads();
  if (a.flag_dep && tid == 0) {
    // this CTA is done with parameters, optimizer state, scratch and inbox: tell the next step
    __threadfence();
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(a.sync + 2) : "memory");
  }
  PDM_TRACE(4);
#undef PDM_TRACE
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
constexpr int kRT = 8;    // rows per tile of the evaluation kernel and default of the training kernel

// Rows per tile of the training kernel.  4 for small batches (twice the CTAs, half the phase-1
// work per CTA: measured 8.75 vs 9.06 us/step at B = 512), 8 once every SM has several tiles
// anyway.  PDM_NET_RT=4|8 forces one.
int train_rt(int batch) {
  static const int forced = [] {
    const char* e = getenv("PDM_NET_RT");
    return (e && *e == '4') ? 4 : ((e && *e == '8') ? 8 : 0);
  }();
  if (forced) return forced;
  return batch <= 1024 ? 4 : 8;
}
// How the training kernel is launched (PDM_NET_LAUNCH):
//   pdl   (default) programmatic dependent launch: the next step's input prefetch overlaps this
//         step's tail (measured 8.75 vs 10.5 us/step).  Not a cooperative launch; the grid barrier
//         is still deadlock-free because (a) the grid never exceeds one CTA per SM of the device
//         (checked at launch, else the cooperative path is taken), (b) the only kernels it can
//         overlap are its predecessor, which finishes without it, and its successor, which is
//         released only after this grid's barrier, i.e. once every CTA is resident, and (c) no
//         other kernel of this library waits on it while holding SMs
//   coop  cooperative launch: co-residency of all CTAs is guaranteed (or refused) by the driver
//   plain ordinary launch (measurement only)
enum { LAUNCH_COOP = 0, LAUNCH_PDL = 1, LAUNCH_PLAIN = 2 };
int launch_mode() {
  static const int mode = [] {
    const char* e = getenv("PDM_NET_LAUNCH");
    if (e && !strcmp(e, "coop")) return (int)LAUNCH_COOP;
    if (e && !strcmp(e, "plain")) return (int)LAUNCH_PLAIN;
    return (int)LAUNCH_PDL;
  }();
  return mode;
}

OptConsts make_consts(const pdm_hyper& h, int world) {
  OptConsts oc;
  memset(&oc, 0, sizeof(oc));
  oc.opt = h.opt;
  oc.first = h.step <= 1;
  oc.lr = h.lr;
  oc.beta1 = h.beta1;
  oc.beta2 = h.beta2;
  oc.eps = h.eps;
  oc.wd = h.weight_decay;
  if (h.opt == PDM_OPT_ADAM) {
    const double step = (double)(h.step < 1 ? 1 : h.step);
    const double bc1 = 1.0 - pow((double)h.beta1, step);
    const double bc2 = 1.0 - pow((double)h.beta2, step);
    oc.step_size = (float)((double)h.lr / bc1);
    oc.bc2_sqrt = (float)sqrt(bc2);
  }
  oc.inv_world = world > 1 ? 1.0f / (float)world : 1.0f;
  return oc;
}

// ctas = CTAs launched; active = those that own a slice of the batch (ceil(batch / rows_per_cta)).
// Small batches still launch kMinCtas CTAs: the extra ones skip phase 1 and shorten phase 2,
// where the 1963 float4 of the gradient are divided among ALL launched CTAs.
constexpr int kMinCtas = 64;
void plan_rows(int batch, int want_ctas, int* ctas, int* active, int* rows_per_cta) {
  const int rt = train_rt(batch);
  int g = want_ctas > 0 ? want_ctas : (batch + rt - 1) / rt;
  if (want_ctas <= 0 && g < kMinCtas) g = kMinCtas;
  if (g > PDM_NET_MAX_CTAS) g = PDM_NET_MAX_CTAS;
  if (g < 1) g = 1;
  int rpc = (batch + g - 1) / g;
  if (rpc < rt && batch >= rt) rpc = rt;        // never split a tile's worth of rows further
  if (rpc < 1) rpc = 1;
  int act = (batch + rpc - 1) / rpc;
  if (act < 1) act = 1;
  *ctas = g;
  *active = act;
  *rows_per_cta = rpc;
}

// PDM_NET_TRACE=1: every training launch stamps clock64 at five points per CTA into a buffer
// that pdm_net_trace_read copies out (used by benchmarks/step_trace.py; off by default).
long long* g_trace_buf = nullptr;
long long* trace_buffer() {
  static const bool on = [] { const char* e = getenv("PDM_NET_TRACE"); return e && *e == '1'; }();
  if (!on) return nullptr;
  if (g_trace_buf == nullptr) {
    if (cudaMalloc((void**)&g_trace_buf, PDM_NET_MAX_CTAS * 8 * sizeof(long long)) != cudaSuccess) return nullptr;
    cudaMemset(g_trace_buf, 0, PDM_NET_MAX_CTAS * 8 * sizeof(long long));
  }
  return g_trace_buf;
}

template <int MODE>
cudaError_t launch(const KernelArgs& ka, int x_dtype, int ctas, cudaStream_t stream, bool allow_trace = true) {
  KernelArgs local = ka;
  // allow_trace == false: the allocation-free entry point.  trace_buffer() allocates on first use,
  // so that path never traces, whatever PDM_NET_TRACE says.
  if (MODE == 0) local.trace = allow_trace ? trace_buffer() : nullptr;
  void* params[] = {&local};
  const void* fn;
  if (MODE == 0 && train_rt(ka.batch) == 4)
    fn = x_dtype == PDM_U8 ? (const void*)net_kernel<4, uint8_t, MODE> : (const void*)net_kernel<4, float, MODE>;
  else
    fn = x_dtype == PDM_U8 ? (const void*)net_kernel<kRT, uint8_t, MODE> : (const void*)net_kernel<kRT, float, MODE>;
  if (MODE != 0) return cudaLaunchKernel(fn, dim3(ctas), dim3(kThreads), params, 0, stream);
  int mode = launch_mode();
  if (mode != LAUNCH_COOP) {
    // without the driver's co-residency guarantee the grid must fit one CTA per SM
    static int sm_count[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && sm_count[dev] == 0) cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (dev < 0 || dev >= 64 || ctas > sm_count[dev]) mode = LAUNCH_COOP;
  }
  if (mode == LAUNCH_PDL) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(ctas);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelExC(&cfg, fn, params);
  }
  if (mode == LAUNCH_COOP && ctas > 1)
    return cudaLaunchCooperativeKernel(fn, dim3(ctas), dim3(kThreads), params, 0, stream);
  return cudaLaunchKernel(fn, dim3(ctas), dim3(kThreads), params, 0, stream);
}

// Protocol knobs shared by both training entry points (so that ranks which mix them still agree
// on the exchange form): PDM_NET_DEP=grid|flag, PDM_NET_LL=push|pull, PDM_NET_TRIGGER=barrier|start,
// PDM_NET_POLL_NS=<n>.
// Measured on B200 (profiles/raw/step_sweep_dep.log, step_sweep_multi_ll_2.log), us/step at B = 512:
//   1 GPU : grid 8.7, flag 10.3                      -> grid
//   2 GPUs: push+grid 17.8, push+flag 16.0, pull+grid 15.3, pull+flag 12.6  -> pull + flag
// With peers, whatever waits for this grid's memory traffic to drain (grid completion, or the
// fence before the sync[2] bump) also waits for the acknowledgement of pushed remote stores;
// pulled data leaves nothing in flight.  Unset knobs therefore follow the world size.
void apply_knobs(KernelArgs* ka, const pdm_net_step_args* args, int world, bool allow_flag_dependency) {
  static const int dep_env = [] { const char* e = getenv("PDM_NET_DEP"); return !e ? -1 : (!strcmp(e, "flag") ? 1 : 0); }();
  static const int pull_env = [] { const char* e = getenv("PDM_NET_LL"); return !e ? -1 : (!strcmp(e, "pull") ? 1 : 0); }();
  static const int early = [] { const char* e = getenv("PDM_NET_TRIGGER"); return (e && !strcmp(e, "barrier")) ? 0 : 1; }();
  static const int poll = [] { const char* e = getenv("PDM_NET_POLL_NS"); return e && *e ? atoi(e) : 0; }();
  const int dep_flag = dep_env >= 0 ? dep_env : (world > 1 ? 1 : 0);
  // flag_dependency 1: count this launch's CTAs in sync[2] and wait for the earlier ones through it;
  // 2: count, but wait with griddepcontrol.wait (first step of a run: whatever ran before on the
  // stream may not have been counted)
  ka->flag_dep = (allow_flag_dependency && args->flag_dependency && dep_flag && launch_mode() == LAUNCH_PDL) ? 1 : 0;
  ka->flag_wait = (ka->flag_dep && args->flag_dependency == 1) ? 1 : 0;
  ka->done_before = args->done_before;
  ka->trigger_early = (ka->flag_wait && early) ? 1 : 0;
  ka->poll_ns = poll;
  ka->ll_pull = pull_env >= 0 ? pull_env : (world > 1 ? 1 : 0);
  ka->skew_ns = pdm_skew_ns();
}

}  // namespace

extern "C" int pdm_net_default_ctas(int batch) {
  int g, act, rpc;
  plan_rows(batch, 0, &g, &act, &rpc);
  return g;
}

extern "C" int pdm_net_planned_ctas(int batch, int want_ctas) {
  int g, act, rpc;
  plan_rows(batch, want_ctas, &g, &act, &rpc);
  return g;
}

extern "C" int pdm_net_train_step(const pdm_net_step_args* args, pdm_stream_t stream) {
  PDM_REQUIRE(args != nullptr, PDM_ERR_ARG, "pdm_net_train_step: args is NULL");
  PDM_REQUIRE(args->batch >= 1, PDM_ERR_ARG, "pdm_net_train_step: batch must be >= 1 (got %d)", args->batch);
  PDM_REQUIRE(args->x && args->y && args->params && args->scratch && args->sync, PDM_ERR_ARG,
              "pdm_net_train_step: x, y, params, scratch and sync are required");
  PDM_REQUIRE(args->x_dtype == PDM_U8 || args->x_dtype == PDM_F32, PDM_ERR_ARG,
              "pdm_net_train_step: x_dtype must be PDM_U8 or PDM_F32");
  PDM_REQUIRE(args->y_dtype == PDM_U8 || args->y_dtype == PDM_I64 || args->y_dtype == PDM_I32, PDM_ERR_ARG,
              "pdm_net_train_step: y_dtype must be PDM_U8, PDM_I64 or PDM_I32");
  PDM_REQUIRE(args->ctas >= 0 && args->ctas <= PDM_NET_MAX_CTAS, PDM_ERR_LIMIT,
              "pdm_net_train_step: ctas %d outside [0, %d]", args->ctas, PDM_NET_MAX_CTAS);
  const pdm_hyper& h = args->hyper;
  PDM_REQUIRE(h.opt == PDM_OPT_ADAM || h.opt == PDM_OPT_SGD || h.opt == PDM_OPT_NONE, PDM_ERR_ARG,
              "pdm_net_train_step: unknown optimizer %d", h.opt);
  PDM_REQUIRE(h.opt != PDM_OPT_ADAM || (args->m && args->v), PDM_ERR_ARG,
              "pdm_net_train_step: Adam needs m and v");
  PDM_REQUIRE(!(h.opt == PDM_OPT_SGD && h.beta1 != 0.f) || args->m, PDM_ERR_ARG,
              "pdm_net_train_step: SGD with momentum needs m");
  const int world = args->peers.world > 1 ? args->peers.world : 1;
  PDM_REQUIRE(world <= PDM_MAX_RANKS, PDM_ERR_LIMIT, "pdm_net_train_step: world %d > %d", world, PDM_MAX_RANKS);
  if (world > 1) {
    PDM_REQUIRE(args->peers.seq != 0, PDM_ERR_ARG, "pdm_net_train_step: peers.seq must be non-zero");
    PDM_REQUIRE(args->peers.rank >= 0 && args->peers.rank < world, PDM_ERR_ARG, "pdm_net_train_step: bad rank");
    for (int p = 0; p < world; ++p)
      PDM_REQUIRE(args->peers.inbox[p] != nullptr, PDM_ERR_STATE, "pdm_net_train_step: inbox[%d] not mapped", p);
  }

  KernelArgs ka;
  memset(&ka, 0, sizeof(ka));
  int ctas = 0;
  plan_rows(args->batch, args->ctas, &ctas, &ka.active, &ka.rows_per_cta);
  ka.x = args->x;
  ka.y = args->y;
  ka.y_dtype = args->y_dtype;
  ka.batch = args->batch;
  ka.inv_batch = 1.0f / (float)args->batch;
  ka.params = args->params;
  ka.m = args->m;
  ka.v = args->v;
  ka.grad_out = args->grad_out;
  ka.scratch = args->scratch;
  ka.sync = args->sync;
  ka.metrics = args->metrics;
  ka.step_out = args->step_out;
  ka.row_index = args->row_index;
  apply_knobs(&ka, args, world, /*allow_flag_dependency=*/true);
  ka.inputs_stable = args->inputs_stable != 0;
  ka.oc = make_consts(h, world);
  ka.peers = args->peers;
  ka.peers.world = world;
  PDM_CHECK_CUDA(launch<0>(ka, args->x_dtype, ctas, (cudaStream_t)stream));
  return PDM_OK;
}

static int eval_like(const void* x, int x_dtype, const void* y, int y_dtype, int batch, const float* params,
                     float* logits_out, double* metrics, cudaStream_t stream, const char* who) {
  PDM_REQUIRE(batch >= 1, PDM_ERR_ARG, "%s: batch must be >= 1 (got %d)", who, batch);
  PDM_REQUIRE(x && params, PDM_ERR_ARG, "%s: x and params are required", who);
  PDM_REQUIRE(x_dtype == PDM_U8 || x_dtype == PDM_F32, PDM_ERR_ARG, "%s: x_dtype must be PDM_U8 or PDM_F32", who);
  KernelArgs ka;
  memset(&ka, 0, sizeof(ka));
  ka.x = x;
  ka.y = y;
  ka.y_dtype = y_dtype;
  ka.batch = batch;
  ka.rows_per_cta = kRT;
  int ctas = (batch + kRT - 1) / kRT;
  const int cap = PDM_NET_MAX_CTAS * 8;  // a few waves at most; then each CTA loops
  if (ctas > cap) {
    ka.rows_per_cta = ((batch + cap - 1) / cap + kRT - 1) / kRT * kRT;
    ctas = (batch + ka.rows_per_cta - 1) / ka.rows_per_cta;
  }
  ka.inv_batch = 1.0f / (float)batch;
  ka.params = const_cast<float*>(params);
  ka.metrics = metrics;
  ka.logits_out = logits_out;
  ka.oc.opt = PDM_OPT_NONE;
  PDM_CHECK_CUDA(launch<1>(ka, x_dtype, ctas, stream));
  return PDM_OK;
}

extern "C" int pdm_net_eval_step(const void* x, int x_dtype, const void* y, int y_dtype, int batch,
                                 const float* params, float* logits_out, double* metrics,
                                 pdm_stream_t stream) {
  PDM_REQUIRE(y != nullptr, PDM_ERR_ARG, "pdm_net_eval_step: y is required");
  return eval_like(x, x_dtype, y, y_dtype, batch, params, logits_out, metrics, (cudaStream_t)stream,
                   "pdm_net_eval_step");
}

extern "C" int pdm_net_forward(const void* x, int x_dtype, int batch, const float* params, float* logits,
                               pdm_stream_t stream) {
  PDM_REQUIRE(logits != nullptr, PDM_ERR_ARG, "pdm_net_forward: logits is required");
  return eval_like(x, x_dtype, nullptr, PDM_U8, batch, params, logits, nullptr, (cudaStream_t)stream,
                   "pdm_net_forward");
}

// ---------------------------------------------------------------------------
// pdm_net_train_step_static -- the same fused step for callers that must not allocate.
//
// pdm_net_train_step needs two pieces of device memory besides its inputs and outputs: the
// per-CTA partial-gradient scratch and the two grid-barrier words.  Some hosts of this
// library cannot supply them, because they cannot call cudaMalloc (or anything that does)
// where the step runs:
//   * a process that handed the whole HBM budget to its own pool at start-up and runs with a
//     "no device allocation after init" rule (serving and real-time loops: cudaMalloc and
//     cudaFree also synchronise the device, which stalls every other stream);
//   * code called back from inside another framework's allocator or with that allocator's
//     lock held, where re-entering an allocator deadlocks;
//   * recovery paths that run after an out-of-memory error, when an allocation would fail again.
// For them the scratch lives in this file as a fixed-size __device__ array that is part of the
// loaded module: every GPU that loads the library holds PDM_NET_STATIC_SCRATCH_BYTES (about
// 1.9 MiB) plus 256 bytes of barrier words until the process exits, whether or not this entry
// point is ever called.  The size fixes the limit: at most PDM_NET_STATIC_MAX_CTAS (64) CTAs;
// a larger request is refused with PDM_ERR_LIMIT.  Any batch size still works -- fewer CTAs
// just take more rows each.
//
// Set-up: the first call on each device resolves the device addresses of the arrays with
// cudaGetSymbolAddress (the array names are host-side handles in host code, not device
// pointers) while that device is current, and caches them.  Make that first call outside any
// stream capture.  No call here allocates or frees device or host memory (clock tracing,
// PDM_NET_TRACE, is therefore ignored on this path: its buffer is allocated on first use).
//
// Sharing: all calls on a device use the same array, which keeps the last call's partial
// gradients.  The step overwrites every element it later reads, and the barrier words -- the
// only part that must start at zero -- are cleared in-stream before each launch, so stale
// contents are harmless; but two calls that overlap on the same GPU (different streams,
// different host threads) race on the array and must be serialised by the caller.
// ---------------------------------------------------------------------------
namespace {

constexpr int kStaticMaxCtas = PDM_NET_STATIC_MAX_CTAS;
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) float g_static_scratch[(size_t)kStaticMaxCtas * PDM_NET_SCRATCH_STRIDE];
static __device__ __align__(256) uint32_t g_static_sync[64];

static_assert(sizeof(float) * kStaticMaxCtas * PDM_NET_SCRATCH_STRIDE == PDM_NET_STATIC_SCRATCH_BYTES,
              "PDM_NET_STATIC_SCRATCH_BYTES is out of date");

struct StaticAddrs {
  float* scratch = nullptr;
  uint32_t* sync = nullptr;
  bool ready = false;
};
StaticAddrs g_static_addrs[kStaticMaxDevices];
pthread_mutex_t g_static_mu = PTHREAD_MUTEX_INITIALIZER;

// Device addresses of the static arrays on the CURRENT device (resolved once per device).
int static_addrs(StaticAddrs* out) {
  int dev = -1;
  PDM_CHECK_CUDA(cudaGetDevice(&dev));
  PDM_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, PDM_ERR_LIMIT,
              "pdm_net_train_step_static: device ordinal %d outside [0, %d)", dev, kStaticMaxDevices);
  pthread_mutex_lock(&g_static_mu);
  StaticAddrs cur = g_static_addrs[dev];
  if (!cur.ready) {
    void* ps = nullptr;
    void* pb = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ps, g_static_scratch);
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&pb, g_static_sync);
    if (e != cudaSuccess || ps == nullptr || pb == nullptr) {
      pthread_mutex_unlock(&g_static_mu);
      pdm_set_error("pdm_net_train_step_static: cudaGetSymbolAddress on device %d -> %s", dev, cudaGetErrorString(e));
      return PDM_ERR_CUDA;
    }
    cur.scratch = static_cast<float*>(ps);
    cur.sync = static_cast<uint32_t*>(pb);
    cur.ready = true;
    g_static_addrs[dev] = cur;
  }
  pthread_mutex_unlock(&g_static_mu);
  *out = cur;
  return PDM_OK;
}

}  // namespace

extern "C" int pdm_net_train_step_static(const pdm_net_step_args* args, pdm_stream_t stream) {
  PDM_REQUIRE(args != nullptr, PDM_ERR_ARG, "pdm_net_train_step_static: args is NULL");
  PDM_REQUIRE(args->batch >= 1, PDM_ERR_ARG, "pdm_net_train_step_static: batch must be >= 1 (got %d)", args->batch);
  PDM_REQUIRE(args->x && args->y && args->params, PDM_ERR_ARG,
              "pdm_net_train_step_static: x, y and params are required");
  PDM_REQUIRE(args->x_dtype == PDM_U8 || args->x_dtype == PDM_F32, PDM_ERR_ARG,
              "pdm_net_train_step_static: x_dtype must be PDM_U8 or PDM_F32");
  PDM_REQUIRE(args->y_dtype == PDM_U8 || args->y_dtype == PDM_I64 || args->y_dtype == PDM_I32, PDM_ERR_ARG,
              "pdm_net_train_step_static: y_dtype must be PDM_U8, PDM_I64 or PDM_I32");
  PDM_REQUIRE(args->ctas >= 0, PDM_ERR_ARG, "pdm_net_train_step_static: ctas must be >= 0 (got %d)", args->ctas);
  PDM_REQUIRE(args->ctas <= kStaticMaxCtas, PDM_ERR_LIMIT,
              "pdm_net_train_step_static: %d CTAs requested but the built-in scratch holds at most %d "
              "(PDM_NET_STATIC_MAX_CTAS); pass ctas <= %d, or use pdm_net_train_step with your own scratch",
              args->ctas, kStaticMaxCtas, kStaticMaxCtas);
  const pdm_hyper& h = args->hyper;
  PDM_REQUIRE(h.opt == PDM_OPT_ADAM || h.opt == PDM_OPT_SGD || h.opt == PDM_OPT_NONE, PDM_ERR_ARG,
              "pdm_net_train_step_static: unknown optimizer %d", h.opt);
  PDM_REQUIRE(h.opt != PDM_OPT_ADAM || (args->m && args->v), PDM_ERR_ARG,
              "pdm_net_train_step_static: Adam needs m and v");
  PDM_REQUIRE(!(h.opt == PDM_OPT_SGD && h.beta1 != 0.f) || args->m, PDM_ERR_ARG,
              "pdm_net_train_step_static: SGD with momentum needs m");
  const int world = args->peers.world > 1 ? args->peers.world : 1;
  PDM_REQUIRE(world <= PDM_MAX_RANKS, PDM_ERR_LIMIT, "pdm_net_train_step_static: world %d > %d", world, PDM_MAX_RANKS);
  if (world > 1) {
    PDM_REQUIRE(args->peers.seq != 0, PDM_ERR_ARG, "pdm_net_train_step_static: peers.seq must be non-zero");
    PDM_REQUIRE(args->peers.rank >= 0 && args->peers.rank < world, PDM_ERR_ARG, "pdm_net_train_step_static: bad rank");
    for (int p = 0; p < world; ++p)
      PDM_REQUIRE(args->peers.inbox[p] != nullptr, PDM_ERR_STATE, "pdm_net_train_step_static: inbox[%d] not mapped", p);
  }

  StaticAddrs addrs;
  int rc = static_addrs(&addrs);
  if (rc != PDM_OK) return rc;

  // CTA count: the caller's, or the usual choice for this batch clipped to what the array holds.
  int want = args->ctas;
  if (want == 0) {
    want = pdm_net_default_ctas(args->batch);
    if (want > kStaticMaxCtas) want = kStaticMaxCtas;
  }
  KernelArgs ka;
  memset(&ka, 0, sizeof(ka));
  int ctas = 0;
  plan_rows(args->batch, want, &ctas, &ka.active, &ka.rows_per_cta);
  PDM_REQUIRE(ctas >= 1 && ctas <= kStaticMaxCtas, PDM_ERR_LIMIT,
              "pdm_net_train_step_static: planned %d CTAs, limit %d", ctas, kStaticMaxCtas);
  ka.x = args->x;
  ka.y = args->y;
  ka.y_dtype = args->y_dtype;
  ka.batch = args->batch;
  ka.inv_batch = 1.0f / (float)args->batch;
  ka.params = args->params;
  ka.m = args->m;
  ka.v = args->v;
  ka.grad_out = args->grad_out;
  ka.scratch = addrs.scratch;      // device address from cudaGetSymbolAddress, never the symbol itself
  ka.sync = addrs.sync;
  ka.metrics = args->metrics;
  ka.step_out = args->step_out;
  ka.row_index = args->row_index;
  // Same exchange form (push / pull) as pdm_net_train_step would pick, so ranks may mix the two entry
  // points; the step-to-step order is always grid completion here (the built-in barrier words are
  // re-zeroed per call, which a running count in sync[2] would not survive).
  apply_knobs(&ka, args, world, /*allow_flag_dependency=*/false);
  ka.inputs_stable = args->inputs_stable != 0;
  ka.oc = make_consts(h, world);
  ka.peers = args->peers;
  ka.peers.world = world;
  // The barrier words are the only scratch that must start at zero (an aborted earlier launch
  // could have left an arrival count behind): clear them in stream order.
  PDM_CHECK_CUDA(cudaMemsetAsync(addrs.sync, 0, 2 * sizeof(uint32_t), (cudaStream_t)stream));
  PDM_CHECK_CUDA(launch<0>(ka, args->x_dtype, ctas, (cudaStream_t)stream, /*allow_trace=*/false));
  return PDM_OK;
}

// Copies the last training launch's per-CTA clock64 stamps ([PDM_NET_MAX_CTAS][8]) to `out`.
// Only meaningful when the process runs with PDM_NET_TRACE=1; synchronises the device.
extern "C" int pdm_net_trace_read(long long* out) {
  PDM_REQUIRE(out != nullptr, PDM_ERR_ARG, "pdm_net_trace_read: out is NULL");
  PDM_REQUIRE(g_trace_buf != nullptr, PDM_ERR_STATE, "pdm_net_trace_read: tracing is off (set PDM_NET_TRACE=1)");
  PDM_CHECK_CUDA(cudaDeviceSynchronize());
  PDM_CHECK_CUDA(cudaMemcpy(out, g_trace_buf, PDM_NET_MAX_CTAS * 8 * sizeof(long long), cudaMemcpyDeviceToHost));
  return PDM_OK;
}
