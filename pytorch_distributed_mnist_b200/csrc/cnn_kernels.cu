// cnn_kernels.cu -- the non-GEMM layers of the CNN extension,
// each fused with its neighbours so an activation is read once:
//
//   conv1_fwd      1->C1 3x3 direct convolution + bias + ReLU, NHWC output [B*26*26, C1]
//                  (K = 9: not GEMM-shaped, FMA is the right tool; input uint8 or fp32)
//   pool_fwd       max-pool 2x2 + bias + ReLU + dropout over the raw conv2 accumulators
//                  (relu(max(z)+b) == max(relu(z+b))), NHWC in -> NCHW-flattened out so fc1
//                  consumes PyTorch's weight layout unchanged; 1 byte/element of replay code
//   pool_bwd       scatter of dPooled through argmax/ReLU/dropout into dY2 (NHWC, zero at the
//                  discarded positions) + conv2 bias gradient
//   fcout_fwd/bwd  dropout + Linear(H,10): logits; backward produces dW, db, and the gradient
//                  of fc1's PRE-activation (dropout mask and ReLU mask folded in) + fc1 bias grad
//   conv1_bwd      weight/bias gradient of conv1 (reduction over all pixels)
//   relayout       conv weight [Cout][Cin][3][3] <-> GEMM operand layouts [Cout][tap][Cin],
//                  [Cin][tap][Cout]
//
// Dropout uses Philox4x32-10 keyed on (seed, element index): the mask is recomputed in the
// backward kernels instead of being stored.
#include "common.cuh"

namespace {

constexpr float kMean = 0.1307f, kStd = 0.3081f;

// ------------------------------------------------------------------ Philox4x32-10
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
// Uniform in [0,1) for element `idx` of stream `seed`.
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t idx) {
  uint32_t c[4] = {(uint32_t)(idx >> 2), (uint32_t)(idx >> 34), 0u, 0u};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return (float)(c[idx & 3] >> 8) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ float keep_scale(uint64_t seed, uint64_t idx, float p, float inv_keep) {
  if (p <= 0.f) return 1.f;
  return philox_uniform(seed, idx) >= p ? inv_keep : 0.f;
}

// ------------------------------------------------------------------ conv1 forward
// grid-stride over (pixel, channel quad); x is [B,784] uint8 (normalised here) or fp32.
template <typename XT>
__global__ void __launch_bounds__(256) conv1_fwd_kernel(const XT* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ bias, float* __restrict__ out,
                                                        int batch, int c1) {
  extern __shared__ float sw[];                    // [c1][9] weights then [c1] bias
  for (int i = threadIdx.x; i < c1 * 9; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < c1; i += blockDim.x) sw[c1 * 9 + i] = bias[i];
  __syncthreads();
  const int quads = c1 / 4;
  const long long total = (long long)batch * 676 * quads;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int cq = (int)(t % quads);
    const long long pix = t / quads;
    const int n = (int)(pix / 676), r = (int)(pix % 676), h = r / 26, wv = r % 26;
    float xv[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const XT raw = x[(size_t)n * 784 + (h + k / 3) * 28 + (wv + k % 3)];
      xv[k] = sizeof(XT) == 1 ? __fdiv_rn(__fdiv_rn((float)raw, 255.0f) - kMean, kStd) : (float)raw;
    }
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = cq * 4 + j;
      float acc = sw[c1 * 9 + c];
#pragma unroll
      for (int k = 0; k < 9; ++k) acc = fmaf(xv[k], sw[c * 9 + k], acc);
      o[j] = fmaxf(acc, 0.f);
    }
    *reinterpret_cast<float4*>(out + pix * c1 + cq * 4) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------ conv1 backward (wgrad + bias grad)
// dw[c][k] += sum_q dz[q][c] * x[q + tap k],  db[c] += sum_q dz[q][c]   (dz already ReLU-masked).
// One CTA walks a contiguous range of pixels in chunks of 64: the 9 input taps of every pixel are
// computed once into smem, then each warp takes 8 pixels and lane = channel (mod 32) accumulates
// its 10 sums in registers for every channel group.  One block-level reduction and c1*10 atomics
// per CTA at the end; dz is read exactly once, coalesced.
constexpr int kC1Chunk = 64;        // pixels per chunk
constexpr int kC1MaxGroups = 8;     // channel groups of 32 held in registers (c1 <= 256)

template <typename XT>
__global__ void __launch_bounds__(256) conv1_bwd_kernel(const XT* __restrict__ x, const float* __restrict__ dz,
                                                        float* __restrict__ dw, float* __restrict__ db, int batch, int c1,
                                                        int pixels_per_cta) {
  __shared__ float patch[kC1Chunk][9];
  __shared__ float red[8][32 * 10];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long pixels = (long long)batch * 676;
  const long long p_begin = (long long)blockIdx.x * pixels_per_cta;
  const long long p_end = min(pixels, p_begin + pixels_per_cta);
  const int groups = (c1 + 31) / 32;
  float acc[kC1MaxGroups][10];
#pragma unroll
  for (int g = 0; g < kC1MaxGroups; ++g)
#pragma unroll
    for (int k = 0; k < 10; ++k) acc[g][k] = 0.f;

  for (long long c0 = p_begin; c0 < p_end; c0 += kC1Chunk) {
    const int n_here = (int)min((long long)kC1Chunk, p_end - c0);
    __syncthreads();
    for (int t = threadIdx.x; t < kC1Chunk * 9; t += blockDim.x) {
      const int pi = t / 9, k = t % 9;
      float v = 0.f;
      if (pi < n_here) {
        const long long pix = c0 + pi;
        const int n = (int)(pix / 676), r = (int)(pix % 676), h = r / 26, wv = r % 26;
        const XT raw = x[(size_t)n * 784 + (h + k / 3) * 28 + (wv + k % 3)];
        v = sizeof(XT) == 1 ? __fdiv_rn(__fdiv_rn((float)raw, 255.0f) - kMean, kStd) : (float)raw;
      }
      patch[pi][k] = v;
    }
    __syncthreads();
    for (int pi = warp; pi < n_here; pi += 8) {
      float xv[9];
#pragma unroll
      for (int k = 0; k < 9; ++k) xv[k] = patch[pi][k];
      const float* row = dz + (c0 + pi) * c1;
#pragma unroll
      for (int g = 0; g < kC1MaxGroups; ++g) {
        if (g < groups) {
          const int c = g * 32 + lane;
          const float d = c < c1 ? __ldg(row + c) : 0.f;
#pragma unroll
          for (int k = 0; k < 9; ++k) acc[g][k] = fmaf(d, xv[k], acc[g][k]);
          acc[g][9] += d;
        }
      }
    }
  }
  // block reduction, one channel group at a time
#pragma unroll
  for (int g = 0; g < kC1MaxGroups; ++g) {
    if (g < groups) {
      __syncthreads();
#pragma unroll
      for (int k = 0; k < 10; ++k) red[warp][lane * 10 + k] = acc[g][k];
      __syncthreads();
      for (int t = threadIdx.x; t < 320; t += blockDim.x) {
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) sum += red[w][t];
        const int c = g * 32 + t / 10, k = t % 10;
        if (c < c1) {
          if (k < 9) atomicAdd(dw + c * 9 + k, sum);
          else atomicAdd(db + c, sum);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ pool forward
// One CTA per (image, pooled row).  z [B*676, C] NHWC raw conv accumulators (26-wide rows, valid 24x24).
// out [B, C*144] NCHW-flattened.  code: bits 0-1 argmax in the 2x2 window, bit 2 = gradient flows.
__global__ void __launch_bounds__(256) pool_fwd_kernel(const float* __restrict__ z, const float* __restrict__ bias,
                                                       float* __restrict__ out, uint8_t* __restrict__ code, int C,
                                                       float p, float inv_keep, uint64_t seed) {
  extern __shared__ float tile[];                  // [C][13] (padded) pooled values of this row
  const int n = blockIdx.x / 12, ph = blockIdx.x % 12;
  const size_t base = ((size_t)n * 676 + (size_t)(2 * ph) * 26) * C;
  uint8_t* ctile = reinterpret_cast<uint8_t*>(tile + C * 13);
  for (int t = threadIdx.x; t < 12 * C; t += blockDim.x) {
    const int c = t % C, pw = t / C;
    const float* q = z + base + (size_t)(2 * pw) * C + c;
    const float v0 = q[0], v1 = q[C], v2 = q[26 * C], v3 = q[27 * C];
    float m = v0;
    int arg = 0;
    if (v1 > m) { m = v1; arg = 1; }
    if (v2 > m) { m = v2; arg = 2; }
    if (v3 > m) { m = v3; arg = 3; }
    const float a = m + bias[c];
    const size_t oidx = (size_t)n * C * 144 + (size_t)c * 144 + ph * 12 + pw;
    const float ks = keep_scale(seed, oidx, p, inv_keep);
    const bool live = a > 0.f && ks != 0.f;
    tile[c * 13 + pw] = live ? a * ks : 0.f;
    ctile[c * 12 + pw] = (uint8_t)(arg | (live ? 4 : 0));
  }
  __syncthreads();
  for (int t = threadIdx.x; t < 12 * C; t += blockDim.x) {
    const int pw = t % 12, c = t / 12;
    const size_t oidx = (size_t)n * C * 144 + (size_t)c * 144 + ph * 12 + pw;
    out[oidx] = tile[c * 13 + pw];
    code[oidx] = ctile[c * 12 + pw];
  }
}

// ------------------------------------------------------------------ pool backward
// dp [B, C*144] -> dy [B*676, C] (every position of rows 2ph, 2ph+1 written; rows 24,25 zeroed by ph == 11).
__global__ void __launch_bounds__(256) pool_bwd_kernel(const float* __restrict__ dp, const uint8_t* __restrict__ code,
                                                       float* __restrict__ dy, float* __restrict__ dbias, int C,
                                                       float p, float inv_keep, uint64_t seed) {
  extern __shared__ float tile[];                  // [12][C] gradient, then codes
  const int n = blockIdx.x / 12, ph = blockIdx.x % 12;
  uint8_t* ctile = reinterpret_cast<uint8_t*>(tile + 12 * C);
  for (int t = threadIdx.x; t < 12 * C; t += blockDim.x) {
    const int pw = t % 12, c = t / 12;
    const size_t oidx = (size_t)n * C * 144 + (size_t)c * 144 + ph * 12 + pw;
    const uint8_t cd = code[oidx];
    float g = 0.f;
    if (cd & 4) g = dp[oidx] * (p > 0.f ? inv_keep : 1.f);   // live implies the element was kept
    tile[pw * C + c] = g;
    ctile[pw * C + c] = cd;
  }
  __syncthreads();
  const size_t base = ((size_t)n * 676 + (size_t)(2 * ph) * 26) * C;
  const int rows = ph == 11 ? 4 : 2;               // the last pooled row also clears image rows 24, 25
  for (int t = threadIdx.x; t < rows * 26 * C; t += blockDim.x) {
    const int c = t % C, pos = t / C, dyr = pos / 26, wv = pos % 26;
    float g = 0.f;
    if (dyr < 2 && wv < 24) {
      const int pw = wv >> 1, which = dyr * 2 + (wv & 1);
      if ((ctile[pw * C + c] & 3) == which) g = tile[pw * C + c];
    }
    dy[base + (size_t)pos * C + c] = g;
  }
  if (dbias != nullptr) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float s = 0.f;
      for (int pw = 0; pw < 12; ++pw) s += tile[pw * C + c];
      atomicAdd(dbias + c, s);
    }
  }
}

// ------------------------------------------------------------------ dropout + Linear(H, 10)
constexpr int kCls = 10;

// One warp per row.  h [B,H] (post-ReLU), w [10][H], logits [B,10].
__global__ void __launch_bounds__(256) fcout_fwd_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                                        const float* __restrict__ bias, float* __restrict__ logits,
                                                        int batch, int H, float p, float inv_keep, uint64_t seed) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  for (int row = gw; row < batch; row += nw) {
    float acc[kCls];
#pragma unroll
    for (int c = 0; c < kCls; ++c) acc[c] = 0.f;
    for (int j = lane; j < H; j += 32) {
      const size_t idx = (size_t)row * H + j;
      const float v = h[idx] * keep_scale(seed, idx, p, inv_keep);
#pragma unroll
      for (int c = 0; c < kCls; ++c) acc[c] = fmaf(v, __ldg(w + c * H + j), acc[c]);
    }
#pragma unroll
    for (int c = 0; c < kCls; ++c)
      for (int off = 16; off >= 1; off >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], off);
    if (lane < kCls) {
      float v = 0.f;
#pragma unroll
      for (int c = 0; c < kCls; ++c) v = lane == c ? acc[c] : v;
      logits[(size_t)row * kCls + lane] = v + bias[lane];
    }
  }
}

// dl [B,10].  Produces dz1 [B,H] = (dl . W) * dropout * (h > 0), and accumulates (atomically)
// dw [10][H], db [10], db1 [H] = colsum(dz1).
__global__ void __launch_bounds__(256) fcout_bwd_kernel(const float* __restrict__ dl, const float* __restrict__ h,
                                                        const float* __restrict__ w, float* __restrict__ dz1,
                                                        float* __restrict__ dw, float* __restrict__ db,
                                                        float* __restrict__ db1, int batch, int H, float p,
                                                        float inv_keep, uint64_t seed) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nw = gridDim.x * (blockDim.x >> 5);
  for (int j0 = 0; j0 < H; j0 += 32) {             // lane owns feature j0 + lane for this pass
    const int j = j0 + lane;
    float wcol[kCls], gw_acc[kCls];
#pragma unroll
    for (int c = 0; c < kCls; ++c) { wcol[c] = j < H ? __ldg(w + c * H + j) : 0.f; gw_acc[c] = 0.f; }
    float gb1 = 0.f;
    float gb[kCls];
#pragma unroll
    for (int c = 0; c < kCls; ++c) gb[c] = 0.f;
    for (int row = gw; row < batch; row += nw) {
      float d[kCls];
#pragma unroll
      for (int c = 0; c < kCls; ++c) d[c] = __ldg(dl + (size_t)row * kCls + c);
      if (j < H) {
        const size_t idx = (size_t)row * H + j;
        const float hv = h[idx];
        const float ks = keep_scale(seed, idx, p, inv_keep);
        const float hd = hv * ks;
        float g = 0.f;
#pragma unroll
        for (int c = 0; c < kCls; ++c) { g = fmaf(d[c], wcol[c], g); gw_acc[c] = fmaf(d[c], hd, gw_acc[c]); }
        g = hv > 0.f ? g * ks : 0.f;
        dz1[idx] = g;
        gb1 += g;
      }
      if (j0 == 0 && lane == 0) {
#pragma unroll
        for (int c = 0; c < kCls; ++c) gb[c] += d[c];
      }
    }
    if (j < H) {
#pragma unroll
      for (int c = 0; c < kCls; ++c) atomicAdd(dw + c * H + j, gw_acc[c]);
      if (db1 != nullptr) atomicAdd(db1 + j, gb1);
    }
    if (j0 == 0 && lane == 0) {
#pragma unroll
      for (int c = 0; c < kCls; ++c) atomicAdd(db + c, gb[c]);
    }
  }
}

// ------------------------------------------------------------------ conv weight re-layouts
// dir 0: w[Co][Ci][9] -> a[Co][9][Ci] and b[Ci][9][Co]      (before the conv GEMMs)
// dir 1: g[Co][9][Ci] -> dw[Co][Ci][9] (+=)                  (after the wgrad GEMM)
__global__ void relayout_kernel(const float* __restrict__ src, float* __restrict__ a, float* __restrict__ b, int co, int ci, int dir) {
  const int total = co * ci * 9;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    if (dir == 0) {
      const int k = t % 9, i = (t / 9) % ci, o = t / (9 * ci);
      const float v = src[t];
      a[(o * 9 + k) * ci + i] = v;
      if (b != nullptr) b[(i * 9 + k) * co + o] = v;
    } else {
      const int i = t % ci, k = (t / ci) % 9, o = t / (9 * ci);
      a[(o * ci + i) * 9 + k] += src[t];
    }
  }
}

// z[r][c] = relu(z[r][c] + bias[c]) in place (epilogue of a split-K GEMM).
__global__ void bias_relu_kernel(float* __restrict__ z, const float* __restrict__ bias, long long rows, int cols) {
  const long long total = rows * cols;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x)
    z[t] = fmaxf(z[t] + bias[(int)(t % cols)], 0.f);
}

int grid_for(long long work, int per_block, int cap) {
  long long g = (work + per_block - 1) / per_block;
  if (g < 1) g = 1;
  if (g > cap) g = cap;
  return (int)g;
}

}  // namespace

extern "C" {

int pdm_conv1_fwd(const void* x, int x_dtype, const float* w, const float* bias, float* out, int batch, int c1,
                  pdm_stream_t stream) {
  PDM_REQUIRE(x && w && bias && out && batch >= 1, PDM_ERR_ARG, "pdm_conv1_fwd: bad arguments");
  PDM_REQUIRE(c1 % 4 == 0 && c1 >= 4 && c1 <= 1024, PDM_ERR_LIMIT, "pdm_conv1_fwd: channels %d must be a multiple of 4 in [4, 1024]", c1);
  PDM_REQUIRE(x_dtype == PDM_U8 || x_dtype == PDM_F32, PDM_ERR_ARG, "pdm_conv1_fwd: x_dtype must be PDM_U8 or PDM_F32");
  const int grid = grid_for((long long)batch * 676 * (c1 / 4), 256, 148 * 16);
  const size_t smem = (size_t)c1 * 10 * sizeof(float);
  if (x_dtype == PDM_U8)
    conv1_fwd_kernel<uint8_t><<<grid, 256, smem, (cudaStream_t)stream>>>((const uint8_t*)x, w, bias, out, batch, c1);
  else
    conv1_fwd_kernel<float><<<grid, 256, smem, (cudaStream_t)stream>>>((const float*)x, w, bias, out, batch, c1);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_conv1_bwd(const void* x, int x_dtype, const float* dz, float* dw, float* db, int batch, int c1, pdm_stream_t stream) {
  PDM_REQUIRE(x && dz && dw && db && batch >= 1 && c1 >= 1, PDM_ERR_ARG, "pdm_conv1_bwd: bad arguments");
  PDM_REQUIRE(c1 <= 32 * kC1MaxGroups, PDM_ERR_LIMIT, "pdm_conv1_bwd: at most %d channels (got %d)", 32 * kC1MaxGroups, c1);
  const long long pixels = (long long)batch * 676;
  int per_cta = 512;
  while ((pixels + per_cta - 1) / per_cta > 148 * 8) per_cta *= 2;
  const int grid = (int)((pixels + per_cta - 1) / per_cta);
  if (x_dtype == PDM_U8)
    conv1_bwd_kernel<uint8_t><<<grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)x, dz, dw, db, batch, c1, per_cta);
  else
    conv1_bwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, dz, dw, db, batch, c1, per_cta);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_pool_fwd(const float* z, const float* bias, float* out, uint8_t* code, int batch, int C, float p, uint64_t seed,
                 pdm_stream_t stream) {
  PDM_REQUIRE(z && bias && out && code && batch >= 1 && C >= 1, PDM_ERR_ARG, "pdm_pool_fwd: bad arguments");
  PDM_REQUIRE(C <= 2048 && p >= 0.f && p < 1.f, PDM_ERR_LIMIT, "pdm_pool_fwd: C <= 2048 and 0 <= p < 1 required");
  const size_t smem = (size_t)C * 13 * sizeof(float) + (size_t)C * 12;
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    PDM_CHECK_CUDA(cudaFuncSetAttribute(pool_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  pool_fwd_kernel<<<batch * 12, 256, smem, (cudaStream_t)stream>>>(z, bias, out, code, C, p, 1.f / (1.f - p), seed);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_pool_bwd(const float* dp, const uint8_t* code, float* dy, float* dbias, int batch, int C, float p, uint64_t seed,
                 pdm_stream_t stream) {
  PDM_REQUIRE(dp && code && dy && batch >= 1 && C >= 1, PDM_ERR_ARG, "pdm_pool_bwd: bad arguments");
  PDM_REQUIRE(C <= 2048 && p >= 0.f && p < 1.f, PDM_ERR_LIMIT, "pdm_pool_bwd: C <= 2048 and 0 <= p < 1 required");
  const size_t smem = (size_t)C * 12 * sizeof(float) + (size_t)C * 12;
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) {
    PDM_CHECK_CUDA(cudaFuncSetAttribute(pool_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr = smem;
  }
  pool_bwd_kernel<<<batch * 12, 256, smem, (cudaStream_t)stream>>>(dp, code, dy, dbias, C, p, 1.f / (1.f - p), seed);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_fcout_fwd(const float* h, const float* w, const float* bias, float* logits, int batch, int H, float p,
                  uint64_t seed, pdm_stream_t stream) {
  PDM_REQUIRE(h && w && bias && logits && batch >= 1 && H >= 1, PDM_ERR_ARG, "pdm_fcout_fwd: bad arguments");
  PDM_REQUIRE(p >= 0.f && p < 1.f, PDM_ERR_ARG, "pdm_fcout_fwd: 0 <= p < 1 required");
  fcout_fwd_kernel<<<grid_for(batch, 8, 148 * 4), 256, 0, (cudaStream_t)stream>>>(h, w, bias, logits, batch, H, p,
                                                                                  1.f / (1.f - p), seed);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_fcout_bwd(const float* dl, const float* h, const float* w, float* dz1, float* dw, float* db, float* db1,
                  int batch, int H, float p, uint64_t seed, pdm_stream_t stream) {
  PDM_REQUIRE(dl && h && w && dz1 && dw && db && batch >= 1 && H >= 1, PDM_ERR_ARG, "pdm_fcout_bwd: bad arguments");
  PDM_REQUIRE(p >= 0.f && p < 1.f, PDM_ERR_ARG, "pdm_fcout_bwd: 0 <= p < 1 required");
  fcout_bwd_kernel<<<grid_for(batch, 8 * 4, 148 * 2), 256, 0, (cudaStream_t)stream>>>(dl, h, w, dz1, dw, db, db1, batch, H, p,
                                                                                      1.f / (1.f - p), seed);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_bias_relu(float* z, const float* bias, int64_t rows, int cols, pdm_stream_t stream) {
  PDM_REQUIRE(z && bias && rows >= 1 && cols >= 1, PDM_ERR_ARG, "pdm_bias_relu: bad arguments");
  bias_relu_kernel<<<grid_for(rows * cols, 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(z, bias, rows, cols);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

int pdm_conv_relayout(const float* src, float* a, float* b, int cout, int cin, int dir, pdm_stream_t stream) {
  PDM_REQUIRE(src && a && cout >= 1 && cin >= 1 && (dir == 0 || dir == 1), PDM_ERR_ARG, "pdm_conv_relayout: bad arguments");
  relayout_kernel<<<grid_for((long long)cout * cin * 9, 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(src, a, b, cout, cin, dir);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

}  // extern "C"
