// optim.cu -- flat fused optimizers and the fused cross-entropy kernel.
//
// pdm_optim_step replaces torch.optim.Adam's foreach path (ref:92; SURVEY
// N11/K11: lerp_, mul_, addcmul_, sqrt, div_, add_, addcdiv_ = 8 multi-tensor
// launches) and the commented-out SGD (ref:192-194) with ONE launch over the
// model's flat parameter buffer.  The math mirrors torch's operation order so
// trajectories agree to rounding.
//
// pdm_cross_entropy replaces F.cross_entropy forward + backward and the
// argmax/eq/sum of Accuracy.update (ref:88, ref:61-62).
#include "common.cuh"

#include <math.h>

namespace {

struct OptK {
  int opt, first;
  float lr, beta1, beta2, eps, wd, step_size, bc2_sqrt, gscale;
};

__device__ __forceinline__ void upd(const OptK& k, float g, float& p, float& m, float& v) {
  g *= k.gscale;
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

// One float4 per thread per iteration; grid-stride.  n4 = n / 4, tail handled scalar.
__global__ void __launch_bounds__(256) optim_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                    float* __restrict__ m, float* __restrict__ v, size_t n,
                                                    const OptK k) {
  const size_t n4 = n / 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const bool has_m = m != nullptr, has_v = v != nullptr;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 p4 = reinterpret_cast<float4*>(p)[i];
    const float4 g4 = __ldcs(reinterpret_cast<const float4*>(g) + i);
    float4 m4 = make_float4(0.f, 0.f, 0.f, 0.f), v4 = m4;
    if (has_m) m4 = reinterpret_cast<float4*>(m)[i];
    if (has_v) v4 = reinterpret_cast<float4*>(v)[i];
    upd(k, g4.x, p4.x, m4.x, v4.x);
    upd(k, g4.y, p4.y, m4.y, v4.y);
    upd(k, g4.z, p4.z, m4.z, v4.z);
    upd(k, g4.w, p4.w, m4.w, v4.w);
    reinterpret_cast<float4*>(p)[i] = p4;
    if (has_m) reinterpret_cast<float4*>(m)[i] = m4;
    if (has_v) reinterpret_cast<float4*>(v)[i] = v4;
  }
  if (blockIdx.x == 0) {
    for (size_t i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
      float pp = p[i], mm = has_m ? m[i] : 0.f, vv = has_v ? v[i] : 0.f;
      upd(k, g[i], pp, mm, vv);
      p[i] = pp;
      if (has_m) m[i] = mm;
      if (has_v) v[i] = vv;
    }
  }
}

constexpr int kCeThreads = 256;

// One thread per row, C <= 32 logits held in registers via a bounded loop.
__global__ void __launch_bounds__(kCeThreads) ce_kernel(const float* __restrict__ logits, const void* y, int y_dtype,
                                                        int batch, int classes, float inv_batch,
                                                        float* loss_mean, float* __restrict__ dlogits,
                                                        double* metrics, int single_block) {
  __shared__ float s_loss[kCeThreads / 32];
  __shared__ float s_hit[kCeThreads / 32];
  float loss = 0.f, hits = 0.f;
  for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < batch; row += gridDim.x * blockDim.x) {
    const float* z = logits + (size_t)row * classes;
    int label;
    if (y_dtype == PDM_U8) label = ((const uint8_t*)y)[row];
    else if (y_dtype == PDM_I64) label = (int)((const long long*)y)[row];
    else label = ((const int*)y)[row];
    label = min(max(label, 0), classes - 1);
    float mx = -INFINITY;
    int arg = 0;
    for (int c = 0; c < classes; ++c) {
      const float t = z[c];
      if (t > mx) { mx = t; arg = c; }
    }
    float s = 0.f;
    for (int c = 0; c < classes; ++c) s += expf(z[c] - mx);
    loss += (mx + logf(s)) - z[label];
    hits += (arg == label) ? 1.f : 0.f;
    if (dlogits != nullptr) {
      const float inv_s = 1.f / s;
      float* d = dlogits + (size_t)row * classes;
      for (int c = 0; c < classes; ++c) d[c] = (expf(z[c] - mx) * inv_s - (c == label ? 1.f : 0.f)) * inv_batch;
    }
  }
  for (int off = 16; off >= 1; off >>= 1) {
    loss += __shfl_xor_sync(0xffffffffu, loss, off);
    hits += __shfl_xor_sync(0xffffffffu, hits, off);
  }
  if ((threadIdx.x & 31) == 0) { s_loss[threadIdx.x >> 5] = loss; s_hit[threadIdx.x >> 5] = hits; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, h = 0.f;
    for (int i = 0; i < kCeThreads / 32; ++i) { l += s_loss[i]; h += s_hit[i]; }
    if (single_block) {
      if (loss_mean != nullptr) *loss_mean = l * inv_batch;
      if (metrics != nullptr) { metrics[0] += (double)l; metrics[1] += (double)h; metrics[2] += (double)batch; }
    } else {
      if (loss_mean != nullptr) atomicAdd(loss_mean, l * inv_batch);
      if (metrics != nullptr) {
        atomicAdd(metrics + 0, (double)l);
        atomicAdd(metrics + 1, (double)h);
        if (blockIdx.x == 0) atomicAdd(metrics + 2, (double)batch);
      }
    }
  }
}

}  // namespace

extern "C" int pdm_optim_step(float* params, const float* grad, float* m, float* v, size_t n,
                              const pdm_hyper* h, float grad_scale, pdm_stream_t stream) {
  PDM_REQUIRE(params && grad && h, PDM_ERR_ARG, "pdm_optim_step: params, grad and hyper are required");
  PDM_REQUIRE(h->opt == PDM_OPT_ADAM || h->opt == PDM_OPT_SGD, PDM_ERR_ARG, "pdm_optim_step: unknown optimizer %d", h->opt);
  PDM_REQUIRE(h->opt != PDM_OPT_ADAM || (m && v), PDM_ERR_ARG, "pdm_optim_step: Adam needs m and v");
  PDM_REQUIRE(!(h->opt == PDM_OPT_SGD && h->beta1 != 0.f) || m, PDM_ERR_ARG, "pdm_optim_step: SGD momentum needs m");
  PDM_REQUIRE(((uintptr_t)params | (uintptr_t)grad | (uintptr_t)m | (uintptr_t)v) % 16 == 0, PDM_ERR_ARG,
              "pdm_optim_step: buffers must be 16-byte aligned");
  if (n == 0) return PDM_OK;
  OptK k;
  k.opt = h->opt;
  k.first = h->step <= 1;
  k.lr = h->lr; k.beta1 = h->beta1; k.beta2 = h->beta2; k.eps = h->eps; k.wd = h->weight_decay;
  k.step_size = 0.f; k.bc2_sqrt = 1.f;
  k.gscale = grad_scale;
  if (h->opt == PDM_OPT_ADAM) {
    const double step = (double)(h->step < 1 ? 1 : h->step);
    k.step_size = (float)((double)h->lr / (1.0 - pow((double)h->beta1, step)));
    k.bc2_sqrt = (float)sqrt(1.0 - pow((double)h->beta2, step));
  }
  size_t blocks = (n / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  optim_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, grad, m, v, n, k);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}

extern "C" int pdm_cross_entropy(const float* logits, const void* y, int y_dtype, int batch, int classes,
                                 float* loss_mean, float* dlogits, double* metrics, pdm_stream_t stream) {
  PDM_REQUIRE(logits && y, PDM_ERR_ARG, "pdm_cross_entropy: logits and y are required");
  PDM_REQUIRE(batch >= 1 && classes >= 1, PDM_ERR_ARG, "pdm_cross_entropy: bad shape [%d, %d]", batch, classes);
  PDM_REQUIRE(classes <= 1024, PDM_ERR_LIMIT, "pdm_cross_entropy: classes %d > 1024", classes);
  int blocks = (batch + kCeThreads - 1) / kCeThreads;
  if (blocks > 148 * 4) blocks = 148 * 4;
  const int single = blocks == 1;
  cudaStream_t s = (cudaStream_t)stream;
  if (!single && loss_mean != nullptr) PDM_CHECK_CUDA(cudaMemsetAsync(loss_mean, 0, sizeof(float), s));
  ce_kernel<<<blocks, kCeThreads, 0, s>>>(logits, y, y_dtype, batch, classes, 1.0f / (float)batch, loss_mean,
                                          dlogits, metrics, single);
  PDM_CHECK_CUDA(cudaGetLastError());
  return PDM_OK;
}
