// core.cu -- error reporting, version, device probe.
#include "common.cuh"

static thread_local char g_err[1024] = "";

void pdm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* pdm_last_error(void) { return g_err; }

#include <stdlib.h>
unsigned pdm_skew_ns(void) {
  static const unsigned v = [] { const char* e = getenv("PDM_SKEW_NS"); return e && *e ? (unsigned)atoi(e) : 0u; }();
  return v;
}

extern "C" int pdm_version(void) { return 100; }

// sizeof() of every struct that crosses the C ABI, so that a binding (ctypes, cffi, another compiler)
// can check its own declaration against this build.  Order = order of appearance in pdmnist.h.
extern "C" size_t pdm_abi_sizeof(int which) {
  switch (which) {
    case 0: return sizeof(pdm_hyper);
    case 1: return sizeof(pdm_peers);
    case 2: return sizeof(pdm_net_step_args);
    case 3: return sizeof(pdm_net_tc_args);
    case 4: return sizeof(pdm_gemm_args);
    case 5: return sizeof(pdm_cnn_params);
    case 6: return sizeof(pdm_cnn_plan);
    case 7: return sizeof(pdm_reducer_config);
    default: return 0;
  }
}

extern "C" int pdm_device_ok(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
    cudaGetLastError();
    return 0;
  }
  int major = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return major == 10 ? 1 : 0;
}

// Bucket assignment of the gradient reducer (the role of c10d's
// _compute_bucket_assignment_by_size): walk the parameters in REVERSE order (gradients
// become ready back to front), close a bucket once it reaches its byte cap; the first bucket uses
// the small cap so communication starts early.  out_bucket[i] = bucket index of parameter i.
// Returns the number of buckets, or a negative status.
extern "C" int pdm_assign_buckets(const int64_t* numels, int n, int64_t first_cap_bytes, int64_t cap_bytes,
                                  int elem_size, int32_t* out_bucket) {
  PDM_REQUIRE(numels && out_bucket && n >= 1 && elem_size >= 1 && first_cap_bytes >= 1 && cap_bytes >= 1, PDM_ERR_ARG,
              "pdm_assign_buckets: bad arguments");
  int bucket = 0;
  int64_t bytes = 0, cap = first_cap_bytes;
  bool open = false;
  for (int i = n - 1; i >= 0; --i) {
    PDM_REQUIRE(numels[i] >= 0, PDM_ERR_ARG, "pdm_assign_buckets: negative size at %d", i);
    out_bucket[i] = bucket;
    bytes += numels[i] * elem_size;
    open = true;
    if (bytes >= cap) {
      bucket += 1;
      bytes = 0;
      cap = cap_bytes;
      open = false;
    }
  }
  return bucket + (open ? 1 : 0);
}
