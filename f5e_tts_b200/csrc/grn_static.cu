// f5e_grn_static: GRN (model/modules.py:225-234) for callers that may not allocate device memory - code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  Same two kernels as f5e_grn (embed.cu: grn_sumsq_kernel,
// grn_apply_kernel); the slab partial sums live in a file-scope __device__ array of this translation unit instead of a
// caller-provided buffer.  No counterpart in the reference.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its device address is looked up once per
// device with cudaGetSymbolAddress (with that device current) and passed to the kernels as an ordinary pointer.
// Every GPU that loads the library holds the array (F5E_GRN_STATIC_MAX_FLOATS * 4 bytes = 4 MiB) until the process exits.
// Every call shares it: calls that overlap on the GPU (different streams, or the same stream from two threads without
// ordering) race on it.  Nothing in it has to start at zero - grn_sumsq writes every partial grn_apply reads - so the
// contents a previous call leaves behind are harmless.
#include "../../include/f5e_b200.h"
#include "host.h"
#include <mutex>

using namespace f5e;

static_assert(F5E_GRN_STATIC_MAX_FLOATS == (1 << 20), "keep the header's limit and the array in step");
static __device__ __align__(256) float g_grn_scratch[F5E_GRN_STATIC_MAX_FLOATS];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
float* g_scratch_addr[MAX_DEVICES] = {};  // device address of g_grn_scratch on device i, nullptr = not looked up yet

int scratch_for_current_device(float** out) {
  int dev = -1;
  F5E_CUDA_CHECK(cudaGetDevice(&dev));
  F5E_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "f5e_grn_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[dev]) {
    float* p = nullptr;
    F5E_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_grn_scratch));  // resolves on the CURRENT device; allocates nothing
    F5E_REQUIRE(p != nullptr, "f5e_grn_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" {

int64_t f5e_grn_static_scratch_floats(int32_t samples, int32_t n_pad, int32_t C) {
  if (samples <= 0 || n_pad <= 0 || C <= 0) return 0;
  return (int64_t)samples * ((n_pad + 63) / 64) * C;
}

int f5e_grn_static(const void* x_bf16, const float* gamma, const float* beta, void* out_bf16, int32_t samples,
                   int32_t n_pad, int32_t C, void* stream) {
  F5E_REQUIRE(x_bf16 && gamma && beta && out_bf16, "f5e_grn_static: null argument");
  F5E_REQUIRE(samples > 0 && n_pad > 0 && C > 0, "f5e_grn_static: bad shape (samples=%d n_pad=%d C=%d)", samples, n_pad, C);
  const int64_t need = f5e_grn_static_scratch_floats(samples, n_pad, C);
  F5E_REQUIRE(need <= (int64_t)F5E_GRN_STATIC_MAX_FLOATS,
              "f5e_grn_static: samples * ceil(n_pad / 64) * C = %lld floats of scratch exceed the built-in %d "
              "(F5E_GRN_STATIC_MAX_FLOATS); use f5e_grn with a caller-provided scratch buffer",
              (long long)need, (int)F5E_GRN_STATIC_MAX_FLOATS);
  float* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc) return rc;
  return grn((cudaStream_t)stream, (const __nv_bfloat16*)x_bf16, gamma, beta, (__nv_bfloat16*)out_bf16, scratch, samples,
             n_pad, C);
}

}  // extern "C"
