// Host helpers: last-error string, SM count, TMA tensor-map encoding through the driver entry point
// (resolved at run time so the library does not link against libcuda directly).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <mutex>
#include <vector>
#include "../../include/f5e_b200.h"
#include "host.h"

namespace f5e {

static thread_local char g_err[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// Programmatic dependent launch: F5E_PDL=0 / 1 in the environment forces it off / on; otherwise it is on inside a PdlScope
// that asks for it (dit.cu: solves below PDL_MAX_ROWS rows, where the kernels are short enough for their prologues and the
// launch latency to matter; at config-2 size it costs 2 %, measured).
static int pdl_mode() {  // -1 auto, 0 off, 1 on, 2 light kernels only, 3 heavy kernels only (2 / 3: experiments)
  static const int mode = [] {
    const char* e = getenv("F5E_PDL");
    return e ? (e[0] >= '0' && e[0] <= '3' ? e[0] - '0' : 1) : -1;
  }();
  return mode;
}
static thread_local bool g_pdl_scope = false;
bool pdl_enabled(bool heavy) {
  const int m = pdl_mode();
  return m == 1 || (m == 2 && !heavy) || (m == 3 && heavy) || (m < 0 && g_pdl_scope);
}
PdlScope::PdlScope(bool on) : prev(g_pdl_scope) { g_pdl_scope = on; }
PdlScope::~PdlScope() { g_pdl_scope = prev; }

int ensure_configured() {
  static std::mutex mu;
  static bool done[64] = {};
  int dev = 0;
  F5E_CUDA_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  if (dev >= 0 && dev < 64 && done[dev]) return 0;
  int rc = gemm_configure();
  if (!rc) rc = attention_configure();
  if (!rc) rc = conv_pos_configure();
  if (!rc && dev >= 0 && dev < 64) done[dev] = true;
  return rc;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int encode(CUtensorMap* out, const void* base, int elem_bytes, int rank, const cuuint64_t* dims,
                  const cuuint64_t* strides, const cuuint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  F5E_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available (no CUDA driver?)");
  F5E_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base address must be 16-byte aligned");
  F5E_REQUIRE(box[0] * elem_bytes == 128, "TMA box inner extent must span exactly 128 bytes");
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, rank, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  F5E_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu %llu)", (int)r,
              rank, (unsigned long long)dims[0], (unsigned long long)dims[1]);
  return 0;
}

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                 uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer) {
  F5E_REQUIRE(row_stride_bytes % 16 == 0, "TMA row stride (%llu B) must be a multiple of 16",
              (unsigned long long)row_stride_bytes);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  return encode(out, base, elem_bytes, 2, dims, strides, box);
}

int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t d1, uint64_t d2,
                 uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box_inner, uint32_t box1,
                 uint32_t box2) {
  F5E_REQUIRE(stride1_bytes % 16 == 0 && stride2_bytes % 16 == 0, "TMA strides must be multiples of 16 bytes");
  cuuint64_t dims[3] = {inner, d1, d2};
  cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
  cuuint32_t box[3] = {box_inner, box1, box2};
  return encode(out, base, elem_bytes, 3, dims, strides, box);
}

}  // namespace f5e

// ---------------------------------------------------------------------------------------------
// launch accounting and in-stream timing
// ---------------------------------------------------------------------------------------------
namespace f5e {
namespace {
struct ProfRec { cudaEvent_t a, b; int cls; long long d[3]; };
std::mutex g_prof_mu;
std::vector<ProfRec> g_prof;
std::vector<cudaEvent_t> g_event_pool;
std::atomic<int> g_prof_mask{0};
std::atomic<long long> g_launches{0};

cudaEvent_t take_event() {
  if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

LaunchScope::LaunchScope(cudaStream_t s, int cls, long long a, long long b, long long c) : st(s), rec(-1) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (!(g_prof_mask.load(std::memory_order_relaxed) & cls)) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;  // not inside a graph capture
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r{take_event(), take_event(), cls, {a, b, c}};
  if (!r.a || !r.b) return;
  cudaEventRecord(r.a, s);
  g_prof.push_back(r);
  rec = (int)g_prof.size() - 1;
}
long long launch_count() { return g_launches.load(std::memory_order_relaxed); }
void add_launches(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
LaunchScope::~LaunchScope() {
  if (rec < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (rec < (int)g_prof.size()) cudaEventRecord(g_prof[rec].b, st);
}
}  // namespace f5e

extern "C" {
int64_t f5e_launch_count(void) { return f5e::g_launches.load(); }

int f5e_prof_begin(int32_t class_mask) {
  std::lock_guard<std::mutex> lk(f5e::g_prof_mu);
  for (auto& r : f5e::g_prof) { f5e::g_event_pool.push_back(r.a); f5e::g_event_pool.push_back(r.b); }
  f5e::g_prof.clear();
  f5e::g_prof_mask.store(class_mask);
  return 0;
}

int64_t f5e_prof_end(void) {
  f5e::g_prof_mask.store(0);
  std::lock_guard<std::mutex> lk(f5e::g_prof_mu);
  for (auto& r : f5e::g_prof) cudaEventSynchronize(r.b);
  return (int64_t)f5e::g_prof.size();
}

int f5e_prof_read(int64_t index, int32_t* cls, float* ms, int64_t* dims3) {
  std::lock_guard<std::mutex> lk(f5e::g_prof_mu);
  F5E_REQUIRE(index >= 0 && index < (int64_t)f5e::g_prof.size(), "f5e_prof_read: index %lld out of range", (long long)index);
  const auto& r = f5e::g_prof[index];
  float t = 0.f;
  F5E_CUDA_CHECK(cudaEventElapsedTime(&t, r.a, r.b));
  if (cls) *cls = r.cls;
  if (ms) *ms = t;
  if (dims3) { dims3[0] = r.d[0]; dims3[1] = r.d[1]; dims3[2] = r.d[2]; }
  return 0;
}
}  // extern "C"
