// extern "C" surface of libf5e_b200.so (declared in include/f5e_b200.h).
#include "../../include/f5e_b200.h"
#include "host.h"

using namespace f5e;

static EpiParams to_epi(const f5e_epilogue* e, int M) {
  EpiParams p{};
  p.out = e->out; p.ldo = e->ldo; p.out_f32 = e->out_f32;
  p.bias = e->bias; p.act = e->act;
  p.gate = e->gate; p.gate_stride = e->gate_stride;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(e->residual); p.ldr = e->ldr;
  p.row_len = e->row_len; p.n_pad = e->n_pad > 0 ? e->n_pad : M;
  p.rope_inner = e->rope_inner; p.rope_heads = e->rope_heads;
  p.res_f32 = e->residual_f32; p.split3 = e->out_split3; p.row_info = e->row_info;
  p.rope_inv_freq = e->rope_inv_freq;
  p.ln_rstd = e->ln_rstd; p.ln_a2 = reinterpret_cast<const __nv_bfloat16*>(e->ln_a2);
  p.stats_out = reinterpret_cast<float2*>(e->stats_out); p.stats_slots = e->stats_slots;
  return p;
}

extern "C" {

const char* f5e_last_error(void) { return last_error(); }
#ifndef F5E_BUILD_STAMP
#define F5E_BUILD_STAMP "unstamped"
#endif
const char* f5e_build_stamp(void) { return F5E_BUILD_STAMP; }
int f5e_version(void) { return 100; }

int f5e_check_device(void) {
  int dev = 0;
  F5E_CUDA_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  F5E_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
  F5E_REQUIRE(prop.major == 10, "device %d (%s) is sm_%d%d; libf5e_b200 needs sm_100a (B200)", dev, prop.name,
              prop.major, prop.minor);
  return 0;
}

int f5e_gemm_bf16(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                  const f5e_epilogue* ep, void* stream) {
  F5E_REQUIRE(A && W && ep, "f5e_gemm_bf16: null argument");
  F5E_REQUIRE(ep->out_f32 ? ep->ldo % 4 == 0 : ep->ldo % 8 == 0, "f5e_gemm_bf16: ldo=%d breaks 16-byte row alignment", ep->ldo);
  F5E_REQUIRE(ep->residual == nullptr || ep->ldr % (ep->residual_f32 ? 4 : 8) == 0,
              "f5e_gemm_bf16: ldr=%d breaks 16-byte row alignment", ep->ldr);
  F5E_REQUIRE(ep->out_split3 == 0 || (N % 32 == 0 && ep->out_split3 % 8 == 0 && !ep->out_f32),
              "f5e_gemm_bf16: out_split3 needs N %% 32 == 0, a bf16 output and an 8-aligned split stride");
  return gemm_bf16((cudaStream_t)stream, (const __nv_bfloat16*)A, lda, (const __nv_bfloat16*)W, ldw, M, N, K,
                   to_epi(ep, M));
}

int f5e_norm_mod(const void* x, void* out, int32_t M, int32_t D, const float* mul, const float* add,
                 int32_t mod_stride, int32_t n_pad, float plus_one, float eps, void* stream) {
  F5E_REQUIRE(x && out && mul && add, "f5e_norm_mod: null argument");
  return norm_mod((cudaStream_t)stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)out, M, D, mul, add, mod_stride,
                  n_pad, plus_one, eps);
}

int f5e_row_stats(const void* x, float* rstd, void* a2, int32_t M, int32_t D, float eps, void* stream) {
  F5E_REQUIRE(x && rstd && a2, "f5e_row_stats: null argument");
  return row_stats((cudaStream_t)stream, (const __nv_bfloat16*)x, rstd, (__nv_bfloat16*)a2, M, D, eps);
}

int f5e_ln_finalize(const float* stats, int32_t slots, int32_t M, int32_t D, float eps, float* rstd, void* a2, void* stream) {
  F5E_REQUIRE(stats && rstd && a2, "f5e_ln_finalize: null argument");
  return ln_finalize((cudaStream_t)stream, reinterpret_cast<const float2*>(stats), slots, M, D, eps, rstd, (__nv_bfloat16*)a2);
}

int f5e_fold_ln_weights(const f5e_fold_job* jobs_dev, int32_t n_jobs, int32_t total_rows, int32_t D, const float* mod,
                        void* stream) {
  F5E_REQUIRE(jobs_dev && mod, "f5e_fold_ln_weights: null argument");
  return fold_ln_weights((cudaStream_t)stream, reinterpret_cast<const FoldJob*>(jobs_dev), n_jobs, total_rows, D, mod);
}

int f5e_gather_rows(const void* src, void* dst, const int32_t* row_map, int32_t rows, int32_t C, void* stream) {
  F5E_REQUIRE(src && dst && row_map && rows > 0 && C % 8 == 0, "f5e_gather_rows: bad arguments");
  return gather_rows((cudaStream_t)stream, (const __nv_bfloat16*)src, (__nv_bfloat16*)dst, row_map, rows, C);
}

int f5e_cfg_euler(const float* pred, int32_t ld_pred, float* y, void* x_cat, int32_t ld_cat, float* traj,
                  int32_t rows, int32_t C, float dt, float cfg_strength, int32_t round_bf16, void* stream) {
  F5E_REQUIRE(pred && y, "f5e_cfg_euler: null argument");
  const int nb = cfg_strength >= 1e-5f ? 2 : 1;
  return guided_euler((cudaStream_t)stream, pred, ld_pred, y, (__nv_bfloat16*)x_cat, ld_cat, traj, rows, C, dt, nb,
                      cfg_strength, 0.0f, round_bf16);
}

int f5e_guided_euler(const float* pred, int32_t ld_pred, float* y, void* x_cat, int32_t ld_cat, float* traj,
                     int32_t rows, int32_t C, float dt, const f5e_guidance* g, int32_t round_bf16, void* stream) {
  F5E_REQUIRE(pred && y && g, "f5e_guided_euler: null argument");
  return guided_euler((cudaStream_t)stream, pred, ld_pred, y, (__nv_bfloat16*)x_cat, ld_cat, traj, rows, C, dt,
                      g->branches, g->c0, g->c1, round_bf16);
}

int f5e_attention(const void* qkv, int32_t ld_qkv, void* out, int32_t ldo, const int32_t* klen, int32_t samples,
                  int32_t n_pad, int32_t heads, float scale, void* stream) {
  F5E_REQUIRE(qkv && out, "f5e_attention: null argument");
  return attention((cudaStream_t)stream, (const __nv_bfloat16*)qkv, ld_qkv, (__nv_bfloat16*)out, ldo, klen, samples,
                   n_pad, heads, scale);
}

int f5e_attention_bias(const void* qkv, int32_t ld_qkv, void* out, int32_t ldo, const int32_t* klen, const float* key_bias,
                       int32_t ld_bias, int32_t samples, int32_t n_pad, int32_t heads, float scale, void* stream) {
  F5E_REQUIRE(qkv && out && key_bias, "f5e_attention_bias: null argument");
  return attention((cudaStream_t)stream, (const __nv_bfloat16*)qkv, ld_qkv, (__nv_bfloat16*)out, ldo, klen, samples,
                   n_pad, heads, scale, nullptr, 0, key_bias, ld_bias);
}

int f5e_conv_pos(const void* x, const void* w_packed, int32_t samples, int32_t n_pad, int32_t dim, int32_t groups,
                 const f5e_epilogue* ep, void* stream) {
  F5E_REQUIRE(x && w_packed && ep && ep->out, "f5e_conv_pos: null argument");
  F5E_REQUIRE(ep->ldo % 8 == 0 && (ep->residual == nullptr || ep->ldr % 8 == 0), "f5e_conv_pos: unaligned leading dim");
  EpiParams p = to_epi(ep, samples * n_pad);
  p.n_pad = n_pad;
  return conv_pos((cudaStream_t)stream, (const __nv_bfloat16*)x, (const __nv_bfloat16*)w_packed, samples, n_pad, dim,
                  groups, p);
}

}  // extern "C"
