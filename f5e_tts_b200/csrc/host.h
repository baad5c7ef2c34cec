// Host-side internal interfaces between the translation units of libf5e_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "epilogue.cuh"

namespace f5e {

int num_sms();

// 2-D row-major tensor map with 128-byte swizzle; `elem_bytes` 2 (bf16) or 4 (fp32/tf32).
// box_inner * elem_bytes must be 128.
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t outer,
                 uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer);
// 3-D map over [d2][d1][inner] (used for [sample][frame][channel] views where out-of-range frames
// of ONE sample must read as zero instead of bleeding into the neighbouring sample).
int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t d1,
                 uint64_t d2, uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box_inner,
                 uint32_t box1, uint32_t box2);

// C[M,N] = epilogue(A[M,K] @ W[N,K]^T)   (bf16 operands, fp32 accumulation in TMEM)
int gemm_bf16(cudaStream_t st, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int ldw, int M,
              int N, int K, const EpiParams& ep);

bool gemm_can_fuse_stats(int M, int N);
// (mean, M2) partials of gemm_bf16's stats_out ([M, slots], 128 columns each) -> rstd [M] and a2 [M, 64] of the folded LayerNorm
int ln_finalize(cudaStream_t st, const float2* stats, int slots, int M, int D, float eps, float* rstd, __nv_bfloat16* a2);
int gemm_vq(cudaStream_t st, const __nv_bfloat16* A, int lda, const __nv_bfloat16* Wp, int ldw, const float* bias_p, int M, int K,
            const VqEpi& vq);
int vq_refine(cudaStream_t st, const float* x, int ld_x, int dim, const float* W, const float* bias, const VqEpi& vq,
              float* perplexity, int M);

int norm_mod(cudaStream_t st, const __nv_bfloat16* x, __nv_bfloat16* out, int M, int D, const float* mul,
             const float* add, int mod_stride, int n_pad, float plus_one, float eps, const int* row_info = nullptr);
// LayerNorm folded into the next GEMM (elementwise.cu): per-row (mean, rstd) of x; per-step rewrite of the weights
struct FoldJob {  // device-resident job list of fold_ln_weights; mirrors the public f5e_fold_job
  const __nv_bfloat16* w;   // [N, D] original weight
  const float* bias;        // [N] or null
  __nv_bfloat16* wf;        // [N, D + 64] out: bf16(w * (1 + scale)) | column sum as (hi, hi, lo, 0 ...)
  float* c;                 // [N] out: row sums of wf
  float* d;                 // [N] out: bias + w . shift
  int N, row0, scale_off, shift_off;  // rows of this job, first launch row, offsets of scale / shift in the step's mod vector
};
int row_stats(cudaStream_t st, const __nv_bfloat16* x, float* rstd, __nv_bfloat16* a2, int M, int D, float eps);
int fold_ln_weights(cudaStream_t st, const FoldJob* jobs_dev, int n_jobs, int total_rows, int D, const float* mod);
int guided_euler(cudaStream_t st, const float* pred, int ld_pred, float* y, __nv_bfloat16* x_cat, int ld_cat,
                 float* traj, int rows, int C, float dt, int branches, float c0, float c1, int round_bf16,
                 const int* row_map = nullptr, int rows_state = 0);
int qk_norm_rope(cudaStream_t st, __nv_bfloat16* qkv, int ld, int rows, int inner, int heads, const float* qw, const float* kw,
                 float eps, const float* inv_freq, int rope_heads, int n_pad, const int* row_info);
int gather_rows(cudaStream_t st, const __nv_bfloat16* src, __nv_bfloat16* dst, const int* row_map, int rows, int C);
// o[rows, inner] = softmax(q k^T * scale, keys < klen[sample]) v over qkv[rows, 3*inner] (Q|K|V column blocks)
// row_start (optional): first row of every sample when the rows are PACKED (valid frames only); total_rows = rows of qkv
int attention(cudaStream_t st, const __nv_bfloat16* qkv, int ld_qkv, __nv_bfloat16* out, int ldo, const int* klen,
              int samples, int n_pad, int heads, float scale, const int* row_start = nullptr, int total_rows = 0,
              const float* key_bias = nullptr, int ld_bias = 0);
// grouped conv k=31 as implicit GEMM; wp = packed weights [31*dim, 64]
int conv_pos(cudaStream_t st, const __nv_bfloat16* x, const __nv_bfloat16* wp, int samples, int n_pad, int dim,
             int groups, const EpiParams& ep);
int pack_cols(cudaStream_t st, __nv_bfloat16* dst, int ld_dst, int col0, const void* src, int src_is_f32, int ld_src,
              int rows, int cols);
int text_gather(cudaStream_t st, const int64_t* text, int samples, int nt, const __nv_bfloat16* table, const float* pos,
                int max_pos, __nv_bfloat16* out, int n_pad, int dim, int drop, int mask_padding, int add_pos);
int dwconv7(cudaStream_t st, const __nv_bfloat16* x, const float* w, const float* b, __nv_bfloat16* out, int samples,
            int n_pad, int C);
int grn(cudaStream_t st, const __nv_bfloat16* x, const float* gamma, const float* beta, __nv_bfloat16* out,
        float* scratch, int samples, int n_pad, int C);
int unfold_frames(cudaStream_t st, const __nv_bfloat16* x, __nv_bfloat16* out, int samples, int n_pad, int C, int taps);
// ---- fp32 DSP front/back end and the Vocos / VQ helpers (dsp.cu, vq.cu) ----
int melspec(cudaStream_t st, const float* wav, long long ld_wav, int batch, int nw, const float* window, const float* tw,
            const float* fbT, const int* fb_lo, const int* fb_hi, int n_mels, float* out);
int istft_head(cudaStream_t st, const float* spec, long long ld_spec, int batch, int t_pad, const int* n_frames,
               const float* window, const float* tw, float* out, long long ld_out, int out_len);
int layernorm_f32(cudaStream_t st, const float* x, int M, int D, const float* w, const float* b, float eps,
                  const int* row_len, int n_pad, float* out_f32, __nv_bfloat16* out_split);
int dwconv7_f32(cudaStream_t st, const float* x, const float* w, const float* b, float* out, int samples, int n_pad, int C);
int unfold_mel_split3(cudaStream_t st, const float* mel, long long ld_b, long long ld_c, long long ld_t, const int* t_off,
                      const int* n_frames, int batch, int t_pad, int C, int taps, int kpad, __nv_bfloat16* out);
int split3(cudaStream_t st, const float* x, long long ld, int M, int K, __nv_bfloat16* out);
int softmax_split3(cudaStream_t st, const float* logits, int ld, int M, int V, int kpad, __nv_bfloat16* out);
int vq_select(cudaStream_t st, const float* logits, int ld_logits, const float* x, int ld_x, int dim, const float* W,
              const float* bias, const float* vars, int M, int G, int V, int vd, int combine_groups, float refine_tol,
              long long* targets, void* xq, int xq_bf16, float* sums, int* hist, int* refined, float* perplexity);
// one-time per-device kernel attribute setup (opt-in shared memory sizes); idempotent
int ensure_configured();
int gemm_configure();
int attention_configure();
int conv_pos_configure();
const char* last_error();

// ---- launch accounting / in-stream kernel timing (bench.py's roofline leg) --------------------------
// Every launcher opens a LaunchScope: it bumps the library's launch counter and, when profiling of its
// class is enabled (f5e_prof_begin), brackets the launch with a CUDA event pair on the launching stream.
enum ProfClass : int { PROF_GEMM = 1, PROF_ATTN = 2, PROF_CONV = 4, PROF_NORM = 8, PROF_ELEM = 16, PROF_DSP = 32,
                       PROF_VQ = 64 };
// Programmatic dependent launch of the Euler-loop kernels (common.cuh: pdl_wait / pdl_trigger): on inside a PdlScope(true)
// of the calling thread (dit.cu opens one for small solves); F5E_PDL=0 / 1 in the environment (read once) forces it.
bool pdl_enabled(bool heavy = false);  // heavy: GEMM / attention / conv_pos (tcgen05 kernels with a TMEM prologue)
struct PdlScope {
  explicit PdlScope(bool on);
  ~PdlScope();
  bool prev;
};
// <<<grid, block, smem, st>>> with the programmatic-stream-serialization attribute when pdl_enabled().  ONLY for kernels
// whose every thread calls pdl_wait() before its first global-memory access.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(smem > 48 * 1024) ? 1 : 0;  // the heavy kernels are the ones with opt-in shared memory
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

long long launch_count();         // kernels enqueued so far (f5e_launch_count)
void add_launches(long long n);   // a replayed CUDA graph accounts for the kernels it holds
struct LaunchScope {
  LaunchScope(cudaStream_t st, int cls, long long a = 0, long long b = 0, long long c = 0);
  ~LaunchScope();
  cudaStream_t st;
  int rec;
};

}  // namespace f5e
