// Host orchestration + C ABI of the fp32 front / back ends: MelSpec, Vocos decode, VQ quantize.
// The Vocos backbone keeps fp32 activations; every GEMM runs on tcgen05 as an fp32-accurate 3-term bf16 split
// (A' = [A_hi | A_hi | A_lo], W' = [W_hi | W_lo | W_hi] along K), so the waveform matches an fp32 run to ~1e-5.
#include "../../include/f5e_b200.h"
#include "host.h"

using namespace f5e;

#define F5E_TRY(expr)       \
  do {                      \
    int _rc = (expr);       \
    if (_rc) return _rc;    \
  } while (0)

extern "C" {

int f5e_melspec(const float* wav, int64_t ld_wav, int32_t batch, int32_t n_samples, const float* window,
                const float* twiddles, const float* fb_t, const int32_t* fb_lo, const int32_t* fb_hi, int32_t n_mels,
                float* out, void* stream) {
  F5E_REQUIRE(wav && window && twiddles && fb_t && fb_lo && fb_hi && out, "f5e_melspec: null argument");
  return melspec((cudaStream_t)stream, wav, ld_wav, batch, n_samples, window, twiddles, fb_t, fb_lo, fb_hi, n_mels, out);
}

int f5e_istft_head(const float* spec, int64_t ld_spec, int32_t batch, int32_t t_pad, const int32_t* n_frames,
                   const float* window, const float* twiddles, float* out, int64_t ld_out, int32_t out_len, void* stream) {
  F5E_REQUIRE(spec && window && twiddles && out, "f5e_istft_head: null argument");
  return istft_head((cudaStream_t)stream, spec, ld_spec, batch, t_pad, n_frames, window, twiddles, out, ld_out, out_len);
}

int f5e_vocos_decode(const f5e_vocos_model* m, const f5e_vocos_buffers* b, const float* mel, int64_t ld_b, int64_t ld_c,
                     int64_t ld_t, const int32_t* t_off, const int32_t* n_frames, int32_t batch, int32_t t_pad, float* out,
                     int64_t ld_out, int32_t out_len, void* stream) {
  F5E_REQUIRE(m && b && mel && out && m->blocks, "f5e_vocos_decode: null argument");
  F5E_REQUIRE(b->a_embed && b->x && b->h && b->a1 && b->a2 && b->spec, "f5e_vocos_decode: null scratch buffer");
  F5E_REQUIRE(m->dim % 64 == 0 && m->inter % 32 == 0 && m->k_embed % 64 == 0 && m->k_embed >= 7 * m->in_ch,
              "f5e_vocos_decode: unsupported widths (dim=%d inter=%d k_embed=%d)", m->dim, m->inter, m->k_embed);
  F5E_REQUIRE(m->n_fft == 1024 && m->hop == 256, "f5e_vocos_decode: the ISTFT head is specialised for n_fft 1024 / hop 256");
  cudaStream_t st = (cudaStream_t)stream;
  const int rows = batch * t_pad, D = m->dim, I = m->inter, NS = m->n_fft + 2, LDS = (NS + 3) & ~3;
  auto bf = [](const void* p) { return reinterpret_cast<const __nv_bfloat16*>(p); };
  auto bfm = [](void* p) { return reinterpret_cast<__nv_bfloat16*>(p); };
  // backbone.embed: Conv1d(in_ch, dim, 7, padding=3) as one GEMM over the unfolded mel; padded frames -> 0
  F5E_TRY(unfold_mel_split3(st, mel, ld_b, ld_c, ld_t, t_off, n_frames, batch, t_pad, m->in_ch, 7, m->k_embed, bfm(b->a_embed)));
  {
    EpiParams ep{};
    ep.out = b->h; ep.ldo = D; ep.out_f32 = 1; ep.bias = m->b_embed; ep.row_len = n_frames; ep.n_pad = t_pad;
    F5E_TRY(gemm_bf16(st, bf(b->a_embed), 3 * m->k_embed, bf(m->w_embed), 3 * m->k_embed, rows, D, 3 * m->k_embed, ep));
  }
  F5E_TRY(layernorm_f32(st, b->h, rows, D, m->ln0_w, m->ln0_b, 1e-6f, n_frames, t_pad, b->x, nullptr));
  for (int l = 0; l < m->n_layers; ++l) {  // ConvNeXtBlock: dwconv k7 -> LN -> Linear -> GELU(erf) -> Linear -> gamma -> + residual
    const f5e_vocos_block& B = m->blocks[l];
    F5E_TRY(dwconv7_f32(st, b->x, B.dw_w, B.dw_b, b->h, batch, t_pad, D));
    F5E_TRY(layernorm_f32(st, b->h, rows, D, B.ln_w, B.ln_b, 1e-6f, nullptr, t_pad, nullptr, bfm(b->a1)));
    {
      EpiParams ep{};
      ep.out = b->a2; ep.ldo = 3 * I; ep.split3 = I; ep.bias = B.b1; ep.act = ACT_GELU_ERF; ep.n_pad = t_pad;
      F5E_TRY(gemm_bf16(st, bf(b->a1), 3 * D, bf(B.w1), 3 * D, rows, I, 3 * D, ep));
    }
    {
      EpiParams ep{};
      ep.out = b->x; ep.ldo = D; ep.out_f32 = 1; ep.bias = B.b2; ep.row_len = n_frames; ep.n_pad = t_pad;
      ep.gate = B.gamma; ep.residual = reinterpret_cast<const __nv_bfloat16*>(b->x); ep.ldr = D; ep.res_f32 = 1;
      F5E_TRY(gemm_bf16(st, bf(b->a2), 3 * I, bf(B.w2), 3 * I, rows, D, 3 * I, ep));
    }
  }
  F5E_TRY(layernorm_f32(st, b->x, rows, D, m->lnf_w, m->lnf_b, 1e-6f, nullptr, t_pad, nullptr, bfm(b->a1)));
  {  // head.out: Linear(dim, n_fft + 2) -> (log-magnitude | phase)
    EpiParams ep{};
    ep.out = b->spec; ep.ldo = LDS; ep.out_f32 = 1; ep.bias = m->b_head; ep.n_pad = t_pad;
    F5E_TRY(gemm_bf16(st, bf(b->a1), 3 * D, bf(m->w_head), 3 * D, rows, NS, 3 * D, ep));
  }
  return istft_head(st, b->spec, LDS, batch, t_pad, n_frames, m->window, m->twiddles, out, ld_out, out_len);
}

int f5e_vq_quantize(const float* x, int32_t ld_x, int32_t M, int32_t dim, const void* w_split3, const float* w_f32,
                    const float* bias, const float* vars, int32_t groups, int32_t num_vars, int32_t var_dim,
                    int32_t combine_groups, float refine_tol, void* a_split3, float* logits, int32_t ld_logits,
                    int64_t* targets, void* xq, int32_t xq_bf16, float* prob_sums, int32_t* hist, int32_t* n_refined,
                    float* perplexity, void* stream) {
  F5E_REQUIRE(x && w_split3 && w_f32 && vars && a_split3 && logits && targets && xq && prob_sums && hist,
              "f5e_vq_quantize: null argument");
  F5E_REQUIRE(dim % 64 == 0 && ld_logits % 4 == 0 && ld_logits >= groups * num_vars, "f5e_vq_quantize: bad leading dims");
  cudaStream_t st = (cudaStream_t)stream;
  F5E_TRY(split3(st, x, ld_x, M, dim, reinterpret_cast<__nv_bfloat16*>(a_split3)));
  EpiParams ep{};
  ep.out = logits; ep.ldo = ld_logits; ep.out_f32 = 1; ep.bias = bias; ep.n_pad = M;
  F5E_TRY(gemm_bf16(st, reinterpret_cast<const __nv_bfloat16*>(a_split3), 3 * dim, reinterpret_cast<const __nv_bfloat16*>(w_split3),
                    3 * dim, M, groups * num_vars, 3 * dim, ep));
  return vq_select(st, logits, ld_logits, x, ld_x, dim, w_f32, bias, vars, M, groups, num_vars, var_dim, combine_groups,
                   refine_tol, reinterpret_cast<long long*>(targets), xq, xq_bf16, prob_sums, hist, n_refined, perplexity);
}

int f5e_vq_quantize_fused(const float* x, int32_t ld_x, int32_t M, int32_t dim, const void* w_split3_padded,
                          const float* bias_padded, const float* w_f32, const float* bias, const float* vars, int32_t groups,
                          int32_t num_vars, int32_t var_dim, int32_t combine_groups, float refine_tol, void* a_split3,
                          void* tie_list, int64_t* targets, void* xq, int32_t xq_bf16, float* prob_sums, int32_t* hist,
                          int32_t* n_refined, float* perplexity, void* stream) {
  F5E_REQUIRE(x && w_split3_padded && bias_padded && w_f32 && vars && a_split3 && tie_list && targets && xq && prob_sums &&
                  hist && n_refined,
              "f5e_vq_quantize_fused: null argument");
  F5E_REQUIRE(dim % 64 == 0 && groups >= 1 && groups <= 2 && num_vars >= 1 && num_vars <= 128,
              "f5e_vq_quantize_fused: needs dim %% 64 == 0 and 1-2 groups of at most 128 entries (else f5e_vq_quantize)");
  cudaStream_t st = (cudaStream_t)stream;
  F5E_TRY(split3(st, x, ld_x, M, dim, reinterpret_cast<__nv_bfloat16*>(a_split3)));
  F5E_CUDA_CHECK(cudaMemsetAsync(prob_sums, 0, (size_t)groups * num_vars * sizeof(float), st));
  F5E_CUDA_CHECK(cudaMemsetAsync(hist, 0, (size_t)groups * num_vars * sizeof(int), st));
  F5E_CUDA_CHECK(cudaMemsetAsync(n_refined, 0, sizeof(int), st));
  VqEpi vq{};
  vq.targets = reinterpret_cast<long long*>(targets); vq.xq = xq; vq.vars = vars; vq.sums = prob_sums; vq.hist = hist;
  vq.tie_list = reinterpret_cast<int2*>(tie_list); vq.tie_count = n_refined; vq.refine_tol = refine_tol;
  vq.G = groups; vq.V = num_vars; vq.vd = var_dim; vq.combine_groups = combine_groups; vq.xq_bf16 = xq_bf16;
  F5E_TRY(gemm_vq(st, reinterpret_cast<const __nv_bfloat16*>(a_split3), 3 * dim, reinterpret_cast<const __nv_bfloat16*>(w_split3_padded),
                  3 * dim, bias_padded, M, 3 * dim, vq));
  return vq_refine(st, x, ld_x, dim, w_f32, bias, vq, perplexity, M);
}

int f5e_softmax_split3(const float* logits, int32_t ld, int32_t M, int32_t V, int32_t kpad, void* out_bf16, void* stream) {
  F5E_REQUIRE(logits && out_bf16, "f5e_softmax_split3: null argument");
  return softmax_split3((cudaStream_t)stream, logits, ld, M, V, kpad, reinterpret_cast<__nv_bfloat16*>(out_bf16));
}

int f5e_split3(const float* x, int64_t ld, int32_t M, int32_t K, void* out_bf16, void* stream) {
  F5E_REQUIRE(x && out_bf16, "f5e_split3: null argument");
  return split3((cudaStream_t)stream, x, ld, M, K, reinterpret_cast<__nv_bfloat16*>(out_bf16));
}

}  // extern "C"
