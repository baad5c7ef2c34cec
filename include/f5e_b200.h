/* libf5e_b200.so - C ABI of the B200-native (sm_100a) F5E-TTS inference hot path.
 *
 * The reference (kaleo996/F5E-TTS) has no FFI: its boundary is the Python object protocol
 * (CFM.sample, DiT.sample, MelSpec.forward, GumbelVectorQuantizer.forward, vocoder.decode).  The
 * Python mirror of those objects lives in f5e_tts_b200/ and reaches the GPU ONLY through the entry
 * points below (ctypes).  Every entry point cites the reference code it replaces (paths relative to
 * the reference's src/f5_tts/).
 *
 * Conventions
 *  - all pointers are DEVICE pointers unless the name ends in _host; the caller owns every buffer;
 *  - `stream` is a cudaStream_t passed as void*; no call synchronises unless documented;
 *  - return 0 on success, negative on error; f5e_last_error() gives the message (thread local);
 *  - there is NO CPU fallback: a missing/incompatible device is an error.
 */
#ifndef F5E_B200_H
#define F5E_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* f5e_last_error(void);
/* sha256 over the sources and compiler flags this library was built from (f5e_tts_b200/build.py); the ctypes loader
 * refuses a library whose stamp differs from the sources next to it. */
const char* f5e_build_stamp(void);
/* 0 when the current device is compute capability 10.x, else an error (message says what it is). */
int f5e_check_device(void);
int f5e_version(void);

/* ---- launch accounting and in-stream kernel timing (used by bench.py; nothing here is on the data path) ----
 * Every kernel launch of the library bumps a counter.  f5e_prof_begin(mask) additionally brackets each launch whose
 * class is in `mask` with a CUDA event pair on the launching stream (skipped inside a graph capture);
 * f5e_prof_end() synchronises those events and returns the number of records; f5e_prof_read() returns one record:
 * class, elapsed ms and three launch dimensions (GEMM: M,N,K; attention: samples,n_pad,heads; ...). */
enum { F5E_PROF_GEMM = 1, F5E_PROF_ATTN = 2, F5E_PROF_CONV = 4, F5E_PROF_NORM = 8, F5E_PROF_ELEM = 16, F5E_PROF_DSP = 32,
       F5E_PROF_VQ = 64 };
int64_t f5e_launch_count(void);
int f5e_prof_begin(int32_t class_mask);
int64_t f5e_prof_end(void);
int f5e_prof_read(int64_t index, int32_t* cls, float* ms, int64_t* dims3);

/* ---- activation ids for f5e_epilogue.act ---- */
enum { F5E_ACT_NONE = 0, F5E_ACT_GELU_TANH = 1, F5E_ACT_GELU_ERF = 2, F5E_ACT_MISH = 3, F5E_ACT_RELU = 4,
       F5E_ACT_SILU = 5 };

/* Fused GEMM epilogue:  v = acc + bias ; RoPE(v) ; act(v) ; row-mask ; v *= gate ; v += residual ; store */
typedef struct f5e_epilogue {
  void* out;               /* [M, ldo] bf16, or fp32 when out_f32 != 0 */
  int32_t ldo;
  int32_t out_f32;
  const float* bias;       /* [N] or NULL */
  int32_t act;
  const float* gate;       /* [N] (or [samples, gate_stride]) or NULL */
  int32_t gate_stride;
  const void* residual;    /* bf16 [M, ldr] or NULL */
  int32_t ldr;
  const int32_t* row_len;  /* [M / n_pad] valid frames per sample, or NULL (rows beyond are zeroed) */
  int32_t n_pad;           /* frames per sample */
  int32_t rope_inner;      /* heads * 64 */
  int32_t rope_heads;      /* rotate heads [0, rope_heads) of Q and K */
  int32_t residual_f32;    /* residual is fp32 [M, ldr] */
  int32_t out_split3;      /* > 0: bf16 output as the 3-term split [hi | hi | lo] at columns c, c+out_split3,
                              c+2*out_split3: the A operand of a following fp32-accurate "bf16x3" GEMM whose weight
                              is packed [W_hi | W_lo | W_hi] along K (used by the fp32 Vocos / VQ paths) */
  const int32_t* row_info; /* [M] sample << 16 | frame for packed rows (see f5e_dit_packing), or NULL */
  const float* rope_inv_freq; /* [32] or NULL: interleaved-pair RoPE on the Q and K column blocks; the angles
                              frame * inv_freq[j] are evaluated in the epilogue (x_transformers RotaryEmbedding.inv_freq,
                              reference backbones/dit.py:245, modules.py:470-480) */
  const float* ln_rstd;    /* [M] or NULL: LayerNorm folded into this GEMM (see f5e_fold_job): v = acc * rstd[row] + bias[col] */
  const void* ln_a2;       /* bf16 [M, 64], with ln_rstd: A operand of the LAST K block (K = dim + 64), the row's -mean
                              as (hi, lo, hi, 0 ...); both written by f5e_row_stats */
  float* stats_out;        /* fp32 [M, stats_slots, 2] or NULL: for x += gate * (acc + bias) IN PLACE (residual == out, bf16) on the
                              CTA-pair kernel (M * N large enough: >= 74 tiles of 256 x 256, N % 256 == 0): the residual tile is
                              read (TMA-staged) instead of reduce-added and every epilogue thread writes the (mean, M2) of its
                              128 NEW output columns; f5e_ln_finalize turns the slots into ln_rstd / ln_a2 */
  int32_t stats_slots;     /* N / 128 */
} f5e_epilogue;

/* C[M,N] = epilogue(A[M,K] @ W[N,K]^T); bf16 operands, tcgen05 / TMEM fp32 accumulation.
 * Replaces F.linear at model/modules.py:442-449, :494, :342-353; backbones/dit.py:169-177, :470. */
int f5e_gemm_bf16(const void* A, int32_t lda, const void* W, int32_t ldw, int32_t M, int32_t N, int32_t K,
                  const f5e_epilogue* ep, void* stream);

/* out[r,:] = LN(x[r,:]; eps, no affine) * (plus_one + mul[s,:]) + add[s,:], s = r / n_pad when
 * mod_stride != 0 else shared.  bf16 in/out, fp32 statistics.  D must be a multiple of 256.
 * Replaces AdaLayerNorm / AdaLayerNorm_Final / ff_norm modulate (model/modules.py:301-336, :636)
 * and, with plus_one = 0, nn.LayerNorm with affine (modules.py:258). */
int f5e_norm_mod(const void* x, void* out, int32_t M, int32_t D, const float* mul, const float* add,
                 int32_t mod_stride, int32_t n_pad, float plus_one, float eps, void* stream);

/* Fused classifier-free-guidance combine + Euler update (model/cfm.py:431-447 + torchdiffeq Euler):
 *   v = pc + (pc - pu) * cfg ;  y += dt * v
 * pred: fp32 [2, rows, ld_pred] (cond half, then uncond half; when cfg_strength < 1e-5 only the
 * first half is read).  y: fp32 [rows, C] state, updated in place.  x_cat: bf16 [2*rows, ld_cat]
 * concat buffer whose first C columns receive the new state for BOTH halves (NULL to skip).
 * traj: optional fp32 [rows, C] snapshot.  round_bf16 != 0 reproduces a bf16-carried state. */
int f5e_cfg_euler(const float* pred, int32_t ld_pred, float* y, void* x_cat, int32_t ld_cat, float* traj,
                  int32_t rows, int32_t C, float dt, float cfg_strength, int32_t round_bf16, void* stream);

/* Guidance of the sampler: how the flow predictions of the `branches` stacked evaluations are combined.
 *   branches 1: v = p0                                   (cfg_strength < 1e-5, model/cfm.py:440-441)
 *   branches 2: v = p0 + (p0 - p1) * c0                  (CFM.sample: p0 = conditional, p1 = null, c0 = cfg_strength;
 *                                                         model/cfm.py:431-447)
 *   branches 3: v = c0 * (p0 - p1) + c1 * (p1 - p2) + p2 (CFM.sample_tts / sample_vc: p0 = speaker + content,
 *                                                         p1 = content only, p2 = null, c0 = alpha_spk,
 *                                                         c1 = alpha_txt / alpha_ppg; model/cfm.py:166-185, :298-317) */
typedef struct f5e_guidance {
  int32_t branches;
  float c0, c1;
} f5e_guidance;

/* f5e_cfg_euler for any guidance: pred fp32 [branches, rows, ld_pred]; the new state is written into the first C
 * columns of x_cat (bf16 [branches*rows, ld_cat]) for every branch. */
int f5e_guided_euler(const float* pred, int32_t ld_pred, float* y, void* x_cat, int32_t ld_cat, float* traj,
                     int32_t rows, int32_t C, float dt, const f5e_guidance* g, int32_t round_bf16, void* stream);

/* Non-causal flash attention, dim_head 64, keys n >= klen[sample] masked (klen NULL = no mask).
 * qkv: bf16 [samples*n_pad, ld_qkv] with column blocks Q | K | V of heads*64 each (RoPE already applied);
 * out: bf16 [samples*n_pad, ldo] (heads merged).  Query tiles entirely beyond klen are not written.
 * Replaces F.scaled_dot_product_attention + expanded bool mask at model/modules.py:483-492. */
int f5e_attention(const void* qkv, int32_t ld_qkv, void* out, int32_t ldo, const int32_t* klen, int32_t samples,
                  int32_t n_pad, int32_t heads, float scale, void* stream);

/* f5e_attention with a per-key additive term: scores = q.k_j + key_bias[(sample * heads + head) * ld_bias + j], then * scale.
 * key_bias fp32, 16-byte aligned, ld_bias a multiple of 4 covering whole 128-key blocks (>= ceil(n_pad / 128) * 128).
 * With f5e_relpos_prepare this is RelPositionMultiHeadedAttention.forward of the PPG Conformer
 * (ppg/wenet/transformer/attention.py:179-222; this wenet version applies no rel_shift). */
int f5e_attention_bias(const void* qkv, int32_t ld_qkv, void* out, int32_t ldo, const int32_t* klen, const float* key_bias,
                       int32_t ld_bias, int32_t samples, int32_t n_pad, int32_t heads, float scale, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * PPG extractor: streaming kernels of the Conformer encoder (ppg/wenet/transformer/)
 * ------------------------------------------------------------------------------------------------- */
/* Conv2dSubsampling2.conv (subsampling.py:81-84, 114-117): Conv2d(1, C, 3, stride 2) + ReLU over (frames, mel bins).
 * feats fp32 [batch, T, F] (utterance stride ld_feats), n_frames[b] valid frames (NULL: T); w fp32 [C, 9], bias [C].
 * out bf16 [batch * t_pad, Fo * C], Fo = (F - 3) / 2 + 1, column f * C + c (the Linear that follows takes its weight
 * columns in that order); rows >= (n_frames[b] - 3) / 2 + 1 of an utterance are zeros. */
int f5e_subsample_conv(const float* feats, int64_t ld_feats, const int32_t* n_frames, int32_t batch, int32_t T, int32_t F,
                       const float* w, const float* bias, int32_t C, void* out, int32_t t_pad, void* stream);

/* Rel-pos attention, rewritten: (q+u).k_j + (q+v).p_j = q.(k_j + p_j) + (u.k_j + v.p_j).  In place on the K block of
 * qkv (bf16 [samples * n_pad, 3 * heads * 64]): k_j += pos[j]; key_bias[(s * heads + h) * ld_bias + j] = u_h.k_j + v_h.pos_j.
 * pos fp32 [n_pad, heads * 64] = linear_pos(pos_emb) (input independent, built once per layer); u, v fp32 [heads, 64]. */
int f5e_relpos_prepare(void* qkv, int32_t ld_qkv, const float* pos, const float* u, const float* v, float* key_bias,
                       int32_t ld_bias, int32_t samples, int32_t n_pad, int32_t heads, void* stream);

/* ConvolutionModule between its two pointwise convs (convolution.py:104-127): GLU over channels, depthwise Conv1d(taps)
 * with left_pad frames of left context (symmetric: (taps - 1) / 2; causal: taps - 1), norm, swish.
 * g bf16 [samples * n_pad, 2 C] (value | gate) -> out bf16 [samples * n_pad, C].  The GLU output reads as zero at frames
 * >= len[s] and as pad_val[c] at frames < 0 (NULL: zero; the causal module pads BEFORE its biased pointwise conv, so its left
 * context is glu(bias): the caller passes that vector).  norm = eval BatchNorm1d: folded into w / b by the caller (ln_w NULL);
 * norm = LayerNorm over channels: ln_w / ln_b fp32 [C], ln_eps, C <= 256.  Rows >= len[s] are written as zeros. */
int f5e_glu_dwconv(const void* g, const float* w, const float* b, const int32_t* len, void* out, int32_t samples,
                   int32_t n_pad, int32_t C, int32_t taps, int32_t left_pad, const float* pad_val, const float* ln_w,
                   const float* ln_b, float ln_eps, void* stream);

/* Grouped Conv1d(dim, dim, k=31, groups, padding=15) over frames, per sample, as a tcgen05 implicit GEMM with the
 * fused epilogue (bias, Mish, residual).  x: bf16 [samples*n_pad, dim] contiguous; w_packed: bf16 [31*dim, 64]
 * with w_packed[tap*dim + co, ci] = weight[co, ci, tap] (ci >= dim/groups zero).  Frames outside [0, n_pad) of a
 * sample read as zero.  Replaces ConvPositionEmbedding's Conv1d+Mish pairs (model/modules.py:167-190). */
int f5e_conv_pos(const void* x, const void* w_packed, int32_t samples, int32_t n_pad, int32_t dim, int32_t groups,
                 const f5e_epilogue* ep, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Conditioning kernels (run twice per ODE solve: cond / uncond)
 * ------------------------------------------------------------------------------------------------- */
/* dst[r, col0 + c] = bf16(src[r, c]) for c < cols; src NULL writes zeros.  Builds the InputEmbedding
 * concat buffer [x | cond | text_embed | ppg_embed] (backbones/dit.py:169-175). */
int f5e_pack_cols(void* dst, int32_t ld_dst, int32_t col0, const void* src, int32_t src_is_f32, int32_t ld_src,
                  int32_t rows, int32_t cols, void* stream);
/* TextEmbedding front (backbones/dit.py:54-80): ids+1, truncate/pad to n_pad with the filler 0, embedding
 * gather, + sinusoid position table row min(n, max_pos-1), zero where (mask_padding && filler).  text: int64
 * [samples, nt] padded with -1.  drop != 0 embeds the filler everywhere (the mask still comes from the text). */
int f5e_text_gather(const int64_t* text, int32_t samples, int32_t nt, const void* table_bf16, const float* pos_table,
                    int32_t max_pos, void* out_bf16, int32_t n_pad, int32_t dim, int32_t drop, int32_t mask_padding,
                    int32_t add_pos, void* stream);
/* Depthwise Conv1d(C, C, k=7, padding=3) + bias over frames per sample (model/modules.py:250-252). */
int f5e_dwconv7(const void* x_bf16, const float* w, const float* b, void* out_bf16, int32_t samples, int32_t n_pad,
                int32_t C, void* stream);
/* GRN (model/modules.py:225-234) over all n_pad frames of each sample; scratch: fp32 [samples, ceil(n_pad / 64), C]
 * (slab partial sums, added in a fixed order: bit-reproducible wherever the buffers live). */
int f5e_grn(const void* x_bf16, const float* gamma, const float* beta, void* out_bf16, float* scratch, int32_t samples,
            int32_t n_pad, int32_t C, void* stream);
/* f5e_grn without a scratch argument and without any allocation (no counterpart in the reference; added
 * for callers that may not call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The slab partials live in a file-scope __device__ array of csrc/grn_static.cu: F5E_GRN_STATIC_MAX_FLOATS
 * floats = 4 MiB, held by every GPU that loads the library until the process exits.  Same kernels, same results (bit for bit)
 * as f5e_grn.  Returns -1 with a clear f5e_last_error() when samples * ceil(n_pad / 64) * C (= f5e_grn_static_scratch_floats)
 * exceeds the limit: use f5e_grn then.  Every call shares the array: calls that overlap on the GPU (two streams, or two
 * threads without ordering) RACE - serialise them.  Nothing in the array needs clearing between calls (every partial read
 * is written first by the same call).  The first call on a device looks the array's address up (cudaGetSymbolAddress; with
 * CUDA's lazy module loading the driver materialises the array at that moment): make one call outside any stream capture. */
#define F5E_GRN_STATIC_MAX_FLOATS (1 << 20)
int f5e_grn_static(const void* x_bf16, const float* gamma, const float* beta, void* out_bf16, int32_t samples,
                   int32_t n_pad, int32_t C, void* stream);
int64_t f5e_grn_static_scratch_floats(int32_t samples, int32_t n_pad, int32_t C);

/* out[s, n, tap*C + c] = x[s, n + tap - taps/2, c] (zero outside the sample): turns the dense k=5 Conv1d of
 * PPGEmbedding (backbones/dit.py:124-135) into one GEMM with K = taps*C. */
int f5e_unfold_frames(const void* x_bf16, void* out_bf16, int32_t samples, int32_t n_pad, int32_t C, int32_t taps,
                      void* stream);

/* ---------------------------------------------------------------------------------------------------
 * DiT evaluation and the Euler/CFG ODE loop
 * ------------------------------------------------------------------------------------------------- */
typedef struct f5e_dit_layer {     /* one DiTBlock (model/modules.py:610-641); weights bf16 [out, in], biases fp32 */
  const void* w_qkv;  const float* b_qkv;   /* to_q | to_k | to_v stacked: [3*inner, dim] */
  const void* w_out;  const float* b_out;   /* attn.to_out.0 : [dim, inner] */
  const void* w_ff1;  const float* b_ff1;   /* ff.ff.0.0     : [ff_dim, dim] */
  const void* w_ff2;  const float* b_ff2;   /* ff.ff.2       : [dim, ff_dim] */
  /* Optional scratch of the folded-LayerNorm path of f5e_ode_solve (all NULL = off; see f5e_fold_job): per-step
   * rewritten weights bf16 [3*inner, dim + 64] / [ff_dim, dim + 64] (ZERO-initialised by the caller; the 64 extra K
   * columns carry the row sum) and their fp32 column vectors c (row sums) and d (bias). */
  void* wf_qkv;  float* c_qkv;  float* d_qkv;
  void* wf_ff1;  float* c_ff1;  float* d_ff1;
  /* qk_norm = "rms_norm" (model/modules.py:391-398, 463-467): fp32 [64] weights of attn.q_norm / attn.k_norm, both NULL = off.
   * RMSNorm over dim_head on every head of Q and K before the rotary embedding. */
  const float* q_norm_w;  const float* k_norm_w;
} f5e_dit_layer;

/* One entry of the DEVICE-resident job list of the folded-LayerNorm path.  In the Euler loop every sample shares the
 * time step, so AdaLayerNorm's (1 + scale) is one vector per layer and step (model/modules.py:310-315, :636) and
 *   (LN(x) * (1 + scale) + shift) @ W^T + b  ==  rstd_r * (x_r . W'_n - mean_r * c_n) + d_n,
 *   W' = bf16(W * (1 + scale)),  c_n = sum_k W'_nk,  d_n = b_n + sum_k shift_k W_nk :
 * the QKV / FF1 GEMMs read the residual stream itself and the normalised copy is never written; the mean term is one
 * more UMMA_K step ((-mean) split in bf16 against c_n split in bf16), the epilogue is v * rstd + d.  Once per step ONE
 * kernel rewrites W', c, d of every job from that step's modulation vector (scale_off / shift_off index into it). */
typedef struct f5e_fold_job {
  const void* w;      /* bf16 [N, dim] */
  const float* bias;  /* [N] or NULL  */
  void* wf;           /* bf16 [N, dim + 64] out: W' | (c_hi, c_hi, c_lo, 0 ...) */
  float* c;           /* [N] out */
  float* d;           /* [N] out */
  int32_t N, row0, scale_off, shift_off;  /* row0 = first row of this job in the launch (prefix sum of N) */
} f5e_fold_job;

typedef struct f5e_dit_model {
  int32_t dim, depth, heads, ff_dim, in_dim, mel_dim, rope_heads, conv_groups;
  const void* w_in;    const float* b_in;     /* input_embed.proj [dim, in_dim] */
  const void* w_conv1; const float* b_conv1;  /* conv_pos_embed.conv1d.0, packed (see f5e_conv_pos) */
  const void* w_conv2; const float* b_conv2;  /* conv_pos_embed.conv1d.2, packed */
  const f5e_dit_layer* layers;                /* HOST array [depth] */
  const void* w_proj;  const float* b_proj;   /* proj_out [mel_dim, dim] */
  const float* rope_inv_freq;                 /* rotary_embed.inv_freq [32] fp32 (NULL only when rope_heads == 0) */
  const f5e_fold_job* fold_jobs;              /* DEVICE array [fold_n_jobs] (2 per layer: QKV, FF1) or NULL = no folding */
  int32_t fold_n_jobs, fold_rows;             /* fold_rows = sum of N over the jobs */
  const void* w_long_skip;                    /* long_skip_connection.weight bf16 [dim, 2 * dim] or NULL (backbones/dit.py:264,
                                               * 456-467): x = W @ cat(x_after_blocks, x_before_blocks) ahead of norm_out */
  int32_t isolate_samples;                    /* != 0 (needs klen): frames >= klen[s] are ZEROED at the inputs of both
                                               * ConvPositionEmbedding convs, so a sample's result does not depend on
                                               * what it is batched with (a padded batch then equals the per-sample
                                               * solves of infer/utils_infer.py:474-482).  0 = the reference's batched
                                               * behaviour: padded frames leak into the last 15 valid ones */
} f5e_dit_model;

typedef struct f5e_dit_buffers {   /* caller-allocated scratch, rows = samples * n_pad */
  void* x_cat;    /* bf16 [rows, in_dim]  concat input (columns [0, mel_dim) = current state)      */
  void* h0;       /* bf16 [rows, dim]                                                              */
  void* h1;       /* bf16 [rows, dim]                                                              */
  void* x;        /* bf16 [rows, dim]     residual stream                                          */
  void* xn;       /* bf16 [rows, dim]                                                              */
  void* qkv;      /* bf16 [rows, 3*heads*64]                                                       */
  void* attn;     /* bf16 [rows, heads*64]                                                         */
  void* ff;       /* bf16 [rows, ff_dim]                                                           */
  float* pred;    /* fp32 [rows, mel_dim]                                                          */
  float* stats;   /* fp32 [rows] rstd of the folded-LayerNorm path; NULL when fold_jobs is NULL                    */
  void* stats_a;  /* bf16 [rows, 64], ZERO-initialised by the caller: (-mean_hi, -mean_lo, -mean_hi, 0 ...) per row */
  float* stats_part; /* fp32 [rows, dim / 128, 2] or NULL: (mean, M2) partials written by the epilogues of the out-projection /
                      * FF2 GEMMs of the folded path (they then take the next LayerNorm's statistics from the rows they
                      * produce, and the residual stream is not re-read for them); NULL = one row_stats pass per norm */
} f5e_dit_buffers;

/* Optional row packing: after the (unmasked, hence padded-layout) ConvPositionEmbedding nothing but the key-masked
 * attention mixes frames, so the DiT blocks can run on a PERMUTED row order with the valid frames first.
 * rows_packed = sum of klen; row_map[r] = row of packed row r in the padded [samples * n_pad] layout,
 * row_info[r] = sample << 16 | frame, row_start[s] = first packed row of sample s.
 * Two forms:
 *   rows_total == 0 (or == rows_packed): the padded frames are DROPPED.  Exact when no padded frame can reach a valid one
 *     (isolate_samples = 1, or all klen equal); otherwise the padded frames of the state stay at their initial value instead
 *     of evolving, which the reference's unmasked convolutions leak into the last ~30 valid frames of the shorter samples.
 *   rows_total == samples * n_pad: row_map / row_info list the valid rows first (sample after sample) and then the padded
 *     rows (sample after sample).  The attention sub-block (QKV projection, attention, out-projection: its result on a
 *     padded row is masked to zero by the reference, modules.py:499-501, so the row is unchanged) runs on the first
 *     rows_packed rows only; the row-local ops (LayerNorm, feed-forward, final norm + projection) on all of them.
 *     Bit-identical to the padded layout, without the attention-side GEMM FLOPs on padding.
 * With a packing, buf->pred holds [rows_total or rows_packed, mel_dim] in packed order, and in f5e_ode_solve the first
 * 1 / branches of the valid part (and of the padded part) of row_map addresses the state rows.
 * NULL = padded layout everywhere. */
typedef struct f5e_dit_packing {
  const int32_t* row_map;
  const int32_t* row_info;
  const int32_t* row_start;
  int32_t rows_packed;
  int32_t rows_total;
} f5e_dit_packing;

/* The two helper kernels of the folded-LayerNorm path (exported for unit tests; f5e_ode_solve calls them itself).
 * f5e_row_stats: per row of x (bf16 [M, D], D % 256 == 0): rstd[r] = 1 / sqrt(var + eps) and a2[r, 0..2] = bf16 split of
 * -mean as (hi, lo, hi), a2[r, 3..63] untouched (zero-initialised by the caller).  F.layer_norm's statistics,
 * model/modules.py:310-315.  f5e_fold_ln_weights: one launch rewrites W' | c | d of every job from `mod`. */
int f5e_row_stats(const void* x, float* rstd, void* a2, int32_t M, int32_t D, float eps, void* stream);
/* stats [M, slots, 2] written through f5e_epilogue.stats_out (slots * 128 == D) -> the same rstd / a2 f5e_row_stats gives. */
int f5e_ln_finalize(const float* stats, int32_t slots, int32_t M, int32_t D, float eps, float* rstd, void* a2, void* stream);
int f5e_fold_ln_weights(const f5e_fold_job* jobs_dev, int32_t n_jobs, int32_t total_rows, int32_t D, const float* mod,
                        void* stream);
/* dst[r, :] = src[row_map[r], :] (bf16, C % 8 == 0): the packing step of f5e_dit_packing. */
int f5e_gather_rows(const void* src, void* dst, const int32_t* row_map, int32_t rows, int32_t C, void* stream);

/* One flow evaluation over `samples` sequences of n_pad frames (DiT.sample after the embeddings:
 * backbones/dit.py:450-470).  mod: fp32 modulation vectors for this time step, laid out
 * [depth][shift_msa|scale_msa|gate_msa|shift_mlp|scale_mlp|gate_mlp][dim] then [scale|shift][dim] of norm_out
 * (AdaLayerNorm / AdaLayerNorm_Final chunk order, model/modules.py:310-315, 331-336); mod_stride = 0 when every
 * sample shares the time step, else the per-sample stride in floats.  klen: int32 [samples] valid frames
 * (attention key mask + zeroing of the attention branch on padded rows), NULL = no mask (batch 1).
 * Result in buf->pred. */
int f5e_dit_forward(const f5e_dit_model* model, const f5e_dit_buffers* buf, const float* mod, int32_t mod_stride,
                    const int32_t* klen, int32_t samples, int32_t n_pad, const f5e_dit_packing* pack, void* stream);

/* The whole sampler loop (model/cfm.py:431-447, 471 with torchdiffeq fixed-grid Euler): for i in [0, steps):
 *   pred = DiT(x_cat; mod[i]) on the doubled batch (cond half then uncond half; a single half when
 *   cfg_strength < 1e-5);  y += dt[i] * (pc + (pc - pu) * cfg);  state written back into x_cat for both halves.
 * y: fp32 [batch*n_pad, mel_dim] state (y0 on entry, y(1) on exit).  mod: fp32 [steps, mod_size].  dt_host: HOST
 * float [steps].  traj: optional fp32 [steps, batch*n_pad, mel_dim] (state after each step).  klen: int32
 * [2*batch] (both halves) or NULL.  use_graph != 0 runs the loop as ONE CUDA graph launch: the instantiated graph is
 * cached (8 entries, LRU) on every argument it bakes in - pointers, sizes, dt values, the model / layer / buffer structs -
 * so a caller that keeps its buffers pays capture + instantiation once per shape; no host synchronisation either way.
 * Solves below 6144 rows (branches * batch * n_pad) launch every kernel of the loop with programmatic dependent launch (each
 * kernel's prologue runs under the tail of its predecessor; results are bit-identical); the environment variable F5E_PDL,
 * read once per process, forces it off (0) or on (1) for every size.  f5e_launch_count() counts a replayed graph's kernels. */
int f5e_ode_solve(const f5e_dit_model* model, const f5e_dit_buffers* buf, float* y, const float* mod, int32_t mod_size,
                  const float* dt_host, int32_t steps, float cfg_strength, const int32_t* klen, int32_t batch,
                  int32_t n_pad, float* traj, int32_t round_bf16, int32_t use_graph, const f5e_dit_packing* pack,
                  void* stream);

/* The same loop under any guidance (CFM.sample_tts / CFM.sample_vc, model/cfm.py:93-346: three evaluations per step on
 * a TRIPLED batch, branch order speaker+content | content | null).  x_cat, klen and the packing cover
 * g->branches * batch samples; everything else as f5e_ode_solve. */
int f5e_ode_solve_guided(const f5e_dit_model* model, const f5e_dit_buffers* buf, float* y, const float* mod,
                         int32_t mod_size, const float* dt_host, int32_t steps, const f5e_guidance* g,
                         const int32_t* klen, int32_t batch, int32_t n_pad, float* traj, int32_t round_bf16,
                         int32_t use_graph, const f5e_dit_packing* pack, void* stream);

/* Destroys every cached Euler-loop graph (after a device synchronise); returns how many there were.  Call it before
 * freeing buffers or weights a cached graph refers to if the addresses may be reused for something else. */
int f5e_graph_cache_clear(void);

/* ---------------------------------------------------------------------------------------------------
 * Prompt preparation / chunk assembly around the sampler
 * ------------------------------------------------------------------------------------------------- */
/* Prompt loudness (eval/utils_eval.py:120-123, infer/utils_infer.py:441-443):
 *   rms[b] = sqrt(mean(wav[b, :lens[b]]^2));  out = rms < target_rms ? wav * target_rms / rms : wav  (zero past lens[b]).
 * wav / out: fp32 [batch, n_max] with row strides ld_wav / ld_out (out may be NULL: RMS only); lens NULL = n_max each;
 * scratch: f5e_rms_scratch_floats(batch, n_max) floats.  The sum of squares is reduced in a fixed order (deterministic). */
int f5e_rms_normalize(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max,
                      float target_rms, float* scratch, float* rms_out, float* out, int64_t ld_out, void* stream);
int f5e_rms_scratch_floats(int32_t batch, int32_t n_max);

/* torchaudio.transforms.Resample(orig, new) as called at eval/utils_eval.py:125-127 and infer/utils_infer.py:444-446
 * (third-party torchaudio; sinc_interp_hann, lowpass_filter_width 6, rolloff 0.99), as a polyphase FIR:
 *   out[b, q*up + p] = sum_k kern[p, k] * wav[b, q*down + k - width],  up = new/gcd, down = orig/gcd, taps = 2*width + down,
 * samples outside [0, lens[b]) read as zero; ceil(up * lens[b] / down) samples are produced per row, the rest of
 * out[b, :out_max] is zero.  kern: fp32 [up, taps], built by the caller like torchaudio's _get_sinc_resample_kernel. */
int f5e_resample(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max, const float* kern,
                 int32_t taps, int32_t down, int32_t up, int32_t width, float* out, int64_t ld_out, int32_t out_max,
                 void* stream);

/* One step of the chunk cross-fade chain of infer/utils_infer.py:525-556 (numpy, float64): the float64 wave `acc`
 * (n_acc samples) is extended in place by the fp32 chunk `next` (n_next samples): cf = min(cf_samples, n_acc, n_next),
 * the last cf samples become acc * linspace(1, 0, cf) + next[:cf] * linspace(0, 1, cf), the rest of `next` is appended;
 * *n_out = n_acc + n_next - cf (HOST).  apply_gain != 0 first scales the chunk like `generated_wave * rms / target_rms`
 * (utils_infer.py:495-496) with gain_num = rms, gain_den = target_rms in fp32.  Bit-identical to the numpy chain. */
int f5e_crossfade_append(double* acc, int64_t n_acc, const float* next, int32_t n_next, int32_t cf_samples,
                         float gain_num, float gain_den, int32_t apply_gain, int64_t* n_out, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * fp32 front end / back end / codebook
 * ------------------------------------------------------------------------------------------------- */
/* MelSpec.forward (model/modules.py:104-143 -> get_vocos_mel_spectrogram :75-101): reflect pad 512, periodic Hann 1024,
 * hop 256, |rFFT|, HTK mel filterbank, log(clamp(1e-5)).  wav: fp32 [batch, n_samples] (row stride ld_wav);
 * out: fp32 [batch, n_mels, n_samples / 256 + 1].  Tables (built by the caller once): window [1024]; twiddles
 * float2[512] = exp(-2 pi i j / 1024); fb_t [n_mels, 513] transposed filterbank with its non-zero bin range
 * [fb_lo[m], fb_hi[m]) per filter. */
int f5e_melspec(const float* wav, int64_t ld_wav, int32_t batch, int32_t n_samples, const float* window,
                const float* twiddles, const float* fb_t, const int32_t* fb_lo, const int32_t* fb_hi, int32_t n_mels,
                float* out, void* stream);

/* Front end of the PPG extractor: kaldiFbank.forward (ppg/wenet/dataset/feats.py:49-83) = wave * 2^15 ->
 * torchaudio.compliance.kaldi.fbank(num_mel_bins 80, frame_length 25, frame_shift 10, dither 0, energy_floor 0) and,
 * optionally fused, GlobalCMVN (ppg/wenet/transformer/cmvn.py:9-47; encoder.py:171-172): (x - mean) * istd.
 * wav fp32 [batch, n_max] (row stride ld_wav), lens[b] valid samples (NULL: n_max).  Per utterance
 * T = 1 + (len - win) / shift frames (snip_edges; 0 if len < win): DC removal, pre-emphasis, window[win] (povey, built
 * by the caller), zero pad to 512, |rFFT|^2, fb_t [n_mels, 257] mel banks (non-zero bins [fb_lo, fb_hi) per bank),
 * log(max(., FLT_EPSILON)).  twiddles: float2[256] = exp(-2 pi i j / 512).  out fp32 [batch, t_max, n_mels] (utterance
 * stride ld_out); frames >= T are zeros.  cmvn_mean / cmvn_istd: [n_mels] or both NULL. */
int f5e_kaldi_fbank(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max, float scale,
                    float preemph, int32_t win, int32_t shift, const float* window, const float* twiddles,
                    const float* fb_t, const int32_t* fb_lo, const int32_t* fb_hi, int32_t n_mels,
                    const float* cmvn_mean, const float* cmvn_istd, float* out, int64_t ld_out, int32_t t_max,
                    void* stream);

/* Vocos ISTFTHead after its Linear (published vocos architecture; in-tree corroboration
 * runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:45-59, scripts/conv_stft.py:193-234):
 * spec fp32 [batch, t_pad, ld_spec] rows = (log-magnitude[513] | phase[513]); frames >= n_frames[b] ignored (NULL: all).
 * out fp32 [batch, ld_out]: samples [0, (n_frames[b]-1)*256) = torch.istft(center=True) of exp/clip/cos/sin, rest of
 * [0, out_len) zero. */
int f5e_istft_head(const float* spec, int64_t ld_spec, int32_t batch, int32_t t_pad, const int32_t* n_frames,
                   const float* window, const float* twiddles, float* out, int64_t ld_out, int32_t out_len, void* stream);

typedef struct f5e_vocos_block {   /* ConvNeXtBlock; w1 / w2 are bf16 3-term splits [W_hi | W_lo | W_hi] along K */
  const float* dw_w; const float* dw_b;   /* dwconv [dim, 7], [dim] */
  const float* ln_w; const float* ln_b;
  const void* w1; const float* b1;        /* pwconv1 [inter, 3*dim] */
  const void* w2; const float* b2;        /* pwconv2 [dim, 3*inter] */
  const float* gamma;                     /* layer scale [dim] */
} f5e_vocos_block;

typedef struct f5e_vocos_model {
  int32_t dim, inter, n_layers, in_ch, k_embed, n_fft, hop;
  const void* w_embed; const float* b_embed;   /* backbone.embed as [dim, 3*k_embed], column tap*in_ch + c, zero padded */
  const float* ln0_w; const float* ln0_b;      /* backbone.norm */
  const f5e_vocos_block* blocks;               /* HOST array [n_layers] */
  const float* lnf_w; const float* lnf_b;      /* backbone.final_layer_norm */
  const void* w_head; const float* b_head;     /* head.out [n_fft + 2, 3*dim] */
  const float* window; const float* twiddles;
} f5e_vocos_model;

typedef struct f5e_vocos_buffers {   /* rows = batch * t_pad */
  void* a_embed;   /* bf16 [rows, 3*k_embed] */
  float* x;        /* fp32 [rows, dim] */
  float* h;        /* fp32 [rows, dim] */
  void* a1;        /* bf16 [rows, 3*dim] */
  void* a2;        /* bf16 [rows, 3*inter] */
  float* spec;     /* fp32 [rows, round_up(n_fft + 2, 4)] */
} f5e_vocos_buffers;

/* vocoder.decode (call sites infer/utils_infer.py:488-489, eval/eval_infer_batch.py:211-212) for a whole batch of
 * variable-length spans in one go: sample b decodes mel frames [t_off[b], t_off[b] + n_frames[b]) of
 * mel[b, c, t] = mel[b*ld_b + c*ld_c + t*ld_t] (so both [B,100,T] and [B,T,100] layouts work without a copy);
 * every span is decoded exactly as if it were alone (zero "same" padding at ITS ends).  out as in f5e_istft_head. */
int f5e_vocos_decode(const f5e_vocos_model* model, const f5e_vocos_buffers* buf, const float* mel, int64_t ld_b,
                     int64_t ld_c, int64_t ld_t, const int32_t* t_off, const int32_t* n_frames, int32_t batch,
                     int32_t t_pad, float* out, int64_t ld_out, int32_t out_len, void* stream);

/* GumbelVectorQuantizer.forward / .forward_idx in eval mode (model/modules.py:877-950): logits = x W^T + b on tcgen05
 * (bf16x3 split, fp32 accumulate), per-group first-max argmax (near ties re-decided with sequential fp32 FMAs so indices
 * are those of an fp32 implementation), code-vector gather, histogram / mean-softmax perplexities.
 * x fp32 [M, dim]; w_split3 bf16 [groups*num_vars, 3*dim]; w_f32 [groups*num_vars, dim]; vars fp32 [(combine ? 1 : groups)
 * * num_vars, var_dim]; scratch a_split3 bf16 [M, 3*dim], logits fp32 [M, ld_logits], prob_sums fp32 / hist int32
 * [groups*num_vars]; outputs targets int64 [M, groups], xq [M, groups*var_dim] (fp32 or bf16), perplexity[2] =
 * (code_perplexity, prob_perplexity), n_refined = number of near-tie re-decisions. */
int f5e_vq_quantize(const float* x, int32_t ld_x, int32_t M, int32_t dim, const void* w_split3, const float* w_f32,
                    const float* bias, const float* vars, int32_t groups, int32_t num_vars, int32_t var_dim,
                    int32_t combine_groups, float refine_tol, void* a_split3, float* logits, int32_t ld_logits,
                    int64_t* targets, void* xq, int32_t xq_bf16, float* prob_sums, int32_t* hist, int32_t* n_refined,
                    float* perplexity, void* stream);

/* fp32 [M, K] -> bf16 [M, 3K] = [hi | hi | lo]: the A operand of a bf16x3 GEMM (test hook / building block). */
/* The same quantize call with the decision FUSED into the epilogue of the logit GEMM (north_star (3): "distance GEMM fused
 * with argmin"): for 1-2 groups of at most 128 entries.  w_split3_padded: bf16 [256, 3*dim], the 3-term split of
 * weight_proj with every group padded to 128 rows (group g at rows [128 g, 128 g + num_vars), zeros beyond);
 * bias_padded: fp32 [256], the bias there and -1e30 on the pad rows.  An epilogue thread owns one row and one group: it
 * takes the arg-max (first maximum) straight out of TMEM, adds the row's softmax to the column sums, and the warp gathers
 * the code vectors - NO logits are written.  Rows whose top-2 margin is within refine_tol are appended to tie_list
 * (int32 pairs [M * groups, 2], scratch) and decided by a second small kernel in sequential fp32, exactly like
 * f5e_vq_quantize; *n_refined (device) is their count.  Indices are bit-identical to f5e_vq_quantize's. */
int f5e_vq_quantize_fused(const float* x, int32_t ld_x, int32_t M, int32_t dim, const void* w_split3_padded,
                          const float* bias_padded, const float* w_f32, const float* bias, const float* vars, int32_t groups,
                          int32_t num_vars, int32_t var_dim, int32_t combine_groups, float refine_tol, void* a_split3,
                          void* tie_list, int64_t* targets, void* xq, int32_t xq_bf16, float* prob_sums, int32_t* hist,
                          int32_t* n_refined, float* perplexity, void* stream);

int f5e_split3(const float* x, int64_t ld, int32_t M, int32_t K, void* out_bf16, void* stream);

/* Row softmax over the first V columns of fp32 logits [M, ld] -> bf16 [M, 3 * kpad] as the split [hi | hi | lo] (columns
 * >= V zero; kpad % 8 == 0): the A operand of the fp32-accurate GEMM with the phone-centre table in
 * PPGModelWapper.ppg_to_target, output_type "map" (ppg/ppg_model.py:113-118: logit.softmax(-1) @ global_phn_center). */
int f5e_softmax_split3(const float* logits, int32_t ld, int32_t M, int32_t V, int32_t kpad, void* out_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* F5E_B200_H */
