// Host orchestration of one DiT flow evaluation and of the Euler/CFG ODE loop.  Every step below is one
// of the hand-written kernels of this library; there are no library GEMMs and no host synchronisation.
#include "../../include/f5e_b200.h"
#include "host.h"
#include <mutex>
#include <vector>

using namespace f5e;

#define F5E_TRY(expr)       \
  do {                      \
    int _rc = (expr);       \
    if (_rc) return _rc;    \
  } while (0)

static_assert(sizeof(f5e_fold_job) == sizeof(FoldJob), "f5e_fold_job must mirror FoldJob");
#ifndef F5E_PDL_MAX_ROWS
#define F5E_PDL_MAX_ROWS 6144
#endif

// ---- cache of instantiated Euler-loop graphs (f5e_ode_solve_guided, use_graph != 0) ----------------------------------
namespace {
struct GraphKey {  // two independent 64-bit FNV-1a style hashes over the bytes of every argument the graph bakes in
  uint64_t h = 1469598103934665603ull, h2 = 0x9E3779B97F4A7C15ull;
  void add(const void* p, size_t n) {
    const unsigned char* c = static_cast<const unsigned char*>(p);
    for (size_t i = 0; i < n; ++i) {
      h = (h ^ c[i]) * 1099511628211ull;
      h2 = (h2 + c[i] + 0x632BE59BD9B4E019ull) * 0xFF51AFD7ED558CCDull;
      h2 ^= h2 >> 29;
    }
  }
};
struct GraphEntry {
  uint64_t key, key2;
  cudaGraphExec_t exec;
  uint64_t last_use;
  long long launches;  // kernel nodes (counted while capturing): every replay adds them to f5e_launch_count
};
constexpr size_t GRAPH_CACHE_ENTRIES = 8;
std::mutex g_graph_mu;
std::vector<GraphEntry> g_graphs;
uint64_t g_graph_clock = 0;
}  // namespace

static int dit_forward_impl(cudaStream_t st, const f5e_dit_model* m, const f5e_dit_buffers* b, const float* mod,
                            int mod_stride, const int* klen, int samples, int n_pad, const f5e_dit_packing* pack,
                            bool folded = false) {
  const int D = m->dim, inner = m->heads * 64;
  int rows = samples * n_pad;
  auto bf = [](const void* p) { return reinterpret_cast<const __nv_bfloat16*>(p); };
  auto bfm = [](void* p) { return reinterpret_cast<__nv_bfloat16*>(p); };

  // InputEmbedding.proj over the concat buffer (backbones/dit.py:172-175)
  {
    EpiParams ep{};
    ep.out = b->h0; ep.ldo = D; ep.bias = m->b_in; ep.n_pad = n_pad;
    if (m->isolate_samples) ep.row_len = klen;  // padded frames read as zeros by the convolutions below
    F5E_TRY(gemm_bf16(st, bf(b->x_cat), m->in_dim, bf(m->w_in), m->in_dim, rows, D, m->in_dim, ep));
  }
  // ConvPositionEmbedding: conv+Mish, conv+Mish, + residual (modules.py:167-190, dit.py:176)
  {
    EpiParams ep{};
    ep.out = b->h1; ep.ldo = D; ep.bias = m->b_conv1; ep.act = ACT_MISH; ep.n_pad = n_pad;
    if (m->isolate_samples) ep.row_len = klen;
    F5E_TRY(conv_pos(st, bf(b->h0), bf(m->w_conv1), samples, n_pad, D, m->conv_groups, ep));
    ep.out = b->x; ep.bias = m->b_conv2; ep.residual = bf(b->h0); ep.ldr = D;
    F5E_TRY(conv_pos(st, bf(b->h1), bf(m->w_conv2), samples, n_pad, D, m->conv_groups, ep));
  }
  // Everything up to here runs on the padded [samples, n_pad] layout (the reference's unmasked convolution sees the
  // padded frames).  From here on no op mixes frames except attention, which masks padded keys: with a packing the 22
  // blocks run on a permuted row order, valid frames first.  `rows` row-local work (norms, feed-forward), `rows_att`
  // rows of attention-side work (QKV, attention, out-projection): the reference zeroes that branch on padded rows, so
  // skipping it there leaves them exactly as they are.  A dropping packing (rows_total == 0) has rows == rows_att.
  void* xs = b->x;  // residual stream
  const int* row_info = nullptr;
  const int* row_start = nullptr;
  const int* row_len = klen;
  int rows_att = rows;
  if (pack) {
    rows = pack->rows_total > 0 ? pack->rows_total : pack->rows_packed;
    rows_att = pack->rows_packed;
    F5E_TRY(gather_rows(st, bf(b->x), bfm(b->h0), pack->row_map, rows, D));
    xs = b->h0;
    row_info = pack->row_info;
    row_start = pack->row_start;
    row_len = nullptr;  // the out-projection never touches a padded row
  }
  if (m->w_long_skip)  // the residual of the long skip: the stream as it enters the blocks (h1 is free after the convs)
    F5E_CUDA_CHECK(cudaMemcpyAsync(b->h1, xs, (size_t)rows * D * sizeof(__nv_bfloat16), cudaMemcpyDeviceToDevice, st));
  // Folded path: the out-projection / FF2 GEMMs take the LayerNorm statistics of the rows they produce in their epilogue
  // (EpiParams::stats_out) whenever they run on the CTA-pair kernel; row_stats then reads the residual stream only once per
  // evaluation (before the first block) instead of twice per block.
  // Only FF2 by default: its K = 2 * dim main loop hides the heavier epilogue completely; the out-projection (K = dim) is
  // epilogue-bound already and slows down by more than the row_stats pass it would save (measured: +12 us per launch).
#ifndef DIT_FUSE_STATS_OUT
#define DIT_FUSE_STATS_OUT 0
#endif
  const int slots = D / 128;
  const bool fuse_stats = folded && b->stats_part && D % 256 == 0 && gemm_can_fuse_stats(rows, D);
  const bool fuse_stats_out = fuse_stats && DIT_FUSE_STATS_OUT && rows_att == rows;
  auto block_stats = [&](bool from_gemm) -> int {
    if (from_gemm)
      return ln_finalize(st, reinterpret_cast<const float2*>(b->stats_part), slots, rows, D, 1e-6f, b->stats, bfm(b->stats_a));
    return row_stats(st, bf(xs), b->stats, bfm(b->stats_a), rows, D, 1e-6f);
  };
  for (int l = 0; l < m->depth; ++l) {
    const f5e_dit_layer& L = m->layers[l];
    const float* ml = mod + (size_t)l * 6 * D;
    const float *shift_msa = ml, *scale_msa = ml + D, *gate_msa = ml + 2 * D;
    const float *shift_mlp = ml + 3 * D, *scale_mlp = ml + 4 * D, *gate_mlp = ml + 5 * D;
    // adaLN-modulated LayerNorm (modules.py:310-315): materialised (norm_mod) or folded into the GEMM (row_stats + W')
    if (folded) F5E_TRY(block_stats(fuse_stats && l > 0));
    else F5E_TRY(norm_mod(st, bf(xs), bfm(b->xn), rows, D, scale_msa, shift_msa, mod_stride, n_pad, 1.0f, 1e-6f, row_info));
    {  // fused q/k/v projection + RoPE (modules.py:442-480)
      EpiParams ep{};
      ep.out = b->qkv; ep.ldo = 3 * inner; ep.bias = L.b_qkv; ep.n_pad = n_pad; ep.row_info = row_info;
      const bool qk_norm = L.q_norm_w != nullptr;  // then RMSNorm and the rotation run as one in-place pass after the GEMM
      if (m->rope_heads > 0 && !qk_norm) {
        ep.rope_inv_freq = m->rope_inv_freq; ep.rope_inner = inner; ep.rope_heads = m->rope_heads;
      }
      if (folded) {
        ep.ln_rstd = b->stats; ep.ln_a2 = bf(b->stats_a); ep.bias = L.d_qkv;
        F5E_TRY(gemm_bf16(st, bf(xs), D, bf(L.wf_qkv), D + 64, rows_att, 3 * inner, D + 64, ep));
      } else {
        F5E_TRY(gemm_bf16(st, bf(b->xn), D, bf(L.w_qkv), D, rows_att, 3 * inner, D, ep));
      }
    }
    if (L.q_norm_w)
      F5E_TRY(qk_norm_rope(st, bfm(b->qkv), 3 * inner, rows_att, inner, m->heads, L.q_norm_w, L.k_norm_w, 1e-6f,
                           m->rope_heads > 0 ? m->rope_inv_freq : nullptr, m->rope_heads, n_pad, row_info));
    F5E_TRY(attention(st, bf(b->qkv), 3 * inner, bfm(b->attn), inner, klen, samples, n_pad, m->heads, 0.125f, row_start, rows_att));
    {  // to_out + masked_fill + gate_msa * attn + residual (modules.py:494-501, :634)
      EpiParams ep{};
      ep.out = xs; ep.ldo = D; ep.bias = L.b_out; ep.n_pad = n_pad; ep.row_info = row_info; ep.row_len = row_len;
      ep.gate = gate_msa; ep.gate_stride = mod_stride; ep.residual = bf(xs); ep.ldr = D;
      if (fuse_stats_out) { ep.stats_out = reinterpret_cast<float2*>(b->stats_part); ep.stats_slots = slots; }
      F5E_TRY(gemm_bf16(st, bf(b->attn), inner, bf(L.w_out), inner, rows_att, D, inner, ep));
    }
    if (folded) F5E_TRY(block_stats(fuse_stats_out));
    else F5E_TRY(norm_mod(st, bf(xs), bfm(b->xn), rows, D, scale_mlp, shift_mlp, mod_stride, n_pad, 1.0f, 1e-6f, row_info));
    {  // FeedForward (modules.py:342-353) with GELU(tanh) fused, then gate_mlp * ff + residual (:638)
      EpiParams ep{};
      ep.out = b->ff; ep.ldo = m->ff_dim; ep.bias = L.b_ff1; ep.act = ACT_GELU_TANH; ep.n_pad = n_pad; ep.row_info = row_info;
      if (folded) {
        ep.ln_rstd = b->stats; ep.ln_a2 = bf(b->stats_a); ep.bias = L.d_ff1;
        F5E_TRY(gemm_bf16(st, bf(xs), D, bf(L.wf_ff1), D + 64, rows, m->ff_dim, D + 64, ep));
      } else {
        F5E_TRY(gemm_bf16(st, bf(b->xn), D, bf(L.w_ff1), D, rows, m->ff_dim, D, ep));
      }
      EpiParams e2{};
      e2.out = xs; e2.ldo = D; e2.bias = L.b_ff2; e2.n_pad = n_pad; e2.row_info = row_info;
      e2.gate = gate_mlp; e2.gate_stride = mod_stride; e2.residual = bf(xs); e2.ldr = D;
      if (fuse_stats && l + 1 < m->depth) { e2.stats_out = reinterpret_cast<float2*>(b->stats_part); e2.stats_slots = slots; }
      F5E_TRY(gemm_bf16(st, bf(b->ff), m->ff_dim, bf(L.w_ff2), m->ff_dim, rows, D, m->ff_dim, e2));
    }
  }
  if (m->w_long_skip) {  // x = W[:, :D] x + W[:, D:] residual  (Linear(2 dim -> dim, no bias) over the concatenation)
    const __nv_bfloat16* W = bf(m->w_long_skip);
    EpiParams e1{};
    e1.out = b->xn; e1.ldo = D; e1.n_pad = n_pad;
    F5E_TRY(gemm_bf16(st, bf(b->h1), D, W + D, 2 * D, rows, D, D, e1));
    EpiParams e2{};
    e2.out = b->h1; e2.ldo = D; e2.n_pad = n_pad; e2.residual = bf(b->xn); e2.ldr = D;
    F5E_TRY(gemm_bf16(st, bf(xs), D, W, 2 * D, rows, D, D, e2));
    xs = b->h1;
  }
  // AdaLayerNorm_Final (chunk order scale, shift: modules.py:331-336) + proj_out (dit.py:469-470)
  const float* mf = mod + (size_t)m->depth * 6 * D;
  F5E_TRY(norm_mod(st, bf(xs), bfm(b->xn), rows, D, mf, mf + D, mod_stride, n_pad, 1.0f, 1e-6f, row_info));
  {
    EpiParams ep{};
    ep.out = b->pred; ep.ldo = m->mel_dim; ep.out_f32 = 1; ep.bias = m->b_proj; ep.n_pad = n_pad; ep.row_info = row_info;
    F5E_TRY(gemm_bf16(st, bf(b->xn), D, bf(m->w_proj), D, rows, m->mel_dim, D, ep));
  }
  return 0;
}

static int check_model(const f5e_dit_model* m, const f5e_dit_buffers* b) {
  F5E_REQUIRE(m && b && m->layers, "dit: null model/buffers");
  F5E_REQUIRE(m->dim % 256 == 0, "dit: dim %d must be a multiple of 256", m->dim);
  F5E_REQUIRE(m->in_dim % 8 == 0 && m->mel_dim % 4 == 0, "dit: in_dim %d must be a multiple of 8, mel_dim %d of 4",
              m->in_dim, m->mel_dim);
  F5E_REQUIRE(m->rope_heads == 0 || m->rope_inv_freq, "dit: rope_inv_freq is null");
  F5E_REQUIRE(m->isolate_samples == 0 || m->isolate_samples == 1, "dit: isolate_samples must be 0 or 1");
  F5E_REQUIRE(b->x_cat && b->h0 && b->h1 && b->x && b->xn && b->qkv && b->attn && b->ff && b->pred, "dit: null scratch buffer");
  return 0;
}

static int check_pack(const f5e_dit_packing* p, const int* klen, int samples, int n_pad) {
  if (!p) return 0;
  F5E_REQUIRE(klen != nullptr, "dit: a row packing needs klen");
  F5E_REQUIRE(p->row_map && p->row_info && p->row_start, "dit: null packing array");
  F5E_REQUIRE(p->rows_packed > 0 && p->rows_packed <= samples * n_pad, "dit: rows_packed=%d out of range", p->rows_packed);
  F5E_REQUIRE(p->rows_total == 0 || p->rows_total == p->rows_packed || p->rows_total == samples * n_pad,
              "dit: rows_total=%d must be 0, rows_packed or samples * n_pad = %d", p->rows_total, samples * n_pad);
  F5E_REQUIRE(n_pad <= 65535, "dit: packed rows need n_pad <= 65535");
  return 0;
}

extern "C" {

int f5e_dit_forward(const f5e_dit_model* model, const f5e_dit_buffers* buf, const float* mod, int32_t mod_stride,
                    const int32_t* klen, int32_t samples, int32_t n_pad, const f5e_dit_packing* pack, void* stream) {
  F5E_TRY(check_model(model, buf));
  F5E_REQUIRE(mod && samples > 0 && n_pad > 0, "f5e_dit_forward: bad arguments");
  F5E_TRY(check_pack(pack, klen, samples, n_pad));
  return dit_forward_impl((cudaStream_t)stream, model, buf, mod, mod_stride, klen, samples, n_pad, pack);
}

int f5e_ode_solve(const f5e_dit_model* model, const f5e_dit_buffers* buf, float* y, const float* mod, int32_t mod_size,
                  const float* dt_host, int32_t steps, float cfg_strength, const int32_t* klen, int32_t batch,
                  int32_t n_pad, float* traj, int32_t round_bf16, int32_t use_graph, const f5e_dit_packing* pack,
                  void* stream) {
  f5e_guidance g{cfg_strength >= 1e-5f ? 2 : 1, cfg_strength, 0.0f};
  return f5e_ode_solve_guided(model, buf, y, mod, mod_size, dt_host, steps, &g, klen, batch, n_pad, traj, round_bf16,
                              use_graph, pack, stream);
}

int f5e_ode_solve_guided(const f5e_dit_model* model, const f5e_dit_buffers* buf, float* y, const float* mod,
                         int32_t mod_size, const float* dt_host, int32_t steps, const f5e_guidance* g,
                         const int32_t* klen, int32_t batch, int32_t n_pad, float* traj, int32_t round_bf16,
                         int32_t use_graph, const f5e_dit_packing* pack, void* stream) {
  F5E_TRY(check_model(model, buf));
  F5E_REQUIRE(y && mod && dt_host && g && steps > 0 && batch > 0 && n_pad > 0, "f5e_ode_solve: bad arguments");
  F5E_REQUIRE(g->branches >= 1 && g->branches <= 3, "f5e_ode_solve: guidance.branches=%d (1, 2 or 3)", g->branches);
  cudaStream_t st = (cudaStream_t)stream;
  F5E_TRY(ensure_configured());  // attribute setup must not happen inside a stream capture
  const int nb = g->branches;
  const int samples = nb * batch;
  const int rows_state = batch * n_pad;
  F5E_TRY(check_pack(pack, klen, samples, n_pad));
  F5E_REQUIRE(!pack || pack->rows_packed % nb == 0, "f5e_ode_solve: every guidance branch must pack to the same rows");
  // packed: pred holds the valid frames back to back, branch after branch; row_map[0 .. rows_upd) maps the first
  // branch onto the padded state rows.  A split packing appends the padded frames the same way (rows_tail per branch).
  const int rows_upd = pack ? pack->rows_packed / nb : rows_state;
  const int* upd_map = pack ? pack->row_map : nullptr;
  const int rows_tail_all = pack && pack->rows_total > pack->rows_packed ? pack->rows_total - pack->rows_packed : 0;
  F5E_REQUIRE(rows_tail_all % nb == 0, "f5e_ode_solve: every guidance branch must pad to the same rows");
  const int rows_tail = rows_tail_all / nb;
  // every sample shares the time step here, so the LayerNorms can be folded into the QKV / FF1 GEMMs
  const bool folded = model->fold_jobs != nullptr;
  if (folded) {
    F5E_REQUIRE(buf->stats && buf->stats_a && model->fold_n_jobs == 2 * model->depth && model->fold_rows > 0, "f5e_ode_solve: bad fold set-up");
    for (int l = 0; l < model->depth; ++l) {
      const f5e_dit_layer& L = model->layers[l];
      F5E_REQUIRE(L.wf_qkv && L.c_qkv && L.d_qkv && L.wf_ff1 && L.c_ff1 && L.d_ff1, "f5e_ode_solve: layer %d lacks fold scratch", l);
      F5E_REQUIRE((L.q_norm_w == nullptr) == (L.k_norm_w == nullptr), "f5e_ode_solve: layer %d has only one of q_norm / k_norm", l);
    }
  }
  // Small solves (the batch-1 shape of infer_cli.py: ~160 launches of 5-25 us per NFE step): programmatic dependent launch,
  // every kernel's prologue and launch latency under the tail of its predecessor (2.36 -> 1.99 ms per NFE step on the stream,
  // 2.01 -> 1.93 in the graph).  Large solves lose 2 % to it (config 2: 17.5 -> 17.9 ms), hence the threshold.
  constexpr int PDL_MAX_ROWS = F5E_PDL_MAX_ROWS;
  const bool use_pdl = (pack ? (pack->rows_total > 0 ? pack->rows_total : pack->rows_packed) : samples * n_pad) < PDL_MAX_ROWS;
  auto enqueue = [&](cudaStream_t es) -> int {
    PdlScope pdl(use_pdl);
    for (int i = 0; i < steps; ++i) {
      const float* mod_i = mod + (size_t)i * mod_size;
      if (folded)
        F5E_TRY(fold_ln_weights(es, reinterpret_cast<const FoldJob*>(model->fold_jobs), model->fold_n_jobs,
                                model->fold_rows, model->dim, mod_i));
      F5E_TRY(dit_forward_impl(es, model, buf, mod_i, 0, klen, samples, n_pad, pack, folded));
      float* tr = traj ? traj + (size_t)i * rows_state * model->mel_dim : nullptr;
      F5E_TRY(guided_euler(es, buf->pred, model->mel_dim, y, (__nv_bfloat16*)buf->x_cat, model->in_dim, tr, rows_upd,
                           model->mel_dim, dt_host[i], nb, g->c0, g->c1, round_bf16, upd_map, rows_state));
      if (rows_tail > 0)
        F5E_TRY(guided_euler(es, buf->pred + (size_t)pack->rows_packed * model->mel_dim, model->mel_dim, y,
                             (__nv_bfloat16*)buf->x_cat, model->in_dim, tr, rows_tail, model->mel_dim, dt_host[i], nb, g->c0,
                             g->c1, round_bf16, upd_map + pack->rows_packed, rows_state));
    }
    return 0;
  };
  if (!use_graph) return enqueue(st);

  // CUDA-graph path.  The instantiated graph is CACHED on everything it bakes in (every pointer and scalar argument of
  // every launch derives from the arguments hashed below), so a caller that keeps its buffers (model.py does, per shape)
  // pays capture + instantiation once and afterwards ONE cudaGraphLaunch per solve with no host synchronisation:
  // this is what makes the batch-1 latency shape GPU-bound instead of launch-bound (~170 launches per NFE step).
  GraphKey key;
  key.add(model, sizeof(*model));
  key.add(model->layers, sizeof(f5e_dit_layer) * (size_t)model->depth);
  key.add(buf, sizeof(*buf));
  key.add(&y, sizeof(y)); key.add(&mod, sizeof(mod)); key.add(&mod_size, sizeof(mod_size));
  key.add(dt_host, sizeof(float) * (size_t)steps); key.add(&steps, sizeof(steps));
  key.add(g, sizeof(*g)); key.add(&klen, sizeof(klen)); key.add(&batch, sizeof(batch)); key.add(&n_pad, sizeof(n_pad));
  key.add(&traj, sizeof(traj)); key.add(&round_bf16, sizeof(round_bf16));
  if (pack) key.add(pack, sizeof(*pack));
  int device = 0;
  F5E_CUDA_CHECK(cudaGetDevice(&device));
  key.add(&device, sizeof(device));
  std::lock_guard<std::mutex> lock(g_graph_mu);
  for (auto& e : g_graphs)
    if (e.key == key.h && e.key2 == key.h2) {
      e.last_use = ++g_graph_clock;
      F5E_CUDA_CHECK(cudaGraphLaunch(e.exec, st));
      add_launches(e.launches);
      return 0;
    }
  // Capture on a private stream (the caller's may be the legacy default stream, which cannot be captured);
  // nothing executes during capture.
  cudaStream_t cap = nullptr;
  F5E_CUDA_CHECK(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaError_t ce = cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal);
  if (ce != cudaSuccess) {
    cudaStreamDestroy(cap);
    F5E_CUDA_CHECK(ce);
  }
  const long long launches0 = launch_count();
  int rc = enqueue(cap);
  const long long launches = launch_count() - launches0;  // (other threads' launches in between would be miscounted: accounting only)
  ce = cudaStreamEndCapture(cap, &graph);
  cudaStreamDestroy(cap);
  if (rc) {
    if (graph) cudaGraphDestroy(graph);
    return rc;
  }
  F5E_CUDA_CHECK(ce);
  ce = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  F5E_CUDA_CHECK(ce);
  if (g_graphs.size() >= GRAPH_CACHE_ENTRIES) {  // evict the least recently used one; it may still be running on `st`
    size_t lru = 0;
    for (size_t i = 1; i < g_graphs.size(); ++i)
      if (g_graphs[i].last_use < g_graphs[lru].last_use) lru = i;
    cudaDeviceSynchronize();
    cudaGraphExecDestroy(g_graphs[lru].exec);
    g_graphs.erase(g_graphs.begin() + (long)lru);
  }
  g_graphs.push_back(GraphEntry{key.h, key.h2, exec, ++g_graph_clock, launches});
  F5E_CUDA_CHECK(cudaGraphLaunch(exec, st));
  return 0;
}

int f5e_graph_cache_clear(void) {
  std::lock_guard<std::mutex> lock(g_graph_mu);
  cudaDeviceSynchronize();
  for (auto& e : g_graphs) cudaGraphExecDestroy(e.exec);
  const int n = (int)g_graphs.size();
  g_graphs.clear();
  return n;
}

int f5e_pack_cols(void* dst, int32_t ld_dst, int32_t col0, const void* src, int32_t src_is_f32, int32_t ld_src,
                  int32_t rows, int32_t cols, void* stream) {
  F5E_REQUIRE(dst, "f5e_pack_cols: null dst");
  return pack_cols((cudaStream_t)stream, (__nv_bfloat16*)dst, ld_dst, col0, src, src_is_f32, ld_src, rows, cols);
}
int f5e_text_gather(const int64_t* text, int32_t samples, int32_t nt, const void* table_bf16, const float* pos_table,
                    int32_t max_pos, void* out_bf16, int32_t n_pad, int32_t dim, int32_t drop, int32_t mask_padding,
                    int32_t add_pos, void* stream) {
  F5E_REQUIRE(table_bf16 && out_bf16 && (text || nt == 0) && (pos_table || !add_pos), "f5e_text_gather: null argument");
  return text_gather((cudaStream_t)stream, text, samples, nt, (const __nv_bfloat16*)table_bf16, pos_table, max_pos,
                     (__nv_bfloat16*)out_bf16, n_pad, dim, drop, mask_padding, add_pos);
}
int f5e_dwconv7(const void* x_bf16, const float* w, const float* b, void* out_bf16, int32_t samples, int32_t n_pad,
                int32_t C, void* stream) {
  F5E_REQUIRE(x_bf16 && w && b && out_bf16, "f5e_dwconv7: null argument");
  return dwconv7((cudaStream_t)stream, (const __nv_bfloat16*)x_bf16, w, b, (__nv_bfloat16*)out_bf16, samples, n_pad, C);
}
int f5e_grn(const void* x_bf16, const float* gamma, const float* beta, void* out_bf16, float* scratch, int32_t samples,
            int32_t n_pad, int32_t C, void* stream) {
  F5E_REQUIRE(x_bf16 && gamma && beta && out_bf16 && scratch, "f5e_grn: null argument");
  return grn((cudaStream_t)stream, (const __nv_bfloat16*)x_bf16, gamma, beta, (__nv_bfloat16*)out_bf16, scratch, samples,
             n_pad, C);
}

int f5e_unfold_frames(const void* x_bf16, void* out_bf16, int32_t samples, int32_t n_pad, int32_t C, int32_t taps,
                      void* stream) {
  F5E_REQUIRE(x_bf16 && out_bf16, "f5e_unfold_frames: null argument");
  return unfold_frames((cudaStream_t)stream, (const __nv_bfloat16*)x_bf16, (__nv_bfloat16*)out_bf16, samples, n_pad, C, taps);
}

}  // extern "C"
