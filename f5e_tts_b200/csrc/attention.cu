// Non-causal flash attention for sm_100a, dim_head 64, per-sample key length (padding mask), bf16 I/O.
//
// Replaces F.scaled_dot_product_attention with the expanded [B,H,N,N] bool mask of reference
// model/modules.py:483-492: keys n >= klen[sample] are excluded, nothing of size N x N is ever stored.
//
// At dim_head 64 a 128 x 128 score block costs 512 tensor cycles but 1024 cycles of MUFU.EX2 (16 lanes / clk / SM,
// tools/ubench/tmem_bw.cu).  Design:
//   * PERSISTENT CTAs, two per SM.  A CTA walks a static list of work items (sample, head, 128-query tile); barriers, the
//     K/V ring and TMEM live across items, the next item's Q / K tiles are prefetched while the current item drains.
//     Each SM sub-partition hosts one softmax warp of either CTA: ONE warp reaches only ~70 % of the MUFU rate
//     (1450-1550 cycles per 128-exponential block), two sharing the pipe reach 96 % (1056 each) - which is why
//     making the exponential phases of the two warps mutually exclusive (FlashAttention-4's role split with a
//     ping-pong, built and measured in commit ebfddec: 0.277 ms against 0.243 here) is the wrong schedule for this chip.
//   * S_j is read from TMEM exactly once (128 fp32 registers, one wait) and released immediately so QK_{j+1} overlaps
//     the softmax math; O accumulates ACROSS key blocks in TMEM (tcgen05.mma accumulate) and is rescaled lazily
//     (FlashAttention-4 style: only when a row maximum grew by more than 2^8; P and the row sum use the same stale
//     reference, so O / l is exact).
//   * The last key block of a sample is narrowed to the 32-key chunks that hold real keys (QK with N = nc, PV with
//     K = nc, nc exponentials per row): 0.261 -> 0.243 ms at 1030 frames.
//   * Ragged batches: the work items are enumerated densely over the samples sorted by descending key length (AttSched,
//     built in shared memory by every CTA), so the static round-robin shares of the CTAs differ by one cheap item at most:
//     0.144 -> 0.128 ms for lengths in [343, 1030], 17.35 -> 17.10 ms per NFE step at config 2 (same box, same bits).
//   * (Evaluating part of the exponentials as a polynomial on the FMA pipe costs MORE time and more joules here -
//     0.270 / 0.288 ms sustained for 1/4 / 1/2 of them against 0.261 - although a MUFU.EX2 is ~64 pJ:
//     profiles/r02_attention.md.)
// Roles inside a CTA (single-thread issuers chosen with elect.sync: `if (lane == 0)` made the compiler wrap every
// UTCHMMA / UTMALDG in an ELECT / BRA.U.ANY loop, ~130 cycles per instruction):
//   warps 0-3 softmax: thread == query row; row max, ex2, row sum, bf16 P written back to TMEM (tcgen05.st).
//   warp 4    TMA producer: Q tile per item, K_j / V_j tiles (128 keys x 64, 128B swizzle) through a 4-slot ring.
//   warp 5    MMA issuer: S = Q K_j^T (UMMA 128 x nc x 16, K-major operands) into TMEM cols [0,128);
//             O += P_j V_j (UMMA 128x64x16, A = P from TMEM cols [192,256), B = V as an MN-major smem operand)
//             into TMEM cols [128,192).
// Input layout: qkv [rows, 3*inner] with Q | K | V column blocks (what the fused QKV GEMM writes,
// RoPE already applied); output [rows, inner].
#include "host.h"
#include <type_traits>

namespace f5e {

constexpr int ATT_BM = 128;   // queries per work item
constexpr int ATT_BN = 128;   // keys per block
constexpr int ATT_DH = 64;
constexpr int ATT_THREADS = 192;
constexpr int ATT_WARP_TMA = 4, ATT_WARP_MMA = 5;  // (issuers below or above the softmax warps: measured, no difference)
constexpr int ATT_KV_SLOTS = 4;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB
constexpr int ATT_SMEM_Q = 0;
constexpr int ATT_SMEM_KV = ATT_TILE_BYTES;
constexpr int ATT_SMEM_BAR = ATT_SMEM_KV + ATT_KV_SLOTS * ATT_TILE_BYTES;
constexpr int ATT_SMEM_SCHED = ATT_SMEM_BAR + 128;  // int [3 * 256 + 1]: clipped key lengths, sorted order, tile prefix
constexpr int ATT_SMEM_TOTAL = ATT_SMEM_SCHED + 3 * 1024 + 16 + 1024;
constexpr uint32_t ATT_TMEM_COLS = 256;
constexpr uint32_t ATT_TMEM_O = 128;   // S [0,128) fp32 | O [128,192) fp32 | P [192,256) bf16 pairs (128 keys)
constexpr uint32_t ATT_TMEM_P = 192;

// ATT_KO_* (developer knobs, WRONG results on purpose): knock one unit out to see what it costs in time and in joules
// (tools/attn_lab.py --sustain; profiles/r02_attention.md).
__device__ __forceinline__ float ex2_approx(float x) {
#ifdef ATT_KO_MUFU
  return x * 0.001f;
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}

#ifdef ATT_TRACE
// developer tool (tools/attn_lab.py --trace): per-block phase timestamps of the softmax warps and the MMA issuer
__device__ long long* g_att_trace = nullptr;
constexpr int ATT_TRACE_BLOCKS = 48, ATT_TRACE_SLOTS = 8;
__device__ __forceinline__ void att_trace(int role, int blk, int slot) {
  if (g_att_trace && blk < ATT_TRACE_BLOCKS)
    g_att_trace[(((size_t)blockIdx.x * 5 + role) * ATT_TRACE_BLOCKS + blk) * ATT_TRACE_SLOTS + slot] = clock64();
}
#define ATT_T(role, blk, slot) do { if (lane == 0) att_trace(role, blk, slot); } while (0)
#else
#define ATT_T(role, blk, slot) do { } while (0)
#endif

struct AttItem {
  int sample, head, q0, klen, nb, row_base, q_limit;
};
// Item schedule of a ragged batch (built in shared memory by every CTA, see the kernel prologue): the samples sorted by
// descending key length and, over that order, the running count of their 128-query tiles.  Items are enumerated DENSELY
// (no slots for the all-padding query tiles of the shorter samples) and longest first, so the CTAs' static round-robin
// shares differ by at most one item of the cheapest kind; the plain (sample, head, tile) enumeration left the busiest CTA
// ~25 % above the mean at config 2's lengths (850 - 1030 frames), a tail during which most SMs idle.
constexpr int ATT_SCHED_MAX = 256;  // samples; larger batches (or klen == NULL: nothing is ragged) use the plain enumeration
struct AttSched {
  const int* tab;  // shared memory: [ATT_SCHED_MAX] clipped key lengths | [ATT_SCHED_MAX] sample of rank r | [n + 1] tiles before rank r
  int n;           // 0 = plain enumeration
};
// Work item w -> (sample, head, query tile); false when the tile holds padded queries only (skipped by every role).
__device__ __forceinline__ bool att_decode(int w, int qtiles, int heads, const int* __restrict__ klen_ptr,
                                           const int* __restrict__ row_start, int n_pad, const AttSched& sc, AttItem& it) {
  int qt;
  if (sc.n > 0) {
    const int th = w / heads;  // running tile index over the sorted samples; heads of one tile are neighbours
    int lo = 0, hi = sc.n;     // largest r with prefix[r] <= th (ranks with zero tiles are stepped over)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (sc.tab[2 * ATT_SCHED_MAX + mid] <= th) lo = mid; else hi = mid;
    }
    it.sample = sc.tab[ATT_SCHED_MAX + lo];
    it.head = w - th * heads;
    qt = th - sc.tab[2 * ATT_SCHED_MAX + lo];
  } else {
    qt = w % qtiles;
    const int sh = w / qtiles;
    it.head = sh % heads;
    it.sample = sh / heads;
  }
  it.klen = klen_ptr ? min(__ldg(klen_ptr + it.sample), n_pad) : n_pad;
  it.q0 = qt * ATT_BM;
  it.nb = (it.klen + ATT_BN - 1) / ATT_BN;
  // padded layout: sample s owns rows [s * n_pad, (s+1) * n_pad); packed layout: rows [row_start[s], row_start[s] + klen)
  it.row_base = row_start ? __ldg(row_start + it.sample) : it.sample * n_pad;
  it.q_limit = row_start ? it.klen : n_pad;  // query rows at or beyond this are not stored
  return it.q0 < it.klen;
}

// Computed key columns of a block that holds `kvalid` real keys: whole 32-key chunks, at most the block.
__device__ __forceinline__ int att_ncols(int kvalid) {
#ifdef ATT_NO_NARROW
  return ATT_BN;
#else
  return min(ATT_BN, (kvalid + 31) & ~31);
#endif
}

// KEY_BIAS: scores get key_bias[(sample * heads + head) * ld_bias + key] added BEFORE the scale (the position term of the
// Conformer's rel-pos attention, see conformer.cu); ld_bias covers whole 128-key blocks.
template <bool KEY_BIAS>
__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out, int ldo,
                 const int* __restrict__ klen_ptr, const int* __restrict__ row_start, int n_pad, int inner, int heads,
                 int qtiles, int total_items, float scale_log2, const float* __restrict__ key_bias, int ld_bias,
                 int samples) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ATT_SMEM_BAR);
  uint64_t* q_full = bars;           // 1
  uint64_t* q_empty = bars + 1;      // 1
  uint64_t* kv_full = bars + 2;                  // ATT_KV_SLOTS
  uint64_t* kv_empty = kv_full + ATT_KV_SLOTS;   // ATT_KV_SLOTS
  uint64_t* s_full = kv_empty + ATT_KV_SLOTS;
  uint64_t* s_free = s_full + 1;
  uint64_t* p_full = s_full + 2;
  uint64_t* o_full = s_full + 3;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(s_full + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == ATT_WARP_TMA && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int i = 0; i < ATT_KV_SLOTS; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == ATT_WARP_MMA) tmem_alloc(tmem_holder, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  pdl_wait();     // programmatic dependent launch: the prologue above may overlap the tail of the QKV GEMM
  pdl_trigger();

  // Item schedule (AttSched): rank every sample by (key length descending, index) and count the tiles before it.
  AttSched sched{reinterpret_cast<const int*>(smem + ATT_SMEM_SCHED), 0};
#ifndef ATT_PLAIN_ORDER
  if (klen_ptr != nullptr && samples > 1 && samples <= ATT_SCHED_MAX) {  // CTA-uniform
    int* sk = reinterpret_cast<int*>(smem + ATT_SMEM_SCHED);
    int* so = sk + ATT_SCHED_MAX;
    int* sp = so + ATT_SCHED_MAX;
    for (int t = threadIdx.x; t < samples; t += ATT_THREADS) sk[t] = min(__ldg(klen_ptr + t), n_pad);
    __syncthreads();
    for (int t = threadIdx.x; t < samples; t += ATT_THREADS) {
      const int kt = sk[t];
      int rank = 0, pre = 0;
      for (int u = 0; u < samples; ++u) {
        const int ku = sk[u];
        const bool before = ku > kt || (ku == kt && u < t);
        rank += before ? 1 : 0;
        pre += before ? (ku + ATT_BM - 1) / ATT_BM : 0;
      }
      so[rank] = t;
      sp[rank] = pre;
      if (rank == samples - 1) sp[samples] = pre + (kt + ATT_BM - 1) / ATT_BM;
    }
    __syncthreads();
    sched.n = samples;
    total_items = sp[samples] * heads;
  }
#endif

  if (warp == ATT_WARP_TMA) {
    if (elect_one()) {
      int slot = 0, item = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
        AttItem it;
        if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, sched, it)) continue;
        const int row_base = it.row_base;
        if (item > 0) mbar_wait(q_empty, (item - 1) & 1);  // every QK of the previous item has read Q
        mbar_expect_tx(q_full, ATT_TILE_BYTES);
        tma_load_2d(smem + ATT_SMEM_Q, &tmQKV, q_full, it.head * ATT_DH, row_base + it.q0);
        for (int j = 0; j < it.nb; ++j) {
#pragma unroll
          for (int kv = 0; kv < 2; ++kv) {  // K_j then V_j
            mbar_wait(&kv_empty[slot], phase ^ 1);
#ifdef ATT_KO_TMA
            mbar_arrive(&kv_full[slot]);
#else
            mbar_expect_tx(&kv_full[slot], ATT_TILE_BYTES);
            tma_load_2d(smem + ATT_SMEM_KV + slot * ATT_TILE_BYTES, &tmQKV, &kv_full[slot],
                        (kv + 1) * inner + it.head * ATT_DH, row_base + j * ATT_BN);
#endif
            if (++slot == ATT_KV_SLOTS) { slot = 0; phase ^= 1; }
          }
        }
        ++item;
      }
    }
  } else if (warp == ATT_WARP_MMA) {
    if (elect_one()) {
      constexpr uint32_t idesc_pv = umma_idesc(kFmtBF16, ATT_BM, ATT_DH, 0, 1);
      const uint64_t qdesc = umma_desc_kmajor_sw128(smem_u32(smem + ATT_SMEM_Q));
      // K/V tiles are numbered over the CTA's whole life: block b -> K tile 2b, V tile 2b+1; tile t -> slot t % SLOTS,
      // phase (t / SLOTS) & 1.  S / P / O barriers complete one phase per block: parity = b & 1.
      // `nc` = key columns of the block that are computed at all (a multiple of 32): the last block of a sample is
      // narrowed to the 32-key chunks that hold real keys (QK with N = nc, PV with K = nc).
      auto tile_slot = [](int t) { return t % ATT_KV_SLOTS; };
      auto tile_phase = [](int t) { return (uint32_t)((t / ATT_KV_SLOTS) & 1); };
      auto do_qk = [&](int b, int nc) {
        const int t = 2 * b, sl = tile_slot(t);
        mbar_wait(&kv_full[sl], tile_phase(t));
        tc_fence_after();
        const uint64_t kdesc = umma_desc_kmajor_sw128(smem_u32(smem + ATT_SMEM_KV + sl * ATT_TILE_BYTES));
        const uint32_t idesc_qk = umma_idesc(kFmtBF16, ATT_BM, (uint32_t)nc, 0, 0);
#ifndef ATT_KO_MMA
#pragma unroll
        for (int k = 0; k < ATT_DH / 16; ++k) umma_f16_ss(tmem_base, qdesc + 2 * k, kdesc + 2 * k, idesc_qk, k != 0);
#endif
        umma_commit(&kv_empty[sl]);
        umma_commit(s_full);
        ATT_T(4, b, 0);
      };
      auto do_pv = [&](int b, bool first, int nc) {
        const int t = 2 * b + 1, sl = tile_slot(t);
        mbar_wait(&kv_full[sl], tile_phase(t));
        tc_fence_after();
        // 16 keys = 2048 bytes of the MN-major V tile per K step: the descriptor's address field advances by 2048 >> 4
        const uint64_t vdesc = umma_desc_mnmajor_sw128(smem_u32(smem + ATT_SMEM_KV + sl * ATT_TILE_BYTES), 1024, 1024);
#ifndef ATT_KO_MMA
        for (int k = 0; k < nc / 16; ++k)  // A = P straight from TMEM: 16 bf16 keys = 8 columns per K step
          umma_f16_ts(tmem_base + ATT_TMEM_O, tmem_base + ATT_TMEM_P + k * 8, vdesc + k * 128, idesc_pv, !(first && k == 0));
#endif
        umma_commit(&kv_empty[sl]);
        umma_commit(o_full);
        ATT_T(4, b, 2);
      };
      int item = 0, blk = 0;
      for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
        AttItem it;
        if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, sched, it)) continue;
        auto ncols = [&](int j) { return att_ncols(it.klen - j * ATT_BN); };
        mbar_wait(q_full, item & 1);
        if (blk > 0) {  // the softmax warps hold the previous item's last S block in registers
          mbar_wait(s_free, (blk - 1) & 1);
          tc_fence_after();
        }
        do_qk(blk, ncols(0));
        for (int j = 0; j < it.nb; ++j) {
          const int b = blk + j;
          if (j + 1 < it.nb) {
            mbar_wait(s_free, b & 1);  // softmax finished reading S_b
            tc_fence_after();
            do_qk(b + 1, ncols(j + 1));
            if (j + 2 == it.nb) umma_commit(q_empty);  // last QK of the item issued: Q may be replaced once it retires
          } else if (it.nb == 1) {
            umma_commit(q_empty);
          }
          mbar_wait(p_full, b & 1);  // P_b in TMEM, O rescaled if needed
          tc_fence_after();
          ATT_T(4, b, 1);
          do_pv(b, j == 0, ncols(j));
        }
        blk += it.nb;
        ++item;
      }
    }
  } else {
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;  // query row inside the tile == TMEM lane
    const uint32_t t_s = tmem_base + (uint32_t(quarter * 32) << 16);
    const uint32_t t_o = t_s + ATT_TMEM_O;
    const uint32_t t_p = t_s + ATT_TMEM_P;
    int blk = 0;
    for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
      AttItem it;
      if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, sched, it)) continue;
      float m_ref = -INFINITY;  // exponent reference (scaled log2 units); lags the true row maximum by at most 8
      float l = 0.f;
      // One key block of NC computed columns (NC = 128 except in the last block of a sample, see `ncols` above).
      auto softmax_block = [&](auto nc_tag, const int j) {
        constexpr int NC = decltype(nc_tag)::value;
        const int b = blk + j;
        const int kvalid = it.klen - j * ATT_BN;  // keys [0, kvalid) of this block are real
        ATT_T(quarter, b, 0);
        mbar_wait(s_full, b & 1);
        tc_fence_after();
        float s[NC];
        {
          uint32_t* raw = reinterpret_cast<uint32_t*>(s);
#pragma unroll
          for (int c = 0; c < NC / 32; ++c) tmem_ld_32x32(t_s + 32 * c, *reinterpret_cast<uint32_t(*)[32]>(raw + 32 * c));
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);  // S_b is in registers: the next QK may overwrite the TMEM columns now
        ATT_T(quarter, b, 2);
        if constexpr (KEY_BIAS) {
          const float4* kb = reinterpret_cast<const float4*>(key_bias + ((size_t)it.sample * heads + it.head) * ld_bias +
                                                             (size_t)j * ATT_BN);
#pragma unroll
          for (int i = 0; i < NC / 4; ++i) {
            const float4 v4 = __ldg(kb + i);
            s[4 * i] += v4.x;
            s[4 * i + 1] += v4.y;
            s[4 * i + 2] += v4.z;
            s[4 * i + 3] += v4.w;
          }
        }
        if (kvalid < NC) {  // last key block of the sample only; kept a real (warp-uniform) branch
#pragma unroll
          for (int i = 0; i < NC; ++i)
            if (i >= kvalid) s[i] = -INFINITY;
        }
        float mx4[4];  // four independent chains instead of one 64-deep dependency chain
#pragma unroll
        for (int c = 0; c < 4; ++c) mx4[c] = fmaxf(s[2 * c], s[2 * c + 1]);
#pragma unroll
        for (int i = 8; i < NC; i += 8)
#pragma unroll
          for (int c = 0; c < 4; ++c) mx4[c] = fmaxf(mx4[c], fmaxf(s[i + 2 * c], s[i + 2 * c + 1]));
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_new = fmaxf(m_ref, mx * scale_log2);
        const bool grow = m_new - m_ref > 8.0f;  // always true in the first block (m_ref = -inf)
        if (j > 0) {  // PV_{b-1} finished: O is complete up to block b-1 and P (single-buffered in TMEM) may be rewritten.
          // (Deferring this wait to just before the P store was measured: slower, the exponentials of the two
          // co-resident CTAs then collide more often.)
          mbar_wait(o_full, (b - 1) & 1);
          tc_fence_after();
        }
        if (__any_sync(0xffffffffu, grow)) {
          const float alpha = grow ? ex2_approx(m_ref - m_new) : 1.0f;  // first block: ex2(-inf) = 0
          if (grow) m_ref = m_new;
          l *= alpha;
          if (j > 0) {
#pragma unroll 1
            for (int c = 0; c < ATT_DH / 16; ++c) {
              uint32_t o16[16];
              tmem_ld_32x16(t_o + c * 16, o16);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) o16[i] = __float_as_uint(__uint_as_float(o16[i]) * alpha);
              tmem_st_32x16(t_o + c * 16, o16);
            }
            tmem_st_wait();
          }
        }
        ATT_T(quarter, b, 5);
        // p = 2^(s * scale - m_ref): packed fp32x2 FMAs / adds (2 elements per issue slot), masked keys give ex2(-inf) = 0.
        // The bf16 pairs are packed as they are produced so the fp32 values die early (register pressure).
        const float2 sc2 = make_float2(scale_log2, scale_log2), nm2 = make_float2(-m_ref, -m_ref);
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        uint32_t pk[NC / 2];
#pragma unroll
        for (int i = 0; i < NC; i += 2) {
          const float2 a = __ffma2_rn(make_float2(s[i], s[i + 1]), sc2, nm2);
          float2 e;
          e.x = ex2_approx(a.x);
          e.y = ex2_approx(a.y);
          sum2[(i >> 1) & 1] = __fadd2_rn(sum2[(i >> 1) & 1], e);
          pk[i >> 1] = pack_bf16x2(e.x, e.y);
        }
        l += (sum2[0].x + sum2[1].x) + (sum2[0].y + sum2[1].y);
        ATT_T(quarter, b, 6);
        // P (bf16, two keys per 32-bit column) goes back to TMEM: the PV MMA reads its A operand from there, so P never
        // touches shared memory (the SS form was bound by shared-memory bandwidth: P write + P read + V read per block)
#pragma unroll
        for (int h = 0; h < NC / 32; ++h) tmem_st_32x16(t_p + h * 16, *reinterpret_cast<uint32_t(*)[16]>(&pk[h * 16]));
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
        ATT_T(quarter, b, 7);
      };
      for (int j = 0; j + 1 < it.nb; ++j) softmax_block(std::integral_constant<int, 128>{}, j);
      {
        const int j = it.nb - 1;
        const int nc = att_ncols(it.klen - j * ATT_BN);
        if (nc == 128) softmax_block(std::integral_constant<int, 128>{}, j);
        else if (nc == 96) softmax_block(std::integral_constant<int, 96>{}, j);
        else if (nc == 64) softmax_block(std::integral_constant<int, 64>{}, j);
        else softmax_block(std::integral_constant<int, 32>{}, j);
      }
      blk += it.nb;
      mbar_wait(o_full, (blk - 1) & 1);
      tc_fence_after();
      const int q = it.q0 + r;
      const float inv = 1.0f / l;
      __nv_bfloat16* orow = out + (size_t)(it.row_base + q) * ldo + it.head * ATT_DH;
#pragma unroll 1
      for (int c = 0; c < ATT_DH / 32; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(t_o + c * 32, raw);
        tmem_ld_wait();
        if (q < it.q_limit) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 u;
            u.x = pack_bf16x2(__uint_as_float(raw[i]) * inv, __uint_as_float(raw[i + 1]) * inv);
            u.y = pack_bf16x2(__uint_as_float(raw[i + 2]) * inv, __uint_as_float(raw[i + 3]) * inv);
            u.z = pack_bf16x2(__uint_as_float(raw[i + 4]) * inv, __uint_as_float(raw[i + 5]) * inv);
            u.w = pack_bf16x2(__uint_as_float(raw[i + 6]) * inv, __uint_as_float(raw[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c * 32 + i) = u;
          }
        }
      }
      tc_fence_before();  // O has been read: the next item's first PV (issued after our next p_full arrive) may overwrite it
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == ATT_WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ATT_TMEM_COLS);
  }
}

#ifdef ATT_TRACE
extern "C" int f5e_attention_trace(long long* buf) {
  F5E_CUDA_CHECK(cudaMemcpyToSymbol(g_att_trace, &buf, sizeof(buf)));
  return 0;
}
#endif

int attention_configure() {
  F5E_CUDA_CHECK(cudaFuncSetAttribute(attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM_TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM_TOTAL));
  return 0;
}

int attention(cudaStream_t st, const __nv_bfloat16* qkv, int ld_qkv, __nv_bfloat16* out, int ldo, const int* klen,
              int samples, int n_pad, int heads, float scale, const int* row_start, int total_rows, const float* key_bias,
              int ld_bias) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(samples > 0 && n_pad > 0 && heads > 0, "attention: empty problem");
  F5E_REQUIRE(ldo % 8 == 0, "attention: ldo must be a multiple of 8");
  F5E_REQUIRE(row_start == nullptr || (klen != nullptr && total_rows > 0), "attention: packed rows need klen and total_rows");
  F5E_REQUIRE(key_bias == nullptr || (row_start == nullptr && ld_bias % 4 == 0 && ld_bias >= (n_pad + ATT_BN - 1) / ATT_BN * ATT_BN &&
                                      (reinterpret_cast<uintptr_t>(key_bias) & 15) == 0),
              "attention: key_bias needs padded rows and a 16-byte aligned stride covering whole 128-key blocks");
  const int inner = heads * ATT_DH;
  CUtensorMap tm;
  int rc = make_tmap_2d(&tm, qkv, 2, (uint64_t)3 * inner, (uint64_t)(row_start ? total_rows : samples * n_pad), (uint64_t)ld_qkv * 2,
                        ATT_DH, 128);
  if (rc) return rc;
  const int qtiles = (n_pad + ATT_BM - 1) / ATT_BM;
  const long long total = (long long)samples * heads * qtiles;
  F5E_REQUIRE(total < (1ll << 30), "attention: too many work items");
  const int grid = (int)(total < 2 * num_sms() ? total : 2 * num_sms());
  LaunchScope ls(st, PROF_ATTN, samples, n_pad, heads);
  const float sl2 = scale * 1.4426950408889634f;
  if (key_bias)
    F5E_CUDA_CHECK(launch_pdl(attention_kernel<true>, dim3(grid), dim3(ATT_THREADS), ATT_SMEM_TOTAL, st, tm, out, ldo, klen,
                              row_start, n_pad, inner, heads, qtiles, (int)total, sl2, key_bias, ld_bias, samples));
  else
    F5E_CUDA_CHECK(launch_pdl(attention_kernel<false>, dim3(grid), dim3(ATT_THREADS), ATT_SMEM_TOTAL, st, tm, out, ldo, klen,
                              row_start, n_pad, inner, heads, qtiles, (int)total, sl2, (const float*)nullptr, 0, samples));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
