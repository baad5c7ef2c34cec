// Non-causal flash attention for sm_100a, dim_head 64, per-sample key length (padding mask), bf16 I/O.
//
// Replaces F.scaled_dot_product_attention with the expanded [B,H,N,N] bool mask of reference
// model/modules.py:483-492: keys n >= klen[sample] are excluded, nothing of size N x N is ever stored.
//
// At dim_head 64 a 128 x 128 score block costs 512 tensor cycles but 1024 cycles of MUFU.EX2 (16 lanes / clk / SM,
// tools/ubench/tmem_bw.cu), so the design goal is to keep the MUFU pipe of every SM sub-partition busy.
// ONE persistent CTA per SM (512 threads, all 512 TMEM columns) runs TWO independent streams of work items
// (sample, head, 128-query tile), FlashAttention-4 style:
//   warps 0-3   softmax of stream A     thread == query row: S from TMEM (once), row max, ex2, row sum, bf16 P back to TMEM
//   warps 4-7   softmax of stream B     (warp i of A and warp i of B share SM sub-partition i)
//   warps 8-11  epilogue warpgroup      O / l of a finished item -> global memory, for both streams; the softmax warps
//                                       never wait for the last PV of an item and never touch global memory
//   warp 12/13  TMA producers (A / B)   Q tile per item, K_j / V_j tiles through a 4-slot ring per stream
//   warp 14/15  MMA issuers (A / B)     S = Q K_j^T (SS) into the stream's S columns; O += P_j V_j (TS: P read from TMEM)
// Registers are re-divided with setmaxnreg (softmax 184, epilogue 96, issuers 48 = 64 K).
// PING-PONG: the exponential phases of the two softmax warps of a sub-partition are made mutually exclusive with a
// pair of named barriers (A's turn / B's turn): while one warp occupies the MUFU pipe the other runs its MUFU-free
// phase (S load, row max, P store, barrier hops).  Unsynchronised (two co-resident CTAs, round 1) the two warps drift
// into overlapping exponential phases and the pipe idles for the rest of the period (profiles/r02_attention.md).
//   * S_j is released right after it is in registers, so QK_{j+1} overlaps the softmax math; O accumulates ACROSS key
//     blocks in TMEM and is rescaled lazily (only when a row maximum grew by more than 2^8; P and the row sum use the
//     same stale reference, so O / l is exact).
//   * The last key block of a sample is narrowed to the 32-key chunks that hold real keys (QK with N = nc, PV with
//     K = nc, nc exponentials per row instead of 128).
//   * tcgen05 / TMA instructions are issued by a thread chosen with elect.sync (the compiler then emits them without a
//     per-instruction serialisation loop, see gemm.cu).
// Input layout: qkv [rows, 3*inner] with Q | K | V column blocks (what the fused QKV GEMM writes, RoPE already
// applied); output [rows, inner].
#include "host.h"
#include <type_traits>

namespace f5e {

constexpr int ATT_BM = 128;   // queries per work item
constexpr int ATT_BN = 128;   // keys per block
constexpr int ATT_DH = 64;
constexpr int ATT_THREADS = 512;
constexpr int ATT_KV_SLOTS = 4;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB
constexpr int ATT_SMEM_STREAM = ATT_TILE_BYTES * (1 + ATT_KV_SLOTS);  // Q + K/V ring of one stream
constexpr int ATT_SMEM_L = 2 * ATT_SMEM_STREAM;                       // row sums handed to the epilogue: float [2][2][128]
constexpr int ATT_SMEM_BAR = ATT_SMEM_L + 2 * 2 * 128 * 4;
constexpr int ATT_SMEM_TOTAL = ATT_SMEM_BAR + 512 + 1024;
constexpr uint32_t ATT_TMEM_COLS = 512;
constexpr uint32_t ATT_TMEM_STREAM = 256;  // per stream: S [0,128) fp32 | O [128,192) fp32 | P [192,256) bf16 pairs (128 keys)
constexpr uint32_t ATT_TMEM_O = 128;
constexpr uint32_t ATT_TMEM_P = 192;
constexpr int ATT_WARP_EPI = 8, ATT_WARP_TMA = 12, ATT_WARP_MMA = 14;
constexpr int ATT_BARS_PER_STREAM = 16;

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

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {  // non-blocking
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

#ifdef ATT_TRACE
// developer tool (tools/attn_lab.py --trace): per-block phase timestamps of the softmax warps and the MMA issuers
__device__ long long* g_att_trace = nullptr;
constexpr int ATT_TRACE_BLOCKS = 48, ATT_TRACE_SLOTS = 8, ATT_TRACE_ROLES = 10;
__device__ __forceinline__ void att_trace(int role, int blk, int slot) {
  if (g_att_trace && blk < ATT_TRACE_BLOCKS)
    g_att_trace[(((size_t)blockIdx.x * ATT_TRACE_ROLES + role) * ATT_TRACE_BLOCKS + blk) * ATT_TRACE_SLOTS + slot] = clock64();
}
#define ATT_T(role, blk, slot) do { if (lane == 0) att_trace(role, blk, slot); } while (0)
#else
#define ATT_T(role, blk, slot) do { } while (0)
#endif

struct AttItem {
  int sample, head, q0, klen, nb, row_base, q_limit;
};
// Work item w -> (sample, head, query tile); false when the tile holds padded queries only (skipped by every role).
__device__ __forceinline__ bool att_decode(int w, int qtiles, int heads, const int* __restrict__ klen_ptr,
                                           const int* __restrict__ row_start, int n_pad, AttItem& it) {
  const int qt = w % qtiles, sh = w / qtiles;
  it.head = sh % heads;
  it.sample = sh / heads;
  it.klen = klen_ptr ? min(__ldg(klen_ptr + it.sample), n_pad) : n_pad;
  it.q0 = qt * ATT_BM;
  it.nb = (it.klen + ATT_BN - 1) / ATT_BN;
  // padded layout: sample s owns rows [s * n_pad, (s+1) * n_pad); packed layout: rows [row_start[s], row_start[s] + klen)
  it.row_base = row_start ? __ldg(row_start + it.sample) : it.sample * n_pad;
  it.q_limit = row_start ? it.klen : n_pad;  // query rows at or beyond this are not stored
  return it.q0 < it.klen;
}
// Key blocks and valid items of the whole item list of one stream (items first, first + stride, ...), lane-parallel.
__device__ __forceinline__ void att_stream_totals(int first, int stride, int total_items, int qtiles, int heads,
                                                  const int* __restrict__ klen_ptr, int n_pad, int lane, int& blocks, int& items) {
  int nb = 0, ni = 0;
  for (long long w = first + (long long)lane * stride; w < total_items; w += 32ll * stride) {
    const int qt = (int)(w % qtiles), sample = (int)(w / qtiles) / heads;
    const int klen = klen_ptr ? min(__ldg(klen_ptr + sample), n_pad) : n_pad;
    if (qt * ATT_BM < klen) {
      nb += (klen + ATT_BN - 1) / ATT_BN;
      ++ni;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nb += __shfl_xor_sync(0xffffffffu, nb, o);
    ni += __shfl_xor_sync(0xffffffffu, ni, o);
  }
  blocks = nb;
  items = ni;
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
__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out, int ldo,
                 const int* __restrict__ klen_ptr, const int* __restrict__ row_start, int n_pad, int inner, int heads,
                 int qtiles, int total_items, float scale_log2, const float* __restrict__ key_bias, int ld_bias) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* all_bars = reinterpret_cast<uint64_t*>(smem + ATT_SMEM_BAR);
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(all_bars + 2 * ATT_BARS_PER_STREAM);
  float* l_buf = reinterpret_cast<float*>(smem + ATT_SMEM_L);  // [stream][item parity][row]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == ATT_WARP_TMA && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int s = 0; s < 2; ++s) {
      uint64_t* b = all_bars + s * ATT_BARS_PER_STREAM;
      mbar_init(b + 0, 1);  // q_full
      mbar_init(b + 1, 1);  // q_empty
      for (int i = 0; i < ATT_KV_SLOTS; ++i) {
        mbar_init(b + 2 + i, 1);                 // kv_full
        mbar_init(b + 2 + ATT_KV_SLOTS + i, 1);  // kv_empty
      }
      mbar_init(b + 10, 1);  // s_full
      mbar_init(b + 11, 4);  // s_free
      mbar_init(b + 12, 4);  // p_full
      mbar_init(b + 13, 1);  // o_full  (every PV)
      mbar_init(b + 14, 1);  // o_final (last PV of an item)
      mbar_init(b + 15, 4);  // o_free  (the epilogue has read O)
    }
    mbar_fence_init();
  }
  if (warp == ATT_WARP_MMA) tmem_alloc(tmem_holder, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_cta = *tmem_holder;
  const int stride = 2 * gridDim.x;  // item stride of a stream; stream g = 2 * blockIdx.x + s starts at item g

  if (warp >= ATT_WARP_TMA) {
    // ============================== issuer warpgroup: TMA producers and MMA issuers ==============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 48;" ::: "memory");
    const int s = warp & 1;  // warps 12 / 14 -> stream A, 13 / 15 -> stream B
    uint8_t* sm = smem + s * ATT_SMEM_STREAM;
    uint64_t* bars = all_bars + s * ATT_BARS_PER_STREAM;
    uint64_t *q_full = bars, *q_empty = bars + 1, *kv_full = bars + 2, *kv_empty = bars + 2 + ATT_KV_SLOTS;
    uint64_t *s_full = bars + 10, *s_free = bars + 11, *p_full = bars + 12, *o_full = bars + 13, *o_final = bars + 14,
             *o_free = bars + 15;
    const int first = 2 * blockIdx.x + s;
    if (warp < ATT_WARP_MMA) {
      if (elect_one()) {
        int slot = 0, item = 0;
        uint32_t phase = 0;
        for (int w = first; w < total_items; w += stride) {
          AttItem it;
          if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, it)) continue;
          const int row_base = it.row_base;
          if (item > 0) mbar_wait(q_empty, (item - 1) & 1);  // every QK of the previous item has read Q
          mbar_expect_tx(q_full, ATT_TILE_BYTES);
          tma_load_2d(sm, &tmQKV, q_full, it.head * ATT_DH, row_base + it.q0);
          for (int j = 0; j < it.nb; ++j) {
#pragma unroll
            for (int kv = 0; kv < 2; ++kv) {  // K_j then V_j
              mbar_wait(&kv_empty[slot], phase ^ 1);
#ifdef ATT_KO_TMA
              mbar_arrive(&kv_full[slot]);
#else
              mbar_expect_tx(&kv_full[slot], ATT_TILE_BYTES);
              tma_load_2d(sm + ATT_TILE_BYTES * (1 + slot), &tmQKV, &kv_full[slot], (kv + 1) * inner + it.head * ATT_DH,
                          row_base + j * ATT_BN);
#endif
              if (++slot == ATT_KV_SLOTS) { slot = 0; phase ^= 1; }
            }
          }
          ++item;
        }
      }
    } else {
      if (elect_one()) {
        const uint32_t tmem_base = tmem_cta + s * ATT_TMEM_STREAM;
        constexpr uint32_t idesc_pv = umma_idesc(kFmtBF16, ATT_BM, ATT_DH, 0, 1);
        const uint64_t qdesc = umma_desc_kmajor_sw128(smem_u32(sm));
        // K/V tiles are numbered over the stream's whole life: block b -> K tile 2b, V tile 2b+1; tile t -> slot t % SLOTS,
        // phase (t / SLOTS) & 1.  S / P / O barriers complete one phase per block: parity = b & 1; o_final / o_free one per item.
        // `nc` = key columns of the block that are computed at all (a multiple of 32, see att_ncols).
        auto tile_slot = [](int t) { return t % ATT_KV_SLOTS; };
        auto tile_phase = [](int t) { return (uint32_t)((t / ATT_KV_SLOTS) & 1); };
        auto do_qk = [&](int b, int nc) {
          const int t = 2 * b, sl = tile_slot(t);
          mbar_wait(&kv_full[sl], tile_phase(t));
          tc_fence_after();
          const uint64_t kdesc = umma_desc_kmajor_sw128(smem_u32(sm + ATT_TILE_BYTES * (1 + sl)));
          const uint32_t idesc_qk = umma_idesc(kFmtBF16, ATT_BM, (uint32_t)nc, 0, 0);
#ifndef ATT_KO_MMA
#pragma unroll
          for (int k = 0; k < ATT_DH / 16; ++k) umma_f16_ss(tmem_base, qdesc + 2 * k, kdesc + 2 * k, idesc_qk, k != 0);
#endif
          umma_commit(&kv_empty[sl]);
          umma_commit(s_full);
          ATT_T(8 + s, b, 0);
        };
        auto do_pv = [&](int b, bool first_blk, bool last_blk, int nc) {
          const int t = 2 * b + 1, sl = tile_slot(t);
          mbar_wait(&kv_full[sl], tile_phase(t));
          tc_fence_after();
          // 16 keys = 2048 bytes of the MN-major V tile per K step: the descriptor's address field advances by 2048 >> 4
          const uint64_t vdesc = umma_desc_mnmajor_sw128(smem_u32(sm + ATT_TILE_BYTES * (1 + sl)), 1024, 1024);
#ifndef ATT_KO_MMA
          for (int k = 0; k < nc / 16; ++k)  // A = P straight from TMEM: 16 bf16 keys = 8 columns per K step
            umma_f16_ts(tmem_base + ATT_TMEM_O, tmem_base + ATT_TMEM_P + k * 8, vdesc + k * 128, idesc_pv, !(first_blk && k == 0));
#endif
          umma_commit(&kv_empty[sl]);
          umma_commit(o_full);
          if (last_blk) umma_commit(o_final);
          ATT_T(8 + s, b, 2);
        };
        int item = 0, blk = 0;
        for (int w = first; w < total_items; w += stride) {
          AttItem it;
          if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, it)) continue;
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
            if (j == 0 && item > 0) mbar_wait(o_free, (item - 1) & 1);  // the epilogue has taken the previous item's O out
            tc_fence_after();
            ATT_T(8 + s, b, 1);
            do_pv(b, j == 0, j + 1 == it.nb, ncols(j));
          }
          blk += it.nb;
          ++item;
        }
      }
    }
  } else if (warp >= ATT_WARP_EPI) {
    // ============================== epilogue warpgroup: O / l -> global, both streams ==============================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 96;" ::: "memory");
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;
    int n_items[2], done[2] = {0, 0}, next_w[2];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int nb_unused;
      next_w[s] = 2 * blockIdx.x + s;
      att_stream_totals(next_w[s], stride, total_items, qtiles, heads, klen_ptr, n_pad, lane, nb_unused, n_items[s]);
    }
    while (done[0] < n_items[0] || done[1] < n_items[1]) {
      bool progressed = false;
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (done[s] >= n_items[s]) continue;
        uint64_t* bars = all_bars + s * ATT_BARS_PER_STREAM;
        if (!mbar_test(bars + 14, done[s] & 1)) continue;  // o_final: the item's last PV has retired
        tc_fence_after();
        AttItem it;
        while (!att_decode(next_w[s], qtiles, heads, klen_ptr, row_start, n_pad, it)) next_w[s] += stride;
        next_w[s] += stride;
        const uint32_t t_o = tmem_cta + s * ATT_TMEM_STREAM + ATT_TMEM_O + (uint32_t(quarter * 32) << 16);
        uint32_t o0[32], o1[32];
        tmem_ld_32x32(t_o, o0);
        tmem_ld_32x32(t_o + 32, o1);
        tmem_ld_wait();
        const float inv = 1.0f / l_buf[(s * 2 + (done[s] & 1)) * 128 + r];
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bars + 15);  // o_free: the stream's next item may overwrite O (and, two items on, l)
        const int q = it.q0 + r;
        if (q < it.q_limit) {
          __nv_bfloat16* orow = out + (size_t)(it.row_base + q) * ldo + it.head * ATT_DH;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 u;
            u.x = pack_bf16x2(__uint_as_float(o0[i]) * inv, __uint_as_float(o0[i + 1]) * inv);
            u.y = pack_bf16x2(__uint_as_float(o0[i + 2]) * inv, __uint_as_float(o0[i + 3]) * inv);
            u.z = pack_bf16x2(__uint_as_float(o0[i + 4]) * inv, __uint_as_float(o0[i + 5]) * inv);
            u.w = pack_bf16x2(__uint_as_float(o0[i + 6]) * inv, __uint_as_float(o0[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + i) = u;
          }
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 u;
            u.x = pack_bf16x2(__uint_as_float(o1[i]) * inv, __uint_as_float(o1[i + 1]) * inv);
            u.y = pack_bf16x2(__uint_as_float(o1[i + 2]) * inv, __uint_as_float(o1[i + 3]) * inv);
            u.z = pack_bf16x2(__uint_as_float(o1[i + 4]) * inv, __uint_as_float(o1[i + 5]) * inv);
            u.w = pack_bf16x2(__uint_as_float(o1[i + 6]) * inv, __uint_as_float(o1[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + 32 + i) = u;
          }
        }
        ++done[s];
        progressed = true;
      }
      if (!progressed) __nanosleep(200);
    }
  } else {
    // ============================== softmax warpgroups (stream = warpgroup index) ==============================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 184;" ::: "memory");
    const int s = warp >> 2;
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;  // query row inside the tile == TMEM lane
    uint64_t* bars = all_bars + s * ATT_BARS_PER_STREAM;
    uint64_t *s_full = bars + 10, *s_free = bars + 11, *p_full = bars + 12, *o_full = bars + 13;
    const uint32_t t_s = tmem_cta + s * ATT_TMEM_STREAM + (uint32_t(quarter * 32) << 16);
    const uint32_t t_o = t_s + ATT_TMEM_O;
    const uint32_t t_p = t_s + ATT_TMEM_P;
    // ping-pong with the partner warp (same quarter, other stream): named barrier 1 + stream * 4 + quarter is "stream's
    // turn"; 64 threads = this warp's bar.sync + the partner's bar.arrive.  Turns alternate A, B, A, B, ... for
    // T = max(blocks of A, blocks of B) rounds; the stream with fewer blocks passes its remaining turns on.
    const int my_turn = 1 + s * 4 + quarter, partner_turn = 1 + (1 - s) * 4 + quarter;
    int n_blk[2], ni_unused;
    att_stream_totals(2 * blockIdx.x, stride, total_items, qtiles, heads, klen_ptr, n_pad, lane, n_blk[0], ni_unused);
    att_stream_totals(2 * blockIdx.x + 1, stride, total_items, qtiles, heads, klen_ptr, n_pad, lane, n_blk[1], ni_unused);
    const int turns = max(n_blk[0], n_blk[1]);
    int turn = 0;
#ifndef ATT_NO_PINGPONG
    if (s == 1) named_bar_arrive(partner_turn, 64);  // stream A moves first
#endif
    auto take_turn = [&]() {
#ifndef ATT_NO_PINGPONG
      named_bar_sync(my_turn, 64);
#endif
    };
    auto pass_turn = [&]() {
      ++turn;
#ifndef ATT_NO_PINGPONG
      if (!(s == 1 && turn == turns)) named_bar_arrive(partner_turn, 64);  // nobody waits for A's turn after the last round
#endif
    };
    int blk = 0, item = 0;
    for (int w = 2 * blockIdx.x + s; w < total_items; w += stride) {
      AttItem it;
      if (!att_decode(w, qtiles, heads, klen_ptr, row_start, n_pad, it)) continue;
      float m_ref = -INFINITY;  // exponent reference (scaled log2 units); lags the true row maximum by at most 8
      float l = 0.f;
      // One key block of NC computed columns (NC = 128 except in the last block of a sample, see att_ncols).
      auto softmax_block = [&](auto nc_tag, const int j) {
        constexpr int NC = decltype(nc_tag)::value;
        const int b = blk + j;
        const int kvalid = it.klen - j * ATT_BN;  // keys [0, kvalid) of this block are real
        ATT_T(warp, b, 0);
        mbar_wait(s_full, b & 1);
        tc_fence_after();
        float sc[NC];
        {
          uint32_t* raw = reinterpret_cast<uint32_t*>(sc);
#pragma unroll
          for (int c = 0; c < NC / 32; ++c) tmem_ld_32x32(t_s + 32 * c, *reinterpret_cast<uint32_t(*)[32]>(raw + 32 * c));
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);  // S_b is in registers: the next QK may overwrite the TMEM columns now
        ATT_T(warp, b, 2);
        if constexpr (KEY_BIAS) {
          const float4* kb = reinterpret_cast<const float4*>(key_bias + ((size_t)it.sample * heads + it.head) * ld_bias +
                                                             (size_t)j * ATT_BN);
#pragma unroll
          for (int i = 0; i < NC / 4; ++i) {
            const float4 v4 = __ldg(kb + i);
            sc[4 * i] += v4.x;
            sc[4 * i + 1] += v4.y;
            sc[4 * i + 2] += v4.z;
            sc[4 * i + 3] += v4.w;
          }
        }
        if (kvalid < NC) {  // last key block of the sample only; kept a real (warp-uniform) branch
#pragma unroll
          for (int i = 0; i < NC; ++i)
            if (i >= kvalid) sc[i] = -INFINITY;
        }
        float mx4[4];  // four independent chains instead of one 64-deep dependency chain
#pragma unroll
        for (int c = 0; c < 4; ++c) mx4[c] = fmaxf(sc[2 * c], sc[2 * c + 1]);
#pragma unroll
        for (int i = 8; i < NC; i += 8)
#pragma unroll
          for (int c = 0; c < 4; ++c) mx4[c] = fmaxf(mx4[c], fmaxf(sc[i + 2 * c], sc[i + 2 * c + 1]));
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_new = fmaxf(m_ref, mx * scale_log2);
        const bool grow = m_new - m_ref > 8.0f;  // always true in the first block (m_ref = -inf)
        if (b > 0) {  // PV_{b-1} (of this item or, for j == 0, the last one of the previous item) has retired:
                      // P (single-buffered in TMEM) may be rewritten and, for j > 0, O is complete up to block b-1
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
        take_turn();  // ---- the MUFU pipe of this sub-partition is ours until pass_turn() ----
        ATT_T(warp, b, 5);
        // p = 2^(s * scale - m_ref): packed fp32x2 FMAs / adds (2 elements per issue slot), masked keys give ex2(-inf) = 0.
        // Every 32 columns the 16 packed bf16 pairs go back to TMEM (tcgen05.st is asynchronous): the PV MMA reads its A
        // operand from there, so P never touches shared memory.
        const float2 sc2 = make_float2(scale_log2, scale_log2), nm2 = make_float2(-m_ref, -m_ref);
        float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
        for (int h = 0; h < NC / 32; ++h) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float2 a = __ffma2_rn(make_float2(sc[32 * h + i], sc[32 * h + i + 1]), sc2, nm2);
            float2 e;
            e.x = ex2_approx(a.x);
            e.y = ex2_approx(a.y);
            sum2[(i >> 1) & 1] = __fadd2_rn(sum2[(i >> 1) & 1], e);
            pk[i >> 1] = pack_bf16x2(e.x, e.y);
          }
          tmem_st_32x16(t_p + h * 16, pk);
        }
        pass_turn();
        ATT_T(warp, b, 6);
        l += (sum2[0].x + sum2[1].x) + (sum2[0].y + sum2[1].y);
        if (j + 1 == it.nb) l_buf[(s * 2 + (item & 1)) * 128 + r] = l;  // handed to the epilogue with the last p_full
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
        ATT_T(warp, b, 7);
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
      ++item;
    }
    while (turn < turns) {  // the other stream still has blocks: keep handing the turn back
      take_turn();
      pass_turn();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == ATT_WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_cta, ATT_TMEM_COLS);
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
  const long long pairs = (total + 1) / 2;  // two streams per CTA
  const int grid = (int)(pairs < num_sms() ? pairs : num_sms());
  LaunchScope ls(st, PROF_ATTN, samples, n_pad, heads);
  if (key_bias)
    attention_kernel<true><<<grid, ATT_THREADS, ATT_SMEM_TOTAL, st>>>(tm, out, ldo, klen, row_start, n_pad, inner, heads,
                                                                      qtiles, (int)total, scale * 1.4426950408889634f,
                                                                      key_bias, ld_bias);
  else
    attention_kernel<false><<<grid, ATT_THREADS, ATT_SMEM_TOTAL, st>>>(tm, out, ldo, klen, row_start, n_pad, inner, heads,
                                                                       qtiles, (int)total, scale * 1.4426950408889634f,
                                                                       nullptr, 0);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
