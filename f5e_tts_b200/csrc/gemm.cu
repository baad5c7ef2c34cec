// Dense bf16 GEMM for sm_100a:  C[M,N] = epilogue(A[M,K] @ W[N,K]^T), fp32 accumulation.
//
// Persistent, warp-specialised kernel (one CTA per SM, 320 threads):
//   warp 0   : TMA producer  - streams 128x64 A tiles and BNx64 W tiles (128B swizzle) through a
//              STAGES-deep shared-memory ring guarded by full/empty mbarriers.
//   warp 1   : MMA issuer    - one thread issues tcgen05.mma (UMMA 128 x BN x 16) into one of TWO
//              TMEM accumulator stages, tcgen05.commit releases ring slots / publishes accumulators.
//   warps 2-9: epilogue      - warp w owns TMEM lanes 32*(w%4).. (its 32 rows) and one half of the tile's columns,
//              64 columns at a time: tcgen05.ld -> registers -> fused epilogue (bias / RoPE / activation / row mask /
//              gate) -> 128B-swizzled staging block in shared memory -> ONE TMA store per 32x64 block.  A residual that
//              aliases the output (x += gate * branch, every DiT block) is not read at all: the block is sent with a
//              TMA reduce-add (cp.reduce.async.bulk.tensor .add), so the read-modify-write happens in L2.
//              The epilogue of tile i overlaps the main loop of tile i+1 through the second TMEM stage.
//
// Replaces the F.linear (cuBLAS) calls of reference model/modules.py:442-449 (to_q/k/v), :494
// (to_out), :342-353 (FeedForward), dit.py:169-177 (InputEmbedding.proj), dit.py:470 (proj_out).
#include "host.h"

namespace f5e {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;
constexpr int GEMM_STAGING_BYTES = 4096;  // 32 rows x 128 bytes per epilogue warp
#ifndef GEMM_PAIR
#define GEMM_PAIR 1                       // 0: never use CTA pairs (A/B switch for measurements)
#endif
// 1: 256 x 128 pair tiles where 256 x 256 tiles quantise badly over the 74 pairs (out-projection / FF2 at the bench shape).
// Measured 8.7 % slower per step, so it is off; kept as a switch for the next look at it.
#ifndef GEMM_NARROW_PAIR
#define GEMM_NARROW_PAIR 0
#endif
#ifndef GEMM_NARROW_STAGES
#define GEMM_NARROW_STAGES 6
#endif

// CG = CTAs per tile: 1, or 2 = a CTA PAIR (cta_group::2) computing a 256 x BN tile; every CTA then stages only half of
// the W tile, which halves the shared-memory operand traffic per MMA: a single-CTA 128x256 tile needs 192 B/clk of
// operand reads (A 4 KB + B 8 KB per 64-cycle UMMA), above the 128 B/clk an SM's shared memory delivers.
template <int BN, int STAGES, int CG = 1>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = (BN / CG) * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGING_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int NBUF = CG;  // staging blocks per epilogue warp: the pair kernel has room to double-buffer them
  static constexpr int BAR_OFFSET = STAGING_OFFSET + GEMM_EPI_WARPS * NBUF * GEMM_STAGING_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;  // barriers + alignment slack
};

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

// Send the warp's staged 32 x 128-byte block to global memory (plain store or reduce-add).  ONE elected thread issues:
// elect.sync picks the same leader for the same (full) member mask every time, so the bulk groups it commits are the
// ones it later waits on, and - unlike `lane == 0` - the compiler issues UTMASTG / UTMAREDG without a serialisation loop.
__device__ __forceinline__ void staging_flush(const CUtensorMap* tmC, const uint8_t* stage, int c0, int c1, bool accumulate,
                                              int lane) {
  fence_proxy_async_smem();
  __syncwarp();
  if (elect_one()) {
    if (accumulate) tma_reduce_add_2d(tmC, stage, c0, c1);
    else tma_store_2d(tmC, stage, c0, c1);
    tma_store_commit();
  }
}
// The staging block may be rewritten once the previous bulk store has finished READING it.
template <int NBUF>
__device__ __forceinline__ void staging_acquire(int lane) {
  if (elect_one()) tma_store_wait_read<NBUF - 1>();  // the block flushed NBUF flushes ago has been read out
  __syncwarp();
}
// 16-byte slot j (0..7) of row r inside a 128B-swizzled 32-row block
__device__ __forceinline__ uint4* staging_slot(uint8_t* stage, int r, int j) {
  return reinterpret_cast<uint4*>(stage + r * 128 + ((j ^ (r & 7)) << 4));
}

// VQ = true: the codebook instantiation (BN = 256 = two groups x 128 padded columns, CG = 1): the epilogue takes the
// per-row, per-group arg-max, the softmax column sums and the code-vector gather instead of storing anything (see VqEpi).
// STATS = true (CG = 2 only): x += gate * (acc + bias) with the LayerNorm statistics of the new rows taken on the way
// (EpiParams::stats_out).
template <int BN, int STAGES, int CG, bool VQ = false, bool STATS = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmA2, int M, int N, int K,
                 EpiParams ep, int accumulate, VqEpi vq) {
  using S = GemmSmem<BN, STAGES, CG>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint64_t* ld_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET + 128);  // STATS: 2 per epilogue warp (residual tiles)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int TILE_M = GEMM_BM * CG;
  const int m_blocks = (M + TILE_M - 1) / TILE_M;
  const int n_blocks = (N + BN - 1) / BN;
  const int num_tiles = m_blocks * n_blocks;
  const int nkb = (K + GEMM_BK - 1) / GEMM_BK;
  constexpr uint32_t TMEM_COLS = 2 * BN;  // 256 or 512: power of two
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;   // 0 = leader: issues the MMAs of the pair
  const int tile0 = blockIdx.x / CG, tile_stride = gridDim.x / CG;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (ep.ln_a2) tma_prefetch_desc(&tmA2);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], GEMM_EPI_WARPS * CG);  // the leader's barrier collects the epilogue warps of both CTAs
    }
    if (STATS)
      for (int i = 0; i < 2 * GEMM_EPI_WARPS; ++i) mbar_init(&ld_bar[i], 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    if (CG == 2) tmem_alloc_2cta(tmem_holder, TMEM_COLS);
    else tmem_alloc(tmem_holder, TMEM_COLS);
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) touched no global memory and may
  // have run under the tail of the previous kernel; from here on its results are needed.
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    if (elect_one()) {  // elect.sync, not `lane == 0`: the compiler then knows ONE thread is active and issues
                        // UTMALDG / UTCHMMA directly instead of wrapping each in an ELECT / BRA.U.ANY loop
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          // folded LayerNorm: the LAST K block of A comes from the per-row statistics matrix (-mean split in bf16),
          // against the column sums appended to the folded weight, so the tensor core does the mean correction
          const bool from_a2 = ep.ln_a2 && kb == nkb - 1;
          const CUtensorMap* ta = from_a2 ? &tmA2 : &tmA;
          const int ka = from_a2 ? 0 : kb * GEMM_BK;
          if (CG == 2) {  // both CTAs' bytes complete on the leader's barrier
#if defined(GEMM_DUP_A) || defined(GEMM_DUP_B)  // energy experiment: one operand tile fetched twice (same bytes, same slot)
#ifdef GEMM_DUP_A
            constexpr uint32_t DUP_BYTES = S::A_BYTES;
#else
            constexpr uint32_t DUP_BYTES = S::STAGE_BYTES - S::A_BYTES;
#endif
            if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * (S::STAGE_BYTES + DUP_BYTES));
#ifdef GEMM_DUP_A
            tma_load_2d_2cta(sa, ta, &full_bar[stage], ka, m_blk * TILE_M + rank * GEMM_BM);
#else
            tma_load_2d_2cta(sb, &tmB, &full_bar[stage], kb * GEMM_BK, n_blk * BN + rank * (BN / 2));
#endif
#else
            if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * S::STAGE_BYTES);
#endif
            tma_load_2d_2cta(sa, ta, &full_bar[stage], ka, m_blk * TILE_M + rank * GEMM_BM);
            tma_load_2d_2cta(sb, &tmB, &full_bar[stage], kb * GEMM_BK, n_blk * BN + rank * (BN / 2));
          } else {
            mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
            tma_load_2d(sa, ta, &full_bar[stage], ka, m_blk * GEMM_BM);
            tma_load_2d(sb, &tmB, &full_bar[stage], kb * GEMM_BK, n_blk * BN);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = umma_idesc(kFmtBF16, TILE_M, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::STAGE_BYTES);
          const uint64_t adesc = umma_desc_kmajor_sw128(sa);
          const uint64_t bdesc = umma_desc_kmajor_sw128(sa + S::A_BYTES);
          const int ksteps = (ep.ln_a2 && kb == nkb - 1) ? 1 : GEMM_BK / 16;  // the statistics block is 16 columns deep
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {  // +32 bytes (2 x 16B units) per UMMA_K inside the swizzle atom
            if (k >= ksteps) break;
            if (CG == 2) umma_f16_ss_2cta(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
            else umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          if (CG == 2) umma_commit_2cta(&empty_bar[stage], 3);  // frees the slot in both CTAs
          else umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (CG == 2) umma_commit_2cta(&tfull_bar[acc], 3);
        else umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    const int quarter = warp & 3;            // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;        // which half of the tile's columns
    constexpr int WCOLS = BN / 2;            // columns per epilogue warp
    uint8_t* const stage_base = smem + S::STAGING_OFFSET + (warp - 2) * S::NBUF * GEMM_STAGING_BYTES;
    uint8_t* stage_buf = stage_base;
    int sbuf = 0;
    auto acquire = [&]() {  // next staging block of this warp (round robin), free once its last bulk store has read it
      stage_buf = stage_base + sbuf * GEMM_STAGING_BYTES;
      sbuf = (sbuf + 1) % S::NBUF;
      staging_acquire<S::NBUF>(lane);
    };
    int acc = 0;
    uint32_t acc_phase = 0, ld_phase = 0;
    for (int tile = tile0; tile < num_tiles; tile += tile_stride) {
      const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
      const int row0 = m_blk * TILE_M + rank * GEMM_BM + quarter * 32;
      const int row = row0 + lane;
      const int rowc = row < M ? row : M - 1;  // clamped row for table look-ups (stores are clipped by TMA)
      int sample, frame;
      if (ep.row_info) {
        const int ri = __ldg(ep.row_info + rowc);
        sample = ri >> 16;
        frame = ri & 0xffff;
      } else {
        sample = rowc / ep.n_pad;
        frame = rowc - sample * ep.n_pad;
      }
      const bool row_masked = ep.row_len && frame >= __ldg(ep.row_len + sample);
      const int wcol0 = n_blk * BN + half * WCOLS;
      // RoPE: angle(frame, j) = frame * inv_freq[j], shared by the interleaved pair (2j, 2j+1) of every head.  The angles
      // are evaluated per 64-column chunk (= one head): no table traffic and nothing live across chunks.  (Reading
      // per-row cos/sin tables cost 28 us per QKV launch: 512 scattered LSU wavefronts per warp and 64 live registers.)
      const bool tile_rope = ep.rope_inv_freq && wcol0 < 2 * ep.rope_inner && wcol0 < N;
      const float frame_f = (float)frame;
      const float* gate = ep.gate ? ep.gate + (ep.gate_stride ? (size_t)sample * ep.gate_stride : 0) : nullptr;

      if constexpr (STATS) {
        // The residual tiles of this warp's 32 rows x 128 columns travel (TMA, same map and swizzle as the store) while the
        // main loop of the tile still runs.  Both staging blocks must be free: every earlier bulk store has read them.
        __syncwarp();
        if (elect_one()) {
          tma_store_wait_read<0>();
          uint64_t* lb = &ld_bar[(warp - 2) * 2];
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            mbar_expect_tx(&lb[cc], GEMM_STAGING_BYTES);
            tma_load_2d(stage_base + cc * GEMM_STAGING_BYTES, &tmC, &lb[cc], wcol0 + cc * 64, row0);
          }
        }
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BN + half * WCOLS;
      if constexpr (STATS) {
        static_assert(!STATS || (CG == 2 && BN == 256), "the statistics epilogue needs two staging blocks per warp");
        float cmean[2], cm2[2];
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int col0 = wcol0 + cc * 64;
          float v[64];
          {
            uint32_t r0[32], r1[32];
            tmem_ld_32x32(taddr + cc * 64, r0);
            tmem_ld_32x32(taddr + cc * 64 + 32, r1);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) { v[i] = __uint_as_float(r0[i]); v[32 + i] = __uint_as_float(r1[i]); }
          }
#pragma unroll
          for (int i = 0; i < 64; i += 4) {
            float4 b = ep.bias ? __ldg(reinterpret_cast<const float4*>(ep.bias + col0 + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 gt = gate ? __ldg(reinterpret_cast<const float4*>(gate + col0 + i)) : make_float4(1.f, 1.f, 1.f, 1.f);
            v[i] = row_masked ? 0.f : (v[i] + b.x) * gt.x;
            v[i + 1] = row_masked ? 0.f : (v[i + 1] + b.y) * gt.y;
            v[i + 2] = row_masked ? 0.f : (v[i + 2] + b.z) * gt.z;
            v[i + 3] = row_masked ? 0.f : (v[i + 3] + b.w) * gt.w;
          }
          uint8_t* buf = stage_base + cc * GEMM_STAGING_BYTES;
          mbar_wait(&ld_bar[(warp - 2) * 2 + cc], ld_phase);
          // new row = bf16(old + branch); the statistics are those of the ROUNDED values, taken in the same pass as shifted
          // sums (shift = the chunk's first value: sum d, sum d^2 with d = y - shift lose nothing to cancellation)
          float shift = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4* slot = staging_slot(buf, lane, j);
            const uint4 r = *slot;
            const float2 a0 = unpack_bf16x2(r.x), a1 = unpack_bf16x2(r.y), a2 = unpack_bf16x2(r.z), a3 = unpack_bf16x2(r.w);
            uint4 o;
            o.x = pack_bf16x2(a0.x + v[8 * j], a0.y + v[8 * j + 1]);
            o.y = pack_bf16x2(a1.x + v[8 * j + 2], a1.y + v[8 * j + 3]);
            o.z = pack_bf16x2(a2.x + v[8 * j + 4], a2.y + v[8 * j + 5]);
            o.w = pack_bf16x2(a3.x + v[8 * j + 6], a3.y + v[8 * j + 7]);
            *slot = o;
            const float2 y0 = unpack_bf16x2(o.x), y1 = unpack_bf16x2(o.y), y2 = unpack_bf16x2(o.z), y3 = unpack_bf16x2(o.w);
            if (j == 0) shift = y0.x;
            const float d[8] = {y0.x - shift, y0.y - shift, y1.x - shift, y1.y - shift, y2.x - shift, y2.y - shift, y3.x - shift, y3.y - shift};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              s1 += d[i];
              s2 = fmaf(d[i], d[i], s2);
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            tma_store_2d(&tmC, buf, col0, row0);
            tma_store_commit();
          }
          cmean[cc] = shift + s1 * (1.0f / 64.0f);
          cm2[cc] = fmaxf(s2 - s1 * s1 * (1.0f / 64.0f), 0.f);
        }
        ld_phase ^= 1;
        if (row < M) {  // two equal halves: Chan's update with n0 = n1 = 64
          const float d = cmean[1] - cmean[0];
          ep.stats_out[(size_t)row * ep.stats_slots + (wcol0 >> 7)] =
              make_float2(0.5f * (cmean[0] + cmean[1]), cm2[0] + cm2[1] + 32.0f * d * d);
        }
      } else if constexpr (VQ) {
        // ---- codebook decision (reference model/modules.py:889-901, 940-943): this thread owns row `row`, group `half` ----
        const int g = half;
        if (g < vq.G) {  // warp-uniform
          const float* bias_g = ep.bias + g * WCOLS;  // padded bias: -1e30 beyond the group's V columns
          auto load32 = [&](int c, float (&v)[32]) {
            uint32_t raw[32];
            tmem_ld_32x32(taddr + c * 32, raw);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(bias_g + c * 32 + i));
              v[i] = __uint_as_float(raw[i]) + b.x;
              v[i + 1] = __uint_as_float(raw[i + 1]) + b.y;
              v[i + 2] = __uint_as_float(raw[i + 2]) + b.z;
              v[i + 3] = __uint_as_float(raw[i + 3]) + b.w;
            }
          };
          float best = -INFINITY, second = -INFINITY;
          int bi = 0;
#pragma unroll 1
          for (int c = 0; c < WCOLS / 32; ++c) {  // first maximum wins ties: strict > while scanning upwards
            float v[32];
            load32(c, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (v[i] > best) { second = best; best = v[i]; bi = c * 32 + i; }
              else second = fmaxf(second, v[i]);
            }
          }
          float se = 0.f;
#pragma unroll 1
          for (int c = 0; c < WCOLS / 32; ++c) {
            float v[32];
            load32(c, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) se += __expf(v[i] - best);
          }
          const float inv = row < M ? 1.0f / se : 0.f;
          // softmax column sums over this warp's 32 rows: transpose through the warp's staging block, one column per lane
          float* stg = reinterpret_cast<float*>(stage_base);
#pragma unroll 1
          for (int c = 0; c < WCOLS / 32; ++c) {
            if (c * 32 >= vq.V) break;
            float v[32];
            load32(c, v);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 32; ++i) stg[lane * 32 + ((i + lane) & 31)] = __expf(v[i] - best) * inv;  // skewed: no bank conflicts
            __syncwarp();
            float colsum = 0.f;
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) colsum += stg[rr * 32 + ((lane + rr) & 31)];
            if (c * 32 + lane < vq.V) atomicAdd(vq.sums + g * vq.V + c * 32 + lane, colsum);
          }
          const bool tie = row < M && vq.V > 1 && best - second <= vq.refine_tol * fmaxf(1.0f, fabsf(best));
          if (row < M) {
            if (tie) {  // a tensor-core rounding could flip this one: vq_refine_kernel decides it in sequential fp32
              const int slot = atomicAdd(vq.tie_count, 1);
              vq.tie_list[slot] = make_int2(row, g);
            } else {
              vq.targets[(size_t)row * vq.G + g] = bi;
              atomicAdd(vq.hist + g * vq.V + bi, 1);
            }
          }
          // gather of the selected code vectors, one row at a time with the whole warp (coalesced)
          const int pick = (row < M && !tie) ? bi : -1;
#pragma unroll 1
          for (int rr = 0; rr < 32; ++rr) {
            const int idx = __shfl_sync(0xffffffffu, pick, rr);
            if (idx < 0) continue;  // warp-uniform
            const float* src = vq.vars + (size_t)((vq.combine_groups ? 0 : g * vq.V) + idx) * vq.vd;
            const size_t o = (size_t)(row0 + rr) * vq.G * vq.vd + (size_t)g * vq.vd;
            if (vq.xq_bf16) {
              __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(vq.xq) + o;
              for (int cidx = lane * 2; cidx < vq.vd; cidx += 64)
                *reinterpret_cast<uint32_t*>(dst + cidx) = pack_bf16x2(__ldg(src + cidx), __ldg(src + cidx + 1));
            } else {
              float* dst = reinterpret_cast<float*>(vq.xq) + o;
              for (int cidx = lane * 2; cidx < vq.vd; cidx += 64)
                *reinterpret_cast<float2*>(dst + cidx) = __ldg(reinterpret_cast<const float2*>(src + cidx));
            }
          }
        }
      } else
#pragma unroll 1
      for (int cc = 0; cc < WCOLS / 64; ++cc) {
        const int col0 = wcol0 + cc * 64;
        if (col0 >= N) break;  // warp-uniform
        float v[64];
        {
          uint32_t r0[32], r1[32];
          tmem_ld_32x32(taddr + cc * 64, r0);
          tmem_ld_32x32(taddr + cc * 64 + 32, r1);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) { v[i] = __uint_as_float(r0[i]); v[32 + i] = __uint_as_float(r1[i]); }
        }
        const bool full = col0 + 64 <= N;
        if (ep.ln_rstd) {  // folded LayerNorm: acc already holds x . W' - mean * colsum(W'); scale by rstd, add d (as bias)
          const float rs = __ldg(ep.ln_rstd + rowc);
#pragma unroll
          for (int i = 0; i < 64; i += 4) {
            if (col0 + i < N) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(ep.bias + col0 + i));
              v[i] = fmaf(v[i], rs, b.x); v[i + 1] = fmaf(v[i + 1], rs, b.y);
              v[i + 2] = fmaf(v[i + 2], rs, b.z); v[i + 3] = fmaf(v[i + 3], rs, b.w);
            }
          }
        } else if (ep.bias) {
          if (full) {
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(ep.bias + col0 + i));
              v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (col0 + i < N) v[i] += __ldg(ep.bias + col0 + i);
          }
        }
        if (tile_rope && col0 < 2 * ep.rope_inner && ((col0 % ep.rope_inner) >> 6) < ep.rope_heads) {
          {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              // x = frame * inv_freq[j] is the same fp32 product the reference's table holds; reduce to [-pi, pi] with a
              // two-term 2*pi (error ~1e-7 rad), then the hardware sin / cos (abs error ~5e-7): far below bf16 rounding
              const float x = frame_f * __ldg(ep.rope_inv_freq + j);
              const float k = rintf(x * 0.15915494309189535f);
              float r = fmaf(k, -6.2831854820251465f, x);
              r = fmaf(k, 1.7484555e-7f, r);
              float sn, cs;
              asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(r));
              asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(r));
              const float a = v[2 * j], b = v[2 * j + 1];
              v[2 * j] = a * cs - b * sn;
              v[2 * j + 1] = b * cs + a * sn;
            }
          }
        }
        switch (ep.act) {
          case ACT_GELU_TANH:
#pragma unroll
            for (int i = 0; i < 64; i += 2) {
              const float2 g = gelu_tanh_fast2(make_float2(v[i], v[i + 1]));
              v[i] = g.x;
              v[i + 1] = g.y;
            }
            break;
          case ACT_GELU_ERF:
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = gelu_erf(v[i]);
            break;
          case ACT_MISH:
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = mish(v[i]);
            break;
          case ACT_RELU:
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = fmaxf(v[i], 0.0f);
            break;
          case ACT_SILU:
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = silu(v[i]);
            break;
          default:
            break;
        }
        if (row_masked) {
#pragma unroll
          for (int i = 0; i < 64; ++i) v[i] = 0.0f;
        }
        if (gate) {
          if (full) {
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
              const float4 g = __ldg(reinterpret_cast<const float4*>(gate + col0 + i));
              v[i] *= g.x; v[i + 1] *= g.y; v[i + 2] *= g.z; v[i + 3] *= g.w;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (col0 + i < N) v[i] *= __ldg(gate + col0 + i);
          }
        }
        if (ep.residual && !accumulate && row < M) {  // residual that does NOT alias the output: read it here
          if (ep.res_f32) {
            const float* r = reinterpret_cast<const float*>(ep.residual) + (size_t)row * ep.ldr + col0;
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (col0 + i < N) v[i] += r[i];
          } else {
            const __nv_bfloat16* r = ep.residual + (size_t)row * ep.ldr + col0;
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (col0 + i < N) v[i] += __bfloat162float(r[i]);
          }
        }
        // ---- registers -> swizzled staging -> TMA store / reduce-add
        if (ep.out_f32) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {  // 32 fp32 columns = 128 bytes per row per block
            if (col0 + 32 * h >= N) break;
            acquire();
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *staging_slot(stage_buf, lane, j) =
                  make_uint4(__float_as_uint(v[32 * h + 4 * j]), __float_as_uint(v[32 * h + 4 * j + 1]),
                             __float_as_uint(v[32 * h + 4 * j + 2]), __float_as_uint(v[32 * h + 4 * j + 3]));
            staging_flush(&tmC, stage_buf, col0 + 32 * h, row0, accumulate != 0, lane);
          }
        } else if (ep.split3 > 0) {
          uint4 lo[8];
          acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 hi;
            uint32_t* hp = &hi.x;
            uint32_t* lp = &lo[j].x;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float a = v[8 * j + 2 * q], b = v[8 * j + 2 * q + 1];
              const float ah = __bfloat162float(__float2bfloat16(a)), bh = __bfloat162float(__float2bfloat16(b));
              hp[q] = pack_bf16x2(ah, bh);
              lp[q] = pack_bf16x2(a - ah, b - bh);
            }
            *staging_slot(stage_buf, lane, j) = hi;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            tma_store_2d(&tmC, stage_buf, col0, row0);
            tma_store_2d(&tmC, stage_buf, col0 + ep.split3, row0);
            tma_store_commit();
          }
          acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j) *staging_slot(stage_buf, lane, j) = lo[j];
          staging_flush(&tmC, stage_buf, col0 + 2 * ep.split3, row0, false, lane);
        } else {
          acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *staging_slot(stage_buf, lane, j) =
                make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                           pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
          staging_flush(&tmC, stage_buf, col0, row0, accumulate != 0, lane);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_cluster(&tempty_bar[acc], 0);
        else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    __syncwarp();
    if (elect_one()) tma_store_wait<0>();  // every bulk store has completed before the CTA's shared memory goes away
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all();  // no CTA leaves while its peer may still signal its barriers / read its smem
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_2cta(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int STAGES, int CG, bool VQ = false, bool STATS = false>
static int launch_gemm(cudaStream_t st, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                       const CUtensorMap& tmA2, int M, int N, int K, const EpiParams& ep, int accumulate,
                       const VqEpi& vq = VqEpi{}) {
  using S = GemmSmem<BN, STAGES, CG>;
  const int tiles = ((M + GEMM_BM * CG - 1) / (GEMM_BM * CG)) * ((N + BN - 1) / BN);
  const int slots = num_sms() / CG;  // CTAs (CG = 1) or CTA pairs (CG = 2) that can be resident
  const int grid = (tiles < slots ? tiles : slots) * CG;
  LaunchScope ls(st, PROF_GEMM, M, N, K);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = S::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled(true) ? 2 : 1;
  F5E_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<BN, STAGES, CG, VQ, STATS>, tmA, tmB, tmC, tmA2, M, N, K, ep, accumulate, vq));
  return 0;
}

// True when gemm_bf16 runs this shape on the CTA-pair kernel, the only one whose epilogue can take the LayerNorm statistics
// (two staging blocks per warp for the TMA-staged residual tiles).
bool gemm_can_fuse_stats(int M, int N) {
  return GEMM_PAIR && N > 128 && N % 256 == 0 && !GEMM_NARROW_PAIR && ((M + 255) / 256) * (N / 256) >= num_sms() / 2;
}

int gemm_configure() {
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<256, 4, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<256, 4, 1>::TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<256, 5, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<256, 5, 2>::TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<128, 6, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<128, 6, 1>::TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<256, 4, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<256, 4, 1>::TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<256, 5, 2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<256, 5, 2>::TOTAL));
#if GEMM_NARROW_PAIR
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<128, GEMM_NARROW_STAGES, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<128, GEMM_NARROW_STAGES, 2>::TOTAL));
#endif
  return 0;
}

int gemm_bf16(cudaStream_t st, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int ldw, int M, int N,
              int K, const EpiParams& ep) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_bf16: empty problem (M=%d N=%d K=%d)", M, N, K);
  F5E_REQUIRE(ep.out != nullptr, "gemm_bf16: null output");
  F5E_REQUIRE(ep.n_pad > 0, "gemm_bf16: n_pad must be positive");
  F5E_REQUIRE((ep.ln_rstd == nullptr) == (ep.ln_a2 == nullptr), "gemm_bf16: folded LayerNorm needs ln_rstd and ln_a2");
  F5E_REQUIRE(!ep.ln_rstd || (ep.bias && N % 4 == 0 && K % GEMM_BK == 0 && K > GEMM_BK),
              "gemm_bf16: folded LayerNorm needs a bias (d), N %% 4 == 0 and K = dim + 64");
  F5E_REQUIRE(!(ep.split3 > 0) || (N % 64 == 0 && !ep.out_f32), "gemm_bf16: split3 output needs N %% 64 == 0 and bf16");
  const bool wide = N > 128;
  // x += epilogue(...) with the residual aliasing the output: the read-modify-write is a TMA reduce-add
  const bool alias = ep.residual && (const void*)ep.residual == (const void*)ep.out && ep.ldr == ep.ldo &&
                     (ep.res_f32 != 0) == (ep.out_f32 != 0) && ep.split3 == 0;
  CUtensorMap tmA, tmB, tmC, tmA2;
  // folded LayerNorm: K counts the 64-column statistics block, which lives in its own [M, 64] matrix
  int rc = make_tmap_2d(&tmA, A, 2, (uint64_t)(ep.ln_a2 ? K - GEMM_BK : K), (uint64_t)M, (uint64_t)lda * 2, GEMM_BK, GEMM_BM);
  if (rc) return rc;
  if (ep.ln_a2) {
    rc = make_tmap_2d(&tmA2, ep.ln_a2, 2, (uint64_t)GEMM_BK, (uint64_t)M, (uint64_t)GEMM_BK * 2, GEMM_BK, GEMM_BM);
    if (rc) return rc;
  } else {
    tmA2 = tmA;
  }
  // CTA pairs when there are enough 256-row tiles to fill the 74 pairs; each CTA then loads a 128-row half of the W tile
  const bool pair = wide && GEMM_PAIR && ((M + 255) / 256) * ((N + 255) / 256) >= num_sms() / 2;
  if (ep.stats_out) {
    F5E_REQUIRE(gemm_can_fuse_stats(M, N) && alias && !ep.out_f32 && ep.act == ACT_NONE && !ep.rope_inv_freq && !ep.ln_rstd &&
                    ep.stats_slots == N / 128,
                "gemm_bf16: stats_out needs the CTA-pair kernel (gemm_can_fuse_stats), x += gate * branch in place, bf16, no "
                "activation / RoPE / folded LayerNorm, stats_slots == N / 128");
  }
#if GEMM_NARROW_PAIR
  auto wave_eff = [&](int bn) {
    const long long t = (long long)((M + 255) / 256) * ((N + bn - 1) / bn);
    const long long slots = num_sms() / 2, waves = (t + slots - 1) / slots;
    return (double)t / (double)(waves * slots);
  };
  const bool narrow = pair && N % 128 == 0 && wave_eff(128) > wave_eff(256) + 0.04;
#else
  const bool narrow = false;
#endif
  // Latency shapes (batch 1: M ~ 2 k rows): 128 x 256 tiles leave most SMs idle (N = 1024: 60 tiles on 148 SMs); 128 x 128 tiles fill
  // them.  Only when the wide tiling cannot occupy the machine - at full occupancy the wide tile is the faster one.
#ifndef GEMM_SMALL_M_TILES
#define GEMM_SMALL_M_TILES 1
#endif
  auto fill = [&](int bn) {  // fraction of SM-waves doing work under the static tile schedule
    const long long t = (long long)((M + GEMM_BM - 1) / GEMM_BM) * ((N + bn - 1) / bn);
    const long long waves = (t + num_sms() - 1) / num_sms();
    return (double)t / (double)(waves * num_sms());
  };
  // the 128 x 128 tile is smem-bandwidth-bound (~0.85 of the wide tile's rate): it must win by more than that
  const bool small = GEMM_SMALL_M_TILES && wide && !pair && N % 128 == 0 && 0.85 * fill(128) > fill(256);
  rc = make_tmap_2d(&tmB, W, 2, (uint64_t)K, (uint64_t)N, (uint64_t)ldw * 2, GEMM_BK,
                    narrow ? 64 : ((wide && !pair && !small) ? 256 : 128));
  if (rc) return rc;
  if (ep.out_f32)
    rc = make_tmap_2d(&tmC, ep.out, 4, (uint64_t)N, (uint64_t)M, (uint64_t)ep.ldo * 4, 32, 32);
  else
    rc = make_tmap_2d(&tmC, ep.out, 2, (uint64_t)(ep.split3 > 0 ? 3 * ep.split3 : N), (uint64_t)M, (uint64_t)ep.ldo * 2, 64, 32);
  if (rc) return rc;
#if GEMM_NARROW_PAIR
  if (narrow) return launch_gemm<128, GEMM_NARROW_STAGES, 2>(st, tmA, tmB, tmC, tmA2, M, N, K, ep, alias ? 1 : 0);
#endif
  if (pair && ep.stats_out) return launch_gemm<256, 5, 2, false, true>(st, tmA, tmB, tmC, tmA2, M, N, K, ep, 0);
  if (pair) return launch_gemm<256, 5, 2>(st, tmA, tmB, tmC, tmA2, M, N, K, ep, alias ? 1 : 0);
  if (wide && !small) return launch_gemm<256, 4, 1>(st, tmA, tmB, tmC, tmA2, M, N, K, ep, alias ? 1 : 0);
  return launch_gemm<128, 6, 1>(st, tmA, tmB, tmC, tmA2, M, N, K, ep, alias ? 1 : 0);
}

// Logit GEMM of the codebook (A [M, K] bf16x3-split activations, Wp [256, K] = the projection with every group padded to
// 128 rows, bias_p [256] with -1e30 on the pad columns) with the decision fused into the epilogue: nothing is stored but
// targets / x_q / histogram / softmax sums / the near-tie list (VqEpi).
int gemm_vq(cudaStream_t st, const __nv_bfloat16* A, int lda, const __nv_bfloat16* Wp, int ldw, const float* bias_p, int M, int K,
            const VqEpi& vq) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(M > 0 && K > 0 && K % 8 == 0 && vq.G >= 1 && vq.G <= 2 && vq.V >= 1 && vq.V <= 128 && vq.vd % 2 == 0,
              "gemm_vq: needs 1-2 groups of at most 128 entries (G=%d V=%d)", vq.G, vq.V);
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, 2, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, GEMM_BK, GEMM_BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tmB, Wp, 2, (uint64_t)K, (uint64_t)256, (uint64_t)ldw * 2, GEMM_BK, 256);
  if (rc) return rc;
  EpiParams ep{};
  ep.bias = bias_p;
  ep.n_pad = M;
  return launch_gemm<256, 4, 1, true>(st, tmA, tmB, tmA /* no output map */, tmA, M, 256, K, ep, 0, vq);
}

}  // namespace f5e
