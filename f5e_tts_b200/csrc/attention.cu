// Non-causal flash attention for sm_100a, dim_head 64, per-sample key length (padding mask), bf16 I/O.
//
// Replaces F.scaled_dot_product_attention with the expanded [B,H,N,N] bool mask of reference
// model/modules.py:483-492: keys n >= klen[sample] are excluded, nothing of size N x N is ever stored.
//
// One CTA = 128 queries of one (sample, head); two CTAs are resident per SM so one CTA's softmax
// overlaps the other's MMAs.  Roles inside a CTA:
//   warp 0    TMA producer: Q tile once, then K_j / V_j tiles (128 keys x 64, 128B swizzle) through a
//             3-slot ring.
//   warp 1    MMA issuer: S = Q K_j^T (UMMA 128x128x16, K-major operands) into TMEM cols [0,128);
//             O_j = P_j V_j (UMMA 128x64x16, A = P from smem, B = V as an MN-major operand) into
//             TMEM cols [128,192).
//             O accumulates ACROSS key blocks in TMEM (tcgen05.mma accumulate), it is never read back per block.
//   warps 2-5 softmax: thread == query row.  S_j is read from TMEM exactly once (128 fp32 registers, one wait) and
//             released immediately so QK_{j+1} overlaps the softmax math; row max, ex2, row sum and the bf16 P tile
//             (UMMA 128B-swizzled layout in shared memory) come from registers.  The running maximum is LAZY
//             (FlashAttention-4 style): O in TMEM is rescaled (tcgen05.ld / tcgen05.st) only when a row's maximum grew
//             by more than 2^8 since the last rescale, which after the first block is rare; P and the row sum use
//             the same stale reference so the final O / l is exact.  TMEM reads per 128x128 block drop from
//             160 KB (S twice + O) to 64 KB - TMEM read bandwidth was the limiter of the first version.
// Input layout: qkv [rows, 3*inner] with Q | K | V column blocks (what the fused QKV GEMM writes,
// RoPE already applied); output [rows, inner].
#include "host.h"

namespace f5e {

constexpr int ATT_BM = 128;   // queries per CTA
constexpr int ATT_BN = 128;   // keys per block
constexpr int ATT_DH = 64;
constexpr int ATT_THREADS = 192;
constexpr int ATT_KV_SLOTS = 3;
constexpr int ATT_TILE_BYTES = 128 * 64 * 2;  // 16 KB
constexpr int ATT_SMEM_Q = 0;
constexpr int ATT_SMEM_KV = ATT_TILE_BYTES;
constexpr int ATT_SMEM_P = ATT_SMEM_KV + ATT_KV_SLOTS * ATT_TILE_BYTES;
constexpr int ATT_SMEM_BAR = ATT_SMEM_P + 2 * ATT_TILE_BYTES;
constexpr int ATT_SMEM_TOTAL = ATT_SMEM_BAR + 128 + 1024;
constexpr uint32_t ATT_TMEM_COLS = 256;
constexpr uint32_t ATT_TMEM_O = 128;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out, int ldo,
                 const int* __restrict__ klen_ptr, int n_pad, int inner, float scale_log2) {
  const int qt = blockIdx.x, head = blockIdx.y, sample = blockIdx.z;
  const int klen = klen_ptr ? min(__ldg(klen_ptr + sample), n_pad) : n_pad;
  const int q0 = qt * ATT_BM;
  if (q0 >= klen) return;  // padded queries only: their rows are zeroed by the out-projection row mask
  const int nb = (klen + ATT_BN - 1) / ATT_BN;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ATT_SMEM_BAR);
  uint64_t* bar_q = bars;            // 1
  uint64_t* kv_full = bars + 1;      // 3
  uint64_t* kv_empty = bars + 4;     // 3
  uint64_t* s_full = bars + 7;
  uint64_t* s_free = bars + 8;
  uint64_t* p_full = bars + 9;
  uint64_t* o_full = bars + 10;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 11);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    mbar_init(bar_q, 1);
    for (int i = 0; i < ATT_KV_SLOTS; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_holder, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const int row_base = sample * n_pad;  // first row of this sample in the [rows, 3*inner] buffer

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar_q, ATT_TILE_BYTES);
      tma_load_2d(smem + ATT_SMEM_Q, &tmQKV, bar_q, head * ATT_DH, row_base + q0);
      int slot = 0;
      uint32_t phase = 0;
      for (int j = 0; j < nb; ++j) {
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {  // K_j then V_j
          mbar_wait(&kv_empty[slot], phase ^ 1);
          mbar_expect_tx(&kv_full[slot], ATT_TILE_BYTES);
          tma_load_2d(smem + ATT_SMEM_KV + slot * ATT_TILE_BYTES, &tmQKV, &kv_full[slot],
                      (kv + 1) * inner + head * ATT_DH, row_base + j * ATT_BN);
          if (++slot == ATT_KV_SLOTS) { slot = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc(kFmtBF16, ATT_BM, ATT_BN, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc(kFmtBF16, ATT_BM, ATT_DH, 0, 1);
      const uint64_t qdesc = umma_desc_kmajor_sw128(smem_u32(smem + ATT_SMEM_Q));
      const uint32_t p_addr = smem_u32(smem + ATT_SMEM_P);
      mbar_wait(bar_q, 0);
      // Ring order is K0 V0 K1 V1 ...; issue order is QK0, [QK1, PV0], [QK2, PV1], ... so tiles are
      // addressed by their index t (slot t % 3, phase (t / 3) & 1) rather than by a running cursor.
      auto tile_slot = [](int t) { return t % ATT_KV_SLOTS; };
      auto tile_phase = [](int t) { return (uint32_t)((t / ATT_KV_SLOTS) & 1); };
      auto do_qk = [&](int j) {
        const int t = 2 * j, sl = tile_slot(t);
        mbar_wait(&kv_full[sl], tile_phase(t));
        tc_fence_after();
        const uint64_t kdesc = umma_desc_kmajor_sw128(smem_u32(smem + ATT_SMEM_KV + sl * ATT_TILE_BYTES));
#pragma unroll
        for (int k = 0; k < ATT_DH / 16; ++k) umma_f16_ss(tmem_base, qdesc + 2 * k, kdesc + 2 * k, idesc_qk, k != 0);
        umma_commit(&kv_empty[sl]);
        umma_commit(s_full);
      };
      auto do_pv = [&](int j) {
        const int t = 2 * j + 1, sl = tile_slot(t);
        mbar_wait(&kv_full[sl], tile_phase(t));
        tc_fence_after();
        const uint32_t v_addr = smem_u32(smem + ATT_SMEM_KV + sl * ATT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < ATT_BN / 16; ++k) {
          const uint64_t pdesc = umma_desc_kmajor_sw128(p_addr + (k >> 2) * ATT_TILE_BYTES) + 2 * (k & 3);
          const uint64_t vdesc = umma_desc_mnmajor_sw128(v_addr + k * 2048, 1024, 1024);
          umma_f16_ss(tmem_base + ATT_TMEM_O, pdesc, vdesc, idesc_pv, (j | k) != 0);  // O accumulates over key blocks
        }
        umma_commit(&kv_empty[sl]);
        umma_commit(o_full);
      };
      do_qk(0);
      for (int j = 0; j < nb; ++j) {
        if (j + 1 < nb) {
          mbar_wait(s_free, j & 1);  // softmax finished reading S_j
          tc_fence_after();
          do_qk(j + 1);
        }
        mbar_wait(p_full, j & 1);  // P_j in smem, O rescaled if needed
        tc_fence_after();
        do_pv(j);
      }
    }
  } else {
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;  // query row inside the tile == TMEM lane
    const uint32_t t_s = tmem_base + (uint32_t(quarter * 32) << 16);
    const uint32_t t_o = t_s + ATT_TMEM_O;
    uint8_t* p_row = smem + ATT_SMEM_P + r * 128;
    const int sw = r & 7;
    float m_ref = -INFINITY;  // exponent reference (scaled log2 units); lags the true row maximum by at most 8
    float l = 0.f;

    for (int j = 0; j < nb; ++j) {
      const int kvalid = klen - j * ATT_BN;  // keys [0, kvalid) of this block are real (>= 128 except in the last block)
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float s[ATT_BN];
      {
        uint32_t* raw = reinterpret_cast<uint32_t*>(s);
        tmem_ld_32x32(t_s, *reinterpret_cast<uint32_t(*)[32]>(raw));
        tmem_ld_32x32(t_s + 32, *reinterpret_cast<uint32_t(*)[32]>(raw + 32));
        tmem_ld_32x32(t_s + 64, *reinterpret_cast<uint32_t(*)[32]>(raw + 64));
        tmem_ld_32x32(t_s + 96, *reinterpret_cast<uint32_t(*)[32]>(raw + 96));
        tmem_ld_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free);  // S_j is in registers: QK_{j+1} may overwrite the TMEM columns now
      if (kvalid < ATT_BN) {
#pragma unroll
        for (int i = 0; i < ATT_BN; ++i)
          if (i >= kvalid) s[i] = -INFINITY;
      }
      float mx = s[0];
#pragma unroll
      for (int i = 1; i < ATT_BN; ++i) mx = fmaxf(mx, s[i]);
      const float m_new = fmaxf(m_ref, mx * scale_log2);
      const bool grow = m_new - m_ref > 8.0f;  // always true in the first block (m_ref = -inf)
      if (j > 0) {  // PV_{j-1} finished: O_{j-1} is complete and the P buffer may be rewritten
        mbar_wait(o_full, (j - 1) & 1);
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
      const float neg_m = -m_ref;
      float rowsum = 0.f;
#pragma unroll
      for (int i = 0; i < ATT_BN; ++i) {
        s[i] = ex2_approx(fmaf(s[i], scale_log2, neg_m));  // masked keys: ex2(-inf) = 0
        rowsum += s[i];
      }
      l += rowsum;
      // 128 keys = 16 chunks of 16 B; chunk g lives in atom g>>3, position (g&7)^sw
#pragma unroll
      for (int g = 0; g < 16; ++g) {
        uint4 u;
        u.x = pack_bf16x2(s[g * 8 + 0], s[g * 8 + 1]);
        u.y = pack_bf16x2(s[g * 8 + 2], s[g * 8 + 3]);
        u.z = pack_bf16x2(s[g * 8 + 4], s[g * 8 + 5]);
        u.w = pack_bf16x2(s[g * 8 + 6], s[g * 8 + 7]);
        *reinterpret_cast<uint4*>(p_row + (g >> 3) * ATT_TILE_BYTES + (((g & 7) ^ sw) << 4)) = u;
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }
    mbar_wait(o_full, (nb - 1) & 1);
    tc_fence_after();
    const int q = q0 + r;
    const float inv = 1.0f / l;
    __nv_bfloat16* orow = out + (size_t)(row_base + q) * ldo + head * ATT_DH;
#pragma unroll 1
    for (int c = 0; c < ATT_DH / 32; ++c) {
      uint32_t raw[32];
      tmem_ld_32x32(t_o + c * 32, raw);
      tmem_ld_wait();
      if (q < n_pad) {
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
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ATT_TMEM_COLS);
  }
}

int attention_configure() {
  F5E_CUDA_CHECK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM_TOTAL));
  return 0;
}

int attention(cudaStream_t st, const __nv_bfloat16* qkv, int ld_qkv, __nv_bfloat16* out, int ldo, const int* klen,
              int samples, int n_pad, int heads, float scale) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(samples > 0 && n_pad > 0 && heads > 0, "attention: empty problem");
  F5E_REQUIRE(ldo % 8 == 0, "attention: ldo must be a multiple of 8");
  const int inner = heads * ATT_DH;
  CUtensorMap tm;
  int rc = make_tmap_2d(&tm, qkv, 2, (uint64_t)3 * inner, (uint64_t)samples * n_pad, (uint64_t)ld_qkv * 2, ATT_DH, 128);
  if (rc) return rc;
  dim3 grid((n_pad + ATT_BM - 1) / ATT_BM, heads, samples);
  LaunchScope ls(st, PROF_ATTN, samples, n_pad, heads);
  attention_kernel<<<grid, ATT_THREADS, ATT_SMEM_TOTAL, st>>>(tm, out, ldo, klen, n_pad, inner,
                                                              scale * 1.4426950408889634f);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
