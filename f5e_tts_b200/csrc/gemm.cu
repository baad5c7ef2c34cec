// Dense bf16 GEMM for sm_100a:  C[M,N] = epilogue(A[M,K] @ W[N,K]^T), fp32 accumulation.
//
// Persistent, warp-specialised kernel (one CTA per SM):
//   warp 0   : TMA producer  - streams 128x64 A tiles and BNx64 W tiles (128B swizzle) through a
//              STAGES-deep shared-memory ring guarded by full/empty mbarriers.
//   warp 1   : MMA issuer    - one thread issues tcgen05.mma (UMMA 128 x BN x 16) into one of TWO
//              TMEM accumulator stages, tcgen05.commit releases ring slots / publishes accumulators.
//   warps 2-5: epilogue      - tcgen05.ld the accumulator (thread == row, 32 columns at a time),
//              apply the fused epilogue (bias / RoPE / activation / row mask / gate / residual)
//              and store.  It overlaps with the next tile's main loop thanks to the second TMEM stage.
//
// Replaces the F.linear (cuBLAS) calls of reference model/modules.py:442-449 (to_q/k/v), :494
// (to_out), :342-353 (FeedForward), dit.py:169-177 (InputEmbedding.proj), dit.py:470 (proj_out).
#include "host.h"

namespace f5e {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 192;

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = BN * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;  // barriers + alignment slack
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N,
                 int K, EpiParams ep) {
  using S = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_blocks = (M + GEMM_BM - 1) / GEMM_BM;
  const int n_blocks = (N + BN - 1) / BN;
  const int num_tiles = m_blocks * n_blocks;
  const int nkb = (K + GEMM_BK - 1) / GEMM_BK;
  constexpr uint32_t TMEM_COLS = 2 * BN;  // 256 or 512: power of two

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_holder, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
          tma_load_2d(sa, &tmA, &full_bar[stage], kb * GEMM_BK, m_blk * GEMM_BM);
          tma_load_2d(sb, &tmB, &full_bar[stage], kb * GEMM_BK, n_blk * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(kFmtBF16, GEMM_BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::STAGE_BYTES);
          const uint64_t adesc = umma_desc_kmajor_sw128(sa);
          const uint64_t bdesc = umma_desc_kmajor_sw128(sa + S::A_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)  // +32 bytes (2 x 16B units) per UMMA_K inside the swizzle atom
            umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    const int quarter = warp & 3;  // TMEM lane quarter this warp may read
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / n_blocks, n_blk = tile - m_blk * n_blocks;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int row = m_blk * GEMM_BM + quarter * 32 + lane;
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(taddr + c * 32, raw);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
        epilogue_chunk(v, row, n_blk * BN + c * 32, M, N, ep);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, int STAGES>
static int launch_gemm(cudaStream_t st, const CUtensorMap& tmA, const CUtensorMap& tmB, int M, int N, int K,
                       const EpiParams& ep) {
  using S = GemmSmem<BN, STAGES>;
  const int tiles = ((M + GEMM_BM - 1) / GEMM_BM) * ((N + BN - 1) / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  LaunchScope ls(st, PROF_GEMM, M, N, K);
  gemm_bf16_kernel<BN, STAGES><<<grid, GEMM_THREADS, S::TOTAL, st>>>(tmA, tmB, M, N, K, ep);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int gemm_configure() {
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<256, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<256, 4>::TOTAL));
  F5E_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<128, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      GemmSmem<128, 6>::TOTAL));
  return 0;
}

int gemm_bf16(cudaStream_t st, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int ldw, int M, int N,
              int K, const EpiParams& ep) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_bf16: empty problem (M=%d N=%d K=%d)", M, N, K);
  F5E_REQUIRE(ep.out != nullptr, "gemm_bf16: null output");
  const bool wide = N > 128;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, 2, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, GEMM_BK, GEMM_BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tmB, W, 2, (uint64_t)K, (uint64_t)N, (uint64_t)ldw * 2, GEMM_BK, wide ? 256 : 128);
  if (rc) return rc;
  if (wide) return launch_gemm<256, 4>(st, tmA, tmB, M, N, K, ep);
  return launch_gemm<128, 6>(st, tmA, tmB, M, N, K, ep);
}

}  // namespace f5e
