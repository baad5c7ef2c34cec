// Grouped 1-D convolution (kernel 31, 16 groups, "same" zero padding) as an implicit GEMM on tcgen05.
//
// Replaces nn.Conv1d(dim, dim, 31, groups=16, padding=15) + Mish of ConvPositionEmbedding
// (reference model/modules.py:167-190, called UNMASKED from backbones/dit.py:176).
//
//   out[s, n, g*cpg + co] = act( bias + sum_{tap, ci} x[s, n + tap - 15, g*cpg + ci] * w[g*cpg + co, ci, tap] )
//
// For one (sample, 128-frame tile, group) the accumulator is a 128 x cpg TMEM tile; every tap adds one
// K=64 GEMM step whose A operand is the activation tile shifted by (tap - 15) frames.  The shift is done
// by the TMA coordinates on a 3-D [sample][frame][channel] tensor map, so frames outside [0, n_pad) read
// as zeros (the conv's zero padding) and never bleed into the neighbouring sample.
// Weights are pre-packed to Wp[tap][dim][64] bf16 (input channels >= cpg zero-filled) so that the A tile
// can always be 64 channels wide (cpg = 48 for dim 768 reads 16 channels of the next group against zeros).
//
// Same warp-specialised persistent structure as gemm.cu (TMA producer / MMA issuer / 4 epilogue warps,
// double-buffered TMEM accumulator).
#include "host.h"

namespace f5e {

constexpr int CV_BM = 128;
constexpr int CV_TAPS = 31;
constexpr int CV_STAGES = 6;
constexpr int CV_A_BYTES = 128 * 64 * 2;
constexpr int CV_B_BYTES = 64 * 64 * 2;
constexpr int CV_STAGE_BYTES = CV_A_BYTES + CV_B_BYTES;
constexpr int CV_BAR_OFFSET = CV_STAGES * CV_STAGE_BYTES;
constexpr int CV_SMEM_TOTAL = CV_BAR_OFFSET + 256 + 1024;
constexpr int CV_THREADS = 192;

__global__ void __launch_bounds__(CV_THREADS, 1)
conv_pos_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, int samples,
                int n_pad, int dim, int cpg, EpiParams ep) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + CV_BAR_OFFSET);
  uint64_t* empty_bar = full_bar + CV_STAGES;
  uint64_t* tfull_bar = empty_bar + CV_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = dim / cpg;
  const int mtiles = (n_pad + CV_BM - 1) / CV_BM;
  const int num_tiles = samples * mtiles * groups;
  constexpr uint32_t TMEM_COLS = 128;  // 2 accumulator stages x 64 columns
  const uint32_t b_bytes = (uint32_t)cpg * 128;  // bytes the W box really transfers

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    for (int i = 0; i < CV_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
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
        const int g = tile % groups, sm = tile / groups, mt = sm % mtiles, s = sm / mtiles;
        for (int tap = 0; tap < CV_TAPS; ++tap) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * CV_STAGE_BYTES;
          mbar_expect_tx(&full_bar[stage], CV_A_BYTES + b_bytes);
          tma_load_3d(sa, &tmX, &full_bar[stage], g * cpg, mt * CV_BM + tap - CV_TAPS / 2, s);
          tma_load_2d(sa + CV_A_BYTES, &tmW, &full_bar[stage], 0, tap * dim + g * cpg);
          if (++stage == CV_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(kFmtBF16, CV_BM, (uint32_t)cpg, 0, 0);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 64;
        for (int tap = 0; tap < CV_TAPS; ++tap) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * CV_STAGE_BYTES);
          const uint64_t adesc = umma_desc_kmajor_sw128(sa);
          const uint64_t bdesc = umma_desc_kmajor_sw128(sa + CV_A_BYTES);
          const int ksteps = (cpg + 15) / 16;
          for (int k = 0; k < ksteps; ++k)
            umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (tap | k) != 0);
          umma_commit(&empty_bar[stage]);
          if (++stage == CV_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int g = tile % groups, sm = tile / groups, mt = sm % mtiles, s = sm / mtiles;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int frame = mt * CV_BM + quarter * 32 + lane;
      const int row = s * n_pad + frame;
      const int row_limit = frame < n_pad ? s * n_pad + n_pad : 0;  // frames past the sample end: skip
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * 64;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(taddr + c * 32, raw);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
        epilogue_chunk(v, row, g * cpg + c * 32, row_limit, (g + 1) * cpg, ep);
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

int conv_pos_configure() {
  F5E_CUDA_CHECK(cudaFuncSetAttribute(conv_pos_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CV_SMEM_TOTAL));
  return 0;
}

// x: bf16 [samples*n_pad, dim] (contiguous); wp: bf16 [31*dim, 64] packed weights.
int conv_pos(cudaStream_t st, const __nv_bfloat16* x, const __nv_bfloat16* wp, int samples, int n_pad, int dim,
             int groups, const EpiParams& ep) {
  { int rc0 = ensure_configured(); if (rc0) return rc0; }
  F5E_REQUIRE(samples > 0 && n_pad > 0, "conv_pos: empty problem");
  F5E_REQUIRE(dim % groups == 0, "conv_pos: dim %d not divisible by groups %d", dim, groups);
  const int cpg = dim / groups;
  F5E_REQUIRE(cpg % 16 == 0 && cpg <= 64, "conv_pos: channels per group %d must be a multiple of 16, at most 64", cpg);
  CUtensorMap tmX, tmW;
  int rc = make_tmap_3d(&tmX, x, 2, (uint64_t)dim, (uint64_t)n_pad, (uint64_t)samples, (uint64_t)dim * 2,
                        (uint64_t)n_pad * dim * 2, 64, 128, 1);
  if (rc) return rc;
  rc = make_tmap_2d(&tmW, wp, 2, 64, (uint64_t)CV_TAPS * dim, 128, 64, (uint32_t)cpg);
  if (rc) return rc;
  const int tiles = samples * ((n_pad + CV_BM - 1) / CV_BM) * groups;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  LaunchScope ls(st, PROF_CONV, samples, n_pad, dim);
  conv_pos_kernel<<<grid, CV_THREADS, CV_SMEM_TOTAL, st>>>(tmX, tmW, samples, n_pad, dim, cpg, ep);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
