// Grouped 1-D convolution (kernel 31, 16 groups, "same" zero padding) as an implicit GEMM on tcgen05.
//
// Replaces nn.Conv1d(dim, dim, 31, groups=16, padding=15) + Mish of ConvPositionEmbedding
// (reference model/modules.py:167-190, called UNMASKED from backbones/dit.py:176).
//
//   out[s, n, g*cpg + co] = act( bias + sum_{tap, ci} x[s, n + tap - 15, g*cpg + ci] * w[g*cpg + co, ci, tap] )
//
// Work unit = (sample, PAIR of 128-frame tiles, group).  The activation tile is loaded ONCE with its halo (256 + 30
// frames x 64 channels, 128B-swizzled rows, frames outside [0, n_pad) zero-filled by the 3-D tensor map = the conv's
// padding, and never bleeding into the neighbouring sample); tap t of the kernel is the same tile read `t` rows further
// down, which costs nothing: the UMMA shared-memory descriptor's start address advances by t * 128 bytes.  Only the
// weights stream: one 64 x cpg tap (8 KB) per K step, shared by the two 128-frame accumulators (284 KB of L2 traffic
// per 130 MFLOP; the first version re-fetched the A tile for every tap, 744 KB).
// Measured bound (profiles/r01_conv_pos.md): the 128x64x16 MMAs themselves, ~125 cycles each whatever feeds them
// (removing the W loads or the epilogue changes nothing; removing the MMAs gives 0.09 ms): at N = 64 an UMMA costs as
// much as at N = 256, so this kernel tops out near 430 TFLOP/s with the group width the model fixes.
// Weights are pre-packed to Wp[tap][dim][64] bf16 (input channels >= cpg zero-filled) so that the A tile
// can always be 64 channels wide (cpg = 48 for dim 768 reads 16 channels of the next group against zeros).
//
// Warp-specialised persistent structure as gemm.cu (TMA producer / MMA issuer / 4 epilogue warps, double-buffered
// A tile and TMEM accumulators).
#include "host.h"

namespace f5e {

constexpr int CV_BM = 128;
constexpr int CV_TAPS = 31;
constexpr int CV_PAIR = 2 * CV_BM;              // frames per work unit
constexpr int CV_A_ROWS = CV_PAIR + 32;         // 288 >= 256 + 30, a whole number of 8-row swizzle atoms
constexpr int CV_A_BYTES = CV_A_ROWS * 128;     // 36 KB
constexpr int CV_W_STAGES = 8;
constexpr int CV_W_BYTES = 64 * 64 * 2;         // one tap: up to 64 output channels x 64 input channels
constexpr int CV_SMEM_W = 2 * CV_A_BYTES;
constexpr int CV_BAR_OFFSET = CV_SMEM_W + CV_W_STAGES * CV_W_BYTES;
constexpr int CV_SMEM_TOTAL = CV_BAR_OFFSET + 256 + 1024;
constexpr int CV_THREADS = 192;

__global__ void __launch_bounds__(CV_THREADS, 1)
conv_pos_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmX32,
                const __grid_constant__ CUtensorMap tmW, int samples, int n_pad, int dim, int cpg, EpiParams ep) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + CV_BAR_OFFSET);  // 2
  uint64_t* a_empty = a_full + 2;                                        // 2
  uint64_t* w_full = a_empty + 2;                                        // CV_W_STAGES
  uint64_t* w_empty = w_full + CV_W_STAGES;                              // CV_W_STAGES
  uint64_t* tfull_bar = w_empty + CV_W_STAGES;                           // 2
  uint64_t* tempty_bar = tfull_bar + 2;                                  // 2
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int groups = dim / cpg;
  const int mpairs = (n_pad + CV_PAIR - 1) / CV_PAIR;
  const int num_units = samples * mpairs * groups;
  constexpr uint32_t TMEM_COLS = 256;  // 2 accumulator stages x 2 frame tiles x 64 columns
  const uint32_t w_bytes = (uint32_t)cpg * 128;  // bytes the W box really transfers

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmX32);
    tma_prefetch_desc(&tmW);
    for (int i = 0; i < 2; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < CV_W_STAGES; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_holder, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  pdl_wait();  // programmatic dependent launch: the prologue above touches no global memory
  pdl_trigger();

  if (warp == 0) {
    if (elect_one()) {  // see gemm.cu: elect.sync lets the compiler issue UTMALDG / UTCHMMA without a serialisation loop
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x, ++it) {
        const int g = u % groups, sm = u / groups, mp = sm % mpairs, s = sm / mpairs;
        const int buf = it & 1;
        mbar_wait(&a_empty[buf], ((it >> 1) & 1) ^ 1);
        uint8_t* sa = smem + buf * CV_A_BYTES;
        // halo tile: frames [mp * 256 - 15, mp * 256 + 273) in two boxes (256 + 32 rows; a TMA box holds at most 256)
        mbar_expect_tx(&a_full[buf], CV_A_BYTES);
        tma_load_3d(sa, &tmX, &a_full[buf], g * cpg, mp * CV_PAIR - CV_TAPS / 2, s);
        tma_load_3d(sa + 256 * 128, &tmX32, &a_full[buf], g * cpg, mp * CV_PAIR - CV_TAPS / 2 + 256, s);
        for (int tap = 0; tap < CV_TAPS; ++tap) {
          mbar_wait(&w_empty[stage], phase ^ 1);
          mbar_expect_tx(&w_full[stage], w_bytes);
          tma_load_2d(smem + CV_SMEM_W + stage * CV_W_BYTES, &tmW, &w_full[stage], 0, tap * dim + g * cpg);
          if (++stage == CV_W_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(kFmtBF16, CV_BM, (uint32_t)cpg, 0, 0);
      const int ksteps = (cpg + 15) / 16;
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int u = blockIdx.x; u < num_units; u += gridDim.x, ++it) {
        const int mp = (u / groups) % mpairs;
        const int ntile = (mp * CV_PAIR + CV_BM < n_pad) ? 2 : 1;  // the second frame tile may lie past the sample end
        const int buf = it & 1;
        mbar_wait(&tempty_bar[buf], ((it >> 1) & 1) ^ 1);
        mbar_wait(&a_full[buf], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + buf * CV_A_BYTES);
        for (int tap = 0; tap < CV_TAPS; ++tap) {
          mbar_wait(&w_full[stage], phase);
          tc_fence_after();
          const uint64_t bdesc = umma_desc_kmajor_sw128(smem_u32(smem + CV_SMEM_W + stage * CV_W_BYTES));
          for (int t = 0; t < ntile; ++t) {
            // rows [t * 128 + tap, +128) of the halo tile: the start address is `tap` rows off the 1024-byte atom
            // boundary; the 128B swizzle is a function of the absolute shared-memory address, so the descriptor's
            // base-offset field stays 0 (setting it to tap & 7 gives wrong results - measured)
            const uint64_t adesc = umma_desc_kmajor_sw128(sa + (uint32_t)(t * CV_BM + tap) * 128);
            const uint32_t d_tmem = tmem_base + buf * 128 + t * 64;
            for (int k = 0; k < ksteps; ++k) umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (tap | k) != 0);
          }
          umma_commit(&w_empty[stage]);
          if (++stage == CV_W_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&a_empty[buf]);
        umma_commit(&tfull_bar[buf]);
      }
    }
  } else {
    const int quarter = warp & 3;
    int it = 0;
    for (int u = blockIdx.x; u < num_units; u += gridDim.x, ++it) {
      const int g = u % groups, sm = u / groups, mp = sm % mpairs, s = sm / mpairs;
      const int ntile = (mp * CV_PAIR + CV_BM < n_pad) ? 2 : 1;
      const int buf = it & 1;
      mbar_wait(&tfull_bar[buf], (it >> 1) & 1);
      tc_fence_after();
      for (int t = 0; t < ntile; ++t) {
        const int frame = mp * CV_PAIR + t * CV_BM + quarter * 32 + lane;
        const int row = s * n_pad + frame;
        const int row_limit = frame < n_pad ? s * n_pad + n_pad : 0;  // frames past the sample end: skip
        const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + buf * 128 + t * 64;
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);
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
  CUtensorMap tmX, tmX32, tmW;
  int rc = make_tmap_3d(&tmX, x, 2, (uint64_t)dim, (uint64_t)n_pad, (uint64_t)samples, (uint64_t)dim * 2,
                        (uint64_t)n_pad * dim * 2, 64, 256, 1);
  if (rc) return rc;
  rc = make_tmap_3d(&tmX32, x, 2, (uint64_t)dim, (uint64_t)n_pad, (uint64_t)samples, (uint64_t)dim * 2,
                    (uint64_t)n_pad * dim * 2, 64, 32, 1);
  if (rc) return rc;
  rc = make_tmap_2d(&tmW, wp, 2, 64, (uint64_t)CV_TAPS * dim, 128, 64, (uint32_t)cpg);
  if (rc) return rc;
  const int tiles = samples * ((n_pad + CV_PAIR - 1) / CV_PAIR) * groups;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  LaunchScope ls(st, PROF_CONV, samples, n_pad, dim);
  F5E_CUDA_CHECK(launch_pdl(conv_pos_kernel, dim3(grid), dim3(CV_THREADS), CV_SMEM_TOTAL, st, tmX, tmX32, tmW, samples, n_pad, dim,
                            cpg, ep));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
