// Bandwidth-bound kernels of the DiT step: modulated LayerNorm and the fused CFG + Euler update.
// All accesses are 16-byte vectorised and coalesced; reductions are warp shuffles (one warp per row).
#include "host.h"

namespace f5e {

// ---------------------------------------------------------------------------------------------
// norm_mod: out = LN(x) * (plus_one + mul) + add      (reference model/modules.py:301-336, :636)
// Persistent warps: a warp walks rows (row = first, first + stride, ...), keeps the modulation vectors of the current
// sample in registers (they are shared by all of its rows) and prefetches the next row while it reduces the current
// one, so per row only the 2*D bytes of x and 2*D bytes of out move.  D = 256 * CH, each lane owns CH chunks of 8.
// Algorithmic bytes: 2 * M * D * 2.
// ---------------------------------------------------------------------------------------------
constexpr int NM_THREADS = 128;
template <int CH>
__global__ void __launch_bounds__(NM_THREADS)
norm_mod_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int M,
                const float* __restrict__ mul, const float* __restrict__ add, int mod_stride, int n_pad,
                float plus_one, float eps, const int* __restrict__ row_info) {
  constexpr int D = CH * 256;
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int wstride = gridDim.x * (NM_THREADS / 32);
  int row = blockIdx.x * (NM_THREADS / 32) + (threadIdx.x >> 5);
  if (row >= M) return;
  float gm[CH][8], ga[CH][8];  // (plus_one + mul), add for this lane's columns
  int cur_sample = -1;
  uint4 nxt[CH];
  {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * D);
#pragma unroll
    for (int c = 0; c < CH; ++c) nxt[c] = xr[c * 32 + lane];
  }
  for (; row < M; row += wstride) {
    float v[CH][8];
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const uint4 u = nxt[c];
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      v[c][0] = a.x; v[c][1] = a.y; v[c][2] = b.x; v[c][3] = b.y;
      v[c][4] = cc.x; v[c][5] = cc.y; v[c][6] = d.x; v[c][7] = d.y;
#pragma unroll
      for (int i = 0; i < 8; ++i) sum += v[c][i];
    }
    if (row + wstride < M) {  // prefetch the next row of this warp
      const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)(row + wstride) * D);
#pragma unroll
      for (int c = 0; c < CH; ++c) nxt[c] = xr[c * 32 + lane];
    }
    const int sample = mod_stride ? (row_info ? (__ldg(row_info + row) >> 16) : row / n_pad) : 0;
    if (sample != cur_sample) {
      cur_sample = sample;
      const float4* mp = reinterpret_cast<const float4*>(mul + (size_t)sample * mod_stride);
      const float4* ap = reinterpret_cast<const float4*>(add + (size_t)sample * mod_stride);
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int e4 = (c * 32 + lane) * 2;
        const float4 m0 = __ldg(mp + e4), m1 = __ldg(mp + e4 + 1), a0 = __ldg(ap + e4), a1 = __ldg(ap + e4 + 1);
        gm[c][0] = plus_one + m0.x; gm[c][1] = plus_one + m0.y; gm[c][2] = plus_one + m0.z; gm[c][3] = plus_one + m0.w;
        gm[c][4] = plus_one + m1.x; gm[c][5] = plus_one + m1.y; gm[c][6] = plus_one + m1.z; gm[c][7] = plus_one + m1.w;
        ga[c][0] = a0.x; ga[c][1] = a0.y; ga[c][2] = a0.z; ga[c][3] = a0.w;
        ga[c][4] = a1.x; ga[c][5] = a1.y; ga[c][6] = a1.z; ga[c][7] = a1.w;
      }
    }
    const float mean = warp_sum(sum) * (1.0f / D);
    float sq = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = v[c][i] - mean;
        sq += d * d;
      }
    const float rstd = rsqrtf(warp_sum(sq) * (1.0f / D) + eps);
    uint4* orow = reinterpret_cast<uint4*>(out + (size_t)row * D);
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      float r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = (v[c][i] - mean) * rstd * gm[c][i] + ga[c][i];
      uint4 u;
      u.x = pack_bf16x2(r[0], r[1]); u.y = pack_bf16x2(r[2], r[3]);
      u.z = pack_bf16x2(r[4], r[5]); u.w = pack_bf16x2(r[6], r[7]);
      orow[c * 32 + lane] = u;
    }
  }
}

int norm_mod(cudaStream_t st, const __nv_bfloat16* x, __nv_bfloat16* out, int M, int D, const float* mul,
             const float* add, int mod_stride, int n_pad, float plus_one, float eps, const int* row_info) {
  F5E_REQUIRE(M > 0, "norm_mod: M must be positive");
  F5E_REQUIRE(D % 256 == 0 && D >= 256 && D <= 2048, "norm_mod: D=%d must be a multiple of 256 in [256,2048]", D);
  if (n_pad <= 0) n_pad = M;
  const int wpb = NM_THREADS / 32;
  int grid = (M + wpb - 1) / wpb;
  const int cap = num_sms() * 3;  // persistent: 3 resident CTAs per SM (140 registers x 128 threads)
  if (grid > cap) grid = cap;
  LaunchScope ls(st, PROF_NORM, M, D, 0);
#define F5E_NM(CH)                                                                                                    \
  case CH:                                                                                                            \
    F5E_CUDA_CHECK(launch_pdl(norm_mod_kernel<CH>, dim3(grid), dim3(NM_THREADS), 0, st, x, out, M, mul, add, mod_stride,   \
                              n_pad, plus_one, eps, row_info));                                                       \
    break;
  switch (D / 256) {
    F5E_NM(1) F5E_NM(2) F5E_NM(3) F5E_NM(4) F5E_NM(5) F5E_NM(6) F5E_NM(7) F5E_NM(8)
  }
#undef F5E_NM
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// LayerNorm folded into the following GEMM (the Euler loop: every sample shares the time step, so the adaLN
// (1 + scale) is one vector per layer and step):
//   LN(x) * (1 + scale) + shift, then @ W^T + b   ==   rstd_r * (x_r . W'_n  -  mu_r * c_n) + d_n
//   W'_n = bf16(W_n * (1 + scale)),  c_n = sum_k W'_nk,  d_n = b_n + sum_k shift_k W_nk
// so the GEMM reads the residual stream x itself and the normalised copy (2 * M * D bytes written and read back per
// norm) is never materialised: row_stats reads x once (M * D * 2 bytes), fold_ln_weights rewrites the weights of all
// layers once per step (a few hundred MB).  The mean term runs on the tensor core too: W' carries 64 extra K columns
// (c_hi, c_hi, c_lo, 0...) and the GEMM takes its last K block of A from the [M, 64] matrix row_stats writes,
// (-mu_hi, -mu_lo, -mu_hi, 0...) (bf16 hi / lo splits: 2^-17 relative), one UMMA_K = 16 step deep; the epilogue is then
// a single FMA per element, v * rstd_r + d_n, where the plain GEMM has its bias add.  (Applying mean and c_n in the
// epilogue instead cost two more instructions per element and made the GEMMs 10 % slower: they are epilogue-latency
// sensitive.)  c_n is summed from the ROUNDED W' so the mean term cancels whatever the rounding of W' did.
// ---------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(NM_THREADS)
row_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ rstd_out, __nv_bfloat16* __restrict__ a2, int M,
                 float eps) {
  constexpr int D = CH * 256;
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int wstride = gridDim.x * (NM_THREADS / 32);
  for (int row = blockIdx.x * (NM_THREADS / 32) + (threadIdx.x >> 5); row < M; row += wstride) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * D);
    float v[CH][8];
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const uint4 u = xr[c * 32 + lane];
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      v[c][0] = a.x; v[c][1] = a.y; v[c][2] = b.x; v[c][3] = b.y;
      v[c][4] = cc.x; v[c][5] = cc.y; v[c][6] = d.x; v[c][7] = d.y;
#pragma unroll
      for (int i = 0; i < 8; ++i) sum += v[c][i];
    }
    const float mean = warp_sum(sum) * (1.0f / D);  // same statistics as norm_mod_kernel
    float sq = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = v[c][i] - mean;
        sq += d * d;
      }
    const float rstd = rsqrtf(warp_sum(sq) * (1.0f / D) + eps);
    if (lane == 0) {
      rstd_out[row] = rstd;
      const __nv_bfloat16 hi = __float2bfloat16(-mean);
      const __nv_bfloat16 lo = __float2bfloat16(-mean - __bfloat162float(hi));
      uint2 u;
      u.x = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16);
      u.y = (uint32_t)__bfloat16_as_ushort(hi);
      *reinterpret_cast<uint2*>(a2 + (size_t)row * 64) = u;  // columns 4..63 stay zero (allocated zeroed, never written)
    }
  }
}

int row_stats(cudaStream_t st, const __nv_bfloat16* x, float* rstd, __nv_bfloat16* a2, int M, int D, float eps) {
  F5E_REQUIRE(M > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "row_stats: bad shape M=%d D=%d", M, D);
  const int wpb = NM_THREADS / 32;
  int grid = (M + wpb - 1) / wpb;
  const int cap = num_sms() * 8;
  if (grid > cap) grid = cap;
  LaunchScope ls(st, PROF_NORM, M, D, 1);
#define F5E_RS(CH)                                                                              \
  case CH:                                                                                      \
    F5E_CUDA_CHECK(launch_pdl(row_stats_kernel<CH>, dim3(grid), dim3(NM_THREADS), 0, st, x, rstd, a2, M, eps)); \
    break;
  switch (D / 256) {
    F5E_RS(1) F5E_RS(2) F5E_RS(3) F5E_RS(4) F5E_RS(5) F5E_RS(6) F5E_RS(7) F5E_RS(8)
  }
#undef F5E_RS
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Partial statistics written by the producing GEMM's epilogue (EpiParams::stats_out: (mean, M2) of 128 columns per slot) ->
// what the folded GEMM consumes: rstd and the bf16 split of -mean.  Chan's parallel update over equal-sized groups, in a
// fixed order (deterministic).  One thread per row; 8 slots for dim 1024.
__global__ void __launch_bounds__(256)
ln_finalize_kernel(const float2* __restrict__ stats, int slots, int M, float inv_d, float eps, float* __restrict__ rstd_out,
                   __nv_bfloat16* __restrict__ a2) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= M) return;
  const float2* p = stats + (size_t)row * slots;
  float mean = 0.f;
  for (int i = 0; i < slots; ++i) mean += p[i].x;
  mean /= (float)slots;
  float m2 = 0.f;
  for (int i = 0; i < slots; ++i) {
    const float d = p[i].x - mean;
    m2 += p[i].y + 128.0f * d * d;
  }
  rstd_out[row] = rsqrtf(m2 * inv_d + eps);
  const __nv_bfloat16 hi = __float2bfloat16(-mean);
  const __nv_bfloat16 lo = __float2bfloat16(-mean - __bfloat162float(hi));
  uint2 u;
  u.x = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16);
  u.y = (uint32_t)__bfloat16_as_ushort(hi);
  *reinterpret_cast<uint2*>(a2 + (size_t)row * 64) = u;
}

int ln_finalize(cudaStream_t st, const float2* stats, int slots, int M, int D, float eps, float* rstd, __nv_bfloat16* a2) {
  F5E_REQUIRE(stats && rstd && a2 && M > 0 && slots > 0 && slots * 128 == D, "ln_finalize: bad arguments (slots=%d D=%d)", slots, D);
  LaunchScope ls(st, PROF_NORM, M, D, 3);
  F5E_CUDA_CHECK(launch_pdl(ln_finalize_kernel, dim3((M + 255) / 256), dim3(256), 0, st, stats, slots, M, 1.0f / (float)D, eps,
                            rstd, a2));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// qk_norm = "rms_norm" (reference model/modules.py:391-398, 463-467; RMSNorm.forward :282-294): RMSNorm over dim_head with a
// learned weight on every head of Q and K, BEFORE the rotary embedding.  No shipped config turns it on, so it is a small
// in-place pass over the Q and K column blocks of the QKV GEMM's output (the GEMM then runs without its RoPE epilogue and
// the rotation happens here): one warp per (row, q|k, head), a lane holds one interleaved RoPE pair.
__global__ void __launch_bounds__(256)
qk_norm_rope_kernel(__nv_bfloat16* __restrict__ qkv, int ld, int rows, int inner, int heads, const float* __restrict__ qw,
                    const float* __restrict__ kw, float eps, const float* __restrict__ inv_freq, int rope_heads, int n_pad,
                    const int* __restrict__ row_info) {
  const int lane = threadIdx.x & 31;
  const long long total = (long long)rows * 2 * heads;
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  for (long long it = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); it < total; it += (long long)gridDim.x * 8) {
    const int head = (int)(it % heads), which = (int)((it / heads) & 1), row = (int)(it / (2 * heads));
    __nv_bfloat16* p = qkv + (size_t)row * ld + which * inner + head * 64 + 2 * lane;
    float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(p));
    const float ms = warp_sum(v.x * v.x + v.y * v.y) * (1.0f / 64.0f);
    const float r = rsqrtf(ms + eps);
    const float* w = which ? kw : qw;
    v.x = v.x * r * __ldg(w + 2 * lane);
    v.y = v.y * r * __ldg(w + 2 * lane + 1);
    if (inv_freq && head < rope_heads) {  // same angle evaluation as the GEMM epilogue (gemm.cu)
      const int frame = row_info ? (__ldg(row_info + row) & 0xffff) : row % n_pad;
      const float x = (float)frame * __ldg(inv_freq + lane);
      const float k = rintf(x * 0.15915494309189535f);
      float a = fmaf(k, -6.2831854820251465f, x);
      a = fmaf(k, 1.7484555e-7f, a);
      float sn, cs;
      asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(a));
      asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(a));
      const float x0 = v.x, x1 = v.y;
      v.x = x0 * cs - x1 * sn;
      v.y = x1 * cs + x0 * sn;
    }
    *reinterpret_cast<uint32_t*>(p) = pack_bf16x2(v.x, v.y);
  }
}

int qk_norm_rope(cudaStream_t st, __nv_bfloat16* qkv, int ld, int rows, int inner, int heads, const float* qw, const float* kw,
                 float eps, const float* inv_freq, int rope_heads, int n_pad, const int* row_info) {
  F5E_REQUIRE(qkv && qw && kw && rows > 0 && inner == heads * 64 && ld % 2 == 0 && n_pad > 0, "qk_norm_rope: bad arguments");
  const long long total = (long long)rows * 2 * heads;
  long long grid = (total + 7) / 8;
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  LaunchScope ls(st, PROF_NORM, rows, inner, 4);
  F5E_CUDA_CHECK(launch_pdl(qk_norm_rope_kernel, dim3((int)grid), dim3(256), 0, st, qkv, ld, rows, inner, heads, qw, kw, eps,
                            inv_freq, rope_heads, n_pad, row_info));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// one warp per weight row; `jobs` (device) lists the matrices of every layer: rows [row0, row0 + N) of the launch
__global__ void __launch_bounds__(256)
fold_ln_weights_kernel(const FoldJob* __restrict__ jobs, int n_jobs, int total_rows, int D, const float* __restrict__ mod) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int grow = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (grow >= total_rows) return;
  int j = 0;
  while (j + 1 < n_jobs && grow >= jobs[j + 1].row0) ++j;  // a few dozen jobs: linear scan
  const FoldJob job = jobs[j];
  const int n = grow - job.row0;
  const float* scale = mod + job.scale_off;
  const float* shift = mod + job.shift_off;
  const __nv_bfloat16* w = job.w + (size_t)n * D;
  __nv_bfloat16* wf = job.wf + (size_t)n * (D + 64);  // 64 extra K columns hold the column sum (hi, hi, lo, 0...)
  float c = 0.f, d = 0.f;
  for (int k0 = lane * 8; k0 < D; k0 += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(w + k0);
    const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + k0)), s1 = __ldg(reinterpret_cast<const float4*>(scale + k0 + 4));
    const float4 h0 = __ldg(reinterpret_cast<const float4*>(shift + k0)), h1 = __ldg(reinterpret_cast<const float4*>(shift + k0 + 4));
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
    const float wv[8] = {a.x, a.y, b.x, b.y, cc.x, cc.y, e.x, e.y};
    const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const float hv[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
      const uint32_t pk = pack_bf16x2(wv[i] * (1.0f + sv[i]), wv[i + 1] * (1.0f + sv[i + 1]));
      o[i >> 1] = pk;
      const float2 r = unpack_bf16x2(pk);  // the rounded values are what the GEMM multiplies
      c += r.x + r.y;
      d = fmaf(hv[i], wv[i], d);
      d = fmaf(hv[i + 1], wv[i + 1], d);
    }
    *reinterpret_cast<uint4*>(wf + k0) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  c = warp_sum(c);
  d = warp_sum(d);
  if (lane == 0) {
    job.c[n] = c;
    job.d[n] = d + (job.bias ? job.bias[n] : 0.f);
    const __nv_bfloat16 hi = __float2bfloat16(c);
    const __nv_bfloat16 lo = __float2bfloat16(c - __bfloat162float(hi));
    uint2 u;
    u.x = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
    u.y = (uint32_t)__bfloat16_as_ushort(lo);
    *reinterpret_cast<uint2*>(wf + D) = u;  // pairs with (-mu_hi, -mu_lo, -mu_hi): -(mu_hi c_hi + mu_lo c_hi + mu_hi c_lo)
  }
}

int fold_ln_weights(cudaStream_t st, const FoldJob* jobs_dev, int n_jobs, int total_rows, int D, const float* mod) {
  F5E_REQUIRE(jobs_dev && n_jobs > 0 && total_rows > 0 && D % 256 == 0, "fold_ln_weights: bad arguments");
  LaunchScope ls(st, PROF_NORM, total_rows, D, 2);
  F5E_CUDA_CHECK(launch_pdl(fold_ln_weights_kernel, dim3((total_rows + 7) / 8), dim3(256), 0, st, jobs_dev, n_jobs, total_rows, D,
                            mod));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// guided_euler: combine the flow predictions of the guidance branches, then one Euler step y += dt * v
//   1 branch : v = p0
//   2 branches (CFM.sample, reference model/cfm.py:431-447):        v = p0 + (p0 - p1) * c0
//   3 branches (CFM.sample_tts / sample_vc, cfm.py:166-185, 298-317): v = c0 * (p0 - p1) + c1 * (p1 - p2) + p2
// (same operation order as the reference expressions, in fp32)
// =====================================================================================================
template <int NB>
__global__ void __launch_bounds__(256)
guided_euler_kernel(const float* __restrict__ pred, int ld_pred, float* __restrict__ y,
                    __nv_bfloat16* __restrict__ x_cat, int ld_cat, float* __restrict__ traj, int rows, int C,
                    float dt, float c0, float c1, int round_bf16, const int* __restrict__ row_map, int rows_state) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_trigger();
  const int quads = C >> 2;  // C % 4 == 0
  const long long total = (long long)rows * quads;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx / quads), q = (int)(idx - (long long)r * quads);
    const float4 p0 = *reinterpret_cast<const float4*>(pred + (size_t)r * ld_pred + q * 4);
    float4 v = p0;
    if (NB == 2) {
      const float4 p1 = *reinterpret_cast<const float4*>(pred + ((size_t)rows + r) * ld_pred + q * 4);
      v.x = p0.x + (p0.x - p1.x) * c0; v.y = p0.y + (p0.y - p1.y) * c0;
      v.z = p0.z + (p0.z - p1.z) * c0; v.w = p0.w + (p0.w - p1.w) * c0;
    } else if (NB == 3) {
      const float4 p1 = *reinterpret_cast<const float4*>(pred + ((size_t)rows + r) * ld_pred + q * 4);
      const float4 p2 = *reinterpret_cast<const float4*>(pred + (2 * (size_t)rows + r) * ld_pred + q * 4);
      v.x = c0 * (p0.x - p1.x) + c1 * (p1.x - p2.x) + p2.x; v.y = c0 * (p0.y - p1.y) + c1 * (p1.y - p2.y) + p2.y;
      v.z = c0 * (p0.z - p1.z) + c1 * (p1.z - p2.z) + p2.z; v.w = c0 * (p0.w - p1.w) + c1 * (p1.w - p2.w) + p2.w;
    }
    // packed mode: pred rows are the valid frames back to back, row_map gives their row in the padded state / concat buffer
    const int sr = row_map ? __ldg(row_map + r) : r;
    float4* yp = reinterpret_cast<float4*>(y + (size_t)sr * C + q * 4);
    float4 yy = *yp;
    yy.x += dt * v.x; yy.y += dt * v.y; yy.z += dt * v.z; yy.w += dt * v.w;
    if (round_bf16) {
      yy.x = __bfloat162float(__float2bfloat16(yy.x)); yy.y = __bfloat162float(__float2bfloat16(yy.y));
      yy.z = __bfloat162float(__float2bfloat16(yy.z)); yy.w = __bfloat162float(__float2bfloat16(yy.w));
    }
    *yp = yy;
    if (traj) *reinterpret_cast<float4*>(traj + (size_t)sr * C + q * 4) = yy;
    if (x_cat) {  // the new state goes into the concat buffer of every branch
      uint2 u;
      u.x = pack_bf16x2(yy.x, yy.y);
      u.y = pack_bf16x2(yy.z, yy.w);
#pragma unroll
      for (int b = 0; b < NB; ++b)
        *reinterpret_cast<uint2*>(x_cat + ((size_t)b * rows_state + sr) * ld_cat + q * 4) = u;
    }
  }
}

int guided_euler(cudaStream_t st, const float* pred, int ld_pred, float* y, __nv_bfloat16* x_cat, int ld_cat,
                 float* traj, int rows, int C, float dt, int branches, float c0, float c1, int round_bf16,
                 const int* row_map, int rows_state) {
  F5E_REQUIRE(rows > 0 && C > 0 && C % 4 == 0, "guided_euler: need rows>0 and C%%4==0 (rows=%d C=%d)", rows, C);
  F5E_REQUIRE(ld_pred % 4 == 0 && (x_cat == nullptr || ld_cat % 4 == 0), "guided_euler: leading dims must be multiples of 4");
  F5E_REQUIRE(branches >= 1 && branches <= 3, "guided_euler: branches=%d (1, 2 or 3)", branches);
  const long long total = (long long)rows * (C / 4);
  int grid = (int)((total + 255) / 256);
  const int cap = num_sms() * 8;
  if (grid > cap) grid = cap;
  LaunchScope ls(st, PROF_ELEM, rows, C, 0);
  const int rs = row_map ? rows_state : rows;
  if (branches == 1)
    F5E_CUDA_CHECK(launch_pdl(guided_euler_kernel<1>, dim3(grid), dim3(256), 0, st, pred, ld_pred, y, x_cat, ld_cat, traj, rows, C, dt,
                              c0, c1, round_bf16, row_map, rs));
  else if (branches == 2)
    F5E_CUDA_CHECK(launch_pdl(guided_euler_kernel<2>, dim3(grid), dim3(256), 0, st, pred, ld_pred, y, x_cat, ld_cat, traj, rows, C, dt,
                              c0, c1, round_bf16, row_map, rs));
  else
    F5E_CUDA_CHECK(launch_pdl(guided_euler_kernel<3>, dim3(grid), dim3(256), 0, st, pred, ld_pred, y, x_cat, ld_cat, traj, rows, C, dt,
                              c0, c1, round_bf16, row_map, rs));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
