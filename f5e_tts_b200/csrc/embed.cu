// Once-per-solve conditioning kernels: text embedding gather, depthwise conv k7, GRN, column packing.
// These run twice per ODE solve (cond / uncond), not per NFE step; they are plain coalesced, vectorised
// bandwidth kernels.
#include "host.h"

namespace f5e {

// ---------------------------------------------------------------------------------------------
// pack_cols: dst[r, col0 + c] = (bf16) src[r, c]   (src fp32 or bf16, or zeros when src == null)
// Builds the [x | cond | text | ppg] concat buffer of InputEmbedding (reference dit.py:169-175).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
pack_cols_kernel(__nv_bfloat16* __restrict__ dst, int ld_dst, int col0, const T* __restrict__ src, int ld_src,
                 int rows, int cols) {
  const int pairs = cols >> 1;  // cols even
  const long long total = (long long)rows * pairs;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx / pairs), c = (int)(idx - (long long)r * pairs) * 2;
    float a = 0.f, b = 0.f;
    if (src) {
      a = (float)src[(size_t)r * ld_src + c];
      b = (float)src[(size_t)r * ld_src + c + 1];
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)r * ld_dst + col0 + c) = pack_bf16x2(a, b);
  }
}

int pack_cols(cudaStream_t st, __nv_bfloat16* dst, int ld_dst, int col0, const void* src, int src_is_f32, int ld_src,
              int rows, int cols) {
  F5E_REQUIRE(rows > 0 && cols > 0 && cols % 2 == 0 && col0 % 2 == 0 && ld_dst % 2 == 0,
              "pack_cols: cols/col0/ld_dst must be even (cols=%d col0=%d ld=%d)", cols, col0, ld_dst);
  const long long total = (long long)rows * (cols / 2);
  int grid = (int)((total + 255) / 256);
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  LaunchScope ls(st, PROF_ELEM, rows, cols, 1);
  if (src_is_f32)
    pack_cols_kernel<float><<<grid, 256, 0, st>>>(dst, ld_dst, col0, (const float*)src, ld_src, rows, cols);
  else
    pack_cols_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(dst, ld_dst, col0, (const __nv_bfloat16*)src, ld_src, rows, cols);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// text_gather: out[s, n, :] = table[id(s,n)] + pos[min(n, max_pos-1)], zero where masked
//   id(s,n) = drop ? 0 : (n < nt ? text[s,n] + 1 : 0);  masked iff mask_padding && (n >= nt || text[s,n]+1 == 0)
// Reference backbones/dit.py:54-80 (the "+1 / truncate / pad with 0 / filler mask" logic) and
// model/modules.py:196-219 (position table: pos[p] = [cos(p*f_j) | sin(p*f_j)]).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
text_gather_kernel(const int64_t* __restrict__ text, int nt, const __nv_bfloat16* __restrict__ table,
                   const float* __restrict__ pos, int max_pos, __nv_bfloat16* __restrict__ out, int n_pad, int dim,
                   int drop, int mask_padding, int add_pos) {
  const int n = blockIdx.x, s = blockIdx.y;
  long long id = 0;
  if (n < nt) id = text[(size_t)s * nt + n] + 1;
  const bool masked = mask_padding && id == 0;
  if (drop) id = 0;
  const __nv_bfloat16* trow = table + (size_t)id * dim;
  const float* prow = pos + (size_t)min(n, max_pos - 1) * dim;
  __nv_bfloat16* orow = out + ((size_t)s * n_pad + n) * dim;
  for (int c = threadIdx.x * 2; c < dim; c += blockDim.x * 2) {
    float a = 0.f, b = 0.f;
    if (!masked) {
      float2 t = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(trow + c));
      a = t.x; b = t.y;
      if (add_pos) {
        // reference adds in model dtype: bf16(text) + bf16(pos) -> bf16
        a += __bfloat162float(__float2bfloat16(prow[c]));
        b += __bfloat162float(__float2bfloat16(prow[c + 1]));
      }
    }
    *reinterpret_cast<uint32_t*>(orow + c) = pack_bf16x2(a, b);
  }
}

int text_gather(cudaStream_t st, const int64_t* text, int samples, int nt, const __nv_bfloat16* table, const float* pos,
                int max_pos, __nv_bfloat16* out, int n_pad, int dim, int drop, int mask_padding, int add_pos) {
  F5E_REQUIRE(samples > 0 && n_pad > 0 && dim % 2 == 0, "text_gather: bad shape");
  dim3 grid(n_pad, samples);
  LaunchScope ls(st, PROF_ELEM, samples, n_pad, 2);
  text_gather_kernel<<<grid, 128, 0, st>>>(text, nt, table, pos, max_pos, out, n_pad, dim, drop, mask_padding, add_pos);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// dwconv7: depthwise Conv1d(C, C, k=7, padding=3, groups=C) over frames, per sample, + bias.
// x/out: bf16 [samples, n_pad, C]; w: fp32 [C, 7]; b: fp32 [C].   (model/modules.py:250-252, :261-263)
// Algorithmic bytes 2 * S * N * C * 2; a thread owns 2 channels and slides along 16 frames so each
// input element is read ~1.4 times from L1/L2.
// ---------------------------------------------------------------------------------------------
constexpr int DW_FRAMES = 16;
__global__ void __launch_bounds__(256)
dwconv7_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
               __nv_bfloat16* __restrict__ out, int n_pad, int C) {
  const int s = blockIdx.z, n0 = blockIdx.y * DW_FRAMES;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (c >= C) return;
  float w0[7], w1[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) { w0[k] = __ldg(w + (size_t)c * 7 + k); w1[k] = __ldg(w + (size_t)(c + 1) * 7 + k); }
  const float b0 = __ldg(b + c), b1 = __ldg(b + c + 1);
  const __nv_bfloat16* xs = x + (size_t)s * n_pad * C + c;
  float2 win[DW_FRAMES + 6];
#pragma unroll
  for (int i = 0; i < DW_FRAMES + 6; ++i) {
    const int n = n0 + i - 3;
    win[i] = (n >= 0 && n < n_pad) ? unpack_bf16x2(*reinterpret_cast<const uint32_t*>(xs + (size_t)n * C))
                                    : make_float2(0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < DW_FRAMES; ++i) {
    const int n = n0 + i;
    if (n >= n_pad) break;
    float a0 = b0, a1 = b1;
#pragma unroll
    for (int k = 0; k < 7; ++k) { a0 = fmaf(win[i + k].x, w0[k], a0); a1 = fmaf(win[i + k].y, w1[k], a1); }
    *reinterpret_cast<uint32_t*>(out + ((size_t)s * n_pad + n) * C + c) = pack_bf16x2(a0, a1);
  }
}

int dwconv7(cudaStream_t st, const __nv_bfloat16* x, const float* w, const float* b, __nv_bfloat16* out, int samples,
            int n_pad, int C) {
  F5E_REQUIRE(samples > 0 && n_pad > 0 && C % 2 == 0, "dwconv7: bad shape");
  dim3 grid((C / 2 + 255) / 256, (n_pad + DW_FRAMES - 1) / DW_FRAMES, samples);
  LaunchScope ls(st, PROF_ELEM, samples, n_pad, 3);
  dwconv7_kernel<<<grid, 256, 0, st>>>(x, w, b, out, n_pad, C);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// GRN (model/modules.py:225-234):  Gx[s,c] = ||x[s,:,c]||_2 over ALL n_pad frames;
//   Nx = Gx / (mean_c Gx + 1e-6);  out = gamma * (x * Nx) + beta + x
// grn_sumsq writes the sum of squares of every 64-frame slab per (sample, slab, channel); grn_apply adds the slabs in a
// FIXED order (no atomics: the result must not depend on where the allocator put the buffers - float atomics arrive in an
// address-dependent order), recomputes the tiny per-sample mean in shared memory and applies.
// ---------------------------------------------------------------------------------------------
constexpr int GRN_SLAB = 64;
__global__ void __launch_bounds__(256)
grn_sumsq_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sumsq, int n_pad, int C) {
  const int s = blockIdx.z, n0 = blockIdx.y * GRN_SLAB;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (c >= C) return;
  float a0 = 0.f, a1 = 0.f;
  const int n1 = min(n0 + GRN_SLAB, n_pad);
  for (int n = n0; n < n1; ++n) {
    float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + ((size_t)s * n_pad + n) * C + c));
    a0 = fmaf(v.x, v.x, a0);
    a1 = fmaf(v.y, v.y, a1);
  }
  float* dst = sumsq + ((size_t)s * gridDim.y + blockIdx.y) * C + c;
  dst[0] = a0;
  dst[1] = a1;
}

__global__ void __launch_bounds__(256)
grn_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ sumsq, const float* __restrict__ gamma,
                 const float* __restrict__ beta, __nv_bfloat16* __restrict__ out, int n_pad, int C) {
  extern __shared__ float sh[];  // scale[C] then 8 partials
  float* scale = sh;
  float* part = sh + C;
  const int s = blockIdx.z, n0 = blockIdx.y * GRN_SLAB;
  float loc = 0.f;
  const int slabs = gridDim.y;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float tot_c = 0.f;
    for (int k = 0; k < slabs; ++k) tot_c += sumsq[((size_t)s * slabs + k) * C + c];
    const float g = sqrtf(tot_c);
    scale[c] = g;
    loc += g;
  }
  loc = warp_sum(loc);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = loc;
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) tot += part[i];
  const float inv = 1.0f / (tot / C + 1e-6f);
  const int n1 = min(n0 + GRN_SLAB, n_pad);
  for (int c = threadIdx.x * 2; c < C; c += blockDim.x * 2) {
    const float k0 = fmaf(__ldg(gamma + c), scale[c] * inv, 1.0f), k1 = fmaf(__ldg(gamma + c + 1), scale[c + 1] * inv, 1.0f);
    const float b0 = __ldg(beta + c), b1 = __ldg(beta + c + 1);
    for (int n = n0; n < n1; ++n) {
      const size_t off = ((size_t)s * n_pad + n) * C + c;
      float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + off));
      *reinterpret_cast<uint32_t*>(out + off) = pack_bf16x2(fmaf(v.x, k0, b0), fmaf(v.y, k1, b1));
    }
  }
}

// scratch: fp32 [samples, ceil(n_pad / 64), C] slab partials, fully written by grn_sumsq.
int grn(cudaStream_t st, const __nv_bfloat16* x, const float* gamma, const float* beta, __nv_bfloat16* out,
        float* scratch, int samples, int n_pad, int C) {
  F5E_REQUIRE(samples > 0 && n_pad > 0 && C % 2 == 0 && C <= 8192, "grn: bad shape");
  dim3 g1((C / 2 + 255) / 256, (n_pad + GRN_SLAB - 1) / GRN_SLAB, samples);
  {
    LaunchScope ls(st, PROF_ELEM, samples, n_pad, 4);
    grn_sumsq_kernel<<<g1, 256, 0, st>>>(x, scratch, n_pad, C);
  }
  dim3 g2(1, (n_pad + GRN_SLAB - 1) / GRN_SLAB, samples);
  LaunchScope ls(st, PROF_ELEM, samples, n_pad, 5);
  grn_apply_kernel<<<g2, 256, (C + 8) * sizeof(float), st>>>(x, scratch, gamma, beta, out, n_pad, C);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// unfold_frames: out[s, n, tap*C + c] = x[s, n + tap - taps/2, c] (zero outside the sample), so that a dense
// Conv1d(C, C', k=taps, padding="same") becomes ONE GEMM with K = taps*C (PPGEmbedding's k=5 convs,
// backbones/dit.py:124-135; runs twice per solve).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
unfold_frames_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int n_pad, int C, int taps) {
  const int n = blockIdx.x, s = blockIdx.y;
  const int half = taps / 2, vec = C / 8;
  for (int i = threadIdx.x; i < taps * vec; i += blockDim.x) {
    const int tap = i / vec, c8 = i - tap * vec;
    const int src = n + tap - half;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (src >= 0 && src < n_pad) v = *reinterpret_cast<const uint4*>(x + ((size_t)s * n_pad + src) * C + c8 * 8);
    *reinterpret_cast<uint4*>(out + ((size_t)s * n_pad + n) * taps * C + (size_t)tap * C + c8 * 8) = v;
  }
}

int unfold_frames(cudaStream_t st, const __nv_bfloat16* x, __nv_bfloat16* out, int samples, int n_pad, int C, int taps) {
  F5E_REQUIRE(samples > 0 && n_pad > 0 && C % 8 == 0 && taps % 2 == 1, "unfold_frames: bad shape");
  dim3 grid(n_pad, samples);
  LaunchScope ls(st, PROF_ELEM, samples, n_pad, 6);
  unfold_frames_kernel<<<grid, 256, 0, st>>>(x, out, n_pad, C, taps);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// gather_rows: dst[r, :] = src[row_map[r], :]  (bf16, C % 8 == 0).  Drops the padded frames after the (unmasked, hence
// padded-layout) ConvPositionEmbedding so that the 22 DiT blocks only see valid frames.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_rows_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, const int* __restrict__ row_map,
                   int rows, int C) {
  pdl_wait();  // programmatic dependent launch (common.cuh)
  pdl_trigger();
  const int vec = C / 8;
  const long long total = (long long)rows * vec;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / vec), c = (int)(i - (long long)r * vec);
    reinterpret_cast<uint4*>(dst + (size_t)r * C)[c] = reinterpret_cast<const uint4*>(src + (size_t)__ldg(row_map + r) * C)[c];
  }
}

int gather_rows(cudaStream_t st, const __nv_bfloat16* src, __nv_bfloat16* dst, const int* row_map, int rows, int C) {
  F5E_REQUIRE(rows > 0 && C % 8 == 0, "gather_rows: bad shape");
  const long long total = (long long)rows * (C / 8);
  int grid = (int)((total + 255) / 256);
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  LaunchScope ls(st, PROF_ELEM, rows, C, 10);
  F5E_CUDA_CHECK(launch_pdl(gather_rows_kernel, dim3(grid), dim3(256), 0, st, src, dst, row_map, rows, C));
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
