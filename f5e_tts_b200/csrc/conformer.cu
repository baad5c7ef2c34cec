// Streaming kernels of the PPG extractor's Conformer encoder (second half); its Linears run on gemm.cu, its
// LayerNorms on norm_mod, its attention on attention.cu (key-bias form).  Reference: ppg/wenet/transformer/
//   subsample_conv : Conv2dSubsampling2.conv (subsampling.py:81-84, 114-117): Conv2d(1, C, 3, stride 2) + ReLU
//   relpos_prepare : RelPositionMultiHeadedAttention.forward (attention.py:179-222) rewritten as
//                    (q+u).k_j + (q+v).p_j = q.(k_j + p_j) + (u.k_j + v.p_j): keys get p_j added, the rest is a per-key bias
//   glu_dwconv     : ConvolutionModule (convolution.py:116-127): GLU, depthwise Conv1d(k), BatchNorm1d(eval, folded), swish
// All HBM-bound: every element read once, written once.
#include "../../include/f5e_b200.h"
#include "host.h"

namespace f5e {

// feats fp32 [B, T, F] (utterance stride ld_feats), n_frames[b] valid frames -> out bf16 [B * t_pad, Fo * C], column f' * C + c,
// Fo = (F - 3) / 2 + 1; rows t' >= (n_frames[b] - 3) / 2 + 1 are zeros.  w [C, 9] (3x3, time-major), bias [C].
__global__ void __launch_bounds__(256)
subsample_conv_kernel(const float* __restrict__ feats, long long ld_feats, const int* __restrict__ n_frames, int T, int F,
                      const float* __restrict__ w, const float* __restrict__ bias, int C, __nv_bfloat16* __restrict__ out,
                      int t_pad, int Fo) {
  extern __shared__ float rows[];  // [3][F]
  const int b = blockIdx.y, tp = blockIdx.x;
  const int nf = n_frames ? min(n_frames[b], T) : T;
  const int t_out = nf >= 3 ? (nf - 3) / 2 + 1 : 0;
  __nv_bfloat16* o = out + ((size_t)b * t_pad + tp) * ((size_t)Fo * C);
  if (tp >= t_out) {
    for (int i = threadIdx.x; i < Fo * C; i += blockDim.x) o[i] = __float2bfloat16(0.f);
    return;
  }
  const float* x = feats + (size_t)b * ld_feats + (size_t)(2 * tp) * F;
  for (int i = threadIdx.x; i < 3 * F; i += blockDim.x) rows[i] = x[i];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float wk[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) wk[i] = __ldg(w + c * 9 + i);
    const float bc = __ldg(bias + c);
    for (int f = 0; f < Fo; ++f) {
      float acc = bc;
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) acc = fmaf(wk[i * 3 + j], rows[i * F + 2 * f + j], acc);
      o[(size_t)f * C + c] = __float2bfloat16(fmaxf(acc, 0.f));
    }
  }
}

// qkv bf16 [samples * n_pad, 3 * heads * 64]: K block (columns [inner, 2 inner)) += pos[t] in place; key_bias[(s * heads + h) *
// ld_bias + t] = sum_d u[h, d] * k[t, h, d] + v[h, d] * pos[t, h, d]   (k as the GEMM wrote it, fp32 accumulation).
// One warp per (row, head); pos fp32 [n_pad, inner] = linear_pos(pos_emb), u / v fp32 [heads, 64].
__global__ void __launch_bounds__(256)
relpos_prepare_kernel(__nv_bfloat16* __restrict__ qkv, int ld_qkv, const float* __restrict__ pos, const float* __restrict__ u,
                      const float* __restrict__ v, float* __restrict__ key_bias, int ld_bias, int rows, int n_pad, int heads) {
  const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (wid >= rows * heads) return;
  const int row = wid / heads, h = wid - row * heads;
  const int s = row / n_pad, t = row - s * n_pad;
  const int inner = heads * 64;
  __nv_bfloat162* kp = reinterpret_cast<__nv_bfloat162*>(qkv + (size_t)row * ld_qkv + inner + h * 64) + lane;
  const float2 k = __bfloat1622float2(*kp);
  const float2 p = *reinterpret_cast<const float2*>(pos + (size_t)t * inner + h * 64 + 2 * lane);
  const float2 uu = *reinterpret_cast<const float2*>(u + h * 64 + 2 * lane);
  const float2 vv = *reinterpret_cast<const float2*>(v + h * 64 + 2 * lane);
  float acc = uu.x * k.x + uu.y * k.y + vv.x * p.x + vv.y * p.y;
  acc = warp_sum(acc);
  *kp = __floats2bfloat162_rn(k.x + p.x, k.y + p.y);
  if (lane == 0) key_bias[((size_t)s * heads + h) * ld_bias + t] = acc;
}

// g bf16 [samples * n_pad, 2 C] (value | gate) -> out bf16 [samples * n_pad, C]:
//   y[t, c] = swish(norm(b[c] + sum_k w[c, k] * glu[t + k - lpad, c])),  glu = value * sigmoid(gate)
// glu reads as zero at t >= len[s] and as pad_val[c] at t < 0 (NULL: zero).  The reference pads the causal left context BEFORE
// its biased pointwise conv (convolution.py:104-113), so that context is glu(bias) - pad_val carries it; the symmetric module
// pads the depthwise conv itself with zeros.  norm: eval BatchNorm1d arrives folded into w / b; LayerNorm over the C channels
// of a frame (ln_w != NULL; needs C <= blockDim.x, one block per frame tile) is applied here.  Rows t >= len[s] are zeros.
constexpr int DW_TT = 16;

__device__ __forceinline__ void block_sum_tt(float (&v)[DW_TT], float* red) {  // red: [8][DW_TT]; all threads take part
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
#pragma unroll
  for (int i = 0; i < DW_TT; ++i) v[i] = warp_sum(v[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < DW_TT; ++i) red[warp * DW_TT + i] = v[i];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < DW_TT; ++i) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += red[w * DW_TT + i];
    v[i] = t;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256)
glu_dwconv_kernel(const __nv_bfloat16* __restrict__ g, const float* __restrict__ w, const float* __restrict__ b,
                  const int* __restrict__ len, __nv_bfloat16* __restrict__ out, int n_pad, int C, int taps, int lpad,
                  const float* __restrict__ pad_val, const float* __restrict__ ln_w, const float* __restrict__ ln_b, float ln_eps) {
  extern __shared__ float tile[];  // [DW_TT + taps - 1][blockDim.x]
  __shared__ float red[8 * DW_TT];
  const int s = blockIdx.y, t0 = blockIdx.x * DW_TT, c0 = blockIdx.z * blockDim.x;
  const int cw = min((int)blockDim.x, C - c0);
  const int n = len ? min(len[s], n_pad) : n_pad;
  const int span = DW_TT + taps - 1;
  const int c = threadIdx.x;
  const bool active = c < cw;
  float acc[DW_TT];
#pragma unroll
  for (int i = 0; i < DW_TT; ++i) acc[i] = 0.f;
  if (active) {
    const float pv = pad_val ? __ldg(pad_val + c0 + c) : 0.f;
    for (int r = 0; r < span; ++r) {
      const int t = t0 + r - lpad;
      float val = t < 0 ? pv : 0.f;
      if (t >= 0 && t < n) {
        const __nv_bfloat16* row = g + ((size_t)s * n_pad + t) * (2 * (size_t)C) + c0 + c;
        const float a = __bfloat162float(row[0]), gt = __bfloat162float(row[C]);
        val = a / (1.f + __expf(-gt));
      }
      tile[r * blockDim.x + c] = val;
    }
    // a thread only reads the column it wrote: no barrier needed
    const float bc = __ldg(b + c0 + c);
#pragma unroll
    for (int i = 0; i < DW_TT; ++i) acc[i] = bc;
    for (int k = 0; k < taps; ++k) {
      const float wk = __ldg(w + (size_t)(c0 + c) * taps + k);
#pragma unroll
      for (int i = 0; i < DW_TT; ++i) acc[i] = fmaf(wk, tile[(i + k) * blockDim.x + c], acc[i]);
    }
  }
  if (ln_w) {  // LayerNorm over the channels of each frame (two-pass, fp32); the launcher guarantees gridDim.z == 1
    float m[DW_TT];
#pragma unroll
    for (int i = 0; i < DW_TT; ++i) m[i] = acc[i];  // inactive threads hold zeros
    block_sum_tt(m, red);
    float q[DW_TT];
#pragma unroll
    for (int i = 0; i < DW_TT; ++i) {
      m[i] /= (float)C;
      const float d = active ? acc[i] - m[i] : 0.f;
      q[i] = d * d;
    }
    block_sum_tt(q, red);
    if (active) {
      const float lw = __ldg(ln_w + c), lb = __ldg(ln_b + c);
#pragma unroll
      for (int i = 0; i < DW_TT; ++i) acc[i] = (acc[i] - m[i]) * rsqrtf(q[i] / (float)C + ln_eps) * lw + lb;
    }
  }
  if (active) {
#pragma unroll
    for (int i = 0; i < DW_TT; ++i) {
      const int t = t0 + i;
      if (t < n_pad) {
        const float y = acc[i] / (1.f + __expf(-acc[i]));
        out[((size_t)s * n_pad + t) * C + c0 + c] = __float2bfloat16(t < n ? y : 0.f);
      }
    }
  }
}

}  // namespace f5e

using namespace f5e;

extern "C" {

int f5e_subsample_conv(const float* feats, int64_t ld_feats, const int32_t* n_frames, int32_t batch, int32_t T, int32_t F,
                       const float* w, const float* bias, int32_t C, void* out, int32_t t_pad, void* stream) {
  F5E_REQUIRE(feats && w && bias && out && batch > 0 && T >= 3 && F >= 3 && C > 0 && t_pad > 0, "f5e_subsample_conv: bad arguments");
  const int Fo = (F - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, PROF_CONV, batch, T, C);
  subsample_conv_kernel<<<dim3(t_pad, batch), 256, 3 * F * sizeof(float), st>>>(feats, ld_feats, n_frames, T, F, w, bias, C,
                                                                               (__nv_bfloat16*)out, t_pad, Fo);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int f5e_relpos_prepare(void* qkv, int32_t ld_qkv, const float* pos, const float* u, const float* v, float* key_bias,
                       int32_t ld_bias, int32_t samples, int32_t n_pad, int32_t heads, void* stream) {
  F5E_REQUIRE(qkv && pos && u && v && key_bias && samples > 0 && n_pad > 0 && heads > 0, "f5e_relpos_prepare: bad arguments");
  F5E_REQUIRE(ld_qkv % 2 == 0 && ld_bias >= n_pad, "f5e_relpos_prepare: bad leading dimensions");
  const int rows = samples * n_pad;
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, PROF_ELEM, rows, heads, 0);
  relpos_prepare_kernel<<<(rows * heads + 7) / 8, 256, 0, st>>>((__nv_bfloat16*)qkv, ld_qkv, pos, u, v, key_bias, ld_bias,
                                                                rows, n_pad, heads);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int f5e_glu_dwconv(const void* g, const float* w, const float* b, const int32_t* len, void* out, int32_t samples,
                   int32_t n_pad, int32_t C, int32_t taps, int32_t left_pad, const float* pad_val, const float* ln_w,
                   const float* ln_b, float ln_eps, void* stream) {
  F5E_REQUIRE(g && w && b && out && samples > 0 && n_pad > 0 && C > 0, "f5e_glu_dwconv: bad arguments");
  F5E_REQUIRE(taps >= 1 && taps <= 64 && left_pad >= 0 && left_pad < taps, "f5e_glu_dwconv: taps must be in [1, 64]");
  F5E_REQUIRE((ln_w == nullptr) == (ln_b == nullptr), "f5e_glu_dwconv: LayerNorm weight and bias go together");
  F5E_REQUIRE(ln_w == nullptr || C <= 256, "f5e_glu_dwconv: the LayerNorm variant needs C <= 256 (one block per frame tile), C=%d", C);
  const int threads = C < 256 ? ((C + 31) / 32) * 32 : 256;
  const size_t smem = (size_t)(DW_TT + taps - 1) * threads * sizeof(float);
  if (smem > 48 * 1024)
    F5E_CUDA_CHECK(cudaFuncSetAttribute(glu_dwconv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaStream_t st = (cudaStream_t)stream;
  LaunchScope ls(st, PROF_CONV, samples, n_pad, C);
  glu_dwconv_kernel<<<dim3((n_pad + DW_TT - 1) / DW_TT, samples, (C + threads - 1) / threads), threads, smem, st>>>(
      (const __nv_bfloat16*)g, w, b, len, (__nv_bfloat16*)out, n_pad, C, taps, left_pad, pad_val, ln_w, ln_b, ln_eps);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
