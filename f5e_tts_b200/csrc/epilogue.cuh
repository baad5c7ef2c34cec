// Fused GEMM epilogue shared by the dense GEMM and the grouped-conv implicit GEMM.
// One thread owns one output row and handles 32 consecutive fp32 accumulator columns at a time
// (that is exactly what tcgen05.ld 32x32b.x32 hands it), so everything row-wise (RoPE pair
// rotation, row masks) is register-local.
#pragma once
#include "common.cuh"

namespace f5e {

enum Act : int { ACT_NONE = 0, ACT_GELU_TANH = 1, ACT_GELU_ERF = 2, ACT_MISH = 3, ACT_RELU = 4, ACT_SILU = 5 };

struct EpiParams {
  void* out;                       // [M, ldo] bf16 (or fp32 when out_f32)
  int ldo;
  int out_f32;
  const float* bias;               // [N] or null
  int act;                         // Act
  const float* gate;               // [N] or null : v *= gate[sample * gate_stride + col]
  int gate_stride;                 // 0 = one gate vector shared by every sample
  const __nv_bfloat16* residual;   // [M, ldr] or null : v += residual
  int ldr;
  const float* ln_rstd;            // [M] or null: folded LayerNorm (gemm_bf16 only): v = v * rstd[row] + bias[col]
  const __nv_bfloat16* ln_a2;      // [M, 64] with ln_rstd: A operand of the LAST K block (K = dim + 64): -mean as (hi, lo, hi)
  const int* row_len;              // [M / n_pad] or null : rows with (row % n_pad) >= row_len[row / n_pad] -> 0
  int n_pad;                       // frames per sample (row -> (sample, frame)); needed by row_len / rope
  int rope_inner;                  // heads*64 (columns [0, 2*rope_inner) hold Q then K)
  int rope_heads;                  // heads [0, rope_heads) of Q and of K get rotated
  const float* rope_inv_freq;      // [32] or null: interleaved-pair RoPE (dim_head 64) on Q and K; the angles
                                   // frame * inv_freq[j] are evaluated in the epilogue (Cody-Waite reduction + sin/cos.approx)
  const int* row_info;             // [M] or null: (sample << 16 | frame) per row, for PACKED rows (valid frames of all
                                   // samples back to back, no padding); replaces the row / n_pad arithmetic
  int res_f32;                     // residual is fp32 [M, ldr] instead of bf16
  float2* stats_out;               // [M, stats_slots] or null (gemm_bf16, CTA-pair kernel, x += gate * branch only): the residual is
  int stats_slots;                 // READ (TMA-staged) instead of reduce-added, and every epilogue thread writes the (mean, M2)
                                   // of its 128 NEW output columns: slot = column / 128.  ln_finalize turns them into rstd / a2
                                   // of the next folded LayerNorm, so row_stats never re-reads the residual stream.
  int split3;                      // > 0: bf16 output written as the 3-term split [hi | hi | lo] at columns
                                   // c, c + split3, c + 2*split3 (A operand of a following fp32-accurate bf16x3 GEMM)
};

// Codebook decision fused into the epilogue of the logit GEMM (gemm.cu, VQ instantiation; vq.cu holds the launcher):
// the weight is laid out with every group padded to 128 rows (pad bias = -1e30), so an epilogue thread (== one row) sees the
// whole group of its column half and takes the arg-max straight out of TMEM; no logits are written.
struct VqEpi {
  long long* targets;   // [M, G]
  void* xq;             // [M, G * vd] bf16 or fp32
  const float* vars;    // [G * V (or V when combine_groups), vd]
  float* sums;          // [G * V] softmax column sums (prob_perplexity), zeroed by the launcher
  int* hist;            // [G * V] code histogram (code_perplexity), zeroed by the launcher
  int2* tie_list;       // [M * G] (row, group) pairs whose top-2 margin is below the tolerance: decided by vq_refine_kernel
  int* tie_count;       // zeroed by the launcher
  float refine_tol;
  int G, V, vd, combine_groups, xq_bf16;
};

// v: 32 accumulator values for (row, col0 .. col0+31).  M/N are the logical output extents.
__device__ __forceinline__ void epilogue_chunk(float (&v)[32], int row, int col0, int M, int N,
                                               const EpiParams& p) {
  if (row >= M || col0 >= N) return;
  const bool full = (col0 + 32 <= N);

  if (p.bias) {
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i));
        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) v[i] += __ldg(p.bias + col0 + i);
    }
  }

  switch (p.act) {
    case ACT_GELU_TANH:
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = gelu_tanh(v[i]);
      break;
    case ACT_GELU_ERF:
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = gelu_erf(v[i]);
      break;
    case ACT_MISH:
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = mish(v[i]);
      break;
    case ACT_RELU:
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
      break;
    case ACT_SILU:
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = silu(v[i]);
      break;
    default:
      break;
  }

  if (p.row_len) {
    const int s = row / p.n_pad, n = row - s * p.n_pad;
    if (n >= __ldg(p.row_len + s)) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 0.0f;
    }
  }

  if (p.gate) {
    const float* gate = p.gate + (p.gate_stride ? (size_t)(row / p.n_pad) * p.gate_stride : 0);
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 g = __ldg(reinterpret_cast<const float4*>(gate + col0 + i));
        v[i] *= g.x; v[i + 1] *= g.y; v[i + 2] *= g.z; v[i + 3] *= g.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) v[i] *= __ldg(gate + col0 + i);
    }
  }

  if (p.residual && p.res_f32) {
    const float* r = reinterpret_cast<const float*>(p.residual) + (size_t)row * p.ldr + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 u = *reinterpret_cast<const float4*>(r + i);
        v[i] += u.x; v[i + 1] += u.y; v[i + 2] += u.z; v[i + 3] += u.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) v[i] += r[i];
    }
  } else if (p.residual) {
    const __nv_bfloat16* r = p.residual + (size_t)row * p.ldr + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 u = *reinterpret_cast<const uint4*>(r + i);
        float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
               d = unpack_bf16x2(u.w);
        v[i] += a.x; v[i + 1] += a.y; v[i + 2] += b.x; v[i + 3] += b.y;
        v[i + 4] += c.x; v[i + 5] += c.y; v[i + 6] += d.x; v[i + 7] += d.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) v[i] += __bfloat162float(r[i]);
    }
  }

  if (p.split3 > 0) {  // requires full chunks (N % 32 == 0, checked by the launcher)
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + col0;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 hi, lo;
      uint32_t* h = &hi.x;
      uint32_t* l = &lo.x;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float a = v[i + 2 * q], b = v[i + 2 * q + 1];
        const __nv_bfloat16 ah = __float2bfloat16(a), bh = __float2bfloat16(b);
        h[q] = pack_bf16x2(__bfloat162float(ah), __bfloat162float(bh));
        l[q] = pack_bf16x2(a - __bfloat162float(ah), b - __bfloat162float(bh));
      }
      *reinterpret_cast<uint4*>(o + i) = hi;
      *reinterpret_cast<uint4*>(o + p.split3 + i) = hi;
      *reinterpret_cast<uint4*>(o + 2 * p.split3 + i) = lo;
    }
  } else if (p.out_f32) {
    float* o = reinterpret_cast<float*>(p.out) + (size_t)row * p.ldo + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) o[i] = v[i];
    }
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 u;
        u.x = pack_bf16x2(v[i], v[i + 1]);
        u.y = pack_bf16x2(v[i + 2], v[i + 3]);
        u.z = pack_bf16x2(v[i + 4], v[i + 5]);
        u.w = pack_bf16x2(v[i + 6], v[i + 7]);
        *reinterpret_cast<uint4*>(o + i) = u;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (col0 + i < N) o[i] = __float2bfloat16(v[i]);
    }
  }
}

}  // namespace f5e
