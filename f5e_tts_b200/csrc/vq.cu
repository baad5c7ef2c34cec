// Codebook selection of GumbelVectorQuantizer.forward in eval mode (reference model/modules.py:881-950):
//   logits = weight_proj(x)  [M, groups*num_vars]      -> computed by the tcgen05 GEMM (gemm.cu) as an fp32-accurate
//                                                          3-term bf16 split ("bf16x3"), fp32 accumulation
//   k_g    = argmax_j logits[g*V + j]  (first maximum)  -> vq_select_kernel below
//   x_q    = concat_g vars[g*V + k_g]                   (the reference's one-hot * vars sum is exactly this gather)
//   code_perplexity / prob_perplexity                   -> histogram + softmax column sums, vq_perplexity_kernel
//
// Bit-exact indices: the split GEMM is accurate to ~1e-6 relative; whenever the top-2 margin of a group is below
// `refine_tol * max(1, |best|)` the group's logits are recomputed for that row with plain sequential fp32 FMAs (the same
// arithmetic a scalar fp32 implementation performs) before the decision is taken, so a tensor-core rounding can never
// flip an index.
#include "host.h"

namespace f5e {

__device__ __forceinline__ void argmax_pair(float& v, int& i, float ov, int oi) {
  if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

// one warp per row; 8 rows per CTA.  sums_g / hist_g: [G*V] global accumulators (zeroed by the launcher).
__global__ void __launch_bounds__(256)
vq_select_kernel(const float* __restrict__ logits, int ld_logits, const float* __restrict__ x, int ld_x, int dim,
                 const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ vars, int M, int G,
                 int V, int vd, int combine_groups, float refine_tol, long long* __restrict__ targets, void* __restrict__ xq,
                 int xq_bf16, float* __restrict__ sums_g, int* __restrict__ hist_g, int* __restrict__ refined) {
  extern __shared__ float vq_sums[];  // [G*V]
  for (int i = threadIdx.x; i < G * V; i += blockDim.x) vq_sums[i] = 0.f;
  __syncthreads();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row < M) {
    for (int g = 0; g < G; ++g) {
      const float* lr = logits + (size_t)row * ld_logits + g * V;
      float l[8];  // V <= 256
      const int nj = (V + 31) / 32;
      float best = -INFINITY;
      int bi = 0x7fffffff;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int j = q * 32 + lane;
        l[q] = (q < nj && j < V) ? lr[j] : -INFINITY;
        if (l[q] > best) { best = l[q]; bi = j; }
      }
      // warp arg-max (first index on ties) and runner-up value
      float wb = best;
      int wi = bi;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) argmax_pair(wb, wi, __shfl_xor_sync(0xffffffffu, wb, o), __shfl_xor_sync(0xffffffffu, wi, o));
      float second = -INFINITY;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (q * 32 + lane != wi) second = fmaxf(second, l[q]);
      second = warp_max(second);
      if (V > 1 && wb - second <= refine_tol * fmaxf(1.0f, fabsf(wb))) {
        // near tie: exact sequential fp32 recomputation of this group's logits for this row
        const float* xr = x + (size_t)row * ld_x;
        best = -INFINITY;
        bi = 0x7fffffff;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int j = q * 32 + lane;
          if (q < nj && j < V) {
            const float* wr = W + (size_t)(g * V + j) * dim;
            float acc = 0.f;
            for (int k = 0; k < dim; ++k) acc = fmaf(xr[k], wr[k], acc);
            acc += bias ? bias[g * V + j] : 0.f;
            l[q] = acc;
            if (acc > best) { best = acc; bi = j; }
          }
        }
        wb = best;
        wi = bi;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) argmax_pair(wb, wi, __shfl_xor_sync(0xffffffffu, wb, o), __shfl_xor_sync(0xffffffffu, wi, o));
        if (lane == 0 && refined) atomicAdd(refined, 1);
      }
      if (lane == 0) {
        targets[(size_t)row * G + g] = wi;
        atomicAdd(hist_g + g * V + wi, 1);
      }
      // softmax over the group (fp32), accumulated per column for prob_perplexity
      float se = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        l[q] = (q * 32 + lane < V && q < nj) ? expf(l[q] - wb) : 0.f;
        se += l[q];
      }
      const float inv = 1.0f / warp_sum(se);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (q < nj && q * 32 + lane < V) atomicAdd(vq_sums + g * V + q * 32 + lane, l[q] * inv);
      // gather the selected code vector
      const float* src = vars + (size_t)((combine_groups ? 0 : g * V) + wi) * vd;
      if (xq_bf16) {
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(xq) + (size_t)row * G * vd + (size_t)g * vd;
        for (int c = lane * 2; c < vd; c += 64) *reinterpret_cast<uint32_t*>(dst + c) = pack_bf16x2(src[c], src[c + 1]);
      } else {
        float* dst = reinterpret_cast<float*>(xq) + (size_t)row * G * vd + (size_t)g * vd;
        for (int c = lane * 2; c < vd; c += 64) *reinterpret_cast<float2*>(dst + c) = *reinterpret_cast<const float2*>(src + c);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < G * V; i += blockDim.x)
    if (vq_sums[i] != 0.f) atomicAdd(sums_g + i, vq_sums[i]);
}

// out[0] = code_perplexity = sum_g exp(-sum_j p log(p + 1e-7)), p = hist / M;  out[1] = prob_perplexity (same on the mean
// softmax).  One CTA, thread j strides over the V entries of a group.   (modules.py:902-913)
__global__ void __launch_bounds__(256)
vq_perplexity_kernel(const int* __restrict__ hist, const float* __restrict__ sums, int M, int G, int V, float* __restrict__ out) {
  __shared__ float red[2][8];
  float tot_c = 0.f, tot_p = 0.f;
  for (int g = 0; g < G; ++g) {
    float ec = 0.f, ep = 0.f;
    for (int j = threadIdx.x; j < V; j += blockDim.x) {
      const float pc = (float)hist[g * V + j] / (float)M, pp = sums[g * V + j] / (float)M;
      ec += pc * logf(pc + 1e-7f);
      ep += pp * logf(pp + 1e-7f);
    }
    ec = warp_sum(ec);
    ep = warp_sum(ep);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ec; red[1][threadIdx.x >> 5] = ep; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float a = 0.f, b = 0.f;
      for (int i = 0; i < 8; ++i) { a += red[0][i]; b += red[1][i]; }
      tot_c += expf(-a);
      tot_p += expf(-b);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out[0] = tot_c; out[1] = tot_p; }
}

// softmax over the first V columns of every row of fp32 logits [M, ld], written as the 3-term bf16 split [hi | hi | lo]
// of width kpad (columns >= V zero): the A operand of a following fp32-accurate bf16x3 GEMM.  One warp per row.
// (PPGModelWapper.ppg_to_target, output_type "map": logit.softmax(-1) @ global_phn_center, ppg/ppg_model.py:113-118)
__global__ void __launch_bounds__(256)
softmax_split3_kernel(const float* __restrict__ logits, int ld, int M, int V, int kpad, __nv_bfloat16* __restrict__ out) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float* lr = logits + (size_t)row * ld;
  float mx = -INFINITY;
  for (int j = lane; j < V; j += 32) mx = fmaxf(mx, lr[j]);
  mx = warp_max(mx);
  float se = 0.f;
  for (int j = lane; j < V; j += 32) se += expf(lr[j] - mx);
  const float inv = 1.0f / warp_sum(se);
  __nv_bfloat16* orow = out + (size_t)row * 3 * kpad;
  for (int j = lane; j < kpad; j += 32) {
    const float p = j < V ? expf(lr[j] - mx) * inv : 0.f;
    const __nv_bfloat16 h = __float2bfloat16(p);
    orow[j] = h;
    orow[kpad + j] = h;
    orow[2 * kpad + j] = __float2bfloat16(p - __bfloat162float(h));
  }
}

int softmax_split3(cudaStream_t st, const float* logits, int ld, int M, int V, int kpad, __nv_bfloat16* out) {
  F5E_REQUIRE(M > 0 && V > 0 && kpad >= V && kpad % 8 == 0 && ld >= V, "softmax_split3: bad shape (M=%d V=%d kpad=%d ld=%d)", M, V, kpad, ld);
  LaunchScope ls(st, PROF_VQ, M, V, -2);
  softmax_split3_kernel<<<(M + 7) / 8, 256, 0, st>>>(logits, ld, M, V, kpad, out);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Near-tie (row, group) pairs the fused epilogue of gemm_vq left undecided: the group's logits recomputed with plain
// sequential fp32 FMAs (the arithmetic a scalar fp32 implementation performs), then arg-max (first maximum), histogram
// and code-vector gather.  One warp per pair; the pair count is read on the device (no host round trip).
__global__ void __launch_bounds__(256)
vq_refine_kernel(const float* __restrict__ x, int ld_x, int dim, const float* __restrict__ W, const float* __restrict__ bias,
                 VqEpi vq) {
  const int lane = threadIdx.x & 31;
  const int n = *vq.tie_count;
  for (int e = blockIdx.x * 8 + (threadIdx.x >> 5); e < n; e += gridDim.x * 8) {
    const int2 rg = vq.tie_list[e];
    const int row = rg.x, g = rg.y;
    const float* xr = x + (size_t)row * ld_x;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int j = lane; j < vq.V; j += 32) {
      const float* wr = W + (size_t)(g * vq.V + j) * dim;
      float acc = 0.f;
      for (int k = 0; k < dim; ++k) acc = fmaf(xr[k], wr[k], acc);
      acc += bias ? bias[g * vq.V + j] : 0.f;
      if (acc > best) { best = acc; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) argmax_pair(best, bi, __shfl_xor_sync(0xffffffffu, best, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (lane == 0) {
      vq.targets[(size_t)row * vq.G + g] = bi;
      atomicAdd(vq.hist + g * vq.V + bi, 1);
    }
    const float* src = vq.vars + (size_t)((vq.combine_groups ? 0 : g * vq.V) + bi) * vq.vd;
    const size_t o = (size_t)row * vq.G * vq.vd + (size_t)g * vq.vd;
    if (vq.xq_bf16) {
      __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(vq.xq) + o;
      for (int c = lane * 2; c < vq.vd; c += 64) *reinterpret_cast<uint32_t*>(dst + c) = pack_bf16x2(src[c], src[c + 1]);
    } else {
      float* dst = reinterpret_cast<float*>(vq.xq) + o;
      for (int c = lane * 2; c < vq.vd; c += 64) *reinterpret_cast<float2*>(dst + c) = *reinterpret_cast<const float2*>(src + c);
    }
  }
}

int vq_refine(cudaStream_t st, const float* x, int ld_x, int dim, const float* W, const float* bias, const VqEpi& vq,
              float* perplexity, int M) {
  {
    LaunchScope ls(st, PROF_VQ, M, vq.G, -3);
    vq_refine_kernel<<<64, 256, 0, st>>>(x, ld_x, dim, W, bias, vq);
  }
  if (perplexity) {
    LaunchScope ls(st, PROF_VQ, M, vq.G, -1);
    vq_perplexity_kernel<<<1, 256, 0, st>>>(vq.hist, vq.sums, M, vq.G, vq.V, perplexity);
  }
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int vq_select(cudaStream_t st, const float* logits, int ld_logits, const float* x, int ld_x, int dim, const float* W,
              const float* bias, const float* vars, int M, int G, int V, int vd, int combine_groups, float refine_tol,
              long long* targets, void* xq, int xq_bf16, float* sums, int* hist, int* refined, float* perplexity) {
  F5E_REQUIRE(M > 0 && G > 0 && V > 0 && V <= 256 && vd % 2 == 0, "vq_select: bad shape (M=%d G=%d V=%d vd=%d)", M, G, V, vd);
  F5E_CUDA_CHECK(cudaMemsetAsync(sums, 0, (size_t)G * V * sizeof(float), st));
  F5E_CUDA_CHECK(cudaMemsetAsync(hist, 0, (size_t)G * V * sizeof(int), st));
  if (refined) F5E_CUDA_CHECK(cudaMemsetAsync(refined, 0, sizeof(int), st));
  {
    LaunchScope ls(st, PROF_VQ, M, G, V);
    vq_select_kernel<<<(M + 7) / 8, 256, (size_t)G * V * sizeof(float), st>>>(logits, ld_logits, x, ld_x, dim, W, bias, vars, M,
                                                                             G, V, vd, combine_groups, refine_tol, targets, xq,
                                                                             xq_bf16, sums, hist, refined);
  }
  if (perplexity) {
    LaunchScope ls(st, PROF_VQ, M, G, -1);
    vq_perplexity_kernel<<<1, 256, 0, st>>>(hist, sums, M, G, V, perplexity);
  }
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
