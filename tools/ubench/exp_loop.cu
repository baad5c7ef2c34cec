// Micro-benchmark of the attention softmax inner loop (128 scores per thread -> ex2 -> row sum -> bf16 pairs), registers
// only: cycles per block for 1, 2 and 4 warps per SM sub-partition and several instruction mixes (developer tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/exp_loop tools/ubench/exp_loop.cu && /tmp/exp_loop
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// MODE 0: packed ffma2 / fadd2 + pack (the kernel's loop)   1: scalar fma / add + pack   2: packed, no bf16 pack
//      3: MUFU only (sum with scalar adds)                   4: packed, pack, but half of the exps as a degree-3 polynomial
template <int MODE>
__global__ void k(const float* __restrict__ in, int iters, long long* cycles, float* sink, float scale, float mref) {
  float s[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) s[i] = in[(threadIdx.x * 128 + i) & 4095];
  float l = 0.f;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const float2 sc2 = make_float2(scale, scale), nm2 = make_float2(-mref, -mref);
    float2 sum2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
    uint32_t pk[64];
#pragma unroll
    for (int i = 0; i < 128; i += 2) {
      float2 a;
      if (MODE == 1 || MODE == 3) {
        a.x = fmaf(s[i], scale, -mref);
        a.y = fmaf(s[i + 1], scale, -mref);
      } else {
        a = __ffma2_rn(make_float2(s[i], s[i + 1]), sc2, nm2);
      }
      float2 e;
      if (MODE == 4 && (i & 2)) {  // 2^x = 2^floor(x) * p(frac): Cody-Waite + degree-3 polynomial on the FMA pipe
        const float fx = floorf(a.x), fy = floorf(a.y);
        const float rx = a.x - fx, ry = a.y - fy;
        float px = fmaf(fmaf(fmaf(0.0555041f, rx, 0.2402265f), rx, 0.6931472f), rx, 1.0f);
        float py = fmaf(fmaf(fmaf(0.0555041f, ry, 0.2402265f), ry, 0.6931472f), ry, 1.0f);
        e.x = __int_as_float(__float_as_int(px) + ((int)fx << 23));
        e.y = __int_as_float(__float_as_int(py) + ((int)fy << 23));
      } else {
        e.x = ex2_approx(a.x);
        e.y = ex2_approx(a.y);
      }
      if (MODE == 1 || MODE == 3) {
        sum2[(i >> 1) & 1].x += e.x;
        sum2[(i >> 1) & 1].y += e.y;
      } else {
        sum2[(i >> 1) & 1] = __fadd2_rn(sum2[(i >> 1) & 1], e);
      }
      if (MODE != 2 && MODE != 3) pk[i >> 1] = pack_bf16x2(e.x, e.y);
      else pk[i >> 1] = __float_as_uint(e.x) ^ __float_as_uint(e.y);
    }
    l += (sum2[0].x + sum2[1].x) + (sum2[0].y + sum2[1].y);
#pragma unroll
    for (int i = 0; i < 64; ++i) acc ^= pk[i];
    mref += 1e-7f * l;  // loop-carried so iterations are not hoisted
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = l + __uint_as_float(acc & 0x3f800000u);
}

template <int MODE>
void run(const float* in, long long* cyc, float* sink) {
  const int iters = 500;
  for (int warps : {4, 8, 16}) {
    k<MODE><<<1, warps * 32>>>(in, iters, cyc, sink, 0.18f, 3.0f);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("mode %d, %d warps/SMSP: %.0f cycles per 128-score block per warp, %.2f exp lanes/clk/SM  (%s)\n", MODE, warps / 4,
           (double)h / iters, warps * 32 * 128.0 * iters / h, cudaGetErrorString(cudaGetLastError()));
  }
}

int main() {
  float* in; long long* cyc; float* sink;
  cudaMalloc(&in, 4096 * 4); cudaMalloc(&cyc, 1024); cudaMalloc(&sink, 4 * 1024 * 1024);
  float h[4096];
  for (int i = 0; i < 4096; ++i) h[i] = (float)((i * 37) % 101) * 0.1f - 5.f;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  run<0>(in, cyc, sink); run<1>(in, cyc, sink); run<2>(in, cyc, sink); run<3>(in, cyc, sink); run<4>(in, cyc, sink);
  return 0;
}
