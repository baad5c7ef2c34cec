// Micro-benchmark: board power of a B200 whose SMs are awake but do (almost) no work, against SMs saturating one pipe
// (developer tool; run next to `nvidia-smi --query-gpu=power.draw,clocks.sm -lms 100`).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/spin_power tools/ubench/spin_power.cu && /tmp/spin_power
// mode 0: every warp sleeps in __nanosleep; 1: one warp per SM sub-partition spins on clock64; 2: 8 warps per sub-partition
// issue dependent FFMAs; 3: 2 warps per sub-partition issue MUFU.EX2 back to back; 4: 2 warps per sub-partition packed FFMA2.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void k(int mode, long long cycles, float* sink) {
  const long long t0 = clock64();
  float a = threadIdx.x * 1e-3f, b = 1.0001f, c = 0.f, d = 0.5f;
  if (mode == 0) {
    while (clock64() - t0 < cycles) __nanosleep(2000);
  } else if (mode == 1) {
    while (clock64() - t0 < cycles) {}
  } else if (mode == 2) {
    while (clock64() - t0 < cycles) {
#pragma unroll
      for (int i = 0; i < 64; ++i) { a = fmaf(a, b, c); d = fmaf(d, b, a); }
    }
  } else if (mode == 3) {
    while (clock64() - t0 < cycles) {
#pragma unroll
      for (int i = 0; i < 64; ++i) { asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(a) : "f"(a)); asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(d)); }
    }
  } else {
    float2 x = make_float2(a, d), y = make_float2(b, b), z = make_float2(c, c);
    float2 u = make_float2(d, a);
    while (clock64() - t0 < cycles) {
#pragma unroll
      for (int i = 0; i < 64; ++i) { x = __ffma2_rn(x, y, z); u = __ffma2_rn(u, y, x); }
    }
    a = x.x + x.y + u.x + u.y;
  }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = a + d;
}

int main(int argc, char** argv) {
  const double seconds = argc > 1 ? atof(argv[1]) : 2.5;
  float* sink; cudaMalloc(&sink, 4 << 20);
  const int threads[5] = {128, 128, 1024, 256, 256};
  for (int mode = 0; mode < 5; ++mode) {
    printf("MODE %d start\n", mode); fflush(stdout);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    const int launches = (int)(seconds * 1000 / 1.0);  // ~1 ms per launch at 1.9 GHz
    for (int i = 0; i < launches; ++i) k<<<148, threads[mode]>>>(mode, 1900000, sink);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("MODE %d end: %.1f ms for %d launches (%s)\n", mode, ms, launches, cudaGetErrorString(cudaGetLastError())); fflush(stdout);
  }
  return 0;
}
