// Micro-benchmark: tcgen05.ld throughput / latency and MUFU.EX2 throughput per SM (developer tool).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_bw tools/ubench/tmem_bw.cu && /tmp/tmem_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr) : "memory");
}

// mode 0: LDTM only; mode 1: MUFU only; mode 2: both in the same warp (dependent: exp of loaded data)
__global__ void k(int mode, int iters, long long* cycles, float* sink) {
  __shared__ uint32_t holder;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&holder)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = holder + (uint32_t((warp & 3) * 32) << 16);
  float acc = threadIdx.x * 1e-9f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t v[32];
    if (mode == 0 || mode == 2) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld32(base + ((it * 4 + c) & 3) * 32 + (warp >> 2) * 128, v);
        if (mode == 0) { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); acc += __uint_as_float(v[0] & 0x3f800000u); }
        else {
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int i = 0; i < 32; ++i) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(__uint_as_float(v[i] & 0x3fffffffu))); acc += y; }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 128; ++i) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(acc * 1e-3f + i)); acc += y * 1e-9f; }
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(holder) : "memory");
}

int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 8 * 1024); cudaMalloc(&sink, 4 * 1024 * 1024);
  const int iters = 2000;
  for (int mode = 0; mode < 3; ++mode)
    for (int warps : {1, 4, 8, 16}) {
      k<<<1, warps * 32>>>(mode, iters, cyc, sink);
      cudaDeviceSynchronize();
      long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      double per_iter = (double)h / iters;  // one iteration = 4 x LDTM.x32 (16 KB per warp) and/or 128 MUFU per thread
      printf("mode %d warps %2d: %.1f cyc/iter/warp  -> %.1f B/clk/SM LDTM, %.2f MUFU lanes/clk/SM  (%s)\n", mode, warps, per_iter,
             mode != 1 ? warps * 16384.0 / per_iter : 0.0, mode != 0 ? warps * 32 * 128.0 / per_iter : 0.0, cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
