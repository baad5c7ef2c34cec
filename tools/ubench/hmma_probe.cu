// Micro-benchmark (developer tool): (1) rate of the legacy warp-level mma.sync.m16n8k16 bf16 path on sm_100a,
// (2) which element each thread gets from tcgen05.ld.16x256b (is it the mma.sync accumulator fragment?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/hmma_probe tools/ubench/hmma_probe.cu && /tmp/hmma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__global__ void hmma_rate(int iters, long long* cycles, float* sink) {
  float acc[8][4] = {};
  uint32_t a[4] = {0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u}, b[2] = {0x3f803f80u + threadIdx.x, 0x3f803f80u};
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < 8; ++j) mma16816(acc[j], a, b);
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int j = 0; j < 8; ++j) s += acc[j][0] + acc[j][3];
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void ld_layout(float* out) {  // one warp
  __shared__ uint32_t holder;
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"((uint32_t)__cvta_generic_to_shared(&holder)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncwarp();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = holder;
  const int lane = threadIdx.x;
  // value(row, col) = row * 100 + col written with the thread == row layout (32x32b), 16 columns
  for (int c = 0; c < 16; ++c) {
    uint32_t v = __float_as_uint((float)(lane * 100 + c));
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(base + c), "r"(v) : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(base) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 8; ++i) out[lane * 8 + i] = __uint_as_float(r[i]);
  uint32_t q[8];
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7])
               : "r"(base + (16u << 16)) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int i = 0; i < 8; ++i) out[256 + lane * 8 + i] = __uint_as_float(q[i]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncwarp();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(holder) : "memory");
}

int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 1024 * 8); cudaMalloc(&sink, 1024 * 1024 * 4);
  const int iters = 20000;
  for (int warps : {1, 2, 4, 8, 16, 32}) {
    for (int ctas_per_sm : {1, 2}) {
      if (warps * ctas_per_sm > 32) continue;
      hmma_rate<<<148 * ctas_per_sm, warps * 32>>>(iters, cyc, sink);
      cudaDeviceSynchronize();
      long long h[296]; cudaMemcpy(h, cyc, 148 * ctas_per_sm * 8, cudaMemcpyDeviceToHost);
      double mean = 0; for (int i = 0; i < 148 * ctas_per_sm; ++i) mean += h[i]; mean /= 148 * ctas_per_sm;
      double flop_per_sm = (double)iters * 8 * 4096 * warps * ctas_per_sm;
      printf("hmma m16n8k16 bf16: %2d warps x %d CTAs/SM: %.0f FLOP/clk/SM (%.2f cycles per mma per warp)\n", warps, ctas_per_sm,
             flop_per_sm / mean, mean / (iters * 8.0));
    }
  }
  float* out; cudaMalloc(&out, 512 * 4);
  ld_layout<<<1, 32>>>(out);
  cudaError_t e = cudaDeviceSynchronize();
  printf("ld_layout: %s\n", cudaGetErrorString(e));
  float h[512]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  for (int half = 0; half < 2; ++half)
    for (int lane : {0, 1, 2, 3, 4, 5, 31}) {
      printf("16x256b.x2 lane-offset %2d, thread %2d:", half * 16, lane);
      for (int i = 0; i < 8; ++i) printf(" (r%d c%d)", (int)h[half * 256 + lane * 8 + i] / 100, (int)h[half * 256 + lane * 8 + i] % 100);
      printf("\n");
    }
  return 0;
}
