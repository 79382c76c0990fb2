// Throughput of the legacy warp-level tensor path on this GPU: mma.sync.aligned.m16n8k8 tf32 (fp32 accumulate),
// 8 independent accumulator tiles per warp, all SMs busy.  Prints TFLOP/s for a few resident-warp counts.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, int iters) {
  unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 * 3, b1 = a0 * 5;
  float c[8][4];
#pragma unroll
  for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 16 * 1024 * 4);
  for (int threads : {128, 256, 512, 1024}) {
    for (int blocks_per_sm : {1, 2}) {
      int grid = 148 * blocks_per_sm, iters = 20000;
      k<<<grid, threads>>>(out, 100);
      cudaDeviceSynchronize();
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0); k<<<grid, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double flops = (double)grid * (threads / 32) * iters * 8.0 * 16 * 8 * 8 * 2;
      printf("threads %4d x %d blocks/SM: %.1f TFLOP/s tf32 mma.sync (%.3f ms)\n", threads, blocks_per_sm, flops / ms / 1e9, ms);
    }
  }
  return 0;
}
