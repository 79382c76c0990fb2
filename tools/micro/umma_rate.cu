// Issue rate of the tcgen05 MMA the whole-network MLP kernels are built on (kind::tf32, both operands in shared memory,
// 128-byte swizzle, K = 8 per instruction), measured in SM cycles per instruction on every SM at once:
//   * cta_group::2, M = 256 (128 rows per CTA), N = 256 or 128 — the shape mlp_fused_kernel issues;
//   * cta_group::1, M = 128, N = 256;
//   * the same with 0 / 4 / 8 other warps of the CTA storing to shared memory as fast as they can (what the epilogue
//     warps and the TMA weight stream do to the shared-memory port while the MMAs run), or draining the other
//     accumulator with tcgen05.ld, and with one or two tcgen05.commit per 4 MMAs;
//   * short bursts from an idle pipe: cycles from the first issue until the commit's arrival is seen.
// Operand contents are irrelevant (zeros).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I trase_b200/csrc
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace tb::ptx;

static __device__ int ITERS_D = 4096;
static int ITERS = 4096;      // k-blocks of 4 MMAs

template <int CG, int N>
__global__ void __launch_bounds__(384, 1) rate_kernel(long long* out, int noisy_warps, int noisy_bytes_per_iter, int commits, int noisy_tmem) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                 // 128 rows x 32 fp32, 4 k-steps
  uint8_t* sB = smem + 16384;         // 128 (CG 2) or 256 (CG 1) rows
  uint8_t* sN = smem + 16384 + 32768; // 64 KB the noisy warps write
  __shared__ uint64_t bar;
  __shared__ uint64_t sink[2];      // absorbs the per-k-block commits (count far above what arrives)
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (16384 + 32768 + 65536) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&sink[0], 1u << 20); mbar_init(&sink[1], 1u << 20); stop = 0; fence_barrier_init(); }
  if (warp == 2) {
    if (CG == 2) { tmem_alloc_2cta(&tmem_slot, 512); tmem_relinquish_2cta(); }
    else { tmem_alloc(&tmem_slot, 512); tmem_relinquish(); }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync();
  tc_fence_after();
  const uint32_t tmem_d = tmem_slot;
  const bool leader = CG == 1 || cluster_ctarank() == 0;
  if (warp == 1 && lane == 0) {
    if (leader) {
      constexpr uint32_t idesc = make_idesc_tf32(CG == 2 ? 256 : 128, N, 0, 0);
      const uint32_t a = smem_u32(sA), b = smem_u32(sB);
      const long long t0 = clock64();
      const int iters = ITERS_D;
      for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const uint64_t da = make_smem_desc_sw128(a + k * 32, 16, 1024);
          const uint64_t db = make_smem_desc_sw128(b + k * 32, 16, 1024);
          if (CG == 2) mma_tf32_ss_2cta(tmem_d, da, db, idesc, (it | k) != 0);
          else mma_tf32_ss(tmem_d, da, db, idesc, (it | k) != 0);
        }
        for (int c = 0; c < commits; c++) { if (CG == 2) mma_commit_2cta(&sink[c], 3); else mma_commit(&sink[c]); }
      }
      if (CG == 2) mma_commit_2cta(&bar, 3); else mma_commit(&bar);
      const long long t1 = clock64();
      mbar_wait(&bar, 0);
      const long long t2 = clock64();
      out[blockIdx.x * 2] = t2 - t0;
      out[blockIdx.x * 2 + 1] = t1 - t0;
    } else {
      mbar_wait(&bar, 0);
    }
    stop = 1;
  } else if (warp >= 4 && warp < 4 + noisy_warps) {
    // each noisy warp owns 8 KB and rewrites noisy_bytes_per_iter of it per loop trip with STS.128
    uint8_t* mine = sN + (warp - 4) * 8192;
    uint32_t x = lane;
    if (noisy_tmem) {
      // drain the OTHER accumulator (columns 256..511) over and over, as the epilogue warps do
      uint32_t v[32], acc = 0;
      const uint32_t t = tmem_d + ((uint32_t)((warp & 3) * 32) << 16) + 256;
      while (!stop) {
        for (int s8 = 0; s8 < 8; s8++) { tmem_ld_32x32(t + s8 * 32, v); tmem_ld_wait(); acc += v[0] ^ v[31]; }
      }
      if (acc == 0x12345u) out[0] = acc;
    } else
    while (!stop) {
      for (int o = 0; o < noisy_bytes_per_iter; o += 512) *reinterpret_cast<uint4*>(mine + o + lane * 16) = make_uint4(x, x, x, x);
      x++;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync();
  if (warp == 2) { if (CG == 2) tmem_dealloc_2cta(tmem_d, 512); else tmem_dealloc(tmem_d, 512); }
}

static double last_total, last_issue;
template <int CG, int N>
static void run(const char* tag, int noisy, long long* d_out, int commits = 0, int noisy_tmem = 0) {
  const int smem = 16384 + 32768 + 65536;
  cudaFuncSetAttribute(rate_kernel<CG, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(148);
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  for (int rep = 0; rep < 2; rep++) {
    cudaError_t e = cudaLaunchKernelEx(&cfg, rate_kernel<CG, N>, d_out, noisy, 8192, commits, noisy_tmem);
    if (e != cudaSuccess || (e = cudaDeviceSynchronize()) != cudaSuccess) { printf("%s: %s\n", tag, cudaGetErrorString(e)); return; }
  }
  long long h[296];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  double tot = 0, iss = 0; int n = 0;
  for (int b = 0; b < 148; b += CG) { tot += h[2 * b]; iss += h[2 * b + 1]; n++; }
  const double per = tot / n / (ITERS * 4.0);
  last_total = tot / n; last_issue = iss / n;
  const double flop = 2.0 * (CG == 2 ? 256 : 128) * N * 8;
  printf("%-34s commits/k-block %d, %d noisy warps (%s): %7.1f cycles per MMA (issue loop alone %7.1f)  = %6.1f flop/cycle/SM\n", tag, commits, noisy, noisy_tmem ? "tcgen05.ld" : "STS.128", per,
         iss / n / (ITERS * 4.0), flop / per / CG);
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 296 * sizeof(long long));
  for (int noisy : {0, 4, 8}) {
    run<2, 256>("cta_group::2 M=256 N=256 K=8 tf32", noisy, d_out);
    run<2, 128>("cta_group::2 M=256 N=128 K=8 tf32", noisy, d_out);
    run<1, 256>("cta_group::1 M=128 N=256 K=8 tf32", noisy, d_out);
  }
  // a commit (or two) after every 4 MMAs, as the ring of the real kernel needs; epilogue-like TMEM reads alongside
  for (int commits : {1, 2}) run<2, 256>("cta_group::2 M=256 N=256 K=8 tf32", 0, d_out, commits, 0);
  for (int noisy : {4, 8}) run<2, 256>("cta_group::2 M=256 N=256 K=8 tf32", noisy, d_out, 0, 1);
  run<2, 256>("cta_group::2 M=256 N=256 K=8 tf32", 8, d_out, 2, 1);
  // latency from the first issue to the commit's arrival being seen by the waiting thread, for short bursts
  for (int n : {1, 2, 4, 8, 16}) {
    ITERS = n;
    cudaMemcpyToSymbol(ITERS_D, &n, sizeof(int));
    char tag[64];
    snprintf(tag, sizeof(tag), "burst of %d MMAs, cta_group::2 N=256", 4 * n);
    run<2, 256>(tag, 0, d_out);
    printf("    -> %0.f cycles from first issue to completion seen, %0.f cycles to issue\n", last_total, last_issue);
  }
  return 0;
}
