// Small helper kernels around the MLP GEMMs: TF32 rounding / transposition of the
// weights, packing of the three output heads into one 16-row matrix, packing of the
// three upstream gradients into one [M][16] operand (+ their column sums = head bias
// gradients).  Each replaces a handful of elementwise torch kernels per layer.
#include "common.cuh"
#include <cuda_bf16.h>

namespace tb {
namespace mlp {

__device__ __forceinline__ float rna(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

// Wr[r][c] = rna(W[r][c]) for all columns; WrT[c - col0][r] = rna(W[r][c]) for c in [col0, col0 + ncols)
__global__ void __launch_bounds__(256) round_transpose_kernel(const float* __restrict__ W, int rows, int cols, int col0,
                                                              int ncols, float* __restrict__ Wr,
                                                              float* __restrict__ WrT) {
  pdl_wait();
  __shared__ float tile[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;   // bx: column block, by: row block
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
  for (int j = ty; j < 32; j += 8) {
    const int r = by + j, c = bx + tx;
    float v = 0.f;
    if (r < rows && c < cols) {
      v = rna(W[(size_t)r * cols + c]);
      if (Wr) Wr[(size_t)r * cols + c] = v;
    }
    tile[j][tx] = v;
  }
  __syncthreads();
  if (WrT) {
    for (int j = ty; j < 32; j += 8) {
      const int c = bx + j, r = by + tx;                 // output row = c - col0, output column = r
      if (c >= col0 && c < col0 + ncols && r < rows) WrT[(size_t)(c - col0) * rows + r] = tile[tx][j];
    }
  }
}

// wh[16][K], whT[K][16], bh[16]: rows 0-2 warp, 3-6 rotation, 7-9 scaling, 10-15 zero
__global__ void __launch_bounds__(256) pack_heads_kernel(const float* __restrict__ w0, const float* __restrict__ b0,
                                                         const float* __restrict__ w1, const float* __restrict__ b1,
                                                         const float* __restrict__ w2, const float* __restrict__ b2,
                                                         int K, float* __restrict__ wh, float* __restrict__ whT,
                                                         float* __restrict__ bh) {
  pdl_wait();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < 16 * K) {
    const int r = e / K, c = e - r * K;
    float v = 0.f;
    if (r < 3) v = w0[r * K + c];
    else if (r < 7) v = w1[(r - 3) * K + c];
    else if (r < 10) v = w2[(r - 7) * K + c];
    v = rna(v);
    wh[e] = v;
    whT[c * 16 + r] = v;
  }
  if (e < 16) bh[e] = e < 3 ? b0[e] : (e < 7 ? b1[e - 3] : (e < 10 ? b2[e - 7] : 0.f));
}

// dout[m][16] = rna([g0[m][0..2], g1[m][0..3], g2[m][0..2], 0...]);  colsum[j] += sum_m (unrounded) values;
// dout16[m][32] (optional) = the same 16 values as bf16 hi and lo parts
__global__ void __launch_bounds__(256) pack_dout_kernel(int M, const float* __restrict__ g0, const float* __restrict__ g1,
                                                        const float* __restrict__ g2, float* __restrict__ dout,
                                                        __nv_bfloat16* __restrict__ dout16, float* __restrict__ colsum) {
  pdl_wait();
  __shared__ float s_sum[8][16];
  const int lane16 = threadIdx.x & 15;             // column
  const int rsub = threadIdx.x >> 4;               // 16 rows per block iteration
  float acc = 0.f;
  for (int m = blockIdx.x * 16 + rsub; m < M; m += gridDim.x * 16) {
    float v = 0.f;
    if (lane16 < 3) v = g0 ? g0[(size_t)m * 3 + lane16] : 0.f;
    else if (lane16 < 7) v = g1 ? g1[(size_t)m * 4 + (lane16 - 3)] : 0.f;
    else if (lane16 < 10) v = g2 ? g2[(size_t)m * 3 + (lane16 - 7)] : 0.f;
    acc += v;
    dout[(size_t)m * 16 + lane16] = rna(v);
    if (dout16) {
      // operand of the bf16 head weight gradient, as TWO bf16 terms [hi(16) | lo(16)] per row: this factor comes out of
      // the rasterizer's atomics (last-bit differences from run to run), and a single 8-bit rounding of it turns such a
      // difference into a 0.4 % step of an element — visible in the head gradients' cancelling sums
      const __nv_bfloat16 hi = __float2bfloat16_rn(v);
      dout16[(size_t)m * 32 + lane16] = hi;
      dout16[(size_t)m * 32 + 16 + lane16] = __float2bfloat16_rn(v - __bfloat162float(hi));
    }
  }
  // rows live in lanes {l, l+16} of each warp and in the 8 warps
  acc += __shfl_xor_sync(0xffffffffu, acc, 16);
  if ((threadIdx.x & 31) < 16) s_sum[threadIdx.x >> 5][lane16] = acc;
  __syncthreads();
  if (threadIdx.x < 16) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; w++) t += s_sum[w][threadIdx.x];
    atomicAdd(colsum + threadIdx.x, t);
  }
}

// Head gradients straight into the parameters' gradient buffers: gw_h[j][i] += sum over the `halves` 16-column blocks
// of dwhT[i][16 b + col0_h + j], gb_h[j] += dbh[col0_h + j] for the three heads (3 / 4 / 3 rows, columns 0 / 3 / 7 of
// the packed head matrix).  One launch instead of the hi + lo add and six gradient accumulations.
__global__ void __launch_bounds__(256) heads_grad_accum_kernel(const float* __restrict__ dwhT, int ld, int halves,
                                                               const float* __restrict__ dbh, float* gw0, float* gb0,
                                                               float* gw1, float* gb1, float* gw2, float* gb2, int K) {
  pdl_wait();
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < 10 * K; e += gridDim.x * blockDim.x) {
    const int j = e / K, i = e - j * K;                 // row j of the packed head matrix, input feature i
    float v = 0.f;
    for (int b = 0; b < halves; b++) v += dwhT[(size_t)i * ld + 16 * b + j];
    float* dst = j < 3 ? gw0 + (size_t)j * K : (j < 7 ? gw1 + (size_t)(j - 3) * K : gw2 + (size_t)(j - 7) * K);
    dst[i] += v;
    if (i == 0) {
      float* bd = j < 3 ? gb0 + j : (j < 7 ? gb1 + (j - 3) : gb2 + (j - 7));
      *bd += dbh[j];
    }
  }
}

}  // namespace mlp
}  // namespace tb

using namespace tb;

// dwhT: [K][ld] (ld = 16, halves = 1: fp32 path; ld = 32, halves = 2: the bf16 path's hi and lo products), dbh: [16]
// head bias gradient; gw*/gb*: the three heads' weight [rows][K] and bias gradients, ACCUMULATED into.
extern "C" TRASE_API int trase_b200_mlp_heads_grad_accumulate(const float* dwhT, int ld, int halves, const float* dbh,
                                                              float* gw0, float* gb0, float* gw1, float* gb1,
                                                              float* gw2, float* gb2, int K, void* stream) {
  if (!dwhT || !dbh || !gw0 || !gb0 || !gw1 || !gb1 || !gw2 || !gb2 || K <= 0 || halves < 1 || ld < 16 * halves) {
    set_error("trase_b200_mlp_heads_grad_accumulate: bad arguments");
    return 1;
  }
  launch_k(mlp::heads_grad_accum_kernel, (10 * K + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream), dwhT, ld,
           halves, dbh, gw0, gb0, gw1, gb1, gw2, gb2, K);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_heads_grad_accumulate: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API int trase_b200_mlp_round_transpose(const float* W, int rows, int cols, int col0, int ncols,
                                                        float* Wr, float* WrT, void* stream) {
  if (!W || rows <= 0 || cols <= 0 || col0 < 0 || ncols < 0 || col0 + ncols > cols) {
    set_error("trase_b200_mlp_round_transpose: bad arguments");
    return 1;
  }
  dim3 grid((cols + 31) / 32, (rows + 31) / 32);
  launch_k(mlp::round_transpose_kernel, grid, 256, 0, reinterpret_cast<cudaStream_t>(stream), W, rows, cols, col0, ncols, Wr, WrT);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_round_transpose: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API int trase_b200_mlp_pack_heads(const float* w0, const float* b0, const float* w1, const float* b1,
                                                   const float* w2, const float* b2, int K, float* wh, float* whT,
                                                   float* bh, void* stream) {
  if (!w0 || !b0 || !w1 || !b1 || !w2 || !b2 || !wh || !whT || !bh || K <= 0) {
    set_error("trase_b200_mlp_pack_heads: bad arguments");
    return 1;
  }
  launch_k(mlp::pack_heads_kernel, (16 * K + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream), w0, b0, w1, b1, w2, b2, K,
                                                                                               wh, whT, bh);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_pack_heads: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

static int pack_dout_impl(int M, const float* g0, const float* g1, const float* g2, float* dout, void* dout16,
                          float* colsum, void* stream);
extern "C" TRASE_API int trase_b200_mlp_pack_dout(int M, const float* g0, const float* g1, const float* g2, float* dout,
                                                  float* colsum, void* stream) {
  return pack_dout_impl(M, g0, g1, g2, dout, nullptr, colsum, stream);
}
// same, also writing the [M][32] bf16 copy (hi | lo parts) that the bf16 head weight gradient multiplies with the last
// activations
extern "C" TRASE_API int trase_b200_mlp_pack_dout_bf16(int M, const float* g0, const float* g1, const float* g2,
                                                       float* dout, void* dout16, float* colsum, void* stream) {
  if (M > 0 && !dout16) { set_error("trase_b200_mlp_pack_dout_bf16: bad arguments"); return 1; }
  return pack_dout_impl(M, g0, g1, g2, dout, dout16, colsum, stream);
}
static int pack_dout_impl(int M, const float* g0, const float* g1, const float* g2, float* dout, void* dout16,
                          float* colsum, void* stream) {
  if (M <= 0) return 0;
  if (!dout || !colsum) { set_error("trase_b200_mlp_pack_dout: bad arguments"); return 1; }
  int grid = (M + 15) / 16;
  if (grid > 148 * 8) grid = 148 * 8;
  launch_k(mlp::pack_dout_kernel, grid, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      M, g0, g1, g2, dout, reinterpret_cast<__nv_bfloat16*>(dout16), colsum);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_pack_dout: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
