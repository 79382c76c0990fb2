// Fused L1 + SSIM image loss, forward and backward.
// ref: utils/loss_utils.py:30-31 (l1_loss), :46-86 (gaussian / create_window / ssim / _ssim) and the
// GAUSSIAN-state loss of train.py:235-243, loss = (1 - l) * L1 + l * (1 - SSIM).
//
// The reference evaluates SSIM with five 11x11 depthwise convolutions (zero padding 5) plus ~15
// image-sized elementwise kernels, and autograd runs as many again.  Here one kernel per direction:
//   forward : per 32x32 tile, stage the 42x42 halo of both images in shared memory, run the separable
//             11-tap Gaussian over {x, y, x^2, y^2, xy}, form the SSIM map and its three partial
//             derivatives (w.r.t. mu1, E[x^2], E[xy]) and reduce |x-y| and SSIM to per-block partials;
//   backward: convolve the three derivative maps with the same (symmetric) window and combine:
//             dL/dx = k_l1 * sign(x-y) + k_ssim * (G*Dmu + 2x * G*D11 + y * G*D12),
//             k_* = (upstream gradient of each term) / N.
// Partials are reduced in a fixed order by a one-block kernel, so the loss value is deterministic.
#include "common.cuh"

#include <math.h>

namespace tb {

constexpr int LW = 11, LR = 5;                           // window size and radius
constexpr int LTX = 32, LTY = 32;                        // output tile
constexpr int LHX = LTX + 2 * LR, LHY = LTY + 2 * LR;    // tile + halo
constexpr int LOSS_THREADS = 256;
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;

struct LossWindow {
  float g[LW];
};

// the reference's window: fp32 of exp(-(x-5)^2 / (2 * 1.5^2)), normalised by its fp32 sum
static LossWindow make_window() {
  LossWindow w;
  float sum = 0.f;
  for (int i = 0; i < LW; i++) {
    w.g[i] = (float)exp(-(double)((i - LR) * (i - LR)) / (2.0 * 1.5 * 1.5));
    sum += w.g[i];
  }
  for (int i = 0; i < LW; i++) w.g[i] /= sum;
  return w;
}

__device__ __forceinline__ float block_sum(float v, float* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 0; w < LOSS_THREADS / 32; w++) t += s_red[w];
  }
  return t;   // valid on thread 0
}

__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_fwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, const LossWindow win,
                   int do_ssim, float* __restrict__ maps, float* __restrict__ partials) {
  pdl_wait();
  __shared__ float sx[LHY][LHX + 1], sy[LHY][LHX + 1];
  __shared__ __align__(16) float sh[5][LHY][LTX];
  __shared__ float s_red[LOSS_THREADS / 32];
  const int tx0 = blockIdx.x * LTX, ty0 = blockIdx.y * LTY;
  const size_t plane = (size_t)H * W;
  const float* x = img + blockIdx.z * plane;
  const float* y = gt + blockIdx.z * plane;

  for (int i = threadIdx.x; i < LHY * LHX; i += LOSS_THREADS) {
    const int r = i / LHX, c = i - r * LHX;
    const int gy = ty0 + r - LR, gx = tx0 + c - LR;
    const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
    sx[r][c] = in ? x[(size_t)gy * W + gx] : 0.f;
    sy[r][c] = in ? y[(size_t)gy * W + gx] : 0.f;
  }
  __syncthreads();
  if (do_ssim) {
    // horizontal pass, four adjacent outputs per work item: 14 + 14 loads feed 4 x 11 taps of all five moments
    // (same tap order per output as a one-output loop, so the sums are unchanged)
    for (int i = threadIdx.x; i < LHY * (LTX / 4); i += LOSS_THREADS) {
      const int r = i / (LTX / 4), c0 = (i - r * (LTX / 4)) * 4;
      float xs[LW + 3], ys[LW + 3];
#pragma unroll
      for (int k = 0; k < LW + 3; k++) { xs[k] = sx[r][c0 + k]; ys[k] = sy[r][c0 + k]; }
      float a[5][4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        a[0][j] = a[1][j] = a[2][j] = a[3][j] = a[4][j] = 0.f;
#pragma unroll
        for (int k = 0; k < LW; k++) {
          const float g = win.g[k], xv = xs[j + k], yv = ys[j + k];
          a[0][j] = fmaf(g, xv, a[0][j]);
          a[1][j] = fmaf(g, yv, a[1][j]);
          a[2][j] = fmaf(g, xv * xv, a[2][j]);
          a[3][j] = fmaf(g, yv * yv, a[3][j]);
          a[4][j] = fmaf(g, xv * yv, a[4][j]);
        }
      }
#pragma unroll
      for (int m = 0; m < 5; m++)
        *reinterpret_cast<float4*>(&sh[m][r][c0]) = make_float4(a[m][0], a[m][1], a[m][2], a[m][3]);
    }
    __syncthreads();
  }
  float l1 = 0.f, ss = 0.f;
  const int c = threadIdx.x & 31;
  constexpr int RPT = LTY / (LOSS_THREADS / 32);          // 4 vertically adjacent outputs per thread
  const int r0 = (threadIdx.x >> 5) * RPT;
  // vertical pass: 14 loads per moment feed the 4 x 11 taps of this thread's four rows
  float mom[5][RPT];
  if (do_ssim) {
#pragma unroll
    for (int m = 0; m < 5; m++) {
      float col[LW + RPT - 1];
#pragma unroll
      for (int k = 0; k < LW + RPT - 1; k++) col[k] = sh[m][r0 + k][c];
#pragma unroll
      for (int j = 0; j < RPT; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < LW; k++) acc = fmaf(win.g[k], col[j + k], acc);
        mom[m][j] = acc;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < RPT; j++) {
    const int r = r0 + j;
    const int gy = ty0 + r, gx = tx0 + c;
    if (gy >= H || gx >= W) continue;
    const float xv = sx[r + LR][c + LR], yv = sy[r + LR][c + LR];
    l1 += fabsf(xv - yv);
    if (!do_ssim) continue;
    const float mu1 = mom[0][j], mu2 = mom[1][j], e11 = mom[2][j], e22 = mom[3][j], e12 = mom[4][j];
    const float mu1sq = mu1 * mu1, mu2sq = mu2 * mu2, mu12 = mu1 * mu2;
    const float s1 = e11 - mu1sq, s2 = e22 - mu2sq, s12 = e12 - mu12;
    const float A1 = 2.f * mu12 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
    const float B1 = mu1sq + mu2sq + SSIM_C1, B2 = s1 + s2 + SSIM_C2;
    const float inv = 1.f / (B1 * B2);
    const float S = A1 * A2 * inv;
    ss += S;
    if (maps) {
      // S as a function of (mu1, E[x^2], E[xy]) with s1 = E[x^2] - mu1^2, s12 = E[xy] - mu1 mu2
      const size_t o = blockIdx.z * plane + (size_t)gy * W + gx;
      const size_t n = (size_t)gridDim.z * plane;
      maps[o] = 2.f * (mu2 * (A2 - A1) - mu1 * S * (B2 - B1)) * inv;
      maps[n + o] = -S / B2;
      maps[2 * n + o] = 2.f * A1 * inv;
    }
  }
  const float tl1 = block_sum(l1, s_red);
  const float tss = block_sum(ss, s_red);
  if (threadIdx.x == 0) {
    const size_t b = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    partials[2 * b] = tl1;
    partials[2 * b + 1] = tss;
  }
}

// fixed-order reduction of the per-block partials: out = {mean |x-y|, mean SSIM, w_l1 * L1 + w_ssim * SSIM + bias}
__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_finalize_kernel(int nblocks, const float* __restrict__ partials, double inv_n, float w_l1, float w_ssim,
                        float bias, float* __restrict__ out) {
  pdl_wait();
  __shared__ double s0[LOSS_THREADS], s1[LOSS_THREADS];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += LOSS_THREADS) {
    a += (double)partials[2 * i];
    b += (double)partials[2 * i + 1];
  }
  s0[threadIdx.x] = a;
  s1[threadIdx.x] = b;
  __syncthreads();
  for (int o = LOSS_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      s0[threadIdx.x] += s0[threadIdx.x + o];
      s1[threadIdx.x] += s1[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float l1 = (float)(s0[0] * inv_n), ss = (float)(s1[0] * inv_n);
    out[0] = l1;
    out[1] = ss;
    out[2] = w_l1 * l1 + w_ssim * ss + bias;
  }
}

__global__ void __launch_bounds__(LOSS_THREADS)
l1_ssim_bwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ gt,
                   const float* __restrict__ maps, const LossWindow win, float w_l1, float w_ssim, float inv_n,
                   const float* __restrict__ grad_out3, float* __restrict__ d_img) {
  pdl_wait();
  __shared__ float sm[3][LHY][LHX + 1];
  __shared__ __align__(16) float sh[3][LHY][LTX];
  const int tx0 = blockIdx.x * LTX, ty0 = blockIdx.y * LTY;
  const size_t plane = (size_t)H * W;
  const size_t n = (size_t)gridDim.z * plane;
  const size_t base = blockIdx.z * plane;
  // upstream gradients of {L1, SSIM, combined}: effective weights of the two terms
  const float g0 = grad_out3 ? grad_out3[0] : 0.f, g1 = grad_out3 ? grad_out3[1] : 0.f;
  const float g2 = grad_out3 ? grad_out3[2] : 1.f;
  const float k_l1 = (g2 * w_l1 + g0) * inv_n, k_ssim = (g2 * w_ssim + g1) * inv_n;
  const bool do_ssim = maps != nullptr;
  if (do_ssim) {
    for (int i = threadIdx.x; i < LHY * LHX; i += LOSS_THREADS) {
      const int r = i / LHX, c = i - r * LHX;
      const int gy = ty0 + r - LR, gx = tx0 + c - LR;
      const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
      const size_t o = base + (size_t)gy * W + gx;
      sm[0][r][c] = in ? maps[o] : 0.f;
      sm[1][r][c] = in ? maps[n + o] : 0.f;
      sm[2][r][c] = in ? maps[2 * n + o] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < LHY * (LTX / 4); i += LOSS_THREADS) {
      const int r = i / (LTX / 4), c0 = (i - r * (LTX / 4)) * 4;
#pragma unroll
      for (int m = 0; m < 3; m++) {
        float v[LW + 3];
#pragma unroll
        for (int k = 0; k < LW + 3; k++) v[k] = sm[m][r][c0 + k];
        float a[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          a[j] = 0.f;
#pragma unroll
          for (int k = 0; k < LW; k++) a[j] = fmaf(win.g[k], v[j + k], a[j]);
        }
        *reinterpret_cast<float4*>(&sh[m][r][c0]) = make_float4(a[0], a[1], a[2], a[3]);
      }
    }
    __syncthreads();
  }
  const int c = threadIdx.x & 31;
  constexpr int RPT = LTY / (LOSS_THREADS / 32);
  const int r0 = (threadIdx.x >> 5) * RPT;
  float cv[3][RPT];
  if (do_ssim) {
#pragma unroll
    for (int m = 0; m < 3; m++) {
      float col[LW + RPT - 1];
#pragma unroll
      for (int k = 0; k < LW + RPT - 1; k++) col[k] = sh[m][r0 + k][c];
#pragma unroll
      for (int j = 0; j < RPT; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < LW; k++) acc = fmaf(win.g[k], col[j + k], acc);
        cv[m][j] = acc;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < RPT; j++) {
    const int r = r0 + j;
    const int gy = ty0 + r, gx = tx0 + c;
    if (gy >= H || gx >= W) continue;
    const size_t o = base + (size_t)gy * W + gx;
    const float xv = img[o], yv = gt[o];
    const float d = xv - yv;
    float g = k_l1 * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
    if (do_ssim) g += k_ssim * (cv[0][j] + 2.f * xv * cv[1][j] + yv * cv[2][j]);
    d_img[o] = g;
  }
}

// ---- feature-norm regulariser (train.py:280-281): reg = (1 - mean_p ||F[:, p]||_2)^2 ------------------
constexpr int FN_MAX_BLOCKS = 148 * 8;

__global__ void __launch_bounds__(LOSS_THREADS)
feat_norm_fwd_kernel(int C, long long N, const float* __restrict__ feat, float* __restrict__ norm,
                     float* __restrict__ partials) {
  pdl_wait();
  __shared__ float s_red[LOSS_THREADS / 32];
  float acc = 0.f;
  for (long long p = (long long)blockIdx.x * LOSS_THREADS + threadIdx.x; p < N; p += (long long)gridDim.x * LOSS_THREADS) {
    float s = 0.f;
    for (int c = 0; c < C; c++) {
      const float f = feat[(size_t)c * N + p];
      s = fmaf(f, f, s);
    }
    const float n = sqrtf(s);
    norm[p] = n;
    acc += n;
  }
  const float t = block_sum(acc, s_red);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void __launch_bounds__(LOSS_THREADS)
feat_norm_finalize_kernel(int nblocks, const float* __restrict__ partials, double inv_n, float* __restrict__ out2) {
  pdl_wait();
  __shared__ double s0[LOSS_THREADS];
  double a = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += LOSS_THREADS) a += (double)partials[i];
  s0[threadIdx.x] = a;
  __syncthreads();
  for (int o = LOSS_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) s0[threadIdx.x] += s0[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float m = (float)(s0[0] * inv_n);
    out2[0] = m;
    out2[1] = (1.f - m) * (1.f - m);
  }
}

// d reg / d F[c, p] = (g_mean + g_reg * 2 (mean - 1)) / N * F[c, p] / ||F[:, p]||   (0 where the norm is 0)
__global__ void __launch_bounds__(LOSS_THREADS)
feat_norm_bwd_kernel(int C, long long N, const float* __restrict__ feat, const float* __restrict__ norm,
                     const float* __restrict__ out2, const float* __restrict__ grad_out2, float inv_n,
                     float* __restrict__ d_feat) {
  pdl_wait();
  const float g_mean = grad_out2 ? grad_out2[0] : 0.f, g_reg = grad_out2 ? grad_out2[1] : 1.f;
  const float k = (g_mean + g_reg * 2.f * (out2[0] - 1.f)) * inv_n;
  for (long long p = (long long)blockIdx.x * LOSS_THREADS + threadIdx.x; p < N; p += (long long)gridDim.x * LOSS_THREADS) {
    const float n = norm[p];
    const float w = n > 0.f ? k / n : 0.f;
    for (int c = 0; c < C; c++) d_feat[(size_t)c * N + p] = w * feat[(size_t)c * N + p];
  }
}

static int feat_norm_blocks(long long N) {
  const long long b = (N + LOSS_THREADS - 1) / LOSS_THREADS;
  return (int)(b < FN_MAX_BLOCKS ? (b > 0 ? b : 1) : FN_MAX_BLOCKS);
}

static bool loss_dims_ok(int C, int H, int W) {
  return C > 0 && H > 0 && W > 0 && C <= 65535 && (H + LTY - 1) / LTY <= 65535;
}

}  // namespace tb

using namespace tb;

extern "C" TRASE_API size_t trase_b200_loss_scratch_bytes(int C, int H, int W) {
  if (!loss_dims_ok(C, H, W)) return 0;
  return (size_t)C * ((H + LTY - 1) / LTY) * ((W + LTX - 1) / LTX) * 2 * sizeof(float);
}

extern "C" TRASE_API int trase_b200_l1_ssim_forward(int C, int H, int W, const float* img, const float* gt, float w_l1,
                                                    float w_ssim, float bias, float* maps, float* scratch,
                                                    float* out3, void* stream) {
  if (!loss_dims_ok(C, H, W)) { set_error("trase_b200_l1_ssim_forward: bad image shape %dx%dx%d", C, H, W); return 1; }
  if (!img || !gt || !scratch || !out3) { set_error("trase_b200_l1_ssim_forward: null pointer"); return 1; }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  static const LossWindow win = make_window();
  const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, C);
  const int nblocks = (int)(grid.x * grid.y * grid.z);
  launch_k(l1_ssim_fwd_kernel, grid, LOSS_THREADS, 0, s, H, W, img, gt, win, w_ssim != 0.f ? 1 : 0, maps, scratch);
  launch_k(l1_ssim_finalize_kernel, 1, LOSS_THREADS, 0, s, nblocks, scratch, 1.0 / ((double)C * H * W), w_l1, w_ssim, bias,
                                                     out3);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_l1_ssim_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(2);
  return 0;
}

extern "C" TRASE_API int trase_b200_l1_ssim_backward(int C, int H, int W, const float* img, const float* gt,
                                                     const float* maps, float w_l1, float w_ssim,
                                                     const float* grad_out3, float* d_img, void* stream) {
  if (!loss_dims_ok(C, H, W)) { set_error("trase_b200_l1_ssim_backward: bad image shape %dx%dx%d", C, H, W); return 1; }
  if (!img || !gt || !d_img || (w_ssim != 0.f && !maps)) { set_error("trase_b200_l1_ssim_backward: null pointer"); return 1; }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  static const LossWindow win = make_window();
  const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, C);
  launch_k(l1_ssim_bwd_kernel, grid, LOSS_THREADS, 0, s, H, W, img, gt, maps, win, w_l1, w_ssim,
                                                   (float)(1.0 / ((double)C * H * W)), grad_out3, d_img);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_l1_ssim_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API size_t trase_b200_feature_norm_scratch_bytes(int64_t N) {
  return N > 0 ? (size_t)feat_norm_blocks(N) * sizeof(float) : 0;
}

extern "C" TRASE_API int trase_b200_feature_norm_reg_forward(int C, int64_t N, const float* feat, float* norm,
                                                             float* scratch, float* out2, void* stream) {
  if (C <= 0 || N <= 0) { set_error("trase_b200_feature_norm_reg_forward: empty feature image"); return 1; }
  if (!feat || !norm || !scratch || !out2) { set_error("trase_b200_feature_norm_reg_forward: null pointer"); return 1; }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int nb = feat_norm_blocks(N);
  launch_k(feat_norm_fwd_kernel, nb, LOSS_THREADS, 0, s, C, N, feat, norm, scratch);
  launch_k(feat_norm_finalize_kernel, 1, LOSS_THREADS, 0, s, nb, scratch, 1.0 / (double)N, out2);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_feature_norm_reg_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(2);
  return 0;
}

extern "C" TRASE_API int trase_b200_feature_norm_reg_backward(int C, int64_t N, const float* feat, const float* norm,
                                                              const float* out2, const float* grad_out2,
                                                              float* d_feat, void* stream) {
  if (C <= 0 || N <= 0) { set_error("trase_b200_feature_norm_reg_backward: empty feature image"); return 1; }
  if (!feat || !norm || !out2 || !d_feat) { set_error("trase_b200_feature_norm_reg_backward: null pointer"); return 1; }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  launch_k(feat_norm_bwd_kernel, feat_norm_blocks(N), LOSS_THREADS, 0, s, C, N, feat, norm, out2, grad_out2,
                                                                    (float)(1.0 / (double)N), d_feat);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_feature_norm_reg_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// ---- feature-loss front end: bilinear resize + sampled-pixel gather + L2 normalise ----
// ref: train.py:283  rendered_features = F.interpolate(rendered_features[None], sam_masks.shape[-2:], mode='bilinear')[0]
//      utils/feature_utils.py:50-53  f = normalize(rendered_features[:, sampled_pixel].permute(1, 0), dim=-1)
// The reference resizes all C x H x W planes to the mask resolution and then keeps ~5000 pixels; its backward
// writes full gradient planes several times.  Here one warp per SAMPLED pixel evaluates the four bilinear taps
// for all channels (lane = channel), normalises, and the backward scatters to the 4 x C touched entries only.
namespace tb {

// torch's area_pixel_compute_source_index (align_corners = False): src = scale * (dst + 0.5) - 0.5, clamped at 0
__device__ __forceinline__ void bilinear_taps(int dst, int in_size, float scale, int& i0, int& i1, float& l1) {
  float src = scale * ((float)dst + 0.5f) - 0.5f;
  src = src < 0.f ? 0.f : src;
  i0 = (int)src;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = src - (float)i0;
}

constexpr int SF_MAX_C = 64;   // channels per sample: two per lane at most

__global__ void __launch_bounds__(256) sample_features_fwd_kernel(int C, int H, int W, int Hm, int Wm, int N,
                                                                  const float* __restrict__ feat,
                                                                  const long long* __restrict__ idx, int normalize,
                                                                  float* __restrict__ out, float* __restrict__ inv_norm) {
  pdl_wait();
  const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (n >= N) return;
  const long long p = idx[n];
  const int ym = (int)(p / Wm), xm = (int)(p - (long long)ym * Wm);
  int y0, y1, x0, x1;
  float ly, lx;
  bilinear_taps(ym, H, (float)H / (float)Hm, y0, y1, ly);
  bilinear_taps(xm, W, (float)W / (float)Wm, x0, x1, lx);
  const float hy = 1.f - ly, hx = 1.f - lx;
  const size_t plane = (size_t)H * W;
  float v[2] = {0.f, 0.f};
  float ss = 0.f;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int c = lane + 32 * r;
    if (c < C) {
      const float* f = feat + c * plane;
      // same association as torch's upsample_bilinear2d: h0 * (w0 * a + w1 * b) + h1 * (w0 * c + w1 * d)
      v[r] = hy * (hx * f[(size_t)y0 * W + x0] + lx * f[(size_t)y0 * W + x1]) +
             ly * (hx * f[(size_t)y1 * W + x0] + lx * f[(size_t)y1 * W + x1]);
      ss += v[r] * v[r];
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  // F.normalize: v / max(||v||, 1e-12)
  const float inv = normalize ? 1.f / fmaxf(sqrtf(ss), 1e-12f) : 1.f;
  if (lane == 0 && inv_norm) inv_norm[n] = inv;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int c = lane + 32 * r;
    if (c < C) out[(size_t)n * C + c] = v[r] * inv;
  }
}

__global__ void __launch_bounds__(256) sample_features_bwd_kernel(int C, int H, int W, int Hm, int Wm, int N,
                                                                  const long long* __restrict__ idx, int normalize,
                                                                  const float* __restrict__ out,
                                                                  const float* __restrict__ inv_norm,
                                                                  const float* __restrict__ g_out,
                                                                  float* __restrict__ d_feat) {
  pdl_wait();
  const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (n >= N) return;
  const long long p = idx[n];
  const int ym = (int)(p / Wm), xm = (int)(p - (long long)ym * Wm);
  int y0, y1, x0, x1;
  float ly, lx;
  bilinear_taps(ym, H, (float)H / (float)Hm, y0, y1, ly);
  bilinear_taps(xm, W, (float)W / (float)Wm, x0, x1, lx);
  const float hy = 1.f - ly, hx = 1.f - lx;
  const size_t plane = (size_t)H * W;
  float g[2] = {0.f, 0.f}, o[2] = {0.f, 0.f};
  float dot = 0.f;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int c = lane + 32 * r;
    if (c < C) {
      g[r] = g_out[(size_t)n * C + c];
      o[r] = out[(size_t)n * C + c];
      dot += g[r] * o[r];
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, s);
  const float inv = normalize ? inv_norm[n] : 1.f;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int c = lane + 32 * r;
    if (c < C) {
      // d/dv of v / ||v||: (g - o (o . g)) / ||v||     (o = normalised output)
      const float gv = normalize ? (g[r] - o[r] * dot) * inv : g[r];
      float* d = d_feat + c * plane;
      atomicAdd(d + (size_t)y0 * W + x0, hy * hx * gv);
      atomicAdd(d + (size_t)y0 * W + x1, hy * lx * gv);
      atomicAdd(d + (size_t)y1 * W + x0, ly * hx * gv);
      atomicAdd(d + (size_t)y1 * W + x1, ly * lx * gv);
    }
  }
}

}  // namespace tb

extern "C" TRASE_API int trase_b200_sample_features_forward(int C, int H, int W, int Hm, int Wm, int N,
                                                            const float* feat, const int64_t* idx, int normalize,
                                                            float* out, float* inv_norm, void* stream) {
  if (N <= 0) return 0;
  if (C <= 0 || C > SF_MAX_C || H <= 0 || W <= 0 || Hm <= 0 || Wm <= 0 || !feat || !idx || !out) {
    set_error("trase_b200_sample_features_forward: bad arguments (C=%d, at most %d channels)", C, SF_MAX_C);
    return 1;
  }
  launch_k(sample_features_fwd_kernel, (N + 7) / 8, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      C, H, W, Hm, Wm, N, feat, reinterpret_cast<const long long*>(idx), normalize, out, inv_norm);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_sample_features_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// d_feat ([C][H][W]) must be zero-initialised (or hold a gradient to accumulate into).
extern "C" TRASE_API int trase_b200_sample_features_backward(int C, int H, int W, int Hm, int Wm, int N,
                                                             const int64_t* idx, int normalize, const float* out,
                                                             const float* inv_norm, const float* g_out,
                                                             float* d_feat, void* stream) {
  if (N <= 0) return 0;
  if (C <= 0 || C > SF_MAX_C || !idx || !out || !g_out || !d_feat || (normalize && !inv_norm)) {
    set_error("trase_b200_sample_features_backward: bad arguments");
    return 1;
  }
  launch_k(sample_features_bwd_kernel, (N + 7) / 8, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      C, H, W, Hm, Wm, N, reinterpret_cast<const long long*>(idx), normalize, out, inv_norm, g_out, d_feat);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_sample_features_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
