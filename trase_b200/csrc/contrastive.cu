// FEATURE-state contrastive loss on the sampled pixels, without the N x N matrices (last part).
//
// ref: train.py:253-296 with the default contrastive_mode = 'soft':
//   get_pixel_mask_correspondence_matrix, get_pixel_weights            utils/feature_utils.py:28-48
//   get_features_correspondence_matrix (C_F = F F^T)                    utils/feature_utils.py:50-55
//   pixel_mask_correspondence_loss_soft_hard_positive / _soft_negative  utils/loss_utils.py:304-349
// The reference materialises four N x N tensors (N ~ 5000 sampled pixels: C, weights, C_F and boolean masks, ~100 MB
// each) and runs ~40 elementwise / reduction kernels plus two host synchronisations per loss over them, forward
// and again in autograd.  Everything those matrices hold is a function of per-pixel data:
//   C[i][j]      = 1 iff pixels i and j share a sampled mask  = (bits_i & bits_j) != 0   (bit m = "in sampled mask m")
//   weight[i][j] = affine map of clamp(top / (a_i a_j), 1) onto [1, 10], a_i = mean area of the masks covering pixel i;
//                  its minimum over all pairs is exactly 1 (the pair of the largest a with itself) and its maximum
//                  clamp(top / (a_min a_min), 1), so the min-max normalisation needs only a_min and a_max
//   C_F[i][j]    = f_i . f_j  (32 channels)
// so three sweeps over the pair space recompute them on the fly, each block owning 64 pixels and looping over all
// 64-pixel tiles of partners (one wave of ~80 blocks, no atomics, fixed summation order = deterministic):
//   FLAGS: kept_pos[j] = any_i (C_F < pos_th and same mask), kept_neg[j] = any_i (C_F > neg_th and different mask)
//          — over ALL i, the diagonal included, exactly like torch.any(..., dim=0);
//   LOSS : pairs i < j of kept columns: sum of -w C_F (same mask) and of w relu(C_F) (different mask), pair counts;
//          a last tiny kernel forms  n_kept = sum_j kept[j] * j  and  loss = sum / n_kept  (0 when no pair was selected);
//   GRAD : d loss / d f_i = sum_j G[i][j] f_j  with the pair coefficient of {i, j} taken at column max(i, j).
#include "common.cuh"

namespace tb {
namespace contrastive {

constexpr int CT = 64;            // pixels per tile
constexpr int CC = 32;            // feature channels
constexpr int CT_THREADS = 256;   // 16 x 16 threads, each a 4 x 4 patch of the 64 x 64 pair tile
constexpr int MAXW = 4;           // up to 256 masks

struct Params {
  int N, nwords, use_weights;
  const float* f;                 // [N][32] normalised sampled features
  const unsigned long long* bits; // [N][nwords]
  const float* a;                 // [N] mean mask area per sampled pixel
  const float* a_minmax;          // {min a, max a} (device)
  float pos_th, neg_th;
  unsigned char* kept;            // [2][N]  kept_pos, kept_neg
  double* partial;                // [blocks][4] pos_sum, neg_sum, pos_pairs, neg_pairs
  float* result;                  // [8]: loss_pos, loss_neg, n_pos, n_neg, pos_pairs, neg_pairs
  const float* upstream;          // [2] d/d loss_pos, d/d loss_neg (GRAD)
  float* df;                      // [N][32] (GRAD)
};

struct WeightCtx {
  float top, hi;                  // (max a)^2 and the largest clamped ratio
  bool on;
};
__device__ __forceinline__ WeightCtx weight_ctx(const Params& p) {
  WeightCtx w;
  w.on = p.use_weights != 0;
  if (w.on) {
    const float amin = p.a_minmax[0], amax = p.a_minmax[1];
    w.top = amax * amax;
    float pr = amin * amin;
    if (pr == 0.f) pr = 1e10f;
    w.hi = fmaxf(w.top / pr, 1.0f);
  } else {
    w.top = w.hi = 1.f;
  }
  return w;
}
// (clamp(top / (a_i a_j), 1) - 1) / (hi - 1) * 9 + 1     (feature_utils.py:31-37; the minimum over pairs is 1)
__device__ __forceinline__ float pair_weight(const WeightCtx& w, float ai, float aj) {
  if (!w.on) return 1.f;
  float pr = ai * aj;
  if (pr == 0.f) pr = 1e10f;
  const float r = fmaxf(w.top / pr, 1.0f);
  return (r - 1.0f) / (w.hi - 1.0f) * 9.0f + 1.0f;
}

enum { MODE_FLAGS = 0, MODE_LOSS = 1, MODE_GRAD = 2 };

template <int MODE>
__global__ void __launch_bounds__(CT_THREADS) contrastive_kernel(const Params p) {
  __shared__ float s_own[CT][CC + 1];
  __shared__ float s_oth[CT][CC + 1];
  __shared__ unsigned long long s_bown[CT][MAXW], s_both[CT][MAXW];
  __shared__ float s_aown[CT], s_aoth[CT];
  __shared__ unsigned char s_kown[2][CT], s_koth[2][CT];
  __shared__ float s_G[MODE == MODE_GRAD ? CT : 1][CT + 1];
  __shared__ double s_red[4][CT_THREADS / 32];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int own0 = blockIdx.x * CT;
  const WeightCtx wc = weight_ctx(p);
  for (int e = tid; e < CT * CC; e += CT_THREADS) {
    const int r = e / CC, c = e - r * CC;
    s_own[r][c] = own0 + r < p.N ? p.f[(size_t)(own0 + r) * CC + c] : 0.f;
  }
  for (int e = tid; e < CT * MAXW; e += CT_THREADS) {
    const int r = e / MAXW, w = e - r * MAXW;
    s_bown[r][w] = (own0 + r < p.N && w < p.nwords) ? p.bits[(size_t)(own0 + r) * p.nwords + w] : 0ull;
  }
  if (tid < CT) {
    s_aown[tid] = (own0 + tid < p.N && p.use_weights) ? p.a[own0 + tid] : 1.f;
    if (MODE != MODE_FLAGS) {
      s_kown[0][tid] = own0 + tid < p.N ? p.kept[own0 + tid] : 0;
      s_kown[1][tid] = own0 + tid < p.N ? p.kept[p.N + own0 + tid] : 0;
    }
  }
  float inv_npos = 0.f, inv_nneg = 0.f;
  if (MODE == MODE_GRAD) {
    // the loss is sum / n_kept when a pair was selected and 0 otherwise
    inv_npos = p.result[4] > 0.f ? p.upstream[0] / p.result[2] : 0.f;
    inv_nneg = p.result[5] > 0.f ? p.upstream[1] / p.result[3] : 0.f;
  }

  unsigned flag_pos = 0, flag_neg = 0;                    // FLAGS: per own row (bit r of the thread's 4 rows)
  double sum_pos = 0, sum_neg = 0, cnt_pos = 0, cnt_neg = 0;
  const int gr = tid >> 2, gc = (tid & 3) * 8;            // GRAD phase 2: row and first of 8 channels
  float dacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};

  const int ntiles = (p.N + CT - 1) / CT;
  // LOSS only needs partners i < j: tiles up to the own one
  const int last_tile = MODE == MODE_LOSS ? blockIdx.x : ntiles - 1;
  for (int t = 0; t <= last_tile; t++) {
    const int oth0 = t * CT;
    __syncthreads();
    for (int e = tid; e < CT * CC; e += CT_THREADS) {
      const int r = e / CC, c = e - r * CC;
      s_oth[r][c] = oth0 + r < p.N ? p.f[(size_t)(oth0 + r) * CC + c] : 0.f;
    }
    for (int e = tid; e < CT * MAXW; e += CT_THREADS) {
      const int r = e / MAXW, w = e - r * MAXW;
      s_both[r][w] = (oth0 + r < p.N && w < p.nwords) ? p.bits[(size_t)(oth0 + r) * p.nwords + w] : 0ull;
    }
    if (tid < CT) {
      s_aoth[tid] = (oth0 + tid < p.N && p.use_weights) ? p.a[oth0 + tid] : 1.f;
      if (MODE == MODE_GRAD) {
        s_koth[0][tid] = oth0 + tid < p.N ? p.kept[oth0 + tid] : 0;
        s_koth[1][tid] = oth0 + tid < p.N ? p.kept[p.N + oth0 + tid] : 0;
      }
    }
    __syncthreads();
    // 4 x 4 patch of similarities: own rows ty*4.., partner rows tx*4..
    float c[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) c[a][b] = 0.f;
#pragma unroll 8
    for (int k = 0; k < CC; k++) {
      float o[4], q[4];
#pragma unroll
      for (int a = 0; a < 4; a++) { o[a] = s_own[ty * 4 + a][k]; q[a] = s_oth[tx * 4 + a][k]; }
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) c[a][b] = fmaf(o[a], q[b], c[a][b]);
    }
#pragma unroll
    for (int a = 0; a < 4; a++) {
      const int ro = ty * 4 + a, own = own0 + ro;
#pragma unroll
      for (int b = 0; b < 4; b++) {
        const int rq = tx * 4 + b, oth = oth0 + rq;
        const bool valid = own < p.N && oth < p.N;
        bool same = false;
#pragma unroll
        for (int w = 0; w < MAXW; w++) same = same || ((s_bown[ro][w] & s_both[rq][w]) != 0ull);
        const float cv = c[a][b];
        if (MODE == MODE_FLAGS) {
          if (valid && same && cv < p.pos_th) flag_pos |= 1u << a;
          if (valid && !same && cv > p.neg_th) flag_neg |= 1u << a;
        } else if (MODE == MODE_LOSS) {
          // own = column j, partner = row i < j
          if (valid && oth < own) {
            const float w = pair_weight(wc, s_aown[ro], s_aoth[rq]);
            if (same && s_kown[0][ro]) { sum_pos += (double)(-w * cv); cnt_pos += 1.0; }
            if (!same && s_kown[1][ro]) { sum_neg += (double)(w * fmaxf(cv, 0.f)); cnt_neg += 1.0; }
          }
        } else {
          float g = 0.f;
          if (valid && own != oth) {
            const bool own_is_col = own > oth;
            const unsigned char kp = own_is_col ? s_kown[0][ro] : s_koth[0][rq];
            const unsigned char kn = own_is_col ? s_kown[1][ro] : s_koth[1][rq];
            const float w = pair_weight(wc, s_aown[ro], s_aoth[rq]);
            if (same && kp) g = -w * inv_npos;
            if (!same && kn && cv > 0.f) g = w * inv_nneg;
          }
          s_G[ro][rq] = g;
        }
      }
    }
    if (MODE == MODE_GRAD) {
      __syncthreads();
#pragma unroll 4
      for (int j = 0; j < CT; j++) {
        const float g = s_G[gr][j];
#pragma unroll
        for (int q = 0; q < 8; q++) dacc[q] = fmaf(g, s_oth[j][gc + q], dacc[q]);
      }
    }
  }

  if (MODE == MODE_FLAGS) {
    // OR over the 16 threads (tx) that share the own rows: they are the 16 lanes of a half warp
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
      flag_pos |= __shfl_xor_sync(0xffffffffu, flag_pos, o);
      flag_neg |= __shfl_xor_sync(0xffffffffu, flag_neg, o);
    }
    if (tx == 0) {
#pragma unroll
      for (int a = 0; a < 4; a++) {
        const int own = own0 + ty * 4 + a;
        if (own < p.N) {
          p.kept[own] = (flag_pos >> a) & 1u;
          p.kept[p.N + own] = (flag_neg >> a) & 1u;
        }
      }
    }
  } else if (MODE == MODE_LOSS) {
    double v[4] = {sum_pos, sum_neg, cnt_pos, cnt_neg};
#pragma unroll
    for (int q = 0; q < 4; q++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
      if ((tid & 31) == 0) s_red[q][tid >> 5] = v[q];
    }
    __syncthreads();
    if (tid < 4) {
      double tot = 0;
      for (int w = 0; w < CT_THREADS / 32; w++) tot += s_red[tid][w];
      p.partial[(size_t)blockIdx.x * 4 + tid] = tot;
    }
  } else {
    const int own = own0 + gr;
    if (own < p.N) {
      float4* dst = reinterpret_cast<float4*>(p.df + (size_t)own * CC + gc);
      dst[0] = make_float4(dacc[0], dacc[1], dacc[2], dacc[3]);
      dst[1] = make_float4(dacc[4], dacc[5], dacc[6], dacc[7]);
    }
  }
}

// n_kept = sum_j kept[j] * j (the number of rows i < j of each kept column), loss = sum / n_kept
__global__ void __launch_bounds__(256) contrastive_finish_kernel(const Params p, int nblocks) {
  __shared__ double s_n[2][8];
  double n[2] = {0, 0};
  for (int j = threadIdx.x; j < p.N; j += 256) {
    if (p.kept[j]) n[0] += (double)j;
    if (p.kept[p.N + j]) n[1] += (double)j;
  }
#pragma unroll
  for (int q = 0; q < 2; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n[q] += __shfl_xor_sync(0xffffffffu, n[q], o);
    if ((threadIdx.x & 31) == 0) s_n[q][threadIdx.x >> 5] = n[q];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double np = 0, nn = 0, v[4] = {0, 0, 0, 0};
    for (int w = 0; w < 8; w++) { np += s_n[0][w]; nn += s_n[1][w]; }
    for (int b = 0; b < nblocks; b++)
      for (int q = 0; q < 4; q++) v[q] += p.partial[(size_t)b * 4 + q];
    p.result[0] = v[2] > 0 ? (float)(v[0] / np) : 0.f;
    p.result[1] = v[3] > 0 ? (float)(v[1] / nn) : 0.f;
    p.result[2] = (float)np;
    p.result[3] = (float)nn;
    p.result[4] = (float)v[2];
    p.result[5] = (float)v[3];
  }
}

// Per sampled pixel: membership bits over the sampled masks and the mean area of ALL masks covering it
// (feature_utils.py:29-32, 41-44).  masks [n_masks][HW] uint8, area [n_masks] int64 (pixels per mask).
__global__ void __launch_bounds__(256) pixel_mask_info_kernel(int N, int n_masks, long long HW, int nwords,
                                                              const unsigned char* __restrict__ masks,
                                                              const long long* __restrict__ area,
                                                              const unsigned char* __restrict__ sampled_mask,
                                                              const long long* __restrict__ pix,
                                                              unsigned long long* __restrict__ bits,
                                                              float* __restrict__ a) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const long long px = pix[n];
  unsigned long long b[MAXW] = {0, 0, 0, 0};
  long long sum_area = 0, cover = 0;
  for (int m = 0; m < n_masks; m++) {
    const unsigned char v = masks[(size_t)m * HW + px];
    if (v) {
      sum_area += area[m] * (long long)v;
      cover += v;
      if (sampled_mask[m]) b[m >> 6] |= 1ull << (m & 63);
    }
  }
  for (int w = 0; w < nwords; w++) bits[(size_t)n * nwords + w] = b[w];
  a[n] = (float)sum_area / ((float)cover + 1e-9f);
}

}  // namespace contrastive
}  // namespace tb

using namespace tb;

extern "C" TRASE_API size_t trase_b200_contrastive_scratch_bytes(int N) {
  const size_t nb = (size_t)((N + contrastive::CT - 1) / contrastive::CT);
  return align_up(2 * (size_t)(N > 0 ? N : 0), 256) + align_up(nb * 4 * sizeof(double), 256);
}

extern "C" TRASE_API int trase_b200_pixel_mask_info(int N, int n_masks, int64_t HW, const unsigned char* masks,
                                                    const int64_t* area, const unsigned char* sampled_mask,
                                                    const int64_t* pix, uint64_t* bits, float* a, void* stream) {
  clear_error();
  if (N <= 0) return 0;
  if (n_masks <= 0 || n_masks > 64 * contrastive::MAXW || !masks || !area || !sampled_mask || !pix || !bits || !a) {
    set_error("trase_b200_pixel_mask_info: bad arguments (n_masks=%d, at most %d)", n_masks, 64 * contrastive::MAXW);
    return 1;
  }
  const int nwords = (n_masks + 63) / 64;
  contrastive::pixel_mask_info_kernel<<<(N + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      N, n_masks, (long long)HW, nwords, masks, reinterpret_cast<const long long*>(area), sampled_mask,
      reinterpret_cast<const long long*>(pix), reinterpret_cast<unsigned long long*>(bits), a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_pixel_mask_info: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

static int contrastive_fill(contrastive::Params& p, const char* who, int N, int nwords, const float* f,
                            const uint64_t* bits, const float* a, const float* a_minmax, float pos_th, float neg_th,
                            void* scratch, float* result) {
  if (nwords < 1 || nwords > contrastive::MAXW || !f || !bits || !scratch || !result || (a && !a_minmax) ||
      ((uintptr_t)f & 15)) {
    set_error("%s: bad arguments (nwords=%d)", who, nwords);
    return 1;
  }
  p.N = N; p.nwords = nwords; p.use_weights = a != nullptr;
  p.f = f; p.bits = reinterpret_cast<const unsigned long long*>(bits); p.a = a; p.a_minmax = a_minmax;
  p.pos_th = pos_th; p.neg_th = neg_th;
  p.kept = reinterpret_cast<unsigned char*>(scratch);
  p.partial = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + align_up(2 * (size_t)N, 256));
  p.result = result; p.upstream = nullptr; p.df = nullptr;
  return 0;
}

extern "C" TRASE_API int trase_b200_contrastive_soft_forward(int N, int nwords, const float* f, const uint64_t* bits,
                                                             const float* a, const float* a_minmax, float pos_th,
                                                             float neg_th, void* scratch, float* result, void* stream) {
  clear_error();
  if (N <= 0) return 0;
  contrastive::Params p;
  if (int rc = contrastive_fill(p, "trase_b200_contrastive_soft_forward", N, nwords, f, bits, a, a_minmax, pos_th,
                                neg_th, scratch, result))
    return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int nb = (N + contrastive::CT - 1) / contrastive::CT;
  contrastive::contrastive_kernel<contrastive::MODE_FLAGS><<<nb, contrastive::CT_THREADS, 0, s>>>(p);
  contrastive::contrastive_kernel<contrastive::MODE_LOSS><<<nb, contrastive::CT_THREADS, 0, s>>>(p);
  contrastive::contrastive_finish_kernel<<<1, 256, 0, s>>>(p, nb);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_contrastive_soft_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(3);
  return 0;
}

extern "C" TRASE_API int trase_b200_contrastive_soft_backward(int N, int nwords, const float* f, const uint64_t* bits,
                                                              const float* a, const float* a_minmax, float pos_th,
                                                              float neg_th, const void* scratch, const float* result,
                                                              const float* upstream, float* df, void* stream) {
  clear_error();
  if (N <= 0) return 0;
  contrastive::Params p;
  if (int rc = contrastive_fill(p, "trase_b200_contrastive_soft_backward", N, nwords, f, bits, a, a_minmax, pos_th,
                                neg_th, const_cast<void*>(scratch), const_cast<float*>(result)))
    return rc;
  if (!upstream || !df || ((uintptr_t)df & 15)) { set_error("trase_b200_contrastive_soft_backward: null / unaligned pointer"); return 1; }
  p.upstream = upstream; p.df = df;
  const int nb = (N + contrastive::CT - 1) / contrastive::CT;
  contrastive::contrastive_kernel<contrastive::MODE_GRAD><<<nb, contrastive::CT_THREADS, 0,
                                                           reinterpret_cast<cudaStream_t>(stream)>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_contrastive_soft_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
