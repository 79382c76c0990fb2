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
// so three sweeps over the pair space recompute them on the fly.  A block owns 64 pixels and walks 64-pixel tiles of
// partners (the tiles of one owner are split over a few blocks so that two waves of the SMs are busy); similarities
// come from channel-major shared-memory tiles with two 16-byte reads per channel for a 4 x 4 patch per thread:
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

// Pair weight (feature_utils.py:31-37):  (clamp(top / (a_i a_j), 1) - 1) / (hi - 1) * 9 + 1, top = (max a)^2,
// hi = clamp(top / (min a)^2, 1) (the minimum of the clamped ratio over all pairs is exactly 1: the largest a with
// itself).  Evaluated as (max(ra_i * ra_j, 1) - 1) * k9 + 1 with ra = max a / a per pixel and k9 = 9 / (hi - 1): no
// division per pair (differs from the reference's rounding by an ulp or two; all-equal areas give hi = 1, k9 = inf
// and NaN weights, as in the reference).
struct WeightCtx {
  float amax, k9;
  bool on;
};
__device__ __forceinline__ WeightCtx weight_ctx(const Params& p) {
  WeightCtx w;
  w.on = p.use_weights != 0;
  w.amax = 1.f;
  w.k9 = 0.f;
  if (w.on) {
    const float amin = p.a_minmax[0], amax = p.a_minmax[1];
    float pr = amin * amin;
    if (pr == 0.f) pr = 1e10f;
    const float hi = fmaxf((amax * amax) / pr, 1.0f);
    w.amax = amax;
    w.k9 = 9.0f / (hi - 1.0f);
  }
  return w;
}
__device__ __forceinline__ float pixel_ratio(const WeightCtx& w, const float* a, int idx, int N) {
  if (!w.on || idx >= N) return 1.f;
  const float v = a[idx];
  return v > 0.f ? w.amax / v : 0.f;          // a == 0: the reference replaces the product by 1e10 -> ratio clamps to 1
}
__device__ __forceinline__ float pair_weight(const WeightCtx& w, float rai, float raj) {
  if (!w.on) return 1.f;
  return (fmaxf(rai * raj, 1.0f) - 1.0f) * w.k9 + 1.0f;
}

// acc = a * b + acc on two packed fp32 lanes (FFMA2); bb holds the second factor of both lanes
__device__ __forceinline__ void ffma2(float2& acc, const float2 a, const unsigned long long bb) {
  unsigned long long av, cv;
  asm("mov.b64 %0, {%1, %2};" : "=l"(av) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(cv) : "f"(acc.x), "f"(acc.y));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(cv) : "l"(av), "l"(bb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(acc.x), "=f"(acc.y) : "l"(cv));
}

enum { MODE_FLAGS = 0, MODE_LOSS = 1, MODE_GRAD = 2 };
constexpr int LD = CT + 4;         // row pitch of the channel-major tiles (float4 reads of 4 consecutive pixels)

// grid = (own 64-pixel blocks, partner splits): block (x, y) pairs its own pixels with partner tiles y, y + S, ...
// NW = mask words per pixel the kernel is built for: 1 (up to 64 sampled masks, the usual case: a quarter of the bit
// logic per pair) or MAXW.
template <int MODE, int NW>
__global__ void __launch_bounds__(CT_THREADS) contrastive_kernel(const Params p) {
  pdl_wait();
  // channel-major: s[k][pixel].  The own tile holds every value TWICE, (v, v) side by side: a 16-byte read is then two
  // own pixels as ready-made packed operands of FFMA2 (two partner pixels per instruction), with no register moves
  __shared__ __align__(16) float2 s_own[CC][LD];
  __shared__ __align__(16) float s_oth[CC][LD];
  __shared__ unsigned long long s_bown[CT][NW], s_both[CT][NW];
  __shared__ float s_raown[CT], s_raoth[CT];
  __shared__ unsigned char s_kown[2][CT], s_koth[2][CT];
  __shared__ __align__(16) float s_Gt[MODE == MODE_GRAD ? CT : 1][LD];   // G transposed: s_Gt[partner][own]
  __shared__ double s_red[4][CT_THREADS / 32];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int own0 = blockIdx.x * CT;
  const WeightCtx wc = weight_ctx(p);
  for (int e = tid; e < CT * CC; e += CT_THREADS) {
    const int r = e / CC, c = e - r * CC;            // coalesced global read, transposed shared write
    const float v = own0 + r < p.N ? p.f[(size_t)(own0 + r) * CC + c] : 0.f;
    s_own[c][r] = make_float2(v, v);
  }
  for (int e = tid; e < CT * NW; e += CT_THREADS) {
    const int r = e / NW, w = e - r * NW;
    s_bown[r][w] = (own0 + r < p.N && w < p.nwords) ? p.bits[(size_t)(own0 + r) * p.nwords + w] : 0ull;
  }
  if (tid < CT) {
    s_raown[tid] = pixel_ratio(wc, p.a, own0 + tid, p.N);
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

  unsigned flag_pos = 0, flag_neg = 0;                    // FLAGS: per own row (bit a of the thread's 4 rows)
  float sum_pos = 0.f, sum_neg = 0.f, cnt_pos = 0.f, cnt_neg = 0.f;   // per thread: at most ~1000 terms each
  const int rg = (tid >> 4) * 4, cp = (tid & 15) * 2;     // GRAD phase 2: 4 own rows x 2 channels per thread
  float dacc[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};

  const int ntiles = (p.N + CT - 1) / CT;
  // LOSS only needs partners i < j: tiles up to the own one
  const int last_tile = MODE == MODE_LOSS ? (int)blockIdx.x : ntiles - 1;
  for (int t = blockIdx.y; t <= last_tile; t += gridDim.y) {
    const int oth0 = t * CT;
    __syncthreads();
    for (int e = tid; e < CT * CC; e += CT_THREADS) {
      const int r = e / CC, c = e - r * CC;
      s_oth[c][r] = oth0 + r < p.N ? p.f[(size_t)(oth0 + r) * CC + c] : 0.f;
    }
    for (int e = tid; e < CT * NW; e += CT_THREADS) {
      const int r = e / NW, w = e - r * NW;
      s_both[r][w] = (oth0 + r < p.N && w < p.nwords) ? p.bits[(size_t)(oth0 + r) * p.nwords + w] : 0ull;
    }
    if (tid < CT) {
      s_raoth[tid] = pixel_ratio(wc, p.a, oth0 + tid, p.N);
      if (MODE == MODE_GRAD) {
        s_koth[0][tid] = oth0 + tid < p.N ? p.kept[oth0 + tid] : 0;
        s_koth[1][tid] = oth0 + tid < p.N ? p.kept[p.N + oth0 + tid] : 0;
      }
    }
    __syncthreads();
    // 4 x 4 patch of similarities: own pixels ty*4.., partner pixels tx*4..; two 16-byte reads per channel
    // c2[a][h] = similarities of own pixel a with partner pixels 2h, 2h + 1 (each lane of the packed FMA is the same
    // fp32 fma, in the same order over k, as the scalar form: identical bits)
    float2 c2[4][2];
#pragma unroll
    for (int a = 0; a < 4; a++) c2[a][0] = c2[a][1] = make_float2(0.f, 0.f);
#pragma unroll 8
    for (int k = 0; k < CC; k++) {
      const ulonglong2 o01 = *reinterpret_cast<const ulonglong2*>(&s_own[k][ty * 4]);
      const ulonglong2 o23 = *reinterpret_cast<const ulonglong2*>(&s_own[k][ty * 4 + 2]);
      const float4 q4 = *reinterpret_cast<const float4*>(&s_oth[k][tx * 4]);
      const unsigned long long oo[4] = {o01.x, o01.y, o23.x, o23.y};
      const float2 q01 = make_float2(q4.x, q4.y), q23 = make_float2(q4.z, q4.w);
#pragma unroll
      for (int a = 0; a < 4; a++) {
        ffma2(c2[a][0], q01, oo[a]);
        ffma2(c2[a][1], q23, oo[a]);
      }
    }
    float c[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++) {
      c[a][0] = c2[a][0].x; c[a][1] = c2[a][0].y; c[a][2] = c2[a][1].x; c[a][3] = c2[a][1].y;
    }
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int rq = tx * 4 + b, oth = oth0 + rq;
      float gcol[4];
#pragma unroll
      for (int a = 0; a < 4; a++) {
        const int ro = ty * 4 + a, own = own0 + ro;
        const bool valid = own < p.N && oth < p.N;
        bool same = false;
#pragma unroll
        for (int w = 0; w < NW; w++) same = same || ((s_bown[ro][w] & s_both[rq][w]) != 0ull);
        const float cv = c[a][b];
        if (MODE == MODE_FLAGS) {
          if (valid && same && cv < p.pos_th) flag_pos |= 1u << a;
          if (valid && !same && cv > p.neg_th) flag_neg |= 1u << a;
        } else if (MODE == MODE_LOSS) {
          // own = column j, partner = row i < j
          if (valid && oth < own) {
            const float w = pair_weight(wc, s_raown[ro], s_raoth[rq]);
            if (same && s_kown[0][ro]) { sum_pos -= w * cv; cnt_pos += 1.f; }
            if (!same && s_kown[1][ro]) { sum_neg += w * fmaxf(cv, 0.f); cnt_neg += 1.f; }
          }
        } else {
          float g = 0.f;
          if (valid && own != oth) {
            const bool own_is_col = own > oth;
            const unsigned char kp = own_is_col ? s_kown[0][ro] : s_koth[0][rq];
            const unsigned char kn = own_is_col ? s_kown[1][ro] : s_koth[1][rq];
            const float w = pair_weight(wc, s_raown[ro], s_raoth[rq]);
            if (same && kp) g = -w * inv_npos;
            if (!same && kn && cv > 0.f) g = w * inv_nneg;
          }
          gcol[a] = g;
        }
      }
      if (MODE == MODE_GRAD)
        *reinterpret_cast<float4*>(&s_Gt[rq][ty * 4]) = make_float4(gcol[0], gcol[1], gcol[2], gcol[3]);
    }
    if (MODE == MODE_GRAD) {
      __syncthreads();
      // d f_own[r][ch] += sum_j G[r][j] f_j[ch]: 4 own rows x 2 channels per thread
#pragma unroll 4
      for (int j = 0; j < CT; j++) {
        const float4 g4 = *reinterpret_cast<const float4*>(&s_Gt[j][rg]);
        const float f0 = s_oth[cp][j], f1 = s_oth[cp + 1][j];
        dacc[0][0] = fmaf(g4.x, f0, dacc[0][0]); dacc[0][1] = fmaf(g4.x, f1, dacc[0][1]);
        dacc[1][0] = fmaf(g4.y, f0, dacc[1][0]); dacc[1][1] = fmaf(g4.y, f1, dacc[1][1]);
        dacc[2][0] = fmaf(g4.z, f0, dacc[2][0]); dacc[2][1] = fmaf(g4.z, f1, dacc[2][1]);
        dacc[3][0] = fmaf(g4.w, f0, dacc[3][0]); dacc[3][1] = fmaf(g4.w, f1, dacc[3][1]);
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
        // the flag arrays start at zero; blocks of different partner splits only ever write 1
        if (own < p.N && ((flag_pos >> a) & 1u)) p.kept[own] = 1;
        if (own < p.N && ((flag_neg >> a) & 1u)) p.kept[p.N + own] = 1;
      }
    }
  } else if (MODE == MODE_LOSS) {
    double v[4] = {(double)sum_pos, (double)sum_neg, (double)cnt_pos, (double)cnt_neg};
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
      p.partial[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 4 + tid] = tot;
    }
  } else {
#pragma unroll
    for (int a = 0; a < 4; a++) {
      const int own = own0 + rg + a;
      if (own < p.N) {
        float* dst = p.df + (size_t)own * CC + cp;
        if (gridDim.y == 1) { dst[0] = dacc[a][0]; dst[1] = dacc[a][1]; }
        else { atomicAdd(dst, dacc[a][0]); atomicAdd(dst + 1, dacc[a][1]); }     // df arrives zero-filled
      }
    }
  }
}

// n_kept = sum_j kept[j] * j (the number of rows i < j of each kept column), loss = sum / n_kept
__global__ void __launch_bounds__(256) contrastive_finish_kernel(const Params p, int nblocks) {
  pdl_wait();
  __shared__ double s_v[6][8];
  double v[6] = {0, 0, 0, 0, 0, 0};            // n_pos, n_neg, pos_sum, neg_sum, pos_pairs, neg_pairs
  for (int j = threadIdx.x; j < p.N; j += 256) {
    if (p.kept[j]) v[0] += (double)j;
    if (p.kept[p.N + j]) v[1] += (double)j;
  }
  for (int b = threadIdx.x; b < nblocks; b += 256)
#pragma unroll
    for (int q = 0; q < 4; q++) v[2 + q] += p.partial[(size_t)b * 4 + q];
#pragma unroll
  for (int q = 0; q < 6; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    if ((threadIdx.x & 31) == 0) s_v[q][threadIdx.x >> 5] = v[q];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t[6];
    for (int q = 0; q < 6; q++) {
      t[q] = 0;
      for (int w = 0; w < 8; w++) t[q] += s_v[q][w];
    }
    p.result[0] = t[4] > 0 ? (float)(t[2] / t[0]) : 0.f;
    p.result[1] = t[5] > 0 ? (float)(t[3] / t[1]) : 0.f;
    p.result[2] = (float)t[0];
    p.result[3] = (float)t[1];
    p.result[4] = (float)t[4];
    p.result[5] = (float)t[5];
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
  pdl_wait();
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

// partner splits per own block: enough blocks for two waves of the 148 SMs
static int contrastive_splits(int nb) {
  int s = (2 * 148 + nb - 1) / nb;
  return s < 1 ? 1 : (s > 8 ? 8 : s);
}
extern "C" TRASE_API size_t trase_b200_contrastive_scratch_bytes(int N) {
  const size_t nb = (size_t)((N + contrastive::CT - 1) / contrastive::CT);
  return align_up(2 * (size_t)(N > 0 ? N : 0), 256) + align_up(nb * 8 * 4 * sizeof(double), 256);
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
  launch_k(contrastive::pixel_mask_info_kernel, (N + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
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
  const dim3 grid(nb, contrastive_splits(nb));
  cudaError_t e = cudaMemsetAsync(p.kept, 0, 2 * (size_t)N, s);           // the mined-column flags are OR-ed into
  if (e != cudaSuccess) { set_error("trase_b200_contrastive_soft_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  if (nwords == 1) {
    launch_k(contrastive::contrastive_kernel<contrastive::MODE_FLAGS, 1>, grid, contrastive::CT_THREADS, 0, s, p);
    launch_k(contrastive::contrastive_kernel<contrastive::MODE_LOSS, 1>, grid, contrastive::CT_THREADS, 0, s, p);
  } else {
    launch_k(contrastive::contrastive_kernel<contrastive::MODE_FLAGS, contrastive::MAXW>, grid, contrastive::CT_THREADS, 0, s, p);
    launch_k(contrastive::contrastive_kernel<contrastive::MODE_LOSS, contrastive::MAXW>, grid, contrastive::CT_THREADS, 0, s, p);
  }
  launch_k(contrastive::contrastive_finish_kernel, 1, 256, 0, s, p, (int)(grid.x * grid.y));
  e = cudaGetLastError();
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
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const dim3 grid(nb, contrastive_splits(nb));
  cudaError_t e = cudaMemsetAsync(df, 0, (size_t)N * contrastive::CC * sizeof(float), s);   // split blocks add into it
  if (e == cudaSuccess) {
    if (nwords == 1) launch_k(contrastive::contrastive_kernel<contrastive::MODE_GRAD, 1>, grid, contrastive::CT_THREADS, 0, s, p);
    else launch_k(contrastive::contrastive_kernel<contrastive::MODE_GRAD, contrastive::MAXW>, grid, contrastive::CT_THREADS, 0, s, p);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) { set_error("trase_b200_contrastive_soft_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
