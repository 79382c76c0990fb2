// Fused parameter activation + rasterizer-input assembly, forward and backward
// (SURVEY.md §8(f) rank 1).  Replaces the ~20 elementwise torch kernels (and as
// many again in autograd) that the reference's glue launches per view:
//   means3D  = xyz + d_xyz                                   gaussian_renderer/__init__.py:82
//   scales   = exp(_scaling) + d_scaling                      :95, scene/gaussian_model.py:183
//   rotations= normalize(_rotation) + d_rotation              :96, scene/gaussian_model.py:187
//   opacity  = sigmoid(_opacity)                              scene/gaussian_model.py:201
//   shs      = cat(f_dc, f_rest, dim=1)                        scene/gaussian_model.py:195-197
//   sh_objs  = feat / (||feat|| + 1e-9)                        gaussian_renderer/__init__.py:116-121
// One thread per Gaussian for the small vectors; the 48-float SH row and the
// 32-float feature row are handled by 4 / 8 cooperating lanes with 16-byte accesses.
#include "common.cuh"

namespace tb {

struct AssembleFwd {
  int P, Mrest;   // Mrest = SH coefficients in f_rest (15 for degree 3)
  const float *xyz, *d_xyz, *scaling, *d_scaling, *rotation, *d_rotation, *opacity, *f_dc, *f_rest, *feat;
  float *means3D, *scales, *rotations, *opac_out, *shs, *sh_objs;
};

__global__ void __launch_bounds__(256) assemble_fwd_kernel(const AssembleFwd a) {
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.P) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      a.means3D[3 * i + c] = a.xyz[3 * i + c] + a.d_xyz[3 * i + c];
      a.scales[3 * i + c] = expf(a.scaling[3 * i + c]) + a.d_scaling[3 * i + c];
    }
    const float4 q = reinterpret_cast<const float4*>(a.rotation)[i];
    const float4 dq = reinterpret_cast<const float4*>(a.d_rotation)[i];
    // F.normalize: v / max(||v||, 1e-12)
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    reinterpret_cast<float4*>(a.rotations)[i] = make_float4(q.x / n + dq.x, q.y / n + dq.y, q.z / n + dq.z, q.w / n + dq.w);
    a.opac_out[i] = 1.f / (1.f + expf(-a.opacity[i]));
  }
}

// SH rows: out[i][0] = f_dc[i][0], out[i][1..] = f_rest[i][..]; flat float copy, coalesced
__global__ void __launch_bounds__(256) assemble_sh_kernel(int P, int Mrest, const float* __restrict__ f_dc,
                                                          const float* __restrict__ f_rest, float* __restrict__ shs) {
  pdl_wait();
  const int M3 = (Mrest + 1) * 3;
  const size_t total = (size_t)P * M3;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t i = e / M3;
    const int c = (int)(e - i * M3);
    shs[e] = c < 3 ? f_dc[i * 3 + c] : f_rest[i * (size_t)(Mrest * 3) + (c - 3)];
  }
}

// feature rows: 8 lanes per Gaussian, float4 each; norm by shuffle over the 8 lanes
__global__ void __launch_bounds__(256) assemble_feat_kernel(int P, const float4* __restrict__ feat,
                                                            float4* __restrict__ out) {
  pdl_wait();
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < (size_t)P * 8;
  float4 f = valid ? feat[e] : make_float4(0, 0, 0, 0);
  float s = f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  const float inv = 1.f / (sqrtf(s) + 1e-9f);
  if (valid) out[e] = make_float4(f.x * inv, f.y * inv, f.z * inv, f.w * inv);
}

struct AssembleBwd {
  int P, Mrest;
  const float *scaling, *rotation, *opacity, *feat;                          // raw parameters
  const float *g_means3D, *g_scales, *g_rotations, *g_opac, *g_shs, *g_objs; // upstream gradients (may be null)
  float *g_xyz, *g_dxyz, *g_scaling, *g_dscaling, *g_rotation, *g_drotation, *g_opacity, *g_fdc, *g_frest, *g_feat;
  int acc;   // bit 0 xyz, 1 scaling, 2 rotation, 3 opacity, 4 SH: that PARAMETER gradient is added to, not overwritten
};

__global__ void __launch_bounds__(256) assemble_bwd_kernel(const AssembleBwd a) {
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.P) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float gm = a.g_means3D ? a.g_means3D[3 * i + c] : 0.f;
      a.g_xyz[3 * i + c] = (a.acc & 1) ? a.g_xyz[3 * i + c] + gm : gm;
      a.g_dxyz[3 * i + c] = gm;
      const float gs = a.g_scales ? a.g_scales[3 * i + c] : 0.f;
      const float gsc = gs * expf(a.scaling[3 * i + c]);
      a.g_scaling[3 * i + c] = (a.acc & 2) ? a.g_scaling[3 * i + c] + gsc : gsc;
      a.g_dscaling[3 * i + c] = gs;
    }
    const float4 q = reinterpret_cast<const float4*>(a.rotation)[i];
    const float4 g = a.g_rotations ? reinterpret_cast<const float4*>(a.g_rotations)[i] : make_float4(0, 0, 0, 0);
    const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float4 gq;
    if (n > 1e-12f) {
      // d/dq (q/|q|) : (g - r (r.g)) / |q| with r = q/|q|
      const float inv = 1.f / n;
      const float rx = q.x * inv, ry = q.y * inv, rz = q.z * inv, rw = q.w * inv;
      const float rg = rx * g.x + ry * g.y + rz * g.z + rw * g.w;
      gq = make_float4((g.x - rx * rg) * inv, (g.y - ry * rg) * inv, (g.z - rz * rg) * inv, (g.w - rw * rg) * inv);
    } else {
      gq = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);   // clamped branch of F.normalize
    }
    if (a.acc & 4) {
      const float4 o = reinterpret_cast<float4*>(a.g_rotation)[i];
      gq = make_float4(o.x + gq.x, o.y + gq.y, o.z + gq.z, o.w + gq.w);
    }
    reinterpret_cast<float4*>(a.g_rotation)[i] = gq;
    reinterpret_cast<float4*>(a.g_drotation)[i] = g;
    const float sgm = 1.f / (1.f + expf(-a.opacity[i]));
    const float gop = (a.g_opac ? a.g_opac[i] : 0.f) * sgm * (1.f - sgm);
    a.g_opacity[i] = (a.acc & 8) ? a.g_opacity[i] + gop : gop;
  }
}

// block (64, 4): x = coefficient*3 + channel of one Gaussian's SH row (at most 48 used), y = one of 4 rows per trip
__global__ void __launch_bounds__(256) assemble_sh_bwd_kernel(int P, int Mrest, const float* __restrict__ g_shs,
                                                              float* __restrict__ g_fdc, float* __restrict__ g_frest,
                                                              int accumulate) {
  pdl_wait();
  const int M3 = (Mrest + 1) * 3, R3 = Mrest * 3;
  const int c = threadIdx.x;
  if (c >= M3) return;
  // eight rows per thread in flight: the loop is a chain of dependent load / load / store otherwise
  constexpr int U = 8;
  const int step = gridDim.x * 4;
  for (int i0 = blockIdx.x * 4 + threadIdx.y; i0 < P; i0 += step * U) {
    float v[U], o[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * step;
      v[u] = (g_shs && i < P) ? g_shs[(size_t)i * M3 + c] : 0.f;
      o[u] = 0.f;
      if (accumulate && i < P) o[u] = c < 3 ? g_fdc[(size_t)i * 3 + c] : g_frest[(size_t)i * R3 + (c - 3)];
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = i0 + u * step;
      if (i < P) {
        float* dst = c < 3 ? g_fdc + (size_t)i * 3 + c : g_frest + (size_t)i * R3 + (c - 3);
        *dst = o[u] + v[u];
      }
    }
  }
}

// y = f / (n + eps):  df = g / (n + eps) - f (f.g) / (n (n + eps)^2)
__global__ void __launch_bounds__(256) assemble_feat_bwd_kernel(int P, const float4* __restrict__ feat,
                                                                const float4* __restrict__ g_objs,
                                                                float4* __restrict__ g_feat) {
  pdl_wait();
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < (size_t)P * 8;
  const float4 f = valid ? feat[e] : make_float4(0, 0, 0, 0);
  const float4 g = (valid && g_objs) ? g_objs[e] : make_float4(0, 0, 0, 0);
  float s = f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
  float d = f.x * g.x + f.y * g.y + f.z * g.z + f.w * g.w;
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    d += __shfl_xor_sync(0xffffffffu, d, o);
  }
  const float n = sqrtf(s);
  const float inv = 1.f / (n + 1e-9f);
  const float k = (n > 0.f) ? d * inv * inv / n : 0.f;
  if (valid) g_feat[e] = make_float4(g.x * inv - f.x * k, g.y * inv - f.y * k, g.z * inv - f.z * k, g.w * inv - f.w * k);
}


// ---------------------------------------------------------------------------
// kNN-smoothed features (scene/gaussian_model.py:79-104): mean over a subset of each Gaussian's K nearest
// neighbours of their L2-normalised (F.normalize, eps 1e-12) feature rows.  One warp per Gaussian, lane = channel;
// a neighbour row is one coalesced 128-byte read.  The neighbour index map is the caller's (the search itself is
// outside this path); `sel` lists which of the map's K columns take part (the reference draws int(K*dropout) of
// them per call with torch.randperm).
constexpr int SMOOTH_MAX_SEL = 32;
struct SmoothSel { int n; unsigned char col[SMOOTH_MAX_SEL]; };

__global__ void __launch_bounds__(256) smooth_feat_fwd_kernel(int P, int K, SmoothSel sel,
                                                              const float* __restrict__ feat,
                                                              const long long* __restrict__ idx, float* __restrict__ out) {
  pdl_wait();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  float acc = 0.f;
  for (int k = 0; k < sel.n; k++) {
    const long long j = __ldg(idx + (size_t)p * K + sel.col[k]);
    const float v = __ldg(feat + (size_t)j * NUM_OBJECTS + lane);
    float n2 = v * v;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    acc += v / fmaxf(sqrtf(n2), 1e-12f);
  }
  out[(size_t)p * NUM_OBJECTS + lane] = acc / (float)sel.n;
}

// Scatter of the mean's gradient back through each neighbour's normalisation: d/dv (v/|v|) = (g - u (u.g)) / |v|.
__global__ void __launch_bounds__(256) smooth_feat_bwd_kernel(int P, int K, SmoothSel sel,
                                                              const float* __restrict__ feat,
                                                              const long long* __restrict__ idx,
                                                              const float* __restrict__ g_out, float* __restrict__ g_feat) {
  pdl_wait();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  const float g = g_out[(size_t)p * NUM_OBJECTS + lane] / (float)sel.n;
  for (int k = 0; k < sel.n; k++) {
    const long long j = __ldg(idx + (size_t)p * K + sel.col[k]);
    const float v = __ldg(feat + (size_t)j * NUM_OBJECTS + lane);
    float n2 = v * v;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    const float n = sqrtf(n2);
    float gv;
    if (n > 1e-12f) {
      const float u = v / n;
      float ug = u * g;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ug += __shfl_xor_sync(0xffffffffu, ug, o);
      gv = (g - u * ug) / n;
    } else {
      gv = g * 1e12f;                       // clamped branch of F.normalize
    }
    atomicAdd(g_feat + (size_t)j * NUM_OBJECTS + lane, gv);
  }
}

// The same two passes with each feature row normalised ONCE instead of once per use:
//   forward : unit[j] = normalize(feat[j]) (smooth_unit_kernel), then out[p] = mean_k unit[idx[p][sel_k]] — the
//             neighbour loads of a Gaussian are independent of each other (no per-neighbour norm reduction);
//   backward: the normalisation is linear in its upstream gradient, so the mean's gradient is first summed per
//             neighbour, G[j] = (1/n) sum over the Gaussians p that list j in a selected column of g_out[p], and the
//             normalisation backward is applied once per j.  The sum is a GATHER over the reverse of the neighbour map
//             (rev_ptr [P+1], rev_ent [P*K]: the entries (p << 8 | column) of row j, built once per map by the
//             caller) — no atomics, no memset, every g_feat row written exactly once.
__global__ void __launch_bounds__(256) smooth_unit_kernel(int P, const float* __restrict__ feat, float* __restrict__ unit) {
  pdl_wait();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  const float v = feat[(size_t)p * NUM_OBJECTS + lane];
  float n2 = v * v;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
  unit[(size_t)p * NUM_OBJECTS + lane] = v / fmaxf(sqrtf(n2), 1e-12f);
}
__global__ void __launch_bounds__(256) smooth_gather_fwd_kernel(int P, int K, SmoothSel sel,
                                                                const float* __restrict__ unit,
                                                                const long long* __restrict__ idx, float* __restrict__ out) {
  pdl_wait();
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  // lane k fetches the k-th selected neighbour's index; the rows are then read with all loads in flight together
  const long long mine = lane < sel.n ? __ldg(idx + (size_t)p * K + sel.col[lane]) : 0;
  float acc = 0.f;
#pragma unroll 8
  for (int k = 0; k < sel.n; k++) {
    const long long j = __shfl_sync(0xffffffffu, mine, k);
    acc += __ldg(unit + (size_t)j * NUM_OBJECTS + lane);      // same order of additions as the per-use kernel
  }
  out[(size_t)p * NUM_OBJECTS + lane] = acc / (float)sel.n;
}
__global__ void __launch_bounds__(256) smooth_feat_bwd_csr_kernel(int P, unsigned long long colmask, float inv_n,
                                                                  const float* __restrict__ feat,
                                                                  const int* __restrict__ rev_ptr,
                                                                  const int* __restrict__ rev_ent,
                                                                  const float* __restrict__ g_out,
                                                                  float* __restrict__ g_feat) {
  pdl_wait();
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= P) return;
  const int b = __ldg(rev_ptr + j), e = __ldg(rev_ptr + j + 1);
  float acc0 = 0.f, acc1 = 0.f;
  for (int t = b; t < e; t += 32) {
    const int ent = (t + lane < e) ? __ldg(rev_ent + t + lane) : -1;
    // entries whose column takes part in this call
    unsigned live = __ballot_sync(0xffffffffu, ent >= 0 && ((colmask >> (ent & 255)) & 1ull));
    while (live) {
      const int q0 = __ffs(live) - 1;
      live &= live - 1;
      const int p0 = __shfl_sync(0xffffffffu, ent, q0) >> 8;
      acc0 += __ldg(g_out + (size_t)p0 * NUM_OBJECTS + lane);
      if (live) {
        const int q1 = __ffs(live) - 1;
        live &= live - 1;
        const int p1 = __shfl_sync(0xffffffffu, ent, q1) >> 8;
        acc1 += __ldg(g_out + (size_t)p1 * NUM_OBJECTS + lane);
      }
    }
  }
  const float g = (acc0 + acc1) * inv_n;
  const float v = feat[(size_t)j * NUM_OBJECTS + lane];
  float n2 = v * v;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
  const float n = sqrtf(n2);
  float gv;
  if (n > 1e-12f) {
    const float u = v / n;
    float ug = u * g;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ug += __shfl_xor_sync(0xffffffffu, ug, o);
    gv = (g - u * ug) / n;
  } else {
    gv = g * 1e12f;                       // clamped branch of F.normalize
  }
  g_feat[(size_t)j * NUM_OBJECTS + lane] = gv;
}

static bool smooth_args_ok(const char* who, int P, int K, const int* sel_cols, int n_sel, const void* a, const void* b,
                           const void* c, SmoothSel& sel) {
  if (!a || !b || !c || !sel_cols || K <= 0 || n_sel <= 0 || n_sel > SMOOTH_MAX_SEL || K > 255) {
    set_error("%s: bad arguments (K=%d, n_sel=%d; at most %d selected columns)", who, K, n_sel, SMOOTH_MAX_SEL);
    return false;
  }
  sel.n = n_sel;
  for (int i = 0; i < n_sel; i++) {
    if (sel_cols[i] < 0 || sel_cols[i] >= K) { set_error("%s: selected column %d outside [0, %d)", who, sel_cols[i], K); return false; }
    sel.col[i] = (unsigned char)sel_cols[i];
  }
  (void)P;
  return true;
}

}  // namespace tb

using namespace tb;

extern "C" TRASE_API int trase_b200_smooth_features_forward(int P, int K, const int* sel_cols, int n_sel,
                                                            const float* feat, const long long* knn_idx, float* out,
                                                            float* unit_scratch, void* stream) {
  clear_error();
  if (P <= 0) return 0;
  SmoothSel sel;
  if (!smooth_args_ok("trase_b200_smooth_features_forward", P, K, sel_cols, n_sel, feat, knn_idx, out, sel)) return 1;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = (int)(((size_t)P * 32 + 255) / 256);
  if (unit_scratch) {
    launch_k(smooth_unit_kernel, grid, 256, 0, s, P, feat, unit_scratch);
    launch_k(smooth_gather_fwd_kernel, grid, 256, 0, s, P, K, sel, unit_scratch, knn_idx, out);
  } else {
    launch_k(smooth_feat_fwd_kernel, grid, 256, 0, s, P, K, sel, feat, knn_idx, out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_smooth_features_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(unit_scratch ? 2 : 1);
  return 0;
}

extern "C" TRASE_API int trase_b200_smooth_features_backward_csr(int P, int K, const int* sel_cols, int n_sel,
                                                                 const float* feat, const int* rev_ptr,
                                                                 const int* rev_ent, const float* g_out, float* g_feat,
                                                                 void* stream) {
  clear_error();
  if (P <= 0) return 0;
  SmoothSel sel;
  if (!g_out || !rev_ent || K > 64 || P >= (1 << 23) ||
      !smooth_args_ok("trase_b200_smooth_features_backward_csr", P, K, sel_cols, n_sel, feat, rev_ptr, g_feat, sel)) {
    if (!g_out || !rev_ent) set_error("trase_b200_smooth_features_backward_csr: null pointer");
    else if (K > 64 || P >= (1 << 23)) set_error("trase_b200_smooth_features_backward_csr: K <= 64 and P < 2^23 only");
    return 1;
  }
  unsigned long long colmask = 0;
  for (int i = 0; i < sel.n; i++) {
    if ((colmask >> sel.col[i]) & 1ull) {
      set_error("trase_b200_smooth_features_backward_csr: column %d selected twice", (int)sel.col[i]);
      return 1;
    }
    colmask |= 1ull << sel.col[i];
  }
  launch_k(smooth_feat_bwd_csr_kernel, (int)(((size_t)P * 32 + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      P, colmask, 1.f / (float)sel.n, feat, rev_ptr, rev_ent, g_out, g_feat);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_smooth_features_backward_csr: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API int trase_b200_smooth_features_backward(int P, int K, const int* sel_cols, int n_sel,
                                                             const float* feat, const long long* knn_idx,
                                                             const float* g_out, float* g_feat, void* stream) {
  clear_error();
  if (P <= 0) return 0;
  SmoothSel sel;
  if (!g_out || !smooth_args_ok("trase_b200_smooth_features_backward", P, K, sel_cols, n_sel, feat, knn_idx, g_feat, sel)) {
    if (!g_out) set_error("trase_b200_smooth_features_backward: null pointer");
    return 1;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(g_feat, 0, (size_t)P * NUM_OBJECTS * 4, s);      // accumulated into below
  if (e == cudaSuccess) {
    launch_k(smooth_feat_bwd_kernel, (int)(((size_t)P * 32 + 255) / 256), 256, 0, s, P, K, sel, feat, knn_idx, g_out, g_feat);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) { set_error("trase_b200_smooth_features_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}


extern "C" TRASE_API int trase_b200_assemble_forward(int P, int Mrest, const float* xyz, const float* d_xyz,
                                                     const float* scaling, const float* d_scaling,
                                                     const float* rotation, const float* d_rotation,
                                                     const float* opacity, const float* f_dc, const float* f_rest,
                                                     const float* feat, float* means3D, float* scales,
                                                     float* rotations, float* opac_out, float* shs, float* sh_objs,
                                                     void* stream) {
  if (P <= 0) return 0;
  // shs == NULL: the caller's f_dc / f_rest already alias one [P][Mrest+1][3] block, no SH copy is made
  if (!xyz || !d_xyz || !scaling || !d_scaling || !rotation || !d_rotation || !opacity || !feat ||
      !means3D || !scales || !rotations || !opac_out || !sh_objs || Mrest < 0 ||
      (shs && (!f_dc || (Mrest > 0 && !f_rest)))) {
    set_error("trase_b200_assemble_forward: null pointer");
    return 1;
  }
  if (((uintptr_t)rotation | (uintptr_t)d_rotation | (uintptr_t)rotations | (uintptr_t)feat | (uintptr_t)sh_objs) & 15) {
    set_error("trase_b200_assemble_forward: rotation and feature buffers must be 16-byte aligned");
    return 1;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  AssembleFwd a{P, Mrest, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, f_dc, f_rest, feat,
                means3D, scales, rotations, opac_out, shs, sh_objs};
  launch_k(assemble_fwd_kernel, (P + 255) / 256, 256, 0, s, a);
  const size_t tot = (size_t)P * (Mrest + 1) * 3;
  int g = (int)((tot + 255) / 256 < 148 * 32 ? (tot + 255) / 256 : 148 * 32);
  if (shs) launch_k(assemble_sh_kernel, g, 256, 0, s, P, Mrest, f_dc, f_rest, shs);
  const int n_launched_f = shs ? 3 : 2;
  launch_k(assemble_feat_kernel, (int)(((size_t)P * 8 + 255) / 256), 256, 0, s, P, reinterpret_cast<const float4*>(feat),
                                                                       reinterpret_cast<float4*>(sh_objs));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_assemble_forward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(n_launched_f);
  return 0;
}

extern "C" TRASE_API int trase_b200_assemble_backward(int P, int Mrest, const float* scaling, const float* rotation,
                                                      const float* opacity, const float* feat, const float* g_means3D,
                                                      const float* g_scales, const float* g_rotations,
                                                      const float* g_opac, const float* g_shs, const float* g_objs,
                                                      float* g_xyz, float* g_dxyz, float* g_scaling, float* g_dscaling,
                                                      float* g_rotation, float* g_drotation, float* g_opacity,
                                                      float* g_fdc, float* g_frest, float* g_feat, int accumulate_mask,
                                                      void* stream) {
  if (P <= 0) return 0;
  // the seven geometry gradients come together or not at all (all NULL: every geometry parameter is frozen, FEATURE state)
  const bool geom = g_xyz || g_dxyz || g_scaling || g_dscaling || g_rotation || g_drotation || g_opacity;
  if (!scaling || !rotation || !opacity || !feat ||
      (geom && (!g_xyz || !g_dxyz || !g_scaling || !g_dscaling || !g_rotation || !g_drotation || !g_opacity)) ||
      (g_fdc && Mrest > 0 && !g_frest)) {
    set_error("trase_b200_assemble_backward: null pointer");
    return 1;
  }
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  AssembleBwd a{P, Mrest, scaling, rotation, opacity, feat, g_means3D, g_scales, g_rotations, g_opac, g_shs, g_objs,
                g_xyz, g_dxyz, g_scaling, g_dscaling, g_rotation, g_drotation, g_opacity, g_fdc, g_frest, g_feat,
                accumulate_mask};
  if (geom) launch_k(assemble_bwd_kernel, (P + 255) / 256, 256, 0, s, a);
  if (g_fdc) {
    if ((Mrest + 1) * 3 > 64) { set_error("trase_b200_assemble_backward: at most 21 SH coefficients"); return 1; }
    const int gb = (P + 3) / 4 < 148 * 16 ? (P + 3) / 4 : 148 * 16;
    launch_k(assemble_sh_bwd_kernel, gb, dim3(64, 4), 0, s, P, Mrest, g_shs, g_fdc, g_frest, (accumulate_mask >> 4) & 1);
  }   // NULL: caller slices g_shs itself
  if (g_feat)      // NULL: the feature parameter is frozen (GAUSSIAN state), nobody wants this gradient
    launch_k(assemble_feat_bwd_kernel, (int)(((size_t)P * 8 + 255) / 256), 256, 0, s, 
        P, reinterpret_cast<const float4*>(feat), reinterpret_cast<const float4*>(g_objs),
        reinterpret_cast<float4*>(g_feat));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_assemble_backward: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches((geom ? 1 : 0) + (g_fdc ? 1 : 0) + (g_feat ? 1 : 0));
  return 0;
}
