// Per-Gaussian kernels: forward preprocess, backward preprocess (conic → cov3D
// → scale/rotation, projection, SH), frustum test.
//
// Parity notes.  The reference computes these with glm (column-major) in fp32
// with nvcc's default FMA contraction; integer outputs (radii, tile rects)
// are decided by those floats, so the expression *shapes* below follow the
// math of the reference exactly (operand order of every 3x3 product, the
// double-precision ndc→pixel step, 1/(w+1e-7), un-normalised quaternion, ...):
//   forward : DGR/cuda_rasterizer/forward.cu:23-259, auxiliary.h:39-52,134-159
//   backward: DGR/cuda_rasterizer/backward.cu:18-389
// What is new: one fused backward kernel instead of two, block-staged 16-byte
// loads of the AoS inputs, a fused per-block tile count (feeds the scan), and
// SH gradients written once (no read-modify-write of dL_dmeans).
#include "common.cuh"
#include <cstdio>

namespace tb {

// ---- SH constants (values of the real SH basis; same literals as any 3DGS code)
__device__ __constant__ float kSH_C0 = 0.28209479177387814f;
__device__ __constant__ float kSH_C1 = 0.4886025119029199f;
__device__ __constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                           -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                           0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                           -0.5900435899266435f};

// ---- pinned fp32 sequences -------------------------------------------------
// ptxas is free to re-fuse un-annotated mul/add pairs, and does so differently
// in differently shaped kernels.  Everything that decides an integer output of
// the forward pass (radius, tile rectangle, depth key) is therefore written
// with explicit round-to-nearest intrinsics, in the operation order read off
// the reference's compiled sm_100 kernel (cuobjdump -sass of oracle/_ref):
//   a1*b1 + a2*b2 + a3*b3        -> fma(a3,b3, fma(a1,b1, a2*b2))
//   ... + m                      -> fadd(that, m)
__device__ __forceinline__ float dot3p(float a1, float b1, float a2, float b2, float a3, float b3) {
  return __fmaf_rn(a3, b3, __fmaf_rn(a1, b1, __fmul_rn(a2, b2)));
}
__device__ __forceinline__ float3 xform4x3p(const float3& p, const float* __restrict__ m) {
  return float3{__fadd_rn(dot3p(m[0], p.x, m[4], p.y, m[8], p.z), m[12]),
                __fadd_rn(dot3p(m[1], p.x, m[5], p.y, m[9], p.z), m[13]),
                __fadd_rn(dot3p(m[2], p.x, m[6], p.y, m[10], p.z), m[14])};
}
__device__ __forceinline__ float4 xform4x4p(const float3& p, const float* __restrict__ m) {
  return float4{__fadd_rn(dot3p(m[0], p.x, m[4], p.y, m[8], p.z), m[12]),
                __fadd_rn(dot3p(m[1], p.x, m[5], p.y, m[9], p.z), m[13]),
                __fadd_rn(dot3p(m[2], p.x, m[6], p.y, m[10], p.z), m[14]),
                __fadd_rn(dot3p(m[3], p.x, m[7], p.y, m[11], p.z), m[15])};
}

__device__ __forceinline__ float3 xform4x3(const float3& p, const float* __restrict__ m) {
  float3 t = {
      m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
      m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
      m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
  };
  return t;
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* __restrict__ m) {
  float4 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
              m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
  return t;
}
__device__ __forceinline__ float3 xform_vec4x3_T(const float3& p, const float* __restrict__ m) {
  float3 t = {
      m[0] * p.x + m[1] * p.y + m[2] * p.z,
      m[4] * p.x + m[5] * p.y + m[6] * p.z,
      m[8] * p.x + m[9] * p.y + m[10] * p.z,
  };
  return t;
}

// NDC → pixel centre; evaluated in double and narrowed (auxiliary.h:39-42).
__device__ __forceinline__ float ndc2pix(float v, int S) {
  return (float)__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5);
}

// Tile rectangle of a splat (auxiliary.h:44-52): float division, truncation, clamp.
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, int grid_x, int grid_y, uint2& rmin,
                                          uint2& rmax) {
  rmin = {(unsigned)min(grid_x, max((int)0, (int)((p.x - max_radius) / TILE))),
          (unsigned)min(grid_y, max((int)0, (int)((p.y - max_radius) / TILE)))};
  rmax = {(unsigned)min(grid_x, max((int)0, (int)((p.x + max_radius + TILE - 1) / TILE))),
          (unsigned)min(grid_y, max((int)0, (int)((p.y + max_radius + TILE - 1) / TILE)))};
}

// Stage `n_items` rows of `FL` floats (AoS, row-major) for a block into shared
// memory with 16-byte loads when the slab is aligned, scalar loads otherwise.
template <int FL>
__device__ __forceinline__ void stage_rows(const float* __restrict__ src, int first_row, int n_rows, float* s_dst) {
  const float* base = src + (size_t)first_row * FL;
  const int n = n_rows * FL;
  if ((((uintptr_t)base) & 15) == 0) {
    const int n4 = n >> 2;
    const float4* b4 = reinterpret_cast<const float4*>(base);
    for (int i = threadIdx.x; i < n4; i += blockDim.x) reinterpret_cast<float4*>(s_dst)[i] = __ldg(b4 + i);
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) s_dst[i] = __ldg(base + i);
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_dst[i] = __ldg(base + i);
  }
}

struct PreFwdParams {
  int P, D, M, W, H, grid_x, grid_y;
  float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
  int prefiltered;
  const float* means3D; const float* scales; const float* rotations; const float* opacities;
  const float* shs; const float* cov3D_precomp; const float* colors_precomp;
  const float* view; const float* proj; const float* campos;
  int* radii; float2* means2D; float* depths; float* cov3Ds; float4* rgbd; float4* conic_opacity;
  uint8_t* clamped; uint32_t* tiles_touched; uint32_t* block_sums;
};

// SH → RGB (forward.cu:23-74).  `sh` points at this Gaussian's [M][3] block.
// `avail` = floats readable from shp to the end of the SH tensor (bounds the rounded-up vector loads).
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float3 pos, const float3 campos,
                                            const float* __restrict__ shp, long long avail, uint8_t& clamp_bits) {
  float3 dir = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
  // glm::length = sqrt(dot(v,v)); dot is x*x + y*y + z*z evaluated left to right
  float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  dir.x = dir.x / len; dir.y = dir.y / len; dir.z = dir.z / len;

  // 16-byte loads of the (up to) 48 floats of this Gaussian's SH block when it
  // is 16-byte aligned (M = 4 or 16 on an aligned tensor), scalar loads otherwise.
  float shv[48];
  const int nfl = (deg + 1) * (deg + 1) * 3;
  const int nvec = (nfl + 3) >> 2;
  if ((((uintptr_t)shp) & 15) == 0 && 4ll * nvec <= avail) {
    const float4* s4 = reinterpret_cast<const float4*>(shp);
#pragma unroll
    for (int i = 0; i < 12; i++) {
      if (i < nvec) {
        float4 v = __ldg(s4 + i);
        shv[4 * i] = v.x; shv[4 * i + 1] = v.y; shv[4 * i + 2] = v.z; shv[4 * i + 3] = v.w;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 48; i++)
      if (i < nfl) shv[i] = __ldg(shp + i);
  }
  auto SH = [&](int i) -> float3 { return float3{shv[3 * i], shv[3 * i + 1], shv[3 * i + 2]}; };
  float3 c0 = SH(0);
  float3 result = {kSH_C0 * c0.x, kSH_C0 * c0.y, kSH_C0 * c0.z};
  if (deg > 0) {
    float x = dir.x, y = dir.y, z = dir.z;
    float3 s1 = SH(1), s2 = SH(2), s3 = SH(3);
    result.x = result.x - kSH_C1 * y * s1.x + kSH_C1 * z * s2.x - kSH_C1 * x * s3.x;
    result.y = result.y - kSH_C1 * y * s1.y + kSH_C1 * z * s2.y - kSH_C1 * x * s3.y;
    result.z = result.z - kSH_C1 * y * s1.z + kSH_C1 * z * s2.z - kSH_C1 * x * s3.z;
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z;
      float xy = x * y, yz = y * z, xz = x * z;
      float3 s4 = SH(4), s5 = SH(5), s6 = SH(6), s7 = SH(7), s8 = SH(8);
      result.x = result.x + kSH_C2[0] * xy * s4.x + kSH_C2[1] * yz * s5.x + kSH_C2[2] * (2.0f * zz - xx - yy) * s6.x +
                 kSH_C2[3] * xz * s7.x + kSH_C2[4] * (xx - yy) * s8.x;
      result.y = result.y + kSH_C2[0] * xy * s4.y + kSH_C2[1] * yz * s5.y + kSH_C2[2] * (2.0f * zz - xx - yy) * s6.y +
                 kSH_C2[3] * xz * s7.y + kSH_C2[4] * (xx - yy) * s8.y;
      result.z = result.z + kSH_C2[0] * xy * s4.z + kSH_C2[1] * yz * s5.z + kSH_C2[2] * (2.0f * zz - xx - yy) * s6.z +
                 kSH_C2[3] * xz * s7.z + kSH_C2[4] * (xx - yy) * s8.z;
      if (deg > 2) {
        float3 s9 = SH(9), s10 = SH(10), s11 = SH(11), s12 = SH(12), s13 = SH(13), s14 = SH(14), s15 = SH(15);
        result.x = result.x + kSH_C3[0] * y * (3.0f * xx - yy) * s9.x + kSH_C3[1] * xy * z * s10.x +
                   kSH_C3[2] * y * (4.0f * zz - xx - yy) * s11.x +
                   kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s12.x +
                   kSH_C3[4] * x * (4.0f * zz - xx - yy) * s13.x + kSH_C3[5] * z * (xx - yy) * s14.x +
                   kSH_C3[6] * x * (xx - 3.0f * yy) * s15.x;
        result.y = result.y + kSH_C3[0] * y * (3.0f * xx - yy) * s9.y + kSH_C3[1] * xy * z * s10.y +
                   kSH_C3[2] * y * (4.0f * zz - xx - yy) * s11.y +
                   kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s12.y +
                   kSH_C3[4] * x * (4.0f * zz - xx - yy) * s13.y + kSH_C3[5] * z * (xx - yy) * s14.y +
                   kSH_C3[6] * x * (xx - 3.0f * yy) * s15.y;
        result.z = result.z + kSH_C3[0] * y * (3.0f * xx - yy) * s9.z + kSH_C3[1] * xy * z * s10.z +
                   kSH_C3[2] * y * (4.0f * zz - xx - yy) * s11.z +
                   kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s12.z +
                   kSH_C3[4] * x * (4.0f * zz - xx - yy) * s13.z + kSH_C3[5] * z * (xx - yy) * s14.z +
                   kSH_C3[6] * x * (xx - 3.0f * yy) * s15.z;
      }
    }
  }
  result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
  clamp_bits = (uint8_t)((result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0));
  return float3{(result.x < 0.0f) ? 0.0f : result.x, (result.y < 0.0f) ? 0.0f : result.y,
                (result.z < 0.0f) ? 0.0f : result.z};
}

// Σ = (S R)^T (S R) (forward.cu:121-154); quaternion (r,x,y,z), not normalised.
// S is diagonal, so M[c][r] = s_r * R[c][r] (the zero terms of the full product
// add ±0 and do not change the rounding).
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 rot, float* cov3D) {
  const float sx = __fmul_rn(mod, scale.x), sy = __fmul_rn(mod, scale.y), sz = __fmul_rn(mod, scale.z);
  const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
  const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
  const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
  // R[c][r], column-major
  const float R00 = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(yy, zz)));
  const float R01 = __fmul_rn(2.f, __fmaf_rn(x, y, -rz));
  const float R02 = __fmul_rn(2.f, __fmaf_rn(r, y, xz));
  const float R10 = __fmul_rn(2.f, __fmaf_rn(x, y, rz));
  const float R11 = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, zz)));
  const float R12 = __fmul_rn(2.f, __fmaf_rn(y, z, -rx));
  const float R20 = __fmul_rn(2.f, __fmaf_rn(-r, y, xz));
  const float R21 = __fmul_rn(2.f, __fmaf_rn(y, z, rx));
  const float R22 = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, yy)));
  const float M00 = __fmul_rn(sx, R00), M01 = __fmul_rn(sy, R01), M02 = __fmul_rn(sz, R02);
  const float M10 = __fmul_rn(sx, R10), M11 = __fmul_rn(sy, R11), M12 = __fmul_rn(sz, R12);
  const float M20 = __fmul_rn(sx, R20), M21 = __fmul_rn(sy, R21), M22 = __fmul_rn(sz, R22);
  // Sigma[c][r] = dot(column r of M, column c of M)
  cov3D[0] = dot3p(M00, M00, M01, M01, M02, M02);
  cov3D[1] = dot3p(M10, M00, M11, M01, M12, M02);
  cov3D[2] = dot3p(M20, M00, M21, M01, M22, M02);
  cov3D[3] = dot3p(M10, M10, M11, M11, M12, M12);
  cov3D[4] = dot3p(M20, M10, M21, M11, M22, M12);
  cov3D[5] = dot3p(M20, M20, M21, M21, M22, M22);
}

// EWA projection of Σ to the image plane (forward.cu:77-116):
// cov = T^T Vrk^T T with T = W J, +0.3 on the diagonal.
__device__ __forceinline__ float3 cov2d_from_cov3d(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                                   float tan_fovy, const float* c, const float* __restrict__ v) {
  float3 t = xform4x3p(mean, v);
  const float limx = __fmul_rn(1.3f, tan_fovx);
  const float limy = __fmul_rn(1.3f, tan_fovy);
  const float txtz = __fdiv_rn(t.x, t.z);
  const float tytz = __fdiv_rn(t.y, t.z);
  t.x = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), t.z);
  t.y = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), t.z);

  const float tz2 = __fmul_rn(t.z, t.z);
  const float J00 = __fdiv_rn(focal_x, t.z);
  const float J02 = __fdiv_rn(-__fmul_rn(focal_x, t.x), tz2);
  const float J11 = __fdiv_rn(focal_y, t.z);
  const float J12 = __fdiv_rn(-__fmul_rn(focal_y, t.y), tz2);
  // T = W J with W = [v0 v4 v8 | v1 v5 v9 | v2 v6 v10] (columns) and
  // J = [J00 0 J02 | 0 J11 J12 | 0 0 0]: T[c][r] = W[0][r] J[c][0] + W[1][r] J[c][1] + W[2][r] J[c][2]
  const float T00 = __fmaf_rn(v[2], J02, __fmul_rn(v[0], J00));
  const float T01 = __fmaf_rn(v[6], J02, __fmul_rn(v[4], J00));
  const float T02 = __fmaf_rn(v[10], J02, __fmul_rn(v[8], J00));
  const float T10 = __fmaf_rn(v[2], J12, __fmul_rn(v[1], J11));
  const float T11 = __fmaf_rn(v[6], J12, __fmul_rn(v[5], J11));
  const float T12 = __fmaf_rn(v[10], J12, __fmul_rn(v[9], J11));
  // X = T^T Vrk^T : X[c][r] = T[r][0] V[0][c] + T[r][1] V[1][c] + T[r][2] V[2][c]   (rows r = 0,1 needed)
  const float X00 = dot3p(T00, c[0], T01, c[1], T02, c[2]);
  const float X10 = dot3p(T00, c[1], T01, c[3], T02, c[4]);
  const float X20 = dot3p(T00, c[2], T01, c[4], T02, c[5]);
  const float X01 = dot3p(T10, c[0], T11, c[1], T12, c[2]);
  const float X11 = dot3p(T10, c[1], T11, c[3], T12, c[4]);
  const float X21 = dot3p(T10, c[2], T11, c[4], T12, c[5]);
  // cov = X T : cov[c][r] = X[0][r] T[c][0] + X[1][r] T[c][1] + X[2][r] T[c][2]
  const float cov00 = dot3p(T00, X00, T01, X10, T02, X20);
  const float cov01 = dot3p(T00, X01, T01, X11, T02, X21);
  const float cov11 = dot3p(T10, X01, T11, X11, T12, X21);
  return {__fadd_rn(cov00, 0.3f), cov01, __fadd_rn(cov11, 0.3f)};
}

__global__ void __launch_bounds__(PRE_BLOCK) preprocess_fwd_kernel(const PreFwdParams q) {
  pdl_wait();
  __shared__ float s_mean[PRE_BLOCK * 3];
  __shared__ float s_scale[PRE_BLOCK * 3];
  __shared__ uint32_t s_warp_sum[PRE_BLOCK / 32];

  const int first = blockIdx.x * PRE_BLOCK;
  const int n_rows = min(PRE_BLOCK, q.P - first);
  stage_rows<3>(q.means3D, first, n_rows, s_mean);
  if (q.cov3D_precomp == nullptr) stage_rows<3>(q.scales, first, n_rows, s_scale);
  __syncthreads();

  const int idx = first + threadIdx.x;
  uint32_t my_tiles = 0;
  if (idx < q.P) {
    int out_radius = 0;
    do {
      const float3 p_orig = {s_mean[3 * threadIdx.x], s_mean[3 * threadIdx.x + 1], s_mean[3 * threadIdx.x + 2]};
      // near cull (auxiliary.h:134-159): view-space z <= 0.2 is invisible; no x/y cull
      const float3 p_view = xform4x3p(p_orig, q.view);
      if (p_view.z <= 0.2f) {
        if (q.prefiltered) {
          printf("Point is filtered although prefiltered is set. This shouldn't happen!");
          __trap();
        }
        break;
      }
      float4 p_hom = xform4x4p(p_orig, q.proj);
      float p_w = __frcp_rn(__fadd_rn(p_hom.w, 0.0000001f));
      float3 p_proj = {__fmul_rn(p_hom.x, p_w), __fmul_rn(p_hom.y, p_w), __fmul_rn(p_hom.z, p_w)};

      float cov3D_local[6];
      const float* cov3D;
      if (q.cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; i++) cov3D_local[i] = __ldg(q.cov3D_precomp + (size_t)idx * 6 + i);
        cov3D = cov3D_local;
      } else {
        const float3 sc = {s_scale[3 * threadIdx.x], s_scale[3 * threadIdx.x + 1], s_scale[3 * threadIdx.x + 2]};
        const float4 rot = __ldg(reinterpret_cast<const float4*>(q.rotations) + idx);
        cov3d_from_scale_rot(sc, q.scale_modifier, rot, cov3D_local);
#pragma unroll
        for (int i = 0; i < 6; i++) q.cov3Ds[(size_t)idx * 6 + i] = cov3D_local[i];
        cov3D = cov3D_local;
      }

      float3 cov = cov2d_from_cov3d(p_orig, q.focal_x, q.focal_y, q.tan_fovx, q.tan_fovy, cov3D, q.view);

      float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
      if (det == 0.0f) break;
      float det_inv = __frcp_rn(det);
      float3 conic = {__fmul_rn(cov.z, det_inv), -__fmul_rn(cov.y, det_inv), __fmul_rn(cov.x, det_inv)};

      float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
      float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
      float lambda1 = __fadd_rn(mid, sq);
      float lambda2 = __fsub_rn(mid, sq);
      float my_radius = ceilf(__fmul_rn(3.f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
      float2 point_image = {ndc2pix(p_proj.x, q.W), ndc2pix(p_proj.y, q.H)};
      uint2 rmin, rmax;
      tile_rect(point_image, (int)my_radius, q.grid_x, q.grid_y, rmin, rmax);
      if ((rmax.x - rmin.x) * (rmax.y - rmin.y) == 0) break;

      float3 rgb;
      if (q.colors_precomp == nullptr) {
        uint8_t cb;
        const float3 cam = {q.campos[0], q.campos[1], q.campos[2]};
        rgb = sh_to_rgb(q.D, p_orig, cam, q.shs + (size_t)idx * q.M * 3, (long long)(q.P - idx) * q.M * 3, cb);
        q.clamped[idx] = cb;
      } else {
        rgb = {__ldg(q.colors_precomp + (size_t)idx * 3), __ldg(q.colors_precomp + (size_t)idx * 3 + 1),
               __ldg(q.colors_precomp + (size_t)idx * 3 + 2)};
      }

      q.depths[idx] = p_view.z;
      out_radius = (int)my_radius;
      q.means2D[idx] = point_image;
      q.conic_opacity[idx] = {conic.x, conic.y, conic.z, __ldg(q.opacities + idx)};
      q.rgbd[idx] = {rgb.x, rgb.y, rgb.z, p_view.z};
      my_tiles = (rmax.y - rmin.y) * (rmax.x - rmin.x);
    } while (false);
    q.radii[idx] = out_radius;
    q.tiles_touched[idx] = my_tiles;
  }

  // per-block tile count → feeds the block-offset scan (no separate reduce pass)
  uint32_t v = my_tiles;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) s_warp_sum[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < PRE_BLOCK / 32; w++) t += s_warp_sum[w];
    q.block_sums[blockIdx.x] = t;
  }
}

cudaError_t launch_preprocess_forward(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s) {
  PreFwdParams q;
  q.P = a.P; q.D = a.sh_degree; q.M = a.M; q.W = a.W; q.H = a.H;
  q.grid_x = (a.W + TILE - 1) / TILE; q.grid_y = (a.H + TILE - 1) / TILE;
  q.tan_fovx = a.tan_fovx; q.tan_fovy = a.tan_fovy;
  q.focal_y = a.H / (2.0f * a.tan_fovy);
  q.focal_x = a.W / (2.0f * a.tan_fovx);
  q.scale_modifier = a.scale_modifier; q.prefiltered = a.prefiltered;
  q.means3D = a.means3D; q.scales = a.scales; q.rotations = a.rotations; q.opacities = a.opacities;
  q.shs = a.shs; q.cov3D_precomp = a.cov3D_precomp; q.colors_precomp = a.colors_precomp;
  q.view = a.viewmatrix; q.proj = a.projmatrix; q.campos = a.campos;
  q.radii = a.radii; q.means2D = p.means2D; q.depths = p.depths; q.cov3Ds = p.cov3D; q.rgbd = p.rgbd;
  q.conic_opacity = p.conic_opacity; q.clamped = p.clamped; q.tiles_touched = p.tiles_touched;
  q.block_sums = p.block_sums;
  int nblk = (a.P + PRE_BLOCK - 1) / PRE_BLOCK;
  launch_k(preprocess_fwd_kernel, nblk, PRE_BLOCK, 0, s, q);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Frustum test (rasterizer_impl.cu:57-69)
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  pdl_wait();
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  float3 pv = xform4x3(p, view);
  present[idx] = (pv.z <= 0.2f) ? 0 : 1;
}
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  launch_k(mark_visible_kernel, (P + 255) / 256, 256, 0, s, P, means3D, view, present);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Backward preprocess: one kernel for what the reference does in two
// (computeCov2DCUDA, backward.cu:139-269; preprocessCUDA, backward.cu:339-389).
struct PreBwdParams {
  int P, D, M;
  float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
  const float* means3D; const int* radii; const float* shs; const uint8_t* clamped;
  const float* scales; const float* rotations; const float* cov3Ds;  // cov3Ds: precomp or geom
  const float* view; const float* proj; const float* campos;
  const float* dL_dmean2D;  // [P][3]
  const float* dL_dconic;   // [P][4]
  const float* dL_dcolor;   // [P][3]
  float* dL_dmeans3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscale; float* dL_drot;
};

// Gradient through v -> v/|v|:  J^T g = (g |v|^2 - v (v.g)) / |v|^3.
__device__ __forceinline__ float3 unit_vector_grad(const float3 v, const float3 g) {
  const float n2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float vg = v.x * g.x + v.y * g.y + v.z * g.z;
  const float inv_n3 = 1.0f / (n2 * sqrtf(n2));
  return float3{(g.x * n2 - v.x * vg) * inv_n3, (g.y * n2 - v.y * vg) * inv_n3, (g.z * n2 - v.z * vg) * inv_n3};
}

__device__ __forceinline__ float dot3(const float3 a, const float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 axpby(float a, const float3 x, float b, const float3 y) {
  return float3{a * x.x + b * y.x, a * x.y + b * y.y, a * x.z + b * y.z};
}
// symmetric 3x3 (xx, xy, xz, yy, yz, zz) times vector
__device__ __forceinline__ float3 sym3_mul(const float* s, const float3 v) {
  return float3{s[0] * v.x + s[1] * v.y + s[2] * v.z, s[1] * v.x + s[3] * v.y + s[4] * v.z,
                s[2] * v.x + s[4] * v.y + s[5] * v.z};
}

// Backward of the EWA projection, in the kernel's own terms.  The forward pass is
//   cov2D = A Sigma A^T + 0.3 I,   A = J Wr  (2x3),   conic = cov2D^-1,
// with Wr the rotation rows w0,w1,w2 of the view matrix and J's two rows (j00,0,j02), (0,j11,j12).  With the rows
// a0 = j00 w0 + j02 w2 and a1 = j11 w1 + j12 w2 of A and u_i = Sigma a_i:
//   cov2D = [[a0.u0 + .3, a0.u1], [a0.u1, a1.u1 + .3]]
//   dL/dcov2D = -adj G adj / (det^2 + 1e-7)      (G = the symmetric conic gradient, adj = adjugate of cov2D;
//                                                 the 1e-7 is the reference's guard, backward.cu:196-198)
//   dL/dSigma = A^T D A  (returned with doubled off-diagonals: one parameter stands for both entries)
//   dL/da0 = 2 Daa u0 + 2 Dab u1,  dL/da1 = 2 Dcc u1 + 2 Dab u0
//   dL/dj** = the w rows dotted with those, and from there to the view-space mean t and back through Wr^T.
// Same mathematics as computeCov2DCUDA (DGR/cuda_rasterizer/backward.cu:139-269); gradients only have to agree
// to 1e-3, so the evaluation order is this kernel's own.
__device__ __forceinline__ float3 ewa_projection_backward(const float3 mean, const float* __restrict__ view,
                                                          float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                                                          const float* sigma, const float3 g_conic, float* dsigma) {
  const float3 w0 = {view[0], view[4], view[8]}, w1 = {view[1], view[5], view[9]}, w2 = {view[2], view[6], view[10]};
  float tx = dot3(w0, mean) + view[12], ty = dot3(w1, mean) + view[13];
  const float tz = dot3(w2, mean) + view[14];
  const float iz = 1.0f / tz;
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float rx = tx * iz, ry = ty * iz;
  const bool x_inside = !(rx < -limx || rx > limx), y_inside = !(ry < -limy || ry > limy);
  const float cx = fminf(limx, fmaxf(-limx, rx)), cy = fminf(limy, fmaxf(-limy, ry));   // clamped t.x/t.z, t.y/t.z
  const float j00 = focal_x * iz, j11 = focal_y * iz;
  const float j02 = -j00 * cx, j12 = -j11 * cy;                                        // -f t.x / t.z^2

  const float3 a0 = axpby(j00, w0, j02, w2), a1 = axpby(j11, w1, j12, w2);
  const float3 u0 = sym3_mul(sigma, a0), u1 = sym3_mul(sigma, a1);
  const float a = dot3(a0, u0) + 0.3f, b = dot3(a0, u1), c = dot3(a1, u1) + 0.3f;

  const float det = a * c - b * b;
  const float k = 1.0f / (det * det + 0.0000001f);
  float Daa = 0.f, Dab = 0.f, Dcc = 0.f;          // dL/dcov2D as a symmetric matrix
  if (k != 0.f) {
    // adj = [[c,-b],[-b,a]];  Q = adj G adj
    const float p0 = c * g_conic.x - b * g_conic.y, p1 = c * g_conic.y - b * g_conic.z;   // first row of adj G
    const float q0 = a * g_conic.y - b * g_conic.x, q1 = a * g_conic.z - b * g_conic.y;   // second row of adj G
    Daa = -k * (p0 * c - p1 * b);
    Dab = -k * (p1 * a - p0 * b);
    Dcc = -k * (q1 * a - q0 * b);
  }
  // dL/dSigma = Daa a0 a0^T + Dab (a0 a1^T + a1 a0^T) + Dcc a1 a1^T
  const float3 e0 = axpby(Daa, a0, Dab, a1), e1 = axpby(Dab, a0, Dcc, a1);    // dSigma = a0 e0^T + a1 e1^T
  dsigma[0] = a0.x * e0.x + a1.x * e1.x;
  dsigma[3] = a0.y * e0.y + a1.y * e1.y;
  dsigma[5] = a0.z * e0.z + a1.z * e1.z;
  dsigma[1] = (a0.x * e0.y + a1.x * e1.y) + (a0.y * e0.x + a1.y * e1.x);
  dsigma[2] = (a0.x * e0.z + a1.x * e1.z) + (a0.z * e0.x + a1.z * e1.x);
  dsigma[4] = (a0.y * e0.z + a1.y * e1.z) + (a0.z * e0.y + a1.z * e1.y);

  const float3 g0 = axpby(2.f * Daa, u0, 2.f * Dab, u1), g1 = axpby(2.f * Dcc, u1, 2.f * Dab, u0);  // dL/da0, dL/da1
  const float dj00 = dot3(w0, g0), dj02 = dot3(w2, g0), dj11 = dot3(w1, g1), dj12 = dot3(w2, g1);
  // j00 = fx/tz, j02 = -fx tx/tz^2 (tx clamped: no gradient through x when the clamp is active), same for y
  const float dtx = x_inside ? -j00 * iz * dj02 : 0.f;
  const float dty = y_inside ? -j11 * iz * dj12 : 0.f;
  const float dtz = -iz * (j00 * dj00 + j11 * dj11 + 2.f * (j02 * dj02 + j12 * dj12));
  return float3{w0.x * dtx + w1.x * dty + w2.x * dtz, w0.y * dtx + w1.y * dty + w2.y * dtz,
                w0.z * dtx + w1.z * dty + w2.z * dtz};
}

// Backward of Sigma = sum_k s_k^2 r_k r_k^T (r_k = columns of the rotation built from the UN-normalised quaternion
// (r,x,y,z), s = scale_modifier * scale).  With G = dL/dSigma as a symmetric matrix:
//   dL/ds_k = 2 s_k r_k^T G r_k              (the reference does not apply the scale_modifier chain factor; kept)
//   h_k = dL/dr_k = 2 s_k^2 G r_k, and the quaternion gradient collects the symmetric and antisymmetric parts of
//   H = [h_0 h_1 h_2].  Same mathematics as computeCov3D's backward (backward.cu:273-334), own evaluation order.
__device__ __forceinline__ void scale_rotation_backward(const float3 scale, float mod, const float4 q,
                                                        const float* dsigma, float3& dscale, float4& dquat) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const float3 r0 = {1.f - 2.f * (y * y + z * z), 2.f * (x * y + r * z), 2.f * (x * z - r * y)};
  const float3 r1 = {2.f * (x * y - r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + r * x)};
  const float3 r2 = {2.f * (x * z + r * y), 2.f * (y * z - r * x), 1.f - 2.f * (x * x + y * y)};
  // symmetric-matrix form of the stored gradient (off-diagonals were doubled)
  const float G[6] = {dsigma[0], 0.5f * dsigma[1], 0.5f * dsigma[2], dsigma[3], 0.5f * dsigma[4], dsigma[5]};
  const float3 s = {mod * scale.x, mod * scale.y, mod * scale.z};
  const float3 t0 = sym3_mul(G, r0), t1 = sym3_mul(G, r1), t2 = sym3_mul(G, r2);
  dscale = float3{2.f * s.x * dot3(r0, t0), 2.f * s.y * dot3(r1, t1), 2.f * s.z * dot3(r2, t2)};
  const float k0 = 2.f * s.x * s.x, k1 = 2.f * s.y * s.y, k2 = 2.f * s.z * s.z;
  const float3 h0 = {k0 * t0.x, k0 * t0.y, k0 * t0.z}, h1 = {k1 * t1.x, k1 * t1.y, k1 * t1.z},
               h2 = {k2 * t2.x, k2 * t2.y, k2 * t2.z};
  // H[j][k] = h_k[j]: axial vector of H - H^T and the symmetric sums
  const float ax = h1.z - h2.y, ay = h2.x - h0.z, az = h0.y - h1.x;
  const float sxy = h0.y + h1.x, sxz = h0.z + h2.x, syz = h1.z + h2.y;
  dquat.x = 2.f * (x * ax + y * ay + z * az);
  dquat.y = 2.f * (r * ax + y * sxy + z * sxz) - 4.f * x * (h1.y + h2.z);
  dquat.z = 2.f * (r * ay + x * sxy + z * syz) - 4.f * y * (h0.x + h2.z);
  dquat.w = 2.f * (r * az + x * sxz + y * syz) - 4.f * z * (h0.x + h1.y);
}

__global__ void __launch_bounds__(PRE_BLOCK) preprocess_bwd_kernel(const PreBwdParams q) {
  pdl_wait();
  const int idx = blockIdx.x * PRE_BLOCK + threadIdx.x;
  // dL/dSH of a Gaussian is (basis value i) x (dL/dRGB): 16 + 3 numbers.  They are staged per warp and written
  // by the whole warp as one contiguous block (a thread writing its own 192 bytes with 4-byte stores at a
  // 192-byte stride kept the L2 at 64 % of its peak on partial sectors).
  __shared__ float s_shw[PRE_BLOCK / 32][32][17];
  __shared__ float s_shrgb[PRE_BLOCK / 32][32][3];
  __shared__ unsigned char s_shact[PRE_BLOCK / 32][32];
  float shw[16];
#pragma unroll
  for (int i = 0; i < 16; i++) shw[i] = 0.f;
  float3 sh_rgb = {0.f, 0.f, 0.f};
  const bool active = idx < q.P && q.radii[idx] > 0;
  if (active) {
  const float3 mean = {q.means3D[3 * idx], q.means3D[3 * idx + 1], q.means3D[3 * idx + 2]};

  // ---- conic -> cov2D -> Sigma, and the Jacobian's share of dL/dmean
  float sigma[6], dcov[6];
#pragma unroll
  for (int i = 0; i < 6; i++) sigma[i] = q.cov3Ds[(size_t)idx * 6 + i];
  const float4 dLc4 = reinterpret_cast<const float4*>(q.dL_dconic)[idx];   // (x, y, unused, w)
  float3 dL_dmean_acc = ewa_projection_backward(mean, q.view, q.focal_x, q.focal_y, q.tan_fovx, q.tan_fovy, sigma,
                                                float3{dLc4.x, dLc4.y, dLc4.w}, dcov);
#pragma unroll
  for (int i = 0; i < 6; i++) q.dL_dcov3D[(size_t)idx * 6 + i] = dcov[i];

  // ---- screen-space mean: ndc = h.xy / (h.w + 1e-7), h = proj p  =>  d ndc_x / d p_k = (P[k][0] - ndc_x P[k][3]) / w
  {
    const float* pm = q.proj;
    const float4 h = xform4x4(mean, pm);
    const float iw = 1.0f / (h.w + 0.0000001f);
    const float nx = h.x * iw, ny = h.y * iw;
    const float gx = q.dL_dmean2D[3 * idx] * iw, gy = q.dL_dmean2D[3 * idx + 1] * iw;
    dL_dmean_acc.x += (pm[0] - nx * pm[3]) * gx + (pm[1] - ny * pm[3]) * gy;
    dL_dmean_acc.y += (pm[4] - nx * pm[7]) * gx + (pm[5] - ny * pm[7]) * gy;
    dL_dmean_acc.z += (pm[8] - nx * pm[11]) * gx + (pm[9] - ny * pm[11]) * gy;
  }

  // ---- SH backward (backward.cu:18-134)
  if (q.shs) {
    const float3 campos = {q.campos[0], q.campos[1], q.campos[2]};
    const float3 dir_orig = {mean.x - campos.x, mean.y - campos.y, mean.z - campos.z};
    float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    const float* shp = q.shs + (size_t)idx * q.M * 3;
    auto SH = [&](int i) -> float3 {
      return float3{__ldg(shp + 3 * i), __ldg(shp + 3 * i + 1), __ldg(shp + 3 * i + 2)};
    };
    const uint8_t cb = q.clamped[idx];
    float3 dL_dRGB = {q.dL_dcolor[3 * idx], q.dL_dcolor[3 * idx + 1], q.dL_dcolor[3 * idx + 2]};
    dL_dRGB.x *= (cb & 1) ? 0 : 1;
    dL_dRGB.y *= (cb & 2) ? 0 : 1;
    dL_dRGB.z *= (cb & 4) ? 0 : 1;

    float3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
    sh_rgb = dL_dRGB;
    auto put = [&](int i, float w) { shw[i] = w; };
    auto axpy = [](float3& acc, float w, const float3& v) {
      acc.x += w * v.x; acc.y += w * v.y; acc.z += w * v.z;
    };
    put(0, kSH_C0);
    if (q.D > 0) {
      float3 s1 = SH(1), s2 = SH(2), s3 = SH(3);
      put(1, -kSH_C1 * y);
      put(2, kSH_C1 * z);
      put(3, -kSH_C1 * x);
      dRGBdx = {-kSH_C1 * s3.x, -kSH_C1 * s3.y, -kSH_C1 * s3.z};
      dRGBdy = {-kSH_C1 * s1.x, -kSH_C1 * s1.y, -kSH_C1 * s1.z};
      dRGBdz = {kSH_C1 * s2.x, kSH_C1 * s2.y, kSH_C1 * s2.z};
      if (q.D > 1) {
        float xx = x * x, yy = y * y, zz = z * z;
        float xy = x * y, yz = y * z, xz = x * z;
        float3 s4 = SH(4), s5 = SH(5), s6 = SH(6), s7 = SH(7), s8 = SH(8);
        put(4, kSH_C2[0] * xy);
        put(5, kSH_C2[1] * yz);
        put(6, kSH_C2[2] * (2.f * zz - xx - yy));
        put(7, kSH_C2[3] * xz);
        put(8, kSH_C2[4] * (xx - yy));
        axpy(dRGBdx, kSH_C2[0] * y, s4); axpy(dRGBdx, kSH_C2[2] * 2.f * -x, s6);
        axpy(dRGBdx, kSH_C2[3] * z, s7); axpy(dRGBdx, kSH_C2[4] * 2.f * x, s8);
        axpy(dRGBdy, kSH_C2[0] * x, s4); axpy(dRGBdy, kSH_C2[1] * z, s5);
        axpy(dRGBdy, kSH_C2[2] * 2.f * -y, s6); axpy(dRGBdy, kSH_C2[4] * 2.f * -y, s8);
        axpy(dRGBdz, kSH_C2[1] * y, s5); axpy(dRGBdz, kSH_C2[2] * 2.f * 2.f * z, s6);
        axpy(dRGBdz, kSH_C2[3] * x, s7);
        if (q.D > 2) {
          float3 s9 = SH(9), s10 = SH(10), s11 = SH(11), s12 = SH(12), s13 = SH(13), s14 = SH(14), s15 = SH(15);
          put(9, kSH_C3[0] * y * (3.f * xx - yy));
          put(10, kSH_C3[1] * xy * z);
          put(11, kSH_C3[2] * y * (4.f * zz - xx - yy));
          put(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
          put(13, kSH_C3[4] * x * (4.f * zz - xx - yy));
          put(14, kSH_C3[5] * z * (xx - yy));
          put(15, kSH_C3[6] * x * (xx - 3.f * yy));
          axpy(dRGBdx, kSH_C3[0] * 3.f * 2.f * xy, s9); axpy(dRGBdx, kSH_C3[1] * yz, s10);
          axpy(dRGBdx, kSH_C3[2] * -2.f * xy, s11); axpy(dRGBdx, kSH_C3[3] * -3.f * 2.f * xz, s12);
          axpy(dRGBdx, kSH_C3[4] * (-3.f * xx + 4.f * zz - yy), s13); axpy(dRGBdx, kSH_C3[5] * 2.f * xz, s14);
          axpy(dRGBdx, kSH_C3[6] * 3.f * (xx - yy), s15);
          axpy(dRGBdy, kSH_C3[0] * 3.f * (xx - yy), s9); axpy(dRGBdy, kSH_C3[1] * xz, s10);
          axpy(dRGBdy, kSH_C3[2] * (-3.f * yy + 4.f * zz - xx), s11); axpy(dRGBdy, kSH_C3[3] * -3.f * 2.f * yz, s12);
          axpy(dRGBdy, kSH_C3[4] * -2.f * xy, s13); axpy(dRGBdy, kSH_C3[5] * -2.f * yz, s14);
          axpy(dRGBdy, kSH_C3[6] * -3.f * 2.f * xy, s15);
          axpy(dRGBdz, kSH_C3[1] * xy, s10); axpy(dRGBdz, kSH_C3[2] * 4.f * 2.f * yz, s11);
          axpy(dRGBdz, kSH_C3[3] * 3.f * (2.f * zz - xx - yy), s12); axpy(dRGBdz, kSH_C3[4] * 4.f * 2.f * xz, s13);
          axpy(dRGBdz, kSH_C3[5] * (xx - yy), s14);
        }
      }
    }
    float3 dL_ddir = {dRGBdx.x * dL_dRGB.x + dRGBdx.y * dL_dRGB.y + dRGBdx.z * dL_dRGB.z,
                      dRGBdy.x * dL_dRGB.x + dRGBdy.y * dL_dRGB.y + dRGBdy.z * dL_dRGB.z,
                      dRGBdz.x * dL_dRGB.x + dRGBdz.y * dL_dRGB.y + dRGBdz.z * dL_dRGB.z};
    float3 dm = unit_vector_grad(dir_orig, dL_ddir);
    dL_dmean_acc.x += dm.x; dL_dmean_acc.y += dm.y; dL_dmean_acc.z += dm.z;
  }
  q.dL_dmeans3D[3 * idx] = dL_dmean_acc.x;
  q.dL_dmeans3D[3 * idx + 1] = dL_dmean_acc.y;
  q.dL_dmeans3D[3 * idx + 2] = dL_dmean_acc.z;

  // ---- Sigma -> scale / rotation
  if (q.scales) {
    const float3 scale = {q.scales[3 * idx], q.scales[3 * idx + 1], q.scales[3 * idx + 2]};
    const float4 rot = reinterpret_cast<const float4*>(q.rotations)[idx];
    float3 dscale;
    float4 dq;
    scale_rotation_backward(scale, q.scale_modifier, rot, dcov, dscale, dq);
    q.dL_dscale[3 * idx] = dscale.x; q.dL_dscale[3 * idx + 1] = dscale.y; q.dL_dscale[3 * idx + 2] = dscale.z;
    reinterpret_cast<float4*>(q.dL_drot)[idx] = dq;
  }
  }  // active

  if (q.shs) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < 16; i++) s_shw[warp][lane][i] = shw[i];
    s_shrgb[warp][lane][0] = sh_rgb.x; s_shrgb[warp][lane][1] = sh_rgb.y; s_shrgb[warp][lane][2] = sh_rgb.z;
    s_shact[warp][lane] = active ? 1 : 0;
    __syncwarp();
    const int M3 = q.M * 3;
    const int ncoef = min((q.D + 1) * (q.D + 1), 16);
    float* dst = q.dL_dsh + (size_t)(blockIdx.x * PRE_BLOCK + warp * 32) * M3;
    for (int f = lane; f < 32 * M3; f += 32) {
      const int g = f / M3, e = f - g * M3;
      const int i = e / 3, c = e - 3 * i;
      // coefficients beyond the active degree are not written (the buffer arrives zero-initialised), as in the reference
      if (s_shact[warp][g] && i < ncoef) dst[f] = s_shw[warp][g][i] * s_shrgb[warp][g][c];
    }
  }
}

cudaError_t launch_preprocess_backward(const trase_b200_bwd_args& a, const FwdPointers& p, cudaStream_t s) {
  PreBwdParams q;
  q.P = a.P; q.D = a.sh_degree; q.M = a.M;
  q.tan_fovx = a.tan_fovx; q.tan_fovy = a.tan_fovy;
  q.focal_y = a.H / (2.0f * a.tan_fovy);
  q.focal_x = a.W / (2.0f * a.tan_fovx);
  q.scale_modifier = a.scale_modifier;
  q.means3D = a.means3D; q.radii = a.radii; q.shs = a.shs; q.clamped = p.clamped;
  q.scales = a.scales; q.rotations = a.rotations;
  q.cov3Ds = (a.cov3D_precomp != nullptr) ? a.cov3D_precomp : p.cov3D;
  q.view = a.viewmatrix; q.proj = a.projmatrix; q.campos = a.campos;
  q.dL_dmean2D = a.dL_dmeans2D; q.dL_dconic = a.dL_dconic; q.dL_dcolor = a.dL_dcolors;
  q.dL_dmeans3D = a.dL_dmeans3D; q.dL_dcov3D = a.dL_dcov3D; q.dL_dsh = a.dL_dsh;
  q.dL_dscale = a.dL_dscales; q.dL_drot = a.dL_drotations;
  int nblk = (a.P + PRE_BLOCK - 1) / PRE_BLOCK;
  launch_k(preprocess_bwd_kernel, nblk, PRE_BLOCK, 0, s, q);
  return cudaGetLastError();
}

}  // namespace tb
