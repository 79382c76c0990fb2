// Pieces shared by the compositing kernels (render.cu, render_mma.cu): the per-splat reach / per-warp culling tests
// and the forward parameter block.
#pragma once
#include "common.cuh"

namespace tb {

__device__ __forceinline__ float splat_power(const float2 xy, const float4 co, float pxf, float pyf, float& dx,
                                             float& dy) {
  dx = xy.x - pxf;
  dy = xy.y - pyf;
  const float S = __fmaf_rn(dx, __fmul_rn(dx, co.x), __fmul_rn(__fmul_rn(dy, co.z), dy));
  return __fmaf_rn(S, -0.5f, -__fmul_rn(__fmul_rn(dx, co.y), dy));
}

// Conservative half-extents (in pixels) of the axis-aligned box outside of which this splat cannot
// reach alpha >= 1/255:  o*exp(-q/2) >= 1/255  <=>  q = d^T C d <= 2 ln(255 o), and over that ellipse
// |dx| <= sqrt(2 ln(255 o) * (C^-1)_xx).  Used only to skip whole (warp, splat) pairs early; the exact
// per-pixel test still decides every pair that is evaluated, so results are unchanged.
// .z = a power threshold: o*exp(power) < 1/255 for every power below it (with slack), so such pairs skip
// the expf; pairs at or above it go through the exact test.
__device__ __forceinline__ float3 splat_reach(const float4 co) {
  const float t = __logf(255.f * co.w);
  const float t2 = 2.f * t;
  const float det = co.x * co.z - co.y * co.y;
  const float thr = (t == t) ? -(t + 1e-3f * (1.f + fabsf(t))) : -3.0e38f;
  if (!(t2 > 0.f)) return make_float3(-1.f, -1.f, thr);       // opacity below 1/255: never blends
  if (!(det > 0.f) || !isfinite(det)) return make_float3(3.0e38f, 3.0e38f, thr);
  const float inv = 1.f / det;
  float3 r;
  r.z = thr;
  r.x = sqrtf(t2 * co.z * inv) * 1.002f + 0.05f;              // generous margin over fp32 rounding
  r.y = sqrtf(t2 * co.x * inv) * 1.002f + 0.05f;
  if (!(r.x == r.x)) r.x = 3.0e38f;
  if (!(r.y == r.y)) r.y = 3.0e38f;
  return r;
}

// Smallest value of q(d) = a dx^2 + 2 b dx dy + c dy^2 (power = -q/2) over the axis-aligned rectangle
// dx in [u0,u1], dy in [v0,v1] (offsets of a warp's pixel block from the splat centre), for a positive
// definite conic.  A pair can blend only if q <= -2*thr somewhere in the block, so a splat whose minimum
// exceeds that can be dropped for the whole warp before any per-pixel work.
__device__ __forceinline__ float min_q_over_rect(float a, float b, float c, float u0, float u1, float v0, float v1) {
  if (u0 <= 0.f && u1 >= 0.f && v0 <= 0.f && v1 >= 0.f) return 0.f;
  float best = 3.0e38f;
#pragma unroll
  for (int s = 0; s < 2; s++) {
    const float U = s ? u1 : u0;                       // edge dx = U
    const float v = fminf(fmaxf(-b * U / c, v0), v1);
    best = fminf(best, a * U * U + 2.f * b * U * v + c * v * v);
    const float V = s ? v1 : v0;                       // edge dy = V
    const float u = fminf(fmaxf(-b * V / a, u0), u1);
    best = fminf(best, a * u * u + 2.f * b * u * V + c * V * V);
  }
  return best;
}
// true if the splat may blend into some pixel of the block [wx0,wx1] x [wy0,wy1]
__device__ __forceinline__ bool splat_may_touch(const float2 xy, const float4 co, const float3 reach, float wx0,
                                                float wx1, float wy0, float wy1) {
  if (fmaxf(fmaxf(wx0 - xy.x, xy.x - wx1), 0.f) > reach.x || fmaxf(fmaxf(wy0 - xy.y, xy.y - wy1), 0.f) > reach.y)
    return false;
  if (reach.x > 1.0e37f) return true;                  // degenerate conic: no tight test
  const float qmin = min_q_over_rect(co.x, co.y, co.z, wx0 - xy.x, wx1 - xy.x, wy0 - xy.y, wy1 - xy.y);
  return qmin * 0.998f <= -2.f * reach.z + 1e-3f;
}

struct RenderFwdParams {
  const uint2* ranges; const uint32_t* point_list;
  int W, H;
  const float2* means2D; const float4* rgbd; const float4* feat4;  // feat4: sh_objs as float4[P][8]
  const float4* conic_opacity;
  float* final_T; uint32_t* n_contrib; const float* bg;
  float* out_color; float* out_objects; float* out_depth;
};


cudaError_t launch_render_forward_mma(const RenderFwdParams& q, dim3 grid, int min_blocks, cudaStream_t s);

}  // namespace tb
