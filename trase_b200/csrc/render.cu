// Tile compositing: forward (front-to-back alpha blending of RGB + 32 feature
// channels + depth) and backward (back-to-front, warp-reduced gradients).
//
// Reference counterparts:
//   forward  renderCUDA  DGR/cuda_rasterizer/forward.cu:264-396
//   backward renderCUDA  DGR/cuda_rasterizer/backward.cu:392-574
//
// Bit-exactness.  n_contrib (and the set of blended pairs) depends on the exact
// fp32 value of alpha and of the running transmittance.  The alpha chain below
// is pinned with explicit round-to-nearest intrinsics to the operation order
// the reference's compiled kernel uses (checked against its SASS):
//   S     = fma(dx, dx*a, (dy*c)*dy)
//   power = fma(S, -0.5, -((dx*b)*dy))
//   alpha = min(0.99, o * expf(power));   T' = T * (1 - alpha)
// Colour/feature accumulation order is free (tolerance 1e-4 abs).
//
// What differs from the reference:
//  - forward stages the 32-channel feature row, colour and depth of every
//    Gaussian of a batch in shared memory once per tile (the reference re-reads
//    36 floats from global memory per pixel per Gaussian);
//  - each thread can own PPT vertically adjacent pixels, so every shared-memory
//    broadcast read and every cross-lane reduction is amortised over PPT pixels;
//  - backward replaces the per-channel accum_rec/last_color recursions by one
//    scalar recursion on dot(colour, dL/dpixel) (linear in the channels, so the
//    result is the same up to rounding) — this removes 70 live registers;
//  - backward reduces the 41 per-Gaussian partial gradients across the warp
//    (transposed butterfly: 31 shuffles for the 32 feature channels) and issues
//    one coalesced reduction per warp instead of 41 atomics per pixel;
//  - backward starts at the tile's largest contributor count instead of the end
//    of the tile's list.
#include "common.cuh"
#include "render_common.cuh"

namespace tb {

struct Tuning {
  int fwd_ppt = 2;
  int bwd_ppt = 2;
  // 3 | 4: forward contraction on the warp-level tensor cores (render_mma.cu) with that many CTAs per SM.  Measured
  // slower than the SIMT kernel at config 2 (1.01 / 1.16 ms against 0.67 ms, profiles/r2_s_*): 3xTF32 on the legacy
  // mma.sync path (278 TFLOP/s measured) is 93 TFLOP/s effective, no better than the 74 TFLOP/s fp32 pipe that the SIMT
  // kernel drives with packed FFMA2 at half the issue slots: nothing left to pay for staging and splits.  Off.
  int fwd_mma = 0;
};
static Tuning g_tuning;
void set_tuning(const char* key, int value);
extern int g_sort_staged;
namespace mlp { extern int g_reserve_sms; }

// Blackwell packed FP32: one FFMA2 performs two independent IEEE fp32 FMAs on register pairs (same results as two
// FFMAs, half the issue slots — the compositing kernels are bound by instruction issue, not by the FMA pipe).
__device__ __forceinline__ void ffma2(float2& acc, const float2 a, const unsigned long long bb) {
  unsigned long long c = *reinterpret_cast<unsigned long long*>(&acc);
  const unsigned long long av = *reinterpret_cast<const unsigned long long*>(&a);
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(av), "l"(bb));
  acc = *reinterpret_cast<float2*>(&c);
}
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}

constexpr int FWD_BATCH = 128;
constexpr int BWD_BATCH = 256;
constexpr unsigned FULL = 0xffffffffu;

#ifndef TB_FWD_MINB
#define TB_FWD_MINB 4
#endif
template <int PPT>
__global__ void __launch_bounds__(256 / PPT, PPT == 2 ? TB_FWD_MINB : 2) render_fwd_kernel(const RenderFwdParams q) {
  pdl_wait();
  constexpr int NT = 256 / PPT;
  __shared__ float2 s_xy[FWD_BATCH];
  __shared__ float4 s_co[FWD_BATCH];
  __shared__ float4 s_rgbd[FWD_BATCH];
  __shared__ float4 s_feat[FWD_BATCH * 8];
  __shared__ float3 s_reach[FWD_BATCH];

  const int tile = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = q.ranges[tile];
  // each warp owns a compact 8 x (4*PPT) pixel block (small extent = few splats survive the per-warp
  // test); lane l sits in column l&7 and owns PPT vertically adjacent pixels
  const int wbx = ((threadIdx.x >> 5) & 1) * 8, wby = (threadIdx.x >> 6) * 4 * PPT;
  const int px = blockIdx.x * TILE + wbx + (threadIdx.x & 7);
  const int py0 = blockIdx.y * TILE + wby + ((threadIdx.x >> 3) & 3) * PPT;
  const float pxf = (float)px;
  const float wx0 = (float)(blockIdx.x * TILE + wbx), wx1 = wx0 + 7.f;
  const float wy0 = (float)(blockIdx.y * TILE + wby), wy1 = wy0 + (float)(4 * PPT - 1);

  float pyf[PPT], T[PPT], C[PPT][3], D[PPT];
  float2 O[PPT][NUM_OBJECTS / 2];      // feature accumulators as channel pairs (FFMA2)
  bool done[PPT], inside[PPT];
  uint32_t last[PPT];
#pragma unroll
  for (int p = 0; p < PPT; p++) {
    pyf[p] = (float)(py0 + p);
    inside[p] = (px < q.W) && (py0 + p < q.H);
    done[p] = !inside[p];
    T[p] = 1.0f; D[p] = 0.f; last[p] = 0;
    C[p][0] = C[p][1] = C[p][2] = 0.f;
#pragma unroll
    for (int ch = 0; ch < NUM_OBJECTS / 2; ch++) O[p][ch] = make_float2(0.f, 0.f);
  }

  for (uint32_t base = range.x; base < range.y; base += FWD_BATCH) {
    bool all_done = true;
#pragma unroll
    for (int p = 0; p < PPT; p++) all_done = all_done && done[p];
    if (__syncthreads_and(all_done)) break;

    const int n = min((uint32_t)FWD_BATCH, range.y - base);
    for (int k = threadIdx.x; k < n; k += NT) {
      const uint32_t id = q.point_list[base + k];
      s_xy[k] = q.means2D[id];
      const float4 co = q.conic_opacity[id];
      s_co[k] = co;
      s_reach[k] = splat_reach(co);
      s_rgbd[k] = q.rgbd[id];
    }
    for (int e = threadIdx.x; e < n * 8; e += NT) {
      const uint32_t id = q.point_list[base + (e >> 3)];
      s_feat[e] = __ldg(q.feat4 + (size_t)id * 8 + (e & 7));
    }
    __syncthreads();

    if (!__all_sync(0xffffffffu, all_done)) {   // warp-uniform: the block below uses full-warp ballots
      // 32 splats at a time: lane l tests splat sub+l against this warp's pixel block, survivors are
      // then composited pixel-parallel in list order
      const int lane = threadIdx.x & 31;
      bool still = true;
      for (int sub = 0; sub < n && still; sub += 32) {
        const int jj = sub + lane;
        bool cand = false;
        if (jj < n) cand = splat_may_touch(s_xy[jj], s_co[jj], s_reach[jj], wx0, wx1, wy0, wy1);
        unsigned cmask = __ballot_sync(0xffffffffu, cand);
        if (!cmask) continue;
        // software pipeline over the survivors: the exponent chain of the NEXT survivor (loads, quadratic form,
        // expf) is issued before the blend of the current one, so its latency hides under the 72 FMAs
        auto stage_a = [&](int jx, float (&ex)[PPT], float (&pw)[PPT], float& opac, float& thr) {
          const float2 xy = s_xy[jx];
          const float4 co = s_co[jx];
          opac = co.w;
          thr = s_reach[jx].z;
#pragma unroll
          for (int p = 0; p < PPT; p++) {
            float dx, dy;
            pw[p] = splat_power(xy, co, pxf, pyf[p], dx, dy);
            ex[p] = expf(pw[p]);
          }
        };
        int j = sub + __ffs(cmask) - 1;
        float ex[PPT], pw[PPT], opac, thr;
        stage_a(j, ex, pw, opac, thr);
        while (true) {
          cmask &= cmask - 1;
          const bool has_next = cmask != 0;
          int jn = j;
          float exn[PPT], pwn[PPT], opacn = 0.f, thrn = 0.f;
#pragma unroll
          for (int p = 0; p < PPT; p++) { exn[p] = 0.f; pwn[p] = 0.f; }
          if (has_next) {
            jn = sub + __ffs(cmask) - 1;
            stage_a(jn, exn, pwn, opacn, thrn);
          }
        float w[PPT];
        bool any = false;
        still = false;
#pragma unroll
        for (int p = 0; p < PPT; p++) {
          w[p] = 0.f;
          if (!done[p]) {
            if (!(pw[p] > 0.0f) && !(pw[p] < thr)) {
              const float alpha = fminf(0.99f, opac * ex[p]);
              if (!(alpha < 1.0f / 255.0f)) {
                const float test_T = T[p] * (1 - alpha);
                if (test_T < 0.0001f) {
                  done[p] = true;
                } else {
                  w[p] = alpha * T[p];
                  T[p] = test_T;
                  last[p] = (base - range.x) + j + 1;
                  any = true;
                }
              }
            }
          }
          still = still || !done[p];
        }
        if (any) {
          const float4 cd = s_rgbd[j];
#pragma unroll
          for (int p = 0; p < PPT; p++) {
            C[p][0] = fmaf(cd.x, w[p], C[p][0]);
            C[p][1] = fmaf(cd.y, w[p], C[p][1]);
            C[p][2] = fmaf(cd.z, w[p], C[p][2]);
            D[p] = fmaf(cd.w, w[p], D[p]);
          }
          unsigned long long ww[PPT];
#pragma unroll
          for (int p = 0; p < PPT; p++) ww[p] = pack2(w[p], w[p]);
#pragma unroll
          for (int c4 = 0; c4 < 8; c4++) {
            const float4 f = s_feat[j * 8 + c4];
#pragma unroll
            for (int p = 0; p < PPT; p++) {
              ffma2(O[p][2 * c4 + 0], make_float2(f.x, f.y), ww[p]);
              ffma2(O[p][2 * c4 + 1], make_float2(f.z, f.w), ww[p]);
            }
          }
        }
        // the candidate loop is warp-uniform (ballots): leave it only when the whole warp is finished
        still = __any_sync(0xffffffffu, still);
        if (!still || !has_next) break;
        j = jn; opac = opacn; thr = thrn;
#pragma unroll
        for (int p = 0; p < PPT; p++) { ex[p] = exn[p]; pw[p] = pwn[p]; }
        }
      }
    }
  }

  const size_t HW = (size_t)q.H * q.W;
  const float bg0 = q.bg[0], bg1 = q.bg[1], bg2 = q.bg[2];
#pragma unroll
  for (int p = 0; p < PPT; p++) {
    if (inside[p]) {
      const size_t pix = (size_t)q.W * (py0 + p) + px;
      q.final_T[pix] = T[p];
      q.n_contrib[pix] = last[p];
      q.out_color[pix] = C[p][0] + T[p] * bg0;
      q.out_color[HW + pix] = C[p][1] + T[p] * bg1;
      q.out_color[2 * HW + pix] = C[p][2] + T[p] * bg2;
#pragma unroll
      for (int ch = 0; ch < NUM_OBJECTS / 2; ch++) {
        q.out_objects[(2 * ch) * HW + pix] = O[p][ch].x;
        q.out_objects[(2 * ch + 1) * HW + pix] = O[p][ch].y;
      }
      q.out_depth[pix] = D[p];
    }
  }
}

cudaError_t launch_render_forward(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s) {
  RenderFwdParams q;
  q.ranges = p.ranges; q.point_list = p.point_list; q.W = a.W; q.H = a.H;
  q.means2D = p.means2D; q.rgbd = p.rgbd; q.feat4 = reinterpret_cast<const float4*>(a.sh_objs);
  q.conic_opacity = p.conic_opacity; q.final_T = p.final_T; q.n_contrib = p.n_contrib; q.bg = a.background;
  q.out_color = a.out_color; q.out_objects = a.out_objects; q.out_depth = a.out_depth;
  dim3 grid((a.W + TILE - 1) / TILE, (a.H + TILE - 1) / TILE, 1);
  if (g_tuning.fwd_mma) return launch_render_forward_mma(q, grid, g_tuning.fwd_mma, s);
  if (g_tuning.fwd_ppt == 1)
    launch_k(render_fwd_kernel<1>, grid, 256, 0, s, q);
  else
    launch_k(render_fwd_kernel<2>, grid, 128, 0, s, q);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Backward
struct RenderBwdParams {
  const uint2* ranges; const uint32_t* point_list;
  int W, H;
  const float* bg;
  const float2* means2D; const float4* conic_opacity; const float4* rgbd; const float4* feat4;
  const float* final_T; const uint32_t* n_contrib;
  const float* dL_dpix; const float* dL_dpix_obj;
  float* dL_dmean2D;  // [P][3]
  float* dL_dconic;   // [P][4]
  float* dL_dopacity; float* dL_dcolors; float* dL_dobjects;
};

// Sum v[i] over the 32 lanes for i = 0..31; on return lane l holds the total of
// channel l in v[0].  31 shuffles instead of 32*5.
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const bool hi = lane & 16;
    const float send = hi ? v[i] : v[i + 16];
    const float keep = hi ? v[i + 16] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const bool hi = lane & 8;
    const float send = hi ? v[i] : v[i + 8];
    const float keep = hi ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const bool hi = lane & 4;
    const float send = hi ? v[i] : v[i + 4];
    const float keep = hi ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const bool hi = lane & 2;
    const float send = hi ? v[i] : v[i + 2];
    const float keep = hi ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 2);
  }
  {
    const bool hi = lane & 1;
    const float send = hi ? v[0] : v[1];
    const float keep = hi ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(FULL, send, 1);
  }
  return v[0];
}
// Same for 16 values: lane l ends with the total of value (l & 15) … both
// halves of the warp hold a copy after the final full exchange.
__device__ __forceinline__ float warp_transpose_reduce16(float (&v)[16], int lane) {
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const bool hi = lane & 8;
    const float send = hi ? v[i] : v[i + 8];
    const float keep = hi ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const bool hi = lane & 4;
    const float send = hi ? v[i] : v[i + 4];
    const float keep = hi ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const bool hi = lane & 2;
    const float send = hi ? v[i] : v[i + 2];
    const float keep = hi ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(FULL, send, 2);
  }
  {
    const bool hi = lane & 1;
    const float send = hi ? v[0] : v[1];
    const float keep = hi ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(FULL, send, 1);
  }
  v[0] += __shfl_xor_sync(FULL, v[0], 16);
  return v[0];
}

// Which gradients the caller wants decides how much of the per-pair work exists at all (the reference always does
// all of it, backward.cu:392-574):
//   GEOM   : gradients of colour, opacity, 2D mean and conic (everything that needs dL/dalpha: the dot products with
//            the upstream gradient, the scalar recursion, the nine scalar reductions);
//   DF     : gradient of the per-Gaussian feature row, dF[j][c] += w * dL/dO[p][c] (32 FMAs per pair and the
//            32-channel cross-lane reduction) — not wanted while the feature parameter is frozen (GAUSSIAN state);
//   OBJ_UP : an upstream gradient of the feature image exists (false: the loss read the colour image only, the 32
//            feature channels drop out of dL/dalpha and the feature rows are not even staged).
// FEATURE state (only the feature parameter trains) is GEOM = false: alpha and T are recomputed and dF accumulated,
// nothing else.  Every variant evaluates alpha and T with the same pinned sequence, so which pairs blend never
// depends on the variant.
// GAUSSIAN-state variant: 128 registers without spills -> 4 CTAs = 16 warps per SM: 1.62 -> 1.43 ms at config 2
// (5 or 6 CTAs with a smaller staging batch spill and are slower: profiles/r2_aa_bwd_occupancy.txt)
#ifndef TB_BWD_G_MINB
#define TB_BWD_G_MINB 4
#endif
// FEATURE-state variant: 128 registers without spills as well: 0.44 -> 0.385 ms per 536x960 view
#ifndef TB_BWD_F_MINB
#define TB_BWD_F_MINB 4
#endif
#ifndef TB_BWD_C_MINB
#define TB_BWD_C_MINB 6
#endif
#ifndef TB_BWD_G_BATCH
#define TB_BWD_G_BATCH 256
#endif
// resident CTAs per SM the register allocation is held to (the kernel is bound by fixed-latency dependencies of the
// per-pixel chain, so resident warps are what buys throughput)
constexpr int bwd_min_blocks(int PPT, bool GEOM, bool DF, bool OBJ_UP) {
  if (PPT != 2) return (GEOM && !OBJ_UP) ? 3 : 2;
  if (GEOM && !OBJ_UP) return TB_BWD_C_MINB;   // colour-only loss: ~80 registers
  if (GEOM && !DF) return TB_BWD_G_MINB;    // GAUSSIAN state (no feature-row gradient)
  if (!GEOM) return TB_BWD_F_MINB;          // FEATURE state (feature-row gradient only)
  return 3;                                 // every gradient: 168 registers (128 would spill)
}
template <int PPT, bool GEOM, bool DF, bool OBJ_UP>
__global__ void __launch_bounds__(256 / PPT, bwd_min_blocks(PPT, GEOM, DF, OBJ_UP))
render_bwd_kernel(const RenderBwdParams q) {
  pdl_wait();      // the colour-only variant needs ~80 registers: twice the resident warps
  constexpr int NT = 256 / PPT;
  constexpr int BWD_BATCH = (GEOM && !DF && OBJ_UP && PPT == 2) ? TB_BWD_G_BATCH : tb::BWD_BATCH;
  constexpr bool NEED_FEAT = GEOM && OBJ_UP;        // feature rows are only read for dL/dalpha
  constexpr bool NEED_DLO = DF || NEED_FEAT;        // upstream feature gradient planes
  __shared__ uint32_t s_id[BWD_BATCH];
  __shared__ float2 s_xy[BWD_BATCH];
  __shared__ float4 s_co[BWD_BATCH];
  __shared__ float4 s_rgbd[GEOM ? BWD_BATCH : 1];
  __shared__ float4 s_feat[NEED_FEAT ? BWD_BATCH * 8 : 1];
  __shared__ float3 s_reach[BWD_BATCH];
  __shared__ uint32_t s_max[NT / 32];

  const int tile = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = q.ranges[tile];
  const int lane = threadIdx.x & 31;
  // each warp owns a compact 8 x (4*PPT) pixel block (small extent = few splats survive the per-warp
  // test); lane l sits in column l&7 and owns PPT vertically adjacent pixels
  const int wbx = ((threadIdx.x >> 5) & 1) * 8, wby = (threadIdx.x >> 6) * 4 * PPT;
  const int px = blockIdx.x * TILE + wbx + (threadIdx.x & 7);
  const int py0 = blockIdx.y * TILE + wby + ((threadIdx.x >> 3) & 3) * PPT;
  const float pxf = (float)px;
  const size_t HW = (size_t)q.H * q.W;
  const float wx0 = (float)(blockIdx.x * TILE + wbx), wx1 = wx0 + 7.f;
  const float wy0 = (float)(blockIdx.y * TILE + wby), wy1 = wy0 + (float)(4 * PPT - 1);

  float pyf[PPT], T[PPT], T_final[PPT], S[PPT], last_alpha[PPT], last_dot[PPT], bg_dot[PPT];
  float dLc[PPT][3];
  float2 dLo[PPT][NUM_OBJECTS / 2];     // upstream feature gradients as channel pairs (FFMA2 operands)
  uint32_t last[PPT];
  uint32_t my_max = 0;
#pragma unroll
  for (int p = 0; p < PPT; p++) {
    pyf[p] = (float)(py0 + p);
    const bool inside = (px < q.W) && (py0 + p < q.H);
    const size_t pix = (size_t)q.W * (py0 + p) + px;
    T_final[p] = inside ? q.final_T[pix] : 0.f;
    T[p] = T_final[p];
    last[p] = inside ? q.n_contrib[pix] : 0u;
    my_max = max(my_max, last[p]);
    S[p] = 0.f; last_alpha[p] = 0.f; last_dot[p] = 0.f;
#pragma unroll
    for (int c = 0; c < 3; c++) dLc[p][c] = (GEOM && inside) ? q.dL_dpix[c * HW + pix] : 0.f;
#pragma unroll
    for (int ch = 0; ch < NUM_OBJECTS / 2; ch++)
      dLo[p][ch] = (NEED_DLO && inside)
                       ? make_float2(q.dL_dpix_obj[(2 * ch) * HW + pix], q.dL_dpix_obj[(2 * ch + 1) * HW + pix])
                       : make_float2(0.f, 0.f);
    bg_dot[p] = GEOM ? q.bg[0] * dLc[p][0] + q.bg[1] * dLc[p][1] + q.bg[2] * dLc[p][2] : 0.f;
  }
  // largest contributor count of the tile: nothing behind it received any weight
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_max = max(my_max, __shfl_xor_sync(FULL, my_max, o));
  const uint32_t warp_last = my_max;   // nothing at or beyond this index contributes to any pixel of this warp
  if (lane == 0) s_max[threadIdx.x >> 5] = my_max;
  __syncthreads();
  uint32_t tile_max = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; w++) tile_max = max(tile_max, s_max[w]);
  const uint32_t n_total = min(tile_max, range.y - range.x);

  const float ddelx_dx = 0.5 * q.W;
  const float ddely_dy = 0.5 * q.H;

  // walk the list back to front in batches; element e (0-based in the tile's
  // list) is visited by pixel p iff e < last[p]
  for (int hi = (int)n_total; hi > 0; hi -= BWD_BATCH) {
    const int n = min(BWD_BATCH, hi);
    __syncthreads();
    // s_*[k] holds element e = hi - 1 - k
    for (int k = threadIdx.x; k < n; k += NT) {
      const uint32_t id = q.point_list[range.x + hi - 1 - k];
      s_id[k] = id;
      s_xy[k] = q.means2D[id];
      const float4 co = q.conic_opacity[id];
      s_co[k] = co;
      s_reach[k] = splat_reach(co);
      if (GEOM) s_rgbd[k] = q.rgbd[id];
    }
    if (NEED_FEAT) {
      for (int e = threadIdx.x; e < n * 8; e += NT) {
        const uint32_t id = q.point_list[range.x + hi - 1 - (e >> 3)];
        s_feat[e] = __ldg(q.feat4 + (size_t)id * 8 + (e & 7));
      }
    }
    __syncthreads();

    for (int sub = 0; sub < n; sub += 32) {
      // lane l tests splat sub+l (list element e = hi-1-(sub+l)) against this warp's pixel block
      bool cand = false;
      {
        const int kk = sub + lane;
        if (kk < n && (uint32_t)(hi - 1 - kk) < warp_last)
          cand = splat_may_touch(s_xy[kk], s_co[kk], s_reach[kk], wx0, wx1, wy0, wy1);
      }
      unsigned cmask = __ballot_sync(FULL, cand);
      while (cmask) {
      const int k = sub + __ffs(cmask) - 1;
      cmask &= cmask - 1;
      const uint32_t e = (uint32_t)(hi - 1 - k);
      const float2 xy = s_xy[k];
      const float3 reach = s_reach[k];
      const float4 co = s_co[k];
      float w[PPT], G[PPT], dx[PPT], dy[PPT], alpha[PPT];
      bool act[PPT];
      bool any = false;
#pragma unroll
      for (int p = 0; p < PPT; p++) {
        act[p] = false;
        if (e < last[p]) {
          const float power = splat_power(xy, co, pxf, pyf[p], dx[p], dy[p]);
          if (!(power > 0.0f) && !(power < reach.z)) {
            G[p] = expf(power);
            alpha[p] = fminf(0.99f, co.w * G[p]);
            if (!(alpha[p] < 1.0f / 255.0f)) {
              T[p] = T[p] / (1.f - alpha[p]);
              w[p] = alpha[p] * T[p];
              act[p] = true;
              any = true;
            }
          }
        }
      }
      if (!__any_sync(FULL, any)) continue;

      float2 vobj2[16];
      float vs[16];
#pragma unroll
      for (int i = 0; i < 16; i++) vobj2[i] = make_float2(0.f, 0.f);
#pragma unroll
      for (int i = 0; i < 16; i++) vs[i] = 0.f;

      if (any && !GEOM) {
        // feature gradient only: dF[j][c] += w * dL/dO[p][c]
#pragma unroll
        for (int p = 0; p < PPT; p++) {
          if (act[p]) {
            const unsigned long long ww = pack2(w[p], w[p]);
#pragma unroll
            for (int ch = 0; ch < NUM_OBJECTS / 2; ch++) ffma2(vobj2[ch], dLo[p][ch], ww);
          }
        }
      }
      if (any && GEOM) {
        const float4 cd = s_rgbd[k];
        float dot[PPT];
#pragma unroll
        for (int p = 0; p < PPT; p++) {
          dot[p] = cd.x * dLc[p][0];
          dot[p] = fmaf(cd.y, dLc[p][1], dot[p]);
          dot[p] = fmaf(cd.z, dLc[p][2], dot[p]);
        }
        if (NEED_FEAT) {
          // two independent packed accumulators per pixel: dependent-FMA chains of 8 instead of 16 (the kernel is
          // bound by fixed-latency dependencies at 12 warps per SM, not by the FMA pipe)
          float2 dot2[PPT], dot2b[PPT];
#pragma unroll
          for (int p = 0; p < PPT; p++) { dot2[p] = make_float2(0.f, 0.f); dot2b[p] = make_float2(0.f, 0.f); }
#pragma unroll
          for (int c4 = 0; c4 < 8; c4++) {
            const float4 f = s_feat[k * 8 + c4];
#pragma unroll
            for (int p = 0; p < PPT; p++) {
              ffma2(dot2[p], make_float2(f.x, f.y), *reinterpret_cast<const unsigned long long*>(&dLo[p][2 * c4 + 0]));
              ffma2(dot2b[p], make_float2(f.z, f.w), *reinterpret_cast<const unsigned long long*>(&dLo[p][2 * c4 + 1]));
            }
          }
#pragma unroll
          for (int p = 0; p < PPT; p++) dot[p] += (dot2[p].x + dot2b[p].x) + (dot2[p].y + dot2b[p].y);
        }
#pragma unroll
        for (int p = 0; p < PPT; p++) {
          if (act[p]) {
            // scalar form of accum_rec: S = last_alpha*last_dot + (1-last_alpha)*S
            S[p] = last_alpha[p] * last_dot[p] + (1.f - last_alpha[p]) * S[p];
            last_dot[p] = dot[p];
            float dL_dalpha = (dot[p] - S[p]) * T[p];
            last_alpha[p] = alpha[p];
            dL_dalpha += (-T_final[p] / (1.f - alpha[p])) * bg_dot[p];

            const float dL_dG = co.w * dL_dalpha;
            const float gdx = G[p] * dx[p];
            const float gdy = G[p] * dy[p];
            const float dG_ddelx = -gdx * co.x - gdy * co.y;
            const float dG_ddely = -gdy * co.z - gdx * co.y;
            vs[0] += w[p] * dLc[p][0];
            vs[1] += w[p] * dLc[p][1];
            vs[2] += w[p] * dLc[p][2];
            vs[3] += dL_dG * dG_ddelx * ddelx_dx;
            vs[4] += dL_dG * dG_ddely * ddely_dy;
            vs[5] += -0.5f * gdx * dx[p] * dL_dG;
            vs[6] += -0.5f * gdx * dy[p] * dL_dG;
            vs[7] += -0.5f * gdy * dy[p] * dL_dG;
            vs[8] += G[p] * dL_dalpha;
            if (DF) {
              const unsigned long long ww = pack2(w[p], w[p]);
#pragma unroll
              for (int ch = 0; ch < NUM_OBJECTS / 2; ch++) ffma2(vobj2[ch], dLo[p][ch], ww);
            }
          }
        }
      }
      const uint32_t id = s_id[k];
      if (DF) {
        float vobj[32];
#pragma unroll
        for (int i = 0; i < 16; i++) { vobj[2 * i] = vobj2[i].x; vobj[2 * i + 1] = vobj2[i].y; }
        const float tot_obj = warp_transpose_reduce32(vobj, lane);
        atomicAdd(q.dL_dobjects + (size_t)id * NUM_OBJECTS + lane, tot_obj);
      }
      if (!GEOM) continue;
      const float tot_s = warp_transpose_reduce16(vs, lane);
      if (lane < 9) {
        float* dst;
        if (lane < 3) dst = q.dL_dcolors + (size_t)id * 3 + lane;
        else if (lane < 5) dst = q.dL_dmean2D + (size_t)id * 3 + (lane - 3);
        else if (lane < 7) dst = q.dL_dconic + (size_t)id * 4 + (lane - 5);
        else if (lane == 7) dst = q.dL_dconic + (size_t)id * 4 + 3;
        else dst = q.dL_dopacity + id;
        atomicAdd(dst, tot_s);
      }
      }
    }
  }
}

cudaError_t launch_render_backward(const trase_b200_bwd_args& a, const FwdPointers& p, cudaStream_t s) {
  RenderBwdParams q;
  q.ranges = p.ranges; q.point_list = p.point_list; q.W = a.W; q.H = a.H; q.bg = a.background;
  q.means2D = p.means2D; q.conic_opacity = p.conic_opacity; q.rgbd = p.rgbd;
  q.feat4 = reinterpret_cast<const float4*>(a.sh_objs);
  q.final_T = p.final_T; q.n_contrib = p.n_contrib;
  q.dL_dpix = a.dL_dout_color; q.dL_dpix_obj = a.dL_dout_objects;
  q.dL_dmean2D = a.dL_dmeans2D; q.dL_dconic = a.dL_dconic; q.dL_dopacity = a.dL_dopacity;
  q.dL_dcolors = a.dL_dcolors; q.dL_dobjects = a.dL_dobjects;
  dim3 grid((a.W + TILE - 1) / TILE, (a.H + TILE - 1) / TILE, 1);
  // what the caller asked for (NULL = not wanted / not provided; api.cu has validated the combinations)
  const bool geom = a.dL_dmeans2D != nullptr, df = a.dL_dobjects != nullptr, obj_up = a.dL_dout_objects != nullptr;
  if (!geom && (!df || !obj_up)) return cudaSuccess;      // nothing to compute
  const bool one = g_tuning.bwd_ppt == 1;
#define TB_BWD(G_, D_, O_)                                                    \
  do {                                                                        \
    if (one) launch_k(render_bwd_kernel<1, G_, D_, O_>, grid, 256, 0, s, q);        \
    else launch_k(render_bwd_kernel<2, G_, D_, O_>, grid, 128, 0, s, q);            \
  } while (0)
  if (!geom) TB_BWD(false, true, true);
  else if (df && obj_up) TB_BWD(true, true, true);
  else if (obj_up) TB_BWD(true, false, true);
  else TB_BWD(true, false, false);        // no upstream feature gradient: dF would be zero as well
#undef TB_BWD
  return cudaGetLastError();
}

void set_tuning(const char* key, int value) {
  auto eq = [](const char* a, const char* b) {
    while (*a && *b && *a == *b) { a++; b++; }
    return *a == 0 && *b == 0;
  };
  if (eq(key, "fwd_ppt")) g_tuning.fwd_ppt = value;
  if (eq(key, "bwd_ppt")) g_tuning.bwd_ppt = value;
  if (eq(key, "fwd_mma")) g_tuning.fwd_mma = value;
  if (eq(key, "sort_staged")) g_sort_staged = value;
  if (eq(key, "reserve_sms")) mlp::g_reserve_sms = value < 0 ? 0 : value;
}

}  // namespace tb
