// Forward compositing with the per-channel contraction on the warp-level tensor cores.
//
// Reference counterpart: renderCUDA forward, DGR/cuda_rasterizer/forward.cu:264-396.
//
// The blend of a tile is C[p][c] = sum_j w[p][j] * F[j][c] over the tile's splats j, with w = alpha * T from the
// per-pixel sequential chain (which stays exactly as in render_fwd_kernel: same pinned alpha sequence, same thresholds,
// same early termination, so n_contrib and final_T are bit-identical) and F[j] the 36 blended values of splat j
// (rgb, depth, 32 features).  render_fwd_kernel spends 40 FMA issue slots and 8 broadcast LDS.128 per (warp, splat) on
// that product; here a warp collects the weights of 16 surviving splats for its 64 pixels in a 4 KB staging buffer and
// issues the [64 x 16] x [16 x 40] product as mma.sync.m16n8k8 TF32 instructions, accumulators in registers.
// Precision: TF32 carries 11 significant bits, the image tolerance is 1e-4 absolute, so both factors are split into a
// TF32 "hi" part and a TF32 remainder (3 MMAs per product: hi*hi + lo*hi + hi*lo, error ~ 2^-21 relative); the
// feature rows are split once when the batch is staged.
// (tcgen05 is not used here on purpose — DESIGN.md §3f: K = 16, N = 40 tiles with operands produced by a sequential
// SIMT chain cannot feed it.)
//
// MEASURED RESULT (config 2, profiles/r2_s_*): 1.01 ms (3 CTAs/SM) / 1.16 ms (4 CTAs/SM, spills) against 0.67 ms for
// render_fwd_kernel — results identical within the tolerances (all parity tests pass with this kernel selected), but
// slower.  mma.sync TF32 peaks at 278 TFLOP/s on this GPU (tools/micro/mma_sync_rate.cu, profiles/r2_r_mma_sync_rate.txt);
// with the 3-way split and half-filled batches at the end of every staged block the kernel issues ~60 M MMAs = 0.12
// TFLOP per view, 0.44 ms of tensor time alone, and 3xTF32 at 278 TFLOP/s is 93 TFLOP/s effective — no more than the
// 74 TFLOP/s fp32 pipe that the SIMT kernel drives with packed FFMA2 at half the issue slots, so nothing is left to
// pay for the staging, the splits and the fragment loads.  The kernel stays selectable
// (trase_b200_set_tuning("fwd_mma", 3|4)) and tested, off by default.
#include "common.cuh"
#include "render_common.cuh"

namespace tb {

namespace fmma {

constexpr int BATCH = 64;          // splats staged per block iteration
constexpr int NCH = 40;            // 3 colour + depth + 32 features, padded to 5 n-tiles of 8
constexpr int LDF = 40;            // row pitch of the feature tiles: 40 = 8 (mod 32) -> conflict-free B fragments
constexpr int KB = 16;             // splats per MMA batch (2 k-steps of 8)
constexpr int LDW = 72;            // row pitch of the weight staging [16][64]: 72 = 8 (mod 32) -> conflict-free A fragments
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float tf32_hi(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const float (&a)[4], const float b0, const float b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
        "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

// acc[mt][nt][4] += W^T (64 pixels x 16 splats, from wbuf) * F (16 splats x 40 channels, rows slot -> s_f*[wj[slot]])
__device__ __forceinline__ void flush_batch(float (&acc)[4][5][4], const float* __restrict__ wbuf,
                                            const int* __restrict__ wj, const float* __restrict__ s_fhi,
                                            const float* __restrict__ s_flo, int lane) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int ks = 0; ks < 2; ks++) {
    const int j0 = wj[ks * 8 + t], j1 = wj[ks * 8 + t + 4];           // B[row = splat slot][col = channel]
#pragma unroll
    for (int mh = 0; mh < 2; mh++) {                                  // two m-tiles at a time: 16 A registers live
      float ahi[2][4], alo[2][4];
#pragma unroll
      for (int m2 = 0; m2 < 2; m2++) {
        const float* w0 = wbuf + (ks * 8 + t) * LDW + (mh * 2 + m2) * 16 + g;   // A[row = pixel][col = splat slot]
        const float v[4] = {w0[0], w0[8], w0[4 * LDW], w0[4 * LDW + 8]};
#pragma unroll
        for (int i = 0; i < 4; i++) {
          ahi[m2][i] = tf32_hi(v[i]);
          alo[m2][i] = tf32_hi(v[i] - ahi[m2][i]);
        }
      }
#pragma unroll
      for (int nt = 0; nt < 5; nt++) {
        const float bh0 = s_fhi[j0 * LDF + nt * 8 + g], bh1 = s_fhi[j1 * LDF + nt * 8 + g];
        const float bl0 = s_flo[j0 * LDF + nt * 8 + g], bl1 = s_flo[j1 * LDF + nt * 8 + g];
#pragma unroll
        for (int m2 = 0; m2 < 2; m2++) {
          mma_tf32(acc[mh * 2 + m2][nt], ahi[m2], bh0, bh1);
          mma_tf32(acc[mh * 2 + m2][nt], alo[m2], bh0, bh1);
          mma_tf32(acc[mh * 2 + m2][nt], ahi[m2], bl0, bl1);
        }
      }
    }
  }
}

constexpr int SMEM_BYTES = BATCH * (8 + 16 + 12) + 2 * BATCH * LDF * 4 + 4 * (KB * LDW * 4 + KB * 4) + 4 * 64 * 4;

template <int MINB>
__global__ void __launch_bounds__(128, MINB) render_fwd_mma_kernel(const RenderFwdParams q) {
  extern __shared__ __align__(16) unsigned char smem[];
  float4* s_co = reinterpret_cast<float4*>(smem);                               // [BATCH]
  float2* s_xy = reinterpret_cast<float2*>(s_co + BATCH);                        // [BATCH]
  float3* s_reach = reinterpret_cast<float3*>(s_xy + BATCH);                     // [BATCH]
  float* s_fhi = reinterpret_cast<float*>(s_reach + BATCH);                      // [BATCH][LDF]
  float* s_flo = s_fhi + BATCH * LDF;                                            // [BATCH][LDF]
  float* s_w = s_flo + BATCH * LDF;                                              // [4 warps][KB][LDW]
  int* s_wj = reinterpret_cast<int*>(s_w + 4 * KB * LDW);                        // [4 warps][KB]
  float* s_T = reinterpret_cast<float*>(s_wj + 4 * KB);                          // [4 warps][64] final T per pixel

  const int tile = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = q.ranges[tile];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // each warp owns a compact 8 x 8 pixel block; lane l sits in column l&7 and owns 2 vertically adjacent pixels;
  // the warp's pixel number (row of the MMA's A operand) is  pq = lane + 32 * p
  const int wbx = (warp & 1) * 8, wby = (warp >> 1) * 8;
  const int px = blockIdx.x * TILE + wbx + (lane & 7);
  const int py0 = blockIdx.y * TILE + wby + ((lane >> 3) & 3) * 2;
  const float pxf = (float)px;
  const float wx0 = (float)(blockIdx.x * TILE + wbx), wx1 = wx0 + 7.f;
  const float wy0 = (float)(blockIdx.y * TILE + wby), wy1 = wy0 + 7.f;
  float* wbuf = s_w + warp * KB * LDW;
  int* wj = s_wj + warp * KB;

  float pyf[2], T[2];
  bool done[2], inside[2];
  uint32_t last[2];
#pragma unroll
  for (int p = 0; p < 2; p++) {
    pyf[p] = (float)(py0 + p);
    inside[p] = (px < q.W) && (py0 + p < q.H);
    done[p] = !inside[p];
    T[p] = 1.0f;
    last[p] = 0;
  }
  float acc[4][5][4];
#pragma unroll
  for (int mt = 0; mt < 4; mt++)
#pragma unroll
    for (int nt = 0; nt < 5; nt++)
#pragma unroll
      for (int i = 0; i < 4; i++) acc[mt][nt][i] = 0.f;
  for (int e = lane; e < KB * LDW; e += 32) wbuf[e] = 0.f;          // unused slots multiply as zero weights
  if (lane < KB) wj[lane] = 0;
  int cnt = 0;                                                      // filled slots of the warp's current MMA batch

  for (uint32_t base = range.x; base < range.y; base += BATCH) {
    const bool all_done = done[0] && done[1];
    if (__syncthreads_and(all_done)) break;
    const int n = min((uint32_t)BATCH, range.y - base);
    for (int k = threadIdx.x; k < n; k += 128) {
      const uint32_t id = q.point_list[base + k];
      s_xy[k] = q.means2D[id];
      const float4 co = q.conic_opacity[id];
      s_co[k] = co;
      s_reach[k] = splat_reach(co);
      const float4 cd = q.rgbd[id];
      const float v[4] = {cd.x, cd.y, cd.z, cd.w};
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const float h = tf32_hi(v[c]);
        s_fhi[k * LDF + c] = h;
        s_flo[k * LDF + c] = tf32_hi(v[c] - h);
        s_fhi[k * LDF + 36 + c] = 0.f;                              // padding channels 36..39
        s_flo[k * LDF + 36 + c] = 0.f;
      }
    }
    for (int e = threadIdx.x; e < n * 8; e += 128) {
      const int k = e >> 3, c4 = e & 7;
      const uint32_t id = q.point_list[base + k];
      const float4 f = __ldg(q.feat4 + (size_t)id * 8 + c4);
      const float v[4] = {f.x, f.y, f.z, f.w};
      float4 h4, l4;
      float* hp = reinterpret_cast<float*>(&h4);
      float* lp = reinterpret_cast<float*>(&l4);
#pragma unroll
      for (int c = 0; c < 4; c++) { hp[c] = tf32_hi(v[c]); lp[c] = tf32_hi(v[c] - hp[c]); }
      *reinterpret_cast<float4*>(s_fhi + k * LDF + 4 + c4 * 4) = h4;
      *reinterpret_cast<float4*>(s_flo + k * LDF + 4 + c4 * 4) = l4;
    }
    __syncthreads();

    if (!__all_sync(FULL, all_done)) {   // warp-uniform
      bool still = true;
      for (int sub = 0; sub < n && still; sub += 32) {
        const int jj = sub + lane;
        bool cand = false;
        if (jj < n) cand = splat_may_touch(s_xy[jj], s_co[jj], s_reach[jj], wx0, wx1, wy0, wy1);
        unsigned cmask = __ballot_sync(FULL, cand);
        while (cmask && still) {
          const int j = sub + __ffs(cmask) - 1;
          cmask &= cmask - 1;
          const float2 xy = s_xy[j];
          const float4 co = s_co[j];
          const float thr = s_reach[j].z;
          float w[2];
          bool any = false;
          still = false;
#pragma unroll
          for (int p = 0; p < 2; p++) {
            w[p] = 0.f;
            if (!done[p]) {
              float dx_, dy_;
              const float pw = splat_power(xy, co, pxf, pyf[p], dx_, dy_);
              if (!(pw > 0.0f) && !(pw < thr)) {
                const float alpha = fminf(0.99f, co.w * expf(pw));
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
          if (__any_sync(FULL, any)) {
            wbuf[cnt * LDW + lane] = w[0];
            wbuf[cnt * LDW + 32 + lane] = w[1];
            if (lane == 0) wj[cnt] = j;
            cnt++;
            if (cnt == KB) {
              __syncwarp();
              flush_batch(acc, wbuf, wj, s_fhi, s_flo, lane);
              __syncwarp();
              cnt = 0;
            }
          }
          still = __any_sync(FULL, still);
        }
      }
    }
    // the staged feature rows are about to be replaced: multiply what the warp has collected so far
    if (cnt > 0) {
      for (int e = cnt * LDW + lane; e < KB * LDW; e += 32) wbuf[e] = 0.f;      // unused slots: zero weights ...
      if (lane >= cnt && lane < KB) wj[lane] = 0;                               // ... on a valid feature row
      __syncwarp();
      flush_batch(acc, wbuf, wj, s_fhi, s_flo, lane);
      __syncwarp();
      cnt = 0;
    }
    __syncthreads();
  }

  // ---- outputs.  final_T / n_contrib from the pixel owners; the blended channels from the accumulator fragments:
  // lane (g = lane >> 2, t = lane & 3) holds pixel rows mt*16 + g (+8), channels nt*8 + 2t (+1)
  const size_t HW = (size_t)q.H * q.W;
  float* sT = s_T + warp * 64;
#pragma unroll
  for (int p = 0; p < 2; p++) {
    sT[lane + 32 * p] = T[p];
    if (inside[p]) {
      const size_t pix = (size_t)q.W * (py0 + p) + px;
      q.final_T[pix] = T[p];
      q.n_contrib[pix] = last[p];
    }
  }
  __syncwarp();
  const float bg[3] = {q.bg[0], q.bg[1], q.bg[2]};
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int mt = 0; mt < 4; mt++) {
#pragma unroll
    for (int half = 0; half < 2; half++) {
      const int pq = mt * 16 + g + half * 8;                 // warp pixel number -> owner lane pq & 31, p = pq >> 5
      const int ol = pq & 31, op = pq >> 5;
      const int x = blockIdx.x * TILE + wbx + (ol & 7);
      const int y = blockIdx.y * TILE + wby + ((ol >> 3) & 3) * 2 + op;
      if (x >= q.W || y >= q.H) continue;
      const size_t pix = (size_t)q.W * y + x;
      const float Tf = sT[pq];
#pragma unroll
      for (int nt = 0; nt < 5; nt++) {
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const int ch = nt * 8 + 2 * t + i;
          const float v = acc[mt][nt][half * 2 + i];
          if (ch < 3) q.out_color[ch * HW + pix] = v + Tf * bg[ch];
          else if (ch == 3) q.out_depth[pix] = v;
          else if (ch < 36) q.out_objects[(size_t)(ch - 4) * HW + pix] = v;
        }
      }
    }
  }
}

}  // namespace fmma

cudaError_t launch_render_forward_mma(const RenderFwdParams& q, dim3 grid, int min_blocks, cudaStream_t s) {
  // min_blocks: 3 = 168 registers, no spills, 12 warps per SM; 4 = 128 registers, 16 warps per SM
  if (min_blocks >= 4)
    fmma::render_fwd_mma_kernel<4><<<grid, 128, fmma::SMEM_BYTES, s>>>(q);
  else
    fmma::render_fwd_mma_kernel<3><<<grid, 128, fmma::SMEM_BYTES, s>>>(q);
  return cudaGetLastError();
}

}  // namespace tb
