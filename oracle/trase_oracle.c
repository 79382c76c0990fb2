/*
 * TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
 *
 * CPU restatement (plain C, fp32, OpenMP over Gaussians / tiles) of the
 * reference rasterizer's algorithm.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / reference legs may load this; the product path
 * (trase_b200/) never does.
 *
 * Follows, function by function (paths under
 * submodules/diff-gaussian-rasterization/cuda_rasterizer/):
 *   oracle_preprocess_forward   forward.cu:23-259, auxiliary.h:39-52,134-159
 *   oracle_bin                  rasterizer_impl.cu:73-141, 284-326 (scan, keys, stable sort, ranges)
 *   oracle_render_forward       forward.cu:264-396
 *   oracle_render_backward      backward.cu:392-574
 *   oracle_preprocess_backward  backward.cu:18-389
 *   oracle_mark_visible         rasterizer_impl.cu:57-69
 *
 * Parity pinning: this file is checked against golden vectors produced by the
 * UNMODIFIED reference CUDA rasterizer run on a B200 (tests/golden/, generated
 * by tests/golden/make_golden.py from oracle/_ref).  Floating point: compiled
 * with -ffp-contract=off; fmaf() is written exactly where the reference's
 * compiled kernel fuses, so that radii / tile rects / keys come out identical.
 * expf() is libm's (CUDA's is a 2-ulp approximation), so per-pixel contributor
 * counts can differ from the GPU on isolated pixels; tests bound that.
 * Gradient accumulation across pixels is done in double (the reference uses
 * float atomics in nondeterministic order).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16
#define NOBJ 32

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                               -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

static inline float dot3p(float a1, float b1, float a2, float b2, float a3, float b3) {
  return fmaf(a3, b3, fmaf(a1, b1, a2 * b2));
}
static inline void xform4x3(const float* p, const float* m, float* o) {
  o[0] = dot3p(m[0], p[0], m[4], p[1], m[8], p[2]) + m[12];
  o[1] = dot3p(m[1], p[0], m[5], p[1], m[9], p[2]) + m[13];
  o[2] = dot3p(m[2], p[0], m[6], p[1], m[10], p[2]) + m[14];
}
static inline void xform4x4(const float* p, const float* m, float* o) {
  o[0] = dot3p(m[0], p[0], m[4], p[1], m[8], p[2]) + m[12];
  o[1] = dot3p(m[1], p[0], m[5], p[1], m[9], p[2]) + m[13];
  o[2] = dot3p(m[2], p[0], m[6], p[1], m[10], p[2]) + m[14];
  o[3] = dot3p(m[3], p[0], m[7], p[1], m[11], p[2]) + m[15];
}
static inline float ndc2pix(float v, int S) { return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5); }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

static void tile_rect(float px, float py, int r, int gx, int gy, int* x0, int* y0, int* x1, int* y1) {
  *x0 = imin(gx, imax(0, (int)((px - r) / TILE)));
  *y0 = imin(gy, imax(0, (int)((py - r) / TILE)));
  *x1 = imin(gx, imax(0, (int)((px + r + TILE - 1) / TILE)));
  *y1 = imin(gy, imax(0, (int)((py + r + TILE - 1) / TILE)));
}

/* cov3D = upper triangle of (S R)^T (S R); quaternion (r,x,y,z) not normalised */
static void cov3d_from_scale_rot(const float* sc, float mod, const float* q, float* c) {
  float sx = mod * sc[0], sy = mod * sc[1], sz = mod * sc[2];
  float r = q[0], x = q[1], y = q[2], z = q[3];
  float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
  float R00 = 1.f - 2.f * (yy + zz), R01 = 2.f * fmaf(x, y, -rz), R02 = 2.f * fmaf(r, y, xz);
  float R10 = 2.f * fmaf(x, y, rz), R11 = 1.f - 2.f * fmaf(x, x, zz), R12 = 2.f * fmaf(y, z, -rx);
  float R20 = 2.f * fmaf(-r, y, xz), R21 = 2.f * fmaf(y, z, rx), R22 = 1.f - 2.f * fmaf(x, x, yy);
  float M00 = sx * R00, M01 = sy * R01, M02 = sz * R02;
  float M10 = sx * R10, M11 = sy * R11, M12 = sz * R12;
  float M20 = sx * R20, M21 = sy * R21, M22 = sz * R22;
  c[0] = dot3p(M00, M00, M01, M01, M02, M02);
  c[1] = dot3p(M10, M00, M11, M01, M12, M02);
  c[2] = dot3p(M20, M00, M21, M01, M22, M02);
  c[3] = dot3p(M10, M10, M11, M11, M12, M12);
  c[4] = dot3p(M20, M10, M21, M11, M22, M12);
  c[5] = dot3p(M20, M20, M21, M21, M22, M22);
}

/* T = W J and cov2D = T^T Vrk^T T (+0.3 on the diagonal); also returns T (2x3) */
static void cov2d_from_cov3d(const float* mean, float fx, float fy, float tanx, float tany, const float* c,
                             const float* v, float* cov, float* Tm, float* t_out, float* txtz_o, float* tytz_o) {
  float t[3];
  xform4x3(mean, v, t);
  float limx = 1.3f * tanx, limy = 1.3f * tany;
  float txtz = t[0] / t[2], tytz = t[1] / t[2];
  t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
  t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
  float tz2 = t[2] * t[2];
  float J00 = fx / t[2], J02 = -(fx * t[0]) / tz2, J11 = fy / t[2], J12 = -(fy * t[1]) / tz2;
  float T00 = fmaf(v[2], J02, v[0] * J00), T01 = fmaf(v[6], J02, v[4] * J00), T02 = fmaf(v[10], J02, v[8] * J00);
  float T10 = fmaf(v[2], J12, v[1] * J11), T11 = fmaf(v[6], J12, v[5] * J11), T12 = fmaf(v[10], J12, v[9] * J11);
  float X00 = dot3p(T00, c[0], T01, c[1], T02, c[2]);
  float X10 = dot3p(T00, c[1], T01, c[3], T02, c[4]);
  float X20 = dot3p(T00, c[2], T01, c[4], T02, c[5]);
  float X01 = dot3p(T10, c[0], T11, c[1], T12, c[2]);
  float X11 = dot3p(T10, c[1], T11, c[3], T12, c[4]);
  float X21 = dot3p(T10, c[2], T11, c[4], T12, c[5]);
  cov[0] = dot3p(T00, X00, T01, X10, T02, X20) + 0.3f;
  cov[1] = dot3p(T00, X01, T01, X11, T02, X21);
  cov[2] = dot3p(T10, X01, T11, X11, T12, X21) + 0.3f;
  if (Tm) { Tm[0] = T00; Tm[1] = T01; Tm[2] = T02; Tm[3] = T10; Tm[4] = T11; Tm[5] = T12; }
  if (t_out) { t_out[0] = t[0]; t_out[1] = t[1]; t_out[2] = t[2]; }
  if (txtz_o) *txtz_o = txtz;
  if (tytz_o) *tytz_o = tytz;
}

static void sh_to_rgb(int deg, const float* pos, const float* cam, const float* sh, float* rgb, uint8_t* clamped) {
  float d[3] = {pos[0] - cam[0], pos[1] - cam[1], pos[2] - cam[2]};
  float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  float x = d[0] / len, y = d[1] / len, z = d[2] / len;
  for (int c = 0; c < 3; c++) {
#define S(i) sh[3 * (i) + c]
    float res = SH_C0 * S(0);
    if (deg > 0) {
      res = res - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = res + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
              SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          res = res + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
                SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
#undef S
    res += 0.5f;
    clamped[c] = res < 0;
    rgb[c] = res < 0 ? 0.0f : res;
  }
}

/* forward.cu:157-259.  All outputs sized P; entries of culled Gaussians are
 * left untouched except radii / tiles_touched which are zeroed (forward.cu:191-192). */
void oracle_preprocess_forward(int P, int D, int M, const float* means, const float* scales, float mod,
                               const float* rots, const float* opac, const float* shs, const float* cov3D_precomp,
                               const float* colors_precomp, const float* view, const float* proj, const float* campos,
                               int W, int H, float tanx, float tany, int* radii, float* means2D, float* depths,
                               float* cov3Ds, float* rgb, float* conic_opacity, uint32_t* tiles_touched,
                               uint8_t* clamped) {
  const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    radii[i] = 0;
    tiles_touched[i] = 0;
    const float* p = means + 3 * i;
    float pv[3], ph[4];
    xform4x3(p, view, pv);
    if (pv[2] <= 0.2f) continue;
    xform4x4(p, proj, ph);
    float pw = 1.0f / (ph[3] + 0.0000001f);
    float projx = ph[0] * pw, projy = ph[1] * pw;
    float c3[6];
    if (cov3D_precomp) memcpy(c3, cov3D_precomp + 6 * i, sizeof(c3));
    else { cov3d_from_scale_rot(scales + 3 * i, mod, rots + 4 * i, c3); memcpy(cov3Ds + 6 * i, c3, sizeof(c3)); }
    float cov[3];
    cov2d_from_cov3d(p, fx, fy, tanx, tany, c3, view, cov, 0, 0, 0, 0);
    float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
    if (det == 0.0f) continue;
    float det_inv = 1.f / det;
    float conic[3] = {cov[2] * det_inv, -(cov[1] * det_inv), cov[0] * det_inv};
    float mid = 0.5f * (cov[0] + cov[2]);
    float sq = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
    float l1 = mid + sq, l2 = mid - sq;
    float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
    float px = ndc2pix(projx, W), py = ndc2pix(projy, H);
    int x0, y0, x1, y1;
    tile_rect(px, py, (int)my_radius, gx, gy, &x0, &y0, &x1, &y1);
    if ((x1 - x0) * (y1 - y0) == 0) continue;
    if (!colors_precomp) sh_to_rgb(D, p, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
    depths[i] = pv[2];
    radii[i] = (int)my_radius;
    means2D[2 * i] = px; means2D[2 * i + 1] = py;
    conic_opacity[4 * i] = conic[0]; conic_opacity[4 * i + 1] = conic[1]; conic_opacity[4 * i + 2] = conic[2];
    conic_opacity[4 * i + 3] = opac[i];
    tiles_touched[i] = (uint32_t)((y1 - y0) * (x1 - x0));
  }
}

void oracle_mark_visible(int P, const float* means, const float* view, uint8_t* present) {
  for (int i = 0; i < P; i++) {
    float pv[3];
    xform4x3(means + 3 * i, view, pv);
    present[i] = pv[2] <= 0.2f ? 0 : 1;
  }
}

/* ---- binning: rasterizer_impl.cu:73-141, 284-326 --------------------------
 * Returns R.  keys/vals must hold sum(tiles_touched) entries (call once with
 * keys == NULL to get R).  Stable LSD radix sort, 16-bit digits. */
int64_t oracle_bin(int P, const int* radii, const float* means2D, const float* depths,
                   const uint32_t* tiles_touched, int W, int H, uint64_t* keys, uint32_t* vals, uint32_t* ranges) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  int64_t R = 0;
  for (int i = 0; i < P; i++) R += tiles_touched[i];
  if (!keys) return R;
  int64_t off = 0;
  for (int i = 0; i < P; i++) {
    if (radii[i] > 0) {
      int x0, y0, x1, y1;
      tile_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
      uint32_t dbits;
      memcpy(&dbits, depths + i, 4);
      for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
          uint64_t key = (uint64_t)(y * gx + x);
          key <<= 32;
          key |= dbits;
          keys[off] = key;
          vals[off] = (uint32_t)i;
          off++;
        }
    }
  }
  /* stable sort by key */
  uint64_t* k2 = (uint64_t*)malloc((size_t)(R ? R : 1) * 8);
  uint32_t* v2 = (uint32_t*)malloc((size_t)(R ? R : 1) * 4);
  size_t* cnt = (size_t*)malloc(65537 * sizeof(size_t));
  uint64_t *ka = keys, *kb = k2;
  uint32_t *va = vals, *vb = v2;
  for (int pass = 0; pass < 4; pass++) {
    int shift = pass * 16;
    memset(cnt, 0, 65537 * sizeof(size_t));
    for (int64_t i = 0; i < R; i++) cnt[((ka[i] >> shift) & 0xffff) + 1]++;
    for (int d = 0; d < 65536; d++) cnt[d + 1] += cnt[d];
    for (int64_t i = 0; i < R; i++) {
      size_t pos = cnt[(ka[i] >> shift) & 0xffff]++;
      kb[pos] = ka[i];
      vb[pos] = va[i];
    }
    uint64_t* tk = ka; ka = kb; kb = tk;
    uint32_t* tv = va; va = vb; vb = tv;
  }
  /* 4 passes: result is back in keys/vals */
  free(k2); free(v2); free(cnt);
  memset(ranges, 0, (size_t)gx * gy * 8);
  for (int64_t i = 0; i < R; i++) {
    uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
      if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
    }
    if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
  }
  return R;
}

/* alpha chain in the operation order of the reference's compiled kernel */
static inline float splat_power(float mx, float my, float a, float b, float c, float px, float py, float* dx,
                                float* dy) {
  *dx = mx - px; *dy = my - py;
  float S = fmaf(*dx, *dx * a, (*dy * c) * *dy);
  return fmaf(S, -0.5f, -((*dx * b) * *dy));
}

/* forward.cu:264-396: per pixel front-to-back compositing */
void oracle_render_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                           const float* colors /*[P][3]*/, const float* feats /*[P][32]*/, const float* depths,
                           const float* conic_opacity, const float* bg, float* final_T, uint32_t* n_contrib,
                           float* out_color, float* out_objects, float* out_depth) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < TILE; ly++)
      for (int lx = 0; lx < TILE; lx++) {
        const int px = tx * TILE + lx, py = ty * TILE + ly;
        if (px >= W || py >= H) continue;
        const float pxf = (float)px, pyf = (float)py;
        float T = 1.0f, C[3] = {0, 0, 0}, O[NOBJ], Dp = 0;
        memset(O, 0, sizeof(O));
        uint32_t contributor = 0, last = 0;
        for (uint32_t k = r0; k < r1; k++) {
          contributor++;
          const uint32_t id = point_list[k];
          const float* co = conic_opacity + 4 * id;
          float dx, dy;
          float power = splat_power(means2D[2 * id], means2D[2 * id + 1], co[0], co[1], co[2], pxf, pyf, &dx, &dy);
          if (power > 0.0f) continue;
          float alpha = fminf(0.99f, co[3] * expf(power));
          if (alpha < 1.0f / 255.0f) continue;
          float test_T = T * (1 - alpha);
          if (test_T < 0.0001f) break;
          for (int ch = 0; ch < 3; ch++) C[ch] += colors[3 * id + ch] * alpha * T;
          for (int ch = 0; ch < NOBJ; ch++) O[ch] += feats[NOBJ * id + ch] * alpha * T;
          Dp += depths[id] * alpha * T;
          T = test_T;
          last = contributor;
        }
        const size_t pix = (size_t)W * py + px;
        final_T[pix] = T;
        n_contrib[pix] = last;
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix] = C[ch] + T * bg[ch];
        for (int ch = 0; ch < NOBJ; ch++) out_objects[ch * HW + pix] = O[ch];
        out_depth[pix] = Dp;
      }
  }
}

/* backward.cu:392-574.  Per-Gaussian sums in double (dL_* arrays are double). */
void oracle_render_backward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                            const float* means2D, const float* conic_opacity, const float* colors,
                            const float* feats, const float* final_T, const uint32_t* n_contrib,
                            const float* dL_dpix, const float* dL_dpix_obj, double* dL_dmean2D /*[P][3]*/,
                            double* dL_dconic /*[P][4]*/, double* dL_dopacity, double* dL_dcolors /*[P][3]*/,
                            double* dL_dobjects /*[P][32]*/) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t HW = (size_t)H * W;
  const float ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < TILE; ly++)
      for (int lx = 0; lx < TILE; lx++) {
        const int px = tx * TILE + lx, py = ty * TILE + ly;
        if (px >= W || py >= H) continue;
        const size_t pix = (size_t)W * py + px;
        const float pxf = (float)px, pyf = (float)py;
        const float T_final = final_T[pix];
        float T = T_final;
        const uint32_t last_contributor = n_contrib[pix];
        float accum_rec[3] = {0, 0, 0}, accum_rec_obj[NOBJ], last_color[3] = {0, 0, 0}, last_object[NOBJ];
        float dLp[3], dLo[NOBJ];
        memset(accum_rec_obj, 0, sizeof(accum_rec_obj));
        memset(last_object, 0, sizeof(last_object));
        for (int c = 0; c < 3; c++) dLp[c] = dL_dpix[c * HW + pix];
        for (int c = 0; c < NOBJ; c++) dLo[c] = dL_dpix_obj[c * HW + pix];
        float last_alpha = 0;
        uint32_t contributor = r1 - r0;
        for (uint32_t k = r1; k-- > r0;) {
          contributor--;
          if (contributor >= last_contributor) continue;
          const uint32_t id = point_list[k];
          const float* co = conic_opacity + 4 * id;
          float dx, dy;
          float power = splat_power(means2D[2 * id], means2D[2 * id + 1], co[0], co[1], co[2], pxf, pyf, &dx, &dy);
          if (power > 0.0f) continue;
          const float G = expf(power);
          const float alpha = fminf(0.99f, co[3] * G);
          if (alpha < 1.0f / 255.0f) continue;
          T = T / (1.f - alpha);
          const float dchannel_dcolor = alpha * T;
          float dL_dalpha = 0.0f;
          for (int ch = 0; ch < 3; ch++) {
            const float c = colors[3 * id + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = c;
            dL_dalpha += (c - accum_rec[ch]) * dLp[ch];
            const double g = dchannel_dcolor * dLp[ch];
#pragma omp atomic
            dL_dcolors[3 * id + ch] += g;
          }
          for (int ch = 0; ch < NOBJ; ch++) {
            const float o = feats[NOBJ * id + ch];
            accum_rec_obj[ch] = last_alpha * last_object[ch] + (1.f - last_alpha) * accum_rec_obj[ch];
            last_object[ch] = o;
            dL_dalpha += (o - accum_rec_obj[ch]) * dLo[ch];
            const double g = dchannel_dcolor * dLo[ch];
#pragma omp atomic
            dL_dobjects[NOBJ * id + ch] += g;
          }
          dL_dalpha *= T;
          last_alpha = alpha;
          float bg_dot = 0;
          for (int i = 0; i < 3; i++) bg_dot += bg[i] * dLp[i];
          dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
          const float dL_dG = co[3] * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * co[0] - gdy * co[1];
          const float dG_ddely = -gdy * co[2] - gdx * co[1];
          const double g0 = dL_dG * dG_ddelx * ddelx_dx, g1 = dL_dG * dG_ddely * ddely_dy;
          const double g2 = -0.5f * gdx * dx * dL_dG, g3 = -0.5f * gdx * dy * dL_dG, g4 = -0.5f * gdy * dy * dL_dG;
          const double g5 = G * dL_dalpha;
#pragma omp atomic
          dL_dmean2D[3 * id] += g0;
#pragma omp atomic
          dL_dmean2D[3 * id + 1] += g1;
#pragma omp atomic
          dL_dconic[4 * id] += g2;
#pragma omp atomic
          dL_dconic[4 * id + 1] += g3;
#pragma omp atomic
          dL_dconic[4 * id + 3] += g4;
#pragma omp atomic
          dL_dopacity[id] += g5;
        }
      }
  }
}

/* ---- preprocess backward: backward.cu:139-389 ------------------------------
 * Evaluated in double from the fp32 inputs (a stricter checker than fp32; the
 * golden vectors from the reference pin it to within fp32 rounding).
 * Inputs dL_dmean2D [P][3], dL_dconic [P][4] (x,y,_,w), dL_dcolor [P][3] are
 * the double sums produced by oracle_render_backward.
 * m3[c][r] column-major 3x3 like the reference's glm matrices. */
typedef struct { double m[3][3]; } M3;
static M3 m3_mul(const M3* A, const M3* B) {
  M3 R;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) R.m[c][r] = A->m[0][r] * B->m[c][0] + A->m[1][r] * B->m[c][1] + A->m[2][r] * B->m[c][2];
  return R;
}
static M3 m3_t(const M3* A) {
  M3 R;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) R.m[c][r] = A->m[r][c];
  return R;
}

void oracle_preprocess_backward(int P, int D, int M, const float* means, const int* radii, const float* shs,
                                const uint8_t* clamped /*[P][3]*/, const float* scales, const float* rots, float mod,
                                const float* cov3Ds, const float* view, const float* proj, float fx, float fy,
                                float tanx, float tany, const float* campos, const double* dL_dmean2D,
                                const double* dL_dconic, const double* dL_dcolor, double* dL_dmeans /*[P][3]*/,
                                double* dL_dcov3D /*[P][6]*/, double* dL_dsh /*[P][M][3]*/, double* dL_dscale,
                                double* dL_drot) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    if (!(radii[i] > 0)) continue;
    const float* mean = means + 3 * i;
    const float* c3 = cov3Ds + 6 * i;
    /* --- computeCov2DCUDA --- */
    double t[3];
    for (int r = 0; r < 3; r++)
      t[r] = (double)view[r] * mean[0] + (double)view[4 + r] * mean[1] + (double)view[8 + r] * mean[2] + view[12 + r];
    const double limx = 1.3 * (double)tanx, limy = 1.3 * (double)tany;
    /* clamp decisions are taken on the fp32 ratios, as the kernel does */
    float tf[3];
    xform4x3(mean, view, tf);
    const float txtz_f = tf[0] / tf[2], tytz_f = tf[1] / tf[2];
    const float limxf = 1.3f * tanx, limyf = 1.3f * tany;
    const double x_grad_mul = (txtz_f < -limxf || txtz_f > limxf) ? 0 : 1;
    const double y_grad_mul = (tytz_f < -limyf || tytz_f > limyf) ? 0 : 1;
    double txtz = t[0] / t[2], tytz = t[1] / t[2];
    if (x_grad_mul == 0) txtz = txtz_f < 0 ? -limx : limx;
    if (y_grad_mul == 0) tytz = tytz_f < 0 ? -limy : limy;
    t[0] = txtz * t[2];
    t[1] = tytz * t[2];
    M3 J = {{{fx / t[2], 0, -(fx * t[0]) / (t[2] * t[2])}, {0, fy / t[2], -(fy * t[1]) / (t[2] * t[2])}, {0, 0, 0}}};
    M3 Wm = {{{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}}};
    M3 Vrk = {{{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}}};
    M3 T = m3_mul(&Wm, &J);
    M3 Tt = m3_t(&T), Vt = m3_t(&Vrk);
    M3 X = m3_mul(&Tt, &Vt);
    M3 cov2D = m3_mul(&X, &T);
    const double a = cov2D.m[0][0] + (double)0.3f, b = cov2D.m[0][1], c = cov2D.m[1][1] + (double)0.3f;
    const double denom = a * c - b * b;
    const double denom2inv = 1.0 / (denom * denom + (double)0.0000001f);
    const double gcx = dL_dconic[4 * i], gcy = dL_dconic[4 * i + 1], gcz = dL_dconic[4 * i + 3];
    double dL_da = 0, dL_db = 0, dL_dc = 0;
    double* dc = dL_dcov3D + 6 * i;
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * gcx + 2 * b * c * gcy + (denom - a * c) * gcz);
      dL_dc = denom2inv * (-a * a * gcz + 2 * a * b * gcy + (denom - a * c) * gcx);
      dL_db = denom2inv * 2 * (b * c * gcx - (denom + 2 * b * b) * gcy + a * b * gcz);
      dc[0] = T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc;
      dc[3] = T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc;
      dc[5] = T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc;
      dc[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
              2 * T.m[1][0] * T.m[1][1] * dL_dc;
      dc[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
              2 * T.m[1][0] * T.m[1][2] * dL_dc;
      dc[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
              2 * T.m[1][1] * T.m[1][2] * dL_dc;
    } else {
      for (int k = 0; k < 6; k++) dc[k] = 0;
    }
    double dT[2][3];
    for (int k = 0; k < 3; k++) {
      const double tv0 = T.m[0][0] * Vrk.m[k][0] + T.m[0][1] * Vrk.m[k][1] + T.m[0][2] * Vrk.m[k][2];
      const double tv1 = T.m[1][0] * Vrk.m[k][0] + T.m[1][1] * Vrk.m[k][1] + T.m[1][2] * Vrk.m[k][2];
      dT[0][k] = 2 * tv0 * dL_da + tv1 * dL_db;
      dT[1][k] = 2 * tv1 * dL_dc + tv0 * dL_db;
    }
    const double dJ00 = Wm.m[0][0] * dT[0][0] + Wm.m[0][1] * dT[0][1] + Wm.m[0][2] * dT[0][2];
    const double dJ02 = Wm.m[2][0] * dT[0][0] + Wm.m[2][1] * dT[0][1] + Wm.m[2][2] * dT[0][2];
    const double dJ11 = Wm.m[1][0] * dT[1][0] + Wm.m[1][1] * dT[1][1] + Wm.m[1][2] * dT[1][2];
    const double dJ12 = Wm.m[2][0] * dT[1][0] + Wm.m[2][1] * dT[1][1] + Wm.m[2][2] * dT[1][2];
    const double tz = 1.0 / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    const double dtx = x_grad_mul * -fx * tz2 * dJ02;
    const double dty = y_grad_mul * -fy * tz2 * dJ12;
    const double dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t[0]) * tz3 * dJ02 + (2 * fy * t[1]) * tz3 * dJ12;
    double dmean[3] = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
                       view[8] * dtx + view[9] * dty + view[10] * dtz};
    /* --- projection (backward.cu:361-377) --- */
    {
      const double mx = mean[0], my = mean[1], mz = mean[2];
      const double hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
      const double m_w = 1.0 / (hw + (double)0.0000001f);
      const double mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
      const double mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
      const double gx2 = dL_dmean2D[3 * i], gy2 = dL_dmean2D[3 * i + 1];
      dmean[0] += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
      dmean[1] += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
      dmean[2] += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;
    }
    /* --- SH (backward.cu:18-134) --- */
    if (shs) {
      const double d0[3] = {(double)mean[0] - campos[0], (double)mean[1] - campos[1], (double)mean[2] - campos[2]};
      const double len = sqrt(d0[0] * d0[0] + d0[1] * d0[1] + d0[2] * d0[2]);
      const double x = d0[0] / len, y = d0[1] / len, z = d0[2] / len;
      const float* sh = shs + (size_t)i * M * 3;
      double* dsh = dL_dsh + (size_t)i * M * 3;
      double g[3];
      for (int ch = 0; ch < 3; ch++) g[ch] = clamped[3 * i + ch] ? 0.0 : dL_dcolor[3 * i + ch];
      double w[16] = {0}, wx[16] = {0}, wy[16] = {0}, wz[16] = {0}; /* basis value and its d/dx, d/dy, d/dz */
      const double xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      w[0] = SH_C0;
      if (D > 0) {
        w[1] = -SH_C1 * y; wy[1] = -SH_C1;
        w[2] = SH_C1 * z;  wz[2] = SH_C1;
        w[3] = -SH_C1 * x; wx[3] = -SH_C1;
      }
      if (D > 1) {
        w[4] = SH_C2[0] * xy; wx[4] = SH_C2[0] * y; wy[4] = SH_C2[0] * x;
        w[5] = SH_C2[1] * yz; wy[5] = SH_C2[1] * z; wz[5] = SH_C2[1] * y;
        w[6] = SH_C2[2] * (2 * zz - xx - yy); wx[6] = SH_C2[2] * -2 * x; wy[6] = SH_C2[2] * -2 * y; wz[6] = SH_C2[2] * 4 * z;
        w[7] = SH_C2[3] * xz; wx[7] = SH_C2[3] * z; wz[7] = SH_C2[3] * x;
        w[8] = SH_C2[4] * (xx - yy); wx[8] = SH_C2[4] * 2 * x; wy[8] = SH_C2[4] * -2 * y;
      }
      if (D > 2) {
        w[9] = SH_C3[0] * y * (3 * xx - yy); wx[9] = SH_C3[0] * 6 * xy; wy[9] = SH_C3[0] * 3 * (xx - yy);
        w[10] = SH_C3[1] * xy * z; wx[10] = SH_C3[1] * yz; wy[10] = SH_C3[1] * xz; wz[10] = SH_C3[1] * xy;
        w[11] = SH_C3[2] * y * (4 * zz - xx - yy); wx[11] = SH_C3[2] * -2 * xy; wy[11] = SH_C3[2] * (-3 * yy + 4 * zz - xx);
        wz[11] = SH_C3[2] * 8 * yz;
        w[12] = SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy); wx[12] = SH_C3[3] * -6 * xz; wy[12] = SH_C3[3] * -6 * yz;
        wz[12] = SH_C3[3] * 3 * (2 * zz - xx - yy);
        w[13] = SH_C3[4] * x * (4 * zz - xx - yy); wx[13] = SH_C3[4] * (-3 * xx + 4 * zz - yy); wy[13] = SH_C3[4] * -2 * xy;
        wz[13] = SH_C3[4] * 8 * xz;
        w[14] = SH_C3[5] * z * (xx - yy); wx[14] = SH_C3[5] * 2 * xz; wy[14] = SH_C3[5] * -2 * yz; wz[14] = SH_C3[5] * (xx - yy);
        w[15] = SH_C3[6] * x * (xx - 3 * yy); wx[15] = SH_C3[6] * 3 * (xx - yy); wy[15] = SH_C3[6] * -6 * xy;
      }
      const int nc = (D + 1) * (D + 1);
      double ddir[3] = {0, 0, 0};
      for (int k = 0; k < nc; k++)
        for (int ch = 0; ch < 3; ch++) {
          dsh[3 * k + ch] = w[k] * g[ch];
          ddir[0] += wx[k] * sh[3 * k + ch] * g[ch];
          ddir[1] += wy[k] * sh[3 * k + ch] * g[ch];
          ddir[2] += wz[k] * sh[3 * k + ch] * g[ch];
        }
      /* through dir = d0/|d0| */
      const double s2 = len * len, inv = 1.0 / (s2 * len);
      dmean[0] += ((s2 - d0[0] * d0[0]) * ddir[0] - d0[1] * d0[0] * ddir[1] - d0[2] * d0[0] * ddir[2]) * inv;
      dmean[1] += (-d0[0] * d0[1] * ddir[0] + (s2 - d0[1] * d0[1]) * ddir[1] - d0[2] * d0[1] * ddir[2]) * inv;
      dmean[2] += (-d0[0] * d0[2] * ddir[0] - d0[1] * d0[2] * ddir[1] + (s2 - d0[2] * d0[2]) * ddir[2]) * inv;
    }
    for (int k = 0; k < 3; k++) dL_dmeans[3 * i + k] = dmean[k];
    /* --- cov3D -> scale, rotation (backward.cu:273-334) --- */
    if (scales) {
      const double r = rots[4 * i], x = rots[4 * i + 1], y = rots[4 * i + 2], z = rots[4 * i + 3];
      M3 R = {{{1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)},
               {2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)},
               {2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)}}};
      const double s[3] = {(double)mod * scales[3 * i], (double)mod * scales[3 * i + 1], (double)mod * scales[3 * i + 2]};
      M3 S = {{{s[0], 0, 0}, {0, s[1], 0}, {0, 0, s[2]}}};
      M3 Mm = m3_mul(&S, &R);
      M3 dSig = {{{dc[0], 0.5 * dc[1], 0.5 * dc[2]}, {0.5 * dc[1], dc[3], 0.5 * dc[4]}, {0.5 * dc[2], 0.5 * dc[4], dc[5]}}};
      M3 M2 = Mm;
      for (int cc = 0; cc < 3; cc++) for (int rr = 0; rr < 3; rr++) M2.m[cc][rr] *= 2.0;
      M3 dM = m3_mul(&M2, &dSig);
      M3 Rt = m3_t(&R), dMt = m3_t(&dM);
      for (int k = 0; k < 3; k++)
        dL_dscale[3 * i + k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
      for (int k = 0; k < 3; k++) for (int rr = 0; rr < 3; rr++) dMt.m[k][rr] *= s[k];
      double* dq = dL_drot + 4 * i;
      dq[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
      dq[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) -
              4 * x * (dMt.m[2][2] + dMt.m[1][1]);
      dq[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) -
              4 * y * (dMt.m[2][2] + dMt.m[0][0]);
      dq[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) -
              4 * z * (dMt.m[1][1] + dMt.m[0][0]);
    }
  }
}
