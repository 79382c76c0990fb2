// Fused Adam over the flat parameter / gradient arenas and the densification statistics
// (first part).
// ref: torch.optim.Adam(groups, lr=0.0, eps=1e-15) as TRASE configures it — one optimizer with a group per
// Gaussian attribute (scene/gaussian_model.py:259-289) and one for the DeformNetwork (scene/deform_model.py:38-43),
// each stepping ~10-20 tensors with a handful of elementwise kernels per tensor — and
// GaussianModel.add_densification_stats (scene/gaussian_model.py:637-639).
// Here: parameters, gradients and both moments live in flat fp32 buffers (the gradient buffer is the
// all-reduced arena of trase_b200/distributed.py), segments carry their own step size, one kernel per step.
#include "common.cuh"

namespace tb {

constexpr int ADAM_MAX_SEGS = 64;
struct AdamTable {
  long long end[ADAM_MAX_SEGS];     // exclusive end offset of each segment (ascending)
  float step_size[ADAM_MAX_SEGS];   // lr / (1 - beta1^t)
  float bc2_sqrt[ADAM_MAX_SEGS];    // sqrt(1 - beta2^t)
  int nseg;
};

// torch's single-tensor Adam, op for op (torch/optim/adam.py _single_tensor_adam):
//   m.lerp_(g, 1 - b1);  v.mul_(b2).addcmul_(g, g, value = 1 - b2);
//   denom = v.sqrt() / sqrt(1 - b2^t) + eps;  p.addcdiv_(m, denom, value = -lr / (1 - b1^t))
__global__ void __launch_bounds__(256) adam_step_kernel(long long n, float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        const AdamTable t, float beta1, float beta2, float eps) {
  pdl_wait();
  const float w1 = 1.f - beta1, w2 = 1.f - beta2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    int lo = 0, hi = t.nseg - 1;            // first segment whose end is > i
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (t.end[mid] > i) hi = mid; else lo = mid + 1;
    }
    const float step = t.step_size[lo];
    if (step == 0.f && t.bc2_sqrt[lo] == 0.f) continue;      // frozen segment (not stepped this call)
    const float gi = g[i];
    const float mi = fmaf(w1, gi - m[i], m[i]);
    const float vi = fmaf(w2 * gi, gi, v[i] * beta2);
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / t.bc2_sqrt[lo] + eps;
    p[i] = fmaf(-step, mi / denom, p[i]);
  }
}

__global__ void __launch_bounds__(256) densification_stats_kernel(int P, const float* __restrict__ grad3,
                                                                  const bool* __restrict__ filter,
                                                                  float* __restrict__ accum, float* __restrict__ denom) {
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P && filter[i]) {
    const float x = grad3[3 * i], y = grad3[3 * i + 1];
    accum[i] += sqrtf(x * x + y * y);
    denom[i] += 1.f;
  }
}

// Densification / pruning as ONE gather pass over the flat arenas (ref: the per-tensor torch.cat / boolean-index
// re-allocation of every parameter and both Adam moments in cat_tensors_to_optimizer / _prune_optimizer,
// scene/gaussian_model.py:472-534).  Every per-Gaussian segment s is a [P][width_s] row-major block; output row r of
// every segment takes input row src[r]; rows flagged `fresh` (clones, split children) start with zero moments.
constexpr int GATHER_MAX_SEGS = 16;
struct GatherTable {
  long long old_off[GATHER_MAX_SEGS], new_off[GATHER_MAX_SEGS];
  int width[GATHER_MAX_SEGS];
  int nseg;
};
__global__ void __launch_bounds__(256) arena_gather_rows_kernel(const GatherTable t, long long P_new,
                                                                const long long* __restrict__ src,
                                                                const unsigned char* __restrict__ fresh,
                                                                const float* __restrict__ old_p,
                                                                const float* __restrict__ old_m,
                                                                const float* __restrict__ old_v, float* __restrict__ new_p,
                                                                float* __restrict__ new_m, float* __restrict__ new_v) {
  pdl_wait();
  const int sgi = blockIdx.y;
  const int w = t.width[sgi];
  const long long total = P_new * w;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long r = e / w;
    const int c = (int)(e - r * w);
    const long long from = t.old_off[sgi] + src[r] * w + c, to = t.new_off[sgi] + e;
    new_p[to] = old_p[from];
    if (old_m) {
      const bool z = fresh[r] != 0;
      new_m[to] = z ? 0.f : old_m[from];
      new_v[to] = z ? 0.f : old_v[from];
    }
  }
}

}  // namespace tb

using namespace tb;

extern "C" TRASE_API int trase_b200_arena_gather_rows(int nseg, const int64_t* old_off, const int64_t* new_off,
                                                      const int* width, int64_t P_new, const int64_t* src_rows,
                                                      const unsigned char* fresh, const float* old_p,
                                                      const float* old_m, const float* old_v, float* new_p, float* new_m,
                                                      float* new_v, void* stream) {
  clear_error();
  if (P_new <= 0 || nseg <= 0) return 0;
  if (nseg > GATHER_MAX_SEGS || !old_off || !new_off || !width || !src_rows || !old_p || !new_p ||
      ((old_m != nullptr) != (old_v != nullptr)) || (old_m && (!new_m || !new_v || !fresh))) {
    set_error("trase_b200_arena_gather_rows: bad arguments (nseg=%d, at most %d segments)", nseg, GATHER_MAX_SEGS);
    return 1;
  }
  GatherTable t;
  t.nseg = nseg;
  int wmax = 1;
  for (int i = 0; i < nseg; i++) {
    if (width[i] <= 0) { set_error("trase_b200_arena_gather_rows: segment width must be positive"); return 1; }
    t.old_off[i] = old_off[i]; t.new_off[i] = new_off[i]; t.width[i] = width[i];
    if (width[i] > wmax) wmax = width[i];
  }
  long long blocks = (P_new * wmax + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  launch_k(arena_gather_rows_kernel, dim3((unsigned)blocks, (unsigned)nseg), 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      t, P_new, reinterpret_cast<const long long*>(src_rows), fresh, old_p, old_m, old_v, new_p, new_m, new_v);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_arena_gather_rows: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API int trase_b200_adam_step(int64_t n, float* params, const float* grads, float* exp_avg,
                                              float* exp_avg_sq, int nseg, const int64_t* seg_end,
                                              const float* step_size, const float* bc2_sqrt, float beta1, float beta2,
                                              float eps, void* stream) {
  if (n <= 0) return 0;
  if (!params || !grads || !exp_avg || !exp_avg_sq || !seg_end || !step_size || !bc2_sqrt || nseg <= 0 ||
      nseg > ADAM_MAX_SEGS) {
    set_error("trase_b200_adam_step: bad arguments (n=%lld nseg=%d, at most %d segments)", (long long)n, nseg,
              ADAM_MAX_SEGS);
    return 1;
  }
  AdamTable t;
  t.nseg = nseg;
  long long prev = 0;
  for (int i = 0; i < nseg; i++) {
    if (seg_end[i] < prev) { set_error("trase_b200_adam_step: segment ends must ascend"); return 1; }
    prev = t.end[i] = seg_end[i];
    t.step_size[i] = step_size[i];
    t.bc2_sqrt[i] = bc2_sqrt[i];
  }
  if (prev != n) { set_error("trase_b200_adam_step: segments must cover the buffer"); return 1; }
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  launch_k(adam_step_kernel, (int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream), n, params, grads, exp_avg, exp_avg_sq,
                                                                                  t, beta1, beta2, eps);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_adam_step: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

extern "C" TRASE_API int trase_b200_densification_stats(int P, const float* viewspace_grad, const unsigned char* filter,
                                                        float* grad_accum, float* denom, void* stream) {
  if (P <= 0) return 0;
  if (!viewspace_grad || !filter || !grad_accum || !denom) { set_error("trase_b200_densification_stats: null pointer"); return 1; }
  launch_k(densification_stats_kernel, (P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      P, viewspace_grad, reinterpret_cast<const bool*>(filter), grad_accum, denom);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_densification_stats: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
