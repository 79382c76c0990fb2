// C-ABI entry points (see include/trase_b200.h) and host-side orchestration of
// one forward / backward pass.  No device memory is allocated here and nothing
// synchronises with the host unless `debug` is set.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <utility>
#include <vector>

namespace tb {

static thread_local char g_err[512] = {0};
std::atomic<unsigned long long> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void clear_error() { g_err[0] = 0; }

static int fail_cuda(const char* where, cudaError_t e) {
  set_error("trase_b200: CUDA error in %s: %s", where, cudaGetErrorString(e));
  return 2;
}

// ---- optional per-stage timing (bench.py's roofline leg) --------------------
// CUDA events recorded on the launching stream around each stage; off by default.
enum { ST_PREPROCESS_FWD = 0, ST_SCAN_DUP, ST_SORT, ST_RANGES, ST_RENDER_FWD, ST_RENDER_BWD, ST_PREPROCESS_BWD,
       ST_OTHER, ST_COUNT };
struct ProfRec { int stage; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_live;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_prof_free;
static std::mutex g_prof_mu;

static void prof_begin(int stage, cudaStream_t s, ProfRec& r) {
  r.stage = -1;
  if (!g_prof_on || stage < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof_free.empty()) { r.a = g_prof_free.back().first; r.b = g_prof_free.back().second; g_prof_free.pop_back(); }
  else { if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return; }
  r.stage = stage;
  cudaEventRecord(r.a, s);
}
static void prof_end(cudaStream_t s, ProfRec& r) {
  if (r.stage < 0) return;
  cudaEventRecord(r.b, s);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_live.push_back(r);
}

// debug mode: synchronise and check after each launch (auxiliary.h:161-171)
#define TB_STEP_ST(stage, name, expr)                              \
  do {                                                             \
    ProfRec pr__;                                                  \
    prof_begin(stage, stream, pr__);                               \
    cudaError_t e__ = (expr);                                      \
    prof_end(stream, pr__);                                        \
    if (e__ != cudaSuccess) return fail_cuda(name, e__);           \
    if (debug) {                                                   \
      e__ = cudaStreamSynchronize(stream);                         \
      if (e__ != cudaSuccess) return fail_cuda(name, e__);         \
    }                                                              \
  } while (0)
#define TB_STEP(name, expr) TB_STEP_ST(-1, name, expr)

static FwdPointers carve(void* geom, int P, void* img, int W, int H, void* binning, int64_t capacity, void* scratch) {
  FwdPointers p;
  memset(&p, 0, sizeof(p));
  char* g = reinterpret_cast<char*>(geom);
  if (g) {
    GeomLayout L = geom_layout(P);
    p.depths = reinterpret_cast<float*>(g + L.off[TRASE_GEOM_DEPTHS]);
    p.means2D = reinterpret_cast<float2*>(g + L.off[TRASE_GEOM_MEANS2D]);
    p.conic_opacity = reinterpret_cast<float4*>(g + L.off[TRASE_GEOM_CONIC_OPACITY]);
    p.rgbd = reinterpret_cast<float4*>(g + L.off[TRASE_GEOM_RGBD]);
    p.cov3D = reinterpret_cast<float*>(g + L.off[TRASE_GEOM_COV3D]);
    p.clamped = reinterpret_cast<uint8_t*>(g + L.off[TRASE_GEOM_CLAMPED]);
    p.tiles_touched = reinterpret_cast<uint32_t*>(g + L.off[TRASE_GEOM_TILES_TOUCHED]);
    p.point_offsets = reinterpret_cast<uint32_t*>(g + L.off[TRASE_GEOM_POINT_OFFSETS]);
    p.block_sums = reinterpret_cast<uint32_t*>(g + L.off[TRASE_GEOM_BLOCK_SUMS]);
    p.block_offsets = p.block_sums + ((size_t)(P + PRE_BLOCK - 1) / PRE_BLOCK + 1);
    p.counters = reinterpret_cast<uint32_t*>(g + L.off[TRASE_GEOM_COUNTERS]);
  }
  char* im = reinterpret_cast<char*>(img);
  if (im) {
    ImgLayout L = img_layout(W, H);
    p.final_T = reinterpret_cast<float*>(im + L.off[TRASE_IMG_FINAL_T]);
    p.n_contrib = reinterpret_cast<uint32_t*>(im + L.off[TRASE_IMG_N_CONTRIB]);
    p.ranges = reinterpret_cast<uint2*>(im + L.off[TRASE_IMG_RANGES]);
  }
  char* b = reinterpret_cast<char*>(binning);
  if (b) {
    BinLayout L = bin_layout(capacity);
    p.point_list = reinterpret_cast<uint32_t*>(b + L.off[TRASE_BIN_POINT_LIST]);
    p.point_list_keys = reinterpret_cast<uint64_t*>(b + L.off[TRASE_BIN_POINT_LIST_KEYS]);
  }
  char* sc = reinterpret_cast<char*>(scratch);
  if (sc) {
    ScratchLayout L = scratch_layout(capacity);
    p.keys_unsorted = reinterpret_cast<uint64_t*>(sc + L.keys_unsorted);
    p.vals_unsorted = reinterpret_cast<uint32_t*>(sc + L.vals_unsorted);
    p.sort_scratch = sc + L.sort;
    p.sort_scratch_bytes = sort_scratch_bytes(capacity);
  }
  return p;
}

static int check_fwd_args(const trase_b200_fwd_args* a) {
  if (!a) { set_error("trase_b200: null argument struct"); return 1; }
  if (a->P < 0 || a->W <= 0 || a->H <= 0) { set_error("trase_b200: bad sizes P=%d W=%d H=%d", a->P, a->W, a->H); return 1; }
  if (a->P == 0) return 0;
  if (!a->means3D || !a->opacities || !a->background || !a->viewmatrix || !a->projmatrix || !a->campos) {
    set_error("trase_b200: means3D, opacities, background, viewmatrix, projmatrix and campos are required");
    return 1;
  }
  if (!a->sh_objs) { set_error("trase_b200: sh_objs (per-Gaussian 32-channel features) is required"); return 1; }
  if (((uintptr_t)a->sh_objs & 15) || (a->rotations && ((uintptr_t)a->rotations & 15))) {
    set_error("trase_b200: sh_objs and rotations must be 16-byte aligned");
    return 1;
  }
  if (!a->shs && !a->colors_precomp) {
    // the reference throws for non-RGB without precomputed colours (rasterizer_impl.cu:248-251);
    // with 3 channels it would dereference a null SH pointer — reject instead.
    set_error("trase_b200: provide either SHs or precomputed colors");
    return 1;
  }
  if (a->shs && !a->colors_precomp) {
    const int need = (a->sh_degree + 1) * (a->sh_degree + 1);
    if (a->sh_degree < 0 || a->sh_degree > 3 || a->M < need) {
      set_error("trase_b200: sh_degree=%d needs %d coefficients, M=%d", a->sh_degree, need, a->M);
      return 1;
    }
  }
  if (!a->cov3D_precomp && (!a->scales || !a->rotations)) {
    set_error("trase_b200: provide either scales+rotations or a precomputed 3D covariance");
    return 1;
  }
  if (!a->out_color || !a->out_objects || !a->out_depth || !a->radii) {
    set_error("trase_b200: output buffers are required");
    return 1;
  }
  if (!a->geom || a->geom_bytes < geom_layout(a->P).total || !a->img || a->img_bytes < img_layout(a->W, a->H).total) {
    set_error("trase_b200: geom/img workspace missing or too small");
    return 1;
  }
  if (a->capacity < 0 || (a->capacity > 0 && (!a->binning || a->binning_bytes < bin_layout(a->capacity).total ||
                                               !a->scratch || a->scratch_bytes < scratch_layout(a->capacity).total))) {
    set_error("trase_b200: binning/scratch workspace missing or too small for capacity %lld", (long long)a->capacity);
    return 1;
  }
  if (a->capacity > 0xFFFFFFF0ll) { set_error("trase_b200: capacity exceeds 32-bit instance indices"); return 1; }
  if (a->P > 0 && a->capacity < a->P) {
    set_error("trase_b200: capacity (%lld) must be at least P (%d): the depth sort of the Gaussians borrows the "
              "instance buffers", (long long)a->capacity, a->P);
    return 1;
  }
  return 0;
}

static int forward_impl(const trase_b200_fwd_args* a, bool with_preprocess) {
  g_err[0] = 0;
  if (int rc = check_fwd_args(a)) return rc;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(a->stream);
  const int debug = a->debug;
  const size_t HW = (size_t)a->W * a->H;
  if (a->P == 0) {
    // reference: zero images, nothing launched (rasterize_points.cu:75-90)
    TB_STEP("memset", cudaMemsetAsync(a->out_color, 0, 3 * HW * 4, stream));
    TB_STEP("memset", cudaMemsetAsync(a->out_objects, 0, NUM_OBJECTS * HW * 4, stream));
    TB_STEP("memset", cudaMemsetAsync(a->out_depth, 0, HW * 4, stream));
    if (a->num_rendered_host) *a->num_rendered_host = 0;
    if (a->readback_event) TB_STEP("record readback event", cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->readback_event), stream));
    return 0;
  }
  FwdPointers p = carve(a->geom, a->P, a->img, a->W, a->H, a->binning, a->capacity, a->scratch);
  const int grid_x = (a->W + TILE - 1) / TILE, grid_y = (a->H + TILE - 1) / TILE;
  const int end_bit = 32 + (int)higher_msb((uint32_t)(grid_x * grid_y));
  // Instances are sorted on their tile bits only (the Gaussians were put in depth order before emission, see
  // binning.cu).  The sort ping-pongs between the binning buffers and the scratch buffers; start on the side
  // that makes the last pass land in the binning buffers.
  const bool even = (sort_num_passes(32, end_bit) & 1) == 0;
  uint64_t* keys_a = even ? p.point_list_keys : p.keys_unsorted;
  uint32_t* vals_a = even ? p.point_list : p.vals_unsorted;
  uint64_t* keys_b = even ? p.keys_unsorted : p.point_list_keys;
  uint32_t* vals_b = even ? p.vals_unsorted : p.point_list;
  FwdPointers pd = p;
  pd.keys_unsorted = keys_a;
  pd.vals_unsorted = vals_a;

  if (with_preprocess) TB_STEP_ST(ST_PREPROCESS_FWD, "preprocess_forward", launch_preprocess_forward(*a, p, stream));
  // depth order of the Gaussians (P-sized sort on the other ping-pong side), block scan, key emission
  // R reaches the host without a copy-engine operation in the stream when the caller's word is device-visible
  // (pinned memory under unified addressing): the scan kernel stores it there and the event is recorded behind that
  // kernel.  Otherwise a 4-byte copy follows the emission as before.
  int* host_total_dev = nullptr;
  if (a->num_rendered_host) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, a->num_rendered_host) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
        attr.devicePointer)
      host_total_dev = static_cast<int*>(attr.devicePointer);
    else
      cudaGetLastError();      // not a registered pointer: clear the sticky error of the query
  }
  TB_STEP_ST(ST_SCAN_DUP, "scan_and_duplicate",
             launch_scan_and_duplicate(*a, pd, keys_b, vals_b, host_total_dev,
                                       reinterpret_cast<cudaEvent_t>(a->readback_event), stream));
  if (!host_total_dev) {
    if (a->num_rendered_host)
      TB_STEP("copy num_rendered",
              cudaMemcpyAsync(a->num_rendered_host, p.counters, sizeof(int), cudaMemcpyDeviceToHost, stream));
    if (a->readback_event) TB_STEP("record readback event", cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->readback_event), stream));
  }
  if (a->capacity > 0)
    TB_STEP_ST(ST_SORT, "sort_pairs", launch_sort_pingpong(keys_a, vals_a, keys_b, vals_b, p.counters, a->capacity, 32,
                                               end_bit, p.sort_scratch, p.sort_scratch_bytes, stream));
  TB_STEP_ST(ST_RANGES, "identify_ranges", launch_identify_ranges(*a, p, stream));
  TB_STEP_ST(ST_RENDER_FWD, "render_forward", launch_render_forward(*a, p, stream));
  // preprocess, block-sum scan, key duplication, 3 kernels per radix pass, ranges, compositing
  count_launches((with_preprocess ? 1 : 0) + 2);   // preprocess, ranges, compositing (binning counts its own)
  return 0;
}

}  // namespace tb

using namespace tb;

extern "C" {

int trase_b200_version(void) { return 100; }
const char* trase_b200_last_error(void) { return g_err; }

size_t trase_b200_geom_bytes(int P) { return geom_layout(P).total; }
size_t trase_b200_img_bytes(int W, int H) { return img_layout(W, H).total; }
size_t trase_b200_binning_bytes(int64_t capacity) { return bin_layout(capacity).total; }
size_t trase_b200_scratch_bytes(int P, int64_t capacity) {
  // the P-sized depth sort borrows this workspace: size it for whichever is larger
  return scratch_layout(capacity > (int64_t)P ? capacity : (int64_t)P).total;
}

void trase_b200_geom_offsets(int P, size_t* out) {
  GeomLayout L = geom_layout(P);
  for (int i = 0; i < TRASE_GEOM_NUM_ARRAYS; i++) out[i] = L.off[i];
}
void trase_b200_img_offsets(int W, int H, size_t* out) {
  ImgLayout L = img_layout(W, H);
  for (int i = 0; i < TRASE_IMG_NUM_ARRAYS; i++) out[i] = L.off[i];
}
void trase_b200_binning_offsets(int64_t capacity, size_t* out) {
  BinLayout L = bin_layout(capacity);
  for (int i = 0; i < TRASE_BIN_NUM_ARRAYS; i++) out[i] = L.off[i];
}

int trase_b200_rasterize_forward(const trase_b200_fwd_args* a) { return forward_impl(a, true); }
int trase_b200_rasterize_forward_rebin(const trase_b200_fwd_args* a) { return forward_impl(a, false); }

static int backward_impl(const trase_b200_bwd_args* a, int stages) {
  g_err[0] = 0;
  if (!a) { set_error("trase_b200: null argument struct"); return 1; }
  if (a->P < 0 || a->W <= 0 || a->H <= 0) { set_error("trase_b200: bad sizes"); return 1; }
  if (a->P == 0) return 0;
  if (!a->means3D || !a->background || !a->viewmatrix || !a->projmatrix || !a->campos || !a->radii || !a->geom ||
      !a->img || !a->sh_objs) {
    set_error("trase_b200: backward is missing a required input");
    return 1;
  }
  // Which gradients are wanted: either the whole geometry/colour group (NULL for none of them), and/or dL_dobjects.
  const bool any_geom = a->dL_dmeans2D || a->dL_dconic || a->dL_dopacity || a->dL_dcolors || a->dL_dmeans3D ||
                        a->dL_dcov3D || a->dL_dsh || a->dL_dscales || a->dL_drotations;
  const bool want_geom = a->dL_dmeans2D && a->dL_dconic && a->dL_dopacity && a->dL_dcolors && a->dL_dmeans3D && a->dL_dcov3D;
  if (any_geom && !want_geom) {
    set_error("trase_b200: backward needs dL_dmeans2D, dL_dconic, dL_dopacity, dL_dcolors, dL_dmeans3D and dL_dcov3D "
              "together (pass NULL for all gradient buffers except dL_dobjects to get the feature gradient only)");
    return 1;
  }
  if (!want_geom && !a->dL_dobjects) { set_error("trase_b200: backward was asked for no gradient at all"); return 1; }
  if (want_geom && !a->dL_dout_color) {
    set_error("trase_b200: dL_dout_color is required when geometry/colour gradients are wanted");
    return 1;
  }
  if (!want_geom && !a->dL_dout_objects) {
    set_error("trase_b200: dL_dout_objects is required when only the feature gradient is wanted");
    return 1;
  }
  if (want_geom) {
    if (a->shs && !a->colors_precomp && !a->dL_dsh) { set_error("trase_b200: dL_dsh required when SHs are used"); return 1; }
    if (a->scales && (!a->rotations || !a->dL_dscales || !a->dL_drotations)) {
      set_error("trase_b200: rotations/dL_dscales/dL_drotations required with scales");
      return 1;
    }
  }
  if (a->capacity > 0 && !a->binning) { set_error("trase_b200: binning workspace missing"); return 1; }
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(a->stream);
  const int debug = a->debug;
  FwdPointers p = carve(const_cast<void*>(a->geom), a->P, const_cast<void*>(a->img), a->W, a->H,
                        const_cast<void*>(a->binning), a->capacity, nullptr);
  if (stages & 1) TB_STEP_ST(ST_RENDER_BWD, "render_backward", launch_render_backward(*a, p, stream));
  if (!want_geom) stages &= 1;                    // feature gradient only: nothing flows into the per-Gaussian geometry
  if (stages & 2) TB_STEP_ST(ST_PREPROCESS_BWD, "preprocess_backward", launch_preprocess_backward(*a, p, stream));
  count_launches(((stages & 1) ? 1 : 0) + ((stages & 2) ? 1 : 0));
  return 0;
}

int trase_b200_rasterize_backward(const trase_b200_bwd_args* a) { return backward_impl(a, 3); }
int trase_b200_rasterize_backward_stages(const trase_b200_bwd_args* a, int stages) { return backward_impl(a, stages); }

unsigned long long trase_b200_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }

void trase_b200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
}
int trase_b200_profile_read(double* total_ms /*[8]*/, int* count /*[8]*/) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int i = 0; i < ST_COUNT; i++) { total_ms[i] = 0; count[i] = 0; }
  for (auto& r : g_prof_live) {
    float ms = 0.f;
    if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      total_ms[r.stage] += ms;
      count[r.stage]++;
    }
    g_prof_free.push_back({r.a, r.b});
  }
  g_prof_live.clear();
  return ST_COUNT;
}

int trase_b200_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                            uint8_t* present, void* stream) {
  g_err[0] = 0;
  (void)projmatrix;  // the reference's test only uses the view-space z (auxiliary.h:149)
  if (P < 0) { set_error("trase_b200: bad P"); return 1; }
  if (P == 0) return 0;
  if (!means3D || !viewmatrix || !present) { set_error("trase_b200: null pointer"); return 1; }
  cudaError_t e = launch_mark_visible(P, means3D, viewmatrix, present, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("mark_visible", e);
  return 0;
}

size_t trase_b200_sort_scratch_bytes(int64_t capacity) { return sort_scratch_bytes(capacity); }

int trase_b200_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                          const uint32_t* count_dev, int64_t capacity, int end_bit, void* sort_scratch,
                          size_t sort_scratch_bytes_, void* stream) {
  g_err[0] = 0;
  if (capacity < 0 || end_bit < 1 || end_bit > 64) { set_error("trase_b200: bad sort arguments"); return 1; }
  if (capacity == 0) return 0;
  if (!keys_in || !vals_in || !keys_out || !vals_out || !count_dev || !sort_scratch ||
      sort_scratch_bytes_ < sort_scratch_bytes(capacity)) {
    set_error("trase_b200: sort buffers missing or scratch too small");
    return 1;
  }
  cudaError_t e = launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, count_dev, capacity, end_bit, sort_scratch,
                                    sort_scratch_bytes_, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("sort_pairs", e);
  return 0;
}

uint32_t trase_b200_higher_msb(uint32_t n) { return higher_msb(n); }

void trase_b200_set_tuning(const char* key, int value) { set_tuning(key, value); }

}  // extern "C"
