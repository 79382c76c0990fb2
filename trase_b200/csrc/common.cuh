// Shared definitions for the trase_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/trase_b200.h"

namespace tb {

constexpr int TILE = TRASE_B200_TILE;           // 16x16 pixel tiles
constexpr int NUM_OBJECTS = TRASE_B200_NUM_OBJECTS;
constexpr int PRE_BLOCK = 256;                  // threads per preprocess block

constexpr size_t WS_ALIGN = 256;

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- workspace layouts (see include/trase_b200.h) --------------------------
struct GeomLayout {
  size_t off[TRASE_GEOM_NUM_ARRAYS];
  size_t total;
};
inline GeomLayout geom_layout(int P) {
  GeomLayout L;
  size_t p = (size_t)(P > 0 ? P : 0);
  size_t nblk = (p + PRE_BLOCK - 1) / PRE_BLOCK;
  size_t cur = 0;
  auto take = [&](int which, size_t bytes) {
    L.off[which] = cur;
    cur = align_up(cur + bytes, WS_ALIGN);
  };
  take(TRASE_GEOM_DEPTHS, p * 4);
  take(TRASE_GEOM_MEANS2D, p * 8);
  take(TRASE_GEOM_CONIC_OPACITY, p * 16);
  take(TRASE_GEOM_RGBD, p * 16);
  take(TRASE_GEOM_COV3D, p * 24);
  take(TRASE_GEOM_CLAMPED, p);
  take(TRASE_GEOM_TILES_TOUCHED, p * 4);
  take(TRASE_GEOM_POINT_OFFSETS, p * 4);
  take(TRASE_GEOM_BLOCK_SUMS, 2 * (nblk + 1) * 4);  // raw sums, then scanned offsets
  take(TRASE_GEOM_COUNTERS, 8 * 4);
  L.total = cur;
  return L;
}
struct ImgLayout {
  size_t off[TRASE_IMG_NUM_ARRAYS];
  size_t total;
};
inline ImgLayout img_layout(int W, int H) {
  ImgLayout L;
  size_t n = (size_t)W * (size_t)H;
  size_t tiles = (size_t)((W + TILE - 1) / TILE) * (size_t)((H + TILE - 1) / TILE);
  size_t cur = 0;
  auto take = [&](int which, size_t bytes) {
    L.off[which] = cur;
    cur = align_up(cur + bytes, WS_ALIGN);
  };
  take(TRASE_IMG_FINAL_T, n * 4);
  take(TRASE_IMG_N_CONTRIB, n * 4);
  take(TRASE_IMG_RANGES, tiles * 8);
  L.total = cur;
  return L;
}
struct BinLayout {
  size_t off[TRASE_BIN_NUM_ARRAYS];
  size_t total;
};
inline BinLayout bin_layout(int64_t capacity) {
  BinLayout L;
  size_t c = (size_t)(capacity > 0 ? capacity : 0);
  size_t cur = 0;
  auto take = [&](int which, size_t bytes) {
    L.off[which] = cur;
    cur = align_up(cur + bytes, WS_ALIGN);
  };
  take(TRASE_BIN_POINT_LIST, c * 4);
  take(TRASE_BIN_POINT_LIST_KEYS, c * 8);
  L.total = cur;
  return L;
}

// ---- radix sort geometry ---------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;                       // keys per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // keys per block
constexpr int SORT_RADIX_BITS = 9;   // 13 tile bits + 32 depth bits = 45 = 5 passes of 9 (8-bit digits needed 6)
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_MAX_PASSES = 8;

// small sorts (the P-sized depth sort at config 2: 300 k keys) use 1024-key tiles: 4x the blocks, a quarter of the
// serial work per block — at 74 blocks of 4096 keys a pass is bound by memory latency, not bandwidth
constexpr int SORT_ITEMS_SMALL = 4;
constexpr int SORT_TILE_SMALL = SORT_THREADS * SORT_ITEMS_SMALL;
constexpr int64_t SORT_SMALL_MAX = 1 << 20;
inline bool sort_is_small(int64_t capacity) { return capacity <= SORT_SMALL_MAX; }
inline size_t sort_num_blocks(int64_t capacity) {
  const int tile = sort_is_small(capacity) ? SORT_TILE_SMALL : SORT_TILE;
  return (size_t)((capacity + tile - 1) / tile);
}
// scratch: [block_hist: RADIX * nblocks u32][digit_totals: MAX_PASSES * RADIX u32]
// Monotone in `capacity`: a workspace sized for `capacity` serves every sort of n <= capacity keys, whichever tile
// size n selects (the P-sized depth sort borrows the workspace sized for the instance capacity, and a capacity just
// above SORT_SMALL_MAX needs fewer 4096-key blocks than a smaller sort needs 1024-key blocks).
inline size_t sort_scratch_bytes(int64_t capacity) {
  const int64_t small_part = capacity < SORT_SMALL_MAX ? capacity : SORT_SMALL_MAX;
  const size_t nb_small = (size_t)((small_part + SORT_TILE_SMALL - 1) / SORT_TILE_SMALL);
  const size_t nb_big = capacity > SORT_SMALL_MAX ? (size_t)((capacity + SORT_TILE - 1) / SORT_TILE) : 0;
  size_t nb = nb_small > nb_big ? nb_small : nb_big;
  if (nb == 0) nb = 1;
  return align_up(nb * SORT_RADIX * 4, WS_ALIGN) + align_up(SORT_MAX_PASSES * SORT_RADIX * 4, WS_ALIGN);
}
// forward scratch: unsorted keys (u64), unsorted values (u32), sort scratch
struct ScratchLayout {
  size_t keys_unsorted, vals_unsorted, sort, total;
};
inline ScratchLayout scratch_layout(int64_t capacity) {
  ScratchLayout L;
  size_t c = (size_t)(capacity > 0 ? capacity : 0);
  size_t cur = 0;
  L.keys_unsorted = cur; cur = align_up(cur + c * 8, WS_ALIGN);
  L.vals_unsorted = cur; cur = align_up(cur + c * 4, WS_ALIGN);
  L.sort = cur; cur = align_up(cur + sort_scratch_bytes(capacity), WS_ALIGN);
  L.total = cur;
  return L;
}

// number of kernels this library has launched (bench.py's gpu_launches)
extern std::atomic<unsigned long long> g_launch_count;
inline void count_launches(int n) { g_launch_count.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// host-side error recording (api.cu); the message is per host thread
void set_error(const char* fmt, ...);
void clear_error();

// Function attributes (the opt-in to more than 48 KB of dynamic shared memory) belong to the (function, device)
// pair, not to the process: one process driving several GPUs has to set them on each.  One flag per device.
constexpr int TB_MAX_DEVICES = 64;
struct PerDeviceOnce {
  std::atomic<unsigned char> done[TB_MAX_DEVICES];
};
template <typename K>
inline cudaError_t ensure_max_dyn_smem(PerDeviceOnce& once, K kernel, int bytes) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const bool tracked = dev >= 0 && dev < TB_MAX_DEVICES;
  if (tracked && once.done[dev].load(std::memory_order_acquire)) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && tracked) once.done[dev].store(1, std::memory_order_release);
  return e;
}

// launchers (one per .cu)
struct FwdPointers {
  // geom
  float* depths; float2* means2D; float4* conic_opacity; float4* rgbd; float* cov3D;
  uint8_t* clamped; uint32_t* tiles_touched; uint32_t* point_offsets; uint32_t* block_sums; uint32_t* block_offsets;
  uint32_t* counters;
  // img
  float* final_T; uint32_t* n_contrib; uint2* ranges;
  // binning
  uint32_t* point_list; uint64_t* point_list_keys;
  // scratch
  uint64_t* keys_unsorted; uint32_t* vals_unsorted; void* sort_scratch; size_t sort_scratch_bytes;
};

// Programmatic dependent launch.  A step is 60-110 launches, most of them small, and the gap between two dependent
// kernels of a stream (previous grid drained -> next grid's blocks scheduled) is ~2.3 us.  With TB_PDL every kernel of
// the library is launched with cudaLaunchAttributeProgrammaticStreamSerialization, so its blocks are scheduled while the
// previous kernel of the stream drains, and every kernel starts with griddepcontrol.wait (= the previous grid is
// complete and its memory visible) before it touches global memory.  ~1 us per launch: binning 0.28 -> 0.23 ms,
// config 2 4.55 -> 4.47 ms, config 3 17.25 -> 16.68 ms (profiles/r2_ah_pdl.txt).  -DTB_PDL=0 restores plain launches.
// Rule for new kernels: launch through launch_k() and call pdl_wait() first thing in the kernel — one without the
// other is either a wasted attribute or a race.
#ifndef TB_PDL
#define TB_PDL 1
#endif
__device__ __forceinline__ void pdl_wait() {
#if TB_PDL
  // (no early griddepcontrol.launch_dependents: dependents that become resident during this grid's last wave take
  // slots from it — binning 0.23 -> 0.26 ms, measured in profiles/r2_ah_pdl.txt)
  asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
template <typename K, typename... Args>
inline cudaError_t launch_k(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
#if TB_PDL
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
#else
  kernel<<<grid, block, smem, s>>>(args...);
  return cudaGetLastError();
#endif
}

cudaError_t launch_preprocess_forward(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s);
cudaError_t launch_scan_and_duplicate(const trase_b200_fwd_args& a, const FwdPointers& p, uint64_t* ok_keys,
                                      uint32_t* ok_vals, int* host_total_dev, cudaEvent_t readback_event,
                                      cudaStream_t s);
cudaError_t launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                              const uint32_t* count_dev, int64_t capacity, int end_bit, void* scratch,
                              size_t scratch_bytes, cudaStream_t s);
int sort_num_passes(int begin_bit, int end_bit);
cudaError_t launch_sort_pingpong(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b,
                                 const uint32_t* count_dev, int64_t capacity, int begin_bit, int end_bit,
                                 void* scratch, size_t scratch_bytes, cudaStream_t s);
cudaError_t launch_identify_ranges(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s);
cudaError_t launch_render_forward(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s);
cudaError_t launch_render_backward(const trase_b200_bwd_args& a, const FwdPointers& p, cudaStream_t s);
cudaError_t launch_preprocess_backward(const trase_b200_bwd_args& a, const FwdPointers& p, cudaStream_t s);
cudaError_t launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);

uint32_t higher_msb(uint32_t n);
void set_tuning(const char* key, int value);

}  // namespace tb
