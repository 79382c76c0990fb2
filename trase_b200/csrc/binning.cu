// Binning: block-offset scan, tile|depth key duplication, stable LSD radix sort
// (hand-written, no CUB), tile-range identification.
//
// Reference counterparts:
//   cub::DeviceScan::InclusiveSum      DGR/cuda_rasterizer/rasterizer_impl.cu:284
//   duplicateWithKeys                  DGR/cuda_rasterizer/rasterizer_impl.cu:73-114
//   cub::DeviceRadixSort::SortPairs    DGR/cuda_rasterizer/rasterizer_impl.cu:310-316
//   identifyTileRanges                 DGR/cuda_rasterizer/rasterizer_impl.cu:119-141
//
// Sort strategy: the reference sorts all R tile instances on the 32 depth bits AND the tile bits (45 bits at
// config 2).  Depth is a property of the Gaussian, not of the instance, so here the P Gaussians are sorted by
// depth first (stable, 4 cheap passes over P keys), instances are emitted in that order, and the R instances
// only need a stable sort on the tile bits (2 passes instead of 5 over 12 bytes x R).  Within a tile the
// result is ordered by (depth, Gaussian index) — exactly what the reference's stable 45-bit sort of the
// index-ordered emission yields.
//
// Design differences: the instance count R never goes to the host between
// kernels (the reference blocks on a 4-byte D2H at rasterizer_impl.cu:288);
// every kernel reads R from device memory and grids are sized for the
// caller's capacity.  The prefix sum is folded into the producer (per-block
// sums come out of the preprocess kernel) and the consumer (the duplicate
// kernel scans its own 256 counts), so only the tiny block-sum scan is a
// kernel of its own.
#include "common.cuh"

namespace tb {

__device__ __forceinline__ void tile_rect_dev(const float2 p, int max_radius, int grid_x, int grid_y, uint2& rmin,
                                              uint2& rmax) {
  rmin = {(unsigned)min(grid_x, max((int)0, (int)((p.x - max_radius) / TILE))),
          (unsigned)min(grid_y, max((int)0, (int)((p.y - max_radius) / TILE)))};
  rmax = {(unsigned)min(grid_x, max((int)0, (int)((p.x + max_radius + TILE - 1) / TILE))),
          (unsigned)min(grid_y, max((int)0, (int)((p.y + max_radius + TILE - 1) / TILE)))};
}

// ---------------------------------------------------------------------------
// Exclusive scan of the per-block tile counts (one block; nblk <= a few
// thousand).  Also publishes R = total and the overflow flag.
constexpr int SCAN_THREADS = 1024;
__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums_kernel(const uint32_t* __restrict__ block_sums,
                                                                       uint32_t* __restrict__ block_offsets, int nblk,
                                                                       uint32_t* __restrict__ counters,
                                                                       unsigned long long capacity,
                                                                       int* __restrict__ host_total) {
  pdl_wait();
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < nblk; base += SCAN_THREADS) {
    const int i = base + threadIdx.x;
    const uint32_t v = (i < nblk) ? block_sums[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0;
    for (int w = 0; w < warp; w++) warp_off += s_warp[w];
    const uint32_t excl = s_carry + warp_off + incl - v;
    if (i < nblk) block_offsets[i] = excl;
    __syncthreads();
    if (threadIdx.x == SCAN_THREADS - 1) s_carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const uint32_t total = s_carry;
    block_offsets[nblk] = total;
    counters[0] = total;
    counters[1] = ((unsigned long long)total > capacity) ? 1u : 0u;
    if (host_total) {          // R straight into the caller's pinned host word (zero-copy): no copy-engine operation
      *host_total = (int)total;  // sits in the stream between the emission and the sort
      __threadfence_system();
    }
  }
}

// ---------------------------------------------------------------------------
// Depth keys of the Gaussians (input of the P-sized sort): visible ones by their view-space depth bits,
// the rest behind everything (they emit no instance, their place does not matter).
__global__ void __launch_bounds__(PRE_BLOCK) gaussian_depth_keys_kernel(int P, const float* __restrict__ depths,
                                                                         const uint32_t* __restrict__ tiles_touched,
                                                                         uint64_t* __restrict__ keys,
                                                                         uint32_t* __restrict__ vals,
                                                                         uint32_t* __restrict__ counters) {
  pdl_wait();
  const int i = blockIdx.x * PRE_BLOCK + threadIdx.x;
  if (i < P) {
    keys[i] = tiles_touched[i] ? (uint64_t)__float_as_uint(depths[i]) : 0xffffffffull;
    vals[i] = (uint32_t)i;
  }
  if (i == 0) counters[2] = (uint32_t)P;      // element count of the P-sized sort, read from device memory like R
}

// Tile counts summed per block of PRE_BLOCK Gaussians taken in depth order.
__global__ void __launch_bounds__(PRE_BLOCK) ordered_block_sums_kernel(int P, const uint32_t* __restrict__ order,
                                                                        const uint32_t* __restrict__ tiles_touched,
                                                                        uint32_t* __restrict__ block_sums) {
  pdl_wait();
  __shared__ uint32_t s_warp[PRE_BLOCK / 32];
  const int pos = blockIdx.x * PRE_BLOCK + threadIdx.x;
  uint32_t v = (pos < P) ? tiles_touched[order[pos]] : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < PRE_BLOCK / 32; w++) t += s_warp[w];
    block_sums[blockIdx.x] = t;
  }
}

// ---------------------------------------------------------------------------
// Key duplication.  One thread per Gaussian, taken in depth order (`order`), computes its offset (block scan +
// scanned block offset); the instances of a warp's 32 Gaussians are then written by the whole warp.
__global__ void __launch_bounds__(PRE_BLOCK) duplicate_keys_kernel(
    int P, const uint32_t* __restrict__ order, const float2* __restrict__ means2D, const float* __restrict__ depths,
    const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ block_offsets,
    uint32_t* __restrict__ point_offsets, const uint32_t* __restrict__ counters, uint64_t* __restrict__ keys,
    uint32_t* __restrict__ vals, const int* __restrict__ radii, int grid_x, int grid_y) {
  pdl_wait();
  __shared__ uint32_t s_warp[PRE_BLOCK / 32];
  const int pos = blockIdx.x * PRE_BLOCK + threadIdx.x;
  const int idx = (pos < P) ? (int)order[pos] : P;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t cnt = (idx < P) ? tiles_touched[idx] : 0u;
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0;
#pragma unroll
  for (int w = 0; w < PRE_BLOCK / 32; w++)
    if (w < warp) warp_off += s_warp[w];
  const uint32_t off_incl = block_offsets[blockIdx.x] + warp_off + incl;
  if (idx < P) point_offsets[idx] = off_incl;
  if (counters[1]) return;  // capacity overflow: emit nothing (host retries)

  const uint32_t off = off_incl - cnt;
  uint2 rmin = {0, 0}, rmax = {0, 0};
  uint32_t depth_bits = 0;
  if (cnt > 0) {
    tile_rect_dev(means2D[idx], radii[idx], grid_x, grid_y, rmin, rmax);
    depth_bits = __float_as_uint(depths[idx]);
  }
  // Emission by the whole warp: the warp's instances form one contiguous output range (the offsets are a prefix
  // sum), lane l writes instances l, l+32, ... of that range and finds the owning Gaussian by a binary search
  // over the warp's 32 offsets — consecutive lanes write consecutive keys (a lane writing its own run of ~12
  // keys scattered every store over partial sectors).
  __shared__ uint32_t s_start[PRE_BLOCK / 32][33];
  __shared__ uint4 s_rec[PRE_BLOCK / 32][32];      // x0, y0, width, depth bits
  __shared__ uint32_t s_gid[PRE_BLOCK / 32][32];
  const uint32_t warp_first = __shfl_sync(0xffffffffu, off, 0);
  s_start[warp][lane] = off - warp_first;
  if (lane == 31) s_start[warp][32] = off + cnt - warp_first;
  s_rec[warp][lane] = make_uint4(rmin.x, rmin.y, rmax.x - rmin.x, depth_bits);
  s_gid[warp][lane] = (uint32_t)idx;
  __syncwarp();
  const uint32_t total = s_start[warp][32];
  for (uint32_t t = lane; t < total; t += 32) {
    int lo = 0;                                   // last lane whose start is <= t (empty Gaussians share a start:
#pragma unroll                                    //  the search lands on the last of them, i.e. the non-empty one)
    for (int step = 16; step > 0; step >>= 1)
      if (s_start[warp][lo + step] <= t) lo += step;
    const uint4 rec = s_rec[warp][lo];
    const uint32_t k = t - s_start[warp][lo];
    const uint32_t y = rec.y + k / rec.z, x = rec.x + k % rec.z;
    uint64_t key = (uint64_t)(y * (uint32_t)grid_x + x);
    key <<= 32;
    key |= rec.w;
    keys[warp_first + t] = key;
    vals[warp_first + t] = s_gid[warp][lo];
  }
}

// Depth order of the Gaussians -> block sums in that order -> R and block offsets -> instance keys.
// `p.keys_unsorted / p.vals_unsorted` is where the instances go (the side the R-sized sort starts from);
// (ok_keys, ok_vals) is the OTHER ping-pong side: the P-sized sort starts and (4 passes) ends there, so the
// order array is not overwritten while the instances are being written.
// host_total_dev (optional): device-visible address of the caller's pinned host word for R; readback_event (optional)
// is recorded right behind the scan that writes it, before the key emission is even launched.
cudaError_t launch_scan_and_duplicate(const trase_b200_fwd_args& a, const FwdPointers& p, uint64_t* ok_keys,
                                      uint32_t* ok_vals, int* host_total_dev, cudaEvent_t readback_event,
                                      cudaStream_t s) {
  const int nblk = (a.P + PRE_BLOCK - 1) / PRE_BLOCK;
  launch_k(gaussian_depth_keys_kernel, nblk, PRE_BLOCK, 0, s, a.P, p.depths, p.tiles_touched, ok_keys, ok_vals, p.counters);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = launch_sort_pingpong(ok_keys, ok_vals, p.keys_unsorted, p.vals_unsorted, p.counters + 2, a.P, 0, 32,
                           p.sort_scratch, p.sort_scratch_bytes, s);
  if (e != cudaSuccess) return e;
  static_assert(((32 + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS) % 2 == 0,
                "the depth sort must take an even number of passes so that it ends on the side it started from");
  const uint32_t* order = ok_vals;
  launch_k(ordered_block_sums_kernel, nblk, PRE_BLOCK, 0, s, a.P, order, p.tiles_touched, p.block_sums);
  launch_k(scan_block_sums_kernel, 1, SCAN_THREADS, 0, s, p.block_sums, p.block_offsets, nblk, p.counters,
                                                   (unsigned long long)a.capacity, host_total_dev);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (host_total_dev && readback_event) {
    e = cudaEventRecord(readback_event, s);
    if (e != cudaSuccess) return e;
  }
  const int grid_x = (a.W + TILE - 1) / TILE, grid_y = (a.H + TILE - 1) / TILE;
  launch_k(duplicate_keys_kernel, nblk, PRE_BLOCK, 0, s, a.P, order, p.means2D, p.depths, p.tiles_touched, p.block_offsets,
                                                   p.point_offsets, p.counters, p.keys_unsorted, p.vals_unsorted,
                                                   a.radii, grid_x, grid_y);
  count_launches(4);      // depth keys, ordered block sums, block-sum scan, key emission (the sort counts its own)
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// LSD radix sort, 9-bit digits (SORT_RADIX_BITS), three kernels per pass:
//   upsweep  : per-block digit histogram  → hist[digit][block], digit totals
//   scan     : one block per digit: exclusive scan along blocks + digit base
//   downsweep: stable in-block ranking (warp match-any) and scatter
// Keys are laid out warp-striped inside a block tile (warp w owns 512
// consecutive keys, item j / lane l ↔ key w*512 + j*32 + l) so loads are
// coalesced and "earlier in memory" == (warp, item, lane) lexicographic order,
// which is what the ranking uses to stay stable.
__device__ __forceinline__ uint32_t sort_count(const uint32_t* count_dev, unsigned long long capacity) {
  unsigned long long r = *count_dev;
  return (uint32_t)(r < capacity ? r : capacity);
}

template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS) radix_upsweep_kernel(const uint64_t* __restrict__ keys,
                                                                     const uint32_t* __restrict__ count_dev,
                                                                     unsigned long long capacity, int shift,
                                                                     uint32_t* __restrict__ hist, int nblk_stride,
                                                                     uint32_t* __restrict__ digit_total) {
  pdl_wait();
  const uint32_t R = sort_count(count_dev, capacity);
  const uint32_t block_start = blockIdx.x * (SORT_THREADS * ITEMS);
  if (block_start >= R) return;
  __shared__ uint32_t s_hist[SORT_RADIX];
  for (int d = threadIdx.x; d < SORT_RADIX; d += SORT_THREADS) s_hist[d] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t warp_base = block_start + warp * (32 * ITEMS);
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    uint32_t i = warp_base + j * 32 + lane;
    if (i < R) {
      uint32_t d = (uint32_t)(keys[i] >> shift) & (SORT_RADIX - 1);
      atomicAdd(&s_hist[d], 1u);
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < SORT_RADIX; d += SORT_THREADS) {
    const uint32_t c = s_hist[d];
    hist[(size_t)d * nblk_stride + blockIdx.x] = c;
    if (c) atomicAdd(&digit_total[d], c);
  }
}

__global__ void __launch_bounds__(SORT_THREADS) radix_scan_kernel(uint32_t* __restrict__ hist, int nblk_stride,
                                                                  const uint32_t* __restrict__ digit_total,
                                                                  const uint32_t* __restrict__ count_dev,
                                                                  unsigned long long capacity, int tile_keys) {
  pdl_wait();
  const uint32_t R = sort_count(count_dev, capacity);
  const int nblk = (int)((R + tile_keys - 1) / tile_keys);
  const int d = blockIdx.x;
  __shared__ uint32_t s_warp[SORT_THREADS / 32];
  __shared__ uint32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // digit base = number of keys with a smaller digit
  uint32_t v = 0;
  for (int x = threadIdx.x; x < d; x += SORT_THREADS) v += digit_total[x];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < SORT_THREADS / 32; w++) t += s_warp[w];
    s_carry = t;
  }
  __syncthreads();
  uint32_t* row = hist + (size_t)d * nblk_stride;
  for (int base = 0; base < nblk; base += SORT_THREADS) {
    const int i = base + threadIdx.x;
    const uint32_t x = (i < nblk) ? row[i] : 0u;
    uint32_t incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0;
#pragma unroll
    for (int w = 0; w < SORT_THREADS / 32; w++)
      if (w < warp) warp_off += s_warp[w];
    const uint32_t carry = s_carry;
    const uint32_t excl = carry + warp_off + incl - x;
    if (i < nblk) row[i] = excl;
    __syncthreads();
    if (threadIdx.x == SORT_THREADS - 1) s_carry = excl + x;
    __syncthreads();
  }
}

template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS) radix_downsweep_kernel(
    const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint64_t* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, const uint32_t* __restrict__ count_dev, unsigned long long capacity, int shift,
    const uint32_t* __restrict__ hist, int nblk_stride) {
  pdl_wait();
  const uint32_t R = sort_count(count_dev, capacity);
  const uint32_t block_start = blockIdx.x * (SORT_THREADS * ITEMS);
  if (block_start >= R) return;
  constexpr int NW = SORT_THREADS / 32;
  __shared__ uint32_t s_cnt[NW][SORT_RADIX];
  for (int d = threadIdx.x; d < NW * SORT_RADIX; d += SORT_THREADS) (&s_cnt[0][0])[d] = 0;
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lane_lt = (1u << lane) - 1u;
  const uint32_t warp_base = block_start + warp * (32 * ITEMS);
  uint64_t key[ITEMS];
  uint32_t rank[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    key[j] = (i < R) ? keys_in[i] : ~0ull;
  }
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    const bool valid = i < R;
    const uint32_t d = (uint32_t)(key[j] >> shift) & (SORT_RADIX - 1);
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    const int leader = __ffs(peers) - 1;
    uint32_t c = 0;
    if (valid && lane == leader) c = s_cnt[warp][d];
    c = __shfl_sync(0xffffffffu, c, leader);
    rank[j] = c + __popc(peers & lane_lt);
    if (valid && lane == leader) s_cnt[warp][d] = c + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  for (int d = threadIdx.x; d < SORT_RADIX; d += SORT_THREADS) {
    // per-digit exclusive offsets across warps, plus this block's global base
    uint32_t run = hist[(size_t)d * nblk_stride + blockIdx.x];
#pragma unroll
    for (int w = 0; w < NW; w++) {
      uint32_t t = s_cnt[w][d];
      s_cnt[w][d] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    if (i < R) {
      const uint32_t d = (uint32_t)(key[j] >> shift) & (SORT_RADIX - 1);
      const uint32_t pos = s_cnt[warp][d] + rank[j];
      keys_out[pos] = key[j];
      vals_out[pos] = vals_in[i];
    }
  }
}

// Downsweep with a local reorder: the block's keys are first placed in shared memory in digit order, then
// written out, so that each digit's run leaves as one contiguous piece instead of 8-byte stores scattered
// over up to 32 sectors per request (the tile-bit passes of depth-ordered instances are fully scattered).
// Same ranks, same stable order as radix_downsweep_kernel.  Dynamic shared memory: keys (aliasing the
// per-warp counters once every thread holds its local position), values, three digit tables.
// 3 CTAs per SM (80 registers, 24 bytes of spill) instead of the 2 that 116 registers allow: tile passes 0.142 -> 0.132 ms
#ifndef TB_SORT_MINB
#define TB_SORT_MINB 3
#endif
template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, TB_SORT_MINB) radix_downsweep_staged_kernel(
    const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint64_t* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, const uint32_t* __restrict__ count_dev, unsigned long long capacity, int shift,
    const uint32_t* __restrict__ hist, int nblk_stride) {
  pdl_wait();
  const uint32_t R = sort_count(count_dev, capacity);
  constexpr int TILE_KEYS = SORT_THREADS * ITEMS;
  const uint32_t block_start = blockIdx.x * TILE_KEYS;
  if (block_start >= R) return;
  constexpr int NW = SORT_THREADS / 32;
  constexpr int DPT = SORT_RADIX / SORT_THREADS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint64_t* s_keys = reinterpret_cast<uint64_t*>(s_raw);                       // [TILE_KEYS]  (>= NW*RADIX*4 bytes)
  uint32_t (*s_cnt)[SORT_RADIX] = reinterpret_cast<uint32_t (*)[SORT_RADIX]>(s_raw);   // aliases s_keys
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_raw + (size_t)TILE_KEYS * 8);        // [TILE_KEYS]
  uint32_t* s_lbase = s_vals + TILE_KEYS;                                      // [RADIX] local exclusive base
  uint32_t* s_gdelta = s_lbase + SORT_RADIX;                                   // [RADIX] global base - local base
  __shared__ uint32_t s_warp[NW];
  static_assert((size_t)TILE_KEYS * 8 >= (size_t)NW * SORT_RADIX * 4, "key staging must cover the counters");
  for (int d = threadIdx.x; d < NW * SORT_RADIX; d += SORT_THREADS) (&s_cnt[0][0])[d] = 0;
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lane_lt = (1u << lane) - 1u;
  const uint32_t warp_base = block_start + warp * (32 * ITEMS);
  uint64_t key[ITEMS];
  uint32_t rank[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    key[j] = (i < R) ? keys_in[i] : ~0ull;
  }
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    const bool valid = i < R;
    const uint32_t d = (uint32_t)(key[j] >> shift) & (SORT_RADIX - 1);
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    const int leader = __ffs(peers) - 1;
    uint32_t c = 0;
    if (valid && lane == leader) c = s_cnt[warp][d];
    c = __shfl_sync(0xffffffffu, c, leader);
    rank[j] = c + __popc(peers & lane_lt);
    if (valid && lane == leader) s_cnt[warp][d] = c + __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  // per-warp exclusive offsets inside each digit (local), block totals, global bases
  uint32_t tot[DPT], gbase[DPT];
#pragma unroll
  for (int q = 0; q < DPT; q++) {
    const int d = threadIdx.x + q * SORT_THREADS;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) {
      const uint32_t t = s_cnt[w][d];
      s_cnt[w][d] = run;
      run += t;
    }
    tot[q] = run;
    gbase[q] = hist[(size_t)d * nblk_stride + blockIdx.x];
  }
  // exclusive scan of the block's digit totals in digit order (thread t holds digits t, t + 256)
  uint32_t carry = 0;
#pragma unroll
  for (int q = 0; q < DPT; q++) {
    uint32_t incl = tot[q];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t woff = 0, all = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) {
      if (w < warp) woff += s_warp[w];
      all += s_warp[w];
    }
    const uint32_t lb = carry + woff + incl - tot[q];
    s_lbase[threadIdx.x + q * SORT_THREADS] = lb;
    s_gdelta[threadIdx.x + q * SORT_THREADS] = gbase[q] - lb;
    carry += all;
    __syncthreads();
  }
  // local positions (registers), then the counters' memory becomes the key staging area
  uint32_t lp[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t d = (uint32_t)(key[j] >> shift) & (SORT_RADIX - 1);
    lp[j] = s_lbase[d] + s_cnt[warp][d] + rank[j];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = warp_base + j * 32 + lane;
    if (i < R) {
      s_keys[lp[j]] = key[j];
      s_vals[lp[j]] = vals_in[i];
    }
  }
  __syncthreads();
  const uint32_t n_valid = min((uint32_t)TILE_KEYS, R - block_start);
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    const uint32_t i = j * SORT_THREADS + threadIdx.x;
    if (i < n_valid) {
      const uint64_t k = s_keys[i];
      const uint32_t d = (uint32_t)(k >> shift) & (SORT_RADIX - 1);
      const uint32_t pos = s_gdelta[d] + i;
      keys_out[pos] = k;
      vals_out[pos] = s_vals[i];
    }
  }
}
constexpr int SORT_STAGED_SMEM = SORT_TILE * 12 + 2 * SORT_RADIX * 4;

__global__ void copy_pairs_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                  uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                  const uint32_t* __restrict__ count_dev, unsigned long long capacity) {
  pdl_wait();
  const uint32_t R = sort_count(count_dev, capacity);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < R; i += gridDim.x * blockDim.x) {
    keys_out[i] = keys_in[i];
    vals_out[i] = vals_in[i];
  }
}

int g_sort_staged = 1;   // trase_b200_set_tuning("sort_staged", 0|1)

// Ping-pong sort.  Data starts in (keys_a, vals_a); after `passes` passes it is
// in A again when `passes` is even, in B when odd (see sort_result_in_b).
int sort_num_passes(int begin_bit, int end_bit) {
  int passes = (end_bit - begin_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  if (passes < 1) passes = 1;
  if (passes > SORT_MAX_PASSES) passes = SORT_MAX_PASSES;
  return passes;
}
cudaError_t launch_sort_pingpong(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b,
                                 const uint32_t* count_dev, int64_t capacity, int begin_bit, int end_bit,
                                 void* scratch, size_t scratch_bytes, cudaStream_t s) {
  if (capacity <= 0) return cudaSuccess;
  if (scratch_bytes < sort_scratch_bytes(capacity)) return cudaErrorInvalidValue;
  const int nblk = (int)sort_num_blocks(capacity);
  uint32_t* hist = reinterpret_cast<uint32_t*>(scratch);
  uint32_t* digit_total =
      reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(scratch) + align_up((size_t)nblk * SORT_RADIX * 4, WS_ALIGN));
  const int passes = sort_num_passes(begin_bit, end_bit);
  cudaError_t e = cudaMemsetAsync(digit_total, 0, (size_t)SORT_MAX_PASSES * SORT_RADIX * 4, s);
  if (e != cudaSuccess) return e;
  uint64_t* kin = keys_a; uint32_t* vin = vals_a;
  uint64_t* kout = keys_b; uint32_t* vout = vals_b;
  const unsigned long long cap = (unsigned long long)capacity;
  // (A one-launch-per-pass variant for small sorts — histogram, grid-wide barrier, prefix, scatter in one resident
  // kernel — was measured for the P-sized depth sort: same 27 us per pass as these three launches, both are bound
  // by memory latency at 74 blocks, so the simpler form stays.)
  for (int pass = 0; pass < passes; pass++) {
    const int shift = begin_bit + pass * SORT_RADIX_BITS;
    uint32_t* tot = digit_total + pass * SORT_RADIX;
    if (sort_is_small(capacity)) {
      launch_k(radix_upsweep_kernel<SORT_ITEMS_SMALL>, nblk, SORT_THREADS, 0, s, kin, count_dev, cap, shift, hist, nblk, tot);
      launch_k(radix_scan_kernel, SORT_RADIX, SORT_THREADS, 0, s, hist, nblk, tot, count_dev, cap, SORT_TILE_SMALL);
      launch_k(radix_downsweep_kernel<SORT_ITEMS_SMALL>, nblk, SORT_THREADS, 0, s, kin, vin, kout, vout, count_dev, cap, shift,
                                                                            hist, nblk);
    } else {
      launch_k(radix_upsweep_kernel<SORT_ITEMS>, nblk, SORT_THREADS, 0, s, kin, count_dev, cap, shift, hist, nblk, tot);
      launch_k(radix_scan_kernel, SORT_RADIX, SORT_THREADS, 0, s, hist, nblk, tot, count_dev, cap, SORT_TILE);
      if (g_sort_staged) {
        static PerDeviceOnce attr_once;
        e = ensure_max_dyn_smem(attr_once, radix_downsweep_staged_kernel<SORT_ITEMS>, SORT_STAGED_SMEM);
        if (e != cudaSuccess) return e;
        launch_k(radix_downsweep_staged_kernel<SORT_ITEMS>, nblk, SORT_THREADS, SORT_STAGED_SMEM, s, 
            kin, vin, kout, vout, count_dev, cap, shift, hist, nblk);
      } else {
        launch_k(radix_downsweep_kernel<SORT_ITEMS>, nblk, SORT_THREADS, 0, s, kin, vin, kout, vout, count_dev, cap, shift, hist,
                                                                        nblk);
      }
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    uint64_t* tk = kin; kin = kout; kout = tk;
    uint32_t* tv = vin; vin = vout; vout = tv;
  }
  count_launches(3 * passes);
  return cudaSuccess;
}

cudaError_t launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                              const uint32_t* count_dev, int64_t capacity, int end_bit, void* scratch,
                              size_t scratch_bytes, cudaStream_t s) {
  if (capacity <= 0) return cudaSuccess;
  cudaError_t e = launch_sort_pingpong(keys_in, vals_in, keys_out, vals_out, count_dev, capacity, 0, end_bit, scratch,
                                       scratch_bytes, s);
  if (e != cudaSuccess) return e;
  if ((sort_num_passes(0, end_bit) & 1) == 0) {  // result sits in the input buffers
    int g = (int)((capacity + 1023) / 1024);
    if (g < 1) g = 1;
    launch_k(copy_pairs_kernel, g, SORT_THREADS, 0, s, keys_in, vals_in, keys_out, vals_out, count_dev,
                                                 (unsigned long long)capacity);
    e = cudaGetLastError();
  }
  return e;
}

// ---------------------------------------------------------------------------
// Tile ranges: ranges[t] = [first, last) positions of tile t in the sorted list.
__global__ void identify_ranges_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ counters,
                                       unsigned long long capacity, uint2* __restrict__ ranges) {
  pdl_wait();
  if (counters[1]) return;  // overflow: leave every range empty
  const uint32_t L = sort_count(counters, capacity);
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < L; idx += gridDim.x * blockDim.x) {
    const uint32_t currtile = (uint32_t)(keys[idx] >> 32);
    if (idx == 0)
      ranges[currtile].x = 0;
    else {
      const uint32_t prevtile = (uint32_t)(keys[idx - 1] >> 32);
      if (currtile != prevtile) {
        ranges[prevtile].y = idx;
        ranges[currtile].x = idx;
      }
    }
    if (idx == L - 1) ranges[currtile].y = L;
  }
}

cudaError_t launch_identify_ranges(const trase_b200_fwd_args& a, const FwdPointers& p, cudaStream_t s) {
  const int grid_x = (a.W + TILE - 1) / TILE, grid_y = (a.H + TILE - 1) / TILE;
  cudaError_t e = cudaMemsetAsync(p.ranges, 0, (size_t)grid_x * grid_y * sizeof(uint2), s);
  if (e != cudaSuccess) return e;
  if (a.capacity <= 0) return cudaSuccess;
  long long g = (a.capacity + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  launch_k(identify_ranges_kernel, (int)g, 256, 0, s, p.point_list_keys, p.counters, (unsigned long long)a.capacity,
                                                p.ranges);
  return cudaGetLastError();
}

// Number of tile-id bits the sort has to cover; equals what the reference's
// getHigherMsb returns (rasterizer_impl.cu:38-53): the bit length of n, at least 1.
uint32_t higher_msb(uint32_t n) {
  uint32_t bits = 0;
  while (bits < 32 && (n >> bits)) bits++;
  return bits < 1 ? 1 : bits;
}

}  // namespace tb
