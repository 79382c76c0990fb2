// trase_b200_sort_pairs_static — the key/value sort for callers that must not
// allocate device memory.
//
// Who needs this.  Some hosts cannot call cudaMalloc (or anything that does so
// on their behalf) at the point where they sort:
//   * an engine that took the whole device for its own arena at start-up, or a
//     process running under a hard memory cap: a later cudaMalloc simply fails;
//   * a frame or training loop with a latency budget: cudaMalloc / cudaFree take
//     the driver's allocator lock and can synchronise the whole device, which
//     stalls every stream the host is running;
//   * a plug-in called from inside another runtime's callback that forbids
//     allocation while it is active.
// The regular entry point (trase_b200_sort_pairs) asks such a caller for a
// histogram workspace it has no way to provide.  This variant keeps that
// workspace inside the library instead: a fixed array in device global memory
// that exists from the moment the library is loaded on a GPU until the process
// exits.  Keys, values and the element count are still the caller's buffers.
//
// Limits and rules (also in include/trase_b200.h):
//   * at most TRASE_B200_STATIC_SORT_MAX pairs per call; larger requests are
//     refused with an error, nothing is launched;
//   * the workspace costs about 2 MiB on every GPU that loads the library;
//   * every call on a GPU uses the same array and finds the previous call's
//     numbers in it.  The part that has to start from zero (the per-digit
//     totals) is cleared on the caller's stream at the start of each call; the
//     per-block histogram is fully rewritten before it is read;
//   * because the array is shared, two calls whose GPU work overlaps (different
//     streams, or different host threads) race on it.  Run them one after the
//     other;
//   * the first call on each GPU looks up the array's device address
//     (cudaGetSymbolAddress).  Make that first call before starting any stream
//     capture on that GPU.
//   * no cudaMalloc, cudaFree, cudaMallocAsync or allocating library routine is
//     called here (cudaMemsetAsync and kernel launches only).
#include "common.cuh"
#include <mutex>

namespace tb {

constexpr int64_t kStaticSortMax = TRASE_B200_STATIC_SORT_MAX;
// the largest block count any admissible capacity needs: big sorts use 4096-key tiles, sorts up to
// SORT_SMALL_MAX keys 1024-key tiles
constexpr size_t kStaticBlocksBig = (size_t)((kStaticSortMax + SORT_TILE - 1) / SORT_TILE);
constexpr size_t kStaticBlocksSmall = (size_t)((SORT_SMALL_MAX + SORT_TILE_SMALL - 1) / SORT_TILE_SMALL);
constexpr size_t kStaticBlocks = kStaticBlocksBig > kStaticBlocksSmall ? kStaticBlocksBig : kStaticBlocksSmall;
// same layout as sort_scratch_bytes(kStaticSortMax): block histograms, then digit totals
constexpr size_t kStaticBytes =
    ((kStaticBlocks * SORT_RADIX * 4 + WS_ALIGN - 1) / WS_ALIGN) * WS_ALIGN +
    (((size_t)SORT_MAX_PASSES * SORT_RADIX * 4 + WS_ALIGN - 1) / WS_ALIGN) * WS_ALIGN;

// The workspace.  In host code this name is only a handle for the runtime, not
// a device address — it must never be passed to a kernel as a pointer.
static __device__ __align__(256) unsigned char g_static_sort_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
static void* g_scratch_addr[kMaxDevices] = {nullptr};
static std::mutex g_scratch_mu;

// Device address of the workspace on the CURRENT device; resolved once per device.
static cudaError_t static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_static_sort_scratch);
    if (e != cudaSuccess) return e;
    if (ptr == nullptr) return cudaErrorInvalidSymbol;
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return cudaSuccess;
}

}  // namespace tb

using namespace tb;

extern "C" TRASE_API int trase_b200_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out,
                                                      uint32_t* vals_out, const uint32_t* count_dev,
                                                      int64_t capacity, int end_bit, void* stream) {
  clear_error();
  if (capacity < 0 || end_bit < 1 || end_bit > 64) {
    set_error("trase_b200_sort_pairs_static: bad arguments (capacity=%lld, end_bit=%d)", (long long)capacity, end_bit);
    return 1;
  }
  if (capacity > kStaticSortMax) {
    set_error("trase_b200_sort_pairs_static: %lld pairs requested but the built-in workspace holds at most %lld; "
              "use trase_b200_sort_pairs with a caller-provided workspace for larger sorts",
              (long long)capacity, (long long)kStaticSortMax);
    return 1;
  }
  if (capacity == 0) return 0;
  if (!keys_in || !vals_in || !keys_out || !vals_out || !count_dev) {
    set_error("trase_b200_sort_pairs_static: null buffer");
    return 1;
  }
  void* scratch = nullptr;
  cudaError_t e = static_scratch_for_current_device(&scratch);
  if (e != cudaSuccess) {
    set_error("trase_b200_sort_pairs_static: cannot resolve the built-in workspace: %s", cudaGetErrorString(e));
    return 2;
  }
  // launch_sort_pairs clears the digit totals (the only part that must start at
  // zero) with cudaMemsetAsync on `stream`, then runs the same passes as the
  // regular entry point.
  e = launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, count_dev, capacity, end_bit, scratch, kStaticBytes,
                        reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    set_error("trase_b200_sort_pairs_static: CUDA error: %s", cudaGetErrorString(e));
    return 2;
  }
  return 0;
}
