// TF32 tensor-core GEMMs for the DeformNetwork MLP (reference: utils/time_utils.py:106-131,
// eleven nn.Linear layers evaluated by cuBLAS SGEMM on CUDA cores).
//
//   gemm_tn :  D[M,N] = epilogue( A[M,K] · B[N,K]^T )      A, B row-major (K contiguous)
//              epilogue: (+ bias[n]) -> (ReLU) -> (· mask bit) ; side outputs: ReLU bit mask, column sums
//     forward  layer : A = activations, B = weight [out,in], bias, ReLU, writes 1 bit/element (output > 0)
//     backward dgrad : A = dZ [M,out], B = weight^T [in,out], mask = that bit mask of the layer input,
//                      column sums = bias gradient of the previous layer
//
// Blackwell structure: persistent CTAs (one per SM), TMA 128-byte-swizzled tiles
// into a 3-stage shared-memory ring, a single elected thread issues
// tcgen05.mma.kind::tf32 (128 x BN x 8) with the fp32 accumulator in TMEM
// (double-buffered: 2 x BN columns), four epilogue warps drain TMEM with
// tcgen05.ld, apply the epilogue, transpose 32x32 sub-slabs through swizzled shared memory and
// write full 128-byte lines,
// overlapping the next tile's MMAs.  K and M tails are handled by TMA
// zero-fill / clipping, so no operand is ever padded in memory.
#include "common.cuh"
#include "ptx.cuh"
#include <cuda_bf16.h>

namespace tb {
namespace mlp {

using namespace ptx;

constexpr int BM = 128;          // rows per tile (= TMEM lanes)
constexpr int BK = 32;           // tf32 elements per k-block = one 128-byte swizzle row
constexpr int UMMA_K = 8;        // tf32 elements per tcgen05.mma
constexpr int STAGES = 3;
constexpr int SLAB = 32;         // output columns per epilogue slab
constexpr int EPI_WARPS = 8;      // two warps per TMEM lane quarter, alternating slabs: hides TMEM/LDS/TMA-store latency
constexpr int GEMM_THREADS = 128 + 32 * EPI_WARPS;

struct GemmEpilogue {
  const float* bias;  // [N] or null
  int relu;
  const uint32_t* mask_in;   // [M][N/32] bit j of word s: keep column 32s+j (ReLU derivative), or null
  uint32_t* mask_out;        // [M][N/32] written as (output > 0), or null
  float* colsum;      // [N] or null: += column sums of the written values (bias gradient)
  float* D; int ldd;  // output, row pitch in floats
  int round_tf32;     // round outputs to TF32 (nearest) so the next GEMM's operand truncation is exact
  int M, N, K;
};

// (A variant that kept the weights resident in shared memory, an L2 prefetch of the activation stream and a
// serpentine tile order were measured and dropped: 154 us vs 120 us per layer, no gain, no gain.)
template <int BN>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 4;       // 16 KB
  static constexpr int B_BYTES = BN * BK * 4;       // one k-block of B: 32 KB (BN = 256), 16 KB (BN = 128)
  static constexpr int OUT_BYTES = BM * SLAB * 4;   // 16 KB
  static constexpr int OFF_A = 0;
  static constexpr int OFF_B = OFF_A + STAGES * A_BYTES;
  static constexpr int OFF_OUT = OFF_B + STAGES * B_BYTES;
  static constexpr int OFF_BIAS = OFF_OUT + EPI_WARPS * 4096;   // per epilogue warp: 32 rows x 128 B staging
  static constexpr int OFF_COL = OFF_BIAS + (BN < 32 ? 32 : BN) * 4;   // [4 warps][BN] bias-gradient partial sums
  static constexpr int OFF_BAR = OFF_COL + 4 * (BN < 32 ? 32 : BN) * 4;
  static constexpr int TOTAL = OFF_BAR + 256 + 1024;  // barriers + slack for 1024-byte alignment
};

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tn_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmEpilogue ep) {
  pdl_wait();
  using L = GemmSmem<BN>;
  // 1024-byte alignment (128-byte swizzle atoms) comes from the declaration: aligning the pointer
  // by integer arithmetic would hide the shared address space from the compiler and turn every
  // LDS/STS of the epilogue into a generic LD/ST.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u)) __trap();
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sB = smem + L::OFF_B;
  uint8_t* sOut = smem + L::OFF_OUT;
  float* sBias = reinterpret_cast<float*>(smem + L::OFF_BIAS);
  float* sCol = reinterpret_cast<float*>(smem + L::OFF_COL);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::OFF_BAR);
  uint64_t* full = bars;                   // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;         // [STAGES]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * STAGES; // [2]       MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;    // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = (ep.M + BM - 1) / BM;
  const int num_kb = (ep.K + BK - 1) / BK;
  constexpr int n_off = 0;
  const int tile0 = (int)blockIdx.x;
  const int tile_step = (int)gridDim.x;
#define TB_TILE(t) (t)
  constexpr uint32_t TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;  // power of two >= 32 (BN = 16 -> 32, 256 -> 512)

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < (BN < 32 ? 32 : BN); i += GEMM_THREADS) sBias[i] = (ep.bias && n_off + i < ep.N) ? ep.bias[n_off + i] : 0.f;
  for (int i = threadIdx.x; i < 4 * (BN < 32 ? 32 : BN); i += GEMM_THREADS) sCol[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], L::A_BYTES + L::B_BYTES);
          tma_load_2d(sA + stage * L::A_BYTES, &tmA, &full[stage], kb * BK, TB_TILE(tile) * BM);
          tma_load_2d(sB + stage * L::B_BYTES, &tmB, &full[stage], kb * BK, 0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(BM, BN, 0, 0);
      uint32_t stage = 0, phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * L::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * L::B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; k++) {
            // K-major, 128-byte swizzle: rows are 128 B apart, 8-row groups 1024 B apart;
            // stepping K by 8 tf32 inside the swizzle row = +32 bytes on the start address
            const uint64_t da = make_smem_desc_sw128(a_addr + k * UMMA_K * 4, 16, 1024);
            const uint64_t db = make_smem_desc_sw128(b_addr + k * UMMA_K * 4, 16, 1024);
            mma_tf32_ss(tmem_d, da, db, idesc, (kb | k) != 0);
          }
          mma_commit(&empty[stage]);  // frees the smem stage when these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: 4 independent warps; warp (w & 3) owns TMEM lanes / tile rows 32*(w&3) .. +31,
    // stages its own 32x32 sub-slab (4 KB, 128-byte swizzled) and stores it with its own TMA store,
    // so the warps never synchronise with each other =====
    const int ew = warp & 3;
    const int eh = (warp - 4) >> 2;           // which half of the slabs this warp takes
    constexpr int CW = (BN < 32 ? 32 : BN);
    uint8_t* out = sOut + (warp - 4) * 4096;
    float* sColW = sCol + ew * CW;
    uint32_t acc = 0, acc_phase = 0;
    constexpr int NSLAB = (BN + SLAB - 1) / SLAB;
    const int words = (ep.N + 31) >> 5;      // mask words per row
    for (int tile = tile0; tile < num_tiles; tile += tile_step) {
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int rtile = TB_TILE(tile);
      const int grow = rtile * BM + ew * 32 + lane;   // global row of this thread
#pragma unroll 1
      for (int s = eh; s < NSLAB; s += EPI_WARPS / 4) {
        if (n_off + s * SLAB >= ep.N) break;
        const int ws = (n_off >> 5) + s;            // mask word / global slab index
        uint32_t keep = 0xffffffffu;
        if (ep.mask_in) keep = (grow < ep.M) ? __ldg(ep.mask_in + (size_t)grow * words + ws) : 0u;
        uint32_t v[32];
        constexpr int W = (BN < SLAB) ? BN : SLAB;  // columns actually present in this slab
        if constexpr (W == 32) {
          tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + acc * BN + s * SLAB, v);
        } else {
          uint32_t v16[16];
          tmem_ld_32x16(tmem_base + ((uint32_t)(ew * 32) << 16) + acc * BN + s * SLAB, v16);
#pragma unroll
          for (int i = 0; i < 16; i++) { v[i] = v16[i]; v[i + 16] = 0; }
        }
        // bias of this slab, fetched while the TMEM load is in flight
        float4 bq[8];
#pragma unroll
        for (int c = 0; c < 8; c++) bq[c] = *reinterpret_cast<const float4*>(sBias + s * SLAB + c * 4);
        tmem_ld_wait();
        __syncwarp();
        uint32_t pos = 0;
        // element (row = lane, 16-byte chunk c) lives at lane*128 + ((c ^ (lane & 7)) << 4)
#pragma unroll
        for (int c = 0; c < 8; c++) {
          const float bb[4] = {bq[c].x, bq[c].y, bq[c].z, bq[c].w};
          float4 o;
          float* of = reinterpret_cast<float*>(&o);
#pragma unroll
          for (int j = 0; j < 4; j++) {
            float x = __uint_as_float(v[c * 4 + j]) + bb[j];
            if (ep.relu) x = fmaxf(x, 0.f);
            if (!((keep >> (c * 4 + j)) & 1u)) x = 0.f;
            if (ep.round_tf32) {
              uint32_t u;
              asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
              x = __uint_as_float(u);
            }
            pos |= (x > 0.f ? 1u : 0u) << (c * 4 + j);
            of[j] = x;
          }
          *reinterpret_cast<float4*>(out + lane * 128 + ((c ^ (lane & 7)) << 4)) = o;
        }
        if (ep.mask_out && grow < ep.M) ep.mask_out[(size_t)grow * words + ws] = pos;
        __syncwarp();
        // copy-out: 8 lanes cover one 128-byte row segment, 4 rows per instruction -> full-line stores
        {
          const int chunk = lane & 7;
          const int col0 = n_off + s * SLAB + chunk * 4;
#pragma unroll
          for (int it = 0; it < 8; it++) {
            const int r = it * 4 + (lane >> 3);
            const float4 val = *reinterpret_cast<const float4*>(out + r * 128 + ((chunk ^ (r & 7)) << 4));
            const int gr = rtile * BM + ew * 32 + r;
            if (gr < ep.M && col0 < ep.N) {
              float* dst = ep.D + (size_t)gr * ep.ldd + col0;
              if (col0 + 3 < ep.N) {
                *reinterpret_cast<float4*>(dst) = val;
              } else {
                const float vv[4] = {val.x, val.y, val.z, val.w};
                for (int j = 0; j < 4 && col0 + j < ep.N; j++) dst[j] = vv[j];
              }
            }
          }
        }
        if (ep.colsum) {
          // bias gradient: lane c sums column c over this warp's 32 staged rows
          float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
          for (int r = 0; r < 32; r += 4) {
            a0 += *reinterpret_cast<const float*>(out + (r + 0) * 128 + (((lane >> 2) ^ ((r + 0) & 7)) << 4) + ((lane & 3) << 2));
            a1 += *reinterpret_cast<const float*>(out + (r + 1) * 128 + (((lane >> 2) ^ ((r + 1) & 7)) << 4) + ((lane & 3) << 2));
            a2 += *reinterpret_cast<const float*>(out + (r + 2) * 128 + (((lane >> 2) ^ ((r + 2) & 7)) << 4) + ((lane & 3) << 2));
            a3 += *reinterpret_cast<const float*>(out + (r + 3) * 128 + (((lane >> 2) ^ ((r + 3) & 7)) << 4) + ((lane & 3) << 2));
          }
          sColW[s * SLAB + lane] += (a0 + a1) + (a2 + a3);
        }
        __syncwarp();   // staging buffer is rewritten by the next slab
      }
      // all TMEM reads of this accumulator are done: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (ep.colsum) {
      // each warp flushes the columns of the slabs it owned
      for (int s = eh; s < NSLAB; s += EPI_WARPS / 4)
        if (n_off + s * SLAB + lane < ep.N) atomicAdd(ep.colsum + n_off + s * SLAB + lane, sColW[s * SLAB + lane]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
#undef TB_TILE
}

// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D fp32 tensor [rows][cols] with row pitch `ld` floats; box = box_rows x box_cols, 128-byte swizzle
bool make_tmap_2d(CUtensorMap* m, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                  uint32_t box_cols, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 4};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

static int g_num_sms = 0;
// SMs the persistent MLP kernels leave free (trase_b200_set_tuning "reserve_sms").  With view-parallel training the
// NCCL all-reduce of the Gaussians' gradients runs under the MLP backward; a persistent kernel that wants every
// SM would find some taken by NCCL's CTAs and run its last CTAs as a second wave.
int g_reserve_sms = 0;
int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  const int n = g_num_sms - g_reserve_sms;
  return n > 2 ? n : 2;
}

template <int BN>
static cudaError_t launch_gemm_tn(const CUtensorMap& tA, const CUtensorMap& tB, const GemmEpilogue& ep,
                                  cudaStream_t s) {
  static PerDeviceOnce attr_once;
  const int smem = GemmSmem<BN>::TOTAL;
  {
    cudaError_t e = ensure_max_dyn_smem(attr_once, gemm_tn_tf32_kernel<BN>, smem);
    if (e != cudaSuccess) return e;
  }
  const int tiles = (ep.M + BM - 1) / BM;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  launch_k(gemm_tn_tf32_kernel<BN>, grid, GEMM_THREADS, smem, s, tA, tB, ep);
  return cudaGetLastError();
}

}  // namespace mlp
}  // namespace tb

using namespace tb;

// D[M,N] = epilogue(A[M,K] · B[N,K]^T); lda/ldb/ldd are row pitches in floats (multiples of 4), all base
// pointers 16-byte aligned, N <= 256.  mask_in / mask_out: [M][ceil(N/32)] bit masks (ReLU derivative).
extern "C" TRASE_API int trase_b200_gemm_tf32(const float* A, int lda, const float* B, int ldb, float* D, int ldd,
                                              int M, int N, int K, const float* bias, int relu,
                                              const uint32_t* mask_in, uint32_t* mask_out, float* colsum,
                                              int round_tf32, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0 || N > 256 || !A || !B || !D) {
    set_error("trase_b200_gemm_tf32: bad arguments (M=%d N=%d K=%d)", M, N, K);
    return 1;
  }
  if ((lda & 3) || (ldb & 3) || (ldd & 3) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) || ((uintptr_t)D & 15)) {
    set_error("trase_b200_gemm_tf32: pointers must be 16-byte aligned and row pitches multiples of 4 floats");
    return 1;
  }
  const bool small = N <= 16;
  CUtensorMap tA, tB;
  const bool ok = mlp::make_tmap_2d(&tA, A, M, K, lda, mlp::BM, mlp::BK) &&
                  mlp::make_tmap_2d(&tB, B, N, K, ldb, small ? 16 : 256, mlp::BK);
  if (!ok) {
    set_error("trase_b200_gemm_tf32: cuTensorMapEncodeTiled failed");
    return 2;
  }
  mlp::GemmEpilogue ep;
  ep.bias = bias; ep.relu = relu; ep.mask_in = mask_in; ep.mask_out = mask_out; ep.colsum = colsum;
  ep.round_tf32 = round_tf32; ep.D = D; ep.ldd = ldd; ep.M = M; ep.N = N; ep.K = K;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  cudaError_t e = small ? mlp::launch_gemm_tn<16>(tA, tB, ep, st) : mlp::launch_gemm_tn<256>(tA, tB, ep, st);
  if (e != cudaSuccess) {
    set_error("trase_b200_gemm_tf32: CUDA error: %s", cudaGetErrorString(e));
    return 2;
  }
  count_launches(1);
  return 0;
}

// ===========================================================================
// Weight gradient:  D[i, j] += sum_m P[m, i] * Q[m, j]      (dW = dZ^T X)
//
// Both operands are read with the non-reduced dimension contiguous ("MN-major"
// for the tensor core), straight from the row-major activations: no transposed
// copies.  The m range is split across the persistent CTAs; each keeps the whole
// I x J (<= 256 x 256) fp32 result in TMEM (2 x 128 lanes x J columns), then adds
// it to global memory with 16-byte vector reductions.
namespace tb {
namespace mlp {

constexpr int WG_ROWS = 32;     // m rows per stage (4 MMAs of K = 8)
constexpr int WG_STAGES = 3;
constexpr int WG_BOX_BYTES = WG_ROWS * 128;  // one 32-row x 32-col fp32 box

struct WgradParams {
  int M, I, J;      // reduction length, output rows (<= 256), output cols (<= 256)
  float* D; int ldd;
};

constexpr int WG_THREADS = 256;
__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_tf32_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ,
                  const WgradParams p) {
  pdl_wait();
  // 1024-byte alignment (128-byte swizzle atoms) comes from the declaration: aligning the pointer
  // by integer arithmetic would hide the shared address space from the compiler and turn every
  // LDS/STS of the epilogue into a generic LD/ST.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u)) __trap();
  constexpr int OP_BYTES = 8 * WG_BOX_BYTES;            // up to 256 columns per operand = 32 KB
  uint8_t* sP = smem;
  uint8_t* sQ = smem + WG_STAGES * OP_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * WG_STAGES * OP_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + WG_STAGES;
  uint64_t* done = bars + 2 * WG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_halves = (p.I + 127) / 128;
  const int i_boxes = n_halves * 4;                      // 32-column boxes of P
  const int j_boxes = (p.J + 31) / 32;
  const int jpad = ((p.J + 15) / 16) * 16;               // N of the MMA
  const int nblk = (p.M + WG_ROWS - 1) / WG_ROWS;
  const int per = (nblk + gridDim.x - 1) / gridDim.x;
  const int blk0 = blockIdx.x * per;
  const int blk1 = min(nblk, blk0 + per);
#define WG_BLK(b) (b)
  constexpr uint32_t TMEM_COLS = 512;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmP); prefetch_tmap(&tmQ);
    for (int i = 0; i < WG_STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int b = blk0; b < blk1; b++) {
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_expect_tx(&full[stage], (i_boxes + j_boxes) * WG_BOX_BYTES);
        for (int c = 0; c < i_boxes; c++)
          tma_load_2d(sP + stage * OP_BYTES + c * WG_BOX_BYTES, &tmP, &full[stage], c * 32, WG_BLK(b) * WG_ROWS);
        for (int c = 0; c < j_boxes; c++)
          tma_load_2d(sQ + stage * OP_BYTES + c * WG_BOX_BYTES, &tmQ, &full[stage], c * 32, WG_BLK(b) * WG_ROWS);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, jpad, 1, 1);
      uint32_t stage = 0, phase = 0;
      for (int b = blk0; b < blk1; b++) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t pa = smem_u32(sP + stage * OP_BYTES);
        const uint32_t qa = smem_u32(sQ + stage * OP_BYTES);
#pragma unroll
        for (int kg = 0; kg < WG_ROWS / UMMA_K; kg++) {
          // MN-major TF32 needs the 128-byte swizzle with 32-byte atoms (TMA: SWIZZLE_128B_ATOM_32B):
          // 32-float chunks along MN are WG_BOX_BYTES apart (LBO), 4-row groups along the
          // reduction are 512 B apart (SBO); one MMA (K = 8) spans two such groups
          const uint64_t db = make_smem_desc_sw128(qa + kg * 1024, WG_BOX_BYTES, 512, 1);
          for (int h = 0; h < n_halves; h++) {
            const uint64_t da = make_smem_desc_sw128(pa + h * 4 * WG_BOX_BYTES + kg * 1024, WG_BOX_BYTES, 512, 1);
            mma_tf32_ss(tmem_base + h * 256, da, db, idesc, (b != blk0) || (kg != 0));
          }
        }
        mma_commit(&empty[stage]);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
      mma_commit(done);
    }
  } else if (warp >= 4 && blk0 < blk1) {
    const int ew = warp & 3;
    mbar_wait(done, 0);
    tc_fence_after();
    for (int h = 0; h < n_halves; h++) {
      const int i = h * 128 + ew * 32 + lane;
      for (int c0 = 0; c0 < p.J; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + h * 256 + c0, v);
        tmem_ld_wait();
        if (i < p.I) {
          float* dst = p.D + (size_t)i * p.ldd + c0;
#pragma unroll
          for (int q = 0; q < 8; q++) {
            if (c0 + q * 4 < p.J) {
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + q * 4),
                           "f"(__uint_as_float(v[q * 4])), "f"(__uint_as_float(v[q * 4 + 1])),
                           "f"(__uint_as_float(v[q * 4 + 2])), "f"(__uint_as_float(v[q * 4 + 3]))
                           : "memory");
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace mlp
}  // namespace tb

// D[i][j] += sum_m P[m][i] * Q[m][j];  P [M][I] (pitch ldp), Q [M][J] (pitch ldq), D [I][J] (pitch ldd).
// I <= 256, J <= 256 and a multiple of 4; pitches multiples of 4 floats; 16-byte aligned bases.
extern "C" TRASE_API int trase_b200_wgrad_tf32(const float* P, int ldp, int I, const float* Q, int ldq, int J, int M,
                                               float* D, int ldd, void* stream) {
  using namespace tb;
  if (M <= 0 || I <= 0 || J <= 0 || I > 256 || J > 256 || (J & 3) || !P || !Q || !D) {
    set_error("trase_b200_wgrad_tf32: bad arguments (M=%d I=%d J=%d)", M, I, J);
    return 1;
  }
  if ((ldp & 3) || (ldq & 3) || (ldd & 3) || ((uintptr_t)P & 15) || ((uintptr_t)Q & 15) || ((uintptr_t)D & 15)) {
    set_error("trase_b200_wgrad_tf32: pointers must be 16-byte aligned and row pitches multiples of 4 floats");
    return 1;
  }
  CUtensorMap tP, tQ;
  if (!mlp::make_tmap_2d(&tP, P, M, I, ldp, mlp::WG_ROWS, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) ||
      !mlp::make_tmap_2d(&tQ, Q, M, J, ldq, mlp::WG_ROWS, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) {
    set_error("trase_b200_wgrad_tf32: cuTensorMapEncodeTiled failed");
    return 2;
  }
  static PerDeviceOnce attr_once;
  const int smem = 2 * mlp::WG_STAGES * 8 * mlp::WG_BOX_BYTES + 256 + 1024;
  {
    cudaError_t e = ensure_max_dyn_smem(attr_once, mlp::wgrad_tf32_kernel, smem);
    if (e != cudaSuccess) { set_error("trase_b200_wgrad_tf32: %s", cudaGetErrorString(e)); return 2; }
  }
  mlp::WgradParams wp;
  wp.M = M; wp.I = I; wp.J = J; wp.D = D; wp.ldd = ldd;
  const int nblk = (M + mlp::WG_ROWS - 1) / mlp::WG_ROWS;
  const int grid = nblk < mlp::num_sms() ? nblk : mlp::num_sms();
  launch_k(mlp::wgrad_tf32_kernel, grid, mlp::WG_THREADS, smem, reinterpret_cast<cudaStream_t>(stream), tP, tQ, wp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_wgrad_tf32: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// ===========================================================================
// Weight gradient from BF16 operands:  D[i, j] += sum_m P[m, i] * Q[m, j], P and Q row-major bf16 (the side copies
// the fused forward / backward kernels write, whose only reader this is), D fp32.
// Same structure as wgrad_tf32_kernel — operands MN-major straight from the row-major tensors, the m range split over
// persistent CTAs, the whole I x J result in TMEM — with half the bytes per element and K = 16 per MMA:
//   TMA box = 64 rows (m) x 64 columns, 128-byte swizzle: the canonical MN-major layout of 16-bit operands
//   (128-byte rows of 64 elements along MN, 8-row swizzle atoms along the reduction: SBO = 1024, next 64 MN elements
//   one box further: LBO = box size).
namespace tb {
namespace mlp {

constexpr int WH_ROWS = 64;                       // m rows per stage (4 MMAs of K = 16)
constexpr int WH_STAGES = 3;
constexpr int WH_BOX_BYTES = WH_ROWS * 128;       // one 64-row x 64-col bf16 box
constexpr int WH_OP_BYTES = 4 * WH_BOX_BYTES;     // up to 256 columns per operand = 32 KB

// One CTA's share of one product: row blocks [blk0, blk1) of 64 rows each, result added to p.D.
__device__ __forceinline__ void wgrad_bf16_body(const CUtensorMap* tmP, const CUtensorMap* tmQ, const WgradParams& p,
                                                const int blk0, const int blk1) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u)) __trap();
  uint8_t* sP = smem;
  uint8_t* sQ = smem + WH_STAGES * WH_OP_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * WH_STAGES * WH_OP_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + WH_STAGES;
  uint64_t* done = bars + 2 * WH_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_halves = (p.I + 127) / 128;
  const int i_boxes = n_halves * 2;                      // 64-column boxes of P
  const int j_boxes = (p.J + 63) / 64;
  const int jpad = ((p.J + 15) / 16) * 16;               // N of the MMA
  constexpr uint32_t TMEM_COLS = 512;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(tmP); prefetch_tmap(tmQ);
    for (int i = 0; i < WH_STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int b = blk0; b < blk1; b++) {
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_expect_tx(&full[stage], (i_boxes + j_boxes) * WH_BOX_BYTES);
        for (int c = 0; c < i_boxes; c++)
          tma_load_2d(sP + stage * WH_OP_BYTES + c * WH_BOX_BYTES, tmP, &full[stage], c * 64, b * WH_ROWS);
        for (int c = 0; c < j_boxes; c++)
          tma_load_2d(sQ + stage * WH_OP_BYTES + c * WH_BOX_BYTES, tmQ, &full[stage], c * 64, b * WH_ROWS);
        if (++stage == WH_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, jpad, 1, 1);
      uint32_t stage = 0, phase = 0;
      for (int b = blk0; b < blk1; b++) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t pa = smem_u32(sP + stage * WH_OP_BYTES);
        const uint32_t qa = smem_u32(sQ + stage * WH_OP_BYTES);
#pragma unroll
        for (int kg = 0; kg < WH_ROWS / 16; kg++) {
          // one MMA (K = 16) = two 8-row swizzle atoms, 2048 bytes into the box
          const uint64_t db = make_smem_desc_sw128(qa + kg * 2048, WH_BOX_BYTES, 1024, 2);
          for (int h = 0; h < n_halves; h++) {
            const uint64_t da = make_smem_desc_sw128(pa + h * 2 * WH_BOX_BYTES + kg * 2048, WH_BOX_BYTES, 1024, 2);
            mma_bf16_ss(tmem_base + h * 256, da, db, idesc, (b != blk0) || (kg != 0));
          }
        }
        mma_commit(&empty[stage]);
        if (++stage == WH_STAGES) { stage = 0; phase ^= 1; }
      }
      mma_commit(done);
    }
  } else if (warp >= 4 && blk0 < blk1) {
    const int ew = warp & 3;
    mbar_wait(done, 0);
    tc_fence_after();
    for (int h = 0; h < n_halves; h++) {
      const int i = h * 128 + ew * 32 + lane;
      for (int c0 = 0; c0 < p.J; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(ew * 32) << 16) + h * 256 + c0, v);
        tmem_ld_wait();
        if (i < p.I) {
          float* dst = p.D + (size_t)i * p.ldd + c0;
#pragma unroll
          for (int q = 0; q < 8; q++) {
            if (c0 + q * 4 < p.J) {
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + q * 4),
                           "f"(__uint_as_float(v[q * 4])), "f"(__uint_as_float(v[q * 4 + 1])),
                           "f"(__uint_as_float(v[q * 4 + 2])), "f"(__uint_as_float(v[q * 4 + 3]))
                           : "memory");
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_bf16_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ, const WgradParams p) {
  pdl_wait();
  const int nblk = (p.M + WH_ROWS - 1) / WH_ROWS;
  const int per = (nblk + gridDim.x - 1) / gridDim.x;
  const int blk0 = blockIdx.x * per;
  wgrad_bf16_body(&tmP, &tmQ, p, blk0, min(nblk, blk0 + per));
}

// All weight-gradient products of one backward pass in ONE launch: the CTAs are dealt out to the products in
// proportion to the bytes each has to stream, every CTA reduces its row slice of ONE product in TMEM and adds the
// I x J result to D once.  Against one launch per product (every one of which ends with all 148 CTAs adding a full
// I x J tile): a tenth of the reduction traffic, one drain instead of ten, no launch gaps.
constexpr int WGM_MAX = 12;
struct WgradMultiParams {
  CUtensorMap mP[WGM_MAX], mQ[WGM_MAX];
  float* D[WGM_MAX];
  int I[WGM_MAX], J[WGM_MAX], ldd[WGM_MAX];
  int cta_first[WGM_MAX + 1];     // CTAs [cta_first[j], cta_first[j + 1]) work on product j
  int M, njobs;
};
__global__ void __launch_bounds__(WG_THREADS, 1) wgrad_bf16_multi_kernel(const __grid_constant__ WgradMultiParams q) {
  pdl_wait();
  int job = 0;
  while (job + 1 < q.njobs && (int)blockIdx.x >= q.cta_first[job + 1]) job++;
  const int ci = blockIdx.x - q.cta_first[job], nc = q.cta_first[job + 1] - q.cta_first[job];
  WgradParams p;
  p.M = q.M; p.I = q.I[job]; p.J = q.J[job]; p.D = q.D[job]; p.ldd = q.ldd[job];
  const int nblk = (p.M + WH_ROWS - 1) / WH_ROWS;
  const int per = (nblk + nc - 1) / nc;
  const int blk0 = ci * per;
  wgrad_bf16_body(&q.mP[job], &q.mQ[job], p, blk0, min(nblk, blk0 + per));
}

typedef CUresult (*EncodeTiledFn16)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static bool make_tmap_2d_bf16(CUtensorMap* m, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                              uint32_t box_rows, uint32_t box_cols) {
  static EncodeTiledFn16 fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) != cudaSuccess ||
        r != cudaDriverEntryPointSuccess)
      return false;
    fn = reinterpret_cast<EncodeTiledFn16>(q);
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// dst[r][c] = bf16(src[r][c]) for c < cols; columns [cols, ld_dst) of dst are zero-filled
__global__ void __launch_bounds__(256) f32_to_bf16_kernel(int rows, int cols, const float* __restrict__ src, int ld_src,
                                                          __nv_bfloat16* __restrict__ dst, int ld_dst) {
  pdl_wait();
  const size_t total = (size_t)rows * ld_dst;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / ld_dst;
    const int c = (int)(e - r * ld_dst);
    dst[e] = __float2bfloat16_rn(c < cols ? src[r * ld_src + c] : 0.f);
  }
}

}  // namespace mlp
}  // namespace tb

// D[i][j] += sum_m P[m][i] * Q[m][j];  P [M][I], Q [M][J] row-major BF16 (pitches ldp / ldq in elements, multiples
// of 8; 16-byte aligned bases), D [I][J] fp32 (pitch ldd, multiple of 4).  I <= 256, J <= 256, J a multiple of 4.
extern "C" TRASE_API int trase_b200_wgrad_bf16(const void* P, int ldp, int I, const void* Q, int ldq, int J, int M,
                                               float* D, int ldd, void* stream) {
  using namespace tb;
  if (M <= 0 || I <= 0 || J <= 0 || I > 256 || J > 256 || (J & 3) || !P || !Q || !D) {
    set_error("trase_b200_wgrad_bf16: bad arguments (M=%d I=%d J=%d)", M, I, J);
    return 1;
  }
  if ((ldp & 7) || (ldq & 7) || (ldd & 3) || ((uintptr_t)P & 15) || ((uintptr_t)Q & 15) || ((uintptr_t)D & 15)) {
    set_error("trase_b200_wgrad_bf16: pointers must be 16-byte aligned, bf16 pitches multiples of 8, ldd of 4");
    return 1;
  }
  CUtensorMap tP, tQ;
  if (!mlp::make_tmap_2d_bf16(&tP, P, M, I, ldp, mlp::WH_ROWS, 64) ||
      !mlp::make_tmap_2d_bf16(&tQ, Q, M, J, ldq, mlp::WH_ROWS, 64)) {
    set_error("trase_b200_wgrad_bf16: cuTensorMapEncodeTiled failed");
    return 2;
  }
  static PerDeviceOnce attr_once;
  const int smem = 2 * mlp::WH_STAGES * mlp::WH_OP_BYTES + 256 + 1024;
  {
    cudaError_t e = ensure_max_dyn_smem(attr_once, mlp::wgrad_bf16_kernel, smem);
    if (e != cudaSuccess) { set_error("trase_b200_wgrad_bf16: %s", cudaGetErrorString(e)); return 2; }
  }
  mlp::WgradParams wp;
  wp.M = M; wp.I = I; wp.J = J; wp.D = D; wp.ldd = ldd;
  const int nblk = (M + mlp::WH_ROWS - 1) / mlp::WH_ROWS;
  const int grid = nblk < mlp::num_sms() ? nblk : mlp::num_sms();
  launch_k(mlp::wgrad_bf16_kernel, grid, mlp::WG_THREADS, smem, reinterpret_cast<cudaStream_t>(stream), tP, tQ, wp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_wgrad_bf16: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// njobs products D_j[i][k] += sum_m P_j[m][i] * Q_j[m][k] over the same M rows in one launch (arguments per product
// as for trase_b200_wgrad_bf16; at most 12 products).
extern "C" TRASE_API int trase_b200_wgrad_bf16_multi(int njobs, const void* const* P, const int* ldp, const int* I,
                                                     const void* const* Q, const int* ldq, const int* J, int M,
                                                     float* const* D, const int* ldd, void* stream) {
  using namespace tb;
  if (njobs <= 0 || M <= 0) return 0;
  if (njobs > mlp::WGM_MAX || !P || !ldp || !I || !Q || !ldq || !J || !D || !ldd) {
    set_error("trase_b200_wgrad_bf16_multi: bad arguments (njobs=%d, at most %d)", njobs, mlp::WGM_MAX);
    return 1;
  }
  mlp::WgradMultiParams q;
  int cost[mlp::WGM_MAX], total_cost = 0;
  for (int j = 0; j < njobs; j++) {
    if (I[j] <= 0 || J[j] <= 0 || I[j] > 256 || J[j] > 256 || (J[j] & 3) || !P[j] || !Q[j] || !D[j] || (ldp[j] & 7) ||
        (ldq[j] & 7) || (ldd[j] & 3) || ((uintptr_t)P[j] & 15) || ((uintptr_t)Q[j] & 15) || ((uintptr_t)D[j] & 15)) {
      set_error("trase_b200_wgrad_bf16_multi: product %d: bad shape, pitch or alignment (I=%d J=%d)", j, I[j], J[j]);
      return 1;
    }
    if (!mlp::make_tmap_2d_bf16(&q.mP[j], P[j], M, I[j], ldp[j], mlp::WH_ROWS, 64) ||
        !mlp::make_tmap_2d_bf16(&q.mQ[j], Q[j], M, J[j], ldq[j], mlp::WH_ROWS, 64)) {
      set_error("trase_b200_wgrad_bf16_multi: cuTensorMapEncodeTiled failed");
      return 2;
    }
    q.D[j] = D[j]; q.I[j] = I[j]; q.J[j] = J[j]; q.ldd[j] = ldd[j];
    cost[j] = ((I[j] + 127) / 128) * 2 + (J[j] + 63) / 64;       // 64-column boxes streamed per row block
    total_cost += cost[j];
  }
  q.M = M; q.njobs = njobs;
  // CTAs per product in proportion to its bytes, at least one each, the grid exactly the SMs available
  const int sms = mlp::num_sms();
  if (sms < njobs) { set_error("trase_b200_wgrad_bf16_multi: fewer SMs (%d) than products (%d)", sms, njobs); return 1; }
  int n[mlp::WGM_MAX], used = 0;
  for (int j = 0; j < njobs; j++) {
    n[j] = (int)((long long)sms * cost[j] / total_cost);
    if (n[j] < 1) n[j] = 1;
    used += n[j];
  }
  while (used != sms) {
    // hand the spare CTAs to (take the excess from) the product with the most (least) work per CTA
    int best = -1;
    for (int j = 0; j < njobs; j++) {
      if (used > sms && n[j] <= 1) continue;
      if (best < 0) { best = j; continue; }
      const long long a = (long long)cost[j] * n[best], b = (long long)cost[best] * n[j];     // cost/n compared
      if (used < sms ? a > b : a < b) best = j;
    }
    n[best] += used < sms ? 1 : -1;
    used += used < sms ? 1 : -1;
  }
  q.cta_first[0] = 0;
  for (int j = 0; j < njobs; j++) q.cta_first[j + 1] = q.cta_first[j] + n[j];
  static PerDeviceOnce attr_once;
  const int smem = 2 * mlp::WH_STAGES * mlp::WH_OP_BYTES + 256 + 1024;
  {
    cudaError_t e = ensure_max_dyn_smem(attr_once, mlp::wgrad_bf16_multi_kernel, smem);
    if (e != cudaSuccess) { set_error("trase_b200_wgrad_bf16_multi: %s", cudaGetErrorString(e)); return 2; }
  }
  launch_k(mlp::wgrad_bf16_multi_kernel, sms, mlp::WG_THREADS, smem, reinterpret_cast<cudaStream_t>(stream), q);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_wgrad_bf16_multi: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// dst [rows][ld_dst] bf16 <- src [rows][ld_src] fp32 (first `cols` columns; the rest of each dst row is zeroed).
// Used for the two small wgrad operands no fused kernel produces: the positional encoding and the packed upstream
// gradient of the heads.
extern "C" TRASE_API int trase_b200_f32_to_bf16(int rows, int cols, const float* src, int ld_src, void* dst,
                                                int ld_dst, void* stream) {
  using namespace tb;
  if (rows <= 0) return 0;
  if (!src || !dst || cols <= 0 || ld_src < cols || ld_dst < cols) {
    set_error("trase_b200_f32_to_bf16: bad arguments");
    return 1;
  }
  const size_t total = (size_t)rows * ld_dst;
  const int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  launch_k(mlp::f32_to_bf16_kernel, grid, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      rows, cols, src, ld_src, reinterpret_cast<__nv_bfloat16*>(dst), ld_dst);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_f32_to_bf16: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// ===========================================================================
// Positional encoding of (xyz, t) written straight into the first 84 columns of
// the [M][ld] buffer that also receives layer 4's output (the skip-connection
// input of layer 5), so neither torch.cat of the reference (utils/time_utils.py:110-117)
// costs a copy.  Column order = the reference's Embedder (time_utils.py:27-57):
//   [x(3), sin(2^0 x)(3), cos(2^0 x)(3), ..., cos(2^9 x)(3) | t, sin(2^0 t), cos(2^0 t), ..., cos(2^(Lt-1) t)]
namespace tb {
namespace mlp {

__device__ __forceinline__ float rna_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
// One block per 64 rows.  The sines and cosines of the tile (see the two loops below for how few transcendental
// evaluations that takes) go into a shared-memory copy of the tile laid out like the output rows; the tile then leaves with 16-byte coalesced stores — fp32 rounded to TF32, and the bf16 copy — when the
// output rows are contiguous (the native DeformNetwork's buffers), column by column otherwise.
constexpr int EMB_MAX_COLS = 128;
constexpr int EMB_ROWS = 64;
__global__ void __launch_bounds__(256) deform_embed_kernel(int M, const float* __restrict__ xyz,
                                                           const float* __restrict__ t, int t_stride, int Lx, int Lt,
                                                           float* __restrict__ out, int ld,
                                                           __nv_bfloat16* __restrict__ out16, int ld16) {
  pdl_wait();
  __shared__ __align__(16) float s_tile[EMB_ROWS * EMB_MAX_COLS];
  __shared__ float s_x[EMB_ROWS * 3];
  __shared__ float s_t[EMB_MAX_COLS];
  const int nx = 3 + 6 * Lx;           // columns of the xyz part
  const int ncol = nx + 1 + 2 * Lt;
  const int m0 = blockIdx.x * EMB_ROWS;
  const int rows = min(EMB_ROWS, M - m0);
  for (int i = threadIdx.x; i < rows * 3; i += blockDim.x) s_x[i] = xyz[(size_t)m0 * 3 + i];
  __syncthreads();
  // xyz part: item = (row, comp).  sin / cos of 2^k x for k = 0 .. Lx-1 by the double-angle recurrence in fp64 from one
  // fp64 sincos of x (2^k x is exact, so these ARE the reference's arguments; the recurrence doubles a 1e-16 error per
  // step, far below the fp32 rounding of the result): one transcendental evaluation per (row, component) instead of Lx
  for (int e = threadIdx.x; e < rows * 3; e += blockDim.x) {
    const int r = e / 3, comp = e - 3 * r;
    double sv, cv;
    sincos((double)s_x[e], &sv, &cv);
    float* dst = s_tile + r * ncol + 3 + comp;
    for (int k = 0; k < Lx; k++) {
      dst[6 * k] = (float)sv;
      dst[6 * k + 3] = (float)cv;
      const double s2 = 2.0 * sv * cv;
      cv = (cv - sv) * (cv + sv);
      sv = s2;
    }
  }
  for (int i = threadIdx.x; i < rows * 3; i += blockDim.x) s_tile[(i / 3) * ncol + (i % 3)] = s_x[i];
  // time part: one time stamp for every row (the usual case: one view = one t) is evaluated once per block and copied
  // into the rows; per-row time stamps cost Lt + 1 evaluations per row
  if (t_stride == 0) {
    for (int k = threadIdx.x; k <= Lt; k += blockDim.x) {
      const float tv = t[0];
      if (k == 0) {
        s_t[0] = tv;
      } else {
        float sv, cv;
        sincosf(tv * exp2f((float)(k - 1)), &sv, &cv);
        s_t[2 * k - 1] = sv;
        s_t[2 * k] = cv;
      }
    }
    __syncthreads();
    const int nt = 2 * Lt + 1;
    for (int e = threadIdx.x; e < rows * nt; e += blockDim.x) {
      const int r = e / nt, c = e - r * nt;
      s_tile[r * ncol + nx + c] = s_t[c];
    }
  } else {
    const int per_row_t = Lt + 1;
    for (int e = threadIdx.x; e < rows * per_row_t; e += blockDim.x) {
      const int r = e / per_row_t, k = e - r * per_row_t - 1;
      const float tv = t[(size_t)(m0 + r) * t_stride];
      if (k < 0) {
        s_tile[r * ncol + nx] = tv;
      } else {
        float sv, cv;
        sincosf(tv * exp2f((float)k), &sv, &cv);
        s_tile[r * ncol + nx + 1 + 2 * k] = sv;
        s_tile[r * ncol + nx + 2 + 2 * k] = cv;
      }
    }
  }
  __syncthreads();
  float* o = out + (size_t)m0 * ld;
  if (ld == ncol && (ncol & 3) == 0 && ((uintptr_t)out & 15) == 0) {
    // the tile is one contiguous piece of the output
    const float4* src = reinterpret_cast<const float4*>(s_tile);
    float4* dst = reinterpret_cast<float4*>(o);
    for (int i = threadIdx.x; i < rows * ncol / 4; i += blockDim.x) {
      const float4 v = src[i];
      dst[i] = make_float4(rna_tf32(v.x), rna_tf32(v.y), rna_tf32(v.z), rna_tf32(v.w));
    }
  } else {
    for (int i = threadIdx.x; i < rows * ncol; i += blockDim.x) {
      const int r = i / ncol, c = i - r * ncol;
      o[(size_t)r * ld + c] = rna_tf32(s_tile[i]);
    }
  }
  if (out16) {         // bf16 copy for the weight-gradient products (columns up to ld16 zero-filled), 8 values per store
    uint4* dst = reinterpret_cast<uint4*>(out16 + (size_t)m0 * ld16);
    const int per8 = ld16 >> 3;
    for (int i = threadIdx.x; i < rows * per8; i += blockDim.x) {
      const int r = i / per8, c0 = (i - r * per8) << 3;
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int c = c0 + 2 * q;
        const __nv_bfloat162 h2 = __floats2bfloat162_rn(c < ncol ? s_tile[r * ncol + c] : 0.f,
                                                        c + 1 < ncol ? s_tile[r * ncol + c + 1] : 0.f);
        w[q] = *reinterpret_cast<const uint32_t*>(&h2);
      }
      dst[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

}  // namespace mlp
}  // namespace tb

// t_stride = 0 broadcasts a single time value to all rows (the usual case: one time stamp per view).
static int deform_embed_impl(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt, float* out, int ld,
                             void* out16, int ld16, void* stream);
extern "C" TRASE_API int trase_b200_deform_embed(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt,
                                                 float* out, int ld, void* stream) {
  return deform_embed_impl(M, xyz, t, t_stride, Lx, Lt, out, ld, nullptr, 0, stream);
}
// same, also writing a bf16 copy out16 [M][ld16] (16-byte aligned, ld16 >= the column count, a multiple of 8; padding
// zero-filled):
// the operand of the bf16 weight-gradient products of the two layers that read the embedding
extern "C" TRASE_API int trase_b200_deform_embed_bf16(int M, const float* xyz, const float* t, int t_stride, int Lx,
                                                      int Lt, float* out, int ld, void* out16, int ld16, void* stream) {
  if (M > 0 && (!out16 || ((uintptr_t)out16 & 15) || (ld16 & 7) || ld16 < 3 + 6 * Lx + 1 + 2 * Lt || ld16 > tb::mlp::EMB_MAX_COLS)) {
    tb::set_error("trase_b200_deform_embed_bf16: bad arguments");
    return 1;
  }
  return deform_embed_impl(M, xyz, t, t_stride, Lx, Lt, out, ld, out16, ld16, stream);
}
static int deform_embed_impl(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt, float* out, int ld,
                             void* out16, int ld16, void* stream) {
  using namespace tb;
  if (M <= 0) return 0;
  if (!xyz || !t || !out || Lx < 0 || Lt < 0 || ld < 3 + 6 * Lx + 1 + 2 * Lt ||
      3 + 6 * Lx + 1 + 2 * Lt > mlp::EMB_MAX_COLS) {
    set_error("trase_b200_deform_embed: bad arguments");
    return 1;
  }
  const int blocks = (M + mlp::EMB_ROWS - 1) / mlp::EMB_ROWS;
  launch_k(mlp::deform_embed_kernel, blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream), 
      M, xyz, t, t_stride, Lx, Lt, out, ld, reinterpret_cast<__nv_bfloat16*>(out16), ld16);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_deform_embed: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}
