// TF32 tensor-core GEMMs for the DeformNetwork MLP (reference: utils/time_utils.py:106-131,
// eleven nn.Linear layers evaluated by cuBLAS SGEMM on CUDA cores).
//
//   gemm_tn :  D[M,N] = epilogue( A[M,K] · B[N,K]^T )      A, B row-major (K contiguous)
//              epilogue: (+ bias[n]) -> (ReLU) -> (· [Z[m,n] > 0])
//     forward  layer : A = activations, B = weight [out,in], bias, ReLU
//     backward dgrad : A = dZ [M,out], B = weight^T [in,out], mask Z = the layer input (post-ReLU)
//
// Blackwell structure: persistent CTAs (one per SM), TMA 128-byte-swizzled tiles
// into a 4-stage shared-memory ring, a single elected thread issues
// tcgen05.mma.kind::tf32 (128 x BN x 8) with the fp32 accumulator in TMEM
// (double-buffered: 2 x BN columns), four epilogue warps drain TMEM with
// tcgen05.ld, apply the epilogue and write 128x32 slabs back with TMA stores,
// overlapping the next tile's MMAs.  K and M tails are handled by TMA
// zero-fill / clipping, so no operand is ever padded in memory.
#include "common.cuh"
#include "ptx.cuh"

namespace tb {
namespace mlp {

using namespace ptx;

constexpr int BM = 128;          // rows per tile (= TMEM lanes)
constexpr int BK = 32;           // tf32 elements per k-block = one 128-byte swizzle row
constexpr int UMMA_K = 8;        // tf32 elements per tcgen05.mma
constexpr int STAGES = 4;
constexpr int SLAB = 32;         // output columns per epilogue slab
constexpr int GEMM_THREADS = 256;

struct GemmEpilogue {
  const float* bias;  // [N] or null
  int relu;
  int use_mask;       // multiply by (Z > 0), Z read through tmZ
  float* colsum;      // [N] or null: += column sums of the written values (bias gradient)
  int round_tf32;     // round outputs to TF32 (nearest) so the next GEMM's operand truncation is exact
  int M, N, K;
};

template <int BN>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 4;       // 16 KB
  static constexpr int B_BYTES = BN * BK * 4;       // 32 KB (BN = 256)
  static constexpr int OUT_BYTES = BM * SLAB * 4;   // 16 KB
  static constexpr int OFF_A = 0;
  static constexpr int OFF_B = OFF_A + STAGES * A_BYTES;
  static constexpr int OFF_OUT = OFF_B + STAGES * B_BYTES;
  static constexpr int OFF_BIAS = OFF_OUT + 2 * OUT_BYTES;
  static constexpr int OFF_BAR = OFF_BIAS + (BN < 32 ? 32 : BN) * 4;
  static constexpr int TOTAL = OFF_BAR + 256 + 1024;  // barriers + slack for 1024-byte alignment
};

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tn_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmZ,
                    const GemmEpilogue ep) {
  using L = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem + L::OFF_A;
  uint8_t* sB = smem + L::OFF_B;
  uint8_t* sOut = smem + L::OFF_OUT;
  float* sBias = reinterpret_cast<float*>(smem + L::OFF_BIAS);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::OFF_BAR);
  uint64_t* full = bars;                   // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;         // [STAGES]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * STAGES; // [2]       MMA -> epilogue
  uint64_t* tmem_empty = tmem_full + 2;    // [2]       epilogue -> MMA
  uint64_t* zbar = tmem_empty + 2;         // [2]       mask tile landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(zbar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = (ep.M + BM - 1) / BM;
  const int num_kb = (ep.K + BK - 1) / BK;
  constexpr uint32_t TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;  // power of two >= 32 (BN = 16 -> 32, 256 -> 512)

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmB); prefetch_tmap(&tmD);
    if (ep.use_mask) prefetch_tmap(&tmZ);
    for (int i = 0; i < STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); mbar_init(&zbar[i], 1); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  for (int i = threadIdx.x; i < (BN < 32 ? 32 : BN); i += GEMM_THREADS) sBias[i] = (ep.bias && i < ep.N) ? ep.bias[i] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < num_kb; kb++) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_expect_tx(&full[stage], L::A_BYTES + L::B_BYTES);
          tma_load_2d(sA + stage * L::A_BYTES, &tmA, &full[stage], kb * BK, tile * BM);
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
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
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
    // ===== epilogue: 4 warps, warp (w & 3) owns TMEM lanes 32*(w&3) .. +31 =====
    const int ew = warp & 3;
    const int row = ew * 32 + lane;        // row inside the tile
    const int et = threadIdx.x - 128;      // 0..127
    uint32_t acc = 0, acc_phase = 0;
    uint32_t zphase[2] = {0, 0};
    uint32_t slab_counter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      constexpr int NSLAB = (BN + SLAB - 1) / SLAB;
      for (int s = 0; s < NSLAB; s++) {
        if (s * SLAB >= ep.N) break;
        const uint32_t buf = slab_counter & 1;
        uint8_t* out = sOut + buf * L::OUT_BYTES;
        // the TMA store that last read this buffer (two slabs ago) must have finished reading it
        if (et == 0) tma_store_wait_read<1>();
        named_bar_sync(1, 128);
        if (ep.use_mask) {
          if (et == 0) {
            mbar_expect_tx(&zbar[buf], L::OUT_BYTES);
            tma_load_2d(out, &tmZ, &zbar[buf], s * SLAB, tile * BM);
          }
          mbar_wait(&zbar[buf], zphase[buf]);
          zphase[buf] ^= 1;
        }
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
        tmem_ld_wait();
        // 128-byte swizzled staging tile: element (row, 16-byte chunk c) at row*128 + ((c ^ (row & 7)) << 4)
#pragma unroll
        for (int c = 0; c < 8; c++) {
          float4 o;
          float* of = reinterpret_cast<float*>(&o);
          float4* dst = reinterpret_cast<float4*>(out + row * 128 + ((c ^ (row & 7)) << 4));
          float4 z = make_float4(1.f, 1.f, 1.f, 1.f);
          if (ep.use_mask) z = *dst;
          const float* zf = reinterpret_cast<const float*>(&z);
#pragma unroll
          for (int j = 0; j < 4; j++) {
            float x = __uint_as_float(v[c * 4 + j]) + sBias[s * SLAB + c * 4 + j];
            if (ep.relu) x = fmaxf(x, 0.f);
            if (ep.use_mask) x = (zf[j] > 0.f) ? x : 0.f;
            if (ep.round_tf32) {
              uint32_t u;
              asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
              x = __uint_as_float(u);
            }
            of[j] = x;
          }
          *dst = o;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (et == 0) {
          tma_store_2d(&tmD, out, s * SLAB, tile * BM);
          tma_store_commit();
        }
        if (ep.colsum) {
          // bias gradient: thread (g, c) sums column c over rows 32g..32g+31 of the staged slab
          const int c = et & 31, g = et >> 5;
          float acc_c = 0.f;
#pragma unroll 8
          for (int r = g * 32; r < g * 32 + 32; r++)
            acc_c += *reinterpret_cast<const float*>(out + r * 128 + (((c >> 2) ^ (r & 7)) << 4) + ((c & 3) << 2));
          if (s * SLAB + c < ep.N) atomicAdd(ep.colsum + s * SLAB + c, acc_c);
        }
        slab_counter++;
      }
      // all TMEM reads of this accumulator are done: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (et == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
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
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <int BN>
static cudaError_t launch_gemm_tn(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD,
                                  const CUtensorMap& tZ, const GemmEpilogue& ep, cudaStream_t s) {
  static bool attr_set = false;
  const int smem = GemmSmem<BN>::TOTAL;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tn_tf32_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int tiles = (ep.M + BM - 1) / BM;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  gemm_tn_tf32_kernel<BN><<<grid, GEMM_THREADS, smem, s>>>(tA, tB, tD, tZ, ep);
  return cudaGetLastError();
}

}  // namespace mlp
}  // namespace tb

using namespace tb;

// D[M,N] = epilogue(A[M,K] · B[N,K]^T); lda/ldb/ldd/ldz are row pitches in floats (multiples of 4),
// all base pointers 16-byte aligned.  N <= 256.  colsum (optional) accumulates the column sums of D.
extern "C" TRASE_API int trase_b200_gemm_tf32(const float* A, int lda, const float* B, int ldb, float* D, int ldd,
                                              int M, int N, int K, const float* bias, int relu, const float* Z,
                                              int ldz, float* colsum, int round_tf32, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0 || N > 256 || !A || !B || !D) {
    set_error("trase_b200_gemm_tf32: bad arguments (M=%d N=%d K=%d)", M, N, K);
    return 1;
  }
  if ((lda & 3) || (ldb & 3) || (ldd & 3) || (Z && (ldz & 3)) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) ||
      ((uintptr_t)D & 15) || ((uintptr_t)Z & 15)) {
    set_error("trase_b200_gemm_tf32: pointers must be 16-byte aligned and row pitches multiples of 4 floats");
    return 1;
  }
  const bool small = N <= 16;
  CUtensorMap tA, tB, tD, tZ;
  bool ok = mlp::make_tmap_2d(&tA, A, M, K, lda, mlp::BM, mlp::BK) &&
            mlp::make_tmap_2d(&tB, B, N, K, ldb, small ? 16 : 256, mlp::BK) &&
            mlp::make_tmap_2d(&tD, D, M, N, ldd, mlp::BM, mlp::SLAB);
  if (ok) ok = Z ? mlp::make_tmap_2d(&tZ, Z, M, N, ldz, mlp::BM, mlp::SLAB) : mlp::make_tmap_2d(&tZ, D, M, N, ldd, mlp::BM, mlp::SLAB);
  if (!ok) {
    set_error("trase_b200_gemm_tf32: cuTensorMapEncodeTiled failed");
    return 2;
  }
  mlp::GemmEpilogue ep;
  ep.bias = bias; ep.relu = relu; ep.use_mask = Z != nullptr; ep.colsum = colsum; ep.round_tf32 = round_tf32; ep.M = M; ep.N = N; ep.K = K;
  cudaError_t e = small ? mlp::launch_gemm_tn<16>(tA, tB, tD, tZ, ep, reinterpret_cast<cudaStream_t>(stream))
                        : mlp::launch_gemm_tn<256>(tA, tB, tD, tZ, ep, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    set_error("trase_b200_gemm_tf32: CUDA error: %s", cudaGetErrorString(e));
    return 2;
  }
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

__global__ void __launch_bounds__(GEMM_THREADS, 1)
wgrad_tf32_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmQ,
                  const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
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
          tma_load_2d(sP + stage * OP_BYTES + c * WG_BOX_BYTES, &tmP, &full[stage], c * 32, b * WG_ROWS);
        for (int c = 0; c < j_boxes; c++)
          tma_load_2d(sQ + stage * OP_BYTES + c * WG_BOX_BYTES, &tmQ, &full[stage], c * 32, b * WG_ROWS);
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
  static bool attr_set = false;
  const int smem = 2 * mlp::WG_STAGES * 8 * mlp::WG_BOX_BYTES + 256 + 1024;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp::wgrad_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("trase_b200_wgrad_tf32: %s", cudaGetErrorString(e)); return 2; }
    attr_set = true;
  }
  mlp::WgradParams wp;
  wp.M = M; wp.I = I; wp.J = J; wp.D = D; wp.ldd = ldd;
  const int nblk = (M + mlp::WG_ROWS - 1) / mlp::WG_ROWS;
  const int grid = nblk < mlp::num_sms() ? nblk : mlp::num_sms();
  mlp::wgrad_tf32_kernel<<<grid, mlp::GEMM_THREADS, smem, reinterpret_cast<cudaStream_t>(stream)>>>(tP, tQ, wp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_wgrad_tf32: CUDA error: %s", cudaGetErrorString(e)); return 2; }
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
__global__ void __launch_bounds__(256) deform_embed_kernel(int M, const float* __restrict__ xyz,
                                                           const float* __restrict__ t, int t_stride, int Lx, int Lt,
                                                           float* __restrict__ out, int ld) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  float* o = out + (size_t)m * ld;
  const float x[3] = {xyz[3 * m], xyz[3 * m + 1], xyz[3 * m + 2]};
  int c = 0;
  o[c++] = rna_tf32(x[0]); o[c++] = rna_tf32(x[1]); o[c++] = rna_tf32(x[2]);
  float f = 1.f;
  for (int k = 0; k < Lx; k++, f *= 2.f) {
    float s0, c0, s1, c1, s2, c2;
    sincosf(x[0] * f, &s0, &c0);
    sincosf(x[1] * f, &s1, &c1);
    sincosf(x[2] * f, &s2, &c2);
    o[c++] = rna_tf32(s0); o[c++] = rna_tf32(s1); o[c++] = rna_tf32(s2);
    o[c++] = rna_tf32(c0); o[c++] = rna_tf32(c1); o[c++] = rna_tf32(c2);
  }
  const float tv = t[(size_t)m * t_stride];
  o[c++] = rna_tf32(tv);
  f = 1.f;
  for (int k = 0; k < Lt; k++, f *= 2.f) {
    float s, cs;
    sincosf(tv * f, &s, &cs);
    o[c++] = rna_tf32(s); o[c++] = rna_tf32(cs);
  }
}

}  // namespace mlp
}  // namespace tb

// t_stride = 0 broadcasts a single time value to all rows (the usual case: one time stamp per view).
extern "C" TRASE_API int trase_b200_deform_embed(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt,
                                                 float* out, int ld, void* stream) {
  using namespace tb;
  if (M <= 0) return 0;
  if (!xyz || !t || !out || Lx < 0 || Lt < 0 || ld < 3 + 6 * Lx + 1 + 2 * Lt) {
    set_error("trase_b200_deform_embed: bad arguments");
    return 1;
  }
  mlp::deform_embed_kernel<<<(M + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(M, xyz, t, t_stride, Lx,
                                                                                              Lt, out, ld);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_deform_embed: CUDA error: %s", cudaGetErrorString(e)); return 2; }
  return 0;
}
