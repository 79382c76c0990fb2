// DeformNetwork forward as ONE persistent kernel (reference: utils/time_utils.py:106-131 — eight
// nn.Linear + ReLU with a skip concatenation, then three linear heads; the reference launches one
// cuBLAS SGEMM and two elementwise kernels per layer).
//
// The layer-by-layer path (mlp_gemm.cu) is bound by HBM: every layer reads the previous activations
// (M x 256 fp32) and writes its own.  Here each CTA carries a 128-row tile through all layers: the
// activations stay in shared memory as the next layer's A operand and are written to HBM only once
// (the backward pass needs them); nothing is read back.
//
// CTA pairs (cluster of 2, tcgen05 cta_group::2): one MMA spans both SMs — M = 256 (each CTA's own
// 128-row tile), N = 256 with each CTA holding only ITS half of the weight k-block.  With fp32 (TF32)
// operands a single-CTA 128x256x8 MMA reads 12 KB of shared memory, which together with the TMA writes
// and the epilogue saturates the 128 B/clk port; the pair halves the B traffic and the weight bytes
// each SM pulls from L2, and lets a 6-deep ring of 16 KB stages cover the L2 latency (a single-CTA
// version with 32 KB stages spent 30 % of its time waiting for weights).  Per pair, tile and layer:
//   * both CTAs' producers TMA their half of every pre-packed, TF32-rounded weight k-block
//     ([k-block][256 rows][32]) into their ring; the follower relays "landed" to the leader;
//   * the leader's MMA thread issues tcgen05.mma.cta_group::2.kind::tf32 256x256x8 into one of two
//     TMEM accumulators (each CTA's TMEM receives its own 128 rows) and commits to both CTAs' barriers;
//   * eight epilogue warps per CTA drain the accumulator (tcgen05.ld), add the bias, apply ReLU, round to
//     TF32, record the ReLU bit mask and write the result IN PLACE over the previous activations, already
//     in the 128-byte-swizzled K-major layout the next layer's MMA reads;
//   * each 32-column chunk is released to the MMA thread through its own mbarrier (arrivals from both
//     CTAs), so the next layer's MMAs start on chunk 0 while chunks 1..7 are still being drained;
//   * warp 1 of each CTA also TMA-stores every finished chunk to the activation tensor in HBM (direct
//     16-byte stores from the epilogue registers were measured 50 % slower).
// The positional encoding (84 columns -> three zero-padded chunks) is the A operand of layer 0 and of
// the first k-blocks of the layer after the skip; its chunks travel through the same ring.  The heads
// run as a last pseudo-layer with N = 32 whose first 16 output columns go straight to global memory.
#include "common.cuh"
#include "ptx.cuh"
#include <cuda_bf16.h>

namespace tb {
namespace mlp {

using namespace ptx;

bool make_tmap_2d(CUtensorMap* m, const float* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                  uint32_t box_cols, CUtensorMapSwizzle swz);
int num_sms();
extern int g_reserve_sms;

constexpr int FU_BM = 128;
constexpr int FU_W = 256;                 // hidden width
constexpr int FU_CHUNK = FU_BM * 128;     // 128 rows x 32 fp32 = 16 KB: one A chunk, or half a weight k-block
constexpr int FU_HCH = FU_W / 32;         // 8 chunks of hidden activations
constexpr int FU_ECH = 3;                 // up to 96 embedding columns
// ring stages of 16 KB (this CTA's half of a weight k-block, or a chunk of the first operand);
// the backward chain gives one stage up for its bias-gradient accumulators
// A16 (bf16 side output, see below) gives one more stage up: 8 epilogue warps x 2 KB of store staging
template <bool BWD, bool A16> struct FuCfg { static constexpr int STAGES = (BWD ? 5 : 6) - (A16 ? 1 : 0); };
// TB_FU_PROF: the stall-cycle counters read by tools/fused_mlp_stalls.py (build with
// TRASE_B200_EXTRA_NVCC_FLAGS=-DTB_FU_PROF=1).  Off in the shipped library: the clock64() pairs around every wait of the
// MMA-issuing thread cost 4-5 % of the kernel (profiles/r2_u_mlp_epilogue_variants.txt).
#ifndef TB_FU_PROF
#define TB_FU_PROF 0
#endif
#ifndef TB_FU_EPI_WARPS
#define TB_FU_EPI_WARPS 8
#endif
constexpr int FU_EPI_WARPS = TB_FU_EPI_WARPS;            // two per TMEM lane quarter, alternating 32-column slabs (12 warps: no gain, the MMA side is the bound)
constexpr int FU_THREADS = 128 + 32 * FU_EPI_WARPS;
constexpr int FU_MAX_LAYERS = 12;
constexpr int FU_CLUSTER = 2;             // CTA pair: one tcgen05.mma.cta_group::2 spans both         // hidden layers + the head pseudo-layer

constexpr int FU_OFF_H = 0;
constexpr int FU_OFF_W = FU_OFF_H + FU_HCH * FU_CHUNK;
constexpr int FU_OFF_BAR = FU_OFF_W + 6 * FU_CHUNK;
constexpr int FU_OFF_COL = FU_OFF_W + 5 * FU_CHUNK;   // backward: [layers][256] bias-gradient partial sums (<= 16 KB)
constexpr int FU_SMEM = FU_OFF_BAR + 512;   // 33 mbarriers + the TMEM slot

struct FusedParams {
  int M, nl, skip, eb;               // rows, hidden layers, index of the skip layer, embedding chunks
  int blk_base[FU_MAX_LAYERS + 1];   // first packed k-block of each layer (layer nl = heads)
  const float* bias;                 // [nl][256], then the 16 head biases
  uint32_t* masks;                   // [nl][M][8]  ReLU bit masks
  float* out16;                      // [M][16] (or NULL when the three heads go to o3)
  float* o3[3];                      // optional separate head outputs: d_xyz [M][3], d_rotation [M][4], d_scaling [M][3]
  float* acts;                       // [nl][M][256]  (backward: dZ of every layer, in backward order)
  int acts16;                        // A16 kernels: the bf16 tensor map (tmActs) is live
  float* colsum[FU_MAX_LAYERS];      // backward: bias gradient [256] of every forward layer (+=)
  unsigned long long* prof;          // optional [grid][8] stall-cycle counters (tools/check_fused_mlp.py)
};

__device__ __forceinline__ int fu_emb_blocks(const FusedParams& p, int L) {
  return (L == 0 || (L < p.nl && L - 1 == p.skip)) ? p.eb : 0;
}
// Column sums over the warp's 32 rows: lane l returns the total of v[l] (31 shuffles).
__device__ __forceinline__ float fu_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const bool hi = lane & w;
#pragma unroll
    for (int i = 0; i < w; i++) {
      const float send = hi ? v[i] : v[i + w];
      const float keep = hi ? v[i + w] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return v[0];
}
__device__ __forceinline__ int fu_h_blocks(int L) { return L == 0 ? 0 : FU_HCH; }

// BWD = false: forward (bias + ReLU epilogue, heads pseudo-layer).
// BWD = true : the data-gradient chain dZ_{l-1} = (dZ_l W_l) * relu'(h_{l-1}): the same pipeline with the packed
// upstream gradient [M][16] as first operand, transposed weights, an epilogue that applies the stored ReLU
// bit mask and accumulates the bias gradients (column sums), and a final drain pseudo-layer without MMAs.
//
// A16 = true: what goes to HBM for the backward pass is a BF16 copy of each layer's output instead of the fp32 one.
// Nothing in the chain itself reads those tensors back (the next layer's operand stays in shared memory, fp32 rounded
// to TF32; the data-gradient chain needs only the ReLU bit masks): their only consumer is the weight-gradient
// product dW = dZ^T X, which streams them once from HBM and is bound by exactly that traffic.  Half the bytes
// written here, half the bytes read there.  Each epilogue warp converts its 32x32 slab (from the value BEFORE the
// TF32 rounding), stages it in 2 KB of its own (64-byte-swizzled rows) and TMA-stores it itself; warp 1 stores nothing.
template <bool BWD, bool A16>
__global__ void __launch_bounds__(FU_THREADS, 1)
mlp_fused_kernel(const __grid_constant__ CUtensorMap tmEmb, const __grid_constant__ CUtensorMap tmW,
                         const __grid_constant__ CUtensorMap tmWh, const __grid_constant__ CUtensorMap tmActs,
                         const FusedParams p) {
  pdl_wait();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u)) __trap();
  uint8_t* sH = smem + FU_OFF_H;
  uint8_t* sW = smem + FU_OFF_W;
  float* sCol = reinterpret_cast<float*>(smem + FU_OFF_COL);
  constexpr int FU_STAGES = FuCfg<BWD, A16>::STAGES;
  uint8_t* sStore = sW + FU_STAGES * FU_CHUNK;          // A16: per-epilogue-warp bf16 store staging (the stage given up)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FU_OFF_BAR);
  uint64_t* full = bars;                        // [STAGES] ring stage landed (leader: in both CTAs)
  uint64_t* empty = full + FU_STAGES;           // [STAGES] ring stage consumed by the pair's MMAs
  uint64_t* a_ready = empty + FU_STAGES;        // [8] leader only: activation chunk written in both CTAs
  uint64_t* a_local = a_ready + FU_HCH;         // [8] activation chunk written in this CTA (for its HBM store)
  uint64_t* tmem_full = a_local + FU_HCH;       // [2]
  uint64_t* tmem_empty = tmem_full + 2;         // [2] leader only: accumulator drained in both CTAs
  uint64_t* h_free = tmem_empty + 2;            // activation chunks may be overwritten (their HBM stores have read them)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = (p.M + FU_BM - 1) / FU_BM;
  const uint32_t crank = cluster_ctarank();
  const bool leader = crank == 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmEmb); prefetch_tmap(&tmW); prefetch_tmap(&tmWh); prefetch_tmap(&tmActs);
    for (int i = 0; i < FU_STAGES; i++) {
      // leader: its own TMA (expect_tx arrival) + the follower's relay; follower: its own TMA
      mbar_init(&full[i], leader ? 2 : 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < FU_HCH; i++) { mbar_init(&a_ready[i], 4 * FU_CLUSTER); mbar_init(&a_local[i], 4); }
    for (int i = 0; i < 2; i++) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], FU_EPI_WARPS * FU_CLUSTER); }
    mbar_init(h_free, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc_2cta(tmem_slot, 512); tmem_relinquish_2cta(); }
  if (BWD) for (int i = threadIdx.x; i < p.nl * FU_W; i += FU_THREADS) sCol[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  cluster_sync();            // the peer's barriers exist before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint16_t CMASK = (1u << FU_CLUSTER) - 1;
  // the two CTAs of a pair walk the same number of tiles (a tile index past the end is a dummy: its loads are
  // zero-filled and its stores clipped): every MMA spans both
  const int iters = (num_tiles + gridDim.x - 1) / gridDim.x;

  if (warp == 0) {
    // ===== TMA producer: this CTA's half of every weight k-block, and its own embedding chunks =====
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int it = 0, tile = blockIdx.x; it < iters; it++, tile += gridDim.x) {
        for (int L = 0; L < (BWD ? p.nl : p.nl + 1); L++) {
          const int ne = fu_emb_blocks(p, L), nkb = ne + fu_h_blocks(L);
          for (int kb = 0; kb < nkb; kb++) {
            if (kb < ne) {
              mbar_wait(&empty[stage], phase ^ 1);
              mbar_expect_tx(&full[stage], FU_CHUNK);
              tma_load_2d(sW + stage * FU_CHUNK, &tmEmb, &full[stage], kb * 32, tile * FU_BM);
              if (++stage == FU_STAGES) { stage = 0; phase ^= 1; }
            }
            mbar_wait(&empty[stage], phase ^ 1);
            if (BWD || L < p.nl) {
              // output features [128 crank, 128 crank + 128) of this k-block
              mbar_expect_tx(&full[stage], FU_CHUNK);
              tma_load_2d(sW + stage * FU_CHUNK, &tmW, &full[stage], 0, (p.blk_base[L] + kb) * FU_W + crank * 128);
            } else {
              // heads: N = 32 over the pair, 16 rows per CTA (rows 16..31 of the packed block are zero)
              mbar_expect_tx(&full[stage], 16 * 128);
              tma_load_2d(sW + stage * FU_CHUNK, &tmWh, &full[stage], 0, (p.blk_base[L] + kb) * FU_W + crank * 16);
            }
            if (++stage == FU_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== leader: MMA issuer for the pair.  follower: relays "stage landed" to the leader.
    // Both: TMA-store each finished activation chunk of their own tile to HBM =====
    if (lane == 0) {
      constexpr uint32_t idesc_h = make_idesc_tf32(2 * FU_BM, 256, 0, 0);
      constexpr uint32_t idesc_o = make_idesc_tf32(2 * FU_BM, 32, 0, 0);
      uint32_t stage = 0, phase = 0, a_phase = 0, lc = 0;
      long long w_acc = 0, w_a = 0, w_full = 0, w_st = 0, w_iss = 0, t_all = clock64(), t0 = 0;
#if TB_FU_PROF
#define FU_T0() (t0 = clock64())
#define FU_ADD(x) ((x) += clock64() - t0)
#else
#define FU_T0() ((void)0)
#define FU_ADD(x) ((void)(x), (void)t0)
#endif
      for (int it = 0, tile = blockIdx.x; it < iters; it++, tile += gridDim.x) {
        for (int L = 0; L <= p.nl; L++) {
          const bool mma_layer = !BWD || L < p.nl;         // backward: layer nl only drains the last chunks to HBM
          const uint32_t acc = lc & 1, acc_phase = (lc >> 1) & 1;
          if (leader && mma_layer) {
            FU_T0();
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            FU_ADD(w_acc);
            tc_fence_after();
          }
          const uint32_t tmem_d = tmem_base + acc * FU_W;
          const int ne = fu_emb_blocks(p, L), nkb = ne + fu_h_blocks(L);
          for (int kb = 0; kb < nkb; kb++) {
            uint32_t a_addr;
            int a_stage = -1;
            if (kb < ne) {
              if (leader) {
                FU_T0();
                mbar_wait(&full[stage], phase);
                FU_ADD(w_full);
                tc_fence_after();
              }
              a_stage = (int)stage;
              a_addr = smem_u32(sW + stage * FU_CHUNK);
              if (++stage == FU_STAGES) { stage = 0; phase ^= 1; }
            } else {
              const int j = kb - ne;
              FU_T0();
              mbar_wait(leader ? &a_ready[j] : &a_local[j], a_phase);
              FU_ADD(w_a);
              tc_fence_after();
              a_addr = smem_u32(sH + j * FU_CHUNK);
              if (!A16 && p.acts) tma_store_3d(&tmActs, sH + j * FU_CHUNK, j * 32, tile * FU_BM, L - 1);   // h_{L-1}[:, 32j .. 32j+31]
            }
            if (!mma_layer) continue;
            if (leader) {
              FU_T0();
              mbar_wait(&full[stage], phase);
              FU_ADD(w_full);
              tc_fence_after();
              const uint32_t b_addr = smem_u32(sW + stage * FU_CHUNK);
              FU_T0();
#pragma unroll
              for (int k = 0; k < 4; k++) {
                const uint64_t da = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
                const uint64_t db = make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
                mma_tf32_ss_2cta(tmem_d, da, db, (BWD || L < p.nl) ? idesc_h : idesc_o, (kb | k) != 0);
              }
              mma_commit_2cta(&empty[stage], CMASK);
              if (a_stage >= 0) mma_commit_2cta(&empty[a_stage], CMASK);   // the embedding chunks' stages
              FU_ADD(w_iss);
            }
            if (++stage == FU_STAGES) { stage = 0; phase ^= 1; }
          }
          if (L > 0) a_phase ^= 1;
          if (leader && mma_layer) mma_commit_2cta(&tmem_full[acc], CMASK);
          if (mma_layer) lc++;
          // the epilogue of this layer overwrites the chunks in place: the stores above must have read them first
          // (the epilogue waits for h_free only before its first shared-memory write)
          tma_store_commit();
          FU_T0();
          tma_store_wait_read<0>();
          FU_ADD(w_st);
          mbar_arrive(h_free);
        }
      }
      tma_store_wait<0>();
      if (p.prof) {
        unsigned long long* q = p.prof + blockIdx.x * 8;
        q[0] = clock64() - t_all; q[1] = w_acc; q[2] = w_a; q[3] = w_full; q[4] = w_st; q[7] = w_iss;
      }
    }
  } else if (warp == 3) {
    // ===== follower only: tell the leader's MMA thread, stage by stage, that this CTA's operands have landed =====
    if (lane == 0 && !leader) {
      uint32_t stage = 0, phase = 0;
      for (int it = 0; it < iters; it++)
        for (int L = 0; L < (BWD ? p.nl : p.nl + 1); L++) {
          const int n = 2 * fu_emb_blocks(p, L) + fu_h_blocks(L);     // ring entries of this layer
          for (int e = 0; e < n; e++) {
            mbar_wait(&full[stage], phase);
            mbar_arrive_cluster(&full[stage], 0);
            if (++stage == FU_STAGES) { stage = 0; phase ^= 1; }
          }
        }
    }
  } else if (warp >= 4) {
    // ===== epilogue: warp (w & 3) owns TMEM lanes / tile rows 32 (w & 3) .. +31; the two warps of a lane
    // quarter take alternate 32-column slabs =====
    const int ew = warp & 3, eh = (warp - 4) >> 2;
    const int r = ew * 32 + lane;                 // row inside the tile
    uint32_t lc = 0;
    long long e_full = 0, e_free = 0, t0 = 0;
    for (int it = 0, tile = blockIdx.x; it < iters; it++, tile += gridDim.x) {
      const int grow = tile * FU_BM + r;
      for (int L = 0; L < (BWD ? p.nl : p.nl + 1); L++, lc++) {
        const uint32_t acc = lc & 1, acc_phase = (lc >> 1) & 1;
        const uint32_t hfree_parity = (uint32_t)(it * (p.nl + 1) + L) & 1u;   // warp 1 signals h_free nl + 1 times per tile
        FU_T0();
        mbar_wait(&tmem_full[acc], acc_phase);
        FU_ADD(e_full);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + ((uint32_t)(ew * 32) << 16) + acc * FU_W;
        if (BWD || L < p.nl) {
          // forward layer this step belongs to: its bias (forward) or its ReLU mask and bias gradient (backward)
          const int fl = BWD ? p.nl - 1 - L : L;
          const float* bias = p.bias + fl * FU_W;
          uint32_t* mask = p.masks + ((size_t)fl * p.M + (grow < p.M ? grow : 0)) * (FU_W / 32);
#pragma unroll 1
          for (int s = eh; s < FU_HCH; s += FU_EPI_WARPS / 4) {
            uint32_t v[32];
            tmem_ld_32x32(t_addr + s * 32, v);
            float4 bq[8];
            uint32_t keep = 0;
            if (BWD) {
              keep = grow < p.M ? __ldg(mask + s) : 0u;
            } else {
#pragma unroll
              for (int c = 0; c < 8; c++) bq[c] = __ldg(reinterpret_cast<const float4*>(bias + s * 32) + c);
            }
            tmem_ld_wait();
            if (s == eh) { FU_T0(); mbar_wait(h_free, hfree_parity); FU_ADD(e_free); }
            uint8_t* dst = sH + s * FU_CHUNK + r * 128;
            uint32_t pos = 0;
            uint8_t* st16 = sStore + (warp - 4) * 2048;
#pragma unroll
            for (int c = 0; c < 8; c++) {
              const float bb[4] = {bq[c].x, bq[c].y, bq[c].z, bq[c].w};
              float4 o;
              float* of = reinterpret_cast<float*>(&o);
#pragma unroll
              for (int j = 0; j < 4; j++) {
                // forward: bias + ReLU; backward: ReLU derivative from the stored bit mask.  Then round to
                // TF32 (nearest, ties away: + half ulp, truncate — what cvt.rna.tf32 computes for finite
                // values, on the integer pipe instead of the quarter-rate converter)
                float x = __uint_as_float(v[c * 4 + j]);
                if (BWD) x = ((keep >> (c * 4 + j)) & 1u) ? x : 0.f;
                else x = fmaxf(x + bb[j], 0.f);
                const uint32_t u = (__float_as_uint(x) + 0x1000u) & 0xFFFFE000u;
                pos |= (u != 0u ? 1u : 0u) << (c * 4 + j);
                of[j] = __uint_as_float(u);
                if (BWD) v[c * 4 + j] = u;
                if (A16 && !BWD) v[c * 4 + j] = __float_as_uint(x);      // unrounded value for the bf16 copy
              }
              *reinterpret_cast<float4*>(dst + ((c ^ (r & 7)) << 4)) = o;
            }
            // the chunk goes to the next layer's MMAs first; what only the backward pass needs (bf16 copy, ReLU mask)
            // follows, off the critical path of the layer hand-over
            fence_proxy_async_smem();     // generic-proxy writes -> visible to the MMA / TMA (async proxy)
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(&a_local[s]);
              if (leader) mbar_arrive(&a_ready[s]); else mbar_arrive_cluster(&a_ready[s], 0);
            }
            if (!BWD && p.masks && grow < p.M) mask[s] = pos;
            if (A16 && p.acts16) {
              // this warp's previous slab store must have read its staging before it is overwritten
              if (lane == 0) tma_store_wait_read<0>();
              __syncwarp();
              // 32 values -> 64 bytes of bf16, row `lane` of the warp's staging, 16-byte chunks XOR-swizzled like
              // CU_TENSOR_MAP_SWIZZLE_64B (chunk ^ bits 7..8 of the address = (row >> 1) & 3): conflict-free STS.128
#pragma unroll
              for (int c = 0; c < 4; c++) {
                uint4 o16;
                uint32_t* ow = reinterpret_cast<uint32_t*>(&o16);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                  const __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(v[c * 8 + 2 * j]),
                                                                  __uint_as_float(v[c * 8 + 2 * j + 1]));
                  ow[j] = *reinterpret_cast<const uint32_t*>(&h2);
                }
                *reinterpret_cast<uint4*>(st16 + lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4)) = o16;
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_3d(&tmActs, st16, s * 32, tile * FU_BM + ew * 32, L);     // rows past M are clipped
                tma_store_commit();
              }
            }
            if (BWD) {
              // bias gradient of forward layer fl: column sums of the values just written (rows past M are 0)
              float cv[32];
#pragma unroll
              for (int i = 0; i < 32; i++) cv[i] = __uint_as_float(v[i]);
              const float tot = fu_transpose_reduce32(cv, lane);
              atomicAdd(sCol + fl * FU_W + s * 32 + lane, tot);
            }
          }
        } else if (eh == 0) {
          uint32_t v[16];
          tmem_ld_32x16(t_addr, v);
          const float* bias = p.bias + p.nl * FU_W;
          float4 bq[4];
#pragma unroll
          for (int c = 0; c < 4; c++) bq[c] = __ldg(reinterpret_cast<const float4*>(bias) + c);
          tmem_ld_wait();
          if (grow < p.M) {
            float4 o[4];
#pragma unroll
            for (int c = 0; c < 4; c++)
              o[c] = make_float4(__uint_as_float(v[c * 4]) + bq[c].x, __uint_as_float(v[c * 4 + 1]) + bq[c].y,
                                 __uint_as_float(v[c * 4 + 2]) + bq[c].z, __uint_as_float(v[c * 4 + 3]) + bq[c].w);
            if (p.o3[0]) {
              // the three heads as the contiguous tensors the caller hands on (columns 0-2, 3-6, 7-9 of the packed row)
              float* x = p.o3[0] + (size_t)grow * 3;
              x[0] = o[0].x; x[1] = o[0].y; x[2] = o[0].z;
              *reinterpret_cast<float4*>(p.o3[1] + (size_t)grow * 4) = make_float4(o[0].w, o[1].x, o[1].y, o[1].z);
              float* sc = p.o3[2] + (size_t)grow * 3;
              sc[0] = o[1].w; sc[1] = o[2].x; sc[2] = o[2].y;
            } else {
              float4* dst = reinterpret_cast<float4*>(p.out16 + (size_t)grow * 16);
#pragma unroll
              for (int c = 0; c < 4; c++) dst[c] = o[c];
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (leader) mbar_arrive(&tmem_empty[acc]); else mbar_arrive_cluster(&tmem_empty[acc], 0); }
      }
    }
    if (A16 && lane == 0) tma_store_wait<0>();
    if (p.prof && warp == 4 && lane == 0) { p.prof[blockIdx.x * 8 + 5] = e_full; p.prof[blockIdx.x * 8 + 6] = e_free; }
  }
  tc_fence_before();
  __syncthreads();
  if (BWD) for (int i = threadIdx.x; i < p.nl * FU_W; i += FU_THREADS) atomicAdd(p.colsum[i / FU_W] + (i % FU_W), sCol[i]);
  cluster_sync();            // no CTA leaves while its peer may still signal its barriers or read its shared memory
  if (warp == 2) tmem_dealloc_2cta(tmem_base, 512);
}

unsigned long long* g_fused_prof = nullptr;   // set through trase_b200_set_tuning_ptr (debug)

// ---- weight packing: [k-block][256 rows][32] fp32, TF32-rounded, zero-padded ---------------------------
constexpr int FU_MAX_SEGS = 16;
struct PackSeg {
  const float* src;   // tr == 0: element (n, k) = src[n * ld + k0 + k];  tr == 1: src[(k0 + k) * ld + n0 + n]
  int rows, ld, k0, kcount, blk0, nblk, tr, n0;
};
struct PackTable {
  PackSeg seg[FU_MAX_SEGS];
  int nseg;
};

__global__ void __launch_bounds__(256) mlp_pack_fused_kernel(const PackTable t, float* __restrict__ out) {
  pdl_wait();
  const int blk = blockIdx.x;
  int si = 0;
  while (si + 1 < t.nseg && blk >= t.seg[si + 1].blk0) si++;
  const PackSeg sg = t.seg[si];
  const int b = blk - sg.blk0;
  for (int e = threadIdx.x; e < FU_W * 32; e += 256) {
    const int n = e >> 5, c = e & 31;
    const int k = b * 32 + c;
    float v = 0.f;
    if (n < sg.rows && k < sg.kcount) {
      uint32_t u;
      const size_t at = sg.tr ? (size_t)(sg.k0 + k) * sg.ld + sg.n0 + n : (size_t)n * sg.ld + sg.k0 + k;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(sg.src[at]));
      v = __uint_as_float(u);
    }
    out[(size_t)blk * FU_W * 32 + e] = v;
  }
}

static int fu_layout(int nl, int skip, int E, int* blk_base) {
  const int eb = (E + 31) / 32;
  int n = 0;
  for (int L = 0; L <= nl; L++) {
    blk_base[L] = n;
    n += (L == 0) ? eb : ((L < nl && L - 1 == skip) ? eb + FU_HCH : FU_HCH);
  }
  blk_base[nl + 1] = n;
  return n;
}
static bool fu_shape_ok(int nl, int skip, int E) {
  return nl >= 2 && nl + 1 <= FU_MAX_LAYERS && E > 0 && E <= 32 * FU_ECH && (E & 3) == 0 && skip < nl - 1;
}

typedef CUresult (*EncodeTiledFn3)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static bool make_tmap_acts(CUtensorMap* m, float* base, uint64_t M, uint64_t nl) {
  static EncodeTiledFn3 fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) != cudaSuccess ||
        r != cudaDriverEntryPointSuccess)
      return false;
    fn = reinterpret_cast<EncodeTiledFn3>(q);
  }
  cuuint64_t dims[3] = {(cuuint64_t)FU_W, M, nl};
  cuuint64_t strides[2] = {(cuuint64_t)FU_W * 4, M * FU_W * 4};
  cuuint32_t box[3] = {32, FU_BM, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// bf16 [nl][M][256] side output of the A16 kernels: one 32-row x 32-column box per epilogue-warp slab
static bool make_tmap_acts16(CUtensorMap* m, void* base, uint64_t M, uint64_t nl) {
  static EncodeTiledFn3 fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) != cudaSuccess ||
        r != cudaDriverEntryPointSuccess)
      return false;
    fn = reinterpret_cast<EncodeTiledFn3>(q);
  }
  cuuint64_t dims[3] = {(cuuint64_t)FU_W, M, nl};
  cuuint64_t strides[2] = {(cuuint64_t)FU_W * 2, M * FU_W * 2};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace mlp
}  // namespace tb

using namespace tb;

// number of floats of the packed weight buffer for a network of nl hidden layers of width 256, skip
// concatenation after layer `skip` (-1: none), embedding width E (multiple of 4, <= 96)
extern "C" TRASE_API size_t trase_b200_mlp_fused_pack_floats(int nl, int skip, int E) {
  if (!mlp::fu_shape_ok(nl, skip, E)) return 0;
  int base[mlp::FU_MAX_LAYERS + 2];
  return (size_t)mlp::fu_layout(nl, skip, E, base) * mlp::FU_W * 32;
}

// weights: host array of nl device pointers, weights[i] = nn.Linear weight of layer i, [256][in_i] with
// in_0 = E, in_{skip+1} = E + 256 (embedding columns first), 256 otherwise; wh: [16][256] packed heads.
extern "C" TRASE_API int trase_b200_mlp_pack_fused(int nl, int skip, int E, const float* const* weights,
                                                   const float* wh, float* wpack, void* stream) {
  if (!mlp::fu_shape_ok(nl, skip, E) || !weights || !wh || !wpack) {
    set_error("trase_b200_mlp_pack_fused: unsupported network shape (nl=%d skip=%d E=%d) or null pointer", nl, skip, E);
    return 1;
  }
  int base[mlp::FU_MAX_LAYERS + 2];
  const int total = mlp::fu_layout(nl, skip, E, base);
  const int eb = (E + 31) / 32;
  mlp::PackTable t;
  t.nseg = 0;
  for (int L = 0; L <= nl; L++) {
    const float* src = L < nl ? weights[L] : wh;
    if (!src) { set_error("trase_b200_mlp_pack_fused: null weight pointer"); return 1; }
    const int rows = L < nl ? mlp::FU_W : 16;
    if (L == 0) {
      t.seg[t.nseg++] = {src, rows, E, 0, E, base[L], eb, 0, 0};
    } else if (L < nl && L - 1 == skip) {
      t.seg[t.nseg++] = {src, rows, E + mlp::FU_W, 0, E, base[L], eb, 0, 0};
      t.seg[t.nseg++] = {src, rows, E + mlp::FU_W, E, mlp::FU_W, base[L] + eb, mlp::FU_HCH, 0, 0};
    } else {
      t.seg[t.nseg++] = {src, rows, mlp::FU_W, 0, mlp::FU_W, base[L], mlp::FU_HCH, 0, 0};
    }
  }
  launch_k(mlp::mlp_pack_fused_kernel, total, 256, 0, reinterpret_cast<cudaStream_t>(stream), t, wpack);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_pack_fused: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

namespace tb {
namespace mlp {
template <bool BWD, bool A16>
static int launch_fused(const char* who, const CUtensorMap& tE, const CUtensorMap& tW, const CUtensorMap& tWh,
                        const CUtensorMap& tA, const FusedParams& p, cudaStream_t stream) {
  static PerDeviceOnce attr_once;
  {
    cudaError_t e = ensure_max_dyn_smem(attr_once, mlp_fused_kernel<BWD, A16>, FU_SMEM);
    if (e != cudaSuccess) { set_error("%s: %s", who, cudaGetErrorString(e)); return 2; }
  }
  // SMs available to the persistent MLP kernels, whole pairs.  "reserve_sms" leaves SMs to a gradient all-reduce that
  // runs under the MLP BACKWARD (chain + weight gradients); nothing overlaps the forward, which takes them all.
  const int sms = (BWD ? num_sms() : num_sms() + g_reserve_sms) & ~1;
  const int tiles = (p.M + FU_BM - 1) / FU_BM;
  int grid = tiles < sms ? tiles : sms;
  grid = (grid + FU_CLUSTER - 1) / FU_CLUSTER * FU_CLUSTER;
  if (grid > sms) grid -= FU_CLUSTER;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(FU_THREADS);
  cfg.dynamicSmemBytes = FU_SMEM;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = FU_CLUSTER;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
#if TB_PDL
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.numAttrs = 2;
#endif
  cudaError_t le = cudaLaunchKernelEx(&cfg, mlp_fused_kernel<BWD, A16>, tE, tW, tWh, tA, p);
  if (le != cudaSuccess) { set_error("%s: launch failed: %s", who, cudaGetErrorString(le)); return 2; }
  count_launches(1);
  return 0;
}
}  // namespace mlp
}  // namespace tb

// emb: [M][ld_emb] (first E columns used, ld_emb multiple of 4), bias: nl*256 + 16 floats,
// acts: [nl][M][256] (written), masks: [nl][M][8] (written), out16: [M][16] (written).
// acts == masks == NULL: inference, nothing is kept for a backward pass (no activation traffic to HBM at all).
static int mlp_forward_fused_impl(int M, int nl, int skip, int E, const float* emb, int ld_emb, const float* wpack,
                                  const float* bias, float* acts, void* acts16, uint32_t* masks, float* out16,
                                  float* d_xyz, float* d_rot, float* d_scale, void* stream);
// d_xyz [M][3], d_rot [M][4] (16-byte aligned), d_scale [M][3]: optional (all three or none); with them the three heads
// are written as those contiguous tensors and out16 may be NULL.
extern "C" TRASE_API int trase_b200_mlp_forward_fused(int M, int nl, int skip, int E, const float* emb, int ld_emb,
                                                      const float* wpack, const float* bias, float* acts,
                                                      uint32_t* masks, float* out16, float* d_xyz, float* d_rot,
                                                      float* d_scale, void* stream) {
  return mlp_forward_fused_impl(M, nl, skip, E, emb, ld_emb, wpack, bias, acts, nullptr, masks, out16, d_xyz, d_rot,
                                d_scale, stream);
}
// Same network, but the activations kept for the backward pass are written as BF16 (acts16: [nl][M][256] bf16,
// 16-byte aligned): their only reader is the weight-gradient product (trase_b200_wgrad_bf16).
extern "C" TRASE_API int trase_b200_mlp_forward_fused_bf16(int M, int nl, int skip, int E, const float* emb,
                                                           int ld_emb, const float* wpack, const float* bias,
                                                           void* acts16, uint32_t* masks, float* out16, float* d_xyz,
                                                           float* d_rot, float* d_scale, void* stream) {
  if (!acts16 || !masks) { set_error("trase_b200_mlp_forward_fused_bf16: acts16 and masks are required"); return 1; }
  return mlp_forward_fused_impl(M, nl, skip, E, emb, ld_emb, wpack, bias, nullptr, acts16, masks, out16, d_xyz, d_rot,
                                d_scale, stream);
}
static int mlp_forward_fused_impl(int M, int nl, int skip, int E, const float* emb, int ld_emb, const float* wpack,
                                  const float* bias, float* acts, void* acts16, uint32_t* masks, float* out16,
                                  float* d_xyz, float* d_rot, float* d_scale, void* stream) {
  if (M <= 0) return 0;
  const bool split = d_xyz || d_rot || d_scale;
  if (split && (!d_xyz || !d_rot || !d_scale || ((uintptr_t)d_rot & 15))) {
    set_error("trase_b200_mlp_forward_fused: d_xyz / d_rot / d_scale come together, d_rot 16-byte aligned");
    return 1;
  }
  if (!mlp::fu_shape_ok(nl, skip, E) || !emb || !wpack || !bias || (!out16 && !split) ||
      (((acts == nullptr) && (acts16 == nullptr)) != (masks == nullptr)) || ((uintptr_t)acts16 & 15) || (ld_emb & 3) ||
      ld_emb < E || ((uintptr_t)emb & 15) || ((uintptr_t)wpack & 15) || ((uintptr_t)acts & 15) ||
      ((uintptr_t)bias & 15) || ((uintptr_t)out16 & 15)) {
    set_error("trase_b200_mlp_forward_fused: bad arguments (M=%d nl=%d skip=%d E=%d)", M, nl, skip, E);
    return 1;
  }
  mlp::FusedParams p;
  p.M = M; p.nl = nl; p.skip = skip; p.eb = (E + 31) / 32;
  int base[mlp::FU_MAX_LAYERS + 2];
  const int total = mlp::fu_layout(nl, skip, E, base);
  for (int L = 0; L <= nl; L++) p.blk_base[L] = base[L];
  p.bias = bias; p.masks = masks; p.out16 = out16; p.acts = acts; p.acts16 = acts16 ? 1 : 0;
  p.o3[0] = d_xyz; p.o3[1] = d_rot; p.o3[2] = d_scale;
  for (int i = 0; i < mlp::FU_MAX_LAYERS; i++) p.colsum[i] = nullptr;
  p.prof = mlp::g_fused_prof;
  CUtensorMap tE, tW, tWh, tA;
  if (!mlp::make_tmap_2d(&tE, emb, M, E, ld_emb, mlp::FU_BM, 32, CU_TENSOR_MAP_SWIZZLE_128B) ||
      !mlp::make_tmap_2d(&tW, wpack, (uint64_t)total * mlp::FU_W, 32, 32, 128, 32, CU_TENSOR_MAP_SWIZZLE_128B) ||
      !mlp::make_tmap_2d(&tWh, wpack, (uint64_t)total * mlp::FU_W, 32, 32, 16, 32, CU_TENSOR_MAP_SWIZZLE_128B) ||
      !(acts16 ? mlp::make_tmap_acts16(&tA, acts16, M, nl)
               : mlp::make_tmap_acts(&tA, acts ? acts : const_cast<float*>(emb), M, acts ? nl : 1))) {   // inference: map unused, must still be valid
    set_error("trase_b200_mlp_forward_fused: cuTensorMapEncodeTiled failed");
    return 2;
  }
  if (acts16)
    return mlp::launch_fused<false, true>("trase_b200_mlp_forward_fused_bf16", tE, tW, tWh, tA, p,
                                          reinterpret_cast<cudaStream_t>(stream));
  return mlp::launch_fused<false, false>("trase_b200_mlp_forward_fused", tE, tW, tWh, tA, p,
                                         reinterpret_cast<cudaStream_t>(stream));
}

// ---- backward data-gradient chain -----------------------------------------------------------------------
// packed operand of the chain: one k-block from whT [256][16] (head weights transposed), then eight k-blocks
// per hidden layer nl-1 .. 1: B[n][k] = W_l[k][off_l + n], read (transposed) straight from the raw nn.Linear
// weights; off_l = E for the layer behind the skip (its first E input columns are the embedding), else 0
extern "C" TRASE_API size_t trase_b200_mlp_fused_pack_bwd_floats(int nl) {
  if (nl < 2 || nl + 1 > mlp::FU_MAX_LAYERS) return 0;
  return (size_t)(1 + mlp::FU_HCH * (nl - 1)) * mlp::FU_W * 32;
}

extern "C" TRASE_API int trase_b200_mlp_pack_fused_bwd(int nl, int skip, int E, const float* whT,
                                                       const float* const* weights, float* wpack, void* stream) {
  if (nl < 2 || nl + 1 > mlp::FU_MAX_LAYERS || !whT || !weights || !wpack || E < 0) {
    set_error("trase_b200_mlp_pack_fused_bwd: bad arguments");
    return 1;
  }
  mlp::PackTable t;
  t.nseg = 0;
  t.seg[t.nseg++] = {whT, mlp::FU_W, 16, 0, 16, 0, 1, 0, 0};
  for (int L = 1; L < nl; L++) {
    const int fl = nl - L;                      // backward step L differentiates forward layer nl - L
    const float* src = weights[fl];
    if (!src) { set_error("trase_b200_mlp_pack_fused_bwd: null weight pointer"); return 1; }
    const int off = (fl - 1 == skip) ? E : 0;
    t.seg[t.nseg++] = {src, mlp::FU_W, off + mlp::FU_W, 0, mlp::FU_W, 1 + (L - 1) * mlp::FU_HCH, mlp::FU_HCH, 1, off};
  }
  const int total = 1 + mlp::FU_HCH * (nl - 1);
  launch_k(mlp::mlp_pack_fused_kernel, total, 256, 0, reinterpret_cast<cudaStream_t>(stream), t, wpack);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("trase_b200_mlp_pack_fused_bwd: %s", cudaGetErrorString(e)); return 2; }
  count_launches(1);
  return 0;
}

// dout: [M][16] packed upstream gradient of the heads; masks: [nl][M][8] from the forward pass;
// dzs: [nl][M][256] (written): dzs[L] = dZ of forward layer nl-1-L; colsum: host array of nl device pointers,
// colsum[l] = 256 floats that receive += the bias gradient of forward layer l.
static int mlp_backward_fused_impl(int M, int nl, const float* dout, const float* wpack, const uint32_t* masks,
                                   float* dzs, void* dzs16, float* const* colsum, void* stream);
extern "C" TRASE_API int trase_b200_mlp_backward_fused(int M, int nl, const float* dout, const float* wpack,
                                                       const uint32_t* masks, float* dzs, float* const* colsum,
                                                       void* stream) {
  if (!dzs) { set_error("trase_b200_mlp_backward_fused: bad arguments (dzs)"); return 1; }
  return mlp_backward_fused_impl(M, nl, dout, wpack, masks, dzs, nullptr, colsum, stream);
}
// Same chain with the layer gradients dZ written as BF16 (dzs16: [nl][M][256] bf16) for trase_b200_wgrad_bf16.
extern "C" TRASE_API int trase_b200_mlp_backward_fused_bf16(int M, int nl, const float* dout, const float* wpack,
                                                            const uint32_t* masks, void* dzs16, float* const* colsum,
                                                            void* stream) {
  if (!dzs16) { set_error("trase_b200_mlp_backward_fused_bf16: bad arguments (dzs16)"); return 1; }
  return mlp_backward_fused_impl(M, nl, dout, wpack, masks, nullptr, dzs16, colsum, stream);
}
static int mlp_backward_fused_impl(int M, int nl, const float* dout, const float* wpack, const uint32_t* masks,
                                   float* dzs, void* dzs16, float* const* colsum, void* stream) {
  if (M <= 0) return 0;
  if (nl < 2 || nl + 1 > mlp::FU_MAX_LAYERS || !dout || !wpack || !masks || !colsum ||
      ((uintptr_t)dout & 15) || ((uintptr_t)wpack & 15) || ((uintptr_t)dzs & 15) || ((uintptr_t)dzs16 & 15)) {
    set_error("trase_b200_mlp_backward_fused: bad arguments (M=%d nl=%d)", M, nl);
    return 1;
  }
  mlp::FusedParams p;
  p.M = M; p.nl = nl; p.skip = -1; p.eb = 1;
  p.blk_base[0] = 0;
  for (int L = 1; L <= nl; L++) p.blk_base[L] = 1 + (L - 1) * mlp::FU_HCH;
  p.bias = nullptr; p.masks = const_cast<uint32_t*>(masks); p.out16 = nullptr; p.acts = dzs; p.acts16 = dzs16 ? 1 : 0;
  p.o3[0] = p.o3[1] = p.o3[2] = nullptr;
  for (int i = 0; i < nl; i++) {
    if (!colsum[i]) { set_error("trase_b200_mlp_backward_fused: null bias-gradient pointer"); return 1; }
    p.colsum[i] = colsum[i];
  }
  p.prof = mlp::g_fused_prof;
  const int total = 1 + mlp::FU_HCH * (nl - 1);
  CUtensorMap tE, tW, tA;
  if (!mlp::make_tmap_2d(&tE, dout, M, 16, 16, mlp::FU_BM, 32, CU_TENSOR_MAP_SWIZZLE_128B) ||
      !mlp::make_tmap_2d(&tW, wpack, (uint64_t)total * mlp::FU_W, 32, 32, 128, 32, CU_TENSOR_MAP_SWIZZLE_128B) ||
      !(dzs16 ? mlp::make_tmap_acts16(&tA, dzs16, M, nl) : mlp::make_tmap_acts(&tA, dzs, M, nl))) {
    set_error("trase_b200_mlp_backward_fused: cuTensorMapEncodeTiled failed");
    return 2;
  }
  if (dzs16)
    return mlp::launch_fused<true, true>("trase_b200_mlp_backward_fused_bf16", tE, tW, tW, tA, p,
                                         reinterpret_cast<cudaStream_t>(stream));
  return mlp::launch_fused<true, false>("trase_b200_mlp_backward_fused", tE, tW, tW, tA, p,
                                        reinterpret_cast<cudaStream_t>(stream));
}

// debug: device buffer of [grid][8] cycle counters filled by the fused kernel (NULL = off)
extern "C" TRASE_API void trase_b200_mlp_fused_set_prof(unsigned long long* dev_buf) { tb::mlp::g_fused_prof = dev_buf; }
