/*
 * trase_b200 — C ABI of the B200-native differentiable Gaussian rasterizer.
 *
 * This is the drop-in boundary for TRASE's rasterizer hot path.  Every entry
 * point takes plain pointers and sizes (no torch types); all device buffers
 * are owned by the caller (in the Python host: torch tensors), the library
 * never allocates device memory.  Kernels are launched on the CUDA stream the
 * caller passes (`stream`, a cudaStream_t cast to void*).
 *
 * "ref:" comments cite the reference interface each declaration replaces;
 * paths are relative to the reference repository root, DGR/ =
 * submodules/diff-gaussian-rasterization/.
 *
 * Return convention: 0 = ok, non-zero = error; the message is available from
 * trase_b200_last_error() (thread-local).  The Python host turns a non-zero
 * status into RuntimeError, mirroring AT_ERROR / std::runtime_error in the
 * reference (DGR/rasterize_points.cu:63-66, DGR/cuda_rasterizer/rasterizer_impl.cu:248-251).
 *
 * Layout conventions (identical to the reference): row-major contiguous fp32;
 * images planar CHW; SH coefficients [P][M][3]; features [P][32]; a NULL
 * pointer means "input not provided" (DGR/diff_gaussian_rasterization/__init__.py:206-219).
 */
#ifndef TRASE_B200_H_
#define TRASE_B200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define TRASE_API __attribute__((visibility("default")))
#else
#define TRASE_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define TRASE_B200_NUM_CHANNELS 3   /* ref: DGR/cuda_rasterizer/config.h:13 */
#define TRASE_B200_NUM_OBJECTS 32   /* ref: DGR/cuda_rasterizer/config.h:15 */
#define TRASE_B200_TILE 16          /* ref: DGR/cuda_rasterizer/config.h:16-17 */

/* Library version (major*10000 + minor*100 + patch). */
TRASE_API int trase_b200_version(void);

/* Message of the last failing call on this host thread ("" if none). */
TRASE_API const char* trase_b200_last_error(void);

/* ------------------------------------------------------------------------
 * Workspace sizing.  ref: CudaRasterizer::required<GeometryState|ImageState|
 * BinningState>() (DGR/cuda_rasterizer/rasterizer_impl.h:72-77) and the three
 * fromChunk() carvers (DGR/cuda_rasterizer/rasterizer_impl.cu:158-197).
 *
 *  geom    : per-Gaussian state written by forward, read by backward.
 *  img     : per-pixel final transmittance + contributor count, per-tile ranges.
 *  binning : depth-sorted instance list (Gaussian index per tile instance) and
 *            the sorted 64-bit tile|depth keys; `capacity` instances.
 *  scratch : sort ping-pong buffers and histograms; only live during forward.
 * ---------------------------------------------------------------------- */
TRASE_API size_t trase_b200_geom_bytes(int P);
TRASE_API size_t trase_b200_img_bytes(int W, int H);
TRASE_API size_t trase_b200_binning_bytes(int64_t capacity);
TRASE_API size_t trase_b200_scratch_bytes(int P, int64_t capacity);

/* Byte offsets of the arrays inside the three persistent workspaces, so that a
 * host (and the parity tests) can view them the way the reference's tests
 * would view geomBuffer/binningBuffer/imgBuffer.  Each call fills `out[n]`. */
enum {
  TRASE_GEOM_DEPTHS = 0,        /* f32[P]   view-space z            */
  TRASE_GEOM_MEANS2D,           /* f32[P][2] pixel centre            */
  TRASE_GEOM_CONIC_OPACITY,     /* f32[P][4] conic a,b,c + opacity   */
  TRASE_GEOM_RGBD,              /* f32[P][4] rgb (SH or precomp) + z */
  TRASE_GEOM_COV3D,             /* f32[P][6]                         */
  TRASE_GEOM_CLAMPED,           /* u8[P] bit c set: channel c clamped */
  TRASE_GEOM_TILES_TOUCHED,     /* u32[P]                            */
  TRASE_GEOM_POINT_OFFSETS,     /* u32[P] inclusive prefix sum       */
  TRASE_GEOM_BLOCK_SUMS,        /* u32[ceil(P/256)+1]                */
  TRASE_GEOM_COUNTERS,          /* u32[8]: [0]=num_rendered [1]=overflow */
  TRASE_GEOM_NUM_ARRAYS
};
enum {
  TRASE_IMG_FINAL_T = 0,        /* f32[H*W]                          */
  TRASE_IMG_N_CONTRIB,          /* u32[H*W]                          */
  TRASE_IMG_RANGES,             /* u32[T][2]                         */
  TRASE_IMG_NUM_ARRAYS
};
enum {
  TRASE_BIN_POINT_LIST = 0,     /* u32[capacity] sorted Gaussian ids */
  TRASE_BIN_POINT_LIST_KEYS,    /* u64[capacity] sorted keys         */
  TRASE_BIN_NUM_ARRAYS
};
TRASE_API void trase_b200_geom_offsets(int P, size_t* out /*[TRASE_GEOM_NUM_ARRAYS]*/);
TRASE_API void trase_b200_img_offsets(int W, int H, size_t* out /*[TRASE_IMG_NUM_ARRAYS]*/);
TRASE_API void trase_b200_binning_offsets(int64_t capacity, size_t* out /*[TRASE_BIN_NUM_ARRAYS]*/);

/* ------------------------------------------------------------------------
 * Forward.  ref: RasterizeGaussiansCUDA (DGR/rasterize_points.cu:40-127) →
 * CudaRasterizer::Rasterizer::forward (DGR/cuda_rasterizer/rasterizer_impl.cu:201-349).
 *
 * Differences from the reference, none visible in the results:
 *  - no host synchronisation inside the call: the instance count R is produced
 *    on the device; the caller provides a binning workspace of `capacity`
 *    instances and (optionally) a pinned host int that receives R with an
 *    asynchronous copy on `stream`.  If R > capacity the tile ranges are left
 *    empty, counters[1] is set and the caller must retry with a larger
 *    workspace (the Python host does this).
 *  - output images are written for every pixel (no pre-zeroing needed).
 * ---------------------------------------------------------------------- */
typedef struct trase_b200_fwd_args {
  int P;                 /* number of Gaussians                                  */
  int sh_degree;         /* active SH degree 0..3         (ref arg "degree")     */
  int M;                 /* SH coefficients stored per Gaussian (0 if no SH)     */
  int W, H;              /* image size                                            */
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int prefiltered;       /* ref: in_frustum trap, DGR/cuda_rasterizer/auxiliary.h:151-155 */
  int debug;             /* !=0: synchronise and check after every launch
                            (ref: CHECK_CUDA, DGR/cuda_rasterizer/auxiliary.h:161-171) */
  const float* background;      /* [3]                                            */
  const float* means3D;         /* [P][3]                                         */
  const float* shs;             /* [P][M][3] or NULL                              */
  const float* sh_objs;         /* [P][32]  (mandatory, as in the reference)      */
  const float* colors_precomp;  /* [P][3] or NULL                                 */
  const float* opacities;       /* [P]                                            */
  const float* scales;          /* [P][3] or NULL                                 */
  const float* rotations;       /* [P][4] or NULL                                 */
  const float* cov3D_precomp;   /* [P][6] or NULL                                 */
  const float* viewmatrix;      /* [16]                                           */
  const float* projmatrix;      /* [16]                                           */
  const float* campos;          /* [3]                                            */
  float* out_color;             /* [3][H][W]                                      */
  float* out_objects;           /* [32][H][W]                                     */
  float* out_depth;             /* [1][H][W]                                      */
  int* radii;                   /* [P]                                            */
  void* geom;    size_t geom_bytes;
  void* img;     size_t img_bytes;
  void* binning; size_t binning_bytes;
  void* scratch; size_t scratch_bytes;
  int64_t capacity;             /* instances the binning/scratch buffers hold;    */
                                /* at least P (the P-sized depth sort borrows them) */
  int* num_rendered_host;       /* pinned host int or NULL: receives R (stored by the scan kernel when the word is
                                   device-visible, i.e. pinned under unified addressing; else a 4-byte async copy) */
  void* stream;                 /* cudaStream_t                                   */
  void* readback_event;         /* cudaEvent_t or NULL: recorded right after the   */
                                /* D2H of R is queued, i.e. BEFORE the sort and    */
                                /* the compositing kernel - a host that waits on   */
                                /* it instead of the stream learns R (and whether  */
                                /* the workspace overflowed) while the rest of the */
                                /* forward pass is still running                   */
} trase_b200_fwd_args;

TRASE_API int trase_b200_rasterize_forward(const trase_b200_fwd_args* args);

/* Re-runs only key generation → sort → ranges → compositing on an existing
 * geom workspace (used after a capacity overflow). */
TRASE_API int trase_b200_rasterize_forward_rebin(const trase_b200_fwd_args* args);

/* ------------------------------------------------------------------------
 * Backward.  ref: RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:129-214)
 * → CudaRasterizer::Rasterizer::backward (DGR/cuda_rasterizer/rasterizer_impl.cu:353-455).
 * All dL_* outputs must be zero-filled by the caller (the reference does the
 * same with torch::zeros, DGR/rasterize_points.cu:165-174); dL_dconic [P][4]
 * is the internal float4 the reference keeps (DGR/cuda_rasterizer/backward.cu:566-568).
 *
 * Partial gradients (the reference always computes all of them): a NULL output
 * pointer means "not wanted", and the kernels skip the work behind it.
 *  - dL_dobjects == NULL: the per-Gaussian feature gradient is not accumulated
 *    (the feature parameter is frozen in TRASE's GAUSSIAN training state,
 *    scene/gaussian_model.py:303-315).
 *  - all other dL_d* outputs NULL, dL_dobjects set: feature gradient only (the
 *    FEATURE state, where nothing else trains): no dL/dalpha chain, no preprocess
 *    backward.  The others must be given or omitted together.
 *  - dL_dout_objects == NULL / dL_dout_color == NULL: that upstream gradient is
 *    zero (the loss did not read that image).  dL_dout_color is required when the
 *    geometry group is wanted, dL_dout_objects when only dL_dobjects is.
 * ---------------------------------------------------------------------- */
typedef struct trase_b200_bwd_args {
  int P, sh_degree, M, W, H;
  float tan_fovx, tan_fovy, scale_modifier;
  int debug;
  const float* background;
  const float* means3D;
  const float* shs;
  const float* sh_objs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int* radii;
  const void* geom;
  const void* img;
  const void* binning;
  int64_t capacity;
  const float* dL_dout_color;    /* [3][H][W]  */
  const float* dL_dout_objects;  /* [32][H][W] */
  float* dL_dmeans2D;   /* [P][3] */
  float* dL_dconic;     /* [P][4] */
  float* dL_dopacity;   /* [P]    */
  float* dL_dcolors;    /* [P][3] */
  float* dL_dobjects;   /* [P][32]*/
  float* dL_dmeans3D;   /* [P][3] */
  float* dL_dcov3D;     /* [P][6] */
  float* dL_dsh;        /* [P][M][3] or NULL */
  float* dL_dscales;    /* [P][3] or NULL */
  float* dL_drotations; /* [P][4] or NULL */
  void* stream;
} trase_b200_bwd_args;

TRASE_API int trase_b200_rasterize_backward(const trase_b200_bwd_args* args);

/* Same, restricted to a subset of its two kernels: bit 0 = compositing backward
 * (ref: BACKWARD::render, DGR/cuda_rasterizer/backward.cu:639-679), bit 1 =
 * preprocess backward (ref: BACKWARD::preprocess, backward.cu:576-637).  For
 * tests and profiling of one stage at a time. */
TRASE_API int trase_b200_rasterize_backward_stages(const trase_b200_bwd_args* args, int stages);

/* Per-stage device timing with CUDA events on the launching stream (the
 * reference has none; its CHECK_CUDA debug mode only synchronises,
 * DGR/cuda_rasterizer/auxiliary.h:161-171).  After enable(1), every forward /
 * backward call records an event pair around each stage; read() waits for the
 * recorded events, returns the number of stages and fills total_ms[i] /
 * count[i] (i: 0 preprocess_fwd, 1 scan+duplicate, 2 sort, 3 ranges,
 * 4 render_fwd, 5 render_bwd, 6 preprocess_bwd, 7 unused), then clears. */
/* Number of CUDA kernels this library has launched in this process so far (no
 * reference counterpart; bench.py reports the difference over its timed region). */
TRASE_API unsigned long long trase_b200_launch_count(void);
TRASE_API void trase_b200_profile_enable(int on);
TRASE_API int trase_b200_profile_read(double* total_ms, int* count);

/* ref: markVisible (DGR/rasterize_points.cu:216-235) → checkFrustum
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:57-69).  present[i] = 1 if the
 * view-space z of means3D[i] is > 0.2. */
TRASE_API int trase_b200_mark_visible(int P, const float* means3D, const float* viewmatrix,
                            const float* projmatrix, uint8_t* present, void* stream);

/* ------------------------------------------------------------------------
 * Stable key/value radix sort on the low `end_bit` bits of 64-bit keys.
 * ref: cub::DeviceRadixSort::SortPairs call at DGR/cuda_rasterizer/rasterizer_impl.cu:310-316.
 * `count_dev` points to the number of pairs (device u32, clamped to
 * `capacity`).  The sorted result ends up in (keys_out, vals_out); keys_in /
 * vals_in are clobbered.  Exposed so that the sort can be tested and profiled
 * on its own.
 * ---------------------------------------------------------------------- */
TRASE_API size_t trase_b200_sort_scratch_bytes(int64_t capacity);
TRASE_API int trase_b200_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out,
                          uint32_t* vals_out, const uint32_t* count_dev, int64_t capacity,
                          int end_bit, void* sort_scratch, size_t sort_scratch_bytes,
                          void* stream);

/* ------------------------------------------------------------------------
 * Fused parameter activation + rasterizer-input assembly and its backward
 * (SURVEY.md §8(f) rank 1).  ref: the elementwise torch glue of
 * gaussian_renderer/__init__.py:82-121 and scene/gaussian_model.py:181-205:
 *   means3D = xyz + d_xyz;  scales = exp(scaling) + d_scaling;
 *   rotations = normalize(rotation) + d_rotation;  opacity = sigmoid(opacity);
 *   shs = cat(f_dc[P,1,3], f_rest[P,Mrest,3]);  sh_objs = feat / (||feat|| + 1e-9).
 * Backward takes the six upstream gradients (NULL = zero) and writes the
 * gradients of the ten inputs.  rotation / feature buffers 16-byte aligned.
 * shs == NULL (forward) / g_fdc == NULL (backward): the caller's f_dc and f_rest
 * are views of one [P][Mrest+1][3] block already, so the SH copy and its
 * backward split are skipped (the caller passes that block to the rasterizer and
 * slices its gradient).  g_feat == NULL (backward): the feature parameter is frozen,
 * its gradient is not computed; the seven geometry gradients g_xyz .. g_opacity all NULL:
 * every geometry parameter is frozen (FEATURE state), that kernel is skipped.
 * accumulate_mask (backward): bit 0 xyz, 1 scaling, 2 rotation, 3 opacity, 4 SH (g_fdc and
 * g_frest) — that PARAMETER gradient is added to what the buffer holds instead of
 * overwriting it (the caller points it at the parameter's own gradient storage); the
 * gradients of the deformation inputs (g_dxyz, g_dscaling, g_drotation) are always written.
 * ---------------------------------------------------------------------- */
TRASE_API int trase_b200_assemble_forward(int P, int Mrest, const float* xyz, const float* d_xyz, const float* scaling,
                                          const float* d_scaling, const float* rotation, const float* d_rotation,
                                          const float* opacity, const float* f_dc, const float* f_rest,
                                          const float* feat, float* means3D, float* scales, float* rotations,
                                          float* opac_out, float* shs, float* sh_objs, void* stream);
TRASE_API int trase_b200_assemble_backward(int P, int Mrest, const float* scaling, const float* rotation,
                                           const float* opacity, const float* feat, const float* g_means3D,
                                           const float* g_scales, const float* g_rotations, const float* g_opac,
                                           const float* g_shs, const float* g_objs, float* g_xyz, float* g_dxyz,
                                           float* g_scaling, float* g_dscaling, float* g_rotation, float* g_drotation,
                                           float* g_opacity, float* g_fdc, float* g_frest, float* g_feat,
                                           int accumulate_mask, void* stream);

/* FEATURE-state contrastive loss on the sampled pixels without the N x N matrices.
 * ref: train.py:253-296 in the default 'soft' mode: get_pixel_mask_correspondence_matrix and get_pixel_weights
 * (utils/feature_utils.py:28-48), C_F = F F^T (:50-55), pixel_mask_correspondence_loss_soft_hard_positive /
 * _soft_negative (utils/loss_utils.py:304-349).  The reference materialises four N x N tensors (N ~ 5000) and runs
 * ~40 kernels plus two host synchronisations per loss over them; here the mask relation, the pair weights and the
 * similarities are recomputed from per-pixel data in three sweeps over the pair space.
 *   trase_b200_pixel_mask_info: masks [n_masks][HW] uint8 (n_masks <= 256), area[m] = pixels of mask m (int64),
 *     sampled_mask [n_masks] (0/1), pix [N] = flat indices of the sampled pixels  ->  bits [N][ceil(n_masks/64)]
 *     (bit m: pixel lies in sampled mask m) and a [N] = mean area of all masks covering the pixel.
 *   trase_b200_contrastive_soft_forward: f [N][32] L2-normalised sampled features; a / a_minmax ({min a, max a}, device)
 *     may be NULL (no pair weights); writes result[0..5] = {loss_pos, loss_neg, n_kept_pos, n_kept_neg, selected
 *     positive pairs, selected negative pairs} and keeps the mined-column flags in `scratch`
 *     (trase_b200_contrastive_scratch_bytes) for the backward.
 *   trase_b200_contrastive_soft_backward: df [N][32] = upstream[0] * d loss_pos/df + upstream[1] * d loss_neg/df
 *     (upstream: two device floats), from the same inputs, scratch and result. */
TRASE_API size_t trase_b200_contrastive_scratch_bytes(int N);
TRASE_API int trase_b200_pixel_mask_info(int N, int n_masks, int64_t HW, const unsigned char* masks, const int64_t* area,
                                         const unsigned char* sampled_mask, const int64_t* pix, uint64_t* bits, float* a,
                                         void* stream);
TRASE_API int trase_b200_contrastive_soft_forward(int N, int nwords, const float* f, const uint64_t* bits, const float* a,
                                                  const float* a_minmax, float pos_th, float neg_th, void* scratch,
                                                  float* result, void* stream);
TRASE_API int trase_b200_contrastive_soft_backward(int N, int nwords, const float* f, const uint64_t* bits,
                                                   const float* a, const float* a_minmax, float pos_th, float neg_th,
                                                   const void* scratch, const float* result, const float* upstream,
                                                   float* df, void* stream);

/* kNN-smoothed per-Gaussian features.  ref: GaussianModel.get_smoothed_gaussian_features
 * (scene/gaussian_model.py:79-104), the FEATURE state's default input to render()
 * (train.py:274-275 with smooth_K = 16, gaussian_renderer/__init__.py:116-119):
 *   out[p] = mean over the selected neighbours j of normalize(feat[j])     (F.normalize, eps 1e-12)
 * knn_idx [P][K] int64 is the caller's neighbour map (the reference gets it from pytorch3d.ops.knn_points and caches
 * it; the search is not part of this library); sel_cols (HOST array of n_sel <= 32 column numbers in [0, K)) are the
 * neighbours that take part in this call — the reference draws int(K*dropout) of them with torch.randperm.
 * feat, out, g_out, g_feat: [P][32] fp32 on the device.  backward overwrites g_feat.
 * unit_scratch (forward, optional [P][32] fp32): with it every row is normalised once and the neighbours' unit rows
 * are gathered (same values, same order of additions); NULL normalises each neighbour row at every use.
 * backward_csr: the same gradient as a gather over the REVERSE of the neighbour map, which the caller builds once per
 * map: rev_ptr [P+1] int32 row starts, rev_ent [P*K] int32, the entries (p << 8 | column) with knn_idx[p][column] == j
 * listed for every j (any order inside a row).  No atomics and no pre-zeroing; K <= 64, P < 2^23, no column selected
 * twice. */
TRASE_API int trase_b200_smooth_features_forward(int P, int K, const int* sel_cols, int n_sel, const float* feat,
                                                 const long long* knn_idx, float* out, float* unit_scratch,
                                                 void* stream);
TRASE_API int trase_b200_smooth_features_backward_csr(int P, int K, const int* sel_cols, int n_sel, const float* feat,
                                                      const int* rev_ptr, const int* rev_ent, const float* g_out,
                                                      float* g_feat, void* stream);
TRASE_API int trase_b200_smooth_features_backward(int P, int K, const int* sel_cols, int n_sel, const float* feat,
                                                  const long long* knn_idx, const float* g_out, float* g_feat,
                                                  void* stream);

/* ------------------------------------------------------------------------
 * DeformNetwork MLP building block.  ref: the eleven nn.Linear layers of
 * DeformNetwork.forward (utils/time_utils.py:106-131), which the reference runs
 * as cuBLAS SGEMM + separate bias/ReLU kernels, and their autograd backward.
 *
 *   D[M,N] = epilogue( A[M,K] · B[N,K]^T ),  epilogue = (+bias[n]) → (ReLU) → (· mask_in bit)
 *
 * A, B, D row-major fp32 with row pitches lda/ldb/ldd in floats (multiples of
 * 4; base pointers 16-byte aligned), N <= 256.  TF32 tensor-core products
 * (tcgen05, fp32 accumulation in TMEM).
 *   Forward layer: B = weight [out,in], bias, relu=1; mask_out (optional,
 *   [M][ceil(N/32)] uint32) receives one bit per element, set where the output
 *   is > 0, so the backward pass never re-reads activations for the ReLU
 *   derivative.
 *   Backward data gradient: A = dZ, B = weight^T, mask_in = the bit mask of the
 *   layer's input; colsum[n] += sum_m D[m,n] yields the bias gradient of the
 *   previous layer in the same pass.
 * round_tf32 != 0 rounds D to TF32 (nearest) on the way out: the tensor core
 * truncates fp32 operands to TF32, so a chain of layers stays at
 * round-to-nearest accuracy when every producer rounds.  bias, mask_in,
 * mask_out and colsum may be NULL.
 * ---------------------------------------------------------------------- */
TRASE_API int trase_b200_gemm_tf32(const float* A, int lda, const float* B, int ldb, float* D, int ldd, int M, int N,
                                   int K, const float* bias, int relu, const uint32_t* mask_in, uint32_t* mask_out,
                                   float* colsum, int round_tf32, void* stream);

/* Positional encoding of (xyz, t) — ref: Embedder.embed (utils/time_utils.py:27-57)
 * and the torch.cat at utils/time_utils.py:110 — written into columns
 * [0, 3+6*Lx+1+2*Lt) of the row-major buffer out[M][ld].  t_stride = 0 broadcasts
 * one time stamp to every row, t_stride = 1 reads t[m]. */
TRASE_API int trase_b200_deform_embed(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt,
                                      float* out, int ld, void* stream);

/* Helpers around the MLP GEMMs (each replaces several elementwise torch kernels of
 * the reference's eager path, utils/time_utils.py:106-131 and its autograd):
 *  round_transpose: Wr = tf32-round(W) [rows][cols] and WrT = tf32-round(W[:, col0:col0+ncols])^T
 *                   [ncols][rows] (either output may be NULL);
 *  pack_heads:      the three head weights [3|4|3][K] and biases -> wh [16][K], whT [K][16], bh [16];
 *  pack_dout:       the three upstream gradients [M][3|4|3] (NULL = zero) -> dout [M][16] (tf32-rounded),
 *                   colsum[16] += their column sums (= head bias gradients). */
TRASE_API int trase_b200_mlp_round_transpose(const float* W, int rows, int cols, int col0, int ncols, float* Wr,
                                             float* WrT, void* stream);
TRASE_API int trase_b200_mlp_pack_heads(const float* w0, const float* b0, const float* w1, const float* b1,
                                        const float* w2, const float* b2, int K, float* wh, float* whT, float* bh,
                                        void* stream);
TRASE_API int trase_b200_mlp_pack_dout(int M, const float* g0, const float* g1, const float* g2, float* dout,
                                       float* colsum, void* stream);

/* Weight gradient of one nn.Linear: D[i][j] += sum_m P[m][i] * Q[m][j]
 * (dW = dZ^T X; ref: autograd backward of utils/time_utils.py:113, cuBLAS SGEMM
 * in the reference).  P [M][I], Q [M][J], D [I][J] row-major fp32; I, J <= 256,
 * J a multiple of 4.  D is ACCUMULATED into (zero it first).  TF32 tcgen05
 * products read both operands straight from the row-major activations
 * (MN-major descriptors), the row range is split over the SMs. */
TRASE_API int trase_b200_wgrad_tf32(const float* P, int ldp, int I, const float* Q, int ldq, int J, int M, float* D,
                                    int ldd, void* stream);

/* The same sort for callers that cannot allocate device memory.  It has no
 * counterpart in the reference (which sizes a CUB temp buffer and has torch
 * allocate it on every call, DGR/rasterize_points.cu:30-38).
 *
 * Meant for hosts where cudaMalloc is off the table at sort time: an engine
 * that claimed the whole GPU for its own arena, a process under a hard memory
 * cap, a frame loop that may not take the driver's allocator lock (cudaMalloc
 * and cudaFree can stall every stream), or a plug-in running inside another
 * runtime's no-allocation callback.  The histogram workspace lives inside the
 * library in a fixed ~2 MiB device array that every GPU loading the library
 * holds until the process exits; the caller passes only its own key/value
 * buffers and the device count.
 *
 *  - capacity may not exceed TRASE_B200_STATIC_SORT_MAX; beyond that the call
 *    returns an error and launches nothing.
 *  - All calls on a GPU share the one workspace and see the previous call's
 *    contents; what must start at zero is cleared on `stream` at the start of
 *    every call.  Calls whose GPU work overlaps race on it: issue them one at
 *    a time (same stream, or ordered with events).
 *  - The first call on each GPU resolves the workspace's device address with
 *    that GPU current; make that first call before any stream capture begins.
 *  - Calls no cudaMalloc/cudaFree or other allocating routine. */
#define TRASE_B200_STATIC_SORT_MAX (1 << 22)
TRASE_API int trase_b200_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out,
                                           uint32_t* vals_out, const uint32_t* count_dev, int64_t capacity,
                                           int end_bit, void* stream);

/* ------------------------------------------------------------------------
 * Fused L1 + SSIM image loss and its backward.
 * ref: l1_loss (utils/loss_utils.py:30-31), ssim / _ssim with the 11x11
 * Gaussian window, sigma 1.5, zero padding 5, C1 = 0.01^2, C2 = 0.03^2
 * (utils/loss_utils.py:46-86), and the GAUSSIAN-state loss of train.py:235-243,
 * (1 - l) * L1 + l * (1 - SSIM).  The reference runs five depthwise cuDNN
 * convolutions plus ~15 elementwise kernels forward and as many again in
 * autograd.
 *
 * img, gt: [C][H][W] fp32 on the device.
 * forward writes out3 = { mean|img-gt|, mean SSIM, w_l1*L1 + w_ssim*SSIM + bias }
 * (device floats; fixed-order reduction, deterministic).  `maps` (3*C*H*W
 * floats, may be NULL when w_ssim == 0) receives the SSIM map's partial
 * derivatives for backward; `scratch` holds trase_b200_loss_scratch_bytes().
 * backward writes d_img = sum_k grad_out3[k] * d(out3[k])/d(img) (grad_out3:
 * three device floats, NULL = {0,0,1}; a term evaluated with w_ssim == 0 and
 * maps == NULL has no SSIM gradient).  No gradient is produced for gt (the
 * reference never asks for one).
 * ---------------------------------------------------------------------- */
TRASE_API size_t trase_b200_loss_scratch_bytes(int C, int H, int W);
TRASE_API int trase_b200_l1_ssim_forward(int C, int H, int W, const float* img, const float* gt, float w_l1,
                                         float w_ssim, float bias, float* maps, float* scratch, float* out3,
                                         void* stream);
TRASE_API int trase_b200_l1_ssim_backward(int C, int H, int W, const float* img, const float* gt, const float* maps,
                                          float w_l1, float w_ssim, const float* grad_out3, float* d_img,
                                          void* stream);

/* Feature-norm regulariser of the FEATURE training state (SURVEY.md §8(f) rank 3,
 * first part).  ref: train.py:280-281,
 *   rendered_feature_norm = rendered_features.norm(dim=0, p=2).mean();  reg = (1 - that)^2.
 * feat: [C][N] fp32 (N = H*W pixels).  forward writes the per-pixel norm (N
 * floats, kept for backward) and out2 = { mean norm, reg }; backward writes
 * d_feat = grad_out2[0] * d(mean)/dF + grad_out2[1] * d(reg)/dF (NULL = {0,1}),
 * zero where a pixel's norm is zero (as torch's norm backward). */
TRASE_API size_t trase_b200_feature_norm_scratch_bytes(int64_t N);
TRASE_API int trase_b200_feature_norm_reg_forward(int C, int64_t N, const float* feat, float* norm, float* scratch,
                                                  float* out2, void* stream);
TRASE_API int trase_b200_feature_norm_reg_backward(int C, int64_t N, const float* feat, const float* norm,
                                                   const float* out2, const float* grad_out2, float* d_feat,
                                                   void* stream);

/* ------------------------------------------------------------------------
 * DeformNetwork forward as one persistent kernel.  ref: DeformNetwork.forward
 * (utils/time_utils.py:106-131): nl hidden nn.Linear(256) + ReLU layers, the
 * input embedding re-concatenated after layer `skip` (embedding columns first,
 * time_utils.py:117), then the three linear heads (:119-129).
 * A 128-row tile is carried through all layers on chip; activations are written
 * to HBM once (for the backward pass) and never read back.
 *   trase_b200_mlp_fused_pack_floats: size of the packed weight buffer.
 *   trase_b200_mlp_pack_fused: weights[i] (host array of nl device pointers) =
 *     layer i's [256][in_i] matrix, wh = the packed [16][256] head matrix
 *     (trase_b200_mlp_pack_heads) -> TF32-rounded k-block layout.
 *   trase_b200_mlp_forward_fused: emb [M][ld_emb] (E columns used), bias =
 *     nl*256 hidden biases then 16 head biases; writes acts [nl][M][256]
 *     (TF32-rounded post-ReLU activations), masks [nl][M][8] (bit = output > 0)
 *     and out16 [M][16] (heads: 3 warp, 4 rotation, 3 scaling, 6 zero).  acts and masks
 *     both NULL = inference: nothing is kept for a backward pass.  d_xyz [M][3],
 *     d_rot [M][4] (16-byte aligned), d_scale [M][3]: optional, all three or none; with
 *     them the heads are written as those three contiguous tensors (what the caller hands
 *     to the rasterizer glue) instead of the packed rows, and out16 may be NULL.
 * Restrictions: width 256, 2 <= nl <= 11, E <= 96 and a multiple of 4,
 * skip < nl - 1 (or -1 for none).
 * ---------------------------------------------------------------------- */
TRASE_API size_t trase_b200_mlp_fused_pack_floats(int nl, int skip, int E);
TRASE_API int trase_b200_mlp_pack_fused(int nl, int skip, int E, const float* const* weights, const float* wh,
                                        float* wpack, void* stream);
TRASE_API int trase_b200_mlp_forward_fused(int M, int nl, int skip, int E, const float* emb, int ld_emb,
                                           const float* wpack, const float* bias, float* acts, uint32_t* masks,
                                           float* out16, float* d_xyz, float* d_rot, float* d_scale, void* stream);
/* The matching data-gradient chain of autograd's backward through the same layers,
 *   dZ_{l-1} = (dZ_l W_l[:, hidden columns]) * relu'(h_{l-1}),  db_l = column sums of dZ_l,
 * again one persistent kernel: the gradient tile stays on chip between layers and is
 * written to HBM once per layer (the weight-gradient GEMMs read it from there).
 *   trase_b200_mlp_pack_fused_bwd: whT = packed head matrix transposed [256][16]
 *     (trase_b200_mlp_pack_heads), weights = the same host array of nl raw weight
 *     pointers as trase_b200_mlp_pack_fused (read transposed, hidden columns only).
 *   trase_b200_mlp_backward_fused: dout [M][16] (trase_b200_mlp_pack_dout), masks from the
 *     forward; writes dzs [nl][M][256] with dzs[L] = dZ of layer nl-1-L, and adds the bias
 *     gradient of layer l to colsum[l] (host array of nl device pointers, 256 floats each). */
TRASE_API size_t trase_b200_mlp_fused_pack_bwd_floats(int nl);
TRASE_API int trase_b200_mlp_pack_fused_bwd(int nl, int skip, int E, const float* whT, const float* const* weights,
                                            float* wpack, void* stream);
TRASE_API int trase_b200_mlp_backward_fused(int M, int nl, const float* dout, const float* wpack,
                                            const uint32_t* masks, float* dzs, float* const* colsum, void* stream);
/* BF16 side outputs for the weight gradients.  ref: the autograd backward of the eleven nn.Linear layers of
 * utils/time_utils.py:106-131 (dW = dZ^T X, cuBLAS SGEMM in the reference).  The activations X_l and the layer
 * gradients dZ_l that the two whole-network kernels write to HBM have exactly one reader, the weight-gradient
 * product, which is bound by streaming them.  These variants write them as BF16 ([nl][M][256], 16-byte aligned)
 * instead of fp32 — half the bytes written and read; the layer chain itself is unchanged (TF32 operands in shared
 * memory, fp32 accumulation) — and trase_b200_wgrad_bf16 consumes them (BF16 products, fp32 accumulation in TMEM,
 * D accumulated into like trase_b200_wgrad_tf32).  trase_b200_f32_to_bf16 converts the two small operands no
 * fused kernel produces (the positional encoding, the packed head gradient): dst [rows][ld_dst] bf16, columns
 * >= cols zero-filled. */
TRASE_API int trase_b200_mlp_forward_fused_bf16(int M, int nl, int skip, int E, const float* emb, int ld_emb,
                                                const float* wpack, const float* bias, void* acts16,
                                                uint32_t* masks, float* out16, float* d_xyz, float* d_rot,
                                                float* d_scale, void* stream);
TRASE_API int trase_b200_mlp_backward_fused_bf16(int M, int nl, const float* dout, const float* wpack,
                                                 const uint32_t* masks, void* dzs16, float* const* colsum,
                                                 void* stream);
TRASE_API int trase_b200_wgrad_bf16(const void* P, int ldp, int I, const void* Q, int ldq, int J, int M, float* D,
                                    int ldd, void* stream);
TRASE_API int trase_b200_f32_to_bf16(int rows, int cols, const float* src, int ld_src, void* dst, int ld_dst,
                                     void* stream);
/* All weight-gradient products of one backward pass (ref: the same eleven autograd GEMMs) in ONE launch:
 * D_j[i][k] += sum_m P_j[m][i] * Q_j[m][k] for j < njobs (<= 12) over the same M rows, per-product arguments as
 * for trase_b200_wgrad_bf16 (host arrays of njobs entries; the pointers in them are device pointers).  The SMs are
 * dealt out to the products in proportion to the bytes each streams; each CTA reduces its row slice of one product
 * in TMEM and adds it to D once. */
TRASE_API int trase_b200_wgrad_bf16_multi(int njobs, const void* const* P, const int* ldp, const int* I,
                                          const void* const* Q, const int* ldq, const int* J, int M, float* const* D,
                                          const int* ldd, void* stream);
/* The two producers of those small operands can write the bf16 copy themselves (no separate conversion pass):
 * the positional encoding (out16 [M][ld16], 16-byte aligned, ld16 a multiple of 8 and >= the column count, padding
 * zero-filled) and the
 * packed head gradient, as two bf16 terms per value: dout16 [M][32] = [hi(16) | lo(16)], hi = bf16(v),
 * lo = bf16(v - hi).  The head weight gradient is then the product with all 32 columns, D [256][32], and
 * dWh^T = D[:, :16] + D[:, 16:] (the upstream gradient comes out of atomics; one 8-bit rounding of it would turn
 * last-bit run-to-run differences into 0.4 % steps). */
TRASE_API int trase_b200_deform_embed_bf16(int M, const float* xyz, const float* t, int t_stride, int Lx, int Lt,
                                           float* out, int ld, void* out16, int ld16, void* stream);
TRASE_API int trase_b200_mlp_pack_dout_bf16(int M, const float* g0, const float* g1, const float* g2, float* dout,
                                            void* dout16, float* colsum, void* stream);
/* Head gradients straight into the parameters' gradient buffers (ref: autograd's accumulation of the three heads'
 * weight and bias gradients, utils/time_utils.py:119-129): dwhT [K][ld] is the head weight gradient transposed (ld = 16,
 * halves = 1; or ld = 32, halves = 2 for the bf16 path's hi and lo products), dbh [16] the bias gradient; the three
 * heads' [rows][K] weight gradients and bias gradients are ACCUMULATED into.  One launch. */
TRASE_API int trase_b200_mlp_heads_grad_accumulate(const float* dwhT, int ld, int halves, const float* dbh, float* gw0,
                                                   float* gb0, float* gw1, float* gb1, float* gw2, float* gb2, int K,
                                                   void* stream);
/* Debug only (no reference counterpart): device buffer of [grid][8] uint64 stall-cycle counters that the
 * whole-network kernels fill when set (NULL = off); read by tools/check_fused_mlp.py. */
TRASE_API void trase_b200_mlp_fused_set_prof(unsigned long long* dev_buf);

/* ------------------------------------------------------------------------
 * Fused Adam over flat arenas and the densification statistics (SURVEY.md
 * §8(f) rank 4).  ref: torch.optim.Adam(groups, lr=0.0, eps=1e-15) as set up in
 * scene/gaussian_model.py:259-289 and scene/deform_model.py:38-43 (torch's
 * single-tensor update, op for op), and GaussianModel.add_densification_stats
 * (scene/gaussian_model.py:637-639).
 * params / grads / exp_avg / exp_avg_sq: n fp32 each.  Segments (host arrays of
 * nseg <= 64 entries): seg_end[i] = exclusive end offset, step_size[i] =
 * lr_i / (1 - beta1^t_i), bc2_sqrt[i] = sqrt(1 - beta2^t_i); a segment with
 * step_size == 0 and bc2_sqrt == 0 is left untouched.
 * densification_stats: grad_accum[i] += ||viewspace_grad[i][0:2]||, denom[i] += 1
 * where filter[i] (bool bytes). */
TRASE_API int trase_b200_adam_step(int64_t n, float* params, const float* grads, float* exp_avg, float* exp_avg_sq,
                                   int nseg, const int64_t* seg_end, const float* step_size, const float* bc2_sqrt,
                                   float beta1, float beta2, float eps, void* stream);
TRASE_API int trase_b200_densification_stats(int P, const float* viewspace_grad, const unsigned char* filter,
                                             float* grad_accum, float* denom, void* stream);
/* Densification / pruning over the flat arenas.  ref: cat_tensors_to_optimizer, _prune_optimizer, prune_points and
 * densification_postfix (scene/gaussian_model.py:472-561), which re-allocate every parameter tensor and both Adam
 * moments one by one with torch.cat / boolean indexing.  Here the host builds ONE row map (output row r <- input row
 * src_rows[r]; fresh[r] = 1 for clones and split children, whose moments start at zero) and one launch gathers all
 * per-Gaussian segments of the parameter buffer and of both moment buffers.  Segment s is a [P][width[s]] row-major
 * block at float offset old_off[s] (input) / new_off[s] (output); old_m/old_v/new_m/new_v may all be NULL. */
TRASE_API int trase_b200_arena_gather_rows(int nseg, const int64_t* old_off, const int64_t* new_off, const int* width,
                                           int64_t P_new, const int64_t* src_rows, const unsigned char* fresh,
                                           const float* old_p, const float* old_m, const float* old_v, float* new_p,
                                           float* new_m, float* new_v, void* stream);


/* Feature-loss front end (second part): bilinear resize of the
 * rendered feature image to the mask resolution, gather of the sampled pixels and L2
 * normalisation, without materialising the resized image.
 * ref: train.py:283 (F.interpolate(..., mode='bilinear'), align_corners=False) and
 * utils/feature_utils.py:50-53 (rendered_features[:, sampled_pixel] -> normalize).
 * feat [C][H][W] (C <= 64); idx: N flat indices (row-major, int64) into the Hm x Wm
 * mask grid; out [N][C]; inv_norm [N] kept for backward (may be NULL when
 * normalize == 0).  backward ADDS into d_feat [C][H][W] (zero it first). */
TRASE_API int trase_b200_sample_features_forward(int C, int H, int W, int Hm, int Wm, int N, const float* feat,
                                                 const int64_t* idx, int normalize, float* out, float* inv_norm,
                                                 void* stream);
TRASE_API int trase_b200_sample_features_backward(int C, int H, int W, int Hm, int Wm, int N, const int64_t* idx,
                                                  int normalize, const float* out, const float* inv_norm,
                                                  const float* g_out, float* d_feat, void* stream);

/* ref: getHigherMsb (DGR/cuda_rasterizer/rasterizer_impl.cu:38-53). */
TRASE_API uint32_t trase_b200_higher_msb(uint32_t n);

/* Kernel-variant selection for benchmarking (no reference counterpart; the
 * reference fixes its launch shapes at compile time, DGR/cuda_rasterizer/config.h).
 * Keys: "fwd_ppt", "bwd_ppt" (pixels per thread: 1 or 2), "sort_staged" (0 | 1:
 * shared-memory reorder in the radix downsweep), "reserve_sms" (SMs the persistent MLP
 * kernels leave free for a concurrent collective). Results do not depend
 * on the variant beyond floating-point summation order. */
TRASE_API void trase_b200_set_tuning(const char* key, int value);

#ifdef __cplusplus
}
#endif
#endif /* TRASE_B200_H_ */
