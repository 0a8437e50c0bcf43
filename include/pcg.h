/* pcg.h -- C ABI of libpcg.so: the B200 (sm_100a) point-cloud-codec graft.
 *
 * One shared object of hand-written CUDA kernels that replaces, for the learned codec's hot
 * path, the native extensions the reference reaches through two un-vendored Python packages:
 *   - MinkowskiEngine's pybind `_C` module (coordinate maps, kernel maps, sparse convolution,
 *     pruning, interpolation)            -- reference call sites: model/transforms.py:32-166,
 *     model/entropy_models.py:28-191,294,381,446, model/blocks.py:53, model/model.py:66-70,
 *     147-161,227, utils.py:159, loss.py:105,185-188,221-271
 *   - compressai's entropy-model arithmetic (EntropyBottleneck / GaussianConditional
 *     likelihood, quantise, CDF index)   -- model/entropy_models.py:272-327,371-400,438-484
 * Each entry point cites the reference interface it replaces.  See INTEGRATION.md for the
 * ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name starts with h_; the caller allocates
 *     every buffer (no ownership transfer); data-dependent sizes use a device counter
 *     (`d_n_out`) the caller reads back, with the output buffer sized for the stated bound;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value 0 = ok, non-zero = error; pcg_last_error() gives a thread-local message;
 *   - coordinates are rows of 4 x int32 (batch, x, y, z), 16-byte aligned, with
 *     batch in [0, 32767] and x,y,z in [-32768, 32767] (they are packed into one 64-bit key).
 *     MinkowskiEngine takes any int32; here a row outside that range is NOT inserted and bit 2
 *     (value 2) of the call's `d_status` word is raised, which the host layer turns into an error
 *     (a wrapped key would silently merge two voxels);
 *   - features are row-major fp32 (N x C) unless a bf16 shadow is named;
 *   - kernel offsets follow MinkowskiEngine: odd k -> {-(k/2)..k/2}, even k -> {0..k-1},
 *     offset index i = ix + k*iy + k*k*iz;
 *   - a neighbour table `nbr` is int32 [K][pitch], pitch = pcg_nbr_pitch(n_out) (n_out rounded up to 128 so
 *     that every 128-row tile of every kernel offset is one aligned 512-byte line): nbr[i][o] = the input row
 *     feeding output row o through kernel offset i, or -1; the padding entries are -1.
 */
#ifndef PCG_H
#define PCG_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCG_ACT_NONE 0
#define PCG_ACT_RELU 1
#define PCG_ACT_LEAKY_RELU 2 /* slope 0.01 (ME.MinkowskiLeakyReLU default) */

const char *pcg_last_error(void);
int pcg_version(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
long pcg_launch_count(void);

/* ---- scratch sizing (host helpers, no GPU work) ------------------------------------------ */
long pcg_hash_capacity(long n_keys);            /* power of two >= 2*n_keys (min 1024) */
long pcg_nbr_pitch(long n_out);                 /* row pitch (elements) of a neighbour table over n_out output rows */
size_t pcg_scratch_bytes(long n_items);         /* scratch needed by any compaction/select over n_items */

/* ---- coordinate hash: replaces ME CoordinateMapGPU::insert_and_map (ME1) -------------------
 * keys: uint64[cap], vals: int32[cap].  After build, lookup(c) = smallest row with coordinate c.
 * d_status (optional int32 word): OR-ed with 2 when a row lies outside the packing range; this call never clears it
 * (a caller may keep one sticky word per stream and read it at its next synchronisation point). */
int pcg_hash_build(const int32_t *coords, long n, uint64_t *keys, int32_t *vals, long cap, int32_t *d_status, void *stream);
int pcg_hash_lookup(const uint64_t *keys, const int32_t *vals, long cap, const int32_t *query, long nq,
                    int32_t *rows_out, void *stream);

/* First-occurrence unique rows (ME SparseTensor ctor dedupe; ME.utils.sparse_quantize, ME9).
 * out_coords: up to n rows; out_index: source row of every kept row; table ends up mapping
 * unique coordinate -> new row. */
int pcg_unique_first(const int32_t *coords, long n, uint64_t *keys, int32_t *vals, long cap,
                     int32_t *out_coords, int32_t *out_index, int32_t *d_n_out, int32_t *d_status, void *scratch,
                     size_t scratch_bytes, void *stream);    /* d_status (optional): cleared, then 2 = row out of range */

/* Strided output coordinates floor(c/step)*step, first-occurrence order, + their hash table
 * (ME CoordinateMapGPU::stride; used by every stride-2 MinkowskiConvolution and by
 * transforms.py:166 down_conv whose output coordinates alone are consumed, a6). */
int pcg_stride_coords(const int32_t *coords, long n, int step, uint64_t *keys, int32_t *vals, long cap,
                      int32_t *out_coords, int32_t *d_n_out, int32_t *d_status, void *scratch, size_t scratch_bytes,
                      void *stream);                         /* d_status (optional): cleared, then 2 = row out of range */

/* Generative up-sampling coordinates unique{p + off_i*ts_out} over all rows and all ksize^3
 * offsets + their hash table.  Row order: first occurrence in an enumeration that groups the offsets by
 * their residue class modulo `stride` (the child's parity class for stride 2), classes ascending, and
 * inside a class input-row major / offset minor -- the SET equals ME's, the order is chosen so that a
 * 128-row tile of the following convolution touches one class only
 * (ME CoordinateMapGPU::stride_region; MinkowskiGenerativeConvolutionTranspose, ME3).
 * out_coords must hold max_out rows; *d_n_out > max_out or a full table sets d_status[0]=1. */
int pcg_generate_coords(const int32_t *coords, long n, int ksize, int stride, int ts_out, uint64_t *keys,
                        int32_t *vals, long cap, int32_t *out_coords, long max_out, int32_t *d_n_out,
                        int32_t *d_status, void *scratch, size_t scratch_bytes, void *stream);

/* Kernel map (ME CoordinateMapGPU::kernel_map): nbr[i][o] = row of (out_coords[o] +/- off_i*ts)
 * in the table of the INPUT coordinates.  transposed=0: in = out + off*ts (ts = input stride);
 * transposed=1: in = out - off*ts (ts = output stride).  tile_mask (optional, may be NULL):
 * uint32 [ceil(n_out/128)][4], bit i set when any row of the 128-row tile has a neighbour at
 * offset i (ksize^3 <= 128 only).  in_lattice > 1 declares that input coordinates are multiples of
 * in_lattice (transposed conv: the input tensor stride), letting impossible offsets skip the probe. */
int pcg_kernel_map(const uint64_t *keys, const int32_t *vals, long cap, const int32_t *out_coords, long n_out,
                   int ksize, int ts, int transposed, int in_lattice, int32_t *nbr, long nbr_pitch, uint32_t *tile_mask,
                   void *stream);

/* Row sort permutation.  morton=0: lexicographic (b,x,y,z) -- utils.py:142-187 sort_tensor / sort_points
 * (a11), the order that defines the bitstream's symbol order.  morton=1: (b, Z-order of x,y,z) -- used once
 * per block at the codec's entry so that every derived row order is spatially coherent. */
int pcg_sort_rows(const int32_t *coords, long n, int morton, int32_t *perm, void *scratch, size_t scratch_bytes,
                  void *stream);
/* The same stable sort restricted to key bits [begin_bit, end_bit) (keys: morton = 0: b:16|x:16|y:16|z:16, coordinates biased by
 * 32768; morton = 1: b:16 | 48 interleaved bits).  morton = 1 with bits [0, 36) orders by the low 12 bits of every axis: a
 * locality-only Z-order (row order is free in MinkowskiEngine semantics) in 5 radix passes instead of 8. */
int pcg_sort_rows_bits(const int32_t *coords, long n, int morton, int begin_bit, int end_bit, int32_t *perm, void *scratch,
                       size_t scratch_bytes, void *stream);
size_t pcg_sort_rows_scratch_bytes(long n);

/* Block entry (model/model.py:142-161): (n,6) float rows [x y z r g b] -> coords[i] = (batch, floor(xyz[perm[i]])),
 * feats[i] = (1, rgb[perm[i]]); perm may be NULL (identity). */
int pcg_gather_block(const float *cloud, const int32_t *perm, long n, int batch, int32_t *coords, float *feats, void *stream);

/* ---- sparse convolution: replaces ME ConvolutionForwardGPU / ConvolutionTransposeForwardGPU
 * (ME2-ME4).  out[o] = act(bias + sum_i F[nbr[i][o]] @ W[i]); accumulation order is fixed
 * (offset index ascending, then input channel), so results are bit-reproducible run to run. */
int pcg_conv_fwd_f32(const float *feats, long n_in, int cin, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                     const float *weight, const float *bias, int cout, int act, float *out, void *stream);

/* Weight gradient dW[k] = sum_o F[nbr[k][o]]^T (x) G[o] (ME ConvolutionBackwardGPU; train.py:195-234), fp32, fixed
 * summation order.  tile_mask (optional, K <= 128) skips empty (tile, offset) combinations.  The input gradient is the
 * forward convolution of G over the inverse kernel map with W[k]^T and needs no entry point of its own. */
int pcg_conv_bwd_weight_f32(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, const uint32_t *tile_mask,
                            int K, long n_out, const float *gout, int cout, float *dweight, void *stream);
/* The same with the tiles split into parallel slices (scratch of pcg_conv_bwd_weight_f32_scratch_bytes() bytes; partial sums
 * added in ascending slice order: still one fixed summation order).  Serves the layers the tensor-core kernel does not take. */
size_t pcg_conv_bwd_weight_f32_scratch_bytes(int K, int cin, int cout, long n_out);
int pcg_conv_bwd_weight_f32_sliced(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, const uint32_t *tile_mask,
                                   int K, long n_out, const float *gout, int cout, void *scratch, size_t scratch_bytes,
                                   float *dweight, void *stream);
/* The same gradient on the tensor cores (tcgen05, bf16 operands, fp32 accumulation in TMEM): per offset a GEMM whose reduction
 * dimension is the row index, both operands taken MN-major straight from the gathered row tiles.  feats_bf16: [n_in][roundup64(cin)],
 * gout_bf16: [n_out][roundup64(cout)] (both zero padded); cin <= 256, cout <= 256, K <= 128, tile_mask required.  Tiles are
 * split into slices whose partial sums (scratch) are added in ascending order: deterministic.  dweight: fp32 [K][cin][cout]. */
size_t pcg_conv_bwd_weight_tc_scratch_bytes(int K, int cin, int cout, long n_out);
int pcg_conv_bwd_weight_tc(const void *feats_bf16, int cin, const void *gout_bf16, int cout, const int32_t *nbr, long nbr_pitch,
                           const uint32_t *tile_mask, int K, long n_out, void *scratch, size_t scratch_bytes, float *dweight,
                           void *stream);

/* bf16 tensor-core path (tcgen05.mma, fp32 accumulation in TMEM, operands gathered by TMA tile::gather4).
 * 1 <= cout <= 256, K <= 128, any cin >= 1, n_in >= 1.  feats_bf16: n_in rows of pcg_tc_cin_padded(cin) bf16
 * (16, 32 or a multiple of 64 channels, zero padded; pcg_f32_to_bf16_rows makes them); wpacked from
 * pcg_pack_weights_bf16.  cout is padded to the 16-column MMA granule internally, so the occupancy / colour
 * heads and the 4-channel input layer run on the same kernel.  out_f32 and/or out_bf16 (cout % 16 == 0 only)
 * may be NULL. */
int pcg_tc_cin_padded(int cin);
size_t pcg_packed_weight_bytes(int K, int cin, int cout);
int pcg_pack_weights_bf16(const float *weight, int K, int cin, int cout, void *wpacked, void *stream);
int pcg_conv_fwd_tc(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch,
                    const uint32_t *tile_mask, int K, long n_out, const void *wpacked, const float *bias, int cout,
                    int act, float *out_f32, void *out_bf16, void *stream);
/* Occupancy head (model/transforms.py:146-160: conv k3 C -> cout, ReLU, conv k3 cout -> 1) with the second layer taken
 * GEMM-first out of the first layer's epilogue: pcg_conv_fwd_tc_proj is pcg_conv_fwd_tc (cout = 16, 32 or 64; cin >= 17)
 * that also writes proj_out[k][o] = act(bias + conv(o)) . proj_w[k][:] for the proj_k = 27 offsets of the second layer
 * (proj_w: the second layer's kernel as [27][cout] fp32, and bias_host: the FIRST layer's bias as [cout] fp32 or NULL, both in
 * HOST memory -- they are copied into the launch's kernel parameters so that the epilogue's multiply-adds read them from the
 * constant bank; every other pointer of this header is device memory; proj_out: [27][proj_pitch] fp32, proj_pitch >= n_out),
 * from the fp32 accumulator row; out_f32 / out_bf16 may both be NULL (the hidden rows are then never written).  cout = 16, 32 or 64.
 * pcg_gather_sum finishes the second layer on the same 3x3x3 neighbour table: out[o] = act(bias[0] + sum_k planes[k][nbr[k][o]]),
 * k ascending, absent neighbours skipped (K must be 27; bias: device pointer or NULL). */
int pcg_conv_fwd_tc_proj(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch,
                         const uint32_t *tile_mask, int K, long n_out, const void *wpacked, const float *bias_host, int cout,
                         int act, float *out_f32, void *out_bf16, const float *proj_w, int proj_k, float *proj_out,
                         long proj_pitch, void *stream);
int pcg_gather_sum(const float *planes, long plane_pitch, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                   const float *bias, int act, float *out, void *stream);
/* Layers with 128-byte operand rows (cin >= 33) and cout padded to a multiple of 32 run on the LSU-gather kernel
 * (k_conv_tc_mt: 16-byte cp.async gathers, one or two output tiles per CTA sharing each weight slice); grids of at least
 * `tiles` 128-row tiles take the two-tile form (default 4 x 148, env PCG_TC_MT_TILES).  0 = use the TMA gather4 kernel
 * for everything; < 0 = query only.  Returns the previous threshold.  Results do not depend on it (same per-element
 * accumulation order in every variant). */
long pcg_tc_set_mt_min_tiles(long tiles);
int pcg_f32_to_bf16(const float *src, void *dst, long n, void *stream);
/* n rows of c floats -> n rows of c_padded >= c bf16, zero padded. */
int pcg_f32_to_bf16_rows(const float *src, long n, int c, int c_padded, void *dst, void *stream);

/* im2col for layers with a tiny C_in (the 4-channel input layer, model/transforms.py:33): col[o][k*cin + c] =
 * F[nbr[k][o]][c] as bf16, zeros where absent, row_elems (a multiple of cin, >= K*cin) elements per row.  The layer
 * is then the dense product col @ W.reshape(K*cin, cout): pcg_conv_fwd_tc with K = 1 and the identity map. */
int pcg_im2col_bf16(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, int K, long n_out, int row_elems,
                    void *col_bf16, void *stream);

/* Narrow occupancy-head layers on bf16 rows gathered by 8/16-byte loads straight into mma.sync A fragments
 * (csrc/pcg_conv_narrow.cu; model/transforms.py:146-160).  Same arithmetic contract as pcg_conv_fwd_tc (bf16 operands,
 * fp32 accumulation, fixed order).  C_in <= 64 (rows of pcg_tc_cin_padded(cin) bf16), C_out <= 16; wfrag from
 * pcg_pack_weights_mma; out_bf16 (optional) is written as rows of 16 bf16, zero padded. */
size_t pcg_mma_weight_bytes(int K, int cin, int cout);
int pcg_pack_weights_mma(const float *weight, int K, int cin, int cout, void *wfrag, void *stream);
int pcg_conv_fwd_mma(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                     const void *wfrag, const float *bias, int cout, int act, float *out_f32, void *out_bf16, void *stream);

/* Dense bf16 product on tcgen05 (csrc/pcg_gemm_tc.cu): out_bf16[n][0..n_cols) = feats[n][:] @ Wmat[:, 0..n_cols)
 * (fp32 accumulation in TMEM, bf16 result, row pitch n_cols).  feats_bf16: n rows of roundup64(cin) bf16, 33 <= cin <= 128.
 * Wmat's columns are given in groups of 256: wpacked = pcg_pack_weights_bf16 of a (G, cin, 256) tensor,
 * G = ceil(n_cols/256), zero padded; n_cols % 16 == 0.
 * First half of the GEMM-first generative transposed convolution: T[p][k][:] = F[p] @ W[k] for all k at once
 * (ME ConvolutionTransposeForwardGPU, ME3: every (parent, offset) pair hits, so nothing needs gathering). */
int pcg_gemm_bf16_grouped(const void *feats_bf16, long n, int cin, const void *wpacked, long n_cols, void *out_bf16,
                          void *stream);

/* ---- octree-cell generative up-sampling, kernel 5 / stride 2 (ME3; model/transforms.py:129-137) -------------
 * Children (stride ts_out = ts/2) of parents (stride ts) grouped by CELL = ts-aligned cube of <= 8 children, child
 * bit b = bx + 2 by + 4 bz.  Build order:
 *   1. cells       = pcg_generate_coords(parents, ksize 3, stride 1, ts_out = ts)     (+ the cell table)
 *   2. pnbr        = pcg_kernel_map(parent table, cells, ksize 3, ts, transposed 0)   (27 x pitch(n_cells))
 *   3. pcg_cells_info  -> cell_mask (existing children), cell_base (first child row), *d_n_out (number of children)
 *   4. pcg_cells_emit  -> child coordinates (cell-major, bit order) and out_cell (cell of every child)
 * The child SET equals ME's unique{p + off*ts_out : off in [-2,2]^3}; the row order is libpcg's own. */
/* pnbr[e][cell] (27 x pnbr_pitch, -1 = absent) by scatter: one probe of the cell table per (parent, offset) pair instead of
 * 27 probes of the parent table per cell -- same table as pcg_kernel_map(parents table, cells, 3) produces.  cell_keys /
 * cell_vals: coordinate hash of the cells (pcg_generate_coords(ksize 3, stride 1) of the parents).  *d_status bit 0: a
 * dilated parent position is missing from the cell table (the cells are not the 3^3 dilation of these parents). */
int pcg_cells_parent_map(const int32_t *parent_coords, long n_par, int ts, const uint64_t *cell_keys, const int32_t *cell_vals,
                         long cell_cap, long n_cells, int32_t *pnbr, long pnbr_pitch, int32_t *d_status, void *stream);
size_t pcg_cells_scratch_bytes(long n_cells);
int pcg_cells_info(const int32_t *pnbr, long pnbr_pitch, long n_cells, uint8_t *cell_mask, int32_t *cell_base,
                   int32_t *d_n_out, void *scratch, size_t scratch_bytes, void *stream);
int pcg_cells_emit(const int32_t *cell_coords, const uint8_t *cell_mask, const int32_t *cell_base, long n_cells, int ts_out,
                   int32_t *out_coords, int32_t *out_cell, void *stream);
/* 3x3x3 stride-1 kernel map of the children among themselves -- identical to pcg_kernel_map over a table of the
 * children, derived from 7 probes of the (8x smaller) cell table per child. */
int pcg_cells_kernel_map3(const uint64_t *cell_keys, const int32_t *cell_vals, long cell_cap, const int32_t *out_coords,
                          const int32_t *out_cell, long n_out, int ts_out, const uint8_t *cell_mask, const int32_t *cell_base,
                          int32_t *nbr, long nbr_pitch, uint32_t *tile_mask, void *stream);
/* Second half of the GEMM-first transposed convolution: out[child] = act(bias + sum_e T[pnbr[e][cell]][k(b - 2e)][:]),
 * e ascending over the parents that reach the child (fixed order, fp32 accumulation).  T: bf16 [n_parents][125][cout],
 * cout in {32, 64, 128, 256}.  Child rows are addressed through cell_base / cell_mask (row = base + rank of bit b). */
int pcg_cells_reduce(const void *t_bf16, int cout, const int32_t *pnbr, long pnbr_pitch, const uint8_t *cell_mask,
                     const int32_t *cell_base, long n_cells, const float *bias, int act, float *out_f32, void *out_bf16,
                     void *stream);

/* Channel-wise convolution out[o][c] = sum_i F[nbr[i][o]][c] * w[i][c or 0]
 * (ME.MinkowskiChannelwiseConvolution, loss.py:185-188; w_cols = C or 1). */
int pcg_channelwise_conv(const float *feats, int c, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                         const float *w, int w_cols, float *out, void *stream);

/* Elementwise activation on features (ME.MinkowskiReLU / MinkowskiLeakyReLU, ME7). */
int pcg_activation(const float *x, long n, int act, float *out, void *stream);

/* GDN1 / inverse GDN1 (model/blocks.py:26-57): norm = beta + |x| @ gamma^T; out = x/norm or x*norm.
 * beta (C), gamma (C x C) are the re-parametrised values. */
int pcg_gdn_fwd(const float *x, long n, int c, const float *beta, const float *gamma, int inverse, float *out,
                void *stream);
/* The same layer at the precision of the tensor-core mode: |x| gamma^T as one bf16 product with fp32 accumulation (c = 32, 64, 128),
 * and optionally (out_bf16 != NULL) the output rows a second time as the bf16 operand image of the next tensor-core layer. */
int pcg_gdn_fwd_bf16(const float *x, long n, int c, const float *beta, const float *gamma, int inverse, float *out, void *out_bf16,
                     void *stream);

/* Decoder output rows (model/model.py:247-250): out[i] = [x, y, z, clamp(round(f * 255), 0, 255) / 255], (n, 3 + c) fp32. */
int pcg_finalize(const int32_t *coords, const float *feats, int c, long n, float *out, void *stream);

/* ---- occupancy top-k + prune (transforms.py:228-282, ME.MinkowskiPruning ME5) ---------------
 * mask[i] = 1 for the k rows of batch `batch` with the largest logits (ties: smaller row first);
 * rows of other batches are left untouched. */
int pcg_topk_mask(const float *logits, const int32_t *coords, long n, int batch, long k, uint8_t *mask,
                  void *scratch, size_t scratch_bytes, void *stream);
/* The same selection for ALL batch items in one set of passes (n_batches <= 32; h_k: HOST array of n_batches counts; a row's
 * batch item is its first coordinate; rows of other batch indices are left untouched).  One histogram pass reads every row once
 * whatever n_batches is; ties at an item's threshold are taken in row order.  scratch >= pcg_topk_mask_batched_scratch_bytes(n). */
size_t pcg_topk_mask_batched_scratch_bytes(long n);
int pcg_topk_mask_batched(const float *logits, const int32_t *coords, long n, int n_batches, const long *h_k, uint8_t *mask,
                          void *scratch, size_t scratch_bytes, void *stream);
/* Order-preserving row compaction of coords (may be NULL) and features (may be NULL) under mask. */
int pcg_prune_rows(const int32_t *coords, const float *feats, long n, int c, const uint8_t *mask,
                   int32_t *out_coords, float *out_feats, int32_t *d_n_out, void *scratch,
                   size_t scratch_bytes, void *stream);
/* Index form of the prune: ascending source rows of the kept entries of `mask` (n_out <= n), so that the caller can
 * compose the prune with a re-ordering of the survivors (transforms.py:257-282 followed by a row permutation is free in
 * ME semantics) and move the feature rows once with pcg_gather_rows.  scratch >= pcg_scratch_bytes(n). */
int pcg_mask_to_rows(const uint8_t *mask, long n, int32_t *out_rows, int32_t *d_n_out, void *scratch, size_t scratch_bytes,
                     void *stream);
/* The same compaction WITHOUT caller scratch and without any allocation (no cudaMalloc / cudaFree / allocating library call): the
 * scratch is a file-scope device array inside libpcg.so, sized for n <= PCG_MASK_TO_ROWS_STATIC_MAX items (4.0 MiB, held by every
 * GPU that loads the library until exit); larger n is rejected with an error.  For callers that cannot allocate: code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc.  Output identical to pcg_mask_to_rows.  All calls share
 * the array (its zero-initialised words are cleared on `stream` at every call): calls that OVERLAP on the GPU race -- keep them on
 * one stream or order them.  No counterpart in the reference. */
#define PCG_MASK_TO_ROWS_STATIC_MAX (4L * 1024 * 1024)
int pcg_mask_to_rows_static(const uint8_t *mask, long n, int32_t *out_rows, int32_t *d_n_out, void *stream);
/* out_coords[r] = coords[rows[r]], out_feats[r] = feats[rows[r]] (c fp32 channels) for r < n_rows; either pair may be NULL. */
int pcg_gather_rows(const int32_t *coords, const float *feats, int c, const int32_t *rows, long n_rows, int32_t *out_coords,
                    float *out_feats, void *stream);
/* The same gather from bf16 source rows (c % 8 == 0) into fp32 output rows: survivors of a prune whose candidate rows exist
 * only as the bf16 operand image the up-sampling kernel emitted (no fp32 copy of the ~20x larger candidate set is ever written). */
int pcg_gather_rows_bf16(const int32_t *coords, const void *feats_bf16, int c, const int32_t *rows, long n_rows,
                         int32_t *out_coords, float *out_feats, void *stream);

/* ---- interpolation: SparseTensor.features_at_coordinates (ME6) ---------------------------------
 * Trilinear over the 2^3 corners at spacing ts; absent corners contribute 0.  query: nq x 4 fp32. */
int pcg_features_at(const uint64_t *keys, const int32_t *vals, long cap, const float *feats, int c, int ts,
                    const float *query, long nq, float *out, void *stream);

/* ---- entropy models (compressai 1.2.4, CAI1/CAI2) ------------------------------------------------
 * Element (n, ch) of every operand lives at base[n*sn + ch*sc] (strides in elements), so both the
 * (N,C) feature layout and the reference's (1,C,N) layout are addressable without a copy.
 * EntropyBottleneck: eb_params is 58 floats per channel for filters (3,3,3,3):
 *   per layer l: softplus(matrix_l) row-major, bias_l, tanh(factor_l) (no factor for the last). */
int pcg_eb_likelihood(const float *x, long n, int c, long sn, long sc, const float *eb_params,
                      float bound, float *lik, void *stream);
/* mode 0: out_f = round(x - med) + med ("dequantize"); mode 1: out_i = (int) round(x - med). */
int pcg_quantize(const float *x, const float *means, long n, int c, long sn, long sc, long msn, long msc,
                 int mode, float *out_f, int32_t *out_i, void *stream);
/* out = (float) sym + means  (EntropyModel.dequantize). */
int pcg_dequantize(const int32_t *sym, const float *means, long n, int c, long sn, long sc, long msn, long msc,
                   float *out, void *stream);
/* GaussianConditional._likelihood: v=|x-mean|, s=max(scale,bound_scale);
 * lik = max(Phi((.5-v)/s) - Phi((-.5-v)/s), bound). means may be NULL. */
int pcg_gc_likelihood(const float *x, const float *scales, const float *means, long n, int c, long sn, long sc,
                      float scale_bound, float lik_bound, float *lik, void *stream);
/* GaussianConditional.build_indexes: idx = (T-1) - #{t in table[:-1] : max(s,bound) <= t}. */
int pcg_gc_build_indexes(const float *scales, long n, int c, long sn, long sc, const float *table, int T,
                         float scale_bound, int32_t *idx, void *stream);
/* Fused reconstruction with learned offsets (entropy_models.py:218-233,469-481; a8):
 *   sd = max(scales_hat*gain, bound); off = -quant_nn([gain, sd]); off = 0 where |q| < 1e-4;
 *   y_hat = sign(q) * (|q| + off) * rescale + mean.
 * q, scales_hat, means_hat, y_hat share (sn, sc); gain/rescale are per (row, channel) with strides
 * (gsn, gsc).  mlp: 2->10->10->1 weights packed [W1(10x2) b1(10) W2(10x10) b2(10) W3(1x10) b3(1)]. */
int pcg_gc_dequant_offsets(const float *q, const float *scales_hat, const float *means_hat, const float *gain,
                           const float *rescale, long n, int c, long sn, long sc, long gsn, long gsc,
                           const float *mlp, float scale_bound, float *y_hat, void *stream);

/* ---- PSNR metric: nearest-neighbour association of two voxel clouds (metrics/metric.py:36-43,107-166) ----
 * keys/vals/cap: coordinate hash of cloud B (pcg_hash_build).  For every row of a_coords (int32 (b,x,y,z)) finds the
 * nearest row of B by probing growing Chebyshev shells (at most max_radius): nn_index (smallest row among equidistant
 * neighbours), nn_d2 (integer squared distance), and -- when yuv_sqerr, a_colors and b_colors (fp32 rgb in [0,1], 3 per
 * row) are given -- the per-point squared BT.709 YUV error after the reference's 8-bit rounding, with the colours of
 * equidistant neighbours averaged as metric.py:127-146 does (first neighbour counted twice, five nearest at most).
 * *d_status bit 0: some point had no neighbour within max_radius (its nn_index / nn_d2 are -1). */
int pcg_nn_search(const uint64_t *keys, const int32_t *vals, long cap, const float *b_colors, const int32_t *a_coords,
                  const float *a_colors, long na, int max_radius, int32_t *nn_index, int32_t *nn_d2, float *yuv_sqerr,
                  int32_t *d_status, void *stream);

/* ---- training-time augmentation of a cube (data/transform.py:33-124) -------------------------
 * pcg_color_jitter: torchvision.transforms.ColorJitter semantics on n x 3 fp32 colour rows in [0,1] seen as a 3 x n x 1
 * image (transform.py:50-53): the four ops (0 brightness, 1 contrast, 2 saturation, 3 hue) applied in the order op0..op3
 * with the given factors; the contrast op blends with the mean grey level of the whole sample at that point of the chain.
 * scratch: pcg_color_jitter_scratch_bytes() device bytes. */
size_t pcg_color_jitter_scratch_bytes(void);
int pcg_color_jitter(const float *colors, long n, int op0, int op1, int op2, int op3, float brightness, float contrast,
                     float saturation, float hue, float *out, void *scratch, void *stream);
/* pcg_rotate_floor: row i of (points - half) @ R^T + half, floored, as [0,x,y,z] int32 rows (transform.py:78-99 up to the
 * de-duplication, which is pcg_unique_first).  rotation_host9: R row-major in HOST memory.  keep (optional): 1 where the
 * rotated position lies in [0, crop_size)^3 (transform.py:86-89), all ones for crop_size <= 0. */
int pcg_rotate_floor(const float *points, long n, const float *rotation_host9, float half, float crop_size,
                     int32_t *out_coords, uint8_t *keep, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PCG_H */
