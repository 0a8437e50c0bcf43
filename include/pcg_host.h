/* pcg_host.h -- C ABI of libpcg_host.so: the HOST side of the codec (entropy coders and the latent-coordinate side
 * channel).  North-star: these stay on the host; they consume / produce the int32 buffers of the GPU path (pcg.h).
 * Plain pointers and sizes, caller allocates, no state.  Sizes unknown in advance: the call returns -(needed) and the
 * caller retries.
 *
 * Replaces, in the reference:
 *   pcg_rans_encode / pcg_rans_decode   compressai `RansEncoder.encode_with_indexes` / `RansDecoder.decode_with_indexes`
 *                                       (call sites model/entropy_models.py:371-372,397-400,438,471,484)
 *   pcg_pmf_to_quantized_cdf            compressai `pmf_to_quantized_cdf` (via EntropyBottleneck.update /
 *                                       GaussianConditional.update_scale_table, model/model.py:30-34)
 *   pcg_octree_encode / _decode         the tmc3 (G-PCC) subprocess + ASCII PLY round trip of the stride-8 latent
 *                                       coordinates (model/model.py:388-486 `gpcc_encode` / `gpcc_decode`)
 */
#ifndef PCG_HOST_H
#define PCG_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* symbols / indexes: n int32 each; cdfs: [n_tables][cdf_stride] int32; cdf_sizes / offsets: per table.
 * Returns the bytes written, or -(bytes needed) when out_cap is too small. */
long pcg_rans_encode(const int32_t *symbols, const int32_t *indexes, long n, const int32_t *cdfs, int cdf_stride,
                     const int32_t *cdf_sizes, const int32_t *offsets, uint8_t *out, long out_cap);
/* Returns 0, or non-zero for a corrupt stream. */
int pcg_rans_decode(const uint8_t *enc, long enc_bytes, const int32_t *indexes, long n, const int32_t *cdfs,
                    int cdf_stride, const int32_t *cdf_sizes, const int32_t *offsets, int32_t *symbols_out);
/* cdf_out: n + 2 int32 (leading 0, tail entry last), total 2^precision. */
int pcg_pmf_to_quantized_cdf(const float *pmf, int n, int precision, int32_t *cdf_out);

/* Lossless octree occupancy coder for a SET of lattice points.  xyz: n x 3 int32, every coordinate = (per-axis minimum)
 * + stride * (0 .. 2^21 - 1); duplicates are coded once.  Returns the bytes written, -(bytes needed) when out_cap is too
 * small, LONG_MIN for points off the stride lattice / extent beyond 21 bits. */
long pcg_octree_encode(const int32_t *xyz, long n, int stride, uint8_t *out, long out_cap);
/* Point count recorded in a stream's header; LONG_MIN when the header is not this coder's. */
long pcg_octree_count(const uint8_t *in, long nbytes);
/* Points in Morton order (order of the input is not preserved: it is a set).  Returns the count, -(count) when cap is
 * too small, LONG_MIN for a corrupt stream. */
long pcg_octree_decode(const uint8_t *in, long nbytes, int32_t *xyz_out, long cap);

#ifdef __cplusplus
}
#endif
#endif
