/* CPU restatement of compressai 1.2.4's range coder.  TEST INFRASTRUCTURE (oracle).
 *
 * Restates cpp_exts/rans/rans_interface.cpp (BufferedRansEncoder::encode_with_indexes /
 * flush, RansDecoder::decode_with_indexes) over ryg's public-domain rans64.h scheme:
 * 64-bit state, 32-bit renormalisation words, 16-bit probability precision, 4-bit bypass
 * digits for out-of-range symbols.  Call sites in the reference:
 * model/entropy_models.py:371-372,397-400,438,471,484 (through EntropyModel.compress /
 * decompress).  PARITY UNPINNED: compressai is not installable here; round-trip identity
 * is what tests/test_oracle_entropy.py checks.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PRECISION 16
#define BYPASS_PRECISION 4
#define MAX_BYPASS_VAL ((1 << BYPASS_PRECISION) - 1)
#define RANS64_L (1ull << 31)

typedef struct { uint16_t start; uint16_t range; uint8_t bypass; } sym_t;

static inline void enc_put(uint64_t *r, uint32_t **pp, uint32_t start, uint32_t freq, uint32_t scale_bits) {
    uint64_t x = *r;
    uint64_t x_max = ((RANS64_L >> scale_bits) << 32) * freq;
    if (x >= x_max) { *pp -= 1; **pp = (uint32_t)x; x >>= 32; }
    *r = ((x / freq) << scale_bits) + (x % freq) + start;
}

static inline void enc_put_bits(uint64_t *r, uint32_t **pp, uint32_t val, uint32_t nbits) {
    uint64_t x = *r;
    uint32_t freq = 1u << (16 - nbits);
    uint64_t x_max = ((RANS64_L >> 16) << 32) * freq;
    if (x >= x_max) { *pp -= 1; **pp = (uint32_t)x; x >>= 32; }
    *r = (x << nbits) | val;
}

/* Returns number of bytes written to `out` (capacity out_cap bytes), or -1 on overflow. */
long rans_encode_with_indexes(const int32_t *symbols, const int32_t *indexes, long n,
                              const int32_t *cdfs, int cdf_stride, const int32_t *cdf_sizes,
                              const int32_t *offsets, uint8_t *out, long out_cap) {
    long cap = n * 12 + 16, ns = 0;
    sym_t *syms = (sym_t *)malloc(sizeof(sym_t) * (size_t)cap);
    if (!syms) return -1;
    for (long i = 0; i < n; ++i) {
        int32_t ci = indexes[i];
        const int32_t *cdf = cdfs + (long)ci * cdf_stride;
        int32_t max_value = cdf_sizes[ci] - 2;
        int32_t value = symbols[i] - offsets[ci];
        uint32_t raw_val = 0;
        if (value < 0) { raw_val = (uint32_t)(-2 * value - 1); value = max_value; }
        else if (value >= max_value) { raw_val = (uint32_t)(2 * (value - max_value)); value = max_value; }
        if (ns + 12 >= cap) { cap *= 2; syms = (sym_t *)realloc(syms, sizeof(sym_t) * (size_t)cap); }
        syms[ns++] = (sym_t){(uint16_t)cdf[value], (uint16_t)(cdf[value + 1] - cdf[value]), 0};
        if (value == max_value) {
            int32_t n_bypass = 0;
            while ((raw_val >> (n_bypass * BYPASS_PRECISION)) != 0) ++n_bypass;
            int32_t val = n_bypass;
            while (val >= MAX_BYPASS_VAL) {
                syms[ns++] = (sym_t){MAX_BYPASS_VAL, MAX_BYPASS_VAL + 1, 1};
                val -= MAX_BYPASS_VAL;
            }
            syms[ns++] = (sym_t){(uint16_t)val, (uint16_t)(val + 1), 1};
            for (int32_t j = 0; j < n_bypass; ++j) {
                int32_t v = (raw_val >> (j * BYPASS_PRECISION)) & MAX_BYPASS_VAL;
                syms[ns++] = (sym_t){(uint16_t)v, (uint16_t)(v + 1), 1};
            }
        }
    }
    uint32_t *buf = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(ns + 2));
    uint32_t *end = buf + ns + 2, *ptr = end;
    uint64_t r = RANS64_L;
    for (long i = ns - 1; i >= 0; --i) {
        if (!syms[i].bypass) enc_put(&r, &ptr, syms[i].start, syms[i].range, PRECISION);
        else enc_put_bits(&r, &ptr, syms[i].start, BYPASS_PRECISION);
    }
    ptr -= 2; ptr[0] = (uint32_t)r; ptr[1] = (uint32_t)(r >> 32);
    long nbytes = (long)(end - ptr) * 4;
    long ret = -1;
    if (nbytes <= out_cap) { memcpy(out, ptr, (size_t)nbytes); ret = nbytes; }
    free(buf); free(syms);
    return ret;
}

static inline uint32_t dec_get_bits(uint64_t *r, const uint32_t **pp, uint32_t nbits) {
    uint64_t x = *r;
    uint32_t val = (uint32_t)(x & ((1u << nbits) - 1));
    x >>= nbits;
    if (x < RANS64_L) { x = (x << 32) | **pp; *pp += 1; }
    *r = x;
    return val;
}

void rans_decode_with_indexes(const uint8_t *encoded, const int32_t *indexes, long n,
                              const int32_t *cdfs, int cdf_stride, const int32_t *cdf_sizes,
                              const int32_t *offsets, int32_t *output) {
    const uint32_t *ptr = (const uint32_t *)encoded;
    uint64_t r = (uint64_t)ptr[0] | ((uint64_t)ptr[1] << 32);
    ptr += 2;
    const uint32_t mask = (1u << PRECISION) - 1;
    for (long i = 0; i < n; ++i) {
        int32_t ci = indexes[i];
        const int32_t *cdf = cdfs + (long)ci * cdf_stride;
        int32_t max_value = cdf_sizes[ci] - 2;
        uint32_t cum = (uint32_t)(r & mask);
        int32_t s = 0, size = cdf_sizes[ci];
        while (s < size && (uint32_t)cdf[s] <= cum) ++s;
        s -= 1;
        {   /* Rans64DecAdvance */
            uint64_t x = r;
            uint32_t start = (uint32_t)cdf[s], freq = (uint32_t)(cdf[s + 1] - cdf[s]);
            x = freq * (x >> PRECISION) + (x & mask) - start;
            if (x < RANS64_L) { x = (x << 32) | *ptr; ptr += 1; }
            r = x;
        }
        int32_t value = s;
        if (value == max_value) {
            int32_t val = (int32_t)dec_get_bits(&r, &ptr, BYPASS_PRECISION);
            int32_t n_bypass = val;
            while (val == MAX_BYPASS_VAL) { val = (int32_t)dec_get_bits(&r, &ptr, BYPASS_PRECISION); n_bypass += val; }
            int32_t raw_val = 0;
            for (int32_t j = 0; j < n_bypass; ++j) {
                val = (int32_t)dec_get_bits(&r, &ptr, BYPASS_PRECISION);
                raw_val |= val << (j * BYPASS_PRECISION);
            }
            value = raw_val >> 1;
            if (raw_val & 1) value = -value - 1; else value += max_value;
        }
        output[i] = value + offsets[ci];
    }
}
