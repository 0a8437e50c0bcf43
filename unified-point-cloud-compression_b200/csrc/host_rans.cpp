// Host range coder for the symbol streams the GPU entropy kernels produce.
//
// The north-star keeps the arithmetic coder on the host ("stays the reference's host coder");
// compressai's `ans` extension is not installable here, so this is a from-scratch host coder that
// writes the SAME bitstream format (rANS, 64-bit state, 32-bit words, 16-bit probabilities, 4-bit
// bypass digits for out-of-table symbols) so strings interoperate with
// compressai.ans.RansEncoder/RansDecoder (call sites: model/entropy_models.py:371-372,397-400,438,
// 471,484).  Unlike the reference's coder it does not buffer a symbol list: rANS consumes symbols
// last-to-first, so the encoder walks the int32 symbol array backwards and emits directly.
#include <cstdint>
#include <climits>
#include <cstring>
#include <cmath>
#include <vector>

namespace {
constexpr uint32_t kProbBits = 16, kBypassBits = 4, kBypassMax = (1u << kBypassBits) - 1;
constexpr uint64_t kLow = 1ull << 31;

struct Writer {            // grows towards lower addresses
    std::vector<uint32_t> buf;
    size_t head;
    explicit Writer(size_t words) : buf(words), head(words) {}
    inline void push(uint32_t w) {
        if (head == 0) {
            size_t old = buf.size();
            std::vector<uint32_t> nb(old * 2);
            std::memcpy(nb.data() + old, buf.data(), old * sizeof(uint32_t));
            buf.swap(nb);
            head = old;
        }
        buf[--head] = w;
    }
};

inline void put(uint64_t &x, Writer &w, uint32_t start, uint32_t freq, uint32_t bits) {
    const uint64_t x_max = ((kLow >> bits) << 32) * freq;
    if (x >= x_max) { w.push((uint32_t)x); x >>= 32; }
    x = ((x / freq) << bits) + (x % freq) + start;
}
inline void put_raw(uint64_t &x, Writer &w, uint32_t val) {   // kBypassBits literal bits
    const uint64_t x_max = ((kLow >> kProbBits) << 32) * (uint64_t)(1u << (kProbBits - kBypassBits));
    if (x >= x_max) { w.push((uint32_t)x); x >>= 32; }
    x = (x << kBypassBits) | val;
}
}  // namespace

extern "C" {

// symbols/indexes: n int32 each (host).  cdfs: [n_tables][cdf_stride] int32.  Returns bytes written or
// -(needed bytes) when out_cap is too small.
long pcg_rans_encode(const int32_t *symbols, const int32_t *indexes, long n, const int32_t *cdfs, int cdf_stride,
                     const int32_t *cdf_sizes, const int32_t *offsets, uint8_t *out, long out_cap) {
    Writer w((size_t)(n / 2 + 64));
    uint64_t x = kLow;
    for (long i = n - 1; i >= 0; --i) {
        const int32_t t = indexes[i];
        const int32_t *cdf = cdfs + (long)t * cdf_stride;
        const int32_t top = cdf_sizes[t] - 2;           // escape slot
        int32_t v = symbols[i] - offsets[t];
        if (v >= 0 && v < top) {
            put(x, w, (uint32_t)cdf[v], (uint32_t)(cdf[v + 1] - cdf[v]), kProbBits);
            continue;
        }
        // escape: [escape symbol][unary-ish digit count][raw digits low->high]; emitted in reverse
        const uint32_t raw = v < 0 ? (uint32_t)(-2 * v - 1) : (uint32_t)(2 * (v - top));
        int digits = 0;
        while ((raw >> (digits * kBypassBits)) != 0) ++digits;
        for (int j = digits - 1; j >= 0; --j) put_raw(x, w, (raw >> (j * kBypassBits)) & kBypassMax);
        // count is coded as (count / 15) saturated digits followed by the remainder
        int full = digits / (int)kBypassMax, rem = digits % (int)kBypassMax;
        put_raw(x, w, (uint32_t)rem);
        for (int j = 0; j < full; ++j) put_raw(x, w, kBypassMax);
        put(x, w, (uint32_t)cdf[top], (uint32_t)(cdf[top + 1] - cdf[top]), kProbBits);
    }
    w.push((uint32_t)(x >> 32));
    w.push((uint32_t)x);
    const long nbytes = (long)(w.buf.size() - w.head) * 4;
    if (nbytes > out_cap) return -nbytes;
    std::memcpy(out, w.buf.data() + w.head, (size_t)nbytes);
    return nbytes;
}

int pcg_rans_decode(const uint8_t *enc, long enc_bytes, const int32_t *indexes, long n, const int32_t *cdfs,
                    int cdf_stride, const int32_t *cdf_sizes, const int32_t *offsets, int32_t *out) {
    if (enc_bytes < 8) return 1;
    const uint32_t *p = reinterpret_cast<const uint32_t *>(enc);
    const uint32_t *end = p + enc_bytes / 4;
    uint64_t x = (uint64_t)p[0] | ((uint64_t)p[1] << 32);
    p += 2;
    auto refill = [&]() { if (x < kLow) { x = (x << 32) | (p < end ? *p : 0u); ++p; } };
    auto raw4 = [&]() { uint32_t v = (uint32_t)(x & kBypassMax); x >>= kBypassBits; refill(); return v; };
    const uint32_t mask = (1u << kProbBits) - 1;
    for (long i = 0; i < n; ++i) {
        const int32_t t = indexes[i];
        const int32_t *cdf = cdfs + (long)t * cdf_stride;
        const int32_t size = cdf_sizes[t], top = size - 2;
        const uint32_t cum = (uint32_t)(x & mask);
        int32_t lo = 0, hi = size - 1;                 // largest s with cdf[s] <= cum
        while (hi - lo > 1) { int32_t mid = (lo + hi) >> 1; if ((uint32_t)cdf[mid] <= cum) lo = mid; else hi = mid; }
        const int32_t s = lo;
        x = (uint64_t)(uint32_t)(cdf[s + 1] - cdf[s]) * (x >> kProbBits) + (x & mask) - (uint32_t)cdf[s];
        refill();
        int32_t v = s;
        if (s == top) {
            uint32_t d = raw4();
            int32_t digits = (int32_t)d;
            while (d == kBypassMax) { d = raw4(); digits += (int32_t)d; }
            uint32_t raw = 0;
            for (int32_t j = 0; j < digits; ++j) raw |= raw4() << (j * kBypassBits);
            v = (int32_t)(raw >> 1);
            v = (raw & 1u) ? -v - 1 : v + top;
        }
        out[i] = v + offsets[t];
    }
    return 0;
}

}  // extern "C"

// 16-bit quantised CDF from a pmf (+ tail mass as the last entry), the table format the coder above
// consumes.  Restates the published rule of compressai's `pmf_to_quantized_cdf` (round to 2^precision,
// renormalise, then make every symbol codable by taking one count from the cheapest donor).
extern "C" int pcg_pmf_to_quantized_cdf(const float *pmf, int n, int precision, int32_t *cdf_out) {
    std::vector<uint64_t> freq(n + 1, 0);
    uint64_t total = 0;
    for (int i = 0; i < n; ++i) {
        if (!(pmf[i] >= 0.f) || pmf[i] > 3.0e38f) return 1;
        freq[i + 1] = (uint64_t)std::round(pmf[i] * (float)(1 << precision));     // compressai: std::round (fp32 product, half away from zero)
        total += freq[i + 1];
    }
    if (total == 0) return 2;
    std::vector<int64_t> cdf(n + 1);
    int64_t run = 0;
    for (int i = 0; i <= n; ++i) { run += (int64_t)(((uint64_t)(1 << precision) * freq[i]) / total); cdf[i] = run; }
    cdf[n] = 1 << precision;
    for (int i = 0; i < n; ++i) {
        if (cdf[i] != cdf[i + 1]) continue;
        int64_t best = INT64_MAX;
        int donor = -1;
        for (int j = 0; j < n; ++j) {
            int64_t f = cdf[j + 1] - cdf[j];
            if (f > 1 && f < best) { best = f; donor = j; }
        }
        if (donor < 0) return 3;
        if (donor < i) for (int j = donor + 1; j <= i; ++j) --cdf[j];
        else for (int j = i + 1; j <= donor; ++j) ++cdf[j];
    }
    for (int i = 0; i <= n; ++i) cdf_out[i] = (int32_t)cdf[i];
    return 0;
}
