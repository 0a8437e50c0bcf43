// Lossless coder for the latent coordinate set (the stride-8 voxels of y) -- SURVEY.md §8 (f2).
//
// The reference writes these points to an ASCII PLY and runs the external G-PCC binary tmc3 on it
// (/root/reference/model/model.py:388-486, `gpcc_encode` / `gpcc_decode`); tmc3 is not part of the reference tree and
// cannot be built here, so the side channel is restated as a native octree occupancy coder with the same contract:
// a SET of integer points in, bytes out, the same set back (order not preserved -- the decoder re-derives the
// canonical row order, as it does after tmc3).
//
//   points -> (p - origin) / stride -> 63-bit Morton codes, sorted, unique
//   breadth-first octree: every internal node emits its 8-bit child-occupancy mask, nodes of a level in Morton order
//   each mask is coded MSB-first through a 255-node binary tree of adaptive 11-bit probabilities (one tree per
//   context = {levels left: 0, 1, >= 2} x {children of the parent: <= 1, 2, 3-4, >= 5}) with a carry-propagating
//   32-bit binary range coder (the LZMA rc construction).
//
// Host code by design (north-star: entropy coders stay on the host; the set has ~14 k points for a 10-bit frame).
#include <algorithm>
#include <climits>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

constexpr uint32_t kTop = 1u << 24;
constexpr int kProbBits = 11, kMove = 5, kCtx = 12;
constexpr long kHeader = 24;

struct Encoder {
    std::vector<uint8_t> &out;
    uint64_t low = 0;
    uint32_t range = 0xFFFFFFFFu;
    uint8_t cache = 0;
    uint64_t cache_size = 1;
    explicit Encoder(std::vector<uint8_t> &o) : out(o) {}
    void shift_low() {
        if ((uint32_t)low < 0xFF000000u || (low >> 32) != 0) {
            uint8_t carry = (uint8_t)(low >> 32), b = cache;
            do {
                out.push_back((uint8_t)(b + carry));
                b = 0xFF;
            } while (--cache_size != 0);
            cache = (uint8_t)((uint32_t)low >> 24);
        }
        ++cache_size;
        low = (uint64_t)((uint32_t)low << 8);
    }
    void bit(uint16_t &p, int b) {
        const uint32_t bound = (range >> kProbBits) * p;
        if (b == 0) {
            range = bound;
            p = (uint16_t)(p + (((1u << kProbBits) - p) >> kMove));
        } else {
            low += bound;
            range -= bound;
            p = (uint16_t)(p - (p >> kMove));
        }
        while (range < kTop) {
            range <<= 8;
            shift_low();
        }
    }
    void flush() {
        for (int i = 0; i < 5; ++i) shift_low();
    }
};

struct Decoder {
    const uint8_t *in;
    long n, pos = 0;
    uint32_t range = 0xFFFFFFFFu, code = 0;
    bool overrun = false;
    Decoder(const uint8_t *p, long nb) : in(p), n(nb) {
        for (int i = 0; i < 5; ++i) code = (code << 8) | next();
    }
    uint8_t next() {
        if (pos < n) return in[pos++];
        overrun = true;
        return 0;
    }
    int bit(uint16_t &p) {
        const uint32_t bound = (range >> kProbBits) * p;
        int b;
        if (code < bound) {
            range = bound;
            p = (uint16_t)(p + (((1u << kProbBits) - p) >> kMove));
            b = 0;
        } else {
            code -= bound;
            range -= bound;
            p = (uint16_t)(p - (p >> kMove));
            b = 1;
        }
        while (range < kTop) {
            range <<= 8;
            code = (code << 8) | next();
        }
        return b;
    }
};

struct Models {
    uint16_t p[kCtx][256];
    Models() {
        for (auto &t : p)
            for (auto &v : t) v = 1u << (kProbBits - 1);
    }
};

inline int context(int levels_left, int parent_children) {
    const int lb = levels_left > 2 ? 2 : levels_left;
    const int cb = parent_children <= 1 ? 0 : (parent_children == 2 ? 1 : (parent_children <= 4 ? 2 : 3));
    return lb * 4 + cb;
}

inline uint64_t spread3(uint64_t v) {          // 21 bits -> every third bit
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x1F00000000FFFFull;
    v = (v | (v << 16)) & 0x1F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
inline uint32_t compact3(uint64_t v) {
    v &= 0x1249249249249249ull;
    v = (v | (v >> 2)) & 0x10C30C30C30C30C3ull;
    v = (v | (v >> 4)) & 0x100F00F00F00F00Full;
    v = (v | (v >> 8)) & 0x1F0000FF0000FFull;
    v = (v | (v >> 16)) & 0x1F00000000FFFFull;
    v = (v | (v >> 32)) & 0x1FFFFFull;
    return (uint32_t)v;
}

inline void put32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }
inline uint32_t get32(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

}  // namespace

extern "C" {

// xyz: n x 3 int32 (host), every coordinate = origin + stride * (0 .. 2^21 - 1).  Duplicates are coded once.
// Returns the bytes written, -(bytes needed) when out_cap is too small, LONG_MIN for coordinates off the stride
// lattice or an extent beyond 21 bits per axis.
long pcg_octree_encode(const int32_t *xyz, long n, int stride, uint8_t *out, long out_cap) {
    if (n < 0 || stride < 1 || stride > 0xFFFF) return LONG_MIN;
    int32_t lo[3] = {0, 0, 0};
    for (long i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a)
            if (i == 0 || xyz[3 * i + a] < lo[a]) lo[a] = xyz[3 * i + a];
    std::vector<uint64_t> codes((size_t)n);
    uint32_t extent = 0;
    for (long i = 0; i < n; ++i) {
        uint64_t c = 0;
        for (int a = 0; a < 3; ++a) {
            const int64_t d = (int64_t)xyz[3 * i + a] - lo[a];
            if (d % stride != 0 || d / stride >= (1 << 21)) return LONG_MIN;
            const uint32_t v = (uint32_t)(d / stride);
            extent = std::max(extent, v);
            c |= spread3(v) << a;
        }
        codes[(size_t)i] = c;
    }
    std::sort(codes.begin(), codes.end());
    codes.erase(std::unique(codes.begin(), codes.end()), codes.end());
    int depth = 0;
    while (depth < 21 && (extent >> depth) != 0) ++depth;

    std::vector<uint8_t> body;
    body.reserve(codes.size() + 64);
    Encoder enc(body);
    Models models;
    // level L: nodes = distinct prefixes code >> 3*(depth-L); a node's children are the distinct prefixes one level down
    std::vector<uint8_t> parent_children(1, 0), next_children;
    for (int L = 0; L < depth && !codes.empty(); ++L) {
        const int s = 3 * (depth - L);
        next_children.clear();
        size_t i = 0, node = 0;
        while (i < codes.size()) {
            const uint64_t prefix = codes[i] >> s;
            unsigned mask = 0;
            size_t j = i;
            for (; j < codes.size() && (codes[j] >> s) == prefix; ++j) mask |= 1u << ((codes[j] >> (s - 3)) & 7u);
            uint16_t *tree = models.p[context(depth - 1 - L, parent_children[node])];
            unsigned m = 1;
            for (int b = 7; b >= 0; --b) {
                const int bitv = (mask >> b) & 1;
                enc.bit(tree[m], bitv);
                m = (m << 1) | (unsigned)bitv;
            }
            const int cnt = __builtin_popcount(mask);
            for (int c = 0; c < cnt; ++c) next_children.push_back((uint8_t)cnt);
            i = j;
            ++node;
        }
        parent_children.swap(next_children);
    }
    enc.flush();
    const long total = kHeader + (long)body.size();
    if (total > out_cap) return -total;
    std::memcpy(out, "PCO1", 4);
    out[4] = (uint8_t)stride; out[5] = (uint8_t)(stride >> 8); out[6] = (uint8_t)depth; out[7] = 0;
    put32(out + 8, (uint32_t)codes.size());
    for (int a = 0; a < 3; ++a) put32(out + 12 + 4 * a, (uint32_t)lo[a]);
    std::memcpy(out + kHeader, body.data(), body.size());
    return total;
}

// Number of points of a stream (from its header), or LONG_MIN when the header is not one of ours.
long pcg_octree_count(const uint8_t *in, long nbytes) {
    if (nbytes < kHeader || std::memcmp(in, "PCO1", 4) != 0 || in[6] > 21) return LONG_MIN;
    return (long)get32(in + 8);
}

// Points in Morton order into xyz_out (cap points).  Returns the point count, -(count) when cap is too small,
// LONG_MIN for a corrupt stream (bad header, zero occupancy mask, payload overrun, count mismatch).
long pcg_octree_decode(const uint8_t *in, long nbytes, int32_t *xyz_out, long cap) {
    const long n = pcg_octree_count(in, nbytes);
    if (n == LONG_MIN) return LONG_MIN;
    if (n > cap) return -n;
    const int stride = (int)in[4] | ((int)in[5] << 8), depth = in[6];
    if (stride < 1) return LONG_MIN;
    int32_t lo[3];
    for (int a = 0; a < 3; ++a) lo[a] = (int32_t)get32(in + 12 + 4 * a);
    if (n == 0) return 0;
    Decoder dec(in + kHeader, nbytes - kHeader);
    Models models;
    std::vector<uint64_t> nodes(1, 0), next_nodes;
    std::vector<uint8_t> parent_children(1, 0), next_children;
    for (int L = 0; L < depth; ++L) {
        next_nodes.clear();
        next_children.clear();
        for (size_t node = 0; node < nodes.size(); ++node) {
            uint16_t *tree = models.p[context(depth - 1 - L, parent_children[node])];
            unsigned m = 1;
            for (int b = 0; b < 8; ++b) m = (m << 1) | (unsigned)dec.bit(tree[m]);
            const unsigned mask = m & 0xFFu;
            if (mask == 0 || dec.overrun) return LONG_MIN;
            const int cnt = __builtin_popcount(mask);
            for (unsigned c = 0; c < 8; ++c)
                if ((mask >> c) & 1u) {
                    next_nodes.push_back((nodes[node] << 3) | c);
                    next_children.push_back((uint8_t)cnt);
                }
            if ((long)next_nodes.size() > n) return LONG_MIN;
        }
        nodes.swap(next_nodes);
        parent_children.swap(next_children);
    }
    if ((long)nodes.size() != n) return LONG_MIN;
    for (long i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a)
            xyz_out[3 * i + a] = (int32_t)((int64_t)lo[a] + (int64_t)stride * compact3(nodes[(size_t)i] >> a));
    return n;
}

}  // extern "C"
