// CPU restatement of MinkowskiEngine's CPU sparse-convolution algorithm.
// TEST INFRASTRUCTURE (oracle) and the timed CPU baseline of bench.py.
//
// Restates (from memory of NVIDIA/MinkowskiEngine master, un-vendored, PARITY UNPINNED):
//   src/coordinate_map_cpu.hpp   hash map of (b,x,y,z) rows -> row index, `stride`,
//                                `stride_region` (generative) and `kernel_map`
//   src/convolution_kernel.cpp   per kernel offset: gather rows -> cblas_sgemm -> scatter-add
// built `--blas=openblas` in the reference (README.md:52).  BLAS here is the OpenBLAS that
// ships inside scipy (`scipy_cblas_sgemm`), bound at run time through `ref_set_sgemm`.
// Call sites in the reference: every ME.MinkowskiConvolution /
// MinkowskiGenerativeConvolutionTranspose in model/transforms.py:32-166 and
// model/entropy_models.py:178-191.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef void (*sgemm_fn)(int order, int ta, int tb, int M, int N, int K, float alpha,
                         const float *A, int lda, const float *B, int ldb, float beta,
                         float *C, int ldc);
static sgemm_fn g_sgemm = nullptr;

static inline uint64_t pack(int32_t b, int32_t x, int32_t y, int32_t z) {
    return ((uint64_t)(uint16_t)b << 48) | ((uint64_t)(uint16_t)(x + 32768) << 32) |
           ((uint64_t)(uint16_t)(y + 32768) << 16) | (uint64_t)(uint16_t)(z + 32768);
}
static inline uint64_t mix(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

struct CoordMap {  // open addressing, linear probing (stands in for ME's robin-hood map)
    std::vector<uint64_t> keys;
    std::vector<int32_t> vals;
    uint64_t mask;
    explicit CoordMap(size_t n) {
        size_t cap = 16;
        while (cap < 2 * n + 2) cap <<= 1;
        keys.assign(cap, ~0ull);
        vals.assign(cap, -1);
        mask = cap - 1;
    }
    // returns true if inserted (first occurrence)
    bool insert(uint64_t k, int32_t v) {
        uint64_t s = mix(k) & mask;
        while (true) {
            if (keys[s] == ~0ull) { keys[s] = k; vals[s] = v; return true; }
            if (keys[s] == k) return false;
            s = (s + 1) & mask;
        }
    }
    int32_t find(uint64_t k) const {
        uint64_t s = mix(k) & mask;
        while (true) {
            if (keys[s] == k) return vals[s];
            if (keys[s] == ~0ull) return -1;
            s = (s + 1) & mask;
        }
    }
};

static void offsets_for(int ksize, std::vector<int> &off) {
    int lo = (ksize % 2 == 1) ? -(ksize / 2) : 0;
    off.clear();
    for (int iz = 0; iz < ksize; ++iz)
        for (int iy = 0; iy < ksize; ++iy)
            for (int ix = 0; ix < ksize; ++ix) {
                off.push_back(lo + ix); off.push_back(lo + iy); off.push_back(lo + iz);
            }
}

static inline int32_t floordiv(int32_t a, int32_t b) {
    int32_t q = a / b;
    return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q;
}

extern "C" {

void ref_set_sgemm(void *fn) { g_sgemm = (sgemm_fn)fn; }
void ref_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#endif
    (void)n;
}

// first-occurrence unique of floor(c / s) * s; returns n_out, writes out rows
long ref_stride_coords(const int32_t *C, long n, int s, int32_t *out) {
    CoordMap m((size_t)n);
    long no = 0;
    for (long i = 0; i < n; ++i) {
        int32_t b = C[4 * i], x = floordiv(C[4 * i + 1], s) * s, y = floordiv(C[4 * i + 2], s) * s,
                z = floordiv(C[4 * i + 3], s) * s;
        if (m.insert(pack(b, x, y, z), (int32_t)no)) {
            out[4 * no] = b; out[4 * no + 1] = x; out[4 * no + 2] = y; out[4 * no + 3] = z; ++no;
        }
    }
    return no;
}

// generative transposed conv output coords: unique{p + off*ts_out}; out capacity n*K rows
long ref_generate_coords(const int32_t *C, long n, int ksize, int ts_out, int32_t *out) {
    std::vector<int> off; offsets_for(ksize, off);
    int K = ksize * ksize * ksize;
    CoordMap m((size_t)n * (size_t)std::min(K, 32));
    long no = 0;
    for (long i = 0; i < n; ++i)
        for (int k = 0; k < K; ++k) {
            int32_t b = C[4 * i], x = C[4 * i + 1] + off[3 * k] * ts_out, y = C[4 * i + 2] + off[3 * k + 1] * ts_out,
                    z = C[4 * i + 3] + off[3 * k + 2] * ts_out;
            if ((size_t)no * 2 + 2 >= m.keys.size()) {  // grow
                CoordMap g((size_t)no * 2);
                for (long j = 0; j < no; ++j) g.insert(pack(out[4 * j], out[4 * j + 1], out[4 * j + 2], out[4 * j + 3]), (int32_t)j);
                m = std::move(g);
            }
            if (m.insert(pack(b, x, y, z), (int32_t)no)) {
                out[4 * no] = b; out[4 * no + 1] = x; out[4 * no + 2] = y; out[4 * no + 3] = z; ++no;
            }
        }
    return no;
}

// neighbour table nbr[K][n_out]: conv: in = out + off*ts ; transposed: in = out - off*ts
void ref_kernel_map(const int32_t *Cin, long n_in, const int32_t *Cout, long n_out, int ksize, int ts,
                    int transposed, int32_t *nbr) {
    CoordMap m((size_t)n_in);
    for (long i = 0; i < n_in; ++i) m.insert(pack(Cin[4 * i], Cin[4 * i + 1], Cin[4 * i + 2], Cin[4 * i + 3]), (int32_t)i);
    std::vector<int> off; offsets_for(ksize, off);
    int K = ksize * ksize * ksize;
    int sgn = transposed ? -1 : 1;
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k) {
        int dx = sgn * off[3 * k] * ts, dy = sgn * off[3 * k + 1] * ts, dz = sgn * off[3 * k + 2] * ts;
        int32_t *row = nbr + (long)k * n_out;
        for (long o = 0; o < n_out; ++o)
            row[o] = m.find(pack(Cout[4 * o], Cout[4 * o + 1] + dx, Cout[4 * o + 2] + dy, Cout[4 * o + 3] + dz));
    }
}

// out[n_out][cout] = bias + sum_k scatter( gather(F, in_k) @ W[k] )   (fp32, ME CPU algorithm)
// act: 0 none, 1 relu, 2 leaky-relu(0.01).
// Threading: ME runs the offsets one after another and leaves the threading to OpenBLAS, which does not pay
// for the small per-offset GEMMs; to give the CPU baseline every core, output rows are split into blocks and
// each OpenMP thread runs the whole gather -> sgemm -> scatter-add sequence for its block with a
// single-threaded sgemm (callers set the BLAS thread count to 1).  Per-row arithmetic order is unchanged.
int ref_conv(const float *F, long n_in, int cin, const int32_t *nbr, int K, long n_out, const float *W,
             const float *bias, int cout, int act, float *out) {
    if (!g_sgemm) return 1;
    (void)n_in;
    const long BLK = 2048;
    const long n_blocks = (n_out + BLK - 1) / BLK;
#pragma omp parallel
    {
        std::vector<int32_t> in_rows, out_rows;
        std::vector<float> A, Y;
#pragma omp for schedule(dynamic, 1)
        for (long blk = 0; blk < n_blocks; ++blk) {
            const long o0 = blk * BLK, o1 = std::min(n_out, o0 + BLK);
            for (long o = o0; o < o1; ++o)
                for (int c = 0; c < cout; ++c) out[o * cout + c] = bias ? bias[c] : 0.f;
            for (int k = 0; k < K; ++k) {
                const int32_t *row = nbr + (long)k * n_out;
                in_rows.clear(); out_rows.clear();
                for (long o = o0; o < o1; ++o)
                    if (row[o] >= 0) { in_rows.push_back(row[o]); out_rows.push_back((int32_t)o); }
                const long P = (long)in_rows.size();
                if (P == 0) continue;
                A.resize((size_t)P * cin); Y.resize((size_t)P * cout);
                for (long p = 0; p < P; ++p) memcpy(&A[(size_t)p * cin], F + (size_t)in_rows[p] * cin, sizeof(float) * cin);
                g_sgemm(101 /*RowMajor*/, 111, 111, (int)P, cout, cin, 1.f, A.data(), cin, W + (size_t)k * cin * cout, cout,
                        0.f, Y.data(), cout);
                for (long p = 0; p < P; ++p) {
                    float *dst = out + (size_t)out_rows[p] * cout;
                    const float *src = &Y[(size_t)p * cout];
                    for (int c = 0; c < cout; ++c) dst[c] += src[c];
                }
            }
            if (act)
                for (long i = o0 * cout; i < o1 * (long)cout; ++i) {
                    float v = out[i];
                    out[i] = v > 0.f ? v : (act == 1 ? 0.f : 0.01f * v);
                }
        }
    }
    return 0;
}

// GDN1 / IGDN1: norm = beta + |x| @ gamma^T ; out = x / norm (or x * norm)   (model/blocks.py:26-57)
int ref_gdn(const float *X, long n, int c, const float *beta, const float *gamma, int inverse, float *out) {
    if (!g_sgemm) return 1;
    std::vector<float> A((size_t)n * c), Nrm((size_t)n * c);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n * (long)c; ++i) A[i] = std::fabs(X[i]);
    // Nrm = A (n x c) @ gamma^T (c x c)
    g_sgemm(101, 111, 112, (int)n, c, c, 1.f, A.data(), c, gamma, c, 0.f, Nrm.data(), c);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i)
        for (int j = 0; j < c; ++j) {
            float nv = Nrm[i * c + j] + beta[j];
            out[i * c + j] = inverse ? X[i * c + j] * nv : X[i * c + j] * (1.0f / nv);
        }
    return 0;
}

}  // extern "C"
