// Nearest-neighbour association of two voxel clouds and the per-point error terms of the reference's PSNR metric
// (metrics/metric.py:36-43 2-NN association, :107-166 D1 / colour errors) -- SURVEY.md §8 row a15.
//
// The reference walks a KD-tree per point in Python.  Both clouds here live on the integer voxel grid, so the
// neighbour of a query is found by probing the other cloud's coordinate hash in growing Chebyshev shells: shell r holds
// the lattice points with max(|dx|,|dy|,|dz|) = r, all at squared distance >= r^2, so the search stops after shell r as
// soon as (r+1)^2 > best (ties at exactly `best` are still collected: the reference averages the colours of
// equidistant neighbours).  On a decoded surface the neighbour sits 0-2 voxels away: 1-125 probes per point.
//
// Tie rule (metric.py:127-146): when the two nearest neighbours are equidistant the reference averages the colours of
// the equidistant points among the FIVE nearest (its search loop always stops at max_search = 5) with the first
// neighbour counted twice; which of the tied points a KD-tree lists first is arbitrary.  Here the tied points are taken
// in ascending row order (first = smallest row, at most five), which is one of the orders the reference may see.
#include "pcg_common.cuh"

struct NnBest {
    int d2;          // best squared distance (integer lattice)
    int n;           // tied points kept (<= 5)
    int row[5];      // their rows, ascending
};

__device__ __forceinline__ void nn_offer(NnBest &b, int d2, int row) {
    if (d2 < b.d2) { b.d2 = d2; b.n = 1; b.row[0] = row; return; }
    if (d2 != b.d2) return;
    // insert in ascending row order, keep the five smallest
    int pos = b.n < 5 ? b.n : 5;
    while (pos > 0 && b.row[pos - 1] > row) --pos;
    if (pos >= 5) return;
    const int last = b.n < 5 ? b.n : 4;
    for (int j = last; j > pos; --j) b.row[j] = b.row[j - 1];
    b.row[pos] = row;
    if (b.n < 5) ++b.n;
}

// metric.py:148-149 + :171-189: 8-bit rounding, the truncating uint8 cast, BT.709 -- in double like numpy
__device__ __forceinline__ void nn_yuv(const double (&rgb)[3], float (&yuv)[3]) {
    double q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        double v = rint(rgb[c] * 255.0) / 255.0;
        v = v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v);
        q[c] = (double)(unsigned char)(v * 255.0);
    }
    const float y = (float)(0.2126 * q[0] + 0.7152 * q[1] + 0.0722 * q[2]);
    const float u = (float)(-0.1146 * q[0] - 0.3854 * q[1] + 0.5 * q[2]);
    const float v = (float)(0.5 * q[0] - 0.4542 * q[1] - 0.0458 * q[2]);
    yuv[0] = y / 255.0f;
    yuv[1] = u / 255.0f + 0.5f;
    yuv[2] = v / 255.0f + 0.5f;
}

__global__ void k_nn_voxel(const uint64_t *__restrict__ keys, const int32_t *__restrict__ vals, uint64_t mask,
                           const float *__restrict__ b_colors, const int32_t *__restrict__ a_coords,
                           const float *__restrict__ a_colors, long na, int max_radius, int32_t *__restrict__ nn_index,
                           int32_t *__restrict__ nn_d2, float *__restrict__ yuv_sqerr, int32_t *__restrict__ d_status) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na) return;
    const int4 a = pcg_ld_coord(a_coords, i);
    NnBest best;
    best.d2 = 0x7fffffff; best.n = 0;
    for (int r = 0; r <= max_radius; ++r) {
        if (best.n > 0 && (long)r * r > (long)best.d2) break;          // shell r is entirely farther than the best so far
        for (int dx = -r; dx <= r; ++dx)
            for (int dy = -r; dy <= r; ++dy) {
                const bool face = (dx == -r || dx == r || dy == -r || dy == r);
                const int step = face ? 1 : (r > 0 ? 2 * r : 1);        // interior of the slab: only dz = -r and +r
                for (int dz = -r; dz <= r; dz += step) {
                    const int d2 = dx * dx + dy * dy + dz * dz;
                    if (d2 > best.d2) continue;
                    const int x = a.y + dx, y = a.z + dy, z = a.w + dz;
                    if (!pcg_in_range(a.x, x, y, z)) continue;
                    const int row = pcg_table_find(keys, vals, mask, pcg_pack(a.x, x, y, z));
                    if (row >= 0) nn_offer(best, d2, row);
                }
            }
    }
    if (best.n == 0) {
        atomicOr(d_status, 1);                                       // nothing within max_radius
        nn_index[i] = -1; nn_d2[i] = -1;
        if (yuv_sqerr) { yuv_sqerr[3 * i] = 0.f; yuv_sqerr[3 * i + 1] = 0.f; yuv_sqerr[3 * i + 2] = 0.f; }
        return;
    }
    nn_index[i] = best.row[0];
    nn_d2[i] = best.d2;
    if (!yuv_sqerr) return;
    double ca[3], cb[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        ca[c] = (double)__ldg(a_colors + 3 * i + c);
        cb[c] = (double)__ldg(b_colors + 3 * (long)best.row[0] + c);
    }
    if (best.n > 1) {                                                // first neighbour counted twice (metric.py:143-145)
        for (int j = 0; j < best.n; ++j)
#pragma unroll
            for (int c = 0; c < 3; ++c) cb[c] += (double)__ldg(b_colors + 3 * (long)best.row[j] + c);
#pragma unroll
        for (int c = 0; c < 3; ++c) cb[c] /= (double)(best.n + 1);
    }
    float ya[3], yb[3];
    nn_yuv(ca, ya);
    nn_yuv(cb, yb);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float d = ya[c] - yb[c];
        yuv_sqerr[3 * i + c] = d * d;
    }
}

extern "C" int pcg_nn_search(const uint64_t *keys, const int32_t *vals, long cap, const float *b_colors,
                             const int32_t *a_coords, const float *a_colors, long na, int max_radius, int32_t *nn_index,
                             int32_t *nn_d2, float *yuv_sqerr, int32_t *d_status, void *stream) {
    PCG_CHECK_ARG(cap > 0 && (cap & (cap - 1)) == 0, "pcg_nn_search: cap=%ld must be a power of two", cap);
    PCG_CHECK_ARG(max_radius >= 0 && max_radius <= 1024, "pcg_nn_search: max_radius=%d", max_radius);
    PCG_CHECK_ARG(yuv_sqerr == nullptr || (a_colors && b_colors), "pcg_nn_search: colour errors need both colour arrays");
    cudaStream_t st = as_stream(stream);
    PCG_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    if (na <= 0) return 0;
    k_nn_voxel<<<(unsigned)cdiv(na, 128), 128, 0, st>>>(keys, vals, (uint64_t)cap - 1, b_colors, a_coords, a_colors, na, max_radius,
                                                       nn_index, nn_d2, yuv_sqerr, d_status);
    PCG_LAUNCHED();
    return 0;
}
