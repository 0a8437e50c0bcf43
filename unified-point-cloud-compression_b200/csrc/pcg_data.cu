// Training-time augmentation of a cube on the GPU -- SURVEY.md §8 (f4): data/transform.py:33-55 (ColorJitter, a wrapper of
// torchvision.transforms.ColorJitter over the colours seen as a 3 x N x 1 image) and data/transform.py:57-124 (RandomRotate:
// rotate about the cube centre, floor, de-duplicate).  The reference runs these per sample on 12 CPU workers; here a sample
// is two launches (+ the existing unique-first kernels for the de-duplication), so the 8-GPU step is not loader bound.
#include "pcg_common.cuh"

// ---------------------------------------------------------------------------- colour jitter
// torchvision semantics (functional tensor ops on float images in [0,1]):
//   brightness  clamp(f * x)                          contrast    clamp(f * x + (1-f) * mean(gray(x)))   (mean over the image)
//   saturation  clamp(f * x + (1-f) * gray(x))        hue         rgb -> hsv, h = (h + f) mod 1, hsv -> rgb
//   gray = 0.2989 r + 0.587 g + 0.114 b;  the four ops run in a random order drawn per sample.
// The contrast op needs a mean over ALL points of the image as it stands when the op is reached: pass 1 applies the ops before
// it and writes per-block partial sums (fixed order, double), one block folds them, pass 2 redoes the cheap prefix and finishes.
struct JitterParams {
    int order[4];        // op ids in application order: 0 brightness, 1 contrast, 2 saturation, 3 hue
    float factor[4];     // indexed by op id
};

__device__ __forceinline__ float jit_clamp01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }
__device__ __forceinline__ float jit_gray(const float (&c)[3]) { return 0.2989f * c[0] + 0.587f * c[1] + 0.114f * c[2]; }

__device__ __forceinline__ void jit_blend(float (&c)[3], float other, float f) {
#pragma unroll
    for (int j = 0; j < 3; ++j) c[j] = jit_clamp01(f * c[j] + (1.0f - f) * other);
}

__device__ __forceinline__ void jit_hue(float (&c)[3], float shift) {
    const float r = c[0], g = c[1], b = c[2];
    const float maxc = fmaxf(r, fmaxf(g, b)), minc = fminf(r, fminf(g, b));
    const bool eq = maxc == minc;
    const float cr = maxc - minc;
    const float s = cr / (eq ? 1.f : maxc);
    const float div = eq ? 1.f : cr;
    const float rc = (maxc - r) / div, gc = (maxc - g) / div, bc = (maxc - b) / div;
    float h = 0.f;
    if (maxc == r) h = bc - gc;
    else if (maxc == g) h = 2.0f + rc - bc;
    else h = 4.0f + gc - rc;
    h = fmodf(h / 6.0f + 1.0f, 1.0f);
    // python's % on floats: result takes the sign of the divisor
    h = h + shift;
    h = h - floorf(h);
    const float v = maxc;
    const float h6 = h * 6.0f;
    const float fl = floorf(h6);
    const float f = h6 - fl;
    int i = (int)fl % 6;
    if (i < 0) i += 6;
    const float p = jit_clamp01(v * (1.0f - s));
    const float q = jit_clamp01(v * (1.0f - s * f));
    const float t = jit_clamp01(v * (1.0f - s * (1.0f - f)));
    switch (i) {
        case 0: c[0] = v; c[1] = t; c[2] = p; break;
        case 1: c[0] = q; c[1] = v; c[2] = p; break;
        case 2: c[0] = p; c[1] = v; c[2] = t; break;
        case 3: c[0] = p; c[1] = q; c[2] = v; break;
        case 4: c[0] = t; c[1] = p; c[2] = v; break;
        default: c[0] = v; c[1] = p; c[2] = q; break;
    }
}

// applies ops [first, last) of the order; `mean` is used by the contrast op
__device__ __forceinline__ void jit_apply(float (&c)[3], const JitterParams &p, int first, int last, float mean) {
    for (int s = first; s < last; ++s) {
        const int op = p.order[s];
        const float f = p.factor[op];
        if (op == 0) jit_blend(c, 0.f, f);
        else if (op == 1) jit_blend(c, mean, f);
        else if (op == 2) jit_blend(c, jit_gray(c), f);
        else jit_hue(c, f);
    }
}

__global__ void __launch_bounds__(256) k_jitter_partial(const float *__restrict__ colors, long n, JitterParams p, int contrast_at,
                                                        double *__restrict__ partials) {
    __shared__ double s_sum[256];
    double acc = 0.0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        float c[3] = {colors[3 * i], colors[3 * i + 1], colors[3 * i + 2]};
        jit_apply(c, p, 0, contrast_at, 0.f);
        acc += (double)jit_gray(c);
    }
    s_sum[threadIdx.x] = acc;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) s_sum[threadIdx.x] += s_sum[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = s_sum[0];
}

__global__ void k_jitter_fold(double *__restrict__ partials, int n_partials, long n) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < n_partials; ++i) s += partials[i];
        partials[n_partials] = s / (double)n;
    }
}

__global__ void __launch_bounds__(256) k_jitter_apply(const float *__restrict__ colors, long n, JitterParams p,
                                                      const double *__restrict__ mean_slot, float *__restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float mean = mean_slot ? (float)*mean_slot : 0.f;
    float c[3] = {colors[3 * i], colors[3 * i + 1], colors[3 * i + 2]};
    jit_apply(c, p, 0, 4, mean);      // the prefix before the contrast op is recomputed: cheaper than a round trip through HBM
    out[3 * i] = c[0]; out[3 * i + 1] = c[1]; out[3 * i + 2] = c[2];
}

static constexpr int JIT_BLOCKS = 296;        // 2 per SM

extern "C" size_t pcg_color_jitter_scratch_bytes(void) { return (JIT_BLOCKS + 1) * sizeof(double); }

extern "C" int pcg_color_jitter(const float *colors, long n, int op0, int op1, int op2, int op3, float brightness, float contrast,
                                float saturation, float hue, float *out, void *scratch, void *stream) {
    JitterParams p;
    p.order[0] = op0; p.order[1] = op1; p.order[2] = op2; p.order[3] = op3;
    p.factor[0] = brightness; p.factor[1] = contrast; p.factor[2] = saturation; p.factor[3] = hue;
    int seen = 0, contrast_at = -1;
    for (int s = 0; s < 4; ++s) {
        PCG_CHECK_ARG(p.order[s] >= 0 && p.order[s] < 4, "pcg_color_jitter: op id %d", p.order[s]);
        seen |= 1 << p.order[s];
        if (p.order[s] == 1) contrast_at = s;
    }
    PCG_CHECK_ARG(seen == 15, "pcg_color_jitter: the order must be a permutation of 0..3 (mask %d)", seen);
    PCG_CHECK_ARG(scratch != nullptr, "pcg_color_jitter: scratch of pcg_color_jitter_scratch_bytes() is required");
    if (n <= 0) return 0;
    cudaStream_t st = as_stream(stream);
    double *partials = reinterpret_cast<double *>(scratch);
    const int blocks = (int)(cdiv(n, 256) < JIT_BLOCKS ? cdiv(n, 256) : JIT_BLOCKS);
    k_jitter_partial<<<blocks, 256, 0, st>>>(colors, n, p, contrast_at, partials);
    PCG_LAUNCHED();
    k_jitter_fold<<<1, 32, 0, st>>>(partials, blocks, n);
    PCG_LAUNCHED();
    k_jitter_apply<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(colors, n, p, partials + blocks, out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- rotate + floor
// data/transform.py:78-99: p' = (p - half) @ R^T + half, then ME.utils.sparse_quantize (floor, first occurrence kept).
// Writes the floored rows as [0, x, y, z] int32 (the layout pcg_unique_first consumes) and, with crop_size > 0, a keep flag
// for rows whose rotated position lies in [0, crop_size)^3 (transform.py:86-89).
__global__ void k_rotate_floor(const float *__restrict__ points, long n, float r00, float r01, float r02, float r10, float r11,
                               float r12, float r20, float r21, float r22, float half, float crop_size,
                               int32_t *__restrict__ out_coords, uint8_t *__restrict__ keep) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = points[3 * i] - half, y = points[3 * i + 1] - half, z = points[3 * i + 2] - half;
    // row i of (P - half) @ R^T: component a = sum_k p_k * R[a][k], accumulated k = 0, 1, 2 without contraction
    const float qx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, r00), __fmul_rn(y, r01)), __fmul_rn(z, r02)), half);
    const float qy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, r10), __fmul_rn(y, r11)), __fmul_rn(z, r12)), half);
    const float qz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, r20), __fmul_rn(y, r21)), __fmul_rn(z, r22)), half);
    reinterpret_cast<int4 *>(out_coords)[i] = make_int4(0, (int)floorf(qx), (int)floorf(qy), (int)floorf(qz));
    if (keep) keep[i] = (crop_size <= 0.f || (qx >= 0.f && qx < crop_size && qy >= 0.f && qy < crop_size && qz >= 0.f && qz < crop_size)) ? 1 : 0;
}

extern "C" int pcg_rotate_floor(const float *points, long n, const float *rotation_host9, float half, float crop_size,
                                int32_t *out_coords, uint8_t *keep, void *stream) {
    PCG_CHECK_ARG(rotation_host9 != nullptr, "pcg_rotate_floor: rotation matrix (9 floats, host, row-major) is required");
    PCG_CHECK_ARG(((uintptr_t)out_coords & 15) == 0, "pcg_rotate_floor: out_coords must be 16-byte aligned");
    if (n <= 0) return 0;
    const float *r = rotation_host9;
    k_rotate_floor<<<(unsigned)cdiv(n, 256), 256, 0, as_stream(stream)>>>(points, n, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], r[8],
                                                                        half, crop_size, out_coords, keep);
    PCG_LAUNCHED();
    return 0;
}
