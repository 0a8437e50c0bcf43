// Activation, GDN1, occupancy top-k (radix select), trilinear feature lookup, bf16 cast.
#include "pcg_compact.cuh"

// ---------------------------------------------------------------------------- activation / cast
__global__ void k_activation(const float *__restrict__ x, long n, int act, float *__restrict__ out) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = pcg_apply_act(x[i], act);
}

extern "C" int pcg_activation(const float *x, long n, int act, float *out, void *stream) {
    if (n <= 0) return 0;
    long blocks = cdiv(n, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_activation<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(x, n, act, out);
    PCG_LAUNCHED();
    return 0;
}

__global__ void k_f32_to_bf16(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst, long n) {
    long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    long stride = (long)gridDim.x * blockDim.x * 4;
    for (; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = __ldg(reinterpret_cast<const float4 *>(src + i));
            __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t *>(&lo);
            pk.y = *reinterpret_cast<uint32_t *>(&hi);
            *reinterpret_cast<uint2 *>(dst + i) = pk;
        } else {
            for (long j = i; j < n; ++j) dst[j] = __float2bfloat16_rn(src[j]);
        }
    }
}

extern "C" int pcg_f32_to_bf16(const float *src, void *dst, long n, void *stream) {
    if (n <= 0) return 0;
    PCG_CHECK_ARG(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 7) == 0, "pcg_f32_to_bf16: unaligned buffers");
    long blocks = cdiv(n, 256 * 4);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_f32_to_bf16<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(src, reinterpret_cast<__nv_bfloat16 *>(dst), n);
    PCG_LAUNCHED();
    return 0;
}

// rows of c floats -> rows of cp = roundup8(c) bf16 (zero padded): operand layout of pcg_conv_fwd_tc for odd widths
__global__ void k_f32_to_bf16_rows(const float *__restrict__ src, long n, int c, int cp, __nv_bfloat16 *__restrict__ dst) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long total = n * cp;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        long r = i / cp;
        int j = (int)(i - r * cp);
        dst[i] = __float2bfloat16_rn(j < c ? __ldg(src + r * c + j) : 0.f);
    }
}

extern "C" int pcg_f32_to_bf16_rows(const float *src, long n, int c, int cp, void *dst, void *stream) {
    if (n <= 0) return 0;
    PCG_CHECK_ARG(c >= 1 && cp >= c, "pcg_f32_to_bf16_rows: c=%d c_padded=%d", c, cp);
    if (cp == c) return pcg_f32_to_bf16(src, dst, n * c, stream);
    long blocks = cdiv(n * cp, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_f32_to_bf16_rows<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(src, n, c, cp, reinterpret_cast<__nv_bfloat16 *>(dst));
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- GDN1
// norm[n][j] = beta[j] + sum_i gamma[j][i] * |x[n][i]|;  out = x * (1/norm)  or  x * norm.
// 64x64 output tile per CTA, 16-deep smem stages, fixed ascending-i accumulation.
__global__ void __launch_bounds__(256) k_gdn(const float *__restrict__ x, long n, int c, const float *__restrict__ beta,
                                             const float *__restrict__ gamma, int inverse, float *__restrict__ out) {
    __shared__ __align__(16) float As[16][64 + 4];
    __shared__ __align__(16) float Bs[16][64 + 4];
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const long row0 = (long)blockIdx.x * 64;
    const int col0 = blockIdx.y * 64;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < c; k0 += 16) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int idx = t + 256 * e;
            {   // A: |x[row0 + r][k0 + kk]|, idx -> (r = idx/16, kk = idx%16) coalesced along kk
                int r = idx >> 4, kk = idx & 15;
                long o = row0 + r;
                As[kk][r] = (o < n && k0 + kk < c) ? fabsf(__ldg(x + o * c + k0 + kk)) : 0.f;
            }
            {   // B[kk][j] = gamma[col0 + j][k0 + kk]
                int j = idx >> 4, kk = idx & 15;
                Bs[kk][j] = (col0 + j < c && k0 + kk < c) ? __ldg(gamma + (long)(col0 + j) * c + k0 + kk) : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float4 a = *reinterpret_cast<const float4 *>(&As[kk][ty * 4]);
            float4 b = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        long o = row0 + ty * 4 + i;
        if (o >= n) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int col = col0 + tx * 4 + j;
            if (col < c) {
                float nv = acc[i][j] + __ldg(beta + col);
                float xv = __ldg(x + o * c + col);
                out[o * c + col] = inverse ? xv * nv : xv * (1.0f / nv);
            }
        }
    }
}

// Tensor-core version for C a multiple of 32 (<= 128): norm = beta + |x| @ gamma^T on mma.sync.m16n8k8 TF32 with the 3xTF32
// split (a_hi*b_hi + a_lo*b_hi + a_hi*b_lo, fp32 accumulate: ~2^-21 relative, i.e. fp32-grade -- a plain bf16/TF32
// product would perturb every quantised symbol downstream).  One warp = 16 rows x all C columns; lane (g, t) loads
// channels [32q+8t, +8) of rows g and g+8 as two float4 and uses them as the A fragments of four k8 steps (k-step j:
// slot t <- channel 32q+8t+j, slot t+4 <- channel 32q+8t+4+j); gamma sits in shared memory in the matching fragment
// order.  The op is bound by reading and writing x (8 B per element), not by the 2*N*C^2 FLOPs.
__device__ __forceinline__ uint32_t f2tf32(float f) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(f));
    return r;
}
__device__ __forceinline__ void mma_1688_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int C>
__global__ void __launch_bounds__(256) k_gdn_mma(const float *__restrict__ x, long n, const float *__restrict__ beta,
                                                 const float *__restrict__ gamma, int inverse, float *__restrict__ out) {
    constexpr int Q = C / 32, NT = C / 8;
    extern __shared__ float2 bs[];                  // [Q][4][NT][32]: {gamma[n][32q+8t+j], gamma[n][32q+8t+4+j]}, n = nt*8+g
    for (int i = threadIdx.x; i < Q * 4 * NT * 32; i += blockDim.x) {
        const int lane = i & 31, nt = (i >> 5) % NT, j = ((i >> 5) / NT) & 3, q = (i >> 5) / (NT * 4);
        const int g = lane >> 2, t = lane & 3, nn = nt * 8 + g;
        bs[i] = make_float2(__ldg(gamma + (long)nn * C + 32 * q + 8 * t + j), __ldg(gamma + (long)nn * C + 32 * q + 8 * t + 4 + j));
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // persistent: one CTA per SM walks the 128-row tiles, so the 64 KiB fragment image of gamma is built once per SM
    for (long tile = blockIdx.x; tile * 128 < n; tile += gridDim.x) {
    const long row0 = (tile * 8 + (threadIdx.x >> 5)) * 16;
    if (row0 >= n) continue;
    const long r0 = row0 + g, r1 = row0 + g + 8;
    const bool ok0 = r0 < n, ok1 = r1 < n;
    float acc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[nt][c] = 0.f;
#pragma unroll 1
    for (int q = 0; q < Q; ++q) {
        float xa[2][8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 u0 = ok0 ? __ldg(reinterpret_cast<const float4 *>(x + r0 * C + 32 * q + 8 * t) + h) : z;
            const float4 u1 = ok1 ? __ldg(reinterpret_cast<const float4 *>(x + r1 * C + 32 * q + 8 * t) + h) : z;
            xa[0][4 * h] = fabsf(u0.x); xa[0][4 * h + 1] = fabsf(u0.y); xa[0][4 * h + 2] = fabsf(u0.z); xa[0][4 * h + 3] = fabsf(u0.w);
            xa[1][4 * h] = fabsf(u1.x); xa[1][4 * h + 1] = fabsf(u1.y); xa[1][4 * h + 2] = fabsf(u1.z); xa[1][4 * h + 3] = fabsf(u1.w);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            // A fragment: a0 = (row g, slot t), a1 = (row g+8, slot t), a2 = (row g, slot t+4), a3 = (row g+8, slot t+4)
            const float af[4] = {xa[0][j], xa[1][j], xa[0][4 + j], xa[1][4 + j]};
            uint32_t ah[4], al[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ah[i] = f2tf32(af[i]);
                al[i] = f2tf32(af[i] - __uint_as_float(ah[i]));
            }
            const float2 *bq = bs + ((q * 4 + j) * NT) * 32 + lane;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const float2 b = bq[nt * 32];
                const uint32_t bh0 = f2tf32(b.x), bh1 = f2tf32(b.y);
                const uint32_t bl0 = f2tf32(b.x - __uint_as_float(bh0)), bl1 = f2tf32(b.y - __uint_as_float(bh1));
                mma_1688_tf32(acc[nt], al[0], al[1], al[2], al[3], bh0, bh1);
                mma_1688_tf32(acc[nt], ah[0], ah[1], ah[2], ah[3], bl0, bl1);
                mma_1688_tf32(acc[nt], ah[0], ah[1], ah[2], ah[3], bh0, bh1);
            }
        }
    }
    // C fragment: c0,c1 = (row g, cols 2t,2t+1), c2,c3 = (row g+8, same), n tile nt adds 8 columns
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int col = nt * 8 + 2 * t;
        const float2 bt = __ldg(reinterpret_cast<const float2 *>(beta + col));
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const long r = h ? r1 : r0;
            if (!(h ? ok1 : ok0)) continue;
            const float2 xv = __ldg(reinterpret_cast<const float2 *>(x + r * C + col));
            const float n0 = acc[nt][2 * h] + bt.x, n1 = acc[nt][2 * h + 1] + bt.y;
            float2 o;
            o.x = inverse ? xv.x * n0 : xv.x * (1.0f / n0);
            o.y = inverse ? xv.y * n1 : xv.y * (1.0f / n1);
            *reinterpret_cast<float2 *>(out + r * C + col) = o;
        }
    }
    }
}

// Same op on mma.sync.m16n8k16 with the bf16x3 split (x = x_hi + x_lo, gamma = g_hi + g_lo in bf16; x_lo*g_hi + x_hi*g_lo +
// x_hi*g_hi, fp32 accumulate: ~2^-16 relative per term, still far inside fp32-conv noise).  Per 16 reduction channels this is
// 3 MMAs + one 16-byte fragment load, against 6 MMAs + 12 conversion instructions for 3xTF32 whose gamma split was redone in
// the inner loop.  Lane (g, t) owns channels [32q+8t, +8) of rows g and g+8; slab s of the pair takes its channels 4s..4s+3
// as k-slots {2t, 2t+1, 2t+8, 2t+9}; the gamma fragments in shared memory follow the same permutation.
__device__ __forceinline__ void mma_16816_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void split_bf16x2(float v0, float v1, uint32_t &hi, uint32_t &lo) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v0, v1);
    const __nv_bfloat162 l = __floats2bfloat162_rn(v0 - __low2float(h), v1 - __high2float(h));
    hi = *reinterpret_cast<const uint32_t *>(&h);
    lo = *reinterpret_cast<const uint32_t *>(&l);
}

// SPLIT = 3: bf16x3 (fp32-grade); SPLIT = 1: the hi x hi product only (bf16-grade norm: the mode whose convolutions run on bf16
// operands anyway).  out_bf16 (optional): the same rows also as the bf16 operand image of the next tensor-core layer.
template <int C, int SPLIT>
__global__ void __launch_bounds__(256) k_gdn_mma16(const float *__restrict__ x, long n, const float *__restrict__ beta,
                                                   const float *__restrict__ gamma, int inverse, float *__restrict__ out,
                                                   __nv_bfloat16 *__restrict__ out_bf16) {
    constexpr int Q = C / 32, NT = C / 8;
    extern __shared__ uint4 bq_s[];                 // [Q][2][NT][32]: {b0_hi, b1_hi, b0_lo, b1_lo} of output channel nt*8+g
    for (int i = threadIdx.x; i < Q * 2 * NT * 32; i += blockDim.x) {
        const int lane = i & 31, nt = (i >> 5) % NT, sl = ((i >> 5) / NT) & 1, q = (i >> 5) / (NT * 2);
        const int g = lane >> 2, t = lane & 3;
        const float *gr = gamma + (long)(nt * 8 + g) * C + 32 * q + 8 * t + 4 * sl;
        uint4 v;
        split_bf16x2(__ldg(gr), __ldg(gr + 1), v.x, v.z);
        split_bf16x2(__ldg(gr + 2), __ldg(gr + 3), v.y, v.w);
        bq_s[i] = v;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    for (long tile = blockIdx.x; tile * 128 < n; tile += gridDim.x) {
        const long row0 = (tile * 8 + (threadIdx.x >> 5)) * 16;
        if (row0 >= n) continue;
        const long r0 = row0 + g, r1 = row0 + g + 8;
        const bool ok0 = r0 < n, ok1 = r1 < n;
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[nt][c] = 0.f;
#pragma unroll 1
        for (int q = 0; q < Q; ++q) {
#pragma unroll
            for (int sl = 0; sl < 2; ++sl) {
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 u0 = ok0 ? __ldg(reinterpret_cast<const float4 *>(x + r0 * C + 32 * q + 8 * t) + sl) : z;
                const float4 u1 = ok1 ? __ldg(reinterpret_cast<const float4 *>(x + r1 * C + 32 * q + 8 * t) + sl) : z;
                uint32_t ah[4], al[4];      // a0 = (row g, slots 2t,2t+1), a1 = (row g+8, same), a2 = (row g, slots 2t+8,2t+9), a3 = (row g+8, same)
                split_bf16x2(fabsf(u0.x), fabsf(u0.y), ah[0], al[0]);
                split_bf16x2(fabsf(u1.x), fabsf(u1.y), ah[1], al[1]);
                split_bf16x2(fabsf(u0.z), fabsf(u0.w), ah[2], al[2]);
                split_bf16x2(fabsf(u1.z), fabsf(u1.w), ah[3], al[3]);
                const uint4 *bq = bq_s + ((q * 2 + sl) * NT) * 32 + lane;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const uint4 b = bq[nt * 32];
                    if (SPLIT == 3) {
                        mma_16816_bf16(acc[nt], al[0], al[1], al[2], al[3], b.x, b.y);
                        mma_16816_bf16(acc[nt], ah[0], ah[1], ah[2], ah[3], b.z, b.w);
                    }
                    mma_16816_bf16(acc[nt], ah[0], ah[1], ah[2], ah[3], b.x, b.y);
                }
            }
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            const float2 bt = __ldg(reinterpret_cast<const float2 *>(beta + col));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const long r = h ? r1 : r0;
                if (!(h ? ok1 : ok0)) continue;
                const float2 xv = __ldg(reinterpret_cast<const float2 *>(x + r * C + col));
                const float n0 = acc[nt][2 * h] + bt.x, n1 = acc[nt][2 * h + 1] + bt.y;
                float2 o;
                o.x = inverse ? xv.x * n0 : xv.x * (1.0f / n0);
                o.y = inverse ? xv.y * n1 : xv.y * (1.0f / n1);
                *reinterpret_cast<float2 *>(out + r * C + col) = o;
                if (out_bf16) *reinterpret_cast<__nv_bfloat162 *>(out_bf16 + r * C + col) = __floats2bfloat162_rn(o.x, o.y);
            }
        }
    }
}

template <int C, int SPLIT>
static int launch_gdn_mma16(const float *x, long n, const float *beta, const float *gamma, int inverse, float *out, __nv_bfloat16 *out_bf16,
                            cudaStream_t st) {
    const size_t smem = (size_t)(C / 32) * 2 * (C / 8) * 32 * sizeof(uint4);
    PCG_CUDA(cudaFuncSetAttribute(k_gdn_mma16<C, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long tiles = cdiv(n, 128);
    k_gdn_mma16<C, SPLIT><<<(unsigned)(tiles < 2 * PCG_SMS ? tiles : 2 * PCG_SMS), 256, smem, st>>>(x, n, beta, gamma, inverse, out, out_bf16);
    PCG_LAUNCHED();
    return 0;
}

template <int C>
static int launch_gdn_mma(const float *x, long n, const float *beta, const float *gamma, int inverse, float *out, cudaStream_t st) {
    const size_t smem = (size_t)(C / 32) * 4 * (C / 8) * 32 * sizeof(float2);
    PCG_CUDA(cudaFuncSetAttribute(k_gdn_mma<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long tiles = cdiv(n, 128);
    k_gdn_mma<C><<<(unsigned)(tiles < 2 * PCG_SMS ? tiles : 2 * PCG_SMS), 256, smem, st>>>(x, n, beta, gamma, inverse, out);
    PCG_LAUNCHED();
    return 0;
}

// bf16-grade variant for the tensor-core mode: one bf16 product for |x| gamma^T (relative error of the norm ~2^-9, the precision
// of the convolutions around it) and, optionally, the result also as bf16 operand rows for the next layer (c in 32 / 64 / 128).
extern "C" int pcg_gdn_fwd_bf16(const float *x, long n, int c, const float *beta, const float *gamma, int inverse, float *out,
                                void *out_bf16, void *stream) {
    if (n <= 0) return 0;
    cudaStream_t st = as_stream(stream);
    PCG_CHECK_ARG(c == 128 || c == 64 || c == 32, "pcg_gdn_fwd_bf16: c=%d must be 32, 64 or 128", c);
    PCG_CHECK_ARG((((uintptr_t)x | (uintptr_t)out) & 15) == 0 && ((uintptr_t)beta & 7) == 0 && ((uintptr_t)out_bf16 & 3) == 0, "pcg_gdn_fwd_bf16: misaligned buffers");
    __nv_bfloat16 *ob = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    if (c == 128) return launch_gdn_mma16<128, 1>(x, n, beta, gamma, inverse, out, ob, st);
    if (c == 64) return launch_gdn_mma16<64, 1>(x, n, beta, gamma, inverse, out, ob, st);
    return launch_gdn_mma16<32, 1>(x, n, beta, gamma, inverse, out, ob, st);
}

extern "C" int pcg_gdn_fwd(const float *x, long n, int c, const float *beta, const float *gamma, int inverse, float *out,
                           void *stream) {
    if (n <= 0) return 0;
    cudaStream_t st = as_stream(stream);
    const bool aligned = (((uintptr_t)x | (uintptr_t)out) & 15) == 0 && ((uintptr_t)beta & 7) == 0;
    static const bool tf32 = getenv("PCG_GDN") && !strcmp(getenv("PCG_GDN"), "tf32");     // A/B knob: the 3xTF32 version
    if (aligned && !tf32) {
        if (c == 128) return launch_gdn_mma16<128, 3>(x, n, beta, gamma, inverse, out, nullptr, st);
        if (c == 64) return launch_gdn_mma16<64, 3>(x, n, beta, gamma, inverse, out, nullptr, st);
        if (c == 32) return launch_gdn_mma16<32, 3>(x, n, beta, gamma, inverse, out, nullptr, st);
    }
    if (aligned && c == 128) return launch_gdn_mma<128>(x, n, beta, gamma, inverse, out, st);
    if (aligned && c == 64) return launch_gdn_mma<64>(x, n, beta, gamma, inverse, out, st);
    if (aligned && c == 32) return launch_gdn_mma<32>(x, n, beta, gamma, inverse, out, st);
    dim3 grid((unsigned)cdiv(n, 64), (unsigned)cdiv(c, 64));
    k_gdn<<<grid, 256, 0, st>>>(x, n, c, beta, gamma, inverse, out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- top-k mask (radix select)
struct TopkState {
    uint32_t prefix;       // selected high digits so far
    uint32_t prefix_mask;  // which bits of prefix are decided
    long long k_rem;       // how many still to take among keys matching prefix
    int hist[256];
    unsigned done;         // CTAs of the current histogram pass that have added their counts
};

__device__ __forceinline__ uint32_t pcg_order_key(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // larger float -> larger key
}

__global__ void k_topk_init(TopkState *s, long k) {
    if (threadIdx.x == 0) { s->prefix = 0; s->prefix_mask = 0; s->k_rem = k; s->done = 0; }
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s->hist[i] = 0;
}

// One radix pass: histogram of the digit at `shift` over the keys that match the digits chosen so far, then the LAST CTA to add
// its counts (ticket in s->done) picks the digit holding the k_rem-th largest key and clears the state for the next pass.  The
// counts are integers, so the result does not depend on which CTA comes last.  blockDim.x must be 256 (one thread per digit).
__global__ void __launch_bounds__(256) k_topk_hist(const float *__restrict__ logits, const int32_t *__restrict__ coords, long n, int batch,
                                                   int shift, TopkState *s) {
    __shared__ int h[256];
    __shared__ long long suf[257];
    __shared__ bool last;
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t prefix = s->prefix, pmask = s->prefix_mask;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        if (coords && __ldg(coords + 4 * i) != batch) continue;
        uint32_t key = pcg_order_key(__ldg(logits + i));
        if ((key & pmask) == prefix) atomicAdd(&h[(key >> shift) & 255], 1);
    }
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&s->hist[threadIdx.x], h[threadIdx.x]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&s->done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    // digit d = threadIdx.x; read through L2 (the line holding prefix / hist[0..] may sit stale in this SM's L1) and clear
    const int d = threadIdx.x;
    const long long cnt = atomicExch(&s->hist[d], 0);
    long long k = s->k_rem;                // read by every thread before the chosen one rewrites it (barriers below)
    suf[d] = cnt;
    if (d == 0) suf[256] = 0;
    __syncthreads();
    // suffix sums suf[d] = number of matching keys with digit >= d (Hillis-Steele over 256 entries)
    for (int off = 1; off < 256; off <<= 1) {
        long long add = d + off < 256 ? suf[d + off] : 0;
        __syncthreads();
        suf[d] += add;
        __syncthreads();
    }
    if (k > suf[0]) k = suf[0];            // fewer rows than k: take them all
    // the largest digit with at least k keys at or above it (k == 0 -> digit 255, nothing above)
    if (suf[d] >= k && (d == 255 || suf[d + 1] < k)) {
        s->prefix = prefix | ((uint32_t)d << shift);
        s->prefix_mask = pmask | (255u << shift);
        s->k_rem = k - suf[d + 1];         // to take among keys with this digit (0 when k == 0)
        s->done = 0;
    }
}

__global__ void k_topk_above(const float *__restrict__ logits, const int32_t *__restrict__ coords, long n, int batch,
                             const TopkState *s, uint8_t *mask) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (coords && __ldg(coords + 4 * i) != batch) return;
    uint32_t key = pcg_order_key(__ldg(logits + i));
    // strictly above the threshold key; with k_rem == 0 and the all-ones prefix nothing qualifies
    mask[i] = (key > s->prefix) ? 1 : 0;
}

struct FlagTie {
    const float *logits;
    const int32_t *coords;
    int batch;
    const TopkState *s;
    __device__ bool operator()(long i) const {
        if (coords && __ldg(coords + 4 * i) != batch) return false;
        return pcg_order_key(__ldg(logits + i)) == s->prefix;
    }
};
struct EmitTie {
    const TopkState *s;
    uint8_t *mask;
    __device__ void operator()(long i, int rank) const {
        if ((long long)rank < s->k_rem) mask[i] = 1;
    }
};

extern "C" int pcg_topk_mask(const float *logits, const int32_t *coords, long n, int batch, long k, uint8_t *mask,
                             void *scratch, size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    if (n <= 0) return 0;
    PCG_CHECK_ARG(k >= 0, "pcg_topk_mask: k=%ld", k);
    size_t need = pcg_compact_scratch_bytes(n) + 2048;
    PCG_CHECK_ARG(scratch && scratch_bytes >= need, "pcg_topk_mask: scratch too small (%zu < %zu)", scratch_bytes, need);
    PcgScratch cs;
    pcg_carve_scratch(scratch, scratch_bytes, n, cs);
    uintptr_t sp = (uintptr_t)scratch + pcg_compact_scratch_bytes(n);
    sp = (sp + 255) & ~(uintptr_t)255;
    TopkState *s = reinterpret_cast<TopkState *>(sp);
    int32_t *d_dummy = reinterpret_cast<int32_t *>(sp + 1536);
    k_topk_init<<<1, 256, 0, st>>>(s, k);
    PCG_LAUNCHED();
    long blocks = cdiv(n, 256 * 4);
    if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
    for (int shift = 24; shift >= 0; shift -= 8) {
        k_topk_hist<<<(unsigned)blocks, 256, 0, st>>>(logits, coords, n, batch, shift, s);
        PCG_LAUNCHED();
    }
    k_topk_above<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(logits, coords, n, batch, s, mask);
    PCG_LAUNCHED();
    return pcg_compact(n, FlagTie{logits, coords, batch, s}, EmitTie{s, mask}, d_dummy, cs, st);
}

// ---------------------------------------------------------------------------- decoder output rows
// out[i] = [x, y, z, clamp(round(f * 255), 0, 255) / 255] (model/model.py:247-250: 8-bit colours kept as floats; round = half to
// even like torch.round, true division like torch's `/ 255`).  One pass instead of round / clamp / div / cast / cat.
__global__ void k_finalize(const int32_t *__restrict__ coords, const float *__restrict__ feats, int c, long n, float *__restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int4 p = pcg_ld_coord(coords, i);
    float *o = out + i * (3 + c);
    o[0] = (float)p.y; o[1] = (float)p.z; o[2] = (float)p.w;
    for (int j = 0; j < c; ++j) o[3 + j] = fminf(fmaxf(rintf(__ldg(feats + i * c + j) * 255.0f), 0.0f), 255.0f) / 255.0f;
}

extern "C" int pcg_finalize(const int32_t *coords, const float *feats, int c, long n, float *out, void *stream) {
    PCG_CHECK_ARG(c >= 1 && c <= 16, "pcg_finalize: c=%d", c);
    if (n <= 0) return 0;
    k_finalize<<<(unsigned)cdiv(n, 256), 256, 0, as_stream(stream)>>>(coords, feats, c, n, out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- top-k for all batch items at once
// The same radix select with one TopkState per batch item: a pass reads every row once (its batch index = first coordinate),
// whatever the number of batch items (a frame coded as B batched blocks, or a training batch, would otherwise pay B x 7 launches
// and B full passes per level).  Ties at a batch item's threshold are ranked in row order like the single-batch path: one
// compaction lists the rows whose key equals their item's threshold, and one thread per batch item walks that (normally
// B-entry) list and takes the first k_rem.
#define TOPKB_MAX 32
struct TopkKs {
    long long k[TOPKB_MAX];
};

__global__ void k_topkb_init(TopkState *s, TopkKs ks, int nb) {
    const int b = blockIdx.x;
    if (b >= nb) return;
    if (threadIdx.x == 0) { s[b].prefix = 0; s[b].prefix_mask = 0; s[b].k_rem = ks.k[b]; s[b].done = 0; }
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s[b].hist[i] = 0;
}

__global__ void __launch_bounds__(256) k_topkb_hist(const float *__restrict__ logits, const int32_t *__restrict__ coords, long n, int nb,
                                                    int shift, TopkState *s) {
    __shared__ int h[TOPKB_MAX][256];
    __shared__ uint32_t s_prefix[TOPKB_MAX], s_pmask[TOPKB_MAX];
    __shared__ long long suf[257];
    __shared__ bool last;
    for (int b = 0; b < nb; ++b) h[b][threadIdx.x] = 0;
    if (threadIdx.x < nb) { s_prefix[threadIdx.x] = s[threadIdx.x].prefix; s_pmask[threadIdx.x] = s[threadIdx.x].prefix_mask; }
    __syncthreads();
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long stride = (long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int b = __ldg(coords + 4 * i);
        if (b < 0 || b >= nb) continue;
        const uint32_t key = pcg_order_key(__ldg(logits + i));
        if ((key & s_pmask[b]) == s_prefix[b]) atomicAdd(&h[b][(key >> shift) & 255], 1);
    }
    __syncthreads();
    for (int b = 0; b < nb; ++b)
        if (h[b][threadIdx.x]) atomicAdd(&s[b].hist[threadIdx.x], h[b][threadIdx.x]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&s[0].done, 1u) == gridDim.x - 1;       // one ticket for the whole pass
    __syncthreads();
    if (!last) return;
    __threadfence();
    const int d = threadIdx.x;
    for (int b = 0; b < nb; ++b) {
        const long long cnt = atomicExch(&s[b].hist[d], 0);
        long long k = s[b].k_rem;
        __syncthreads();
        suf[d] = cnt;
        if (d == 0) suf[256] = 0;
        __syncthreads();
        for (int off = 1; off < 256; off <<= 1) {
            long long add = d + off < 256 ? suf[d + off] : 0;
            __syncthreads();
            suf[d] += add;
            __syncthreads();
        }
        if (k > suf[0]) k = suf[0];
        if (suf[d] >= k && (d == 255 || suf[d + 1] < k)) {
            s[b].prefix = s_prefix[b] | ((uint32_t)d << shift);
            s[b].prefix_mask = s_pmask[b] | (255u << shift);
            s[b].k_rem = k - suf[d + 1];
        }
        __syncthreads();
    }
    if (d == 0) s[0].done = 0;
}

__global__ void k_topkb_above(const float *__restrict__ logits, const int32_t *__restrict__ coords, long n, int nb, const TopkState *s,
                              uint8_t *mask) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = __ldg(coords + 4 * i);
    if (b < 0 || b >= nb) return;
    mask[i] = (pcg_order_key(__ldg(logits + i)) > s[b].prefix) ? 1 : 0;
}

struct FlagTieB {
    const float *logits;
    const int32_t *coords;
    int nb;
    const TopkState *s;
    __device__ bool operator()(long i) const {
        const int b = __ldg(coords + 4 * i);
        return b >= 0 && b < nb && pcg_order_key(__ldg(logits + i)) == s[b].prefix;
    }
};
struct EmitTieB {
    int32_t *list;
    __device__ void operator()(long i, int rank) const { list[rank] = (int)i; }
};

// CTA b walks the (row-ordered) tie list in chunks of 256 and takes the first k_rem rows of batch item b: rank of an entry =
// ties of item b before it (running count + ballot ranks inside the chunk); stops as soon as k_rem are taken
__global__ void __launch_bounds__(256) k_topkb_ties(const int32_t *__restrict__ list, const int32_t *__restrict__ d_n_ties,
                                                    const int32_t *__restrict__ coords, const TopkState *s, uint8_t *mask) {
    __shared__ int warp_cnt[8];
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long want = s[b].k_rem;
    const int n = d_n_ties[0];
    long long taken = 0;
    for (int base = 0; base < n && taken < want; base += 256) {
        const int j = base + threadIdx.x;
        const int i = j < n ? list[j] : -1;
        const bool mine = i >= 0 && __ldg(coords + 4 * (long)i) == b;
        const unsigned bal = __ballot_sync(0xffffffffu, mine);
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            if (w < warp) before += warp_cnt[w];
            total += warp_cnt[w];
        }
        if (mine && taken + before + __popc(bal & ((1u << lane) - 1u)) < want) mask[i] = 1;
        taken += total;
        __syncthreads();
    }
}

extern "C" size_t pcg_topk_mask_batched_scratch_bytes(long n) {
    return pcg_compact_scratch_bytes(n) + 256 + TOPKB_MAX * sizeof(TopkState) + 256 + (size_t)(n > 0 ? n : 1) * sizeof(int32_t) + 256;
}

extern "C" int pcg_topk_mask_batched(const float *logits, const int32_t *coords, long n, int n_batches, const long *h_k, uint8_t *mask,
                                     void *scratch, size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    if (n <= 0) return 0;
    PCG_CHECK_ARG(n_batches >= 1 && n_batches <= TOPKB_MAX && h_k != nullptr && coords != nullptr, "pcg_topk_mask_batched: n_batches=%d must be in [1,%d]", n_batches, TOPKB_MAX);
    PCG_CHECK_ARG(n < 2147483647L, "pcg_topk_mask_batched: n=%ld exceeds int32", n);
    PCG_CHECK_ARG(scratch && scratch_bytes >= pcg_topk_mask_batched_scratch_bytes(n), "pcg_topk_mask_batched: scratch too small");
    TopkKs ks;
    for (int b = 0; b < TOPKB_MAX; ++b) ks.k[b] = b < n_batches ? (long long)h_k[b] : 0;
    for (int b = 0; b < n_batches; ++b) PCG_CHECK_ARG(h_k[b] >= 0, "pcg_topk_mask_batched: k[%d]=%ld", b, h_k[b]);
    PcgScratch cs;
    pcg_carve_scratch(scratch, scratch_bytes, n, cs);
    uintptr_t sp = ((uintptr_t)scratch + pcg_compact_scratch_bytes(n) + 255) & ~(uintptr_t)255;
    TopkState *s = reinterpret_cast<TopkState *>(sp);
    sp = (sp + TOPKB_MAX * sizeof(TopkState) + 255) & ~(uintptr_t)255;
    int32_t *d_n_ties = reinterpret_cast<int32_t *>(sp);
    int32_t *list = reinterpret_cast<int32_t *>(sp + 256);
    k_topkb_init<<<n_batches, 256, 0, st>>>(s, ks, n_batches);
    PCG_LAUNCHED();
    long blocks = cdiv(n, 256 * 4);
    if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
    for (int shift = 24; shift >= 0; shift -= 8) {
        k_topkb_hist<<<(unsigned)blocks, 256, 0, st>>>(logits, coords, n, n_batches, shift, s);
        PCG_LAUNCHED();
    }
    k_topkb_above<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(logits, coords, n, n_batches, s, mask);
    PCG_LAUNCHED();
    if (int rc = pcg_compact(n, FlagTieB{logits, coords, n_batches, s}, EmitTieB{list}, d_n_ties, cs, st)) return rc;
    k_topkb_ties<<<n_batches, 256, 0, st>>>(list, d_n_ties, coords, s, mask);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- features_at_coordinates
// one warp per query; lanes 0..7 resolve the 8 corners, then all lanes stride the channels
__global__ void k_features_at(const uint64_t *__restrict__ keys, const int32_t *__restrict__ vals, uint64_t mask,
                              const float *__restrict__ feats, int c, int ts, const float *__restrict__ query, long nq,
                              float *__restrict__ out) {
    long q = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    float4 qv = __ldg(reinterpret_cast<const float4 *>(query) + q);
    int row = -1;
    float w = 0.f;
    if (lane < 8) {
        const float fts = (float)ts;
        float bx = floorf(qv.y / fts) * fts, by = floorf(qv.z / fts) * fts, bz = floorf(qv.w / fts) * fts;
        float fx = (qv.y - bx) / fts, fy = (qv.z - by) / fts, fz = (qv.w - bz) / fts;
        int dx = lane & 1, dy = (lane >> 1) & 1, dz = (lane >> 2) & 1;
        w = (dx ? fx : 1.f - fx) * (dy ? fy : 1.f - fy) * (dz ? fz : 1.f - fz);
        int b = (int)qv.x, x = (int)bx + dx * ts, y = (int)by + dy * ts, z = (int)bz + dz * ts;
        if (w != 0.f && pcg_in_range(b, x, y, z)) row = pcg_table_find(keys, vals, mask, pcg_pack(b, x, y, z));
    }
    for (int ch = lane; ch < ((c + 31) & ~31); ch += 32) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int rj = __shfl_sync(0xffffffffu, row, j);
            float wj = __shfl_sync(0xffffffffu, w, j);
            if (rj >= 0 && ch < c) acc += __ldg(feats + (long)rj * c + ch) * wj;
        }
        if (ch < c) out[q * c + ch] = acc;
    }
}

extern "C" int pcg_features_at(const uint64_t *keys, const int32_t *vals, long cap, const float *feats, int c, int ts,
                               const float *query, long nq, float *out, void *stream) {
    if (nq <= 0) return 0;
    PCG_CHECK_ARG(cap > 0 && (cap & (cap - 1)) == 0, "pcg_features_at: cap=%ld must be a power of two", cap);
    k_features_at<<<(unsigned)cdiv(nq * 32, 256), 256, 0, as_stream(stream)>>>(keys, vals, (uint64_t)cap - 1, feats, c, ts,
                                                                            query, nq, out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- im2col for tiny C_in
// The first analysis layer has C_in = 4 and 125 offsets: gathered row by row it costs 125 row requests of 8 useful bytes
// per output.  Instead, lay the whole receptive field of an output row out as ONE dense bf16 row
// col[o][k*cin + c] = F[nbr[k][o]][c] (zeros where absent, zero padded to a multiple of 64), after which the layer is a
// plain dense product col @ W.reshape(K*cin, cout) on the tensor-core kernel with the identity map.
// One thread = one output row x 4 consecutive offsets (a full 32-byte sector of the col row for cin = 4).
template <int CIN>
__global__ void k_im2col_bf16(const float *__restrict__ feats, const int32_t *__restrict__ nbr, long pitch, int K, long n_out,
                              int row_elems, __nv_bfloat16 *__restrict__ col) {
    const int kgroups = row_elems / (4 * CIN);             // groups of 4 offsets, including the zero padding
    long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long o = t % n_out;                                    // consecutive threads -> consecutive rows: coalesced nbr reads
    int kg = (int)(t / n_out);
    if (kg >= kgroups) return;
    __align__(16) __nv_bfloat16 v[4 * CIN];
    int idx[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) idx[j] = (kg * 4 + j < K) ? __ldg(nbr + (long)(kg * 4 + j) * pitch + o) : -1;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < CIN; ++c) v[j * CIN + c] = __float2bfloat16_rn(idx[j] >= 0 ? __ldg(feats + (long)idx[j] * CIN + c) : 0.f);
    uint2 *dst = reinterpret_cast<uint2 *>(col + o * row_elems + (long)kg * 4 * CIN);
    const uint2 *src = reinterpret_cast<const uint2 *>(v);
#pragma unroll
    for (int q = 0; q < CIN; ++q) dst[q] = src[q];         // 8*CIN bytes per thread (one 32-byte sector for CIN = 4)
}

extern "C" int pcg_im2col_bf16(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, int K, long n_out, int row_elems,
                               void *col_bf16, void *stream) {
    PCG_CHECK_ARG((cin == 1 || cin == 2 || cin == 4 || cin == 8) && K >= 1 && row_elems >= K * cin && row_elems % (4 * cin) == 0,
                  "pcg_im2col_bf16: cin=%d (1, 2, 4 or 8) K=%d row_elems=%d (a multiple of 4*cin, >= K*cin)", cin, K, row_elems);
    if (n_out <= 0) return 0;
    const long kgroups = row_elems / (4 * cin);
    const unsigned blocks = (unsigned)cdiv(n_out * kgroups, 256);
    __nv_bfloat16 *col = reinterpret_cast<__nv_bfloat16 *>(col_bf16);
    cudaStream_t st = as_stream(stream);
    switch (cin) {
        case 1: k_im2col_bf16<1><<<blocks, 256, 0, st>>>(feats, nbr, nbr_pitch, K, n_out, row_elems, col); break;
        case 2: k_im2col_bf16<2><<<blocks, 256, 0, st>>>(feats, nbr, nbr_pitch, K, n_out, row_elems, col); break;
        case 4: k_im2col_bf16<4><<<blocks, 256, 0, st>>>(feats, nbr, nbr_pitch, K, n_out, row_elems, col); break;
        default: k_im2col_bf16<8><<<blocks, 256, 0, st>>>(feats, nbr, nbr_pitch, K, n_out, row_elems, col); break;
    }
    PCG_LAUNCHED();
    return 0;
}
