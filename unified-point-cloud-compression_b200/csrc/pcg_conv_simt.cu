// fp32 SIMT sparse convolution (output-stationary implicit GEMM, no atomics).
//
// out[o][n] = act(bias[n] + sum_{k<K} sum_{c<cin} F[nbr[k][o]][c] * W[k][c][n])
// is a GEMM with an implicitly gathered A = [n_out x (K*cin)] and B = W viewed as
// [(K*cin) x cout]; each output element is accumulated by exactly one thread in ascending
// (k, c) order, so results are bit-reproducible and independent of the tile they fall in.
// Used for the shapes tensor-core tiles would mostly pad (cin = 4, cout in {1,3}) and as the
// fp32 reference mode of the tcgen05 path (pcg_conv_tc.cu).
#include "pcg_common.cuh"

#define TM 64
#define TN 64
#define KC 16

__global__ void __launch_bounds__(256) k_conv_tile_f32(const float *__restrict__ feats, int cin,
                                                       const int32_t *__restrict__ nbr, long pitch, int K, long n_out,
                                                       const float *__restrict__ weight, const float *__restrict__ bias,
                                                       int cout, int act, float *__restrict__ out) {
    __shared__ __align__(16) float As[KC][TM + 4];
    __shared__ __align__(16) float Bs[KC][TN + 4];
    const int t = threadIdx.x;
    const long row0 = (long)blockIdx.x * TM;
    const int col0 = blockIdx.y * TN;
    const int ty = t >> 4, tx = t & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const bool vec_a = (cin % KC) == 0;
    const bool vec_b = (cout & 3) == 0;
    const long flat_total = (long)K * cin;

    if (vec_a) {
        // A loader: thread -> (row = t/4, quad = t%4): one float4 of a gathered row per chunk
        const int a_row = t >> 2, a_quad = t & 3;
        const long o = row0 + a_row;
        for (int k = 0; k < K; ++k) {
            int nb = (o < n_out) ? __ldg(nbr + (long)k * pitch + o) : -1;
            if (!__syncthreads_or(nb >= 0)) continue;  // no row of this tile has offset k
            const float *src = nb >= 0 ? feats + (long)nb * cin : nullptr;
            for (int c0 = 0; c0 < cin; c0 += KC) {
                float4 av = src ? __ldg(reinterpret_cast<const float4 *>(src + c0) + a_quad) : make_float4(0.f, 0.f, 0.f, 0.f);
                As[a_quad * 4 + 0][a_row] = av.x;
                As[a_quad * 4 + 1][a_row] = av.y;
                As[a_quad * 4 + 2][a_row] = av.z;
                As[a_quad * 4 + 3][a_row] = av.w;
                {   // B loader: thread -> (jj = t/16, col4 = (t%16)*4)
                    const int jj = t >> 4, cc = (t & 15) * 4;
                    const float *wrow = weight + ((long)k * cin + c0 + jj) * cout + col0 + cc;
                    float4 bv;
                    if (vec_b && col0 + cc + 3 < cout) bv = __ldg(reinterpret_cast<const float4 *>(wrow));
                    else {
                        bv.x = (col0 + cc + 0 < cout) ? __ldg(wrow + 0) : 0.f;
                        bv.y = (col0 + cc + 1 < cout) ? __ldg(wrow + 1) : 0.f;
                        bv.z = (col0 + cc + 2 < cout) ? __ldg(wrow + 2) : 0.f;
                        bv.w = (col0 + cc + 3 < cout) ? __ldg(wrow + 3) : 0.f;
                    }
                    *reinterpret_cast<float4 *>(&Bs[jj][cc]) = bv;
                }
                __syncthreads();
#pragma unroll
                for (int kk = 0; kk < KC; ++kk) {
                    float4 a = *reinterpret_cast<const float4 *>(&As[kk][ty * 4]);
                    float4 b = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
                    const float av4[4] = {a.x, a.y, a.z, a.w};
                    const float bv4[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av4[i], bv4[j], acc[i][j]);
                }
                __syncthreads();
            }
        }
    } else {
        // generic path: flat index j = k*cin + c, scalar gathers (cin = 1, 3, 4, ...)
        for (long j0 = 0; j0 < flat_total; j0 += KC) {
            int any = 0;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                int idx = t + 256 * e;           // 0..1023
                int r = idx & (TM - 1), jj = idx >> 6;
                long j = j0 + jj, o = row0 + r;
                float v = 0.f;
                if (j < flat_total && o < n_out) {
                    int k = (int)(j / cin), c = (int)(j - (long)k * cin);
                    int nb = __ldg(nbr + (long)k * pitch + o);
                    if (nb >= 0) { v = __ldg(feats + (long)nb * cin + c); any = 1; }
                }
                As[jj][r] = v;
            }
            if (!__syncthreads_or(any)) { __syncthreads(); continue; }
            {
                const int jj = t >> 4, cc = (t & 15) * 4;
                long j = j0 + jj;
                float bv[4] = {0.f, 0.f, 0.f, 0.f};
                if (j < flat_total) {
                    const float *wrow = weight + j * cout + col0 + cc;
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (col0 + cc + q < cout) bv[q] = __ldg(wrow + q);
                }
                *reinterpret_cast<float4 *>(&Bs[jj][cc]) = make_float4(bv[0], bv[1], bv[2], bv[3]);
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
                float4 a = *reinterpret_cast<const float4 *>(&As[kk][ty * 4]);
                float4 b = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
                const float av4[4] = {a.x, a.y, a.z, a.w};
                const float bv4[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av4[i], bv4[j], acc[i][j]);
            }
            __syncthreads();
        }
    }

#pragma unroll
    for (int i = 0; i < 4; ++i) {
        long o = row0 + ty * 4 + i;
        if (o >= n_out) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int n = col0 + tx * 4 + j;
            if (n < cout) {
                float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
                out[o * cout + n] = pcg_apply_act(v, act);
            }
        }
    }
}

// Narrow-output kernel (cout <= 4, cin % 4 == 0): LPR lanes share one output row, each lane owns
// cin/LPR channels (float4 gathers), fixed-order xor-shuffle reduction.
template <int LPR>
__global__ void __launch_bounds__(256) k_conv_narrow_f32(const float *__restrict__ feats, int cin,
                                                         const int32_t *__restrict__ nbr, long pitch, int K, long n_out,
                                                         const float *__restrict__ weight, const float *__restrict__ bias,
                                                         int cout, int act, float *__restrict__ out) {
    extern __shared__ __align__(16) float wsm[];  // [K][cin][4] (cout padded to 4)
    for (int i = threadIdx.x; i < K * cin * 4; i += blockDim.x) {
        int n = i & 3, kc = i >> 2;
        wsm[i] = n < cout ? __ldg(weight + (long)kc * cout + n) : 0.f;
    }
    __syncthreads();
    const int sub = threadIdx.x % LPR;
    const long rows_per_block = blockDim.x / LPR;
    const int vec_per_row = cin >> 2;  // float4 per row
    for (long base = (long)blockIdx.x * rows_per_block; base < n_out; base += (long)gridDim.x * rows_per_block) {
        const long o = base + threadIdx.x / LPR;
        const bool live = o < n_out;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < K; ++k) {
            int nb = live ? __ldg(nbr + (long)k * pitch + o) : -1;
            if (nb < 0) continue;
            const float4 *src = reinterpret_cast<const float4 *>(feats + (long)nb * cin);
            for (int v = sub; v < vec_per_row; v += LPR) {
                float4 f = __ldg(src + v);
                const float4 *w = reinterpret_cast<const float4 *>(wsm + ((long)k * cin + v * 4) * 4);
                float4 w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3];
                acc[0] = fmaf(f.x, w0.x, acc[0]); acc[1] = fmaf(f.x, w0.y, acc[1]); acc[2] = fmaf(f.x, w0.z, acc[2]); acc[3] = fmaf(f.x, w0.w, acc[3]);
                acc[0] = fmaf(f.y, w1.x, acc[0]); acc[1] = fmaf(f.y, w1.y, acc[1]); acc[2] = fmaf(f.y, w1.z, acc[2]); acc[3] = fmaf(f.y, w1.w, acc[3]);
                acc[0] = fmaf(f.z, w2.x, acc[0]); acc[1] = fmaf(f.z, w2.y, acc[1]); acc[2] = fmaf(f.z, w2.z, acc[2]); acc[3] = fmaf(f.z, w2.w, acc[3]);
                acc[0] = fmaf(f.w, w3.x, acc[0]); acc[1] = fmaf(f.w, w3.y, acc[1]); acc[2] = fmaf(f.w, w3.z, acc[2]); acc[3] = fmaf(f.w, w3.w, acc[3]);
            }
        }
#pragma unroll
        for (int d = LPR >> 1; d >= 1; d >>= 1) {
#pragma unroll
            for (int n = 0; n < 4; ++n) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], d);
        }
        if (live && sub == 0) {
            for (int n = 0; n < cout; ++n) out[o * cout + n] = pcg_apply_act(acc[n] + (bias ? __ldg(bias + n) : 0.f), act);
        }
    }
}

template <int LPR>
static int launch_narrow(const float *feats, int cin, const int32_t *nbr, long pitch, int K, long n_out, const float *weight,
                         const float *bias, int cout, int act, float *out, cudaStream_t st) {
    size_t smem = (size_t)K * cin * 4 * sizeof(float);
    PCG_CUDA(cudaFuncSetAttribute(k_conv_narrow_f32<LPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long rows_per_block = 256 / LPR;
    long blocks = cdiv(n_out, rows_per_block);
    if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
    k_conv_narrow_f32<LPR><<<(unsigned)blocks, 256, smem, st>>>(feats, cin, nbr, pitch, K, n_out, weight, bias, cout, act, out);
    PCG_LAUNCHED();
    return 0;
}

extern "C" int pcg_conv_fwd_f32(const float *feats, long n_in, int cin, const int32_t *nbr, long pitch, int K, long n_out,
                                const float *weight, const float *bias, int cout, int act, float *out, void *stream) {
    (void)n_in;
    PCG_CHECK_ARG(cin >= 1 && cout >= 1 && K >= 1, "pcg_conv_fwd_f32: cin=%d cout=%d K=%d", cin, cout, K);
    if (n_out <= 0) return 0;
    cudaStream_t st = as_stream(stream);
    if (cout <= 4 && (cin & 3) == 0 && (size_t)K * cin * 16 <= 200 * 1024) {
        if (cin >= 32) return launch_narrow<8>(feats, cin, nbr, pitch, K, n_out, weight, bias, cout, act, out, st);
        if (cin >= 16) return launch_narrow<4>(feats, cin, nbr, pitch, K, n_out, weight, bias, cout, act, out, st);
        return launch_narrow<1>(feats, cin, nbr, pitch, K, n_out, weight, bias, cout, act, out, st);
    }
    dim3 grid((unsigned)cdiv(n_out, TM), (unsigned)cdiv(cout, TN));
    k_conv_tile_f32<<<grid, 256, 0, st>>>(feats, cin, nbr, pitch, K, n_out, weight, bias, cout, act, out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- channel-wise conv
// out[o][c] = sum_k F[nbr[k][o]][c] * w[k][c or 0]   (ME.MinkowskiChannelwiseConvolution)
__global__ void k_channelwise(const float *__restrict__ feats, int c, const int32_t *__restrict__ nbr, long pitch, int K, long n_out,
                              const float *__restrict__ w, int w_cols, float *__restrict__ out) {
    long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long o = idx / c;
    int ch = (int)(idx - o * c);
    if (o >= n_out) return;
    float acc = 0.f;
    for (int k = 0; k < K; ++k) {
        int nb = __ldg(nbr + (long)k * pitch + o);
        if (nb >= 0) acc = fmaf(__ldg(feats + (long)nb * c + ch), __ldg(w + (long)k * w_cols + (w_cols == 1 ? 0 : ch)), acc);
    }
    out[idx] = acc;
}

extern "C" int pcg_channelwise_conv(const float *feats, int c, const int32_t *nbr, long pitch, int K, long n_out,
                                    const float *w, int w_cols, float *out, void *stream) {
    PCG_CHECK_ARG(w_cols == 1 || w_cols == c, "pcg_channelwise_conv: w_cols=%d must be 1 or C=%d", w_cols, c);
    if (n_out <= 0) return 0;
    k_channelwise<<<(unsigned)cdiv(n_out * c, 256), 256, 0, as_stream(stream)>>>(feats, c, nbr, pitch, K, n_out, w, w_cols, out);
    PCG_LAUNCHED();
    return 0;
}
