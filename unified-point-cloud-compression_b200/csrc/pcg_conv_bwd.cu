// Weight gradient of the sparse convolution (training step, /root/reference/train.py:195-234 through ME's
// ConvolutionBackwardGPU):  dW[k] = sum_o F[nbr[k][o]]^T (x) G[o]   -- fp32, deterministic.
// One CTA owns a (k, 16 input channels, 64 output channels) block of dW and walks the output rows in 128-row tiles in
// ascending order (fixed summation order, no atomics); tiles whose mask bit k is clear are skipped, which removes the
// empty (tile, offset) combinations of strided and generative maps.  The input gradient needs no kernel of its own:
// it is the forward convolution of G over the inverse map with W[k]^T (autograd.py).
#include "pcg_common.cuh"

#define BW_ROWS 128
#define BW_CI 16
#define BW_CO 64

__global__ void __launch_bounds__(256) k_conv_bwd_weight(const float *__restrict__ feats, int cin, const int32_t *__restrict__ nbr,
                                                         long pitch, const uint32_t *__restrict__ tile_mask, long n_out,
                                                         const float *__restrict__ gout, int cout, float *__restrict__ dw) {
    __shared__ float xs[BW_ROWS][BW_CI];
    __shared__ __align__(16) float gs[BW_ROWS][BW_CO];
    __shared__ int idx_s[BW_ROWS];
    const int k = blockIdx.x, ci0 = blockIdx.y * BW_CI, co0 = blockIdx.z * BW_CO;
    const int tid = threadIdx.x, ci = tid >> 4, cq = (tid & 15) * 4;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const long tiles = (n_out + BW_ROWS - 1) / BW_ROWS;
    for (long t = 0; t < tiles; ++t) {
        if (tile_mask && !((__ldg(tile_mask + t * 4 + (k >> 5)) >> (k & 31)) & 1u)) continue;     // uniform
        const long o0 = t * BW_ROWS;
        if (tid < BW_ROWS) idx_s[tid] = (o0 + tid < n_out) ? __ldg(nbr + (long)k * pitch + o0 + tid) : -1;
        __syncthreads();
        for (int e = tid; e < BW_ROWS * BW_CI; e += 256) {
            const int r = e / BW_CI, c = e % BW_CI;
            const int idx = idx_s[r];
            xs[r][c] = (idx >= 0 && ci0 + c < cin) ? __ldg(feats + (long)idx * cin + ci0 + c) : 0.f;
        }
        for (int e = tid; e < BW_ROWS * BW_CO; e += 256) {
            const int r = e / BW_CO, c = e % BW_CO;
            gs[r][c] = (idx_s[r] >= 0 && co0 + c < cout) ? __ldg(gout + (o0 + r) * cout + co0 + c) : 0.f;
        }
        __syncthreads();
#pragma unroll 8
        for (int r = 0; r < BW_ROWS; ++r) {
            const float x = xs[r][ci];
            const float4 g = *reinterpret_cast<const float4 *>(&gs[r][cq]);
            acc[0] = fmaf(x, g.x, acc[0]); acc[1] = fmaf(x, g.y, acc[1]);
            acc[2] = fmaf(x, g.z, acc[2]); acc[3] = fmaf(x, g.w, acc[3]);
        }
        __syncthreads();
    }
    if (ci0 + ci < cin)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (co0 + cq + j < cout) dw[((long)k * cin + ci0 + ci) * cout + co0 + cq + j] = acc[j];
}

extern "C" int pcg_conv_bwd_weight_f32(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, const uint32_t *tile_mask,
                                       int K, long n_out, const float *gout, int cout, float *dweight, void *stream) {
    PCG_CHECK_ARG(K >= 1 && K <= 65535 && cin >= 1 && cout >= 1, "pcg_conv_bwd_weight_f32: K=%d cin=%d cout=%d", K, cin, cout);
    PCG_CHECK_ARG(tile_mask == nullptr || K <= 128, "pcg_conv_bwd_weight_f32: tile_mask needs K <= 128");
    dim3 grid((unsigned)K, (unsigned)cdiv(cin, BW_CI), (unsigned)cdiv(cout, BW_CO));
    k_conv_bwd_weight<<<grid, 256, 0, as_stream(stream)>>>(feats, cin, nbr, nbr_pitch, tile_mask, n_out, gout, cout, dweight);
    PCG_LAUNCHED();
    return 0;
}
