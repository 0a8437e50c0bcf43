// Narrow sparse convolutions of the occupancy heads (model/transforms.py:146-160) on bf16 rows.
//
// These layers (32 -> 16 and 16 / 64 -> 1 over every candidate voxel, 27 offsets) carry < 3 % of the FLOPs but touch
// 27 neighbour rows of 32-128 bytes per output: they are bound by the row-gather rate, not by math.  The TMA
// tile::gather4 path of k_conv_tc costs ~4 cycles per gathered row per SM whatever the row width, so short rows
// starve it.  Here rows are gathered by plain 16-byte loads, LANES lanes per row (coalesced per row, L1-cached across
// the heavily overlapping neighbourhoods of adjacent voxels), straight into registers:
//   k_conv_mma   C_in <= 64, C_out <= 16: the bytes a lane loads ARE its mma.sync.m16n8k16 A fragment (the channel ->
//                k-slot permutation is folded into the weight fragments), so nothing is staged in shared memory and
//                the instruction count per gathered row is ~10x below a scalar dot product.  The legacy warp-level
//                MMA is used on purpose: operands are in registers, the tensor pipe is < 10 % busy either way, and
//                tcgen05 would need the rows written to shared memory first.
// One fixed accumulation order per output element (offsets ascending, slabs ascending): bit-reproducible.
#include "pcg_common.cuh"

// mma.sync.m16n8k16 (bf16 x bf16 -> fp32).  Lane (g = lane/4, t = lane%4) of the A fragment holds, for rows g and g+8,
// k-slots {2t, 2t+1, 2t+8, 2t+9} of each k16 slab.  Four lanes share a row:
//   CIN_P = 16: lane t loads channels [4t, 4t+4) as one uint2 -> slab 0 (.x = slots 2t,2t+1; .y = slots 2t+8,2t+9)
//   CIN_P >= 32: load v of lane t is channels [32v + 8t, +8) as one uint4 -> slab 2v (.x,.y) and slab 2v+1 (.z,.w)
// i.e. slot q of slab s holds channel 4*tt + 2*hi + lo (CIN_P = 16) or 32*(s>>1) + 8*tt + 4*(s&1) + 2*hi + lo, with
// tt = (q&7)>>1, lo = q&1, hi = q>>3.  k_pack_mma lays the weights out with the same permutation.
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

static inline int mma_cin_p(int cin) { return cin <= 16 ? 16 : (cin <= 32 ? 32 : 64); }

// wfrag[k][slab][ntile][lane] = uint2 {b0b1, b2b3}: B fragment (k16 x n8, "col") of W[k] under the channel permutation
__global__ void k_pack_mma(const float *__restrict__ w, int K, int cin, int cout, int cin_p, int nt_n, uint2 *__restrict__ wfrag) {
    const int slabs = cin_p / 16;
    const int total = K * slabs * nt_n * 32;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int lane = i & 31;
        int r = i >> 5;
        const int nt = r % nt_n; r /= nt_n;
        const int slab = r % slabs;
        const int k = r / slabs;
        const int g = lane >> 2, t = lane & 3;
        const int n = nt * 8 + g;
        auto val = [&](int slot) {
            const int tt = (slot & 7) >> 1, lo = slot & 1, hi = slot >> 3;
            const int c = cin_p == 16 ? 4 * tt + 2 * hi + lo : 32 * (slab >> 1) + 8 * tt + 4 * (slab & 1) + 2 * hi + lo;
            return (c < cin && n < cout) ? __ldg(w + ((long)k * cin + c) * cout + n) : 0.f;
        };
        __nv_bfloat162 lo = __floats2bfloat162_rn(val(2 * t), val(2 * t + 1));
        __nv_bfloat162 hi = __floats2bfloat162_rn(val(2 * t + 8), val(2 * t + 9));
        wfrag[i] = make_uint2(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi));
    }
}

extern "C" size_t pcg_mma_weight_bytes(int K, int cin, int cout) {
    return (size_t)K * (mma_cin_p(cin) / 16) * ((cout + 7) / 8) * 32 * sizeof(uint2);
}

extern "C" int pcg_pack_weights_mma(const float *weight, int K, int cin, int cout, void *wfrag, void *stream) {
    PCG_CHECK_ARG(K >= 1 && cin >= 1 && cin <= 64 && cout >= 1 && cout <= 16, "pcg_pack_weights_mma: K=%d cin=%d cout=%d", K, cin, cout);
    const int cin_p = mma_cin_p(cin), nt_n = (cout + 7) / 8;
    k_pack_mma<<<(unsigned)cdiv((long)K * (cin_p / 16) * nt_n * 32, 256), 256, 0, as_stream(stream)>>>(
        weight, K, cin, cout, cin_p, nt_n, reinterpret_cast<uint2 *>(wfrag));
    PCG_LAUNCHED();
    return 0;
}

// One warp = MT m16 tiles (16*MT output rows); U kernel offsets are gathered per batch before any MMA.
template <int CIN_P, int NT, int MT, int U>
__global__ void __launch_bounds__(256) k_conv_mma(const __nv_bfloat16 *__restrict__ feats, const int32_t *__restrict__ nbr, long pitch,
                                                  int K, long n_out, const uint2 *__restrict__ wfrag, const float *__restrict__ bias,
                                                  int cout, int act, float *__restrict__ out_f32, __nv_bfloat16 *__restrict__ out_bf16) {
    constexpr int S = CIN_P / 16;                 // k16 slabs
    constexpr int V = CIN_P >= 32 ? CIN_P / 32 : 1;   // loads per lane per row
    constexpr int R = 2 * MT;                     // rows per lane (g + 8j)
    constexpr int ROWS = 16 * MT;
    extern __shared__ uint2 wf_s[];               // [K][S][NT][32]
    for (int i = threadIdx.x; i < K * S * NT * 32; i += blockDim.x) wf_s[i] = wfrag[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const long row0 = ((long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * ROWS;
    if (row0 >= n_out) return;
    const long my_row = row0 + (lane < ROWS ? lane : ROWS - 1);
    const bool my_ok = lane < ROWS && my_row < n_out;
    float acc[MT][NT][4];
#pragma unroll
    for (int a = 0; a < MT; ++a)
#pragma unroll
        for (int b = 0; b < NT; ++b)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[a][b][c] = 0.f;
    // The neighbour table is streamed from DRAM (no reuse) and every row gather depends on it: the indices of batch b+1 are
    // requested before the rows of batch b, so the two DRAM latencies of a batch overlap instead of adding up.
    int mine[U];
#pragma unroll
    for (int u = 0; u < U; ++u) mine[u] = (my_ok && u < K) ? __ldg(nbr + (long)u * pitch + my_row) : -1;
    for (int k0 = 0; k0 < K; k0 += U) {
        int nxt[U];
#pragma unroll
        for (int u = 0; u < U; ++u) nxt[u] = (my_ok && k0 + U + u < K) ? __ldg(nbr + (long)(k0 + U + u) * pitch + my_row) : -1;
        uint4 v[U][R][V];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const int idx = __shfl_sync(0xffffffffu, mine[u], g + 8 * j);
#pragma unroll
                for (int q = 0; q < V; ++q) {
                    if (CIN_P == 16) {
                        const uint2 h = idx >= 0 ? __ldg(reinterpret_cast<const uint2 *>(feats + (long)idx * 16) + t) : make_uint2(0, 0);
                        v[u][j][q] = make_uint4(h.x, h.y, 0, 0);
                    } else {
                        v[u][j][q] = idx >= 0 ? __ldg(reinterpret_cast<const uint4 *>(feats + (long)idx * CIN_P + 32 * q) + t)
                                              : make_uint4(0, 0, 0, 0);
                    }
                }
            }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (k0 + u >= K) break;
            const uint2 *wk = wf_s + (k0 + u) * (S * NT * 32) + lane;
#pragma unroll
            for (int s = 0; s < S; ++s) {
                uint2 bf[NT];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) bf[nt] = wk[(s * NT + nt) * 32];
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    const uint4 lo = v[u][2 * m][s >> 1], hi = v[u][2 * m + 1][s >> 1];     // rows g+16m, g+8+16m
                    // a0 = (row g, slots 2t,2t+1), a1 = (row g+8, same), a2 = (row g, slots 2t+8,2t+9), a3 = (row g+8, same)
                    const uint32_t a0 = (s & 1) ? lo.z : lo.x, a1 = (s & 1) ? hi.z : hi.x, a2 = (s & 1) ? lo.w : lo.y,
                                   a3 = (s & 1) ? hi.w : hi.y;
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) mma_16816(acc[m][nt], a0, a1, a2, a3, bf[nt].x, bf[nt].y);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) mine[u] = nxt[u];
    }
    // C fragment: c0,c1 = (row g, cols 2t, 2t+1), c2,c3 = (row g+8, same cols); n tile nt adds 8 columns
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const long o = row0 + g + 8 * (2 * m + h);
            if (o >= n_out) continue;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {        // bf16 rows are always 16 wide (zero padded)
                const int c = nt * 8 + 2 * t;
                float x0 = 0.f, x1 = 0.f;
                if (nt < NT) {
                    if (c < cout) x0 = pcg_apply_act(acc[m][nt < NT ? nt : 0][2 * h] + (bias ? __ldg(bias + c) : 0.f), act);
                    if (c + 1 < cout) x1 = pcg_apply_act(acc[m][nt < NT ? nt : 0][2 * h + 1] + (bias ? __ldg(bias + c + 1) : 0.f), act);
                }
                if (out_f32) {
                    if (c < cout) out_f32[o * cout + c] = x0;
                    if (c + 1 < cout) out_f32[o * cout + c + 1] = x1;
                }
                if (out_bf16) {
                    __nv_bfloat162 pk = __floats2bfloat162_rn(x0, x1);
                    *reinterpret_cast<__nv_bfloat162 *>(out_bf16 + o * 16 + c) = pk;
                }
            }
        }
}

template <int CIN_P, int NT, int MT, int U>
static int launch_mma(const __nv_bfloat16 *f, const int32_t *nbr, long pitch, int K, long n_out, const uint2 *wf, const float *bias,
                      int cout, int act, float *out_f32, __nv_bfloat16 *out_bf16, cudaStream_t st) {
    const size_t smem = (size_t)K * (CIN_P / 16) * NT * 32 * sizeof(uint2);
    PCG_CUDA(cudaFuncSetAttribute(k_conv_mma<CIN_P, NT, MT, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_mma<CIN_P, NT, MT, U><<<(unsigned)cdiv(n_out, 8 * 16 * MT), 256, smem, st>>>(f, nbr, pitch, K, n_out, wf, bias, cout, act,
                                                                                     out_f32, out_bf16);
    PCG_LAUNCHED();
    return 0;
}

extern "C" int pcg_conv_fwd_mma(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                                const void *wfrag, const float *bias, int cout, int act, float *out_f32, void *out_bf16,
                                void *stream) {
    PCG_CHECK_ARG(cin >= 1 && cin <= 64, "pcg_conv_fwd_mma: cin=%d must be <= 64", cin);
    PCG_CHECK_ARG(cout >= 1 && cout <= 16, "pcg_conv_fwd_mma: cout=%d must be <= 16", cout);
    PCG_CHECK_ARG(K >= 1 && pcg_mma_weight_bytes(K, cin, cout) <= 160 * 1024, "pcg_conv_fwd_mma: K=%d too large", K);
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)wfrag & 7) == 0, "pcg_conv_fwd_mma: operands must be aligned");
    if (n_out <= 0) return 0;
    PCG_CHECK_ARG(n_in >= 1, "pcg_conv_fwd_mma: n_in=%ld", n_in);
    cudaStream_t st = as_stream(stream);
    const __nv_bfloat16 *f = reinterpret_cast<const __nv_bfloat16 *>(feats_bf16);
    const uint2 *wf = reinterpret_cast<const uint2 *>(wfrag);
    __nv_bfloat16 *ob = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    const int cin_p = mma_cin_p(cin), nt = (cout + 7) / 8;
#define PCG_MMA(C, N, M, UU) launch_mma<C, N, M, UU>(f, nbr, nbr_pitch, K, n_out, wf, bias, cout, act, out_f32, ob, st)
    if (cin_p == 16) return nt == 1 ? PCG_MMA(16, 1, 2, 4) : PCG_MMA(16, 2, 2, 4);
    if (cin_p == 32) return nt == 1 ? PCG_MMA(32, 1, 2, 3) : PCG_MMA(32, 2, 2, 3);
    return nt == 1 ? PCG_MMA(64, 1, 1, 3) : PCG_MMA(64, 2, 1, 3);
#undef PCG_MMA
}
