// Sparse convolution on the 5th-gen tensor cores (tcgen05, sm_100a): output-stationary
// gather-GEMM with fp32 accumulators in TMEM and no scatter.
//
//   out[o][:] = act(bias + sum_{k active} F[nbr[k][o]][:] @ W[k])       (bf16 x bf16 -> fp32)
//
// One CTA owns one 128-row output tile (UMMA M = 128, N = roundup16(cout) <= 256, K = 16 per instruction):
//   warps 0-3  producers, then epilogue.  The gather is done by the TMA unit: lane l of a producer warp loads the
//              four neighbour indices nbr[k][4l..4l+3] of a kernel offset as one int4 and issues ONE
//              cp.async.bulk.tensor tile::gather4, which fetches those four feature rows straight into the
//              K-major swizzled operand image (absent neighbours are index -1 = out of bounds = zero fill).
//              A warp-wide issue therefore fills a whole 128-row operand sub-tile; the matching pre-swizzled
//              weight slice arrives through a 1-D bulk copy.  Both complete on the stage's mbarrier (expect_tx).
//   warp 4     allocates TMEM, one elected lane issues tcgen05.mma and commits each stage back to the producers.
// Operand rows are RB = 32 / 64 / 128 bytes (cin = 16 / 32 / >= 33 after padding) with the matching 32B / 64B / 128B
// swizzle, so narrow layers (occupancy heads, input layer) move no padding; a 16 KiB stage holds 128/RB kernel
// offsets.  Kernel offsets no row of the tile uses are skipped (tile_mask from pcg_kernel_map).  Every output element
// is one fixed-order chain of MMAs over ascending (k, cin-chunk): bit-reproducible, and independent of which
// tile a row lands in (absent neighbours contribute exact zeros).
#include <stdlib.h>
#include <string.h>

#include "pcg_tc.cuh"

#define TC_ROWS 128
#define TC_KCHUNK 64                       // bf16 elements per 128-byte swizzled row
#define TC_A_BYTES (TC_ROWS * 128)         // 16 KiB per stage
#define TC_LAG 2                           // cp.async groups in flight per producer thread
#define TC_MAX_STAGES 12

struct TcParams {
    const int32_t *nbr;           // [K][nbr_pitch], padding entries -1
    const uint32_t *tile_mask;    // [tiles][4]
    const uint8_t *wpacked;       // [k][chunk][cout_p][RB bytes] pre-swizzled
    const float *bias;
    float *out_f32;
    __nv_bfloat16 *out_bf16;
    const uint8_t *feats;         // bf16 operand rows (LSU gather mode)
    long row_bytes;               // bytes per operand row = 2 * cin_p
    long n_out, nbr_pitch;
    int cout, cout_p, K;
    int chunks;      // RB == 128: 64-wide K chunks per kernel offset; otherwise 1
    int stages;
    int tmem_cols;   // power of two >= max(32, cout_p)
    int act;
    // fused second head layer (pcg_conv_fwd_tc_proj): proj_out[k][o] = act(row o) . w2[k][:] for the TC_PROJ_K offsets of the
    // following cout -> 1 convolution, taken from the fp32 accumulator row in the epilogue (weights: kernel parameter TcProjW)
    float *proj_out;              // [TC_PROJ_K][proj_pitch] fp32
    long proj_pitch;
};
#define TC_PROJ_K 27
// Second-layer weights [TC_PROJ_K][PC] of the fused head, passed BY VALUE as a kernel parameter: they sit in the constant bank, so
// every FFMA of the epilogue projection takes its weight as a constant operand (no load instruction, no latency).
// 27 x 64 x 4 B = 6.9 KB (kernel parameters may be up to 32 KB since CUDA 12.1).  PC = 0: no projection (empty).
// Layout: offset PAIRS, column-major: w2[c * TC_PROJ_KP + kp] = {w[2 kp][c], w[2 kp + 1][c]} (offset 27 = zero), so that one packed
// fp32 multiply-add (fma.rn.f32x2, sm_100) advances two offsets and consecutive instructions feed DIFFERENT accumulators (14
// independent chains per column instead of one 16-deep dependent chain per offset).  bias: the first layer's bias, also as constant
// operands (the ncu source view of the first version showed a third of the projection's stall samples on the bias LDGs).
#define TC_PROJ_KP ((TC_PROJ_K + 1) / 2)
template <int PC>
struct TcProjW {
    float2 w2[TC_PROJ_KP * PC];
    float bias[PC];
};
template <>
struct TcProjW<0> {
    char unused;
};

// epilogue piece shared by the tcgen05 conv kernels: 16 accumulator columns [c0, c0+16) of output row o -> bias, activation, fp32 and/or bf16
__device__ __forceinline__ void tc_store16(const TcParams &p, long o, int c0, const uint32_t (&v)[16]) {
    float f[16];
#pragma unroll
    for (int j = 0; j < 16; ++j)
        f[j] = pcg_apply_act(__uint_as_float(v[j]) + ((p.bias && c0 + j < p.cout) ? __ldg(p.bias + c0 + j) : 0.f), p.act);
    if ((p.cout & 15) == 0) {
        if (p.out_f32) {
            float4 *dst = reinterpret_cast<float4 *>(p.out_f32 + o * p.cout + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
        }
        if (p.out_bf16) {
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
                pk[j] = *reinterpret_cast<uint32_t *>(&h);
            }
            uint4 *dst = reinterpret_cast<uint4 *>(p.out_bf16 + o * p.cout + c0);
            dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
    } else if (p.out_f32) {   // narrow heads (cout = 1, 3, ...): scalar stores of the real columns
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (c0 + j < p.cout) p.out_f32[o * p.cout + c0 + j] = f[j];
    }
}

// 16 already activated columns of row o -> fp32 and/or bf16 rows (cout % 16 == 0)
__device__ __forceinline__ void tc_write16(const TcParams &p, long o, int c0, const float (&f)[16]) {
    if (p.out_f32) {
        float4 *dst = reinterpret_cast<float4 *>(p.out_f32 + o * p.cout + c0);
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
    }
    if (p.out_bf16) {
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
            pk[j] = *reinterpret_cast<uint32_t *>(&h);
        }
        uint4 *dst = reinterpret_cast<uint4 *>(p.out_bf16 + o * p.cout + c0);
        dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
}

// (tc_lsu_gather_tile: the LSU gather of one operand sub-tile lives in pcg_tc.cuh, shared with the weight-gradient kernel)

template <int RB>
__global__ void __launch_bounds__(160) k_conv_tc(const __grid_constant__ CUtensorMap tm, const TcParams p) {
    constexpr int G = 128 / RB;                  // operand sub-tiles (kernel offsets) per 16 KiB stage
    constexpr uint32_t SUB_A = TC_ROWS * RB;     // bytes of one A sub-tile
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte aligned operand ring (swizzle atom alignment), barriers behind it
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const uint32_t sub_b = (uint32_t)p.cout_p * RB;            // bytes of one B sub-tile
    const uint32_t stage_bytes = TC_A_BYTES + G * sub_b;
    uint8_t *ctl = sm + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(ctl);            // [stages]
    uint64_t *empty_bar = full_bar + TC_MAX_STAGES;                    // [stages]
    uint64_t *accum_bar = empty_bar + TC_MAX_STAGES;                   // [1]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_bar + 1); // [1]
    int *n_active_s = reinterpret_cast<int *>(tmem_slot + 1);          // [1]
    uint8_t *active = reinterpret_cast<uint8_t *>(n_active_s + 1);     // [128]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long tile = blockIdx.x;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(smem_u32(full_bar + s), 1);         // one arrive.expect_tx; data arrives as tx bytes
            mbar_init(smem_u32(empty_bar + s), 1);        // one tcgen05.commit
        }
        mbar_init(smem_u32(accum_bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < TC_ROWS) {
        // list of kernel offsets at least one row of this tile uses
        const uint32_t m0 = __ldg(p.tile_mask + tile * 4 + 0), m1 = __ldg(p.tile_mask + tile * 4 + 1),
                       m2 = __ldg(p.tile_mask + tile * 4 + 2), m3 = __ldg(p.tile_mask + tile * 4 + 3);
        const uint32_t w = tid >> 5, b = tid & 31;
        const uint32_t mine = w == 0 ? m0 : (w == 1 ? m1 : (w == 2 ? m2 : m3));
        if (tid < p.K && ((mine >> b) & 1u)) {
            int pos = __popc(mine & ((1u << b) - 1u));
            if (w > 0) pos += __popc(m0);
            if (w > 1) pos += __popc(m1);
            if (w > 2) pos += __popc(m2);
            active[pos] = (uint8_t)tid;
        }
        if (tid == 0) *n_active_s = __popc(m0) + __popc(m1) + __popc(m2) + __popc(m3);
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;
    const int n_active = *n_active_s;
    const int n_sub = n_active * p.chunks;              // operand sub-tiles of this tile
    const int n_items = (n_sub + G - 1) / G;            // pipeline stages to run

    if (warp < 4) {
        // ------------------------------------------------------------------ producers (item it -> warp it % P)
        // P <= stages: a warp is then never two ring wraps ahead of the consumer, which the 1-bit mbarrier
        // phase parity could not tell apart from "slot free".
        const int P = p.stages < 4 ? p.stages : 4;
        const int32_t *nbr_tile = p.nbr + tile * TC_ROWS + lane * 4;
        int s = warp, wrap = 0;                            // stage slot / how often the ring wrapped, for item `it`
        while (s >= p.stages) { s -= p.stages; ++wrap; }
        int a = 0, cc = 0;                                 // RB == 128: (active offset, chunk) of item `it`
        if (RB == 128) { cc = warp; while (cc >= p.chunks) { cc -= p.chunks; ++a; } }
        int4 idx[G];
        auto load_idx = [&](int it_, int a_) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const int aa = RB == 128 ? a_ : it_ * G + g;
                idx[g] = aa < n_active ? __ldg(reinterpret_cast<const int4 *>(nbr_tile + (long)active[aa] * p.nbr_pitch))
                                       : make_int4(-1, -1, -1, -1);
            }
        };
        if (warp < P && warp < n_items) load_idx(warp, a);
        for (int it = warp; warp < P && it < n_items; it += P) {
            const int subs = (n_sub - it * G) < G ? (n_sub - it * G) : G;     // sub-tiles in this item
            const uint32_t fb = smem_u32(full_bar + s);
            const uint32_t a_base = base + (uint32_t)s * stage_bytes;
            mbar_wait(smem_u32(empty_bar + s), (uint32_t)(wrap & 1) ^ 1u);
            // Quads (4 consecutive rows = one gather4) with no neighbour at all cost a TMA slot like any other, and sparse maps
            // (strided / transposed: 12-35 % dense) are mostly such quads: skip their gather and let the warp zero the
            // 4 rows itself (32 lanes x 16 B).  The stage's expect_tx shrinks accordingly; the arrive comes after the
            // generic-proxy zero stores + proxy fence, so the MMA thread's wait covers them too.
            uint32_t tx_bytes = 0;
            bool zeroed = false;
#pragma unroll
            for (int g = 0; g < G; ++g) {
                if (g < subs) {
                    const bool empty_quad = (idx[g].x & idx[g].y & idx[g].z & idx[g].w) < 0;
                    uint32_t em = __ballot_sync(0xffffffffu, empty_quad);
                    tx_bytes += (32u - (uint32_t)__popc(em)) * 4u * RB + sub_b;
                    if (!empty_quad) tma_gather4(a_base + g * SUB_A + (uint32_t)lane * 4u * RB, &tm, cc * TC_KCHUNK, idx[g], fb);
                    zeroed |= em != 0;
                    while (em) {
                        const uint32_t q = (uint32_t)__ffs(em) - 1u;
                        em &= em - 1u;
                        if (lane * 16u < 4u * RB)
                            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(a_base + g * SUB_A + q * 4u * RB + (uint32_t)lane * 16u), "r"(0u) : "memory");
                    }
                }
            }
            if (lane < subs) {
                const int k = (RB == 128 ? active[a] : active[it * G + lane]);
                bulk_g2s(a_base + TC_A_BYTES + lane * sub_b, p.wpacked + ((size_t)k * p.chunks + cc) * sub_b, sub_b, fb);
            }
            if (zeroed) fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(tx_bytes) : "memory");
            }
            // advance this warp's cursor by P items and prefetch the next item's neighbour indices
            s += P;
            while (s >= p.stages) { s -= p.stages; ++wrap; }
            if (RB == 128) { cc += P; while (cc >= p.chunks) { cc -= p.chunks; ++a; } }
            if (it + P < n_items) load_idx(it + P, a);
        }

        // ------------------------------------------------------------------ epilogue
        const long o = tile * TC_ROWS + tid;
        const bool live = o < p.n_out;
        if (n_items > 0) {
            mbar_wait(smem_u32(accum_bar), 0);
            tc_fence_after();
        }
        const uint32_t lane_base = tmem_d + ((uint32_t)(warp * 32) << 16);
        for (int c0 = 0; c0 < p.cout; c0 += 16) {
            uint32_t v[16];
            if (n_items > 0) tmem_ld16(lane_base + (uint32_t)c0, v);
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0u;
            }
            if (live) tc_store16(p, o, c0, v);
        }
        tc_fence_before();
    } else {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.cout_p >> 3) << 17) | ((TC_ROWS >> 4) << 24);
            int s = 0;
            uint32_t ph = 0;
            for (int it = 0; it < n_items; ++it) {
                const int subs = (n_sub - it * G) < G ? (n_sub - it * G) : G;
                mbar_wait(smem_u32(full_bar + s), ph);
                tc_fence_after();
                const uint32_t a_addr = base + (uint32_t)s * stage_bytes;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    if (g < subs) {
                        const uint64_t adesc = umma_desc<RB>(a_addr + g * SUB_A), bdesc = umma_desc<RB>(a_addr + TC_A_BYTES + g * sub_b);
#pragma unroll
                        for (int kk = 0; kk < RB / 32; ++kk)  // +32 B per K=16 step inside the swizzle atom
                            umma_bf16(tmem_d, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (it | g | kk) ? 1u : 0u);
                    }
                }
                umma_commit(smem_u32(empty_bar + s));
                if (++s == p.stages) { s = 0; ph ^= 1u; }
            }
            if (n_items > 0) umma_commit(smem_u32(accum_bar));
        }
        __syncwarp();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols) : "memory");
    }
}

// ---------------------------------------------------------------------------- LSU gather, T output tiles per CTA
// k_conv_tc above is paced by the TMA unit: tile::gather4 retires ~4 cycles per gathered row per SM, 556 cycles per
// 128-row stage, whatever else is idle (ncu: L2 38 %, tensor 23 %, L1 15 %).  Here the rows are gathered by the LSU instead:
// 16-byte cp.async (LDGSTS) straight into the 128B-swizzled operand image, 8 lanes per 128-byte row, completion tracked by
// cp.async.mbarrier.arrive.noinc on the stage's mbarrier, absent neighbours zero-filled by src-size 0 (no request at all).
// The producer warps must be lean: a first version with ~20 SASS instructions per copy was bound by their issue latency
// (profiles/r01_conv_variants.md).  One CTA owns T adjacent tiles: a stage holds T gathered sub-tiles and ONE weight slice
// (for cout_p >= 64 the slice is as many L2 bytes as a tile's rows), warp (g, t) gathers tile t of items g, g+P, ..., the MMA
// thread issues T accumulations (TMEM columns t*cout_p) against the same B descriptor, and a tile that does not use an offset
// skips both its gather and its MMAs.  Per-element accumulation order is k_conv_tc's (ascending offset, chunk): same bits.
// (A hybrid with tile 0 gathered by TMA gather4 and tile 1 by the LSU was measured and is slower than LSU-only: 0.68 vs 0.53 ms
// on the 27x128x64 layer; the gather4 row rate, ~4 cycles per row per SM, paces the pair.)
// PC: 0 = plain; 16 / 32 / 64 = cout of a first head layer whose second layer (cout -> 1, 27 offsets) is folded into the epilogue.
template <int T, int RB, bool CA, int PC = 0>
__global__ void __launch_bounds__(32 * (4 * T + 1), (T == 2 && PC != 0) ? 2 : 0) k_conv_tc_mt(const TcParams p, const TcProjW<PC> pw) {   // T = 2 runs two CTAs per SM: the projection must stay <= 112 registers
    constexpr int MMA_WARP = 4 * T;
    constexpr uint32_t A_BYTES = TC_ROWS * RB;         // one gathered operand sub-tile (RB = 128 / 64 / 32-byte rows: cin_p >= 64 / 32 / 16)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const uint32_t sub_b = (uint32_t)p.cout_p * RB;
    const uint32_t stage_bytes = T * A_BYTES + sub_b;
    uint8_t *ctl = sm + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(ctl);
    uint64_t *empty_bar = full_bar + TC_MAX_STAGES;
    uint64_t *accum_bar = empty_bar + TC_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_bar + 1);
    int *n_active_s = reinterpret_cast<int *>(tmem_slot + 1);
    uint32_t *tmask_s = reinterpret_cast<uint32_t *>(n_active_s + 1);   // [T][4]
    uint8_t *active = reinterpret_cast<uint8_t *>(tmask_s + 4 * T);     // [128]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long tile0 = (long)blockIdx.x * T;
    const long n_tiles = (p.n_out + TC_ROWS - 1) / TC_ROWS;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(smem_u32(full_bar + s), 32 * T + 1);   // one cp.async-completion arrive per producer lane + the weight slice's expect_tx
            mbar_init(smem_u32(empty_bar + s), 1);
        }
        mbar_init(smem_u32(accum_bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < TC_ROWS) {
        uint32_t u[4] = {0u, 0u, 0u, 0u};                 // union of the T tiles' offset masks
#pragma unroll
        for (int t = 0; t < T; ++t)
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const uint32_t m = tile0 + t < n_tiles ? __ldg(p.tile_mask + (tile0 + t) * 4 + w) : 0u;
                u[w] |= m;
                if (tid == 0) tmask_s[t * 4 + w] = m;
            }
        const uint32_t w = tid >> 5, b = tid & 31;
        const uint32_t mine = w == 0 ? u[0] : (w == 1 ? u[1] : (w == 2 ? u[2] : u[3]));
        if (tid < p.K && ((mine >> b) & 1u)) {
            int pos = __popc(mine & ((1u << b) - 1u));
            if (w > 0) pos += __popc(u[0]);
            if (w > 1) pos += __popc(u[1]);
            if (w > 2) pos += __popc(u[2]);
            active[pos] = (uint8_t)tid;
        }
        if (tid == 0) *n_active_s = __popc(u[0]) + __popc(u[1]) + __popc(u[2]) + __popc(u[3]);
    }
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // PC >= 32: the second-layer weights (3.4 / 6.9 KB) are walked once per epilogue warp -- more than the constant cache keeps
    // close (measured: 3x the issue-bound time as constant operands) -- so they are staged in shared memory once per CTA and read
    // as broadcast 16-byte LDS; PC == 16 (1.7 KB) stays in the constant bank.
    float2 *proj_sw = reinterpret_cast<float2 *>(ctl + 512);
    if constexpr (PC >= 32) {
        for (int i = tid; i < TC_PROJ_KP * PC; i += 32 * (4 * T + 1)) proj_sw[i] = pw.w2[i];      // column-major pairs, as in the parameter
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;
    const int n_active = *n_active_s;
    const int n_items = n_active * p.chunks;
    auto uses = [&](int t, int k) -> bool { return (tmask_s[t * 4 + (k >> 5)] >> (k & 31)) & 1u; };

    if (warp < MMA_WARP) {
        // ------------------------------------------------------------------ producers: warp (g, t) gathers tile t of items g, g+P, ...
        const int g = warp & 3, t = warp >> 2;
        const int P = p.stages < 4 ? p.stages : 4;
        int s = g, wrap = 0;
        while (s >= p.stages) { s -= p.stages; ++wrap; }
        int a = 0, cc = g;
        while (cc >= p.chunks) { cc -= p.chunks; ++a; }
        const int32_t *nbr_t = p.nbr + (tile0 + t) * TC_ROWS + lane;
        const uint8_t *fsrc = p.feats + (lane % (RB / 16)) * 16;
        const int row_bytes = (int)p.row_bytes;
        int idx4[4];
        auto load_idx4 = [&](int a_) {
            const int k = active[a_];
            if (uses(t, k)) {
                const int32_t *nb = nbr_t + (long)k * p.nbr_pitch;
#pragma unroll
                for (int u = 0; u < 4; ++u) idx4[u] = __ldg(nb + 32 * u);
            }
        };
        if (g < P && g < n_items) load_idx4(a);
        for (int it = g; g < P && it < n_items; it += P) {
            const int k = active[a];
            const uint32_t fb = smem_u32(full_bar + s);
            const uint32_t a_base = base + (uint32_t)s * stage_bytes;
            mbar_wait(smem_u32(empty_bar + s), (uint32_t)(wrap & 1) ^ 1u);
            if (t == 0 && lane == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(sub_b) : "memory");
                bulk_g2s(a_base + T * A_BYTES, p.wpacked + ((size_t)k * p.chunks + cc) * sub_b, sub_b, fb);
            }
            if (uses(t, k)) tc_lsu_gather_tile<RB, CA>(a_base + (uint32_t)t * A_BYTES, fsrc + cc * 128, row_bytes, idx4, lane);
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(fb) : "memory");
            const int a_prev = a;
            s += P;
            while (s >= p.stages) { s -= p.stages; ++wrap; }
            cc += P;
            while (cc >= p.chunks) { cc -= p.chunks; ++a; }
            if (it + P < n_items && a != a_prev) load_idx4(a);
        }

        // ------------------------------------------------------------------ epilogue: warp (g, t) stores rows 32g..32g+31 of tile t
        if (n_items > 0) {
            mbar_wait(smem_u32(accum_bar), 0);
            tc_fence_after();
        }
        const long o = (tile0 + t) * TC_ROWS + g * 32 + lane;
        const bool live = o < p.n_out;
        const bool has = (tmask_s[t * 4] | tmask_s[t * 4 + 1] | tmask_s[t * 4 + 2] | tmask_s[t * 4 + 3]) != 0u;
        const uint32_t lane_base = tmem_d + ((uint32_t)(g * 32) << 16) + (uint32_t)(t * p.cout_p);
        if (tile0 + t < n_tiles) {
            if constexpr (PC == 0) {
                for (int c0 = 0; c0 < p.cout; c0 += 16) {
                    uint32_t v[16];
                    if (has) tmem_ld16(lane_base + (uint32_t)c0, v);
                    else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = 0u;
                    }
                    if (live) tc_store16(p, o, c0, v);
                }
            } else {
                // fused second layer: ps[k] = act(row) . w2[k][:], columns ascending (one fixed fp32 chain per offset; the packed
                // multiply-add rounds each half like a scalar fma.rn, so pairing offsets does not change a bit)
                float2 ps2[TC_PROJ_KP];
#pragma unroll
                for (int kp = 0; kp < TC_PROJ_KP; ++kp) ps2[kp] = make_float2(0.f, 0.f);
#pragma unroll
                for (int c0 = 0; c0 < PC; c0 += 16) {
                    uint32_t v[16];
                    if (has) tmem_ld16(lane_base + (uint32_t)c0, v);
                    else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = 0u;
                    }
                    float f[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) f[j] = pcg_apply_act(__uint_as_float(v[j]) + pw.bias[c0 + j], p.act);
                    if (live && (p.out_f32 || p.out_bf16)) tc_write16(p, o, c0, f);     // the hidden row, only if a caller wants it
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float2 ff = make_float2(f[j], f[j]);
                        if constexpr (PC == 16) {
#pragma unroll
                            for (int kp = 0; kp < TC_PROJ_KP; ++kp) ps2[kp] = __ffma2_rn(ff, pw.w2[(c0 + j) * TC_PROJ_KP + kp], ps2[kp]);
                        } else {
#pragma unroll
                            for (int kp = 0; kp < TC_PROJ_KP; kp += 2) {
                                const float4 q = *reinterpret_cast<const float4 *>(proj_sw + (c0 + j) * TC_PROJ_KP + kp);   // warp-uniform: broadcast
                                ps2[kp] = __ffma2_rn(ff, make_float2(q.x, q.y), ps2[kp]);
                                ps2[kp + 1] = __ffma2_rn(ff, make_float2(q.z, q.w), ps2[kp + 1]);
                            }
                        }
                    }
                }
                if (live) {                                          // lane = row: one coalesced 128-byte store per warp and offset
#pragma unroll
                    for (int kp = 0; kp < TC_PROJ_KP; ++kp) {
                        p.proj_out[(long)(2 * kp) * p.proj_pitch + o] = ps2[kp].x;
                        if (2 * kp + 1 < TC_PROJ_K) p.proj_out[(long)(2 * kp + 1) * p.proj_pitch + o] = ps2[kp].y;
                    }
                }
            }
        }
        tc_fence_before();
    } else {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.cout_p >> 3) << 17) | ((TC_ROWS >> 4) << 24);
            int s = 0, a = 0, cc = 0;
            uint32_t ph = 0, started = 0;
            for (int it = 0; it < n_items; ++it) {
                const int k = active[a];
                mbar_wait(smem_u32(full_bar + s), ph);
                fence_proxy_async();                 // cp.async wrote the operand rows through the generic proxy
                tc_fence_after();
                const uint32_t a_addr = base + (uint32_t)s * stage_bytes;
                const uint64_t bdesc = umma_desc<RB>(a_addr + T * A_BYTES);
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    if (!uses(t, k)) continue;
                    const uint64_t adesc = umma_desc<RB>(a_addr + t * A_BYTES);
#pragma unroll
                    for (int kk = 0; kk < RB / 32; ++kk)
                        umma_bf16(tmem_d + (uint32_t)(t * p.cout_p), adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc,
                                  (((started >> t) & 1u) | (uint32_t)kk) ? 1u : 0u);
                    started |= 1u << t;
                }
                umma_commit(smem_u32(empty_bar + s));
                if (++s == p.stages) { s = 0; ph ^= 1u; }
                if (++cc == p.chunks) { cc = 0; ++a; }
            }
            if (n_items > 0) umma_commit(smem_u32(accum_bar));
        }
        __syncwarp();
    }
    __syncthreads();
    if (warp == MMA_WARP) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols) : "memory");
    }
}

// ---------------------------------------------------------------------------- weight packing
static inline int tc_cin_p(int cin) { return cin <= 16 ? 16 : (cin <= 32 ? 32 : ((cin + 63) & ~63)); }   // operand row width
static inline int tc_cout_p(int cout) { return (cout + 15) & ~15; }
static inline int tc_rb(int cin_p) { return cin_p == 16 ? 32 : (cin_p == 32 ? 64 : 128); }

// wpacked[(k*chunks + cc)][n][RB/2] bf16 = B operand sub-tile (N rows of RB bytes, K-major, RB-byte swizzle):
// element (n, kk) <- W[k][cc*64 + kk][n] (zero beyond cin / cout), its 16-byte chunk XOR-swizzled with the row.
__global__ void k_pack_weights(const float *__restrict__ w, int K, int cin, int cout, int cout_p, int chunks, int rb,
                               __nv_bfloat16 *__restrict__ out) {
    const int kw = rb / 2;                       // K elements per operand row
    long total = (long)K * chunks * cout_p * kw;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        int kk = (int)(i % kw);
        long t = i / kw;
        int n = (int)(t % cout_p);
        t /= cout_p;
        int cc = (int)(t % chunks);
        int k = (int)(t / chunks);
        int c = cc * 64 + kk;
        float v = (c < cin && n < cout) ? __ldg(w + ((long)k * cin + c) * cout + n) : 0.f;
        long sub = ((long)k * chunks + cc) * cout_p * kw;  // elements
        int xr = rb == 128 ? (n & 7) : (rb == 64 ? ((n >> 1) & 3) : ((n >> 2) & 1));
        int chunk16 = (kk >> 3) ^ xr;
        out[sub + (long)n * kw + chunk16 * 8 + (kk & 7)] = __float2bfloat16_rn(v);
    }
}

extern "C" size_t pcg_packed_weight_bytes(int K, int cin, int cout) {
    const int cin_p = tc_cin_p(cin), cout_p = tc_cout_p(cout), rb = tc_rb(cin_p);
    const int chunks = rb == 128 ? (cin_p + 63) / 64 : 1;
    return (size_t)K * chunks * cout_p * rb;
}

extern "C" int pcg_pack_weights_bf16(const float *weight, int K, int cin, int cout, void *wpacked, void *stream) {
    PCG_CHECK_ARG(K >= 1 && cin >= 1 && cout >= 1, "pcg_pack_weights_bf16: K=%d cin=%d cout=%d", K, cin, cout);
    const int cin_p = tc_cin_p(cin), cout_p = tc_cout_p(cout), rb = tc_rb(cin_p);
    const int chunks = rb == 128 ? (cin_p + 63) / 64 : 1;
    long total = (long)K * chunks * cout_p * (rb / 2);
    long blocks = cdiv(total, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_pack_weights<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(weight, K, cin, cout, cout_p, chunks, rb,
                                                                   reinterpret_cast<__nv_bfloat16 *>(wpacked));
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- host side
PcgEncodeTiled pcg_tensor_map_encoder() {
    static PcgEncodeTiled fn = []() -> PcgEncodeTiled {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess) return nullptr;
        return reinterpret_cast<PcgEncodeTiled>(p);
    }();
    return fn;
}

extern "C" int pcg_tc_cin_padded(int cin) { return tc_cin_p(cin); }

// grids of at least this many 128-row tiles take the two-tiles-per-CTA kernel (0 = never); env PCG_TC_MT_TILES or
// pcg_tc_set_mt_min_tiles() (tests force the path on small inputs)
static long g_mt_min_tiles = -1;
static long tc_mt_min_tiles() {
    if (g_mt_min_tiles < 0) g_mt_min_tiles = getenv("PCG_TC_MT_TILES") ? atol(getenv("PCG_TC_MT_TILES")) : 4L * PCG_SMS;
    return g_mt_min_tiles;
}
extern "C" long pcg_tc_set_mt_min_tiles(long tiles) {
    const long prev = tc_mt_min_tiles();
    if (tiles >= 0) g_mt_min_tiles = tiles;
    return prev;
}

template <int RB>
static int launch_tc(const CUtensorMap &tm, const TcParams &p, size_t smem, cudaStream_t st) {
    PCG_CUDA(cudaFuncSetAttribute(k_conv_tc<RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_tc<RB><<<(unsigned)cdiv(p.n_out, TC_ROWS), 160, smem, st>>>(tm, p);
    PCG_LAUNCHED();
    return 0;
}

static int conv_fwd_tc_impl(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch,
                            const uint32_t *tile_mask, int K, long n_out, const void *wpacked, const float *bias, int cout,
                            int act, float *out_f32, void *out_bf16, const float *proj_w_host, const float *bias_host, float *proj_out,
                            long proj_pitch, void *stream) {
    PCG_CHECK_ARG(cin >= 1 && cin <= 1024, "pcg_conv_fwd_tc: cin=%d", cin);
    PCG_CHECK_ARG(cout >= 1 && cout <= 256, "pcg_conv_fwd_tc: cout=%d must be in [1,256]", cout);
    PCG_CHECK_ARG(out_bf16 == nullptr || (cout & 15) == 0, "pcg_conv_fwd_tc: a bf16 output needs cout %% 16 == 0 (cout=%d)", cout);
    PCG_CHECK_ARG(K >= 1 && K <= 128, "pcg_conv_fwd_tc: K=%d must be <= 128", K);
    PCG_CHECK_ARG(tile_mask != nullptr, "pcg_conv_fwd_tc: tile_mask required");
    PCG_CHECK_ARG(nbr_pitch >= n_out && (nbr_pitch & 127) == 0 && ((uintptr_t)nbr & 15) == 0,
                  "pcg_conv_fwd_tc: nbr must be 16-byte aligned with a pitch that is a multiple of 128 (pitch=%ld)", nbr_pitch);
    PCG_CHECK_ARG(n_in >= 1, "pcg_conv_fwd_tc: n_in=%ld (an empty input has no tensor map)", n_in);
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)wpacked & 15) == 0, "pcg_conv_fwd_tc: operands must be 16-byte aligned");
    if (n_out <= 0) return 0;
    const int cin_p = tc_cin_p(cin), rb = tc_rb(cin_p);
    TcParams p;
    p.nbr = nbr; p.nbr_pitch = nbr_pitch; p.tile_mask = tile_mask;
    p.wpacked = reinterpret_cast<const uint8_t *>(wpacked);
    p.bias = bias; p.out_f32 = out_f32; p.out_bf16 = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    p.n_out = n_out; p.cout = cout; p.cout_p = tc_cout_p(cout); p.K = K;
    p.chunks = rb == 128 ? (cin_p + 63) / 64 : 1;
    p.act = act;
    p.feats = reinterpret_cast<const uint8_t *>(feats_bf16);
    p.row_bytes = (long)cin_p * 2;
    p.proj_out = proj_out; p.proj_pitch = proj_pitch;
    const bool proj = proj_out != nullptr;
    cudaStream_t st = as_stream(stream);
    // (the fused second head layer lives in the LSU-gather kernel only)
    if (proj || (tc_mt_min_tiles() != 0 && !(getenv("PCG_TC_GATHER") && !strcmp(getenv("PCG_TC_GATHER"), "tma")))) {
        // LSU-gather kernel.  Static shape rule (measured with tools/bench_conv.py, profiles/r01_conv_variants.md):
        //   at most one wave of tiles                        one tile per CTA, one CTA per SM, deepest ring
        //   wide outputs (weight slice >= 24 KiB per stage)  two tiles per CTA, one CTA per SM, deep ring
        //   large grids                                      two tiles per CTA, two CTAs per SM (16 producer warps per SM)
        //   other grids                                      one tile per CTA, three CTAs per SM (most independent pipelines)
        //   narrow rows / outputs (heads in "tc" mode)       one tile per CTA, small stages, many CTAs per SM
        static const int mt_t = getenv("PCG_TC_MT_T") ? atoi(getenv("PCG_TC_MT_T")) : 0;          // tuning knobs: force T / the smem budget
        static const int mt_kb = getenv("PCG_TC_MT_SMEM_KB") ? atoi(getenv("PCG_TC_MT_SMEM_KB")) : 0;
        static const int narrow_kb = getenv("PCG_TC_NARROW_KB") ? atoi(getenv("PCG_TC_NARROW_KB")) : 20;
        const long tiles = cdiv(n_out, TC_ROWS);
        const bool wide = rb == 128 && (p.cout_p & 31) == 0;
        int T, kb;
        if (!wide) { T = 1; kb = tiles <= PCG_SMS ? 100 : narrow_kb; }
        else if (tiles <= PCG_SMS) { T = 1; kb = 200; }                             // at most one CTA per SM: deepest ring
        else if (p.cout_p > 128) { T = tiles >= 2 * PCG_SMS ? 2 : 1; kb = 200; }
        else if (tiles >= tc_mt_min_tiles()) { T = 2; kb = 104; }
        else { T = 1; kb = 70; }
        if (wide && (mt_t == 1 || mt_t == 2)) T = mt_t;
        if (wide && mt_kb > 0) kb = mt_kb;
        const size_t stage_bytes = (size_t)T * TC_ROWS * rb + (size_t)p.cout_p * rb;
        int stages = (int)(((size_t)kb * 1024) / stage_bytes);
        if (stages < 2) stages = 2;
        if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
        p.stages = stages;
        p.tmem_cols = 32;
        while (p.tmem_cols < T * p.cout_p) p.tmem_cols <<= 1;
        const size_t smem = 1024 + (size_t)stages * stage_bytes + 512 + (proj && cout >= 32 ? (size_t)TC_PROJ_KP * cout * sizeof(float2) : 0);
        // tuning knobs: PCG_TC_CA bit 0 = narrow rows (RB 32 / 64), bit 1 = 128-byte rows gathered with cp.async.ca (L1-allocating);
        // PCG_TC_CARVEOUT = preferred shared-memory share of the unified L1 in percent (what is left caches the gathered rows)
        static const int ca_bits = getenv("PCG_TC_CA") ? atoi(getenv("PCG_TC_CA")) : 0;
        static const int carveout = getenv("PCG_TC_CARVEOUT") ? atoi(getenv("PCG_TC_CARVEOUT")) : -1;
        const bool ca = rb == 128 ? (ca_bits & 2) != 0 : (ca_bits & 1) != 0;
#define PCG_MT(TT, RBB, CAA)                                                                                              \
    do {                                                                                                                  \
        PCG_CUDA(cudaFuncSetAttribute(k_conv_tc_mt<TT, RBB, CAA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        if (carveout >= 0 && CAA)                                                                                         \
            PCG_CUDA(cudaFuncSetAttribute(k_conv_tc_mt<TT, RBB, CAA>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout)); \
        k_conv_tc_mt<TT, RBB, CAA><<<(unsigned)cdiv(tiles, TT), 32 * (4 * TT + 1), smem, st>>>(p, TcProjW<0>{});          \
    } while (0)
#define PCG_MT2(TT, RBB)                                                                                                  \
    do {                                                                                                                  \
        if (ca) PCG_MT(TT, RBB, true);                                                                                    \
        else PCG_MT(TT, RBB, false);                                                                                      \
    } while (0)
#define PCG_MTP(TT, RBB, PCC)                                                                                             \
    do {                                                                                                                  \
        TcProjW<PCC> pw;                                                                                                  \
        for (int kp = 0; kp < TC_PROJ_KP; ++kp)                                                                           \
            for (int c = 0; c < PCC; ++c)                                                                                 \
                pw.w2[c * TC_PROJ_KP + kp] = make_float2(proj_w_host[(2 * kp) * PCC + c],                                 \
                                                         2 * kp + 1 < TC_PROJ_K ? proj_w_host[(2 * kp + 1) * PCC + c] : 0.f); \
        for (int c = 0; c < PCC; ++c) pw.bias[c] = bias_host ? bias_host[c] : 0.f;                                        \
        PCG_CUDA(cudaFuncSetAttribute(k_conv_tc_mt<TT, RBB, false, PCC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        k_conv_tc_mt<TT, RBB, false, PCC><<<(unsigned)cdiv(tiles, TT), 32 * (4 * TT + 1), smem, st>>>(p, pw);             \
    } while (0)
#define PCG_MTP3(PCC)                                                                                                     \
    do {                                                                                                                  \
        if (rb == 64) PCG_MTP(1, 64, PCC);                                                                                \
        else if (T == 1) PCG_MTP(1, 128, PCC);                                                                            \
        else PCG_MTP(2, 128, PCC);                                                                                        \
    } while (0)
        if (proj) {
            if (cout == 16) PCG_MTP3(16);
            else if (cout == 32) PCG_MTP3(32);
            else PCG_MTP3(64);
        } else if (rb == 32) PCG_MT2(1, 32);
        else if (rb == 64) PCG_MT2(1, 64);
        else if (T == 1) PCG_MT2(1, 128);
        else PCG_MT2(2, 128);
#undef PCG_MTP3
#undef PCG_MTP
#undef PCG_MT2
#undef PCG_MT
        PCG_LAUNCHED();
        return 0;
    }
    // TMA gather4 arm only: the tensor map of the feature rows (the LSU kernel above never reads one, so it is not encoded there)
    PcgEncodeTiled encode = pcg_tensor_map_encoder();
    PCG_CHECK_ARG(encode != nullptr, "pcg_conv_fwd_tc: cuTensorMapEncodeTiled is not available from this driver");
    CUtensorMap tm;
    {
        cuuint64_t dims[2] = {(cuuint64_t)cin_p, (cuuint64_t)n_in};
        cuuint64_t strides[1] = {(cuuint64_t)cin_p * 2};
        cuuint32_t box[2] = {(cuuint32_t)(rb / 2), 1};                  // tile::gather4: one row per index
        cuuint32_t estr[2] = {1, 1};
        CUtensorMapSwizzle sw = rb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (rb == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
        CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(feats_bf16), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        PCG_CHECK_ARG(r == CUDA_SUCCESS, "pcg_conv_fwd_tc: cuTensorMapEncodeTiled failed (%d) for n_in=%ld cin_p=%d", (int)r, n_in, cin_p);
    }
    p.tmem_cols = 32;
    while (p.tmem_cols < p.cout_p) p.tmem_cols <<= 1;
    const size_t stage_bytes = TC_A_BYTES + (size_t)p.cout_p * 128;
    static const int budget_kb = getenv("PCG_TC_SMEM_KB") ? atoi(getenv("PCG_TC_SMEM_KB")) : 108;   // tuning knob (tools/bench_conv.py)
    static const int max_stages = getenv("PCG_TC_STAGES") ? atoi(getenv("PCG_TC_STAGES")) : TC_MAX_STAGES;
    int stages = (int)(((size_t)budget_kb * 1024) / stage_bytes);   // default: two CTAs per SM where the tile allows it
    if (stages < 3) stages = 3;
    if (stages > max_stages) stages = max_stages;
    if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
    p.stages = stages;
    const size_t smem = 1024 + (size_t)stages * stage_bytes + 512;
    if (rb == 32) return launch_tc<32>(tm, p, smem, st);
    if (rb == 64) return launch_tc<64>(tm, p, smem, st);
    return launch_tc<128>(tm, p, smem, st);
}

extern "C" int pcg_conv_fwd_tc(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch,
                               const uint32_t *tile_mask, int K, long n_out, const void *wpacked, const float *bias, int cout,
                               int act, float *out_f32, void *out_bf16, void *stream) {
    return conv_fwd_tc_impl(feats_bf16, n_in, cin, nbr, nbr_pitch, tile_mask, K, n_out, wpacked, bias, cout, act, out_f32, out_bf16,
                            nullptr, nullptr, nullptr, 0, stream);
}

// First layer of an occupancy head with the second (cout -> 1, 27 offsets) folded into its epilogue, GEMM-first:
// proj_out[k][o] = act(bias + conv(o)) . proj_w[k][:] from the fp32 accumulator row (the hidden row is never rounded to bf16 and,
// with out_f32 == out_bf16 == NULL, never written).  proj_w ([27][cout] fp32) and bias_host ([cout] fp32 or NULL) are HOST pointers:
// they are copied into the launch's kernel parameters (constant bank) -- the only arguments of the ABI that are not device memory.
// pcg_gather_sum then forms logit[o] = bias2 + sum_k proj_out[k][nbr[k][o]].
extern "C" int pcg_conv_fwd_tc_proj(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, long nbr_pitch,
                                    const uint32_t *tile_mask, int K, long n_out, const void *wpacked, const float *bias_host, int cout,
                                    int act, float *out_f32, void *out_bf16, const float *proj_w, int proj_k, float *proj_out,
                                    long proj_pitch, void *stream) {
    PCG_CHECK_ARG(proj_w != nullptr && proj_out != nullptr, "pcg_conv_fwd_tc_proj: proj_w / proj_out required");
    PCG_CHECK_ARG(proj_k == TC_PROJ_K, "pcg_conv_fwd_tc_proj: proj_k=%d must be %d (a 3x3x3 second layer)", proj_k, TC_PROJ_K);
    PCG_CHECK_ARG(cout == 16 || cout == 32 || cout == 64, "pcg_conv_fwd_tc_proj: cout=%d must be 16, 32 or 64", cout);
    PCG_CHECK_ARG(cin >= 17, "pcg_conv_fwd_tc_proj: cin=%d must be >= 17 (64- or 128-byte operand rows)", cin);
    PCG_CHECK_ARG(proj_pitch >= n_out, "pcg_conv_fwd_tc_proj: proj_pitch=%ld < n_out=%ld", proj_pitch, n_out);
    return conv_fwd_tc_impl(feats_bf16, n_in, cin, nbr, nbr_pitch, tile_mask, K, n_out, wpacked, nullptr, cout, act, out_f32, out_bf16,
                            proj_w, bias_host, proj_out, proj_pitch, stream);
}

// out[o] = act(bias + sum_k planes[k][nbr[k][o]]), k ascending (one fixed-order fp32 chain per output); absent neighbours add nothing.
// The K index loads of a thread are coalesced lines of nbr, the K plane reads land in the neighbourhood of o in every plane.
template <int K>
__global__ void __launch_bounds__(256) k_gather_sum(const float *__restrict__ planes, long plane_pitch, const int32_t *__restrict__ nbr,
                                                    long nbr_pitch, long n_out, const float *__restrict__ bias, int act,
                                                    float *__restrict__ out) {
    const long o = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n_out) return;
    int idx[K];
#pragma unroll
    for (int k = 0; k < K; ++k) idx[k] = __ldg(nbr + (long)k * nbr_pitch + o);
    float v[K];
#pragma unroll
    for (int k = 0; k < K; ++k) v[k] = idx[k] >= 0 ? __ldg(planes + (long)k * plane_pitch + idx[k]) : 0.f;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) acc += v[k];
    out[o] = pcg_apply_act(acc + (bias ? __ldg(bias) : 0.f), act);
}

extern "C" int pcg_gather_sum(const float *planes, long plane_pitch, const int32_t *nbr, long nbr_pitch, int K, long n_out,
                              const float *bias, int act, float *out, void *stream) {
    PCG_CHECK_ARG(K == TC_PROJ_K, "pcg_gather_sum: K=%d must be %d", K, TC_PROJ_K);
    PCG_CHECK_ARG(nbr_pitch >= n_out && plane_pitch >= 1, "pcg_gather_sum: nbr_pitch=%ld plane_pitch=%ld n_out=%ld", nbr_pitch, plane_pitch, n_out);
    if (n_out <= 0) return 0;
    k_gather_sum<TC_PROJ_K><<<(unsigned)cdiv(n_out, 256), 256, 0, as_stream(stream)>>>(planes, plane_pitch, nbr, nbr_pitch, n_out, bias, act, out);
    PCG_LAUNCHED();
    return 0;
}
