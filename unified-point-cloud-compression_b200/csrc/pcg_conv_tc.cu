// Sparse convolution on the 5th-gen tensor cores (tcgen05, sm_100a): output-stationary
// gather-GEMM with fp32 accumulators in TMEM and no scatter.
//
//   out[o][:] = act(bias + sum_{k active} F[nbr[k][o]][:] @ W[k])       (bf16 x bf16 -> fp32)
//
// One CTA owns one 128-row output tile (UMMA M = 128, N = cout <= 256, K = 16 per instruction):
//   warps 0-3  producers, then epilogue.  Thread r gathers input row nbr[k][r] of every active
//              kernel offset k straight into the K-major SWIZZLE_128B operand image with
//              16-byte cp.async (zero-fill for absent neighbours); thread 0 also starts the
//              bulk (TMA 1-D) copy of the pre-swizzled weight slice.  Completion is published
//              to the MMA warp through mbarriers after fence.proxy.async.
//   warp 4     allocates TMEM, one elected lane issues tcgen05.mma and commits each stage back
//              to the producers (tcgen05.commit -> mbarrier).
// Kernel offsets no row of the tile uses are skipped (tile_mask from pcg_kernel_map).  Every output
// element is one fixed-order chain of MMAs over ascending (k, cin-chunk): bit-reproducible, and
// independent of which tile a row lands in (absent neighbours contribute exact zeros).
#include "pcg_common.cuh"

#define TC_ROWS 128
#define TC_KCHUNK 64                       // bf16 elements per 128-byte swizzled row
#define TC_A_BYTES (TC_ROWS * 128)         // 16 KiB per stage
#define TC_LAG 2                           // cp.async groups in flight per producer thread
#define TC_MAX_STAGES 6

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major SWIZZLE_128B shared-memory matrix descriptor (rows of 128 B, 8-row groups 1024 B apart)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);  // start address
    d |= (uint64_t)1 << 16;                   // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;         // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                   // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                   // SWIZZLE_128B
    return d;
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct TcParams {
    const __nv_bfloat16 *feats;
    const int32_t *nbr;
    const uint32_t *tile_mask;
    const uint8_t *wpacked;
    const float *bias;
    float *out_f32;
    __nv_bfloat16 *out_bf16;
    long n_out;
    int cin, cout, K;
    int chunks;      // ceil(cin / 64)
    int stages;
    int tmem_cols;   // power of two >= max(32, cout)
    int act;
};

__global__ void __launch_bounds__(160) k_conv_tc(const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte aligned operand ring (SWIZZLE_128B requirement), barriers behind it
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const uint32_t b_bytes = (uint32_t)p.cout * 128u;
    const uint32_t stage_bytes = TC_A_BYTES + b_bytes;
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
            mbar_init(smem_u32(full_bar + s), TC_ROWS);   // 128 producer arrivals (+ weight bytes via tx)
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
    const int n_items = n_active * p.chunks;

    if (warp < 4) {
        // ------------------------------------------------------------------ producers
        const int r = tid;
        const long o = tile * TC_ROWS + r;
        const bool live = o < p.n_out;
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t xr = (uint32_t)(r & 7);
        int nb = -1;
        for (int it = 0; it < n_items; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
            const int a = it / p.chunks, cc = it - a * p.chunks;
            const int k = active[a];
            if (cc == 0) nb = live ? __ldg(p.nbr + (long)k * p.n_out + o) : -1;
            mbar_wait(smem_u32(empty_bar + s), ph ^ 1u);
            const uint32_t a_dst = base + (uint32_t)s * stage_bytes + row_off;
            const int c0 = cc * TC_KCHUNK;
            const __nv_bfloat16 *src = p.feats + (nb >= 0 ? (long)nb * p.cin + c0 : 0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool ok = nb >= 0 && (c0 + j * 8) < p.cin;
                cp_async16(a_dst + (((uint32_t)j ^ xr) << 4), ok ? (const void *)(src + j * 8) : (const void *)p.feats, ok ? 16u : 0u);
            }
            cp_async_commit();
            if (tid == 0) {
                const uint32_t fb = smem_u32(full_bar + s);
                mbar_expect_tx(fb, b_bytes);
                bulk_g2s(base + (uint32_t)s * stage_bytes + TC_A_BYTES,
                         p.wpacked + ((size_t)k * p.chunks + cc) * b_bytes, b_bytes, fb);
            }
            if (it >= TC_LAG) {
                cp_async_wait<TC_LAG>();
                fence_proxy_async();
                mbar_arrive(smem_u32(full_bar + (it - TC_LAG) % p.stages));
            }
        }
        // drain the last TC_LAG groups
        cp_async_wait<0>();
        fence_proxy_async();
        for (int it = (n_items > TC_LAG ? n_items - TC_LAG : 0); it < n_items; ++it)
            mbar_arrive(smem_u32(full_bar + it % p.stages));

        // ------------------------------------------------------------------ epilogue
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
            if (live) {
                float f[16];
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    f[j] = pcg_apply_act(__uint_as_float(v[j]) + (p.bias ? __ldg(p.bias + c0 + j) : 0.f), p.act);
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
        }
        tc_fence_before();
    } else {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.cout >> 3) << 17) | ((TC_ROWS >> 4) << 24);
            for (int it = 0; it < n_items; ++it) {
                const int s = it % p.stages;
                const uint32_t ph = (uint32_t)(it / p.stages) & 1u;
                mbar_wait(smem_u32(full_bar + s), ph);
                tc_fence_after();
                const uint32_t a_addr = base + (uint32_t)s * stage_bytes;
                const uint64_t adesc = umma_desc_sw128(a_addr), bdesc = umma_desc_sw128(a_addr + TC_A_BYTES);
#pragma unroll
                for (int kk = 0; kk < TC_KCHUNK / 16; ++kk)  // +32 B per K=16 step inside the swizzle atom
                    umma_bf16(tmem_d, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (it | kk) ? 1u : 0u);
                umma_commit(smem_u32(empty_bar + s));
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

// ---------------------------------------------------------------------------- weight packing
// wpacked[(k*chunks + cc)][n][64] bf16: B operand image (N rows x 64 K-elements, K-major, 128B swizzle)
// element (n, kk) <- W[k][cc*64 + kk][n] (zero beyond cin)
__global__ void k_pack_weights(const float *__restrict__ w, int K, int cin, int cout, int chunks, __nv_bfloat16 *__restrict__ out) {
    long total = (long)K * chunks * cout * 64;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        int kk = (int)(i & 63);
        long t = i >> 6;
        int n = (int)(t % cout);
        t /= cout;
        int cc = (int)(t % chunks);
        int k = (int)(t / chunks);
        int c = cc * 64 + kk;
        float v = c < cin ? __ldg(w + ((long)k * cin + c) * cout + n) : 0.f;
        long stage = ((long)k * chunks + cc) * cout * 64;  // elements
        int chunk16 = (kk >> 3) ^ (n & 7);
        out[stage + (long)n * 64 + chunk16 * 8 + (kk & 7)] = __float2bfloat16_rn(v);
    }
}

extern "C" size_t pcg_packed_weight_bytes(int K, int cin, int cout) {
    int chunks = (cin + 63) / 64;
    return (size_t)K * chunks * cout * 128;
}

extern "C" int pcg_pack_weights_bf16(const float *weight, int K, int cin, int cout, void *wpacked, void *stream) {
    PCG_CHECK_ARG(K >= 1 && cin >= 1 && cout >= 1, "pcg_pack_weights_bf16: K=%d cin=%d cout=%d", K, cin, cout);
    int chunks = (cin + 63) / 64;
    long total = (long)K * chunks * cout * 64;
    long blocks = cdiv(total, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_pack_weights<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(weight, K, cin, cout, chunks,
                                                                   reinterpret_cast<__nv_bfloat16 *>(wpacked));
    PCG_LAUNCHED();
    return 0;
}

extern "C" int pcg_conv_fwd_tc(const void *feats_bf16, long n_in, int cin, const int32_t *nbr, const uint32_t *tile_mask,
                               int K, long n_out, const void *wpacked, const float *bias, int cout, int act,
                               float *out_f32, void *out_bf16, void *stream) {
    (void)n_in;
    PCG_CHECK_ARG(cin % 8 == 0 && cin >= 8, "pcg_conv_fwd_tc: cin=%d must be a multiple of 8", cin);
    PCG_CHECK_ARG(cout % 16 == 0 && cout >= 16 && cout <= 256, "pcg_conv_fwd_tc: cout=%d must be a multiple of 16 in [16,256]", cout);
    PCG_CHECK_ARG(K >= 1 && K <= 128, "pcg_conv_fwd_tc: K=%d must be <= 128", K);
    PCG_CHECK_ARG(tile_mask != nullptr, "pcg_conv_fwd_tc: tile_mask required");
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)wpacked & 15) == 0, "pcg_conv_fwd_tc: operands must be 16-byte aligned");
    if (n_out <= 0) return 0;
    TcParams p;
    p.feats = reinterpret_cast<const __nv_bfloat16 *>(feats_bf16);
    p.nbr = nbr; p.tile_mask = tile_mask;
    p.wpacked = reinterpret_cast<const uint8_t *>(wpacked);
    p.bias = bias; p.out_f32 = out_f32; p.out_bf16 = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    p.n_out = n_out; p.cin = cin; p.cout = cout; p.K = K;
    p.chunks = (cin + 63) / 64;
    p.act = act;
    p.tmem_cols = 32;
    while (p.tmem_cols < cout) p.tmem_cols <<= 1;
    const size_t stage_bytes = TC_A_BYTES + (size_t)cout * 128;
    int stages = (int)((108 * 1024) / stage_bytes);   // two CTAs per SM where the tile allows it
    if (stages < 3) stages = 3;
    if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
    p.stages = stages;
    const size_t smem = 1024 + (size_t)stages * stage_bytes + 512;
    PCG_CUDA(cudaFuncSetAttribute(k_conv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_tc<<<(unsigned)cdiv(n_out, TC_ROWS), 160, smem, as_stream(stream)>>>(p);
    PCG_LAUNCHED();
    return 0;
}
