// Dense bf16 product T = F @ Wmat on tcgen05 -- the first half of the GEMM-first generative transposed convolution
// (T[p][k][:] = F[p] @ W[k] for all 125 offsets at once: every (parent, offset) pair hits, so nothing is gathered).
//
// One CTA owns one 128-row tile of F for its whole life: the tile is loaded once by two plain TMA tile loads
// (128B-swizzled K-major image) and stays in shared memory while the CTA walks its column groups of 256.
//   warp 5     producer: streams the pre-packed weight group (chunks x 32 KiB) through a 2-stage ring with bulk copies
//   warp 4     allocates all 512 TMEM columns = two 128x256 fp32 accumulators; one lane issues the tcgen05.mma chain
//   warps 0-3  epilogue: tcgen05.ld the finished accumulator, round to bf16, stage 32x64 blocks in 128B-swizzled shared
//              memory and hand them to TMA tile stores (full-line writes, rows / columns past the edge clipped by
//              the tensor map) -- overlapped with the MMAs of the next group through the second accumulator
// The kernel is bound by the HBM write of T (2 bytes per 2*cin FLOPs), not by the tensor pipe.
#include "pcg_tc.cuh"

#define GT_ROWS 128
#define GT_NCOLS 256
#define GT_STAGES 2

struct GemmParams {
    const uint8_t *wpacked;     // [group][chunk][256][128 B] pre-swizzled (pcg_pack_weights_bf16 of a (G, cin, 256) tensor)
    __nv_bfloat16 *out;         // [n][n_cols]
    long n, n_cols;
    int chunks;                 // 64-wide K chunks (1 or 2)
    int groups;                 // ceil(n_cols / 256)
};

#define GT_STAGING (4 * 2 * 4096)   // per epilogue warp: two 32-row x 128-byte staging blocks

__global__ void __launch_bounds__(192) k_gemm_first(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tm_out,
                                                    const GemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const uint32_t a_bytes = (uint32_t)p.chunks * GT_ROWS * 128u;            // A tile
    const uint32_t b_bytes = (uint32_t)p.chunks * GT_NCOLS * 128u;           // one weight group
    const uint32_t b_base = base + a_bytes;
    const uint32_t stg_base = b_base + GT_STAGES * b_bytes;
    uint8_t *ctl = sm + a_bytes + (size_t)GT_STAGES * b_bytes + GT_STAGING;
    uint64_t *a_full = reinterpret_cast<uint64_t *>(ctl);
    uint64_t *b_full = a_full + 1;               // [GT_STAGES]
    uint64_t *b_empty = b_full + GT_STAGES;      // [GT_STAGES]
    uint64_t *t_full = b_empty + GT_STAGES;      // [2]
    uint64_t *t_empty = t_full + 2;              // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(t_empty + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long tile = blockIdx.x;
    const int g0 = blockIdx.y, gstep = gridDim.y;
    const int n_it = p.groups > g0 ? (p.groups - g0 + gstep - 1) / gstep : 0;

    if (tid == 0) {
        mbar_init(smem_u32(a_full), 1);
        for (int s = 0; s < GT_STAGES; ++s) {
            mbar_init(smem_u32(b_full + s), 1);
            mbar_init(smem_u32(b_empty + s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(t_full + b), 1);
            mbar_init(smem_u32(t_empty + b), 4);       // one arrive per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 5) {
        // ------------------------------------------------------------------ producer
        if (lane == 0 && n_it > 0) {
            const uint32_t af = smem_u32(a_full);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(af), "r"(a_bytes) : "memory");
            for (int cc = 0; cc < p.chunks; ++cc)
                tma_load_2d(base + (uint32_t)cc * GT_ROWS * 128u, &tm, cc * 64, (int)(tile * GT_ROWS), af);
            for (int it = 0; it < n_it; ++it) {
                const int s = it % GT_STAGES;
                const uint32_t n_use = (uint32_t)(it / GT_STAGES);
                mbar_wait(smem_u32(b_empty + s), (n_use & 1u) ^ 1u);
                const uint32_t fb = smem_u32(b_full + s);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(b_bytes) : "memory");
                const uint8_t *src = p.wpacked + (size_t)(g0 + it * gstep) * b_bytes;
                for (int cc = 0; cc < p.chunks; ++cc)
                    bulk_g2s(b_base + (uint32_t)s * b_bytes + (uint32_t)cc * GT_NCOLS * 128u, src + (size_t)cc * GT_NCOLS * 128u,
                             GT_NCOLS * 128u, fb);
            }
        }
        __syncwarp();
    } else if (warp == 4) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0 && n_it > 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(GT_NCOLS >> 3) << 17) | ((GT_ROWS >> 4) << 24);
            mbar_wait(smem_u32(a_full), 0);
            for (int it = 0; it < n_it; ++it) {
                const int s = it % GT_STAGES, buf = it & 1;
                mbar_wait(smem_u32(b_full + s), (uint32_t)(it / GT_STAGES) & 1u);
                mbar_wait(smem_u32(t_empty + buf), ((uint32_t)(it >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d = tmem_d + (uint32_t)buf * GT_NCOLS;
                for (int cc = 0; cc < p.chunks; ++cc) {
                    const uint64_t adesc = umma_desc<128>(base + (uint32_t)cc * GT_ROWS * 128u);
                    const uint64_t bdesc = umma_desc<128>(b_base + (uint32_t)s * b_bytes + (uint32_t)cc * GT_NCOLS * 128u);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        umma_bf16(d, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, (cc | kk) ? 1u : 0u);
                }
                umma_commit(smem_u32(b_empty + s));
                umma_commit(smem_u32(t_full + buf));
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ epilogue (warp w reads TMEM lanes 32w..32w+31)
        const uint32_t stg = stg_base + (uint32_t)warp * 8192u;
        const uint32_t my_row = (uint32_t)lane * 128u, xr = (uint32_t)(lane & 7);
        int sb = 0;
        for (int it = 0; it < n_it; ++it) {
            const int buf = it & 1;
            const long col0 = (long)(g0 + it * gstep) * GT_NCOLS;
            mbar_wait(smem_u32(t_full + buf), (uint32_t)(it >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_addr = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)buf * GT_NCOLS;
            for (int c0 = 0; c0 < GT_NCOLS && col0 + c0 < p.n_cols; c0 += 64) {
                uint32_t v0[32], v1[32];
                tmem_ld32(t_addr + (uint32_t)c0, v0);
                tmem_ld32(t_addr + (uint32_t)c0 + 32u, v1);
                if (lane == 0) bulk_wait_read<1>();       // the store that last read this staging block is done with it
                __syncwarp();
                const uint32_t dst = stg + (uint32_t)sb * 4096u + my_row;
#pragma unroll
                for (int j = 0; j < 8; ++j) {            // 16-byte chunk j = columns 8j..8j+7 of this 64-column block
                    const uint32_t *src = j < 4 ? v0 + 8 * j : v1 + 8 * (j - 4);
                    uint32_t pk[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(src[2 * q]), __uint_as_float(src[2 * q + 1]));
                        pk[q] = *reinterpret_cast<uint32_t *>(&h);
                    }
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)j ^ xr) << 4)), "r"(pk[0]),
                                 "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                                 : "memory");
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tm_out, stg + (uint32_t)sb * 4096u, (int)(col0 + c0), (int)(tile * GT_ROWS + warp * 32));
                    bulk_commit();
                }
                sb ^= 1;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(t_empty + buf));
        }
        if (lane == 0) bulk_wait_read<0>();               // staging must outlive the last stores' reads
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
    }
}

// out_bf16[n][0..n_cols) = feats[n][:] @ Wmat[:, 0..n_cols); see include/pcg.h.
extern "C" int pcg_gemm_bf16_grouped(const void *feats_bf16, long n, int cin, const void *wpacked, long n_cols, void *out_bf16,
                                     void *stream) {
    PCG_CHECK_ARG(n_cols >= 16 && (n_cols & 15) == 0 && n_cols < (1L << 30), "pcg_gemm_bf16_grouped: n_cols=%ld must be a multiple of 16", n_cols);
    PCG_CHECK_ARG(cin >= 33 && cin <= 128, "pcg_gemm_bf16_grouped: cin=%d must be in [33,128] (64 or 128 padded channels)", cin);
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)wpacked & 15) == 0 && ((uintptr_t)out_bf16 & 15) == 0,
                  "pcg_gemm_bf16_grouped: operands must be 16-byte aligned");
    if (n <= 0) return 0;
    const int cin_p = (cin + 63) & ~63;
    PcgEncodeTiled encode = pcg_tensor_map_encoder();
    PCG_CHECK_ARG(encode != nullptr, "pcg_gemm_bf16_grouped: cuTensorMapEncodeTiled is not available from this driver");
    CUtensorMap tm;
    {
        cuuint64_t dims[2] = {(cuuint64_t)cin_p, (cuuint64_t)n};
        cuuint64_t strides[1] = {(cuuint64_t)cin_p * 2};
        cuuint32_t box[2] = {64, GT_ROWS};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(feats_bf16), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        PCG_CHECK_ARG(r == CUDA_SUCCESS, "pcg_gemm_bf16_grouped: cuTensorMapEncodeTiled failed (%d) for n=%ld cin_p=%d", (int)r, n, cin_p);
    }
    CUtensorMap tm_out;
    {
        cuuint64_t dims[2] = {(cuuint64_t)n_cols, (cuuint64_t)n};
        cuuint64_t strides[1] = {(cuuint64_t)n_cols * 2};
        cuuint32_t box[2] = {64, 32};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&tm_out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out_bf16, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        PCG_CHECK_ARG(r == CUDA_SUCCESS, "pcg_gemm_bf16_grouped: output tensor map failed (%d) for n=%ld n_cols=%ld", (int)r, n, n_cols);
    }
    GemmParams p;
    p.wpacked = reinterpret_cast<const uint8_t *>(wpacked);
    p.out = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    p.n = n; p.n_cols = n_cols;
    p.chunks = cin_p / 64;
    p.groups = (int)cdiv(n_cols, GT_NCOLS);
    const long tiles = cdiv(n, GT_ROWS);
    // split the column groups over blockIdx.y until the grid fills the machine a few times over
    int nsplit = 1;
    while (tiles * nsplit < 4L * PCG_SMS && nsplit * 2 <= p.groups) nsplit *= 2;
    const size_t smem = 1024 + (size_t)p.chunks * GT_ROWS * 128 + (size_t)GT_STAGES * p.chunks * GT_NCOLS * 128 + GT_STAGING + 256;
    PCG_CUDA(cudaFuncSetAttribute(k_gemm_first, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_gemm_first<<<dim3((unsigned)tiles, (unsigned)nsplit), 192, smem, as_stream(stream)>>>(tm, tm_out, p);
    PCG_LAUNCHED();
    return 0;
}
