// Weight gradient of the sparse convolution (training step, /root/reference/train.py:195-234 through ME's
// ConvolutionBackwardGPU):  dW[k] = sum_o F[nbr[k][o]]^T (x) G[o]   -- fp32, deterministic.
// One CTA owns a (k, 16 input channels, 64 output channels) block of dW and walks the output rows in 128-row tiles in
// ascending order (fixed summation order, no atomics); tiles whose mask bit k is clear are skipped, which removes the
// empty (tile, offset) combinations of strided and generative maps.  The input gradient needs no kernel of its own:
// it is the forward convolution of G over the inverse map with W[k]^T (autograd.py).
#include <stdlib.h>

#include "pcg_common.cuh"

#define BW_ROWS 128
#define BW_CI 16
#define BW_CO 64

__global__ void __launch_bounds__(256) k_conv_bwd_weight(const float *__restrict__ feats, int cin, const int32_t *__restrict__ nbr,
                                                         long pitch, const uint32_t *__restrict__ tile_mask, long n_out,
                                                         const float *__restrict__ gout, int cout, float *__restrict__ dw, int K) {
    __shared__ float xs[BW_ROWS][BW_CI];
    __shared__ __align__(16) float gs[BW_ROWS][BW_CO];
    __shared__ int idx_s[BW_ROWS];
    // blockIdx.x = slice * K + k: a slice is a contiguous range of tiles whose partial block goes to dw + slice * K*cin*cout
    const int k = blockIdx.x % K, slice = blockIdx.x / K, ci0 = blockIdx.y * BW_CI, co0 = blockIdx.z * BW_CO;
    dw += (size_t)slice * K * cin * cout;
    const int tid = threadIdx.x, ci = tid >> 4, cq = (tid & 15) * 4;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const long tiles = (n_out + BW_ROWS - 1) / BW_ROWS;
    const long n_slices = gridDim.x / K;            // slice s = tiles s, s + S, ... (interleaved, see k_conv_bwd_weight_tc)
    for (long t = slice; t < tiles; t += n_slices) {
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

static long bwf_slices(int K, int cin, int cout, long tiles) {
    const long ctas = (long)K * cdiv(cin, BW_CI) * cdiv(cout, BW_CO);
    long s = cdiv(4L * PCG_SMS, ctas);        // small (K, cin, cout) blocks (the narrow head layers) would leave most SMs idle
    if (s > tiles) s = tiles;
    if (s > 256) s = 256;
    return s < 1 ? 1 : s;
}

extern "C" size_t pcg_conv_bwd_weight_f32_scratch_bytes(int K, int cin, int cout, long n_out) {
    const long tiles = cdiv(n_out > 0 ? n_out : 1, BW_ROWS);
    const long s = bwf_slices(K, cin, cout, tiles);
    return s > 1 ? (size_t)s * K * cin * cout * sizeof(float) + 256 : 0;
}

__global__ void k_bwd_reduce_partials(const float *__restrict__ partial, int slices, long n, float *__restrict__ dw);

// scratch (optional): with pcg_conv_bwd_weight_f32_scratch_bytes() bytes the tiles are split into slices that run in parallel and
// whose partial sums are added in ascending order afterwards; without it one CTA per (k, channel block) walks every tile.
extern "C" int pcg_conv_bwd_weight_f32_sliced(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, const uint32_t *tile_mask,
                                              int K, long n_out, const float *gout, int cout, void *scratch, size_t scratch_bytes,
                                              float *dweight, void *stream) {
    PCG_CHECK_ARG(K >= 1 && K <= 65535 && cin >= 1 && cout >= 1, "pcg_conv_bwd_weight_f32: K=%d cin=%d cout=%d", K, cin, cout);
    PCG_CHECK_ARG(tile_mask == nullptr || K <= 128, "pcg_conv_bwd_weight_f32: tile_mask needs K <= 128");
    cudaStream_t st = as_stream(stream);
    const long tiles = cdiv(n_out > 0 ? n_out : 1, BW_ROWS);
    long slices = bwf_slices(K, cin, cout, tiles);
    const long n_w = (long)K * cin * cout;
    if (slices > 1 && (!scratch || scratch_bytes < (size_t)slices * n_w * sizeof(float))) slices = 1;
    float *dst = slices > 1 ? reinterpret_cast<float *>(scratch) : dweight;
    dim3 grid((unsigned)(K * slices), (unsigned)cdiv(cin, BW_CI), (unsigned)cdiv(cout, BW_CO));
    k_conv_bwd_weight<<<grid, 256, 0, st>>>(feats, cin, nbr, nbr_pitch, tile_mask, n_out, gout, cout, dst, K);
    PCG_LAUNCHED();
    if (slices > 1) {
        long blocks = cdiv(n_w, 256);
        if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
        k_bwd_reduce_partials<<<(unsigned)blocks, 256, 0, st>>>(dst, (int)slices, n_w, dweight);
        PCG_LAUNCHED();
    }
    return 0;
}

extern "C" int pcg_conv_bwd_weight_f32(const float *feats, int cin, const int32_t *nbr, long nbr_pitch, const uint32_t *tile_mask,
                                       int K, long n_out, const float *gout, int cout, float *dweight, void *stream) {
    return pcg_conv_bwd_weight_f32_sliced(feats, cin, nbr, nbr_pitch, tile_mask, K, n_out, gout, cout, nullptr, 0, dweight, stream);
}

// ---------------------------------------------------------------------------- weight gradient on the tensor cores
// dW[k] = F_k^T . G with F_k = the gathered input rows of offset k (absent neighbours = zero rows): per offset a GEMM whose
// REDUCTION dimension is the row index.  tcgen05 takes both operands "MN-major" for that (instruction-descriptor bits 15 / 16):
// the canonical SWIZZLE_128B MN-major tile is rows (reduction index) of 128 bytes = 64 channels, 8-row groups 1024 bytes apart,
// 64-channel blocks LBO apart -- byte for byte the image the forward kernel's LSU gather writes for a 64-channel chunk of 128
// rows.  So the producers of k_conv_tc_mt are reused unchanged: per 128-row tile they gather cin_p/64 chunk images of F (through
// nbr[k]) and cout_p/64 chunk images of G (identity rows), and one thread issues, per 16 rows, MMAs of M = 128 input channels x
// N = cout into fp32 accumulators in TMEM (one per block of 128 input channels).
// CTA (k, s) walks the tiles of slice s in ascending order (tiles whose mask bit k is clear are skipped) and writes its partial
// [cin][cout] block; k_bwd_reduce_partials then adds the slices in ascending order: one fixed summation order per element, no
// atomics (same determinism argument as the forward path).
#include "pcg_tc.cuh"

#define BWT_ROWS 128
#define BWT_CHUNK_BYTES (BWT_ROWS * 128)       // one 64-channel chunk image of a 128-row tile
#define BWT_MAX_STAGES 4

struct BwtParams {
    const uint8_t *feats;        // bf16 [n_in][cin_p]
    const uint8_t *gout;         // bf16 [n_out][cout_p]
    const int32_t *nbr;          // [K][pitch]
    const uint32_t *tile_mask;   // [tiles][4]
    float *partial;              // [S][K][cin][cout]
    long n_out, pitch, tiles;
    int cin, cout, K;
    int ci_chunks;               // 64-channel chunks of F actually present (cin_p / 64)
    int ci_chunks_p;             // rounded up to a multiple of 2 (blocks of M = 128; the padding chunk is zero-filled)
    int co_chunks;               // cout_p / 64
    int n_mma;                   // N of the instruction = roundup16(cout)
    int stages, tmem_cols;
};

// MN-major SWIZZLE_128B shared-memory matrix descriptor: 64-element blocks along M/N `lbo` bytes apart, 8-row groups along the
// reduction dimension 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr, uint32_t lbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// Warp-wide iterator over the ACTIVE tiles of a slice (tiles t0, t0 + step, ... whose mask bit k is set): the masks of the next 32
// candidate tiles are fetched by the 32 lanes at once, so the walk costs one dependent global load per 32 tiles instead of one per
// tile (the per-tile mask and index loads were ~1.5 us of exposed latency per tile on the 88 k-tile head layers).
struct BwtTileIter {
    const uint32_t *mask;
    long next_base, cur_base, t1, step;
    uint32_t bits;
    int word, bit;
    __device__ __forceinline__ long next(int lane) {         // call with the whole warp; returns -1 at the end
        while (bits == 0u) {
            if (next_base >= t1) return -1;
            const long t = next_base + (long)lane * step;
            const bool a = t < t1 && ((__ldg(mask + t * 4 + word) >> bit) & 1u);
            bits = __ballot_sync(0xffffffffu, a);
            cur_base = next_base;
            next_base += 32 * step;
        }
        const int j = __ffs(bits) - 1;
        bits &= bits - 1;
        return cur_base + (long)j * step;
    }
};

__global__ void __launch_bounds__(160) k_conv_bwd_weight_tc(const BwtParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const int n_chunks = p.ci_chunks_p + p.co_chunks;              // chunk images per stage: F chunks first, then G chunks
    const uint32_t stage_bytes = (uint32_t)n_chunks * BWT_CHUNK_BYTES;
    uint8_t *ctl = sm + (size_t)p.stages * stage_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(ctl);
    uint64_t *empty_bar = full_bar + BWT_MAX_STAGES;
    uint64_t *accum_bar = empty_bar + BWT_MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int k = blockIdx.x;
    // slice s = tiles s, s + S, s + 2S, ... (interleaved: with class-major row orders an offset is only active in one contiguous
    // range of tiles, which a contiguous split would hand to a single slice)
    const long t0 = blockIdx.y, t1 = p.tiles, tstep = gridDim.y;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(smem_u32(full_bar + s), 32 * (n_chunks - (p.ci_chunks_p - p.ci_chunks)));      // one cp.async-completion arrive per lane and GATHERED chunk image
            mbar_init(smem_u32(empty_bar + s), 1);
        }
        mbar_init(smem_u32(accum_bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the zero padding chunk of the last 128-channel block (cin <= 64, 129..192) never changes: written once per stage here
    for (int s = 0; s < p.stages; ++s)
        for (int c = p.ci_chunks; c < p.ci_chunks_p; ++c)
            for (uint32_t off = tid * 16u; off < BWT_CHUNK_BYTES; off += 160u * 16u)
                asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + (uint32_t)s * stage_bytes + (uint32_t)c * BWT_CHUNK_BYTES + off), "r"(0u) : "memory");
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;
    const int m_blocks = p.ci_chunks_p / 2;

    if (warp < 4) {
        // ------------------------------------------------------------------ producers: warp w gathers chunk images w, w+4, ...
        const int row_f = p.ci_chunks * 128, row_g = p.co_chunks * 128;       // bytes per source row
        const uint8_t *fsrc = p.feats + (lane & 7) * 16, *gsrc = p.gout + (lane & 7) * 16;
        int s = 0;
        uint32_t wrap = 0;
        BwtTileIter it{p.tile_mask, t0, t0, t1, tstep, 0u, k >> 5, k & 31};
        auto load_idx = [&](long t, int (&idx)[4]) {
#pragma unroll
            for (int u = 0; u < 4; ++u) idx[u] = __ldg(p.nbr + (long)k * p.pitch + t * BWT_ROWS + lane + 32 * u);   // padded with -1 to 128 rows
        };
        int idx_f[4], idx_n[4];
        long t = it.next(lane);
        if (t >= 0) load_idx(t, idx_f);
        while (t >= 0) {
            const long tn = it.next(lane);
            if (tn >= 0) load_idx(tn, idx_n);                                    // next tile's indices in flight under this tile's gathers
            int idx_g[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long o = t * BWT_ROWS + lane + 32 * u;
                idx_g[u] = (o < p.n_out && idx_f[u] >= 0) ? (int)o : -1;
            }
            const uint32_t fb = smem_u32(full_bar + s);
            const uint32_t st_base = base + (uint32_t)s * stage_bytes;
            mbar_wait(smem_u32(empty_bar + s), (wrap & 1u) ^ 1u);
            for (int c = warp; c < n_chunks; c += 4) {
                const uint32_t dst = st_base + (uint32_t)c * BWT_CHUNK_BYTES;
                if (c < p.ci_chunks) tc_lsu_gather_tile<128, false>(dst, fsrc + c * 128, row_f, idx_f, lane);
                else if (c < p.ci_chunks_p) continue;                          // the padding chunk stays zero
                else tc_lsu_gather_tile<128, false>(dst, gsrc + (c - p.ci_chunks_p) * 128, row_g, idx_g, lane);
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(fb) : "memory");
            }
            if (++s == p.stages) { s = 0; ++wrap; }
#pragma unroll
            for (int u = 0; u < 4; ++u) idx_f[u] = idx_n[u];
            t = tn;
        }
    } else {
        // ------------------------------------------------------------------ MMA issuer (the warp walks the tiles, lane 0 issues)
        // idesc: fp32 accumulate, bf16 x bf16, A and B MN-major (bits 15 / 16), N, M = 128
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(p.n_mma >> 3) << 17) | ((128u >> 4) << 24);
        int s = 0;
        uint32_t ph = 0, started = 0;
        BwtTileIter it{p.tile_mask, t0, t0, t1, tstep, 0u, k >> 5, k & 31};
        for (long t = it.next(lane); t >= 0; t = it.next(lane)) {
            if (lane == 0) {
                mbar_wait(smem_u32(full_bar + s), ph);
                fence_proxy_async();                      // cp.async wrote the images through the generic proxy
                tc_fence_after();
                const uint32_t a_base = base + (uint32_t)s * stage_bytes;
                const uint32_t b_base = a_base + (uint32_t)p.ci_chunks_p * BWT_CHUNK_BYTES;
                for (int mb = 0; mb < m_blocks; ++mb) {
#pragma unroll
                    for (int kk = 0; kk < BWT_ROWS / 16; ++kk) {             // 16 rows (two 8-row groups = 2048 bytes) per instruction
                        const uint64_t adesc = umma_desc_mn_sw128(a_base + (uint32_t)(2 * mb) * BWT_CHUNK_BYTES + kk * 2048, BWT_CHUNK_BYTES);
                        const uint64_t bdesc = umma_desc_mn_sw128(b_base + kk * 2048, BWT_CHUNK_BYTES);
                        umma_bf16(tmem_d + (uint32_t)(mb * p.n_mma), adesc, bdesc, idesc, (started | (uint32_t)kk) ? 1u : 0u);
                    }
                }
                umma_commit(smem_u32(empty_bar + s));
            }
            started = 1u;
            if (++s == p.stages) { s = 0; ph ^= 1u; }
            __syncwarp();
        }
        if (lane == 0) {
            if (started) umma_commit(smem_u32(accum_bar));
            *reinterpret_cast<volatile uint32_t *>(tmem_slot + 1) = started;       // read by the epilogue after the barrier below
        }
    }
    __syncwarp();
    __syncthreads();
    const uint32_t any = *reinterpret_cast<volatile uint32_t *>(tmem_slot + 1);
    if (warp < 4) {
        // ------------------------------------------------------------------ epilogue: thread = input channel (TMEM lane) of each 128-channel block
        if (any) {
            mbar_wait(smem_u32(accum_bar), 0);
            tc_fence_after();
        }
        float *dst_k = p.partial + ((size_t)blockIdx.y * p.K + k) * (size_t)p.cin * p.cout;
        for (int mb = 0; mb < m_blocks; ++mb) {
            const int ci = mb * 128 + tid;
            const uint32_t lane_base = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(mb * p.n_mma);
            for (int c0 = 0; c0 < p.cout; c0 += 16) {
                uint32_t v[16];
                if (any) tmem_ld16(lane_base + (uint32_t)c0, v);
                else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = 0u;
                }
                if (ci < p.cin) {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (c0 + j < p.cout) dst_k[(size_t)ci * p.cout + c0 + j] = __uint_as_float(v[j]);
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)p.tmem_cols) : "memory");
    }
}

// ---------------------------------------------------------------------------- narrow outputs: offsets inner, G loaded once per tile
// For cout <= 16 and K <= 32 (the occupancy-head layers on the candidate sets: 88 k tiles x 27 dense offsets in the training batch)
// the kernel above streams the padded F and G copies 27 times each and is HBM bound by its own padding.  Here a CTA owns a slice of
// TILES and walks the offsets inside: the G tile is gathered once per tile, the F tiles of the 27 offsets of one tile are
// neighbouring rows (L2 hits), and every offset has its own accumulator (TMEM columns 16 k .. 16 k + 15, 432 of 512 columns).
// Same summation order per element: tiles ascending inside a slice, slices added in ascending order.
#define BWK_STAGES 4
__global__ void __launch_bounds__(160) k_conv_bwd_weight_tc_kin(const BwtParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - raw);
    const uint32_t stage_bytes = (uint32_t)p.ci_chunks_p * BWT_CHUNK_BYTES;        // F chunk images of one (tile, offset)
    const uint32_t g_base = base + BWK_STAGES * stage_bytes;                        // two G images (one 64-channel chunk each)
    uint8_t *ctl = sm + (size_t)BWK_STAGES * stage_bytes + 2 * BWT_CHUNK_BYTES;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(ctl);            // [BWK_STAGES]
    uint64_t *empty_bar = full_bar + BWK_STAGES;                       // [BWK_STAGES]
    uint64_t *gfull_bar = empty_bar + BWK_STAGES;                      // [2]
    uint64_t *gempty_bar = gfull_bar + 2;                              // [2]
    uint64_t *accum_bar = gempty_bar + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long t0 = blockIdx.x, tstep = gridDim.x;
    if (tid == 0) {
        for (int s = 0; s < BWK_STAGES; ++s) {
            mbar_init(smem_u32(full_bar + s), 32 * p.ci_chunks);
            mbar_init(smem_u32(empty_bar + s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(gfull_bar + s), 32);
            mbar_init(smem_u32(gempty_bar + s), 1);
        }
        mbar_init(smem_u32(accum_bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int s = 0; s < BWK_STAGES; ++s)                              // the zero padding chunk (cin <= 64): written once
        for (int c = p.ci_chunks; c < p.ci_chunks_p; ++c)
            for (uint32_t off = tid * 16u; off < BWT_CHUNK_BYTES; off += 160u * 16u)
                asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + (uint32_t)s * stage_bytes + (uint32_t)c * BWT_CHUNK_BYTES + off), "r"(0u) : "memory");
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;
    const int row_f = p.ci_chunks * 128, row_g = p.co_chunks * 128;
    const int m_blocks = p.ci_chunks_p / 2;                            // 1 (cin <= 128)

    if (warp < 3) {
        // ------------------------------------------------------------------ F producers: warp w gathers the (tile, offset) items w, w + 3, ...
        const uint8_t *fsrc = p.feats + (lane & 7) * 16;
        long item = 0;                                                  // running index over the active (tile, offset) pairs of this CTA
        for (long t = t0; t < p.tiles; t += tstep) {
            uint32_t m = __ldg(p.tile_mask + t * 4);                   // K <= 32: one mask word
            while (m) {
                const int k = __ffs(m) - 1;
                m &= m - 1;
                if (item % 3 == warp) {
                    int idx[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) idx[u] = __ldg(p.nbr + (long)k * p.pitch + t * BWT_ROWS + lane + 32 * u);
                    const int s = (int)(item % BWK_STAGES);
                    const uint32_t par = (uint32_t)((item / BWK_STAGES) & 1);
                    mbar_wait(smem_u32(empty_bar + s), par ^ 1u);
                    const uint32_t fb = smem_u32(full_bar + s);
                    for (int c = 0; c < p.ci_chunks; ++c) {          // (the padding chunk stays zero)
                        const uint32_t dst = base + (uint32_t)s * stage_bytes + (uint32_t)c * BWT_CHUNK_BYTES;
                        tc_lsu_gather_tile<128, false>(dst, fsrc + c * 128, row_f, idx, lane);
                        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(fb) : "memory");
                    }
                }
                ++item;
            }
        }
    } else if (warp == 3) {
        // ------------------------------------------------------------------ G producer: one image per tile, double buffered
        const uint8_t *gsrc = p.gout + (lane & 7) * 16;
        long n = 0;
        for (long t = t0; t < p.tiles; t += tstep) {
            if (__ldg(p.tile_mask + t * 4) == 0u) continue;
            int idx[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long o = t * BWT_ROWS + lane + 32 * u;
                idx[u] = o < p.n_out ? (int)o : -1;
            }
            const int s = (int)(n & 1);
            mbar_wait(smem_u32(gempty_bar + s), (uint32_t)((n >> 1) & 1) ^ 1u);
            tc_lsu_gather_tile<128, false>(g_base + (uint32_t)s * BWT_CHUNK_BYTES, gsrc, row_g, idx, lane);
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(gfull_bar + s)) : "memory");
            ++n;
        }
    } else {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(16 >> 3) << 17) | ((128u >> 4) << 24);
            long item = 0, n = 0;
            uint32_t started = 0;                                       // bit k: accumulator k holds a value
            for (long t = t0; t < p.tiles; t += tstep) {
                uint32_t m = __ldg(p.tile_mask + t * 4);
                if (m == 0u) continue;
                const int gs = (int)(n & 1);
                mbar_wait(smem_u32(gfull_bar + gs), (uint32_t)((n >> 1) & 1));
                while (m) {
                    const int k = __ffs(m) - 1;
                    m &= m - 1;
                    const int s = (int)(item % BWK_STAGES);
                    mbar_wait(smem_u32(full_bar + s), (uint32_t)((item / BWK_STAGES) & 1));
                    fence_proxy_async();
                    tc_fence_after();
                    const uint32_t a_base = base + (uint32_t)s * stage_bytes;
#pragma unroll
                    for (int kk = 0; kk < BWT_ROWS / 16; ++kk) {
                        const uint64_t adesc = umma_desc_mn_sw128(a_base + kk * 2048, BWT_CHUNK_BYTES);
                        const uint64_t bdesc = umma_desc_mn_sw128(g_base + (uint32_t)gs * BWT_CHUNK_BYTES + kk * 2048, BWT_CHUNK_BYTES);
                        umma_bf16(tmem_d + (uint32_t)(k * 16), adesc, bdesc, idesc, (((started >> k) & 1u) | (uint32_t)kk) ? 1u : 0u);
                    }
                    started |= 1u << k;
                    umma_commit(smem_u32(empty_bar + s));
                    ++item;
                }
                umma_commit(smem_u32(gempty_bar + gs));
                ++n;
            }
            if (started) umma_commit(smem_u32(accum_bar));
            tmem_slot[1] = started;
        }
        __syncwarp();
    }
    __syncthreads();
    const uint32_t started = *reinterpret_cast<volatile uint32_t *>(tmem_slot + 1);
    if (warp < 4) {
        if (started) {
            mbar_wait(smem_u32(accum_bar), 0);
            tc_fence_after();
        }
        float *dst = p.partial + (size_t)blockIdx.x * p.K * (size_t)p.cin * p.cout;
        const int ci = tid;                                              // TMEM lane = input channel (m_blocks == 1)
        const uint32_t lane_base = tmem_d + ((uint32_t)(warp * 32) << 16);
        for (int k = 0; k < p.K; ++k) {
            uint32_t v[16];
            if ((started >> k) & 1u) tmem_ld16(lane_base + (uint32_t)(k * 16), v);      // warp-uniform
            else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = 0u;
            }
            if (ci < p.cin) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (j < p.cout) dst[((size_t)k * p.cin + ci) * p.cout + j] = __uint_as_float(v[j]);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
    }
    (void)m_blocks;
}

// dw[i] = sum_s partial[s][i], s ascending
__global__ void k_bwd_reduce_partials(const float *__restrict__ partial, int slices, long n, float *__restrict__ dw) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long stride = (long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        float acc = 0.f;
        for (int s = 0; s < slices; ++s) acc += __ldg(partial + (long)s * n + i);
        dw[i] = acc;
    }
}

static bool bwt_k_inner(int K, int cin, int cout, long tiles) {      // narrow outputs on many tiles: k_conv_bwd_weight_tc_kin
    const char *e = getenv("PCG_BWT_KIN_MIN_TILES");          // tests force the variant on small inputs
    return cout <= 16 && K <= 32 && cin <= 128 && tiles >= (e ? atol(e) : 4L * PCG_SMS);
}

static long bwt_slices(int K, int cin, int cout, long tiles) {
    if (bwt_k_inner(K, cin, cout, tiles)) return PCG_SMS;      // one CTA per SM, each a strided slice of the tiles
    // One CTA per SM (the operand ring takes most of the shared memory) and no dynamic scheduling: K * slices CTAs run in waves of
    // 148.  Base rule: about two waves.  Large outputs (>= 1024 tiles per slice: a CTA must keep enough ACTIVE tiles to amortise
    // its pipeline fill and its 64 KB partial block) may take more slices, and then the count that fills the last wave best is
    // chosen (K = 125 with 2 slices was 250 CTAs = 1.69 waves: ncu showed the second wave 31 % empty; 13 slices = 10.98 waves:
    // 12.4 -> 10.0 ms on the 125x128x32 layer of the training batch); partial blocks stay below 256 MiB.
    long lo = (2L * PCG_SMS) / K;
    if (lo > tiles) lo = tiles;
    if (lo > 64) lo = 64;
    if (lo < 1) lo = 1;
    long hi = tiles / 1024 < 64 ? tiles / 1024 : 64;
    const long by_mem = (256L << 20) / ((long)K * cin * cout * 4);
    if (hi > by_mem) hi = by_mem;
    if (hi <= lo) return lo;
    long best = lo;
    double best_eff = 0.0;
    for (long c = lo; c <= hi; ++c) {
        const long ctas = (long)K * c;
        const double eff = (double)ctas / (double)(cdiv(ctas, PCG_SMS) * PCG_SMS);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = c; }
    }
    return best;
}

extern "C" size_t pcg_conv_bwd_weight_tc_scratch_bytes(int K, int cin, int cout, long n_out) {
    const long tiles = cdiv(n_out > 0 ? n_out : 1, BWT_ROWS);
    return (size_t)bwt_slices(K, cin, cout, tiles) * K * cin * cout * sizeof(float) + 256;
}

extern "C" int pcg_conv_bwd_weight_tc(const void *feats_bf16, int cin, const void *gout_bf16, int cout, const int32_t *nbr, long nbr_pitch,
                                      const uint32_t *tile_mask, int K, long n_out, void *scratch, size_t scratch_bytes, float *dweight,
                                      void *stream) {
    PCG_CHECK_ARG(K >= 1 && K <= 128, "pcg_conv_bwd_weight_tc: K=%d must be in [1,128] (tile mask)", K);
    PCG_CHECK_ARG(cin >= 1 && cin <= 256 && cout >= 1 && cout <= 256, "pcg_conv_bwd_weight_tc: cin=%d (1..256) cout=%d (1..256)", cin, cout);
    PCG_CHECK_ARG(tile_mask != nullptr && (nbr_pitch & 127) == 0 && nbr_pitch >= n_out, "pcg_conv_bwd_weight_tc: tile_mask / padded neighbour table required");
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)gout_bf16 & 15) == 0, "pcg_conv_bwd_weight_tc: operands must be 16-byte aligned");
    cudaStream_t st = as_stream(stream);
    const long n_w = (long)K * cin * cout;
    if (n_out <= 0) {
        PCG_CUDA(cudaMemsetAsync(dweight, 0, (size_t)n_w * sizeof(float), st));
        return 0;
    }
    BwtParams p;
    p.feats = reinterpret_cast<const uint8_t *>(feats_bf16); p.gout = reinterpret_cast<const uint8_t *>(gout_bf16);
    p.nbr = nbr; p.tile_mask = tile_mask; p.pitch = nbr_pitch; p.n_out = n_out;
    p.tiles = cdiv(n_out, BWT_ROWS);
    const long slices = bwt_slices(K, cin, cout, p.tiles);
    p.cin = cin; p.cout = cout; p.K = K;
    p.ci_chunks = (cin + 63) / 64; p.ci_chunks_p = (p.ci_chunks + 1) & ~1; p.co_chunks = (cout + 63) / 64;
    p.n_mma = (cout + 15) & ~15;
    const int cols = (p.ci_chunks_p / 2) * p.n_mma;
    PCG_CHECK_ARG(cols <= 512, "pcg_conv_bwd_weight_tc: cin=%d x cout=%d needs %d TMEM columns (> 512)", cin, cout, cols);
    p.tmem_cols = 32;
    while (p.tmem_cols < cols) p.tmem_cols <<= 1;
    const size_t stage_bytes = (size_t)(p.ci_chunks_p + p.co_chunks) * BWT_CHUNK_BYTES;
    int stages = (int)((208 * 1024) / stage_bytes);
    PCG_CHECK_ARG(stages >= 1, "pcg_conv_bwd_weight_tc: one stage of %zu bytes does not fit in shared memory", stage_bytes);
    if (stages > BWT_MAX_STAGES) stages = BWT_MAX_STAGES;
    p.stages = stages;
    PCG_CHECK_ARG(scratch && scratch_bytes >= (size_t)slices * n_w * sizeof(float), "pcg_conv_bwd_weight_tc: scratch too small");
    p.partial = reinterpret_cast<float *>(scratch);
    if (bwt_k_inner(K, cin, cout, p.tiles)) {
        const size_t smem_k = 1024 + (size_t)BWK_STAGES * p.ci_chunks_p * BWT_CHUNK_BYTES + 2 * BWT_CHUNK_BYTES + 256;
        PCG_CUDA(cudaFuncSetAttribute(k_conv_bwd_weight_tc_kin, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_k));
        k_conv_bwd_weight_tc_kin<<<(unsigned)slices, 160, smem_k, st>>>(p);
        PCG_LAUNCHED();
        long blocks_k = cdiv(n_w, 256);
        if (blocks_k > PCG_SMS * 8) blocks_k = PCG_SMS * 8;
        k_bwd_reduce_partials<<<(unsigned)blocks_k, 256, 0, st>>>(p.partial, (int)slices, n_w, dweight);
        PCG_LAUNCHED();
        return 0;
    }
    const size_t smem = 1024 + (size_t)stages * stage_bytes + 256;
    PCG_CUDA(cudaFuncSetAttribute(k_conv_bwd_weight_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_bwd_weight_tc<<<dim3((unsigned)K, (unsigned)slices), 160, smem, st>>>(p);
    PCG_LAUNCHED();
    long blocks = cdiv(n_w, 256);
    if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
    k_bwd_reduce_partials<<<(unsigned)blocks, 256, 0, st>>>(p.partial, (int)slices, n_w, dweight);
    PCG_LAUNCHED();
    return 0;
}
