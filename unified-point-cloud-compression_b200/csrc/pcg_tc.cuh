// PTX helpers shared by the tcgen05 kernels (sm_100a): mbarrier, bulk / tensor TMA copies, UMMA descriptors,
// tcgen05.mma / commit / ld.
#pragma once
#include <cuda.h>

#include "pcg_common.cuh"

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
// same copy allocating in L1 (.ca): for gathers whose rows are re-read by the neighbouring kernel offsets of the same tile
__device__ __forceinline__ void cp_async16_ca(uint32_t dst, const void *src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
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

// K-major swizzled shared-memory matrix descriptor for rows of RB bytes (8-row groups 8*RB bytes apart)
template <int RB>
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    constexpr uint64_t layout = RB == 128 ? 2 : (RB == 64 ? 4 : 6);   // SWIZZLE_128B / 64B / 32B
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);  // start address
    d |= (uint64_t)1 << 16;                   // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)((8 * RB) >> 4) << 32;     // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                   // descriptor version (Blackwell)
    d |= layout << 61;
    return d;
}

// four feature rows (row indices r.x..r.w, column `col`) -> four consecutive operand rows at `dst`
__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap *tm, int col, int4 r, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
        ::"r"(dst), "l"(tm), "r"(col), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w), "r"(bar)
        : "memory");
}


// plain 2-D TMA tile load: box (inner c0.., outer c1..) of the tensor map -> shared memory, completes on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}

// 2-D TMA tile store: shared memory box -> global (bulk async group of the issuing thread)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *tm, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tm), "r"(c0), "r"(c1), "r"(src)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// LSU gather of one 128-row x 128-byte operand sub-tile into its 128B-swizzled image at dst_tile.  Lane m holds the neighbour
// indices of rows m, 32+m, 64+m, 96+m in idx[0..3]; a warp instruction copies rows 4j..4j+3 (8 lanes x 16 B each).  A producer
// warp is alone on its scheduler, so every instruction costs its full latency: the copy is kept to shuffle + compare + one
// 32x32->64 multiply-add + cp.async (absent rows: src-size 0 zero-fills and the source address is never dereferenced; the
// destination offsets are two per-lane constants + immediates since (row & 7) = 4 (u & 1) + q).
__device__ __forceinline__ int tc_shfl_idx(int v, int src_lane) {      // volatile: keeps the batch order written below
    int r;
    asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(r) : "r"(v), "r"(src_lane));
    return r;
}
template <int RB, bool CA>
__device__ __forceinline__ void tc_lsu_gather_tile(uint32_t dst_tile, const uint8_t *fcol, int row_bytes, const int (&idx)[4], int lane) {
    // rows of RB bytes: LPR lanes per row, RPI rows per warp instruction.  fcol already points at this lane's 16-byte chunk of
    // the operand columns.  Swizzle of the destination chunk: RB 128: c ^ (r & 7), 64: c ^ ((r >> 1) & 3), 32: c ^ ((r >> 2) & 1);
    // with r = j*RPI + q the term depends on j only for RB = 128 (two per-lane constants), not at all for 64 / 32.
    constexpr int LPR = RB / 16, RPI = 32 / LPR, NI = 128 / RPI;
    const int q = lane / LPR, c16 = lane % LPR;
    uint32_t d0, d1;
    if (RB == 128) {
        d0 = dst_tile + (uint32_t)q * 128u + (uint32_t)((c16 ^ q) << 4);
        d1 = dst_tile + (uint32_t)q * 128u + (uint32_t)((c16 ^ (4 + q)) << 4);
    } else {
        const int x = RB == 64 ? ((q >> 1) & 3) : ((q >> 2) & 1);
        d0 = d1 = dst_tile + (uint32_t)q * RB + (uint32_t)((c16 ^ x) << 4);
    }
    asm volatile("" : "+l"(fcol));        // keep the column base one 64-bit addend of the multiply-add below
#pragma unroll
    for (int jb = 0; jb < NI / 8; ++jb) {
        int src[8];                       // eight shuffles in flight, then eight copies: the latencies overlap instead of adding up
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int r0 = (8 * jb + u) * RPI;                      // first row of this instruction; row = r0 + q
            src[u] = tc_shfl_idx(idx[r0 >> 5], (r0 & 31) + q);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int r0 = (8 * jb + u) * RPI;
            const uint32_t dst = ((RB == 128 && (u & 1)) ? d1 : d0) + (uint32_t)(r0 * RB);
            const uint8_t *from = fcol + (unsigned long long)(unsigned)src[u] * (unsigned)row_bytes;
            if (CA) cp_async16_ca(dst, from, src[u] >= 0 ? 16u : 0u);
            else cp_async16(dst, from, src[u] >= 0 ? 16u : 0u);
        }
    }
}

typedef CUresult (*PcgEncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PcgEncodeTiled pcg_tensor_map_encoder();
