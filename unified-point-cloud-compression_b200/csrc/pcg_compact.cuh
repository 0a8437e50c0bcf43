// Order-preserving stream compaction: flag+count -> scan of block counts -> scatter.
// Items are processed in fixed chunks of PCG_CHUNK so the output order is the input order
// (deterministic; no atomics on the output cursor).
#pragma once
#include "pcg_common.cuh"

#define PCG_CHUNK_THREADS 256
#define PCG_CHUNK_ITEMS 4
#define PCG_CHUNK (PCG_CHUNK_THREADS * PCG_CHUNK_ITEMS)

struct PcgScratch {
    uint8_t *flags;         // n items
    int32_t *block_counts;  // n_blocks + 1
    long n_blocks;
};

static inline size_t pcg_compact_scratch_bytes(long n) {
    long nb = cdiv(n > 0 ? n : 1, PCG_CHUNK);
    size_t flags = ((size_t)(n > 0 ? n : 1) + 255) & ~(size_t)255;
    return flags + (size_t)(nb + 2) * sizeof(int32_t) + 256;
}

static inline bool pcg_carve_scratch(void *scratch, size_t bytes, long n, PcgScratch &s) {
    if (bytes < pcg_compact_scratch_bytes(n) || scratch == nullptr) return false;
    size_t flags = ((size_t)(n > 0 ? n : 1) + 255) & ~(size_t)255;
    s.flags = reinterpret_cast<uint8_t *>(scratch);
    s.block_counts = reinterpret_cast<int32_t *>(s.flags + flags);
    s.n_blocks = cdiv(n > 0 ? n : 1, PCG_CHUNK);
    return true;
}

// Block-wide exclusive scan of one int per thread (256 threads); returns total in `total`.
__device__ __forceinline__ int pcg_block_excl_scan(int v, int &total) {
    __shared__ int warp_sums[PCG_CHUNK_THREADS / 32];
    __shared__ int block_total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = lane < PCG_CHUNK_THREADS / 32 ? warp_sums[lane] : 0;
        int wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += t;
        }
        if (lane < PCG_CHUNK_THREADS / 32) warp_sums[lane] = wi - w;
        if (lane == PCG_CHUNK_THREADS / 32 - 1) block_total = wi;
    }
    __syncthreads();
    int excl = incl - v + warp_sums[warp];
    total = block_total;
    __syncthreads();
    return excl;
}

// Phase 1: flags[i] = fn(i); block_counts[b] = number of set flags in chunk b.
template <class FlagFn>
__global__ void __launch_bounds__(PCG_CHUNK_THREADS) pcg_k_flag_count(long n, FlagFn fn, uint8_t *flags, int32_t *block_counts) {
    long base = (long)blockIdx.x * PCG_CHUNK;
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < PCG_CHUNK_ITEMS; ++j) {
        long i = base + (long)j * PCG_CHUNK_THREADS + threadIdx.x;
        if (i < n) {
            uint8_t f = fn(i) ? 1 : 0;
            flags[i] = f;
            cnt += f;
        }
    }
    int total;
    pcg_block_excl_scan(cnt, total);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

// Phase 3: emit(i, rank) for every flagged i, rank = global output position (input order).
template <class EmitFn>
__global__ void __launch_bounds__(PCG_CHUNK_THREADS) pcg_k_scatter(long n, const uint8_t *flags, const int32_t *block_offsets, EmitFn emit) {
    long base = (long)blockIdx.x * PCG_CHUNK;
    // order inside a chunk: item index = base + j*THREADS + tid  => scan per j-slab, slabs in order
    int run = block_offsets[blockIdx.x];
#pragma unroll
    for (int j = 0; j < PCG_CHUNK_ITEMS; ++j) {
        long i = base + (long)j * PCG_CHUNK_THREADS + threadIdx.x;
        int f = (i < n) ? flags[i] : 0;
        int total;
        int excl = pcg_block_excl_scan(f, total);
        if (f) emit(i, run + excl);
        run += total;
    }
}

template <class FlagFn, class EmitFn>
static int pcg_compact(long n, FlagFn flag, EmitFn emit, int32_t *d_n_out, const PcgScratch &s, cudaStream_t st) {
    if (n <= 0) {
        PCG_CUDA(cudaMemsetAsync(d_n_out, 0, sizeof(int32_t), st));
        return 0;
    }
    pcg_k_flag_count<<<(unsigned)s.n_blocks, PCG_CHUNK_THREADS, 0, st>>>(n, flag, s.flags, s.block_counts);
    PCG_LAUNCHED();
    if (int rc = pcg_scan_block_counts(s.block_counts, s.n_blocks, d_n_out, st)) return rc;
    pcg_k_scatter<<<(unsigned)s.n_blocks, PCG_CHUNK_THREADS, 0, st>>>(n, s.flags, s.block_counts, emit);
    PCG_LAUNCHED();
    return 0;
}
