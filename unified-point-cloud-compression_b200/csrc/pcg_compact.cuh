// Order-preserving stream compaction: single pass with decoupled look-back (default), or flag+count -> scan of block
// counts -> scatter (PCG_COMPACT=3pass).
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

// Single pass (default): one launch per compaction.  A tile (= chunk of PCG_CHUNK items, taken in launch order through an
// atomic ticket so that every lower-numbered tile is already running) evaluates its flags, publishes its count, obtains the
// count of everything before it by a decoupled look-back over the tile states (warp 0 inspects 32 predecessors at a time;
// state word = status << 32 | value, one 64-bit store, so value and status become visible together) and emits.  Items are
// blocked per thread (thread t owns items 4t..4t+3 of the tile), so rank order = input order: same output as the three-pass
// form, without the flag array round trip and with one launch instead of three.
#define PCG_TILE_AGGREGATE 1ull
#define PCG_TILE_INCLUSIVE 2ull

template <class FlagFn, class EmitFn>
__global__ void __launch_bounds__(PCG_CHUNK_THREADS) pcg_k_compact_single(long n, FlagFn flag, EmitFn emit, unsigned long long *tile_state,
                                                                         int32_t *ticket, int32_t *d_n_out, int n_tiles) {
    __shared__ int s_tile, s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1);
    __syncthreads();
    const int tile = s_tile;
    const long first = (long)tile * PCG_CHUNK + (long)threadIdx.x * PCG_CHUNK_ITEMS;
    bool f[PCG_CHUNK_ITEMS];
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < PCG_CHUNK_ITEMS; ++j) {
        f[j] = first + j < n && flag(first + j);
        cnt += f[j] ? 1 : 0;
    }
    int total;
    const int excl = pcg_block_excl_scan(cnt, total);
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        volatile unsigned long long *state = tile_state;
        if (lane == 0 && tile > 0) state[tile] = (PCG_TILE_AGGREGATE << 32) | (unsigned)total;
        int prefix = 0;
        int t = tile - 1;                                  // nearest predecessor of the current window
        bool done = tile == 0;
        while (!done) {
            const int idx = t - lane;
            unsigned long long w = PCG_TILE_INCLUSIVE << 32;    // before tile 0: inclusive prefix 0
            if (idx >= 0) {
                do { w = state[idx]; } while ((w >> 32) == 0ull);
            }
            const unsigned incl = __ballot_sync(0xffffffffu, (w >> 32) == PCG_TILE_INCLUSIVE);
            const int nearest = incl ? __ffs(incl) - 1 : 31;    // lanes 0..nearest contribute
            int v = lane <= nearest ? (int)(unsigned)w : 0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            prefix += v;
            if (incl) done = true;
            else t -= 32;
        }
        if (lane == 0) {
            state[tile] = (PCG_TILE_INCLUSIVE << 32) | (unsigned)(prefix + total);
            s_prefix = prefix;
            if (tile == n_tiles - 1) *d_n_out = prefix + total;
        }
    }
    __syncthreads();
    int rank = s_prefix + excl;
#pragma unroll
    for (int j = 0; j < PCG_CHUNK_ITEMS; ++j)
        if (f[j]) emit(first + j, rank++);
}

template <class FlagFn, class EmitFn>
static int pcg_compact(long n, FlagFn flag, EmitFn emit, int32_t *d_n_out, const PcgScratch &s, cudaStream_t st) {
    if (n <= 0) {
        PCG_CUDA(cudaMemsetAsync(d_n_out, 0, sizeof(int32_t), st));
        return 0;
    }
    static const bool three_pass = getenv("PCG_COMPACT") && !strcmp(getenv("PCG_COMPACT"), "3pass");     // A/B knob
    if (!three_pass && (size_t)s.n_blocks * sizeof(unsigned long long) <= (((size_t)n + 255) & ~(size_t)255) && s.n_blocks < (1L << 30)) {
        // tile states live where the three-pass form keeps its flag bytes (8 bytes per 1024 items), the ticket in block_counts[0]
        unsigned long long *state = reinterpret_cast<unsigned long long *>(s.flags);
        PCG_CUDA(cudaMemsetAsync(state, 0, (size_t)s.n_blocks * sizeof(unsigned long long), st));
        PCG_CUDA(cudaMemsetAsync(s.block_counts, 0, sizeof(int32_t), st));
        pcg_k_compact_single<<<(unsigned)s.n_blocks, PCG_CHUNK_THREADS, 0, st>>>(n, flag, emit, state, s.block_counts, d_n_out, (int)s.n_blocks);
        PCG_LAUNCHED();
        return 0;
    }
    pcg_k_flag_count<<<(unsigned)s.n_blocks, PCG_CHUNK_THREADS, 0, st>>>(n, flag, s.flags, s.block_counts);
    PCG_LAUNCHED();
    if (int rc = pcg_scan_block_counts(s.block_counts, s.n_blocks, d_n_out, st)) return rc;
    pcg_k_scatter<<<(unsigned)s.n_blocks, PCG_CHUNK_THREADS, 0, st>>>(n, s.flags, s.block_counts, emit);
    PCG_LAUNCHED();
    return 0;
}
