// Shared device/host helpers for libpcg.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/pcg.h"

#define PCG_SMS 148  // B200: size grids as a multiple of this where a kernel is persistent

extern thread_local char pcg_err_buf[512];
extern std::atomic<long> pcg_launches;

int pcg_fail(const char *fmt, ...);

#define PCG_CHECK_ARG(cond, ...)            \
    do {                                    \
        if (!(cond)) return pcg_fail(__VA_ARGS__); \
    } while (0)

#define PCG_CUDA(call)                                                                    \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) return pcg_fail("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    } while (0)

// count + check a kernel launch
#define PCG_LAUNCHED()                                                                   \
    do {                                                                                 \
        pcg_launches.fetch_add(1, std::memory_order_relaxed);                            \
        cudaError_t e_ = cudaGetLastError();                                             \
        if (e_ != cudaSuccess) return pcg_fail("%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
    } while (0)

static inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }
static inline long cdiv(long a, long b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------- coordinate keys
#define PCG_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull

__host__ __device__ __forceinline__ uint64_t pcg_pack(int b, int x, int y, int z) {
    return ((uint64_t)(uint16_t)b << 48) | ((uint64_t)(uint16_t)(x + 32768) << 32) |
           ((uint64_t)(uint16_t)(y + 32768) << 16) | (uint64_t)(uint16_t)(z + 32768);
}
__host__ __device__ __forceinline__ bool pcg_in_range(int b, int x, int y, int z) {
    return b >= 0 && b < 32768 && x >= -32768 && x < 32768 && y >= -32768 && y < 32768 && z >= -32768 && z < 32768;
}
__device__ __forceinline__ uint64_t pcg_mix(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

// Lookup in an open-addressing table (linear probing).  Returns value or -1.
__device__ __forceinline__ int pcg_table_find(const uint64_t *__restrict__ keys, const int32_t *__restrict__ vals,
                                              uint64_t mask, uint64_t key) {
    uint64_t s = pcg_mix(key) & mask;
#pragma unroll 1
    for (uint64_t probes = 0; probes <= mask; ++probes) {
        uint64_t k = __ldg(keys + s);
        if (k == key) return __ldg(vals + s);
        if (k == PCG_EMPTY_KEY) return -1;
        s = (s + 1) & mask;
    }
    return -1;
}

// Insert key, keep the minimum value.  Returns slot, or -1 when the table is full.
__device__ __forceinline__ long pcg_table_insert_min(uint64_t *keys, int32_t *vals, uint64_t mask, uint64_t key, int val) {
    uint64_t s = pcg_mix(key) & mask;
#pragma unroll 1
    for (uint64_t probes = 0; probes <= mask; ++probes) {
        unsigned long long prev = atomicCAS((unsigned long long *)(keys + s), (unsigned long long)PCG_EMPTY_KEY,
                                            (unsigned long long)key);
        if (prev == PCG_EMPTY_KEY || prev == key) {
            atomicMin(vals + s, val);
            return (long)s;
        }
        s = (s + 1) & mask;
    }
    return -1;
}

__device__ __forceinline__ int4 pcg_ld_coord(const int32_t *coords, long i) {
    return __ldg(reinterpret_cast<const int4 *>(coords) + i);
}

__device__ __forceinline__ void pcg_kernel_offset(int i, int ksize, int &dx, int &dy, int &dz) {
    int lo = (ksize & 1) ? -(ksize >> 1) : 0;
    dx = lo + i % ksize;
    dy = lo + (i / ksize) % ksize;
    dz = lo + i / (ksize * ksize);
}

__device__ __forceinline__ int pcg_floor_to(int a, int step) {
    int q = a / step;
    if ((a % step != 0) && ((a < 0) != (step < 0))) --q;
    return q * step;
}

__device__ __forceinline__ float pcg_apply_act(float v, int act) {
    if (act == PCG_ACT_RELU) return v > 0.f ? v : 0.f;
    if (act == PCG_ACT_LEAKY_RELU) return v > 0.f ? v : 0.01f * v;
    return v;
}

// ---------------------------------------------------------------------------- compaction helper
// Exclusive scan of per-block counts; implemented in pcg_scan.cu.
// block_counts[n_blocks] -> block_offsets[n_blocks], total -> d_total[0].
int pcg_scan_block_counts(int32_t *block_counts, long n_blocks, int32_t *d_total, cudaStream_t st);
