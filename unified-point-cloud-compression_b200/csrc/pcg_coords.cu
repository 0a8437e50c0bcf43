// Coordinate hash, strided / generative coordinate sets, kernel maps, pruning, lexsort.
// HBM-bound integer work: int4 coordinate loads, 64-bit packed keys, open addressing.
#include <cub/device/device_radix_sort.cuh>

#include "pcg_compact.cuh"

// ---------------------------------------------------------------------------- scan of block counts
__global__ void __launch_bounds__(1024) k_scan_block_counts(int32_t *counts, long n, int32_t *d_total) {
    __shared__ int warp_sums[32];
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long base = 0; base < n; base += 1024) {
        long i = base + threadIdx.x;
        int v = i < n ? counts[i] : 0;
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = warp_sums[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, wi, d);
                if (lane >= d) wi += t;
            }
            warp_sums[lane] = wi - w;
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + warp_sums[warp] + incl - v;
        if (i < n) counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) d_total[0] = carry_s;
}

int pcg_scan_block_counts(int32_t *block_counts, long n_blocks, int32_t *d_total, cudaStream_t st) {
    k_scan_block_counts<<<1, 1024, 0, st>>>(block_counts, n_blocks, d_total);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- hash build / lookup
__global__ void k_table_clear(uint64_t *keys, int32_t *vals, long cap) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < cap; i += stride) {
        keys[i] = PCG_EMPTY_KEY;
        vals[i] = 0x7fffffff;
    }
}

// A coordinate outside the 16-bit-per-axis / 15-bit-batch packing range would wrap onto another voxel's key: it is not inserted
// and bit 2 of *d_status is raised instead (the host side turns that into an error; MinkowskiEngine itself takes any int32).
__global__ void k_hash_insert(const int32_t *__restrict__ coords, long n, uint64_t *keys, int32_t *vals, uint64_t mask, int32_t *d_status) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int4 c = pcg_ld_coord(coords, i);
    if (!pcg_in_range(c.x, c.y, c.z, c.w)) { if (d_status) atomicOr(d_status, 2); return; }
    pcg_table_insert_min(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w), (int)i);
}

__global__ void k_hash_lookup(const uint64_t *__restrict__ keys, const int32_t *__restrict__ vals, uint64_t mask,
                              const int32_t *__restrict__ query, long nq, int32_t *rows) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    int4 c = pcg_ld_coord(query, i);
    rows[i] = pcg_in_range(c.x, c.y, c.z, c.w) ? pcg_table_find(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w)) : -1;
}

static int clear_table(uint64_t *keys, int32_t *vals, long cap, cudaStream_t st) {
    long blocks = cdiv(cap, 256 * 8);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_table_clear<<<(unsigned)blocks, 256, 0, st>>>(keys, vals, cap);
    PCG_LAUNCHED();
    return 0;
}

static bool is_pow2(long v) { return v > 0 && (v & (v - 1)) == 0; }

extern "C" long pcg_hash_capacity(long n_keys) {
    long cap = 1024;
    while (cap < 2 * n_keys) cap <<= 1;
    return cap;
}

extern "C" long pcg_nbr_pitch(long n_out) { return n_out <= 0 ? 128 : (n_out + 127) / 128 * 128; }

extern "C" size_t pcg_scratch_bytes(long n_items) { return pcg_compact_scratch_bytes(n_items) + 8192; }

extern "C" int pcg_hash_build(const int32_t *coords, long n, uint64_t *keys, int32_t *vals, long cap, int32_t *d_status, void *stream) {
    PCG_CHECK_ARG(is_pow2(cap) && cap >= n + 1, "pcg_hash_build: cap=%ld must be a power of two > n=%ld", cap, n);
    cudaStream_t st = as_stream(stream);
    if (int rc = clear_table(keys, vals, cap, st)) return rc;
    if (n > 0) {
        k_hash_insert<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(coords, n, keys, vals, (uint64_t)cap - 1, d_status);
        PCG_LAUNCHED();
    }
    return 0;
}

extern "C" int pcg_hash_lookup(const uint64_t *keys, const int32_t *vals, long cap, const int32_t *query, long nq,
                               int32_t *rows_out, void *stream) {
    PCG_CHECK_ARG(is_pow2(cap), "pcg_hash_lookup: cap=%ld must be a power of two", cap);
    if (nq <= 0) return 0;
    k_hash_lookup<<<(unsigned)cdiv(nq, 256), 256, 0, as_stream(stream)>>>(keys, vals, (uint64_t)cap - 1, query, nq, rows_out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- unique / stride
struct StridedKey {  // key of floor(c/step)*step for row i
    const int32_t *coords;
    int step;
    __device__ __forceinline__ int4 coord(long i) const {
        int4 c = pcg_ld_coord(coords, i);
        if (step > 1) { c.y = pcg_floor_to(c.y, step); c.z = pcg_floor_to(c.z, step); c.w = pcg_floor_to(c.w, step); }
        return c;
    }
};

__global__ void k_insert_strided(StridedKey sk, long n, uint64_t *keys, int32_t *vals, uint64_t mask, int32_t *d_status) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int4 c = sk.coord(i);
    if (!pcg_in_range(c.x, c.y, c.z, c.w)) { if (d_status) atomicOr(d_status, 2); return; }       // see k_hash_insert
    pcg_table_insert_min(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w), (int)i);
}

struct FlagFirstStrided {
    StridedKey sk;
    const uint64_t *keys;
    const int32_t *vals;
    uint64_t mask;
    __device__ bool operator()(long i) const {
        int4 c = sk.coord(i);
        if (!pcg_in_range(c.x, c.y, c.z, c.w)) return false;
        return pcg_table_find(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w)) == (int)i;
    }
};

struct EmitStrided {
    StridedKey sk;
    int32_t *out_coords;
    int32_t *out_index;  // may be null
    __device__ void operator()(long i, int rank) const {
        int4 c = sk.coord(i);
        reinterpret_cast<int4 *>(out_coords)[rank] = c;
        if (out_index) out_index[rank] = (int)i;
    }
};

// after compaction: table value (first source row) -> new row id
__global__ void k_remap_table(const int32_t *__restrict__ out_coords, const int32_t *d_n_out, long max_out,
                              const uint64_t *keys, int32_t *vals, uint64_t mask) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long n = d_n_out[0];
    if (n > max_out) n = max_out;
    long stride = (long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int4 c = pcg_ld_coord(out_coords, i);
        uint64_t key = pcg_pack(c.x, c.y, c.z, c.w);
        uint64_t s = pcg_mix(key) & mask;
        for (uint64_t probes = 0; probes <= mask && keys[s] != key; ++probes) s = (s + 1) & mask;
        if (keys[s] == key) vals[s] = (int)i;
    }
}

static int unique_strided(const int32_t *coords, long n, int step, uint64_t *keys, int32_t *vals, long cap,
                          int32_t *out_coords, int32_t *out_index, int32_t *d_n_out, int32_t *d_status, void *scratch,
                          size_t scratch_bytes, cudaStream_t st) {
    PCG_CHECK_ARG(is_pow2(cap) && cap >= n + 1, "unique/stride: cap=%ld must be a power of two > n=%ld", cap, n);
    PCG_CHECK_ARG(step >= 1, "unique/stride: step=%d", step);
    PcgScratch s;
    PCG_CHECK_ARG(pcg_carve_scratch(scratch, scratch_bytes, n, s), "unique/stride: scratch too small (%zu < %zu)",
                  scratch_bytes, pcg_compact_scratch_bytes(n));
    if (d_status) PCG_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    if (int rc = clear_table(keys, vals, cap, st)) return rc;
    if (n <= 0) {
        PCG_CUDA(cudaMemsetAsync(d_n_out, 0, sizeof(int32_t), st));
        return 0;
    }
    StridedKey sk{coords, step};
    uint64_t mask = (uint64_t)cap - 1;
    k_insert_strided<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(sk, n, keys, vals, mask, d_status);
    PCG_LAUNCHED();
    if (int rc = pcg_compact(n, FlagFirstStrided{sk, keys, vals, mask}, EmitStrided{sk, out_coords, out_index}, d_n_out, s, st))
        return rc;
    long blocks = cdiv(n, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    k_remap_table<<<(unsigned)blocks, 256, 0, st>>>(out_coords, d_n_out, n, keys, vals, mask);
    PCG_LAUNCHED();
    return 0;
}

extern "C" int pcg_unique_first(const int32_t *coords, long n, uint64_t *keys, int32_t *vals, long cap,
                                int32_t *out_coords, int32_t *out_index, int32_t *d_n_out, int32_t *d_status, void *scratch,
                                size_t scratch_bytes, void *stream) {
    return unique_strided(coords, n, 1, keys, vals, cap, out_coords, out_index, d_n_out, d_status, scratch, scratch_bytes,
                          as_stream(stream));
}

extern "C" int pcg_stride_coords(const int32_t *coords, long n, int step, uint64_t *keys, int32_t *vals, long cap,
                                 int32_t *out_coords, int32_t *d_n_out, int32_t *d_status, void *scratch, size_t scratch_bytes,
                                 void *stream) {
    return unique_strided(coords, n, step, keys, vals, cap, out_coords, nullptr, d_n_out, d_status, scratch, scratch_bytes,
                          as_stream(stream));
}

// ---------------------------------------------------------------------------- generative coords
// Candidate enumeration of the generative up-sampling.  Offsets are grouped into classes by their residue
// modulo the up-sampling stride (for stride 2: the parity class of the child voxel), classes in ascending
// order, and inside a class candidates run input-row major / offset minor.  Output rows therefore come out
// grouped by class, so a 128-row tile of the following convolution only ever touches the (k/stride)^3-ish
// offsets its class can reach instead of all k^3 -- and rows of one tile stem from neighbouring parents.
struct GenKey {
    const int32_t *coords;
    int ksize, K, ts_out;
    long n;
    int n_classes;
    int cls_count[65];     // offsets per class
    long cls_base[65];     // first candidate id of the class (= n * offsets in earlier classes)
    int cls_first[65];     // start of the class in `order16`
    unsigned short order16[736];   // offset indices sorted by class (stable)
    __device__ __forceinline__ int4 coord(long cand) const {
        int c = 0;
        while (c + 1 < n_classes && cand >= cls_base[c + 1]) ++c;
        long rel = cand - cls_base[c];
        long row = rel / cls_count[c];
        int t = (int)(rel - row * cls_count[c]);
        int4 v = pcg_ld_coord(coords, row);
        int dx, dy, dz;
        pcg_kernel_offset(k_of(cls_first[c] + t), ksize, dx, dy, dz);
        v.y += dx * ts_out; v.z += dy * ts_out; v.w += dz * ts_out;
        return v;
    }
    __device__ __forceinline__ int k_of(int pos) const { return order16[pos]; }
};

static void gen_key_init(GenKey &g, const int32_t *coords, long n, int ksize, int ts_out, int stride) {
    g.coords = coords; g.ksize = ksize; g.K = ksize * ksize * ksize; g.ts_out = ts_out; g.n = n;
    int s = stride < 1 ? 1 : stride;
    if (s > 4) s = 4;
    g.n_classes = s * s * s;
    const int lo = (ksize & 1) ? -(ksize >> 1) : 0;
    auto md = [s](int v) { int r = v % s; return r < 0 ? r + s : r; };
    int pos = 0;
    long base = 0;
    for (int c = 0; c < g.n_classes; ++c) {
        g.cls_first[c] = pos;
        g.cls_base[c] = base;
        int cnt = 0;
        for (int k = 0; k < g.K; ++k) {
            int dx = lo + k % ksize, dy = lo + (k / ksize) % ksize, dz = lo + k / (ksize * ksize);
            if ((md(dx) * s + md(dy)) * s + md(dz) == c) { g.order16[pos++] = (unsigned short)k; ++cnt; }
        }
        g.cls_count[c] = cnt > 0 ? cnt : 1;
        base += (long)cnt * n;
    }
    g.cls_base[g.n_classes] = base;
}

__global__ void k_insert_generated(GenKey gk, long n_cand, uint64_t *keys, int32_t *vals, uint64_t mask, int32_t *d_status) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cand) return;
    int4 c = gk.coord(i);
    if (!pcg_in_range(c.x, c.y, c.z, c.w)) { atomicOr(d_status, 2); return; }
    // candidate ids can exceed int32 only for n*K > 2^31, rejected on the host side
    if (pcg_table_insert_min(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w), (int)i) < 0) atomicOr(d_status, 1);
}

struct FlagFirstGen {
    GenKey gk;
    const uint64_t *keys;
    const int32_t *vals;
    uint64_t mask;
    __device__ bool operator()(long i) const {
        int4 c = gk.coord(i);
        if (!pcg_in_range(c.x, c.y, c.z, c.w)) return false;
        return pcg_table_find(keys, vals, mask, pcg_pack(c.x, c.y, c.z, c.w)) == (int)i;
    }
};

struct EmitGen {
    GenKey gk;
    int32_t *out_coords;
    long max_out;
    __device__ void operator()(long i, int rank) const {
        if (rank < max_out) reinterpret_cast<int4 *>(out_coords)[rank] = gk.coord(i);
    }
};

__global__ void k_flag_overflow(const int32_t *d_n_out, long max_out, int32_t *d_status) {
    if (d_n_out[0] > max_out) atomicOr(d_status, 1);
}

extern "C" int pcg_generate_coords(const int32_t *coords, long n, int ksize, int stride, int ts_out, uint64_t *keys,
                                   int32_t *vals, long cap, int32_t *out_coords, long max_out, int32_t *d_n_out,
                                   int32_t *d_status, void *scratch, size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    PCG_CHECK_ARG(ksize >= 1 && ksize <= 9 && ts_out >= 1, "pcg_generate_coords: ksize=%d ts_out=%d", ksize, ts_out);
    int K = ksize * ksize * ksize;
    long n_cand = n * (long)K;
    PCG_CHECK_ARG(n_cand < 2147483647L, "pcg_generate_coords: %ld candidates exceed int32", n_cand);
    PCG_CHECK_ARG(is_pow2(cap), "pcg_generate_coords: cap=%ld must be a power of two", cap);
    PcgScratch s;
    PCG_CHECK_ARG(pcg_carve_scratch(scratch, scratch_bytes, n_cand, s), "pcg_generate_coords: scratch too small (%zu < %zu)",
                  scratch_bytes, pcg_compact_scratch_bytes(n_cand));
    PCG_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    if (int rc = clear_table(keys, vals, cap, st)) return rc;
    if (n <= 0) {
        PCG_CUDA(cudaMemsetAsync(d_n_out, 0, sizeof(int32_t), st));
        return 0;
    }
    GenKey gk;
    gen_key_init(gk, coords, n, ksize, ts_out, stride);
    uint64_t mask = (uint64_t)cap - 1;
    k_insert_generated<<<(unsigned)cdiv(n_cand, 256), 256, 0, st>>>(gk, n_cand, keys, vals, mask, d_status);
    PCG_LAUNCHED();
    if (int rc = pcg_compact(n_cand, FlagFirstGen{gk, keys, vals, mask}, EmitGen{gk, out_coords, max_out}, d_n_out, s, st))
        return rc;
    k_flag_overflow<<<1, 1, 0, st>>>(d_n_out, max_out, d_status);
    PCG_LAUNCHED();
    long blocks = cdiv(max_out, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    if (blocks < 1) blocks = 1;
    k_remap_table<<<(unsigned)blocks, 256, 0, st>>>(out_coords, d_n_out, max_out, keys, vals, mask);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- kernel map
// One CTA = one 128-row output tile x S = blockDim.y offset slices: thread (r, s) owns row r and walks offsets s, s+S, ...
// (the coordinate is one int4 load; nbr[k][tile rows] is written as one coalesced 512 B line per offset).  Slicing the
// offsets over threadIdx.y shortens the chain of dependent table probes per thread 4x and fills the machine on the
// small strided sets (54 k / 13 k rows = 425 / 107 tiles), where one thread per row left most SMs idle.
// Small sets (at most two waves of tiles) take 8 slices: 1024 threads per CTA, 16 instead of 32 dependent probes for k = 5.
#define KM_SLICES_MAX 8
__global__ void __launch_bounds__(128 * KM_SLICES_MAX) k_kernel_map(const uint64_t *__restrict__ keys, const int32_t *__restrict__ vals,
                                                                uint64_t mask, const int32_t *__restrict__ out_coords, long n_out,
                                                                int ksize, int K, int step, int lattice, int32_t *__restrict__ nbr,
                                                                long pitch, uint32_t *__restrict__ tile_mask) {
    long o = (long)blockIdx.x * 128 + threadIdx.x;
    bool live = o < n_out;
    int4 c = live ? pcg_ld_coord(out_coords, o) : make_int4(0, 0, 0, 0);
    __shared__ uint32_t smask[4];
    if (threadIdx.y == 0 && threadIdx.x < 4) smask[threadIdx.x] = 0;
    __syncthreads();
    const int lo = (ksize & 1) ? -(ksize >> 1) : 0;
    const int k2 = ksize * ksize;
    for (int k = threadIdx.y; k < K; k += blockDim.y) {
        const int iz = k / k2, iy = (k - iz * k2) / ksize, ix = k - iz * k2 - iy * ksize;
        int row = -1;
        if (live) {
            int x = c.y + (lo + ix) * step, y = c.z + (lo + iy) * step, z = c.w + (lo + iz) * step;
            // lattice > 0: input rows only exist on multiples of `lattice` (transposed conv onto a finer
            // grid), so most offsets cannot hit for a given output parity -- no probe needed
            bool possible = lattice <= 1 || ((x % lattice) == 0 && (y % lattice) == 0 && (z % lattice) == 0);
            if (possible && pcg_in_range(c.x, x, y, z)) row = pcg_table_find(keys, vals, mask, pcg_pack(c.x, x, y, z));
        }
        nbr[(long)k * pitch + o] = row;      // rows past n_out are the -1 padding of the last tile
        if (tile_mask) {
            unsigned any = __ballot_sync(0xffffffffu, row >= 0);
            if (any && (threadIdx.x & 31) == 0) atomicOr(&smask[k >> 5], 1u << (k & 31));
        }
    }
    if (tile_mask) {
        __syncthreads();
        if (threadIdx.y == 0 && threadIdx.x < 4) tile_mask[(long)blockIdx.x * 4 + threadIdx.x] = smask[threadIdx.x];
    }
}

extern "C" int pcg_kernel_map(const uint64_t *keys, const int32_t *vals, long cap, const int32_t *out_coords, long n_out,
                              int ksize, int ts, int transposed, int in_lattice, int32_t *nbr, long nbr_pitch,
                              uint32_t *tile_mask, void *stream) {
    PCG_CHECK_ARG(is_pow2(cap), "pcg_kernel_map: cap=%ld must be a power of two", cap);
    PCG_CHECK_ARG(ksize >= 1 && ksize <= 9 && ts >= 1, "pcg_kernel_map: ksize=%d ts=%d", ksize, ts);
    int K = ksize * ksize * ksize;
    PCG_CHECK_ARG(tile_mask == nullptr || K <= 128, "pcg_kernel_map: tile_mask needs ksize^3 <= 128 (got %d)", K);
    PCG_CHECK_ARG(nbr_pitch == pcg_nbr_pitch(n_out), "pcg_kernel_map: nbr_pitch=%ld must be pcg_nbr_pitch(n_out)=%ld", nbr_pitch,
                  pcg_nbr_pitch(n_out));
    if (n_out <= 0) return 0;
    int step = transposed ? -ts : ts;
    const int slices = (cdiv(n_out, 128) <= 2 * PCG_SMS && K >= 64) ? KM_SLICES_MAX : 4;
    k_kernel_map<<<(unsigned)cdiv(n_out, 128), dim3(128, slices), 0, as_stream(stream)>>>(keys, vals, (uint64_t)cap - 1, out_coords, n_out,
                                                                           ksize, K, step, in_lattice, nbr, nbr_pitch, tile_mask);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- prune rows
struct FlagMask {
    const uint8_t *mask;
    __device__ bool operator()(long i) const { return mask[i] != 0; }
};
struct EmitCoordIndex {
    const int32_t *coords;
    int32_t *out_coords;
    int32_t *rank_of;  // rank_of[rank] = source row
    __device__ void operator()(long i, int rank) const {
        if (coords) reinterpret_cast<int4 *>(out_coords)[rank] = pcg_ld_coord(coords, i);
        rank_of[rank] = (int)i;
    }
};

// copy kept feature rows: one warp per output row, float4 lanes when C % 4 == 0
__global__ void k_gather_rows(const float *__restrict__ feats, int c, const int32_t *__restrict__ src_rows,
                              const int32_t *d_n, float *__restrict__ out) {
    long n = d_n[0];
    long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    long nwarps = ((long)gridDim.x * blockDim.x) >> 5;
    for (long r = warp; r < n; r += nwarps) {
        const float *src = feats + (long)src_rows[r] * c;
        float *dst = out + r * c;
        if ((c & 3) == 0) {
            for (int j = lane; j < (c >> 2); j += 32)
                reinterpret_cast<float4 *>(dst)[j] = __ldg(reinterpret_cast<const float4 *>(src) + j);
        } else {
            for (int j = lane; j < c; j += 32) dst[j] = __ldg(src + j);
        }
    }
}

extern "C" int pcg_prune_rows(const int32_t *coords, const float *feats, long n, int c, const uint8_t *mask,
                              int32_t *out_coords, float *out_feats, int32_t *d_n_out, void *scratch,
                              size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    // scratch: compaction scratch + n int32 source-row list
    size_t need = pcg_compact_scratch_bytes(n) + (size_t)(n > 0 ? n : 1) * sizeof(int32_t);
    PCG_CHECK_ARG(scratch && scratch_bytes >= need, "pcg_prune_rows: scratch too small (%zu < %zu)", scratch_bytes, need);
    PcgScratch s;
    pcg_carve_scratch(scratch, scratch_bytes, n, s);
    int32_t *src_rows = reinterpret_cast<int32_t *>(reinterpret_cast<uint8_t *>(scratch) + pcg_compact_scratch_bytes(n));
    if (int rc = pcg_compact(n, FlagMask{mask}, EmitCoordIndex{coords, out_coords, src_rows}, d_n_out, s, st)) return rc;
    if (feats && n > 0) {
        long blocks = cdiv(n * 32, 256);
        if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
        k_gather_rows<<<(unsigned)blocks, 256, 0, st>>>(feats, c, src_rows, d_n_out, out_feats);
        PCG_LAUNCHED();
    }
    return 0;
}

// Source rows of the kept entries of a mask, ascending (the index form of pcg_prune_rows): lets the caller compose the
// prune with a re-ordering of the survivors and move the feature rows once (pcg_gather_rows) instead of twice.
struct EmitIndex {
    int32_t *rows;
    __device__ void operator()(long i, int rank) const { rows[rank] = (int)i; }
};
extern "C" int pcg_mask_to_rows(const uint8_t *mask, long n, int32_t *out_rows, int32_t *d_n_out, void *scratch, size_t scratch_bytes,
                                void *stream) {
    PcgScratch s;
    PCG_CHECK_ARG(pcg_carve_scratch(scratch, scratch_bytes, n, s), "pcg_mask_to_rows: scratch too small (%zu < %zu)", scratch_bytes,
                  pcg_compact_scratch_bytes(n));
    return pcg_compact(n, FlagMask{mask}, EmitIndex{out_rows}, d_n_out, s, as_stream(stream));
}

// ---------------------------------------------------------------------------- pcg_mask_to_rows without caller scratch
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the same
// compaction as pcg_mask_to_rows with its scratch in a file-scope device array of a stated maximum size.  Every GPU that loads
// the library holds the array (4.0 MiB) until exit.  The array's NAME is a host-side stand-in in host code, not a device address:
// the address is taken once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// All calls share the array: the compaction clears the words that must start at zero (tile states, ticket) on the caller's
// stream at every call, and calls that overlap on the GPU (different streams, no ordering between them) RACE -- serialise them.
#define PCG_STATIC_MAX_ITEMS (4L * 1024 * 1024)
// = pcg_compact_scratch_bytes(PCG_STATIC_MAX_ITEMS) rounded up to 256: item flags / tile states + per-tile counts + slack
#define PCG_STATIC_SCRATCH_BYTES (PCG_STATIC_MAX_ITEMS + (PCG_STATIC_MAX_ITEMS / PCG_CHUNK + 2) * 4 + 256 + 248)
static __device__ __align__(256) unsigned char g_static_scratch[PCG_STATIC_SCRATCH_BYTES];

#include <mutex>
static int static_scratch_address(void **out) {
    static std::mutex mu;
    static void *addr[64] = {nullptr};            // per device ordinal
    int dev = -1;
    PCG_CUDA(cudaGetDevice(&dev));
    PCG_CHECK_ARG(dev >= 0 && dev < 64, "pcg_mask_to_rows_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);       // device `dev` is current: its copy of the array
        if (e != cudaSuccess || p == nullptr)
            return pcg_fail("pcg_mask_to_rows_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

extern "C" int pcg_mask_to_rows_static(const uint8_t *mask, long n, int32_t *out_rows, int32_t *d_n_out, void *stream) {
    PCG_CHECK_ARG(n >= 0, "pcg_mask_to_rows_static: n=%ld", n);
    PCG_CHECK_ARG(n <= PCG_STATIC_MAX_ITEMS, "pcg_mask_to_rows_static: n=%ld exceeds the built-in scratch (at most %ld items; use "
                  "pcg_mask_to_rows with caller scratch beyond that)", n, (long)PCG_STATIC_MAX_ITEMS);
    static_assert(PCG_STATIC_SCRATCH_BYTES % 256 == 0, "scratch array size must keep the 256-byte granularity");
    PCG_CHECK_ARG(pcg_compact_scratch_bytes(PCG_STATIC_MAX_ITEMS) <= (size_t)PCG_STATIC_SCRATCH_BYTES, "pcg_mask_to_rows_static: scratch array too small");
    void *scratch = nullptr;
    if (int rc = static_scratch_address(&scratch)) return rc;
    PcgScratch s;
    PCG_CHECK_ARG(pcg_carve_scratch(scratch, PCG_STATIC_SCRATCH_BYTES, n, s), "pcg_mask_to_rows_static: scratch carve failed");
    return pcg_compact(n, FlagMask{mask}, EmitIndex{out_rows}, d_n_out, s, as_stream(stream));
}

// out[r] = src[rows[r]] for r < n_rows: coordinate rows (int32 x 4, src_coords/out_coords optional) and fp32 feature rows of c
// channels (optional), one warp per row.
// bf16 source rows (c % 8 == 0 channels, 16-byte aligned rows) -> fp32 output rows: the survivors of a prune whose candidates
// were only ever written as bf16 (the generative up-sampling emits the operand image of the occupancy head and nothing else)
__global__ void k_gather_rows_bf16(const int32_t *__restrict__ src_coords, const __nv_bfloat16 *__restrict__ feats, int c,
                                   const int32_t *__restrict__ rows, long n, int32_t *__restrict__ out_coords, float *__restrict__ out) {
    // a group of LPR = min(32, pow2 >= c / 8) lanes per row: narrow rows (32 channels = 4 chunks of 16 bytes) put 8 rows on a warp
    // instead of leaving 28 lanes idle
    const int chunks = c >> 3;
    int lpr = 1;
    while (lpr < chunks && lpr < 32) lpr <<= 1;
    const int rpw = 32 / lpr;
    long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31, sub = lane / lpr, t = lane % lpr;
    long nwarps = ((long)gridDim.x * blockDim.x) >> 5;
    for (long r = warp * rpw + sub; r < n; r += nwarps * rpw) {
        const long sr = rows[r];
        if (src_coords && t == 0) reinterpret_cast<int4 *>(out_coords)[r] = pcg_ld_coord(src_coords, sr);
        const uint4 *src = reinterpret_cast<const uint4 *>(feats + sr * c);
        float4 *dst = reinterpret_cast<float4 *>(out + r * c);
        for (int j = t; j < chunks; j += lpr) {              // 8 channels per lane step: one 16-byte load, two 16-byte stores
            const uint4 v = __ldg(src + j);
            dst[2 * j] = make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xFFFF0000u), __uint_as_float(v.y << 16), __uint_as_float(v.y & 0xFFFF0000u));
            dst[2 * j + 1] = make_float4(__uint_as_float(v.z << 16), __uint_as_float(v.z & 0xFFFF0000u), __uint_as_float(v.w << 16), __uint_as_float(v.w & 0xFFFF0000u));
        }
    }
}

__global__ void k_gather_rows2(const int32_t *__restrict__ src_coords, const float *__restrict__ feats, int c,
                               const int32_t *__restrict__ rows, long n, int32_t *__restrict__ out_coords, float *__restrict__ out) {
    long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    long nwarps = ((long)gridDim.x * blockDim.x) >> 5;
    for (long r = warp; r < n; r += nwarps) {
        const long sr = rows[r];
        if (src_coords && lane == 0) reinterpret_cast<int4 *>(out_coords)[r] = pcg_ld_coord(src_coords, sr);
        if (feats) {
            const float *src = feats + sr * c;
            float *dst = out + r * c;
            if ((c & 3) == 0) {
                for (int j = lane; j < (c >> 2); j += 32)
                    reinterpret_cast<float4 *>(dst)[j] = __ldg(reinterpret_cast<const float4 *>(src) + j);
            } else {
                for (int j = lane; j < c; j += 32) dst[j] = __ldg(src + j);
            }
        }
    }
}
extern "C" int pcg_gather_rows(const int32_t *coords, const float *feats, int c, const int32_t *rows, long n_rows,
                               int32_t *out_coords, float *out_feats, void *stream) {
    PCG_CHECK_ARG(n_rows >= 0 && (feats == nullptr || c >= 1), "pcg_gather_rows: n_rows=%ld c=%d", n_rows, c);
    if (n_rows == 0 || (!coords && !feats)) return 0;
    long blocks = cdiv(n_rows * 32, 256);
    if (blocks > PCG_SMS * 8) blocks = PCG_SMS * 8;
    k_gather_rows2<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(coords, feats, c, rows, n_rows, out_coords, out_feats);
    PCG_LAUNCHED();
    return 0;
}

extern "C" int pcg_gather_rows_bf16(const int32_t *coords, const void *feats_bf16, int c, const int32_t *rows, long n_rows,
                                    int32_t *out_coords, float *out_feats, void *stream) {
    PCG_CHECK_ARG(n_rows >= 0 && feats_bf16 != nullptr && c >= 8 && (c & 7) == 0, "pcg_gather_rows_bf16: n_rows=%ld c=%d (c must be a multiple of 8)", n_rows, c);
    PCG_CHECK_ARG(((uintptr_t)feats_bf16 & 15) == 0 && ((uintptr_t)out_feats & 15) == 0, "pcg_gather_rows_bf16: rows must be 16-byte aligned");
    if (n_rows == 0) return 0;
    int lpr_h = 1;
    while (lpr_h < (c >> 3) && lpr_h < 32) lpr_h <<= 1;
    long blocks = cdiv(n_rows * lpr_h, 256);
    if (blocks > PCG_SMS * 16) blocks = PCG_SMS * 16;
    if (blocks < 1) blocks = 1;
    k_gather_rows_bf16<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(coords, reinterpret_cast<const __nv_bfloat16 *>(feats_bf16), c, rows, n_rows,
                                                                         out_coords, out_feats);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- lexsort
__device__ __forceinline__ uint64_t spread3(uint32_t v) {   // 16 bits -> every third bit of 48
    uint64_t x = v & 0xFFFFu;
    x = (x | (x << 32)) & 0x00FF00000000FFFFull;
    x = (x | (x << 16)) & 0x00FF0000FF0000FFull;
    x = (x | (x << 8)) & 0xF00F00F00F00F00Full;
    x = (x | (x << 4)) & 0x30C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x9249249249249249ull;
    return x;
}

__global__ void k_pack_keys(const int32_t *__restrict__ coords, long n, int morton, uint64_t *keys, int32_t *idx) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int4 c = pcg_ld_coord(coords, i);
    if (morton) {
        uint64_t m = (spread3((uint32_t)(c.y + 32768)) << 2) | (spread3((uint32_t)(c.z + 32768)) << 1) | spread3((uint32_t)(c.w + 32768));
        keys[i] = ((uint64_t)(uint16_t)c.x << 48) | (m & 0xFFFFFFFFFFFFull);
    } else {
        keys[i] = pcg_pack(c.x, c.y, c.z, c.w);
    }
    idx[i] = (int)i;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" size_t pcg_sort_rows_scratch_bytes(long n) {
    size_t temp = 0;
    long m = n > 0 ? n : 1;
    cub::DeviceRadixSort::SortPairs(nullptr, temp, (const uint64_t *)nullptr, (uint64_t *)nullptr, (const int32_t *)nullptr,
                                    (int32_t *)nullptr, (int)m);
    return align256(temp) + 2 * align256((size_t)m * 8) + align256((size_t)m * 4) + 256;
}

extern "C" int pcg_sort_rows_bits(const int32_t *coords, long n, int morton, int begin_bit, int end_bit, int32_t *perm, void *scratch,
                                  size_t scratch_bytes, void *stream);
extern "C" int pcg_sort_rows(const int32_t *coords, long n, int morton, int32_t *perm, void *scratch, size_t scratch_bytes,
                             void *stream) {
    return pcg_sort_rows_bits(coords, n, morton, 0, 64, perm, scratch, scratch_bytes, stream);
}

// Stable sort on key bits [begin_bit, end_bit) only.  Key layout: morton = 0: b:16 | x:16 | y:16 | z:16 (biased by 32768);
// morton = 1: b:16 | 48 bits of x/y/z interleaved (bit 3 i + 2 = x_i, + 1 = y_i, + 0 = z_i).  A Z-order that only serves
// memory locality needs the low coordinate bits alone: [0, 36) = 12 bits per axis is 5 radix passes instead of 8.
extern "C" int pcg_sort_rows_bits(const int32_t *coords, long n, int morton, int begin_bit, int end_bit, int32_t *perm, void *scratch,
                                  size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    if (n <= 0) return 0;
    PCG_CHECK_ARG(0 <= begin_bit && begin_bit < end_bit && end_bit <= 64, "pcg_sort_rows_bits: bits [%d, %d)", begin_bit, end_bit);
    PCG_CHECK_ARG(scratch && scratch_bytes >= pcg_sort_rows_scratch_bytes(n), "pcg_sort_rows: scratch too small");
    uint8_t *p = reinterpret_cast<uint8_t *>(scratch);
    uint64_t *k_in = reinterpret_cast<uint64_t *>(p); p += align256((size_t)n * 8);
    uint64_t *k_out = reinterpret_cast<uint64_t *>(p); p += align256((size_t)n * 8);
    int32_t *i_in = reinterpret_cast<int32_t *>(p); p += align256((size_t)n * 4);
    size_t temp = scratch_bytes - (size_t)(p - reinterpret_cast<uint8_t *>(scratch));
    k_pack_keys<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(coords, n, morton, k_in, i_in);
    PCG_LAUNCHED();
    // plumbing: CUB's radix sort orders the packed 64-bit keys (stable)
    PCG_CUDA(cub::DeviceRadixSort::SortPairs(p, temp, k_in, k_out, i_in, perm, (int)n, begin_bit, end_bit, st));
    pcg_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

// ---------------------------------------------------------------------------- block entry
// One pass from the caller's (n,6) float rows [x y z r g b] to the codec's input tensor in a given row order:
// coords[i] = (batch, floor(xyz[perm[i]])), feats[i] = (1, rgb[perm[i]])  (model/model.py:142-161: occupancy channel +
// colours, integer coordinates).  Replaces four framework gathers / casts / concatenations.
__global__ void k_gather_block(const float *__restrict__ cloud, const int32_t *__restrict__ perm, long n, int batch,
                               int32_t *__restrict__ coords, float *__restrict__ feats) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *src = cloud + (long)(perm ? perm[i] : i) * 6;
    const float2 a = __ldg(reinterpret_cast<const float2 *>(src)), b = __ldg(reinterpret_cast<const float2 *>(src) + 1),
                 c = __ldg(reinterpret_cast<const float2 *>(src) + 2);
    reinterpret_cast<int4 *>(coords)[i] = make_int4(batch, (int)floorf(a.x), (int)floorf(a.y), (int)floorf(b.x));
    reinterpret_cast<float4 *>(feats)[i] = make_float4(1.f, b.y, c.x, c.y);
}

extern "C" int pcg_gather_block(const float *cloud, const int32_t *perm, long n, int batch, int32_t *coords, float *feats, void *stream) {
    PCG_CHECK_ARG(((uintptr_t)cloud & 7) == 0 && ((uintptr_t)coords & 15) == 0 && ((uintptr_t)feats & 15) == 0, "pcg_gather_block: misaligned buffers");
    if (n <= 0) return 0;
    k_gather_block<<<(unsigned)cdiv(n, 256), 256, 0, as_stream(stream)>>>(cloud, perm, n, batch, coords, feats);
    PCG_LAUNCHED();
    return 0;
}
