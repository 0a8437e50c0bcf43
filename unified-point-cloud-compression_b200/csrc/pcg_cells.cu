// Octree-cell organisation of the kernel-5 / stride-2 generative up-sampling (ME3, model/transforms.py:129-137)
// and everything that follows it on the candidate set (3x3x3 kernel map for the occupancy head, the
// transposed convolution itself as GEMM-first + gather-reduce).
//
// A parent at tensor stride ts sits on multiples of ts; its 5^3 children at stride ts/2 are p + off*(ts/2),
// off in [-2,2]^3.  Group the children by CELL = ts-aligned cube (<= 8 children, bit b = bx + 2 by + 4 bz):
//   * cells          = parents dilated by {-1,0,1}^3 * ts     (pcg_generate_coords with ksize 3, stride 1)
//   * pnbr[e][cell]  = parent row at cell + e*ts, e in {-1,0,1}^3  (pcg_kernel_map of the cells on the parents)
//   * child b exists iff some parent neighbour e has, on every axis, e_a in {-1,0,1} (b_a = 0) or {0,1} (b_a = 1)
//     [ off_a = b_a - 2 e_a must lie in [-2,2] ]
// so the candidate set, its parent lists (= the transposed kernel map) and its 27-neighbourhood all follow from
// 27 probes per CELL into the small parent table, instead of 125 inserts per parent + 27 probes per CHILD into a
// table of all children.  Rows come out cell-major: siblings are adjacent and neighbouring cells are near.
#include <cub/device/device_scan.cuh>

#include <stdlib.h>

#include "pcg_common.cuh"

// valid-parent mask over the 27 e-indices (bit i = ex+1 + 3(ey+1) + 9(ez+1)) for child bit pattern b
__device__ __forceinline__ uint32_t cells_valid_e(int b) {
    // axis masks over e in {-1,0,1}: b_a = 0 -> all three, b_a = 1 -> e in {0,1}
    const uint32_t ax = (b & 1) ? 0b110u : 0b111u, ay = (b & 2) ? 0b110u : 0b111u, az = (b & 4) ? 0b110u : 0b111u;
    uint32_t m = 0;
#pragma unroll
    for (int ez = 0; ez < 3; ++ez)
#pragma unroll
        for (int ey = 0; ey < 3; ++ey)
#pragma unroll
            for (int ex = 0; ex < 3; ++ex)
                if (((ax >> ex) & 1u) && ((ay >> ey) & 1u) && ((az >> ez) & 1u)) m |= 1u << (ex + 3 * ey + 9 * ez);
    return m;
}

// ---------------------------------------------------------------------------- parent lists by scatter
// pnbr[e][cell] = parent row at cell + e*ts.  The cells were generated as parent + d*ts, d in {-1,0,1}^3, so every
// (parent, d) pair names exactly one cell: one probe of the CELL table per pair (27 * n_parents probes, all hits) and a
// scatter into pnbr[-d][cell], instead of 27 probes per cell of the parent table (~6.7x more probes, mostly misses that
// walk a probe chain to its end).  Distinct (parent, d) pairs write distinct entries: no race, order free.
__global__ void k_cells_parent_scatter(const int32_t *__restrict__ parent_coords, long n_par, int ts, const uint64_t *__restrict__ ckeys,
                                       const int32_t *__restrict__ cvals, uint64_t cmask, int32_t *__restrict__ pnbr, long pitch,
                                       int32_t *__restrict__ d_status) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_par * 27) return;
    const long p = i / 27;
    const int d = (int)(i - p * 27);
    const int dx = d % 3 - 1, dy = (d / 3) % 3 - 1, dz = d / 9 - 1;
    const int4 c = pcg_ld_coord(parent_coords, p);
    const int x = c.y + dx * ts, y = c.z + dy * ts, z = c.w + dz * ts;
    if (!pcg_in_range(c.x, x, y, z)) return;
    const int cell = pcg_table_find(ckeys, cvals, cmask, pcg_pack(c.x, x, y, z));
    if (cell < 0) { atomicOr(d_status, 1); return; }       // the cell set is not the 3^3 dilation of these parents
    pnbr[(long)(26 - d) * pitch + cell] = (int)p;
}

extern "C" int pcg_cells_parent_map(const int32_t *parent_coords, long n_par, int ts, const uint64_t *cell_keys,
                                    const int32_t *cell_vals, long cell_cap, long n_cells, int32_t *pnbr, long pnbr_pitch,
                                    int32_t *d_status, void *stream) {
    PCG_CHECK_ARG(cell_cap > 0 && (cell_cap & (cell_cap - 1)) == 0, "pcg_cells_parent_map: cell_cap=%ld must be a power of two", cell_cap);
    PCG_CHECK_ARG(pnbr_pitch >= n_cells && ts >= 1, "pcg_cells_parent_map: pitch=%ld n_cells=%ld ts=%d", pnbr_pitch, n_cells, ts);
    PCG_CHECK_ARG(n_par * 27 < 2147483647L, "pcg_cells_parent_map: n_par=%ld too large", n_par);
    cudaStream_t st = as_stream(stream);
    PCG_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    PCG_CUDA(cudaMemsetAsync(pnbr, 0xFF, (size_t)27 * pnbr_pitch * sizeof(int32_t), st));      // -1 everywhere
    if (n_par <= 0 || n_cells <= 0) return 0;
    k_cells_parent_scatter<<<(unsigned)cdiv(n_par * 27, 256), 256, 0, st>>>(parent_coords, n_par, ts, cell_keys, cell_vals,
                                                                           (uint64_t)cell_cap - 1, pnbr, pnbr_pitch, d_status);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- cell info
// occupancy bits of the 27 parent neighbours -> 8-bit child mask + child count
__global__ void k_cells_mask(const int32_t *__restrict__ pnbr, long pitch, long n_cells, uint8_t *__restrict__ cell_mask,
                             int32_t *__restrict__ counts) {
    long c = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cells) return;
    uint32_t occ = 0;
#pragma unroll
    for (int e = 0; e < 27; ++e)
        if (__ldg(pnbr + (long)e * pitch + c) >= 0) occ |= 1u << e;
    uint32_t mask = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b)
        if (occ & cells_valid_e(b)) mask |= 1u << b;
    cell_mask[c] = (uint8_t)mask;
    counts[c] = __popc(mask);
}

__global__ void k_cells_total(const int32_t *__restrict__ base, const uint8_t *__restrict__ cell_mask, long n_cells,
                              int32_t *d_n_out) {
    if (threadIdx.x == 0 && blockIdx.x == 0)
        d_n_out[0] = n_cells > 0 ? base[n_cells - 1] + __popc((uint32_t)cell_mask[n_cells - 1]) : 0;
}

extern "C" size_t pcg_cells_scratch_bytes(long n_cells) {
    size_t temp = 0;
    long m = n_cells > 0 ? n_cells : 1;
    cub::DeviceScan::ExclusiveSum(nullptr, temp, (const int32_t *)nullptr, (int32_t *)nullptr, (int)m);
    return ((temp + 255) & ~(size_t)255) + (size_t)m * sizeof(int32_t) + 512;
}

extern "C" int pcg_cells_info(const int32_t *pnbr, long pnbr_pitch, long n_cells, uint8_t *cell_mask, int32_t *cell_base,
                              int32_t *d_n_out, void *scratch, size_t scratch_bytes, void *stream) {
    cudaStream_t st = as_stream(stream);
    PCG_CHECK_ARG(n_cells < 2147483647L / 8, "pcg_cells_info: n_cells=%ld too large", n_cells);
    PCG_CHECK_ARG(scratch && scratch_bytes >= pcg_cells_scratch_bytes(n_cells), "pcg_cells_info: scratch too small");
    if (n_cells <= 0) {
        PCG_CUDA(cudaMemsetAsync(d_n_out, 0, sizeof(int32_t), st));
        return 0;
    }
    int32_t *counts = reinterpret_cast<int32_t *>(scratch);
    size_t off = ((size_t)n_cells * sizeof(int32_t) + 255) & ~(size_t)255;
    void *temp = reinterpret_cast<uint8_t *>(scratch) + off;
    size_t temp_bytes = scratch_bytes - off;
    k_cells_mask<<<(unsigned)cdiv(n_cells, 256), 256, 0, st>>>(pnbr, pnbr_pitch, n_cells, cell_mask, counts);
    PCG_LAUNCHED();
    // plumbing: CUB exclusive scan of the per-cell child counts
    PCG_CUDA(cub::DeviceScan::ExclusiveSum(temp, temp_bytes, counts, cell_base, (int)n_cells, st));
    pcg_launches.fetch_add(1, std::memory_order_relaxed);
    k_cells_total<<<1, 32, 0, st>>>(cell_base, cell_mask, n_cells, d_n_out);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- children
__global__ void k_cells_emit(const int32_t *__restrict__ cell_coords, const uint8_t *__restrict__ cell_mask,
                             const int32_t *__restrict__ cell_base, long n_cells, int ts_out, int32_t *__restrict__ out_coords,
                             int32_t *__restrict__ out_cell) {
    // 8 threads per cell: thread b writes child b (rows of a cell are adjacent -> coalesced int4 stores)
    long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    long c = t >> 3;
    int b = (int)(t & 7);
    if (c >= n_cells) return;
    uint32_t mask = cell_mask[c];
    if (!((mask >> b) & 1u)) return;
    int4 cc = pcg_ld_coord(cell_coords, c);
    long row = (long)cell_base[c] + __popc(mask & ((1u << b) - 1u));
    cc.y += (b & 1) * ts_out; cc.z += ((b >> 1) & 1) * ts_out; cc.w += ((b >> 2) & 1) * ts_out;
    reinterpret_cast<int4 *>(out_coords)[row] = cc;
    out_cell[row] = (int)c;
}

extern "C" int pcg_cells_emit(const int32_t *cell_coords, const uint8_t *cell_mask, const int32_t *cell_base, long n_cells,
                              int ts_out, int32_t *out_coords, int32_t *out_cell, void *stream) {
    PCG_CHECK_ARG(ts_out >= 1, "pcg_cells_emit: ts_out=%d", ts_out);
    if (n_cells <= 0) return 0;
    k_cells_emit<<<(unsigned)cdiv(n_cells * 8, 256), 256, 0, as_stream(stream)>>>(cell_coords, cell_mask, cell_base, n_cells,
                                                                                 ts_out, out_coords, out_cell);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- 3x3x3 kernel map of the children
// Same result as pcg_kernel_map(table of the children, children, ksize 3, ts_out) without a table of the children:
// the 27 neighbours of a child live in the 2x2x2 cells around it; a neighbour's row is base[cell] + rank of its bit.
__global__ void __launch_bounds__(128) k_cells_kernel_map3(const uint64_t *__restrict__ cell_keys, const int32_t *__restrict__ cell_vals,
                                                           uint64_t cmask, const int32_t *__restrict__ out_coords,
                                                           const int32_t *__restrict__ out_cell, long n_out, int ts_out,
                                                           const uint8_t *__restrict__ cell_mask, const int32_t *__restrict__ cell_base,
                                                           int32_t *__restrict__ nbr, long pitch, uint32_t *__restrict__ tile_mask) {
    __shared__ int s_base[8][128];
    __shared__ uint8_t s_mask[8][128];
    __shared__ uint32_t smask_any;
    const int tid = threadIdx.x;
    long o = (long)blockIdx.x * 128 + tid;
    bool live = o < n_out;
    if (tid == 0) smask_any = 0;
    int4 c = live ? pcg_ld_coord(out_coords, o) : make_int4(0, 0, 0, 0);
    const int ts_in = 2 * ts_out;
    // cell origin and child bits
    int Cx = pcg_floor_to(c.y, ts_in), Cy = pcg_floor_to(c.z, ts_in), Cz = pcg_floor_to(c.w, ts_in);
    int bx = (c.y - Cx) / ts_out, by = (c.z - Cy) / ts_out, bz = (c.w - Cz) / ts_out;
    int own = live ? __ldg(out_cell + o) : -1;
    // the 7 neighbour-cell probes of a child are independent: their probe chains advance in lockstep, so the (dependent) loads of
    // one chain overlap those of the others instead of running one chain after the other (the kernel is latency bound)
    int cellv[8];
    uint64_t key[8], slot[8];
    uint32_t pend = 0;
    cellv[0] = own;
#pragma unroll
    for (int s = 1; s < 8; ++s) {
        cellv[s] = -1;
        const int x = Cx + ((s & 1) ? (bx ? ts_in : -ts_in) : 0);
        const int y = Cy + ((s & 2) ? (by ? ts_in : -ts_in) : 0);
        const int z = Cz + ((s & 4) ? (bz ? ts_in : -ts_in) : 0);
        key[s] = pcg_pack(c.x, x, y, z);
        slot[s] = pcg_mix(key[s]) & cmask;
        if (live && pcg_in_range(c.x, x, y, z)) pend |= 1u << s;
    }
    while (pend) {
        uint64_t got[8];
#pragma unroll
        for (int s = 1; s < 8; ++s)
            if ((pend >> s) & 1u) got[s] = __ldg(cell_keys + slot[s]);
#pragma unroll
        for (int s = 1; s < 8; ++s) {
            if (!((pend >> s) & 1u)) continue;
            if (got[s] == key[s]) { cellv[s] = __ldg(cell_vals + slot[s]); pend &= ~(1u << s); }
            else if (got[s] == PCG_EMPTY_KEY) pend &= ~(1u << s);
            else slot[s] = (slot[s] + 1) & cmask;
        }
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        const int cell = cellv[s];
        s_base[s][tid] = cell >= 0 ? __ldg(cell_base + cell) : 0;
        s_mask[s][tid] = cell >= 0 ? cell_mask[cell] : (uint8_t)0;
    }
    __syncthreads();
    uint32_t any_bits = 0;   // offsets (bit k) for which this warp has a neighbour
    int k = 0;
#pragma unroll
    for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx, ++k) {
                int tx = bx + dx, ty = by + dy, tz = bz + dz;          // in [-1, 2]
                int s = ((tx < 0 || tx > 1) ? 1 : 0) | ((ty < 0 || ty > 1) ? 2 : 0) | ((tz < 0 || tz > 1) ? 4 : 0);
                int bit = (tx & 1) | ((ty & 1) << 1) | ((tz & 1) << 2);
                uint32_t m = s_mask[s][tid];
                int row = (live && ((m >> bit) & 1u)) ? s_base[s][tid] + __popc(m & ((1u << bit) - 1u)) : -1;
                nbr[(long)k * pitch + o] = row;
                if (__ballot_sync(0xffffffffu, row >= 0)) any_bits |= 1u << k;
            }
    if (tile_mask) {
        if ((tid & 31) == 0) atomicOr(&smask_any, any_bits);
        __syncthreads();
        if (tid < 4) tile_mask[(long)blockIdx.x * 4 + tid] = tid == 0 ? smask_any : 0u;
    }
}

extern "C" int pcg_cells_kernel_map3(const uint64_t *cell_keys, const int32_t *cell_vals, long cell_cap, const int32_t *out_coords,
                                     const int32_t *out_cell, long n_out, int ts_out, const uint8_t *cell_mask,
                                     const int32_t *cell_base, int32_t *nbr, long nbr_pitch, uint32_t *tile_mask, void *stream) {
    PCG_CHECK_ARG(cell_cap > 0 && (cell_cap & (cell_cap - 1)) == 0, "pcg_cells_kernel_map3: cell_cap=%ld must be a power of two", cell_cap);
    PCG_CHECK_ARG(nbr_pitch == pcg_nbr_pitch(n_out), "pcg_cells_kernel_map3: nbr_pitch=%ld must be pcg_nbr_pitch(n_out)", nbr_pitch);
    PCG_CHECK_ARG(ts_out >= 1, "pcg_cells_kernel_map3: ts_out=%d", ts_out);
    if (n_out <= 0) return 0;
    k_cells_kernel_map3<<<(unsigned)cdiv(n_out, 128), 128, 0, as_stream(stream)>>>(
        cell_keys, cell_vals, (uint64_t)cell_cap - 1, out_coords, out_cell, n_out, ts_out, cell_mask, cell_base, nbr, nbr_pitch,
        tile_mask);
    PCG_LAUNCHED();
    return 0;
}

// ---------------------------------------------------------------------------- gather-reduce of the per-parent products
// out[child] = act(bias + sum_{e valid} T[pnbr[e][cell]][k(b - 2e)][:]), e ascending: one fixed-order fp32 chain per element.
// Cell-centric: the 8 child slots of a cell x LANES = cout/8 lanes (8 channels = one 16-byte load each) are adjacent
// threads, a warp covers 32/LANES child slots of ONE cell.  The warp reads the cell's 27 parent rows once (lane e loads
// pnbr[e][cell]), ballots which parents exist and walks only those -- a warp-uniform loop, four parents per batch so
// that four independent 16-byte loads of T are in flight per lane; whether parent e reaches child b is a per-lane
// predicate, not a branch.
template <int LANES>
__global__ void __launch_bounds__(256) k_cells_reduce(const __nv_bfloat16 *__restrict__ T, int cout, const int32_t *__restrict__ pnbr,
                                                      long pitch, const uint8_t *__restrict__ cell_mask,
                                                      const int32_t *__restrict__ cell_base, long n_cells,
                                                      const float *__restrict__ bias, int act, float *__restrict__ out_f32,
                                                      __nv_bfloat16 *__restrict__ out_bf16) {
    constexpr int CPW = 32 / LANES;                 // child slots per warp
    constexpr int WPC = 8 / CPW;                    // warps per cell
    const int lane = threadIdx.x & 31;
    const long n_units = n_cells * WPC;             // one unit = the CPW child slots of one cell handled by one warp
    const long n_warps = ((long)gridDim.x * blockDim.x) >> 5;
    long unit = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (unit >= n_units) return;                    // warp-uniform
    const int part = lane % LANES;
    const long t_row = (long)125 * cout;
    float bv[8];                                   // this lane's 8 bias values, loaded once (a lane's channels never change)
#pragma unroll
    for (int j = 0; j < 8; ++j) bv[j] = bias ? __ldg(bias + part * 8 + j) : 0.f;
    // Grid-stride over units with the NEXT unit's parent row / child mask / row base requested before the current unit's rows
    // of T: each unit is a chain of two dependent global loads (parent list -> T rows) with few bytes in flight per warp, so the
    // first link of unit i+1 overlaps the second link of unit i.
    long cell = unit / WPC;
    int my_prow = lane < 27 ? __ldg(pnbr + (long)lane * pitch + cell) : -1;
    uint32_t cmask = cell_mask[cell];
    int cbase = __ldg(cell_base + cell);
    for (; unit < n_units; unit += n_warps) {
        const long nunit = unit + n_warps;
        const bool has_next = nunit < n_units;
        const long ncell = has_next ? nunit / WPC : cell;
        const int nprow = (has_next && lane < 27) ? __ldg(pnbr + (long)lane * pitch + ncell) : -1;
        const uint32_t ncmask = has_next ? cell_mask[ncell] : 0u;
        const int ncbase = has_next ? __ldg(cell_base + ncell) : 0;

        const int b = (int)(unit % WPC) * CPW + lane / LANES;
        const int bx = b & 1, by = (b >> 1) & 1, bz = (b >> 2) & 1;
        uint32_t present = __ballot_sync(0xffffffffu, my_prow >= 0);
        const bool exists = (cmask >> b) & 1u;
        const uint32_t reach = exists ? cells_valid_e(b) : 0u;      // parents e that reach child b
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        while (present) {
            int e[4], pr[4];
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                e[u] = present ? __ffs(present) - 1 : 31;
                present &= present - 1;
                pr[u] = __shfl_sync(0xffffffffu, my_prow, e[u]);        // lane 31 holds -1
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int ex = e[u] % 3 - 1, ey = (e[u] / 3) % 3 - 1, ez = e[u] / 9 - 1;
                const int k = (bx - 2 * ex + 2) + 5 * (by - 2 * ey + 2) + 25 * (bz - 2 * ez + 2);
                const bool ok = pr[u] >= 0 && ((reach >> e[u]) & 1u);
                v[u] = ok ? __ldg(reinterpret_cast<const uint4 *>(T + (long)pr[u] * t_row + (long)k * cout) + part) : make_uint4(0, 0, 0, 0);
                if (!ok) pr[u] = -1;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (pr[u] < 0) continue;                // absent / non-reaching parents add nothing
                const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    acc[2 * j] += __uint_as_float(w[j] << 16);
                    acc[2 * j + 1] += __uint_as_float(w[j] & 0xFFFF0000u);
                }
            }
        }
        if (exists) {
            const long o = (long)cbase + __popc(cmask & ((1u << b) - 1u));
            const int c0 = part * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = pcg_apply_act(acc[j] + bv[j], act);
            if (out_f32) {
                float4 *dst = reinterpret_cast<float4 *>(out_f32 + o * cout + c0);
                dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
            }
            if (out_bf16) {
                uint32_t pk[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
                    pk[j] = *reinterpret_cast<uint32_t *>(&h);
                }
                *reinterpret_cast<uint4 *>(out_bf16 + o * cout + c0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
        }
        cell = ncell; my_prow = nprow; cmask = ncmask; cbase = ncbase;
    }
}

extern "C" int pcg_cells_reduce(const void *t_bf16, int cout, const int32_t *pnbr, long pnbr_pitch, const uint8_t *cell_mask,
                                const int32_t *cell_base, long n_cells, const float *bias, int act, float *out_f32, void *out_bf16,
                                void *stream) {
    PCG_CHECK_ARG(cout == 32 || cout == 64 || cout == 128 || cout == 256, "pcg_cells_reduce: cout=%d must be 32, 64, 128 or 256", cout);
    PCG_CHECK_ARG(((uintptr_t)t_bf16 & 15) == 0, "pcg_cells_reduce: T must be 16-byte aligned");
    if (n_cells <= 0) return 0;
    cudaStream_t st = as_stream(stream);
    const __nv_bfloat16 *T = reinterpret_cast<const __nv_bfloat16 *>(t_bf16);
    __nv_bfloat16 *ob = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    const int lanes = cout / 8;
    long nb = cdiv(n_cells * 8 * lanes, 256);
    static const long cap_blocks = getenv("PCG_REDUCE_BLOCKS") ? atol(getenv("PCG_REDUCE_BLOCKS")) : (long)PCG_SMS * 5 * 8;   // 8 units per warp at 5 blocks/SM
    if (nb > cap_blocks) nb = cap_blocks;
    const unsigned blocks = (unsigned)nb;
#define PCG_REDUCE(L) k_cells_reduce<L><<<blocks, 256, 0, st>>>(T, cout, pnbr, pnbr_pitch, cell_mask, cell_base, n_cells, bias, act, out_f32, ob)
    switch (lanes) {
        case 4: PCG_REDUCE(4); break;
        case 8: PCG_REDUCE(8); break;
        case 16: PCG_REDUCE(16); break;
        default: PCG_REDUCE(32); break;
    }
#undef PCG_REDUCE
    PCG_LAUNCHED();
    return 0;
}
