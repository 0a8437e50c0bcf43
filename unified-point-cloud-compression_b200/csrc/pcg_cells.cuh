// Device pieces of the octree-cell organisation shared by pcg_cells.cu and the occupancy-head kernels (pcg_conv_tc.cu).
#pragma once
#include "pcg_common.cuh"

// What a kernel needs to derive the 3x3x3 neighbourhood of a candidate row from the cell structure (no table of the children)
struct CellsMapArgs {
    const uint64_t *cell_keys;      // hash table of the cells (keys / row of the cell)
    const int32_t *cell_vals;
    uint64_t cmask;                 // capacity - 1
    const int32_t *out_coords;      // candidate rows (b, x, y, z)
    const int32_t *out_cell;        // cell of every candidate row
    const uint8_t *cell_mask;       // children present in a cell (bit = bx + 2 by + 4 bz)
    const int32_t *cell_base;       // first candidate row of a cell
    int ts_out;                     // tensor stride of the candidates
};

// rows[k] = candidate row at (coordinate of row o) + offset k * ts_out, k = (dx+1) + 3 (dy+1) + 9 (dz+1), or -1.
// The 27 neighbours live in the 2x2x2 cells around the child: the own cell is known, the other 7 are probed in the (8x smaller)
// cell table -- their probe chains advanced in lockstep -- and a neighbour's row is base[cell] + rank of its bit in the cell mask.
__device__ __forceinline__ void cells_map3_row(const CellsMapArgs &a, long o, bool live, int (&rows)[27]) {
    const int4 c = live ? pcg_ld_coord(a.out_coords, o) : make_int4(0, 0, 0, 0);
    const int ts_in = 2 * a.ts_out;
    const int Cx = pcg_floor_to(c.y, ts_in), Cy = pcg_floor_to(c.z, ts_in), Cz = pcg_floor_to(c.w, ts_in);
    const int bx = (c.y - Cx) / a.ts_out, by = (c.z - Cy) / a.ts_out, bz = (c.w - Cz) / a.ts_out;
    int cellv[8];
    uint64_t key[8], slot[8];
    uint32_t pend = 0;
    cellv[0] = live ? __ldg(a.out_cell + o) : -1;
#pragma unroll
    for (int s = 1; s < 8; ++s) {
        cellv[s] = -1;
        const int x = Cx + ((s & 1) ? (bx ? ts_in : -ts_in) : 0);
        const int y = Cy + ((s & 2) ? (by ? ts_in : -ts_in) : 0);
        const int z = Cz + ((s & 4) ? (bz ? ts_in : -ts_in) : 0);
        key[s] = pcg_pack(c.x, x, y, z);
        slot[s] = pcg_mix(key[s]) & a.cmask;
        if (live && pcg_in_range(c.x, x, y, z)) pend |= 1u << s;
    }
    while (pend) {
        uint64_t got[8];
#pragma unroll
        for (int s = 1; s < 8; ++s)
            if ((pend >> s) & 1u) got[s] = __ldg(a.cell_keys + slot[s]);
#pragma unroll
        for (int s = 1; s < 8; ++s) {
            if (!((pend >> s) & 1u)) continue;
            if (got[s] == key[s]) { cellv[s] = __ldg(a.cell_vals + slot[s]); pend &= ~(1u << s); }
            else if (got[s] == PCG_EMPTY_KEY) pend &= ~(1u << s);
            else slot[s] = (slot[s] + 1) & a.cmask;
        }
    }
    int base[8];
    uint32_t mask[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        base[s] = cellv[s] >= 0 ? __ldg(a.cell_base + cellv[s]) : 0;
        mask[s] = cellv[s] >= 0 ? (uint32_t)a.cell_mask[cellv[s]] : 0u;
    }
    int k = 0;
#pragma unroll
    for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx, ++k) {
                const int tx = bx + dx, ty = by + dy, tz = bz + dz;          // in [-1, 2]
                const int s = ((tx < 0 || tx > 1) ? 1 : 0) | ((ty < 0 || ty > 1) ? 2 : 0) | ((tz < 0 || tz > 1) ? 4 : 0);
                const int bit = (tx & 1) | ((ty & 1) << 1) | ((tz & 1) << 2);
                // (s is a compile-time-unknown index into 8 registers: resolved by selects after unrolling)
                int bs = 0;
                uint32_t m = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q == s) { bs = base[q]; m = mask[q]; }
                rows[k] = (live && ((m >> bit) & 1u)) ? bs + __popc(m & ((1u << bit) - 1u)) : -1;
            }
}
