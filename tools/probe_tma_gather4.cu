// Probe (run on the GPU box): does cp.async.bulk.tensor.2d tile::gather4 deliver four arbitrary rows of a bf16
// [n][64] matrix into a SWIZZLE_128B operand tile, with out-of-range row indices zero-filled?  Tries box rows 1 and 4.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/probe_tma_gather4.bin tools/probe_tma_gather4.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void k_probe(const __grid_constant__ CUtensorMap tm, const int *rows, int ngroups, int col0, uint32_t tx_per_group,
                        uint4 *out, int *status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    uint8_t *tile = smem + (base - smem_u32(smem));
    for (int i = threadIdx.x; i < 128 * 128 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(tile)[i] = make_uint4(0xDEADBEEFu, 0xDEADBEEFu, 0xDEADBEEFu, 0xDEADBEEFu);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(tx_per_group * ngroups) : "memory");
        for (int g = 0; g < ngroups; ++g) {
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                ::"r"(base + g * 512), "l"(&tm), "r"(col0), "r"(rows[4 * g]), "r"(rows[4 * g + 1]), "r"(rows[4 * g + 2]), "r"(rows[4 * g + 3]),
                "r"(smem_u32(&bar))
                : "memory");
        }
        // bounded wait
        uint32_t done = 0;
        long long t0 = clock64();
        while (!done && clock64() - t0 < 2000000000LL) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        }
        status[0] = done ? 1 : -1;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 128 * 128 / 16; i += blockDim.x) out[i] = reinterpret_cast<uint4 *>(tile)[i];
}

int main() {
    EncodeTiled encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres);
    if (e != cudaSuccess || !encode) { printf("no cuTensorMapEncodeTiled entry point: %s\n", cudaGetErrorString(e)); return 1; }
    const int n = 1000, c = 128;     // 2 chunks of 64
    std::vector<__nv_bfloat16> h((size_t)n * c);
    for (int r = 0; r < n; ++r) for (int j = 0; j < c; ++j) h[(size_t)r * c + j] = __float2bfloat16((float)((r % 250) + j * 0.001953125f * 0 ) );
    // value = r%250 exactly representable; column identity carried separately below via a second pattern
    for (int r = 0; r < n; ++r) for (int j = 0; j < c; ++j) h[(size_t)r * c + j] = __float2bfloat16((float)((r * 7 + j) % 256));
    __nv_bfloat16 *d; cudaMalloc(&d, h.size() * 2); cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    std::vector<int> rows(128);
    for (int i = 0; i < 128; ++i) rows[i] = (i * 37 + 11) % n;
    rows[2] = -1; rows[5] = n + 3; rows[9] = -1;     // absent neighbours: expect zeros
    int *drows; cudaMalloc(&drows, 128 * 4); cudaMemcpy(drows, rows.data(), 128 * 4, cudaMemcpyHostToDevice);
    uint4 *dout; cudaMalloc(&dout, 128 * 128); int *dstat; cudaMalloc(&dstat, 4);
    for (int box_rows = 1; box_rows <= 4; box_rows += 3) {
        CUtensorMap tm;
        cuuint64_t dims[2] = {(cuuint64_t)c, (cuuint64_t)n};
        cuuint64_t strides[1] = {(cuuint64_t)c * 2};
        cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("box_rows=%d encode -> %d\n", box_rows, (int)r);
        if (r != CUDA_SUCCESS) continue;
        for (int col0 = 0; col0 <= 64; col0 += 64) {
            cudaMemset(dstat, 0, 4);
            k_probe<<<1, 128, 128 * 128 + 2048>>>(tm, drows, 32, col0, 512, dout, dstat);
            cudaError_t err = cudaDeviceSynchronize();
            int st = 0; cudaMemcpy(&st, dstat, 4, cudaMemcpyDeviceToHost);
            printf("  col0=%d sync=%s status=%d\n", col0, cudaGetErrorString(err), st);
            if (err != cudaSuccess) return 2;
            std::vector<uint16_t> o(128 * 64);
            cudaMemcpy(o.data(), dout, 128 * 128, cudaMemcpyDeviceToHost);
            // expected: tile row i, 16B chunk j stored at i*128 + ((j ^ (i&7))<<4)
            long bad = 0, untouched = 0;
            for (int i = 0; i < 128; ++i) for (int j = 0; j < 64; ++j) {
                int chunk = j >> 3, within = j & 7;
                uint16_t got = o[i * 64 + ((chunk ^ (i & 7)) << 3) + within];
                int rr = rows[i];
                float want = (rr >= 0 && rr < n) ? (float)((rr * 7 + col0 + j) % 256) : 0.f;
                __nv_bfloat16 wb = __float2bfloat16(want);
                uint16_t wbits = *reinterpret_cast<uint16_t *>(&wb);
                if (got == 0xBEEF || got == 0xDEAD) ++untouched;
                if (got != wbits) { if (bad < 5) printf("    mismatch row %d col %d got %04x want %04x (src row %d)\n", i, j, got, wbits, rr); ++bad; }
            }
            printf("  -> mismatches %ld untouched %ld of %d\n", bad, untouched, 128 * 64);
        }
    }
    return 0;
}
