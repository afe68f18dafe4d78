// Probe: what does one K block of the convolution's main loop cost?  A warp-specialised TMA -> tcgen05.mma pipeline exactly as
// in csrc/conv_pm.cu (one producer thread, one MMA thread, full/empty mbarriers, a commit per K block), with no epilogue,
// run on one CTA and on all SMs, for several N, stage counts and K-block widths (64 or 128 elements per stage).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -I flink-tensorflow_b200/csrc -o tools/probe_mainloop tools/probe_mainloop.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
#include "ptx.cuh"
using namespace ftf;

typedef CUresult (*enc_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint64_t sw128_desc(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}


// Whole-warp variants: every lane executes the instruction stream with warp-uniform operands and one elected lane issues, so
// the compiler can keep descriptors in uniform registers (no per-MMA ELECT / R2UR.BROADCAST waterfall).
__device__ __forceinline__ void umma_bf16_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
                 "l"(da), "l"(db), "r"(idesc), "r"(acc)
                 : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint64_t* bar) {
    asm volatile("{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(ptx::smem_u32(bar))
                 : "memory");
}

struct Params { int N, stages, k_blocks, tiles, sub, spin, whole_warp; };   // spin: warps 4-7 poll the final barrier like epilogue warps do   // sub = 64-element sub-blocks per stage (1 or 2)

__global__ void __launch_bounds__(256, 1) mainloop(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, Params p,
                                                   unsigned long long* out) {
    extern __shared__ __align__(1024) uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    const int a_bytes = 128 * 128 * p.sub, b_bytes = p.N * 128 * p.sub, stage_bytes = a_bytes + b_bytes;
    __shared__ __align__(8) uint64_t full[8], empty[8], done;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
        ptx::mbar_init(&done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<256>(&tmem_slot);
    ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    const int total = p.tiles * p.k_blocks;
    if (warp == 0 && lane == 0) {
        int stage = 0, phase = 0;
        for (int t = 0; t < p.tiles; ++t) {
            const int row0 = (blockIdx.x * p.tiles + t) * 128;
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                ptx::mbar_wait(&empty[stage], phase ^ 1);
                uint8_t* a = smem + (size_t)stage * stage_bytes;
                ptx::mbar_expect_tx(&full[stage], (uint32_t)stage_bytes);
                for (int u = 0; u < p.sub; ++u) {
                    ptx::tma_load_2d(a + u * 128 * 128, &tm_a, &full[stage], (kb * p.sub + u) * 64, row0);
                    ptx::tma_load_2d(a + a_bytes + u * p.N * 128, &tm_b, &full[stage], (kb * p.sub + u) * 64, 0);
                }
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1 && lane == 0 && !p.whole_warp) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        int stage = 0, phase = 0;
        long long t_wait = 0, t_mma = 0, t_commit = 0;
        const long long t0 = clock64();
        for (int i = 0; i < total; ++i) {
            const long long w0 = clock64();
            ptx::mbar_wait(&full[stage], phase);
            const long long w1 = clock64();
            ptx::tc_fence_after();
            const uint32_t a_addr = ptx::smem_u32(smem + (size_t)stage * stage_bytes);
            for (int u = 0; u < p.sub; ++u) {
                const uint64_t da = sw128_desc(a_addr + u * 128 * 128), db = sw128_desc(a_addr + a_bytes + u * p.N * 128);
                for (int k = 0; k < 4; ++k) ptx::umma_bf16(tmem, da + 2 * k, db + 2 * k, idesc, (i | u | k) ? 1u : 0u);
            }
            const long long w2 = clock64();
            ptx::umma_commit(&empty[stage]);
            const long long w3 = clock64();
            t_wait += w1 - w0; t_mma += w2 - w1; t_commit += w3 - w2;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        ptx::umma_commit(&done);
        ptx::mbar_wait(&done, 0);
        const long long t1 = clock64();
        out[blockIdx.x * 4] = (unsigned long long)(t1 - t0);
        out[blockIdx.x * 4 + 1] = (unsigned long long)t_wait;
        out[blockIdx.x * 4 + 2] = (unsigned long long)t_mma;
        out[blockIdx.x * 4 + 3] = (unsigned long long)t_commit;
    }
    else if (warp == 1 && p.whole_warp) {
        // the same loop run by all 32 lanes of the MMA warp
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        int stage = 0, phase = 0;
        long long t_wait = 0, t_mma = 0, t_commit = 0;
        const long long t0 = clock64();
        for (int i = 0; i < total; ++i) {
            const long long w0 = clock64();
            ptx::mbar_wait(&full[stage], phase);
            const long long w1 = clock64();
            ptx::tc_fence_after();
            const uint32_t a_addr = ptx::smem_u32(smem + (size_t)stage * stage_bytes);
            for (int u = 0; u < p.sub; ++u) {
                const uint64_t da = sw128_desc(a_addr + u * 128 * 128), db = sw128_desc(a_addr + a_bytes + u * p.N * 128);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16_elect(tmem, da + 2 * k, db + 2 * k, idesc, (i | u | k) ? 1u : 0u);
            }
            const long long w2 = clock64();
            umma_commit_elect(&empty[stage]);
            const long long w3 = clock64();
            t_wait += w1 - w0; t_mma += w2 - w1; t_commit += w3 - w2;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_elect(&done);
        ptx::mbar_wait(&done, 0);
        const long long t1 = clock64();
        if (lane == 0) {
            out[blockIdx.x * 4] = (unsigned long long)(t1 - t0);
            out[blockIdx.x * 4 + 1] = (unsigned long long)t_wait;
            out[blockIdx.x * 4 + 2] = (unsigned long long)t_mma;
            out[blockIdx.x * 4 + 3] = (unsigned long long)t_commit;
        }
    }
    else if (warp >= 4 && p.spin) {
        if (p.spin == 1) ptx::mbar_wait(&done, 0);                      // plain try_wait loop
        else { while (!ptx::mbar_try_wait(&done, 0)) __nanosleep(200); }   // polite: back off between polls
    }
    ptx::tc_fence_before(); __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<256>(tmem);
}

int main() {
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q); auto enc = (enc_tiled_fn)fp;
    const int K = 64 * 32, NMAX = 256, CTAS = 148, TILES = 4;
    const size_t rows = (size_t)CTAS * TILES * 128;
    __nv_bfloat16 *da, *db; unsigned long long* dout;
    cudaMalloc(&da, rows * K * 2); cudaMalloc(&db, (size_t)NMAX * K * 2); cudaMalloc(&dout, CTAS * 32);
    cudaMemset(da, 0, rows * K * 2); cudaMemset(db, 0, (size_t)NMAX * K * 2);
    cudaFuncSetAttribute(mainloop, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    std::vector<unsigned long long> ho(CTAS * 4);
    printf("%4s %4s %6s %6s\n", "N", "sub", "stages", "ctas");
    for (int N : {160})
        for (int sub : {1, 2})
            for (int stages : {4}) {
                const size_t smem = (size_t)stages * (128 * 128 + N * 128) * sub + 2048;
                if (smem > 215 * 1024) continue;
                CUtensorMap tma, tmb;
                { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)rows}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, 128}, e[2] = {1, 1};
                  enc(&tma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
                { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)NMAX}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, (cuuint32_t)N}, e[2] = {1, 1};
                  enc(&tmb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
                for (int spin : {0, 1})
                for (int ctas : {1, CTAS}) {
                    Params p{N, stages, 32 / sub, TILES, sub, 0, spin};
                    for (int rep = 0; rep < 2; ++rep) mainloop<<<ctas, 256, smem>>>(tma, tmb, p, dout);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("N %d sub %d stages %d: %s\n", N, sub, stages, cudaGetErrorString(e)); return 2; }
                    cudaMemcpy(ho.data(), dout, ctas * 32, cudaMemcpyDeviceToHost);
                    double tot = 0, wait = 0, mma = 0, com = 0;
                    for (int c = 0; c < ctas; ++c) { tot += (double)ho[4 * c]; wait += (double)ho[4 * c + 1]; mma += (double)ho[4 * c + 2]; com += (double)ho[4 * c + 3]; }
                    const double blocks = (double)ctas * TILES * 32;
                    printf("%4d %4d %6d %6d whole-warp %d | total %7.1f | wait full %7.1f | fence+desc+issue MMAs %7.1f | commit %7.1f   (cycles per K=64 block)\n", N, sub, stages, ctas, spin,
                           tot / blocks, wait / blocks, mma / blocks, com / blocks);
                }
            }
    return 0;
}
