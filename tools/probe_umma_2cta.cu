// Probe: tcgen05.mma with cta_group::2 (a CTA pair: M = 256, each CTA holds its own 128 rows of A and HALF of the rows of B).
// Checks the operand/accumulator placement against a CPU product and measures the issue cadence per MMA for
// cta_group::1 (M = 128) and cta_group::2 (M = 256) at several N.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -I flink-tensorflow_b200/csrc -o tools/probe_umma_2cta tools/probe_umma_2cta.cu
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

constexpr uint32_t kPeerMask = 0xFEFFFFFFu;      // clears the CTA-rank bit of a shared::cluster address: -> CTA 0 of the pair

__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                     ptx::smem_u32(dst)),
                 "l"(m), "r"(ptx::smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
                 "l"(da), "l"(db), "r"(idesc), "r"(acc)
                 : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     ptx::smem_u32(bar)),
                 "h"(mask)
                 : "memory");
}
template <int kCols> __device__ __forceinline__ void tmem_alloc_2sm(uint32_t* slot) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(slot)), "n"(kCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int kCols> __device__ __forceinline__ void tmem_dealloc_2sm(uint32_t addr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "n"(kCols) : "memory");
}

__device__ __forceinline__ uint64_t sw128_desc(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// CG = 1: one CTA, M = 128, all N rows of B.  CG = 2: a CTA pair, M = 256, N/2 rows of B per CTA.
template <int CG>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, float* out,
                                             int N, int reps, unsigned long long* cycles, int commit_every, int a_shift_rows, int sw64) {
    extern __shared__ __align__(1024) uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t* a_buf = smem;                        // 128 x 128 B
    uint8_t* b_buf = smem + 128 * 128;            // N/CG x 128 B
    __shared__ __align__(8) uint64_t full, mma_bar, dummy_bar;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    const uint32_t rank = CG == 2 ? cluster_rank() : 0;
    if (threadIdx.x == 0) { ptx::mbar_init(&full, 1); ptx::mbar_init(&mma_bar, 1); ptx::mbar_init(&dummy_bar, 1 << 20); ptx::fence_mbar_init(); }
    if (warp == 0) { if (CG == 2) tmem_alloc_2sm<256>(&tmem_slot); else ptx::tmem_alloc<256>(&tmem_slot); }
    ptx::tc_fence_before();
    if (CG == 2) cluster_sync(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    const int nb = N / CG;                         // rows of B in this CTA
    if (threadIdx.x == 0) {
        const uint32_t bytes = 128 * 128 + (uint32_t)nb * 128;
        if (CG == 2) {
            if (rank == 0) ptx::mbar_expect_tx(&full, 2 * bytes);      // both CTAs' loads complete on the leader's barrier
            tma_load_2d_2sm(a_buf, &tm_a, &full, 0, (int)rank * 128);
            tma_load_2d_2sm(b_buf, &tm_b, &full, 0, (int)rank * nb);
        } else {
            ptx::mbar_expect_tx(&full, bytes);
            ptx::tma_load_2d(a_buf, &tm_a, &full, 0, 0);
            ptx::tma_load_2d(b_buf, &tm_b, &full, 0, 0);
        }
        if (rank == 0) {
            ptx::mbar_wait(&full, 0);
            ptx::tc_fence_after();
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)((128 * CG) >> 4) << 24);
            uint64_t da = sw128_desc(ptx::smem_u32(a_buf) + (uint32_t)a_shift_rows * 128u), db = sw128_desc(ptx::smem_u32(b_buf));
            if (sw64) {   // 64-byte rows: 8-row groups 512 bytes apart, layout type SWIZZLE_64B; two K = 16 steps per row
                auto d64 = [](uint32_t addr) { uint64_t d = 0; d |= (uint64_t)((addr & 0x3FFFF) >> 4); d |= (uint64_t)1 << 16; d |= (uint64_t)(512 >> 4) << 32; d |= (uint64_t)1 << 46; d |= (uint64_t)4 << 61; return d; };
                da = d64(ptx::smem_u32(a_buf) + (uint32_t)a_shift_rows * 64u); db = d64(ptx::smem_u32(b_buf));
            }
            const long long t0 = clock64();
            for (int r = 0; r < reps; ++r) {
                for (int k = 0; k < 4; ++k) {
                    const int kk = sw64 ? (k & 1) : k;            // a 64-byte row holds two K steps
                    if (CG == 2) umma_bf16_2sm(tmem, da + 2 * kk, db + 2 * kk, idesc, (r | k) ? 1u : 0u);
                    else ptx::umma_bf16(tmem, da + 2 * kk, db + 2 * kk, idesc, (r | k) ? 1u : 0u);
                }
                // as a pipelined main loop does: a commit per K block (here onto a barrier nobody waits for)
                if (commit_every && (r + 1) % commit_every == 0) { if (CG == 2) umma_commit_2sm(&dummy_bar, 0b11); else ptx::umma_commit(&dummy_bar); }
            }
            if (CG == 2) umma_commit_2sm(&mma_bar, 0b11); else ptx::umma_commit(&mma_bar);
            ptx::mbar_wait(&mma_bar, 0);
            if (cycles) *cycles = (unsigned long long)(clock64() - t0);
        }
    }
    ptx::mbar_wait(&mma_bar, 0);
    ptx::tc_fence_after();
    const int row = threadIdx.x;
    for (int c = 0; c < N; c += 32) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
        ptx::tmem_ld_wait();
        for (int i = 0; i < 32; ++i) out[((size_t)rank * 128 + row) * N + c + i] = __uint_as_float(v[i]);
    }
    ptx::tc_fence_before();
    if (CG == 2) cluster_sync(); else __syncthreads();
    if (warp == 0) { if (CG == 2) tmem_dealloc_2sm<256>(tmem); else ptx::tmem_dealloc<256>(tmem); }
}

template <int CG>
static cudaError_t launch(const CUtensorMap& tma, const CUtensorMap& tmb, float* out, int N, int reps, unsigned long long* cyc, int commit_every = 0, int a_shift_rows = 0, int sw64 = 0) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(CG); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 128 * 128 + 256 * 128 + 2048;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaFuncSetAttribute(probe<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.dynamicSmemBytes);
    cudaError_t e = cudaLaunchKernelEx(&cfg, probe<CG>, tma, tmb, out, N, reps, cyc, commit_every, a_shift_rows, sw64);
    if (e != cudaSuccess) return e;
    return cudaDeviceSynchronize();
}

int main() {
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q); auto enc = (enc_tiled_fn)fp;
    const int M = 256, K = 64, NMAX = 256;
    std::vector<__nv_bfloat16> ha((size_t)M * K), hb((size_t)NMAX * K);
    std::vector<float> fa((size_t)M * K), fb((size_t)NMAX * K);
    for (int r = 0; r < M; ++r) for (int k = 0; k < K; ++k) { fa[r * K + k] = (float)((r * 7 + k * 3) % 13 - 6); ha[r * K + k] = __float2bfloat16(fa[r * K + k]); }
    for (int n = 0; n < NMAX; ++n) for (int k = 0; k < K; ++k) { fb[n * K + k] = (float)((n * 5 + k * 11) % 7 - 3); hb[n * K + k] = __float2bfloat16(fb[n * K + k]); }
    __nv_bfloat16 *da, *db; float* dout; unsigned long long* dcyc;
    cudaMalloc(&da, ha.size() * 2); cudaMalloc(&db, hb.size() * 2); cudaMalloc(&dout, (size_t)M * NMAX * 4); cudaMalloc(&dcyc, 8);
    cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
    std::vector<float> ho((size_t)M * NMAX);
    int ns[] = {64, 128, 160, 192, 256};
    for (int N : ns) {
        for (int cg = 1; cg <= 2; ++cg) {
            CUtensorMap tma, tmb;
            { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)M}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, 128}, e[2] = {1, 1};
              enc(&tma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
            { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)NMAX}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, (cuuint32_t)(N / cg)}, e[2] = {1, 1};
              enc(&tmb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
            // correctness: one pass over K = 64
            cudaMemset(dout, 0xFF, (size_t)M * NMAX * 4);
            cudaError_t e = cg == 1 ? launch<1>(tma, tmb, dout, N, 1, dcyc) : launch<2>(tma, tmb, dout, N, 1, dcyc);
            if (e != cudaSuccess) { printf("N %d cta_group::%d: kernel error %s\n", N, cg, cudaGetErrorString(e)); return 2; }
            cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost);
            int bad = 0, first = -1;
            const int rows = 128 * cg;
            for (int i = 0; i < rows; ++i) for (int n = 0; n < N; ++n) {
                float want = 0; for (int k = 0; k < K; ++k) want += fa[i * K + k] * fb[n * K + k];
                if (ho[(size_t)i * N + n] != want) { if (first < 0) first = i * N + n; ++bad; }
            }
            // cadence: 256 passes = 1024 MMAs back to back
            unsigned long long cyc = 0;
            e = cg == 1 ? launch<1>(tma, tmb, dout, N, 256, dcyc) : launch<2>(tma, tmb, dout, N, 256, dcyc);
            if (e != cudaSuccess) { printf("N %d cta_group::%d: timing kernel error %s\n", N, cg, cudaGetErrorString(e)); return 2; }
            cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
            unsigned long long cyc1 = 0, cyc4 = 0;
            e = cg == 1 ? launch<1>(tma, tmb, dout, N, 256, dcyc, 1) : launch<2>(tma, tmb, dout, N, 256, dcyc, 1);
            if (e != cudaSuccess) { printf("commit-per-block kernel error %s\n", cudaGetErrorString(e)); return 2; }
            cudaMemcpy(&cyc1, dcyc, 8, cudaMemcpyDeviceToHost);
            e = cg == 1 ? launch<1>(tma, tmb, dout, N, 256, dcyc, 4) : launch<2>(tma, tmb, dout, N, 256, dcyc, 4);
            cudaMemcpy(&cyc4, dcyc, 8, cudaMemcpyDeviceToHost);
            printf("N %3d cta_group::%d (M %3d): %s (%d mismatches%s)  cycles per MMA: %.1f with one commit at the end, %.1f with a commit per 4 MMAs, %.1f per 16 MMAs\n", N, cg, rows,
                   bad ? "MISMATCH" : "ok", bad, first >= 0 ? ", first shown below" : "", (double)cyc / 1024.0, (double)cyc1 / 1024.0, (double)cyc4 / 1024.0);
            if (first >= 0) printf("   first mismatch at row %d col %d: got %.1f\n", first / N, first % N, ho[first]);
        }
    }
    // 64-byte operand rows (SWIZZLE_64B, 32-channel K blocks): the timing only (buffers filled by the 128B maps, values irrelevant)
    for (int N : {32, 64, 160}) {
        CUtensorMap tma, tmb;
        { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)M}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, 128}, e[2] = {1, 1};
          enc(&tma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)NMAX}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, (cuuint32_t)N}, e[2] = {1, 1};
          enc(&tmb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        for (int shift : {0, 3}) {
            unsigned long long cyc = 0;
            cudaError_t e = launch<1>(tma, tmb, dout, N, 256, dcyc, 0, shift, 1);
            if (e != cudaSuccess) { printf("sw64 probe: %s\n", cudaGetErrorString(e)); return 2; }
            cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
            printf("N %3d, SWIZZLE_64B operands, A starting %d rows in: %.1f cycles per MMA\n", N, shift, (double)cyc / 1024.0);
        }
    }
    // an A descriptor that starts at a row which is not a multiple of 8 (the halo-window trick): does the MMA slow down?
    for (int N : {32, 64, 160}) {
        CUtensorMap tma, tmb;
        { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)M}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, 128}, e[2] = {1, 1};
          enc(&tma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        { cuuint64_t d[2] = {(cuuint64_t)K, (cuuint64_t)NMAX}, s[1] = {(cuuint64_t)K * 2}; cuuint32_t b[2] = {64, (cuuint32_t)N}, e[2] = {1, 1};
          enc(&tmb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
        for (int shift : {0, 1, 3, 8, 37}) {
            unsigned long long cyc = 0;
            cudaError_t e = launch<1>(tma, tmb, dout, N, 256, dcyc, 0, shift);
            if (e != cudaSuccess) { printf("shift probe: %s\n", cudaGetErrorString(e)); return 2; }
            cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
            printf("N %3d, A descriptor starting %2d rows into the swizzled buffer: %.1f cycles per MMA\n", N, shift, (double)cyc / 1024.0);
        }
    }
    return 0;
}
