// Probe: can a tcgen05.mma A-operand descriptor start at an arbitrary 128-byte row of a 128B-swizzled buffer that
// TMA filled, and what must the descriptor's base_offset field be?  (Needed to reuse one shared-memory halo
// buffer for every filter tap of a convolution.)  Also probes an im2col tensor map over the PADDED bounding box.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -I flink-tensorflow_b200/csrc -o tools/probe_umma_shift tools/probe_umma_shift.cu
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
typedef CUresult (*enc_im2col_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int R = 384;  // rows in the A buffer

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                                             float* out, int shift, int base_off, int sbo_bytes) {
    extern __shared__ __align__(1024) uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t* a_buf = smem;                       // R x 128 B
    uint8_t* b_buf = smem + R * 128;             // 64 x 128 B
    __shared__ __align__(8) uint64_t bar, mma_bar;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_init(&mma_bar, 1); ptx::fence_mbar_init(); }
    if (warp == 0) ptx::tmem_alloc<64>(&tmem_slot);
    ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        ptx::mbar_expect_tx(&bar, R * 128 + 64 * 128);
        for (int r = 0; r < R; r += 128) ptx::tma_load_2d(a_buf + r * 128, &tm_a, &bar, 0, r);
        ptx::tma_load_2d(b_buf, &tm_b, &bar, 0, 0);
        ptx::mbar_wait(&bar, 0);
        ptx::tc_fence_after();
        auto desc = [&](uint32_t addr, int boff, int sbo) {
            uint64_t d = 0;
            d |= (uint64_t)((addr & 0x3FFFF) >> 4);
            d |= (uint64_t)1 << 16;
            d |= (uint64_t)(sbo >> 4) << 32;
            d |= (uint64_t)1 << 46;
            d |= (uint64_t)(boff & 7) << 49;
            d |= (uint64_t)2 << 61;
            return d;
        };
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint64_t da = desc(ptx::smem_u32(a_buf) + shift * 128, base_off, sbo_bytes);
        const uint64_t db = desc(ptx::smem_u32(b_buf), 0, 1024);
        for (int k = 0; k < 4; ++k) ptx::umma_bf16(tmem, da + 2 * k, db + 2 * k, idesc, k ? 1u : 0u);
        ptx::umma_commit(&mma_bar);
    }
    ptx::mbar_wait(&mma_bar, 0);
    ptx::tc_fence_after();
    const int row = threadIdx.x;
    for (int c = 0; c < 64; c += 32) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c, v);
        ptx::tmem_ld_wait();
        for (int i = 0; i < 32; ++i) out[row * 64 + c + i] = __uint_as_float(v[i]);
    }
    ptx::tc_fence_before(); __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<64>(tmem);
}

int main() {
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q); auto enc = (enc_tiled_fn)fp;
    cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fp, cudaEnableDefault, &q); auto enc_i = (enc_im2col_fn)fp;
    // A[r][k] = r + k/64 (exact in bf16 for r < 256: use r%200 + small), B = identity
    std::vector<__nv_bfloat16> ha((size_t)R * 64), hb(64 * 64);
    for (int r = 0; r < R; ++r) for (int k = 0; k < 64; ++k) ha[r * 64 + k] = __float2bfloat16((float)((r * 7 + k * 3) % 251));
    for (int n = 0; n < 64; ++n) for (int k = 0; k < 64; ++k) hb[n * 64 + k] = __float2bfloat16(n == k ? 1.0f : 0.0f);
    __nv_bfloat16 *da, *db; float* dout;
    cudaMalloc(&da, ha.size() * 2); cudaMalloc(&db, hb.size() * 2); cudaMalloc(&dout, 128 * 64 * 4);
    cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
    CUtensorMap tma, tmb;
    { cuuint64_t d[2] = {64, R}, s[1] = {128}; cuuint32_t b[2] = {64, 128}, e[2] = {1, 1};
      printf("enc a %d\n", (int)enc(&tma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
    { cuuint64_t d[2] = {64, 64}, s[1] = {128}; cuuint32_t b[2] = {64, 64}, e[2] = {1, 1};
      printf("enc b %d\n", (int)enc(&tmb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)); }
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    std::vector<float> ho(128 * 64);
    int shifts[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 16, 37, 149, 255};
    for (int sh : shifts) {
        for (int mode = 0; mode < 2; ++mode) {
            const int boff = mode == 0 ? 0 : (sh & 7);
            cudaMemset(dout, 0xFF, 128 * 64 * 4);
            probe<<<1, 128, R * 128 + 64 * 128 + 2048>>>(tma, tmb, dout, sh, boff, 1024);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("shift %d boff %d: kernel error %s\n", sh, boff, cudaGetErrorString(e)); return 2; }
            cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost);
            int bad = 0, first = -1;
            for (int i = 0; i < 128; ++i) for (int n = 0; n < 64; ++n) {
                const float want = (float)(((sh + i) * 7 + n * 3) % 251);
                if (ho[i * 64 + n] != want) { if (first < 0) first = i * 64 + n; ++bad; }
            }
            printf("shift %3d base_offset %d : %s (%d mismatches, first at row %d col %d: got %.0f)\n", sh, boff, bad ? "MISMATCH" : "ok", bad,
                   first < 0 ? -1 : first / 64, first < 0 ? -1 : first % 64, first < 0 ? 0.f : ho[first]);
        }
    }
    // im2col over the padded bounding box: lower = -pad, upper = +pad, "1x1 filter"
    {
        const int N = 2, H = 4, W = 5, C = 64;
        std::vector<uint16_t> h((size_t)N * H * W * C);
        for (int n = 0; n < N; ++n) for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) for (int c = 0; c < C; ++c)
            h[(((size_t)n * H + y) * W + x) * C + c] = (uint16_t)(0x8000 | n << 12 | y << 8 | x << 4 | (c & 15));
        uint16_t* d; cudaMalloc(&d, h.size() * 2); cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
        CUtensorMap tm; cuuint64_t dims[4] = {C, W, H, N}; cuuint64_t st[3] = {C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
        int lo[2] = {-1, -1}, up[2] = {1, 1}; cuuint32_t es[4] = {1, 1, 1, 1};
        CUresult rc = enc_i(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, d, dims, st, lo, up, C, 64, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("im2col padded-box encode (lower -1, upper +1): rc=%d\n", (int)rc);
    }
    return 0;
}
