// Probe: semantics of TMA im2col mode (cuTensorMapEncodeIm2col + cp.async.bulk.tensor.4d...im2col) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o probe tools/probe_tma_im2col.cu ; run on a B200.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

typedef CUresult (*encode_im2col_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int N = 2, H = 5, W = 6, C = 64, PIX = 32;

__global__ void probe(const __grid_constant__ CUtensorMap tmap, uint16_t* out, int c0, int w0, int h0, int n0,
                      int offw, int offh) {
    __shared__ __align__(128) uint16_t tile[PIX * C];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile);
    for (int i = threadIdx.x; i < PIX * C; i += blockDim.x) tile[i] = 0xEEEE;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(PIX * C * 2));
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
            "l"(&tmap), "r"(bar_a), "r"(c0), "r"(w0), "r"(h0), "r"(n0), "h"((uint16_t)offw), "h"((uint16_t)offh)
            : "memory");
    }
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(bar_a) : "memory");
    for (int i = threadIdx.x; i < PIX * C; i += blockDim.x) out[i] = tile[i];
}

int main() {
    std::vector<uint16_t> h((size_t)N * H * W * C);
    for (int n = 0; n < N; ++n) for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) for (int c = 0; c < C; ++c)
        h[(((size_t)n * H + y) * W + x) * C + c] = (uint16_t)(0x8000 | n << 12 | y << 8 | x << 4 | (c & 15));
    uint16_t *d, *o;
    cudaMalloc(&d, h.size() * 2); cudaMalloc(&o, PIX * C * 2);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fp, cudaEnableDefault, &q);
    if (!fp) { printf("no cuTensorMapEncodeIm2col\n"); return 1; }
    auto enc = (encode_im2col_fn)fp;
    struct Case { const char* name; int lo[2], up[2]; unsigned es[4]; int start[4]; int off[2]; };
    Case cases[] = {
        {"3x3 p1 s1 start(-1,-1,n0) off(0,0)", {-1, -1}, {-1, -1}, {1, 1, 1, 1}, {0, -1, -1, 0}, {0, 0}},
        {"3x3 p1 s1 start(-1,-1,n0) off(1,1)", {-1, -1}, {-1, -1}, {1, 1, 1, 1}, {0, -1, -1, 0}, {1, 1}},
        {"3x3 p1 s1 start(-1,-1,n0) off(2,0)", {-1, -1}, {-1, -1}, {1, 1, 1, 1}, {0, -1, -1, 0}, {2, 0}},
        {"3x3 p1 s1 start(w2,h1,n0) off(1,1)", {-1, -1}, {-1, -1}, {1, 1, 1, 1}, {0, 2, 1, 0}, {1, 1}},
        {"3x3 p1 s1 start(w0,h0,n0) off(1,1) c16", {-1, -1}, {-1, -1}, {1, 1, 1, 1}, {16, 0, 0, 0}, {1, 1}},
        {"3x3 p0 s2 start(0,0,n0) off(0,0)", {0, 0}, {-2, -2}, {1, 2, 2, 1}, {0, 0, 0, 0}, {0, 0}},
        {"3x3 p0 s2 start(0,0,n0) off(2,1)", {0, 0}, {-2, -2}, {1, 2, 2, 1}, {0, 0, 0, 0}, {2, 1}},
        {"3x3 p0 s2 start(w2,h0,n0) off(0,0)", {0, 0}, {-2, -2}, {1, 2, 2, 1}, {0, 2, 0, 0}, {0, 0}},
        {"1x7 p(0,3) s1 start(-3,0,n0) off(0,0)", {-3, 0}, {-3, 0}, {1, 1, 1, 1}, {0, -3, 0, 0}, {0, 0}},
        {"1x7 p(0,3) s1 start(-3,0,n0) off(6,0)", {-3, 0}, {-3, 0}, {1, 1, 1, 1}, {0, -3, 0, 0}, {6, 0}},
        {"1x1 p0 s1 start(0,0,n0) off(0,0)", {0, 0}, {0, 0}, {1, 1, 1, 1}, {0, 0, 0, 0}, {0, 0}},
        {"1x1 p0 s1 start(w4,h4,n0) off(0,0)", {0, 0}, {0, 0}, {1, 1, 1, 1}, {0, 4, 4, 0}, {0, 0}},
        {"1x1 p0 s1 start(w4,h4,n1) off(0,0) (runs past the tensor)", {0, 0}, {0, 0}, {1, 1, 1, 1}, {0, 4, 4, 1}, {0, 0}},
    };
    for (auto& cs : cases) {
        CUtensorMap tm;
        cuuint64_t dims[4] = {C, W, H, N};
        cuuint64_t strides[3] = {C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
        CUresult rc = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, d, dims, strides, cs.lo, cs.up, C, PIX, cs.es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("== %s : encode rc=%d\n", cs.name, (int)rc);
        if (rc) continue;
        cudaMemset(o, 0xAB, PIX * C * 2);
        probe<<<1, 128>>>(tm, o, cs.start[0], cs.start[1], cs.start[2], cs.start[3], cs.off[0], cs.off[1]);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("   kernel error: %s\n", cudaGetErrorString(e)); return 2; }
        std::vector<uint16_t> r(PIX * C);
        cudaMemcpy(r.data(), o, PIX * C * 2, cudaMemcpyDeviceToHost);
        printf("   ");
        for (int p = 0; p < PIX; ++p) {
            uint16_t v = r[p * C];
            if (v == 0) printf("[Z] ");
            else if (v == 0xEEEE) printf("[untouched] ");
            else printf("[n%d h%d w%d c%d] ", (v >> 12) & 7, (v >> 8) & 15, (v >> 4) & 15, v & 15);
        }
        printf("\n   channels of pixel 0: c0=%04x c1=%04x c63=%04x\n", r[0], r[1], r[63]);
    }
    return 0;
}
