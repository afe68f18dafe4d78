// pools.cu -- K4 max pool 3x3/s2, K5 (average pool 3x3/s1 + bias + ReLU) on a 1x1-convolution pre-activation,
// K6 global average pool.  The pooling nodes of the frozen Inception graph the reference runs
// (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67).  All HBM-bound: 16-byte accesses along
// the channel dimension, outputs written straight into the channel slice of the block's concat tensor.
#include "pools.h"
#include "ptx.cuh"
#include <cuda_bf16.h>

namespace ftf {

namespace {

// 3x3 max pool, stride 2, no padding; 8 channels (16 bytes) per thread; writes a channel slice.  32-bit index
// arithmetic (the launcher rejects tensors of 2^31 items), the nine loads issued together.
__global__ void __launch_bounds__(256)
maxpool3x3s2(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int H, int W, int C, int OH, int OW,
             int out_ld, int out_off, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const unsigned cg = (unsigned)C / 8u;
    const unsigned total = (unsigned)n * OH * OW * cg;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const unsigned g = t % cg, pix = t / cg;
        const unsigned ox = pix % (unsigned)OW, rest = pix / (unsigned)OW;
        const unsigned oy = rest % (unsigned)OH, img = rest / (unsigned)OH;
        const __nv_bfloat16* base = in + (((size_t)img * H + 2 * oy) * W + 2 * ox) * C + g * 8;
        uint4 v[9];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) v[ky * 3 + kx] = __ldg(reinterpret_cast<const uint4*>(base + ((size_t)ky * W + kx) * C));
        __nv_bfloat162 m[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = reinterpret_cast<const __nv_bfloat162*>(&v[0])[j];
#pragma unroll
        for (int k = 1; k < 9; ++k)
#pragma unroll
            for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], reinterpret_cast<const __nv_bfloat162*>(&v[k])[j]);
        *reinterpret_cast<uint4*>(out + (size_t)pix * out_ld + out_off + g * 8) = *reinterpret_cast<uint4*>(m);
    }
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

// out = bf16(relu(avgpool3x3_s1_p1(pre) * (1/9) + bias)); pre is the fp32 1x1-conv output [n,H,W,C]; taps summed
// row-major in fp32; divisor 9 at the borders too (count_include_pad).  4 channels per thread; the nine loads are
// issued together, out-of-image taps read the centre pixel and contribute +0.
__global__ void __launch_bounds__(256)
avgpool3x3_bias_relu(const float* __restrict__ pre, const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int n, int H,
                     int W, int C, int out_ld, int out_off, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const unsigned cg = (unsigned)C / 4u;
    const unsigned total = (unsigned)n * H * W * cg;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const unsigned g = t % cg, pix = t / cg;
        const int x = (int)(pix % (unsigned)W);
        const unsigned rest = pix / (unsigned)W;
        const int y = (int)(rest % (unsigned)H);
        const float* centre = pre + (size_t)pix * C + g * 4;
        float4 v[9];
#pragma unroll
        for (int ky = -1; ky <= 1; ++ky)
#pragma unroll
            for (int kx = -1; kx <= 1; ++kx) {
                const bool ok = (unsigned)(y + ky) < (unsigned)H && (unsigned)(x + kx) < (unsigned)W;
                const float4 r = __ldg(reinterpret_cast<const float4*>(ok ? centre + ((ptrdiff_t)ky * W + kx) * C : centre));
                v[(ky + 1) * 3 + kx + 1] = ok ? r : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 9; ++k) { s.x += v[k].x; s.y += v[k].y; s.z += v[k].z; s.w += v[k].w; }
        const float4 b = __ldg(reinterpret_cast<const float4*>(bias + g * 4));
        const float ninth = 1.0f / 9.0f;
        const float r0 = fmaxf(__fadd_rn(__fmul_rn(s.x, ninth), b.x), 0.f), r1 = fmaxf(__fadd_rn(__fmul_rn(s.y, ninth), b.y), 0.f);
        const float r2 = fmaxf(__fadd_rn(__fmul_rn(s.z, ninth), b.z), 0.f), r3 = fmaxf(__fadd_rn(__fmul_rn(s.w, ninth), b.w), 0.f);
        __nv_bfloat162 h0 = __floats2bfloat162_rn(r0, r1), h1 = __floats2bfloat162_rn(r2, r3);
        uint2 o = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
        *reinterpret_cast<uint2*>(out + (size_t)pix * out_ld + out_off + g * 4) = o;
    }
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

// global average pool over HW pixels: fp32 sum in pixel order, times 1/HW, rounded to bf16. One thread per (n, 2 channels).
__global__ void __launch_bounds__(256)
global_avgpool(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int HW, int C, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const int c2 = C / 2;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * c2) { ptx::tl_stamp_max(tl ? tl + 1 : nullptr); return; }
    const int img = idx / c2, c = (idx - img * c2) * 2;
    const __nv_bfloat16* src = in + (size_t)img * HW * C + c;
    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll 8
    for (int p = 0; p < HW; ++p) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(src + (size_t)p * C);
        s0 += __low2float(v);
        s1 += __high2float(v);
    }
    const float inv = 1.0f / (float)HW;
    *reinterpret_cast<__nv_bfloat162*>(out + (size_t)img * C + c) = __floats2bfloat162_rn(__fmul_rn(s0, inv), __fmul_rn(s1, inv));
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

}  // namespace

// ctas_per_sm: 16 fills an otherwise idle GPU; a pooling kernel that runs BESIDE tensor-core convolutions (the pool branch of
// an Inception block, on its own stream) is launched with 1: a convolution CTA needs three quarters of an SM's registers, and
// a pool kernel with thousands of small CTAs floods every SM a convolution has just left, so that the next convolution's CTAs
// wait for the flood to drain (measured: 7-9 us per block boundary in the 17x17 blocks).
// An SM's split between L1 and shared memory is fixed while any CTA runs on it.  A pooling kernel that asks for no shared
// memory gets SMs configured for a large L1, and a convolution CTA (200 KB of shared memory) cannot start on such an SM until
// every pool CTA has left it -- nor can a pool CTA join an SM a convolution has configured.  Measured in the Inception blocks:
// the convolutions behind a pool kernel on a sibling stream started only when the pool kernel had finished, 7-15 us late per
// block.  Asking for the maximum shared-memory carve-out gives both kinds of kernel the same configuration, so they share SMs.
template <typename K>
static void prefer_max_smem(K kernel) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaGetLastError();
}
static long long grid_cap(int ctas_per_sm) {
    int dev = 0;
    cudaGetDevice(&dev);
    return (long long)device_sm_count(dev) * (ctas_per_sm > 0 ? ctas_per_sm : 16);
}

int launch_maxpool3x3s2(const void* in, void* out, int n, int H, int W, int C, int out_ld, int out_off, cudaStream_t s,
                        std::string* err, unsigned long long* tl, int ctas_per_sm) {
    if (n <= 0) return FTF_OK;
    if (H < 3 || W < 3 || C % 8 || out_ld % 8 || out_off % 8) { if (err) *err = "maxpool3x3s2: H,W >= 3 and C, out_ld, out_off multiples of 8"; return FTF_INVALID_ARGUMENT; }
    const int OH = (H - 3) / 2 + 1, OW = (W - 3) / 2 + 1;
    const long long total = (long long)n * OH * OW * (C / 8);
    if (total >= (1ll << 31)) { if (err) *err = "maxpool3x3s2: tensor too large"; return FTF_INVALID_ARGUMENT; }
    const long long want = (total + 255) / 256, cap = grid_cap(ctas_per_sm);
    prefer_max_smem(maxpool3x3s2);
    FTF_CUDA(launch_pdl(maxpool3x3s2, dim3((unsigned)(want < cap ? want : cap)), dim3(256), 0, s, (const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, H, W, C,
                        OH, OW, out_ld, out_off, tl), err);
    count_launch();
    return FTF_OK;
}

int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, void* out, int n, int H, int W, int C, int out_ld,
                                int out_off, cudaStream_t s, std::string* err, unsigned long long* tl, int ctas_per_sm) {
    if (n <= 0) return FTF_OK;
    if (C % 4 || out_ld % 4 || out_off % 4) { if (err) *err = "avgpool3x3: C, out_ld, out_off multiples of 4"; return FTF_INVALID_ARGUMENT; }
    const long long total = (long long)n * H * W * (C / 4);
    if (total >= (1ll << 31)) { if (err) *err = "avgpool3x3: tensor too large"; return FTF_INVALID_ARGUMENT; }
    const long long want = (total + 255) / 256, cap = grid_cap(ctas_per_sm);
    prefer_max_smem(avgpool3x3_bias_relu);
    FTF_CUDA(launch_pdl(avgpool3x3_bias_relu, dim3((unsigned)(want < cap ? want : cap)), dim3(256), 0, s, pre, bias, (__nv_bfloat16*)out, n, H, W, C,
                        out_ld, out_off, tl), err);
    count_launch();
    return FTF_OK;
}

int launch_global_avgpool(const void* in, void* out, int n, int HW, int C, cudaStream_t s, std::string* err, unsigned long long* tl) {
    if (n <= 0) return FTF_OK;
    if (C % 2) { if (err) *err = "global_avgpool: C must be even"; return FTF_INVALID_ARGUMENT; }
    FTF_CUDA(launch_pdl(global_avgpool, dim3((n * (C / 2) + 255) / 256), dim3(256), 0, s, (const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, HW, C, tl), err);
    count_launch();
    return FTF_OK;
}

}  // namespace ftf
