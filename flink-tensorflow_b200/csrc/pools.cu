// pools.cu -- K4 max pool 3x3/s2, K5 (average pool 3x3/s1 + bias + ReLU) on a 1x1-convolution pre-activation,
// K6 global average pool.  The pooling nodes of the frozen Inception graph the reference runs
// (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67).  All HBM-bound: 16-byte accesses along
// the channel dimension, outputs written straight into the channel slice of the block's concat tensor.
#include "pools.h"
#include <cuda_bf16.h>

namespace ftf {

namespace {

// 3x3 max pool, stride 2, no padding; 8 channels (16 bytes) per thread; writes a channel slice.
__global__ void __launch_bounds__(256)
maxpool3x3s2(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int H, int W, int C, int OH, int OW,
             int out_ld, int out_off) {
    const int cg = C / 8;
    const long long total = (long long)n * OH * OW * cg;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int g = (int)(t % cg);
        const long long pix = t / cg;
        const int ox = (int)(pix % OW);
        const int oy = (int)((pix / OW) % OH);
        const int img = (int)(pix / ((long long)OW * OH));
        __nv_bfloat162 m[4];
        bool first = true;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + (((size_t)img * H + 2 * oy + ky) * W + 2 * ox + kx) * C + g * 8));
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
                for (int j = 0; j < 4; ++j) m[j] = first ? h[j] : __hmax2(m[j], h[j]);
                first = false;
            }
        *reinterpret_cast<uint4*>(out + (size_t)pix * out_ld + out_off + g * 8) = *reinterpret_cast<uint4*>(m);
    }
}

// out = bf16(relu(avgpool3x3_s1_p1(pre) * (1/9) + bias)); pre is the fp32 1x1-conv output [n,H,W,C]; taps summed
// row-major in fp32; divisor 9 at the borders too (count_include_pad).  4 channels per thread.
__global__ void __launch_bounds__(256)
avgpool3x3_bias_relu(const float* __restrict__ pre, const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int n, int H,
                     int W, int C, int out_ld, int out_off) {
    const int cg = C / 4;
    const long long total = (long long)n * H * W * cg;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int g = (int)(t % cg);
        const long long pix = t / cg;
        const int x = (int)(pix % W);
        const int y = (int)((pix / W) % H);
        const int img = (int)(pix / ((long long)W * H));
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int ky = -1; ky <= 1; ++ky)
#pragma unroll
            for (int kx = -1; kx <= 1; ++kx) {
                const int iy = y + ky, ix = x + kx;
                if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
                const float4 v = __ldg(reinterpret_cast<const float4*>(pre + (((size_t)img * H + iy) * W + ix) * C + g * 4));
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
        const float4 b = __ldg(reinterpret_cast<const float4*>(bias + g * 4));
        const float ninth = 1.0f / 9.0f;
        const float r0 = fmaxf(__fadd_rn(__fmul_rn(s.x, ninth), b.x), 0.f), r1 = fmaxf(__fadd_rn(__fmul_rn(s.y, ninth), b.y), 0.f);
        const float r2 = fmaxf(__fadd_rn(__fmul_rn(s.z, ninth), b.z), 0.f), r3 = fmaxf(__fadd_rn(__fmul_rn(s.w, ninth), b.w), 0.f);
        __nv_bfloat162 h0 = __floats2bfloat162_rn(r0, r1), h1 = __floats2bfloat162_rn(r2, r3);
        uint2 o = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
        *reinterpret_cast<uint2*>(out + (size_t)pix * out_ld + out_off + g * 4) = o;
    }
}

// global average pool over HW pixels: fp32 sum in pixel order, times 1/HW, rounded to bf16. One thread per (n, c).
__global__ void __launch_bounds__(256)
global_avgpool(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int HW, int C) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * C) return;
    const int img = idx / C, c = idx - img * C;
    float s = 0.0f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(in[((size_t)img * HW + p) * C + c]);
    out[idx] = __float2bfloat16_rn(__fmul_rn(s, 1.0f / (float)HW));
}

}  // namespace

static long long grid_cap(cudaStream_t) {
    int dev = 0;
    cudaGetDevice(&dev);
    return (long long)device_sm_count(dev) * 16;
}

int launch_maxpool3x3s2(const void* in, void* out, int n, int H, int W, int C, int out_ld, int out_off, cudaStream_t s,
                        std::string* err) {
    if (n <= 0) return FTF_OK;
    if (H < 3 || W < 3 || C % 8 || out_ld % 8 || out_off % 8) { if (err) *err = "maxpool3x3s2: H,W >= 3 and C, out_ld, out_off multiples of 8"; return FTF_INVALID_ARGUMENT; }
    const int OH = (H - 3) / 2 + 1, OW = (W - 3) / 2 + 1;
    const long long total = (long long)n * OH * OW * (C / 8);
    const long long want = (total + 255) / 256, cap = grid_cap(s);
    maxpool3x3s2<<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, H, W, C, OH, OW,
                                                                    out_ld, out_off);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, void* out, int n, int H, int W, int C, int out_ld,
                                int out_off, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    if (C % 4 || out_ld % 4 || out_off % 4) { if (err) *err = "avgpool3x3: C, out_ld, out_off multiples of 4"; return FTF_INVALID_ARGUMENT; }
    const long long total = (long long)n * H * W * (C / 4);
    const long long want = (total + 255) / 256, cap = grid_cap(s);
    avgpool3x3_bias_relu<<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>(pre, bias, (__nv_bfloat16*)out, n, H, W, C, out_ld, out_off);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

int launch_global_avgpool(const void* in, void* out, int n, int HW, int C, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    global_avgpool<<<(n * C + 255) / 256, 256, 0, s>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, HW, C);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

}  // namespace ftf
