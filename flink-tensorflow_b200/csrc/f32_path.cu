// f32_path.cu -- the label model's kernels in fp32 (ftf_config.compute_dtype = FTF_FLOAT).
//
// The reference runs its Inception graph in fp32 (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67 feeds a
// FLOAT image tensor into Session.run, flink-tensorflow/.../models/ModelFunction.scala:65); this mode reproduces that
// precision on the GPU: fp32 activations and weights, fp32 FMA accumulation in the oracle's K order (filter row, filter
// column, input channel), so logits agree with the fp32 oracle to accumulation rounding (tests: <= 1e-4 relative).
// It is the precision mode, not the fast one: plain CUDA-core implicit GEMM (64 pixels x 64 channels per CTA, 4 x 4
// outputs per thread, operands staged through shared memory), about 50x slower than the tcgen05 bf16 path.
#include "f32_path.h"
#include <math.h>

namespace ftf {
namespace f32 {

namespace {

constexpr int kBM = 64, kBN = 64, kBK = 16, kThreads = 256;

struct ConvArgs {
    const float* in; const float* w; const float* bias;
    int H, W, Cin, in_ld, Cout, KH, KW, stride, pad_h, pad_w, OH, OW;
    long long M;                        // output pixels
    int n_seg;
    ConvSegment seg[kConvMaxSegments];
};

// D[m, co] = sum over (kh, kw, c) of in[pixel(m) shifted by the tap, c] * w[co][kh][kw][c]
__global__ void __launch_bounds__(kThreads)
conv_f32_kernel(const __grid_constant__ ConvArgs a) {
    __shared__ float As[kBK][kBM + 4];          // [k][pixel], padded against bank conflicts of the transposing stores
    __shared__ float Bs[kBK][kBN + 4];          // [k][cout]
    const int tid = threadIdx.x;
    const long long m0 = (long long)blockIdx.x * kBM;
    const int n0 = blockIdx.y * kBN;
    const int tm = tid / 16, tn = tid % 16;     // this thread's 4 pixels (tm*4..) and 4 channels (tn*4..)
    // loader roles: thread -> (row = tid / 4, four consecutive k = (tid % 4) * 4 ..)
    const int lrow = tid >> 2, lk = (tid & 3) * 4;
    const long long lm = m0 + lrow;
    const bool lm_ok = lm < a.M;
    int l_img = 0, l_oy = 0, l_ox = 0;
    if (lm_ok) {
        l_img = (int)(lm / (a.OH * a.OW));
        const int rem = (int)(lm - (long long)l_img * a.OH * a.OW);
        l_oy = rem / a.OW; l_ox = rem - l_oy * a.OW;
    }
    const int l_co = n0 + lrow;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    for (int kh = 0; kh < a.KH; ++kh) {
        const int iy = l_oy * a.stride - a.pad_h + kh;
        for (int kw = 0; kw < a.KW; ++kw) {
            const int ix = l_ox * a.stride - a.pad_w + kw;
            const bool pix_ok = lm_ok && (unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W;
            const float* src = a.in + (((size_t)l_img * a.H + (pix_ok ? iy : 0)) * a.W + (pix_ok ? ix : 0)) * a.in_ld;
            const float* wsrc = a.w + (((size_t)(l_co < a.Cout ? l_co : 0) * a.KH + kh) * a.KW + kw) * a.Cin;
            for (int c0 = 0; c0 < a.Cin; c0 += kBK) {
                float av[4], bv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = c0 + lk + q;
                    av[q] = (pix_ok && c < a.Cin) ? __ldg(src + c) : 0.0f;
                    bv[q] = (l_co < a.Cout && c < a.Cin) ? __ldg(wsrc + c) : 0.0f;
                }
                __syncthreads();                 // the previous chunk has been consumed
#pragma unroll
                for (int q = 0; q < 4; ++q) { As[lk + q][lrow] = av[q]; Bs[lk + q][lrow] = bv[q]; }
                __syncthreads();
#pragma unroll
                for (int k = 0; k < kBK; ++k) {
                    const float4 x = *reinterpret_cast<const float4*>(&As[k][tm * 4]);
                    const float4 y = *reinterpret_cast<const float4*>(&Bs[k][tn * 4]);
                    const float xa[4] = {x.x, x.y, x.z, x.w}, ya[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xa[i], ya[j], acc[i][j]);
                }
            }
        }
    }
    // epilogue: bias, ReLU, into the segment that owns each output channel
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int gc = n0 + tn * 4 + j;
        if (gc >= a.Cout) continue;
        int s = 0;
        for (int q = 1; q < a.n_seg; ++q) if (a.seg[q].col_begin <= gc) s = q;
        const ConvSegment& sg = a.seg[s];
        const float b = (sg.flags & kSegBias) ? __ldg(a.bias + gc) : 0.0f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const long long m = m0 + tm * 4 + i;
            if (m >= a.M) continue;
            float v = acc[i][j] + b;
            if ((sg.flags & kSegRelu) && !(v > 0.0f)) v = 0.0f;
            ((float*)sg.out)[(size_t)m * sg.ld + sg.coff + (gc - sg.col_begin)] = v;
        }
    }
}

__global__ void __launch_bounds__(256)
maxpool3x3s2_f32(const float* __restrict__ in, float* __restrict__ out, int n, int H, int W, int C, int OH, int OW, int out_ld, int out_off) {
    const long long total = (long long)n * OH * OW * C;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(t % C);
        const long long pix = t / C;
        const int ox = (int)(pix % OW);
        const long long rest = pix / OW;
        const int oy = (int)(rest % OH), img = (int)(rest / OH);
        const float* base = in + (((size_t)img * H + 2 * oy) * W + 2 * ox) * C + c;
        float m = -INFINITY;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) m = fmaxf(m, __ldg(base + ((size_t)ky * W + kx) * C));
        out[(size_t)pix * out_ld + out_off + c] = m;
    }
}

// out = relu(avgpool3x3_s1_p1(pre) * (1/9) + bias), divisor 9 everywhere, taps summed row-major (the oracle's order)
__global__ void __launch_bounds__(256)
avgpool3x3_bias_relu_f32(const float* __restrict__ pre, const float* __restrict__ bias, float* __restrict__ out, int n, int H, int W, int C,
                         int out_ld, int out_off) {
    const long long total = (long long)n * H * W * C;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(t % C);
        const long long pix = t / C;
        const int x = (int)(pix % W);
        const int y = (int)((pix / W) % H);
        float s = 0.0f;
#pragma unroll
        for (int ky = -1; ky <= 1; ++ky)
#pragma unroll
            for (int kx = -1; kx <= 1; ++kx)
                if ((unsigned)(y + ky) < (unsigned)H && (unsigned)(x + kx) < (unsigned)W) s += __ldg(pre + ((size_t)pix + (ptrdiff_t)ky * W + kx) * C + c);
        const float v = __fadd_rn(__fmul_rn(s, 1.0f / 9.0f), __ldg(bias + c));
        out[(size_t)pix * out_ld + out_off + c] = v > 0.0f ? v : 0.0f;
    }
}

__global__ void __launch_bounds__(256)
global_avgpool_f32(const float* __restrict__ in, float* __restrict__ out, int n, int HW, int C) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * C) return;
    const int img = idx / C, c = idx - img * C;
    const float* src = in + (size_t)img * HW * C + c;
    float s = 0.0f;
    for (int p = 0; p < HW; ++p) s += __ldg(src + (size_t)p * C);
    out[idx] = __fmul_rn(s, 1.0f / (float)HW);
}

long long grid_cap() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (long long)device_sm_count(dev) * 16;
}

}  // namespace

int launch_conv(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int n, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    if (n_seg < 1 || n_seg > kConvMaxSegments) { if (err) *err = "conv_f32: segment count"; return FTF_INVALID_ARGUMENT; }
    ConvArgs a{};
    a.in = (const float*)g.in; a.w = (const float*)g.weights; a.bias = g.bias;
    a.H = g.H; a.W = g.W; a.Cin = g.Cin; a.in_ld = g.in_ld; a.Cout = g.Cout; a.KH = g.KH; a.KW = g.KW; a.stride = g.stride;
    a.pad_h = g.pad_h; a.pad_w = g.pad_w;
    a.OH = (g.H + 2 * g.pad_h - g.KH) / g.stride + 1; a.OW = (g.W + 2 * g.pad_w - g.KW) / g.stride + 1;
    if (a.OH <= 0 || a.OW <= 0) { if (err) *err = "conv_f32: empty output"; return FTF_INVALID_ARGUMENT; }
    a.M = (long long)n * a.OH * a.OW;
    a.n_seg = n_seg;
    for (int i = 0; i < n_seg; ++i) a.seg[i] = segs[i];
    const long long m_tiles = (a.M + kBM - 1) / kBM;
    if (m_tiles >= (1ll << 31)) { if (err) *err = "conv_f32: tensor too large"; return FTF_INVALID_ARGUMENT; }
    conv_f32_kernel<<<dim3((unsigned)m_tiles, (unsigned)((g.Cout + kBN - 1) / kBN)), kThreads, 0, s>>>(a);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

int launch_maxpool3x3s2(const float* in, float* out, int n, int H, int W, int C, int out_ld, int out_off, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    const int OH = (H - 3) / 2 + 1, OW = (W - 3) / 2 + 1;
    const long long total = (long long)n * OH * OW * C, want = (total + 255) / 256, cap = grid_cap();
    maxpool3x3s2_f32<<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>(in, out, n, H, W, C, OH, OW, out_ld, out_off);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, float* out, int n, int H, int W, int C, int out_ld, int out_off,
                                cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    const long long total = (long long)n * H * W * C, want = (total + 255) / 256, cap = grid_cap();
    avgpool3x3_bias_relu_f32<<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>(pre, bias, out, n, H, W, C, out_ld, out_off);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

int launch_global_avgpool(const float* in, float* out, int n, int HW, int C, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    global_avgpool_f32<<<(n * C + 255) / 256, 256, 0, s>>>(in, out, n, HW, C);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

}  // namespace f32
}  // namespace ftf
