// pools.cu -- K4 max pool 3x3/s2, K5 (average pool 3x3/s1 + bias + ReLU) on a 1x1-convolution pre-activation,
// K6 global average pool.  The pooling nodes of the frozen Inception graph the reference runs
// (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67).  Inputs staged tile by tile with TMA (K4, K5), 16-byte
// accesses along the channel dimension, outputs written straight into the channel slice of the block's concat tensor.
#include "pools.h"
#include "ptx.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdlib>

namespace ftf {

namespace {

// 3x3 max pool, stride 2, no padding; 8 channels (16 bytes) per thread; writes a channel slice.  32-bit index
// arithmetic (the launcher rejects tensors of 2^31 items), the nine loads issued together.
__global__ void __launch_bounds__(256)
maxpool3x3s2(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int H, int W, int C, int OH, int OW,
             int out_ld, int out_off, int reverse, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const unsigned cg = (unsigned)C / 8u;
    const unsigned total = (unsigned)n * OH * OW * cg;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned t_ = blockIdx.x * blockDim.x + threadIdx.x; t_ < total; t_ += stride) {
        const unsigned t = reverse ? total - 1u - t_ : t_;      // from the last image to the first (Op::reverse)
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

// The same pool with its input staged by TMA: a CTA owns TR x TC outputs of one image and one chunk of 64 (or 32) channels,
// loads the (2 TR + 1) x (2 TC + 1) input pixels under them with ONE 4-D tensor load, and takes every window from shared
// memory.  One thread per output re-reads every input row and column that two windows share from L2 (the rows always: the
// outputs of the next row belong to another CTA; 2.25 x the tensor in the worst case), and a pool that runs right after the
// convolution that wrote its 177 MB input is bound by exactly that L2 -> SM traffic; the tile re-reads only its one-pixel
// rim (1.16 x).  Edge tiles reach outside the image: TMA fills zeros there and those outputs are not stored.
__global__ void __launch_bounds__(256)
maxpool3x3s2_tma(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out, int OH, int OW, int TR, int TC, int tiles_x,
                 int tiles_y, int chunks, int cgc, int out_ld, int out_off, unsigned long long* tl) {
    extern __shared__ __align__(128) unsigned char pool_smem[];
    __shared__ uint64_t bar;
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    unsigned b = blockIdx.x;
    const int chunk = (int)(b % (unsigned)chunks); b /= (unsigned)chunks;
    const int tx = (int)(b % (unsigned)tiles_x); b /= (unsigned)tiles_x;
    const int ty = (int)(b % (unsigned)tiles_y);
    const int img = (int)(b / (unsigned)tiles_y);
    const int IR = 2 * TR + 1, IC = 2 * TC + 1;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    __syncthreads();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    if (threadIdx.x == 0) {
        ptx::mbar_expect_tx(&bar, (uint32_t)(IR * IC * cgc * 16));
        ptx::tma_load_4d(pool_smem, &tm, &bar, chunk * cgc * 8, tx * TC * 2, ty * TR * 2, img);
    }
    ptx::mbar_wait(&bar, 0);
    const int items = TR * TC * cgc;
    const uint4* tile = reinterpret_cast<const uint4*>(pool_smem);
    for (int i = threadIdx.x; i < items; i += 256) {
        const int g = i % cgc, p = i / cgc;
        const int oxl = p % TC, oyl = p / TC;
        const int oy = ty * TR + oyl, ox = tx * TC + oxl;
        if (oy >= OH || ox >= OW) continue;
        const uint4* base = tile + ((2 * oyl) * IC + 2 * oxl) * cgc + g;
        uint4 v[9];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) v[ky * 3 + kx] = base[(ky * IC + kx) * cgc];
        __nv_bfloat162 m[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = reinterpret_cast<const __nv_bfloat162*>(&v[0])[j];
#pragma unroll
        for (int k = 1; k < 9; ++k)
#pragma unroll
            for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], reinterpret_cast<const __nv_bfloat162*>(&v[k])[j]);
        *reinterpret_cast<uint4*>(out + (((size_t)img * OH + oy) * OW + ox) * out_ld + out_off + (chunk * cgc + g) * 8) = *reinterpret_cast<uint4*>(m);
    }
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

// out = bf16(relu(avgpool3x3_s1_p1(pre) * (1/9) + bias)); pre is the fp32 1x1-conv output [n,H,W,C]; taps summed
// row-major in fp32; divisor 9 at the borders too (count_include_pad); out-of-image taps contribute +0.
// A thread owns 4 channels of one image column and walks R output rows down it with the three input rows of the window in
// registers: every step loads ONE new row (3 x 16 bytes) instead of nine taps, and the x-neighbours of a column are the
// adjacent threads' columns, so those loads hit L1.  One thread per output re-read every input nine times from L2 (a CTA's
// 256 threads covered five pixels of one row; the rows above and below belonged to CTAs on other SMs): 18-23 us for a
// 14-20 MB tensor, and inside an Inception block that is time the next convolutions of the sibling branches wait for.
template <int R, bool kSlide>
__global__ void __launch_bounds__(128)
avgpool3x3_bias_relu(const float* __restrict__ pre, const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int n, int H,
                     int W, int C, int out_ld, int out_off, int strips, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const unsigned cg = (unsigned)C / 4u;
    const unsigned total = (unsigned)n * strips * W * cg;
    const unsigned stride = gridDim.x * blockDim.x;
    const float ninth = 1.0f / 9.0f;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const unsigned g = t % cg, r1 = t / cg;
        const int x = (int)(r1 % (unsigned)W);
        const unsigned r2 = r1 / (unsigned)W;
        const int y0 = (int)(r2 % (unsigned)strips) * R;
        const unsigned img = r2 / (unsigned)strips;
        const float* col = pre + ((size_t)img * H * W + x) * C + g * 4;          // row y of this column: + y * W * C
        __nv_bfloat16* ocol = out + ((size_t)img * H * W + x) * out_ld + out_off + g * 4;
        const bool xl = x > 0, xr = x + 1 < W;
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 b = __ldg(reinterpret_cast<const float4*>(bias + g * 4));
        auto load_row = [&](int y, float4* d) {       // refused taps read the column's first pixel and count as zero
            const bool ok = (unsigned)y < (unsigned)H;
            const float* p = col + (ptrdiff_t)(ok ? y : 0) * W * C;
            const float4 l = __ldg(reinterpret_cast<const float4*>(ok && xl ? p - C : col));
            const float4 m = __ldg(reinterpret_cast<const float4*>(p));
            const float4 r = __ldg(reinterpret_cast<const float4*>(ok && xr ? p + C : col));
            d[0] = ok && xl ? l : zero;
            d[1] = ok ? m : zero;
            d[2] = ok && xr ? r : zero;
        };
        // kSlide: four rows in registers (the window and the next row on its way); else every row of the strip loaded up front
        constexpr int kRows = kSlide ? 4 : R + 2;
        float4 w[kRows][3];
#pragma unroll
        for (int i = 0; i < (kSlide ? 3 : R + 2); ++i) load_row(y0 - 1 + i, w[i]);
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int y = y0 + i;
            if (kSlide && i + 1 < R) load_row(y + 2, w[(i + 3) % kRows]);
            float4 s = zero;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const float4 v = w[(i + k / 3) % kRows][k % 3];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            const float r0 = fmaxf(__fadd_rn(__fmul_rn(s.x, ninth), b.x), 0.f), r1f = fmaxf(__fadd_rn(__fmul_rn(s.y, ninth), b.y), 0.f);
            const float r2f = fmaxf(__fadd_rn(__fmul_rn(s.z, ninth), b.z), 0.f), r3 = fmaxf(__fadd_rn(__fmul_rn(s.w, ninth), b.w), 0.f);
            __nv_bfloat162 h0 = __floats2bfloat162_rn(r0, r1f), h1 = __floats2bfloat162_rn(r2f, r3);
            const uint2 o = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
            if (y < H) *reinterpret_cast<uint2*>(ocol + (size_t)y * W * out_ld) = o;
        }
    }
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

// K5 with its input staged by TMA (see maxpool3x3s2_tma): a CTA owns TR x TC outputs of one image and 32 channels, loads the
// (TR + 2) x (TC + 2) pre-activations around them with one tensor load that starts one pixel outside the tile -- TMA fills
// what lies outside the image with zeros, which is exactly the padding -- and a thread walks 4 output rows of one column with
// the window's rows in registers.  No border predicates, no address arithmetic per tap, every input read from L2 1.3 times
// instead of 9; same additions in the same order as the kernel above.
constexpr int kAvgStrip = 4;
__global__ void __launch_bounds__(256)
avgpool3x3_bias_relu_tma(const __grid_constant__ CUtensorMap tm, const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int H, int W,
                         int TR, int TC, int tiles_x, int tiles_y, int chunks, int out_ld, int out_off, unsigned long long* tl) {
    extern __shared__ __align__(128) unsigned char pool_smem[];
    __shared__ uint64_t bar;
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    unsigned b = blockIdx.x;
    const int chunk = (int)(b % (unsigned)chunks); b /= (unsigned)chunks;
    const int tx = (int)(b % (unsigned)tiles_x); b /= (unsigned)tiles_x;
    const int ty = (int)(b % (unsigned)tiles_y);
    const int img = (int)(b / (unsigned)tiles_y);
    const int IR = TR + 2, IC = TC + 2;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    __syncthreads();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    if (threadIdx.x == 0) {
        ptx::mbar_expect_tx(&bar, (uint32_t)(IR * IC * 128));
        ptx::tma_load_4d(pool_smem, &tm, &bar, chunk * 32, tx * TC - 1, ty * TR - 1, img);
    }
    const int strips = (TR + kAvgStrip - 1) / kAvgStrip;
    const int items = strips * TC * 8;
    const float ninth = 1.0f / 9.0f;
    ptx::mbar_wait(&bar, 0);
    const float4* tile = reinterpret_cast<const float4*>(pool_smem);
    for (int i = threadIdx.x; i < items; i += 256) {
        const int g = i & 7, p = i >> 3;
        const int oxl = p % TC, strip = p / TC;
        const int ox = tx * TC + oxl;
        if (ox >= W) continue;
        const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + chunk * 32 + g * 4));
        const int r0 = strip * kAvgStrip;                                  // first output row of the strip inside the tile
        const float4* col = tile + (r0 * IC + oxl) * 8 + g;                // input row r0 - 1 (tile row r0), column ox - 1
        float4 w[kAvgStrip + 2][3];
#pragma unroll
        for (int r = 0; r < kAvgStrip + 2; ++r)
#pragma unroll
            for (int k = 0; k < 3; ++k) w[r][k] = r0 + r < IR ? col[(r * IC + k) * 8] : make_float4(0.f, 0.f, 0.f, 0.f);
        __nv_bfloat16* o = out + (((size_t)img * H + ty * TR + r0) * W + ox) * out_ld + out_off + chunk * 32 + g * 4;
#pragma unroll
        for (int r = 0; r < kAvgStrip; ++r) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const float4 v = w[r + k / 3][k % 3];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            const float q0 = fmaxf(__fadd_rn(__fmul_rn(s.x, ninth), bv.x), 0.f), q1 = fmaxf(__fadd_rn(__fmul_rn(s.y, ninth), bv.y), 0.f);
            const float q2 = fmaxf(__fadd_rn(__fmul_rn(s.z, ninth), bv.z), 0.f), q3 = fmaxf(__fadd_rn(__fmul_rn(s.w, ninth), bv.w), 0.f);
            __nv_bfloat162 h0 = __floats2bfloat162_rn(q0, q1), h1 = __floats2bfloat162_rn(q2, q3);
            const uint2 ov = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
            if (r0 + r < TR && ty * TR + r0 + r < H) *reinterpret_cast<uint2*>(o + (size_t)r * W * out_ld) = ov;
        }
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

// The one-thread-per-output kernels (channel counts the tiled kernels do not take): ctas_per_sm caps their grid (16 fills an
// otherwise idle GPU; FTF_BRANCH_POOL_CTAS for the pool of an Inception branch, a measurement switch).
// An SM's split between L1 and shared memory is fixed while any CTA runs on it.  A kernel that asks for no shared memory
// gets SMs configured for a large L1, and a convolution CTA (200 KB of shared memory) cannot start on such an SM until every
// pool CTA has left it; asking for the maximum shared-memory carve-out gives both kinds of kernel the same configuration.
// What remains, whatever the pool kernel looks like: a convolution CTA needs a whole SM, so the next
// convolution of a sibling branch starts only when the pool CTAs that slipped onto a freed SM have left it.
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
                        std::string* err, unsigned long long* tl, int ctas_per_sm, int reverse) {
    if (n <= 0) return FTF_OK;
    if (H < 3 || W < 3 || C % 8 || out_ld % 8 || out_off % 8) { if (err) *err = "maxpool3x3s2: H,W >= 3 and C, out_ld, out_off multiples of 8"; return FTF_INVALID_ARGUMENT; }
    const int OH = (H - 3) / 2 + 1, OW = (W - 3) / 2 + 1;
    const long long total = (long long)n * OH * OW * (C / 8);
    if (total >= (1ll << 31)) { if (err) *err = "maxpool3x3s2: tensor too large"; return FTF_INVALID_ARGUMENT; }
    // TMA-tiled kernel: channels in chunks of 64 or 32, 16-byte aligned rows
    static const bool no_tma = getenv("FTF_POOL_NO_TMA") != nullptr;       // measurement switch
    const int chunk_c = C % 64 == 0 ? 64 : (C % 32 == 0 ? 32 : 0);
    if (!no_tma && chunk_c && ((uintptr_t)in & 15u) == 0 && n <= 65535) {
        const int cgc = chunk_c / 8, chunks = C / chunk_c;
        // tile: least padding of the output grid first, then the larger tile; at most 40 KB of shared memory (five CTAs per SM)
        int TR = 0, TC = 0;
        double best = 1e30;
        for (int tr = 3; tr <= 8; ++tr)
            for (int tc = 6; tc <= 18; ++tc) {
                if ((2 * tr + 1) * (2 * tc + 1) * chunk_c * 2 > 40 * 1024) continue;
                const double padded = (double)((OH + tr - 1) / tr * tr) * ((OW + tc - 1) / tc * tc) / ((double)OH * OW);
                const double rim = (double)(2 * tr + 1) * (2 * tc + 1) / (4.0 * tr * tc);
                const double cost = padded * rim;                       // input bytes loaded per useful input byte
                if (cost < best - 1e-9) { best = cost; TR = tr; TC = tc; }
            }
        const int tiles_y = (OH + TR - 1) / TR, tiles_x = (OW + TC - 1) / TC;
        const long long ctas = (long long)n * tiles_y * tiles_x * chunks;
        auto encode = get_tensor_map_encode_tiled();
        if (TR && encode && ctas < (1ll << 31)) {
            CUtensorMap tm;
            cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
            cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * W, (cuuint64_t)C * 2 * W * H};
            cuuint32_t box[4] = {(cuuint32_t)chunk_c, (cuuint32_t)(2 * TC + 1), (cuuint32_t)(2 * TR + 1), 1};
            cuuint32_t estr[4] = {1, 1, 1, 1};
            const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (cr != CUDA_SUCCESS) { if (err) *err = "maxpool3x3s2: cuTensorMapEncodeTiled failed, code " + std::to_string((int)cr); return FTF_INTERNAL; }
            const size_t smem = (size_t)(2 * TR + 1) * (2 * TC + 1) * chunk_c * 2;
            FTF_CUDA(launch_pdl(maxpool3x3s2_tma, dim3((unsigned)ctas), dim3(256), smem, s, tm, (__nv_bfloat16*)out, OH, OW, TR, TC, tiles_x, tiles_y,
                                chunks, cgc, out_ld, out_off, tl), err);
            count_launch();
            return FTF_OK;
        }
    }
    const long long want = (total + 255) / 256, cap = grid_cap(ctas_per_sm);
    prefer_max_smem(maxpool3x3s2);
    FTF_CUDA(launch_pdl(maxpool3x3s2, dim3((unsigned)(want < cap ? want : cap)), dim3(256), 0, s, (const __nv_bfloat16*)in, (__nv_bfloat16*)out, n, H, W, C,
                        OH, OW, out_ld, out_off, reverse, tl), err);
    count_launch();
    return FTF_OK;
}

int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, void* out, int n, int H, int W, int C, int out_ld,
                                int out_off, cudaStream_t s, std::string* err, unsigned long long* tl, int ctas_per_sm) {
    if (n <= 0) return FTF_OK;
    if (C % 4 || out_ld % 4 || out_off % 4) { if (err) *err = "avgpool3x3: C, out_ld, out_off multiples of 4"; return FTF_INVALID_ARGUMENT; }
    if ((long long)n * H * W * (C / 4) >= (1ll << 31)) { if (err) *err = "avgpool3x3: tensor too large"; return FTF_INVALID_ARGUMENT; }
    static const bool no_tma = getenv("FTF_POOL_NO_TMA") != nullptr;       // measurement switch
    if (!no_tma && C % 32 == 0 && ((uintptr_t)pre & 15u) == 0) {
        const int chunks = C / 32;
        int TR = 0, TC = 0;
        double best = 1e30;
        for (int tr = 4; tr <= 36; ++tr)
            for (int tc = 4; tc <= 36; ++tc) {
                if (tr > H + 3 || tc > W + 3 || (tr + 2) * (tc + 2) * 128 > 44 * 1024) continue;
                const double padded = (double)((H + tr - 1) / tr * tr) * ((W + tc - 1) / tc * tc) / ((double)H * W);
                const double rim = (double)(tr + 2) * (tc + 2) / ((double)tr * tc);
                const double strip_pad = (double)((tr + kAvgStrip - 1) / kAvgStrip * kAvgStrip) / tr;      // idle rows of the last strip
                const double cost = padded * (rim + 0.5 * strip_pad);
                if (cost < best - 1e-9) { best = cost; TR = tr; TC = tc; }
            }
        const int tiles_y = TR ? (H + TR - 1) / TR : 0, tiles_x = TR ? (W + TC - 1) / TC : 0;
        const long long ctas = (long long)n * tiles_y * tiles_x * chunks;
        auto encode = get_tensor_map_encode_tiled();
        if (TR && encode && ctas < (1ll << 31)) {
            CUtensorMap tm;
            cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
            cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)C * 4 * W, (cuuint64_t)C * 4 * W * H};
            cuuint32_t box[4] = {32, (cuuint32_t)(TC + 2), (cuuint32_t)(TR + 2), 1};
            cuuint32_t estr[4] = {1, 1, 1, 1};
            const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(pre), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (cr != CUDA_SUCCESS) { if (err) *err = "avgpool3x3: cuTensorMapEncodeTiled failed, code " + std::to_string((int)cr); return FTF_INTERNAL; }
            const size_t smem = (size_t)(TR + 2) * (TC + 2) * 128;
            FTF_CUDA(launch_pdl(avgpool3x3_bias_relu_tma, dim3((unsigned)ctas), dim3(256), smem, s, tm, bias, (__nv_bfloat16*)out, H, W, TR, TC, tiles_x,
                                tiles_y, chunks, out_ld, out_off, tl), err);
            count_launch();
            return FTF_OK;
        }
    }
    const long long cap = grid_cap(ctas_per_sm);
    auto go = [&](auto kernel, int rows) -> cudaError_t {
        const int strips = (H + rows - 1) / rows;
        const long long total = (long long)n * strips * W * (C / 4), want = (total + 127) / 128;
        prefer_max_smem(kernel);
        return launch_pdl(kernel, dim3((unsigned)(want < cap ? want : cap)), dim3(128), 0, s, pre, bias, (__nv_bfloat16*)out, n, H, W, C, out_ld,
                          out_off, strips, tl);
    };
    // rows per thread: 6 where that still leaves a hundred thousand threads (17x17 and 35x35 at batch 64), 4 on small tensors
    const bool small = (long long)n * ((H + 5) / 6) * W * (C / 4) < 64 * 1024;
    static const bool slide = getenv("FTF_POOL_UPFRONT") == nullptr;
    FTF_CUDA(slide ? (small ? go(avgpool3x3_bias_relu<4, true>, 4) : go(avgpool3x3_bias_relu<6, true>, 6))
                   : (small ? go(avgpool3x3_bias_relu<4, false>, 4) : go(avgpool3x3_bias_relu<6, false>, 6)), err);
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
