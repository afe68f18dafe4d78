// inception.cu -- the label model: Inception-v3 (inference, BN folded, bf16 activations) as a list of device
// operations over a static activation arena; convolutions run on tcgen05 (conv_tc.cu).
//
// Stands where the reference loads and runs its frozen Inception graph:
// flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples/inception/
// InceptionModel.scala:51-67 (DefaultGraphLoader + ModelFunction[LabelMethod]).  The reference ships no graph
// (models/README.md:5); architecture and seeded weight recipe are those of oracle/inception_oracle.c
// (cross-checked there against torchvision).
//
// Data layout in HBM: every activation is a dense bf16 NHWC tensor; an Inception block's branches write
// directly into channel slices of the block's output tensor (no concat pass).  The 1x1 convolutions of a block
// that read the block input are one GEMM whose column ranges ("segments") go to different tensors; the pool
// branch is conv1x1 -> fp32 scratch -> (avgpool3x3 + bias + ReLU) since pool and 1x1 conv commute.
#include "inception.h"
#include "wire.h"
#include "conv_tc.h"
#include "pools.h"
#include "f32_path.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

namespace ftf {

// ------------------------------------------------------------------------------------------------
// seeded weights (same recipe as the oracle: see oracle/inception_oracle.c header)
// ------------------------------------------------------------------------------------------------
namespace {

inline uint64_t splitmix(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline uint64_t name_hash(const std::string& s) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (unsigned char c : s) { h ^= c; h *= 0x100000001b3ull; }
    return h;
}
struct Stream {
    uint64_t key;
    Stream(uint64_t seed, const std::string& name, int kind) : key(splitmix(seed ^ splitmix(name_hash(name) + (uint64_t)kind))) {}
    double u_open(uint64_t i) const { return (double)((splitmix(key ^ i) >> 11) + 1) * (1.0 / 9007199254740992.0); }
    double u_half(uint64_t i) const { return (double)(splitmix(key ^ i) >> 11) * (1.0 / 9007199254740992.0); }
    double normal(uint64_t i) const { return sqrt(-2.0 * log(u_open(2 * i))) * cos(6.283185307179586476925286766559 * u_half(2 * i + 1)); }
};

void fill_parallel(int64_t n, const std::function<void(int64_t, int64_t)>& fn) {
    unsigned hw = std::thread::hardware_concurrency();
    int t = (int)(hw ? (hw > 16 ? 16 : hw) : 4);
    if (n < 65536) { fn(0, n); return; }
    std::vector<std::thread> th;
    const int64_t per = (n + t - 1) / t;
    for (int i = 0; i < t; ++i) {
        const int64_t lo = i * per, hi = std::min<int64_t>(n, lo + per);
        if (lo < hi) th.emplace_back(fn, lo, hi);
    }
    for (auto& x : th) x.join();
}

// folded conv weights [Cout][KH][KW][Cin] (fp32) and bias
void conv_weights(uint64_t seed, const std::string& name, int cout, int kh, int kw, int cin, std::vector<float>* w,
                  std::vector<float>* b) {
    const Stream sw(seed, name, 0), sb(seed, name, 1);
    const double sd = sqrt(2.0 / ((double)kh * kw * cin)), fold = 1.0 / sqrt(1.0 + 1e-3);
    w->resize((size_t)cout * kh * kw * cin);
    float* wp = w->data();
    fill_parallel((int64_t)w->size(), [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) wp[i] = (float)(sw.normal((uint64_t)i) * sd * fold);
    });
    b->resize(cout);
    for (int o = 0; o < cout; ++o) (*b)[o] = (float)(-0.1 + 0.2 * sb.u_half((uint64_t)o));
}

inline uint16_t bf16_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
inline float bf16_value(float f) {
    uint32_t u = (uint32_t)bf16_bits(f) << 16;
    float r;
    memcpy(&r, &u, 4);
    return r;
}

// ------------------------------------------------------------------------------------------------
// kernels other than the tensor-core convolution
// ------------------------------------------------------------------------------------------------
// Conv2d_1a_3x3: 3 -> 32 channels, 3x3, stride 2, no padding.  K = 27 is far below a tensor-core tile and the
// layer is bound by its 64-byte-per-pixel output: direct convolution, one thread per output pixel, weights as
// kernel parameters (constant-bank operands of the FMAs).
struct StemWeights {
    float w[27][32];   // [kh][kw][cin] x cout, bf16-rounded values
    float b[32];
};
// TIn = float: the records of the label model arrive as fp32 images (the reference's ImageTensor); each value is rounded
// to bf16 (RNE) as it is loaded, which is what a separate cast pass in front of the network would have produced, bit for bit.
__device__ __forceinline__ float stem_load(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float stem_load(const float* p) { return __bfloat162float(__float2bfloat16_rn(__ldg(p))); }
template <typename TIn>
__global__ void __launch_bounds__(128)
stem_conv3x3s2(const TIn* __restrict__ in, __nv_bfloat16* __restrict__ out, int n, int H, int W, int OH, int OW,
               const __grid_constant__ StemWeights sw, unsigned long long* tl) {
    ptx::tl_stamp_min(tl);
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const long long total = (long long)n * OH * OW;
    const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= total) return;
    const int img = (int)(pix / (OH * OW));
    const int rem = (int)(pix - (long long)img * OH * OW);
    const int oy = rem / OW, ox = rem - oy * OW;
    float acc[32];
#pragma unroll
    for (int o = 0; o < 32; ++o) acc[o] = 0.0f;
    const TIn* base = in + (((size_t)img * H + 2 * oy) * W + 2 * ox) * 3;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
        float x[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) x[i] = stem_load(base + (size_t)ky * W * 3 + i);   // 3 pixels x 3 channels, contiguous
#pragma unroll
        for (int i = 0; i < 9; ++i)
#pragma unroll
            for (int o = 0; o < 32; ++o) acc[o] = fmaf(x[i], sw.w[ky * 9 + i][o], acc[o]);
    }
    uint4* op = reinterpret_cast<uint4*>(out + (size_t)pix * 32);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint32_t r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = q * 8 + j * 2;
            __nv_bfloat162 h = __floats2bfloat162_rn(fmaxf(acc[o] + sw.b[o], 0.0f), fmaxf(acc[o + 1] + sw.b[o + 1], 0.0f));
            r[j] = *reinterpret_cast<uint32_t*>(&h);
        }
        op[q] = make_uint4(r[0], r[1], r[2], r[3]);
    }
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

// ------------------------------------------------------------------------------------------------
// Conv2d_1a_3x3 on the tensor cores.  K = 27 (padded to 32) is below what TMA im2col can express -- three channels are six
// bytes, no 16-byte box -- so the A operand is built by the threads: thread t of a 128-thread CTA gathers the 3 x 9 input
// values of output pixel t (each filter row is nine contiguous elements of the NHWC image), rounds fp32 records to bf16 as
// it goes, and writes the pixel's 64-byte K row into shared memory in the 64-byte-swizzled K-major layout tcgen05 reads
// (16-byte chunk j of row r lands at chunk j ^ ((r >> 1) & 3): the pattern TMA's SWIZZLE_64B writes).  The 32 x 32 weight
// tile gets the same layout.  Two tcgen05.mma (M = 128, N = 32, K = 16) fill a 32-column TMEM accumulator; every thread
// then reads back its own row (tcgen05.ld), adds the bias, applies ReLU and stores its pixel's 32 channels as 64 bytes.
// The direct-FMA kernel this replaces (864 FMAs per pixel) was bound by the CUDA cores at twice the layer's HBM time.
// ------------------------------------------------------------------------------------------------
constexpr int kStemCtas = 6;             // CTAs per SM of the stem kernel (85 registers each with the next tile's loads in flight)
struct StemTcParams {
    float bias[32];
};
__device__ __forceinline__ uint32_t stem_pack(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float stem_raw(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float stem_raw(const float* p) { return __ldg(p); }
__device__ __forceinline__ __nv_bfloat16 stem_fetch(const __nv_bfloat16* p) { return *p; }
__device__ __forceinline__ float stem_fetch(const float* p) { return __ldg(p); }
__device__ __forceinline__ __nv_bfloat16 stem_zero(__nv_bfloat16) { return __float2bfloat16(0.0f); }
__device__ __forceinline__ float stem_zero(float) { return 0.0f; }
__device__ __forceinline__ float stem_f(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float stem_f(float v) { return v; }
template <typename TIn>
__global__ void __launch_bounds__(128, kStemCtas)
stem_conv_tc(const TIn* __restrict__ in, __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ w32 /* [32 cout][32 k] */,
             int n, int H, int W, int OH, int OW, const __grid_constant__ StemTcParams prm, unsigned long long* tl) {
    __shared__ __align__(1024) uint8_t a_s[128 * 64];
    __shared__ __align__(1024) uint8_t b_s[32 * 64];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t tmem_slot;
    ptx::tl_stamp_min(tl);
    const int t = (int)threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, t >> 5, 0);
    if (warp == 0) ptx::tmem_alloc<32>(&tmem_slot);
    if (t == 32) { ptx::mbar_init(&done_bar, 1); ptx::fence_mbar_init(); }
    // the weight tile: thread o < 32 writes row o (64 bytes = four swizzled 16-byte chunks)
    if (t < 32) {
        const uint4* src = reinterpret_cast<const uint4*>(w32 + (size_t)t * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(b_s + t * 64 + ((j ^ ((t >> 1) & 3)) << 4)) = __ldg(src + j);
    }
    ptx::pdl_launch_dependents();
    ptx::pdl_wait();
    ptx::tl_stamp_min(tl ? tl + 2 : nullptr);
    const long long total = (long long)n * OH * OW;
    const long long tiles = (total + 127) / 128;
    ptx::tc_fence_before();
    __syncthreads();                           // TMEM address, barrier and weight tile are in place
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    // K-major, 64-byte rows, 8-row groups 512 bytes apart, SWIZZLE_64B; K step 1 starts 32 bytes into the row
    const uint32_t hi = (uint32_t)(512 >> 4) | (1u << 14) | (4u << 29);
    const uint32_t a_lo = ((ptx::smem_u32(a_s) & 0x3FFFF) >> 4) | (1u << 16), b_lo = ((ptx::smem_u32(b_s) & 0x3FFFF) >> 4) | (1u << 16);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    uint32_t parity = 0;
    // Persistent: TMEM is allocated once per CTA (allocations on an SM are serialised; one per 128-pixel tile made them
    // the kernel's critical path) and the CTA walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...; eight CTAs per SM
    // overlap one another's gather, MMA and store phases.
    // The 27 loads of the NEXT tile are issued before this tile's MMA and epilogue and consumed at the top of the next
    // iteration: a CTA has two tiles' worth of loads in flight instead of one serial load -> MMA -> store chain per tile.
    TIn raw[27];
    auto gather = [&](long long tile_) {
        const long long pix_ = tile_ * 128 + t;
        if (tile_ < tiles && pix_ < total) {
            const int img = (int)(pix_ / (OH * OW));
            const int rem = (int)(pix_ - (long long)img * OH * OW);
            const int oy = rem / OW, ox = rem - oy * OW;
            const TIn* base = in + (((size_t)img * H + 2 * oy) * W + 2 * ox) * 3;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int i = 0; i < 9; ++i) raw[ky * 9 + i] = stem_fetch(base + (size_t)ky * W * 3 + i);
        } else {
#pragma unroll
            for (int i = 0; i < 27; ++i) raw[i] = stem_zero(raw[i]);
        }
    };
    gather(blockIdx.x);
    for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const long long pix = tile * 128 + t;
        const bool live = pix < total;
        {
            uint32_t kw[16];                       // 32 bf16: k = ky*9 + kx*3 + c for k < 27, zeros behind
#pragma unroll
            for (int i = 0; i < 13; ++i) kw[i] = stem_pack(stem_f(raw[2 * i]), stem_f(raw[2 * i + 1]));      // RNE to bf16, as a cast pass would
            kw[13] = stem_pack(stem_f(raw[26]), 0.0f);
            kw[14] = 0u; kw[15] = 0u;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(a_s + t * 64 + ((j ^ ((t >> 1) & 3)) << 4)) = make_uint4(kw[4 * j], kw[4 * j + 1], kw[4 * j + 2], kw[4 * j + 3]);
        }
        gather(tile + gridDim.x);
        ptx::fence_proxy_async();                  // the tensor core reads shared memory through the async proxy
        ptx::tc_fence_before();
        __syncthreads();                           // (also: every thread has read the previous tile's accumulator)
        ptx::tc_fence_after();
        if (warp == 0) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                uint64_t da, db;
                asm("mov.b64 %0, {%1, %2};" : "=l"(da) : "r"(a_lo + 2u * k), "r"(hi));
                asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(b_lo + 2u * k), "r"(hi));
                ptx::umma_bf16_elect(tmem_base, da, db, idesc, k ? 1u : 0u);
            }
            ptx::umma_commit_elect(&done_bar);
        }
        ptx::mbar_wait(&done_bar, parity);         // the MMAs have read a_s and written the accumulator
        parity ^= 1u;
        ptx::tc_fence_after();
        uint32_t acc[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(warp * 32) << 16), acc);
        ptx::tmem_ld_wait();
        if (live) {
            uint4* op = reinterpret_cast<uint4*>(out + (size_t)pix * 32);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t r[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int o = q * 8 + j * 2;
                    r[j] = stem_pack(fmaxf(__uint_as_float(acc[o]) + prm.bias[o], 0.0f), fmaxf(__uint_as_float(acc[o + 1]) + prm.bias[o + 1], 0.0f));
                }
                op[q] = make_uint4(r[0], r[1], r[2], r[3]);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<32>(tmem_base);
    ptx::tl_stamp_max(tl ? tl + 1 : nullptr);
}

__global__ void __launch_bounds__(256) cast_f32_bf16(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) out[i] = __float2bfloat16_rn(__ldg(in + i));
}

}  // namespace

int launch_cast_f32_bf16(const float* in, void* out, int64_t n, cudaStream_t s, std::string* err) {
    if (n <= 0) return FTF_OK;
    int dev = 0;
    FTF_CUDA(cudaGetDevice(&dev), err);
    int64_t want = (n + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count(dev) * 8;
    cast_f32_bf16<<<(unsigned)(want < cap ? want : cap), 256, 0, s>>>(in, (__nv_bfloat16*)out, n);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

// ------------------------------------------------------------------------------------------------
// network
// ------------------------------------------------------------------------------------------------
struct Act {
    void* p = nullptr;                   // bf16 (tensor-core mode) or fp32 (precision mode) NHWC
    int h = 0, w = 0, c = 0;
};

// Images per group of the layers ahead of Mixed_5b (Impl::split_early_layers); 0: whole batch, layer by layer.  Measured at
// batch 64: groups of 16 make the two max pools 34 us faster (their input is still in L2) and the four convolutions 60 us
// slower (each launch pays its pipeline fill and its tail again, and 630 tiles fall worse on 148 SMs than 2521), groups of 32
// come out even: off, and kept as a switch (FTF_EARLY_GROUP=<images>).
constexpr int kEarlyGroup = 0;
constexpr int kLanes = 4;                // streams an Inception block's independent branches are spread over

struct Op {
    enum Type { kStem, kConv, kMaxPool, kAvgPoolBiasRelu, kGap, kFork, kJoin, kConvF32 } type;
    int lane = 0;                        // stream the op runs on (0 = the caller's stream)
    int lanes = 0;                       // kFork/kJoin: bit mask of the lanes 1..3 that branch off / come back
    int plan = -1;                       // kConv
    const void* in = nullptr;            // other kernels
    void* out = nullptr;
    const float* bias = nullptr;
    int H = 0, W = 0, C = 0, OH = 0, OW = 0, out_ld = 0, out_off = 0;
    int reverse = 0;                     // walk the tensor from its end (FTF_REVERSE, measurement)
    int first = 0, count = 0;            // count > 0: the op covers images [first, first + count) of the batch only (Impl::split_early_layers)
    std::string tag;                     // layer name(s) for the FTF_TIMELINE dump
};

// one convolution of the fp32 mode: geometry + where its column ranges go
struct ConvF32 {
    ConvGeometry g;
    std::vector<ConvSegment> segs;
};

struct InceptionV3::Impl {
    int device = 0, sms = 148, max_batch = 0, classes = 0;
    // Trained weights (ftf_create("inception_v3:<dir>")): a TensorBundle holding "<layer>/weights" fp32 [Cout,KH,KW,Cin] with
    // the BatchNorm already folded in, "<layer>/biases" fp32 [Cout], "fc/weights" [classes,2048] and "fc/biases" [classes].
    // Empty: the seeded random-init recipe of the oracle.
    std::unique_ptr<SavedModelLite> bundle;
    bool weights_dir_loaded = false;
    struct LayerSpec { std::string name; int cout, kh, kw, cin; };
    std::vector<LayerSpec> layer_specs;   // every convolution in build order (what ftf_label_export_weights writes)
    int layer_weights(const std::string& name, int cout, int kh, int kw, int cin, std::vector<float>* w, std::vector<float>* b) {
        layer_specs.push_back({name, cout, kh, kw, cin});
        if (!bundle) { conv_weights(seed, name, cout, kh, kw, cin, w, b); return FTF_OK; }
        const float *pw = nullptr, *pb = nullptr;
        int r = bundle->tensor_f32(name + "/weights", {cout, kh, kw, cin}, &pw, err);
        if (!r) r = bundle->tensor_f32(name + "/biases", {cout}, &pb, err);
        if (r) return r;
        w->assign(pw, pw + (size_t)cout * kh * kw * cin);
        b->assign(pb, pb + cout);
        return FTF_OK;
    }
    int fc_weights(int cin, std::vector<float>* w, std::vector<float>* b) {       // [classes][cin], [classes]
        if (bundle) {
            const float *pw = nullptr, *pb = nullptr;
            int r = bundle->tensor_f32("fc/weights", {classes, cin}, &pw, err);
            if (!r) r = bundle->tensor_f32("fc/biases", {classes}, &pb, err);
            if (r) return r;
            w->assign(pw, pw + (size_t)classes * cin);
            b->assign(pb, pb + classes);
            return FTF_OK;
        }
        w->resize((size_t)classes * cin);
        b->assign(classes, 0.0f);
        const Stream sw(seed + 90, "fc", 0);      // salted fc stream: well-separated top logits for the default seed (see the oracle)
        float* wp = w->data();
        fill_parallel((int64_t)classes * cin, [&](int64_t lo, int64_t hi) { for (int64_t i = lo; i < hi; ++i) wp[i] = (float)(sw.normal((uint64_t)i) * 0.05); });
        return FTF_OK;
    }
    bool f32 = false;                     // compute_dtype FTF_FLOAT: fp32 activations and weights, CUDA-core kernels (f32_path.cu)
    std::vector<ConvF32> convs_f32;
    uint64_t seed = 0;
    std::vector<void*> allocs;
    std::vector<ConvPlan> plans;
    struct PlanInfo { ConvPlanSet set; ConvGeometry g; std::vector<ConvSegment> segs; int sm = 0; std::string tag; };
    std::vector<PlanInfo> plan_info;      // per plan: every candidate kernel of the layer (for the in-situ refinement)
    std::vector<Op> ops;
    std::map<std::string, Act> named;     // activations by layer / block name (test hook)
    StemWeights stem;
    StemTcParams stem_tc;                 // bias of the tensor-core stem kernel
    __nv_bfloat16* stem_w32 = nullptr;    // its weights, bf16 [32 cout][32 k] (k = ky*9 + kx*3 + c, zeros from 27 on)
    Act image;                            // [max_batch,299,299,3] staging of the caller's images
    float* logits = nullptr;              // [max_batch, classes] fp32
    std::string* err = nullptr;
    int rc = FTF_OK;
    int cur_lane = 0;
    cudaStream_t lane_stream[kLanes] = {nullptr, nullptr, nullptr, nullptr};   // [0] unused: lane 0 is the caller's stream
    cudaEvent_t fork_ev = nullptr, join_ev[kLanes] = {nullptr, nullptr, nullptr, nullptr};
    std::map<int, cudaGraphExec_t> graphs;   // one instantiated graph of the whole forward per batch size
    std::map<int, int> graph_kernels;        // kernels per replay of that graph
    unsigned long long* tl_dev = nullptr;    // FTF_TIMELINE=<k>: 4 stamps per op, dumped after the k-th forward
    int tl_after = 0, tl_calls = 0;

    ~Impl() {
        cudaSetDevice(device);
        for (auto& g : graphs) cudaGraphExecDestroy(g.second);
        for (int i = 1; i < kLanes; ++i) {
            if (lane_stream[i]) cudaStreamDestroy(lane_stream[i]);
            if (join_ev[i]) cudaEventDestroy(join_ev[i]);
        }
        if (fork_ev) cudaEventDestroy(fork_ev);
        for (void* p : allocs) cudaFree(p);
    }
    void fork(int lanes) { Op o{}; o.type = Op::kFork; o.lanes = lanes; ops.push_back(o); }
    void join(int lanes) { Op o{}; o.type = Op::kJoin; o.lanes = lanes; ops.push_back(o); cur_lane = 0; }
    void emit_pools() {      // the (avgpool + bias + ReLU) halves of the pool branches recorded by conv_multi
        for (auto& pp : pending_pool) {
            Op o{}; o.type = Op::kAvgPoolBiasRelu; o.lane = cur_lane; o.in = pp.scratch; o.bias = pp.bias; o.out = pp.dst->p;
            o.H = pp.h; o.W = pp.w; o.C = pp.c; o.out_ld = pp.dst->c; o.out_off = pp.coff;
            ops.push_back(o);
        }
        pending_pool.clear();
    }
    void* dev_alloc(size_t bytes, bool zero) {
        if (rc) return nullptr;
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
        if (e != cudaSuccess) { rc = FTF_RESOURCE_EXHAUSTED; *err = std::string("activation/weight arena: ") + cudaGetErrorString(e); cudaGetLastError(); return nullptr; }
        if (zero) cudaMemset(p, 0, bytes);
        allocs.push_back(p);
        return p;
    }
    Act act(int h, int w, int c) {
        Act a; a.h = h; a.w = w; a.c = c;
        a.p = dev_alloc((size_t)max_batch * h * w * c * (f32 ? 4 : 2), true);
        return a;
    }
    struct Branch { std::string name; int cout; Act* dst; int coff; bool pool; };

    // conv(s) reading `in`: all entries share kernel shape; their outputs are column ranges of one GEMM
    void conv_multi(const Act& in, const std::vector<Branch>& br, int kh, int kw, int s, int ph, int pw) {
        if (rc) return;
        int total = 0;
        for (auto& b : br) total += b.cout;
        std::vector<uint16_t> w16(f32 ? 0 : (size_t)total * kh * kw * in.c);
        std::vector<float> w32(f32 ? (size_t)total * kh * kw * in.c : 0);
        std::vector<float> bias(total);
        std::vector<ConvSegment> segs;
        const int OH = (in.h + 2 * ph - kh) / s + 1, OW = (in.w + 2 * pw - kw) / s + 1;
        int col = 0;
        for (auto& b : br) {
            std::vector<float> w, bb;
            if ((rc = layer_weights(b.name, b.cout, kh, kw, in.c, &w, &bb))) return;
            const size_t per = (size_t)kh * kw * in.c;
            if (f32) memcpy(&w32[(size_t)col * per], w.data(), w.size() * 4);         // the oracle's fp32 weights, unrounded
            else for (size_t i = 0; i < w.size(); ++i) w16[(size_t)col * per + i] = bf16_bits(w[i]);
            for (int o = 0; o < b.cout; ++o) bias[col + o] = bb[o];
            ConvSegment sg{};
            sg.col_begin = col;
            if (b.pool) {   // raw fp32 pre-activation for the pool branch; pooled, biased and activated by the next op
                float* scratch = (float*)dev_alloc((size_t)max_batch * OH * OW * b.cout * 4, true);
                float* dbias = (float*)dev_alloc((size_t)b.cout * 4, false);
                if (rc) return;
                cudaMemcpy(dbias, bb.data(), (size_t)b.cout * 4, cudaMemcpyHostToDevice);
                sg.out = scratch; sg.ld = b.cout; sg.coff = 0; sg.flags = kSegF32;
                pending_pool.push_back({scratch, dbias, b.dst, b.coff, b.cout, OH, OW});
            } else {
                sg.out = b.dst->p; sg.ld = b.dst->c; sg.coff = b.coff; sg.flags = kSegBias | kSegRelu | (f32 ? kSegF32 : 0);
            }
            segs.push_back(sg);
            col += b.cout;
        }
        void* dw = dev_alloc(f32 ? w32.size() * 4 : w16.size() * 2, false);
        float* db = (float*)dev_alloc((size_t)total * 4, false);
        if (rc) return;
        if (f32) cudaMemcpy(dw, w32.data(), w32.size() * 4, cudaMemcpyHostToDevice);
        else cudaMemcpy(dw, w16.data(), w16.size() * 2, cudaMemcpyHostToDevice);
        cudaMemcpy(db, bias.data(), (size_t)total * 4, cudaMemcpyHostToDevice);
        ConvGeometry g{in.p, max_batch, in.h, in.w, in.c, in.c, dw, db, total, kh, kw, s, ph, pw};
        if (f32) {
            convs_f32.push_back(ConvF32{g, segs});
            Op op{}; op.type = Op::kConvF32; op.lane = cur_lane; op.plan = (int)convs_f32.size() - 1; op.tag = br[0].name;
            ops.push_back(op);
            return;
        }
        Op op{}; op.type = Op::kConv; op.lane = cur_lane;
        op.tag = br[0].name + (br.size() > 1 ? "+" + std::to_string(br.size() - 1) : std::string()) + " " + std::to_string(kh) + "x" + std::to_string(kw) +
                 " " + std::to_string(in.c) + "->" + std::to_string(total) + " @" + std::to_string(OH);
        if (defer_plans) {        // planned per group of images by split_early_layers
            deferred.push_back(Deferred{g, segs, OH, OW});
            op.plan = -(int)deferred.size();
            ops.push_back(op);
            return;
        }
        op.plan = add_plan(g, segs, share(), op.tag);
        if (rc) return;
        ops.push_back(op);
    }
    int add_plan(const ConvGeometry& g, const std::vector<ConvSegment>& segs, int sm, const std::string& tag) {
        plan_info.emplace_back();
        PlanInfo& pi = plan_info.back();
        pi.g = g; pi.segs = segs; pi.sm = sm; pi.tag = tag;
        rc = conv_tc_plan_set(g, segs.data(), (int)segs.size(), pi.sm, &pi.set, err, kPlanTune);
        if (rc) return -1;
        plans.push_back(pi.set.cands[pi.set.chosen]);
        return (int)plans.size() - 1;
    }
    // The layers between the first convolution and Mixed_5b move 1.2 GB per 64 images through tensors of 44-177 MB each: run
    // layer by layer over the whole batch, every one of them reads its input from HBM because the producer's output no longer
    // is in the 126 MB L2 when the consumer gets to it.  Here the ops [begin, end) are re-emitted group by group: all of these
    // layers for images [0, sub), then for [sub, 2 sub), ... -- a group's tensors (44 MB at most for 16 images) are still in
    // L2 when the next layer reads them.  Same kernels, same tensors, same bits; a layer's plan is made for `sub` images and
    // its tensor maps start at the group's first image.
    struct Deferred { ConvGeometry g; std::vector<ConvSegment> segs; int OH, OW; };
    std::vector<Deferred> deferred;
    bool defer_plans = false;
    void split_early_layers(size_t begin, int sub) {
        if (rc) return;
        std::vector<Op> layer_ops(ops.begin() + begin, ops.end());
        ops.resize(begin);
        for (int first = 0; first < max_batch && !rc; first += sub) {
            const int count = std::min(sub, max_batch - first);
            for (const Op& lo : layer_ops) {
                Op o = lo;
                o.first = first; o.count = count;
                if (o.type == Op::kConv) {
                    const Deferred& d = deferred[-lo.plan - 1];
                    ConvGeometry g = d.g;
                    g.in = (const char*)g.in + (size_t)first * g.H * g.W * g.in_ld * 2;
                    g.n_max = count;
                    std::vector<ConvSegment> segs = d.segs;
                    for (auto& sg : segs) sg.out = (char*)sg.out + (size_t)first * d.OH * d.OW * sg.ld * ((sg.flags & kSegF32) ? 4 : 2);
                    o.plan = add_plan(g, segs, sms, lo.tag);
                    if (rc) return;
                }
                ops.push_back(o);
            }
        }
        deferred.clear();
    }
    // SMs a convolution may plan for.  The layers of the 17x17 and 8x8 blocks are single-wave kernels whose time is the
    // serial fill -> MMA -> epilogue chain of one tile, and a kernel that spreads over all SMs leaves no room for the
    // sibling branch on another stream.  Inside a fork each branch therefore gets a fixed share of the SMs (its
    // persistent CTAs loop over more tiles, overlapping their epilogues with the next tile) and the branches really
    // run side by side.  FTF_NO_SM_SPLIT=1 plans every layer for the whole GPU (measurement switch).
    int sm_share_den = 1;
    int share() const {
        static const bool off = getenv("FTF_NO_SM_SPLIT") != nullptr;
        return off || sm_share_den <= 1 ? sms : (sms / sm_share_den > 0 ? sms / sm_share_den : 1);
    }
    // the sibling branches of a fork are shorter than this one: from here on it plans for the whole GPU again
    // (FTF_SPLIT_WHOLE_BRANCH=1 keeps the share to the end of the branch: measurement switch)
    void rest_alone() {
        static const bool keep = getenv("FTF_SPLIT_WHOLE_BRANCH") != nullptr;
        if (!keep) sm_share_den = 1;
    }
    struct PendingPool { float* scratch; float* bias; Act* dst; int coff, c, h, w; };
    std::vector<PendingPool> pending_pool;

    Act conv(const std::string& name, const Act& in, int cout, int kh, int kw, int s, int ph, int pw) {
        Act out = act((in.h + 2 * ph - kh) / s + 1, (in.w + 2 * pw - kw) / s + 1, cout);
        conv_multi(in, {{name, cout, &out, 0, false}}, kh, kw, s, ph, pw);
        named[name] = out;
        return out;
    }
    void conv_to(const std::string& name, const Act& in, int cout, int kh, int kw, int s, int ph, int pw, Act* dst, int coff) {
        conv_multi(in, {{name, cout, dst, coff, false}}, kh, kw, s, ph, pw);
    }
    void maxpool_to(const Act& in, Act* dst, int coff) {
        Op o{}; o.type = Op::kMaxPool; o.lane = cur_lane; o.in = in.p; o.out = dst->p; o.H = in.h; o.W = in.w; o.C = in.c;
        o.OH = (in.h - 3) / 2 + 1; o.OW = (in.w - 3) / 2 + 1; o.out_ld = dst->c; o.out_off = coff;
        ops.push_back(o);
    }

    // Each block: the 1x1 convolutions that read the block input run first as one GEMM on lane 0; then the branches
    // (independent of each other, writing disjoint channel slices) run on separate streams and join at the block end.
    Act block_a(const std::string& blk, const Act& x, int pf) {
        Act out = act(x.h, x.w, 224 + pf);
        Act t5 = act(x.h, x.w, 48), t3 = act(x.h, x.w, 64);
        conv_multi(x, {{blk + ".branch1x1", 64, &out, 0, false}, {blk + ".branch5x5_1", 48, &t5, 0, false},
                       {blk + ".branch3x3dbl_1", 64, &t3, 0, false}, {blk + ".branch_pool", pf, &out, 224, true}}, 1, 1, 1, 0, 0);
        fork(0b0110);
        if (pool_first) { cur_lane = 2; emit_pools(); }
        cur_lane = 0; conv_to(blk + ".branch5x5_2", t5, 64, 5, 5, 1, 2, 2, &out, 64);
        cur_lane = 1; Act u = conv(blk + ".branch3x3dbl_2", t3, 96, 3, 3, 1, 1, 1);
        conv_to(blk + ".branch3x3dbl_3", u, 96, 3, 3, 1, 1, 1, &out, 128);
        cur_lane = pool_last ? 0 : 2; emit_pools();
        join(0b0110);
        named[blk] = out;
        return out;
    }
    Act block_b(const std::string& blk, const Act& x) {
        const int oh = (x.h - 3) / 2 + 1, ow = (x.w - 3) / 2 + 1;
        Act out = act(oh, ow, 384 + 96 + x.c);
        fork(0b0110);
        cur_lane = 0; conv_to(blk + ".branch3x3", x, 384, 3, 3, 2, 0, 0, &out, 0);
        cur_lane = 1; Act t = conv(blk + ".branch3x3dbl_1", x, 64, 1, 1, 1, 0, 0);
        Act u = conv(blk + ".branch3x3dbl_2", t, 96, 3, 3, 1, 1, 1);
        conv_to(blk + ".branch3x3dbl_3", u, 96, 3, 3, 2, 0, 0, &out, 384);
        cur_lane = pool_last ? 0 : 2; maxpool_to(x, &out, 480);
        join(0b0110);
        named[blk] = out;
        return out;
    }
    Act block_c(const std::string& blk, const Act& x, int c7) {
        Act out = act(x.h, x.w, 768);
        Act t = act(x.h, x.w, c7), d = act(x.h, x.w, c7);
        conv_multi(x, {{blk + ".branch1x1", 192, &out, 0, false}, {blk + ".branch7x7_1", c7, &t, 0, false},
                       {blk + ".branch7x7dbl_1", c7, &d, 0, false}, {blk + ".branch_pool", 192, &out, 576, true}}, 1, 1, 1, 0, 0);
        fork(0b0110);
        if (pool_first) { cur_lane = 2; emit_pools(); }
        sm_share_den = 2;
        cur_lane = 0; Act u = conv(blk + ".branch7x7_2", t, c7, 1, 7, 1, 0, 3);
        conv_to(blk + ".branch7x7_3", u, 192, 7, 1, 1, 3, 0, &out, 192);
        cur_lane = 1; Act d2 = conv(blk + ".branch7x7dbl_2", d, c7, 7, 1, 1, 3, 0);
        Act d3 = conv(blk + ".branch7x7dbl_3", d2, c7, 1, 7, 1, 0, 3);
        rest_alone();            // the other branch (two convolutions) has finished by now: the rest of this one gets every SM
        Act d4 = conv(blk + ".branch7x7dbl_4", d3, c7, 7, 1, 1, 3, 0);
        conv_to(blk + ".branch7x7dbl_5", d4, 192, 1, 7, 1, 0, 3, &out, 384);
        sm_share_den = 1;
        cur_lane = pool_last ? 0 : 2; emit_pools();
        join(0b0110);
        named[blk] = out;
        return out;
    }
    Act block_d(const std::string& blk, const Act& x) {
        const int oh = (x.h - 3) / 2 + 1, ow = (x.w - 3) / 2 + 1;
        Act out = act(oh, ow, 320 + 192 + x.c);
        Act t = act(x.h, x.w, 192), s = act(x.h, x.w, 192);
        conv_multi(x, {{blk + ".branch3x3_1", 192, &t, 0, false}, {blk + ".branch7x7x3_1", 192, &s, 0, false}}, 1, 1, 1, 0, 0);
        fork(0b0110);
        sm_share_den = 2;
        cur_lane = 0; conv_to(blk + ".branch3x3_2", t, 320, 3, 3, 2, 0, 0, &out, 0);
        cur_lane = 1; Act s2 = conv(blk + ".branch7x7x3_2", s, 192, 1, 7, 1, 0, 3);
        rest_alone();            // the 3x3 branch is a single convolution: after it this branch is alone
        Act s3 = conv(blk + ".branch7x7x3_3", s2, 192, 7, 1, 1, 3, 0);
        conv_to(blk + ".branch7x7x3_4", s3, 192, 3, 3, 2, 0, 0, &out, 320);
        sm_share_den = 1;
        cur_lane = pool_last ? 0 : 2; maxpool_to(x, &out, 512);
        join(0b0110);
        named[blk] = out;
        return out;
    }
    Act block_e(const std::string& blk, const Act& x) {
        Act out = act(x.h, x.w, 2048);
        Act t = act(x.h, x.w, 384), d = act(x.h, x.w, 448);
        conv_multi(x, {{blk + ".branch1x1", 320, &out, 0, false}, {blk + ".branch3x3_1", 384, &t, 0, false},
                       {blk + ".branch3x3dbl_1", 448, &d, 0, false}, {blk + ".branch_pool", 192, &out, 1856, true}}, 1, 1, 1, 0, 0);
        fork(0b1110);
        if (pool_first) { cur_lane = 3; emit_pools(); }
        sm_share_den = 4;
        cur_lane = 0; conv_to(blk + ".branch3x3_2a", t, 384, 1, 3, 1, 0, 1, &out, 320);
        cur_lane = 1; conv_to(blk + ".branch3x3_2b", t, 384, 3, 1, 1, 1, 0, &out, 704);
        sm_share_den = 2;
        cur_lane = 2; Act u = conv(blk + ".branch3x3dbl_2", d, 384, 3, 3, 1, 1, 1);
        rest_alone();            // the 3x3 branches are done before the 448-channel 3x3 is: its two followers get every SM
        conv_to(blk + ".branch3x3dbl_3a", u, 384, 1, 3, 1, 0, 1, &out, 1088);
        conv_to(blk + ".branch3x3dbl_3b", u, 384, 3, 1, 1, 1, 0, &out, 1472);
        sm_share_den = 1;
        cur_lane = pool_last ? 0 : 3; emit_pools();
        join(0b1110);
        named[blk] = out;
        return out;
    }

    void build() {
        for (int i = 1; i < kLanes && !rc; ++i)
            if (cudaStreamCreateWithFlags(&lane_stream[i], cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&join_ev[i], cudaEventDisableTiming) != cudaSuccess) { rc = FTF_INTERNAL; *err = "stream/event creation failed"; }
        if (!rc && cudaEventCreateWithFlags(&fork_ev, cudaEventDisableTiming) != cudaSuccess) { rc = FTF_INTERNAL; *err = "event creation failed"; }
        if (rc) return;
        image = act(299, 299, 3);
        // stem layer 1: direct kernel
        {
            std::vector<float> w, b;
            if (!f32) { if ((rc = layer_weights("Conv2d_1a_3x3", 32, 3, 3, 3, &w, &b))) return; }
            for (int o = 0; o < 32 && !f32; ++o) {
                for (int k = 0; k < 27; ++k) stem.w[k][o] = bf16_value(w[(size_t)o * 27 + k]);
                stem.b[o] = b[o];
                stem_tc.bias[o] = b[o];
            }
            if (!f32) {
                std::vector<uint16_t> w32(32 * 32, 0);
                for (int o = 0; o < 32; ++o)
                    for (int k = 0; k < 27; ++k) w32[(size_t)o * 32 + k] = bf16_bits(w[(size_t)o * 27 + k]);
                stem_w32 = (__nv_bfloat16*)dev_alloc(w32.size() * 2, false);
                if (rc) return;
                cudaMemcpy(stem_w32, w32.data(), w32.size() * 2, cudaMemcpyHostToDevice);
            }
        }
        Act a1;
        if (f32) a1 = conv("Conv2d_1a_3x3", image, 32, 3, 3, 2, 0, 0);        // the generic fp32 kernel takes any channel count
        else {
            a1 = act(149, 149, 32);
            Op o{}; o.type = Op::kStem; o.in = image.p; o.out = a1.p; o.H = 299; o.W = 299; o.OH = 149; o.OW = 149; ops.push_back(o);
            named["Conv2d_1a_3x3"] = a1;
        }
        // images per group of the early layers (see split_early_layers); FTF_EARLY_GROUP=0: whole batch, layer by layer
        const int early_group = f32 ? 0 : (getenv("FTF_EARLY_GROUP") ? atoi(getenv("FTF_EARLY_GROUP")) : kEarlyGroup);
        const size_t early_begin = ops.size();
        defer_plans = early_group > 0 && early_group < max_batch;
        Act a2 = conv("Conv2d_2a_3x3", a1, 32, 3, 3, 1, 0, 0);
        Act a3 = conv("Conv2d_2b_3x3", a2, 64, 3, 3, 1, 1, 1);
        Act p1 = act(73, 73, 64);
        maxpool_to(a3, &p1, 0);
        named["pool1"] = p1;
        Act a4 = conv("Conv2d_3b_1x1", p1, 80, 1, 1, 1, 0, 0);
        Act a5 = conv("Conv2d_4a_3x3", a4, 192, 3, 3, 1, 0, 0);
        Act p2 = act(35, 35, 192);
        maxpool_to(a5, &p2, 0);
        named["pool2"] = p2;
        if (defer_plans) { defer_plans = false; split_early_layers(early_begin, early_group); if (rc) return; }
        Act x = block_a("Mixed_5b", p2, 32);
        x = block_a("Mixed_5c", x, 64);
        x = block_a("Mixed_5d", x, 64);
        x = block_b("Mixed_6a", x);
        x = block_c("Mixed_6b", x, 128);
        x = block_c("Mixed_6c", x, 160);
        x = block_c("Mixed_6d", x, 160);
        x = block_c("Mixed_6e", x, 192);
        x = block_d("Mixed_7a", x);
        x = block_e("Mixed_7b", x);
        x = block_e("Mixed_7c", x);
        Act gap = act(1, 1, x.c);
        { Op o{}; o.type = Op::kGap; o.in = x.p; o.out = gap.p; o.H = x.h; o.W = x.w; o.C = x.c; ops.push_back(o); }
        named["gap"] = gap;
        // fc as a 1x1 convolution over a 1x1 image: N rounded up to 16 columns, fp32 logits, bias, no ReLU
        if (rc) return;
        const int cols = (classes + 15) / 16 * 16;
        logits = (float*)dev_alloc((size_t)max_batch * cols * 4, true);
        std::vector<float> fw, fb;
        if ((rc = fc_weights(x.c, &fw, &fb))) return;
        if (f32) {
            std::vector<float> w32((size_t)cols * x.c, 0.0f);
            memcpy(w32.data(), fw.data(), fw.size() * 4);                       // rows [classes, cols) stay zero
            void* dw = dev_alloc(w32.size() * 4, false);
            float* db = (float*)dev_alloc((size_t)cols * 4, true);
            if (rc) return;
            cudaMemcpy(dw, w32.data(), w32.size() * 4, cudaMemcpyHostToDevice);
            cudaMemcpy(db, fb.data(), fb.size() * 4, cudaMemcpyHostToDevice);
            ConvSegment sg{}; sg.out = logits; sg.ld = cols; sg.coff = 0; sg.col_begin = 0; sg.flags = kSegBias | kSegF32;
            convs_f32.push_back(ConvF32{ConvGeometry{gap.p, max_batch, 1, 1, x.c, x.c, dw, db, cols, 1, 1, 1, 0, 0}, {sg}});
            Op op{}; op.type = Op::kConvF32; op.plan = (int)convs_f32.size() - 1; op.tag = "fc";
            ops.push_back(op);
            logits_ld = cols;
            return;
        }
        std::vector<uint16_t> w16((size_t)cols * x.c, 0);
        std::vector<float> bias(cols, 0.0f);
        {
            uint16_t* wp = w16.data();
            const float* src = fw.data();
            fill_parallel((int64_t)classes * x.c, [&](int64_t lo, int64_t hi) { for (int64_t i = lo; i < hi; ++i) wp[i] = bf16_bits(src[i]); });
            memcpy(bias.data(), fb.data(), fb.size() * 4);
        }
        void* dw = dev_alloc(w16.size() * 2, false);
        float* db = (float*)dev_alloc((size_t)cols * 4, false);
        if (rc) return;
        cudaMemcpy(dw, w16.data(), w16.size() * 2, cudaMemcpyHostToDevice);
        cudaMemcpy(db, bias.data(), (size_t)cols * 4, cudaMemcpyHostToDevice);
        ConvSegment sg{}; sg.out = logits; sg.ld = cols; sg.coff = 0; sg.col_begin = 0; sg.flags = kSegBias | kSegF32;
        ConvGeometry g{gap.p, max_batch, 1, 1, x.c, x.c, dw, db, cols, 1, 1, 1, 0, 0};
        plan_info.emplace_back();
        PlanInfo& pi = plan_info.back();
        pi.g = g; pi.segs = {sg}; pi.sm = sms; pi.tag = "fc";
        rc = conv_tc_plan_set(g, &sg, 1, sms, &pi.set, err, kPlanTune);
        if (rc) return;
        plans.push_back(pi.set.cands[pi.set.chosen]);
        Op op{}; op.type = Op::kConv; op.plan = (int)plans.size() - 1; op.tag = "fc";
        ops.push_back(op);
        logits_ld = cols;
    }
    int logits_ld = 0;
    // CTAs per SM of a pooling kernel on a branch stream (beside convolutions): see pools.cu.  FTF_BRANCH_POOL_CTAS overrides (measurement).
    int branch_pool_ctas = getenv("FTF_BRANCH_POOL_CTAS") ? atoi(getenv("FTF_BRANCH_POOL_CTAS")) : 16;
    int conv_priority = 0;          // launch priority of the convolutions (0: none; FTF_CONV_PRIORITY, a measurement switch)
    bool pool_last = getenv("FTF_POOL_LAST") != nullptr;        // measurement switch: a block's pool kernel behind the convolutions of its shortest branch (lane 0)
    bool pool_first = getenv("FTF_POOL_FIRST") != nullptr;      // measurement switch: issue a block's pool kernel before its convolutions
    int refine(std::string* e);
};

static int run_ops(InceptionV3::Impl& m, int nb, cudaStream_t s, std::string* err, const void* image, int image_dtype, size_t first, size_t last);

// In-situ refinement of the kernel choices.  conv_tc_plan timed each layer's candidates ALONE on the GPU; inside the forward
// pass the layers of an Inception block run side by side on shares of the SMs and compete for the L2 -> SM bandwidth, so the
// kernel that wins alone (say the one that re-reads its operands most but has the shortest critical path) is not always the
// one that wins there.  Here every alternative of every layer is swapped in and the WHOLE forward pass is timed at max_batch;
// a swap stays if the pass gets at least 1 % faster.  Layers of identical geometry (Mixed_6c and 6d, the three Mixed_5 blocks'
// towers) are switched together and share one remembered choice, so that later runners of the process -- which skip all
// timing -- run the same kernels.  All kernel variants produce the same bits (tests/test_gpu_conv.py), so this changes when a
// result is ready, never what it is.  One pass over the layers: about half a second at batch 64.
int InceptionV3::Impl::refine(std::string* e) {
    if (f32 || plans.empty()) return FTF_OK;
    bool any_open = false;
    for (auto& pi : plan_info) if (pi.set.cacheable && !pi.set.from_cache && pi.set.n > 1) any_open = true;
    if (!any_open) return FTF_OK;                     // an earlier runner of this process has done it
    cudaStream_t s = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    FTF_CUDA(cudaStreamSynchronize(cudaStreamLegacy), e);      // the arena's memsets and the weight uploads have finished
    FTF_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking), e);
    FTF_CUDA(cudaEventCreate(&e0), e);
    FTF_CUDA(cudaEventCreate(&e1), e);
    int rcode = FTF_OK;
    auto time_pass = [&]() -> float {
        float t[3];
        for (int i = -1; i < 3; ++i) {               // one warm-up, three timed; the median
            cudaEventRecord(e0, s);
            const int r = run_ops(*this, max_batch, s, e, image.p, FTF_BFLOAT16, 0, ops.size());
            cudaEventRecord(e1, s);
            if (r || cudaEventSynchronize(e1) != cudaSuccess) { rcode = r ? r : FTF_INTERNAL; return -1.0f; }
            if (i >= 0) cudaEventElapsedTime(&t[i], e0, e1);
        }
        if (t[0] > t[1]) std::swap(t[0], t[1]);
        if (t[1] > t[2]) std::swap(t[1], t[2]);
        if (t[0] > t[1]) std::swap(t[0], t[1]);
        return t[1];
    };
    const bool verbose = getenv("FTF_CONV_VERBOSE") != nullptr;
    // layers that share a remembered choice: same geometry, segment layout and SM share
    auto same = [](const PlanInfo& a, const PlanInfo& b) {
        if (a.sm != b.sm || a.segs.size() != b.segs.size() || a.g.n_max != b.g.n_max || a.g.H != b.g.H || a.g.W != b.g.W || a.g.Cin != b.g.Cin ||
            a.g.in_ld != b.g.in_ld || a.g.Cout != b.g.Cout || a.g.KH != b.g.KH || a.g.KW != b.g.KW || a.g.stride != b.g.stride ||
            a.g.pad_h != b.g.pad_h || a.g.pad_w != b.g.pad_w || (a.g.bias != nullptr) != (b.g.bias != nullptr)) return false;
        for (size_t i = 0; i < a.segs.size(); ++i)
            if (a.segs[i].ld != b.segs[i].ld || a.segs[i].coff != b.segs[i].coff || a.segs[i].col_begin != b.segs[i].col_begin || a.segs[i].flags != b.segs[i].flags) return false;
        return true;
    };
    float best = time_pass();
    const float first = best;
    std::vector<char> done(plan_info.size(), 0);
    for (size_t a = 0; a < plan_info.size() && best > 0; ++a) {
        if (done[a]) continue;
        std::vector<size_t> grp;
        for (size_t b = a; b < plan_info.size(); ++b) if (!done[b] && same(plan_info[a], plan_info[b])) { grp.push_back(b); done[b] = 1; }
        PlanInfo& lead = plan_info[a];
        if (!lead.set.cacheable || lead.set.from_cache || lead.set.n < 2) continue;
        int chosen_variant = lead.set.variant[lead.set.chosen];
        for (int c = 0; c < lead.set.n && best > 0; ++c) {
            const int v = lead.set.variant[c];
            if (v == chosen_variant) continue;
            std::vector<ConvPlan> keep;
            bool all = true;
            for (size_t b : grp) {                    // the same variant of every member
                int ci = -1;
                for (int k = 0; k < plan_info[b].set.n; ++k) if (plan_info[b].set.variant[k] == v) ci = k;
                if (ci < 0) { all = false; break; }
            }
            if (!all) continue;
            for (size_t b : grp) {
                keep.push_back(plans[b]);
                for (int k = 0; k < plan_info[b].set.n; ++k) if (plan_info[b].set.variant[k] == v) plans[b] = plan_info[b].set.cands[k];
            }
            const float t = time_pass();
            if (t > 0 && t < 0.99f * best) {
                if (verbose) fprintf(stderr, "refine: %s (x%zu): variant %d -> %d, pass %.1f -> %.1f us\n", lead.tag.c_str(), grp.size(), chosen_variant, v, best * 1e3f, t * 1e3f);
                best = t; chosen_variant = v;
            } else {
                for (size_t i = 0; i < grp.size(); ++i) plans[grp[i]] = keep[i];
            }
        }
        for (size_t b : grp) conv_tc_choice_put(plan_info[b].g, plan_info[b].segs.data(), (int)plan_info[b].segs.size(), plan_info[b].sm, chosen_variant);
    }
    if (verbose) fprintf(stderr, "refine: forward pass %.1f -> %.1f us at batch %d\n", first * 1e3f, best * 1e3f, max_batch);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(s);
    cudaGetLastError();
    return rcode;
}

InceptionV3::~InceptionV3() { delete impl_; }

int InceptionV3::create(int device, int num_classes, uint64_t seed, int max_batch, int compute_dtype,
                        std::unique_ptr<InceptionV3>* out, std::string* err, const char* weights_dir) {
    if (compute_dtype != FTF_BFLOAT16 && compute_dtype != FTF_FLOAT) {
        if (err) *err = "builtin:inception_v3: compute_dtype must be FTF_BFLOAT16 (tensor cores, fp32 accumulate) or FTF_FLOAT (fp32 throughout)";
        return FTF_INVALID_ARGUMENT;
    }
    if (max_batch <= 0 || max_batch > 1024 || num_classes <= 0 || num_classes > 65536) {
        if (err) *err = "inception_v3: max_batch must be in [1,1024] and num_classes in [1,65536]";
        return FTF_INVALID_ARGUMENT;
    }
    FTF_CUDA(cudaSetDevice(device), err);
    std::unique_ptr<InceptionV3> net(new InceptionV3());
    net->impl_ = new Impl();
    Impl& m = *net->impl_;
    m.device = device; m.sms = device_sm_count(device); m.max_batch = max_batch; m.classes = num_classes; m.seed = seed; m.err = err;
    m.f32 = compute_dtype == FTF_FLOAT;
    if (weights_dir && weights_dir[0]) {
        m.bundle.reset(new SavedModelLite());
        const int rc = m.bundle->load_variables_only(weights_dir, err);
        if (rc) return rc;
        m.weights_dir_loaded = true;
    }
    {
        // one label model is built at a time per process: the timing runs of two builds would disturb each other, and the
        // second build takes over the first one's kernel choices anyway
        static std::mutex build_mu;
        std::lock_guard<std::mutex> lock(build_mu);
        {
            int least = 0, greatest = 0;
            if (cudaDeviceGetStreamPriorityRange(&least, &greatest) != cudaSuccess) { cudaGetLastError(); greatest = 0; }
            const char* pe = getenv("FTF_CONV_PRIORITY");
            m.conv_priority = pe ? std::max(greatest, std::min(least, atoi(pe))) : 0;
        }
        m.build();
        if (m.rc) return m.rc;
        if (const char* rv = getenv("FTF_REVERSE")) {      // measurement switch: op indices (as in the FTF_TIMELINE table) to walk backwards
            for (const char* q = rv; *q;) {
                const int i = atoi(q);
                if (i >= 0 && i < (int)m.ops.size()) m.ops[i].reverse = 1;
                while (*q && *q != ',') ++q;
                if (*q) ++q;
            }
        }
        static const bool no_refine = getenv("FTF_NO_REFINE") != nullptr;      // measurement switch: isolated timing only
        if (!no_refine) { const int rr = m.refine(err); if (rr) return rr; }
    }
    if (const char* tl = getenv("FTF_TIMELINE")) {         // diagnostics only: kernels stamp their start/end times
        m.tl_after = atoi(tl);
        if (m.tl_after > 0) m.tl_dev = (unsigned long long*)m.dev_alloc(m.ops.size() * 4 * 8, true);
        if (m.rc) return m.rc;
    }
    // the uploads above went through the legacy default stream; wait for that stream only: a device-wide synchronisation
    // is refused (and breaks the capture) while another runner's thread is recording its CUDA graph on this GPU
    FTF_CUDA(cudaStreamSynchronize(cudaStreamLegacy), err);
    m.err = nullptr;
    m.bundle.reset();                     // the host copy of a weight file is not needed once everything is in HBM
    net->classes_ = num_classes; net->max_batch_ = max_batch;
    *out = std::move(net);
    return FTF_OK;
}

// Enqueues every op of the network for nb images: lane 0 on `s`, the other lanes on the net's own streams,
// fenced by events at each block's fork and join (the same calls record a graph when `s` is being captured).
// ops [first, last) are enqueued; the stem convolution (op 0) reads `image`, the caller's batch, where it lies.
static int run_ops(InceptionV3::Impl& m, int nb_all, cudaStream_t s, std::string* err, const void* image, int image_dtype, size_t first, size_t last) {
    auto stream_of = [&](int lane) { return lane == 0 ? s : m.lane_stream[lane]; };
    for (size_t oi = first; oi < last; ++oi) {
        const Op& op = m.ops[oi];
        const int nb = op.count > 0 ? std::max(0, std::min(op.count, nb_all - op.first)) : nb_all;     // this op's images
        if (nb == 0 && op.count > 0) continue;
        cudaStream_t st = stream_of(op.lane);
        unsigned long long* tl = m.tl_dev ? m.tl_dev + 4 * oi : nullptr;
        switch (op.type) {
        case Op::kFork:
            FTF_CUDA(cudaEventRecord(m.fork_ev, s), err);
            for (int l = 1; l < kLanes; ++l)
                if (op.lanes & (1 << l)) FTF_CUDA(cudaStreamWaitEvent(m.lane_stream[l], m.fork_ev, 0), err);
            break;
        case Op::kJoin:
            for (int l = 1; l < kLanes; ++l)
                if (op.lanes & (1 << l)) {
                    FTF_CUDA(cudaEventRecord(m.join_ev[l], m.lane_stream[l]), err);
                    FTF_CUDA(cudaStreamWaitEvent(s, m.join_ev[l], 0), err);
                }
            break;
        case Op::kStem: {
            const long long pix = (long long)nb * op.OH * op.OW;
            static const bool direct = getenv("FTF_STEM_DIRECT") != nullptr;      // measurement switch: the CUDA-core kernel
            if (!direct) {
                static const int stem_ctas = getenv("FTF_STEM_CTAS") ? std::max(1, std::min(kStemCtas, atoi(getenv("FTF_STEM_CTAS")))) : kStemCtas;    // measurement switch
                const long long tiles = (pix + 127) / 128, cap = (long long)m.sms * stem_ctas;
                const unsigned grid = (unsigned)(tiles < cap ? tiles : cap);
                if (image_dtype == FTF_FLOAT)
                    FTF_CUDA(launch_pdl(stem_conv_tc<float>, dim3(grid), dim3(128), 0, st, (const float*)image,
                                        (__nv_bfloat16*)op.out, (const __nv_bfloat16*)m.stem_w32, nb, op.H, op.W, op.OH, op.OW, m.stem_tc, tl), err);
                else
                    FTF_CUDA(launch_pdl(stem_conv_tc<__nv_bfloat16>, dim3(grid), dim3(128), 0, st, (const __nv_bfloat16*)image,
                                        (__nv_bfloat16*)op.out, (const __nv_bfloat16*)m.stem_w32, nb, op.H, op.W, op.OH, op.OW, m.stem_tc, tl), err);
                count_launch();
                break;
            }
            if (image_dtype == FTF_FLOAT)
                FTF_CUDA(launch_pdl(stem_conv3x3s2<float>, dim3((unsigned)((pix + 127) / 128)), dim3(128), 0, st, (const float*)image,
                                    (__nv_bfloat16*)op.out, nb, op.H, op.W, op.OH, op.OW, m.stem, tl), err);
            else
                FTF_CUDA(launch_pdl(stem_conv3x3s2<__nv_bfloat16>, dim3((unsigned)((pix + 127) / 128)), dim3(128), 0, st, (const __nv_bfloat16*)image,
                                    (__nv_bfloat16*)op.out, nb, op.H, op.W, op.OH, op.OW, m.stem, tl), err);
            count_launch();
            break;
        }
        case Op::kConv: {
            // FTF_CONV_PRIORITY=<p> (measurement switch): convolutions launched at a higher priority than the pool kernel of the
            // sibling branch.  A branch's next convolution (200 KB of shared memory: it needs a whole SM) starts late while a pool
            // kernel has CTAs pending, because those slip onto every SM a finished convolution frees.  With priority the block
            // scheduler prefers the convolution's CTAs and the gap shrinks from 8 to 4-5 us (a pool CTA already on the SM still has
            // to finish), but the step time did not move (1.306 / 1.307 ms), so the default is no priority.
            launch_priority() = m.conv_priority;
            const int rc = conv_tc_launch(m.plans[op.plan], nb, st, err, tl, op.reverse);
            launch_priority() = 0;
            if (rc) return rc;
            break;
        }
        case Op::kConvF32: {
            const ConvF32& c = m.convs_f32[op.plan];
            const int rc = f32::launch_conv(c.g, c.segs.data(), (int)c.segs.size(), nb, st, err);
            if (rc) return rc;
            break;
        }
        case Op::kMaxPool: {
            const char* in = (const char*)op.in + (size_t)op.first * op.H * op.W * op.C * 2;             // first > 0 in the bf16 mode only
            char* out = (char*)op.out + (size_t)op.first * op.OH * op.OW * op.out_ld * 2;
            const int rc = m.f32 ? f32::launch_maxpool3x3s2((const float*)op.in, (float*)op.out, nb, op.H, op.W, op.C, op.out_ld, op.out_off, st, err)
                                 : launch_maxpool3x3s2(in, out, nb, op.H, op.W, op.C, op.out_ld, op.out_off, st, err, tl, op.lane ? m.branch_pool_ctas : 16, op.reverse);
            if (rc) return rc;
            break;
        }
        case Op::kAvgPoolBiasRelu: {
            const int rc = m.f32 ? f32::launch_avgpool3x3_bias_relu((const float*)op.in, op.bias, (float*)op.out, nb, op.H, op.W, op.C, op.out_ld, op.out_off, st, err)
                                 : launch_avgpool3x3_bias_relu((const float*)op.in, op.bias, op.out, nb, op.H, op.W, op.C, op.out_ld, op.out_off, st, err, tl, op.lane ? m.branch_pool_ctas : 16);
            if (rc) return rc;
            break;
        }
        case Op::kGap: {
            const int rc = m.f32 ? f32::launch_global_avgpool((const float*)op.in, (float*)op.out, nb, op.H * op.W, op.C, st, err)
                                 : launch_global_avgpool(op.in, op.out, nb, op.H * op.W, op.C, st, err, tl);
            if (rc) return rc;
            break;
        }
        }
    }
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

// FTF_TIMELINE=<k>: after the k-th forward, print when each kernel of that forward started, got past its dependency
// wait and ended (microseconds from the first start), taken from %globaltimer stamps written by the kernels.
static void dump_timeline(InceptionV3::Impl& m, int nb) {
    std::vector<unsigned long long> t(m.ops.size() * 4);
    if (cudaMemcpy(t.data(), m.tl_dev, t.size() * 8, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return; }
    unsigned long long t0 = ~0ull, t1 = 0;
    for (size_t k = 0; k < m.ops.size(); ++k) {
        if (t[4 * k] < t0) t0 = t[4 * k];
        if (t[4 * k + 1] > t1) t1 = t[4 * k + 1];
    }
    static const char* kNames[] = {"stem", "conv", "maxpool", "avgpool", "gap", "fork", "join"};
    fprintf(stderr, "timeline batch %d: %zu ops, %.1f us first start -> last end\n", nb, m.ops.size(), (t1 - t0) * 1e-3);
    fprintf(stderr, "%4s %4s %-8s %9s %9s %9s %8s  %s\n", "op", "lane", "kind", "start", "go", "end", "busy", "layer");
    double busy_sum = 0;
    for (size_t k = 0; k < m.ops.size(); ++k) {
        const Op& op = m.ops[k];
        if (op.type == Op::kFork || op.type == Op::kJoin) { fprintf(stderr, "%4zu %4s %-8s\n", k, "-", kNames[op.type]); continue; }
        const double a = (t[4 * k] - t0) * 1e-3, g = (t[4 * k + 2] - t0) * 1e-3, b = (t[4 * k + 1] - t0) * 1e-3;
        busy_sum += b - g;
        std::string tag = op.tag;
        if (op.type == Op::kConv) {
            const ConvPlan& pl = m.plans[op.plan];
            const bool hm = pl.kind ? pl.wm.p.halo_mode : pl.pm.p.halo_mode;
            tag += pl.kind ? " [wm" : (pl.pm.p.pair ? " [pm pair" : " [pm");
            tag += hm ? " halo]" : "]";
        }
        fprintf(stderr, "%4zu %4d %-8s %9.1f %9.1f %9.1f %8.1f  %s\n", k, op.lane, kNames[op.type], a, g, b, b - g, tag.c_str());
    }
    fprintf(stderr, "sum of (end - go) over kernels: %.1f us\n", busy_sum);
}

int InceptionV3::forward(const void* images, int in_dtype, int64_t n, float* logits, cudaStream_t s, std::string* err) {
    Impl& m = *impl_;
    if (n < 0 || n > m.max_batch) { if (err) *err = "inception_v3: batch exceeds max_batch"; return FTF_INVALID_ARGUMENT; }
    if (in_dtype != FTF_FLOAT && in_dtype != FTF_BFLOAT16) { if (err) *err = "inception_v3: images must be FLOAT or BFLOAT16"; return FTF_INVALID_ARGUMENT; }
    if (n == 0) return FTF_OK;
    const int nb = (int)n;
    if (m.f32) {
        // precision mode: fp32 images only; they are copied into the network's own input tensor (its convolutions were
        // planned on fixed buffers) and the ops run as plain stream launches
        if (in_dtype != FTF_FLOAT) { if (err) *err = "inception_v3 (compute_dtype FLOAT) takes FLOAT images"; return FTF_INVALID_ARGUMENT; }
        FTF_CUDA(cudaMemcpyAsync(m.image.p, images, (size_t)n * 299 * 299 * 3 * 4, cudaMemcpyDeviceToDevice, s), err);
        const int rc = run_ops(m, nb, s, err, m.image.p, FTF_FLOAT, 0, m.ops.size());
        if (rc) return rc;
        FTF_CUDA(cudaMemcpy2DAsync(logits, (size_t)m.classes * 4, m.logits, (size_t)m.logits_ld * 4, (size_t)m.classes * 4, (size_t)n,
                                   cudaMemcpyDeviceToDevice, s), err);
        return FTF_OK;
    }
    // The stem convolution reads the caller's images in place and is launched on its own, in front of the graph: a graph
    // fixes its kernels' pointers, and copying every batch (34 MB in, 34 MB out) into a fixed buffer first cost about 1 %
    // of the step.
    const size_t n_ops = m.ops.size();
    const size_t stem_ops = (n_ops && m.ops[0].type == Op::kStem) ? 1 : 0;
    // The whole forward (about 80 kernels over 4 streams) is one CUDA graph per batch size: recorded on first use,
    // replayed afterwards.  FTF_NO_GRAPH=1 keeps the plain stream launches (debugging, profilers that want them).
    static const bool use_graph = !(getenv("FTF_NO_GRAPH") && getenv("FTF_NO_GRAPH")[0] == '1');
    const bool tl_now = m.tl_dev && ++m.tl_calls == m.tl_after;
    if (tl_now) {
        std::vector<unsigned long long> init(m.ops.size() * 4);
        for (size_t k = 0; k < m.ops.size(); ++k) { init[4 * k] = ~0ull; init[4 * k + 1] = 0; init[4 * k + 2] = ~0ull; init[4 * k + 3] = 0; }
        FTF_CUDA(cudaStreamSynchronize(s), err);
        FTF_CUDA(cudaMemcpy(m.tl_dev, init.data(), init.size() * 8, cudaMemcpyHostToDevice), err);
    }
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    FTF_CUDA(cudaStreamIsCapturing(s, &cap), err);
    if (!use_graph || cap != cudaStreamCaptureStatusNone) {
        const int rc = run_ops(m, nb, s, err, images, in_dtype, 0, n_ops);
        if (rc) return rc;
    } else {
        { const int rc = run_ops(m, nb, s, err, images, in_dtype, 0, stem_ops); if (rc) return rc; }
        // Micro-batches closed by the deadline or by a blocking poll have arbitrary sizes.  One graph per distinct size would
        // pay a capture + instantiation of ~80 kernels on the record path for every new size and keep up to max_batch graphs
        // alive; instead the sizes are bucketed (at most 16 buckets) and a batch replays the graph of the bucket above it.  The
        // surplus rows are rows of the activation arena that hold an earlier batch's values: they are computed and never
        // read back -- no kernel mixes records, so the real rows do not change by a bit.
        const int step = m.max_batch <= 128 ? 8 : (m.max_batch / 16 + 7) / 8 * 8;
        const int gb = std::min(m.max_batch, (nb + step - 1) / step * step);
        auto it = m.graphs.find(gb);
        if (it == m.graphs.end()) {
            const uint64_t before = launches_so_far();
            FTF_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal), err);
            const int rc = run_ops(m, gb, s, err, images, in_dtype, stem_ops, n_ops);
            cudaGraph_t graph = nullptr;
            const cudaError_t e = cudaStreamEndCapture(s, &graph);
            if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
            if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture", err);
            cudaGraphExec_t exec = nullptr;
            const cudaError_t e2 = cudaGraphInstantiate(&exec, graph, 0);
            cudaGraphDestroy(graph);
            if (e2 != cudaSuccess) return cuda_fail(e2, "cudaGraphInstantiate", err);
            m.graph_kernels[gb] = (int)(launches_so_far() - before);
            count_launch(-m.graph_kernels[gb]);          // recording launched nothing
            it = m.graphs.emplace(gb, exec).first;
        }
        FTF_CUDA(cudaGraphLaunch(it->second, s), err);
        count_launch(m.graph_kernels[gb]);
    }
    if (tl_now) {
        FTF_CUDA(cudaStreamSynchronize(s), err);
        dump_timeline(m, nb);
    }
    // logits: [n, classes] out of the padded [n, logits_ld] result
    FTF_CUDA(cudaMemcpy2DAsync(logits, (size_t)m.classes * 4, m.logits, (size_t)m.logits_ld * 4, (size_t)m.classes * 4, (size_t)n,
                               cudaMemcpyDeviceToDevice, s), err);
    return FTF_OK;
}

int InceptionV3::export_weights(const std::string& dir, std::string* err) {
    Impl& m = *impl_;
    if (m.weights_dir_loaded) { if (err) *err = "export_weights: this network was loaded from a weight file; copy that file instead"; return FTF_FAILED_PRECONDITION; }
    TensorBundleWriter wr;
    int rc = wr.open(dir, err);
    std::vector<float> w, b;
    for (size_t i = 0; i < m.layer_specs.size() && !rc; ++i) {
        const auto& L = m.layer_specs[i];
        conv_weights(m.seed, L.name, L.cout, L.kh, L.kw, L.cin, &w, &b);
        rc = wr.add_f32(L.name + "/weights", {L.cout, L.kh, L.kw, L.cin}, w.data(), err);
        if (!rc) rc = wr.add_f32(L.name + "/biases", {L.cout}, b.data(), err);
    }
    if (!rc) {
        std::string* saved = m.err; m.err = err;
        rc = m.fc_weights(2048, &w, &b);
        m.err = saved;
    }
    if (!rc) rc = wr.add_f32("fc/weights", {m.classes, 2048}, w.data(), err);
    if (!rc) rc = wr.add_f32("fc/biases", {m.classes}, b.data(), err);
    if (!rc) rc = wr.finish(err);
    return rc;
}

int InceptionV3::activation(const std::string& name, int64_t n, float* host_out, int64_t cap, int dims[4], std::string* err) {
    Impl& m = *impl_;
    auto it = m.named.find(name);
    if (it == m.named.end()) { if (err) *err = "no activation named " + name; return FTF_NOT_FOUND; }
    const Act& a = it->second;
    const int64_t elems = n * a.h * a.w * a.c;
    dims[0] = (int)n; dims[1] = a.h; dims[2] = a.w; dims[3] = a.c;
    if (!host_out) return FTF_OK;
    if (elems > cap) { if (err) *err = "activation buffer too small"; return FTF_OUT_OF_RANGE; }
    if (m.f32) { FTF_CUDA(cudaMemcpy(host_out, a.p, (size_t)elems * 4, cudaMemcpyDeviceToHost), err); return FTF_OK; }
    std::vector<uint16_t> tmp((size_t)elems);
    FTF_CUDA(cudaMemcpy(tmp.data(), a.p, (size_t)elems * 2, cudaMemcpyDeviceToHost), err);
    for (int64_t i = 0; i < elems; ++i) {
        const uint32_t u = (uint32_t)tmp[(size_t)i] << 16;
        memcpy(host_out + i, &u, 4);
    }
    return FTF_OK;
}

}  // namespace ftf
