// K2  Cast(u8->f32) -> ResizeBilinear(align_corners=false) -> Sub(mean) -> Div(scale), fused.
//
// Reference: flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples/
// inception/ImageNormalization.scala:57-65 builds Div(Sub(ResizeBilinear(ExpandDims(Cast(DecodeJpeg))))),
// ops from .../examples/common/GraphBuilder.java:21-39; five materialised fp32 tensors per record on the
// reference's CPU path, one pass here.  DecodeJpeg sits before this kernel (raw RGB records).
//
// Arithmetic follows TF-1.1's CPU ResizeBilinear exactly and in the same order, every operation
// individually rounded (__f*_rn intrinsics are never fused), so results are bit-identical:
//   scale = in/(float)out; p = (float)i*scale; lo = (int)p; hi = min(lo+1, in-1); l = p - lo
//   top = tl + (tr-tl)*xl; bot = bl + (br-bl)*xl; v = top + (bot-top)*yl; out = (v-mean)/scale
//
// Layout: in u8 [n,in_h,in_w,3] (HWC rows of in_w*3 bytes), out f32/bf16 [n,out_h,out_w,3].
// HBM-bound: each input byte and each output element crosses HBM once.  A CTA owns a band of `th` (5)
// full-width output rows of one image: the input rows the band touches are fetched by TMA
// (cp.async.bulk.tensor, <=256-byte wide column chunks of a 3-D tensor map) into shared memory, the
// 4-tap lerp reads shared memory, and the band's output is one contiguous, fully coalesced range.
#include "common.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <math.h>

namespace ftf {

namespace {

constexpr int kThreads = 256;
constexpr int kCols = 4;           // output columns a thread carries through a band

// exact u8 -> f32 without the quarter-rate I2F: 2^23 + b has b in its low mantissa bits
__device__ __forceinline__ float u8_to_f32(uint32_t b) { return __fsub_rn(__uint_as_float(0x4B000000u | b), 8388608.0f); }
// one byte of shared memory by its 32-bit shared-window address, zero-extended
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

struct ResizeParams {
    int in_h, in_w, out_h, out_w;
    float hs, ws;            // in/(float)out
    float mean, scale, inv_scale;
    int th;                  // output rows per band
    int bands;               // bands per image
    int br;                  // input rows fetched per band (TMA box height)
    int cw;                  // chunk width in bytes (TMA box width)
    float inv_cw;            // 1/cw
    int n_chunks;
    int chunk_stride;        // bytes between chunk regions in smem (128-byte multiple)
    long long n;             // images
    int stages;              // shared-memory stages per CTA (2 or 3)
    int step_img, step_band; // gridDim.x bands further on: += step_img images, += step_band bands (wrapping at `bands`)
};

template <typename T> __device__ __forceinline__ T to_out(float v);
template <> __device__ __forceinline__ float to_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 to_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <bool kMulRecip>
__device__ __forceinline__ float finish(float top, float bot, float yl, const ResizeParams& p) {
    const float v = __fadd_rn(top, __fmul_rn(__fsub_rn(bot, top), yl));
    const float d = __fsub_rn(v, p.mean);
    return kMulRecip ? __fmul_rn(d, p.inv_scale) : __fdiv_rn(d, p.scale);
}

// ---- TMA-staged band kernel -------------------------------------------------------------------
// Persistent: a CTA of kBS threads walks bands blockIdx.x, blockIdx.x + gridDim.x, ... through two shared-memory stages guarded
// by full/empty mbarriers: warp 0 refills a stage as soon as all warps have released it, and no warp ever waits for another
// one (no __syncthreads in the loop), so the warps drift apart and hide each other's shared-memory latency.
//
// A thread owns kN output columns (col = threadIdx.x + j*kBS, kN warp-uniform): their x terms and shared-memory offsets are
// computed ONCE per kernel -- they do not depend on the band -- and live in registers; the row loop is branch-free
// straight-line code: per output value 4 address adds, 4 byte loads, the conversions, the 11 rounded fp32 operations of the
// reference's formula and one store.  kBS is chosen by the launcher so that out_w*3 columns split into whole groups with as
// few idle lanes as possible among 256, 224 and 192 threads (four CTAs per SM).
// u8 -> f32 uses the 2^23 trick: m(b) = bits 0x4B000000|b = 2^23 + b exactly.  A tap that only enters a difference needs
// no conversion at all, since m(tr) - m(tl) = tr - tl exactly; the left taps are converted with one FADD each.
template <int kN, int kBS, bool kMulRecip, bool kPartial, typename OutT>
__device__ __forceinline__ void run_bands(const ResizeParams& p, OutT* __restrict__ out, long long total_bands, int stage_bytes,
                                          uint32_t smem_a, uint64_t* full, uint64_t* empty, const CUtensorMap* tmap, uint8_t* smem, uint4* rows_sm) {
    const int ow3 = p.out_w * 3;
    const int c0 = (int)threadIdx.x;
    uint32_t off_lo[kN], off_hi[kN];
    float xl[kN];
    bool live[kN];
#pragma unroll
    for (int j = 0; j < kN; ++j) {
        live[j] = c0 + j * kBS < ow3;
        const int col = min(c0 + j * kBS, ow3 - 1);                        // a dead lane recomputes the last column, unstored
        const int px = col / 3, c = col - px * 3;
        const float fx = __fmul_rn((float)px, p.ws);
        const int lo = (int)fx;
        const int hi = min(lo + 1, p.in_w - 1);
        xl[j] = __fsub_rn(fx, (float)lo);
        const int blo = lo * 3 + c, bhi = hi * 3 + c;
        // chunk index = byte / cw without an integer division: (b + 0.5)/cw is at least 0.5/cw away from an integer
        const int klo = (int)(((float)blo + 0.5f) * p.inv_cw), khi = (int)(((float)bhi + 0.5f) * p.inv_cw);
        off_lo[j] = smem_a + (uint32_t)(klo * p.chunk_stride + (blo - klo * p.cw));
        off_hi[j] = smem_a + (uint32_t)(khi * p.chunk_stride + (bhi - khi * p.cw));
    }
    // (image, band) of a CTA's bands advance by host-computed steps: the 64-bit division they replace cost more than a
    // whole output row per band
    auto advance = [&](int& img, int& band) {
        img += p.step_img; band += p.step_band;
        if (band >= p.bands) { band -= p.bands; ++img; }
    };
    auto issue = [&](int img, int band, int stage) {   // warp 0, all lanes, one elected lane per instruction (see ptx.cuh)
        const int iy0 = (int)__fmul_rn((float)(band * p.th), p.hs);
        ptx::mbar_expect_tx_elect(&full[stage], (uint32_t)(p.n_chunks * p.cw * p.br));
        for (int k = 0; k < p.n_chunks; ++k)
            ptx::tma_load_3d_elect(smem + (size_t)stage * stage_bytes + (size_t)k * p.chunk_stride, tmap, &full[stage], k * p.cw, iy0, (int)img);
    };
    int it = 0;
    int img = (int)(blockIdx.x / (unsigned)p.bands), band = (int)(blockIdx.x % (unsigned)p.bands);
    int nimg = img, nband = band;                      // the band p.stages steps ahead, which warp 0 fetches
    for (int s = 0; s < p.stages; ++s) advance(nimg, nband);
    int stage = 0;
    uint32_t phase = 0;
    for (unsigned blk = blockIdx.x; blk < (unsigned)total_bands; blk += gridDim.x, ++it, advance(img, band), advance(nimg, nband)) {
        const int oy0 = band * p.th;
        const int rows = min(p.th, p.out_h - oy0);
        const int iy0 = (int)__fmul_rn((float)oy0, p.hs);  // first input row of the band
        ptx::mbar_wait(&full[stage], phase);
        const uint32_t sbase = (uint32_t)(stage * stage_bytes);
        OutT* orow = out + ((size_t)img * p.out_h + oy0) * ow3 + c0;
        // the row terms of the band (at most 8 rows): lane r computes row r's once and leaves them in the warp's corner of
        // shared memory; the row loop reads them back as broadcasts.  (Fetching them by shuffle made the loop body a
        // possibly-divergent region for the compiler, which then kept uniform values -- the global-memory descriptor among
        // them -- in vector registers and moved them to uniform ones again before every store.)
        if ((threadIdx.x & 31) < 8) {
            const float py = __fmul_rn((float)(oy0 + (int)(threadIdx.x & 31)), p.hs);
            const int ylo = (int)py;
            const int yhi = min(ylo + 1, p.in_h - 1);
            uint4 t;
            t.x = sbase + (uint32_t)((ylo - iy0) * p.cw);
            t.y = sbase + (uint32_t)((yhi - iy0) * p.cw);
            t.z = __float_as_uint(__fsub_rn(py, (float)ylo));
            t.w = 0;
            rows_sm[threadIdx.x & 31] = t;
        }
        __syncwarp();
        // When output row r+1 starts at the input row where row r ended (ylo(r+1) == yhi(r): four rows in ten at this scale),
        // its upper horizontal lerp is the previous row's lower one, computed from the same bytes by the same operations:
        // it is carried over instead of being loaded and computed again.
        float carry[kN];
        uint32_t prev_r1 = 0xffffffffu;
        for (int r = 0; r < rows; ++r, orow += ow3) {
            const uint4 t = rows_sm[r];
            const uint32_t r0 = t.x, r1 = t.y;
            const float yl = __uint_as_float(t.z);
            if (r0 == prev_r1) {
#pragma unroll
                for (int j = 0; j < kN; ++j) {
                    const float mbl = __uint_as_float(0x4B000000u | lds_u8(r1 + off_lo[j])), mbr = __uint_as_float(0x4B000000u | lds_u8(r1 + off_hi[j]));
                    const float bl = __fsub_rn(mbl, 8388608.0f);
                    const float top = carry[j];
                    const float bot = __fadd_rn(bl, __fmul_rn(__fsub_rn(mbr, mbl), xl[j]));
                    carry[j] = bot;
                    const OutT v = to_out<OutT>(finish<kMulRecip>(top, bot, yl, p));
                    if (!kPartial || j < kN - 1 || live[j]) orow[j * kBS] = v;
                }
            } else {
#pragma unroll
                for (int j = 0; j < kN; ++j) {
                    const float mtl = __uint_as_float(0x4B000000u | lds_u8(r0 + off_lo[j])), mtr = __uint_as_float(0x4B000000u | lds_u8(r0 + off_hi[j]));
                    const float mbl = __uint_as_float(0x4B000000u | lds_u8(r1 + off_lo[j])), mbr = __uint_as_float(0x4B000000u | lds_u8(r1 + off_hi[j]));
                    const float tl = __fsub_rn(mtl, 8388608.0f), bl = __fsub_rn(mbl, 8388608.0f);
                    const float top = __fadd_rn(tl, __fmul_rn(__fsub_rn(mtr, mtl), xl[j]));
                    const float bot = __fadd_rn(bl, __fmul_rn(__fsub_rn(mbr, mbl), xl[j]));
                    carry[j] = bot;
                    const OutT v = to_out<OutT>(finish<kMulRecip>(top, bot, yl, p));
                    if (!kPartial || j < kN - 1 || live[j]) orow[j * kBS] = v;     // only a warp's last group can hold dead lanes
                }
            }
            prev_r1 = r1;
        }
        // this warp is done with the stage; warp 0 refills it (band it+stages) once every warp has said so
        __syncwarp();
        if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&empty[stage]);
        if (threadIdx.x < 32 && (unsigned long long)blk + (unsigned long long)p.stages * gridDim.x < (unsigned long long)total_bands) {
            ptx::mbar_wait(&empty[stage], phase);
            issue(nimg, nband, stage);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
}

#ifndef FTF_K2_CTAS
#define FTF_K2_CTAS 4          /* persistent CTAs per SM the band kernel is built for */
#endif
template <int kBS, bool kMulRecip, typename OutT>
__global__ void __launch_bounds__(kBS, (kBS <= 256 ? FTF_K2_CTAS : 3))
resize_norm_band_tma(const __grid_constant__ CUtensorMap tmap, OutT* __restrict__ out, const ResizeParams p, long long total_bands,
                     int stage_bytes) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[3], empty[3];
    __shared__ uint4 rows_all[(kBS / 32) * 8];          // per warp: the row terms of its current band
    // The warp index comes out of a shuffle so that the compiler knows it -- and everything derived from it: the warp's first
    // column, its number of column groups, the switch below -- to be warp-uniform.  Otherwise the whole band loop counts as a
    // possibly divergent region: kernel parameters, shared-memory row offsets and the global-memory descriptor are then kept
    // in vector registers and moved to uniform ones (R2UR) in front of every load and store that wants them there.
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    uint4* rows_sm = rows_all + warp * 8;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 3; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], kBS / 32); }
        ptx::fence_mbar_init();
    }
    __syncthreads();
    if (warp == 0) {                                            // the first bands, one per stage
        for (int s = 0; s < p.stages; ++s) {
            const long long blk = (long long)blockIdx.x + (long long)s * gridDim.x;
            if (blk >= total_bands) break;
            const long long img = blk / p.bands;
            const int band = (int)(blk - img * p.bands);
            const int iy0 = (int)__fmul_rn((float)(band * p.th), p.hs);
            ptx::mbar_expect_tx_elect(&full[s], (uint32_t)(p.n_chunks * p.cw * p.br));
            for (int k = 0; k < p.n_chunks; ++k)
                ptx::tma_load_3d_elect(smem + (size_t)s * stage_bytes + (size_t)k * p.chunk_stride, &tmap, &full[s], k * p.cw, iy0, (int)img);
        }
    }
    const uint32_t smem_a = ptx::smem_u32(smem);             // 32-bit shared-window addresses: no generic-pointer arithmetic
    const int warp_c0 = warp * 32;                            // first column of this warp
    const int ow3 = p.out_w * 3;
    const int nj = warp_c0 < ow3 ? min(kCols, (ow3 - warp_c0 + kBS - 1) / kBS) : 0;   // columns with at least one live lane in this warp
    // a warp whose last column group is cut by the end of the row stores under a predicate; every other warp runs code
    // without one (the predicate cost an ISETP and two R2UR per stored value)
    const bool partial = nj > 0 && warp_c0 + (nj - 1) * kBS + 32 > ow3;
#define FTF_K2_RUN(N)                                                                                                   \
    if (partial) run_bands<N, kBS, kMulRecip, true>(p, out, total_bands, stage_bytes, smem_a, full, empty, &tmap, smem, rows_sm);  \
    else run_bands<N, kBS, kMulRecip, false>(p, out, total_bands, stage_bytes, smem_a, full, empty, &tmap, smem, rows_sm);
    switch (nj) {
    case 4: FTF_K2_RUN(4) break;
    case 3: FTF_K2_RUN(3) break;
    case 2: FTF_K2_RUN(2) break;
    default:                      // also warps without columns: they keep the barriers in step
        run_bands<1, kBS, kMulRecip, true>(p, out, total_bands, stage_bytes, smem_a, full, empty, &tmap, smem, rows_sm);
        break;
    }
#undef FTF_K2_RUN
}

// ---- direct kernel: any alignment, any size; one thread per output element ---------------------
template <bool kMulRecip, typename OutT>
__global__ void __launch_bounds__(kThreads)
resize_norm_direct(const uint8_t* __restrict__ in, OutT* __restrict__ out, const ResizeParams p) {
    const long long ow3 = (long long)p.out_w * 3;
    const long long per_img = (long long)p.out_h * ow3;
    const long long total = p.n * per_img;
    const long long stride = (long long)gridDim.x * kThreads;
    for (long long e = (long long)blockIdx.x * kThreads + threadIdx.x; e < total; e += stride) {
        const long long img = e / per_img;
        const long long rem = e - img * per_img;
        const int oy = (int)(rem / ow3);
        const int col = (int)(rem - (long long)oy * ow3);
        const int px = col / 3, c = col - px * 3;
        const float py = __fmul_rn((float)oy, p.hs);
        const int ylo = (int)py, yhi = min(ylo + 1, p.in_h - 1);
        const float yl = __fsub_rn(py, (float)ylo);
        const float fx = __fmul_rn((float)px, p.ws);
        const int xlo = (int)fx, xhi = min(xlo + 1, p.in_w - 1);
        const float xl = __fsub_rn(fx, (float)xlo);
        const uint8_t* base = in + (size_t)img * p.in_h * p.in_w * 3;
        const uint8_t* r0 = base + (size_t)ylo * p.in_w * 3;
        const uint8_t* r1 = base + (size_t)yhi * p.in_w * 3;
        const float tl = (float)__ldg(r0 + xlo * 3 + c), tr = (float)__ldg(r0 + xhi * 3 + c);
        const float bl = (float)__ldg(r1 + xlo * 3 + c), br = (float)__ldg(r1 + xhi * 3 + c);
        const float top = __fadd_rn(tl, __fmul_rn(__fsub_rn(tr, tl), xl));
        const float bot = __fadd_rn(bl, __fmul_rn(__fsub_rn(br, bl), xl));
        out[e] = to_out<OutT>(finish<kMulRecip>(top, bot, yl, p));
    }
}

bool exact_reciprocal(float s, float* inv) {
    if (!(s > 0.0f) || isinf(s)) return false;
    int e = 0;
    const float m = frexpf(s, &e);
    if (m != 0.5f || e < -60 || e > 60) return false;
    *inv = 1.0f / s;
    return true;
}

// Input rows a band of `th` output rows touches, worst case over the bands of one image.
int band_input_rows(int th, int in_h, int out_h, float hs) {
    int worst = 1;
    for (int oy0 = 0; oy0 < out_h; oy0 += th) {
        const int last = (oy0 + th < out_h ? oy0 + th : out_h) - 1;
        const int lo = (int)((float)oy0 * hs);
        const int l2 = (int)((float)last * hs);
        const int hi = l2 + 1 < in_h - 1 ? l2 + 1 : in_h - 1;
        if (hi - lo + 1 > worst) worst = hi - lo + 1;
    }
    return worst;
}

template <typename OutT>
int launch_typed(const uint8_t* in, ResizeParams p, OutT* out, bool recip, bool use_tma, size_t smem,
                 const CUtensorMap* tmap, int sms, cudaStream_t s, std::string* err) {
    if (use_tma) {
        // block size: out_w*3 columns in kc <= kCols whole groups of bs threads with the fewest idle lanes
        const int ow3 = p.out_w * 3;
        // (320- and 288-thread blocks waste fewer lanes on 897 columns but leave room for only three CTAs per SM: measured slower)
        static const int kSizes[] = {256, 224, 192};
        int bs = 256;
        long long best_waste = 1ll << 40;
        for (int cand : kSizes) {
            const int kc = (ow3 + cand - 1) / cand;
            if (kc > kCols) continue;
            const long long waste = (long long)kc * cand - ow3;
            if (waste * bs < best_waste * cand) { best_waste = waste; bs = cand; }        // relative waste
        }
        const long long bands = p.n * p.bands;
        const long long resident = (long long)sms * FTF_K2_CTAS;                          // persistent CTAs, two stages each
        const unsigned grid = (unsigned)(bands < resident ? bands : resident);
        p.step_img = (int)(grid / (unsigned)p.bands); p.step_band = (int)(grid % (unsigned)p.bands);
#define FTF_K2_LAUNCH(BS)                                                                                              \
        {                                                                                                              \
            auto kern = recip ? resize_norm_band_tma<BS, true, OutT> : resize_norm_band_tma<BS, false, OutT>;          \
            FTF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(p.stages * smem)), err); \
            kern<<<grid, BS, p.stages * smem, s>>>(*tmap, out, p, bands, (int)smem);                                   \
        }
        switch (bs) {
        case 224: FTF_K2_LAUNCH(224) break;
        case 192: FTF_K2_LAUNCH(192) break;
        default: FTF_K2_LAUNCH(256) break;
        }
#undef FTF_K2_LAUNCH
    } else {
        auto kern = recip ? resize_norm_direct<true, OutT> : resize_norm_direct<false, OutT>;
        const long long total = p.n * p.out_h * p.out_w * 3;
        long long want = (total + kThreads - 1) / kThreads;
        const long long cap = (long long)sms * 8 * 4;
        kern<<<(unsigned)(want < cap ? want : cap), kThreads, 0, s>>>(in, out, p);
    }
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

}  // namespace

int launch_resize_norm(const uint8_t* in, int64_t n, int in_h, int in_w, int out_h, int out_w, float mean,
                       float scale, void* out, int out_dtype, cudaStream_t s, std::string* err) {
    if (n < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || (n > 0 && (!in || !out))) {
        if (err) *err = "resize_norm: bad shape or null buffer";
        return FTF_INVALID_ARGUMENT;
    }
    if (out_dtype != FTF_FLOAT && out_dtype != FTF_BFLOAT16) {
        if (err) *err = "resize_norm: output dtype must be FLOAT or BFLOAT16";
        return FTF_INVALID_ARGUMENT;
    }
    if (n == 0) return FTF_OK;
    int dev = 0;
    FTF_CUDA(cudaGetDevice(&dev), err);
    const int sms = device_sm_count(dev);

    ResizeParams p{};
    p.in_h = in_h; p.in_w = in_w; p.out_h = out_h; p.out_w = out_w;
    p.hs = (float)in_h / (float)out_h;
    p.ws = (float)in_w / (float)out_w;
    p.mean = mean; p.scale = scale; p.inv_scale = 0.0f; p.n = n;
    const bool recip = exact_reciprocal(scale, &p.inv_scale);

    // TMA path preconditions: 16-byte aligned base and row pitch, dimensions within tensor-map limits.
    const long long row_bytes = (long long)in_w * 3;
    bool use_tma = ((uintptr_t)in & 15u) == 0 && row_bytes % 16 == 0 && n * (long long)((out_h + 0)) < (1ll << 31) &&
                   n <= 0x7fffffffLL;
    size_t smem = 0;
    CUtensorMap tmap;
    if (use_tma) {
        p.n_chunks = (int)((row_bytes + 255) / 256);
        p.cw = (int)(((row_bytes + p.n_chunks - 1) / p.n_chunks + 15) / 16 * 16);
        bool fits = false;
        // FTF_K2_TH / FTF_K2_STAGES: measurement switches (rows per band, stages per CTA)
        static const int th_max = getenv("FTF_K2_TH") ? atoi(getenv("FTF_K2_TH")) : 5;   // measured on 480->299 rows: 8: 0.101 ms, 7: 0.099, 6: 0.096, 5: 0.092, 4: 0.095, 3: 0.104
        static const int stages = getenv("FTF_K2_STAGES") ? atoi(getenv("FTF_K2_STAGES")) : 2;
        p.stages = stages == 3 ? 3 : 2;
        for (int th = th_max; th >= 1; th /= 2) {
            p.th = th < out_h ? th : out_h;
            p.br = band_input_rows(p.th, in_h, out_h, p.hs);
            p.chunk_stride = (p.cw * p.br + 127) / 128 * 128;
            smem = (size_t)p.chunk_stride * p.n_chunks;
            if (p.br <= 256 && smem * p.stages <= (216 / FTF_K2_CTAS) * 1024) { fits = true; break; }      // stages x CTAs per SM
        }
        p.bands = (out_h + p.th - 1) / p.th;
        p.inv_cw = 1.0f / (float)p.cw;
        if (!fits || n * (long long)p.bands > 0x7fffffffLL || (long long)out_w * 3 > 256ll * kCols) use_tma = false;
    }
    if (use_tma) {
        auto encode = get_tensor_map_encode_tiled();
        if (!encode) { if (err) *err = "cuTensorMapEncodeTiled entry point not found"; return FTF_INTERNAL; }
        cuuint64_t dims[3] = {(cuuint64_t)row_bytes, (cuuint64_t)in_h, (cuuint64_t)n};
        cuuint64_t strides[2] = {(cuuint64_t)row_bytes, (cuuint64_t)row_bytes * (cuuint64_t)in_h};
        cuuint32_t box[3] = {(cuuint32_t)p.cw, (cuuint32_t)p.br, 1};
        cuuint32_t estr[3] = {1, 1, 1};
        CUresult rc = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)in, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) {
            if (err) *err = "cuTensorMapEncodeTiled failed for the image tensor, code " + std::to_string((int)rc);
            return FTF_INTERNAL;
        }
    }
    if (out_dtype == FTF_FLOAT)
        return launch_typed<float>(in, p, (float*)out, recip, use_tma, smem, &tmap, sms, s, err);
    return launch_typed<__nv_bfloat16>(in, p, (__nv_bfloat16*)out, recip, use_tma, smem, &tmap, sms, s, err);
}

}  // namespace ftf
