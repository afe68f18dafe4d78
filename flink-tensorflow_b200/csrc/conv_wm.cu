// K3  convolution + folded-BN bias + ReLU as an implicit GEMM on tcgen05 tensor cores.
//
// Stands where the reference's Session.run executes the Conv2D/BatchNorm/Relu nodes of its frozen
// Inception graph (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67 feeds "input", fetches
// "output"; flink-tensorflow/.../models/ModelFunction.scala:65 runs it).
//
//   D^T[Cout, pixels] = W[Cout, K = KH*KW*Cin] * X[pixels, K]^T            (bf16 in, fp32 accumulate)
//
// Operand roles.  A tcgen05.mma streams the M-side operand one row per cycle past an N x 16 array of MACs, so an
// instruction with M = 128 takes about 128 cycles whatever N is and the tensor pipe runs at N/256 of its rate
// (measured with the role counters below: 126-133 cycles per MMA for N = 32..192).  Inception's output-channel
// counts are small (32..192), its pixel counts large, so the WEIGHTS are the M-side operand (128 output channels per
// instruction) and the PIXELS the N-side operand (256 per instruction): the accumulator is the transposed tile,
// channels on TMEM lanes, pixels on columns.
//
// Pixel operand, two ways (chosen per layer by a cycle/traffic model at plan time):
//  * im2col mode (any stride): per filter tap and channel chunk one TMA *im2col* load brings the tile's np pixels'
//    input vectors -- zero-filled outside the image, wrapping across rows and images -- into a swizzled smem tile.
//  * halo mode (stride 1, KxK): outputs are indexed by the PADDED-linear position q = (img*Hp + hp)*Wp + wp; the real
//    outputs are those with hp < OH and wp < OW, the others are computed and dropped.  In that index space tap
//    (kh, kw) of output q reads padded-input pixel q + kh*Wp + kw -- a pure row shift -- so ONE window of
//    T + (KH-1)*Wp + (KW-1) padded pixels per channel chunk (TMA im2col loads over the padded bounding box) serves
//    every tap: the operand descriptor just starts (kh*Wp + kw) rows further down (the swizzle is address based:
//    tools/probe_umma_shift.cu).  Each input pixel crosses L2->SM about once instead of KH*KW times.
//
// Weights arrive by 2-D tiled TMA.  One thread issues the MMAs into a TMEM accumulator (double buffered when it
// fits); eight epilogue warps read it back (tcgen05.ld: a thread owns one output channel), add the folded-BN bias,
// apply ReLU, round to bf16 and store into the channel slice of the concat tensor the layer feeds ("segments"; a
// warp's 32 channels of one pixel are one 64-byte store).  Persistent, warp-specialised: warp 0 TMA producer,
// warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-11 epilogue.
#include "conv_wm.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>
#include <mutex>

namespace ftf {
namespace wm {

namespace {

constexpr int kWRows = 128;                 // weight rows (output channels) per UMMA: the M side
constexpr int kThreads = 384;              // warps 0-3: producer, MMA issuer, TMEM allocator, spare; warps 4-11: epilogue
constexpr int kEpiWarps = 8;
constexpr int kTmemCols = 512;
constexpr int kMaxStages = 8;
constexpr int kMaxAcc = 2;
constexpr int kStageRow = 40;               // u16 per staged pixel row: 32 channels + 8 of padding (80 bytes: conflict-free 16-byte reads)
constexpr int kMaxTilePixels = 1024;        // S * np upper bound (4 x 256)

typedef CUresult (*encode_im2col_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_im2col_fn get_encode_im2col() {
    static std::once_flag once;
    static encode_im2col_fn fn = nullptr;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (encode_im2col_fn)p;
    });
    return fn;
}

__device__ unsigned long long g_conv_prof[16];

// mbarrier wait that adds the cycles spent waiting to *acc when profiling
__device__ __forceinline__ void timed_wait(uint64_t* bar, uint32_t parity, bool prof, long long* acc) {
    if (!prof) { ptx::mbar_wait(bar, parity); return; }
    const long long t0 = clock64();
    ptx::mbar_wait(bar, parity);
    *acc += clock64() - t0;
}


// The same descriptor in two 32-bit halves: the high word (stride, version, swizzle) never changes inside a kernel, the low
// word is 0x10000 | (address >> 4), so another tap, sub-tile or K step is one 32-bit add (see conv_pm.cu).
__device__ __forceinline__ uint32_t desc_hi(int row_bytes) {
    return (uint32_t)((8 * row_bytes) >> 4) | (1u << 14) | ((uint32_t)(row_bytes == 128 ? 2 : 4) << 29);
}
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFF) >> 4) | (1u << 16); }
__device__ __forceinline__ uint64_t make_desc(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// kNK K steps for each of S sub-tiles (S = 1 outside halo mode), no branches inside.
template <int kNK>
__device__ __forceinline__ void issue_subtiles(int S, uint32_t d, uint32_t d_step, uint32_t w_lo, uint32_t x_lo, uint32_t x_step, uint32_t hi,
                                               uint32_t idesc, uint32_t accumulate_first) {
    for (int s = 0; s < S; ++s, x_lo += x_step, d += d_step) {
#pragma unroll
        for (int k = 0; k < kNK; ++k)
            ptx::umma_bf16_elect(d, make_desc(w_lo + 2u * k, hi), make_desc(x_lo + 2u * k, hi), idesc, k ? 1u : accumulate_first);
    }
}
__device__ __forceinline__ void issue_dispatch(int nk, int S, uint32_t d, uint32_t d_step, uint32_t w_lo, uint32_t x_lo, uint32_t x_step, uint32_t hi,
                                               uint32_t idesc, uint32_t accumulate_first) {
    switch (nk) {
    case 4: issue_subtiles<4>(S, d, d_step, w_lo, x_lo, x_step, hi, idesc, accumulate_first); break;
    case 3: issue_subtiles<3>(S, d, d_step, w_lo, x_lo, x_step, hi, idesc, accumulate_first); break;
    case 2: issue_subtiles<2>(S, d, d_step, w_lo, x_lo, x_step, hi, idesc, accumulate_first); break;
    default: issue_subtiles<1>(S, d, d_step, w_lo, x_lo, x_step, hi, idesc, accumulate_first); break;
    }
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, M x N tile.
__device__ __forceinline__ uint32_t instr_desc_bf16(int m, int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}

struct Bars {
    uint64_t a_full[kMaxStages], a_empty[kMaxStages];      // im2col: (weights|pixels) stages; halo: pixel windows
    uint64_t b_full[kMaxStages], b_empty[kMaxStages];      // halo: weight tiles
    uint64_t tmem_full[kMaxAcc], tmem_empty[kMaxAcc];
    uint32_t tmem_slot;
};

template <bool kHalo>
__global__ void __launch_bounds__(kThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_w,
               const __grid_constant__ ConvParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    ptx::tl_stamp_min(p.tl);
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int rb = 2 * p.kc;                                   // bytes per operand row
    // im2col: [stages x (W | P)]            halo: [2 x window][w_stages x W]
    const int stage_bytes = p.w_stage_bytes + p.p_stage_bytes;
    const int w_tap_bytes = kWRows * rb;                         // one tap's weight tile; a halo-mode W stage holds p.w_group of them
    uint8_t* win_base = smem;
    uint8_t* wt_base = smem + 2 * (size_t)p.p_stage_bytes;
    Bars* bars = (Bars*)(kHalo ? wt_base + (size_t)p.w_stages * p.w_stage_bytes : smem + (size_t)p.stages * stage_bytes);

    __shared__ int pix_tab[2 * kMaxTilePixels];
    __shared__ __align__(16) uint16_t stage_t[kEpiWarps * 32 * kStageRow];   // per-warp 32 pixel x 32 channel transposition tiles                // halo mode: tile column -> output pixel index or -1 (double buffered)
    // The warp index comes out of a shuffle so that the compiler KNOWS it is warp-uniform: the role loops below (TMA
    // producer, MMA issuer) then live entirely on the uniform datapath -- descriptors in uniform registers, stepped by
    // UIADD3, UTCHMMA / UTMALDG issued back to back -- instead of being carried in vector registers and moved over with
    // an ELECT + five R2UR + VOTEU sequence in front of every single tcgen05.mma (245 R2UR in the halo kernel before, 32
    // after), which made the issuing thread, not the tensor pipe, the limit of every tile narrower than 256 columns.
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int taps = p.KH * p.KW;
    const int T = p.S * p.np;                                  // (virtual) pixels per CTA tile
    // 32-bit index arithmetic throughout (the plan refuses layers of 2^31 pixels): a 64-bit division is a subroutine call, and
    // with calls in the kernel the compiler keeps the role loops off the uniform datapath
    const uint32_t span = (uint32_t)(kHalo ? p.Q : p.M);
    const int p_tiles = (int)((span + (uint32_t)T - 1u) / (uint32_t)T);
    const int total_tiles = p_tiles * p.m_tiles;
    const int acc_cols = p.S * p.np;
    const int n_a = kHalo ? 2 : p.stages;
    const bool profiling = p.prof != nullptr;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_p);
        ptx::prefetch_tensormap(&tm_w);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < n_a; ++s) { ptx::mbar_init(&bars->a_full[s], 1); ptx::mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < p.w_stages; ++s) { ptx::mbar_init(&bars->b_full[s], 1); ptx::mbar_init(&bars->b_empty[s], 1); }
        for (int a = 0; a < kMaxAcc; ++a) { ptx::mbar_init(&bars->tmem_full[a], 1); ptx::mbar_init(&bars->tmem_empty[a], kEpiWarps); }
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<kTmemCols>(&bars->tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = bars->tmem_slot;
    ptx::pdl_wait();                   // the previous kernel's outputs (this layer's input) are complete
    ptx::tl_stamp_min(p.tl ? p.tl + 2 : nullptr);

    if (warp == 0) {
        // ===================================== TMA producer =====================================
        long long t_wait = 0;
        const long long t_begin = profiling ? clock64() : 0;
        int as = 0, aph = 0, bs = 0, bph = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int tile_ = p.reverse ? total_tiles - 1 - tile : tile;
            const int p_tile = tile_ / p.m_tiles, m_tile = tile_ - p_tile * p.m_tiles;
            const uint32_t x0 = (uint32_t)p_tile * (uint32_t)T;            // first (virtual) pixel of the tile
            if (!kHalo) {
                const int img = (int)(x0 / (uint32_t)(p.OH * p.OW));
                const int rem = (int)(x0 - (uint32_t)img * (uint32_t)(p.OH * p.OW));
                const int oh = rem / p.OW, ow = rem - oh * p.OW;
                const int base_w = ow * p.stride - p.pad_w, base_h = oh * p.stride - p.pad_h;
                for (int chunk = 0; chunk < p.cin_chunks; ++chunk) {          // K order: chunk outer, tap inner (as in halo mode)
                    for (int tap = 0; tap < taps; ++tap) {
                        const int kh = tap / p.KW, kw = tap - kh * p.KW;
                        timed_wait(&bars->a_empty[as], aph ^ 1, profiling, &t_wait);
                        uint8_t* w_dst = smem + (size_t)as * stage_bytes;
                        ptx::mbar_expect_tx_elect(&bars->a_full[as], (uint32_t)stage_bytes);
                        ptx::tma_load_2d_elect(w_dst, &tm_w, &bars->a_full[as], tap * p.Cin + chunk * p.kc, m_tile * kWRows, false);
                        ptx::tma_load_im2col_elect(w_dst + p.w_stage_bytes, &tm_p, &bars->a_full[as], chunk * p.kc, base_w, base_h, img, kw, kh, false);
                        if (++as == p.stages) { as = 0; aph ^= 1; }
                    }
                }
            } else {
                for (int chunk = 0; chunk < p.cin_chunks; ++chunk) {
                    timed_wait(&bars->a_empty[as], aph ^ 1, profiling, &t_wait);
                    uint8_t* dst = win_base + (size_t)as * p.p_stage_bytes;
                    ptx::mbar_expect_tx_elect(&bars->a_full[as], (uint32_t)p.p_stage_bytes);
                    for (int i = 0; i < p.p_loads; ++i) {                  // the window, kHaloLoadRows padded pixels per load
                        const uint32_t q = x0 + (uint32_t)i * kHaloLoadRows;
                        const int wp = (int)(q % (uint32_t)p.Wp);
                        const uint32_t t = q / (uint32_t)p.Wp;
                        const int hp = (int)(t % (uint32_t)p.Hp);
                        const int img = (int)(t / (uint32_t)p.Hp);
                        ptx::tma_load_im2col_elect(dst + (size_t)i * kHaloLoadRows * rb, &tm_p, &bars->a_full[as], chunk * p.kc, wp - p.pad_w,
                                        hp - p.pad_h, img, 0, 0, false);
                    }
                    for (int tap0 = 0; tap0 < taps; tap0 += p.w_group) {   // w_group taps' weight tiles per stage and barrier round trip
                        const int g_taps = min(p.w_group, taps - tap0);
                        timed_wait(&bars->b_empty[bs], bph ^ 1, profiling, &t_wait);
                        ptx::mbar_expect_tx_elect(&bars->b_full[bs], (uint32_t)(g_taps * w_tap_bytes));
                        for (int t = 0; t < g_taps; ++t)
                            ptx::tma_load_2d_elect(wt_base + (size_t)bs * p.w_stage_bytes + (size_t)t * w_tap_bytes, &tm_w, &bars->b_full[bs],
                                                   (tap0 + t) * p.Cin + chunk * p.kc, m_tile * kWRows, false);
                        if (++bs == p.w_stages) { bs = 0; bph ^= 1; }
                    }
                    if (++as == 2) { as = 0; aph ^= 1; }
                }
            }
        }
        if (profiling && lane == 0) {
            atomicAdd(&p.prof[0], 1ull);
            atomicAdd(&p.prof[1], (unsigned long long)(clock64() - t_begin));
            atomicAdd(&p.prof[2], (unsigned long long)t_wait);
        }
    } else if (warp == 1) {
        // ============== MMA issuer: the whole warp walks the loop, one elected lane issues (see ptx.cuh) ==============
        const uint32_t idesc = instr_desc_bf16(kWRows, p.np);
        const int ksteps = p.kc / 16;
        const uint32_t hi = desc_hi(rb);
        long long w_full = 0, w_acc = 0;
        const long long t_begin = profiling ? clock64() : 0;
        int as = 0, aph = 0, bs = 0, bph = 0, acc = 0, acc_phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            timed_wait(&bars->tmem_empty[acc], acc_phase ^ 1, profiling, &w_acc);     // epilogue has drained this accumulator
            ptx::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * acc_cols);
            if (!kHalo) {
                const int k_blocks = taps * p.cin_chunks;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    timed_wait(&bars->a_full[as], aph, profiling, &w_full);
                    ptx::tc_fence_after();
                    const uint32_t w_addr = ptx::smem_u32(smem + (size_t)as * stage_bytes);
                    const int nk = kb >= (p.cin_chunks - 1) * taps ? p.ksteps_last : ksteps;      // skip the all-zero K steps of the last chunk
                    issue_dispatch(nk, 1, d_tmem, 0u, desc_lo(w_addr), desc_lo(w_addr + (uint32_t)p.w_stage_bytes), 0u, hi, idesc, kb != 0 ? 1u : 0u);
                    ptx::umma_commit_elect(&bars->a_empty[as]);                      // frees the stage when these MMAs retire
                    if (++as == p.stages) { as = 0; aph ^= 1; }
                }
            } else {
                for (int chunk = 0; chunk < p.cin_chunks; ++chunk) {
                    timed_wait(&bars->a_full[as], aph, profiling, &w_full);
                    ptx::tc_fence_after();
                    const uint32_t win_addr = ptx::smem_u32(win_base + (size_t)as * p.p_stage_bytes);
                    const int nk = chunk == p.cin_chunks - 1 ? p.ksteps_last : ksteps;
                    int in_group = 0, kh = 0, kw = 0;
                    for (int tap = 0; tap < taps; ++tap) {
                        if (in_group == 0) {                               // a new group of taps' weights
                            timed_wait(&bars->b_full[bs], bph, profiling, &w_full);
                            ptx::tc_fence_after();
                        }
                        const uint32_t w_lo = desc_lo(ptx::smem_u32(wt_base + (size_t)bs * p.w_stage_bytes + (size_t)in_group * w_tap_bytes));
                        // tap (kh, kw) of sub-tile s: the window entered (s*np + kh*Wp + kw) rows further down
                        const uint32_t x_lo = desc_lo(win_addr) + (uint32_t)(kh * p.Wp + kw) * ((uint32_t)rb >> 4);
                        issue_dispatch(nk, p.S, d_tmem, (uint32_t)p.np, w_lo, x_lo, (uint32_t)(p.np * rb) >> 4, hi, idesc, (chunk | tap) != 0 ? 1u : 0u);
                        if (++kw == p.KW) { kw = 0; ++kh; }
                        if (++in_group == p.w_group || tap + 1 == taps) {  // group consumed: release the stage
                            in_group = 0;
                            ptx::umma_commit_elect(&bars->b_empty[bs]);
                            if (++bs == p.w_stages) { bs = 0; bph ^= 1; }
                        }
                    }
                    ptx::umma_commit_elect(&bars->a_empty[as]);
                    if (++as == 2) { as = 0; aph ^= 1; }
                }
            }
            ptx::umma_commit_elect(&bars->tmem_full[acc]);                           // accumulator complete -> epilogue
            if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1; }
        }
        // all MMAs of this CTA are issued: the next kernel in the stream may now be scheduled onto SMs as they drain.
        // (Triggering at kernel start parks its CTAs -- with their shared memory -- on SMs a sibling branch could use.)
        ptx::pdl_launch_dependents();
        if (profiling && lane == 0) {
            atomicAdd(&p.prof[3], (unsigned long long)(clock64() - t_begin));
            atomicAdd(&p.prof[4], (unsigned long long)w_full);
            atomicAdd(&p.prof[5], (unsigned long long)w_acc);
        }
    } else if (warp >= 4) {
        // ============ epilogue: TMEM -> registers -> bias/ReLU -> bf16/fp32 -> global ============
        // 8 warps: warp w reads TMEM lane quarter w % 4 (a hardware rule) and the column half (w - 4) / 4 of every
        // sub-tile.  A thread owns one output channel; a warp's 32 channels of one pixel form one 64-byte store.
        // Halo mode: the map "column of the tile -> output pixel (or none)" is the same for all 128 channels, so the
        // epilogue threads build it once per tile in shared memory (one padded-linear decode per column) while the
        // MMAs of the tile are still running, and the store loop only looks it up.
        const int e = warp & 3;
        const int half = (warp - 4) >> 2;
        const int row = e * 32 + lane;                             // this thread's output channel within the tile
        const int et = threadIdx.x - 128;                          // 0..255 among the epilogue threads
        const bool prof = profiling && threadIdx.x == 128;
        long long w_epi = 0, n_done = 0;
        const long long t_begin = prof ? clock64() : 0;
        const int hcols = p.np / 2;                                // columns per warp per sub-tile
        int acc = 0, acc_phase = 0, tbuf = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            ++n_done;
            const int tile_ = p.reverse ? total_tiles - 1 - tile : tile;
            const int p_tile = tile_ / p.m_tiles, m_tile = tile_ - p_tile * p.m_tiles;
            const uint32_t tile_x0 = (uint32_t)p_tile * (uint32_t)T;
            if (kHalo) {
                int* tab = pix_tab + tbuf * kMaxTilePixels;
                for (int j = et; j < T; j += kEpiWarps * 32) {
                    const uint32_t q = tile_x0 + (uint32_t)j;
                    const int wp = (int)(q % (uint32_t)p.Wp);
                    const uint32_t t = q / (uint32_t)p.Wp;
                    const int hp = (int)(t % (uint32_t)p.Hp);
                    const bool ok = q < span && wp < p.OW && hp < p.OH;
                    tab[j] = ok ? (int)(((t / (uint32_t)p.Hp) * (uint32_t)p.OH + (uint32_t)hp) * (uint32_t)p.OW + (uint32_t)wp) : -1;
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");   // table complete (and the other buffer free)
            }
            const int ch = m_tile * kWRows + row;
            const bool ch_ok = ch < p.Cout;
            const bool warp_has_work = __any_sync(0xffffffffu, ch_ok);
            // the segment that owns this channel: everything the stores need, in registers
            int si = 0;
#pragma unroll
            for (int q = 1; q < kConvMaxSegments; ++q)
                if (q < p.n_seg && p.seg[q].col_begin <= ch) si = q;
            const int flags = p.seg[si].flags;
            const bool f32 = (flags & kSegF32) != 0, relu = (flags & kSegRelu) != 0;
            const int esz = f32 ? 4 : 2;
            char* const out_base = (char*)p.seg[si].out + ((long long)p.seg[si].coff + (ch - p.seg[si].col_begin)) * esz;
            const long long step = (long long)p.seg[si].ld * esz;
            const float bias = (ch_ok && (flags & kSegBias)) ? __ldg(p.bias + ch) : 0.0f;
            timed_wait(&bars->tmem_full[acc], acc_phase, prof, &w_epi);
            ptx::tc_fence_after();
            // Fast path (bf16 output, all 32 channels of this warp real and in one segment): a 32 pixel x 32 channel block is
            // transposed through shared memory, so that a lane then owns a PIXEL and stores its 32 channels as four 16-byte
            // words -- ~5 instructions per value instead of ~11 for one 2-byte store per value (the weight-major kernel was
            // epilogue-bound on the 35x35 layers: profiles/r01_conv_roles.md).
            // (every shuffle outside the && chain: a short-circuited __shfl_sync would wait for lanes that skipped it)
            const int si_lo = __shfl_sync(0xffffffffu, si, 0), si_hi = __shfl_sync(0xffffffffu, si, 31);
            const bool ch_ok_hi = __shfl_sync(0xffffffffu, (int)ch_ok, 31) != 0;
            const bool fast = si_lo == si_hi && ch_ok_hi && !f32 && p.seg[si].coff % 8 == 0 && p.seg[si].ld % 8 == 0;
            if (fast) {
                const unsigned long long ob = (unsigned long long)out_base;
                char* const warp_base = (char*)(((unsigned long long)__shfl_sync(0xffffffffu, (unsigned)(ob >> 32), 0) << 32) |
                                                (unsigned long long)__shfl_sync(0xffffffffu, (unsigned)ob, 0));
                uint16_t* stg = stage_t + (warp - 4) * 32 * kStageRow;
                for (int s = 0; s < p.S; ++s) {
                    const int j0 = s * p.np + half * hcols;
                    const uint32_t x0 = tile_x0 + (uint32_t)j0;
                    const int* tab = pix_tab + tbuf * kMaxTilePixels + j0;
                    const uint32_t t_row = tmem_base + (uint32_t)(acc * acc_cols + j0) + ((uint32_t)(e * 32) << 16);
                    for (int col = 0; col < hcols; col += 32) {
                        if (x0 + (uint32_t)col >= span) break;                                   // warp-uniform
                        uint32_t v[32];
                        ptx::tmem_ld_32x32(t_row + (uint32_t)col, v);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float f = __uint_as_float(v[j]) + bias;
                            f = relu ? fmaxf(f, 0.0f) : f;
                            stg[j * kStageRow + lane] = __bfloat16_as_ushort(__float2bfloat16_rn(f));
                        }
                        __syncwarp();
                        long long m;                                                   // the output pixel this lane now owns
                        bool ok;
                        if (kHalo) { const int t = tab[col + lane]; ok = t >= 0; m = t; }
                        else { m = (long long)(x0 + (uint32_t)(col + lane)); ok = x0 + (uint32_t)(col + lane) < span; }
                        if (ok) {
                            const uint4* r = reinterpret_cast<const uint4*>(stg + lane * kStageRow);
                            uint4* dst = reinterpret_cast<uint4*>(warp_base + m * step);
                            const uint4 q0 = r[0], q1 = r[1], q2 = r[2], q3 = r[3];
                            dst[0] = q0; dst[1] = q1; dst[2] = q2; dst[3] = q3;
                        }
                        __syncwarp();
                    }
                }
            } else if (warp_has_work) {
                for (int s = 0; s < p.S; ++s) {
                    const int j0 = s * p.np + half * hcols;                            // first column of this warp in the tile
                    const uint32_t x0 = tile_x0 + (uint32_t)j0;
                    char* ptr = out_base + (long long)x0 * step;                                  // im2col mode: pixels are consecutive
                    const int* tab = pix_tab + tbuf * kMaxTilePixels + j0;
                    const uint32_t t_row = tmem_base + (uint32_t)(acc * acc_cols + j0) + ((uint32_t)(e * 32) << 16);
                    for (int col = 0; col < hcols; col += 16) {
                        if (x0 + (uint32_t)col >= span) break;                                   // warp-uniform
                        uint32_t v[16];
                        tmem_ld_x16(t_row + (uint32_t)col, v);
                        ptx::tmem_ld_wait();
                        const long long left = (long long)span - (long long)(x0 + (uint32_t)col);
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            float f = __uint_as_float(v[j]) + bias;
                            f = relu ? fmaxf(f, 0.0f) : f;
                            char* dst;
                            bool ok;
                            if (kHalo) {
                                const int m = tab[col + j];                            // same address for the whole warp: a broadcast
                                ok = ch_ok && m >= 0;
                                dst = out_base + (long long)m * step;
                            } else {
                                ok = ch_ok && j < left;
                                dst = ptr;
                                ptr += step;
                            }
                            if (ok) {
                                if (f32) *reinterpret_cast<float*>(dst) = f;
                                else *reinterpret_cast<__nv_bfloat16*>(dst) = __float2bfloat16_rn(f);
                            }
                        }
                    }
                }
            }
            tbuf ^= 1;
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars->tmem_empty[acc]);
            if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1; }
        }
        if (prof) {
            atomicAdd(&p.prof[6], (unsigned long long)(clock64() - t_begin));
            atomicAdd(&p.prof[7], (unsigned long long)w_epi);
            atomicAdd(&p.prof[8], (unsigned long long)n_done);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kTmemCols>(tmem_base);
    ptx::tl_stamp_max(p.tl ? p.tl + 1 : nullptr);
}

}  // namespace

int plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan, std::string* err,
         double* est_cycles, bool may_use_halo) {
    auto bad = [&](const std::string& m) { if (err) *err = "conv_wm plan: " + m; return FTF_INVALID_ARGUMENT; };
    if (n_seg < 1 || n_seg > kConvMaxSegments) return bad("segment count");
    if (g.Cin % 8 || g.in_ld % 8 || g.Cout % 16) return bad("Cin and the input row stride must be multiples of 8, Cout of 16");
    if (((uintptr_t)g.in | (uintptr_t)g.weights) & 15) return bad("unaligned tensor base");
    const int OH = (g.H + 2 * g.pad_h - g.KH) / g.stride + 1, OW = (g.W + 2 * g.pad_w - g.KW) / g.stride + 1;
    if (OH <= 0 || OW <= 0) return bad("empty output");
    // the kernel indexes (virtual) pixels with 32 bits
    if ((long long)g.n_max * (g.H + 2 * g.pad_h) * (g.W + 2 * g.pad_w) >= (1ll << 31) - 65536 || (long long)g.n_max * OH * OW >= (1ll << 31) - 65536)
        return bad("layer too large (2^31 pixels)");
    for (int i = 0; i < n_seg; ++i)
        if (segs[i].col_begin % 16 || (i && segs[i].col_begin <= segs[i - 1].col_begin))
            return bad("segment alignment / order");
    auto enc_tiled = get_tensor_map_encode_tiled();
    auto enc_im2col = get_encode_im2col();
    if (!enc_tiled || !enc_im2col) { if (err) *err = "cuTensorMapEncode* entry points not found"; return FTF_INTERNAL; }

    ConvParams& p = plan->p;
    p = ConvParams{};
    p.OH = OH; p.OW = OW; p.stride = g.stride; p.pad_h = g.pad_h; p.pad_w = g.pad_w; p.KH = g.KH; p.KW = g.KW;
    p.Cin = g.Cin; p.Cout = g.Cout;
    p.M = (long long)g.n_max * OH * OW;
    p.m_tiles = (g.Cout + kWRows - 1) / kWRows;
    p.bias = g.bias;
    p.n_seg = n_seg;
    for (int i = 0; i < n_seg; ++i) p.seg[i] = segs[i];
    plan->grid = sm_count;
    static const bool want_prof = getenv("FTF_CONV_PROFILE") && getenv("FTF_CONV_PROFILE")[0] == '1';
    if (want_prof) {
        void* sym = nullptr;
        if (cudaGetSymbolAddress(&sym, g_conv_prof) == cudaSuccess) p.prof = (unsigned long long*)sym; else cudaGetLastError();
    }

    // ---- tile shape: (channels per K block, pixels per MMA, im2col or halo window with S sub-tiles) that minimises
    //      waves x (max(MMA cycles, operand bytes / per-SM fill bandwidth) + fixed cost) ----
    const bool env_halo = !(getenv("FTF_NO_HALO") && getenv("FTF_NO_HALO")[0] == '1');
    const bool allow_halo = env_halo && may_use_halo;
    static const bool allow_kc32 = !(getenv("FTF_NO_KC32") && getenv("FTF_NO_KC32")[0] == '1');
    const int taps = g.KH * g.KW;
    const int Hp = g.H + 2 * g.pad_h, Wp = g.W + 2 * g.pad_w;
    const int halo = (g.KH - 1) * Wp + (g.KW - 1);
    const long long Q = (long long)g.n_max * Hp * Wp;
    const double kFill = 36.0, kFixed = 3000.0;                     // bytes/cycle/SM with every SM pulling; cycles per tile
    const double kEpi = 11.0;                                       // epilogue issue slots per pixel (one SMSP serves 32 channels)
    const long long kSmem = 188 * 1024;   // dynamic smem budget (28 KB static table + transposition tiles, barriers and alignment slack on top)
    double best = -1.0;
    for (int kc = 64; kc >= 32; kc -= 32) {
        const int chunks = (g.Cin + kc - 1) / kc, rb = 2 * kc, ksteps = kc / 16;
        const int total_ksteps = (chunks - 1) * ksteps + (g.Cin - kc * (chunks - 1) + 15) / 16;   // all-zero K steps are skipped
        // 32-channel K blocks exactly when they save MMAs (Cin % 64 in 1..32): a fixed rule, so that every candidate
        // kernel of a layer walks K in the same order and produces the same bits
        const bool kc32_saves = allow_kc32 && g.Cin <= 32;       // same rule as the pixel-major kernel (see conv_pm.cu)
        if ((kc == 32) != kc32_saves) continue;
        const double real = (double)g.Cin / (chunks * kc);
        const long long w_bytes = (long long)kWRows * rb;
        for (int np = 256; np >= 128; np -= 128) {
            {   // im2col mode
                const long long p_bytes = (long long)np * rb, stage = w_bytes + p_bytes;
                int stages = (int)(kSmem / stage);
                if (stages > kMaxStages) stages = kMaxStages;
                const long long tiles = (p.M + np - 1) / np * p.m_tiles, waves = (tiles + sm_count - 1) / sm_count;
                const double cyc = (double)taps * total_ksteps * 128.0, bytes = (double)taps * chunks * stage * real;
                double work = cyc > bytes / kFill ? cyc : bytes / kFill;
                if (work < np * kEpi) work = np * kEpi;
                const double t = waves * (work + kFixed);
                if (stages >= 2 && (best < 0 || t < best)) {
                    best = t;
                    p.halo_mode = 0; p.kc = kc; p.cin_chunks = chunks; p.np = np; p.S = 1; p.stages = stages; p.w_stages = 0;
                    p.w_stage_bytes = (int)w_bytes; p.p_stage_bytes = (int)p_bytes; p.acc_stages = 2; p.p_loads = 0;
                }
            }
            if (!(allow_halo && g.stride == 1 && taps > 1)) continue;
            for (int S = 1; S <= 4; S *= 2) {
                if (2 * S * np > kTmemCols) continue;                  // keep two accumulator sets: epilogue overlaps the next tile
                const int T = S * np;
                const int loads = (T + halo + kHaloLoadRows - 1) / kHaloLoadRows;
                const long long p_bytes = (long long)loads * kHaloLoadRows * rb;
                const long long room = kSmem - 2 * p_bytes;
                if (room < 2 * w_bytes) continue;
                const long long tiles = (Q + T - 1) / T * p.m_tiles, waves = (tiles + sm_count - 1) / sm_count;
                const double cyc = (double)total_ksteps * taps * S * 128.0;
                const double bytes = (double)chunks * (p_bytes + (double)taps * w_bytes) * real;
                double work = cyc > bytes / kFill ? cyc : bytes / kFill;
                if (work < T * kEpi) work = T * kEpi;
                const double t = waves * (work + kFixed);
                if (best < 0 || t < best) {
                    best = t;
                    p.halo_mode = 1; p.kc = kc; p.cin_chunks = chunks; p.np = np; p.S = S; p.stages = 0;
                    // weight tiles of several taps share a stage (one barrier round trip per group), as many as leave three stages
                    int grp = (int)(room / (3 * w_bytes));
                    if (grp > taps) grp = taps;
                    if (grp < 1) grp = 1;
                    p.w_group = grp; p.w_stage_bytes = (int)(grp * w_bytes);
                    p.w_stages = (int)(room / p.w_stage_bytes > kMaxStages ? kMaxStages : room / p.w_stage_bytes);
                    if (p.w_stages < 2) { p.w_group = 1; p.w_stage_bytes = (int)w_bytes; p.w_stages = (int)(room / w_bytes > kMaxStages ? kMaxStages : room / w_bytes); }
                    p.p_stage_bytes = (int)p_bytes; p.p_loads = loads;
                    p.acc_stages = 2 * S * np <= kTmemCols ? 2 : 1;
                }
            }
        }
    }
    if (best < 0) return bad("no tile shape fits shared memory");
    if (est_cycles) *est_cycles = best;
    p.ksteps_last = (g.Cin - p.kc * (p.cin_chunks - 1) + 15) / 16;
    p.Hp = Hp; p.Wp = Wp; p.Q = Q;
    plan->smem = (p.halo_mode ? 2 * (size_t)p.p_stage_bytes + (size_t)p.w_stages * p.w_stage_bytes
                              : (size_t)p.stages * (p.w_stage_bytes + p.p_stage_bytes)) + 1024 /*alignment slack*/ + sizeof(Bars);

    const CUtensorMapSwizzle swz = p.kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    // pixels: NHWC input seen through an im2col window. dims fastest-first {C, W, H, N}.
    {
        cuuint64_t dims[4] = {(cuuint64_t)g.Cin, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.n_max};
        cuuint64_t strides[3] = {(cuuint64_t)g.in_ld * 2, (cuuint64_t)g.W * g.in_ld * 2, (cuuint64_t)g.H * g.W * g.in_ld * 2};
        int lower[2] = {-g.pad_w, -g.pad_h};
        int upper[2] = {g.pad_w - (g.KW - 1), g.pad_h - (g.KH - 1)};
        if (p.halo_mode) { upper[0] = g.pad_w; upper[1] = g.pad_h; }      // the whole padded image, one pixel per "filter"
        cuuint32_t estr[4] = {1, (cuuint32_t)g.stride, (cuuint32_t)g.stride, 1};
        CUresult rc = enc_im2col(&plan->tm_p, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)g.in, dims, strides, lower, upper,
                                 (cuuint32_t)p.kc, (cuuint32_t)(p.halo_mode ? kHaloLoadRows : p.np), estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { if (err) *err = "cuTensorMapEncodeIm2col failed, code " + std::to_string((int)rc); return FTF_INTERNAL; }
    }
    // weights [Cout][K] with K = KH*KW*Cin contiguous.
    {
        const cuuint64_t K = (cuuint64_t)g.KH * g.KW * g.Cin;
        cuuint64_t dims[2] = {K, (cuuint64_t)g.Cout};
        cuuint64_t strides[1] = {K * 2};
        cuuint32_t box[2] = {(cuuint32_t)p.kc, (cuuint32_t)kWRows};
        cuuint32_t estr[2] = {1, 1};
        CUresult rc = enc_tiled(&plan->tm_w, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)g.weights, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { if (err) *err = "cuTensorMapEncodeTiled (weights) failed, code " + std::to_string((int)rc); return FTF_INTERNAL; }
    }
    // the attribute is per device: set it whenever a plan is made
    cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 198 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 198 * 1024);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc_kernel)", err);
    return FTF_OK;
}

int profile_read(unsigned long long* out16, int reset) {
    if (cudaMemcpyFromSymbol(out16, g_conv_prof, sizeof(unsigned long long) * 16) != cudaSuccess) { cudaGetLastError(); return FTF_INTERNAL; }
    if (reset) {
        unsigned long long zero[16] = {0};
        if (cudaMemcpyToSymbol(g_conv_prof, zero, sizeof(zero)) != cudaSuccess) { cudaGetLastError(); return FTF_INTERNAL; }
    }
    return FTF_OK;
}

int launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err, unsigned long long* tl, int reverse) {
    ConvParams p = plan.p;
    p.tl = tl;
    p.reverse = reverse;
    p.M = (long long)n * p.OH * p.OW;
    p.Q = (long long)n * p.Hp * p.Wp;
    if (p.M <= 0) return FTF_OK;
    const int T = p.S * p.np;
    const long long span = p.halo_mode ? p.Q : p.M;
    const long long tiles = (span + T - 1) / T * p.m_tiles;
    const int grid = (int)(tiles < plan.grid ? tiles : plan.grid);
    if (p.halo_mode) FTF_CUDA(launch_pdl(conv_tc_kernel<true>, dim3(grid), dim3(kThreads), plan.smem, s, plan.tm_p, plan.tm_w, p), err);
    else FTF_CUDA(launch_pdl(conv_tc_kernel<false>, dim3(grid), dim3(kThreads), plan.smem, s, plan.tm_p, plan.tm_w, p), err);
    count_launch();
    return FTF_OK;
}

}  // namespace wm
}  // namespace ftf
