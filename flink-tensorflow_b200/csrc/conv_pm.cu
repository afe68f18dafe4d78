// K3  convolution + folded-BN bias + ReLU as an implicit GEMM on tcgen05 tensor cores.
//
// Stands where the reference's Session.run executes the Conv2D/BatchNorm/Relu nodes of its frozen
// Inception graph (flink-tensorflow-examples/.../inception/InceptionModel.scala:58-67 feeds "input", fetches
// "output"; flink-tensorflow/.../models/ModelFunction.scala:65 runs it).
//
//   D[M = n*OH*OW pixels, N = Cout] = A[M, K = KH*KW*Cin] * B[N, K]^T      (bf16 in, fp32 accumulate)
//
// * A is never materialised: for every filter tap (kh, kw) and 64-channel chunk, one TMA *im2col* load
//   (cp.async.bulk.tensor.4d...im2col) brings the 128 output pixels' input vectors -- zero-filled outside the
//   image, wrapping across rows and images -- into a 128x64 bf16, 128B-swizzled shared-memory tile.
// * B (weights, [Cout][KH*KW*Cin], K-major) arrives by a 2-D tiled TMA load into the same swizzle.
// * One elected thread issues tcgen05.mma (M=128, N=block_n, K=16) with the fp32 accumulator in TMEM; two
//   accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
// * Epilogue warps read TMEM (tcgen05.ld), add the folded BN bias, apply ReLU, round to bf16 and write
//   straight into the channel slice of the concat buffer the layer feeds ("segments"); sibling 1x1 branches
//   that read the same input are one GEMM with several segments.
// Persistent kernel: grid = min(tiles, SMs); warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warps 4-7 = epilogue.
#include "conv_pm.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>
#include <mutex>

namespace ftf {
namespace pm {

namespace {

constexpr int kBlockM = 128;
constexpr int kThreads = 384;              // warps 0-3: TMA, MMA, TMEM allocator, idle; warps 4-11: epilogue
constexpr int kEpiWarps = 8;                // two per TMEM lane quarter: each takes half of a tile's columns
constexpr int kAccStages = 2;
constexpr int kTmemCols = 512;              // 2 x 256 fp32 columns
constexpr int kMaxStages = 8;

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

// K-major swizzled operand descriptor in two 32-bit halves: rows of row_bytes (128 -> SWIZZLE_128B, 64 -> SWIZZLE_64B), 8-row groups
// contiguous.  hi = stride byte offset (one 8-row group) | descriptor version (bit 46) | layout type (bits 61-63);
// lo = start address in 16-byte units | leading byte offset field = 1.  Only lo changes between MMAs.
__device__ __forceinline__ uint32_t desc_hi(int row_bytes) {
    return (uint32_t)((8 * row_bytes) >> 4) | (1u << 14) | ((uint32_t)(row_bytes == 128 ? 2 : 4) << 29);
}
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFF) >> 4) | (1u << 16); }
__device__ __forceinline__ uint64_t make_desc(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// kNK K steps of one K block (im2col kernel), no branches inside.
template <int kNK, bool kPair>
__device__ __forceinline__ void issue_ksteps(uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t accumulate_first) {
#pragma unroll
    for (int k = 0; k < kNK; ++k) {
        if (kPair) ptx::umma_bf16_pair_elect(d, make_desc(a_lo + 2u * k, hi), make_desc(b_lo + 2u * k, hi), idesc, k ? 1u : accumulate_first);
        else ptx::umma_bf16_elect(d, make_desc(a_lo + 2u * k, hi), make_desc(b_lo + 2u * k, hi), idesc, k ? 1u : accumulate_first);
    }
}

// kNK K steps for each of S sub-tiles of one filter tap (halo kernel), no branches inside.
template <int kNK>
__device__ __forceinline__ void issue_subtiles(int S, uint32_t d, uint32_t d_step, uint32_t a_lo, uint32_t a_step, uint32_t b_lo, uint32_t hi,
                                               uint32_t idesc, uint32_t accumulate_first) {
    for (int s = 0; s < S; ++s, a_lo += a_step, d += d_step) {
#pragma unroll
        for (int k = 0; k < kNK; ++k)
            ptx::umma_bf16_elect(d, make_desc(a_lo + 2u * k, hi), make_desc(b_lo + 2u * k, hi), idesc, k ? 1u : accumulate_first);
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

__device__ __forceinline__ void st_global_v8(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f, uint32_t g, uint32_t h) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d), "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// One 16-column chunk of one accumulator row: bias, ReLU, rounding and the store into the segment that owns
// GEMM column gc (m = output pixel index, row_ok = the row is a real output).
__device__ __forceinline__ void epilogue_chunk_seg(const ConvParams& p, const ConvSegment& sg, const float* bias, int gc, long long m, bool row_ok,
                                                   const uint32_t (&v)[16]);
__device__ __forceinline__ void epilogue_chunk(const ConvParams& p, const float* bias, int gc, long long m, bool row_ok, const uint32_t (&v)[16]) {
    // most layers have one segment: its fields then sit at fixed offsets of the parameter block instead of behind a
    // register-indexed constant load per 16-column chunk
    if (p.n_seg == 1) { epilogue_chunk_seg(p, p.seg[0], bias, gc, m, row_ok, v); return; }
    int s = 0;
#pragma unroll
    for (int q = 1; q < kConvMaxSegments; ++q)
        if (q < p.n_seg && p.seg[q].col_begin <= gc) s = q;
    epilogue_chunk_seg(p, p.seg[s], bias, gc, m, row_ok, v);
}
__device__ __forceinline__ void epilogue_chunk_seg(const ConvParams& p, const ConvSegment& sg, const float* bias, int gc, long long m, bool row_ok,
                                                   const uint32_t (&v)[16]) {
    float f[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]);
    if (sg.flags & kSegBias) {
        const float4* bp = reinterpret_cast<const float4*>(bias + gc);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 b = bp[i];
            f[4 * i] += b.x; f[4 * i + 1] += b.y; f[4 * i + 2] += b.z; f[4 * i + 3] += b.w;
        }
    }
    if (sg.flags & kSegRelu) {
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.0f);
    }
    if (row_ok) {
        // A thread owns a row, so the 32 lanes of a store instruction hit 32 different cache lines and the LSU spends a tag
        // lookup per lane whatever the width: 256-bit stores (one full 32-byte sector per lane) halve the instruction count
        // of the epilogue.  With them the stores cost what their bytes cost (FTF_EPI_NO_STORE: the 1x1 layers of the 35x35 blocks
        // write 38-47 MB and run 5-7 us faster without their stores).  p.wide_stores: every segment keeps 32-byte alignment.
        const size_t off = (size_t)m * sg.ld + sg.coff + (gc - sg.col_begin);
        if (sg.flags & kSegF32) {
            float* o = (float*)sg.out + off;
            if (p.wide_stores) {
                st_global_v8(o, __float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]), __float_as_uint(f[4]),
                             __float_as_uint(f[5]), __float_as_uint(f[6]), __float_as_uint(f[7]));
                st_global_v8(o + 8, __float_as_uint(f[8]), __float_as_uint(f[9]), __float_as_uint(f[10]), __float_as_uint(f[11]), __float_as_uint(f[12]),
                             __float_as_uint(f[13]), __float_as_uint(f[14]), __float_as_uint(f[15]));
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(o)[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
            }
        } else {
            __nv_bfloat16* o = (__nv_bfloat16*)sg.out + off;
            const uint32_t w0 = pack_bf16(f[0], f[1]), w1 = pack_bf16(f[2], f[3]), w2 = pack_bf16(f[4], f[5]), w3 = pack_bf16(f[6], f[7]);
            const uint32_t w4 = pack_bf16(f[8], f[9]), w5 = pack_bf16(f[10], f[11]), w6 = pack_bf16(f[12], f[13]), w7 = pack_bf16(f[14], f[15]);
            if (p.wide_stores) st_global_v8(o, w0, w1, w2, w3, w4, w5, w6, w7);
            else {
                reinterpret_cast<uint4*>(o)[0] = make_uint4(w0, w1, w2, w3);
                reinterpret_cast<uint4*>(o)[1] = make_uint4(w4, w5, w6, w7);
            }
        }
    }
}

// The epilogue warps (threads 128..255) copy the layer's bias vector into shared memory while the first accumulator is
// still being computed; every later read is a broadcast from smem instead of an L2 round trip per 16-column chunk.
constexpr int kBiasCap = 2048;
constexpr int kMaxDynSmem = 227 * 1024 - kBiasCap * 4;       // the bias copy is static shared memory
__device__ __forceinline__ const float* stage_bias(const ConvParams& p, float* bias_s) {
    if (p.N_total > kBiasCap || p.bias == nullptr) return p.bias;
    for (int i = (int)threadIdx.x - 128; i < p.N_total; i += 32 * kEpiWarps) bias_s[i] = __ldg(p.bias + i);
    asm volatile("bar.sync 1, 256;" ::: "memory");
    return bias_s;
}

// All column chunks of one accumulator row block: TMEM loads of 32 columns, double-buffered so that the load of group
// g+1 is in flight while group g is biased, rounded and stored.
__device__ __forceinline__ void tmem_ld_group(uint32_t taddr, int cols, uint32_t (&v)[32]) {
    if (cols > 16) ptx::tmem_ld_32x32(taddr, v);
    else tmem_ld_x16(taddr, reinterpret_cast<uint32_t(&)[16]>(v));
}
__device__ __forceinline__ void epilogue_group(const ConvParams& p, const float* bias, int gc, int cols, long long m, bool row_ok, const uint32_t (&v)[32]) {
    epilogue_chunk(p, bias, gc, m, row_ok, reinterpret_cast<const uint32_t(&)[16]>(v));
    if (cols > 16) epilogue_chunk(p, bias, gc + 16, m, row_ok, *reinterpret_cast<const uint32_t(*)[16]>(&v[16]));
}
__device__ __forceinline__ void epilogue_row(const ConvParams& p, const float* bias, uint32_t t_row, int n_tile, long long m, bool row_ok, int half) {
    const int gc0 = n_tile * p.block_n;
    const int n_all = min(p.block_n, p.N_total - gc0);            // multiple of 16, warp-uniform
    // the two warps of a lane quarter split the columns at a multiple of 32
    const int split = ((n_all / 2 + 31) / 32) * 32;
    const int c_begin = half ? min(split, n_all) : 0, n_cols = (half ? n_all : min(split, n_all)) - c_begin;
    if (n_cols <= 0) return;
    t_row += (uint32_t)c_begin;
    const int gcb = gc0 + c_begin;
    uint32_t va[32], vb[32];
    tmem_ld_group(t_row, n_cols, va);
    ptx::tmem_ld_wait();
    for (int col = 0; col < n_cols; col += 64) {
        const int c1 = col + 32, c2 = col + 64;
        if (c1 < n_cols) tmem_ld_group(t_row + (uint32_t)c1, n_cols - c1, vb);
        epilogue_group(p, bias, gcb + col, n_cols - col, m, row_ok, va);
        ptx::tmem_ld_wait();
        if (c1 < n_cols) {
            if (c2 < n_cols) tmem_ld_group(t_row + (uint32_t)c2, n_cols - c2, va);
            epilogue_group(p, bias, gcb + c1, n_cols - c1, m, row_ok, vb);
            ptx::tmem_ld_wait();
        }
    }
}


// kPair = false: one CTA per 128-pixel tile.  kPair = true: a cluster of two CTAs on one TPC works on 256 pixels with
// cta_group::2 MMAs -- CTA r loads the im2col rows of ITS 128 pixels and rows [r*block_n/2, (r+1)*block_n/2) of the weight
// tile, CTA 0 issues the M = 256 MMA, each CTA's TMEM holds its own 128 accumulator rows and its epilogue warps store
// them.  Per CTA and MMA that is half the weight bytes written to and read from shared memory, which is what bounds the
// one-CTA kernel (every operand byte is written once by TMA and read once by the tensor core: ~64 + N/2 cycles per MMA
// instead of N/2; with the pair ~64 + N/4).
template <bool kPair>
__global__ void __launch_bounds__(kThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
               const __grid_constant__ ConvParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(16) float bias_s[kBiasCap];
    ptx::tl_stamp_min(p.tl);
    // carve: [stages x (A | B)] [barriers] [tmem slot]; identical offsets in both CTAs of a pair
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const int rb = 2 * p.kc;                                   // bytes per operand row
    const int stage_bytes = p.a_tile_bytes + p.b_stage_bytes;
    uint64_t* full_bar = (uint64_t*)(smem + (size_t)p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + kMaxStages;
    uint64_t* tmem_full = empty_bar + kMaxStages;
    uint64_t* tmem_empty = tmem_full + kAccStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_empty + kAccStages);

    // The warp index comes out of a shuffle so that the compiler KNOWS it is warp-uniform: the role loops below (TMA
    // producer, MMA issuer) then live entirely on the uniform datapath -- descriptors in uniform registers, stepped by
    // UIADD3, UTCHMMA / UTMALDG issued back to back -- instead of being carried in vector registers and moved over with
    // an ELECT + five R2UR + VOTEU sequence in front of every single tcgen05.mma (245 R2UR in the halo kernel before, 32
    // after), which made the issuing thread, not the tensor pipe, the limit of every tile narrower than 256 columns.
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int k_blocks = p.KH * p.KW * p.cin_chunks;
    const int rank = kPair ? (int)ptx::cluster_ctarank() : 0;
    // work items: (tile of 128 pixels, or pair of tiles) x n_tile, walked by CTA (or cluster) index
    const int worker = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, n_workers = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int total_items = (kPair ? (p.m_tiles + 1) / 2 : p.m_tiles) * p.n_tiles;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_a);
        ptx::prefetch_tensormap(&tm_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.stages; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < kAccStages; ++a) { ptx::mbar_init(&tmem_full[a], 1); ptx::mbar_init(&tmem_empty[a], kPair ? 2 * kEpiWarps : kEpiWarps); }
        ptx::fence_mbar_init();
    }
    if (warp == 2) { if (kPair) ptx::tmem_alloc_pair<kTmemCols>(tmem_slot); else ptx::tmem_alloc<kTmemCols>(tmem_slot); }
    ptx::tc_fence_before();
    if (kPair) ptx::cluster_sync(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();                   // the previous kernel's outputs (this layer's input) are complete
    ptx::tl_stamp_min(p.tl ? p.tl + 2 : nullptr);

    if (warp == 0) {
        // ===== TMA producer (one warp per CTA; whole-warp issue, one elected lane per instruction) =====
        int stage = 0, phase = 0;
        const bool prof = p.prof != nullptr;
        long long t_wait = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int item = worker; item < total_items; item += n_workers) {
            const int item_ = p.reverse ? total_items - 1 - item : item;
            const int m_grp = item_ / p.n_tiles, n_tile = item_ - m_grp * p.n_tiles;
            // an odd tile count leaves the last pair's second CTA without pixels: it reloads the last tile, stores nothing
            const int m_tile = kPair ? min(2 * m_grp + rank, p.m_tiles - 1) : m_grp;
            const int m0 = m_tile * kBlockM;
            const int img = m0 / (p.OH * p.OW);
            const int rem = m0 - img * (p.OH * p.OW);
            const int oh = rem / p.OW, ow = rem - oh * p.OW;
            const int base_w = ow * p.stride - p.pad_w, base_h = oh * p.stride - p.pad_h;
            const int b_row = n_tile * p.block_n + (kPair ? rank * (p.block_n / 2) : 0);
            for (int kb = 0; kb < k_blocks; ++kb) {
                const int taps_ = p.KH * p.KW;                     // K order: channel chunk outer, filter tap inner (as in halo mode)
                const int chunk = kb / taps_, tap = kb - chunk * taps_;
                const int kh = tap / p.KW, kw = tap - kh * p.KW;
                timed_wait(&empty_bar[stage], phase ^ 1, prof, &t_wait);
                uint8_t* a_dst = smem + (size_t)stage * stage_bytes;
                uint8_t* b_dst = a_dst + p.a_tile_bytes;
                // pair: both CTAs' bytes complete on CTA 0's barrier, which alone carries the expectation
                if (!kPair || rank == 0) ptx::mbar_expect_tx_elect(&full_bar[stage], (kPair ? 2u : 1u) * (uint32_t)stage_bytes);
                ptx::tma_load_im2col_elect(a_dst, &tm_a, &full_bar[stage], chunk * p.kc, base_w, base_h, img, kw, kh, kPair);
                ptx::tma_load_2d_elect(b_dst, &tm_b, &full_bar[stage], tap * p.Cin + chunk * p.kc, b_row, kPair);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
        if (prof && lane == 0) { atomicAdd(&p.prof[0], 1ull); atomicAdd(&p.prof[1], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[2], (unsigned long long)t_wait); }
    } else if (warp == 1 && rank == 0) {
        // ===== MMA issuer: the whole warp walks the loop, one elected lane issues (CTA 0 of a pair issues for both) =====
        const uint32_t idesc = instr_desc_bf16(kPair ? 2 * kBlockM : kBlockM, p.block_n);
        const uint32_t hi = desc_hi(rb);
        const int last_chunk_kb = (p.cin_chunks - 1) * p.KH * p.KW;            // K blocks run chunk-major
        int stage = 0, phase = 0, acc = 0, acc_phase = 0;
        const bool prof = p.prof != nullptr;
        long long w_full = 0, w_acc = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int item = worker; item < total_items; item += n_workers) {
            timed_wait(&tmem_empty[acc], acc_phase ^ 1, prof, &w_acc);       // the epilogue warps (of both CTAs) have drained this accumulator
            ptx::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
            for (int kb = 0; kb < k_blocks; ++kb) {
                timed_wait(&full_bar[stage], phase, prof, &w_full);           // TMA bytes have landed (in both CTAs)
                ptx::tc_fence_after();
                const uint32_t a_addr = ptx::smem_u32(smem + (size_t)stage * stage_bytes);
                const uint32_t a_lo = desc_lo(a_addr), b_lo = desc_lo(a_addr + (uint32_t)p.a_tile_bytes);
                // K=16 steps of this block (32 bytes, two 16-byte units each); the last channel chunk stops where the
                // real channels end -- the rest of its rows is TMA zero fill and would add exact zeros.  Straight-line issue.
                const int nk = kb >= last_chunk_kb ? p.ksteps_last : p.kc / 16;
                const uint32_t first = kb != 0 ? 1u : 0u;
                switch (nk) {
                case 4: issue_ksteps<4, kPair>(d_tmem, a_lo, b_lo, hi, idesc, first); break;
                case 3: issue_ksteps<3, kPair>(d_tmem, a_lo, b_lo, hi, idesc, first); break;
                case 2: issue_ksteps<2, kPair>(d_tmem, a_lo, b_lo, hi, idesc, first); break;
                default: issue_ksteps<1, kPair>(d_tmem, a_lo, b_lo, hi, idesc, first); break;
                }
                // frees the smem stage (in both CTAs) when the MMAs retire
                if (kPair) ptx::umma_commit_pair_elect(&empty_bar[stage]); else ptx::umma_commit_elect(&empty_bar[stage]);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
            // accumulator complete -> epilogue warps (of both CTAs)
            if (kPair) ptx::umma_commit_pair_elect(&tmem_full[acc]); else ptx::umma_commit_elect(&tmem_full[acc]);
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        // all MMAs of this CTA are issued: the next kernel in the stream may now be scheduled onto SMs as they drain.
        // (Triggering at kernel start parks its CTAs -- with their shared memory -- on SMs a sibling branch could use.)
        ptx::pdl_launch_dependents();
        if (prof && lane == 0) { atomicAdd(&p.prof[3], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[4], (unsigned long long)w_full); atomicAdd(&p.prof[5], (unsigned long long)w_acc); }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> bias/ReLU -> global =====
        const float* bias = stage_bias(p, bias_s);
        const int e = (warp - 4) & 3, half = (warp - 4) >> 2;      // TMEM lane quarter (a hardware rule: warp % 4) and column half of this warp
        const int row = e * 32 + lane;
        int acc = 0, acc_phase = 0;
        const bool prof = p.prof != nullptr && threadIdx.x == 128;
        long long w_epi = 0, n_tiles_done = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int item = worker; item < total_items; item += n_workers) {
            ++n_tiles_done;
            const int item_ = p.reverse ? total_items - 1 - item : item;
            const int m_grp = item_ / p.n_tiles, n_tile = item_ - m_grp * p.n_tiles;
            const int m_tile = kPair ? 2 * m_grp + rank : m_grp;
            const long long m = (long long)m_tile * kBlockM + row;
            const bool row_ok = m < p.M && !p.dbg_no_store;            // also false for the surplus CTA of an odd last pair
            timed_wait(&tmem_full[acc], acc_phase, prof, &w_epi);
            ptx::tc_fence_after();
            const uint32_t t_row = tmem_base + (uint32_t)(acc * 256) + ((uint32_t)(e * 32) << 16);
            epilogue_row(p, bias, t_row, n_tile, m, row_ok, half);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (kPair) ptx::mbar_arrive_leader(&tmem_empty[acc]); else ptx::mbar_arrive(&tmem_empty[acc]); }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (prof) { atomicAdd(&p.prof[6], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[7], (unsigned long long)w_epi); atomicAdd(&p.prof[8], (unsigned long long)n_tiles_done); }
    }
    ptx::tc_fence_before();
    if (kPair) ptx::cluster_sync(); else __syncthreads();
    if (warp == 2) { if (kPair) ptx::tmem_dealloc_pair<kTmemCols>(tmem_base); else ptx::tmem_dealloc<kTmemCols>(tmem_base); }
    ptx::tl_stamp_max(p.tl ? p.tl + 1 : nullptr);
}

// ------------------------------------------------------------------------------------------------
// Halo mode: stride-1 KxK convolutions.
//
// Index outputs by the PADDED-linear position q = (img*Hp + hp)*Wp + wp (Hp = H + 2*pad_h, Wp = W + 2*pad_w); the
// real outputs are those with hp < OH and wp < OW, the rest (KW-1 columns per row, KH-1 rows per image) are computed
// and dropped.  In that index space filter tap (kh, kw) of output q reads padded-input pixel q + kh*Wp + kw: a pure
// row shift.  So a CTA tile of T = 128*S consecutive q needs ONE window of T + (KH-1)*Wp + (KW-1) padded-input pixels
// per 64-channel chunk -- fetched once by TMA im2col loads over the padded bounding box (zeros in the padding,
// wrapping rows and images) -- and every tap's A operand is the same 128B-swizzled buffer entered at a different
// row: the UMMA descriptor's start address moves by (kh*Wp + kw)*128 bytes (the swizzle is address based, see
// tools/probe_umma_shift.cu).  L2->SM operand traffic drops from KH*KW reads of every input pixel to about one,
// and the weight tiles are shared by S sub-tiles.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                 const __grid_constant__ ConvParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(16) float bias_s[kBiasCap];
    ptx::tl_stamp_min(p.tl);
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    constexpr int kAStages = 2;
    uint8_t* a_base = smem;
    const int rb = 2 * p.kc;
    uint8_t* b_base = smem + (size_t)kAStages * p.a_stage_bytes;
    uint64_t* a_full = (uint64_t*)(b_base + (size_t)p.b_stages * p.b_stage_bytes);
    uint64_t* a_empty = a_full + kAStages;
    uint64_t* b_full = a_empty + kAStages;
    uint64_t* b_empty = b_full + kMaxStages;
    uint64_t* tmem_full = b_empty + kMaxStages;
    uint64_t* tmem_empty = tmem_full + kAccStages;
    uint32_t* tmem_slot = (uint32_t*)(tmem_empty + kAccStages);

    // The warp index comes out of a shuffle so that the compiler KNOWS it is warp-uniform: the role loops below (TMA
    // producer, MMA issuer) then live entirely on the uniform datapath -- descriptors in uniform registers, stepped by
    // UIADD3, UTCHMMA / UTMALDG issued back to back -- instead of being carried in vector registers and moved over with
    // an ELECT + five R2UR + VOTEU sequence in front of every single tcgen05.mma (245 R2UR in the halo kernel before, 32
    // after), which made the issuing thread, not the tensor pipe, the limit of every tile narrower than 256 columns.
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int taps = p.KH * p.KW;
    const int T = p.S * kBlockM;
    const int q_tiles = (int)((p.Q + T - 1) / T);
    const int total_tiles = q_tiles * p.n_tiles;
    const int acc_cols = p.S * p.block_n;
    const int b_tap_bytes = p.block_n * rb;                      // one tap's weight tile; a B stage holds p.b_group of them

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_a);
        ptx::prefetch_tensormap(&tm_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kAStages; ++s) { ptx::mbar_init(&a_full[s], 1); ptx::mbar_init(&a_empty[s], 1); }
        for (int s = 0; s < p.b_stages; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
        for (int a = 0; a < kAccStages; ++a) { ptx::mbar_init(&tmem_full[a], 1); ptx::mbar_init(&tmem_empty[a], kEpiWarps); }
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<kTmemCols>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();                   // the previous kernel's outputs (this layer's input) are complete
    ptx::tl_stamp_min(p.tl ? p.tl + 2 : nullptr);

    if (warp == 0) {
        // ===== TMA producer =====
        int as = 0, aph = 0, bs = 0, bph = 0;
        const bool prof = p.prof != nullptr;
        long long t_wait = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int tile_ = p.reverse ? total_tiles - 1 - tile : tile;
            const int q_tile = tile_ / p.n_tiles, n_tile = tile_ - q_tile * p.n_tiles;
            // first padded pixel of the window, decoded once per tile (32-bit: the plan keeps Q below 2^31); the loads then
            // step through the padded image incrementally instead of dividing per load
            const uint32_t q0 = (uint32_t)q_tile * (uint32_t)T;
            const int wp0 = (int)(q0 % (uint32_t)p.Wp);
            const uint32_t t0 = q0 / (uint32_t)p.Wp;
            const int hp0 = (int)(t0 % (uint32_t)p.Hp), img0 = (int)(t0 / (uint32_t)p.Hp);
            for (int chunk = 0; chunk < p.cin_chunks; ++chunk) {
                timed_wait(&a_empty[as], aph ^ 1, prof, &t_wait);
                uint8_t* a_dst = a_base + (size_t)as * p.a_stage_bytes;
                ptx::mbar_expect_tx_elect(&a_full[as], (uint32_t)p.a_stage_bytes);
                int wp = wp0, hp = hp0, img = img0;
                for (int i = 0; i < p.a_loads; ++i) {           // the window, kHaloLoadRows padded pixels per load
                    ptx::tma_load_im2col_elect(a_dst + (size_t)i * kHaloLoadRows * rb, &tm_a, &a_full[as], chunk * p.kc, wp - p.pad_w,
                                    hp - p.pad_h, img, 0, 0, false);
                    wp += kHaloLoadRows;
                    while (wp >= p.Wp) { wp -= p.Wp; if (++hp == p.Hp) { hp = 0; ++img; } }
                }
                for (int tap0 = 0; tap0 < taps; tap0 += p.b_group) {       // b_group taps' weight tiles per stage and barrier round trip
                    const int g_taps = min(p.b_group, taps - tap0);
                    timed_wait(&b_empty[bs], bph ^ 1, prof, &t_wait);
                    ptx::mbar_expect_tx_elect(&b_full[bs], (uint32_t)(g_taps * b_tap_bytes));
                    for (int t = 0; t < g_taps; ++t)
                        ptx::tma_load_2d_elect(b_base + (size_t)bs * p.b_stage_bytes + (size_t)t * b_tap_bytes, &tm_b, &b_full[bs],
                                               (tap0 + t) * p.Cin + chunk * p.kc, n_tile * p.block_n, false);
                    if (++bs == p.b_stages) { bs = 0; bph ^= 1; }
                }
                if (++as == kAStages) { as = 0; aph ^= 1; }
            }
        }
        if (prof && lane == 0) { atomicAdd(&p.prof[0], 1ull); atomicAdd(&p.prof[1], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[2], (unsigned long long)t_wait); }
    } else if (warp == 1) {
        // ===== MMA issuer: whole warp, one elected lane issues =====
        const uint32_t idesc = instr_desc_bf16(kBlockM, p.block_n);
        const uint32_t hi = desc_hi(rb), rows16 = (uint32_t)rb >> 4, sub16 = (uint32_t)(kBlockM * rb) >> 4;   // descriptor address units (16 B)
        int as = 0, aph = 0, bs = 0, bph = 0, acc = 0, acc_phase = 0;
        const bool prof = p.prof != nullptr;
        long long w_full = 0, w_acc = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            timed_wait(&tmem_empty[acc], acc_phase ^ 1, prof, &w_acc);
            ptx::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * acc_cols);
            for (int chunk = 0; chunk < p.cin_chunks; ++chunk) {
                timed_wait(&a_full[as], aph, prof, &w_full);
                ptx::tc_fence_after();
                // Descriptors as two 32-bit words: the high word (stride, version, swizzle) never changes, the low word is
                // 0x10000 | (address >> 4), so stepping to another tap, sub-tile or K step is one 32-bit add.  (The loop used
                // to rebuild 64-bit descriptors per MMA: ~50 dependent instructions per MMA made this warp the bottleneck of
                // the 32-channel stem layers -- profiles/r01_conv_roles.md.)
                const uint32_t a_lo0 = desc_lo(ptx::smem_u32(a_base + (size_t)as * p.a_stage_bytes));
                const int nk = chunk == p.cin_chunks - 1 ? p.ksteps_last : p.kc / 16;     // skip the all-zero K steps
                int tap = 0, in_group = 0, kh = 0, kw = 0;
                for (; tap < taps; ++tap) {
                    if (in_group == 0) {                                   // a new group of taps' weights
                        timed_wait(&b_full[bs], bph, prof, &w_full);
                        ptx::tc_fence_after();
                    }
                    const uint32_t b_lo = desc_lo(ptx::smem_u32(b_base + (size_t)bs * p.b_stage_bytes + (size_t)in_group * b_tap_bytes));
                    const uint32_t first = (chunk | tap) != 0 ? 1u : 0u;
                    // tap (kh, kw) of sub-tile s: the window entered (s*128 + kh*Wp + kw) rows further down
                    const uint32_t a_lo = a_lo0 + (uint32_t)(kh * p.Wp + kw) * rows16;
                    // straight-line issue: the number of K steps is resolved once per tap, not per MMA (a conditional
                    // around an elected instruction costs a convergence check each time)
                    switch (nk) {
                    case 4: issue_subtiles<4>(p.S, d_tmem, (uint32_t)p.block_n, a_lo, sub16, b_lo, hi, idesc, first); break;
                    case 3: issue_subtiles<3>(p.S, d_tmem, (uint32_t)p.block_n, a_lo, sub16, b_lo, hi, idesc, first); break;
                    case 2: issue_subtiles<2>(p.S, d_tmem, (uint32_t)p.block_n, a_lo, sub16, b_lo, hi, idesc, first); break;
                    default: issue_subtiles<1>(p.S, d_tmem, (uint32_t)p.block_n, a_lo, sub16, b_lo, hi, idesc, first); break;
                    }
                    if (++kw == p.KW) { kw = 0; ++kh; }
                    if (++in_group == p.b_group || tap + 1 == taps) {      // group consumed: release the stage
                        in_group = 0;
                        ptx::umma_commit_elect(&b_empty[bs]);
                        if (++bs == p.b_stages) { bs = 0; bph ^= 1; }
                    }
                }
                ptx::umma_commit_elect(&a_empty[as]);
                if (++as == kAStages) { as = 0; aph ^= 1; }
            }
            ptx::umma_commit_elect(&tmem_full[acc]);
            if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1; }
        }
        // all MMAs of this CTA are issued: the next kernel in the stream may now be scheduled onto SMs as they drain.
        // (Triggering at kernel start parks its CTAs -- with their shared memory -- on SMs a sibling branch could use.)
        ptx::pdl_launch_dependents();
        if (prof && lane == 0) { atomicAdd(&p.prof[3], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[4], (unsigned long long)w_full); atomicAdd(&p.prof[5], (unsigned long long)w_acc); }
    } else if (warp >= 4) {
        // ===== epilogue =====
        const float* bias = stage_bias(p, bias_s);
        const int e = (warp - 4) & 3, half = (warp - 4) >> 2;
        const int row = e * 32 + lane;
        int acc = 0, acc_phase = 0;
        const bool prof = p.prof != nullptr && threadIdx.x == 128;
        long long w_epi = 0, n_tiles_done = 0;
        const long long t_begin = prof ? clock64() : 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            ++n_tiles_done;
            const int tile_ = p.reverse ? total_tiles - 1 - tile : tile;
            const int q_tile = tile_ / p.n_tiles, n_tile = tile_ - q_tile * p.n_tiles;
            timed_wait(&tmem_full[acc], acc_phase, prof, &w_epi);
            ptx::tc_fence_after();
            // a tile of at most 32 columns cannot be split between the two warps of a lane quarter by columns (TMEM is read 32
            // columns at a time): they take alternate sub-tiles instead, so that all eight epilogue warps work on it
            const bool by_subtile = min(p.block_n, p.N_total - n_tile * p.block_n) <= 32 && p.S > 1;
            for (int s = 0; s < p.S; ++s) {
                if (by_subtile && (s & 1) != half) continue;
                const uint32_t q = (uint32_t)q_tile * (uint32_t)T + (uint32_t)(s * kBlockM + row);      // < 2^31 by plan
                const int wp = (int)(q % (uint32_t)p.Wp);
                const uint32_t t = q / (uint32_t)p.Wp;
                const int hp = (int)(t % (uint32_t)p.Hp);
                const long long img = (long long)(t / (uint32_t)p.Hp);
                const bool row_ok = (long long)q < p.Q && wp < p.OW && hp < p.OH;
                const long long m = (img * p.OH + hp) * p.OW + wp;
                const uint32_t t_row = tmem_base + (uint32_t)(acc * acc_cols + s * p.block_n) + ((uint32_t)(e * 32) << 16);
                epilogue_row(p, bias, t_row, n_tile, m, row_ok, by_subtile ? 0 : half);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty[acc]);
            if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1; }
        }
        if (prof) { atomicAdd(&p.prof[6], (unsigned long long)(clock64() - t_begin)); atomicAdd(&p.prof[7], (unsigned long long)w_epi); atomicAdd(&p.prof[8], (unsigned long long)n_tiles_done); }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kTmemCols>(tmem_base);
    ptx::tl_stamp_max(p.tl ? p.tl + 1 : nullptr);
}

}  // namespace

int plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan, std::string* err,
         double* est_cycles, bool may_use_halo, bool pair, bool wide_halo, int halo_S) {
    auto bad = [&](const std::string& m) { if (err) *err = "conv_pm plan: " + m; return FTF_INVALID_ARGUMENT; };
    if (n_seg < 1 || n_seg > kConvMaxSegments) return bad("segment count");
    if (g.Cin % 8 || g.in_ld % 8 || g.Cout % 16) return bad("Cin and the input row stride must be multiples of 8, Cout of 16");
    if (((uintptr_t)g.in | (uintptr_t)g.weights) & 15) return bad("unaligned tensor base");
    const int OH = (g.H + 2 * g.pad_h - g.KH) / g.stride + 1, OW = (g.W + 2 * g.pad_w - g.KW) / g.stride + 1;
    if (OH <= 0 || OW <= 0) return bad("empty output");
    for (int i = 0; i < n_seg; ++i)
        if (segs[i].col_begin % 16 || segs[i].coff % 8 || segs[i].ld % 8 || (i && segs[i].col_begin <= segs[i - 1].col_begin))
            return bad("segment alignment / order");
    auto enc_tiled = get_tensor_map_encode_tiled();
    auto enc_im2col = get_encode_im2col();
    if (!enc_tiled || !enc_im2col) { if (err) *err = "cuTensorMapEncode* entry points not found"; return FTF_INTERNAL; }

    ConvParams& p = plan->p;
    p = ConvParams{};
    p.OH = OH; p.OW = OW; p.stride = g.stride; p.pad_h = g.pad_h; p.pad_w = g.pad_w; p.KH = g.KH; p.KW = g.KW;
    p.Cin = g.Cin;
    p.N_total = g.Cout;
    p.M = g.n_max * OH * OW;
    p.m_tiles = (p.M + kBlockM - 1) / kBlockM;
    p.bias = g.bias;
    p.n_seg = n_seg;
    for (int i = 0; i < n_seg; ++i) p.seg[i] = segs[i];
    plan->grid = sm_count;
    static const bool want_prof = getenv("FTF_CONV_PROFILE") && getenv("FTF_CONV_PROFILE")[0] == '1';
    if (want_prof) {
        void* sym = nullptr;
        if (cudaGetSymbolAddress(&sym, g_conv_prof) == cudaSuccess) p.prof = (unsigned long long*)sym; else cudaGetLastError();
    }
    // ---- tile shape: (N tile, im2col or halo window with S sub-tiles) that minimises
    //      waves x (max(MMA cycles, operand bytes / per-SM fill bandwidth) + fixed cost).  Every tcgen05.mma with
    //      M = 128 costs about 128 cycles whatever N is (measured), so fewer, wider N tiles are cheaper. ----
    const bool env_halo = !(getenv("FTF_NO_HALO") && getenv("FTF_NO_HALO")[0] == '1');
    const bool allow_halo = env_halo && may_use_halo;
    const int taps = g.KH * g.KW;
    const int Hp = g.H + 2 * g.pad_h, Wp = g.W + 2 * g.pad_w;
    const int halo = (g.KH - 1) * Wp + (g.KW - 1);
    const long long Q = (long long)g.n_max * Hp * Wp;
    static const bool allow_kc32 = !(getenv("FTF_NO_KC32") && getenv("FTF_NO_KC32")[0] == '1');
    const double kFill = 36.0, kFixed = 2500.0;
    // experiment switches: force the number of N tiles / halo sub-tiles the shape search may pick
    const int want_nt = getenv("FTF_PM_NT") ? atoi(getenv("FTF_PM_NT")) : 0;
    const int force_nt = want_nt >= (g.Cout + 255) / 256 && want_nt <= (g.Cout + 31) / 32 ? want_nt : 0;     // only where the layer has that many N tiles
    const int force_S = halo_S ? halo_S : (getenv("FTF_PM_S") ? atoi(getenv("FTF_PM_S")) : 0),
              force_grp = getenv("FTF_PM_BGROUP") ? atoi(getenv("FTF_PM_BGROUP")) : 0;
    const long long kSmem = 208 * 1024;
    double best = -1.0;
    for (int kc = 64; kc >= 32; kc -= 32) {
        const int chunks = (g.Cin + kc - 1) / kc, rb = 2 * kc, ksteps = kc / 16;
        const int total_ksteps = (chunks - 1) * ksteps + (g.Cin - kc * (chunks - 1) + 15) / 16;   // all-zero K steps are skipped
        // 32-channel K blocks (64-byte rows, SWIZZLE_64B) only for Cin <= 32, where they halve the operand window without
        // adding K blocks.  For wider inputs they split K into more blocks, and a block's fixed cost (barrier round trip,
        // commit) outweighs the MMAs saved; the MMA itself runs as fast on 64-byte rows as on 128-byte ones (probed).
        // A fixed rule, so that every candidate kernel of a layer walks K in the same order and produces the same bits.
        const bool kc32_saves = allow_kc32 && g.Cin <= 32;
        if ((kc == 32) != kc32_saves) continue;
        const double real_frac = (double)g.Cin / (chunks * kc);
        const long long a_tile = (long long)kBlockM * rb;
        for (int nt = (g.Cout + 255) / 256; nt <= (g.Cout + 31) / 32; ++nt) {
            const int bn = ((g.Cout + nt - 1) / nt + 15) / 16 * 16;
            if (bn > 256) continue;
            if (force_nt && nt != force_nt) continue;
            const int tiles_n = (g.Cout + bn - 1) / bn;
            const long long b_bytes = (long long)bn * rb;
            if (wide_halo || halo_S) {
                // only the (wide / S-pinned) halo shapes below are candidates
            } else if (!pair) {   // one im2col load per tap.  Every operand byte is written to shared memory once and read once: 64 + N/2 cycles per MMA
                const long long tiles = (long long)p.m_tiles * tiles_n, waves = (tiles + sm_count - 1) / sm_count;
                const double cyc = (double)taps * total_ksteps * (64.0 + bn / 2.0), bytes = (double)taps * chunks * (a_tile + b_bytes) * real_frac;
                const double t = waves * ((cyc > bytes / kFill ? cyc : bytes / kFill) + kFixed);
                if (best < 0 || t < best) {
                    best = t; p.halo_mode = 0; p.block_n = bn; p.n_tiles = tiles_n; p.kc = kc; p.cin_chunks = chunks;
                    p.a_tile_bytes = (int)a_tile; p.b_stage_bytes = (int)b_bytes;
                }
            } else {       // CTA pair: 256 pixels per work item, each CTA holds half of the weight rows
                if (sm_count < 2) continue;
                const long long items = (long long)((p.m_tiles + 1) / 2) * tiles_n, workers = sm_count / 2, waves = (items + workers - 1) / workers;
                // Measured (tools/probe_mainloop.cu): a K block costs the issuing thread ~600 cycles whatever the MMA shape, so a
                // pair -- same K blocks per worker, half the workers -- is modelled slightly SLOWER than one CTA per tile and
                // is only chosen where the plan-time measurement says it wins (less weight traffic from L2).
                const double per_mma = 64.0 + bn / 2.0;
                const double cyc = (double)taps * total_ksteps * per_mma * 1.02, bytes = (double)taps * chunks * (a_tile + b_bytes / 2) * real_frac;
                const double t = waves * ((cyc > bytes / kFill ? cyc : bytes / kFill) + kFixed);
                if (best < 0 || t < best) {
                    best = t; p.halo_mode = 0; p.pair = 1; p.block_n = bn; p.n_tiles = tiles_n; p.kc = kc; p.cin_chunks = chunks;
                    p.a_tile_bytes = (int)a_tile; p.b_stage_bytes = (int)(b_bytes / 2);
                }
                continue;
            }
            if (!(allow_halo && g.stride == 1 && taps > 1) || Q >= (1ll << 31) - 4096) continue;
            for (int S = 1; S <= 4; S *= 2) {
                if (force_S && S != force_S) continue;
                // Two accumulator sets let the epilogue of one tile overlap the MMAs of the next.  The "wide" halo candidate
                // gives that up for twice the sub-tiles per weight tile (one set of S * bn <= 512 columns): half the weight
                // bytes pulled from L2 per pixel, which is what counts where sibling layers saturate the L2 -> SM path and a
                // CTA only has a tile or two to work on anyway.
                const bool one_set = 2 * S * bn > kTmemCols;
                if (S * bn > kTmemCols || one_set != wide_halo) continue;
                const int T = S * kBlockM;
                const int loads = (T + halo + kHaloLoadRows - 1) / kHaloLoadRows;
                const long long a_bytes = (long long)loads * kHaloLoadRows * rb;
                const long long room = kSmem - 2 * a_bytes;
                if (room < 2 * b_bytes) continue;
                const long long tiles = (Q + T - 1) / T * tiles_n, waves = (tiles + sm_count - 1) / sm_count;
                const double cyc = (double)total_ksteps * taps * S * 128.0;
                const double bytes = chunks * (a_bytes + (double)taps * b_bytes) * real_frac;
                const double t = waves * ((cyc > bytes / kFill ? cyc : bytes / kFill) + kFixed);
                if (best < 0 || t < best) {
                    best = t;
                    p.halo_mode = 1; p.Hp = Hp; p.Wp = Wp; p.Q = Q; p.S = S; p.block_n = bn; p.n_tiles = tiles_n; p.kc = kc; p.cin_chunks = chunks;
                    p.a_loads = loads; p.a_stage_bytes = (int)a_bytes;
                    // weight tiles of several taps share a stage (one barrier round trip per group): as many as leave three
                    // stages, at most one filter row... the whole filter when its tiles are tiny
                    int grp = (int)(room / (3 * b_bytes));
                    if (grp > taps) grp = taps;
                    if (grp < 1) grp = 1;
                    if (force_grp) grp = force_grp < taps ? force_grp : taps;
                    p.b_group = grp; p.b_stage_bytes = (int)(grp * b_bytes);
                    p.b_stages = (int)(room / p.b_stage_bytes > kMaxStages ? kMaxStages : room / p.b_stage_bytes);
                    if (p.b_stages < 2) { p.b_group = 1; p.b_stage_bytes = (int)b_bytes; p.b_stages = (int)(room / b_bytes > kMaxStages ? kMaxStages : room / b_bytes); }
                    p.acc_stages = one_set ? 1 : 2;
                }
            }
        }
    }
    if (best < 0) return bad(wide_halo || halo_S ? "no such halo shape for this layer" : "no tile shape fits");
    p.ksteps_last = (g.Cin - p.kc * (p.cin_chunks - 1) + 15) / 16;
    p.wide_stores = 1;
    for (int i = 0; i < n_seg; ++i) {
        const int esz = (segs[i].flags & kSegF32) ? 4 : 2;
        if (((uintptr_t)segs[i].out & 31) || (segs[i].ld * esz) % 32 || (segs[i].coff * esz) % 32) p.wide_stores = 0;
    }
    if (est_cycles) *est_cycles = best;
    if (p.halo_mode) {
        plan->smem = 2 * (size_t)p.a_stage_bytes + (size_t)p.b_stages * p.b_stage_bytes + 1024 + 512;
    } else {
        const int stage_bytes = p.a_tile_bytes + p.b_stage_bytes;
        p.stages = (200 * 1024) / stage_bytes;
        if (p.stages > kMaxStages) p.stages = kMaxStages;
        plan->smem = (size_t)p.stages * stage_bytes + 1024 /*alignment slack*/ + 256 /*barriers*/;
    }

    const CUtensorMapSwizzle swz = p.kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    // A: NHWC input seen through an im2col window. dims fastest-first {C, W, H, N}.
    {
        cuuint64_t dims[4] = {(cuuint64_t)g.Cin, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.n_max};
        cuuint64_t strides[3] = {(cuuint64_t)g.in_ld * 2, (cuuint64_t)g.W * g.in_ld * 2, (cuuint64_t)g.H * g.W * g.in_ld * 2};
        int lower[2] = {-g.pad_w, -g.pad_h};
        int upper[2] = {g.pad_w - (g.KW - 1), g.pad_h - (g.KH - 1)};
        if (p.halo_mode) { upper[0] = g.pad_w; upper[1] = g.pad_h; }      // the whole padded image, one pixel per "filter"
        cuuint32_t estr[4] = {1, (cuuint32_t)g.stride, (cuuint32_t)g.stride, 1};
        CUresult rc = enc_im2col(&plan->tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)g.in, dims, strides, lower, upper,
                                 (cuuint32_t)p.kc, p.halo_mode ? kHaloLoadRows : kBlockM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { if (err) *err = "cuTensorMapEncodeIm2col failed, code " + std::to_string((int)rc); return FTF_INTERNAL; }
    }
    // B: weights [Cout][K] with K = KH*KW*Cin contiguous.
    {
        const cuuint64_t K = (cuuint64_t)g.KH * g.KW * g.Cin;
        cuuint64_t dims[2] = {K, (cuuint64_t)g.Cout};
        cuuint64_t strides[1] = {K * 2};
        cuuint32_t box[2] = {(cuuint32_t)p.kc, (cuuint32_t)(p.pair ? p.block_n / 2 : p.block_n)};
        cuuint32_t estr[2] = {1, 1};
        CUresult rc = enc_tiled(&plan->tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)g.weights, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { if (err) *err = "cuTensorMapEncodeTiled (weights) failed, code " + std::to_string((int)rc); return FTF_INTERNAL; }
    }
    static std::once_flag attr_once;
    static cudaError_t attr_rc = cudaSuccess;
    std::call_once(attr_once, [] {
        attr_rc = cudaFuncSetAttribute(conv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    });
    // the attribute is per device: set it again when planning on another device
    cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc_kernel)", err);
    e = cudaFuncSetAttribute(conv_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc_kernel<pair>)", err);
    e = cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_halo_kernel)", err);
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
    static const bool no_store = getenv("FTF_EPI_NO_STORE") != nullptr;
    p.dbg_no_store = no_store;
    p.M = n * p.OH * p.OW;
    if (p.M <= 0) return FTF_OK;
    p.m_tiles = (p.M + kBlockM - 1) / kBlockM;
    if (p.halo_mode) {
        p.Q = (long long)n * p.Hp * p.Wp;
        const int T = p.S * kBlockM;
        const long long tiles = (p.Q + T - 1) / T * p.n_tiles;
        const int grid = (int)(tiles < plan.grid ? tiles : plan.grid);
        FTF_CUDA(launch_pdl(conv_halo_kernel, dim3(grid), dim3(kThreads), plan.smem, s, plan.tm_a, plan.tm_b, p), err);
        count_launch();
        return FTF_OK;
    }
    if (p.pair) {
        const int items = (p.m_tiles + 1) / 2 * p.n_tiles, workers = plan.grid / 2;
        const int grid = 2 * (items < workers ? items : workers);
        FTF_CUDA(launch_pdl_cluster(conv_tc_kernel<true>, dim3(grid), dim3(kThreads), plan.smem, s, /*cluster=*/2, plan.tm_a, plan.tm_b, p), err);
        count_launch();
        return FTF_OK;
    }
    const int tiles = p.m_tiles * p.n_tiles;
    const int grid = tiles < plan.grid ? tiles : plan.grid;
    FTF_CUDA(launch_pdl(conv_tc_kernel<false>, dim3(grid), dim3(kThreads), plan.smem, s, plan.tm_a, plan.tm_b, p), err);
    count_launch();
    return FTF_OK;
}

}  // namespace pm
}  // namespace ftf
