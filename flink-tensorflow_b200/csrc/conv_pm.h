// conv_pm.h -- implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 + TMEM + TMA im2col).
#pragma once
#include "conv_common.h"

namespace ftf {
namespace pm {   // pixel-major: pixels on the UMMA M side, output channels on the N side

struct ConvParams {
    int M;                      // output pixels = n * OH * OW
    int OH, OW;
    int stride, pad_h, pad_w, KH, KW;
    int Cin, cin_chunks;        // cin_chunks = ceil(Cin / kc)
    int kc;                     // channels per K block: 64 (128-byte rows, SWIZZLE_128B) or 32 (64-byte rows, SWIZZLE_64B)
    int a_tile_bytes;           // im2col mode: 128 rows * 2*kc
    int N_total;                // output channels (GEMM N)
    int block_n;                // N tile: multiple of 16, <= 256
    int m_tiles, n_tiles;
    int stages;                 // smem pipeline depth
    int b_stage_bytes;          // block_n * 128
    const float* bias;          // [N_total]
    int n_seg;
    ConvSegment seg[kConvMaxSegments];
    // ---- halo mode (stride-1 KxK): one shared-memory window of the padded input serves every filter tap ----
    int halo_mode;              // 0: one im2col load per tap (any stride); 1: halo window
    int Hp, Wp;                 // padded input extent: H + 2*pad_h, W + 2*pad_w
    long long Q;                // virtual outputs = n * Hp * Wp (valid where hp < OH and wp < OW)
    int S;                      // 128-row sub-tiles per CTA tile
    int a_loads;                // im2col loads of kHaloLoadRows rows that fill one window
    int a_stage_bytes;          // a_loads * kHaloLoadRows * 128
    int b_stages;               // weight-tile ring depth
    int b_group;                // filter taps whose weight tiles share one stage (halo mode)
    int acc_stages;             // 1 or 2 TMEM accumulator sets of S * block_n columns
    int pair;                   // 1: CTA-pair kernel (cta_group::2): tiles of 256 pixels, each CTA holds block_n/2 weight rows
    int wide_stores;            // 1: every segment's rows and channel offset are 32-byte aligned (256-bit epilogue stores)
    int ksteps_last;            // K=16 steps holding real channels in the LAST channel chunk (the all-zero rest is skipped)
    int dbg_no_store;           // measurement only (FTF_EPI_NO_STORE): the im2col kernel's epilogue computes but does not store
    int reverse;                // 1: walk the tiles from the last image to the first (see InceptionV3: the consumer of a tensor larger than L2's share
                                // walks it in the opposite direction of its producer, so that it starts on what is still cached)
    unsigned long long* prof;   // diagnostic cycle counters (FTF_CONV_PROFILE=1), else NULL
    unsigned long long* tl;     // timeline stamps of this launch (FTF_TIMELINE), else NULL
};

constexpr int kHaloLoadRows = 64;

struct ConvPlan {
    CUtensorMap tm_a;           // im2col view of the NHWC input
    CUtensorMap tm_b;           // weights [Cout, KH*KW*Cin] K-major
    ConvParams p;
    int grid;
    size_t smem;
};

// Encodes the tensor maps and tile shapes for one convolution (called once, at network build).
// *est_cycles: modelled cycles of the whole layer on sm_count SMs (for choosing between the two kernels).
int plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan,
         std::string* err, double* est_cycles, bool may_use_halo = true, bool pair = false, bool wide_halo = false, int halo_S = 0);
// Enqueues the convolution for the first n images.
int launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err, unsigned long long* tl = nullptr, int reverse = 0);
// Diagnostic: copies (and optionally clears) the 16 pipeline cycle counters the kernels accumulate when the
// environment has FTF_CONV_PROFILE=1: [0] CTAs, [1] producer cycles, [2] producer waiting for free stages,
// [3] MMA-issuer cycles, [4] waiting for operands, [5] waiting for a drained accumulator, [6] epilogue cycles,
// [7] epilogue waiting for an accumulator, [8] tiles.
int profile_read(unsigned long long* out16, int reset);

}  // namespace pm
}  // namespace ftf
