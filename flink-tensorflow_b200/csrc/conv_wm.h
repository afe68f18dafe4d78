// conv_wm.h -- implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 + TMEM + TMA im2col).
#pragma once
#include "conv_common.h"

namespace ftf {
namespace wm {   // weight-major: output channels on the UMMA M side, pixels on the N side

struct ConvParams {
    long long M;                // real output pixels = n * OH * OW
    int OH, OW;
    int stride, pad_h, pad_w, KH, KW;
    int Cin;
    int kc;                     // channels per K block: 64 (128-byte rows, SWIZZLE_128B) or 32 (64-byte rows, SWIZZLE_64B)
    int cin_chunks;             // ceil(Cin / kc)
    int Cout;                   // output channels
    int m_tiles;                // ceil(Cout / 128): weight-row tiles (the UMMA M side)
    int np;                     // pixels per UMMA (the N side): 128 or 256
    int stages;                 // im2col mode: smem ring depth of (weights | pixels) stages
    int w_stage_bytes;          // 128 rows * 2*kc
    int p_stage_bytes;          // im2col mode: np rows * 2*kc;  halo mode: window rows * 2*kc
    const float* bias;          // [Cout] or NULL
    int n_seg;
    ConvSegment seg[kConvMaxSegments];
    // ---- halo mode (stride-1 KxK): one shared-memory window of the padded input serves every filter tap ----
    int halo_mode;
    int Hp, Wp;                 // padded input extent: H + 2*pad_h, W + 2*pad_w
    long long Q;                // virtual outputs = n * Hp * Wp (real where hp < OH and wp < OW)
    int S;                      // np-pixel sub-tiles per CTA tile (each with its own accumulator)
    int p_loads;                // im2col loads of kHaloLoadRows rows that fill one window
    int w_stages;               // halo mode: weight-tile ring depth
    int w_group;                // halo mode: filter taps whose weight tiles share one stage
    int acc_stages;             // 1 or 2 TMEM accumulator sets of S * np columns
    int ksteps_last;            // K=16 steps holding real channels in the LAST channel chunk (the all-zero rest is skipped)
    int reverse;                // 1: walk the tiles from the last image to the first (see InceptionV3: the consumer of a tensor larger than L2's share
                                // walks it in the opposite direction of its producer, so that it starts on what is still cached)
    unsigned long long* prof;   // diagnostic cycle counters (FTF_CONV_PROFILE=1), else NULL
    unsigned long long* tl;     // timeline stamps of this launch (FTF_TIMELINE), else NULL
};

constexpr int kHaloLoadRows = 128;

struct ConvPlan {
    CUtensorMap tm_p;           // im2col view of the NHWC input (pixels)
    CUtensorMap tm_w;           // weights [Cout, KH*KW*Cin] K-major
    ConvParams p;
    int grid;
    size_t smem;
};

// Encodes the tensor maps and tile shapes for one convolution (called once, at network build).
// *est_cycles: modelled cycles of the whole layer on sm_count SMs (for choosing between the two kernels).
int plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan,
         std::string* err, double* est_cycles, bool may_use_halo = true);
// Enqueues the convolution for the first n images.
int launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err, unsigned long long* tl = nullptr, int reverse = 0);
// Diagnostic: copies (and optionally clears) the 16 pipeline cycle counters the kernels accumulate when the
// environment has FTF_CONV_PROFILE=1: [0] CTAs, [1] producer cycles, [2] producer waiting for free stages,
// [3] MMA-issuer cycles, [4] waiting for operands, [5] waiting for a drained accumulator, [6] epilogue cycles,
// [7] epilogue waiting for an accumulator, [8] tiles.
int profile_read(unsigned long long* out16, int reset);

}  // namespace wm
}  // namespace ftf
