// conv_tc.h -- implicit-GEMM convolution on the 5th-generation tensor cores (tcgen05 + TMEM + TMA im2col).
#pragma once
#include "common.h"
#include <string>

namespace ftf {

constexpr int kConvMaxSegments = 6;

// Where a range of output channels (GEMM columns) goes.  Column ranges start at multiples of 16.
struct ConvSegment {
    void* out;        // base of the destination tensor [M, ld]
    int ld;           // destination row stride in elements (channels of the concat buffer)
    int coff;         // channel offset inside the destination row
    int col_begin;    // first GEMM column of this segment
    int flags;        // kSegBiasRelu | kSegF32
};
enum { kSegBias = 1, kSegRelu = 2, kSegF32 = 4 };

struct ConvParams {
    int M;                      // output pixels = n * OH * OW
    int OH, OW;
    int stride, pad_h, pad_w, KH, KW;
    int Cin, cin_chunks;        // cin_chunks = ceil(Cin / 64)
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
    int acc_stages;             // 1 or 2 TMEM accumulator sets of S * block_n columns
    unsigned long long* prof;   // diagnostic cycle counters (FTF_CONV_PROFILE=1), else NULL
};

constexpr int kHaloLoadRows = 64;

struct ConvPlan {
    CUtensorMap tm_a;           // im2col view of the NHWC input
    CUtensorMap tm_b;           // weights [Cout, KH*KW*Cin] K-major
    ConvParams p;
    int grid;
    size_t smem;
};

struct ConvGeometry {
    const void* in;             // bf16 NHWC [n_max, H, W, in_ld]; channels [0, Cin) are convolved
    int n_max, H, W, Cin, in_ld;
    const void* weights;        // bf16 [Cout][KH][KW][Cin]
    const float* bias;          // fp32 [Cout]
    int Cout, KH, KW, stride, pad_h, pad_w;
};

// Encodes the tensor maps and tile shapes for one convolution (called once, at network build).
int conv_tc_plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan,
                 std::string* err);
// Enqueues the convolution for the first n images.
int conv_tc_launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err);
// Diagnostic: copies (and optionally clears) the 16 pipeline cycle counters the kernels accumulate when the
// environment has FTF_CONV_PROFILE=1: [0] CTAs, [1] producer cycles, [2] producer waiting for free stages,
// [3] MMA-issuer cycles, [4] waiting for operands, [5] waiting for a drained accumulator, [6] epilogue cycles,
// [7] epilogue waiting for an accumulator, [8] tiles.
int conv_tc_profile_read(unsigned long long* out16, int reset);

}  // namespace ftf
