// conv_common.h -- types shared by the two tensor-core convolution kernels and their dispatcher.
#pragma once
#include "common.h"
#include <string>

namespace ftf {

constexpr int kConvMaxSegments = 6;

// Where a range of output channels goes.  Channel ranges start at multiples of 16.
struct ConvSegment {
    void* out;        // base of the destination tensor [pixels, ld]
    int ld;           // destination row stride in elements (channels of the concat buffer)
    int coff;         // channel offset inside the destination row
    int col_begin;    // first output channel of this segment
    int flags;        // kSegBias | kSegRelu | kSegF32
};
enum { kSegBias = 1, kSegRelu = 2, kSegF32 = 4 };

struct ConvGeometry {
    const void* in;             // bf16 NHWC [n_max, H, W, in_ld]; channels [0, Cin) are convolved
    int n_max, H, W, Cin, in_ld;
    const void* weights;        // bf16 [Cout][KH][KW][Cin]
    const float* bias;          // fp32 [Cout]
    int Cout, KH, KW, stride, pad_h, pad_w;
};

}  // namespace ftf
