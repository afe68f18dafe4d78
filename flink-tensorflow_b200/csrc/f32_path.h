// f32_path.h -- fp32 kernels of the label model (compute_dtype = FTF_FLOAT): the precision mode.  See f32_path.cu.
#pragma once
#include "conv_common.h"

namespace ftf {
namespace f32 {

// Convolution on fp32 NHWC tensors.  g.in / g.weights are fp32 here ([n,H,W,in_ld] and [Cout][KH][KW][Cin]); every
// segment's destination is fp32 (kSegF32 is implied); kSegBias / kSegRelu as in the tensor-core path.
int launch_conv(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int n, cudaStream_t s, std::string* err);
int launch_maxpool3x3s2(const float* in, float* out, int n, int H, int W, int C, int out_ld, int out_off, cudaStream_t s, std::string* err);
int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, float* out, int n, int H, int W, int C, int out_ld, int out_off,
                                cudaStream_t s, std::string* err);
int launch_global_avgpool(const float* in, float* out, int n, int HW, int C, cudaStream_t s, std::string* err);

}  // namespace f32
}  // namespace ftf
