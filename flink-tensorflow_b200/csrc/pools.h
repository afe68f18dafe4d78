// pools.h -- launchers of the pooling kernels (pools.cu).
#pragma once
#include "common.h"
#include <string>

namespace ftf {

// bf16 NHWC [n,H,W,C] -> bf16 [n,OH,OW,out_ld] at channel offset out_off; OH = (H-3)/2+1.
int launch_maxpool3x3s2(const void* in, void* out, int n, int H, int W, int C, int out_ld, int out_off, cudaStream_t s,
                        std::string* err, unsigned long long* tl = nullptr, int ctas_per_sm = 16, int reverse = 0);
// out = bf16(relu(avgpool3x3_s1_p1(pre)/9 + bias)); pre fp32 [n,H,W,C] (a 1x1 convolution's raw output).
int launch_avgpool3x3_bias_relu(const float* pre, const float* bias, void* out, int n, int H, int W, int C, int out_ld,
                                int out_off, cudaStream_t s, std::string* err, unsigned long long* tl = nullptr, int ctas_per_sm = 16);
// bf16 [n,HW,C] -> bf16 [n,C]: fp32 mean over the pixels.
int launch_global_avgpool(const void* in, void* out, int n, int HW, int C, cudaStream_t s, std::string* err, unsigned long long* tl = nullptr);

// ctas_per_sm: grid cap per SM; 1 for a pooling kernel that runs beside convolutions on a sibling stream (see pools.cu).
// tl (all launchers): timeline stamps of this launch for the FTF_TIMELINE diagnostics, else NULL.

}  // namespace ftf
