// inception.cu -- label model (placeholder until the convolution kernels land).
#include "inception.h"
#include <cuda_bf16.h>

namespace ftf {

struct InceptionV3::Impl {};
InceptionV3::~InceptionV3() { delete impl_; }

int InceptionV3::create(int, int, uint64_t, int, int, std::unique_ptr<InceptionV3>*, std::string* err) {
    if (err) *err = "builtin:inception_v3 is not built yet";
    return FTF_UNIMPLEMENTED;
}
int InceptionV3::forward(const void*, int64_t, float*, cudaStream_t, std::string* err) {
    if (err) *err = "builtin:inception_v3 is not built yet";
    return FTF_UNIMPLEMENTED;
}

namespace {
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

}  // namespace ftf
