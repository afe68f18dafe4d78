// common.h -- shared declarations of the runner library (host + device).
#pragma once
#include <utility>
#include <stdlib.h>
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <string>
#include "../../include/ftf.h"

namespace ftf {

// Per-thread message of the last failing call that had no runner to attach it to.
void set_thread_error(const std::string& msg);
const char* thread_error();

// Formats "<what>: <cudaGetErrorString>" into *err and returns FTF_INTERNAL.
int cuda_fail(cudaError_t e, const char* what, std::string* err);

#define FTF_CUDA(call, errp)                                                       \
    do {                                                                           \
        cudaError_t e__ = (call);                                                  \
        if (e__ != cudaSuccess) return ::ftf::cuda_fail(e__, #call, (errp));       \
    } while (0)

// Driver entry points fetched through the runtime (no link-time dependency on libcuda).
typedef CUresult (*tensor_map_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                               const cuuint64_t*, const cuuint64_t*,
                                               const cuuint32_t*, const cuuint32_t*,
                                               CUtensorMapInterleave, CUtensorMapSwizzle,
                                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
tensor_map_encode_tiled_fn get_tensor_map_encode_tiled();

int device_sm_count(int device);

// ---- kernel launchers (one per .cu file); each returns an FTF status and fills *err. ----------
int launch_affine_f32(const float* x, float* y, int64_t n, float a, float b, cudaStream_t s,
                      std::string* err);
int launch_resize_norm(const uint8_t* in, int64_t n, int in_h, int in_w, int out_h, int out_w,
                       float mean, float scale, void* out, int out_dtype, cudaStream_t s,
                       std::string* err);
int launch_softmax_topk(const float* x, int64_t m, int n, int n_labels, int k, int apply_softmax,
                        float* probs, int32_t* idx, float* score, cudaStream_t s, std::string* err);

// Launch with programmatic dependent launch allowed (see ptx::pdl_wait): inside a stream or a captured graph the next
// kernel's prologue overlaps this kernel's tail.
// Priority of the kernels the calling thread launches from here on (0: the stream's own).  The block scheduler hands pending
// CTAs of a higher-priority grid to a freed SM first, whichever grid was launched earlier: InceptionV3 gives its convolutions
// a higher priority than the pooling kernel of the sibling branch (see run_ops).  Also recorded into captured graph nodes.
inline int& launch_priority() { static thread_local int p = 0; return p; }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[3];
    static const bool off = getenv("FTF_NO_PDL") != nullptr;      // A/B switch for measurements
    int n = 0;
    if (!off) { at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[n].val.programmaticStreamSerializationAllowed = 1; ++n; }
    if (cluster > 1) { at[n].id = cudaLaunchAttributeClusterDimension; at[n].val.clusterDim.x = (unsigned)cluster; at[n].val.clusterDim.y = 1; at[n].val.clusterDim.z = 1; ++n; }
    if (launch_priority() != 0) { at[n].id = cudaLaunchAttributePriority; at[n].val.priority = launch_priority(); ++n; }
    cfg.attrs = at; cfg.numAttrs = (unsigned)n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    return launch_pdl_cluster(kernel, grid, block, smem, s, 1, std::forward<Args>(args)...);
}

// Count of this library's kernel launches (process-wide, relaxed).
void count_launch(int n = 1);   // n may be negative (a graph recording counted launches that did not run)
uint64_t launches_so_far();

}  // namespace ftf
