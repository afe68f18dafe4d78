// pool_branch_static.cu -- the pool-branch operation (K5) for callers that must not allocate device memory.
//
// ftf_pool_branch_bf16_dev() needs fp32 scratch for the 1x1 convolution's pre-activation and gets it from
// cudaMallocAsync.  Two kinds of caller of this library cannot do that:
//   * a Flink task thread on the per-record path that shares its GPU with other subtasks: a device allocation
//     (and the matching free) may synchronise the whole device and grow the process's footprint, which stalls the
//     other subtasks' streams in the middle of their batches;
//   * a deployment that hands out all HBM at open() time (one budget per runner) and treats any later allocation
//     as a fault -- there a per-call allocation can simply fail.
// For them the scratch lives in this file as a fixed array in device memory: it is part of the library image, so
// every GPU that loads the library carries these 8 MiB until the process exits, and no call here ever allocates.
//
// The array is shared by all calls on a device and keeps whatever the last call left in it.  Nothing has to start
// at zero: the convolution writes every element the pooling step reads (all n*h*w rows, all cout columns) before
// that step runs on the same stream.  Two calls whose GPU work overlaps (different streams, or different threads
// on one device) would write the same scratch at the same time and race; callers must order them.
#include "conv_tc.h"
#include "pools.h"
#include <mutex>

namespace {

constexpr long long kScratchBytes = FTF_POOL_BRANCH_STATIC_MAX_BYTES;   // 8 MiB = 2 Mi fp32 values, e.g. 17x17x192 for 37 images
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_pool_branch_scratch[kScratchBytes];

// Device address of the array, resolved once per device.  In host code the symbol's name is only a handle: its
// device address has to be asked for (with the device current), it cannot be passed to a kernel directly.
std::mutex g_mu;
float* g_scratch_ptr[kMaxDevices];

int scratch_for_device(int dev, float** out, std::string* err) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (dev < 0 || dev >= kMaxDevices) { *err = "pool_branch_static: device ordinal out of range"; return FTF_INVALID_ARGUMENT; }
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_pool_branch_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            *err = std::string("pool_branch_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
            return FTF_INTERNAL;
        }
        g_scratch_ptr[dev] = static_cast<float*>(p);
    }
    *out = g_scratch_ptr[dev];
    return FTF_OK;
}

}  // namespace

// See include/ftf.h.  The first call on a device resolves the scratch address (set-up work that belongs before any
// stream capture); later calls only encode two tensor maps on the host (or find them in the calling thread's plan table)
// and enqueue two kernels.  The FTF_CONV_TUNE environment switch, which makes other planners time their candidates on
// the device, is ignored here.
extern "C" int ftf_pool_branch_bf16_static_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld,
                                               const void* weights, const float* bias, int32_t cout, void* out,
                                               int32_t out_ld, int32_t out_off, void* stream) {
    using namespace ftf;
    std::string err;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_thread_error("pool_branch_static: cudaGetDevice failed"); return FTF_INTERNAL; }
    if (n <= 0 || h <= 0 || w <= 0 || cout <= 0 || !bias) {
        set_thread_error("pool_branch_static: bad shape or null bias");
        return FTF_INVALID_ARGUMENT;
    }
    const long long need = (long long)n * h * w * cout * (long long)sizeof(float);
    if (need > kScratchBytes) {
        set_thread_error("pool_branch_static: n*h*w*cout*4 = " + std::to_string(need) + " bytes exceeds the built-in scratch of " +
                         std::to_string(kScratchBytes) + " bytes; split the batch or use ftf_pool_branch_bf16_dev");
        return FTF_RESOURCE_EXHAUSTED;
    }
    float* pre = nullptr;
    int rc = scratch_for_device(dev, &pre, &err);
    if (rc) { set_thread_error(err); return rc; }
    cudaStream_t s = (cudaStream_t)stream;
    ConvSegment sg{};
    sg.out = pre; sg.ld = cout; sg.coff = 0; sg.col_begin = 0; sg.flags = kSegF32;
    ConvGeometry g{in, n, h, w, cin, in_ld, weights, nullptr, cout, 1, 1, 1, 0, 0};
    // kPlanModelOnly: planning is host work only (tensor-map encoding), whatever FTF_CONV_TUNE says -- no stream, event,
    // launch or synchronisation; a repeated call finds its plan in the calling thread's table
    const ConvPlan* plan = nullptr;
    rc = conv_tc_plan_cached(g, &sg, 1, device_sm_count(dev), dev, &plan, &err, kPlanModelOnly);
    if (!rc) rc = conv_tc_launch(*plan, n, s, &err);
    if (!rc) rc = launch_avgpool3x3_bias_relu(pre, bias, out, n, h, w, cout, out_ld, out_off, s, &err);
    if (rc) set_thread_error(err);
    return rc;
}
