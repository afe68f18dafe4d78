// K7  row softmax + stable descending top-k (label post-processing).
//
// Reference: flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples/
// inception/InceptionModel.scala:81-89 --
//   row.toList.zip(model.labels).sortWith(_._1 > _._1).take(take)
// zip truncates the row to min(N, #labels) columns; List.sortWith is a stable sort, so equal scores
// keep ascending column order.  The k-th label is therefore the maximum, under (score desc, column asc),
// of the columns that come after the (k-1)-th in that order: k rounds of an arg-max reduction done with
// warp shuffles, no sort and no per-row allocation.  NaN scores (ordering unspecified in the reference)
// are never selected ahead of a number.
// The softmax is the final node ("output") of the frozen graph the reference feeds
// (InceptionModel.scala:58-62); here it is fused in front of the selection: one CTA per row, the row
// staged once in shared memory.
#include "common.h"
#include <math_constants.h>

namespace ftf {

namespace {

constexpr int kThreads = 128;

struct Cand {
    float v;
    int i;
};
// a before b in (score desc, column asc); i < 0 marks "none"
__device__ __forceinline__ bool before(const Cand& a, const Cand& b) {
    if (a.i < 0) return false;
    if (b.i < 0) return true;
    return a.v > b.v || (a.v == b.v && a.i < b.i);
}
__device__ __forceinline__ Cand warp_best(Cand c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Cand d;
        d.v = __shfl_xor_sync(0xffffffffu, c.v, o);
        d.i = __shfl_xor_sync(0xffffffffu, c.i, o);
        if (before(d, c)) c = d;
    }
    return c;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(kThreads)
softmax_topk_rows(const float* __restrict__ x, int n, int cols, int k, int apply_softmax,
                  float* __restrict__ probs, int32_t* __restrict__ idx, float* __restrict__ score) {
    extern __shared__ float row[];                 // n floats
    __shared__ float red_f[kThreads / 32];
    __shared__ Cand red_c[kThreads / 32];
    __shared__ Cand picked;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* xr = x + (size_t)blockIdx.x * n;

    float mx = -CUDART_INF_F;
    for (int c = threadIdx.x; c < n; c += kThreads) {
        const float v = __ldg(xr + c);
        row[c] = v;
        mx = fmaxf(mx, v);
    }
    if (apply_softmax) {
        mx = warp_max(mx);
        if (lane == 0) red_f[warp] = mx;
        __syncthreads();
        mx = red_f[0];
#pragma unroll
        for (int w = 1; w < kThreads / 32; ++w) mx = fmaxf(mx, red_f[w]);
        __syncthreads();
        float s = 0.0f;
        for (int c = threadIdx.x; c < n; c += kThreads) {
            const float e = expf(row[c] - mx);
            row[c] = e;
            s += e;
        }
        s = warp_sum(s);
        if (lane == 0) red_f[warp] = s;
        __syncthreads();
        s = 0.0f;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) s += red_f[w];
        for (int c = threadIdx.x; c < n; c += kThreads) {
            const float pv = __fdiv_rn(row[c], s);
            row[c] = pv;
            if (probs) probs[(size_t)blockIdx.x * n + c] = pv;
        }
    }
    __syncthreads();

    Cand prev;
    prev.v = CUDART_INF_F;
    prev.i = -1;  // nothing picked yet
    for (int j = 0; j < k; ++j) {
        Cand best;
        best.v = 0.0f;
        best.i = -1;
        for (int c = threadIdx.x; c < cols; c += kThreads) {
            Cand cur;
            cur.v = row[c];
            cur.i = c;
            // eligible: strictly after `prev` in (score desc, column asc)
            const bool after_prev = prev.i < 0 ? (cur.v == cur.v)
                                               : (cur.v < prev.v || (cur.v == prev.v && cur.i > prev.i));
            if (after_prev && before(cur, best)) best = cur;
        }
        best = warp_best(best);
        if (lane == 0) red_c[warp] = best;
        __syncthreads();
        if (threadIdx.x == 0) {
            Cand b = red_c[0];
#pragma unroll
            for (int w = 1; w < kThreads / 32; ++w)
                if (before(red_c[w], b)) b = red_c[w];
            picked = b;
            idx[(size_t)blockIdx.x * k + j] = b.i;          // -1 when the row ran out of columns
            score[(size_t)blockIdx.x * k + j] = b.i < 0 ? 0.0f : b.v;
        }
        __syncthreads();
        prev = picked;
        if (prev.i < 0) {  // exhausted: pad the rest
            if (threadIdx.x == 0)
                for (int t = j + 1; t < k; ++t) {
                    idx[(size_t)blockIdx.x * k + t] = -1;
                    score[(size_t)blockIdx.x * k + t] = 0.0f;
                }
            break;
        }
    }
}

}  // namespace

int launch_softmax_topk(const float* x, int64_t m, int n, int n_labels, int k, int apply_softmax,
                        float* probs, int32_t* idx, float* score, cudaStream_t s, std::string* err) {
    if (m < 0 || n < 0 || k < 0 || n_labels < 0 || (m > 0 && n > 0 && !x) || (m > 0 && k > 0 && (!idx || !score))) {
        if (err) *err = "softmax_topk: bad shape or null buffer";
        return FTF_INVALID_ARGUMENT;
    }
    if ((size_t)n * sizeof(float) > 200 * 1024) {
        if (err) *err = "softmax_topk: rows longer than 51200 columns are not supported";
        return FTF_UNIMPLEMENTED;
    }
    if (m == 0 || (k == 0 && !(apply_softmax && probs))) return FTF_OK;
    const int cols = n < n_labels ? n : n_labels;
    const size_t smem = (size_t)(n > 0 ? n : 1) * sizeof(float);
    FTF_CUDA(cudaFuncSetAttribute(softmax_topk_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
    softmax_topk_rows<<<(unsigned)m, kThreads, smem, s>>>(x, n, cols, k, apply_softmax, probs, idx, score);
    count_launch();
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

}  // namespace ftf
