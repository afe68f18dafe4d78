// K1  y = Add(Mul(a, x), b)  -- the arithmetic of the half_plus_two graph.
//
// Reference: nodes "Mul" = Mul(a/read, x) and "y" = Add(Mul, b/read) of
// models/half_plus_two/saved_model.pb, executed by Session.run at
// flink-tensorflow/src/main/scala/org/apache/flink/contrib/tensorflow/models/ModelFunction.scala:65.
// Two separately rounded fp32 operations (__fmul_rn/__fadd_rn are never contracted to FMA), so the
// result is bit-identical to the CPU kernels for every input including NaN/inf/denormals.
//
// HBM-bound: 8 algorithmic bytes per element (4 read + 4 written), no reuse.  128-bit coalesced
// accesses (256-bit where the buffers allow), 4 independent loads in flight per thread, streaming cache hints (read-once,
// write-once), grid = SMs x resident CTAs so every SM holds the same number of CTAs.
#include "common.h"
#include "ptx.cuh"
#include <stdlib.h>

namespace ftf {

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;  // float4 per thread per iteration

// 256-bit accesses are an sm_100 addition (ld/st.global.v8.b32); they also admit the L2 evict-first
// policy, which suits data that is read once and written once.
struct __align__(32) f32x8 { float v[8]; };
struct __align__(16) f32x4 { float v[4]; };

__device__ __forceinline__ f32x8 ld_stream(const f32x8* p) {
    f32x8 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(f32x8* p, const f32x8& r) {
    asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
                 "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
                 : "memory");
}
__device__ __forceinline__ f32x4 ld_stream(const f32x4* p) {
    f32x4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(f32x4* p, const f32x4& r) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]),
                 "f"(r.v[3])
                 : "memory");
}
__device__ __forceinline__ float affine1(float x, float a, float b) {
    return __fadd_rn(__fmul_rn(a, x), b);
}
template <typename V>
__device__ __forceinline__ V affine_vec(V r, float a, float b) {
#pragma unroll
    for (int i = 0; i < (int)(sizeof(V) / 4); ++i) r.v[i] = affine1(r.v[i], a, b);
    return r;
}

// Body: nv vectors of V (32 or 16 bytes), both pointers aligned to sizeof(V).
template <typename V>
__global__ void __launch_bounds__(kThreads) affine_f32_vec(const V* __restrict__ x, V* __restrict__ y, int64_t nv,
                                                           float a, float b) {
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    // main loop: kUnroll independent wide loads before the first use
    for (; i + (kUnroll - 1) * stride < nv; i += kUnroll * stride) {
        V v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) v[u] = ld_stream(x + i + u * stride);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) st_stream(y + i + u * stride, affine_vec(v[u], a, b));
    }
    for (; i < nv; i += stride) st_stream(y + i, affine_vec(ld_stream(x + i), a, b));
}

// Bulk variant for large 16-byte-aligned buffers: the copy engine moves the data (cp.async.bulk global->shared and
// shared->global, no tensor map needed for 1-D ranges), the SM only computes.  A persistent CTA walks 16 KB tiles with a
// ring of kBulkStages shared-memory stages; one thread issues the copies, all 256 threads transform a landed tile in
// place.  No per-element address generation and no LSU queue pressure, which is what limits the register kernel.
constexpr int kBulkTile = 16 * 1024;          // bytes per tile (4096 floats, 16 per thread)
constexpr int kBulkStages = 4;

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ptx::smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(ptx::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ptx::smem_u32(smem_src)), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(kThreads) affine_f32_bulk(const float* __restrict__ x, float* __restrict__ y, int64_t n_tiles,
                                                            float a, float b) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[kBulkStages];
    if (threadIdx.x == 0) {
        for (int s = 0; s < kBulkStages; ++s) ptx::mbar_init(&full[s], 1);
        ptx::fence_mbar_init();
    }
    __syncthreads();
    const int64_t first = blockIdx.x, stride = gridDim.x;
    const int64_t mine = first < n_tiles ? (n_tiles - first + stride - 1) / stride : 0;      // tiles of this CTA
    auto tile_ptr = [&](int64_t i) { return (int64_t)(first + i * stride) * (kBulkTile / 4); };
    if (threadIdx.x == 0) {                                   // prologue: fill all but one stage
        for (int i = 0; i < kBulkStages - 1 && i < mine; ++i) {
            ptx::mbar_expect_tx(&full[i], kBulkTile);
            bulk_load(smem + i * kBulkTile, x + tile_ptr(i), kBulkTile, &full[i]);
        }
    }
    for (int64_t i = 0; i < mine; ++i) {
        const int s = (int)(i % kBulkStages);
        if (threadIdx.x == 0) {
            const int64_t nxt = i + kBulkStages - 1;          // refill the stage whose store was issued in iteration i-1
            if (nxt < mine) {
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");      // that store has finished reading shared memory
                const int ns = (int)(nxt % kBulkStages);
                ptx::mbar_expect_tx(&full[ns], kBulkTile);
                bulk_load(smem + ns * kBulkTile, x + tile_ptr(nxt), kBulkTile, &full[ns]);
            }
        }
        ptx::mbar_wait(&full[s], (uint32_t)((i / kBulkStages) & 1));
        f32x4* t = reinterpret_cast<f32x4*>(smem + s * kBulkTile);
#pragma unroll
        for (int k = 0; k < kBulkTile / 16 / kThreads; ++k) {
            f32x4 v = t[threadIdx.x + k * kThreads];
            t[threadIdx.x + k * kThreads] = affine_vec(v, a, b);
        }
        ptx::fence_proxy_async();                             // make the generic-proxy writes visible to the copy engine
        __syncthreads();
        if (threadIdx.x == 0) {
            bulk_store(y + tile_ptr(i), smem + s * kBulkTile, kBulkTile);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores complete before the CTA exits
}

// Unaligned buffers and the <4-element tail.
__global__ void __launch_bounds__(kThreads) affine_f32_scalar(const float* __restrict__ x,
                                                              float* __restrict__ y, int64_t n, float a,
                                                              float b) {
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride)
        y[i] = affine1(__ldg(x + i), a, b);
}

}  // namespace

int launch_affine_f32(const float* x, float* y, int64_t n, float a, float b, cudaStream_t s,
                      std::string* err) {
    if (n < 0 || (n > 0 && (!x || !y))) {
        if (err) *err = "affine_f32: null buffer or negative length";
        return FTF_INVALID_ARGUMENT;
    }
    if (n == 0) return FTF_OK;
    int dev = 0;
    FTF_CUDA(cudaGetDevice(&dev), err);
    const int sms = device_sm_count(dev);
    const int resident = 2048 / kThreads;  // CTAs per SM at full occupancy
    const uintptr_t bits = (uintptr_t)x | (uintptr_t)y;
    int64_t done = 0;
    auto grid_for = [&](int64_t items) {
        const int64_t want = (items + kThreads - 1) / kThreads;
        return (int)(want < (int64_t)sms * resident ? want : (int64_t)sms * resident);
    };
    static const bool no_bulk = getenv("FTF_K1_NO_BULK") && getenv("FTF_K1_NO_BULK")[0] == '1';
    if (!no_bulk && (bits & 15u) == 0 && n >= (int64_t)(kBulkTile / 4) * sms * 4) {
        const int64_t n_tiles = n / (kBulkTile / 4);
        const size_t smem = (size_t)kBulkStages * kBulkTile;
        static bool attr_set[64] = {false};
        if (dev < 64 && !attr_set[dev]) {
            FTF_CUDA(cudaFuncSetAttribute(affine_f32_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), err);
            attr_set[dev] = true;
        }
        const int64_t cap = (int64_t)sms * 3;                 // 3 CTAs x 64 KB per SM
        affine_f32_bulk<<<(unsigned)(n_tiles < cap ? n_tiles : cap), kThreads, smem, s>>>(x, y, n_tiles, a, b);
        count_launch();
        done = n_tiles * (kBulkTile / 4);
        x += done; y += done; n -= done;
        if (n == 0) { FTF_CUDA(cudaGetLastError(), err); return FTF_OK; }
        done = 0;
    }
    if ((bits & 31u) == 0 && n >= 8) {
        const int64_t nv = n / 8;
        affine_f32_vec<f32x8><<<grid_for(nv), kThreads, 0, s>>>((const f32x8*)x, (f32x8*)y, nv, a, b);
        count_launch();
        done = nv * 8;
    } else if ((bits & 15u) == 0 && n >= 4) {
        const int64_t nv = n / 4;
        affine_f32_vec<f32x4><<<grid_for(nv), kThreads, 0, s>>>((const f32x4*)x, (f32x4*)y, nv, a, b);
        count_launch();
        done = nv * 4;
    }
    if (done < n) {
        const int64_t rem = n - done;
        int64_t want = (rem + kThreads - 1) / kThreads;
        int grid = (int)(want < (int64_t)sms * resident ? want : (int64_t)sms * resident);
        affine_f32_scalar<<<grid, kThreads, 0, s>>>(x + done, y + done, rem, a, b);
        count_launch();
    }
    FTF_CUDA(cudaGetLastError(), err);
    return FTF_OK;
}

}  // namespace ftf
