// runner.cu -- C ABI entry points: lifecycle, ModelFunction.apply (ftf_run), micro-batched records.
//
// Replaces, per operator subtask, the TF-Java objects the reference drives
// (paths relative to /root/reference/flink-tensorflow/src/main/scala/org/apache/flink/contrib/tensorflow):
//   models/generic/GenericModel.scala:22-43, models/savedmodel/TensorFlowModel.scala:36-49  open/close
//   models/ModelFunction.scala:34-78                                                        feed/run/fetch
// Records travel  caller buffer -> pinned slot -> HBM -> kernels -> HBM -> pinned result.  Three slots, one
// CUDA stream each, so the upload of chunk c+1 overlaps the kernels of chunk c and the download of c-1.
#include "runner.h"
#include "jpeg.h"
#include "inception.h"
#include "conv_tc.h"
#include "pools.h"
#include <string.h>
#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>

namespace ftf {

static thread_local std::string g_thread_error;
void set_thread_error(const std::string& msg) { g_thread_error = msg; }
const char* thread_error() { return g_thread_error.c_str(); }

int cuda_fail(cudaError_t e, const char* what, std::string* err) {
    if (err) *err = std::string(what) + ": " + cudaGetErrorString(e);
    return FTF_INTERNAL;
}

static std::atomic<uint64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
uint64_t launches_so_far() { return g_launches.load(std::memory_order_relaxed); }

int device_sm_count(int device) {
    static std::mutex mu;
    static int cache[64];
    std::lock_guard<std::mutex> lock(mu);
    if (device < 0 || device >= 64) return 148;
    if (cache[device] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0) v = 148;
        cache[device] = v;
    }
    return cache[device];
}

tensor_map_encode_tiled_fn get_tensor_map_encode_tiled() {
    static std::once_flag once;
    static tensor_map_encode_tiled_fn fn = nullptr;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (tensor_map_encode_tiled_fn)p;
    });
    return fn;
}

int Buf::ensure(size_t bytes, std::string* err) {
    if (bytes <= cap) return FTF_OK;
    free_now();
    const size_t want = std::max<size_t>(bytes, 256);
    cudaError_t e = host ? cudaHostAlloc(&p, want, cudaHostAllocDefault) : cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        if (err) *err = std::string(host ? "pinned host" : "device") + " allocation of " + std::to_string(want) +
                        " bytes failed: " + cudaGetErrorString(e);
        cudaGetLastError();
        return FTF_RESOURCE_EXHAUSTED;
    }
    cap = want;
    return FTF_OK;
}
void Buf::free_now() {
    if (p) { if (host) cudaFreeHost(p); else cudaFree(p); }
    p = nullptr; cap = 0;
}

}  // namespace ftf

using namespace ftf;

namespace {

constexpr size_t kChunkBytes = 32u << 20;   // staging granularity of ftf_run
constexpr int kInceptionSide = 299;

int fail(ftf_runner* r, int code, const std::string& msg) {
    if (r) r->err = msg; else set_thread_error(msg);
    return code;
}

bool is_label_kind(Kind k) { return k == Kind::kLabel || k == Kind::kNormalizeLabel; }

int64_t numel(const ftf_tensor& t) {
    int64_t n = 1;
    for (int i = 0; i < t.rank; ++i) n *= t.dims[i];
    return n;
}

// Fixes per-record geometry (input bytes, output specs).  Called with the first tensor that determines it.
int fix_geometry(ftf_runner* r, const ftf_tensor& t, bool single_record) {
    std::vector<OutSpec> outs;
    size_t in_bytes = 0;
    switch (r->kind) {
    case Kind::kAffine: {
        in_bytes = 4;
        OutSpec o; o.key = r->affine.output_key; o.dtype = FTF_FLOAT; o.rec_dims = {1}; o.rec_bytes = 4;
        outs.push_back(o);
        break;
    }
    case Kind::kNormalize:
    case Kind::kNormalizeLabel: {
        if (t.dtype == FTF_STRING) {
            // JPEG files, the reference's ImageFileTensor (EX/inception/package.scala:12-29): sizes come from each file
            if (t.rank > 1) return fail(r, FTF_INVALID_ARGUMENT, "normalize expects a scalar or vector of STRING (JPEG files)");
            if (r->kind == Kind::kNormalize) {
                OutSpec o; o.key = "outputs"; o.dtype = FTF_FLOAT;
                o.rec_dims = {r->cfg.out_h, r->cfg.out_w, 3};
                o.rec_bytes = (size_t)r->cfg.out_h * r->cfg.out_w * 3 * 4;
                outs.push_back(o);
            }
            in_bytes = 0;
            break;
        }
        const int base = single_record ? t.rank - 3 : t.rank - 3;
        if (t.dtype != FTF_UINT8 || (t.rank != 3 && t.rank != 4) || t.dims[t.rank - 1] != 3)
            return fail(r, FTF_INVALID_ARGUMENT, "normalize expects a UINT8 tensor shaped [N,H,W,3] or [H,W,3], or a STRING tensor of JPEG files");
        const int64_t h = t.dims[base], w = t.dims[base + 1];
        if (h <= 0 || w <= 0 || h > 65535 || w > 65535)
            return fail(r, FTF_INVALID_ARGUMENT, "normalize: image height/width out of range");
        if (r->in_h && (r->in_h != h || r->in_w != w) && (r->open_slot >= 0 || !r->in_flight.empty()))
            return fail(r, FTF_INVALID_ARGUMENT, "normalize: records of one micro-batch must share one image size");
        r->in_h = (int)h; r->in_w = (int)w;
        in_bytes = (size_t)h * w * 3;
        if (r->kind == Kind::kNormalize) {
            OutSpec o; o.key = "outputs"; o.dtype = FTF_FLOAT;
            o.rec_dims = {r->cfg.out_h, r->cfg.out_w, 3};
            o.rec_bytes = (size_t)r->cfg.out_h * r->cfg.out_w * 3 * 4;
            outs.push_back(o);
        }
        break;
    }
    case Kind::kLabel: {
        const int64_t side = kInceptionSide;
        const bool ok4 = t.rank == 4 && t.dims[1] == side && t.dims[2] == side && t.dims[3] == 3;
        const bool ok3 = t.rank == 3 && t.dims[0] == side && t.dims[1] == side && t.dims[2] == 3;
        if (t.dtype != FTF_FLOAT || !(ok4 || ok3))
            return fail(r, FTF_INVALID_ARGUMENT, "label expects a FLOAT tensor shaped [N,299,299,3]");
        in_bytes = (size_t)side * side * 3 * 4;
        break;
    }
    }
    if (is_label_kind(r->kind)) {
        const int c = r->cfg.num_classes, k = r->cfg.top_k;
        OutSpec p; p.key = "outputs"; p.dtype = FTF_FLOAT; p.rec_dims = {c}; p.rec_bytes = (size_t)c * 4;
        OutSpec l = p; l.key = "logits";
        OutSpec i; i.key = "top_k_indices"; i.dtype = FTF_INT32; i.rec_dims = {k}; i.rec_bytes = (size_t)k * 4;
        OutSpec s; s.key = "top_k_scores"; s.dtype = FTF_FLOAT; s.rec_dims = {k}; s.rec_bytes = (size_t)k * 4;
        outs = {p, l, i, s};
    }
    r->in_rec_bytes = in_bytes;
    r->outs = outs;
    for (auto& s : r->slots) { s.h_out.resize(outs.size()); s.d_out.resize(outs.size()); for (auto& b : s.h_out) b.host = true; }
    return FTF_OK;
}

int wait_slot(ftf_runner* r, Slot& s) {
    if (!s.busy) return FTF_OK;
    cudaError_t e = cudaEventSynchronize(s.done);
    if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "waiting for an in-flight batch", &r->err); }
    s.busy = false;
    return FTF_OK;
}

constexpr int64_t kZeroCopyBytes = 16 << 10;      // affine batches up to this size run on the pinned host buffers themselves

// Enqueues one batch on a slot: upload n records from h_src, run the model, download into h_dst[j].
// h_src must be pinned (the slot's own staging or caller memory from ftf_host_alloc).
// sel: the outputs to download (indices into r->outs); s.downloaded remembers them for ftf_poll.
int enqueue(ftf_runner* r, Slot& s, int64_t n, const void* h_src, void* const* h_dst, const std::vector<int>& sel, bool uploaded = false) {
    std::string* err = &r->err;
    int rc;
    const bool jpg = s.jpeg_n > 0;         // d_in already holds this batch's images, decoded from files by decode_files
    if (jpg) uploaded = true;
    else if ((rc = s.d_in.ensure((size_t)n * r->in_rec_bytes, err))) return rc;
    // resize + normalise of the batch in d_in: one launch, or one per image when decoded files differ in size
    auto normalise = [&](int oh, int ow, void* outp, int out_dtype) -> int {
        if (!jpg) return launch_resize_norm((const uint8_t*)s.d_in.p, n, r->in_h, r->in_w, oh, ow, r->cfg.mean, r->cfg.scale, outp, out_dtype, s.stream, err);
        if (s.jpeg_uniform) return launch_resize_norm((const uint8_t*)s.d_in.p, n, s.jh[0], s.jw[0], oh, ow, r->cfg.mean, r->cfg.scale, outp, out_dtype, s.stream, err);
        const size_t rec = (size_t)oh * ow * 3 * (out_dtype == FTF_FLOAT ? 4 : 2);
        for (int64_t i = 0; i < n; ++i) {
            const int rc1 = launch_resize_norm((const uint8_t*)s.d_in.p + s.joff[i], 1, s.jh[i], s.jw[i], oh, ow, r->cfg.mean, r->cfg.scale,
                                               (uint8_t*)outp + (size_t)i * rec, out_dtype, s.stream, err);
            if (rc1) return rc1;
        }
        return FTF_OK;
    };
    // A handful of values (the reference's per-record call: one Example, one float back): two copies and a kernel are three
    // trips to the device for 8 bytes.  Pinned host memory is mapped into the device's address space, so K1 reads the input
    // where it lies and writes the result where ftf_run hands it out: one launch, one wait (34 -> 20 us per call).
    if (r->kind == Kind::kAffine && !uploaded && n * 4 <= kZeroCopyBytes && sel.size() == 1 && sel[0] == 0) {
        const uint64_t before0 = launches_so_far();
        if ((rc = launch_affine_f32((const float*)h_src, (float*)h_dst[0], n, r->affine.a, r->affine.b, s.stream, err))) return rc;
        r->stats.kernel_launches += launches_so_far() - before0;
        r->stats.h2d_bytes += (uint64_t)n * 4;
        r->stats.d2h_bytes += (uint64_t)n * 4;
        s.downloaded = 1u;
        FTF_CUDA(cudaEventRecord(s.done, s.stream), err);
        s.busy = true;
        r->stats.batches += 1;
        r->stats.records += (uint64_t)n;
        return FTF_OK;
    }
    for (size_t j = 0; j < r->outs.size(); ++j)
        if ((rc = s.d_out[j].ensure((size_t)n * r->outs[j].rec_bytes, err))) return rc;
    if (!uploaded) {     // micro-batched records were uploaded one by one as they were submitted
        FTF_CUDA(cudaMemcpyAsync(s.d_in.p, h_src, (size_t)n * r->in_rec_bytes, cudaMemcpyHostToDevice, s.stream), err);
        r->stats.h2d_bytes += (uint64_t)n * r->in_rec_bytes;
    }
    const uint64_t before = launches_so_far();
    switch (r->kind) {
    case Kind::kAffine:
        rc = launch_affine_f32((const float*)s.d_in.p, (float*)s.d_out[0].p, n, r->affine.a, r->affine.b, s.stream, err);
        break;
    case Kind::kNormalize:
        rc = normalise(r->cfg.out_h, r->cfg.out_w, s.d_out[0].p, FTF_FLOAT);
        break;
    case Kind::kLabel:
    case Kind::kNormalizeLabel: {
        // the activation arena is shared by the runner's batches: the net runs on one stream, fenced by events
        const size_t img_elems = (size_t)kInceptionSide * kInceptionSide * 3;
        const void* img = s.d_in.p;        // fp32 records: the first convolution reads them where they were uploaded
        int img_dtype = FTF_FLOAT;
        if (r->kind == Kind::kNormalizeLabel) {
            // normalised image batch (bf16, or fp32 for a compute_dtype FLOAT network), per slot so that the next batch's
            // normalisation can overlap this one's network
            img_dtype = r->cfg.compute_dtype == FTF_FLOAT ? FTF_FLOAT : FTF_BFLOAT16;
            if ((rc = s.d_img.ensure((size_t)n * img_elems * (img_dtype == FTF_FLOAT ? 4 : 2), err))) return rc;
            if ((rc = normalise(kInceptionSide, kInceptionSide, s.d_img.p, img_dtype))) break;
            img = s.d_img.p;
        }
        FTF_CUDA(cudaEventRecord(s.done, s.stream), err);
        FTF_CUDA(cudaStreamWaitEvent(r->net_stream, s.done, 0), err);
        rc = r->net->forward(img, img_dtype, n, (float*)s.d_out[1].p, r->net_stream, err);
        if (rc) break;
        rc = launch_softmax_topk((const float*)s.d_out[1].p, n, r->cfg.num_classes, r->cfg.num_classes, r->cfg.top_k, 1,
                                 (float*)s.d_out[0].p, (int32_t*)s.d_out[2].p, (float*)s.d_out[3].p, r->net_stream, err);
        if (rc) break;
        FTF_CUDA(cudaEventRecord(r->net_done, r->net_stream), err);
        FTF_CUDA(cudaStreamWaitEvent(s.stream, r->net_done, 0), err);
        break;
    }
    }
    if (rc) return rc;
    r->stats.kernel_launches += launches_so_far() - before;
    s.downloaded = 0;
    for (int j : sel) {
        const size_t bytes = (size_t)n * r->outs[j].rec_bytes;
        FTF_CUDA(cudaMemcpyAsync(h_dst[j], s.d_out[j].p, bytes, cudaMemcpyDeviceToHost, s.stream), err);
        r->stats.d2h_bytes += bytes;
        s.downloaded |= 1u << j;
    }
    FTF_CUDA(cudaEventRecord(s.done, s.stream), err);
    s.busy = true;
    r->stats.batches += 1;
    r->stats.records += (uint64_t)n;
    return FTF_OK;
}

// JPEG records of a slot (J1-J3, csrc/jpeg.cu): files_begin opens the slot's batch, files_add parses one file and stages
// its entropy-coded bytes in pinned memory (the caller's buffer is free again afterwards), files_launch uploads the batch and
// decodes it into s.d_in on the slot's stream.
int files_begin(ftf_runner* r, Slot& s) {
    int rc;
    s.jpeg_n = 0;
    if (!s.jpeg && (rc = ftf_jpeg_batch_create(r->device, &s.jpeg))) return fail(r, rc, thread_error());
    if ((rc = jpeg::batch_begin(s.jpeg))) return fail(r, rc, ftf_jpeg_batch_error(s.jpeg));
    s.joff.clear(); s.jh.clear(); s.jw.clear();
    return FTF_OK;
}
// stable: the caller leaves the file's bytes alone until the batch has been uploaded (a blocking call's inputs; a streamed record
// in pinned memory, which stays untouched until its result has been polled): its staging may run on the batch's worker thread
int files_add(ftf_runner* r, Slot& s, const uint8_t* file, int64_t len, bool stable) {
    int32_t h = 0, w = 0;
    int64_t off = 0;
    const int rc = jpeg::batch_add(s.jpeg, file, len, &h, &w, &off, stable);
    if (rc) return fail(r, rc, ftf_jpeg_batch_error(s.jpeg));
    s.joff.push_back(off); s.jh.push_back(h); s.jw.push_back(w);
    return FTF_OK;
}
int files_launch(ftf_runner* r, Slot& s) {
    int rc;
    const int n = (int)s.jh.size();
    if ((rc = jpeg::batch_commit(s.jpeg, s.stream))) return fail(r, rc, ftf_jpeg_batch_error(s.jpeg));
    const int64_t bytes = ftf_jpeg_batch_rgb_bytes(s.jpeg);
    if ((rc = s.d_in.ensure((size_t)bytes, &r->err))) return rc;
    const uint64_t before = launches_so_far();
    if ((rc = ftf_jpeg_batch_decode_dev(s.jpeg, (uint8_t*)s.d_in.p, bytes, s.stream))) return fail(r, rc, ftf_jpeg_batch_error(s.jpeg));
    r->stats.kernel_launches += launches_so_far() - before;
    r->stats.h2d_bytes += (uint64_t)ftf_jpeg_batch_scan_bytes(s.jpeg);
    s.jpeg_uniform = true;
    for (int i = 0; i < n; ++i)
        if (s.jh[i] != s.jh[0] || s.jw[i] != s.jw[0] || s.joff[i] != (int64_t)i * s.jh[0] * s.jw[0] * 3) s.jpeg_uniform = false;
    s.jpeg_n = n;
    return FTF_OK;
}
int decode_files(ftf_runner* r, Slot& s, int n, const uint8_t* const* files, const int64_t* lens) {
    int rc = files_begin(r, s);
    for (int i = 0; i < n && !rc; ++i) rc = files_add(r, s, files[i], lens[i], /*stable=*/true);
    if (!rc) rc = files_launch(r, s);
    if (rc) jpeg::batch_wait_staged(s.jpeg);      // the inputs are the caller's again when this returns
    return rc;
}

// A finished slot whose input was JPEG files: did the device find a corrupt entropy-coded segment?
int files_status(ftf_runner* r, Slot& s, std::vector<int32_t>* per_record = nullptr) {
    if (s.jpeg_n <= 0 || !s.jpeg) return FTF_OK;
    if (per_record) per_record->assign((size_t)s.jpeg_n, FTF_OK);
    const int rc = ftf_jpeg_batch_status(s.jpeg, per_record ? per_record->data() : nullptr, per_record ? (int32_t)per_record->size() : 0);
    s.jpeg_n = 0;
    if (rc) return fail(r, rc, ftf_jpeg_batch_error(s.jpeg));
    return FTF_OK;
}

// The allocations of ftf_host_alloc: two records may share one upload only if both lie inside ONE of them (a copy whose source
// spans two separate pinned allocations is not a pinned copy any more).
std::mutex g_host_mu;
std::map<uintptr_t, size_t> g_host_ranges;
bool in_one_host_alloc(const void* first, const void* last_end) {
    std::lock_guard<std::mutex> lock(g_host_mu);
    auto it = g_host_ranges.upper_bound((uintptr_t)first);
    if (it == g_host_ranges.begin()) return false;
    --it;
    return (uintptr_t)first >= it->first && (uintptr_t)last_end <= it->first + it->second;
}

// Record uploads are merged: a 1 MB copy moves at 44 GB/s on this link, a 4 MB copy at 51, 16 MB at 53 (tools/probe_h2d.py;
// the per-copy turnaround of the copy engine, which a second stream does not hide), and the records of a stream usually lie
// back to back in the pinned ring their source fills -- or at a constant distance (TensorValue wire images: a header in front
// of every payload), which one 2-D copy covers.  A pending run goes to the copy engine when it reaches kUploadRun, when the
// next record does not continue it, when the slot's previous upload has finished (nothing is gained by waiting then), and when
// the batch closes.  Measured on Inception-v3 labelling with fp32 records
// (1.07 MB each): 41.0 k records/s with one copy per record, 47.9 k with runs of 4 MB, 49.3 k with 8 MB, 50.1 k with 16 MB
// (the network itself: 51.9 k), p50 latency 4.5 -> 3.7 ms.
constexpr size_t kUploadRun = 16u << 20;
size_t upload_run() {       // FTF_UPLOAD_RUN=<bytes>: measurement switch (0: every record its own copy)
    static const size_t v = getenv("FTF_UPLOAD_RUN") ? (size_t)atoll(getenv("FTF_UPLOAD_RUN")) : kUploadRun;
    return v;
}
int flush_upload(ftf_runner* r, Slot& s) {
    if (!s.pend_bytes) return FTF_OK;
    cudaError_t e;
    if (s.pend_n <= 1 || s.pend_stride == s.pend_rec)
        e = cudaMemcpyAsync((uint8_t*)s.d_in.p + s.pend_off, s.pend_src, s.pend_bytes, cudaMemcpyHostToDevice, s.stream);
    else
        e = cudaMemcpy2DAsync((uint8_t*)s.d_in.p + s.pend_off, s.pend_rec, s.pend_src, s.pend_stride, s.pend_rec, (size_t)s.pend_n, cudaMemcpyHostToDevice, s.stream);
    s.pend_bytes = 0; s.pend_src = nullptr; s.pend_n = 0; s.pend_stride = 0;
    if (e == cudaSuccess) e = cudaEventRecord(s.up_done, s.stream);
    s.up_inflight = true;
    if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "record upload", &r->err); }
    return FTF_OK;
}

bool pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

int set_device(ftf_runner* r) {
    if (r->dead) return fail(r, FTF_INTERNAL, "runner is dead after a CUDA error: " + r->err);
    cudaError_t e = cudaSetDevice(r->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", &r->err);
    return FTF_OK;
}

void destroy_runner(ftf_runner* r) {
    cudaSetDevice(r->device);
    // wait for this runner's own streams, not for the device: other runners on the GPU may be recording a CUDA graph, and a
    // device-wide synchronisation from this thread would be refused and would invalidate their capture
    for (auto& s : r->slots) if (s.stream) cudaStreamSynchronize(s.stream);
    if (r->net_stream) cudaStreamSynchronize(r->net_stream);
    r->net.reset();
    for (auto& s : r->slots) {
        s.h_in.free_now(); s.d_in.free_now(); s.d_img.free_now();
        if (s.jpeg) { ftf_jpeg_batch_destroy(s.jpeg); s.jpeg = nullptr; }
        for (auto& b : s.h_out) b.free_now();
        for (auto& b : s.d_out) b.free_now();
        if (s.done) cudaEventDestroy(s.done);
        if (s.up_done) cudaEventDestroy(s.up_done);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    for (auto& b : r->run_out) b.free_now();
    if (r->net_done) cudaEventDestroy(r->net_done);
    if (r->net_stream) cudaStreamDestroy(r->net_stream);
    if (r->ev_begin) cudaEventDestroy(r->ev_begin);
    if (r->ev_end) cudaEventDestroy(r->ev_end);
    cudaGetLastError();
    delete r;
}

// Launches the open micro-batch.
int launch_open(ftf_runner* r) {
    if (r->open_slot < 0) return FTF_OK;
    Slot& s = r->slots[r->open_slot];
    if (s.n == 0) { r->open_slot = -1; return FTF_OK; }
    int rc;
    if ((rc = flush_upload(r, s))) return rc;
    for (size_t j = 0; j < r->outs.size(); ++j)
        if ((rc = s.h_out[j].ensure((size_t)s.n * r->outs[j].rec_bytes, &r->err))) return rc;
    std::vector<void*> dst(r->outs.size());
    for (size_t j = 0; j < dst.size(); ++j) dst[j] = s.h_out[j].p;
    if (s.files_n > 0) {
        if ((rc = files_launch(r, s))) return rc;
        s.files_n = 0;
    } else s.jpeg_n = 0;
    rc = enqueue(r, s, s.n, s.h_in.p, dst.data(), r->poll_sel, /*uploaded=*/true);
    if (rc) { if (rc == FTF_INTERNAL) r->dead = true; return rc; }
    r->in_flight.push_back(r->open_slot);
    r->open_slot = -1;
    return FTF_OK;
}

}  // namespace

// ================================================================================================
extern "C" void ftf_default_config(ftf_config* c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->device = 0; c->max_batch = 64; c->max_delay_us = 2000;
    c->out_h = 224; c->out_w = 224; c->mean = 117.0f; c->scale = 1.0f;   // ImageNormalization.scala:51-54
    c->compute_dtype = FTF_BFLOAT16; c->top_k = 3;                       // InceptionModel.scala:81
    c->num_classes = 1000; c->seed = 2024;
}

extern "C" uint64_t ftf_launch_count(void) { return launches_so_far(); }

extern "C" const char* ftf_version(void) { return "flink-tensorflow_b200 0.1 (sm_100a)"; }

static int create_runner(const char* model, const char* signature_key, const char* method_name,
                         const ftf_config* cfg_in, ftf_runner** out) {
    if (out) *out = nullptr;
    if (!model || !out) return fail(nullptr, FTF_INVALID_ARGUMENT, "a model and an output handle must be provided");
    ftf_config cfg;
    if (cfg_in) cfg = *cfg_in; else ftf_default_config(&cfg);
    if (cfg.max_batch <= 0 || cfg.top_k < 0 || cfg.top_k > 32 || cfg.out_h <= 0 || cfg.out_w <= 0 || cfg.num_classes <= 0)
        return fail(nullptr, FTF_INVALID_ARGUMENT, "bad ftf_config (max_batch > 0, 0 <= top_k <= 32, out_h/out_w/num_classes > 0)");

    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        return fail(nullptr, FTF_FAILED_PRECONDITION,
                    std::string("no CUDA device is visible; this runner has no CPU path (") + cudaGetErrorString(e) + ")");
    }
    if (cfg.device < 0 || cfg.device >= n_dev)
        return fail(nullptr, FTF_INVALID_ARGUMENT, "device ordinal " + std::to_string(cfg.device) + " out of range");

    // from here on the runner is torn down by destroy_runner whichever way this function is left (a return or an exception)
    if ((e = cudaSetDevice(cfg.device)) != cudaSuccess) return fail(nullptr, FTF_INTERNAL, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    std::unique_ptr<ftf_runner, void (*)(ftf_runner*)> r(new ftf_runner(), destroy_runner);
    r->cfg = cfg;
    r->device = cfg.device;
    const std::string m(model);
    std::string err, weights_dir;
    int rc = FTF_OK;
    if (m == "builtin:image_normalization") {
        r->kind = Kind::kNormalize; r->in_key = "inputs";
        r->signature.method_name = "inception/normalize";                    // ImageNormalization.scala:21
    } else if (m == "builtin:inception_v3") {
        r->kind = Kind::kLabel; r->in_key = "inputs";
        r->signature.method_name = "inception/label";                        // InceptionModel.scala:27
    } else if (m == "builtin:normalize_label") {
        r->kind = Kind::kNormalizeLabel; r->in_key = "inputs";
        r->signature.method_name = "inception/label";
    } else if (m.rfind("inception_v3:", 0) == 0 || m.rfind("normalize_label:", 0) == 0) {
        // the label model with TRAINED weights from a directory (where InceptionModel.scala:51-55 loads its frozen graph
        // file): "<kind>:<dir>", dir/variables/* a TensorBundle of BatchNorm-folded fp32 tensors (csrc/inception.h)
        r->kind = m[0] == 'i' ? Kind::kLabel : Kind::kNormalizeLabel; r->in_key = "inputs";
        r->signature.method_name = "inception/label";
        weights_dir = m.substr(m.find(':') + 1);
    } else {
        r->kind = Kind::kAffine;
        if (!signature_key) return fail(nullptr, FTF_INVALID_ARGUMENT, "a signatureDef must be provided");
        rc = r->saved_model.load(m, {"serve"}, &err);                        // HalfPlusTwo.scala: load(path, "serve")
        if (rc) return fail(nullptr, rc, err);
        auto it = r->saved_model.signatures.find(signature_key);
        if (it == r->saved_model.signatures.end())
            return fail(nullptr, FTF_NOT_FOUND, std::string("signature '") + signature_key + "' is not in the SavedModel");
        r->signature = it->second;
        rc = plan_affine(r->saved_model, r->signature, &r->affine, &err);
        if (rc) return fail(nullptr, rc, err);
        r->in_key = r->affine.input_key;
    }
    if (is_label_kind(r->kind)) {
        // synthetic signature, as the reference builds it (InceptionModel.scala:58-62)
        r->signature.inputs = {{"inputs", TensorInfoLite{"input", FTF_FLOAT, {}}}};
        r->signature.outputs = {{"outputs", TensorInfoLite{"output", FTF_FLOAT, {}}}};
    } else if (r->kind == Kind::kNormalize) {
        r->signature.inputs = {{"inputs", TensorInfoLite{"input", FTF_UINT8, {}}}};
        r->signature.outputs = {{"outputs", TensorInfoLite{"Div", FTF_FLOAT, {}}}};
    }
    if (method_name && r->signature.method_name != method_name)             // ModelFunction.scala:37
        return fail(nullptr, FTF_INVALID_ARGUMENT, "requirement failed: method name '" + std::string(method_name) +
                    "' does not match the signature's '" + r->signature.method_name + "'");

    ftf_runner* raw = r.get();
    auto bail = [&](int code, const std::string& msg) { return fail(nullptr, code, msg); };
    for (auto& s : raw->slots) {
        s.h_in.host = true;
        if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&s.up_done, cudaEventDisableTiming)) != cudaSuccess)
            return bail(FTF_INTERNAL, std::string("stream/event creation: ") + cudaGetErrorString(e));
    }
    if ((e = cudaEventCreate(&raw->ev_begin)) != cudaSuccess || (e = cudaEventCreate(&raw->ev_end)) != cudaSuccess)
        return bail(FTF_INTERNAL, std::string("event creation: ") + cudaGetErrorString(e));
    if (is_label_kind(raw->kind)) {
        if ((e = cudaStreamCreateWithFlags(&raw->net_stream, cudaStreamNonBlocking)) != cudaSuccess ||
            (e = cudaEventCreateWithFlags(&raw->net_done, cudaEventDisableTiming)) != cudaSuccess)
            return bail(FTF_INTERNAL, std::string("stream/event creation: ") + cudaGetErrorString(e));
        rc = InceptionV3::create(cfg.device, cfg.num_classes, cfg.seed, cfg.max_batch, cfg.compute_dtype, &raw->net, &err,
                                 weights_dir.empty() ? nullptr : weights_dir.c_str());
        if (rc) return bail(rc, err);
    }
    if (raw->kind == Kind::kAffine) {
        ftf_tensor dummy{};
        fix_geometry(raw, dummy, false);
    }
    *out = r.release();
    return FTF_OK;
}

// No exception may cross the C boundary: a malformed model file that drives a container past its limits (std::length_error,
// std::bad_alloc) becomes a status like every other failure, and the half-built runner is released.
extern "C" int ftf_create(const char* model, const char* signature_key, const char* method_name,
                          const ftf_config* cfg_in, ftf_runner** out) {
    try {
        return create_runner(model, signature_key, method_name, cfg_in, out);
    } catch (const std::bad_alloc&) {
        if (out) *out = nullptr;
        return fail(nullptr, FTF_RESOURCE_EXHAUSTED, "ftf_create: out of host memory");
    } catch (const std::exception& ex) {
        if (out) *out = nullptr;
        return fail(nullptr, FTF_INVALID_ARGUMENT, std::string("ftf_create: malformed model (") + ex.what() + ")");
    } catch (...) {
        if (out) *out = nullptr;
        return fail(nullptr, FTF_INTERNAL, "ftf_create: unexpected failure");
    }
}

extern "C" int ftf_destroy(ftf_runner* r) {
    if (!r) return FTF_OK;
    destroy_runner(r);
    return FTF_OK;
}

extern "C" const char* ftf_last_error(const ftf_runner* r) { return r ? r->err.c_str() : thread_error(); }

extern "C" void* ftf_runner_stream(ftf_runner* r) {
    if (!r) return nullptr;
    return is_label_kind(r->kind) ? (void*)r->net_stream : (void*)r->slots[0].stream;
}

extern "C" const char* ftf_signature_info(const ftf_runner* r, int32_t kind, int32_t i) {
    if (!r || i < 0) return nullptr;
    const auto& s = r->signature;
    switch (kind) {
    case 0: return i == 0 ? s.method_name.c_str() : nullptr;
    case 1: return i < (int)s.inputs.size() ? s.inputs[i].first.c_str() : nullptr;
    case 2: return i < (int)s.inputs.size() ? s.inputs[i].second.name.c_str() : nullptr;
    case 3: return i < (int)s.outputs.size() ? s.outputs[i].first.c_str() : nullptr;
    case 4: return i < (int)s.outputs.size() ? s.outputs[i].second.name.c_str() : nullptr;
    default: return nullptr;
    }
}

extern "C" void* ftf_host_alloc(int64_t nbytes) {
    void* p = nullptr;
    if (nbytes <= 0 || cudaHostAlloc(&p, (size_t)nbytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    { std::lock_guard<std::mutex> lock(g_host_mu); g_host_ranges[(uintptr_t)p] = (size_t)nbytes; }
    return p;
}
extern "C" void ftf_host_free(void* p) {
    if (!p) return;
    { std::lock_guard<std::mutex> lock(g_host_mu); g_host_ranges.erase((uintptr_t)p); }
    cudaFreeHost(p);
}

// ------------------------------------------------------------------------------------------------
extern "C" int ftf_run(ftf_runner* r, int32_t n_in, const char* const* in_keys, const ftf_tensor* in,
                       int32_t n_out, const char* const* out_keys, ftf_tensor* out) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    int rc = set_device(r);
    if (rc) return rc;
    if (n_in < 0 || n_out < 0 || (n_in && (!in_keys || !in)) || (n_out && (!out_keys || !out)))
        return fail(r, FTF_INVALID_ARGUMENT, "Input must be provided");
    if (r->open_slot >= 0 || !r->in_flight.empty())
        return fail(r, FTF_FAILED_PRECONDITION, "ftf_run while micro-batched records are in flight; call ftf_flush and drain ftf_poll first");
    const ftf_tensor* x = nullptr;
    for (int i = 0; i < n_in; ++i)
        if (in_keys[i] && r->in_key == in_keys[i]) x = &in[i];
    if (!x)                                                                     // ModelFunction.scala:52-53
        return fail(r, FTF_INVALID_ARGUMENT, "An input tensor named " + r->in_key + " must be provided for this computation.");
    if (x->rank < 0 || x->rank > 25 || (x->rank && !x->dims) || (x->nbytes && !x->data))
        return fail(r, FTF_INVALID_ARGUMENT, "malformed input tensor");

    // ---- how many records, and where the per-record payload starts ------------------------------
    int64_t n = 0;
    bool files_in = false;                 // the input is a STRING tensor of JPEG files
    std::vector<int64_t> lead_dims;        // output dims before the per-record dims
    if (r->kind == Kind::kAffine) {
        if (r->affine.input_is_examples) {
            if (x->dtype != FTF_STRING || x->rank > 1)
                return fail(r, FTF_INVALID_ARGUMENT, "Expected serialized to be a scalar or vector of STRING, got dtype " +
                            std::to_string(x->dtype) + " rank " + std::to_string(x->rank));
            n = x->rank == 0 ? 1 : x->dims[0];
            lead_dims = {n};
        } else {
            if (x->dtype != FTF_FLOAT) return fail(r, FTF_INVALID_ARGUMENT, "Expected a FLOAT tensor for input '" + r->in_key + "'");
            n = numel(*x);
            if (x->nbytes != n * 4) return fail(r, FTF_INVALID_ARGUMENT, "input payload size does not match its shape");
        }
    } else if (x->dtype == FTF_STRING && r->kind != Kind::kLabel) {
        // JPEG files: ImageNormalization.scala:56-57 feeds the file bytes and the graph starts with DecodeJpeg
        if ((rc = fix_geometry(r, *x, false))) return rc;
        n = x->rank == 0 ? 1 : x->dims[0];
        lead_dims = {n};
        files_in = true;
    } else {
        if ((rc = fix_geometry(r, *x, false))) return rc;
        const int rec_rank = 3;
        n = x->rank == rec_rank ? 1 : x->dims[0];
        if (x->nbytes != (int64_t)(n * r->in_rec_bytes)) return fail(r, FTF_INVALID_ARGUMENT, "input payload size does not match its shape");
        lead_dims = {n};
    }
    if (n < 0) return fail(r, FTF_INVALID_ARGUMENT, "negative batch dimension");

    // ---- outputs requested ----------------------------------------------------------------------
    std::vector<int> want(n_out, -1);
    for (int i = 0; i < n_out; ++i) {
        for (size_t j = 0; j < r->outs.size(); ++j)
            if (out_keys[i] && r->outs[j].key == out_keys[i]) want[i] = (int)j;
        if (want[i] < 0) return fail(r, FTF_NOT_FOUND, std::string("no output named '") + (out_keys[i] ? out_keys[i] : "(null)") + "' in this signature");
    }
    std::vector<int> want_unique;                     // only the fetched outputs travel back (Session.runner().fetch)
    for (int j : want) if (std::find(want_unique.begin(), want_unique.end(), j) == want_unique.end()) want_unique.push_back(j);
    r->run_out.resize(r->outs.size());
    r->run_dims.resize(r->outs.size());
    for (size_t j = 0; j < r->outs.size(); ++j) {
        r->run_out[j].host = true;
        if ((rc = r->run_out[j].ensure((size_t)n * r->outs[j].rec_bytes, &r->err))) return rc;
        if (r->kind == Kind::kAffine && !r->affine.input_is_examples)
            r->run_dims[j].assign(x->dims, x->dims + x->rank);                  // elementwise: same shape out
        else {
            r->run_dims[j] = lead_dims;
            r->run_dims[j].insert(r->run_dims[j].end(), r->outs[j].rec_dims.begin(), r->outs[j].rec_dims.end());
        }
    }

    // ---- chunked pipeline -----------------------------------------------------------------------
    int64_t chunk = std::max<int64_t>(1, (int64_t)(kChunkBytes / std::max<size_t>(1, r->in_rec_bytes)));
    if (is_label_kind(r->kind)) {
        chunk = r->net->max_batch();   // whole micro-batches: tile counts of the convolutions scale with n
        // a call that fits one micro-batch is still split in two halves from 64 images on: the upload of the second half then
        // overlaps the kernels of the first (measured +7 % on a blocking 64-image call; smaller chunks lose more in the
        // convolutions' tile efficiency than the overlap gains)
        if (n <= chunk && n >= 64) chunk = (n / 2 + 7) / 8 * 8;
        if (const char* e = getenv("FTF_RUN_CHUNK")) chunk = std::max<int64_t>(1, std::min<int64_t>(r->net->max_batch(), atoll(e)));   // measurement switch
    }
    if (files_in && !is_label_kind(r->kind)) chunk = 128;
    const bool parse = r->kind == Kind::kAffine && r->affine.input_is_examples;
    const bool direct = !parse && !files_in && n > 0 && pinned(x->data);
    int files_rc = FTF_OK;
    cudaError_t e;
    if ((e = cudaEventRecord(r->ev_begin, r->slots[r->next_slot].stream)) != cudaSuccess) return cuda_fail(e, "cudaEventRecord", &r->err);
    for (int64_t c0 = 0; c0 < n; c0 += chunk) {
        const int64_t cn = std::min(chunk, n - c0);
        Slot& s = r->slots[r->next_slot];
        r->next_slot = (r->next_slot + 1) % kSlots;
        if ((rc = wait_slot(r, s))) return rc;
        if (files_rc == FTF_OK) files_rc = files_status(r, s); else s.jpeg_n = 0;
        const void* src;
        if (files_in) {
            std::vector<const uint8_t*> fp((size_t)cn);
            std::vector<int64_t> fl((size_t)cn);
            for (int64_t i = 0; i < cn; ++i)
                if (string_tensor_item((const uint8_t*)x->data, x->nbytes, n, c0 + i, &fp[i], &fl[i]) != FTF_OK)
                    return fail(r, FTF_INVALID_ARGUMENT, "malformed STRING tensor at element " + std::to_string(c0 + i));
            if ((rc = decode_files(r, s, (int)cn, fp.data(), fl.data()))) return rc;
            src = nullptr;
        } else if (direct) { s.jpeg_n = 0; src = (const uint8_t*)x->data + (size_t)c0 * r->in_rec_bytes; }
        else {
            s.jpeg_n = 0;
            if ((rc = s.h_in.ensure((size_t)cn * r->in_rec_bytes, &r->err))) return rc;
            if (parse) {
                for (int64_t i = 0; i < cn; ++i) {
                    const uint8_t* item; int64_t len;
                    if (string_tensor_item((const uint8_t*)x->data, x->nbytes, n, c0 + i, &item, &len) != FTF_OK)
                        return fail(r, FTF_INVALID_ARGUMENT, "malformed STRING tensor at element " + std::to_string(c0 + i));
                    if (parse_example_f32(item, (size_t)len, r->affine.feature_key, r->affine.has_default,
                                          r->affine.default_value, (float*)s.h_in.p + i) != FTF_OK)
                        return fail(r, FTF_INVALID_ARGUMENT, "Name: <unknown>, Feature: " + r->affine.feature_key +
                                    " is required but could not be found (or the Example is malformed), element " + std::to_string(c0 + i));
                }
            } else memcpy(s.h_in.p, (const uint8_t*)x->data + (size_t)c0 * r->in_rec_bytes, (size_t)cn * r->in_rec_bytes);
            src = s.h_in.p;
        }
        std::vector<void*> dst(r->outs.size());
        for (size_t j = 0; j < dst.size(); ++j) dst[j] = (uint8_t*)r->run_out[j].p + (size_t)c0 * r->outs[j].rec_bytes;
        rc = enqueue(r, s, cn, src, dst.data(), want_unique);
        if (rc) { if (rc == FTF_INTERNAL) r->dead = true; return rc; }
    }
    // join the slot streams on slot 0's stream and stamp the end
    for (int i = 1; i < kSlots; ++i)
        if (r->slots[i].busy && (e = cudaStreamWaitEvent(r->slots[0].stream, r->slots[i].done, 0)) != cudaSuccess)
            return cuda_fail(e, "cudaStreamWaitEvent", &r->err);
    if ((e = cudaEventRecord(r->ev_end, r->slots[0].stream)) != cudaSuccess) return cuda_fail(e, "cudaEventRecord", &r->err);
    if ((e = cudaEventSynchronize(r->ev_end)) != cudaSuccess) { r->dead = true; return cuda_fail(e, "model run", &r->err); }
    std::string files_msg = r->err;
    for (auto& s : r->slots) {
        s.busy = false;
        if (files_rc == FTF_OK) { files_rc = files_status(r, s); files_msg = r->err; } else s.jpeg_n = 0;
    }
    if (files_rc) return fail(r, files_rc, files_msg);
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, r->ev_begin, r->ev_end) == cudaSuccess) r->stats.last_run_device_ms = ms; else cudaGetLastError();
    r->stats.runs += 1;

    for (int i = 0; i < n_out; ++i) {
        const int j = want[i];
        out[i].dtype = r->outs[j].dtype;
        out[i].rank = (int32_t)r->run_dims[j].size();
        out[i].dims = r->run_dims[j].data();
        out[i].data = r->run_out[j].p;
        out[i].nbytes = (int64_t)((size_t)n * r->outs[j].rec_bytes);
    }
    return FTF_OK;
}

extern "C" void ftf_release(ftf_runner* r, ftf_tensor* outs, int32_t n) {
    // results live in runner-owned pinned buffers that are recycled by the next run; releasing only
    // invalidates the caller's view (== Tensor.close() on the fetched outputs, ModelFunction.scala:68-69)
    (void)r;
    for (int i = 0; outs && i < n; ++i) { outs[i].data = nullptr; outs[i].nbytes = 0; outs[i].dims = nullptr; outs[i].rank = 0; }
}

// ------------------------------------------------------------------------------------------------
extern "C" int ftf_submit(ftf_runner* r, uint64_t tag, const ftf_tensor* rec) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    int rc = set_device(r);
    if (rc) return rc;
    if (!rec || (rec->nbytes && !rec->data)) return fail(r, FTF_INVALID_ARGUMENT, "Input must be provided");
    float parsed = 0.0f;
    const void* payload = rec->data;
    size_t bytes = 0;
    const uint8_t* file = nullptr;         // set when the record is a JPEG file
    int64_t file_len = 0;
    if (r->kind == Kind::kAffine) {
        if (r->affine.input_is_examples) {
            const int64_t cnt = rec->rank == 0 ? 1 : (rec->rank == 1 ? rec->dims[0] : -1);
            if (rec->dtype != FTF_STRING || cnt != 1) return fail(r, FTF_INVALID_ARGUMENT, "a record is one serialized Example (STRING scalar or [1])");
            const uint8_t* item; int64_t len;
            if (string_tensor_item((const uint8_t*)rec->data, rec->nbytes, 1, 0, &item, &len) != FTF_OK ||
                parse_example_f32(item, (size_t)len, r->affine.feature_key, r->affine.has_default, r->affine.default_value, &parsed) != FTF_OK)
                return fail(r, FTF_INVALID_ARGUMENT, "Name: <unknown>, Feature: " + r->affine.feature_key + " is required but could not be found (or the Example is malformed)");
            payload = &parsed;
        } else if (rec->dtype != FTF_FLOAT || numel(*rec) != 1 || rec->nbytes != 4)
            return fail(r, FTF_INVALID_ARGUMENT, "a record is one FLOAT value");
        bytes = 4;
    } else if (rec->dtype == FTF_STRING && r->kind != Kind::kLabel) {
        // one JPEG file per record (ImageFileTensor).  Its header is read and its entropy-coded bytes are staged now (a
        // malformed or unsupported file is refused here and joins no batch); the device decodes when the batch closes.
        const int64_t cnt = rec->rank == 0 ? 1 : (rec->rank == 1 ? rec->dims[0] : -1);
        if (cnt != 1) return fail(r, FTF_INVALID_ARGUMENT, "a record is one JPEG file (STRING scalar or [1])");
        if (string_tensor_item((const uint8_t*)rec->data, rec->nbytes, 1, 0, &file, &file_len) != FTF_OK)
            return fail(r, FTF_INVALID_ARGUMENT, "malformed STRING tensor");
        if (r->outs.empty() || r->in_rec_bytes != 0) { if ((rc = fix_geometry(r, *rec, true))) return rc; }
        bytes = 0;
    } else {
        if (rec->rank == 4 && rec->dims[0] != 1) return fail(r, FTF_INVALID_ARGUMENT, "a record is one image");
        if ((rc = fix_geometry(r, *rec, true))) return rc;
        bytes = r->in_rec_bytes;
        if (rec->nbytes != (int64_t)bytes) return fail(r, FTF_INVALID_ARGUMENT, "record payload size does not match its shape");
    }
    if (r->open_slot < 0) {
        Slot& cand = r->slots[r->next_slot];
        if (std::find(r->in_flight.begin(), r->in_flight.end(), r->next_slot) != r->in_flight.end())
            return fail(r, FTF_RESOURCE_EXHAUSTED, "all batch slots hold undelivered results; call ftf_poll");
        if ((rc = wait_slot(r, cand))) return rc;
        r->open_slot = r->next_slot;
        r->next_slot = (r->next_slot + 1) % kSlots;
        cand.n = 0; cand.emitted = 0; cand.tags.clear(); cand.t_submit.clear(); cand.files_n = 0; cand.pend_bytes = 0; cand.pend_src = nullptr; cand.pend_n = 0; cand.pend_stride = 0; cand.up_inflight = false;
        const int cap = is_label_kind(r->kind) ? std::min(r->cfg.max_batch, r->net->max_batch()) : r->cfg.max_batch;
        // device staging of every slot on the first batch (an allocation later would stall the batches in flight); the
        // pinned host staging only if a payload ever arrives in pageable memory (see below)
        for (auto& sl : r->slots) {
            rc = sl.d_in.ensure((size_t)cap * bytes, &r->err);
            for (size_t j = 0; !rc && j < r->outs.size(); ++j) {
                rc = sl.d_out[j].ensure((size_t)cap * r->outs[j].rec_bytes, &r->err);
                if (!rc) rc = sl.h_out[j].ensure((size_t)cap * r->outs[j].rec_bytes, &r->err);
            }
            if (!rc && r->kind == Kind::kNormalizeLabel)
                rc = sl.d_img.ensure((size_t)cap * kInceptionSide * kInceptionSide * 3 * (r->cfg.compute_dtype == FTF_FLOAT ? 4 : 2), &r->err);
            if (rc) { r->open_slot = -1; return rc; }
        }
    }
    Slot& s = r->slots[r->open_slot];
    // The record starts its way to HBM now, on the slot's stream, so that the upload overlaps the batches in flight.
    // A payload in pinned memory (ftf_host_alloc) is read in place by the copy engine -- the caller keeps it untouched
    // until the record's result has been polled; any other payload is first copied into the pinned slot.
    if (s.n > 0 && (file != nullptr) != (s.files_n > 0))
        return fail(r, FTF_INVALID_ARGUMENT, "records of one micro-batch must share one format (raw RGB or JPEG files)");
    if (file) {
        if (s.n == 0 && (rc = files_begin(r, s))) return rc;
        if ((rc = files_add(r, s, file, file_len, pinned(rec->data)))) return rc;      // header trouble: the record joins no batch
        s.files_n += 1;
        s.tags.push_back(tag);
        s.t_submit.push_back(std::chrono::steady_clock::now());
        s.n += 1;
        const int capf = is_label_kind(r->kind) ? std::min(r->cfg.max_batch, r->net->max_batch()) : r->cfg.max_batch;
        if (s.n >= capf) return launch_open(r);
        return FTF_OK;
    }
    const void* src = payload;
    if (!(payload == rec->data && pinned(payload))) {
        const int cap0 = is_label_kind(r->kind) ? std::min(r->cfg.max_batch, r->net->max_batch()) : r->cfg.max_batch;
        if ((rc = s.h_in.ensure((size_t)cap0 * bytes, &r->err))) return rc;
        memcpy((uint8_t*)s.h_in.p + (size_t)s.n * bytes, payload, bytes);
        src = (uint8_t*)s.h_in.p + (size_t)s.n * bytes;
    }
    {
        const uint8_t* sp = (const uint8_t*)src;
        const size_t off = (size_t)s.n * bytes;
        const bool own = src != payload;                       // the slot's staging buffer: one allocation by construction
        // continues the pending run: same record size, next place in d_in, and the payload lies one stride behind the previous
        // one (the second record of a run sets the stride: at least a record, at most two) inside the same pinned allocation
        bool joins = s.pend_bytes && bytes == s.pend_rec && off == s.pend_off + s.pend_bytes && sp > s.pend_src;
        if (joins) {
            const size_t stride = s.pend_n == 1 ? (size_t)(sp - s.pend_src) : s.pend_stride;
            joins = stride >= bytes && stride <= 2 * bytes && sp == s.pend_src + (size_t)s.pend_n * stride && (own || in_one_host_alloc(s.pend_src, sp + bytes));
            if (joins) { s.pend_stride = stride; s.pend_bytes += bytes; s.pend_n += 1; }
        }
        if (!joins) {
            if ((rc = flush_upload(r, s))) return rc;
            s.pend_src = sp; s.pend_off = off; s.pend_bytes = bytes; s.pend_rec = bytes; s.pend_n = 1; s.pend_stride = 0;
        }
        // A run waits only while the slot's previous upload is still on its way (the link is busy anyway): a stream slower
        // than the link sees every record uploaded at once, as before; a stream that saturates the link builds long runs.
        bool go = s.pend_bytes >= upload_run();
        if (!go && s.up_inflight) {
            const cudaError_t q = cudaEventQuery(s.up_done);
            if (q == cudaSuccess) s.up_inflight = false;
            else if (q != cudaErrorNotReady) { r->dead = true; return cuda_fail(q, "record upload", &r->err); }
            else cudaGetLastError();           // "not ready" is an answer, not an error to be found by a later check
        }
        if ((go || !s.up_inflight) && (rc = flush_upload(r, s))) return rc;
        r->stats.h2d_bytes += bytes;
    }
    s.tags.push_back(tag);
    s.t_submit.push_back(std::chrono::steady_clock::now());
    s.n += 1;
    const int cap = is_label_kind(r->kind) ? std::min(r->cfg.max_batch, r->net->max_batch()) : r->cfg.max_batch;
    if (s.n >= cap) return launch_open(r);
    return FTF_OK;
}

extern "C" int ftf_flush(ftf_runner* r) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    int rc = set_device(r);
    if (rc) return rc;
    if ((rc = launch_open(r))) return rc;
    for (int idx : r->in_flight) {
        cudaError_t e = cudaEventSynchronize(r->slots[idx].done);
        if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "flush", &r->err); }
    }
    return FTF_OK;
}

extern "C" int ftf_poll(ftf_runner* r, int32_t block, uint64_t* tags, ftf_tensor* outs, int32_t cap, int32_t* n_ready) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    int rc = set_device(r);
    if (rc) return rc;
    if (!n_ready || cap < 0 || (cap && (!tags || !outs))) return fail(r, FTF_INVALID_ARGUMENT, "poll: null output arrays");
    *n_ready = 0;
    r->poll_status.clear();
    r->poll_error.clear();
    const size_t no = r->outs.size();
    // a batch also closes max_delay_us after its first record (and at once when the caller blocks on it)
    if (r->open_slot >= 0 && r->slots[r->open_slot].n > 0) {
        Slot& s = r->slots[r->open_slot];
        const auto age = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - s.t_submit[0]).count();
        if (age >= r->cfg.max_delay_us || (block && r->in_flight.empty()))
            if ((rc = launch_open(r))) return rc;
    }
    r->poll_dims.resize(no);
    for (size_t j = 0; j < no; ++j) r->poll_dims[j] = r->outs[j].rec_dims;
    bool may_block = block != 0;
    const size_t nsel = r->poll_sel.size();
    while (!r->in_flight.empty() && *n_ready < cap) {
        Slot& s = r->slots[r->in_flight.front()];
        cudaError_t e = may_block ? cudaEventSynchronize(s.done) : cudaEventQuery(s.done);
        if (e == cudaErrorNotReady) { cudaGetLastError(); break; }
        if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "poll", &r->err); }
        // a batch of JPEG records: a corrupt entropy-coded segment is only known now.  It fails that record alone -- the
        // reference fails only the offending record's map() -- so the batch's results are all handed out (the damaged
        // image decodes as far as libjpeg would have got) and ftf_poll_status() says which records are bad.
        if (s.emitted == 0) {
            s.rec_status.clear();
            if (s.jpeg_n > 0 && files_status(r, s, &s.rec_status) != FTF_OK) r->poll_error = r->err;
            // outputs selected after this batch was launched were not downloaded with it: fetch them now
            bool late = false;
            for (int j : r->poll_sel)
                if (!(s.downloaded & (1u << j))) {
                    if ((rc = s.h_out[j].ensure((size_t)s.n * r->outs[j].rec_bytes, &r->err))) return rc;
                    e = cudaMemcpyAsync(s.h_out[j].p, s.d_out[j].p, (size_t)s.n * r->outs[j].rec_bytes, cudaMemcpyDeviceToHost, s.stream);
                    if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "poll (late download)", &r->err); }
                    r->stats.d2h_bytes += (uint64_t)s.n * r->outs[j].rec_bytes;
                    s.downloaded |= 1u << j;
                    late = true;
                }
            if (late && (e = cudaStreamSynchronize(s.stream)) != cudaSuccess) { r->dead = true; return cuda_fail(e, "poll (late download)", &r->err); }
        }
        may_block = false;
        const auto now = std::chrono::steady_clock::now();
        while (s.emitted < s.n && *n_ready < cap) {
            const int64_t i = s.emitted++;
            tags[*n_ready] = s.tags[i];
            for (size_t q = 0; q < nsel; ++q) {               // record i's tensors at outs[i*nsel + q]
                const int j = r->poll_sel[q];
                ftf_tensor& t = outs[(size_t)*n_ready * nsel + q];
                t.dtype = r->outs[j].dtype;
                t.rank = (int32_t)r->poll_dims[j].size();
                t.dims = r->poll_dims[j].data();
                t.data = (uint8_t*)s.h_out[j].p + (size_t)i * r->outs[j].rec_bytes;
                t.nbytes = (int64_t)r->outs[j].rec_bytes;
            }
            r->poll_status.push_back((size_t)i < s.rec_status.size() ? s.rec_status[(size_t)i] : FTF_OK);
            r->latencies_us.push_back(std::chrono::duration<double, std::micro>(now - s.t_submit[i]).count());
            *n_ready += 1;
        }
        if (s.emitted == s.n) { r->in_flight.pop_front(); s.busy = false; }
    }
    if (r->latencies_us.size() > (1u << 20)) r->latencies_us.erase(r->latencies_us.begin(), r->latencies_us.begin() + (1u << 19));
    return FTF_OK;
}

extern "C" int ftf_poll_status(ftf_runner* r, int32_t* status, int32_t cap, const char** message) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    if (cap < 0 || (cap && !status)) return fail(r, FTF_INVALID_ARGUMENT, "poll_status: null status array");
    int first = FTF_OK;
    for (size_t i = 0; i < r->poll_status.size(); ++i) {
        if ((int32_t)i < cap) status[i] = r->poll_status[i];
        if (r->poll_status[i] != FTF_OK && first == FTF_OK) first = r->poll_status[i];
    }
    if (message) *message = first != FTF_OK ? r->poll_error.c_str() : "";
    return first;
}

extern "C" int ftf_poll_select(ftf_runner* r, int32_t n_keys, const char* const* keys) {
    if (!r) return fail(nullptr, FTF_INVALID_ARGUMENT, "a session must be provided");
    if (n_keys <= 0 || !keys) return fail(r, FTF_INVALID_ARGUMENT, "poll_select: at least one output key");
    if (r->outs.empty()) return fail(r, FTF_FAILED_PRECONDITION, "poll_select: submit a record first (outputs are fixed by the first record)");
    std::vector<int> sel;
    for (int i = 0; i < n_keys; ++i) {
        int found = -1;
        for (size_t j = 0; j < r->outs.size(); ++j) if (keys[i] && r->outs[j].key == keys[i]) found = (int)j;
        if (found < 0) return fail(r, FTF_NOT_FOUND, std::string("no output named '") + (keys[i] ? keys[i] : "(null)") + "' in this signature");
        sel.push_back(found);
    }
    r->poll_sel = sel;
    return FTF_OK;
}

extern "C" int ftf_stats_get(const ftf_runner* r, ftf_stats* out) {
    if (!r || !out) return FTF_INVALID_ARGUMENT;
    *out = r->stats;
    if (!r->latencies_us.empty()) {
        std::vector<double> v = r->latencies_us;
        std::sort(v.begin(), v.end());
        out->queue_latency_us_p50 = v[v.size() / 2];
        out->queue_latency_us_p99 = v[std::min(v.size() - 1, (size_t)(v.size() * 0.99))];
    }
    return FTF_OK;
}

// ------------------------------------------------------------------------------------------------
// device-level entry points
// ------------------------------------------------------------------------------------------------
extern "C" int ftf_affine_f32_dev(const float* x, float* y, int64_t n, float a, float b, void* stream) {
    std::string err;
    const int rc = launch_affine_f32(x, y, n, a, b, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}
extern "C" int ftf_resize_norm_dev(const uint8_t* in, int64_t n, int32_t in_h, int32_t in_w, int32_t out_h, int32_t out_w,
                                   float mean, float scale, void* out, int32_t out_dtype, void* stream) {
    std::string err;
    const int rc = launch_resize_norm(in, n, in_h, in_w, out_h, out_w, mean, scale, out, out_dtype, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}
extern "C" int ftf_softmax_topk_dev(const float* x, int64_t m, int32_t n, int32_t n_labels, int32_t k, int32_t apply_softmax,
                                    float* probs, int32_t* idx, float* score, void* stream) {
    std::string err;
    if (k > 32) { set_thread_error("softmax_topk: k must be <= 32"); return FTF_INVALID_ARGUMENT; }
    const int rc = launch_softmax_topk(x, m, n, n_labels, k, apply_softmax, probs, idx, score, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}
extern "C" int ftf_label_dev(ftf_runner* r, const void* images, int32_t in_dtype, int64_t n, float* logits, float* probs,
                             int32_t* idx, float* score) {
    if (!r || !is_label_kind(r->kind)) return fail(r, FTF_INVALID_ARGUMENT, "label_dev needs a label runner");
    int rc = set_device(r);
    if (rc) return rc;
    if (n < 0 || n > r->net->max_batch()) return fail(r, FTF_INVALID_ARGUMENT, "label_dev: batch exceeds the runner's max_batch");
    if (n == 0) return FTF_OK;
    Slot& s = r->slots[0];
    const void* img = images;
    const uint64_t before = launches_so_far();
    if (in_dtype != FTF_FLOAT && in_dtype != FTF_BFLOAT16) return fail(r, FTF_INVALID_ARGUMENT, "label_dev: images must be FLOAT or BFLOAT16");
    float* lg = logits;
    if (!lg) {
        if (s.d_out.size() < 2) s.d_out.resize(2);
        if ((rc = s.d_out[1].ensure((size_t)n * r->cfg.num_classes * 4, &r->err))) return rc;
        lg = (float*)s.d_out[1].p;
    }
    if ((rc = r->net->forward(img, in_dtype, n, lg, r->net_stream, &r->err))) { if (rc == FTF_INTERNAL) r->dead = true; return rc; }
    if (probs || (idx && score))
        rc = launch_softmax_topk(lg, n, r->cfg.num_classes, r->cfg.num_classes, (idx && score) ? r->cfg.top_k : 0, 1, probs, idx, score,
                                 r->net_stream, &r->err);
    r->stats.kernel_launches += launches_so_far() - before;
    return rc;
}

extern "C" int ftf_conv2d_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld, const void* weights,
                                   const float* bias, int32_t cout, int32_t kh, int32_t kw, int32_t stride, int32_t pad_h,
                                   int32_t pad_w, int32_t relu, void* out, int32_t out_dtype, int32_t out_ld, int32_t out_off,
                                   void* stream) {
    std::string err;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_thread_error("cudaGetDevice failed"); return FTF_INTERNAL; }
    if (n <= 0 || stride <= 0 || (out_dtype != FTF_BFLOAT16 && out_dtype != FTF_FLOAT)) {
        set_thread_error("conv2d_bf16: bad batch, stride or output dtype");
        return FTF_INVALID_ARGUMENT;
    }
    ConvSegment sg{};
    sg.out = out; sg.ld = out_ld; sg.coff = out_off; sg.col_begin = 0;
    sg.flags = (bias ? kSegBias : 0) | (relu ? kSegRelu : 0) | (out_dtype == FTF_FLOAT ? kSegF32 : 0);
    ConvGeometry g{in, n, h, w, cin, in_ld, weights, bias, cout, kh, kw, stride, pad_h, pad_w};
    const ConvPlan* plan = nullptr;
    int rc = conv_tc_plan_cached(g, &sg, 1, device_sm_count(dev), dev, &plan, &err);
    if (!rc) rc = conv_tc_launch(*plan, n, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}

extern "C" int ftf_label_export_weights(ftf_runner* r, const char* dir) {
    if (!r || !is_label_kind(r->kind) || !dir) return fail(r, FTF_INVALID_ARGUMENT, "label_export_weights needs a label runner and a directory");
    return r->net->export_weights(dir, &r->err);
}

extern "C" int ftf_debug_activation(ftf_runner* r, const char* name, int64_t n, float* host_out, int64_t cap, int32_t* dims4) {
    if (!r || !is_label_kind(r->kind) || !name || !dims4) return fail(r, FTF_INVALID_ARGUMENT, "debug_activation needs a label runner, a name and dims");
    int rc = set_device(r);
    if (rc) return rc;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { r->dead = true; return cuda_fail(e, "debug_activation", &r->err); }
    int d[4];
    rc = r->net->activation(name, n, host_out, cap, d, &r->err);
    for (int i = 0; i < 4; ++i) dims4[i] = d[i];
    return rc;
}

extern "C" int ftf_maxpool3x3s2_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t c, void* out, int32_t out_ld,
                                         int32_t out_off, void* stream) {
    std::string err;
    const int rc = launch_maxpool3x3s2(in, out, n, h, w, c, out_ld, out_off, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}

extern "C" int ftf_global_avgpool_bf16_dev(const void* in, int32_t n, int32_t hw, int32_t c, void* out, void* stream) {
    std::string err;
    const int rc = launch_global_avgpool(in, out, n, hw, c, (cudaStream_t)stream, &err);
    if (rc) set_thread_error(err);
    return rc;
}

extern "C" int ftf_pool_branch_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld,
                                        const void* weights, const float* bias, int32_t cout, void* out, int32_t out_ld,
                                        int32_t out_off, void* stream) {
    std::string err;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_thread_error("cudaGetDevice failed"); return FTF_INTERNAL; }
    if (n <= 0 || h <= 0 || w <= 0 || cout <= 0 || !bias) { set_thread_error("pool_branch: bad shape or null bias"); return FTF_INVALID_ARGUMENT; }
    cudaStream_t s = (cudaStream_t)stream;
    // scratch: the 1x1 convolution's fp32 pre-activation [n*h*w, cout], allocated in stream order
    float* pre = nullptr;
    const size_t bytes = (size_t)n * h * w * cout * sizeof(float);
    cudaError_t e = cudaMallocAsync((void**)&pre, bytes, s);
    if (e != cudaSuccess) { cudaGetLastError(); set_thread_error(std::string("pool_branch scratch: ") + cudaGetErrorString(e)); return FTF_RESOURCE_EXHAUSTED; }
    ConvSegment sg{};
    sg.out = pre; sg.ld = cout; sg.coff = 0; sg.col_begin = 0; sg.flags = kSegF32;
    ConvGeometry g{in, n, h, w, cin, in_ld, weights, nullptr, cout, 1, 1, 1, 0, 0};
    const ConvPlan* plan = nullptr;
    int rc = conv_tc_plan_cached(g, &sg, 1, device_sm_count(dev), dev, &plan, &err);
    if (!rc) rc = conv_tc_launch(*plan, n, s, &err);
    if (!rc) rc = launch_avgpool3x3_bias_relu(pre, bias, out, n, h, w, cout, out_ld, out_off, s, &err);
    cudaFreeAsync(pre, s);
    if (rc) set_thread_error(err);
    return rc;
}

extern "C" int ftf_debug_conv_profile(uint64_t* out16, int32_t reset) {
    if (!out16) return FTF_INVALID_ARGUMENT;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { set_thread_error(std::string("conv_profile: ") + cudaGetErrorString(e)); return FTF_INTERNAL; }
    return conv_tc_profile_read((unsigned long long*)out16, reset);
}
