// J1-J3  baseline JPEG decoding for batches of files: DecodeJpeg(channels=3), the node the reference's normalisation
// graph starts with (EX/common/GraphBuilder.java:41-47, EX/inception/ImageNormalization.scala:56-57; EX = flink-tensorflow-
// examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples).  TF-Java runs it in libjpeg on one core per record;
// here a batch of files is decoded by three kernels and the RGB images never leave HBM before the resize kernel reads them.
//
//   J1a jpeg_huffman_kernel  one CTA per independently decodable segment (restart interval, else the whole scan), one
//                            thread per 1024-bit chunk of it: every chunk is decoded from a guessed state, the guesses
//                            are corrected along the chain until they are the true states (a Huffman decoder started
//                            in the wrong place falls into step with the true one after a kilobit or two), then a last
//                            pass scatters the coefficients (de-zigzagged) into a zeroed int16 buffer.  Canonical
//                            Huffman decode with 9-bit lookahead tables (copies in shared memory) on a byte-unstuffed
//                            stream (the host removes the stuffing while it copies the file into pinned memory).
//   J1b jpeg_dc_kernel       DC differences -> DC values: warp scan per (segment, component).
//   J2 jpeg_idct_kernel      one thread per 8x8 block: dequantise + libjpeg's integer slow IDCT, entirely in
//                            registers, 16-byte coefficient loads, 8-byte row stores into the component plane.
//   J3 jpeg_colour_kernel    one thread per 8 pixels: planes read as aligned words, triangle-filter chroma upsampling
//                            (2x1, 2x2) with every source sample fetched once, fixed-point YCbCr -> RGB packed with
//                            saturation, six 32-bit stores; 32 registers, eight CTAs per SM (it waits on its loads).
// Integer arithmetic throughout: results are bit-identical to libjpeg(-turbo), checked against oracle/jpeg_oracle.c
// and the Pillow-made fixtures in tests/golden/jpeg.
#include "jpeg.h"
#include "runner.h"
#include <string.h>
#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

using namespace ftf;
using namespace ftf::jpeg;

namespace {

constexpr size_t kHuffSmemPerThread = sizeof(ScanState) * (2 + kCheckpoints) + sizeof(int) + sizeof(uint16_t) * (kCheckpoints + 1);

// J1a: one CTA per segment, one thread per chunk of kChunkBits (see decode_chunk in jpeg_core.h for the scheme).
// Segments with more chunks than the CTA has threads are walked in waves: a wave starts from the true state the
// previous wave ended in.
template <int kMaxThreads, int kMinBlocks>
__global__ void __launch_bounds__(kMaxThreads, kMinBlocks)
jpeg_huffman_kernel(const Segment* __restrict__ segs, const ImageDesc* __restrict__ images,
                    const HuffTable* __restrict__ tables, const uint8_t* __restrict__ scan,
                    int16_t* __restrict__ coef, int16_t* __restrict__ dcv, int32_t* __restrict__ status) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ int s_warp[32];
    // the image's lookahead tables in shared memory ([component][dc, ac]): every symbol starts with a gather from one of
    // them, and 32 lanes gathering 2-byte entries from global memory cost up to 32 L1 wavefronts where shared memory
    // costs a few bank conflicts
    __shared__ uint16_t s_lut[6 << kLutBits];
    const Segment sg = segs[blockIdx.x];
    const ImageDesc& im = images[sg.image];
    const int T = (int)blockDim.x, tid = (int)threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < (im.ncomp * 2) << kLutBits; i += T) {
        const int t = i >> kLutBits, c = t >> 1;
        s_lut[i] = tables[(t & 1) ? im.ac[c] : im.dc[c]].lut[i & ((1 << kLutBits) - 1)];
    }
    __syncthreads();
    // dynamic shared memory, all indexed by the chunk's position in the wave: exit states, entry states, checkpoint states
    // [kCheckpoints][T], blocks completed, checkpoint block counts [kCheckpoints][T], work list of a correction round
    ScanState* s_exit = reinterpret_cast<ScanState*>(smem_raw);
    ScanState* s_in = s_exit + T;
    ScanState* s_cps = s_in + T;
    int* s_n = reinterpret_cast<int*>(s_cps + (size_t)T * kCheckpoints);
    uint16_t* s_cpn = reinterpret_cast<uint16_t*>(s_n + T);
    uint16_t* s_list = s_cpn + (size_t)T * kCheckpoints;
    const int bpm = im.ncomp == 1 ? 1 : im.hs * im.vs + 2;
    const int total_blocks = sg.n_mcus * bpm;
    ScanState wave_entry;
    wave_entry.pos = 0; wave_entry.k = 0; wave_entry.b = 0;
    int blocks_before = 0, bad = 0;
    for (int c0 = 0; c0 < sg.n_chunks; c0 += T) {
        const int j = c0 + tid;
        const bool active = j < sg.n_chunks;
        const uint32_t limit = (uint32_t)(j + 1) * kChunkBits;
        ScanState in;
        in.pos = (uint32_t)j * kChunkBits; in.k = 0; in.b = 0;
        if (tid == 0) in = wave_entry;
        ScanState out = in;
        int n = 0;
        // 1. guess
        if (active) out = decode_chunk<kGuess>(sg, im, tables, scan, in, limit, 0, 0, nullptr, nullptr, &n, nullptr,
                                               Checkpoints{s_cps + tid, s_cpn + tid, T}, ScanState{0, 0, 0}, s_lut);
        s_exit[tid] = out;
        s_n[tid] = n;
        // 2. until every chunk started from the state its predecessor left.  The chunks that have to be decoded again are
        // gathered into a list and handed to the first threads of the CTA, so that late rounds with few such chunks run
        // in few, full warps.
        for (;;) {
            __syncthreads();
            bool changed = false;
            ScanState prev = in;
            if (active && tid > 0) { prev = s_exit[tid - 1]; changed = !same_state(prev, in); }
            const unsigned bal = __ballot_sync(0xffffffffu, changed);
            if (lane == 0) s_warp[warp] = __popc(bal);
            __syncthreads();
            int before = 0, total = 0;
            for (int w = 0; w < (T + 31) / 32; ++w) {
                const int v = s_warp[w];
                if (w < warp) before += v;
                total += v;
            }
            if (total == 0) break;
            if (changed) {
                in = prev;
                s_in[tid] = prev;
                s_list[before + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)tid;
            }
            __syncthreads();
            if (tid < total) {
                const int c = s_list[tid];
                int cn = s_n[c];
                const ScanState e = decode_chunk<kCorrect>(sg, im, tables, scan, s_in[c], (uint32_t)(c0 + c + 1) * kChunkBits, 0, 0, nullptr,
                                                           nullptr, &cn, nullptr, Checkpoints{s_cps + c, s_cpn + c, T}, s_exit[c], s_lut);
                s_exit[c] = e;
                s_n[c] = cn;
            }
        }
        __syncthreads();                   // the round loop's last look at s_warp is over before the scan reuses it
        n = s_n[tid];
        // 3. which block does each chunk start in: exclusive prefix sum of the blocks completed per chunk
        int incl = active ? n : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        int warp_before = 0, wave_total = 0;
        for (int w = 0; w < (T + 31) / 32; ++w) {
            const int v = s_warp[w];
            if (w < warp) warp_before += v;
            wave_total += v;
        }
        const int ordinal = blocks_before + warp_before + incl - (active ? n : 0);
        if (active && ordinal < total_blocks) {
            int n2 = 0;
            decode_chunk<kStore>(sg, im, tables, scan, in, j == sg.n_chunks - 1 ? 0xffffffffu : limit, ordinal, total_blocks, coef, dcv, &n2, &bad, Checkpoints{nullptr, nullptr, 0}, ScanState{0, 0, 0}, s_lut);
        }
        if (c0 + T < sg.n_chunks) {             // another wave follows: it starts where this one ended
            wave_entry = s_exit[T - 1];
            blocks_before += wave_total;
            __syncthreads();
        }
    }
    if (bad) atomicOr(&status[sg.image], 1);
}

// J1b: DC differences -> DC values.  One warp per (segment, component): an inclusive scan over the component's
// contiguous stretch of the DC array, 32 elements per step with a carry.
__global__ void __launch_bounds__(128)
jpeg_dc_kernel(const Segment* __restrict__ segs, int nseg, const ImageDesc* __restrict__ images, int16_t* __restrict__ dcv) {
    const int w = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = (int)(threadIdx.x & 31);
    const int s = w / 3, c = w - s * 3;
    if (s >= nseg) return;
    const Segment sg = segs[s];
    const ImageDesc& im = images[sg.image];
    if (c >= im.ncomp) return;
    int64_t first;
    int count;
    dc_range(sg, im, c, &first, &count);
    int16_t* p = dcv + first;
    int carry = 0;
    for (int t0 = 0; t0 < count; t0 += 32) {
        const int t = t0 + lane;
        int v = t < count ? (int)p[t] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        v += carry;
        if (t < count) p[t] = (int16_t)v;
        carry = __shfl_sync(0xffffffffu, v, 31);
    }
}

#ifndef FTF_J2_MINBLOCKS
#define FTF_J2_MINBLOCKS 1
#endif
__global__ void __launch_bounds__(128, FTF_J2_MINBLOCKS)
jpeg_idct_kernel(const ImageDesc* __restrict__ images, const uint16_t* __restrict__ quant,
                 const int16_t* __restrict__ coef, const int16_t* __restrict__ dcv, uint8_t* __restrict__ planes) {
    const ImageDesc& im = images[blockIdx.y];
    int b = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (b >= im.n_blocks) return;
    int c = 0;
    const int nb0 = im.bw[0] * im.bh[0];
    if (b >= nb0) {
        b -= nb0; c = 1;
        const int nb1 = im.bw[1] * im.bh[1];
        if (b >= nb1) { b -= nb1; c = 2; }
    }
    const int bw = im.bw[c];
    const int by = b / bw, bx = b - by * bw;
    int16_t cf[64];
    uint16_t q[64];
    const int4* cp = reinterpret_cast<const int4*>(coef + im.coef_off[c] + (int64_t)b * 64);
    const int4* qp = reinterpret_cast<const int4*>(quant + (int64_t)im.quant[c] * 64);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        reinterpret_cast<int4*>(cf)[i] = __ldg(cp + i);
        reinterpret_cast<int4*>(q)[i] = __ldg(qp + i);
    }
    cf[0] = dc_of_block(im, dcv, c, by, bx);
    idct_block(cf, q, planes + im.plane_off[c] + ((int64_t)by * 8 * bw + bx) * 8, bw * 8);
}

#ifndef FTF_J3_MINBLOCKS
#define FTF_J3_MINBLOCKS 8          /* 32 registers, no spills: measured 179 -> 141 us against 40 registers / 6 CTAs per SM */
#endif
__global__ void __launch_bounds__(256, FTF_J3_MINBLOCKS)
jpeg_colour_kernel(const ImageDesc* __restrict__ images, const uint8_t* __restrict__ planes, uint8_t* __restrict__ rgb) {
    const ImageDesc& im = images[blockIdx.y];
    const int groups = (im.width + 7) >> 3;
    const int t = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (t >= groups * im.height) return;
    const int y = t / groups, x0 = (t - y * groups) * 8;
    const int nx = im.width - x0 < 8 ? im.width - x0 : 8;
    union { uint8_t b[24]; uint32_t w[6]; } px;
    pixels8_rgb(im, planes, y, x0, px.b);
    uint8_t* dst = rgb + im.rgb_off + ((int64_t)y * im.width + x0) * 3;
    if (nx == 8 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
        for (int i = 0; i < 6; ++i) d32[i] = px.w[i];
    } else {
        for (int i = 0; i < 3 * nx; ++i) dst[i] = px.b[i];
    }
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

// A batch of JPEG files being decoded: host-side plan, pinned staging, device buffers.  Grow-only; reused across loads.
struct ftf_jpeg_batch {
    int device = 0;
    std::string err;
    cudaEvent_t done = nullptr;
    Buf h_scan, h_meta, h_status, d_scan, d_meta, d_coef, d_dc, d_planes, d_status;
    std::vector<ImageDesc> images;
    std::vector<Segment> segments;
    std::vector<HuffTable> tables;
    std::vector<HuffSpec> specs;          // the definitions tables[] were built from, same order
    std::vector<uint16_t> quants;
    size_t scan_bytes = 0, meta_bytes = 0, off_seg = 0, off_tab = 0, off_q = 0;
    int64_t coef_elems = 0, plane_bytes = 0, rgb_bytes = 0, dc_elems = 0;
    int max_blocks = 0, max_groups = 0, max_chunks = 1;
    bool decoded = false, attr_set = false, committed = false;
    Parsed parsed;                        // scratch of batch_add
    // Staging (the one host pass over the compressed bytes: unstuff, split at restart markers, copy into pinned memory) costs
    // 12 us per 77 KB file -- two thirds of what the submitting thread spends per record.  For files whose bytes the caller
    // leaves alone until the result is out (may_defer) it runs on a worker thread of the batch, into a region of h_scan
    // reserved by the submitting thread; batch_commit waits for the worker.  Segments are kept per file and put in image order
    // at commit, so the batch is laid out exactly as if everything had been staged in place.
    struct StageTask { const uint8_t* file; int64_t len; Parsed parsed; ImageDesc im; int image; size_t dst_off; };
    std::deque<std::vector<Segment>> file_segs;     // per image (a deque: the worker holds references while images are added)
    std::vector<std::thread> workers;   // kStageWorkers of them, started with the first deferred file
    std::mutex mu;
    std::condition_variable cv_work, cv_idle;
    std::deque<StageTask> tasks;
    int outstanding = 0;
    bool stop = false, stage_failed = false;
    void worker_loop() {
        std::unique_lock<std::mutex> lock(mu);
        for (;;) {
            cv_work.wait(lock, [&] { return stop || !tasks.empty(); });
            if (stop) return;                   // whatever is still queued is dropped: its buffers may be gone
            StageTask t = std::move(tasks.front());
            tasks.pop_front();
            std::vector<Segment>& out = file_segs[(size_t)t.image];
            uint8_t* dst = (uint8_t*)h_scan.p + t.dst_off;
            lock.unlock();
            bool truncated = false;
            stage_scan(t.file, t.len, t.parsed, t.im, t.image, dst, (int64_t)t.dst_off, &out, &truncated);
            lock.lock();
            if (truncated) stage_failed = true;
            if (--outstanding == 0) cv_idle.notify_all();
        }
    }
    void wait_staged() {
        std::unique_lock<std::mutex> lock(mu);
        cv_idle.wait(lock, [&] { return outstanding == 0; });
    }
};

namespace {

int bfail(ftf_jpeg_batch* b, int code, const std::string& msg) {
    if (b) b->err = msg;
    set_thread_error(msg);
    return code;
}

template <typename T>
int pool_index(std::vector<T>& pool, const T& t) {
    const int n = (int)pool.size();
    for (int i = n - 1; i >= 0 && i >= n - 8; --i)
        if (memcmp(&pool[i], &t, sizeof(T)) == 0) return i;
    pool.push_back(t);
    return n;
}

}  // namespace

extern "C" int ftf_jpeg_info(const uint8_t* file, int64_t len, int32_t* height, int32_t* width, int32_t* components) {
    Parsed* p = new Parsed();
    std::string err;
    const int rc = parse_headers(file, len, p, &err);
    if (rc == FTF_OK) {
        if (height) *height = p->height;
        if (width) *width = p->width;
        if (components) *components = p->ncomp;
    } else set_thread_error(err);
    delete p;
    return rc;
}

extern "C" int ftf_jpeg_batch_create(int32_t device, ftf_jpeg_batch** out) {
    if (!out) return FTF_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return bfail(nullptr, FTF_FAILED_PRECONDITION, "no CUDA device is visible; the JPEG decoder has no CPU path");
    }
    if (device < 0 || device >= count) return bfail(nullptr, FTF_INVALID_ARGUMENT, "device ordinal out of range");
    ftf_jpeg_batch* b = new ftf_jpeg_batch();
    b->device = device;
    b->h_scan.host = b->h_meta.host = b->h_status.host = true;
    std::string err;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming);
    if (e != cudaSuccess) { cuda_fail(e, "ftf_jpeg_batch_create", &err); delete b; return bfail(nullptr, FTF_INTERNAL, err); }
    *out = b;
    return FTF_OK;
}

extern "C" void ftf_jpeg_batch_destroy(ftf_jpeg_batch* b) {
    if (!b) return;
    cudaSetDevice(b->device);
    if (!b->workers.empty()) {
        { std::lock_guard<std::mutex> lock(b->mu); b->stop = true; }
        b->cv_work.notify_all();
        for (auto& t : b->workers) t.join();
    }
    if (b->done) { cudaEventSynchronize(b->done); cudaEventDestroy(b->done); }
    for (Buf* f : {&b->h_scan, &b->h_meta, &b->h_status, &b->d_scan, &b->d_meta, &b->d_coef, &b->d_dc, &b->d_planes, &b->d_status})
        f->free_now();
    cudaGetLastError();
    delete b;
}

extern "C" const char* ftf_jpeg_batch_error(const ftf_jpeg_batch* b) { return b ? b->err.c_str() : thread_error(); }

namespace ftf {
namespace jpeg {

int batch_begin(ftf_jpeg_batch* b) {
    if (!b) return FTF_INVALID_ARGUMENT;
    FTF_CUDA(cudaSetDevice(b->device), &b->err);
    // the previous batch's uploads and kernels may still be reading the staging buffers
    FTF_CUDA(cudaEventSynchronize(b->done), &b->err);
    b->wait_staged();                      // an abandoned batch may still have files on their way into h_scan
    b->file_segs.clear(); b->stage_failed = false;
    b->images.clear(); b->segments.clear(); b->tables.clear(); b->specs.clear(); b->quants.clear();
    b->decoded = false; b->committed = false;
    b->coef_elems = 0; b->plane_bytes = 0; b->rgb_bytes = 0; b->dc_elems = 0; b->max_blocks = 0; b->max_groups = 0;
    b->max_chunks = 1; b->scan_bytes = 0;
    return FTF_OK;
}

int batch_add(ftf_jpeg_batch* b, const uint8_t* file, int64_t len, int32_t* height, int32_t* width, int64_t* rgb_offset, bool may_defer) {
    Parsed& p = b->parsed;
    p = Parsed();
    std::string perr;
    const int i = (int)b->images.size();
    if (len > (int64_t)1 << 28)            // bit positions inside a segment are 32-bit
        return bfail(b, FTF_UNIMPLEMENTED, "image " + std::to_string(i) + ": DecodeJpeg: files larger than 256 MiB are not supported");
    int rc = parse_headers(file, len, &p, &perr);
    if (rc) return bfail(b, rc, "image " + std::to_string(i) + ": " + perr);
    // TF's DecodeJpeg refuses frames of 2^29 or more output bytes before it sizes anything ("Image too large"); a forged
    // 65535x65535 header must not turn into a multi-gigabyte allocation here either
    if ((int64_t)p.width * p.height * 3 >= ((int64_t)1 << 29))
        return bfail(b, FTF_INVALID_ARGUMENT, "image " + std::to_string(i) + ": DecodeJpeg: image too large (" + std::to_string(p.width) + "x" +
                     std::to_string(p.height) + ")");
    ImageDesc im;
    fill_geometry(p, &im);
    // what a refused file must give back: the batch's sizes are only committed once the scan has been staged whole
    const int64_t coef0 = b->coef_elems, plane0 = b->plane_bytes, dc0 = b->dc_elems, rgb0 = b->rgb_bytes;
    const int blocks0 = b->max_blocks, groups0 = b->max_groups;
    // room in the pinned staging buffer (grown with its contents kept; sized generously so that this is rare)
    const size_t need = b->scan_bytes + align_up((size_t)stage_bound(len, p, im), 16);
    if (need > b->h_scan.cap) {
        b->wait_staged();                  // the worker writes into the buffer that is about to move
        Buf bigger;
        bigger.host = true;
        if ((rc = bigger.ensure(std::max<size_t>(need * 2, 8u << 20), &b->err))) return rc;
        if (b->scan_bytes) memcpy(bigger.p, b->h_scan.p, b->scan_bytes);
        b->h_scan.free_now();
        b->h_scan = bigger;
    }
    for (int c = 0; c < p.ncomp; ++c) {
        // tables are pooled per batch: files from one encoder share them, so look at the last few entries
        const uint16_t* q = p.quant[p.tq[c]];
        int found = -1;
        const int nq = (int)(b->quants.size() / 64);
        for (int k = nq - 1; k >= 0 && k >= nq - 8; --k)
            if (memcmp(&b->quants[(size_t)k * 64], q, 128) == 0) { found = k; break; }
        if (found < 0) { found = nq; b->quants.insert(b->quants.end(), q, q + 64); }
        im.quant[c] = found;
        for (int which = 0; which < 2; ++which) {      // decoder tables are built once per distinct definition
            const HuffSpec& sp = which ? p.ac[p.ta[c]] : p.dc[p.td[c]];
            const int idx = pool_index(b->specs, sp);
            if (idx == (int)b->tables.size()) { b->tables.emplace_back(); build_table(sp, &b->tables.back()); }
            (which ? im.ac : im.dc)[c] = idx;
        }
        im.coef_off[c] = b->coef_elems;
        b->coef_elems += (int64_t)im.bw[c] * im.bh[c] * 64;
        im.plane_off[c] = b->plane_bytes;
        b->plane_bytes += (int64_t)im.bw[c] * im.bh[c] * 64;
        im.dc_off[c] = b->dc_elems;
        b->dc_elems += (int64_t)im.bw[c] * im.bh[c];
    }
    im.rgb_off = b->rgb_bytes;
    if (rgb_offset) *rgb_offset = im.rgb_off;
    if (height) *height = im.height;
    if (width) *width = im.width;
    b->rgb_bytes += align_up((size_t)im.width * im.height * 3, 16);
    if (im.n_blocks > b->max_blocks) b->max_blocks = im.n_blocks;
    const int groups = ((im.width + 7) >> 3) * im.height;
    if (groups > b->max_groups) b->max_groups = groups;
    // the one host pass over the compressed bytes: unstuff + split at restart markers + copy into pinned memory
    {
        std::lock_guard<std::mutex> lock(b->mu);
        b->file_segs.emplace_back();
    }
    // A file that ends with its end-of-image marker cannot turn out truncated (the scan below stops at a marker at the
    // latest there), so nothing about it can be refused any more: its staging may run behind the caller's back.
    if (may_defer && len >= 2 && file[len - 2] == 0xFF && file[len - 1] == 0xD9) {
        if (b->workers.empty()) {
            // two threads keep up with a submitting thread that spends 4 us per record on everything else (12 us of staging per
            // 77 KB file); FTF_JPEG_STAGE_THREADS: measurement switch
            static const int n_workers = getenv("FTF_JPEG_STAGE_THREADS") ? std::max(1, std::min(8, atoi(getenv("FTF_JPEG_STAGE_THREADS")))) : 2;
            for (int k = 0; k < n_workers; ++k) b->workers.emplace_back([b] { b->worker_loop(); });
        }
        {
            std::lock_guard<std::mutex> lock(b->mu);
            b->tasks.push_back(ftf_jpeg_batch::StageTask{file, len, p, im, i, b->scan_bytes});
            b->outstanding += 1;
        }
        b->cv_work.notify_one();
        b->scan_bytes += align_up((size_t)stage_bound(len, p, im), 16);      // reserved (the kernels read whole aligned words); the staged streams are a little shorter
        b->images.push_back(im);
        return FTF_OK;
    }
    bool truncated = false;
    std::vector<Segment>& segs = b->file_segs.back();              // (references into a deque stay valid)
    const size_t staged = (size_t)stage_scan(file, len, p, im, i, (uint8_t*)b->h_scan.p + b->scan_bytes, (int64_t)b->scan_bytes, &segs, &truncated);
    if (truncated) {
        // the reference's graph leaves DecodeJpeg's try_recover_truncated at false (GraphBuilder.java:41-47 sets channels
        // only): TF refuses a file that ends before its end-of-image marker instead of padding it
        { std::lock_guard<std::mutex> lock(b->mu); b->file_segs.pop_back(); }
        b->coef_elems = coef0; b->plane_bytes = plane0; b->dc_elems = dc0; b->rgb_bytes = rgb0;
        b->max_blocks = blocks0; b->max_groups = groups0;
        return bfail(b, FTF_INVALID_ARGUMENT, "image " + std::to_string(i) + ": DecodeJpeg: invalid JPEG data (the file ends inside the scan)");
    }
    b->scan_bytes += staged;
    b->images.push_back(im);
    return FTF_OK;
}

void batch_wait_staged(ftf_jpeg_batch* b) { if (b) b->wait_staged(); }

int batch_commit(ftf_jpeg_batch* b, cudaStream_t st) {
    const size_t n = b->images.size();
    b->wait_staged();
    if (b->stage_failed) return bfail(b, FTF_INTERNAL, "DecodeJpeg: a deferred file turned out truncated");
    b->segments.clear();
    b->max_chunks = 1;
    for (const auto& fs : b->file_segs)
        for (const Segment& sg : fs) {
            b->segments.push_back(sg);
            if (sg.n_chunks > b->max_chunks) b->max_chunks = sg.n_chunks;
        }
    if (n == 0) { b->committed = true; return FTF_OK; }
    int rc;
    b->off_seg = align_up(n * sizeof(ImageDesc), 16);
    b->off_tab = align_up(b->off_seg + b->segments.size() * sizeof(Segment), 16);
    b->off_q = align_up(b->off_tab + b->tables.size() * sizeof(HuffTable), 16);
    b->meta_bytes = b->off_q + b->quants.size() * sizeof(uint16_t);
    if ((rc = b->h_meta.ensure(b->meta_bytes, &b->err)) || (rc = b->d_scan.ensure(b->scan_bytes + 16, &b->err)) ||
        (rc = b->d_meta.ensure(b->meta_bytes, &b->err)) || (rc = b->d_coef.ensure((size_t)b->coef_elems * 2, &b->err)) ||
        (rc = b->d_dc.ensure((size_t)b->dc_elems * 2, &b->err)) || (rc = b->d_planes.ensure((size_t)b->plane_bytes, &b->err)) ||
        (rc = b->d_status.ensure(n * 4, &b->err)) || (rc = b->h_status.ensure(n * 4, &b->err)))
        return rc;
    uint8_t* m = (uint8_t*)b->h_meta.p;
    memcpy(m, b->images.data(), n * sizeof(ImageDesc));
    memcpy(m + b->off_seg, b->segments.data(), b->segments.size() * sizeof(Segment));
    memcpy(m + b->off_tab, b->tables.data(), b->tables.size() * sizeof(HuffTable));
    memcpy(m + b->off_q, b->quants.data(), b->quants.size() * sizeof(uint16_t));
    FTF_CUDA(cudaMemcpyAsync(b->d_scan.p, b->h_scan.p, b->scan_bytes, cudaMemcpyHostToDevice, st), &b->err);
    FTF_CUDA(cudaMemcpyAsync(b->d_meta.p, b->h_meta.p, b->meta_bytes, cudaMemcpyHostToDevice, st), &b->err);
    FTF_CUDA(cudaEventRecord(b->done, st), &b->err);
    b->committed = true;
    return FTF_OK;
}

}  // namespace jpeg
}  // namespace ftf

extern "C" int ftf_jpeg_batch_load(ftf_jpeg_batch* b, int32_t n, const uint8_t* const* files, const int64_t* lens,
                                   int64_t* rgb_offsets, int32_t* heights, int32_t* widths, void* stream) {
    if (!b) return FTF_INVALID_ARGUMENT;
    if (n < 0 || (n > 0 && (!files || !lens))) return bfail(b, FTF_INVALID_ARGUMENT, "ftf_jpeg_batch_load: null argument");
    int rc = batch_begin(b);
    for (int i = 0; i < n && rc == FTF_OK; ++i)
        rc = batch_add(b, files[i], lens[i], heights ? heights + i : nullptr, widths ? widths + i : nullptr, rgb_offsets ? rgb_offsets + i : nullptr, /*may_defer=*/true);
    if (rc == FTF_OK) rc = batch_commit(b, (cudaStream_t)stream);
    if (rc) { b->wait_staged(); b->images.clear(); b->segments.clear(); }      // (the caller's buffers are his again on return)
    return rc;
}

extern "C" int64_t ftf_jpeg_batch_rgb_bytes(const ftf_jpeg_batch* b) { return b ? b->rgb_bytes : 0; }
extern "C" int64_t ftf_jpeg_batch_scan_bytes(const ftf_jpeg_batch* b) { return b ? (int64_t)(b->scan_bytes + b->meta_bytes) : 0; }
extern "C" int32_t ftf_jpeg_batch_segments(const ftf_jpeg_batch* b) { return b ? (int32_t)b->segments.size() : 0; }

extern "C" int ftf_jpeg_batch_decode_dev(ftf_jpeg_batch* b, uint8_t* rgb_out, int64_t cap, void* stream) {
    if (!b) return FTF_INVALID_ARGUMENT;
    if (b->images.empty()) return FTF_OK;
    if (!b->committed) return bfail(b, FTF_FAILED_PRECONDITION, "ftf_jpeg_batch_decode_dev: the batch has not been uploaded");
    if (!rgb_out || cap < b->rgb_bytes) return bfail(b, FTF_INVALID_ARGUMENT, "ftf_jpeg_batch_decode_dev: output buffer too small");
    cudaStream_t st = (cudaStream_t)stream;
    FTF_CUDA(cudaSetDevice(b->device), &b->err);
    const int n = (int)b->images.size(), nseg = (int)b->segments.size();
    const uint8_t* m = (const uint8_t*)b->d_meta.p;
    const ImageDesc* d_images = (const ImageDesc*)m;
    const Segment* d_segs = (const Segment*)(m + b->off_seg);
    const HuffTable* d_tabs = (const HuffTable*)(m + b->off_tab);
    const uint16_t* d_quant = (const uint16_t*)(m + b->off_q);
    FTF_CUDA(cudaMemsetAsync(b->d_coef.p, 0, (size_t)b->coef_elems * 2, st), &b->err);
    FTF_CUDA(cudaMemsetAsync(b->d_dc.p, 0, (size_t)b->dc_elems * 2, st), &b->err);
    FTF_CUDA(cudaMemsetAsync(b->d_status.p, 0, (size_t)n * 4, st), &b->err);
    // CTA size: one thread per chunk up to 1024; up to 640 threads the kernel is built for two CTAs per SM (48 registers)
    int threads = (b->max_chunks + 31) / 32 * 32;
    threads = threads > 1024 ? 1024 : threads;
    static const int thread_cap = getenv("FTF_JPEG_THREADS") ? atoi(getenv("FTF_JPEG_THREADS")) / 32 * 32 : 0;      // measurement switch
    if (thread_cap >= 32 && threads > thread_cap) threads = thread_cap;
    const size_t hsm = (size_t)threads * kHuffSmemPerThread;
    if (!b->attr_set) {                  // once per batch object: the attribute belongs to the function on this device
        FTF_CUDA(cudaFuncSetAttribute(jpeg_huffman_kernel<640, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 640 * (int)kHuffSmemPerThread), &b->err);
        FTF_CUDA(cudaFuncSetAttribute(jpeg_huffman_kernel<1024, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 * (int)kHuffSmemPerThread), &b->err);
        b->attr_set = true;
    }
    if (threads <= 640)
        jpeg_huffman_kernel<640, 2><<<nseg, threads, hsm, st>>>(d_segs, d_images, d_tabs, (const uint8_t*)b->d_scan.p, (int16_t*)b->d_coef.p,
                                                             (int16_t*)b->d_dc.p, (int32_t*)b->d_status.p);
    else
        jpeg_huffman_kernel<1024, 1><<<nseg, threads, hsm, st>>>(d_segs, d_images, d_tabs, (const uint8_t*)b->d_scan.p, (int16_t*)b->d_coef.p,
                                                              (int16_t*)b->d_dc.p, (int32_t*)b->d_status.p);
    jpeg_dc_kernel<<<(nseg * 3 + 3) / 4, 128, 0, st>>>(d_segs, nseg, d_images, (int16_t*)b->d_dc.p);
    jpeg_idct_kernel<<<dim3((b->max_blocks + 127) / 128, n), 128, 0, st>>>(d_images, d_quant, (const int16_t*)b->d_coef.p,
                                                                            (const int16_t*)b->d_dc.p, (uint8_t*)b->d_planes.p);
    jpeg_colour_kernel<<<dim3((b->max_groups + 255) / 256, n), 256, 0, st>>>(d_images, (const uint8_t*)b->d_planes.p, rgb_out);
    FTF_CUDA(cudaGetLastError(), &b->err);
    count_launch(4);
    FTF_CUDA(cudaMemcpyAsync(b->h_status.p, b->d_status.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st), &b->err);
    FTF_CUDA(cudaEventRecord(b->done, st), &b->err);
    b->decoded = true;
    return FTF_OK;
}

extern "C" int ftf_jpeg_batch_status(ftf_jpeg_batch* b, int32_t* per_image, int32_t n) {
    if (!b) return FTF_INVALID_ARGUMENT;
    if (!b->decoded) return bfail(b, FTF_FAILED_PRECONDITION, "ftf_jpeg_batch_status: nothing has been decoded yet");
    FTF_CUDA(cudaSetDevice(b->device), &b->err);
    FTF_CUDA(cudaEventSynchronize(b->done), &b->err);
    int rc = FTF_OK;
    const int32_t* s = (const int32_t*)b->h_status.p;
    for (int i = 0; i < (int)b->images.size(); ++i) {
        const int code = s[i] ? FTF_INVALID_ARGUMENT : FTF_OK;
        if (per_image && i < n) per_image[i] = code;
        if (code && rc == FTF_OK) rc = bfail(b, FTF_INVALID_ARGUMENT, "image " + std::to_string(i) + ": DecodeJpeg: invalid JPEG data (corrupt entropy-coded segment)");
    }
    return rc;
}
