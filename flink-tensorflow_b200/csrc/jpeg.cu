// J1-J3  baseline JPEG decoding for batches of files: DecodeJpeg(channels=3), the node the reference's normalisation
// graph starts with (EX/common/GraphBuilder.java:41-47, EX/inception/ImageNormalization.scala:56-57; EX = flink-tensorflow-
// examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples).  TF-Java runs it in libjpeg on one core per record;
// here a batch of files is decoded by three kernels and the RGB images never leave HBM before the resize kernel reads them.
//
//   J1 jpeg_huffman_kernel   one thread per independently decodable segment (restart interval, else whole scan):
//                            canonical Huffman decode with a 9-bit lookahead table, byte unstuffing on the fly,
//                            coefficients scattered (de-zigzagged) into a zeroed int16 buffer.  A latency-bound
//                            kernel: its parallelism is segments per batch, so segments are spread one per warp
//                            until there are more segments than the GPU holds warps.
//   J2 jpeg_idct_kernel      one thread per 8x8 block: dequantise + libjpeg's integer slow IDCT, entirely in
//                            registers, 16-byte coefficient loads, 8-byte row stores into the component plane.
//   J3 jpeg_colour_kernel    one thread per 4 pixels: triangle-filter chroma upsampling (2x1, 2x2), fixed-point
//                            YCbCr -> RGB, three 32-bit stores of packed RGB.
// Integer arithmetic throughout: results are bit-identical to libjpeg(-turbo), checked against oracle/jpeg_oracle.c
// and the Pillow-made fixtures in tests/golden/jpeg.
#include "jpeg.h"
#include "runner.h"
#include <string.h>

using namespace ftf;
using namespace ftf::jpeg;

namespace {

__global__ void __launch_bounds__(128)
jpeg_huffman_kernel(const Segment* __restrict__ segs, int nseg, int lanes, const ImageDesc* __restrict__ images,
                    const HuffTable* __restrict__ tables, const uint8_t* __restrict__ scan,
                    int16_t* __restrict__ coef, int32_t* __restrict__ status) {
    const int warp = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = (int)(threadIdx.x & 31);
    if (lane >= lanes) return;
    const int s = warp * lanes + lane;
    if (s >= nseg) return;
    const Segment sg = segs[s];
    if (decode_segment(sg, images[sg.image], tables, scan, coef)) atomicOr(&status[sg.image], 1);
}

__global__ void __launch_bounds__(128)
jpeg_idct_kernel(const ImageDesc* __restrict__ images, const uint16_t* __restrict__ quant,
                 const int16_t* __restrict__ coef, uint8_t* __restrict__ planes) {
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
    idct_block(cf, q, planes + im.plane_off[c] + ((int64_t)by * 8 * bw + bx) * 8, bw * 8);
}

__global__ void __launch_bounds__(256)
jpeg_colour_kernel(const ImageDesc* __restrict__ images, const uint8_t* __restrict__ planes, uint8_t* __restrict__ rgb) {
    const ImageDesc& im = images[blockIdx.y];
    const int groups = (im.width + 3) >> 2;
    const int t = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (t >= groups * im.height) return;
    const int y = t / groups, x0 = (t - y * groups) * 4;
    const int nx = im.width - x0 < 4 ? im.width - x0 : 4;
    union { uint8_t b[12]; uint32_t w[3]; } px;
    px.w[0] = px.w[1] = px.w[2] = 0;
    for (int i = 0; i < nx; ++i) pixel_rgb(im, planes, y, x0 + i, px.b + 3 * i);
    uint8_t* dst = rgb + im.rgb_off + ((int64_t)y * im.width + x0) * 3;
    if (nx == 4 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
        uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
        d32[0] = px.w[0]; d32[1] = px.w[1]; d32[2] = px.w[2];
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
    Buf h_scan, h_meta, h_status, d_scan, d_meta, d_coef, d_planes, d_status;
    std::vector<ImageDesc> images;
    std::vector<Segment> segments;
    std::vector<HuffTable> tables;
    std::vector<uint16_t> quants;
    size_t scan_bytes = 0, meta_bytes = 0, off_seg = 0, off_tab = 0, off_q = 0;
    int64_t coef_elems = 0, plane_bytes = 0, rgb_bytes = 0;
    int max_blocks = 0, max_groups = 0;
    bool decoded = false;
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
    if (b->done) { cudaEventSynchronize(b->done); cudaEventDestroy(b->done); }
    for (Buf* f : {&b->h_scan, &b->h_meta, &b->h_status, &b->d_scan, &b->d_meta, &b->d_coef, &b->d_planes, &b->d_status})
        f->free_now();
    cudaGetLastError();
    delete b;
}

extern "C" const char* ftf_jpeg_batch_error(const ftf_jpeg_batch* b) { return b ? b->err.c_str() : thread_error(); }

extern "C" int ftf_jpeg_batch_load(ftf_jpeg_batch* b, int32_t n, const uint8_t* const* files, const int64_t* lens,
                                   int64_t* rgb_offsets, int32_t* heights, int32_t* widths, void* stream) {
    if (!b) return FTF_INVALID_ARGUMENT;
    if (n < 0 || (n > 0 && (!files || !lens))) return bfail(b, FTF_INVALID_ARGUMENT, "ftf_jpeg_batch_load: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    FTF_CUDA(cudaSetDevice(b->device), &b->err);
    // the previous load's staging may still be in flight on another stream
    FTF_CUDA(cudaEventSynchronize(b->done), &b->err);
    b->images.clear(); b->segments.clear(); b->tables.clear(); b->quants.clear();
    b->decoded = false;
    b->coef_elems = 0; b->plane_bytes = 0; b->rgb_bytes = 0; b->max_blocks = 0; b->max_groups = 0;
    std::vector<Parsed>* parsed = new std::vector<Parsed>((size_t)n);
    std::vector<int64_t> scan_off((size_t)n);
    size_t scan_total = 0;
    int rc = FTF_OK;
    for (int i = 0; i < n && rc == FTF_OK; ++i) {
        Parsed& p = (*parsed)[i];
        std::string perr;
        rc = parse_headers(files[i], lens[i], &p, &perr);
        if (rc) { bfail(b, rc, "image " + std::to_string(i) + ": " + perr); break; }
        ImageDesc im;
        fill_geometry(p, &im);
        for (int c = 0; c < p.ncomp; ++c) {
            // tables are pooled per batch: files from one encoder share them, so look at the last few entries
            const uint16_t* q = p.quant[p.tq[c]];
            int found = -1;
            const int nq = (int)(b->quants.size() / 64);
            for (int k = nq - 1; k >= 0 && k >= nq - 8; --k)
                if (memcmp(&b->quants[(size_t)k * 64], q, 128) == 0) { found = k; break; }
            if (found < 0) { found = nq; b->quants.insert(b->quants.end(), q, q + 64); }
            im.quant[c] = found;
            im.dc[c] = pool_index(b->tables, p.dc[p.td[c]]);
            im.ac[c] = pool_index(b->tables, p.ac[p.ta[c]]);
            im.coef_off[c] = b->coef_elems;
            b->coef_elems += (int64_t)im.bw[c] * im.bh[c] * 64;
            im.plane_off[c] = b->plane_bytes;
            b->plane_bytes += (int64_t)im.bw[c] * im.bh[c] * 64;
        }
        im.rgb_off = b->rgb_bytes;
        if (rgb_offsets) rgb_offsets[i] = im.rgb_off;
        if (heights) heights[i] = im.height;
        if (widths) widths[i] = im.width;
        b->rgb_bytes += align_up((size_t)im.width * im.height * 3, 16);
        if (im.n_blocks > b->max_blocks) b->max_blocks = im.n_blocks;
        const int groups = ((im.width + 3) >> 2) * im.height;
        if (groups > b->max_groups) b->max_groups = groups;
        scan_off[i] = (int64_t)scan_total;
        scan_total += align_up((size_t)(lens[i] - p.scan_begin), 16);
        const size_t first_seg = b->segments.size();
        find_segments(files[i], lens[i], p, im, i, &b->segments);
        for (size_t k = first_seg; k < b->segments.size(); ++k) {      // file offsets -> offsets into the batch buffer
            b->segments[k].begin += scan_off[i] - p.scan_begin;
            b->segments[k].end += scan_off[i] - p.scan_begin;
        }
        b->images.push_back(im);
    }
    if (rc == FTF_OK && n > 0) {
        b->scan_bytes = scan_total;
        b->off_seg = align_up(b->images.size() * sizeof(ImageDesc), 16);
        b->off_tab = align_up(b->off_seg + b->segments.size() * sizeof(Segment), 16);
        b->off_q = align_up(b->off_tab + b->tables.size() * sizeof(HuffTable), 16);
        b->meta_bytes = b->off_q + b->quants.size() * sizeof(uint16_t);
        if ((rc = b->h_scan.ensure(scan_total + 16, &b->err)) == 0 && (rc = b->h_meta.ensure(b->meta_bytes, &b->err)) == 0 &&
            (rc = b->d_scan.ensure(scan_total + 16, &b->err)) == 0 && (rc = b->d_meta.ensure(b->meta_bytes, &b->err)) == 0 &&
            (rc = b->d_coef.ensure((size_t)b->coef_elems * 2, &b->err)) == 0 &&
            (rc = b->d_planes.ensure((size_t)b->plane_bytes, &b->err)) == 0 &&
            (rc = b->d_status.ensure((size_t)n * 4, &b->err)) == 0 && (rc = b->h_status.ensure((size_t)n * 4, &b->err)) == 0) {
            for (int i = 0; i < n; ++i) {
                const Parsed& p = (*parsed)[i];
                memcpy((uint8_t*)b->h_scan.p + scan_off[i], files[i] + p.scan_begin, (size_t)(lens[i] - p.scan_begin));
            }
            uint8_t* m = (uint8_t*)b->h_meta.p;
            memcpy(m, b->images.data(), b->images.size() * sizeof(ImageDesc));
            memcpy(m + b->off_seg, b->segments.data(), b->segments.size() * sizeof(Segment));
            memcpy(m + b->off_tab, b->tables.data(), b->tables.size() * sizeof(HuffTable));
            memcpy(m + b->off_q, b->quants.data(), b->quants.size() * sizeof(uint16_t));
            cudaError_t e = cudaMemcpyAsync(b->d_scan.p, b->h_scan.p, scan_total, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(b->d_meta.p, b->h_meta.p, b->meta_bytes, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaEventRecord(b->done, st);
            if (e != cudaSuccess) rc = cuda_fail(e, "ftf_jpeg_batch_load upload", &b->err);
        }
    }
    delete parsed;
    if (rc) { b->images.clear(); b->segments.clear(); }
    return rc;
}

extern "C" int64_t ftf_jpeg_batch_rgb_bytes(const ftf_jpeg_batch* b) { return b ? b->rgb_bytes : 0; }
extern "C" int64_t ftf_jpeg_batch_scan_bytes(const ftf_jpeg_batch* b) { return b ? (int64_t)(b->scan_bytes + b->meta_bytes) : 0; }
extern "C" int32_t ftf_jpeg_batch_segments(const ftf_jpeg_batch* b) { return b ? (int32_t)b->segments.size() : 0; }

extern "C" int ftf_jpeg_batch_decode_dev(ftf_jpeg_batch* b, uint8_t* rgb_out, int64_t cap, void* stream) {
    if (!b) return FTF_INVALID_ARGUMENT;
    if (b->images.empty()) return FTF_OK;
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
    FTF_CUDA(cudaMemsetAsync(b->d_status.p, 0, (size_t)n * 4, st), &b->err);
    // one segment per warp while the GPU has warps to spare; beyond that pack lanes
    const int warps_wanted = device_sm_count(b->device) * 48;
    int lanes = (nseg + warps_wanted - 1) / warps_wanted;
    lanes = lanes < 1 ? 1 : lanes > 32 ? 32 : lanes;
    const int warps = (nseg + lanes - 1) / lanes;
    const int wpb = lanes == 1 ? 1 : 4;          // lone lanes: one warp per block, so that they spread over all SMs
    jpeg_huffman_kernel<<<(warps + wpb - 1) / wpb, wpb * 32, 0, st>>>(d_segs, nseg, lanes, d_images, d_tabs, (const uint8_t*)b->d_scan.p,
                                                                   (int16_t*)b->d_coef.p, (int32_t*)b->d_status.p);
    jpeg_idct_kernel<<<dim3((b->max_blocks + 127) / 128, n), 128, 0, st>>>(d_images, d_quant, (const int16_t*)b->d_coef.p,
                                                                            (uint8_t*)b->d_planes.p);
    jpeg_colour_kernel<<<dim3((b->max_groups + 255) / 256, n), 256, 0, st>>>(d_images, (const uint8_t*)b->d_planes.p, rgb_out);
    FTF_CUDA(cudaGetLastError(), &b->err);
    count_launch(3);
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
