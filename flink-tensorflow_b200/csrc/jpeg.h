// jpeg.h -- baseline JPEG decoding on the GPU: the DecodeJpeg(channels=3) node in front of the normalisation graph
// (EX/common/GraphBuilder.java:41-47; EX/inception/ImageNormalization.scala:56-57 feeds it the file bytes read by
// EX/inception/ImageInputFormat.scala:63-80).  The host only reads the marker segments (a few hundred bytes per file);
// entropy decoding, inverse DCT, upsampling and colour conversion run in csrc/jpeg.cu.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include "jpeg_core.h"

namespace ftf {
namespace jpeg {

// A Huffman table as a DHT segment defines it (T.81 B.2.4.2): how many codes of each length, then the values.
struct HuffSpec {
    uint8_t counts[17];
    uint8_t values[256];
    int32_t nvalues;
};
// Decoder form of a (validated) table definition.
void build_table(const HuffSpec& spec, HuffTable* t);

// What the marker segments of one file say (host side).
struct Parsed {
    int width = 0, height = 0, ncomp = 0;
    int hs = 1, vs = 1;
    int restart = 0;                       // MCUs per restart interval, 0 = none
    int tq[3] = {0, 0, 0}, td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
    uint16_t quant[4][64];                 // row-major
    bool quant_present[4] = {false, false, false, false};
    HuffSpec dc[4], ac[4];                 // as the file defines them; build_table makes the decoder form
    bool dc_present[4] = {false, false, false, false}, ac_present[4] = {false, false, false, false};
    int64_t scan_begin = 0;                // first entropy-coded byte
};

// Reads SOI .. SOS of one file.  Returns an FTF status; *err says what was wrong or unsupported.
int parse_headers(const uint8_t* d, int64_t n, Parsed* out, std::string* err);

// Geometry shared by the host planner and the kernels.
void fill_geometry(const Parsed& p, ImageDesc* im);

// Upper bound of what stage_scan writes for this file.
int64_t stage_bound(int64_t n, const Parsed& p, const ImageDesc& im);

// Copies the entropy-coded bytes [p.scan_begin, n) of one file to dst with the 0xFF00 byte stuffing removed, one clean
// stream per restart interval (or one for the whole scan), each 4-byte aligned and followed by >= 16 zero bytes, and
// appends one Segment per stream (begin = dst_offset + position in dst).  Returns the bytes written.  This is the one
// pass the host makes over the compressed data; it doubles as the copy into pinned staging memory.
// *truncated: the file ended inside the scan, before any closing marker (the decoder would read zero bits from there on).
int64_t stage_scan(const uint8_t* d, int64_t n, const Parsed& p, const ImageDesc& im, int image, uint8_t* dst,
                   int64_t dst_offset, std::vector<Segment>* out, bool* truncated = nullptr);

}  // namespace jpeg
}  // namespace ftf

// The pieces of ftf_jpeg_batch_load, for callers that receive their files one at a time (ftf_submit): a file is parsed
// and staged when it arrives, the batch is uploaded when it closes.  batch_add copies what it needs: the caller's buffer
// is free again when it returns (unless the caller allows the staging to be deferred, see may_defer).
struct ftf_jpeg_batch;
struct CUstream_st;
namespace ftf {
namespace jpeg {
int batch_begin(ftf_jpeg_batch* b);
// may_defer: the caller leaves the file's bytes alone until the batch has been committed (batch_commit waits); they may then be
// staged by the batch's worker thread while the caller goes on.
int batch_add(ftf_jpeg_batch* b, const uint8_t* file, int64_t len, int32_t* height, int32_t* width, int64_t* rgb_offset, bool may_defer = false);
int batch_commit(ftf_jpeg_batch* b, CUstream_st* stream);
// Returns when no deferred file is being read any more (a caller that gives up on a batch and takes its buffers back).
void batch_wait_staged(ftf_jpeg_batch* b);
}  // namespace jpeg
}  // namespace ftf
