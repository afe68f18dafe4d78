// jpeg_parse.cpp -- host side of the JPEG decoder: marker segments -> tables and geometry (see jpeg.h).
// Marker syntax follows ITU-T T.81 annex B; what is accepted matches what the kernels decode: 8-bit baseline or
// extended-sequential Huffman files (SOF0/SOF1) with one interleaved scan of 1 or 3 components, luma sampled 1x1, 2x1 or
// 2x2 against 1x1 chroma, optional restart intervals.  Progressive, arithmetic-coded, lossless, CMYK and multi-scan
// files are UNIMPLEMENTED (TF-Java would hand those to libjpeg as well; callers see UnsupportedOperationException).
#include "jpeg.h"
#include <string.h>
#include "../../include/ftf.h"

namespace ftf {
namespace jpeg {

namespace {

inline int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

int bad(std::string* err, int code, const char* msg) {
    if (err) *err = std::string("DecodeJpeg: ") + msg;
    return code;
}

// a prefix code needs no more codes of a length than the shorter codes leave room for
bool valid_spec(const HuffSpec& sp) {
    int code = 0;
    for (int l = 1; l <= 16; ++l) {
        code += sp.counts[l];
        if (code > (1 << l)) return false;
        code <<= 1;
    }
    return true;
}

}  // namespace

void build_table(const HuffSpec& sp, HuffTable* t) {
    memset(t, 0, sizeof(*t));
    memcpy(t->vals, sp.values, (size_t)sp.nvalues);
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        const int cnt = sp.counts[l];
        t->valoff[l] = k - code;
        if (l <= kLutBits) {
            for (int i = 0; i < cnt; ++i) {
                const int first = (code + i) << (kLutBits - l), span = 1 << (kLutBits - l);
                const uint16_t e = (uint16_t)((l << 8) | sp.values[k + i]);
                for (int j = 0; j < span; ++j) t->lut[first + j] = e;
            }
        }
        code += cnt;
        k += cnt;
        t->maxcode[l] = cnt ? code - 1 : -1;
        code <<= 1;
    }
    t->maxcode[0] = -1;
    t->maxcode[17] = 0x7fffffff;
    t->valoff[0] = 0;
}

int parse_headers(const uint8_t* d, int64_t n, Parsed* out, std::string* err) {
    if (!d || n < 4 || d[0] != 0xFF || d[1] != 0xD8) return bad(err, FTF_INVALID_ARGUMENT, "not a JPEG file (no SOI marker)");
    int64_t p = 2;
    bool have_frame = false;
    int comp_id[3] = {0, 0, 0};
    for (;;) {
        while (p < n && d[p] != 0xFF) ++p;
        while (p < n && d[p] == 0xFF) ++p;
        if (p >= n) return bad(err, FTF_INVALID_ARGUMENT, "file ends before the scan");
        const int m = d[p++];
        if (m == 0xD9) return bad(err, FTF_INVALID_ARGUMENT, "end of image before any scan");
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (p + 2 > n) return bad(err, FTF_INVALID_ARGUMENT, "truncated marker segment");
        const int len = be16(d + p);
        if (len < 2 || p + len > n) return bad(err, FTF_INVALID_ARGUMENT, "marker segment runs past the end of the file");
        const uint8_t* q = d + p + 2;
        int rem = len - 2;
        switch (m) {
        case 0xDB:
            while (rem > 0) {
                const int pq = q[0] >> 4, tq = q[0] & 15;
                const int need = 1 + 64 * (pq + 1);
                if (tq > 3 || pq > 1 || rem < need) return bad(err, FTF_INVALID_ARGUMENT, "bad quantisation table segment");
                for (int i = 0; i < 64; ++i) {
                    const int v = pq ? be16(q + 1 + 2 * i) : q[1 + i];
                    out->quant[tq][natural_order(i)] = (uint16_t)v;
                }
                out->quant_present[tq] = true;
                q += need; rem -= need;
            }
            break;
        case 0xC4:
            while (rem > 0) {
                if (rem < 17) return bad(err, FTF_INVALID_ARGUMENT, "bad Huffman table segment");
                const int tc = q[0] >> 4, th = q[0] & 15;
                int cnt = 0;
                for (int i = 1; i <= 16; ++i) cnt += q[i];
                if (tc > 1 || th > 3 || cnt > 256 || rem < 17 + cnt) return bad(err, FTF_INVALID_ARGUMENT, "bad Huffman table segment");
                HuffSpec* t = tc ? &out->ac[th] : &out->dc[th];
                memset(t, 0, sizeof(*t));
                memcpy(t->counts, q, 17);
                t->counts[0] = 0;
                memcpy(t->values, q + 17, (size_t)cnt);
                t->nvalues = cnt;
                if (!valid_spec(*t)) return bad(err, FTF_INVALID_ARGUMENT, "Huffman table is not a prefix code");
                (tc ? out->ac_present : out->dc_present)[th] = true;
                q += 17 + cnt; rem -= 17 + cnt;
            }
            break;
        case 0xC0: case 0xC1: {
            if (have_frame || rem < 6) return bad(err, FTF_INVALID_ARGUMENT, "bad frame header");
            if (q[0] != 8) return bad(err, FTF_UNIMPLEMENTED, "only 8-bit samples are supported");
            out->height = be16(q + 1); out->width = be16(q + 3); out->ncomp = q[5];
            if (out->height <= 0 || out->width <= 0) return bad(err, FTF_INVALID_ARGUMENT, "empty image");
            if (out->ncomp != 1 && out->ncomp != 3) return bad(err, FTF_UNIMPLEMENTED, "only 1- and 3-component files are supported");
            if (rem < 6 + 3 * out->ncomp) return bad(err, FTF_INVALID_ARGUMENT, "bad frame header");
            int h[3] = {1, 1, 1}, v[3] = {1, 1, 1};
            for (int i = 0; i < out->ncomp; ++i) {
                comp_id[i] = q[6 + 3 * i];
                h[i] = q[7 + 3 * i] >> 4; v[i] = q[7 + 3 * i] & 15;
                out->tq[i] = q[8 + 3 * i];
                if (h[i] < 1 || h[i] > 4 || v[i] < 1 || v[i] > 4 || out->tq[i] > 3) return bad(err, FTF_INVALID_ARGUMENT, "bad frame header");
            }
            if (out->ncomp == 1) { out->hs = out->vs = 1; }      // a one-component scan is never interleaved (T.81 A.2.2)
            else {
                if (h[1] != 1 || v[1] != 1 || h[2] != 1 || v[2] != 1 ||
                    !((h[0] == 1 && v[0] == 1) || (h[0] == 2 && v[0] == 1) || (h[0] == 2 && v[0] == 2)))
                    return bad(err, FTF_UNIMPLEMENTED, "chroma subsampling other than 4:4:4, 4:2:2 and 4:2:0");
                out->hs = h[0]; out->vs = v[0];
            }
            have_frame = true;
            break;
        }
        case 0xC2: case 0xC3: case 0xC5: case 0xC6: case 0xC7: case 0xC9: case 0xCA: case 0xCB: case 0xCD: case 0xCE: case 0xCF:
            return bad(err, FTF_UNIMPLEMENTED, "progressive, lossless, hierarchical and arithmetic-coded files are not supported");
        case 0xDD:
            if (rem < 2) return bad(err, FTF_INVALID_ARGUMENT, "bad restart interval segment");
            out->restart = be16(q);
            break;
        case 0xDA: {
            if (!have_frame || rem < 1) return bad(err, FTF_INVALID_ARGUMENT, "scan before the frame header");
            const int ns = q[0];
            if (ns != out->ncomp) return bad(err, FTF_UNIMPLEMENTED, "files with more than one scan are not supported");
            if (rem < 1 + 2 * ns + 3) return bad(err, FTF_INVALID_ARGUMENT, "bad scan header");
            for (int i = 0; i < ns; ++i) {
                if (q[1 + 2 * i] != comp_id[i]) return bad(err, FTF_UNIMPLEMENTED, "scan components out of frame order");
                out->td[i] = q[2 + 2 * i] >> 4; out->ta[i] = q[2 + 2 * i] & 15;
                if (out->td[i] > 3 || out->ta[i] > 3) return bad(err, FTF_INVALID_ARGUMENT, "bad scan header");
                if (!out->dc_present[out->td[i]] || !out->ac_present[out->ta[i]] || !out->quant_present[out->tq[i]])
                    return bad(err, FTF_INVALID_ARGUMENT, "scan refers to a table the file does not define");
            }
            out->scan_begin = p + len;
            return FTF_OK;
        }
        default: break;   // APPn, COM and anything else with a length: skipped
        }
        p += len;
    }
}

void fill_geometry(const Parsed& p, ImageDesc* im) {
    memset(im, 0, sizeof(*im));
    im->width = p.width; im->height = p.height; im->ncomp = p.ncomp;
    im->hs = p.hs; im->vs = p.vs;
    im->mcus_x = (p.width + 8 * p.hs - 1) / (8 * p.hs);
    im->mcus_y = (p.height + 8 * p.vs - 1) / (8 * p.vs);
    im->n_blocks = 0;
    for (int c = 0; c < p.ncomp; ++c) {
        const int h = c == 0 ? p.hs : 1, v = c == 0 ? p.vs : 1;
        im->bw[c] = im->mcus_x * h; im->bh[c] = im->mcus_y * v;
        im->dw[c] = (p.width * h + p.hs - 1) / p.hs;
        im->dh[c] = (p.height * v + p.vs - 1) / p.vs;
        im->n_blocks += im->bw[c] * im->bh[c];
    }
}

static int segment_count(const Parsed& p, const ImageDesc& im) {
    const int total = im.mcus_x * im.mcus_y;
    return (p.restart > 0 && p.restart < total) ? (total + p.restart - 1) / p.restart : 1;
}

int64_t stage_bound(int64_t n, const Parsed& p, const ImageDesc& im) {
    return (n - p.scan_begin) + 24ll * segment_count(p, im) + 16;
}

int64_t stage_scan(const uint8_t* d, int64_t n, const Parsed& p, const ImageDesc& im, int image, uint8_t* dst,
                   int64_t dst_offset, std::vector<Segment>* out, bool* truncated) {
    if (truncated) *truncated = false;
    const int total = im.mcus_x * im.mcus_y;
    const int nseg = segment_count(p, im);
    const int per = nseg > 1 ? p.restart : total;
    int64_t pos = p.scan_begin, w = 0;
    bool ended = false;
    for (int k = 0; k < nseg; ++k) {
        Segment s{};
        s.image = image;
        s.first_mcu = k * per;
        s.n_mcus = total - s.first_mcu < per ? total - s.first_mcu : per;
        s.begin = dst_offset + w;
        const int64_t w0 = w;
        while (!ended) {
            const uint8_t* f = pos < n ? (const uint8_t*)memchr(d + pos, 0xFF, (size_t)(n - pos)) : nullptr;
            const int64_t upto = f ? f - d : n;
            memcpy(dst + w, d + pos, (size_t)(upto - pos));
            w += upto - pos;
            pos = upto;
            if (!f || pos + 1 >= n) {                                 // the file ends inside the scan, without a marker
                ended = true;
                if (truncated) *truncated = true;
                break;
            }
            const int m = d[pos + 1];
            if (m == 0x00) { dst[w++] = 0xFF; pos += 2; continue; }   // stuffed byte
            if (m == 0xFF) { pos += 1; continue; }                    // fill byte in front of a marker
            if (m >= 0xD0 && m <= 0xD7 && nseg > 1) { pos += 2; break; }   // RSTn closes this interval
            ended = true;                                             // EOI or any other marker ends the scan
        }
        s.nbits = (w - w0) * 8;
        s.n_chunks = (int32_t)((s.nbits + kChunkBits - 1) / kChunkBits);
        if (s.n_chunks < 1) s.n_chunks = 1;
        const int64_t padded = (w + 3) / 4 * 4 + 16;
        memset(dst + w, 0, (size_t)(padded - w));
        w = padded;
        out->push_back(s);
    }
    return w;
}

}  // namespace jpeg
}  // namespace ftf
