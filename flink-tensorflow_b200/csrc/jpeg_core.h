// jpeg_core.h -- the per-thread work of the JPEG kernels (csrc/jpeg.cu), written as plain functions over plain
// structs so that tests/jpeg_emulate.cpp can also compile them with g++ and step them thread by thread on a machine
// without a GPU (a test harness for the kernel logic; the library itself never runs them on the host).
//
// Reference: the DecodeJpeg(channels=3) node of the normalisation graph, EX/common/GraphBuilder.java:41-47 (EX =
// flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples), executed by libjpeg inside
// TF-Java 1.1 with its defaults: integer slow IDCT, fancy upsampling, fixed-point YCbCr -> RGB.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define FTF_HD __host__ __device__ __forceinline__
#else
#define FTF_HD inline
#endif
#ifdef __CUDA_ARCH__
#define FTF_LDG(p) __ldg(p)
#else
#define FTF_LDG(p) (*(p))
#endif

namespace ftf {
namespace jpeg {

constexpr int kLutBits = 9;

// One Huffman table, decoder form.  lut: the next kLutBits bits of the stream -> (code length << 8 | symbol), 0 when
// the code is longer than kLutBits; then the canonical-code walk of T.81 F.2.2.3 over maxcode / valoff.
struct HuffTable {
    uint16_t lut[1 << kLutBits];
    int32_t maxcode[18];     // largest code of each length (-1: none); [17] = INT32_MAX ends the walk
    int32_t valoff[17];      // index of the first value of that length minus its first code
    uint8_t vals[256];
};

// One image of a batch.  Offsets are into the batch's buffers; tables are indices into the batch's table pools.
struct ImageDesc {
    int32_t width, height, ncomp;
    int32_t hs, vs;               // luma sampling factors (chroma is 1x1): 1x1, 2x1 or 2x2
    int32_t mcus_x, mcus_y;
    int32_t bw[3], bh[3];         // block grid per component, whole MCUs
    int32_t dw[3], dh[3];         // real (downsampled) sample counts per component
    int32_t quant[3], dc[3], ac[3];
    int32_t first_block;          // index of this image's first block in the batch-wide block list
    int32_t n_blocks;
    int64_t coef_off[3];          // int16 elements
    int64_t plane_off[3];         // bytes
    int64_t rgb_off;              // bytes
};

// One independently decodable stretch of a scan: a restart interval, or the whole scan.
struct Segment {
    int32_t image;
    int32_t first_mcu, n_mcus;
    int32_t pad;
    int64_t begin, end;           // bytes of the batch's scan buffer
};

// zig-zag position -> row-major position (T.81 figure A.6)
#define FTF_JPEG_NATURAL_ORDER                                                                                        \
    {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, \
     35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63}
#ifdef __CUDACC__
static __constant__ uint8_t k_natural_dev[64] = FTF_JPEG_NATURAL_ORDER;
#endif
static const uint8_t k_natural_host[64] = FTF_JPEG_NATURAL_ORDER;
FTF_HD uint8_t natural_order(int k) {
#ifdef __CUDA_ARCH__
    return k_natural_dev[k];
#else
    return k_natural_host[k];
#endif
}

// MSB-first bit reader over entropy-coded bytes with the 0xFF00 stuffing removed on the fly; at a marker (or the end
// of the segment) the stream continues as zero bits, which is what libjpeg feeds its decoder there.
struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t buf;
    int cnt;
    FTF_HD void init(const uint8_t* b, const uint8_t* e) { p = b; end = e; buf = 0; cnt = 0; }
    FTF_HD void fill() {
        while (cnt <= 56) {
            uint32_t b = 0;
            if (p < end) {
                b = FTF_LDG(p);
                if (b == 0xFF) {
                    const uint32_t b2 = (p + 1 < end) ? (uint32_t)FTF_LDG(p + 1) : 0xD9u;
                    if (b2 == 0) p += 2;
                    else { b = 0; end = p; }
                } else ++p;
            }
            buf |= (uint64_t)b << (56 - cnt);
            cnt += 8;
        }
    }
    FTF_HD uint32_t peek(int n) const { return (uint32_t)(buf >> (64 - n)); }
    FTF_HD void skip(int n) { buf <<= n; cnt -= n; }
};

// T.81 F.2.2.3 DECODE; returns the symbol or -1 for a code no table entry matches.  Needs >= 16 bits in the reader.
FTF_HD int decode_symbol(BitReader& br, const HuffTable* __restrict__ t) {
    const uint32_t e = FTF_LDG(&t->lut[br.peek(kLutBits)]);
    if (e) { br.skip((int)(e >> 8)); return (int)(e & 0xFF); }
    const int32_t code16 = (int32_t)br.peek(16);
    int l = kLutBits + 1;
    while (l <= 16 && (code16 >> (16 - l)) > FTF_LDG(&t->maxcode[l])) ++l;
    if (l > 16) return -1;
    br.skip(l);
    return (int)FTF_LDG(&t->vals[(code16 >> (16 - l)) + FTF_LDG(&t->valoff[l])]);
}

// T.81 F.2.2.1 EXTEND applied to the next t bits
FTF_HD int receive_extend(BitReader& br, int t) {
    const int v = (int)br.peek(t);
    br.skip(t);
    return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v;
}

// Decodes one segment: every block of its MCUs, non-zero coefficients stored at their row-major position in the
// (pre-zeroed) coefficient buffer.  Returns 0, or 1 when the data is not a valid Huffman stream for these tables.
FTF_HD int decode_segment(const Segment& sg, const ImageDesc& im, const HuffTable* __restrict__ tables,
                          const uint8_t* __restrict__ scan, int16_t* __restrict__ coef) {
    BitReader br;
    br.init(scan + sg.begin, scan + sg.end);
    int pred[3] = {0, 0, 0};
    int mx = sg.first_mcu % im.mcus_x, my = sg.first_mcu / im.mcus_x;
    for (int m = 0; m < sg.n_mcus; ++m) {
        for (int c = 0; c < im.ncomp; ++c) {
            const HuffTable* dc = tables + im.dc[c];
            const HuffTable* ac = tables + im.ac[c];
            const int ch = c == 0 ? im.hs : 1, cv = c == 0 ? im.vs : 1;
            for (int v = 0; v < cv; ++v)
                for (int h = 0; h < ch; ++h) {
                    int16_t* blk = coef + im.coef_off[c] + ((int64_t)(my * cv + v) * im.bw[c] + (mx * ch + h)) * 64;
                    br.fill();
                    const int t = decode_symbol(br, dc);
                    if (t < 0 || t > 16) return 1;
                    if (t) pred[c] += receive_extend(br, t);
                    blk[0] = (int16_t)pred[c];
                    for (int k = 1; k < 64; ++k) {
                        br.fill();
                        const int rs = decode_symbol(br, ac);
                        if (rs < 0) return 1;
                        const int r = rs >> 4, s = rs & 15;
                        if (s) {
                            k += r;
                            const int val = receive_extend(br, s);
                            if (k > 63) return 1;
                            blk[natural_order(k)] = (int16_t)val;
                        } else {
                            if (r != 15) break;
                            k += 15;
                        }
                    }
                }
        }
        if (++mx == im.mcus_x) { mx = 0; ++my; }
    }
    return 0;
}

// ---- inverse DCT: libjpeg's integer "slow" method (13-bit constants, first pass keeps 2 extra bits) -----------------
FTF_HD void idct_1d(const int32_t in0, const int32_t in1, const int32_t in2, const int32_t in3, const int32_t in4,
                    const int32_t in5, const int32_t in6, const int32_t in7, int32_t (&o)[8]) {
    int32_t z1 = (in2 + in6) * 4433;
    const int32_t tmp2 = z1 + in6 * (-15137);
    const int32_t tmp3 = z1 + in2 * 6270;
    const int32_t tmp0 = (int32_t)((uint32_t)(in0 + in4) << 13);
    const int32_t tmp1 = (int32_t)((uint32_t)(in0 - in4) << 13);
    const int32_t e0 = tmp0 + tmp3, e3 = tmp0 - tmp3, e1 = tmp1 + tmp2, e2 = tmp1 - tmp2;
    int32_t t0 = in7, t1 = in5, t2 = in3, t3 = in1;
    z1 = t0 + t3;
    int32_t z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
    const int32_t z5 = (z3 + z4) * 9633;
    t0 *= 2446; t1 *= 16819; t2 *= 25172; t3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
    o[0] = e0 + t3; o[7] = e0 - t3; o[1] = e1 + t2; o[6] = e1 - t2;
    o[2] = e2 + t1; o[5] = e2 - t1; o[3] = e3 + t0; o[4] = e3 - t0;
}

FTF_HD uint32_t idct_range_limit(int32_t x) {
    // libjpeg's post-IDCT table: the value is taken modulo 1024 around the level shift of 128
    x &= 1023;
    return x < 128 ? (uint32_t)(x + 128) : x < 512 ? 255u : x < 896 ? 0u : (uint32_t)(x - 896);
}

// One 8x8 block: coef (row-major int16), q (row-major u16) -> 8 rows of 8 samples at out (row stride `stride`, 8-byte
// aligned).
FTF_HD void idct_block(const int16_t* __restrict__ coef, const uint16_t* __restrict__ q, uint8_t* __restrict__ out,
                       int stride) {
    int32_t ws[64];
    int32_t in[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) in[i] = (int32_t)coef[i] * (int32_t)q[i];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int32_t o[8];
        idct_1d(in[c], in[8 + c], in[16 + c], in[24 + c], in[32 + c], in[40 + c], in[48 + c], in[56 + c], o);
#pragma unroll
        for (int r = 0; r < 8; ++r) ws[r * 8 + c] = (o[r] + (1 << 10)) >> 11;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int32_t o[8];
        idct_1d(ws[r * 8], ws[r * 8 + 1], ws[r * 8 + 2], ws[r * 8 + 3], ws[r * 8 + 4], ws[r * 8 + 5], ws[r * 8 + 6],
                ws[r * 8 + 7], o);
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            lo |= idct_range_limit((o[c] + (1 << 17)) >> 18) << (8 * c);
            hi |= idct_range_limit((o[4 + c] + (1 << 17)) >> 18) << (8 * c);
        }
        uint32_t* row = reinterpret_cast<uint32_t*>(out + (int64_t)r * stride);
        row[0] = lo;
        row[1] = hi;
    }
}

// ---- chroma upsampling (libjpeg's triangle filters) and colour conversion ---------------------------------------------
// Sample of a chroma plane at full-resolution position (y, x).  pl: the component's plane, st: its row stride.
FTF_HD int chroma_at(const uint8_t* __restrict__ pl, int st, int dw, int dh, int hs, int vs, int y, int x) {
    if (hs == 1) return pl[(int64_t)y * st + x];
    const int cx = x >> 1;
    if (vs == 1) {
        const uint8_t* r = pl + (int64_t)y * st;
        if (dw <= 2) return r[cx];
        if (x & 1) return cx == dw - 1 ? r[cx] : (3 * r[cx] + r[cx + 1] + 2) >> 2;
        return cx == 0 ? r[0] : (3 * r[cx] + r[cx - 1] + 1) >> 2;
    }
    const int cy = y >> 1;
    if (dw <= 2) return pl[(int64_t)cy * st + cx];
    int ny = (y & 1) ? cy + 1 : cy - 1;
    ny = ny < 0 ? 0 : ny > dh - 1 ? dh - 1 : ny;
    const uint8_t* r0 = pl + (int64_t)cy * st;
    const uint8_t* r1 = pl + (int64_t)ny * st;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) return cx == dw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    return cx == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

FTF_HD uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

// libjpeg's 16-bit fixed-point YCbCr -> RGB (its four lookup tables written out as arithmetic)
FTF_HD void ycc_to_rgb(int y, int cb, int cr, uint8_t* rgb) {
    cb -= 128; cr -= 128;
    rgb[0] = clamp_u8(y + ((91881 * cr + 32768) >> 16));
    rgb[1] = clamp_u8(y + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
    rgb[2] = clamp_u8(y + ((116130 * cb + 32768) >> 16));
}

// One output pixel of image `im` at (y, x).
FTF_HD void pixel_rgb(const ImageDesc& im, const uint8_t* __restrict__ planes, int y, int x, uint8_t* rgb) {
    const int yy = planes[im.plane_off[0] + (int64_t)y * (im.bw[0] * 8) + x];
    if (im.ncomp == 1) { rgb[0] = rgb[1] = rgb[2] = (uint8_t)yy; return; }
    const int cb = chroma_at(planes + im.plane_off[1], im.bw[1] * 8, im.dw[1], im.dh[1], im.hs, im.vs, y, x);
    const int cr = chroma_at(planes + im.plane_off[2], im.bw[2] * 8, im.dw[2], im.dh[2], im.hs, im.vs, y, x);
    ycc_to_rgb(yy, cb, cr, rgb);
}

}  // namespace jpeg
}  // namespace ftf
