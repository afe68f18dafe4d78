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
// the code is longer than kLutBits; then the canonical-code walk of T.81 F.2.2.3 over maxcode / valoff.  (Measured: 12-bit
// tables, 8 KB each, lose more to L1 misses than they save on the walk; 9 bits keep all of a batch's tables L1-resident.)
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
    int64_t dc_off[3];            // int16 elements of the batch's DC array: the component's blocks in scan order
    int64_t rgb_off;              // bytes
};

#ifndef FTF_JPEG_CHUNK_BITS
#define FTF_JPEG_CHUNK_BITS 1024
#endif
#ifndef FTF_JPEG_CHECKPOINT_BITS
#define FTF_JPEG_CHECKPOINT_BITS 256      /* measured: 256 (three checkpoints) beats 128 and 512 by 1-2 % */
#endif
constexpr int kChunkBits = FTF_JPEG_CHUNK_BITS;   // the unit of parallel entropy decoding (see decode_chunk)

// One independently decodable stretch of a scan: a restart interval, or the whole scan.  The host stages its bytes with
// the 0xFF00 byte stuffing removed ("clean" stream), 4-byte aligned and followed by at least 16 zero bytes.
struct Segment {
    int32_t image;
    int32_t first_mcu, n_mcus;
    int32_t n_chunks;             // max(1, ceil(nbits / kChunkBits))
    int64_t begin;                // byte offset of the clean stream in the batch's scan buffer
    int64_t nbits;                // bits of clean stream; past them the decoder reads zeros, as libjpeg does at a marker
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

#ifdef __CUDA_ARCH__
#define FTF_BE32(x) __byte_perm((x), 0, 0x0123)
#else
#define FTF_BE32(x) __builtin_bswap32(x)
#endif

// MSB-first view of a clean stream: 64 bits starting at a 32-bit boundary are kept in a register pair and slide forward
// one aligned word at a time, so a symbol costs one load per 32 bits consumed.
struct BitWindow {
    const uint32_t* words;
    uint64_t buf;
    uint32_t base;      // bit position of buf's top bit (a multiple of 32)
    uint32_t nbits;
    FTF_HD void seek(const uint32_t* w, uint32_t nb, uint32_t pos) {
        words = w; nbits = nb; base = pos & ~31u; buf = 0;
        if (pos < nbits) buf = ((uint64_t)FTF_BE32(FTF_LDG(words + (base >> 5))) << 32) | FTF_BE32(FTF_LDG(words + (base >> 5) + 1));
    }
    // the 32 bits that start at pos (pos never goes backwards and advances by less than 32 between calls)
    FTF_HD uint32_t at(uint32_t pos) {
        if (pos >= nbits) return 0;
        if (pos - base >= 32) {
            base += 32;
            buf = (buf << 32) | FTF_BE32(FTF_LDG(words + (base >> 5) + 1));
        }
        return (uint32_t)((buf << (pos - base)) >> 32);
    }
};

// T.81 F.2.2.3 DECODE on a 32-bit window: code length (0 = no such code) and symbol.
FTF_HD int lookup_long_code(uint32_t w, const HuffTable* __restrict__ t, int* sym);
// lut: the table's lookahead entries, possibly a copy in shared memory (hence a plain load, not a read-only-path one).
// The table itself is only needed for codes longer than the lookahead.
FTF_HD int lookup_symbol(uint32_t w, const uint16_t* lut, const HuffTable* __restrict__ t, int* sym) {
    const uint32_t e = lut[w >> (32 - kLutBits)];
    if (e) { *sym = (int)(e & 0xFF); return (int)(e >> 8); }
    return lookup_long_code(w, t, sym);
}
FTF_HD int lookup_long_code(uint32_t w, const HuffTable* __restrict__ t, int* sym) {
    const int32_t code16 = (int32_t)(w >> 16);
    int l = kLutBits + 1;
    while (l <= 16 && (code16 >> (16 - l)) > FTF_LDG(&t->maxcode[l])) ++l;
    if (l > 16) return 0;
    *sym = (int)FTF_LDG(&t->vals[(code16 >> (16 - l)) + FTF_LDG(&t->valoff[l])]);
    return l;
}

// T.81 F.2.2.1 EXTEND of the s (1..15) bits that follow the first len bits of the window
FTF_HD int extend_bits(uint32_t w, int len, int s) {
    const int v = (int)((w << len) >> (32 - s));
    return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
}

// Where an entropy decoder stands between two symbols: the bit it reads next, the zig-zag index it decodes next
// (0: the block's DC symbol) and which block of the MCU it is in.
struct ScanState {
    uint32_t pos;
    uint16_t k, b;
};
FTF_HD bool same_state(const ScanState& a, const ScanState& b) { return a.pos == b.pos && a.k == b.k && a.b == b.b; }

// Component and position inside the MCU of block b (luma blocks first, row-major; then Cb, Cr).
FTF_HD int16_t* block_ptr(const ImageDesc& im, int16_t* __restrict__ coef, int my, int mx, int b) {
    const int nluma = im.hs * im.vs;
    if (b < nluma) {
        const int v = b / im.hs, h = b - v * im.hs;
        return coef + im.coef_off[0] + ((int64_t)(my * im.vs + v) * im.bw[0] + (mx * im.hs + h)) * 64;
    }
    const int c = b - nluma + 1;
    return coef + im.coef_off[c] + ((int64_t)my * im.bw[c] + mx) * 64;
}

// DC terms live apart from the other coefficients, per component in scan order (MCU by MCU, luma blocks row-major inside
// the MCU), so that turning differences into values is a scan over contiguous memory.
FTF_HD int dc_index(const ImageDesc& im, int mcu, int b) {
    const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
    return b < nluma ? mcu * nluma + b : mcu;
}
FTF_HD int16_t* dc_slot(const ImageDesc& im, int16_t* __restrict__ dcv, int mcu, int b) {
    const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
    const int c = b < nluma ? 0 : b - nluma + 1;
    return dcv + im.dc_off[c] + dc_index(im, mcu, b);
}
// the DC value of block (by, bx) of component c, for the inverse DCT
FTF_HD int16_t dc_of_block(const ImageDesc& im, const int16_t* __restrict__ dcv, int c, int by, int bx) {
    const int ch = c == 0 ? im.hs : 1, cv = c == 0 ? im.vs : 1;
    const int my = by / cv, v = by - my * cv, mx = bx / ch, h = bx - mx * ch;
    return dcv[im.dc_off[c] + (int64_t)(my * im.mcus_x + mx) * (ch * cv) + v * ch + h];
}

// Parallel entropy decoding.  A Huffman stream has no index, but a decoder started at a wrong place falls into step with
// the true one after a while (one to two kilobits on photographs: it also has to find its place inside the MCU), so a segment is cut into chunks of kChunkBits and every chunk gets a thread:
//   1. each thread decodes its chunk from a guessed state (the chunk's first bit, start of an MCU) and publishes the state
//      it leaves the chunk in -- chunk 0 starts from the true state;
//   2. a thread whose predecessor published a state other than the one it started from decodes again from that state;
//      repeated until nothing changes, at which point every entry state is the true one (by induction from chunk 0);
//   3. a prefix sum over the blocks completed per chunk tells every thread which block it starts in, and a last pass
//      decodes once more and stores the coefficients.  DC values are stored as differences; dc_* below add them up.
// decode_chunk is one such pass over the symbols that start in [st.pos, limit).  kWrite = false: counts completed blocks
// and returns the exit state; a bit pattern that is no code (possible only from a wrong state or in a corrupt file) is
// skipped one bit at a time.  kWrite = true: also stores coefficients for blocks [ordinal, total_blocks) of the segment,
// (DC differences go to the DC array, see dc_slot) and reports a pattern that is no code, or a run past the 64th
// coefficient, through *bad.
// Correction passes do not always need the whole chunk: once the re-started decoder meets the path of the previous attempt, everything (exit state, blocks completed) is already known.  Every attempt
// therefore leaves its state at kCheckpoints positions of the chunk (the first symbol at or after every kCheckpointBits),
// and a correction pass stops at the first checkpoint where it finds itself in the recorded state.
constexpr int kCheckpointBits = FTF_JPEG_CHECKPOINT_BITS;
constexpr int kCheckpoints = kChunkBits / kCheckpointBits - 1;
struct Checkpoints {
    ScanState* state;     // [kCheckpoints], element i at state[i * stride]
    uint16_t* done;       // blocks completed before the checkpoint
    int stride;
};
enum { kGuess = 0, kCorrect = 1, kStore = 2 };

// kMode kGuess / kCorrect: the counting passes (record / compare-and-record checkpoints; kCorrect takes the previous
// attempt's exit state and block count in *exit_io / *n_done and may return them adjusted).  kStore: the storing pass.
template <int kMode>
FTF_HD ScanState decode_chunk(const Segment& sg, const ImageDesc& im, const HuffTable* __restrict__ tables,
                              const uint8_t* __restrict__ scan, ScanState st, uint32_t limit, int ordinal, int total_blocks,
                              int16_t* __restrict__ coef, int16_t* __restrict__ dcv, int* n_done, int* bad,
                              Checkpoints cp = Checkpoints{nullptr, nullptr, 0}, ScanState old_exit = ScanState{0, 0, 0},
                              const uint16_t* luts = nullptr) {      // luts: [component][dc, ac][1 << kLutBits] copies, or null
    constexpr bool kWrite = kMode == kStore;
    const int bpm = im.ncomp == 1 ? 1 : im.hs * im.vs + 2;
    const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
    BitWindow bw;
    bw.seek(reinterpret_cast<const uint32_t*>(scan + sg.begin), (uint32_t)sg.nbits, st.pos);
    uint32_t pos = st.pos;
    int k = st.k, b = st.b, done = 0;
    uint32_t next_cp = limit - kChunkBits + kCheckpointBits;      // counting passes only
    int ci = 0;
    int mx = 0, my = 0;
    int16_t* blk = nullptr;
    int16_t* dcp = nullptr;
    if (kWrite) {
        const int mcu = sg.first_mcu + ordinal / bpm;
        my = mcu / im.mcus_x; mx = mcu - my * im.mcus_x;
        blk = block_ptr(im, coef, my, mx, b);
        dcp = dc_slot(im, dcv, mcu, b);
    }
    int c = b < nluma ? 0 : b - nluma + 1;
    // between blocks only the component index is carried: with the lookahead tables in shared memory a symbol needs that
    // index alone, and the table pointers (one per component and kind) are looked up by the rare long code
    while (pos < limit && (!kWrite || ordinal < total_blocks)) {
        const uint32_t w = bw.at(pos);
        const bool is_dc = k == 0;
        int sym = 0;
        int len;
        if (luts) {
            const uint32_t e = luts[((2 * c + (is_dc ? 0 : 1)) << kLutBits) + (w >> (32 - kLutBits))];
            if (e) { sym = (int)(e & 0xFF); len = (int)(e >> 8); }
            else len = lookup_long_code(w, tables + (is_dc ? im.dc[c] : im.ac[c]), &sym);
        } else {
            const HuffTable* t = tables + (is_dc ? im.dc[c] : im.ac[c]);
            len = lookup_symbol(w, t->lut, t, &sym);
        }
        if (len == 0 || (is_dc && sym > 15)) {          // no such code / a DC symbol is a bit count
            if (kWrite) { *bad = 1; break; }
            pos += 1;
            continue;
        }
        // one path for both kinds of symbol: a DC symbol is a run of 0 and a size, landing on index 0
        const int r = is_dc ? 0 : sym >> 4, s = is_dc ? sym : sym & 15;
        if (s != 0 || is_dc) {
            k += r;
            if (k > 63) {
                if (kWrite) { *bad = 1; break; }
                k = 63;
            }
            if (kWrite) {
                int16_t* dst = is_dc ? dcp : blk + natural_order(k);
                *dst = (int16_t)(s ? extend_bits(w, len, s) : 0);
            }
            pos += (uint32_t)(len + s);
            k += 1;
        } else {
            pos += (uint32_t)len;
            k = r == 15 ? k + 16 : 64;
        }
        if (k >= 64) {                      // block complete
            k = 0; ++done; ++ordinal;
            if (++b == bpm) {
                b = 0;
                if (kWrite && ++mx == im.mcus_x) { mx = 0; ++my; }
            }
            c = b < nluma ? 0 : b - nluma + 1;
            if (kWrite) {
                blk = block_ptr(im, coef, my, mx, b);
                dcp = dc_slot(im, dcv, my * im.mcus_x + mx, b);
            }
        }
        if (!kWrite && pos >= next_cp && ci < kCheckpoints) {
            ScanState cur;
            cur.pos = pos; cur.k = (uint16_t)k; cur.b = (uint16_t)b;
            if (kMode == kCorrect && same_state(cur, cp.state[ci * cp.stride])) {
                // back on the previous attempt's path: its remainder stands, only the block counts shift
                const int delta = done - (int)cp.done[ci * cp.stride];
                for (int i = ci; i < kCheckpoints; ++i) cp.done[i * cp.stride] = (uint16_t)((int)cp.done[i * cp.stride] + delta);
                *n_done += delta;
                return old_exit;
            }
            cp.state[ci * cp.stride] = cur;
            cp.done[ci * cp.stride] = (uint16_t)done;
            ++ci;
            next_cp += kCheckpointBits;
        }
    }
    *n_done = done;
    ScanState out;
    out.pos = pos; out.k = (uint16_t)k; out.b = (uint16_t)b;
    return out;
}

// DC prediction (T.81 F.2.2.1: DC = previous DC of the component + difference, reset at every restart): an in-place
// inclusive scan over the component's stretch of the DC array that belongs to one segment.  dc_range gives the stretch.
FTF_HD void dc_range(const Segment& sg, const ImageDesc& im, int c, int64_t* first, int* count) {
    const int per = c == 0 ? (im.ncomp == 1 ? 1 : im.hs * im.vs) : 1;
    *first = im.dc_off[c] + (int64_t)sg.first_mcu * per;
    *count = sg.n_mcus * per;
}

// ---- inverse DCT: libjpeg's integer "slow" method (13-bit constants, first pass keeps 2 extra bits) -----------------
// The sums and products are formed on unsigned words: a damaged file can hold coefficient x quantiser products that do not
// fit 32 bits (libjpeg computes in 64), and the result for such a file should be garbage, not undefined behaviour; for every
// value a real image produces the arithmetic is the same.
FTF_HD void idct_1d(const int32_t in0, const int32_t in1, const int32_t in2, const int32_t in3, const int32_t in4,
                    const int32_t in5, const int32_t in6, const int32_t in7, int32_t (&o)[8]) {
    typedef uint32_t u32;
    const u32 x0 = (u32)in0, x1 = (u32)in1, x2 = (u32)in2, x3 = (u32)in3, x4 = (u32)in4, x5 = (u32)in5, x6 = (u32)in6, x7 = (u32)in7;
    u32 z1 = (x2 + x6) * 4433u;
    const u32 tmp2 = z1 + x6 * (u32)(-15137);
    const u32 tmp3 = z1 + x2 * 6270u;
    const u32 tmp0 = (x0 + x4) << 13;
    const u32 tmp1 = (x0 - x4) << 13;
    const u32 e0 = tmp0 + tmp3, e3 = tmp0 - tmp3, e1 = tmp1 + tmp2, e2 = tmp1 - tmp2;
    u32 t0 = x7, t1 = x5, t2 = x3, t3 = x1;
    z1 = t0 + t3;
    u32 z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
    const u32 z5 = (z3 + z4) * 9633u;
    t0 *= 2446u; t1 *= 16819u; t2 *= 25172u; t3 *= 12299u;
    z1 *= (u32)(-7373); z2 *= (u32)(-20995); z3 *= (u32)(-16069); z4 *= (u32)(-3196);
    z3 += z5; z4 += z5;
    t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
    o[0] = (int32_t)(e0 + t3); o[7] = (int32_t)(e0 - t3); o[1] = (int32_t)(e1 + t2); o[6] = (int32_t)(e1 - t2);
    o[2] = (int32_t)(e2 + t1); o[5] = (int32_t)(e2 - t1); o[3] = (int32_t)(e3 + t0); o[4] = (int32_t)(e3 - t0);
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
    for (int i = 0; i < 64; ++i) in[i] = (int32_t)coef[i] * (int32_t)q[i];      // 16 x 16 bits: always fits
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int32_t o[8];
        idct_1d(in[c], in[8 + c], in[16 + c], in[24 + c], in[32 + c], in[40 + c], in[48 + c], in[56 + c], o);
#pragma unroll
        for (int r = 0; r < 8; ++r) ws[r * 8 + c] = (int32_t)((uint32_t)o[r] + (1u << 10)) >> 11;
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        int32_t o[8];
        idct_1d(ws[r * 8], ws[r * 8 + 1], ws[r * 8 + 2], ws[r * 8 + 3], ws[r * 8 + 4], ws[r * 8 + 5], ws[r * 8 + 6],
                ws[r * 8 + 7], o);
        uint32_t lo = 0, hi = 0;
#ifdef __CUDA_ARCH__
        // idct_range_limit without its select chain: the table depends on x modulo 1024 only, and for the representative
        // y of x in [-512, 511] it is clamp(y + 128, 0, 255).  y + 128 = (((x + 512) & 1023) - 384), the +512 rides on the
        // descale's rounding constant, and the clamp is the saturation of the pack instruction.
        int v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c)
            v[c] = (int)((((int32_t)((uint32_t)o[c] + ((1u << 17) + (512u << 18)))) >> 18) & 1023) - 384;
        uint32_t t;
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[3]), "r"(v[2]), "r"(0));
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(lo) : "r"(v[1]), "r"(v[0]), "r"(t));
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v[7]), "r"(v[6]), "r"(0));
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(v[5]), "r"(v[4]), "r"(t));
#else
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            lo |= idct_range_limit((int32_t)((uint32_t)o[c] + (1u << 17)) >> 18) << (8 * c);
            hi |= idct_range_limit((int32_t)((uint32_t)o[4 + c] + (1u << 17)) >> 18) << (8 * c);
        }
#endif
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

FTF_HD uint32_t load_u32(const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); }

// The chroma samples of the 8 pixels (y, x0 .. x0+7), x0 a multiple of 8: chroma_at for each of them, with the plane
// read as aligned words and every source sample fetched once.  The triangle filters replicate the edge sample, so the
// neighbours outside [0, dw) are the clamped ones.
template <int hs, int vs>
FTF_HD void chroma8(const uint8_t* __restrict__ pl, int st, int dw, int dh, int y, int x0, int (&o)[8]) {
    if (hs == 1) {
        const uint8_t* r = pl + (int64_t)y * st + x0;
        const uint32_t a = load_u32(r), b = load_u32(r + 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) { o[i] = (int)((a >> (8 * i)) & 0xFF); o[4 + i] = (int)((b >> (8 * i)) & 0xFF); }
        return;
    }
    if (dw <= 2) {
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = chroma_at(pl, st, dw, dh, hs, vs, y, x0 + i);
        return;
    }
    const int cx0 = x0 >> 1;
    const int cy = vs == 2 ? y >> 1 : y;
    int ny = cy;
    if (vs == 2) {
        ny = (y & 1) ? cy + 1 : cy - 1;
        ny = ny < 0 ? 0 : ny > dh - 1 ? dh - 1 : ny;
    }
    const uint8_t* r0 = pl + (int64_t)cy * st;
    const uint8_t* r1 = pl + (int64_t)ny * st;
    int s[6];                                // column terms of samples cx0-1 .. cx0+4
    if (cx0 + 3 <= dw - 1) {
        // the four samples in the middle as one word; the two outer neighbours by clamped index, without a branch: the
        // first and the last thread of every row would otherwise drag their whole warps through the byte-wise path
        const uint32_t a = load_u32(r0 + cx0), b = load_u32(r1 + cx0);
        const int cl = cx0 >= 1 ? cx0 - 1 : 0, cr = cx0 + 4 <= dw - 1 ? cx0 + 4 : dw - 1;
        if (vs == 2) {
#pragma unroll
            for (int i = 0; i < 4; ++i) s[1 + i] = 3 * (int)((a >> (8 * i)) & 0xFF) + (int)((b >> (8 * i)) & 0xFF);
            s[0] = 3 * r0[cl] + r1[cl];
            s[5] = 3 * r0[cr] + r1[cr];
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) s[1 + i] = (int)((a >> (8 * i)) & 0xFF);
            s[0] = r0[cl];
            s[5] = r0[cr];
        }
    } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            int cx = cx0 - 1 + i;
            cx = cx < 0 ? 0 : cx > dw - 1 ? dw - 1 : cx;
            s[i] = vs == 2 ? 3 * r0[cx] + r1[cx] : (int)r0[cx];
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (vs == 2) {
            o[2 * i] = (3 * s[1 + i] + s[i] + 8) >> 4;
            o[2 * i + 1] = (3 * s[1 + i] + s[2 + i] + 7) >> 4;
        } else {
            o[2 * i] = (3 * s[1 + i] + s[i] + 1) >> 2;
            o[2 * i + 1] = (3 * s[1 + i] + s[2 + i] + 2) >> 2;
        }
    }
}

// The 8 pixels (y, x0 .. x0+7) of image `im`, x0 a multiple of 8, as 24 RGB bytes (pixels past the width are computed
// from the planes' padding and dropped by the caller).  One instance per sampling layout (hs x vs luma samples per chroma
// sample; 0 x 0: one component), so that the layout is not a question asked per sample.
template <int hs, int vs>
FTF_HD void pixels8_rgb_t(const ImageDesc& im, const uint8_t* __restrict__ planes, int y, int x0, uint8_t* rgb) {
    const uint8_t* yp = planes + im.plane_off[0] + (int64_t)y * (im.bw[0] * 8) + x0;
    const uint32_t ya = load_u32(yp), yb = load_u32(yp + 4);
    if (hs == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint8_t v = (uint8_t)(((i < 4 ? ya : yb) >> (8 * (i & 3))) & 0xFF);
            rgb[3 * i] = rgb[3 * i + 1] = rgb[3 * i + 2] = v;
        }
        return;
    }
    int cb[8], cr[8];
    constexpr int chs = hs == 0 ? 1 : hs, cvs = vs == 0 ? 1 : vs;
    chroma8<chs, cvs>(planes + im.plane_off[1], im.bw[1] * 8, im.dw[1], im.dh[1], y, x0, cb);
    chroma8<chs, cvs>(planes + im.plane_off[2], im.bw[2] * 8, im.dw[2], im.dh[2], y, x0, cr);
#ifdef __CUDA_ARCH__
    // The same arithmetic as ycc_to_rgb with fewer instructions: y enters as y << 16 (one byte permute), the -128 level
    // shifts and the rounding term are constants added once, the >> 16 then yields y + (term >> 16) exactly because
    // y << 16 is a multiple of 65536, and the clamps to 0..255 are the saturation of the pack instruction.
    int v[24];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint32_t yw = i < 4 ? ya : yb;
        const int yk = (int)__byte_perm(yw, 0, 0x4044 | ((i & 3) << 8));          // luma byte i in bits 16..23
        v[3 * i] = (91881 * cr[i] + (yk + (32768 - 91881 * 128))) >> 16;
        v[3 * i + 1] = (-22554 * cb[i] + (-46802 * cr[i] + (yk + (32768 + (22554 + 46802) * 128)))) >> 16;
        v[3 * i + 2] = (116130 * cb[i] + (yk + (32768 - 116130 * 128))) >> 16;
    }
    uint32_t* out32 = reinterpret_cast<uint32_t*>(rgb);       // callers pass a 4-byte aligned buffer of 24 bytes
#pragma unroll
    for (int w = 0; w < 6; ++w) {
        uint32_t hi, word;
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(v[4 * w + 3]), "r"(v[4 * w + 2]), "r"(0));
        asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(word) : "r"(v[4 * w + 1]), "r"(v[4 * w]), "r"(hi));
        out32[w] = word;
    }
#else
#pragma unroll
    for (int i = 0; i < 8; ++i) ycc_to_rgb((int)(((i < 4 ? ya : yb) >> (8 * (i & 3))) & 0xFF), cb[i], cr[i], rgb + 3 * i);
#endif
}

FTF_HD void pixels8_rgb(const ImageDesc& im, const uint8_t* __restrict__ planes, int y, int x0, uint8_t* rgb) {
    if (im.ncomp == 1) pixels8_rgb_t<0, 0>(im, planes, y, x0, rgb);
    else if (im.hs == 1) pixels8_rgb_t<1, 1>(im, planes, y, x0, rgb);
    else if (im.vs == 1) pixels8_rgb_t<2, 1>(im, planes, y, x0, rgb);
    else pixels8_rgb_t<2, 2>(im, planes, y, x0, rgb);
}

}  // namespace jpeg
}  // namespace ftf
