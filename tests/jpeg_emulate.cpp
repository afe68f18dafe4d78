// Test harness (CPU): compiles the per-thread functions of the JPEG kernels (csrc/jpeg_core.h) and the host parser
// (csrc/jpeg_parse.cpp) with g++ and runs them "thread" by "thread" in plain loops, so that the kernel logic can be
// checked against the oracle on a machine without a GPU.  Built and used by tests/test_jpeg_cpu.py only; it is not
// part of libftf_b200.so and nothing in the library calls it.
#include "../flink-tensorflow_b200/csrc/jpeg_parse.cpp"
#include <algorithm>
#include <vector>

using namespace ftf::jpeg;

static int g_rounds = 0;
extern "C" int emu_last_rounds() { return g_rounds; }      // correction rounds of the last decode (all segments)

// threads: chunks per wave (the CTA size of J1a); 0 = 1024
extern "C" int emu_jpeg_decode(const uint8_t* file, int64_t len, uint8_t* rgb, int32_t* n_segments, int32_t threads) {
    Parsed* p = new Parsed();
    std::string err;
    int rc = parse_headers(file, len, p, &err);
    if (rc) { delete p; return rc; }
    ImageDesc im;
    fill_geometry(*p, &im);
    std::vector<HuffTable> tables;
    std::vector<uint16_t> quant;
    int64_t coef_elems = 0, plane_bytes = 0;
    for (int c = 0; c < p->ncomp; ++c) {
        im.quant[c] = c;
        quant.insert(quant.end(), p->quant[p->tq[c]], p->quant[p->tq[c]] + 64);
        im.dc[c] = (int)tables.size(); tables.emplace_back(); build_table(p->dc[p->td[c]], &tables.back());
        im.ac[c] = (int)tables.size(); tables.emplace_back(); build_table(p->ac[p->ta[c]], &tables.back());
        im.coef_off[c] = coef_elems; coef_elems += (int64_t)im.bw[c] * im.bh[c] * 64;
        im.plane_off[c] = plane_bytes; plane_bytes += (int64_t)im.bw[c] * im.bh[c] * 64;
        im.dc_off[c] = im.coef_off[c] / 64;
    }
    im.rgb_off = 0;
    std::vector<Segment> segs;
    std::vector<uint8_t> scan((size_t)stage_bound(len, *p, im) + 16);
    bool truncated = false;
    stage_scan(file, len, *p, im, 0, scan.data(), 0, &segs, &truncated);
    if (truncated) { delete p; return 3; }
    if (n_segments) *n_segments = (int)segs.size();
    std::vector<int16_t> coef((size_t)coef_elems, 0), dcv((size_t)coef_elems / 64, 0);
    std::vector<uint8_t> planes((size_t)plane_bytes + 16);
    int bad = 0;
    const int bpm = im.ncomp == 1 ? 1 : im.hs * im.vs + 2;
    g_rounds = 0;
    for (const Segment& sg : segs) {                                  // J1a: one CTA per segment, `threads` chunks per wave
        const int T = threads > 0 ? threads : 1024, total_blocks = sg.n_mcus * bpm;
        ScanState wave_entry{0, 0, 0};
        int blocks_before = 0;
        for (int c0 = 0; c0 < sg.n_chunks; c0 += T) {
            const int m = std::min(T, sg.n_chunks - c0);
            std::vector<ScanState> in((size_t)m), out((size_t)m);
            std::vector<int> n((size_t)m, 0);
            std::vector<ScanState> cps((size_t)m * kCheckpoints);              // [kCheckpoints][m], as in shared memory
            std::vector<uint16_t> cpn((size_t)m * kCheckpoints);
            auto cp_of = [&](int t) { return Checkpoints{cps.data() + t, cpn.data() + t, m}; };
            for (int t = 0; t < m; ++t) {                             // 1. guess
                in[t] = ScanState{(uint32_t)(c0 + t) * kChunkBits, 0, 0};
                if (t == 0) in[t] = wave_entry;
                out[t] = decode_chunk<kGuess>(sg, im, tables.data(), scan.data(), in[t], (uint32_t)(c0 + t + 1) * kChunkBits, 0, 0,
                                              nullptr, nullptr, &n[t], nullptr, cp_of(t));
            }
            for (;;) {                                                // 2. correct along the chain, all threads in lock step
                std::vector<ScanState> prev(out);
                bool any = false;
                for (int t = 1; t < m; ++t)
                    if (!same_state(prev[t - 1], in[t])) {
                        any = true;
                        in[t] = prev[t - 1];
                        out[t] = decode_chunk<kCorrect>(sg, im, tables.data(), scan.data(), in[t], (uint32_t)(c0 + t + 1) * kChunkBits, 0, 0,
                                                        nullptr, nullptr, &n[t], nullptr, cp_of(t), out[t]);
                    }
                if (!any) break;
                ++g_rounds;
            }
            int ordinal = blocks_before;                              // 3. prefix sum + the storing pass
            for (int t = 0; t < m; ++t) {
                if (ordinal < total_blocks) {
                    int n2 = 0;
                    const bool last = c0 + t == sg.n_chunks - 1;
                    decode_chunk<kStore>(sg, im, tables.data(), scan.data(), in[t], last ? 0xffffffffu : (uint32_t)(c0 + t + 1) * kChunkBits,
                                       ordinal, total_blocks, coef.data(), dcv.data(), &n2, &bad);
                }
                ordinal += n[t];
            }
            wave_entry = out[(size_t)m - 1];
            blocks_before = ordinal;
        }
        for (int c = 0; c < im.ncomp; ++c) {                          // J1b: inclusive scan of the segment's DC stretch
            int64_t first;
            int count, carry = 0;
            dc_range(sg, im, c, &first, &count);
            for (int t = 0; t < count; ++t) { carry += dcv[(size_t)first + t]; dcv[(size_t)first + t] = (int16_t)carry; }
        }
    }
    for (int c = 0; c < im.ncomp; ++c)                                                              // J2
        for (int b = 0; b < im.bw[c] * im.bh[c]; ++b) {
            const int by = b / im.bw[c], bx = b % im.bw[c];
            coef[(size_t)(im.coef_off[c] + (int64_t)b * 64)] = dc_of_block(im, dcv.data(), c, by, bx);
            idct_block(coef.data() + im.coef_off[c] + (int64_t)b * 64, quant.data() + im.quant[c] * 64,
                       planes.data() + im.plane_off[c] + ((int64_t)by * 8 * im.bw[c] + bx) * 8, im.bw[c] * 8);
        }
    for (int y = 0; y < im.height; ++y)                                                             // J3: 8 pixels per thread
        for (int x0 = 0; x0 < im.width; x0 += 8) {
            uint8_t px[24];
            pixels8_rgb(im, planes.data(), y, x0, px);
            const int nx = std::min(8, im.width - x0);
            memcpy(rgb + ((int64_t)y * im.width + x0) * 3, px, (size_t)nx * 3);
            uint8_t one[3];                                          // the per-pixel form must agree
            for (int i = 0; i < nx; ++i) {
                pixel_rgb(im, planes.data(), y, x0 + i, one);
                if (memcmp(one, px + 3 * i, 3)) bad |= 2;
            }
        }
    delete p;
    return bad ? 3 : 0;
}
