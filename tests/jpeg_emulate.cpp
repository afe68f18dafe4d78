// Test harness (CPU): compiles the per-thread functions of the JPEG kernels (csrc/jpeg_core.h) and the host parser
// (csrc/jpeg_parse.cpp) with g++ and runs them "thread" by "thread" in plain loops, so that the kernel logic can be
// checked against the oracle on a machine without a GPU.  Built and used by tests/test_jpeg_cpu.py only; it is not
// part of libftf_b200.so and nothing in the library calls it.
#include "../flink-tensorflow_b200/csrc/jpeg_parse.cpp"
#include <vector>

using namespace ftf::jpeg;

extern "C" int emu_jpeg_decode(const uint8_t* file, int64_t len, uint8_t* rgb, int32_t* n_segments) {
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
        im.dc[c] = (int)tables.size(); tables.push_back(p->dc[p->td[c]]);
        im.ac[c] = (int)tables.size(); tables.push_back(p->ac[p->ta[c]]);
        im.coef_off[c] = coef_elems; coef_elems += (int64_t)im.bw[c] * im.bh[c] * 64;
        im.plane_off[c] = plane_bytes; plane_bytes += (int64_t)im.bw[c] * im.bh[c] * 64;
    }
    im.rgb_off = 0;
    std::vector<Segment> segs;
    find_segments(file, len, *p, im, 0, &segs);
    if (n_segments) *n_segments = (int)segs.size();
    std::vector<int16_t> coef((size_t)coef_elems, 0);
    std::vector<uint8_t> planes((size_t)plane_bytes + 8);
    int bad = 0;
    for (const Segment& s : segs) bad |= decode_segment(s, im, tables.data(), file, coef.data());   // J1, one thread each
    for (int c = 0; c < im.ncomp; ++c)                                                              // J2
        for (int b = 0; b < im.bw[c] * im.bh[c]; ++b) {
            const int by = b / im.bw[c], bx = b % im.bw[c];
            idct_block(coef.data() + im.coef_off[c] + (int64_t)b * 64, quant.data() + im.quant[c] * 64,
                       planes.data() + im.plane_off[c] + ((int64_t)by * 8 * im.bw[c] + bx) * 8, im.bw[c] * 8);
        }
    for (int y = 0; y < im.height; ++y)                                                             // J3
        for (int x = 0; x < im.width; ++x) pixel_rgb(im, planes.data(), y, x, rgb + ((int64_t)y * im.width + x) * 3);
    delete p;
    return bad ? 3 : 0;
}
