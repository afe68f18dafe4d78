/*
 * jpeg_oracle.c -- CPU restatement of DecodeJpeg(channels=3), the first node of the reference's normalisation graph
 * (EX/common/GraphBuilder.java:41-47, fed by EX/inception/ImageNormalization.scala:56-57 with the raw file bytes that
 * EX/inception/ImageInputFormat.scala:63-80 reads).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.c): used by tests/, smoke() and bench.py's CPU legs as the checker / timed CPU
 * stand-in, never by the product path.
 *
 * The arithmetic lives in libjpeg inside the un-vendored org.tensorflow:tensorflow:1.1.0 (pom.xml:54); the reference has
 * no JPEG fixture or test (flink-tensorflow-examples has no src/test).  This file restates baseline JPEG decoding as
 * published (ITU-T T.81: marker syntax, Huffman decoding procedure F.2.2) with libjpeg's default choices, which are
 * what TF's DecodeJpeg uses when no attribute overrides them: the integer "slow" inverse DCT (13-bit constants, two
 * passes), triangle-filter ("fancy") chroma upsampling for 2:1 horizontal and 2x2 subsampling, and the 16-bit
 * fixed-point YCbCr -> RGB tables.  Pin status: PINNED against libjpeg-turbo (through Pillow) by the fixtures under
 * tests/golden/jpeg (made by tests/golden/make_jpeg_golden.py); TF-1.1 itself cannot run here.
 *
 * Supported: 8-bit baseline / extended-sequential Huffman (SOF0/SOF1), 1 or 3 components, one interleaved scan,
 * sampling 1x1 (4:4:4), 2x1 (4:2:2), 2x2 (4:2:0), restart intervals.  Progressive and arithmetic files are refused.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { JO_OK = 0, JO_INVALID = 3, JO_UNIMPLEMENTED = 12 };

static const uint8_t k_natural[64] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
    62, 63};

typedef struct {
    int present;
    uint8_t bits[17];
    uint8_t vals[256];
    int32_t mincode[17], maxcode[18], valptr[17];
} jo_huff;

typedef struct {
    int id, h, v, tq, td, ta;
    int bw, bh;          /* block grid (padded to whole MCUs) */
    int dw, dh;          /* downsampled_width / height: the real samples */
    int16_t* coef;       /* [bh][bw][64], natural order */
    uint8_t* plane;      /* [bh*8][bw*8] */
} jo_comp;

typedef struct {
    const uint8_t* d;
    int64_t n, pos;
    int width, height, ncomp, hmax, vmax, mcus_x, mcus_y, restart;
    uint16_t qt[4][64];
    int qt_present[4];
    jo_huff dc[4], ac[4];
    jo_comp comp[3];
    /* bit reader */
    uint32_t bitbuf;
    int bitcnt, hit_marker, truncated;
} jo_dec;

static int u16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

static int build_huff(jo_huff* h) {
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        h->valptr[l] = k;
        h->mincode[l] = code;
        code += h->bits[l];
        k += h->bits[l];
        h->maxcode[l] = h->bits[l] ? code - 1 : -1;
        if (code > (1 << l)) return JO_INVALID;
        code <<= 1;
    }
    h->maxcode[17] = 0x7fffffff;
    if (k > 256) return JO_INVALID;
    h->present = 1;
    return JO_OK;
}

/* T.81 F.2.2.5 NEXTBIT, with byte stuffing (B.1.1.5); past a marker the stream reads as zeros (libjpeg's behaviour) */
static int get_bit(jo_dec* s) {
    if (s->bitcnt == 0) {
        int b = 0;
        if (!s->hit_marker && s->pos < s->n) {
            b = s->d[s->pos];
            if (b == 0xFF) {
                if (s->pos + 1 >= s->n) { s->truncated = 1; s->hit_marker = 1; b = 0; }
                else if (s->d[s->pos + 1] == 0) s->pos += 2;
                else { s->hit_marker = 1; b = 0; }
            } else s->pos += 1;
        } else if (!s->hit_marker) {
            s->truncated = 1;              /* the data ran out before any marker */
        }
        s->bitbuf = (uint32_t)b;
        s->bitcnt = 8;
    }
    s->bitcnt--;
    return (s->bitbuf >> s->bitcnt) & 1;
}

static int get_bits(jo_dec* s, int n) {
    int v = 0;
    while (n--) v = (v << 1) | get_bit(s);
    return v;
}

/* T.81 F.2.2.3 DECODE */
static int decode_sym(jo_dec* s, const jo_huff* h) {
    int code = get_bit(s), l = 1;
    while (l <= 16 && code > h->maxcode[l]) { code = (code << 1) | get_bit(s); ++l; }
    if (l > 16) return -1;
    return h->vals[h->valptr[l] + code - h->mincode[l]];
}

/* T.81 F.2.2.1 EXTEND */
static int extend(int v, int t) { return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v; }

static int decode_block(jo_dec* s, const jo_huff* dc, const jo_huff* ac, int* pred, int16_t* blk) {
    int t = decode_sym(s, dc);
    if (t < 0 || t > 16) return JO_INVALID;
    int diff = t ? extend(get_bits(s, t), t) : 0;
    *pred += diff;
    blk[0] = (int16_t)*pred;
    for (int k = 1; k < 64; ++k) {
        int rs = decode_sym(s, ac);
        if (rs < 0) return JO_INVALID;
        int r = rs >> 4, sz = rs & 15;
        if (sz) {
            k += r;
            int v = extend(get_bits(s, sz), sz);
            if (k > 63) return JO_INVALID;
            blk[k_natural[k]] = (int16_t)v;
        } else {
            if (r != 15) break;
            k += 15;
        }
    }
    return JO_OK;
}

static int parse_headers(jo_dec* s) {
    const uint8_t* d = s->d;
    int64_t n = s->n, p = 2;
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return JO_INVALID;
    int have_sof = 0;
    for (;;) {
        while (p < n && d[p] != 0xFF) ++p;          /* tolerate garbage between segments, as libjpeg's next_marker */
        while (p < n && d[p] == 0xFF) ++p;
        if (p >= n) return JO_INVALID;
        int m = d[p++];
        if (m == 0xD9) return JO_INVALID;            /* EOI before a scan */
        if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (p + 2 > n) return JO_INVALID;
        int len = u16(d + p);
        if (len < 2 || p + len > n) return JO_INVALID;
        const uint8_t* q = d + p + 2;
        int rem = len - 2;
        if (m == 0xDB) {
            while (rem > 0) {
                int pq = q[0] >> 4, tq = q[0] & 15;
                if (tq > 3 || pq > 1 || rem < 1 + 64 * (pq + 1)) return JO_INVALID;
                for (int i = 0; i < 64; ++i)
                    s->qt[tq][k_natural[i]] = (uint16_t)(pq ? u16(q + 1 + 2 * i) : q[1 + i]);
                s->qt_present[tq] = 1;
                q += 1 + 64 * (pq + 1); rem -= 1 + 64 * (pq + 1);
            }
        } else if (m == 0xC4) {
            while (rem > 0) {
                if (rem < 17) return JO_INVALID;
                int tc = q[0] >> 4, th = q[0] & 15;
                if (tc > 1 || th > 3) return JO_INVALID;
                jo_huff* h = tc ? &s->ac[th] : &s->dc[th];
                int cnt = 0;
                h->bits[0] = 0;
                for (int i = 1; i <= 16; ++i) { h->bits[i] = q[i]; cnt += q[i]; }
                if (cnt > 256 || rem < 17 + cnt) return JO_INVALID;
                memset(h->vals, 0, 256);
                memcpy(h->vals, q + 17, (size_t)cnt);
                if (build_huff(h)) return JO_INVALID;
                q += 17 + cnt; rem -= 17 + cnt;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (have_sof || rem < 6) return JO_INVALID;
            if (q[0] != 8) return JO_UNIMPLEMENTED;
            s->height = u16(q + 1); s->width = u16(q + 3); s->ncomp = q[5];
            if (s->height <= 0 || s->width <= 0) return JO_INVALID;
            if (s->ncomp != 1 && s->ncomp != 3) return JO_UNIMPLEMENTED;
            if (rem < 6 + 3 * s->ncomp) return JO_INVALID;
            for (int i = 0; i < s->ncomp; ++i) {
                jo_comp* c = &s->comp[i];
                c->id = q[6 + 3 * i]; c->h = q[7 + 3 * i] >> 4; c->v = q[7 + 3 * i] & 15; c->tq = q[8 + 3 * i];
                if (c->h < 1 || c->h > 4 || c->v < 1 || c->v > 4 || c->tq > 3) return JO_INVALID;
            }
            have_sof = 1;
        } else if (m == 0xC2 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC) || m == 0xC3) {
            return JO_UNIMPLEMENTED;                /* progressive, lossless, arithmetic, hierarchical */
        } else if (m == 0xDD) {
            if (rem < 2) return JO_INVALID;
            s->restart = u16(q);
        } else if (m == 0xDA) {
            if (!have_sof || rem < 1) return JO_INVALID;
            int ns = q[0];
            if (ns != s->ncomp) return JO_UNIMPLEMENTED;   /* one interleaved scan only */
            if (rem < 1 + 2 * ns + 3) return JO_INVALID;
            for (int i = 0; i < ns; ++i) {
                int cs = q[1 + 2 * i], j;
                for (j = 0; j < s->ncomp && s->comp[j].id != cs; ++j) {}
                if (j != i) return j == s->ncomp ? JO_INVALID : JO_UNIMPLEMENTED;
                s->comp[i].td = q[2 + 2 * i] >> 4; s->comp[i].ta = q[2 + 2 * i] & 15;
                if (s->comp[i].td > 3 || s->comp[i].ta > 3) return JO_INVALID;
            }
            s->pos = p + len;
            return JO_OK;
        }
        p += len;
    }
}

static int geometry(jo_dec* s) {
    if (s->ncomp == 1) { s->comp[0].h = s->comp[0].v = 1; }   /* a single-component scan is never interleaved (A.2.2) */
    s->hmax = s->vmax = 1;
    for (int i = 0; i < s->ncomp; ++i) {
        if (s->comp[i].h > s->hmax) s->hmax = s->comp[i].h;
        if (s->comp[i].v > s->vmax) s->vmax = s->comp[i].v;
    }
    if (s->ncomp == 3) {
        const jo_comp* c = s->comp;
        if (c[1].h != 1 || c[1].v != 1 || c[2].h != 1 || c[2].v != 1) return JO_UNIMPLEMENTED;
        if (!((c[0].h == 1 && c[0].v == 1) || (c[0].h == 2 && c[0].v == 1) || (c[0].h == 2 && c[0].v == 2)))
            return JO_UNIMPLEMENTED;
    }
    s->mcus_x = (s->width + 8 * s->hmax - 1) / (8 * s->hmax);
    s->mcus_y = (s->height + 8 * s->vmax - 1) / (8 * s->vmax);
    for (int i = 0; i < s->ncomp; ++i) {
        jo_comp* c = &s->comp[i];
        c->bw = s->mcus_x * c->h; c->bh = s->mcus_y * c->v;
        c->dw = (s->width * c->h + s->hmax - 1) / s->hmax;
        c->dh = (s->height * c->v + s->vmax - 1) / s->vmax;
    }
    return JO_OK;
}

int orc_jpeg_info(const uint8_t* d, int64_t n, int32_t* height, int32_t* width, int32_t* ncomp) {
    jo_dec* s = (jo_dec*)calloc(1, sizeof(jo_dec));
    s->d = d; s->n = n;
    int rc = parse_headers(s);
    if (!rc) rc = geometry(s);
    if (!rc) { *height = s->height; *width = s->width; *ncomp = s->ncomp; }
    free(s);
    return rc;
}

static int decode_scan(jo_dec* s) {
    int pred[3] = {0, 0, 0};
    int todo = s->restart, rst = 0;
    for (int i = 0; i < s->ncomp; ++i) {
        if (!s->dc[s->comp[i].td].present || !s->ac[s->comp[i].ta].present || !s->qt_present[s->comp[i].tq])
            return JO_INVALID;
    }
    for (int my = 0; my < s->mcus_y; ++my)
        for (int mx = 0; mx < s->mcus_x; ++mx) {
            if (s->restart && todo == 0) {
                /* byte-align, step over RSTn (E.2.4), reset the predictors */
                s->bitcnt = 0;
                if (!s->hit_marker) {
                    while (s->pos + 1 < s->n && !(s->d[s->pos] == 0xFF && s->d[s->pos + 1] != 0 && s->d[s->pos + 1] != 0xFF))
                        s->pos++;
                }
                if (s->pos + 1 < s->n && s->d[s->pos] == 0xFF && s->d[s->pos + 1] == 0xD0 + rst) {
                    s->pos += 2; s->hit_marker = 0;
                } else if (s->pos + 1 >= s->n) {
                    s->hit_marker = 1; s->truncated = 1;
                } else if (s->d[s->pos + 1] == 0xD9) {
                    s->hit_marker = 1;              /* EOI early: the rest decodes from zero bits */
                } else return JO_INVALID;
                rst = (rst + 1) & 7;
                pred[0] = pred[1] = pred[2] = 0;
                todo = s->restart;
            }
            for (int i = 0; i < s->ncomp; ++i) {
                jo_comp* c = &s->comp[i];
                for (int v = 0; v < c->v; ++v)
                    for (int h = 0; h < c->h; ++h) {
                        int16_t* blk = c->coef + ((size_t)(my * c->v + v) * c->bw + (mx * c->h + h)) * 64;
                        int rc = decode_block(s, &s->dc[c->td], &s->ac[c->ta], &pred[i], blk);
                        if (rc) return rc;
                    }
            }
            if (s->restart) --todo;
        }
    return JO_OK;
}

/* libjpeg's "slow" integer inverse DCT (Loeffler-Ligtenberg-Moshovitz, 13-bit constants, PASS1_BITS = 2) */
#define CB 13
#define P1 2
#define DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
static uint8_t range_limit(int x) {
    x &= 1023;
    if (x < 128) return (uint8_t)(x + 128);
    if (x < 512) return 255;
    if (x < 896) return 0;
    return (uint8_t)(x - 896);
}

static void idct_1d(const int32_t in[8], int32_t even[4], int32_t odd[4]) {
    int32_t z2 = in[2], z3 = in[6];
    int32_t z1 = (z2 + z3) * 4433;
    int32_t tmp2 = z1 + z3 * (-15137);
    int32_t tmp3 = z1 + z2 * 6270;
    int32_t tmp0 = (int32_t)((uint32_t)(in[0] + in[4]) << CB);
    int32_t tmp1 = (int32_t)((uint32_t)(in[0] - in[4]) << CB);
    even[0] = tmp0 + tmp3; even[3] = tmp0 - tmp3; even[1] = tmp1 + tmp2; even[2] = tmp1 - tmp2;
    int32_t t0 = in[7], t1 = in[5], t2 = in[3], t3 = in[1];
    z1 = t0 + t3; z2 = t1 + t2; z3 = t0 + t2;
    int32_t z4 = t1 + t3;
    int32_t z5 = (z3 + z4) * 9633;
    t0 *= 2446; t1 *= 16819; t2 *= 25172; t3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    odd[0] = t0 + z1 + z3; odd[1] = t1 + z2 + z4; odd[2] = t2 + z2 + z3; odd[3] = t3 + z1 + z4;
}

static void idct_block(const int16_t* coef, const uint16_t* q, uint8_t* out, int stride) {
    int32_t ws[64];
    for (int c = 0; c < 8; ++c) {
        int32_t in[8], e[4], o[4];
        for (int r = 0; r < 8; ++r) in[r] = (int32_t)coef[r * 8 + c] * (int32_t)q[r * 8 + c];
        idct_1d(in, e, o);
        ws[0 * 8 + c] = DESCALE(e[0] + o[3], CB - P1); ws[7 * 8 + c] = DESCALE(e[0] - o[3], CB - P1);
        ws[1 * 8 + c] = DESCALE(e[1] + o[2], CB - P1); ws[6 * 8 + c] = DESCALE(e[1] - o[2], CB - P1);
        ws[2 * 8 + c] = DESCALE(e[2] + o[1], CB - P1); ws[5 * 8 + c] = DESCALE(e[2] - o[1], CB - P1);
        ws[3 * 8 + c] = DESCALE(e[3] + o[0], CB - P1); ws[4 * 8 + c] = DESCALE(e[3] - o[0], CB - P1);
    }
    for (int r = 0; r < 8; ++r) {
        int32_t e[4], o[4];
        idct_1d(ws + r * 8, e, o);
        uint8_t* p = out + (size_t)r * stride;
        p[0] = range_limit(DESCALE(e[0] + o[3], CB + P1 + 3)); p[7] = range_limit(DESCALE(e[0] - o[3], CB + P1 + 3));
        p[1] = range_limit(DESCALE(e[1] + o[2], CB + P1 + 3)); p[6] = range_limit(DESCALE(e[1] - o[2], CB + P1 + 3));
        p[2] = range_limit(DESCALE(e[2] + o[1], CB + P1 + 3)); p[5] = range_limit(DESCALE(e[2] - o[1], CB + P1 + 3));
        p[3] = range_limit(DESCALE(e[3] + o[0], CB + P1 + 3)); p[4] = range_limit(DESCALE(e[3] - o[0], CB + P1 + 3));
    }
}

/* One upsampled chroma sample at full-resolution (y, x); rows/columns beyond the real samples replicate the edge. */
static int chroma_at(const jo_comp* c, int hs, int vs, int y, int x) {
    const int st = c->bw * 8;
    const uint8_t* pl = c->plane;
    if (hs == 1) return pl[(size_t)y * st + x];
    const int cx = x >> 1;
    if (vs == 1) {                                   /* 2:1 horizontally */
        const uint8_t* r = pl + (size_t)y * st;
        if (c->dw <= 2) return r[cx];                /* too narrow for the filter: replicate */
        if (x & 1) return cx == c->dw - 1 ? r[cx] : (3 * r[cx] + r[cx + 1] + 2) >> 2;
        return cx == 0 ? r[0] : (3 * r[cx] + r[cx - 1] + 1) >> 2;
    }
    const int cy = y >> 1;                           /* 2x2 */
    if (c->dw <= 2) return pl[(size_t)cy * st + cx];
    int ny = (y & 1) ? cy + 1 : cy - 1;
    if (ny < 0) ny = 0;
    if (ny > c->dh - 1) ny = c->dh - 1;
    const uint8_t *r0 = pl + (size_t)cy * st, *r1 = pl + (size_t)ny * st;
    const int cur = 3 * r0[cx] + r1[cx];
    if (x & 1) return cx == c->dw - 1 ? (cur * 4 + 7) >> 4 : (cur * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    return cx == 0 ? (cur * 4 + 8) >> 4 : (cur * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

static uint8_t clamp255(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

/* decodes one JPEG file into rgb[height][width][3] (caller sized it from orc_jpeg_info) */
int orc_jpeg_decode(const uint8_t* d, int64_t n, uint8_t* rgb) {
    jo_dec* s = (jo_dec*)calloc(1, sizeof(jo_dec));
    s->d = d; s->n = n;
    int rc = parse_headers(s);
    if (!rc) rc = geometry(s);
    if (!rc) {
        for (int i = 0; i < s->ncomp; ++i) {
            jo_comp* c = &s->comp[i];
            c->coef = (int16_t*)calloc((size_t)c->bw * c->bh * 64, sizeof(int16_t));
            c->plane = (uint8_t*)malloc((size_t)c->bw * c->bh * 64);
        }
        rc = decode_scan(s);
        /* TF's DecodeJpeg runs with try_recover_truncated = false (the reference sets only `channels`,
         * GraphBuilder.java:41-47): a file that ends before its closing marker is an error, not a padded image */
        if (!rc && !s->hit_marker) {          /* libjpeg goes on to read the closing marker */
            int64_t q = s->pos;
            while (q + 1 < s->n && !(s->d[q] == 0xFF && s->d[q + 1] != 0 && s->d[q + 1] != 0xFF)) ++q;
            if (q + 1 >= s->n) s->truncated = 1;
        }
        if (!rc && s->truncated) rc = JO_INVALID;
    }
    if (!rc) {
        for (int i = 0; i < s->ncomp; ++i) {
            jo_comp* c = &s->comp[i];
            for (int by = 0; by < c->bh; ++by)
                for (int bx = 0; bx < c->bw; ++bx)
                    idct_block(c->coef + ((size_t)by * c->bw + bx) * 64, s->qt[c->tq],
                               c->plane + ((size_t)by * 8 * c->bw + bx) * 8, c->bw * 8);
        }
        const int W = s->width, H = s->height;
        if (s->ncomp == 1) {
            const jo_comp* c = &s->comp[0];
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x) {
                    uint8_t v = c->plane[(size_t)y * c->bw * 8 + x];
                    uint8_t* o = rgb + ((size_t)y * W + x) * 3;
                    o[0] = o[1] = o[2] = v;
                }
        } else {
            const int hs = s->comp[0].h, vs = s->comp[0].v;
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x) {
                    const int yy = s->comp[0].plane[(size_t)y * s->comp[0].bw * 8 + x];
                    const int cb = chroma_at(&s->comp[1], hs, vs, y, x) - 128;
                    const int cr = chroma_at(&s->comp[2], hs, vs, y, x) - 128;
                    uint8_t* o = rgb + ((size_t)y * W + x) * 3;
                    o[0] = clamp255(yy + ((91881 * cr + 32768) >> 16));
                    o[1] = clamp255(yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
                    o[2] = clamp255(yy + ((116130 * cb + 32768) >> 16));
                }
        }
    }
    for (int i = 0; i < 3; ++i) { free(s->comp[i].coef); free(s->comp[i].plane); }
    free(s);
    return rc;
}
