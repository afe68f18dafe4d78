/*
 * oracle.c -- CPU restatement of the reference's per-record model-invocation path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under flink-tensorflow_b200/ may include, link or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker or as the timed CPU stand-in.
 *
 * The arithmetic of the reference lives in an un-vendored dependency:
 *   Maven org.tensorflow:tensorflow:1.1.0 (libtensorflow_jni, CPU)      pom.xml:54,
 *   flink-tensorflow/pom.xml:99-103.
 * It cannot be compiled or run here (no JVM, no TensorFlow).  The functions below restate
 * the published TF-1.1 CPU kernels that the reference's graphs execute, and are anchored
 * on the reference's own call sites, fixtures and known-answer tests.  Pin status:
 *   - half_plus_two (orc_affine_f32, orc_parse_example_dense_f32, orc_string_tensor_*):
 *       PINNED by RegressITCase.scala:35,58-60 and the models/half_plus_two fixture.
 *   - TensorValue wire image (orc_tv_*): PINNED by TensorValueTest.java:23-55 (layout) and
 *       TensorValue.java:142-187 (the serializer itself, restated bit for bit).
 *   - resize/normalise (orc_resize_bilinear_norm_u8) and top labels (orc_top_labels):
 *       PARITY UNPINNED -- the reference has no test, model file or labels file for them
 *       (flink-tensorflow-examples has no src/test).  They follow TF-1.1
 *       ResizeBilinear(align_corners=false) and InceptionModel.scala:81-89 as written.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (stock TF-1.1 x86 wheels carry no FMA).
 * All paths are relative to /root/reference.  FT = flink-tensorflow/src/main/scala/org/apache/
 * flink/contrib/tensorflow, EX = flink-tensorflow-examples/src/main/scala/org/apache/flink/
 * contrib/tensorflow/examples.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <pthread.h>
#include <unistd.h>

#define ORC_OK 0
#define ORC_INVALID_ARGUMENT 3
#define ORC_OUT_OF_RANGE 11
#define ORC_UNIMPLEMENTED 12

int orc_num_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

/* Static partition of [0, n) over `threads` pthreads (no OpenMP runtime in this image). */
typedef void (*orc_range_fn)(int64_t lo, int64_t hi, void* ctx);
typedef struct { orc_range_fn fn; void* ctx; int64_t lo, hi; } orc_job;
static void* orc_job_main(void* p) { orc_job* j = (orc_job*)p; j->fn(j->lo, j->hi, j->ctx); return NULL; }
void orc_parallel_for(int64_t n, int threads, orc_range_fn fn, void* ctx) {
    if (threads <= 1 || n <= 1) { fn(0, n, ctx); return; }
    if (threads > 256) threads = 256;
    if ((int64_t)threads > n) threads = (int)n;
    pthread_t tid[256]; orc_job job[256];
    const int64_t per = (n + threads - 1) / threads;
    int started = 0;
    for (int t = 0; t < threads; ++t) {
        job[t].fn = fn; job[t].ctx = ctx; job[t].lo = t * per; job[t].hi = (t + 1) * per < n ? (t + 1) * per : n;
        if (job[t].lo >= job[t].hi) break;
        if (pthread_create(&tid[t], NULL, orc_job_main, &job[t]) != 0) { fn(job[t].lo, job[t].hi, ctx); tid[t] = 0; }
        ++started;
    }
    for (int t = 0; t < started; ++t) if (tid[t]) pthread_join(tid[t], NULL);
}

/* ------------------------------------------------------------------------------------------
 * half_plus_two: y = Add(Mul(a, x), b)
 * Graph decoded from models/half_plus_two/saved_model.pb: node "Mul" = Mul(a/read, x),
 * node "y" = Add(Mul, b/read); y2/y3 use c.  Two separately rounded fp32 ops.
 * Call site: FT/models/ModelFunction.scala:47-65 (feed, fetch, run); known answers
 * RegressITCase.scala:35.
 * ---------------------------------------------------------------------------------------- */
typedef struct { const float* x; float* y; float a, b; } affine_ctx;
static void affine_range(int64_t lo, int64_t hi, void* p) {
    affine_ctx* c = (affine_ctx*)p;
    for (int64_t i = lo; i < hi; ++i) {
        const float m = c->a * c->x[i]; /* Mul, rounded to fp32 (no contraction: -ffp-contract=off) */
        c->y[i] = m + c->b;             /* Add, rounded to fp32 */
    }
}
void orc_affine_f32(const float* x, int64_t n, float a, float b, float* y, int threads) {
    affine_ctx c = { x, y, a, b };
    orc_parallel_for(n, threads, affine_range, &c);
}

/* ------------------------------------------------------------------------------------------
 * ImageNormalization: Cast(u8->f32) -> ExpandDims(0) -> ResizeBilinear([H,W]) -> Sub(mean)
 * -> Div(scale).  EX/inception/ImageNormalization.scala:56-65, ops from
 * EX/common/GraphBuilder.java:21-47 (no align_corners attr => false).
 * TF-1.1 ResizeBilinear CPU kernel, align_corners=false:
 *   scale = in / (float)out;  in_pos = (float)i * scale;  lo = (int64)in_pos;
 *   hi = min(lo + 1, in - 1);  lerp = in_pos - lo;
 *   top = tl + (tr - tl) * xl;  bot = bl + (br - bl) * xl;  v = top + (bot - top) * yl;
 * all in fp32 without contraction; then (v - mean) and a true division by scale.
 * in : [n, in_h, in_w, 3] u8 (HWC);  out: [n, out_h, out_w, 3] f32.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    const uint8_t* in; float* out; int in_h, in_w, ch, out_h, out_w; float hs, mean, scale;
    const int64_t *xlo, *xhi; const float* xl;
} resize_ctx;
static void resize_rows(int64_t lo, int64_t hi, void* p) {
    const resize_ctx* c = (const resize_ctx*)p;
    const int ch = c->ch;
    for (int64_t row = lo; row < hi; ++row) {
        const int64_t b = row / c->out_h; const int y = (int)(row - b * c->out_h);
        const uint8_t* img = c->in + (size_t)b * c->in_h * c->in_w * ch;
        float* o = c->out + (size_t)row * c->out_w * ch;
        const float py = (float)y * c->hs;
        const int64_t ylo = (int64_t)py;
        const int64_t yhi = ylo + 1 < c->in_h - 1 ? ylo + 1 : c->in_h - 1;
        const float yl = py - (float)ylo;
        const uint8_t* r0 = img + (size_t)ylo * c->in_w * ch;
        const uint8_t* r1 = img + (size_t)yhi * c->in_w * ch;
        for (int x = 0; x < c->out_w; ++x) {
            const int64_t x0 = c->xlo[x] * ch, x1 = c->xhi[x] * ch; const float xl = c->xl[x];
            for (int k = 0; k < ch; ++k) {
                const float tl = (float)r0[x0 + k], tr = (float)r0[x1 + k];
                const float bl = (float)r1[x0 + k], br = (float)r1[x1 + k];
                const float top = tl + (tr - tl) * xl;
                const float bot = bl + (br - bl) * xl;
                const float v = top + (bot - top) * yl;
                o[x * ch + k] = (v - c->mean) / c->scale;
            }
        }
    }
}
int orc_resize_bilinear_norm_u8(const uint8_t* in, int64_t n, int in_h, int in_w, int ch,
                                int out_h, int out_w, float mean, float scale, float* out,
                                int threads) {
    if (n < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || ch <= 0)
        return ORC_INVALID_ARGUMENT;
    const float ws = (float)in_w / (float)out_w;
    int64_t* xlo = (int64_t*)malloc(sizeof(int64_t) * (size_t)out_w * 2);
    float* xl = (float*)malloc(sizeof(float) * (size_t)out_w);
    if (!xlo || !xl) { free(xlo); free(xl); return ORC_INVALID_ARGUMENT; }
    int64_t* xhi = xlo + out_w;
    for (int x = 0; x < out_w; ++x) {
        const float p = (float)x * ws;
        const int64_t lo = (int64_t)p;
        xlo[x] = lo; xhi[x] = lo + 1 < in_w - 1 ? lo + 1 : in_w - 1; xl[x] = p - (float)lo;
    }
    resize_ctx c = { in, out, in_h, in_w, ch, out_h, out_w, (float)in_h / (float)out_h, mean, scale, xlo, xhi, xl };
    orc_parallel_for(n * out_h, threads, resize_rows, &c);
    free(xlo); free(xl);
    return ORC_OK;
}

/* Same op sequence but the result rounded to bf16 (round-to-nearest-even), returned as the
 * raw 16-bit pattern: the layout the convolution path consumes.  New behaviour (no reference
 * counterpart); defined as bf16_rne(orc_resize_bilinear_norm_u8(...)). */
static inline uint16_t orc_f32_to_bf16_bits(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40); /* quiet NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
uint16_t orc_bf16_bits(float f) { return orc_f32_to_bf16_bits(f); }
float orc_bf16_round(float f) {
    uint32_t u = (uint32_t)orc_f32_to_bf16_bits(f) << 16; float r; memcpy(&r, &u, 4); return r;
}

/* ------------------------------------------------------------------------------------------
 * toTextLabels: EX/inception/InceptionModel.scala:81-89
 *   require rank 2; per row: row.toList.zip(labels).sortWith(_._1 > _._1).take(take)
 * zip truncates to min(N, #labels); List.sortWith is a stable merge sort, so equal scores
 * keep ascending column order.  NaN ordering is unspecified in the reference; here NaN never
 * compares greater, i.e. it behaves like the comparator `a > b` inside a stable insertion.
 * scores: [m, n] f32.  out_idx/out_score: [m, k]; rows shorter than k are padded with -1/0.
 * ---------------------------------------------------------------------------------------- */
int orc_top_labels(const float* scores, int64_t m, int64_t n, int64_t n_labels, int k,
                   int32_t* out_idx, float* out_score) {
    if (m < 0 || n < 0 || k < 0) return ORC_INVALID_ARGUMENT;
    const int64_t cols = n < n_labels ? n : n_labels;
    for (int64_t r = 0; r < m; ++r) {
        const float* row = scores + r * n;
        int32_t* oi = out_idx + r * k;
        float* os = out_score + r * k;
        int filled = 0;
        for (int64_t c = 0; c < cols; ++c) {
            /* stable: a new element goes after every kept element e with !(row[c] > e) */
            int pos = filled;
            while (pos > 0 && row[c] > os[pos - 1]) --pos;
            if (pos >= k) continue;
            int last = filled < k ? filled : k - 1;
            for (int j = last; j > pos; --j) { os[j] = os[j - 1]; oi[j] = oi[j - 1]; }
            os[pos] = row[c]; oi[pos] = (int32_t)c;
            if (filled < k) ++filled;
        }
        for (int j = filled; j < k; ++j) { oi[j] = -1; os[j] = 0.0f; }
    }
    return ORC_OK;
}

/* Row softmax in fp32: p = exp(x - max) / sum(exp(x - max)), sum accumulated left to right.
 * The reference's frozen graph ends in a Softmax node named "output"
 * (EX/inception/InceptionModel.scala:58-62); unpinned. */
void orc_softmax_rows(const float* x, int64_t m, int64_t n, float* p) {
    for (int64_t r = 0; r < m; ++r) {
        const float* xr = x + r * n; float* pr = p + r * n;
        float mx = -INFINITY;
        for (int64_t c = 0; c < n; ++c) if (xr[c] > mx) mx = xr[c];
        float s = 0.0f;
        for (int64_t c = 0; c < n; ++c) { pr[c] = expf(xr[c] - mx); s += pr[c]; }
        for (int64_t c = 0; c < n; ++c) pr[c] = pr[c] / s;
    }
}

/* ------------------------------------------------------------------------------------------
 * TensorValue wire image: FT/types/TensorValue.java:142-187
 *   0x01 | i32 dtype | i32 rank | rank x i64 dims | i32 nbytes | payload
 * Header integers big-endian (Flink DataOutputView), payload native (little-endian) order.
 * dtype codes FT/util/TFUtils.java:10-32: FLOAT=1 DOUBLE=2 INT32=3 STRING=7 INT64=9 BOOL=10.
 * ---------------------------------------------------------------------------------------- */
static int orc_dtype_known(int32_t d) {
    return d == 1 || d == 2 || d == 3 || d == 7 || d == 9 || d == 10;
}
static void put_be32(uint8_t* p, uint32_t v) { p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }
static void put_be64(uint8_t* p, uint64_t v) { put_be32(p, (uint32_t)(v >> 32)); put_be32(p + 4, (uint32_t)v); }
static uint32_t get_be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
static uint64_t get_be64(const uint8_t* p) { return ((uint64_t)get_be32(p) << 32) | get_be32(p + 4); }

int64_t orc_tv_binary_length(int32_t rank, int64_t nbytes) { /* TensorValue.java:142-147 */
    return 1 + 4 + 4 + 8 * (int64_t)rank + 4 + nbytes;
}
int orc_tv_encode(int32_t dtype, int32_t rank, const int64_t* dims, const void* data,
                  int64_t nbytes, uint8_t* out, int64_t cap) {
    if (!orc_dtype_known(dtype)) return ORC_INVALID_ARGUMENT;   /* TFUtils.java:18 */
    if (rank < 0 || rank > 25) return ORC_UNIMPLEMENTED;        /* TFUtils.java:75 */
    if (nbytes < 0 || nbytes > 0x7fffffffLL) return ORC_INVALID_ARGUMENT;
    if (orc_tv_binary_length(rank, nbytes) > cap) return ORC_OUT_OF_RANGE;
    uint8_t* p = out;
    *p++ = 0x01; put_be32(p, (uint32_t)dtype); p += 4; put_be32(p, (uint32_t)rank); p += 4;
    for (int i = 0; i < rank; ++i) { put_be64(p, (uint64_t)dims[i]); p += 8; }
    put_be32(p, (uint32_t)nbytes); p += 4;
    memcpy(p, data, (size_t)nbytes);
    return ORC_OK;
}
/* returns ORC_OK and fills dtype/rank/dims/payload offset/nbytes; 100 = version mismatch */
int orc_tv_decode(const uint8_t* in, int64_t len, int32_t* dtype, int32_t* rank, int64_t* dims,
                  int64_t* payload_off, int64_t* nbytes) {
    if (len < 13) return ORC_OUT_OF_RANGE;
    if (in[0] != 0x01) return 100;                              /* TensorValue.java:152-154 */
    int32_t d = (int32_t)get_be32(in + 1);
    if (!orc_dtype_known(d)) return ORC_INVALID_ARGUMENT;       /* TFUtils.java:30 */
    int32_t r = (int32_t)get_be32(in + 5);
    if (r < 0 || r > 25) return ORC_UNIMPLEMENTED;
    if (len < 13 + 8 * (int64_t)r) return ORC_OUT_OF_RANGE;
    for (int i = 0; i < r; ++i) dims[i] = (int64_t)get_be64(in + 9 + 8 * i);
    int64_t nb = (int32_t)get_be32(in + 9 + 8 * r);
    if (nb < 0 || len < 13 + 8 * (int64_t)r + nb) return ORC_OUT_OF_RANGE;
    *dtype = d; *rank = r; *payload_off = 13 + 8 * (int64_t)r; *nbytes = nb;
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * TF-1.x STRING tensor [N]: N x u64 LE offsets (relative to the end of the table) followed by
 * N x (varint length | bytes).  FT/types/TensorInjections.scala:49-72 builds exactly this
 * (offset = bb.position() - dataStart, :63).
 * ---------------------------------------------------------------------------------------- */
static int put_varint(uint8_t* p, uint64_t v) { int n = 0; while (v >= 0x80) { p[n++] = (uint8_t)(v | 0x80); v >>= 7; } p[n++] = (uint8_t)v; return n; }
static int get_varint(const uint8_t* p, const uint8_t* end, uint64_t* v) {
    uint64_t r = 0; int s = 0; const uint8_t* q = p;
    while (q < end && s < 64) { uint8_t c = *q++; r |= (uint64_t)(c & 0x7f) << s; s += 7; if (c < 0x80) { *v = r; return (int)(q - p); } }
    return -1;
}
int64_t orc_string_tensor_encode(int64_t n, const uint8_t* const* items, const int64_t* lens,
                                 uint8_t* out, int64_t cap) {
    int64_t need = 8 * n;
    for (int64_t i = 0; i < n; ++i) { uint8_t t[10]; need += put_varint(t, (uint64_t)lens[i]) + lens[i]; }
    if (!out) return need;
    if (need > cap) return -ORC_OUT_OF_RANGE;
    int64_t pos = 8 * n;
    for (int64_t i = 0; i < n; ++i) {
        uint64_t off = (uint64_t)(pos - 8 * n);
        memcpy(out + 8 * i, &off, 8);                            /* native order == LE here */
        pos += put_varint(out + pos, (uint64_t)lens[i]);
        memcpy(out + pos, items[i], (size_t)lens[i]); pos += lens[i];
    }
    return pos;
}
/* item i -> (offset into buf, length); returns ORC_OK or an error */
int orc_string_tensor_item(const uint8_t* buf, int64_t nbytes, int64_t n, int64_t i,
                           int64_t* off, int64_t* len) {
    if (i < 0 || i >= n || nbytes < 8 * n) return ORC_INVALID_ARGUMENT;
    uint64_t o; memcpy(&o, buf + 8 * i, 8);
    if ((int64_t)o < 0 || 8 * n + (int64_t)o >= nbytes) return ORC_INVALID_ARGUMENT;
    uint64_t l; int k = get_varint(buf + 8 * n + o, buf + nbytes, &l);
    if (k < 0 || 8 * n + (int64_t)o + k + (int64_t)l > nbytes) return ORC_INVALID_ARGUMENT;
    *off = 8 * n + (int64_t)o + k; *len = (int64_t)l;
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * ParseExample for one dense float feature of shape [1] (the half_plus_two graph's
 * "ParseExample/ParseExample" node: dense keys "x" (no default => required) and "x2"
 * (default [0.0]); decoded from models/half_plus_two/saved_model.pb).
 * Example{1: Features{1: map<string,Feature>{1:key, 2:Feature{1:bytes_list 2:float_list
 * 3:int64_list}}}}; FloatList{1: repeated float, packed or not}.  Last occurrence of a key
 * wins (protobuf map semantics).  Input construction: FTT/util/TestData.scala:15-34.
 * Returns ORC_OK, or ORC_INVALID_ARGUMENT for: malformed proto, key missing and no default,
 * wrong kind, or number of values != 1.
 * ---------------------------------------------------------------------------------------- */
typedef struct { const uint8_t* p; const uint8_t* end; } span_t;
static int next_field(span_t* s, uint32_t* field, uint32_t* wt, span_t* payload, uint64_t* val) {
    uint64_t tag; int k = get_varint(s->p, s->end, &tag); if (k < 0) return -1; s->p += k;
    *field = (uint32_t)(tag >> 3); *wt = (uint32_t)(tag & 7);
    switch (*wt) {
    case 0: k = get_varint(s->p, s->end, val); if (k < 0) return -1; s->p += k; return 0;
    case 1: if (s->end - s->p < 8) return -1; payload->p = s->p; payload->end = s->p + 8; s->p += 8; return 0;
    case 5: if (s->end - s->p < 4) return -1; payload->p = s->p; payload->end = s->p + 4; s->p += 4; return 0;
    case 2: { uint64_t l; k = get_varint(s->p, s->end, &l); if (k < 0) return -1; s->p += k;
              if ((uint64_t)(s->end - s->p) < l) return -1; payload->p = s->p; payload->end = s->p + l; s->p += l; return 0; }
    default: return -1;
    }
}
int orc_parse_example_dense_f32(const uint8_t* ex, int64_t len, const char* key, int has_default,
                                float def, float* out) {
    span_t top = { ex, ex + len }; size_t klen = strlen(key);
    int found = 0; float value = 0.0f; int nvals = 0; int wrong_kind = 0;
    while (top.p < top.end) {
        uint32_t f, w; span_t pl; uint64_t v;
        if (next_field(&top, &f, &w, &pl, &v)) return ORC_INVALID_ARGUMENT;
        if (f != 1 || w != 2) continue;                 /* Example.features */
        span_t feats = pl;
        while (feats.p < feats.end) {
            if (next_field(&feats, &f, &w, &pl, &v)) return ORC_INVALID_ARGUMENT;
            if (f != 1 || w != 2) continue;             /* Features.feature map entry */
            span_t entry = pl; span_t k = { 0, 0 }, val = { 0, 0 };
            while (entry.p < entry.end) {
                if (next_field(&entry, &f, &w, &pl, &v)) return ORC_INVALID_ARGUMENT;
                if (f == 1 && w == 2) k = pl; else if (f == 2 && w == 2) val = pl;
            }
            if ((size_t)(k.end - k.p) != klen || memcmp(k.p, key, klen) != 0) continue;
            found = 1; nvals = 0; wrong_kind = 0;
            span_t feat = val;
            while (feat.p && feat.p < feat.end) {
                if (next_field(&feat, &f, &w, &pl, &v)) return ORC_INVALID_ARGUMENT;
                if (f != 2) { if (f == 1 || f == 3) wrong_kind = 1; continue; }
                if (w != 2) return ORC_INVALID_ARGUMENT; /* float_list */
                span_t fl = pl;
                while (fl.p < fl.end) {
                    if (next_field(&fl, &f, &w, &pl, &v)) return ORC_INVALID_ARGUMENT;
                    if (f != 1) continue;
                    if (w == 5) { memcpy(&value, pl.p, 4); ++nvals; }
                    else if (w == 2) { if ((pl.end - pl.p) % 4) return ORC_INVALID_ARGUMENT;
                                       for (const uint8_t* q = pl.p; q < pl.end; q += 4) { memcpy(&value, q, 4); ++nvals; } }
                    else return ORC_INVALID_ARGUMENT;
                }
            }
        }
    }
    if (found && wrong_kind) return ORC_INVALID_ARGUMENT;
    if (!found || nvals == 0) { if (!has_default) return ORC_INVALID_ARGUMENT; *out = def; return ORC_OK; }
    if (nvals != 1) return ORC_INVALID_ARGUMENT;
    *out = value; return ORC_OK;
}

/* The whole regress_x_to_y call of RegressITCase.scala:52-61: STRING tensor [N] of serialized
 * Examples -> ParseExample("x") -> Mul(a) -> Add(b) -> f32 [N,1]. */
int orc_regress(const uint8_t* st, int64_t nbytes, int64_t n, const char* key, int has_default,
                float def, float a, float b, float* y) {
    for (int64_t i = 0; i < n; ++i) {
        int64_t off, len; int rc = orc_string_tensor_item(st, nbytes, n, i, &off, &len);
        if (rc) return rc;
        float x; rc = orc_parse_example_dense_f32(st + off, len, key, has_default, def, &x);
        if (rc) return rc;
        orc_affine_f32(&x, 1, a, b, y + i, 1);
    }
    return ORC_OK;
}

/* TensorName: FT/types/TensorName.scala:14-19.  "op:idx" -> (op, idx); no colon => idx 0;
 * more than one colon => INVALID_ARGUMENT.  Scala's split(':') drops trailing empty strings,
 * so "op:" parses as ("op", 0); a non-numeric index throws (NumberFormatException is an
 * IllegalArgumentException). */
int orc_tensor_name(const char* s, char* op, int cap, int* idx) {
    int parts = 1; size_t n = strlen(s); size_t end = n;
    while (end > 0 && s[end - 1] == ':') --end;           /* trailing empties removed */
    for (size_t i = 0; i < end; ++i) if (s[i] == ':') ++parts;
    if (parts > 2) return ORC_INVALID_ARGUMENT;
    size_t oplen = end; *idx = 0;
    if (parts == 2) {
        const char* c = memchr(s, ':', end); oplen = (size_t)(c - s);
        const char* d = c + 1; size_t dl = end - oplen - 1; if (dl == 0) return ORC_INVALID_ARGUMENT;
        long v = 0; size_t i = 0; int neg = 0;
        if (d[0] == '-' || d[0] == '+') { neg = d[0] == '-'; i = 1; if (dl == 1) return ORC_INVALID_ARGUMENT; }
        for (; i < dl; ++i) { if (d[i] < '0' || d[i] > '9') return ORC_INVALID_ARGUMENT; v = v * 10 + (d[i] - '0'); if (v > 0x7fffffffL) return ORC_INVALID_ARGUMENT; }
        *idx = (int)(neg ? -v : v);
    }
    if ((int)oplen + 1 > cap) return ORC_OUT_OF_RANGE;
    memcpy(op, s, oplen); op[oplen] = 0;
    return ORC_OK;
}
