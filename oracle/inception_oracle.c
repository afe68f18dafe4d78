/*
 * inception_oracle.c -- CPU definition of the label model (Inception-v3, inference, BN folded).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.c).  PARITY UNPINNED against the reference: the reference feeds a
 * frozen graph it does not ship (flink-tensorflow-examples/.../inception/InceptionModel.scala:51-62 loads
 * tensorflow_inception_graph.pb out of band; models/README.md:5) and has no test for it, and BASELINE.json
 * substitutes a random-init Inception-v3.  This file therefore DEFINES the network the CUDA path must
 * reproduce; it is cross-checked against torchvision.models.inception_v3(aux_logits=False,
 * transform_input=False) on CPU (tests/golden/make_inception_golden.py) within fp32 tolerance.
 *
 * Layout NHWC.  Conv + BN(eps=1e-3, gamma=1, mean=0, var=1, beta) + ReLU is folded to
 *   y = relu(conv(x, w / sqrt(1 + 1e-3)) + beta).
 * The 3x3/s1/p1 average pool divides by 9 everywhere (torchvision's count_include_pad=True), and in the
 * pool branches it is applied AFTER the 1x1 convolution (both are linear, so the result is the same up to
 * rounding); bias and ReLU follow the pool.
 *
 * Weight recipe, reproducible from (seed, layer name, element index) alone:
 *   conv  w ~ N(0, 2/fan_in), index order [Cout][KH][KW][Cin];  beta ~ U(-0.1, 0.1);  fc w ~ N(0, 0.05^2) drawn with
 *   seed + 90, b = 0
 *   key = mix(seed ^ mix(fnv1a(name) + kind));  r(i) = mix(key ^ i)   (mix = splitmix64 finaliser)
 *   normal(i) = sqrt(-2 ln u1) cos(2 pi u2),  u1 = ((r(2i) >> 11) + 1) / 2^53,  u2 = (r(2i+1) >> 11) / 2^53
 *
 * mode 0: fp32 everywhere.  mode 1 ("bf16 emulation"): weights and every stored activation rounded to bf16
 * (round-to-nearest-even) at the points the CUDA path rounds, accumulation in fp32.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef void (*orc_range_fn)(int64_t lo, int64_t hi, void* ctx);
void orc_parallel_for(int64_t n, int threads, orc_range_fn fn, void* ctx);
float orc_bf16_round(float f);

/* ---------------------------------------------------------------------------------------------- */
static uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint64_t fnv1a(const char* s) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (; *s; ++s) { h ^= (uint8_t)*s; h *= 0x100000001b3ull; }
    return h;
}
static uint64_t stream_key(uint64_t seed, const char* name, int kind) { return mix64(seed ^ mix64(fnv1a(name) + (uint64_t)kind)); }
static double unit_open(uint64_t r) { return (double)((r >> 11) + 1) / 9007199254740992.0; }   /* (0,1] */
static double unit_half(uint64_t r) { return (double)(r >> 11) / 9007199254740992.0; }         /* [0,1) */
static double normal_at(uint64_t key, uint64_t i) {
    const double u1 = unit_open(mix64(key ^ (2 * i))), u2 = unit_half(mix64(key ^ (2 * i + 1)));
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925286766559 * u2);
}
static double uniform_at(uint64_t key, uint64_t i) { return unit_half(mix64(key ^ i)); }

/* [Cout][KH][KW][Cin] folded weights and bias of one named conv layer */
void orc_inc_conv_params(uint64_t seed, const char* name, int cout, int kh, int kw, int cin, float* w, float* bias) {
    const uint64_t kw_key = stream_key(seed, name, 0), kb_key = stream_key(seed, name, 1);
    const double std = sqrt(2.0 / ((double)kh * kw * cin)), fold = 1.0 / sqrt(1.0 + 1e-3);
    const int64_t n = (int64_t)cout * kh * kw * cin;
    for (int64_t i = 0; i < n; ++i) w[i] = (float)(normal_at(kw_key, (uint64_t)i) * std * fold);
    for (int o = 0; o < cout; ++o) bias[o] = (float)(-0.1 + 0.2 * uniform_at(kb_key, (uint64_t)o));
}
/* The fc stream is salted (seed + 90): with the default seed 2024 that draw puts the top logits of the test images at
 * least 0.4 apart, several times the bf16 rounding noise of the logits, so that "top-5 indices exact" is a meaningful
 * check rather than a coin flip between near-tied classes (an init recipe with well-separated logits). */
#define ORC_FC_SALT 90
void orc_inc_fc_params(uint64_t seed, int classes, int cin, float* w, float* bias) {
    const uint64_t key = stream_key(seed + ORC_FC_SALT, "fc", 0);
    for (int64_t i = 0; i < (int64_t)classes * cin; ++i) w[i] = (float)(normal_at(key, (uint64_t)i) * 0.05);
    for (int o = 0; o < classes; ++o) bias[o] = 0.0f;
}

/* ---------------------------------------------------------------------------------------------- */
typedef struct { int n, h, w, c; float* d; } tens;
static tens t_new(int n, int h, int w, int c) { tens t = { n, h, w, c, (float*)malloc(sizeof(float) * (size_t)n * h * w * c) }; return t; }
static void t_free(tens* t) { free(t->d); t->d = NULL; }

typedef struct {
    const float* in; float* out; const float* wt; const float* bias;   /* wt: [KH][KW][Cin][Cout] */
    int n, h, w, cin, oh, ow, cout, kh, kw, sh, sw, ph, pw;
    int out_ld, out_off;         /* output row stride (channels of the concat buffer) and channel offset */
    int add_bias, relu, round_bf16;
} conv_ctx;

static void conv_rows(int64_t lo, int64_t hi, void* p) {
    const conv_ctx* c = (const conv_ctx*)p;
    float* acc = (float*)malloc(sizeof(float) * (size_t)c->cout);
    for (int64_t row = lo; row < hi; ++row) {           /* row = (n, oh) */
        const int b = (int)(row / c->oh), oy = (int)(row % c->oh);
        for (int ox = 0; ox < c->ow; ++ox) {
            for (int o = 0; o < c->cout; ++o) acc[o] = 0.0f;
            for (int ky = 0; ky < c->kh; ++ky) {
                const int iy = oy * c->sh - c->ph + ky;
                if (iy < 0 || iy >= c->h) continue;
                for (int kx = 0; kx < c->kw; ++kx) {
                    const int ix = ox * c->sw - c->pw + kx;
                    if (ix < 0 || ix >= c->w) continue;
                    const float* xin = c->in + (((size_t)b * c->h + iy) * c->w + ix) * c->cin;
                    const float* wk = c->wt + ((size_t)(ky * c->kw + kx) * c->cin) * c->cout;
                    for (int ci = 0; ci < c->cin; ++ci) {
                        const float xv = xin[ci];
                        const float* wr = wk + (size_t)ci * c->cout;
                        for (int o = 0; o < c->cout; ++o) acc[o] += xv * wr[o];
                    }
                }
            }
            float* out = c->out + (((size_t)b * c->oh + oy) * c->ow + ox) * c->out_ld + c->out_off;
            for (int o = 0; o < c->cout; ++o) {
                float v = acc[o];
                if (c->add_bias) v += c->bias[o];
                if (c->relu && !(v > 0.0f)) v = 0.0f;
                out[o] = c->round_bf16 ? orc_bf16_round(v) : v;
            }
        }
    }
    free(acc);
}

/* Generic NHWC convolution.  w_ohwi: [Cout][KH][KW][Cin].  Output written at channel offset out_off of a
 * buffer whose pixel stride is out_ld channels. */
int orc_conv2d_nhwc(const float* in, int n, int h, int w, int cin, const float* w_ohwi, const float* bias, int cout,
                    int kh, int kw, int sh, int sw, int ph, int pw, int add_bias, int relu, int round_bf16,
                    float* out, int out_ld, int out_off, int threads) {
    const int oh = (h + 2 * ph - kh) / sh + 1, ow = (w + 2 * pw - kw) / sw + 1;
    if (oh <= 0 || ow <= 0) return 3;
    float* wt = (float*)malloc(sizeof(float) * (size_t)kh * kw * cin * cout);
    if (!wt) return 3;
    for (int o = 0; o < cout; ++o)
        for (int k = 0; k < kh * kw; ++k)
            for (int ci = 0; ci < cin; ++ci)
                wt[((size_t)k * cin + ci) * cout + o] = w_ohwi[((size_t)o * kh * kw + k) * cin + ci];
    conv_ctx c = { in, out, wt, bias, n, h, w, cin, oh, ow, cout, kh, kw, sh, sw, ph, pw, out_ld, out_off, add_bias, relu, round_bf16 };
    orc_parallel_for((int64_t)n * oh, threads, conv_rows, &c);
    free(wt);
    return 0;
}

/* 3x3 max pool, stride 2, no padding (floor), written into a concat slice */
static void maxpool3s2(const tens* in, float* out, int out_ld, int out_off) {
    const int oh = (in->h - 3) / 2 + 1, ow = (in->w - 3) / 2 + 1;
    for (int b = 0; b < in->n; ++b) for (int y = 0; y < oh; ++y) for (int x = 0; x < ow; ++x) {
        float* o = out + (((size_t)b * oh + y) * ow + x) * out_ld + out_off;
        for (int c = 0; c < in->c; ++c) {
            float m = -INFINITY;
            for (int ky = 0; ky < 3; ++ky) for (int kx = 0; kx < 3; ++kx) {
                const float v = in->d[(((size_t)b * in->h + 2 * y + ky) * in->w + 2 * x + kx) * in->c + c];
                if (v > m) m = v;
            }
            o[c] = m;
        }
    }
}

/* out = [round](relu(avgpool3x3_s1_p1(pre) + bias)), divisor 9 everywhere, taps summed (ky, kx) row-major in fp32 */
static void avgpool_bias_relu(const tens* pre, const float* bias, int round_bf16, float* out, int out_ld, int out_off) {
    const float ninth = 1.0f / 9.0f;
    for (int b = 0; b < pre->n; ++b) for (int y = 0; y < pre->h; ++y) for (int x = 0; x < pre->w; ++x) {
        float* o = out + (((size_t)b * pre->h + y) * pre->w + x) * out_ld + out_off;
        for (int c = 0; c < pre->c; ++c) {
            float s = 0.0f;
            for (int ky = -1; ky <= 1; ++ky) for (int kx = -1; kx <= 1; ++kx) {
                const int iy = y + ky, ix = x + kx;
                if (iy < 0 || iy >= pre->h || ix < 0 || ix >= pre->w) continue;
                s += pre->d[(((size_t)b * pre->h + iy) * pre->w + ix) * pre->c + c];
            }
            float v = s * ninth + bias[c];
            if (!(v > 0.0f)) v = 0.0f;
            o[c] = round_bf16 ? orc_bf16_round(v) : v;
        }
    }
}

/* ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint64_t seed; int classes, mode, threads;
    const char* tap; float* tap_out; int64_t tap_cap; int64_t tap_len; int tap_dims[4];
} net_t;

static void tap(net_t* N, const char* name, const tens* t) {
    if (!N->tap || strcmp(N->tap, name) != 0) return;
    const int64_t len = (int64_t)t->n * t->h * t->w * t->c;
    N->tap_len = len; N->tap_dims[0] = t->n; N->tap_dims[1] = t->h; N->tap_dims[2] = t->w; N->tap_dims[3] = t->c;
    if (N->tap_out && len <= N->tap_cap) memcpy(N->tap_out, t->d, sizeof(float) * (size_t)len);
}

/* Generated layer parameters are kept between forward passes (the product builds its weights once, at create; a CPU
 * arm that regenerated 24 M normal draws per pass would time the generator, not the convolutions).  Keyed by everything
 * that determines the values; single-threaded use (the parallelism is inside a convolution). */
typedef struct wcache { uint64_t seed; char name[96]; int cout, kh, kw, cin, mode; float* w; float* b; struct wcache* next; } wcache;
static wcache* g_wcache = NULL;
static const wcache* layer_params(net_t* N, const char* name, int cout, int kh, int kw, int cin) {
    for (wcache* e = g_wcache; e; e = e->next)
        if (e->seed == N->seed && e->cout == cout && e->kh == kh && e->kw == kw && e->cin == cin && e->mode == N->mode && !strcmp(e->name, name)) return e;
    wcache* e = (wcache*)calloc(1, sizeof(wcache));
    const size_t nw = (size_t)cout * kh * kw * cin;
    e->seed = N->seed; snprintf(e->name, sizeof e->name, "%s", name); e->cout = cout; e->kh = kh; e->kw = kw; e->cin = cin; e->mode = N->mode;
    e->w = (float*)malloc(sizeof(float) * nw);
    e->b = (float*)malloc(sizeof(float) * (size_t)cout);
    orc_inc_conv_params(N->seed, name, cout, kh, kw, cin, e->w, e->b);
    if (N->mode == 1) for (size_t i = 0; i < nw; ++i) e->w[i] = orc_bf16_round(e->w[i]);
    e->next = g_wcache; g_wcache = e;
    return e;
}

/* conv + folded BN + ReLU into a slice of `out`; pre_only: raw conv output (no bias/relu/rounding) */
static void conv_into(net_t* N, const char* name, const tens* in, int cout, int kh, int kw, int s, int ph, int pw,
                      tens* out, int off, int pre_only) {
    const wcache* L = layer_params(N, name, cout, kh, kw, in->c);
    orc_conv2d_nhwc(in->d, in->n, in->h, in->w, in->c, L->w, L->b, cout, kh, kw, s, s, ph, pw, !pre_only, !pre_only,
                    !pre_only && N->mode == 1, out->d, out->c, off, N->threads);
}
static tens conv(net_t* N, const char* name, const tens* in, int cout, int kh, int kw, int s, int ph, int pw) {
    tens out = t_new(in->n, (in->h + 2 * ph - kh) / s + 1, (in->w + 2 * pw - kw) / s + 1, cout);
    conv_into(N, name, in, cout, kh, kw, s, ph, pw, &out, 0, 0);
    tap(N, name, &out);
    return out;
}
static void pool_branch(net_t* N, const char* name, const tens* in, int cout, tens* out, int off) {
    tens pre = t_new(in->n, in->h, in->w, cout);
    conv_into(N, name, in, cout, 1, 1, 1, 0, 0, &pre, 0, 1);
    avgpool_bias_relu(&pre, layer_params(N, name, cout, 1, 1, in->c)->b, N->mode == 1, out->d, out->c, off);
    t_free(&pre);
}
#define NAME(buf, blk, leaf) (snprintf(buf, sizeof buf, "%s.%s", blk, leaf), buf)

static tens inception_a(net_t* N, const char* blk, const tens* x, int pool_features) {
    char nm[96];
    tens out = t_new(x->n, x->h, x->w, 64 + 64 + 96 + pool_features);
    conv_into(N, NAME(nm, blk, "branch1x1"), x, 64, 1, 1, 1, 0, 0, &out, 0, 0);
    tens t = conv(N, NAME(nm, blk, "branch5x5_1"), x, 48, 1, 1, 1, 0, 0);
    conv_into(N, NAME(nm, blk, "branch5x5_2"), &t, 64, 5, 5, 1, 2, 2, &out, 64, 0); t_free(&t);
    t = conv(N, NAME(nm, blk, "branch3x3dbl_1"), x, 64, 1, 1, 1, 0, 0);
    tens u = conv(N, NAME(nm, blk, "branch3x3dbl_2"), &t, 96, 3, 3, 1, 1, 1); t_free(&t);
    conv_into(N, NAME(nm, blk, "branch3x3dbl_3"), &u, 96, 3, 3, 1, 1, 1, &out, 128, 0); t_free(&u);
    pool_branch(N, NAME(nm, blk, "branch_pool"), x, pool_features, &out, 224);
    tap(N, blk, &out);
    return out;
}
static tens inception_b(net_t* N, const char* blk, const tens* x) {
    char nm[96];
    const int oh = (x->h - 3) / 2 + 1, ow = (x->w - 3) / 2 + 1;
    tens out = t_new(x->n, oh, ow, 384 + 96 + x->c);
    conv_into(N, NAME(nm, blk, "branch3x3"), x, 384, 3, 3, 2, 0, 0, &out, 0, 0);
    tens t = conv(N, NAME(nm, blk, "branch3x3dbl_1"), x, 64, 1, 1, 1, 0, 0);
    tens u = conv(N, NAME(nm, blk, "branch3x3dbl_2"), &t, 96, 3, 3, 1, 1, 1); t_free(&t);
    conv_into(N, NAME(nm, blk, "branch3x3dbl_3"), &u, 96, 3, 3, 2, 0, 0, &out, 384, 0); t_free(&u);
    maxpool3s2(x, out.d, out.c, 480);
    tap(N, blk, &out);
    return out;
}
static tens inception_c(net_t* N, const char* blk, const tens* x, int c7) {
    char nm[96];
    tens out = t_new(x->n, x->h, x->w, 768);
    conv_into(N, NAME(nm, blk, "branch1x1"), x, 192, 1, 1, 1, 0, 0, &out, 0, 0);
    tens t = conv(N, NAME(nm, blk, "branch7x7_1"), x, c7, 1, 1, 1, 0, 0);
    tens u = conv(N, NAME(nm, blk, "branch7x7_2"), &t, c7, 1, 7, 1, 0, 3); t_free(&t);
    conv_into(N, NAME(nm, blk, "branch7x7_3"), &u, 192, 7, 1, 1, 3, 0, &out, 192, 0); t_free(&u);
    t = conv(N, NAME(nm, blk, "branch7x7dbl_1"), x, c7, 1, 1, 1, 0, 0);
    u = conv(N, NAME(nm, blk, "branch7x7dbl_2"), &t, c7, 7, 1, 1, 3, 0); t_free(&t);
    t = conv(N, NAME(nm, blk, "branch7x7dbl_3"), &u, c7, 1, 7, 1, 0, 3); t_free(&u);
    u = conv(N, NAME(nm, blk, "branch7x7dbl_4"), &t, c7, 7, 1, 1, 3, 0); t_free(&t);
    conv_into(N, NAME(nm, blk, "branch7x7dbl_5"), &u, 192, 1, 7, 1, 0, 3, &out, 384, 0); t_free(&u);
    pool_branch(N, NAME(nm, blk, "branch_pool"), x, 192, &out, 576);
    tap(N, blk, &out);
    return out;
}
static tens inception_d(net_t* N, const char* blk, const tens* x) {
    char nm[96];
    const int oh = (x->h - 3) / 2 + 1, ow = (x->w - 3) / 2 + 1;
    tens out = t_new(x->n, oh, ow, 320 + 192 + x->c);
    tens t = conv(N, NAME(nm, blk, "branch3x3_1"), x, 192, 1, 1, 1, 0, 0);
    conv_into(N, NAME(nm, blk, "branch3x3_2"), &t, 320, 3, 3, 2, 0, 0, &out, 0, 0); t_free(&t);
    t = conv(N, NAME(nm, blk, "branch7x7x3_1"), x, 192, 1, 1, 1, 0, 0);
    tens u = conv(N, NAME(nm, blk, "branch7x7x3_2"), &t, 192, 1, 7, 1, 0, 3); t_free(&t);
    t = conv(N, NAME(nm, blk, "branch7x7x3_3"), &u, 192, 7, 1, 1, 3, 0); t_free(&u);
    conv_into(N, NAME(nm, blk, "branch7x7x3_4"), &t, 192, 3, 3, 2, 0, 0, &out, 320, 0); t_free(&t);
    maxpool3s2(x, out.d, out.c, 512);
    tap(N, blk, &out);
    return out;
}
static tens inception_e(net_t* N, const char* blk, const tens* x) {
    char nm[96];
    tens out = t_new(x->n, x->h, x->w, 2048);
    conv_into(N, NAME(nm, blk, "branch1x1"), x, 320, 1, 1, 1, 0, 0, &out, 0, 0);
    tens t = conv(N, NAME(nm, blk, "branch3x3_1"), x, 384, 1, 1, 1, 0, 0);
    conv_into(N, NAME(nm, blk, "branch3x3_2a"), &t, 384, 1, 3, 1, 0, 1, &out, 320, 0);
    conv_into(N, NAME(nm, blk, "branch3x3_2b"), &t, 384, 3, 1, 1, 1, 0, &out, 704, 0); t_free(&t);
    t = conv(N, NAME(nm, blk, "branch3x3dbl_1"), x, 448, 1, 1, 1, 0, 0);
    tens u = conv(N, NAME(nm, blk, "branch3x3dbl_2"), &t, 384, 3, 3, 1, 1, 1); t_free(&t);
    conv_into(N, NAME(nm, blk, "branch3x3dbl_3a"), &u, 384, 1, 3, 1, 0, 1, &out, 1088, 0);
    conv_into(N, NAME(nm, blk, "branch3x3dbl_3b"), &u, 384, 3, 1, 1, 1, 0, &out, 1472, 0); t_free(&u);
    pool_branch(N, NAME(nm, blk, "branch_pool"), x, 192, &out, 1856);
    tap(N, blk, &out);
    return out;
}

/* images [n,299,299,3] f32 (mode 1: rounded to bf16 first) -> logits [n,classes].
 * tap_name (optional): copies the named activation (layer or block name, "pool1", "pool2", "gap") into tap_out. */
int orc_inc_forward(uint64_t seed, int classes, int mode, const float* images, int n, int side, float* logits, int threads,
                    const char* tap_name, float* tap_out, int64_t tap_cap, int* tap_dims) {
    net_t net = { seed, classes, mode, threads, tap_name, tap_out, tap_cap, 0, { 0, 0, 0, 0 } };
    net_t* N = &net;
    tens x = t_new(n, side, side, 3);
    for (size_t i = 0; i < (size_t)n * side * side * 3; ++i) x.d[i] = mode == 1 ? orc_bf16_round(images[i]) : images[i];
    tens a = conv(N, "Conv2d_1a_3x3", &x, 32, 3, 3, 2, 0, 0); t_free(&x);
    tens b = conv(N, "Conv2d_2a_3x3", &a, 32, 3, 3, 1, 0, 0); t_free(&a);
    a = conv(N, "Conv2d_2b_3x3", &b, 64, 3, 3, 1, 1, 1); t_free(&b);
    b = t_new(n, (a.h - 3) / 2 + 1, (a.w - 3) / 2 + 1, a.c); maxpool3s2(&a, b.d, b.c, 0); t_free(&a); tap(N, "pool1", &b);
    a = conv(N, "Conv2d_3b_1x1", &b, 80, 1, 1, 1, 0, 0); t_free(&b);
    b = conv(N, "Conv2d_4a_3x3", &a, 192, 3, 3, 1, 0, 0); t_free(&a);
    a = t_new(n, (b.h - 3) / 2 + 1, (b.w - 3) / 2 + 1, b.c); maxpool3s2(&b, a.d, a.c, 0); t_free(&b); tap(N, "pool2", &a);
    b = inception_a(N, "Mixed_5b", &a, 32); t_free(&a);
    a = inception_a(N, "Mixed_5c", &b, 64); t_free(&b);
    b = inception_a(N, "Mixed_5d", &a, 64); t_free(&a);
    a = inception_b(N, "Mixed_6a", &b); t_free(&b);
    b = inception_c(N, "Mixed_6b", &a, 128); t_free(&a);
    a = inception_c(N, "Mixed_6c", &b, 160); t_free(&b);
    b = inception_c(N, "Mixed_6d", &a, 160); t_free(&a);
    a = inception_c(N, "Mixed_6e", &b, 192); t_free(&b);
    b = inception_d(N, "Mixed_7a", &a); t_free(&a);
    a = inception_e(N, "Mixed_7b", &b); t_free(&b);
    b = inception_e(N, "Mixed_7c", &a); t_free(&a);
    /* global average pool: fp32 sum over the pixels in row-major order, times 1/(h*w) */
    tens g = t_new(n, 1, 1, b.c);
    const float inv = 1.0f / (float)(b.h * b.w);
    for (int i = 0; i < n; ++i) for (int c = 0; c < b.c; ++c) {
        float s = 0.0f;
        for (int p = 0; p < b.h * b.w; ++p) s += b.d[((size_t)i * b.h * b.w + p) * b.c + c];
        const float v = s * inv;
        g.d[(size_t)i * b.c + c] = mode == 1 ? orc_bf16_round(v) : v;
    }
    t_free(&b); tap(N, "gap", &g);
    float* w = (float*)malloc(sizeof(float) * (size_t)classes * g.c);
    float* fb = (float*)malloc(sizeof(float) * (size_t)classes);
    orc_inc_fc_params(seed, classes, g.c, w, fb);
    if (mode == 1) for (size_t i = 0; i < (size_t)classes * g.c; ++i) w[i] = orc_bf16_round(w[i]);
    orc_conv2d_nhwc(g.d, n, 1, 1, g.c, w, fb, classes, 1, 1, 1, 1, 0, 0, 1, 0, 0, logits, classes, 0, threads);
    free(w); free(fb); t_free(&g);
    if (tap_dims) { for (int i = 0; i < 4; ++i) tap_dims[i] = net.tap_dims[i]; }
    if (tap_name && net.tap_len == 0) return 5;
    if (tap_name && tap_out && net.tap_len > tap_cap) return 11;
    return 0;
}
