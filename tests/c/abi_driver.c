/*
 * abi_driver.c -- drives libftf_b200.so through include/ftf.h from plain C (C99, -pedantic), the way the JNI shim
 * (jni/ftf_jni.cc) does: create -> run -> submit/poll -> flush -> destroy, plus the host-side codecs.  It proves that the
 * boundary needs nothing but C types, and it is the smallest complete example of a binding.
 *
 *   abi_driver <models/half_plus_two dir> [--no-device]
 *
 * --no-device: the host codecs are exercised and ftf_create must fail with FAILED_PRECONDITION (there is no CPU path).
 * Exit status 0 = every check passed.  The known answers are those of FTT/ml/RegressITCase.scala:35.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "ftf.h"

static int failures = 0;
#define CHECK(cond, what)                                                          \
    do {                                                                           \
        if (!(cond)) { fprintf(stderr, "FAILED: %s (%s:%d)\n", what, __FILE__, __LINE__); ++failures; } \
    } while (0)

/* tf.Example{features{feature{"x": float_list[v]}}}: the 17 bytes of SURVEY 8c with the last four replaced */
static void example_x(float v, uint8_t out[17]) {
    static const uint8_t head[13] = {0x0a, 0x0f, 0x0a, 0x0d, 0x0a, 0x01, 0x78, 0x12, 0x08, 0x12, 0x06, 0x0a, 0x04};
    memcpy(out, head, 13);
    memcpy(out + 13, &v, 4);
}

static void host_codecs(void) {
    /* TensorName.apply */
    char op[32];
    int32_t idx = -1;
    CHECK(ftf_tensor_name_parse("y:0", op, (int32_t)sizeof op, &idx) == FTF_OK && strcmp(op, "y") == 0 && idx == 0, "tensor name y:0");
    CHECK(ftf_tensor_name_parse("tf_example", op, (int32_t)sizeof op, &idx) == FTF_OK && idx == 0, "tensor name without index");
    CHECK(ftf_tensor_name_parse("a:b:c", op, (int32_t)sizeof op, &idx) == FTF_INVALID_ARGUMENT, "tensor name with two colons");
    /* TensorValue wire image of int32 [6] = {1..6} (TensorValueTest.java:40-47) */
    {
        static const uint8_t golden_head[21] = {0x01, 0, 0, 0, 3, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 6, 0, 0, 0, 0x18};
        int32_t v[6] = {1, 2, 3, 4, 5, 6};
        int64_t dims[1] = {6}, written = 0, dbuf[25];
        uint8_t wire[64];
        ftf_tensor t, back;
        t.dtype = FTF_INT32; t.rank = 1; t.dims = dims; t.data = v; t.nbytes = 24;
        CHECK(ftf_tv_binary_length(1, 24) == 45, "binary length");
        CHECK(ftf_tv_encode(&t, wire, (int64_t)sizeof wire, &written) == FTF_OK && written == 45, "tv encode");
        CHECK(memcmp(wire, golden_head, 21) == 0 && memcmp(wire + 21, v, 24) == 0, "tv wire bytes");
        CHECK(ftf_tv_decode(wire, written, &back, dbuf) == FTF_OK && back.dtype == FTF_INT32 && back.rank == 1 && back.dims[0] == 6 &&
              back.nbytes == 24 && (uint8_t*)back.data == wire + 21, "tv decode in place");
        wire[0] = 2;
        CHECK(ftf_tv_decode(wire, written, &back, dbuf) == FTF_VERSION_MISMATCH, "tv version mismatch");
    }
    /* STRING tensor of the four RegressITCase examples and the host ParseExample */
    {
        uint8_t ex[4][17], st[128];
        const uint8_t* items[4];
        int64_t lens[4] = {17, 17, 17, 17}, n;
        float x[4] = {-1, -1, -1, -1};
        int i;
        for (i = 0; i < 4; ++i) { example_x((float)i, ex[i]); items[i] = ex[i]; }
        n = ftf_string_tensor_encode(4, items, lens, NULL, 0);
        CHECK(n == 104, "string tensor size");
        CHECK(ftf_string_tensor_encode(4, items, lens, st, (int64_t)sizeof st) == 104, "string tensor encode");
        CHECK(ftf_parse_examples_f32(st, 104, 4, "x", 0, 0.0f, x) == FTF_OK && x[0] == 0.0f && x[1] == 1.0f && x[2] == 2.0f && x[3] == 3.0f, "parse examples");
        CHECK(ftf_parse_examples_f32(st, 104, 4, "missing", 0, 0.0f, x) == FTF_INVALID_ARGUMENT, "required key missing");
        CHECK(ftf_parse_examples_f32(st, 104, 4, "x2", 1, 7.5f, x) == FTF_OK && x[3] == 7.5f, "default value");
    }
}

static int with_device(const char* model_dir) {
    ftf_config cfg;
    ftf_runner* r = NULL;
    int rc, i;
    ftf_default_config(&cfg);
    cfg.max_batch = 3; cfg.max_delay_us = 1000000;
    /* RichModel.open(): wrong method name first (ModelFunction.scala:37), then the real thing */
    rc = ftf_create(model_dir, "regress_x_to_y", "tensorflow/serving/classify", &cfg, &r);
    CHECK(rc == FTF_INVALID_ARGUMENT && r == NULL && strstr(ftf_last_error(NULL), "method name") != NULL, "method name mismatch");
    rc = ftf_create(model_dir, "regress_x_to_y", "tensorflow/serving/regress", &cfg, &r);
    if (rc != FTF_OK) { fprintf(stderr, "ftf_create: %s\n", ftf_last_error(NULL)); return 1; }
    CHECK(strcmp(ftf_signature_info(r, 1, 0), "inputs") == 0 && strcmp(ftf_signature_info(r, 2, 0), "tf_example:0") == 0 &&
          strcmp(ftf_signature_info(r, 3, 0), "outputs") == 0 && strcmp(ftf_signature_info(r, 4, 0), "y:0") == 0, "signature info");
    /* ModelFunction.apply: feed "inputs", fetch "outputs" (RegressITCase: x in {0,1,2,3} -> y in {2.0,2.5,3.0,3.5}) */
    {
        uint8_t ex[4][17], st[128];
        const uint8_t* items[4];
        int64_t lens[4] = {17, 17, 17, 17}, dims[1] = {4};
        const char* in_keys[1] = {"inputs"};
        const char* out_keys[1] = {"outputs"};
        const char* wrong_keys[1] = {"examples"};
        ftf_tensor in, out;
        for (i = 0; i < 4; ++i) { example_x((float)i, ex[i]); items[i] = ex[i]; }
        ftf_string_tensor_encode(4, items, lens, st, (int64_t)sizeof st);
        in.dtype = FTF_STRING; in.rank = 1; in.dims = dims; in.data = st; in.nbytes = 104;
        rc = ftf_run(r, 1, in_keys, &in, 1, out_keys, &out);
        CHECK(rc == FTF_OK && out.dtype == FTF_FLOAT && out.rank == 2 && out.dims[0] == 4 && out.dims[1] == 1 && out.nbytes == 16, "run: shape of y");
        if (rc == FTF_OK) {
            const float* y = (const float*)out.data;
            CHECK(y[0] == 2.0f && y[1] == 2.5f && y[2] == 3.0f && y[3] == 3.5f, "run: known answers");
            ftf_release(r, &out, 1);
            CHECK(out.data == NULL, "release clears the view");
        }
        rc = ftf_run(r, 1, wrong_keys, &in, 1, out_keys, &out);
        CHECK(rc == FTF_INVALID_ARGUMENT && strstr(ftf_last_error(r), "An input tensor named inputs must be provided") != NULL, "missing input key");
    }
    /* the per-record path: 7 records, micro-batches of 3, results in submission order; flush drains the short batch */
    {
        uint8_t rec[7][32];
        uint64_t tags[8];
        ftf_tensor outs[8];
        int32_t n_ready = 0, got = 0, status[8];
        const char* msg = NULL;
        for (i = 0; i < 7; ++i) {
            uint8_t ex[17];
            const uint8_t* item[1];
            int64_t len[1] = {17}, n;
            ftf_tensor t;
            example_x((float)(10 + i), ex);
            item[0] = ex;
            n = ftf_string_tensor_encode(1, item, len, rec[i], 32);
            t.dtype = FTF_STRING; t.rank = 0; t.dims = NULL; t.data = rec[i]; t.nbytes = n;
            CHECK(ftf_submit(r, 100u + (uint64_t)i, &t) == FTF_OK, "submit");
        }
        CHECK(ftf_flush(r) == FTF_OK, "flush");
        while (got < 7) {
            CHECK(ftf_poll(r, 1, tags, outs, 8, &n_ready) == FTF_OK, "poll");
            CHECK(ftf_poll_status(r, status, 8, &msg) == FTF_OK, "poll status");
            for (i = 0; i < n_ready; ++i, ++got) {
                CHECK(tags[i] == 100u + (uint64_t)got, "submission order");
                CHECK(outs[i].nbytes == 4 && *(const float*)outs[i].data == 0.5f * (float)(10 + got) + 2.0f, "record result");
                CHECK(status[i] == FTF_OK, "record status");
            }
            if (n_ready == 0) break;
        }
        CHECK(got == 7, "every record came back");
    }
    {
        ftf_stats st;
        CHECK(ftf_stats_get(r, &st) == FTF_OK && st.records >= 11 && st.kernel_launches >= 4, "stats");
    }
    CHECK(ftf_destroy(r) == FTF_OK, "destroy");
    CHECK(ftf_destroy(NULL) == FTF_OK, "destroy(NULL)");
    return 0;
}

int main(int argc, char** argv) {
    int no_device = argc > 2 && strcmp(argv[2], "--no-device") == 0;
    if (argc < 2) { fprintf(stderr, "usage: %s <half_plus_two dir> [--no-device]\n", argv[0]); return 2; }
    printf("%s\n", ftf_version());
    host_codecs();
    {   /* SavedModel introspection needs no device */
        int64_t need = 0;
        char* json;
        CHECK(ftf_savedmodel_describe(argv[1], NULL, NULL, 0, &need) == FTF_OK && need > 0, "describe: size query");
        json = (char*)malloc((size_t)need + 1);
        CHECK(json != NULL && ftf_savedmodel_describe(argv[1], "serve", json, need + 1, &need) == FTF_OK &&
              strstr(json, "regress_x_to_y") != NULL && strstr(json, "tf_example:0") != NULL, "describe: signatures");
        free(json);
    }
    if (no_device) {
        ftf_runner* r = (ftf_runner*)1;
        int rc = ftf_create(argv[1], "regress_x_to_y", NULL, NULL, &r);
        CHECK(rc == FTF_FAILED_PRECONDITION && r == NULL && strstr(ftf_last_error(NULL), "no CPU path") != NULL, "no device: create fails loudly");
    } else if (with_device(argv[1])) {
        return 1;
    }
    if (failures) { fprintf(stderr, "%d check(s) failed\n", failures); return 1; }
    printf("abi_driver ok\n");
    return 0;
}
