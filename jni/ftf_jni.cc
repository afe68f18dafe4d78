// ftf_jni.cc -- the JNI side of the boundary: class org.apache.flink.contrib.tensorflow.b200.Native over include/ftf.h.
//
// This file is what replaces TF-Java's libtensorflow_jni for the model-invocation path of flink-tensorflow.  Each native
// method is a thin call into the C ABI; the Scala/Java classes under jvm/ call these where the reference calls
// Session.runner().feed().fetch().runAndFetchMetadata() (FT/models/ModelFunction.scala:46-66), new Session(graph)
// (FT/models/generic/GenericModel.scala:22-32), SavedModelBundle.load (FT/models/savedmodel/DefaultSavedModelLoader.scala:55)
// and Tensor.create / writeTo (FT/types/TensorValue.java:131-133,257-263).
//
// Build (on a machine with a JDK; none exists in the image this repository is developed in, so there the file is only
// type-checked against tests/jni_check/jni.h, a declaration-only subset of the JNI header):
//   g++ -std=c++17 -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude jni/ftf_jni.cc
//       -Lflink-tensorflow_b200 -lftf_b200 -o libftf_jni.so          (one command line)
//
// Conventions: a runner handle travels as a jlong; tensor payloads are DIRECT ByteBuffers (pinned ones from hostAlloc are
// uploaded in place); shapes are long[]; every failing call throws the Java exception TF-Java would have thrown for the
// same TF status (INVALID_ARGUMENT -> IllegalArgumentException, FAILED_PRECONDITION / RESOURCE_EXHAUSTED ->
// IllegalStateException, OUT_OF_RANGE -> IndexOutOfBoundsException, UNIMPLEMENTED -> UnsupportedOperationException).
#if defined(__has_include)
#if __has_include(<jni.h>)
#define FTF_HAVE_JNI 1
#endif
#endif

#ifdef FTF_HAVE_JNI
#include <jni.h>
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include "ftf.h"

namespace {

ftf_runner* handle(jlong h) { return reinterpret_cast<ftf_runner*>(static_cast<intptr_t>(h)); }

// TF status -> the exception class TF-Java maps it to (and Flink's VersionMismatchException for the wire codec)
void throw_status(JNIEnv* env, int rc, const char* msg) {
    const char* cls;
    switch (rc) {
    case FTF_INVALID_ARGUMENT: cls = "java/lang/IllegalArgumentException"; break;
    case FTF_NOT_FOUND: cls = "java/util/NoSuchElementException"; break;
    case FTF_OUT_OF_RANGE: cls = "java/lang/IndexOutOfBoundsException"; break;
    case FTF_UNIMPLEMENTED: cls = "java/lang/UnsupportedOperationException"; break;
    case FTF_VERSION_MISMATCH: cls = "org/apache/flink/types/VersionMismatchException"; break;
    default: cls = "java/lang/IllegalStateException"; break;   // FAILED_PRECONDITION, RESOURCE_EXHAUSTED, INTERNAL
    }
    jclass c = env->FindClass(cls);
    if (c != nullptr) env->ThrowNew(c, msg ? msg : "");
}

// A Java String (may be null) as UTF-8; empty + is_null for null.
struct Utf {
    std::string s;
    bool is_null = true;
    Utf(JNIEnv* env, jstring js) {
        if (js == nullptr) return;
        const char* p = env->GetStringUTFChars(js, nullptr);
        if (p != nullptr) { s = p; is_null = false; env->ReleaseStringUTFChars(js, p); }
    }
    const char* c_str() const { return is_null ? nullptr : s.c_str(); }
};

// One tensor argument: dtype, long[] shape, direct ByteBuffer payload.  Returns false (exception pending) on misuse.
struct TensorArg {
    std::vector<int64_t> dims;
    ftf_tensor t;
};
bool fill_tensor(JNIEnv* env, jint dtype, jlongArray shape, jobject data, TensorArg* out) {
    const jsize rank = shape ? env->GetArrayLength(shape) : 0;
    out->dims.resize(static_cast<size_t>(rank > 0 ? rank : 1));
    if (rank > 0) {
        std::vector<jlong> tmp(static_cast<size_t>(rank));
        env->GetLongArrayRegion(shape, 0, rank, tmp.data());
        for (jsize i = 0; i < rank; ++i) out->dims[static_cast<size_t>(i)] = static_cast<int64_t>(tmp[static_cast<size_t>(i)]);
    }
    void* p = data ? env->GetDirectBufferAddress(data) : nullptr;
    const jlong cap = data ? env->GetDirectBufferCapacity(data) : 0;
    if (data != nullptr && (p == nullptr || cap < 0)) {
        throw_status(env, FTF_INVALID_ARGUMENT, "tensor payloads must be direct ByteBuffers (ByteBuffer.allocateDirect or Native.hostAlloc)");
        return false;
    }
    out->t.dtype = dtype;
    out->t.rank = rank;
    out->t.dims = out->dims.data();
    out->t.data = p;
    out->t.nbytes = static_cast<int64_t>(cap);
    return true;
}

// long[] of an ftf_tensor's dims
jlongArray dims_array(JNIEnv* env, const ftf_tensor& t) {
    jlongArray a = env->NewLongArray(t.rank);
    if (a == nullptr) return nullptr;
    std::vector<jlong> tmp(static_cast<size_t>(t.rank > 0 ? t.rank : 1));
    for (int i = 0; i < t.rank; ++i) tmp[static_cast<size_t>(i)] = static_cast<jlong>(t.dims[i]);
    if (t.rank > 0) env->SetLongArrayRegion(a, 0, t.rank, tmp.data());
    return a;
}

}  // namespace

extern "C" {

#define NATIVE(ret, name) JNIEXPORT ret JNICALL Java_org_apache_flink_contrib_tensorflow_b200_Native_##name

// ---- lifecycle: RichModel.open()/close() ------------------------------------------------------------------------------
NATIVE(jlong, create)(JNIEnv* env, jclass, jstring model, jstring signatureKey, jstring methodName, jint device, jint maxBatch,
                      jint maxDelayUs, jint outH, jint outW, jfloat mean, jfloat scale, jint computeDtype, jint topK, jint classes,
                      jlong seed) {
    Utf m(env, model), sk(env, signatureKey), mn(env, methodName);
    ftf_config cfg;
    ftf_default_config(&cfg);
    cfg.device = device;
    if (maxBatch > 0) cfg.max_batch = maxBatch;
    if (maxDelayUs >= 0) cfg.max_delay_us = maxDelayUs;
    if (outH > 0) cfg.out_h = outH;
    if (outW > 0) cfg.out_w = outW;
    cfg.mean = mean; cfg.scale = scale;
    if (computeDtype > 0) cfg.compute_dtype = computeDtype;
    if (topK >= 0) cfg.top_k = topK;
    if (classes > 0) cfg.num_classes = classes;
    cfg.seed = static_cast<uint64_t>(seed);
    ftf_runner* r = nullptr;
    const int rc = ftf_create(m.c_str(), sk.c_str(), mn.c_str(), &cfg, &r);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(nullptr)); return 0; }
    return static_cast<jlong>(reinterpret_cast<intptr_t>(r));
}

NATIVE(void, destroy)(JNIEnv*, jclass, jlong h) { ftf_destroy(handle(h)); }   // NULL-safe, idempotent per handle value 0

NATIVE(jstring, lastError)(JNIEnv* env, jclass, jlong h) { return env->NewStringUTF(ftf_last_error(handle(h))); }

// kind: 0 method name, 1 input keys, 2 input tensor names, 3 output keys, 4 output tensor names
NATIVE(jobjectArray, signatureInfo)(JNIEnv* env, jclass, jlong h, jint kind) {
    int n = 0;
    while (ftf_signature_info(handle(h), kind, n) != nullptr) ++n;
    jobjectArray a = env->NewObjectArray(n, env->FindClass("java/lang/String"), nullptr);
    for (int i = 0; a != nullptr && i < n; ++i) {
        jstring s = env->NewStringUTF(ftf_signature_info(handle(h), kind, i));
        env->SetObjectArrayElement(a, i, s);
        env->DeleteLocalRef(s);
    }
    return a;
}

NATIVE(jstring, describeSavedModel)(JNIEnv* env, jclass, jstring dir, jstring tagsCsv) {
    Utf d(env, dir), t(env, tagsCsv);
    int64_t need = 0;
    int rc = ftf_savedmodel_describe(d.c_str(), t.c_str(), nullptr, 0, &need);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(nullptr)); return nullptr; }
    std::vector<char> buf(static_cast<size_t>(need) + 1);
    rc = ftf_savedmodel_describe(d.c_str(), t.c_str(), buf.data(), static_cast<int64_t>(buf.size()), &need);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(nullptr)); return nullptr; }
    return env->NewStringUTF(buf.data());
}

// ---- ModelFunction.apply: feed by key, run, fetch by key ---------------------------------------------------------------
// Outputs are written into outDtypes[i] / outShapes[i] / outData[i] (direct buffers over runner-owned pinned memory, valid
// until release() or the next run()).
NATIVE(void, run)(JNIEnv* env, jclass, jlong h, jobjectArray inKeys, jintArray inDtypes, jobjectArray inShapes, jobjectArray inData,
                  jobjectArray outKeys, jintArray outDtypes, jobjectArray outShapes, jobjectArray outData) {
    const jsize n_in = inKeys ? env->GetArrayLength(inKeys) : 0, n_out = outKeys ? env->GetArrayLength(outKeys) : 0;
    std::vector<std::string> ik(static_cast<size_t>(n_in)), ok(static_cast<size_t>(n_out));
    std::vector<const char*> ikp(static_cast<size_t>(n_in)), okp(static_cast<size_t>(n_out));
    std::vector<TensorArg> args(static_cast<size_t>(n_in));
    std::vector<ftf_tensor> in(static_cast<size_t>(n_in)), out(static_cast<size_t>(n_out));
    std::vector<jint> dt(static_cast<size_t>(n_in > 0 ? n_in : 1));
    if (n_in > 0) env->GetIntArrayRegion(inDtypes, 0, n_in, dt.data());
    for (jsize i = 0; i < n_in; ++i) {
        const size_t k = static_cast<size_t>(i);
        jstring js = static_cast<jstring>(env->GetObjectArrayElement(inKeys, i));
        ik[k] = Utf(env, js).s; ikp[k] = ik[k].c_str();
        jlongArray shape = static_cast<jlongArray>(env->GetObjectArrayElement(inShapes, i));
        jobject data = env->GetObjectArrayElement(inData, i);
        if (!fill_tensor(env, dt[k], shape, data, &args[k])) return;
        in[k] = args[k].t;
        env->DeleteLocalRef(js); env->DeleteLocalRef(shape); env->DeleteLocalRef(data);
    }
    for (jsize i = 0; i < n_out; ++i) {
        const size_t k = static_cast<size_t>(i);
        jstring js = static_cast<jstring>(env->GetObjectArrayElement(outKeys, i));
        ok[k] = Utf(env, js).s; okp[k] = ok[k].c_str();
        env->DeleteLocalRef(js);
    }
    const int rc = ftf_run(handle(h), n_in, ikp.data(), in.data(), n_out, okp.data(), out.data());
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(handle(h))); return; }
    for (jsize i = 0; i < n_out; ++i) {
        const ftf_tensor& t = out[static_cast<size_t>(i)];
        const jint d = t.dtype;
        env->SetIntArrayRegion(outDtypes, i, 1, &d);
        jlongArray dims = dims_array(env, t);
        env->SetObjectArrayElement(outShapes, i, dims);
        jobject buf = env->NewDirectByteBuffer(t.data, static_cast<jlong>(t.nbytes));
        env->SetObjectArrayElement(outData, i, buf);
        env->DeleteLocalRef(dims); env->DeleteLocalRef(buf);
    }
}

// ManagedResource release (ModelFunction.scala:68-69): the buffers handed out by run() must not be touched afterwards.
NATIVE(void, release)(JNIEnv*, jclass, jlong h) { ftf_release(handle(h), nullptr, 0); }

// ---- the micro-batched per-record path behind a 1-in/1-out map ----------------------------------------------------------
NATIVE(void, pollSelect)(JNIEnv* env, jclass, jlong h, jobjectArray keys) {
    const jsize n = keys ? env->GetArrayLength(keys) : 0;
    std::vector<std::string> ks(static_cast<size_t>(n));
    std::vector<const char*> kp(static_cast<size_t>(n));
    for (jsize i = 0; i < n; ++i) {
        jstring js = static_cast<jstring>(env->GetObjectArrayElement(keys, i));
        ks[static_cast<size_t>(i)] = Utf(env, js).s;
        kp[static_cast<size_t>(i)] = ks[static_cast<size_t>(i)].c_str();
        env->DeleteLocalRef(js);
    }
    const int rc = ftf_poll_select(handle(h), n, kp.data());
    if (rc != FTF_OK) throw_status(env, rc, ftf_last_error(handle(h)));
}

// Returns false when every batch slot holds undelivered results (RESOURCE_EXHAUSTED): poll, then offer the record again.
NATIVE(jboolean, submit)(JNIEnv* env, jclass, jlong h, jlong tag, jint dtype, jlongArray shape, jobject data) {
    TensorArg a;
    if (!fill_tensor(env, dtype, shape, data, &a)) return JNI_FALSE;
    const int rc = ftf_submit(handle(h), static_cast<uint64_t>(tag), &a.t);
    if (rc == FTF_RESOURCE_EXHAUSTED) return JNI_FALSE;
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(handle(h))); return JNI_FALSE; }
    return JNI_TRUE;
}

// Hands out up to tags.length finished records in submission order.  With k selected outputs, record i's tensors are at
// outDtypes/outShapes/outData[i*k + j]; status[i] is FTF_OK or the status of a record that failed on its own (a corrupt
// JPEG file): that record's map() throws, its neighbours are emitted.  Returns the number of records.
NATIVE(jint, poll)(JNIEnv* env, jclass, jlong h, jboolean block, jint nSelected, jlongArray tags, jintArray status, jintArray outDtypes,
                   jobjectArray outShapes, jobjectArray outData) {
    const jsize cap = env->GetArrayLength(tags);
    const int k = nSelected > 0 ? nSelected : 1;
    std::vector<uint64_t> tg(static_cast<size_t>(cap > 0 ? cap : 1));
    std::vector<ftf_tensor> outs(static_cast<size_t>(cap > 0 ? cap : 1) * static_cast<size_t>(k));
    int32_t n = 0;
    int rc = ftf_poll(handle(h), block ? 1 : 0, tg.data(), outs.data(), cap, &n);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(handle(h))); return 0; }
    std::vector<int32_t> st(static_cast<size_t>(n > 0 ? n : 1), FTF_OK);
    ftf_poll_status(handle(h), st.data(), n, nullptr);
    for (int i = 0; i < n; ++i) {
        const jlong t = static_cast<jlong>(tg[static_cast<size_t>(i)]);
        const jint s = st[static_cast<size_t>(i)];
        env->SetLongArrayRegion(tags, i, 1, &t);
        env->SetIntArrayRegion(status, i, 1, &s);
        for (int j = 0; j < k; ++j) {
            const ftf_tensor& o = outs[static_cast<size_t>(i) * static_cast<size_t>(k) + static_cast<size_t>(j)];
            const jint d = o.dtype;
            env->SetIntArrayRegion(outDtypes, i * k + j, 1, &d);
            jlongArray dims = dims_array(env, o);
            jobject buf = env->NewDirectByteBuffer(o.data, static_cast<jlong>(o.nbytes));
            env->SetObjectArrayElement(outShapes, i * k + j, dims);
            env->SetObjectArrayElement(outData, i * k + j, buf);
            env->DeleteLocalRef(dims); env->DeleteLocalRef(buf);
        }
    }
    return n;
}

// snapshotState() / close(): launch the open batch and wait for everything in flight (results stay to be polled).
NATIVE(void, flush)(JNIEnv* env, jclass, jlong h) {
    const int rc = ftf_flush(handle(h));
    if (rc != FTF_OK) throw_status(env, rc, ftf_last_error(handle(h)));
}

// {records, batches, runs, h2d_bytes, d2h_bytes, kernel_launches, p50 latency us, p99 latency us}
NATIVE(jlongArray, stats)(JNIEnv* env, jclass, jlong h) {
    ftf_stats s;
    memset(&s, 0, sizeof s);
    ftf_stats_get(handle(h), &s);
    const jlong v[8] = {static_cast<jlong>(s.records), static_cast<jlong>(s.batches), static_cast<jlong>(s.runs), static_cast<jlong>(s.h2d_bytes),
                        static_cast<jlong>(s.d2h_bytes), static_cast<jlong>(s.kernel_launches), static_cast<jlong>(s.queue_latency_us_p50),
                        static_cast<jlong>(s.queue_latency_us_p99)};
    jlongArray a = env->NewLongArray(8);
    if (a != nullptr) env->SetLongArrayRegion(a, 0, 8, v);
    return a;
}

// ---- pinned record buffers (TensorValue payloads the copy engine reads in place) ---------------------------------------
NATIVE(jobject, hostAlloc)(JNIEnv* env, jclass, jlong nbytes) {
    void* p = ftf_host_alloc(static_cast<int64_t>(nbytes));
    if (p == nullptr) { throw_status(env, FTF_RESOURCE_EXHAUSTED, "pinned host allocation failed"); return nullptr; }
    return env->NewDirectByteBuffer(p, nbytes);
}
NATIVE(void, hostFree)(JNIEnv* env, jclass, jobject buf) {
    if (buf != nullptr) ftf_host_free(env->GetDirectBufferAddress(buf));
}

// ---- TensorValue wire codec: the native twins of TensorValue.write/read (TensorValue.java:142-187) ----------------------
NATIVE(jlong, tvBinaryLength)(JNIEnv*, jclass, jint rank, jlong nbytes) { return static_cast<jlong>(ftf_tv_binary_length(rank, nbytes)); }

NATIVE(jlong, tvEncode)(JNIEnv* env, jclass, jint dtype, jlongArray shape, jobject data, jobject out) {
    TensorArg a;
    if (!fill_tensor(env, dtype, shape, data, &a)) return 0;
    uint8_t* dst = static_cast<uint8_t*>(env->GetDirectBufferAddress(out));
    int64_t written = 0;
    const int rc = ftf_tv_encode(&a.t, dst, static_cast<int64_t>(env->GetDirectBufferCapacity(out)), &written);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(nullptr)); return 0; }
    return static_cast<jlong>(written);
}

// Decodes in place: header[0] = dtype, header[1] = rank, dims[0..rank), header[2] = payload offset in `wire`, header[3] = payload bytes.
NATIVE(void, tvDecode)(JNIEnv* env, jclass, jobject wire, jlongArray header, jlongArray dims) {
    const uint8_t* src = static_cast<const uint8_t*>(env->GetDirectBufferAddress(wire));
    ftf_tensor t;
    int64_t dbuf[25];
    const int rc = ftf_tv_decode(src, static_cast<int64_t>(env->GetDirectBufferCapacity(wire)), &t, dbuf);
    if (rc != FTF_OK) { throw_status(env, rc, ftf_last_error(nullptr)); return; }
    const jlong hd[4] = {t.dtype, t.rank, static_cast<jlong>(static_cast<const uint8_t*>(t.data) - src), static_cast<jlong>(t.nbytes)};
    env->SetLongArrayRegion(header, 0, 4, hd);
    jlong d[25];
    for (int i = 0; i < t.rank; ++i) d[i] = static_cast<jlong>(dbuf[i]);
    if (t.rank > 0) env->SetLongArrayRegion(dims, 0, t.rank, d);
}

NATIVE(jstring, version)(JNIEnv* env, jclass) { return env->NewStringUTF(ftf_version()); }

}  // extern "C"

#else   // no <jni.h> on this machine
// Nothing to build: the JNI shim needs a JDK's jni.h.  (tests/test_jni_shim_cpu.py type-checks the code above against a
// declaration-only jni.h so that it cannot rot unnoticed.)
#endif
