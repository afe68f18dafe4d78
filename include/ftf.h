/*
 * ftf.h -- C ABI of the B200 runner that replaces TF-Java's Session.run on flink-tensorflow's
 * per-record model-invocation path.  Plain C: pointers and sizes only, no C++/torch/JNI types,
 * no exceptions across the boundary.  Every entry point returns a TF-style status code and
 * records a message retrievable with ftf_last_error().
 *
 * Citations are relative to /root/reference;  FT = flink-tensorflow/src/main/scala/org/apache/
 * flink/contrib/tensorflow,  EX = flink-tensorflow-examples/src/main/scala/org/apache/flink/
 * contrib/tensorflow/examples.  INTEGRATION.md shows the JNI stub that binds each of them.
 *
 * Threading: one runner per operator subtask; a runner is never used from two threads at once
 * (Flink drives open -> map* -> close from one task thread, FT/common/functions/util/
 * ModelAwareFunction.scala:15-17); different runners may be used concurrently.
 */
#ifndef FTF_H_
#define FTF_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Status codes (TF error space; the JNI shim maps them to the Java exceptions TF-Java throws). */
enum {
    FTF_OK = 0,
    FTF_INVALID_ARGUMENT = 3,    /* -> IllegalArgumentException (ModelFunction.scala:35-37,52-53) */
    FTF_NOT_FOUND = 5,
    FTF_RESOURCE_EXHAUSTED = 8,  /* -> IllegalStateException */
    FTF_FAILED_PRECONDITION = 9, /* -> IllegalStateException (TensorFlowModel.scala:40) */
    FTF_OUT_OF_RANGE = 11,       /* -> IndexOutOfBoundsException */
    FTF_UNIMPLEMENTED = 12,      /* -> UnsupportedOperationException (TFUtils.java:75) */
    FTF_INTERNAL = 13,
    FTF_VERSION_MISMATCH = 100   /* -> VersionMismatchException (TensorValue.java:152-154) */
};

/* Tensor dtype codes = TF DataType enum values used by FT/util/TFUtils.java:10-32, plus two
 * the reference never serialises: UINT8 (raw RGB records) and BFLOAT16 (device-chain images). */
enum {
    FTF_FLOAT = 1, FTF_DOUBLE = 2, FTF_INT32 = 3, FTF_UINT8 = 4, FTF_STRING = 7,
    FTF_INT64 = 9, FTF_BOOL = 10, FTF_BFLOAT16 = 14
};

/* == the header fields of FT/types/TensorValue.java:52-54; payload row-major, native-endian. */
typedef struct ftf_tensor {
    int32_t dtype;
    int32_t rank;
    const int64_t* dims;
    void* data;
    int64_t nbytes;
} ftf_tensor;

/* Replaces the constants hard-coded in EX/inception/ImageNormalization.scala:51-54 and the
 * (ignored) Configuration of ModelAwareFunction.open (ModelAwareFunction.scala:15). */
typedef struct ftf_config {
    int32_t device;        /* CUDA ordinal: subtask index % visible GPUs                      */
    int32_t max_batch;     /* micro-batch closes at this many records ...                     */
    int32_t max_delay_us;  /* ... or this long after its first record                         */
    int32_t out_h, out_w;  /* normalise: ResizeBilinear size (reference 224x224)              */
    float mean, scale;     /* normalise: (v - mean) / scale  (reference 117, 1)               */
    int32_t compute_dtype; /* label: FTF_BFLOAT16 (tensor cores) or FTF_FLOAT                 */
    int32_t top_k;         /* labels kept per image (reference toTextLabels take = 3)         */
    int32_t num_classes;   /* label: width of the score row (builtin Inception-v3: 1000)      */
    uint64_t seed;         /* builtin random-init weights                                     */
} ftf_config;

typedef struct ftf_stats {
    uint64_t records, batches, runs;
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t kernel_launches;      /* this library's kernels launched so far on this runner */
    double last_run_device_ms;     /* CUDA-event time of the last ftf_run/batch             */
    double queue_latency_us_p50, queue_latency_us_p99; /* submit -> result ready           */
} ftf_stats;

typedef struct ftf_runner ftf_runner;

void ftf_default_config(ftf_config* cfg);

/* ---- lifecycle:  RichModel.open()/close()  (FT/models/Model.scala:24-44,
 *      FT/models/generic/GenericModel.scala:22-43, FT/models/savedmodel/TensorFlowModel.scala:36-49,
 *      FT/models/savedmodel/DefaultSavedModelLoader.scala:40-56) ------------------------------
 * model: a SavedModel directory in the models/half_plus_two layout (saved_model.pb +
 *        variables/variables.{index,data-00000-of-00001}), or "builtin:image_normalization"
 *        (EX/inception/ImageNormalization.scala:40-85), or "builtin:inception_v3"
 *        (stands where EX/inception/InceptionModel.scala:47-67 loads its frozen graph), or
 *        "builtin:normalize_label" (both chained on the device) -- these two with seeded random-init weights -- or
 *        "inception_v3:<dir>" / "normalize_label:<dir>": the same network with TRAINED weights read from <dir>, where
 *        InceptionModel.scala:51-55 reads its frozen graph file.  <dir>/variables/ is a TensorBundle (the format of
 *        models/half_plus_two/variables) of fp32 tensors with BatchNorm folded in: "<layer>/weights" [Cout,KH,KW,Cin],
 *        "<layer>/biases" [Cout] for every convolution (torchvision's layer names: "Conv2d_1a_3x3", "Mixed_5b.branch1x1",
 *        ...), "fc/weights" [classes,2048], "fc/biases" [classes]; shapes and crc32c are checked, a missing or
 *        misshapen tensor is NOT_FOUND / INVALID_ARGUMENT.  ftf_label_export_weights writes such a directory.
 * signature_key: SignatureDef key for a SavedModel ("regress_x_to_y", "serving_default", ...;
 *        TensorFlowModel.signatureDef, TensorFlowModel.scala:19-26); NULL for builtins.
 * method_name: if non-NULL it must equal the SignatureDef's method name, else INVALID_ARGUMENT
 *        (ModelFunction.scala:37).
 * On failure nothing leaks and *out is NULL (GenericModel.scala:24-31). */
int ftf_create(const char* model, const char* signature_key, const char* method_name,
               const ftf_config* cfg, ftf_runner** out);
/* idempotent, NULL-safe (GenericModel.scala:34-43) */
int ftf_destroy(ftf_runner* r);

/* ---- ModelFunction.apply  (FT/models/ModelFunction.scala:34-78): feed by signature input key,
 * run, fetch by signature output key.  Synchronous; batch = leading dimension.
 * Inputs are borrowed host buffers (pageable, or pinned from ftf_host_alloc for zero staging).
 * Outputs are runner-owned pinned memory, valid until ftf_release (== Tensor.close() in
 * ModelFunction.scala:68-69) or the next ftf_run.  A missing input key is INVALID_ARGUMENT
 * (ModelFunction.scala:52-53); an unknown output key NOT_FOUND. */
int ftf_run(ftf_runner* r, int32_t n_in, const char* const* in_keys, const ftf_tensor* in,
            int32_t n_out, const char* const* out_keys, ftf_tensor* out);
void ftf_release(ftf_runner* r, ftf_tensor* outs, int32_t n);

/* ---- micro-batched per-record path behind a 1-in/1-out map (FT/streaming/package.scala:23-34,
 * FT/common/functions/AbstractFlatMapFunction.scala:12-14).  ftf_submit copies one record into
 * the pinned ring; a batch is launched at max_batch records or max_delay_us; results come back
 * from ftf_poll in submission order.  ftf_flush drains everything in flight (checkpoint/close).
 * outs[i] stays valid until the next ftf_poll or ftf_submit on this runner.  A record whose payload lies in
 * ftf_host_alloc memory is uploaded in place (no staging copy): keep it unchanged until its result has been polled.
 * Consecutive records that lie back to back, or at one constant distance, inside ONE ftf_host_alloc allocation share a
 * single upload (a 1 MB copy runs at 44 GB/s on PCIe 5 x16, a 16 MB copy at 53): fill a pinned ring in order. */
int ftf_submit(ftf_runner* r, uint64_t tag, const ftf_tensor* record);
int ftf_poll(ftf_runner* r, int32_t block, uint64_t* tags, ftf_tensor* outs, int32_t cap,
             int32_t* n_ready);
int ftf_flush(ftf_runner* r);
/* Per-record outcome of the records the LAST ftf_poll handed out, in the same order: FTF_OK, or FTF_INVALID_ARGUMENT for
 * a record whose JPEG file turned out to be corrupt once the device decoded it ("Invalid JPEG data" in TF's DecodeJpeg).
 * ftf_poll itself returns FTF_OK in that case and still hands out every record of the batch: a bad record fails alone,
 * as the reference's per-record map() does (EX/inception/inception.scala:39-46), and its neighbours keep their results.
 * status may hold fewer than n_ready entries (cap); returns the first non-OK status (FTF_OK if none) and, through
 * *message (may be NULL), the text describing it (valid until the next call on this runner). */
int ftf_poll_status(ftf_runner* r, int32_t* status, int32_t cap, const char** message);
/* Which outputs ftf_poll hands out per record (default: the signature's "outputs").  With n_keys > 1 the
 * caller's outs array holds cap * n_keys entries and record i's tensors sit at outs[i*n_keys + j].  Label
 * runners also expose "logits", "top_k_indices", "top_k_scores" (toTextLabels, InceptionModel.scala:81-89). */
int ftf_poll_select(ftf_runner* r, int32_t n_keys, const char* const* keys);

const char* ftf_last_error(const ftf_runner* r); /* r == NULL: last create() error of this thread */
int ftf_stats_get(const ftf_runner* r, ftf_stats* out);

/* SignatureDef introspection (TensorFlowModel.signatureDef + SignatureDef.getInputsMap/
 * getOutputsMap/getMethodName as used by ModelFunction.scala:37,51,59).  kind 0 = method name,
 * 1 = i-th input key, 2 = i-th input tensor name ("op:idx"), 3 = i-th output key, 4 = i-th output
 * tensor name.  Returns NULL past the end. */
const char* ftf_signature_info(const ftf_runner* r, int32_t kind, int32_t i);

/* SavedModel introspection without a device: the MetaGraph's tags, signature map and bundle entries as a
 * JSON document (TensorFlowModel.metagraph / signatureDef, TensorFlowModel.scala:16-26;
 * DefaultSavedModelLoader.scala:22-31 loads a whole second bundle for this).  json_out == NULL: size query
 * through *needed.  tags_csv NULL = "serve". */
int ftf_savedmodel_describe(const char* dir, const char* tags_csv, char* json_out, int64_t cap,
                            int64_t* needed);

/* Pinned host memory for record buffers (the TensorValue payloads of FT/types/TensorValue.java:54
 * when the JNI shim wraps them with NewDirectByteBuffer). */
void* ftf_host_alloc(int64_t nbytes);
void ftf_host_free(void* p);

/* ---- host-side codecs on the path ------------------------------------------------------------
 * TensorValue wire image (TensorValue.java:142-187): 0x01 | i32 dtype | i32 rank | rank x i64 dims
 * | i32 nbytes | payload; header big-endian, payload native order. */
int64_t ftf_tv_binary_length(int32_t rank, int64_t nbytes);
int ftf_tv_encode(const ftf_tensor* t, uint8_t* out, int64_t cap, int64_t* written);
/* decode in place: t->dims points into dims_buf (>= 25 entries), t->data into the wire image */
int ftf_tv_decode(const uint8_t* wire, int64_t len, ftf_tensor* t, int64_t* dims_buf);
/* TensorName.apply (FT/types/TensorName.scala:14-19) */
int ftf_tensor_name_parse(const char* s, char* op, int32_t op_cap, int32_t* index);
/* STRING tensor [n] of byte strings (TensorInjections.scala:49-72).  out == NULL: size query. */
int64_t ftf_string_tensor_encode(int64_t n, const uint8_t* const* items, const int64_t* lens,
                                 uint8_t* out, int64_t cap);
/* Host parse of n serialized tf.Example records into a dense f32 column (the ParseExample node of
 * the half_plus_two graph): writes x[i]; INVALID_ARGUMENT when `key` is missing and has_default==0. */
int ftf_parse_examples_f32(const uint8_t* string_tensor, int64_t nbytes, int64_t n,
                           const char* key, int32_t has_default, float def, float* x);

/* ---- device-level entry points: inputs and outputs already in HBM, launched on `stream`
 * (a cudaStream_t; NULL = default stream).  These are the kernels ftf_run launches; they exist
 * so that callers holding device buffers (and bench.py's device-resident timing) reach them
 * without host copies.  All return after enqueueing. */

/* K1: y = Add(Mul(a, x), b), two fp32 roundings -- the half_plus_two graph nodes "Mul" and "y"
 * (models/half_plus_two/saved_model.pb), reached through ModelFunction.scala:65. */
int ftf_affine_f32_dev(const float* x, float* y, int64_t n, float a, float b, void* stream);

/* K2: Cast -> ResizeBilinear(align_corners=false) -> Sub(mean) -> Div(scale) on raw RGB
 * (EX/inception/ImageNormalization.scala:57-65, EX/common/GraphBuilder.java:21-39).
 * in u8 [n,in_h,in_w,3] -> out [n,out_h,out_w,3] f32 (out_dtype FTF_FLOAT) or bf16. */
int ftf_resize_norm_dev(const uint8_t* in, int64_t n, int32_t in_h, int32_t in_w, int32_t out_h,
                        int32_t out_w, float mean, float scale, void* out, int32_t out_dtype,
                        void* stream);

/* K7: optional row softmax + stable descending top-k (EX/inception/InceptionModel.scala:81-89:
 * zip truncates to n_labels columns, equal scores keep ascending column order).
 * x [m,n] f32 -> probs [m,n] f32 (may be NULL) , idx [m,k] i32, score [m,k] f32.  k <= 32. */
int ftf_softmax_topk_dev(const float* x, int64_t m, int32_t n, int32_t n_labels, int32_t k,
                         int32_t apply_softmax, float* probs, int32_t* idx, float* score,
                         void* stream);

/* K3: one convolution (+ bias, + ReLU) as an implicit GEMM on the tcgen05 tensor cores -- the Conv2D/BatchNorm/
 * Relu nodes of the frozen graph EX/inception/InceptionModel.scala:58-67 runs, with BN folded into weights and
 * bias.  in bf16 NHWC [n,h,w,in_ld] (channels [0,cin) are read), weights bf16 [cout][kh][kw][cin], bias fp32 [cout]
 * or NULL; out bf16 or fp32 [n,oh,ow,out_ld] written at channel offset out_off.  cin, in_ld, out_ld, out_off
 * multiples of 8; cout a multiple of 16; bf16 accumulate in fp32. */
int ftf_conv2d_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld,
                        const void* weights, const float* bias, int32_t cout, int32_t kh, int32_t kw,
                        int32_t stride, int32_t pad_h, int32_t pad_w, int32_t relu, void* out,
                        int32_t out_dtype, int32_t out_ld, int32_t out_off, void* stream);

/* K4: 3x3 max pool, stride 2, no padding (the MaxPool nodes of the same graph).  bf16 NHWC [n,h,w,c] ->
 * [n,(h-3)/2+1,(w-3)/2+1,out_ld] at channel offset out_off; c, out_ld, out_off multiples of 8. */
int ftf_maxpool3x3s2_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t c, void* out,
                              int32_t out_ld, int32_t out_off, void* stream);
/* K5: an Inception block's pool branch, AvgPool(3x3, stride 1, SAME, divisor 9) -> Conv2D 1x1 -> BN -> Relu, computed
 * as relu(avgpool(conv1x1(x)) + bias) (pool and 1x1 convolution commute).  weights bf16 [cout][cin], bias fp32.
 * Uses fp32 scratch of n*h*w*cout elements, allocated and freed in stream order on `stream`. */
int ftf_pool_branch_bf16_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld,
                             const void* weights, const float* bias, int32_t cout, void* out, int32_t out_ld,
                             int32_t out_off, void* stream);
/* K5 without any device allocation.  No counterpart in the reference: it exists for callers of this library that
 * cannot allocate GPU memory when they call -- a Flink task thread on the per-record path that shares its GPU with other
 * subtasks (an allocation or free can synchronise the device and stall their streams), or a deployment that assigns all
 * HBM at open() and treats later allocations as faults.  Same arguments and result as ftf_pool_branch_bf16_dev; the fp32
 * scratch is a fixed FTF_POOL_BRANCH_STATIC_MAX_BYTES array inside the library (resident on every GPU that loads it
 * until the process exits).  n*h*w*cout*4 beyond that size is refused with FTF_RESOURCE_EXHAUSTED.  The first call on a
 * device also looks up the array's device address; make that call before starting any stream capture.  Planning is
 * host work only (the FTF_CONV_TUNE switch that lets other planners time kernels on the device is ignored).  All calls on a
 * device share the array: calls whose GPU work may overlap (other streams, other threads) race and must be ordered by
 * the caller.  Nothing in it needs clearing between calls. */
#define FTF_POOL_BRANCH_STATIC_MAX_BYTES (8ll << 20)
int ftf_pool_branch_bf16_static_dev(const void* in, int32_t n, int32_t h, int32_t w, int32_t cin, int32_t in_ld,
                                    const void* weights, const float* bias, int32_t cout, void* out,
                                    int32_t out_ld, int32_t out_off, void* stream);
/* K6: global average pool, bf16 [n,hw,c] -> bf16 [n,c] (fp32 mean over the pixels). */
int ftf_global_avgpool_bf16_dev(const void* in, int32_t n, int32_t hw, int32_t c, void* out, void* stream);

/* ---- JPEG decoding on the GPU: the DecodeJpeg(channels=3) node the reference's normalisation graph starts with
 * (EX/common/GraphBuilder.java:41-47, EX/inception/ImageNormalization.scala:56-57; the file bytes come from
 * EX/inception/ImageInputFormat.scala:63-80).  8-bit baseline / extended-sequential Huffman files, 1 or 3 components,
 * 4:4:4 / 4:2:2 / 4:2:0, optional restart intervals; results are bit-identical to libjpeg's defaults (integer slow IDCT,
 * fancy upsampling).  Progressive/arithmetic/CMYK/multi-scan files: FTF_UNIMPLEMENTED.  Malformed headers:
 * FTF_INVALID_ARGUMENT at load, and so is a file that ends inside its scan (the reference leaves DecodeJpeg's
 * try_recover_truncated at false); a corrupt entropy-coded segment is found on the device and reported by
 * ftf_jpeg_batch_status ("Invalid JPEG data" in TF).  ftf_run on "builtin:image_normalization" and
 * "builtin:normalize_label" accepts a STRING tensor of files (the reference's ImageFileTensor) and runs these kernels
 * in front of K2. */
/* header fields of one file; host only (no device needed) */
int ftf_jpeg_info(const uint8_t* file, int64_t len, int32_t* height, int32_t* width, int32_t* components);
typedef struct ftf_jpeg_batch ftf_jpeg_batch;
int ftf_jpeg_batch_create(int32_t device, ftf_jpeg_batch** out);
void ftf_jpeg_batch_destroy(ftf_jpeg_batch* b);                  /* NULL-safe */
const char* ftf_jpeg_batch_error(const ftf_jpeg_batch* b);
/* Reads the marker segments of n files (host memory), stages their entropy-coded bytes and starts the upload on
 * `stream`.  Optional outputs, n entries each: byte offset of image i in the decoded output, its height and width.
 * Images are laid out in order, each 16-byte aligned, uint8 [h,w,3]. */
int ftf_jpeg_batch_load(ftf_jpeg_batch* b, int32_t n, const uint8_t* const* files, const int64_t* lens,
                        int64_t* rgb_offsets, int32_t* heights, int32_t* widths, void* stream);
int64_t ftf_jpeg_batch_rgb_bytes(const ftf_jpeg_batch* b);      /* size of the decoded output of the loaded batch   */
int64_t ftf_jpeg_batch_scan_bytes(const ftf_jpeg_batch* b);     /* bytes uploaded for the loaded batch              */
int32_t ftf_jpeg_batch_segments(const ftf_jpeg_batch* b);       /* independently decodable segments (= J1 threads)  */
/* Enqueues J1-J3 on `stream` (the stream of the load, or one ordered after it): rgb_out is device memory of at least
 * ftf_jpeg_batch_rgb_bytes().  May be called repeatedly for one load. */
int ftf_jpeg_batch_decode_dev(ftf_jpeg_batch* b, uint8_t* rgb_out, int64_t cap, void* stream);
/* Waits for the last decode and reports per image FTF_OK or FTF_INVALID_ARGUMENT (per_image may be NULL); returns
 * the first failure. */
int ftf_jpeg_batch_status(ftf_jpeg_batch* b, int32_t* per_image, int32_t n);

/* Label path on device buffers: images [n,299,299,3] (f32 or bf16 per in_dtype) -> logits [n,classes]
 * f32, probs, top-k.  Runs on the runner's stream; the weights are the runner's. Any of
 * logits/probs/idx/score may be NULL. */
int ftf_label_dev(ftf_runner* r, const void* images, int32_t in_dtype, int64_t n, float* logits,
                  float* probs, int32_t* idx, float* score);
/* Writes the weights of a seeded label runner as the weight directory "inception_v3:<dir>" reads (the template of that
 * layout; stands where the reference's users download tensorflow_inception_graph.pb, models/README.md:5). */
int ftf_label_export_weights(ftf_runner* r, const char* dir);
/* Diagnostic (no reference counterpart; TF-Java would fetch an intermediate tensor by name): copies a named
 * activation of a label runner's last forward to host fp32 [n,h,w,c].  Names are torchvision's layer/block names
 * ("Conv2d_2a_3x3", "Mixed_5b", "pool1", "gap", ...).  host_out NULL: dims only. */
int ftf_debug_activation(ftf_runner* r, const char* name, int64_t n, float* host_out, int64_t cap, int32_t* dims4);
/* Diagnostic (no reference counterpart): with FTF_CONV_PROFILE=1 in the environment the convolution kernels add up, per
 * role, the cycles they run and the cycles they wait on each other; this copies the 16 counters (see csrc/conv_tc.h) and
 * optionally clears them.  Synchronises the device. */
int ftf_debug_conv_profile(uint64_t* out16, int32_t reset);
/* Stream the runner launches on (cudaStream_t), for event timing by the caller. */
void* ftf_runner_stream(ftf_runner* r);

/* Number of this library's kernels launched by this process so far (all runners and device-level calls). */
uint64_t ftf_launch_count(void);

const char* ftf_version(void);

#ifdef __cplusplus
}
#endif
#endif /* FTF_H_ */
