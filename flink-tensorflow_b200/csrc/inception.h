// inception.h -- the label model: Inception-v3 (random-init, BN folded) on device buffers.
//
// Stands where the reference loads its frozen graph and fetches "output" for a fed "input":
// flink-tensorflow-examples/src/main/scala/org/apache/flink/contrib/tensorflow/examples/inception/
// InceptionModel.scala:51-67.  The reference ships no graph file (models/README.md:5), so the
// architecture and the weight recipe are defined here and restated by oracle/inception_oracle.c.
#pragma once
#include "common.h"
#include <memory>
#include <string>
#include <vector>

namespace ftf {

class InceptionV3 {
public:
    ~InceptionV3();
    // Builds the network for up to max_batch images: generates the weights (seeded), uploads them,
    // allocates the activation arena.  compute_dtype: FTF_BFLOAT16 (tcgen05 path).
    // weights_dir (optional): a TensorBundle directory with trained, BatchNorm-folded weights ("<layer>/weights" fp32
    // [Cout,KH,KW,Cin], "<layer>/biases", "fc/weights" [classes,2048], "fc/biases"); the seed is then unused.
    static int create(int device, int num_classes, uint64_t seed, int max_batch, int compute_dtype,
                      std::unique_ptr<InceptionV3>* out, std::string* err, const char* weights_dir = nullptr);
    // Writes this (seeded) network's weights as such a directory: the template of the layout and the test of the reader.
    int export_weights(const std::string& dir, std::string* err);
    // images: [n,299,299,3] NHWC in HBM, bf16 or fp32 (in_dtype FTF_BFLOAT16 / FTF_FLOAT; fp32 values are rounded to bf16 as
    // the first convolution loads them); logits: [n,classes] f32.  Enqueues on s.
    int forward(const void* images, int in_dtype, int64_t n, float* logits, cudaStream_t s, std::string* err);
    int num_classes() const { return classes_; }
    int max_batch() const { return max_batch_; }
    // Diagnostic: copies the named activation (layer or block name of oracle/inception_oracle.c) of the last
    // forward to host fp32 [n,h,w,c]; host_out == NULL only reports dims.
    int activation(const std::string& name, int64_t n, float* host_out, int64_t cap, int dims[4], std::string* err);
    // test hooks: copy of a named layer's weights/bias (host fp32, layout [Cout][KH][KW][Cin]) and a
    // device pointer to a named activation of the last forward.
    struct Impl;
    Impl* impl() { return impl_; }

private:
    InceptionV3() = default;
    Impl* impl_ = nullptr;
    int classes_ = 0, max_batch_ = 0;
};

// f32 -> bf16 (RNE) cast of n elements, used when label records arrive as fp32 images.
int launch_cast_f32_bf16(const float* in, void* out_bf16, int64_t n, cudaStream_t s, std::string* err);

}  // namespace ftf
