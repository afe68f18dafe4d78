// wire.h -- protobuf wire-format cursor and the host-side record codecs of the path.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <map>
#include <string>
#include <vector>

namespace ftf {

// Forward-only cursor over a protobuf-encoded byte range.
class ProtoCursor {
public:
    ProtoCursor() : p_(nullptr), end_(nullptr), ok_(true) {}
    ProtoCursor(const uint8_t* p, size_t n) : p_(p), end_(p + n), ok_(true) {}
    bool ok() const { return ok_; }
    bool done() const { return !ok_ || p_ >= end_; }
    // Reads the next tag; returns false at the end or on malformed input (check ok()).
    bool next();
    uint32_t field() const { return field_; }
    uint32_t wire_type() const { return wt_; }
    uint64_t varint() const { return val_; }               // wire type 0
    int64_t sint() const { return (int64_t)val_; }
    const uint8_t* data() const { return dat_; }            // wire types 1, 2, 5
    size_t size() const { return len_; }
    ProtoCursor sub() const { return ProtoCursor(dat_, len_); }
    std::string str() const { return std::string((const char*)dat_, len_); }
    float f32() const { float f; memcpy(&f, dat_, 4); return f; }

private:
    bool read_varint(uint64_t* v);
    const uint8_t* p_;
    const uint8_t* end_;
    bool ok_;
    uint32_t field_ = 0, wt_ = 0;
    uint64_t val_ = 0;
    const uint8_t* dat_ = nullptr;
    size_t len_ = 0;
};

// ---- SavedModel (models/half_plus_two layout) ---------------------------------------------------
struct TensorInfoLite {
    std::string name;   // "op:idx"
    int32_t dtype = 0;
    std::vector<int64_t> dims;
};
struct SignatureLite {
    std::string method_name;
    std::vector<std::pair<std::string, TensorInfoLite>> inputs, outputs;  // proto order
};
struct NodeLite {
    std::string name, op;
    std::vector<std::string> inputs;
    // the few attributes the affine matcher needs
    std::vector<std::string> str_values;   // string contents of a Const "value" tensor
    std::vector<float> f32_values;         // float contents of a Const "value" tensor
    std::map<std::string, int64_t> int_attrs;  // integer attributes (Ndense, Nsparse, ...)
};
struct BundleEntry {
    int32_t dtype = 0;
    std::vector<int64_t> dims;
    int32_t shard = 0;
    int64_t offset = 0, size = 0;
    uint32_t crc32c = 0;
};
struct SavedModelLite {
    std::vector<std::string> tags;
    std::map<std::string, SignatureLite> signatures;
    std::map<std::string, NodeLite> nodes;
    std::map<std::string, BundleEntry> variables;
    std::vector<uint8_t> variable_data;  // shard 0
    // Reads dir/saved_model.pb and dir/variables/*; picks the MetaGraph carrying every tag in `tags`.
    // Returns an FTF status; message in *err.
    int load(const std::string& dir, const std::vector<std::string>& want_tags, std::string* err);
    // float scalar variable by name (checks dtype, size and the masked crc32c of the bundle entry)
    int variable_f32(const std::string& name, float* out, std::string* err) const;
    // Only dir/variables/* (a TensorBundle without a saved_model.pb): the container of a label model's trained weights.
    int load_variables_only(const std::string& dir, std::string* err);
    // A whole fp32 tensor by name: shape and crc32c checked, *out points into variable_data.
    int tensor_f32(const std::string& name, const std::vector<int64_t>& want_dims, const float** out, std::string* err) const;
};

// Writes dir/variables/variables.{index,data-00000-of-00001} (fp32 tensors, one shard) in the TensorBundle format that
// load() reads -- and that TensorFlow's tf.train.load_checkpoint / SavedModel loader read.
struct TensorBundleWriter {
    std::string dir;
    FILE* data = nullptr;
    int64_t offset = 0;
    std::map<std::string, BundleEntry> entries;
    int open(const std::string& dir, std::string* err);
    int add_f32(const std::string& name, const std::vector<int64_t>& dims, const float* v, std::string* err);
    int finish(std::string* err);
    ~TensorBundleWriter() { if (data) fclose(data); }
};

// y = a*x + b recovered from the graph behind one signature output.
struct AffinePlan {
    float a = 0, b = 0;
    std::string a_var, b_var;
    bool input_is_examples = false;   // fed through ParseExample (STRING tensor of tf.Example)
    std::string feature_key;          // dense key parsed for x
    bool has_default = false;
    float default_value = 0;
    std::string input_key, output_key, input_tensor, output_tensor;
};
int plan_affine(const SavedModelLite& m, const SignatureLite& sig, AffinePlan* plan, std::string* err);

uint32_t crc32c(const uint8_t* p, size_t n);
inline uint32_t crc32c_mask(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }

// ---- record codecs ------------------------------------------------------------------------------
bool dtype_known(int32_t d);            // codes of TFUtils.java:10-32 (+ the two extensions)
int64_t dtype_size(int32_t d);          // bytes per element, 0 for STRING
int parse_tensor_name(const char* s, std::string* op, int32_t* index);
int parse_example_f32(const uint8_t* ex, size_t len, const std::string& key, bool has_default, float def,
                      float* out);
int string_tensor_item(const uint8_t* buf, int64_t nbytes, int64_t n, int64_t i, const uint8_t** item,
                       int64_t* len);

}  // namespace ftf
