// wire.cpp -- host-side formats either side of the kernels: protobuf cursor, SavedModel / TensorBundle
// reader (models/half_plus_two layout), TensorValue wire image, STRING tensors, tf.Example parsing.
//
// Reference call sites (relative to /root/reference, FT = flink-tensorflow/src/main/scala/org/apache/
// flink/contrib/tensorflow):
//   FT/models/savedmodel/DefaultSavedModelLoader.scala:40-56   SavedModelBundle.load(dir, tags)
//   FT/models/savedmodel/TensorFlowModel.scala:19-26           signatureDef(key)
//   FT/types/TensorValue.java:142-187                          wire image
//   FT/types/TensorInjections.scala:49-72                      STRING tensor of messages
//   FT/types/TensorName.scala:14-19                            "op:idx"
#include "wire.h"
#include "common.h"
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <sys/stat.h>
#include <algorithm>

namespace ftf {

// ------------------------------------------------------------------------------------------------
bool ProtoCursor::read_varint(uint64_t* v) {
    uint64_t r = 0;
    for (int shift = 0; shift < 64 && p_ < end_; shift += 7) {
        const uint8_t c = *p_++;
        r |= (uint64_t)(c & 0x7f) << shift;
        if (!(c & 0x80)) { *v = r; return true; }
    }
    ok_ = false;
    return false;
}

bool ProtoCursor::next() {
    if (done()) return false;
    uint64_t tag;
    if (!read_varint(&tag)) return false;
    field_ = (uint32_t)(tag >> 3);
    wt_ = (uint32_t)(tag & 7);
    dat_ = nullptr; len_ = 0; val_ = 0;
    switch (wt_) {
    case 0: return read_varint(&val_);
    case 1: len_ = 8; break;
    case 5: len_ = 4; break;
    case 2: {
        uint64_t l;
        if (!read_varint(&l)) return false;
        len_ = (size_t)l;
        break;
    }
    default: ok_ = false; return false;
    }
    if ((size_t)(end_ - p_) < len_) { ok_ = false; return false; }
    dat_ = p_;
    p_ += len_;
    return true;
}

// ------------------------------------------------------------------------------------------------
uint32_t crc32c(const uint8_t* p, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82f63b78u : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    uint32_t c = 0xffffffffu;
    for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xff] ^ (c >> 8);
    return c ^ 0xffffffffu;
}

bool dtype_known(int32_t d) {
    switch (d) {
    case FTF_FLOAT: case FTF_DOUBLE: case FTF_INT32: case FTF_STRING: case FTF_INT64: case FTF_BOOL:
    case FTF_UINT8: case FTF_BFLOAT16: return true;
    default: return false;
    }
}
int64_t dtype_size(int32_t d) {
    switch (d) {
    case FTF_FLOAT: case FTF_INT32: return 4;
    case FTF_DOUBLE: case FTF_INT64: return 8;
    case FTF_BOOL: case FTF_UINT8: return 1;
    case FTF_BFLOAT16: return 2;
    default: return 0;
    }
}

// "op:idx"; Scala's String.split(':') drops trailing empty pieces, so "op:" is ("op", 0).
int parse_tensor_name(const char* s, std::string* op, int32_t* index) {
    if (!s) return FTF_INVALID_ARGUMENT;
    std::string str(s);
    while (!str.empty() && str.back() == ':') str.pop_back();
    const size_t colon = str.find(':');
    if (colon == std::string::npos) { *op = str; *index = 0; return FTF_OK; }
    if (str.find(':', colon + 1) != std::string::npos) return FTF_INVALID_ARGUMENT;
    const std::string num = str.substr(colon + 1);
    size_t i = (num[0] == '-' || num[0] == '+') ? 1 : 0;
    if (i == num.size()) return FTF_INVALID_ARGUMENT;
    long long v = 0;
    for (; i < num.size(); ++i) {
        if (num[i] < '0' || num[i] > '9') return FTF_INVALID_ARGUMENT;
        v = v * 10 + (num[i] - '0');
        if (v > 0x7fffffffLL) return FTF_INVALID_ARGUMENT;
    }
    *op = str.substr(0, colon);
    *index = (int32_t)(num[0] == '-' ? -v : v);
    return FTF_OK;
}

int string_tensor_item(const uint8_t* buf, int64_t nbytes, int64_t n, int64_t i, const uint8_t** item,
                       int64_t* len) {
    if (i < 0 || i >= n || nbytes < 8 * n) return FTF_INVALID_ARGUMENT;
    uint64_t rel;
    memcpy(&rel, buf + 8 * i, 8);
    const int64_t table = 8 * n;
    if (rel >= (uint64_t)(nbytes - table)) return FTF_INVALID_ARGUMENT;
    const uint8_t* p = buf + table + rel;
    const uint8_t* end = buf + nbytes;
    uint64_t l = 0;
    int shift = 0;
    for (;;) {
        if (p >= end || shift > 63) return FTF_INVALID_ARGUMENT;
        const uint8_t c = *p++;
        l |= (uint64_t)(c & 0x7f) << shift;
        shift += 7;
        if (!(c & 0x80)) break;
    }
    if ((uint64_t)(end - p) < l) return FTF_INVALID_ARGUMENT;
    *item = p;
    *len = (int64_t)l;
    return FTF_OK;
}

// Dense float feature of shape [1] out of one serialized tf.Example (see ftf.h).
int parse_example_f32(const uint8_t* ex, size_t len, const std::string& key, bool has_default, float def,
                      float* out) {
    bool found = false, other_kind = false;
    int count = 0;
    float value = 0.0f;
    for (ProtoCursor top(ex, len); top.next();) {
        if (top.field() != 1 || top.wire_type() != 2) continue;            // Example.features
        ProtoCursor feats = top.sub();
        while (feats.next()) {
            if (feats.field() != 1 || feats.wire_type() != 2) continue;    // map entry
            std::string k;
            ProtoCursor feature;
            bool have_feature = false;
            ProtoCursor entry = feats.sub();
            while (entry.next()) {
                if (entry.field() == 1 && entry.wire_type() == 2) k = entry.str();
                else if (entry.field() == 2 && entry.wire_type() == 2) { feature = entry.sub(); have_feature = true; }
            }
            if (!entry.ok()) return FTF_INVALID_ARGUMENT;
            if (k != key) continue;
            found = true; other_kind = false; count = 0;                   // last entry wins
            while (have_feature && feature.next()) {
                if (feature.field() == 1 || feature.field() == 3) { other_kind = true; continue; }
                if (feature.field() != 2) continue;
                if (feature.wire_type() != 2) return FTF_INVALID_ARGUMENT;
                ProtoCursor fl = feature.sub();                            // FloatList
                while (fl.next()) {
                    if (fl.field() != 1) continue;
                    if (fl.wire_type() == 5) { value = fl.f32(); ++count; }
                    else if (fl.wire_type() == 2) {
                        if (fl.size() % 4) return FTF_INVALID_ARGUMENT;
                        for (size_t o = 0; o < fl.size(); o += 4) { memcpy(&value, fl.data() + o, 4); ++count; }
                    } else return FTF_INVALID_ARGUMENT;
                }
                if (!fl.ok()) return FTF_INVALID_ARGUMENT;
            }
            if (have_feature && !feature.ok()) return FTF_INVALID_ARGUMENT;
        }
        if (!feats.ok()) return FTF_INVALID_ARGUMENT;
    }
    {   // a truncated top level is malformed too
        ProtoCursor chk(ex, len);
        while (chk.next()) {}
        if (!chk.ok()) return FTF_INVALID_ARGUMENT;
    }
    if (found && other_kind) return FTF_INVALID_ARGUMENT;
    if (!found || count == 0) {
        if (!has_default) return FTF_INVALID_ARGUMENT;
        *out = def;
        return FTF_OK;
    }
    if (count != 1) return FTF_INVALID_ARGUMENT;
    *out = value;
    return FTF_OK;
}

// ------------------------------------------------------------------------------------------------
// SavedModel
// ------------------------------------------------------------------------------------------------
static bool read_file(const std::string& path, std::vector<uint8_t>* out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? fread(out->data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == out->size();
}

static void parse_shape(ProtoCursor c, std::vector<int64_t>* dims) {
    while (c.next())
        if (c.field() == 2 && c.wire_type() == 2) {
            int64_t size = 0;
            ProtoCursor d = c.sub();
            while (d.next()) if (d.field() == 1 && d.wire_type() == 0) size = d.sint();
            dims->push_back(size);
        }
}

static void parse_tensor_info(ProtoCursor c, TensorInfoLite* t) {
    while (c.next()) {
        if (c.field() == 1 && c.wire_type() == 2) t->name = c.str();
        else if (c.field() == 2 && c.wire_type() == 0) t->dtype = (int32_t)c.varint();
        else if (c.field() == 3 && c.wire_type() == 2) parse_shape(c.sub(), &t->dims);
    }
}

static void parse_signature(ProtoCursor c, SignatureLite* s) {
    while (c.next()) {
        if (c.field() == 3 && c.wire_type() == 2) { s->method_name = c.str(); continue; }
        if ((c.field() != 1 && c.field() != 2) || c.wire_type() != 2) continue;
        std::string key;
        TensorInfoLite info;
        ProtoCursor e = c.sub();
        while (e.next()) {
            if (e.field() == 1 && e.wire_type() == 2) key = e.str();
            else if (e.field() == 2 && e.wire_type() == 2) parse_tensor_info(e.sub(), &info);
        }
        (c.field() == 1 ? s->inputs : s->outputs).emplace_back(key, info);
    }
}

static void parse_const_tensor(ProtoCursor c, NodeLite* n) {
    int32_t dtype = 0;
    std::string content;
    while (c.next()) {
        if (c.field() == 1 && c.wire_type() == 0) dtype = (int32_t)c.varint();
        else if (c.field() == 4 && c.wire_type() == 2) content = c.str();
        else if (c.field() == 5 && c.wire_type() == 5) n->f32_values.push_back(c.f32());
        else if (c.field() == 5 && c.wire_type() == 2)
            for (size_t o = 0; o + 4 <= c.size(); o += 4) { float f; memcpy(&f, c.data() + o, 4); n->f32_values.push_back(f); }
        else if (c.field() == 8 && c.wire_type() == 2) n->str_values.push_back(c.str());
    }
    if (dtype == FTF_FLOAT && !content.empty())
        for (size_t o = 0; o + 4 <= content.size(); o += 4) { float f; memcpy(&f, content.data() + o, 4); n->f32_values.push_back(f); }
}

static void parse_node(ProtoCursor c, NodeLite* n) {
    while (c.next()) {
        if (c.wire_type() != 2) continue;
        if (c.field() == 1) n->name = c.str();
        else if (c.field() == 2) n->op = c.str();
        else if (c.field() == 3) n->inputs.push_back(c.str());
        else if (c.field() == 5) {                                 // attr map entry
            std::string key;
            ProtoCursor val;
            bool have = false;
            ProtoCursor e = c.sub();
            while (e.next()) {
                if (e.field() == 1 && e.wire_type() == 2) key = e.str();
                else if (e.field() == 2 && e.wire_type() == 2) { val = e.sub(); have = true; }
            }
            while (have && val.next()) {
                if (val.field() == 3 && val.wire_type() == 0) n->int_attrs[key] = val.sint();
                else if (val.field() == 8 && val.wire_type() == 2 && key == "value") parse_const_tensor(val.sub(), n);
            }
        }
    }
}

static int load_bundle(const std::string& dir, SavedModelLite* m, std::string* err) {
    std::vector<uint8_t> idx;
    const std::string ipath = dir + "/variables/variables.index";
    if (!read_file(ipath, &idx)) return FTF_OK;        // a SavedModel without variables is legal
    // LevelDB-format table: ... | footer(48) = metaindex handle, index handle, padding, magic
    static const uint8_t magic[8] = {0x57, 0xfb, 0x80, 0x8b, 0x24, 0x75, 0x47, 0xdb};
    if (idx.size() < 48 || memcmp(idx.data() + idx.size() - 8, magic, 8) != 0) {
        *err = ipath + ": not a tensor bundle index (bad table magic)";
        return FTF_INVALID_ARGUMENT;
    }
    auto varint = [&](size_t* pos, uint64_t* v) {
        uint64_t r = 0;
        for (int s = 0; s < 64 && *pos < idx.size(); s += 7) {
            const uint8_t c = idx[(*pos)++];
            r |= (uint64_t)(c & 0x7f) << s;
            if (!(c & 0x80)) { *v = r; return true; }
        }
        return false;
    };
    size_t fp = idx.size() - 48;
    uint64_t mo, ms, io, is;
    if (!varint(&fp, &mo) || !varint(&fp, &ms) || !varint(&fp, &io) || !varint(&fp, &is) ||
        is > idx.size() - 5 || io > idx.size() - 5 - is) {          // no sums: offsets near 2^64 must not wrap past the check
        *err = ipath + ": corrupt table footer";
        return FTF_INVALID_ARGUMENT;
    }
    // Walks one block's entries (prefix-compressed keys; restart array at the tail).
    auto walk = [&](uint64_t off, uint64_t size, auto&& fn) -> bool {
        if (size < 4 || size > idx.size() - 5 || off > idx.size() - 5 - size) return false;      // idx.size() >= 48; no wrapping sums
        if (idx[off + size] != 0) return false;                              // compressed blocks unsupported
        uint32_t stored;
        memcpy(&stored, &idx[off + size + 1], 4);
        if (crc32c_mask(crc32c(&idx[off], size + 1)) != stored) return false;
        uint32_t restarts;
        memcpy(&restarts, &idx[off + size - 4], 4);
        if ((uint64_t)restarts * 4 + 4 > size) return false;
        const size_t limit = off + size - 4 - (size_t)restarts * 4;
        size_t pos = off;
        std::string key;
        while (pos < limit) {
            uint64_t shared, fresh, vlen;
            if (!varint(&pos, &shared) || !varint(&pos, &fresh) || !varint(&pos, &vlen)) return false;
            if (shared > key.size() || fresh > limit - pos || vlen > limit - pos - fresh) return false;
            key.resize(shared);
            key.append((const char*)&idx[pos], fresh);
            pos += fresh;
            fn(key, &idx[pos], (size_t)vlen);
            pos += vlen;
        }
        return true;
    };
    bool ok = true, header_seen = false;
    int shards = 1;
    const bool walked = walk(io, is, [&](const std::string&, const uint8_t* v, size_t n) {
        // value = BlockHandle of a data block
        uint64_t bo = 0, bs = 0;
        size_t pos = (size_t)(v - idx.data());
        const size_t value_end = pos + n;
        if (!varint(&pos, &bo) || !varint(&pos, &bs) || pos > value_end) { ok = false; return; }
        ok = ok && walk(bo, bs, [&](const std::string& key, const uint8_t* val, size_t vn) {
            ProtoCursor c(val, vn);
            if (key.empty()) {                                       // BundleHeaderProto
                header_seen = true;
                while (c.next()) if (c.field() == 1 && c.wire_type() == 0) shards = (int)c.varint();
                return;
            }
            BundleEntry e;                                            // BundleEntryProto
            while (c.next()) {
                if (c.field() == 1 && c.wire_type() == 0) e.dtype = (int32_t)c.varint();
                else if (c.field() == 2 && c.wire_type() == 2) parse_shape(c.sub(), &e.dims);
                else if (c.field() == 3 && c.wire_type() == 0) e.shard = (int32_t)c.varint();
                else if (c.field() == 4 && c.wire_type() == 0) e.offset = c.sint();
                else if (c.field() == 5 && c.wire_type() == 0) e.size = c.sint();
                else if (c.field() == 6 && c.wire_type() == 5) memcpy(&e.crc32c, c.data(), 4);
            }
            m->variables[key] = e;
        });
    });
    if (!walked || !ok || !header_seen) {
        *err = ipath + ": corrupt table block (bad crc, handle or entry)";
        return FTF_INVALID_ARGUMENT;
    }
    if (shards != 1) {
        *err = ipath + ": bundles with more than one shard are not supported";
        return FTF_UNIMPLEMENTED;
    }
    const std::string dpath = dir + "/variables/variables.data-00000-of-00001";
    if (!read_file(dpath, &m->variable_data)) {
        *err = dpath + ": cannot read";
        return FTF_NOT_FOUND;
    }
    return FTF_OK;
}

int SavedModelLite::load(const std::string& dir, const std::vector<std::string>& want_tags, std::string* err) {
    std::vector<uint8_t> pb;
    const std::string path = dir + "/saved_model.pb";
    if (!read_file(path, &pb)) {
        *err = "SavedModel not found: " + path;
        return FTF_NOT_FOUND;
    }
    bool picked = false;
    ProtoCursor top(pb.data(), pb.size());
    while (top.next()) {
        if (top.field() != 2 || top.wire_type() != 2 || picked) continue;     // SavedModel.meta_graphs
        std::vector<std::string> mg_tags;
        ProtoCursor graph_def;
        bool have_graph = false;
        std::vector<ProtoCursor> sig_entries;
        ProtoCursor mg = top.sub();
        while (mg.next()) {
            if (mg.wire_type() != 2) continue;
            if (mg.field() == 1) {                                            // MetaInfoDef
                ProtoCursor mi = mg.sub();
                while (mi.next()) if (mi.field() == 4 && mi.wire_type() == 2) mg_tags.push_back(mi.str());
            } else if (mg.field() == 2) { graph_def = mg.sub(); have_graph = true; }
            else if (mg.field() == 5) sig_entries.push_back(mg.sub());
        }
        bool match = true;
        for (const auto& t : want_tags)
            if (std::find(mg_tags.begin(), mg_tags.end(), t) == mg_tags.end()) match = false;
        if (!match || !mg.ok()) continue;
        picked = true;
        tags = mg_tags;
        for (auto& e : sig_entries) {
            std::string key;
            SignatureLite sig;
            while (e.next()) {
                if (e.field() == 1 && e.wire_type() == 2) key = e.str();
                else if (e.field() == 2 && e.wire_type() == 2) parse_signature(e.sub(), &sig);
            }
            signatures[key] = sig;
        }
        while (have_graph && graph_def.next()) {
            if (graph_def.field() != 1 || graph_def.wire_type() != 2) continue;
            NodeLite n;
            parse_node(graph_def.sub(), &n);
            nodes[n.name] = n;
        }
    }
    if (!top.ok()) {
        *err = path + ": malformed protobuf";
        return FTF_INVALID_ARGUMENT;
    }
    if (!picked) {
        *err = path + ": no MetaGraph carries the requested tags";
        return FTF_NOT_FOUND;
    }
    return load_bundle(dir, this, err);
}

int SavedModelLite::load_variables_only(const std::string& dir, std::string* err) {
    variables.clear(); variable_data.clear();
    const int rc = load_bundle(dir, this, err);
    if (rc) return rc;
    if (variables.empty()) { *err = dir + "/variables/variables.index: not found or empty"; return FTF_NOT_FOUND; }
    return FTF_OK;
}

int SavedModelLite::tensor_f32(const std::string& name, const std::vector<int64_t>& want_dims, const float** out, std::string* err) const {
    auto it = variables.find(name);
    if (it == variables.end()) { *err = "tensor '" + name + "' is not in the bundle"; return FTF_NOT_FOUND; }
    const BundleEntry& e = it->second;
    int64_t elems = 1;
    for (int64_t d : want_dims) elems *= d;
    if (e.dtype != FTF_FLOAT || e.dims != want_dims) {
        std::string have = "[", want = "[";
        for (int64_t d : e.dims) have += std::to_string(d) + ",";
        for (int64_t d : want_dims) want += std::to_string(d) + ",";
        *err = "tensor '" + name + "': expected FLOAT " + want + "], the bundle holds dtype " + std::to_string(e.dtype) + " " + have + "]";
        return FTF_INVALID_ARGUMENT;
    }
    if (e.size != elems * 4 || e.offset < 0 || (uint64_t)e.size > variable_data.size() || (uint64_t)e.offset > variable_data.size() - (uint64_t)e.size) {
        *err = "tensor '" + name + "': entry does not fit the data file";
        return FTF_INVALID_ARGUMENT;
    }
    if (crc32c_mask(crc32c(variable_data.data() + e.offset, (size_t)e.size)) != e.crc32c) {
        *err = "tensor '" + name + "': crc32c mismatch (corrupt variables.data file)";
        return FTF_INVALID_ARGUMENT;
    }
    if (e.offset % 4) { *err = "tensor '" + name + "': unaligned offset"; return FTF_INVALID_ARGUMENT; }
    *out = reinterpret_cast<const float*>(variable_data.data() + e.offset);
    return FTF_OK;
}

// ---- TensorBundle writer (fp32 tensors, one shard): the inverse of load_bundle --------------------------------------------
static void put_varint(std::string* o, uint64_t v) {
    while (v >= 0x80) { o->push_back((char)(v | 0x80)); v >>= 7; }
    o->push_back((char)v);
}
static void put_u32(std::string* o, uint32_t v) { o->append((const char*)&v, 4); }
// one table block: entries (no key sharing, every entry a restart point) + restart array + trailer; returns its handle
static void put_block(std::string* file, const std::vector<std::pair<std::string, std::string>>& kv, uint64_t* off, uint64_t* size) {
    std::string b;
    std::vector<uint32_t> restarts;
    for (auto& e : kv) {
        restarts.push_back((uint32_t)b.size());
        put_varint(&b, 0); put_varint(&b, e.first.size()); put_varint(&b, e.second.size());
        b += e.first; b += e.second;
    }
    if (restarts.empty()) restarts.push_back(0);
    for (uint32_t r : restarts) put_u32(&b, r);
    put_u32(&b, (uint32_t)restarts.size());
    *off = file->size(); *size = b.size();
    b.push_back((char)0);                                                       // block type: uncompressed
    const uint32_t crc = crc32c_mask(crc32c((const uint8_t*)b.data(), b.size()));
    *file += b;
    put_u32(file, crc);
}

int TensorBundleWriter::open(const std::string& dir_, std::string* err) {
    dir = dir_;
    const std::string cmd_dir = dir + "/variables";
    if (mkdir(dir.c_str(), 0777) != 0 && errno != EEXIST) { *err = "cannot create " + dir; return FTF_INVALID_ARGUMENT; }
    if (mkdir(cmd_dir.c_str(), 0777) != 0 && errno != EEXIST) { *err = "cannot create " + cmd_dir; return FTF_INVALID_ARGUMENT; }
    data = fopen((cmd_dir + "/variables.data-00000-of-00001").c_str(), "wb");
    if (!data) { *err = "cannot write " + cmd_dir + "/variables.data-00000-of-00001"; return FTF_INVALID_ARGUMENT; }
    offset = 0;
    entries.clear();
    return FTF_OK;
}
int TensorBundleWriter::add_f32(const std::string& name, const std::vector<int64_t>& dims, const float* v, std::string* err) {
    size_t elems = 1;
    for (int64_t d : dims) elems *= (size_t)d;
    if (!data || name.empty() || entries.count(name)) { *err = "bundle writer: bad or duplicate tensor name '" + name + "'"; return FTF_INVALID_ARGUMENT; }
    if (fwrite(v, 4, elems, data) != elems) { *err = "bundle writer: short write"; return FTF_INTERNAL; }
    BundleEntry e;
    e.dtype = FTF_FLOAT; e.dims = dims; e.offset = offset; e.size = (int64_t)elems * 4;
    e.crc32c = crc32c_mask(crc32c((const uint8_t*)v, elems * 4));
    entries[name] = e;
    offset += e.size;
    return FTF_OK;
}
int TensorBundleWriter::finish(std::string* err) {
    if (!data) { *err = "bundle writer: not open"; return FTF_FAILED_PRECONDITION; }
    const bool ok = fclose(data) == 0;
    data = nullptr;
    if (!ok) { *err = "bundle writer: close failed"; return FTF_INTERNAL; }
    std::vector<std::pair<std::string, std::string>> kv;
    kv.emplace_back(std::string(), std::string("\x08\x01\x1a\x02\x08\x01", 6));    // BundleHeaderProto{num_shards 1, version{producer 1}}
    for (auto& it : entries) {                                                  // std::map: keys in bytewise order, as the table wants
        const BundleEntry& e = it.second;
        std::string shape, pb;
        for (int64_t d : e.dims) { std::string dim; dim.push_back((char)0x08); put_varint(&dim, (uint64_t)d); shape.push_back((char)0x12); put_varint(&shape, dim.size()); shape += dim; }
        pb.push_back((char)0x08); put_varint(&pb, (uint64_t)e.dtype);
        pb.push_back((char)0x12); put_varint(&pb, shape.size()); pb += shape;
        pb.push_back((char)0x20); put_varint(&pb, (uint64_t)e.offset);
        pb.push_back((char)0x28); put_varint(&pb, (uint64_t)e.size);
        pb.push_back((char)0x35); put_u32(&pb, e.crc32c);
        kv.emplace_back(it.first, pb);
    }
    std::string file;
    uint64_t d_off, d_size, m_off, m_size, i_off, i_size;
    put_block(&file, kv, &d_off, &d_size);
    put_block(&file, {}, &m_off, &m_size);                                      // metaindex: empty
    std::string handle;
    put_varint(&handle, d_off); put_varint(&handle, d_size);
    put_block(&file, {{kv.back().first, handle}}, &i_off, &i_size);             // index: one data block
    std::string footer;
    put_varint(&footer, m_off); put_varint(&footer, m_size); put_varint(&footer, i_off); put_varint(&footer, i_size);
    footer.resize(40, (char)0);
    static const uint8_t magic[8] = {0x57, 0xfb, 0x80, 0x8b, 0x24, 0x75, 0x47, 0xdb};
    footer.append((const char*)magic, 8);
    file += footer;
    FILE* f = fopen((dir + "/variables/variables.index").c_str(), "wb");
    if (!f || fwrite(file.data(), 1, file.size(), f) != file.size() || fclose(f) != 0) { *err = "cannot write " + dir + "/variables/variables.index"; return FTF_INTERNAL; }
    return FTF_OK;
}

int SavedModelLite::variable_f32(const std::string& name, float* out, std::string* err) const {
    auto it = variables.find(name);
    if (it == variables.end()) { *err = "variable '" + name + "' is not in the bundle"; return FTF_NOT_FOUND; }
    const BundleEntry& e = it->second;
    int64_t elems = 1;
    for (int64_t d : e.dims) elems *= d;
    if (e.dtype != FTF_FLOAT || elems != 1 || e.size != 4 || e.offset < 0 || variable_data.size() < 4 ||
        (uint64_t)e.offset > (uint64_t)variable_data.size() - 4) {
        *err = "variable '" + name + "' is not a readable float scalar";
        return FTF_INVALID_ARGUMENT;
    }
    if (crc32c_mask(crc32c(variable_data.data() + e.offset, 4)) != e.crc32c) {
        *err = "variable '" + name + "': crc32c mismatch in variables.data";
        return FTF_INVALID_ARGUMENT;
    }
    memcpy(out, variable_data.data() + e.offset, 4);
    return FTF_OK;
}

// ------------------------------------------------------------------------------------------------
static std::string tensor_op(const std::string& t, int* idx) {
    std::string s = t;
    if (!s.empty() && s[0] == '^') s = s.substr(1);
    std::string op;
    int32_t i = 0;
    if (parse_tensor_name(s.c_str(), &op, &i) != FTF_OK) { op = s; i = 0; }
    if (idx) *idx = i;
    return op;
}
static std::string canonical(const std::string& t) {
    int i = 0;
    const std::string op = tensor_op(t, &i);
    return op + ":" + std::to_string(i);
}

// Follows Identity nodes down to a variable; returns its name or "".
static std::string resolve_variable(const SavedModelLite& m, std::string t) {
    for (int hop = 0; hop < 16; ++hop) {
        auto it = m.nodes.find(tensor_op(t, nullptr));
        if (it == m.nodes.end()) return "";
        const NodeLite& n = it->second;
        if (n.op == "VariableV2" || n.op == "Variable") return n.name;
        if (n.op != "Identity" || n.inputs.empty()) return "";
        t = n.inputs[0];
    }
    return "";
}

int plan_affine(const SavedModelLite& m, const SignatureLite& sig, AffinePlan* plan, std::string* err) {
    if (sig.inputs.size() != 1 || sig.outputs.size() != 1) {
        *err = "only single-input single-output signatures are supported by the affine runner";
        return FTF_UNIMPLEMENTED;
    }
    plan->input_key = sig.inputs[0].first;
    plan->output_key = sig.outputs[0].first;
    plan->input_tensor = canonical(sig.inputs[0].second.name);
    plan->output_tensor = canonical(sig.outputs[0].second.name);
    std::string t = plan->output_tensor;
    const NodeLite* add = nullptr;
    for (int hop = 0; hop < 16 && !add; ++hop) {
        auto it = m.nodes.find(tensor_op(t, nullptr));
        if (it == m.nodes.end()) { *err = "output tensor '" + t + "' is not in the graph"; return FTF_NOT_FOUND; }
        if (it->second.op == "Identity" && !it->second.inputs.empty()) { t = it->second.inputs[0]; continue; }
        add = &it->second;
    }
    if (!add || add->op != "Add" || add->inputs.size() != 2) {
        *err = "graph behind '" + plan->output_tensor + "' is not Add(Mul(var, x), var)";
        return FTF_UNIMPLEMENTED;
    }
    const NodeLite* mul = nullptr;
    for (int side = 0; side < 2; ++side) {
        auto it = m.nodes.find(tensor_op(add->inputs[side], nullptr));
        if (it != m.nodes.end() && it->second.op == "Mul" && it->second.inputs.size() == 2) {
            const std::string v = resolve_variable(m, add->inputs[1 - side]);
            if (!v.empty()) { mul = &it->second; plan->b_var = v; break; }
        }
    }
    if (!mul) { *err = "graph behind '" + plan->output_tensor + "' is not Add(Mul(var, x), var)"; return FTF_UNIMPLEMENTED; }
    std::string xt;
    for (int side = 0; side < 2; ++side) {
        const std::string v = resolve_variable(m, mul->inputs[side]);
        if (!v.empty()) { plan->a_var = v; xt = mul->inputs[1 - side]; break; }
    }
    if (plan->a_var.empty()) { *err = "Mul has no variable operand"; return FTF_UNIMPLEMENTED; }
    int rc = m.variable_f32(plan->a_var, &plan->a, err);
    if (rc) return rc;
    rc = m.variable_f32(plan->b_var, &plan->b, err);
    if (rc) return rc;
    // walk the x operand back to where the signature feeds it
    for (int hop = 0; hop < 16; ++hop) {
        if (canonical(xt) == plan->input_tensor) { plan->input_is_examples = false; return FTF_OK; }
        int out_idx = 0;
        auto it = m.nodes.find(tensor_op(xt, &out_idx));
        if (it == m.nodes.end()) break;
        const NodeLite& n = it->second;
        if (n.op == "Identity" && !n.inputs.empty()) { xt = n.inputs[0]; continue; }
        if (n.op == "ParseExample") {
            const auto& ia = n.int_attrs;
            const int64_t nsparse = ia.count("Nsparse") ? ia.at("Nsparse") : 0;
            const int64_t ndense = ia.count("Ndense") ? ia.at("Ndense") : 0;
            const int64_t d = out_idx - 3 * nsparse;
            if (d < 0 || d >= ndense || (int64_t)n.inputs.size() < 2 + nsparse + 2 * ndense) break;
            if (canonical(n.inputs[0]) != plan->input_tensor) break;
            auto key_node = m.nodes.find(tensor_op(n.inputs[2 + nsparse + d], nullptr));
            if (key_node == m.nodes.end() || key_node->second.str_values.size() != 1) break;
            plan->feature_key = key_node->second.str_values[0];
            std::string dt = n.inputs[2 + nsparse + ndense + d];
            for (int h2 = 0; h2 < 8; ++h2) {                  // default: Const, possibly behind Reshape
                auto dn = m.nodes.find(tensor_op(dt, nullptr));
                if (dn == m.nodes.end()) break;
                if (dn->second.op == "Const") {
                    plan->has_default = !dn->second.f32_values.empty();
                    if (plan->has_default) plan->default_value = dn->second.f32_values[0];
                    break;
                }
                if (dn->second.inputs.empty()) break;
                dt = dn->second.inputs[0];
            }
            plan->input_is_examples = true;
            return FTF_OK;
        }
        break;
    }
    *err = "cannot connect signature input '" + plan->input_tensor + "' to the Mul operand";
    return FTF_UNIMPLEMENTED;
}

}  // namespace ftf

// ------------------------------------------------------------------------------------------------
// C ABI: codecs
// ------------------------------------------------------------------------------------------------
using namespace ftf;

static void be32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }
static uint32_t rd32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

extern "C" int64_t ftf_tv_binary_length(int32_t rank, int64_t nbytes) { return 13 + 8 * (int64_t)rank + nbytes; }

extern "C" int ftf_tv_encode(const ftf_tensor* t, uint8_t* out, int64_t cap, int64_t* written) {
    if (!t || !out) { set_thread_error("tv_encode: null argument"); return FTF_INVALID_ARGUMENT; }
    // the reference serialises only the TFUtils codes; the two extension dtypes ride the same format
    if (!dtype_known(t->dtype)) { set_thread_error("tv_encode: unknown dtype code"); return FTF_INVALID_ARGUMENT; }
    if (t->rank < 0 || t->rank > 25) { set_thread_error("unable to create a shape with rank > 25"); return FTF_UNIMPLEMENTED; }
    if (t->nbytes < 0 || t->nbytes > 0x7fffffffLL) { set_thread_error("tv_encode: payload does not fit an i32 length"); return FTF_INVALID_ARGUMENT; }
    const int64_t need = ftf_tv_binary_length(t->rank, t->nbytes);
    if (need > cap) { set_thread_error("tv_encode: output buffer too small"); return FTF_OUT_OF_RANGE; }
    uint8_t* p = out;
    *p++ = 1;
    be32(p, (uint32_t)t->dtype); p += 4;
    be32(p, (uint32_t)t->rank); p += 4;
    for (int i = 0; i < t->rank; ++i) {
        const uint64_t d = (uint64_t)t->dims[i];
        be32(p, (uint32_t)(d >> 32)); be32(p + 4, (uint32_t)d); p += 8;
    }
    be32(p, (uint32_t)t->nbytes); p += 4;
    if (t->nbytes) memcpy(p, t->data, (size_t)t->nbytes);
    if (written) *written = need;
    return FTF_OK;
}

extern "C" int ftf_tv_decode(const uint8_t* wire, int64_t len, ftf_tensor* t, int64_t* dims_buf) {
    if (!wire || !t || !dims_buf) { set_thread_error("tv_decode: null argument"); return FTF_INVALID_ARGUMENT; }
    if (len < 1) { set_thread_error("tv_decode: truncated record"); return FTF_OUT_OF_RANGE; }
    if (wire[0] != 1) { set_thread_error("incompatible tensor value"); return FTF_VERSION_MISMATCH; }
    if (len < 13) { set_thread_error("tv_decode: truncated record"); return FTF_OUT_OF_RANGE; }
    const int32_t dtype = (int32_t)rd32(wire + 1), rank = (int32_t)rd32(wire + 5);
    if (!dtype_known(dtype)) { set_thread_error("tv_decode: unknown dtype code"); return FTF_INVALID_ARGUMENT; }
    if (rank < 0 || rank > 25) { set_thread_error("unable to create a shape with rank > 25"); return FTF_UNIMPLEMENTED; }
    if (len < 13 + 8 * (int64_t)rank) { set_thread_error("tv_decode: truncated record"); return FTF_OUT_OF_RANGE; }
    for (int i = 0; i < rank; ++i)
        dims_buf[i] = (int64_t)(((uint64_t)rd32(wire + 9 + 8 * i) << 32) | rd32(wire + 13 + 8 * i));
    const int64_t nb = (int32_t)rd32(wire + 9 + 8 * rank);
    if (nb < 0 || len < 13 + 8 * (int64_t)rank + nb) { set_thread_error("tv_decode: truncated payload"); return FTF_OUT_OF_RANGE; }
    t->dtype = dtype; t->rank = rank; t->dims = dims_buf;
    t->data = (void*)(wire + 13 + 8 * rank); t->nbytes = nb;
    return FTF_OK;
}

extern "C" int ftf_tensor_name_parse(const char* s, char* op, int32_t op_cap, int32_t* index) {
    std::string o;
    int32_t i = 0;
    if (!s || !op || !index || parse_tensor_name(s, &o, &i) != FTF_OK) {
        set_thread_error(std::string("The value ") + (s ? s : "(null)") +
                         " is not a valid tensor name. Tensor names must be of the form \"<op_name>:<output_index>\".");
        return FTF_INVALID_ARGUMENT;
    }
    if ((int64_t)o.size() + 1 > op_cap) { set_thread_error("tensor_name_parse: op buffer too small"); return FTF_OUT_OF_RANGE; }
    memcpy(op, o.c_str(), o.size() + 1);
    *index = i;
    return FTF_OK;
}

extern "C" int64_t ftf_string_tensor_encode(int64_t n, const uint8_t* const* items, const int64_t* lens,
                                            uint8_t* out, int64_t cap) {
    auto vlen = [](uint64_t v) { int k = 1; while (v >= 0x80) { v >>= 7; ++k; } return k; };
    int64_t need = 8 * n;
    for (int64_t i = 0; i < n; ++i) need += vlen((uint64_t)lens[i]) + lens[i];
    if (!out) return need;
    if (need > cap) { set_thread_error("string_tensor_encode: output buffer too small"); return -FTF_OUT_OF_RANGE; }
    int64_t pos = 8 * n;
    for (int64_t i = 0; i < n; ++i) {
        const uint64_t rel = (uint64_t)(pos - 8 * n);
        memcpy(out + 8 * i, &rel, 8);
        uint64_t v = (uint64_t)lens[i];
        while (v >= 0x80) { out[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
        out[pos++] = (uint8_t)v;
        if (lens[i]) memcpy(out + pos, items[i], (size_t)lens[i]);
        pos += lens[i];
    }
    return pos;
}

extern "C" int ftf_parse_examples_f32(const uint8_t* st, int64_t nbytes, int64_t n, const char* key,
                                      int32_t has_default, float def, float* x) {
    if ((!st && nbytes) || !key || (!x && n)) { set_thread_error("parse_examples: null argument"); return FTF_INVALID_ARGUMENT; }
    const std::string k(key);
    for (int64_t i = 0; i < n; ++i) {
        const uint8_t* item;
        int64_t len;
        if (string_tensor_item(st, nbytes, n, i, &item, &len) != FTF_OK) {
            set_thread_error("parse_examples: malformed STRING tensor at element " + std::to_string(i));
            return FTF_INVALID_ARGUMENT;
        }
        const int rc = parse_example_f32(item, (size_t)len, k, has_default != 0, def, &x[i]);
        if (rc != FTF_OK) {
            set_thread_error("Name: <unknown>, Feature: " + k + " is required but could not be found, or is malformed (example " +
                             std::to_string(i) + ")");
            return rc;
        }
    }
    return FTF_OK;
}

// SavedModel introspection without a device (TensorFlowModel.metagraph / signatureDef,
// TensorFlowModel.scala:16-26): JSON with tags, signatures, bundle entries and, per signature, the affine
// plan when the graph behind it is Add(Mul(var, x), var).
static void json_escape(const std::string& s, std::string* out) {
    for (char ch : s) {
        if (ch == '"' || ch == '\\') { out->push_back('\\'); out->push_back(ch); }
        else if ((unsigned char)ch < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", ch); *out += b; }
        else out->push_back(ch);
    }
}
extern "C" int ftf_savedmodel_describe(const char* dir, const char* tags_csv, char* json_out, int64_t cap,
                                       int64_t* needed) {
    if (!dir) { set_thread_error("savedmodel_describe: null path"); return FTF_INVALID_ARGUMENT; }
    std::vector<std::string> tags;
    std::string t = tags_csv ? tags_csv : "serve";
    for (size_t p = 0; p <= t.size();) {
        const size_t q = t.find(',', p);
        const std::string piece = t.substr(p, q == std::string::npos ? std::string::npos : q - p);
        if (!piece.empty()) tags.push_back(piece);
        if (q == std::string::npos) break;
        p = q + 1;
    }
    SavedModelLite m;
    std::string err;
    int rc = m.load(dir, tags, &err);
    if (rc) { set_thread_error(err); return rc; }
    std::string j = "{\"tags\":[";
    for (size_t i = 0; i < m.tags.size(); ++i) { j += i ? ",\"" : "\""; json_escape(m.tags[i], &j); j += "\""; }
    j += "],\"signatures\":{";
    bool first = true;
    auto infos = [&](const std::vector<std::pair<std::string, TensorInfoLite>>& v) {
        j += "{";
        for (size_t i = 0; i < v.size(); ++i) {
            j += i ? ",\"" : "\""; json_escape(v[i].first, &j); j += "\":{\"name\":\""; json_escape(v[i].second.name, &j);
            j += "\",\"dtype\":" + std::to_string(v[i].second.dtype) + ",\"shape\":[";
            for (size_t d = 0; d < v[i].second.dims.size(); ++d) j += (d ? "," : "") + std::to_string(v[i].second.dims[d]);
            j += "]}";
        }
        j += "}";
    };
    for (const auto& kv : m.signatures) {
        j += first ? "\"" : ",\""; first = false;
        json_escape(kv.first, &j);
        j += "\":{\"method_name\":\""; json_escape(kv.second.method_name, &j); j += "\",\"inputs\":";
        infos(kv.second.inputs); j += ",\"outputs\":"; infos(kv.second.outputs);
        AffinePlan plan; std::string perr;
        if (plan_affine(m, kv.second, &plan, &perr) == FTF_OK) {
            char num[128];
            snprintf(num, sizeof num, "%.9g,\"b\":%.9g", plan.a, plan.b);
            j += std::string(",\"affine\":{\"a\":") + num + ",\"a_var\":\"" + plan.a_var + "\",\"b_var\":\"" + plan.b_var +
                 "\",\"examples\":" + (plan.input_is_examples ? "true" : "false") + ",\"feature_key\":\"";
            json_escape(plan.feature_key, &j);
            snprintf(num, sizeof num, "%.9g", plan.default_value);
            j += std::string("\",\"has_default\":") + (plan.has_default ? "true" : "false") + ",\"default\":" + num + "}";
        }
        j += "}";
    }
    j += "},\"variables\":{";
    first = true;
    for (const auto& kv : m.variables) {
        j += first ? "\"" : ",\""; first = false;
        json_escape(kv.first, &j);
        j += "\":{\"dtype\":" + std::to_string(kv.second.dtype) + ",\"offset\":" + std::to_string(kv.second.offset) +
             ",\"size\":" + std::to_string(kv.second.size) + ",\"crc32c\":" + std::to_string(kv.second.crc32c);
        float f;
        std::string verr;
        if (m.variable_f32(kv.first, &f, &verr) == FTF_OK) { char num[64]; snprintf(num, sizeof num, ",\"value\":%.9g", f); j += num; }
        j += "}";
    }
    j += "},\"nodes\":" + std::to_string(m.nodes.size()) + "}";
    if (needed) *needed = (int64_t)j.size() + 1;
    if (!json_out || cap < (int64_t)j.size() + 1) {
        if (json_out) { set_thread_error("savedmodel_describe: output buffer too small"); return FTF_OUT_OF_RANGE; }
        return FTF_OK;
    }
    memcpy(json_out, j.c_str(), j.size() + 1);
    return FTF_OK;
}
