// runner.h -- the per-subtask runner behind ftf_create/ftf_run/ftf_submit/ftf_poll.
#pragma once
#include "common.h"
#include "wire.h"
#include <chrono>
#include <deque>
#include <memory>
#include <string>
#include <vector>

namespace ftf {

class InceptionV3;  // inception.h

enum class Kind { kAffine, kNormalize, kLabel, kNormalizeLabel };

// grow-only pinned-host or device buffer
struct Buf {
    void* p = nullptr;
    size_t cap = 0;
    bool host = false;
    int ensure(size_t bytes, std::string* err);
    void free_now();
};

// One output of a record: key, dtype, per-record dims and bytes.
struct OutSpec {
    std::string key;
    int32_t dtype = FTF_FLOAT;
    std::vector<int64_t> rec_dims;  // dims after the batch dimension
    size_t rec_bytes = 0;
};

constexpr int kSlots = 3;

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;        // everything enqueued on this slot has finished
    bool busy = false;
    Buf h_in, d_in, d_img;
    std::vector<Buf> h_out, d_out;
    // micro-batch bookkeeping
    int64_t n = 0;
    std::vector<uint64_t> tags;
    std::vector<std::chrono::steady_clock::time_point> t_submit;
    int64_t emitted = 0;
    // record uploads not yet handed to the copy engine: consecutive records that lie back to back in one pinned allocation (or
    // in the slot's own staging) and go to consecutive places of d_in are uploaded by ONE copy (see flush_upload in runner.cu)
    const uint8_t* pend_src = nullptr;
    size_t pend_off = 0, pend_bytes = 0;      // place in d_in and payload bytes of the run
    size_t pend_rec = 0, pend_stride = 0;     // bytes of one record; distance of consecutive payloads in host memory (0: one record so far)
    int64_t pend_n = 0;
    cudaEvent_t up_done = nullptr;     // behind the slot's last record upload
    bool up_inflight = false;
    uint32_t downloaded = 0;           // bit j: output j of the batch in flight was copied to h_out[j]
    std::vector<int32_t> rec_status;   // per record of a finished batch of JPEG files: FTF_OK or what was wrong with the file
    // records that arrive as JPEG files (the reference's ImageFileTensor): decoder state of this slot (it stages the
    // files of the open micro-batch as they arrive), and the geometry of the decoded images sitting in d_in (jpeg_n > 0)
    ftf_jpeg_batch* jpeg = nullptr;
    int64_t files_n = 0;               // how many records of the open micro-batch are files
    int64_t jpeg_n = 0;
    bool jpeg_uniform = true;
    std::vector<int64_t> joff;
    std::vector<int32_t> jh, jw;
};

}  // namespace ftf

struct ftf_runner {
    ftf::Kind kind;
    ftf_config cfg;
    int device = 0;
    std::string err;
    bool dead = false;                 // sticky CUDA error: every later call fails fast

    // model state
    ftf::SavedModelLite saved_model;
    ftf::SignatureLite signature;
    ftf::AffinePlan affine;
    std::unique_ptr<ftf::InceptionV3> net;
    cudaStream_t net_stream = nullptr;
    cudaEvent_t net_done = nullptr;

    // record geometry, fixed by the first record/run that fixes it
    int in_h = 0, in_w = 0;            // normalise input image size
    size_t in_rec_bytes = 0;
    std::vector<ftf::OutSpec> outs;
    std::string in_key;

    ftf::Slot slots[ftf::kSlots];
    int next_slot = 0;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;

    // ftf_run result storage (pinned) + dims handed out to the caller
    std::vector<ftf::Buf> run_out;
    std::vector<std::vector<int64_t>> run_dims;
    std::vector<float> parse_tmp;

    // micro-batch state
    int open_slot = -1;                // slot collecting records
    std::deque<int> in_flight;         // launched slots, oldest first
    std::vector<std::vector<int64_t>> poll_dims;
    std::vector<int> poll_sel{0};      // outputs handed out per record by ftf_poll
    std::vector<int32_t> poll_status;  // status of each record the last ftf_poll handed out (ftf_poll_status)
    std::string poll_error;            // message of the first bad record among them
    std::vector<double> latencies_us;

    ftf_stats stats{};
};
