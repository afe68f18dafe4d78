// conv_tc.cu -- chooses between the pixel-major and the weight-major tensor-core convolution kernels.
#include "conv_tc.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <mutex>
#include <vector>

namespace ftf {

// ------------------------------------------------------------------------------------------------
// What the plan-time timing decided, remembered per process: (device, layer geometry, segment layout, SM share) -> the
// variant that won.  The second runner of a process (another Flink slot on the same GPU, or on another GPU of the box)
// plans the same layers and takes the first runner's choice instead of timing again, so that every runner of a process
// runs the same kernels -- per-record results then do not depend on which subtask served the record -- and ftf_create
// of the later runners does no device work for planning.
// ------------------------------------------------------------------------------------------------
namespace {
struct ChoiceKey {
    std::vector<long long> v;
    bool operator<(const ChoiceKey& o) const { return v < o.v; }
};
std::mutex g_tune_mu;                   // guards g_choice and serialises the timing runs
std::map<ChoiceKey, int> g_choice;      // -> index into kVariants

// FTF_CONV_PLAN_FILE=<path>: the remembered choices also live in a text file, one "key numbers = variant" line per layer.
// A process that finds a layer there takes the choice without timing anything; a process that times a layer appends its
// choice.  This carries a tuned plan from one process to the next: a service restart skips the timing runs, every rank of
// a job can be handed rank 0's file, and a profiler run (whose overhead would disturb the timing) replays exactly the plan
// the unprofiled run chose.  Later lines win.
const char* plan_file() {
    static const char* path = getenv("FTF_CONV_PLAN_FILE");
    return (path && path[0]) ? path : nullptr;
}
void load_plan_file_locked() {
    static bool loaded = false;
    if (loaded || !plan_file()) return;
    loaded = true;
    FILE* f = fopen(plan_file(), "r");
    if (!f) return;
    char line[1024];
    while (fgets(line, sizeof line, f)) {
        ChoiceKey k;
        char* p = line;
        bool ok = false;
        int variant = -1;
        for (;;) {
            while (*p == ' ') ++p;
            if (*p == '=') { variant = atoi(p + 1); ok = true; break; }
            char* end = nullptr;
            const long long v = strtoll(p, &end, 10);
            if (end == p) break;
            k.v.push_back(v);
            p = end;
        }
        if (ok && variant >= 0 && variant < kConvVariants && !k.v.empty()) g_choice[k] = variant;
    }
    fclose(f);
}
void append_plan_file_locked(const ChoiceKey& k, int variant) {
    if (!plan_file()) return;
    std::string line;
    for (long long v : k.v) line += std::to_string(v) + " ";
    line += "= " + std::to_string(variant) + "\n";
    FILE* f = fopen(plan_file(), "a");
    if (!f) return;
    fwrite(line.data(), 1, line.size(), f);      // one short write in append mode: lines of concurrent processes do not interleave
    fclose(f);
}

ChoiceKey choice_key(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count) {
    ChoiceKey k;
    // the device is deliberately NOT part of the key: all GPUs of a box are the same part, and subtasks on different GPUs
    // must agree as well
    k.v = {g.n_max, g.H, g.W, g.Cin, g.in_ld, g.Cout, g.KH, g.KW, g.stride, g.pad_h, g.pad_w, sm_count, n_seg, g.bias != nullptr};
    for (int i = 0; i < n_seg; ++i) { k.v.push_back(segs[i].ld); k.v.push_back(segs[i].coff); k.v.push_back(segs[i].col_begin); k.v.push_back(segs[i].flags); }
    return k;
}
}  // namespace

// Median-of-5 device time of one candidate on its own buffers (ms); < 0 on error.
static float time_candidate(const ConvPlan& cand, int n, cudaStream_t s, cudaEvent_t e0, cudaEvent_t e1) {
    std::string err;
    for (int i = 0; i < 2; ++i)
        if (conv_tc_launch(cand, n, s, &err)) return -1.0f;
    float t[5];
    for (int i = 0; i < 5; ++i) {
        if (cudaEventRecord(e0, s) != cudaSuccess || conv_tc_launch(cand, n, s, &err) || cudaEventRecord(e1, s) != cudaSuccess ||
            cudaEventSynchronize(e1) != cudaSuccess || cudaEventElapsedTime(&t[i], e0, e1) != cudaSuccess) { cudaGetLastError(); return -1.0f; }
    }
    for (int i = 1; i < 5; ++i)                            // insertion sort, return the median
        for (int j = i; j > 0 && t[j] < t[j - 1]; --j) { const float x = t[j]; t[j] = t[j - 1]; t[j - 1] = x; }
    return t[2];
}

int conv_tc_plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan,
                 std::string* err, int mode) {
    ConvPlanSet set;
    const int rc = conv_tc_plan_set(g, segs, n_seg, sm_count, &set, err, mode);
    if (rc) return rc;
    *plan = set.cands[set.chosen];
    return FTF_OK;
}

void conv_tc_choice_put(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, int variant) {
    std::lock_guard<std::mutex> lock(g_tune_mu);
    load_plan_file_locked();
    const ChoiceKey k = choice_key(g, segs, n_seg, sm_count);
    auto it = g_choice.find(k);
    if (it != g_choice.end() && it->second == variant) return;
    g_choice[k] = variant;
    append_plan_file_locked(k, variant);
}

int conv_tc_plan_set(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlanSet* set,
                     std::string* err, int mode) {
    // FTF_CONV_KIND=pm|wm forces one kernel (tests and profiling); FTF_CONV_TUNE=0/1 overrides the caller's choice --
    // except for callers that promise to do no device work while planning (kPlanModelOnly: the per-call entry points)
    const char* force = getenv("FTF_CONV_KIND");
    const char* tenv = getenv("FTF_CONV_TUNE");
    const bool halo_only = getenv("FTF_CONV_HALO_ONLY") != nullptr;     // profiling: drop the im2col candidate where a halo one exists
    bool tune = mode == kPlanTune;
    if (tenv && mode != kPlanModelOnly) tune = tenv[0] == '1';
    ConvPlan* cands = set->cands;
    int* cand_variant = set->variant;
    bool eligible[kConvVariants];                    // a candidate of the isolated timing (the CTA pair only on many-wave layers)
    int n_cand = 0;
    std::string first_err;
    // candidates: pixel-major (im2col / halo window / CTA pair), weight-major (im2col / halo window)
    struct Variant { int kind; bool halo, pair, wide; int S; };
    // 1: the halo shape the cost model likes best; 6-8: the halo window with 1, 2, 4 sub-tiles per CTA tile pinned -- the model
    // cannot see how a layer's tile count falls on the SMs it gets (190 tiles on 148 SMs is two rounds, 380 is 2.6), the timing can
    static const Variant kVariants[kConvVariants] = {{0, false, false, false, 0}, {0, true, false, false, 0}, {0, false, true, false, 0},
                                                     {1, false, false, false, 0}, {1, true, false, false, 0}, {0, true, false, true, 0},
                                                     {0, true, false, false, 1}, {0, true, false, false, 2}, {0, true, false, false, 4}};
    // The CTA-pair kernel halves the weight bytes each SM pulls from L2.  That pays on layers of many waves, which are bound by
    // L2 -> SM operand fill (Conv2d_4a: 155 -> 137 us); single-wave layers are bound by the K-block cadence of the issuing
    // warp and a pair only halves the number of workers.  So it is a candidate from four tiles per SM on (FTF_CONV_PAIR=0/1
    // forces it off/on for every layer, FTF_CONV_KIND=pm2 selects it alone).
    const long long m_tiles = ((long long)g.n_max * ((g.H + 2 * g.pad_h - g.KH) / g.stride + 1) * ((g.W + 2 * g.pad_w - g.KW) / g.stride + 1) + 127) / 128;
    const char* pair_env = getenv("FTF_CONV_PAIR");
    const bool no_pair = pair_env ? pair_env[0] != '1' : m_tiles < 4ll * sm_count;
    // Measured: over the 44 distinct layers of Inception-v3 no pinned shape beat the model's halo shape or another family
    // (neither alone nor swapped into the whole forward pass), so they are candidates only on request.
    static const bool s_pinned = getenv("FTF_CONV_S_CANDIDATES") != nullptr;
    for (int vi = 0; vi < kConvVariants; ++vi) {
        const Variant& v = kVariants[vi];
        if (v.S && !s_pinned) continue;
        if (force) {   // FTF_CONV_KIND: "pm" = one-CTA pixel-major, "pm2" = CTA-pair pixel-major, "pmw" = wide halo (falls back to
                       // the other one-CTA pixel-major shapes where a layer has no wide one), "wm" = weight-major
            const bool want_pair = force[0] == 'p' && force[2] == '2', want_wide = force[0] == 'p' && force[2] == 'w';
            if ((force[0] == 'p') != (v.kind == 0) || (v.kind == 0 && want_pair != v.pair)) continue;
            if (v.wide && !want_wide) continue;
        }
        ConvPlan& c = cands[n_cand];
        c.kind = v.kind;
        std::string e;
        const int rc = v.kind ? wm::plan(g, segs, n_seg, sm_count, &c.wm, &e, &c.est_cycles, v.halo)
                              : pm::plan(g, segs, n_seg, sm_count, &c.pm, &e, &c.est_cycles, v.halo, v.pair, v.wide, v.S);
        if (rc == FTF_INTERNAL) { if (err) *err = e; return rc; }   // a CUDA/driver failure, not "this variant does not apply"
        if (rc) { if (first_err.empty()) first_err = e; continue; }
        const bool is_halo = v.kind ? c.wm.p.halo_mode != 0 : c.pm.p.halo_mode != 0;
        if (v.halo && !is_halo) continue;                      // same plan as the no-halo candidate
        if (v.wide && c.pm.p.acc_stages != 1) continue;        // no wide shape exists for this layer
        if (v.S) {                                             // the same shape as the model's halo candidate: once is enough
            bool dup = false;
            for (int k = 0; k < n_cand; ++k)
                if (cand_variant[k] == 1 && cands[k].pm.p.S == c.pm.p.S && cands[k].pm.p.block_n == c.pm.p.block_n && cands[k].pm.p.kc == c.pm.p.kc) dup = true;
            if (dup || force) continue;
        }
        if (v.pair && !c.pm.p.pair) continue;
        cand_variant[n_cand] = vi;
        eligible[n_cand] = force || !(v.pair && no_pair);
        ++n_cand;
    }
    if (n_cand == 0) { if (err) *err = first_err; return FTF_INVALID_ARGUMENT; }
    if (halo_only) {
        int w = 0;
        for (int i = 0; i < n_cand; ++i) {
            const bool hm = cands[i].kind ? cands[i].wm.p.halo_mode != 0 : cands[i].pm.p.halo_mode != 0;
            if (hm) { cand_variant[w] = cand_variant[i]; eligible[w] = eligible[i]; cands[w++] = cands[i]; }
        }
        if (w > 0) n_cand = w;
    }
    int best = -1;
    for (int i = 0; i < n_cand; ++i)
        if (eligible[i] && (best < 0 || cands[i].est_cycles < cands[best].est_cycles)) best = i;
    if (best < 0) best = 0;
    if (force && force[0] == 'p' && force[2] == 'w')
        for (int i = 0; i < n_cand; ++i) if (cand_variant[i] == 5) best = i;
    // A choice another runner of this process already timed for the same layer is taken as it is.  Timing itself runs
    // under a process-wide lock: two task threads that open their models at the same moment would otherwise disturb each
    // other's measurements and could settle on different kernels.
    const bool cacheable = tune && !force && !halo_only && n_cand > 1;
    ChoiceKey key;
    bool cached = false;
    std::unique_lock<std::mutex> tune_lock(g_tune_mu, std::defer_lock);
    if (cacheable) {
        key = choice_key(g, segs, n_seg, sm_count);
        tune_lock.lock();
        load_plan_file_locked();
        auto it = g_choice.find(key);
        if (it != g_choice.end())
            for (int i = 0; i < n_cand; ++i)
                if (cand_variant[i] == it->second) { best = i; cached = true; }
    }
    if (tune && n_cand > 1 && !cached) {
        cudaStream_t s = nullptr;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess && cudaEventCreate(&e0) == cudaSuccess &&
            cudaEventCreate(&e1) == cudaSuccess) {
            // the model's choice stands unless another candidate is measurably (8 %) faster: near-ties then resolve the
            // same way on every runner
            const int model = best;
            const bool verbose = getenv("FTF_CONV_VERBOSE") != nullptr;
            auto describe = [&](int i, float t) {
                const ConvPlan& c = cands[i];
                const bool hm = c.kind ? c.wm.p.halo_mode != 0 : c.pm.p.halo_mode != 0;
                fprintf(stderr, "   candidate %s%s%s S%d kc %d: %.1f us (model %.0f cycles)%s\n", c.kind ? "wm" : (c.pm.p.pair ? "pm pair" : "pm"), hm ? " halo" : "",
                        (!c.kind && hm && c.pm.p.acc_stages == 1) ? " wide" : "", c.kind ? c.wm.p.S : c.pm.p.S,
                        c.kind ? c.wm.p.kc : c.pm.p.kc, t * 1e3f, c.est_cycles, i == model ? "  <- model" : "");
            };
            float best_t = time_candidate(cands[model], g.n_max, s, e0, e1);
            if (verbose) describe(model, best_t);
            for (int i = 0; i < n_cand; ++i) {
                if (i == model || !eligible[i]) continue;
                const float t = time_candidate(cands[i], g.n_max, s, e0, e1);
                if (verbose) describe(i, t);
                if (t >= 0 && (best_t < 0 || t < 0.92f * best_t)) { best_t = t; best = i; }
            }
        }
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        if (s) cudaStreamDestroy(s);
        cudaGetLastError();
        if (cacheable) { g_choice[key] = cand_variant[best]; append_plan_file_locked(key, cand_variant[best]); }
    }
    set->n = n_cand;
    set->chosen = best;
    set->from_cache = cached;
    set->cacheable = cacheable;
    const ConvPlan* plan = &cands[best];
    if (getenv("FTF_CONV_VERBOSE")) {
        const bool hm = plan->kind ? plan->wm.p.halo_mode : plan->pm.p.halo_mode;
        fprintf(stderr, "conv %dx%d s%d Cin %d Cout %d HxW %dx%d -> %s%s kc %d (%d candidates, %s)\n", g.KH, g.KW, g.stride, g.Cin, g.Cout, g.H, g.W,
                plan->kind ? "weight-major" : (plan->pm.p.pair ? "pixel-major pair" : "pixel-major"), hm ? " halo" : "", plan->kind ? plan->wm.p.kc : plan->pm.p.kc, n_cand, cached ? "choice of an earlier runner" : ((tune && n_cand > 1) ? "timed" : "model"));
    }
    return FTF_OK;
}

// The measurement switches a plan depends on, folded into one number (part of the key below: a test that flips one of them
// between two calls on the same buffers must not be served the other setting's plan).
static unsigned long long env_fingerprint() {
    static const char* const kNames[] = {"FTF_CONV_KIND", "FTF_CONV_TUNE", "FTF_CONV_HALO_ONLY", "FTF_CONV_PAIR", "FTF_NO_HALO", "FTF_NO_KC32",
                                         "FTF_PM_NT", "FTF_PM_S", "FTF_PM_BGROUP", "FTF_CONV_PROFILE"};
    unsigned long long h = 0xcbf29ce484222325ull;
    for (const char* name : kNames) {
        const char* v = getenv(name);
        for (const char* c = v ? v : "\x01"; *c; ++c) { h ^= (unsigned char)*c; h *= 0x100000001b3ull; }
        h ^= 0xff; h *= 0x100000001b3ull;
    }
    return h;
}

int conv_tc_plan_cached(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, int device, const ConvPlan** plan,
                        std::string* err, int mode) {
    // per thread: a runner (and a device-level caller) is driven by one thread, so no lock; bounded, cleared when full
    thread_local std::map<std::vector<unsigned long long>, ConvPlan> cache;
    std::vector<unsigned long long> key = {(unsigned long long)device, (unsigned long long)(uintptr_t)g.in, (unsigned long long)(uintptr_t)g.weights,
                                           (unsigned long long)(uintptr_t)g.bias, (unsigned long long)g.n_max, (unsigned long long)g.H, (unsigned long long)g.W,
                                           (unsigned long long)g.Cin, (unsigned long long)g.in_ld, (unsigned long long)g.Cout, (unsigned long long)g.KH,
                                           (unsigned long long)g.KW, (unsigned long long)g.stride, (unsigned long long)g.pad_h, (unsigned long long)g.pad_w,
                                           (unsigned long long)sm_count, (unsigned long long)mode, env_fingerprint()};
    for (int i = 0; i < n_seg; ++i) {
        key.push_back((unsigned long long)(uintptr_t)segs[i].out); key.push_back((unsigned long long)segs[i].ld); key.push_back((unsigned long long)segs[i].coff);
        key.push_back((unsigned long long)segs[i].col_begin); key.push_back((unsigned long long)segs[i].flags);
    }
    auto it = cache.find(key);
    if (it == cache.end()) {
        ConvPlan fresh;
        const int rc = conv_tc_plan(g, segs, n_seg, sm_count, &fresh, err, mode);
        if (rc) return rc;
        if (cache.size() >= 256) cache.clear();
        it = cache.emplace(std::move(key), fresh).first;
    }
    *plan = &it->second;
    return FTF_OK;
}

int conv_tc_launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err, unsigned long long* tl, int reverse) {
    return plan.kind ? wm::launch(plan.wm, n, s, err, tl, reverse) : pm::launch(plan.pm, n, s, err, tl, reverse);
}

int conv_tc_profile_read(unsigned long long* out16, int reset) {
    unsigned long long a[16], b[16];
    int rc = pm::profile_read(a, reset);
    if (rc) return rc;
    rc = wm::profile_read(b, reset);
    if (rc) return rc;
    for (int i = 0; i < 16; ++i) out16[i] = a[i] + b[i];
    return FTF_OK;
}

}  // namespace ftf
