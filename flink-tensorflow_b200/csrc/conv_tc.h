// conv_tc.h -- the tensor-core convolution (K3): plan once per layer, launch per batch.
//
// Two kernels share the work (conv_pm.cu, conv_wm.cu):
//  * pixel-major  (pm): 128 pixels on the UMMA M side, output channels on the N side.  Cheap epilogue (a thread owns a
//    pixel and stores 16-byte channel runs), but every MMA costs ~128 cycles whatever N is, so the tensor pipe runs at
//    Cout_tile/256 of its rate.
//  * weight-major (wm): 128 output channels on the M side, up to 256 pixels on the N side: full MMA rate for Cout >= 128
//    and twice the pm rate for Cout <= 128; the epilogue (a thread owns a channel, 2-byte stores) costs ~11 issue slots
//    per pixel, so it only pays when K is large.
// Each has an im2col mode and a halo-window mode.  The plan keeps the fastest candidate: timed on the layer's own buffers
// when `tune` is set (network build), else by a cycle model.
#pragma once
#include "conv_pm.h"
#include "conv_wm.h"

namespace ftf {

struct ConvPlan {
    int kind;                   // 0 = pixel-major, 1 = weight-major
    double est_cycles;          // modelled cycles of the chosen kernel
    pm::ConvPlan pm;
    wm::ConvPlan wm;
};

// How conv_tc_plan chooses between the candidate kernels.
enum {
    kPlanModel = 0,       // the cycle model decides (FTF_CONV_TUNE=1 in the environment upgrades this to kPlanTune)
    kPlanTune = 1,        // time the candidates on the layer's own buffers for n_max images and keep the fastest; the
                          // choice is remembered per process, so later runners take it without timing (FTF_CONV_TUNE=0 downgrades)
    kPlanModelOnly = 2,   // the cycle model decides, whatever the environment says: planning then only encodes tensor
                          // maps on the host -- no stream, no event, no launch, no synchronisation (per-call entry points)
};
// Encodes the tensor maps and tile shapes for one convolution (called once, at network build).  Every candidate
// accumulates K in the same order (channel chunk, filter tap, 16-element step).
int conv_tc_plan(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlan* plan,
                 std::string* err, int mode = kPlanModel);
// Every candidate kernel of a layer and which one the planning chose; the network keeps the set so that it can re-time the
// alternatives inside the whole forward pass (InceptionV3's in-situ refinement), where sibling branches share the SMs and the
// L2 -> SM bandwidth, which the isolated timing of conv_tc_plan cannot see.
constexpr int kConvVariants = 9;     // pm im2col, pm halo, pm pair, wm im2col, wm halo, pm halo with one wide accumulator set, pm halo with S = 1 / 2 / 4 pinned
struct ConvPlanSet {
    ConvPlan cands[kConvVariants];
    int variant[kConvVariants];   // which of the kernel variants cands[i] is (the value remembered per process)
    int n = 0, chosen = 0;
    bool from_cache = false;  // the choice was taken over from an earlier runner of this process
    bool cacheable = false;   // the choice is one that gets remembered (timed, no FTF_CONV_KIND override, several candidates)
};
int conv_tc_plan_set(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, ConvPlanSet* set,
                     std::string* err, int mode = kPlanModel);
// Remembers `variant` as this process's choice for the layer (what later runners take over).
void conv_tc_choice_put(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, int variant);
// The same for the per-call entry points (ftf_conv2d_bf16_dev, ftf_pool_branch_bf16*_dev): a call that repeats an earlier one
// of this thread -- same device, buffers, geometry and mode -- gets that call's plan back (no tensor-map encoding).
// *plan stays valid until the thread's 256-entry table is cleared by a later call.
int conv_tc_plan_cached(const ConvGeometry& g, const ConvSegment* segs, int n_seg, int sm_count, int device, const ConvPlan** plan,
                        std::string* err, int mode = kPlanModel);
// Enqueues the convolution for the first n images.
int conv_tc_launch(const ConvPlan& plan, int n, cudaStream_t s, std::string* err, unsigned long long* tl = nullptr, int reverse = 0);
// Diagnostic cycle counters of both kernels, summed (see conv_pm.h).
int conv_tc_profile_read(unsigned long long* out16, int reset);

}  // namespace ftf
