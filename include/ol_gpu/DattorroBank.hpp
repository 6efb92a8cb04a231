// ol_gpu/DattorroBank.hpp -- header-only C++ host mirror of libs/dattorro-verb/verb.h:5-26 for a BANK of plates, over the
// C ABI (include/olbank.h).  The reference is a C API on one object:
//     DattorroVerb *v = DattorroVerb_create();  DattorroVerb_setDecay(v, .8);
//     DattorroVerb_process(v, in);  l = DattorroVerb_getLeft(v);  r = DattorroVerb_getRight(v);
// the bank keeps those names as methods over `instances` plates (mono in, stereo wet out, one float per instance):
//     ol::gpu::DattorroBank plates(8192);  plates.SetDecay(.8f);
//     plates.Process(in, left, right);                 // one frame of every plate
//     plates.ProcessBlock(in, left, right, frames);    // [frames][instances]
// Bit-exact against verb.cpp compiled with -ffp-contract=off.  All DSP runs in libolbank.so on the GPU.
#pragma once
#include <cstdint>
#include <string>

#include "../olbank.h"

namespace ol {
namespace gpu {

#ifndef OL_GPU_T_SAMPLE
#define OL_GPU_T_SAMPLE
typedef float t_sample;  // modules/corelib/ol_corelib.h:23
enum InitStatus { Ok = 0, Error = 1 };  // modules/synthlib/SoundSource.h:9-11
#endif

class DattorroBank {
public:
    // DattorroVerb_create (verb.cpp:225-232): rings cleared, parameters at the defaults of initialize() (verb.cpp:215-221)
    explicit DattorroBank(uint32_t instances, int device = 0) : instances_(instances) { note(olb_dattorro_create(device, instances, &h_)); }
    ~DattorroBank() { olb_dattorro_destroy(h_); }  // DattorroVerb_delete (verb.cpp:235-251)
    DattorroBank(const DattorroBank &) = delete;
    DattorroBank &operator=(const DattorroBank &) = delete;

    InitStatus Init() { return note(olb_dattorro_init(h_)); }  // create() again: state cleared, defaults restored

    // DattorroVerb_set* (verb.cpp:136-170): one value for every plate, or one value per plate of a range
    void SetPreDelay(t_sample v) { set(OLB_PLATE_PREDELAY, v); }
    void SetPreFilter(t_sample v) { set(OLB_PLATE_PREFILTER, v); }
    void SetInputDiffusion1(t_sample v) { set(OLB_PLATE_INPUT_DIFFUSION1, v); }
    void SetInputDiffusion2(t_sample v) { set(OLB_PLATE_INPUT_DIFFUSION2, v); }
    void SetDecay(t_sample v) { set(OLB_PLATE_DECAY, v); }
    void SetDecayDiffusion(t_sample v) { set(OLB_PLATE_DECAY_DIFFUSION, v); }
    void SetDamping(t_sample v) { set(OLB_PLATE_DAMPING, v); }
    void Set(int param, const t_sample *values, uint32_t first, uint32_t count) { note(olb_dattorro_set_params(h_, first, count, param, values, 1)); }

    // DattorroVerb_process + _getLeft + _getRight (verb.cpp:258-326) for one frame of every plate: HOST arrays of
    // `instances` floats.  Correct and slow (one PCIe round trip per frame) -- audio hosts should call ProcessBlock.
    void Process(const t_sample *in, t_sample *left, t_sample *right) { note(olb_dattorro_process_host(h_, in, left, right, 1)); }
    void ProcessBlock(const t_sample *in, t_sample *left, t_sample *right, uint64_t frames) {
        note(olb_dattorro_process_host(h_, in, left, right, frames));
    }
    void ProcessDevice(const t_sample *in_dev, t_sample *left_dev, t_sample *right_dev, uint64_t frames, void *stream = nullptr) {
        note(olb_dattorro_process(h_, in_dev, left_dev, right_dev, frames, stream));
    }

    uint32_t Instances() const { return instances_; }
    int status() const { return status_; }
    std::string error() const { return olb_last_error(); }
    olb_dattorro *handle() { return h_; }

private:
    void set(int param, t_sample v) { note(olb_dattorro_set_param(h_, 0, instances_, param, v)); }
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_dattorro *h_ = nullptr;
    uint32_t instances_;
    int status_ = OLB_OK;
};

}  // namespace gpu
}  // namespace ol
