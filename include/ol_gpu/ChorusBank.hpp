// ol_gpu/ChorusBank.hpp -- header-only C++ host handle of a BANK of mono choruses over the C ABI (include/olbank.h).
// The reference has no chorus class: its chorus is the Max/RNBO graph modules/rnbo/patcher/mono-chorus.rnbopat, whose
// `param` boxes (rate, depth, cutoff, q, mix, phase; all 0..1) are the setters here.  Lifecycle and call shape follow the
// fxlib units (Init(sample_rate) -> setters -> Process(in, out) per frame; modules/fxlib/Fx.h:83-114).
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

class ChorusBank {
public:
    explicit ChorusBank(uint32_t instances, int device = 0) : instances_(instances), device_(device) {}
    ~ChorusBank() { olb_chorus_destroy(h_); }
    ChorusBank(const ChorusBank &) = delete;
    ChorusBank &operator=(const ChorusBank &) = delete;

    InitStatus Init(t_sample sample_rate) {
        if (!h_) return note(olb_chorus_create(device_, instances_, sample_rate, &h_));
        return note(olb_chorus_init(h_, sample_rate));
    }
    void SetRate(t_sample v) { set(OLB_CHORUS_RATE, v); }
    void SetDepth(t_sample v) { set(OLB_CHORUS_DEPTH, v); }
    void SetCutoff(t_sample v) { set(OLB_CHORUS_CUTOFF, v); }
    void SetQ(t_sample v) { set(OLB_CHORUS_Q, v); }
    void SetMix(t_sample v) { set(OLB_CHORUS_MIX, v); }
    void SetPhase(t_sample v) { set(OLB_CHORUS_PHASE, v); }
    void SetPitch(t_sample hz) { set(OLB_CHORUS_PITCH, hz); }  // param pitch 0..3: the gen~ pitch shifter's phasor frequency
    void Set(int param, const t_sample *values, uint32_t first, uint32_t count) { note(olb_chorus_set_params(h_, first, count, param, values, 1)); }

    void Process(const t_sample *frame_in, t_sample *frame_out) { note(olb_chorus_process_host(h_, frame_in, frame_out, 1)); }
    void ProcessBlock(const t_sample *in_host, t_sample *out_host, uint64_t frames) { note(olb_chorus_process_host(h_, in_host, out_host, frames)); }
    void ProcessDevice(const t_sample *in_dev, t_sample *out_dev, uint64_t frames, void *stream = nullptr) {
        note(olb_chorus_process(h_, in_dev, out_dev, frames, stream));
    }

    uint32_t Instances() const { return instances_; }
    int status() const { return status_; }
    std::string error() const { return olb_last_error(); }
    olb_chorus *handle() { return h_; }

private:
    void set(int param, t_sample v) { note(olb_chorus_set_param(h_, 0, instances_, param, v)); }
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_chorus *h_ = nullptr;
    uint32_t instances_;
    int device_;
    int status_ = OLB_OK;
};

// modules/rnbo/patcher/stereo-chorus.rnbopat: per stereo instance two mono choruses, the left one with cycle~ phase 1 and the
// right one with phase 0 (obj-8 `setparam phase 1` -> obj-1 -> out~ 1, obj-18 `setparam phase 0` -> obj-2 -> out~ 2), every
// other param shared through `setparam`.  Frames are [instances][2]: the same memory as a mono bank of 2 x instances units,
// unit 2i the left and 2i + 1 the right channel of instance i -- which is what this class sets up (no kernel of its own).
class StereoChorusBank {
public:
    explicit StereoChorusBank(uint32_t instances, int device = 0) : bank_(2 * instances, device), instances_(instances) {}
    InitStatus Init(t_sample sample_rate) {
        const InitStatus st = bank_.Init(sample_rate);
        if (st == Ok && !phased_) {
            const t_sample one = 1.f, zero = 0.f;
            for (uint32_t i = 0; i < instances_; i++) {
                bank_.Set(OLB_CHORUS_PHASE, &one, 2 * i, 1);
                bank_.Set(OLB_CHORUS_PHASE, &zero, 2 * i + 1, 1);
            }
            phased_ = true;
        }
        return st;
    }
    void SetRate(t_sample v) { bank_.SetRate(v); }
    void SetDepth(t_sample v) { bank_.SetDepth(v); }
    void SetCutoff(t_sample v) { bank_.SetCutoff(v); }
    void SetQ(t_sample v) { bank_.SetQ(v); }
    void SetMix(t_sample v) { bank_.SetMix(v); }
    void SetPitch(t_sample hz) { bank_.SetPitch(hz); }
    // one value per stereo instance: both of its channels get it (the patch's params are shared)
    void Set(int param, const t_sample *values, uint32_t first, uint32_t count) {
        for (uint32_t i = 0; i < count; i++) {
            const t_sample v[2] = {values[i], values[i]};
            bank_.Set(param, v, 2 * (first + i), 2);
        }
    }
    void Process(const t_sample *frame_in, t_sample *frame_out) { bank_.Process(frame_in, frame_out); }  // [instances][2]
    void ProcessBlock(const t_sample *in_host, t_sample *out_host, uint64_t frames) { bank_.ProcessBlock(in_host, out_host, frames); }
    void ProcessDevice(const t_sample *in_dev, t_sample *out_dev, uint64_t frames, void *stream = nullptr) {
        bank_.ProcessDevice(in_dev, out_dev, frames, stream);
    }
    uint32_t Instances() const { return instances_; }
    int status() const { return bank_.status(); }
    ChorusBank &mono() { return bank_; }

private:
    ChorusBank bank_;
    uint32_t instances_;
    bool phased_ = false;
};

}  // namespace gpu
}  // namespace ol
