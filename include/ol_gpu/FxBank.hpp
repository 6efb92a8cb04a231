// ol_gpu/FxBank.hpp -- header-only C++ host mirror of the fxlib units (modules/fxlib/Fx.h) as BANKS of mono instances,
// over the C ABI (include/olbank.h).  Each class keeps the reference class's method names and call order
//     default-construct -> Init(sample_rate) -> UpdateMidiControl / UpdateHardwareControl -> Process(in, out) per frame
// (Fx.h:83-114 FilterFx, :183-217 DelayFx, :287-300 ReverbFx, :416-434 FxRack; caller: test/fx_test.cpp:25-55,
// modules/ol_daisy/app/synth/main.cpp:74-90), so a caller swaps ONE embedded-style object for thousands:
//
//     ol::gpu::FxRackBank rack(8192);                       // 8 192 x ol::fx::FxRack<1> on cuda:0
//     rack.Init(48000);
//     rack.UpdateMidiControl(CC_DELAY_TIME, 64);            // all instances, like the reference's setter
//     rack.UpdateHardwareControl(CC_DELAY_TIME, times, 0, 8192);   // or one value per instance
//     for (;;) rack.Process(frame_in, frame_out);           // one float per instance (N = 1 frames side by side)
//     rack.ProcessBlock(in_host, out_host, frames);         // or a block at a time: [frames][instances]
//
// CC numbers are modules/corelib/cc_map.h; the CC -> parameter maths is the reference's ol::core::scale, run on the
// host inside libolbank.so.  Setters take effect at the next Process call (Update() semantics, Fx.h:110-114).
// All DSP runs in libolbank.so on the GPU; there is no CPU path.
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

template <int KIND>
class FxBank {
public:
    explicit FxBank(uint32_t instances, int device = 0) : instances_(instances), device_(device) {}
    ~FxBank() { olb_fx_destroy(h_); }
    FxBank(const FxBank &) = delete;
    FxBank &operator=(const FxBank &) = delete;

    // Init(sample_rate): constructs the bank on first use; every call clears filter state and delay lines and keeps
    // the parameter members (re-Init: test/fx_test.cpp:32,42).
    InitStatus Init(t_sample sample_rate) {
        if (!h_) return note(olb_fx_create(device_, KIND, instances_, sample_rate, &h_));
        return note(olb_fx_init(h_, sample_rate));
    }
    // Update(): the reference pushes cached members to its primitives here (Fx.h:110-114); the bank does that at the
    // next Process call, so this is a no-op kept for source compatibility.
    void Update() {}
    // Not in the reference: OLB_MODE_EXACT (default, bit-exact against it) or OLB_MODE_FMA (fused, olbank.h); after Init.
    void SetMode(int mode) { note(olb_fx_set_mode(h_, mode)); }
    // UpdateMidiControl(control, value 0..127) -- every instance, or a range
    void UpdateMidiControl(uint8_t control, uint8_t value) { note(olb_fx_update_midi(h_, 0, instances_, control, value)); }
    void UpdateMidiControl(uint8_t control, uint8_t value, uint32_t first, uint32_t count) {
        note(olb_fx_update_midi(h_, first, count, control, value));
    }
    // UpdateHardwareControl(control, value 0..1) -- every instance, or one value per instance of a range
    void UpdateHardwareControl(uint8_t control, t_sample value) { note(olb_fx_update_hw(h_, 0, instances_, control, &value, 0)); }
    void UpdateHardwareControl(uint8_t control, const t_sample *values, uint32_t first, uint32_t count) {
        note(olb_fx_update_hw(h_, first, count, control, values, 1));
    }

    // One frame of every instance: HOST arrays of `instances` floats, in place allowed (main.cpp:82-85).  The reference's
    // per-frame call: correct and slow (one PCIe round trip per frame) -- audio hosts should call ProcessBlock.
    void Process(const t_sample *frame_in, t_sample *frame_out) { note(olb_fx_process_host(h_, frame_in, frame_out, 1)); }
    // Block of frames, HOST buffers [frames][instances].
    void ProcessBlock(const t_sample *in_host, t_sample *out_host, uint64_t frames) { note(olb_fx_process_host(h_, in_host, out_host, frames)); }
    // Block of frames, DEVICE buffers on this bank's GPU, asynchronous on `stream` (a cudaStream_t).
    void ProcessDevice(const t_sample *in_dev, t_sample *out_dev, uint64_t frames, void *stream = nullptr) {
        note(olb_fx_process(h_, in_dev, out_dev, frames, stream));
    }

    uint32_t Instances() const { return instances_; }
    int status() const { return status_; }  // last OLB_* code
    std::string error() const { return olb_last_error(); }
    const char *LastKernel() { return olb_fx_last_kernel(h_); }
    olb_fx *handle() { return h_; }

private:
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_fx *h_ = nullptr;
    uint32_t instances_;
    int device_;
    int status_ = OLB_OK;
};

typedef FxBank<OLB_FX_FILTER> FilterFxBank;  // bank of ol::fx::FilterFx<1>   (Fx.h:65-165)
typedef FxBank<OLB_FX_DELAY> DelayFxBank;    // bank of ol::fx::DelayFx<1>    (Fx.h:170-268)
typedef FxBank<OLB_FX_REVERB> ReverbFxBank;  // bank of ol::fx::ReverbFx<1>   (Fx.h:272-393)
typedef FxBank<OLB_FX_RACK> FxRackBank;      // bank of ol::fx::FxRack<1>     (Fx.h:397-493)
// FxRack<1> with the Dattorro plate in the reverb slot (the wiring sketched in modules/fxlib/ReverbFx.cpp:11-37)
typedef FxBank<OLB_FX_RACK_PLATE> FxRackPlateBank;

// ---- stereo instances: banks of the reference's <2> templates (olb_fx2_*, include/olbank.h) ----------------------------
// Frames are the reference's t_sample[2] per instance, side by side: [instances][2] per bank frame.  Second-channel
// behaviour is the reference's own (FilterFx<2> leaves frame_out[1] alone, DelayFx<2> filters channel 0 only, FxRack<2>'s
// channel 1 is 0 * master_volume): modules/fxlib/Fx.h:88-108, 194-208, 294-300, 426-434; caller main.cpp:83-85.
template <int KIND>
class FxBank2 {
public:
    explicit FxBank2(uint32_t instances, int device = 0) : instances_(instances), device_(device) {}
    ~FxBank2() { olb_fx2_destroy(h_); }
    FxBank2(const FxBank2 &) = delete;
    FxBank2 &operator=(const FxBank2 &) = delete;
    InitStatus Init(t_sample sample_rate) {
        if (!h_) return note(olb_fx2_create(device_, KIND, instances_, sample_rate, &h_));
        return note(olb_fx2_init(h_, sample_rate));
    }
    void Update() {}
    void SetMode(int mode) { note(olb_fx2_set_mode(h_, mode)); }
    void UpdateMidiControl(uint8_t control, uint8_t value) { note(olb_fx2_update_midi(h_, 0, instances_, control, value)); }
    void UpdateMidiControl(uint8_t control, uint8_t value, uint32_t first, uint32_t count) {
        note(olb_fx2_update_midi(h_, first, count, control, value));
    }
    void UpdateHardwareControl(uint8_t control, t_sample value) { note(olb_fx2_update_hw(h_, 0, instances_, control, &value, 0)); }
    void UpdateHardwareControl(uint8_t control, const t_sample *values, uint32_t first, uint32_t count) {
        note(olb_fx2_update_hw(h_, first, count, control, values, 1));
    }
    // one stereo frame of every instance: HOST arrays of 2 * instances floats (in place allowed)
    void Process(const t_sample *frame_in, t_sample *frame_out) { note(olb_fx2_process_host(h_, frame_in, frame_out, 1)); }
    void ProcessBlock(const t_sample *in_host, t_sample *out_host, uint64_t frames) { note(olb_fx2_process_host(h_, in_host, out_host, frames)); }
    void ProcessDevice(const t_sample *in_dev, t_sample *out_dev, uint64_t frames, void *stream = nullptr) {
        note(olb_fx2_process(h_, in_dev, out_dev, frames, stream));
    }
    uint32_t Instances() const { return instances_; }
    int status() const { return status_; }
    std::string error() const { return olb_last_error(); }
    const char *LastKernel() { return olb_fx2_last_kernel(h_); }
    olb_fx2 *handle() { return h_; }

private:
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_fx2 *h_ = nullptr;
    uint32_t instances_;
    int device_;
    int status_ = OLB_OK;
};
typedef FxBank2<OLB_FX_FILTER> FilterFx2Bank;  // bank of ol::fx::FilterFx<2>
typedef FxBank2<OLB_FX_DELAY> DelayFx2Bank;    // bank of ol::fx::DelayFx<2>
typedef FxBank2<OLB_FX_REVERB> ReverbFx2Bank;  // bank of ol::fx::ReverbFx<2>
typedef FxBank2<OLB_FX_RACK> FxRack2Bank;      // bank of ol::fx::FxRack<2>

}  // namespace gpu
}  // namespace ol
