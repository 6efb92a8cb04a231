// ol_gpu/BiquadBank.hpp -- header-only C++ host mirror of the biquad / one-pole cascade bank over the C ABI
// (include/olbank.h).  It keeps the reference's object conventions so a caller written against ol_dsp's units
// (default-construct -> Init(sample_rate) -> setters -> Process per frame; e.g. modules/fxlib/Fx.h:83-114,
// modules/ol_daisy/app/synth/main.cpp:74-90) swaps one embedded-style object for a bank of thousands:
//
//     ol::gpu::BiquadBank bank(65536, 8);            // 65 536 channels x 8 sections on cuda:0
//     bank.Init(48000);
//     bank.SetCoeffs(0, 65536, coefs);               // [channels][sections][5] = b0 b1 b2 a1 a2
//     for (;;) bank.Process(frame_in, frame_out);    // one float per channel, like Process(in, out) per frame
//     bank.ProcessBlock(in_host, out_host, frames);  // or a block at a time
//
// t_sample is float as in modules/corelib/ol_corelib.h:23.  The reference never throws and returns void; this
// wrapper keeps void setters and exposes status()/error() instead (InitStatus style: modules/synthlib/SoundSource.h:9-11).
// All compute happens in libolbank.so on the GPU; there is no CPU path.
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

class BiquadBank {
public:
    BiquadBank(uint32_t channels, uint32_t sections, int device = 0) : channels_(channels), sections_(sections), device_(device) {}
    ~BiquadBank() { olb_biquad_destroy(h_); }
    BiquadBank(const BiquadBank &) = delete;
    BiquadBank &operator=(const BiquadBank &) = delete;

    // Init(sample_rate): creates the device tables on first use, zeroes the recurrence state on every call
    // (re-Init resets state and keeps parameters: test/fx_test.cpp:32,42; Fx.h:83-86).
    InitStatus Init(t_sample sample_rate) {
        if (!h_) return note(olb_biquad_create(device_, channels_, sections_, sample_rate, &h_));
        return note(olb_biquad_init(h_, sample_rate));
    }

    // Setters take effect at the next Process call, as Update() does in the reference (Fx.h:110-114).
    void SetCoeffs(uint32_t first, uint32_t count, const t_sample *coefs5) { note(olb_biquad_set_coeffs(h_, first, count, coefs5)); }
    void SetKinds(const uint8_t *kinds) { note(olb_biquad_set_kinds(h_, kinds, sections_)); }
    void SetMode(int mode) { note(olb_biquad_set_mode(h_, mode)); }
    // RBJ designer for one section of a range of channels (type: 0 LP, 1 HP, 2 peaking, 3 one-pole LP).
    static bool Design(int type, double sample_rate, double f0, double q, double gain_db, t_sample *coefs5) {
        return olb_biquad_design(type, sample_rate, f0, q, gain_db, coefs5) == OLB_OK;
    }

    // One frame for every channel: frame_in / frame_out are HOST arrays of `channels` floats, in place allowed
    // (main.cpp:82-85 calls Process in place).  This is the reference's per-frame call; it is correct and slow
    // (one PCIe round trip per frame) -- audio hosts should call ProcessBlock.
    void Process(const t_sample *frame_in, t_sample *frame_out) { note(olb_biquad_process_host(h_, frame_in, frame_out, 1, OLB_LAYOUT_FRAMES)); }

    // Block of frames, HOST buffers [frames][channels] (or planar [channels][frames]).
    void ProcessBlock(const t_sample *in_host, t_sample *out_host, uint64_t frames, int layout = OLB_LAYOUT_FRAMES) {
        note(olb_biquad_process_host(h_, in_host, out_host, frames, layout));
    }
    // Block of frames, DEVICE buffers on this bank's GPU, asynchronous on `stream` (a cudaStream_t).
    void ProcessDevice(const t_sample *in_dev, t_sample *out_dev, uint64_t frames, int layout = OLB_LAYOUT_FRAMES, void *stream = nullptr) {
        note(olb_biquad_process(h_, in_dev, out_dev, frames, layout, stream));
    }

    void GetState(t_sample *state4) { note(olb_biquad_get_state(h_, state4)); }
    void SetState(const t_sample *state4) { note(olb_biquad_set_state(h_, state4)); }

    uint32_t Channels() const { return channels_; }
    uint32_t Sections() const { return sections_; }
    int status() const { return status_; }              // last OLB_* code
    std::string error() const { return olb_last_error(); }
    const char *LastKernel() { return olb_biquad_last_kernel(h_); }
    olb_biquad *handle() { return h_; }

private:
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_biquad *h_ = nullptr;
    uint32_t channels_, sections_;
    int device_;
    int status_ = OLB_OK;
};

}  // namespace gpu
}  // namespace ol
