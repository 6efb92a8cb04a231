// ol_gpu/SamplerBank.hpp -- header-only C++ host mirror of the reference's sample playback for a BANK of players, over the
// C ABI (include/olbank.h):
//   SamplerBank   `instances` x ol::synth::Sample (modules/synthlib/Sample.h:16-52, Sample.cpp:9-61) behind
//                 ol::synth::SampleSoundSource (SampleSoundSource.h:14-41), reading from a device-resident sample pool
//                 instead of one SampleDataSource::Read per frame (SampleDataSource.h:11-25)
// Method names are the reference's; each takes an instance range (default: the whole bank):
//     ol::gpu::SamplerBank kit(65536, 1);
//     uint32_t kick = kit.LoadSample(pcm, frames);          // into the pool (HBM)
//     kit.SetSample(kick, 0, 512);                          // instances 0..511 play it (fresh Sample objects)
//     kit.SetPlayMode(OLB_SAMPLE_LOOP);  kit.SetLoopEnd(24000);
//     kit.GateOn(0, 512);                                   // SampleSoundSource::GateOn: Seek(0) + Play()
//     kit.ProcessBlock(out_host, frames);                   // [frames][instances * channels], played INTO
// As in the reference, a frame of an instance that is not playing (or is past the end of a one-shot) is left untouched.
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

class SamplerBank {
public:
    static constexpr uint32_t kAll = 0xffffffffu;
    SamplerBank(uint32_t instances, uint32_t channels = 1, int device = 0) : instances_(instances), channels_(channels) {
        note(olb_sampler_create(device, instances, channels, &h_));
    }
    ~SamplerBank() { olb_sampler_destroy(h_); }
    SamplerBank(const SamplerBank &) = delete;
    SamplerBank &operator=(const SamplerBank &) = delete;

    InitStatus Init(t_sample sample_rate) { return note(olb_sampler_init(h_, sample_rate)); }  // Sample::Init
    // [frames][channels] floats into the pool; returns the sample id (0xffffffff on failure)
    uint32_t LoadSample(const t_sample *frames_host, uint64_t frames) {
        uint32_t id = kAll;
        note(olb_sampler_load_sample(h_, frames_host, frames, &id));
        return id;
    }
    void SetSample(uint32_t sample_id, uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_set_sample(h_, first, n(first, count), &sample_id, 0)); }
    void SetSamples(const uint32_t *ids, uint32_t first, uint32_t count) { note(olb_sampler_set_sample(h_, first, count, ids, 1)); }
    void Seek(uint64_t frame_index, uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_seek(h_, first, n(first, count), frame_index)); }
    void SetLoopStart(uint64_t frame_index, uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_set_loop_start(h_, first, n(first, count), frame_index)); }
    void SetLoopEnd(uint64_t frame_index, uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_set_loop_end(h_, first, n(first, count), frame_index)); }
    void SetPlayMode(int mode, uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_set_play_mode(h_, first, n(first, count), mode)); }
    void TogglePlay(uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_toggle_play(h_, first, n(first, count))); }
    void Play(uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_play(h_, first, n(first, count))); }
    void Pause(uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_pause(h_, first, n(first, count))); }
    void GateOn(uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_gate(h_, first, n(first, count), 1)); }   // SampleSoundSource.h:21-24
    void GateOff(uint32_t first = 0, uint32_t count = kAll) { note(olb_sampler_gate(h_, first, n(first, count), 0)); }  // SampleSoundSource.h:26
    uint64_t GetChannelCount() const { return channels_; }
    uint64_t Position(uint32_t instance) {
        uint64_t p = 0;
        note(olb_sampler_position(h_, instance, 1, &p, nullptr));
        return p;
    }

    // Sample::Process for every instance: one frame ([instances * channels] floats, host), played into
    void Process(t_sample *frame_out) { note(olb_sampler_process_host(h_, frame_out, 1)); }
    void ProcessBlock(t_sample *out_host, uint64_t frames) { note(olb_sampler_process_host(h_, out_host, frames)); }
    void ProcessDevice(t_sample *out_dev, uint64_t frames, void *stream = nullptr) { note(olb_sampler_process(h_, out_dev, frames, stream)); }

    uint32_t Instances() const { return instances_; }
    int status() const { return status_; }
    std::string error() const { return olb_last_error(); }

private:
    uint32_t n(uint32_t first, uint32_t count) const { return count == kAll ? instances_ - first : count; }
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_sampler *h_ = nullptr;
    uint32_t instances_, channels_;
    int status_ = OLB_OK;
};

}  // namespace gpu
}  // namespace ol
