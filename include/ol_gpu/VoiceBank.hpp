// ol_gpu/VoiceBank.hpp -- header-only C++ host mirror of the synthlib voice path for a BANK of voices, over the C ABI
// (include/olbank.h):
//   VoiceBank       `voices` x ol::synth::SynthVoice with the reference's default parts (modules/synthlib/SynthVoice.h:16-316);
//                   the Voice interface (modules/synthlib/Voice.h:33-57) per voice range
//   PolyvoiceBank   ol::synth::Polyvoice over a VoiceBank (modules/synthlib/Polyvoice.h:12-86): NoteOn takes the first voice
//                   that is not playing, NoteOff releases the first voice playing that note, Process ADDS every voice
//                   into the caller's frame (the caller zeroes it: main.cpp:79)
//   VoiceMapBank    ol::synth::VoiceMap over a VoiceBank (modules/synthlib/VoiceMap.h:14-90): note -> voice and channel -> voice
//                   tables
// Call order as in the reference (modules/ol_daisy/app/synth/main.cpp:74-90, 114-128, 161-168):
//     ol::gpu::VoiceBank bank(65536);  bank.Init(48000);
//     ol::gpu::PolyvoiceBank poly(bank);
//     poly.UpdateMidiControl(CC_FILTER_CUTOFF, 70);  poly.NoteOn(60, 100);
//     float frame = 0;  poly.Process(&frame);                 // per frame, mono
//     poly.ProcessBlock(bus, frames, 2, 1.0f);                // or a block into an interleaved stereo bus
// All DSP runs in libolbank.so on the GPU; the allocation policy is host logic.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../olbank.h"

namespace ol {
namespace gpu {

#ifndef OL_GPU_T_SAMPLE
#define OL_GPU_T_SAMPLE
typedef float t_sample;  // modules/corelib/ol_corelib.h:23
enum InitStatus { Ok = 0, Error = 1 };  // modules/synthlib/SoundSource.h:9-11
#endif

class VoiceBank {
public:
    explicit VoiceBank(uint32_t voices, int device = 0) : voices_(voices), device_(device) {}
    ~VoiceBank() { olb_voice_destroy(h_); }
    VoiceBank(const VoiceBank &) = delete;
    VoiceBank &operator=(const VoiceBank &) = delete;

    InitStatus Init(t_sample sample_rate) {  // SynthVoice::Init (SynthVoice.h:31-39)
        if (!h_) return note(olb_voice_create(device_, voices_, sample_rate, &h_));
        return note(olb_voice_init(h_, sample_rate));
    }
    void Update() {}  // applied at the next Process call
    // SynthVoice::UpdateMidiControl / UpdateHardwareControl (SynthVoice.h:101-229): all voices, or a range
    void UpdateMidiControl(uint8_t control, uint8_t value) { note(olb_voice_update_midi(h_, 0, voices_, control, value)); }
    void UpdateMidiControl(uint8_t control, uint8_t value, uint32_t first, uint32_t count) { note(olb_voice_update_midi(h_, first, count, control, value)); }
    void UpdateHardwareControl(uint8_t control, t_sample value) { note(olb_voice_update_hw(h_, 0, voices_, control, &value, 0)); }
    void UpdateHardwareControl(uint8_t control, const t_sample *values, uint32_t first, uint32_t count) {
        note(olb_voice_update_hw(h_, first, count, control, values, 1));
    }
    // UpdateConfig(Voice::Config&): 16 floats in the member order of modules/synthlib/Voice.h:14-31
    void UpdateConfig(const t_sample *config16) { note(olb_voice_update_config(h_, 0, voices_, config16)); }
    void UpdateConfig(const t_sample *config16, uint32_t first, uint32_t count) { note(olb_voice_update_config(h_, first, count, config16)); }
    // NoteOn / NoteOff / GateOn / GateOff / SetFrequency (SynthVoice.h:231-267) for one voice or a range
    void NoteOn(uint32_t voice, uint8_t midi_note, uint8_t velocity) { note(olb_voice_note_on(h_, voice, 1, &midi_note, 0, velocity)); }
    void NoteOn(const uint8_t *notes, uint8_t velocity, uint32_t first, uint32_t count) { note(olb_voice_note_on(h_, first, count, notes, 1, velocity)); }
    void NoteOff(uint32_t voice) { note(olb_voice_note_off(h_, voice, 1)); }
    void NoteOff(uint32_t first, uint32_t count) { note(olb_voice_note_off(h_, first, count)); }
    void GateOn(uint32_t first, uint32_t count) { note(olb_voice_set_gate(h_, first, count, 1)); }
    void GateOff(uint32_t first, uint32_t count) { note(olb_voice_set_gate(h_, first, count, 0)); }
    void SetFrequency(uint32_t voice, t_sample freq) { note(olb_voice_set_frequency(h_, voice, 1, &freq, 0)); }
    // Playing(): the note each voice plays, 0 = free (SynthVoice.h:263-265)
    uint8_t Playing(uint32_t voice) {
        uint8_t p = 0;
        note(olb_voice_playing(h_, voice, 1, &p));
        return p;
    }
    void Playing(uint8_t *playing, uint32_t first, uint32_t count) { note(olb_voice_playing(h_, first, count, playing)); }

    // SynthVoice::Process(frame_out) for every voice: HOST array of `voices` floats (SynthVoice.h:41-53).
    void Process(t_sample *frame_out) { note(olb_voice_process_host(h_, frame_out, nullptr, 1, 0, 1.0f, 1)); }
    // Block: out_host [frames][voices] and / or an interleaved bus [frames][bus_channels] (bus += gain * sum of voices
    // when accumulate, the Polyvoice convention).
    void ProcessBlock(t_sample *out_host, t_sample *bus_host, uint32_t bus_channels, bool accumulate, t_sample bus_gain, uint64_t frames) {
        note(olb_voice_process_host(h_, out_host, bus_host, bus_channels, accumulate ? 1 : 0, bus_gain, frames));
    }
    void ProcessDevice(t_sample *out_dev, t_sample *bus_dev, uint32_t bus_channels, bool accumulate, t_sample bus_gain, uint64_t frames,
                       void *stream = nullptr) {
        note(olb_voice_process(h_, out_dev, bus_dev, bus_channels, accumulate ? 1 : 0, bus_gain, nullptr, frames, stream));
    }

    // Sample-accurate control: events (sorted by frame) are applied before the frame they name; voice < 0 routes notes
    // through Polyvoice's allocation (olb_voice_process_events).  Device buffers.
    void ProcessEvents(const olb_voice_event *events, uint32_t n_events, t_sample *out_dev, t_sample *bus_dev, uint32_t bus_channels,
                       bool accumulate, t_sample bus_gain, uint64_t frames, void *stream = nullptr) {
        note(olb_voice_process_events(h_, events, n_events, out_dev, bus_dev, bus_channels, accumulate ? 1 : 0, bus_gain, frames, stream));
    }

    uint32_t Voices() const { return voices_; }
    int status() const { return status_; }
    std::string error() const { return olb_last_error(); }
    olb_voice *handle() { return h_; }

private:
    InitStatus note(int rc) {
        status_ = rc;
        return rc == OLB_OK ? Ok : Error;
    }
    olb_voice *h_ = nullptr;
    uint32_t voices_;
    int device_;
    int status_ = OLB_OK;
};

// The cross-GPU voice bus of a SHARDED bank (one process per GPU, rank r owns a contiguous range of the voices): rank 0
// makes the id (VoiceBusComm::UniqueId), the host's launcher hands it to the other ranks, every rank constructs its
// VoiceBusComm (collective), and PolyvoiceBank::SetComm makes ProcessBlock return the mix of ALL ranks' voices
// (Polyvoice.h:28-33 summed over the shards: one in-place NCCL all-reduce of the partial bus over NVLink).
class VoiceBusComm {
public:
    static bool UniqueId(unsigned char (&id)[OLB_COMM_ID_BYTES]) { return olb_comm_unique_id(id) == OLB_OK; }
    VoiceBusComm(int device, int rank, int world, const unsigned char (&id)[OLB_COMM_ID_BYTES]) { status_ = olb_comm_init(device, rank, world, id, &c_); }
    ~VoiceBusComm() { olb_comm_destroy(c_); }
    VoiceBusComm(const VoiceBusComm &) = delete;
    VoiceBusComm &operator=(const VoiceBusComm &) = delete;
    // device buffers: in place on `stream`, or overlapped on the communicator's own stream (see olbank.h)
    int AllReduceBus(t_sample *bus_dev, uint64_t frames, uint32_t channels, void *stream = nullptr) {
        return status_ = olb_voice_bus_allreduce(c_, bus_dev, frames, channels, stream);
    }
    int AllReduceBusAsync(t_sample *bus_dev, uint64_t frames, uint32_t channels, void *producer_stream, int slot) {
        return status_ = olb_voice_bus_allreduce_async(c_, bus_dev, frames, channels, producer_stream, slot);
    }
    int Wait(int slot, void *stream = nullptr) { return status_ = olb_comm_wait(c_, slot, stream); }
    int status() const { return status_; }
    olb_comm *handle() { return c_; }

private:
    olb_comm *c_ = nullptr;
    int status_ = OLB_OK;
};

class PolyvoiceBank {
public:
    explicit PolyvoiceBank(VoiceBank &bank) : bank_(bank), playing_(bank.Voices()) {}
    // this bank holds ONE SHARD of the voices: Process / ProcessBlock mix the voices of every rank of `comm`
    void SetComm(VoiceBusComm *comm) { comm_ = comm; }
    InitStatus Init(t_sample sample_rate) { return bank_.Init(sample_rate); }  // Polyvoice.h:22-26
    // Polyvoice::NoteOn: the first voice that is not playing; silently dropped when all are busy (Polyvoice.h:35-43)
    void NoteOn(uint8_t midi_note, uint8_t velocity) {
        bank_.Playing(playing_.data(), 0, bank_.Voices());
        for (uint32_t v = 0; v < bank_.Voices(); v++)
            if (!playing_[v]) {
                bank_.NoteOn(v, midi_note, velocity);
                return;
            }
    }
    // Polyvoice::NoteOff: the first voice playing that note (Polyvoice.h:45-51)
    void NoteOff(uint8_t midi_note, uint8_t /*velocity*/) {
        bank_.Playing(playing_.data(), 0, bank_.Voices());
        for (uint32_t v = 0; v < bank_.Voices(); v++)
            if (playing_[v] == midi_note) {
                bank_.NoteOff(v);
                return;
            }
    }
    void UpdateMidiControl(uint8_t control, uint8_t value) { bank_.UpdateMidiControl(control, value); }          // Polyvoice.h:53-57
    void UpdateHardwareControl(uint8_t control, t_sample value) { bank_.UpdateHardwareControl(control, value); }  // Polyvoice.h:59-63
    void UpdateConfig(const t_sample *config16) { bank_.UpdateConfig(config16); }
    // Polyvoice::Process: *frame_out += every voice (Polyvoice.h:28-33); the caller zeroes the frame (main.cpp:79)
    void Process(t_sample *frame_out) { ProcessBlock(frame_out, 1, 1, 1.0f); }
    void ProcessBlock(t_sample *bus_host, uint64_t frames, uint32_t bus_channels = 1, t_sample gain = 1.0f) {
        if (comm_) status_ = olb_voice_process_host_comm(bank_.handle(), comm_->handle(), bus_host, bus_channels, 1, gain, frames);
        else bank_.ProcessBlock(nullptr, bus_host, bus_channels, true, gain, frames);
    }
    int status() const { return status_ != OLB_OK ? status_ : bank_.status(); }

private:
    VoiceBank &bank_;
    VoiceBusComm *comm_ = nullptr;
    int status_ = OLB_OK;
    std::vector<uint8_t> playing_;
};

// ol::synth::VoiceMap (modules/synthlib/VoiceMap.h:14-90) over a VoiceBank: a note -> voice table (a pad kit: every mapped
// note triggers ITS voice) and a channel -> voice table for MIDI CCs (the voice last set on that channel).  Unmapped notes
// and channels are ignored as in the reference; Process adds the bank into the caller's frame (voices that were never
// triggered are silent).  One deviation: the reference walks its note table in Process, so a voice set under two notes
// would be advanced twice per frame there -- SetVoice refuses that (returns false).
class VoiceMapBank {
public:
    explicit VoiceMapBank(VoiceBank &bank) : bank_(bank) {
        for (int &v : note2voice_) v = -1;
        for (int &v : channel2voice_) v = -1;
    }
    bool SetVoice(uint8_t channel, uint8_t note, uint32_t voice) {  // VoiceMap.h:45-59
        if (note >= 128 || channel >= 16) return true;               // silently ignored, as upstream
        if (voice >= bank_.Voices()) return false;
        for (int n = 0; n < 128; n++)
            if (n != note && note2voice_[n] == (int) voice) return false;
        note2voice_[note] = (int) voice;
        channel2voice_[channel] = (int) voice;
        return true;
    }
    InitStatus Init(t_sample sample_rate) { return bank_.Init(sample_rate); }  // VoiceMap.h:61-63
    void NoteOn(uint8_t note, uint8_t velocity) {                              // VoiceMap.h:27-34
        if (note < 128 && note2voice_[note] >= 0) bank_.NoteOn((uint32_t) note2voice_[note], note, velocity);
    }
    void NoteOff(uint8_t note, uint8_t /*velocity*/) {  // VoiceMap.h:36-43
        if (note < 128 && note2voice_[note] >= 0) bank_.NoteOff((uint32_t) note2voice_[note]);
    }
    void UpdateMidiControl(uint8_t channel, uint8_t control, uint8_t value) {  // VoiceMap.h:80-87
        if (channel < 16 && channel2voice_[channel] >= 0) bank_.UpdateMidiControl(control, value, (uint32_t) channel2voice_[channel], 1);
    }
    // VoiceMap::Process: frame_out[0] += every mapped voice (VoiceMap.h:65-78)
    void Process(t_sample *frame_out) { bank_.ProcessBlock(nullptr, frame_out, 1, true, 1.0f, 1); }
    void ProcessBlock(t_sample *bus_host, uint64_t frames, uint32_t bus_channels = 1, t_sample gain = 1.0f) {
        bank_.ProcessBlock(nullptr, bus_host, bus_channels, true, gain, frames);
    }
    int status() const { return bank_.status(); }

private:
    VoiceBank &bank_;
    int note2voice_[128], channel2voice_[16];
};

}  // namespace gpu
}  // namespace ol
