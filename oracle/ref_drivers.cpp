// TEST INFRASTRUCTURE ONLY (oracle/).  Builds oracle/_ref/libolref.so.
//
// Thin extern "C" drivers around the REFERENCE'S OWN headers, included from where they lie under
// $OL_DSP_REF (default /root/reference) -- nothing of the reference is copied into this repo:
//   modules/fxlib/Fx.h            FilterFx / DelayFx / ReverbFx / FxRack      (Fx.h:65-493)
//   modules/fxlib/Reverb.h        stub daisysp::ReverbSc + DaisyVerb          (Reverb.h:13-125)
//   modules/synthlib/*.h          SynthVoice / Polyvoice and the Daisy adapters
//   modules/corelib/ol_corelib.h  ol::core::scale                             (ol_corelib.h:27-44)
//   libs/dattorro-verb/verb.cpp   compiled UNMODIFIED as its own translation unit by oracle/Makefile
// DaisySP itself is absent (see daisysp_shim/daisysp.h): Svf/Oscillator/Adsr/DelayLine arithmetic
// comes from our restatement, so results through those are "reference headers + DaisySP restatement".
// The Dattorro drivers cross no shim: they are the true reference.
//
// Flags (oracle/Makefile): g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math
//                          -include cstdint -include cstdio -include vector
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <vector>

// SynthVoice.h:13 defines PRINTF as printf and prints from every MIDI setter; silence it for this TU.
#define printf(...) ((void) 0)
#include "daisysp.h"
#include "fxlib/Fx.h"
#include "synthlib/ol_synthlib.h"
#include "ctllib/Control.h"
#include "iolib/Serializer.h"
#undef printf

#include "oracle_abi.h"

using ol::fx::DaisyVerb;
using ol::fx::DelayFx;
using ol::fx::FilterFx;
using ol::fx::FxRack;
using ol::fx::ReverbFx;

typedef daisysp::DelayLine<t_sample, MAX_DELAY> RefDelayLine;

namespace {

// Apply every event scheduled for `frame`; returns the index of the first event not yet due.
template <class Unit>
size_t apply_fx_events(Unit &u, const olb_event *ev, size_t n_ev, size_t next, uint32_t frame) {
    while (next < n_ev && ev[next].frame <= frame) {
        const olb_event &e = ev[next++];
        switch (e.kind) {
            case OLB_EV_MIDI_CC: u.UpdateMidiControl((uint8_t) e.a, (uint8_t) e.v); break;
            case OLB_EV_HW_CC: u.UpdateHardwareControl((uint8_t) e.a, e.v); break;
            case OLB_EV_INIT: u.Init(e.v); break;
            default: break;
        }
    }
    return next;
}

template <class V>
size_t apply_voice_events(V &v, const olb_event *ev, size_t n_ev, size_t next, uint32_t frame) {
    while (next < n_ev && ev[next].frame <= frame) {
        const olb_event &e = ev[next++];
        switch (e.kind) {
            case OLB_EV_MIDI_CC: v.UpdateMidiControl((uint8_t) e.a, (uint8_t) e.v); break;
            case OLB_EV_HW_CC: v.UpdateHardwareControl((uint8_t) e.a, e.v); break;
            case OLB_EV_INIT: v.Init(e.v); break;
            case OLB_EV_NOTE_ON: v.NoteOn((uint8_t) e.a, (uint8_t) e.v); break;
            case OLB_EV_NOTE_OFF: v.NoteOff((uint8_t) e.a, (uint8_t) e.v); break;
            case OLB_EV_SET_FREQ: v.SetFrequency(e.v); break;
            case OLB_EV_GATE_ON: v.GateOn(); break;
            case OLB_EV_GATE_OFF: v.GateOff(); break;
            default: break;
        }
    }
    return next;
}

template <int N>
void run_filterfx(float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    FilterFx<N> fx;
    fx.Init(sr);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        next = apply_fx_events(fx, ev, n_ev, next, (uint32_t) t);
        // FilterFx writes channel 0 only (Fx.h:88-108); the caller pre-fills `out` to see that.
        fx.Process(in + t * N, out + t * N);
    }
}

template <int N>
void run_delayfx(float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    std::vector<RefDelayLine *> lines;
    for (int i = 0; i < N; i++) lines.push_back(new RefDelayLine());
    {
        DelayFx<N> fx(lines);
        fx.Init(sr);
        size_t next = 0;
        for (size_t t = 0; t < frames; t++) {
            next = apply_fx_events(fx, ev, n_ev, next, (uint32_t) t);
            fx.Process(in + t * N, out + t * N);
        }
    }
    for (auto l : lines) delete l;
}

template <int N>
void run_reverbfx(float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    daisysp::ReverbSc sc;
    DaisyVerb<N> dv(sc);
    ReverbFx<N> fx(dv);
    fx.Init(sr);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        next = apply_fx_events(fx, ev, n_ev, next, (uint32_t) t);
        fx.Process(in + t * N, out + t * N);
    }
}

template <int N>
void run_fxrack(float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    std::vector<RefDelayLine *> lines;
    for (int i = 0; i < N; i++) lines.push_back(new RefDelayLine());
    {
        DelayFx<N> delay(lines);
        daisysp::ReverbSc sc;
        DaisyVerb<N> dv(sc);
        ReverbFx<N> reverb(dv);
        FilterFx<N> filter;
        FxRack<N> rack(delay, reverb, filter);
        rack.Init(sr);
        size_t next = 0;
        for (size_t t = 0; t < frames; t++) {
            next = apply_fx_events(rack, ev, n_ev, next, (uint32_t) t);
            rack.Process(in + t * N, out + t * N);
        }
    }
    for (auto l : lines) delete l;
}

ol::synth::SynthVoice *new_voice() { return new ol::synth::SynthVoice(); }

}  // namespace

extern "C" {

const char *ref_describe(void) {
    return "reference headers (fxlib/Fx.h, synthlib/*.h, corelib) + DaisySP restatement shim; "
           "libs/dattorro-verb/verb.cpp compiled unmodified";
}

float ref_scale(float in, float inlow, float inhigh, float outlow, float outhigh, float power) {
    return ol::core::scale(in, inlow, inhigh, outlow, outhigh, power);
}

int ref_filterfx_run(int channels, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out,
                     size_t frames) {
    if (channels == 1) run_filterfx<1>(sr, ev, n_ev, in, out, frames);
    else if (channels == 2) run_filterfx<2>(sr, ev, n_ev, in, out, frames);
    else return -1;
    return 0;
}

int ref_delayfx_run(int channels, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out,
                    size_t frames) {
    if (channels == 1) run_delayfx<1>(sr, ev, n_ev, in, out, frames);
    else if (channels == 2) run_delayfx<2>(sr, ev, n_ev, in, out, frames);
    else return -1;
    return 0;
}

int ref_reverbfx_run(int channels, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out,
                     size_t frames) {
    if (channels == 1) run_reverbfx<1>(sr, ev, n_ev, in, out, frames);
    else if (channels == 2) run_reverbfx<2>(sr, ev, n_ev, in, out, frames);
    else return -1;
    return 0;
}

int ref_fxrack_run(int channels, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out,
                   size_t frames) {
    if (channels == 1) run_fxrack<1>(sr, ev, n_ev, in, out, frames);
    else if (channels == 2) run_fxrack<2>(sr, ev, n_ev, in, out, frames);
    else return -1;
    return 0;
}

// libs/dattorro-verb: params = {preDelay, preFilter, inputDiffusion1, inputDiffusion2, decay,
// decayDiffusion, damping}; a NaN entry keeps the library default (verb.cpp:215-221).
// out_l/out_r are read after each process() call, as DattorroVerb_getLeft/Right are meant to be used.
int ref_dattorro_run(const float *params7, const float *in, float *out_l, float *out_r, size_t frames) {
    sDattorroVerb *v = DattorroVerb_create();
    if (!v) return -1;
    if (params7) {
        if (!std::isnan(params7[0])) DattorroVerb_setPreDelay(v, params7[0]);
        if (!std::isnan(params7[1])) DattorroVerb_setPreFilter(v, params7[1]);
        if (!std::isnan(params7[2])) DattorroVerb_setInputDiffusion1(v, params7[2]);
        if (!std::isnan(params7[3])) DattorroVerb_setInputDiffusion2(v, params7[3]);
        if (!std::isnan(params7[4])) DattorroVerb_setDecay(v, params7[4]);
        if (!std::isnan(params7[5])) DattorroVerb_setDecayDiffusion(v, params7[5]);
        if (!std::isnan(params7[6])) DattorroVerb_setDamping(v, params7[6]);
    }
    for (size_t t = 0; t < frames; t++) {
        DattorroVerb_process(v, in[t]);
        out_l[t] = DattorroVerb_getLeft(v);
        out_r[t] = DattorroVerb_getRight(v);
    }
    DattorroVerb_delete(v);
    return 0;
}

// One SynthVoice (SynthVoice.h:16-316).  config16 (may be null) is applied with UpdateConfig after Init.
// trace (may be null) receives 4 uint32 per frame: amp-env segment, filter-env segment are not
// reachable through the reference's private members, so the trace is the output bit pattern only.
int ref_voice_run(float sr, const float *config16, const olb_event *ev, size_t n_ev, float *out, size_t frames) {
    ol::synth::SynthVoice *v = new_voice();
    v->Init(sr);
    if (config16) {
        ol::synth::Voice::Config c;
        memcpy(&c, config16, sizeof(c));
        v->UpdateConfig(c);
    }
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        next = apply_voice_events(*v, ev, n_ev, next, (uint32_t) t);
        float frame = 0.f;
        v->Process(&frame);
        out[t] = frame;
    }
    // SynthVoice never frees what its default arguments allocate (SynthVoice.h:20-24); leak is the reference's.
    delete v;
    return 0;
}

// Polyvoice over n SynthVoices (Polyvoice.h:12-86).  Events go to the Polyvoice (first-free
// allocation, Polyvoice.h:35-51) unless ev.voice >= 0, in which case they go to that voice directly.
// out[t] = sum over voices in index order, accumulated into a zeroed frame (main.cpp:79-80).
int ref_polyvoice_run(int n_voices, float sr, const float *config16, const olb_event *ev, size_t n_ev, float *out,
                      size_t frames) {
    std::vector<ol::synth::Voice *> voices;
    for (int i = 0; i < n_voices; i++) voices.push_back(new_voice());
    ol::synth::Polyvoice poly(voices);
    poly.Init(sr);
    if (config16) {
        ol::synth::Voice::Config c;
        memcpy(&c, config16, sizeof(c));
        poly.UpdateConfig(c);
    }
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        while (next < n_ev && ev[next].frame <= t) {
            const olb_event &e = ev[next];
            if (e.voice >= 0 && e.voice < n_voices) {
                apply_voice_events(*voices[e.voice], &e, 1, 0, (uint32_t) t);
            } else {
                apply_voice_events(poly, &e, 1, 0, (uint32_t) t);
            }
            next++;
        }
        float frame = 0.f;
        poly.Process(&frame);
        out[t] = frame;
    }
    for (auto v : voices) delete v;
    return 0;
}

// VoiceMap over n SynthVoices (VoiceMap.h:14-90): map3 = n_map x (channel, note, voice index) fed to SetVoice in order.
// Note events go to the voice mapped to that NOTE (VoiceMap.h:27-43), MIDI CCs to the voice last set on the CHANNEL
// carried in ev.voice (VoiceMap.h:80-87).  out[t] = the mapped voices in note order, accumulated into a zeroed frame.
int ref_voicemap_run(int n_voices, float sr, const float *config16, const int32_t *map3, size_t n_map, const olb_event *ev,
                     size_t n_ev, float *out, size_t frames) {
    std::vector<ol::synth::SynthVoice *> voices;
    for (int i = 0; i < n_voices; i++) voices.push_back(new_voice());
    ol::synth::VoiceMap<1> map;
    for (size_t m = 0; m < n_map; m++) {
        const int32_t v = map3[3 * m + 2];
        if (v < 0 || v >= n_voices) return 1;
        map.SetVoice((uint8_t) map3[3 * m], (uint8_t) map3[3 * m + 1], voices[v]);
    }
    map.Init(sr);
    if (config16) {
        ol::synth::Voice::Config c;
        memcpy(&c, config16, sizeof(c));
        for (auto v : voices) v->UpdateConfig(c);
    }
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        while (next < n_ev && ev[next].frame <= t) {
            const olb_event &e = ev[next++];
            switch (e.kind) {
                case OLB_EV_NOTE_ON: map.NoteOn((uint8_t) e.a, (uint8_t) e.v); break;
                case OLB_EV_NOTE_OFF: map.NoteOff((uint8_t) e.a, (uint8_t) e.v); break;
                case OLB_EV_MIDI_CC: map.UpdateMidiControl((uint8_t) e.voice, (uint8_t) e.a, (uint8_t) e.v); break;
                default: break;
            }
        }
        float frame = 0.f;
        map.Process(&frame);
        out[t] = frame;
    }
    for (auto v : voices) delete v;
    return 0;
}

// ol::synth::Sample + SampleSoundSource<CH> (Sample.cpp, SampleSoundSource.h) over an in-memory SampleDataSource (the
// reference's only implementation decodes files through miniaudio, workouts/workout_buddy.h:18-42; the interface is
// SampleDataSource.h:11-25: Read one frame, return frames read, Seek to a frame).  `out` [frames][channels] arrives
// pre-filled by the caller: Sample::Process leaves a frame untouched when it is not playing or nothing was read.
}  // extern "C"
namespace {
class MemDataSource : public ol::synth::SampleDataSource {
public:
    MemDataSource(const float *data, uint64_t frames, uint64_t channels) : data_(data), frames_(frames), channels_(channels) {}
    uint64_t GetChannelCount() override { return channels_; }
    ol::synth::InitStatus Init(t_sample, const char *) override { return ol::synth::Ok; }
    void Seek(uint64_t frame_index) override { pos_ = frame_index; }
    uint64_t Read(t_sample *out) override {
        if (pos_ >= frames_) return 0;
        for (uint64_t c = 0; c < channels_; c++) out[c] = data_[pos_ * channels_ + c];
        pos_++;
        return 1;
    }

private:
    const float *data_;
    uint64_t frames_, channels_, pos_ = 0;
};

template <int CH>
void run_sampler(const float *data, uint64_t sample_frames, const olb_event *ev, size_t n_ev, float *out, size_t frames) {
    MemDataSource src(data, sample_frames, CH);
    ol::synth::Sample sample(src);
    ol::synth::SampleSoundSource<CH> source(&sample);
    source.Init(48000.f);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        while (next < n_ev && ev[next].frame <= t) {
            const olb_event &e = ev[next++];
            switch (e.kind) {
                case OLB_EV_GATE_ON: source.GateOn(); break;
                case OLB_EV_GATE_OFF: source.GateOff(); break;
                case OLB_EV_SAMPLE_SEEK: sample.Seek(e.a); break;
                case OLB_EV_SAMPLE_LOOP_START: sample.SetLoopStart(e.a); break;
                case OLB_EV_SAMPLE_LOOP_END: sample.SetLoopEnd(e.a); break;
                case OLB_EV_SAMPLE_MODE: sample.SetPlayMode(e.a ? ol::synth::Loop : ol::synth::OneShot); break;
                case OLB_EV_SAMPLE_TOGGLE: sample.TogglePlay(); break;
                case OLB_EV_SAMPLE_PLAY: sample.Play(); break;
                case OLB_EV_SAMPLE_PAUSE: sample.Pause(); break;
                default: break;
            }
        }
        source.Process(out + t * CH);
    }
}
}  // namespace
extern "C" {
int ref_sampler_run(int channels, const float *data, uint64_t sample_frames, const olb_event *ev, size_t n_ev, float *out, size_t frames) {
    if (channels == 1) run_sampler<1>(data, sample_frames, ev, n_ev, out, frames);
    else if (channels == 2) run_sampler<2>(data, sample_frames, ev, n_ev, out, frames);
    else return 1;
    return 0;
}

// Replays test/fx_test.cpp:25-55 (TEST(FX, Delay)); returns the number of NaN outputs (reference expects 0).
int ref_test_fx_delay(float *out_48000) {
    RefDelayLine *dl = new RefDelayLine();
    std::vector<RefDelayLine *> delay_lines{dl};
    int nans = 0;
    {
        DelayFx<1> delay(delay_lines);
        delay.Init(128);
        t_sample in = 1;
        t_sample out = 0;
        delay.Process(&in, &out);
        t_sample sample_rate = 48000;
        t_sample freq = 20000;
        delay.UpdateHardwareControl(CC_DELAY_TIME, 0.5f);
        delay.Update();
        daisysp::Oscillator osc;
        delay.Init(sample_rate);
        osc.Init(sample_rate);
        osc.SetFreq(freq);
        for (int i = 0; i < sample_rate; i++) {
            in = osc.Process();
            delay.Process(&in, &out);
            if (std::isnan(out)) nans++;
            if (out_48000) out_48000[i] = out;
        }
    }
    delete dl;
    return nans;
}

// Replays test/synth_test.cpp:102-148 (TEST(Synth, VoiceDefaultConstructor)).
// frames4 receives the four frame_out values the test inspects; returns a bitmask of FAILED expectations.
int ref_test_synth_voice(float *frames4) {
    using namespace ol::synth;
    int fail = 0;
    SynthVoice *v = new SynthVoice();
    t_sample sample_rate = 48000;
    v->Init(sample_rate);
    if (v->Playing() != 0) fail |= 1;
    if (v->Gate()) fail |= 2;
    uint8_t midi_note = 60, velocity = 100;
    v->NoteOn(midi_note, velocity);
    if (v->Playing() != midi_note) fail |= 4;
    if (!v->Gate()) fail |= 8;
    v->NoteOff(midi_note, velocity);
    if (v->Playing() != 0) fail |= 16;
    if (v->Gate()) fail |= 32;

    t_sample frame_out = 1;
    v->Process(&frame_out);
    frames4[0] = frame_out;
    if (!(frame_out == 0)) fail |= 64;

    v->NoteOn(midi_note, velocity);
    v->Process(&frame_out);
    frames4[1] = frame_out;
    if (frame_out == 0 || frame_out == 1) fail |= 128;
    v->NoteOff(midi_note, velocity);

    Voice::Config config = {};
    config.amp_env_amount = 0;
    v->UpdateConfig(config);
    v->NoteOn(midi_note, velocity);
    v->Process(&frame_out);
    frames4[2] = frame_out;
    if (!(frame_out == 0)) fail |= 256;
    v->NoteOff(midi_note, velocity);

    config.amp_env_amount = 1;
    v->UpdateConfig(config);
    v->NoteOn(midi_note, velocity);
    v->Process(&frame_out);
    frames4[3] = frame_out;
    if (frame_out == 0) fail |= 512;
    v->NoteOff(midi_note, velocity);
    delete v;
    return fail;
}

// ---- control wire format: the reference's own SimpleSerializer::SerializeControl
// (modules/iolib/Serializer.h:113-133) and ol::ctl::Control conversions (modules/ctllib/Control.h:16-115) -------------
struct RefNullSerial : public ol::io::Serial {
    int Write(const char *, int size) override { return size; }
    int Write(const uint8_t *, int size) override { return size; }
    int Write(std::vector<uint8_t>, int size) override { return size; }
};
// which: 0 = Control(controller, MIDI value), 1 = Control(controller, float value), 2 = Control(controller, ADC value).
// out4 = controller, adc, midi, float bits; wire (>= 64 bytes) receives SerializeControl's bytes; returns their count.
int ref_control_wire(int which, uint32_t controller, double value, uint32_t *out4, uint8_t *wire) {
    ol::ctl::Control c;
    if (which == 0) c = ol::ctl::Control((ol::ctl::Control::CONTROLLER_TYPE) controller, (ol::ctl::Control::MIDI_TYPE) value);
    else if (which == 1) c = ol::ctl::Control((ol::ctl::Control::CONTROLLER_TYPE) controller, (ol::ctl::Control::FLOAT_TYPE) value);
    else c = ol::ctl::Control((ol::ctl::Control::CONTROLLER_TYPE) controller, (ol::ctl::Control::ADC_TYPE) value);
    out4[0] = c.GetController();
    out4[1] = c.GetAdcValue();
    out4[2] = c.GetMidiValue();
    const float f = c.GetFloatValue();
    memcpy(&out4[3], &f, 4);
    RefNullSerial serial;
    ol::io::SimpleSerializer ser(serial);
    std::vector<uint8_t> buf;
    ser.SerializeControl(c, buf);
    memcpy(wire, buf.data(), buf.size());
    return (int) buf.size();
}

}  // extern "C"
