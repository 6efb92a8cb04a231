// host_params.h -- setter-time parameter maths of the banks, run on the HOST (the reference also computes its
// coefficients in setters, never per sample: SURVEY.md sections 2.3, 3.4).  The device only ever sees the derived
// coefficients.  Follows, by file:line under the reference tree:
//   ol::core::scale / safediv           modules/corelib/ol_corelib.h:27-44
//   CC numbers                          modules/corelib/cc_map.h:7-70
//   FilterFx setters                    modules/fxlib/Fx.h:110-164
//   DelayFx setters / Init              modules/fxlib/Fx.h:183-192, 210-266
//   ReverbFx setters                    modules/fxlib/Fx.h:302-392
//   FxRack setters / Init               modules/fxlib/Fx.h:416-424, 436-492
// and, for the un-vendored DaisySP primitives (upstream Source/Filters/svf.cpp, Source/Utility/delayline.h), the
// published algorithm as summarised in SURVEY.md section 2.3 (parity unpinned at that boundary).
#pragma once
#include <math.h>
#include <stdint.h>

namespace olb {
namespace host {

enum {
    CC_CTL_PORTAMENTO = 5, CC_CTL_VOLUME = 7,
    CC_REVERB_TIME = 32, CC_REVERB_CUTOFF = 33, CC_REVERB_BALANCE = 34,
    CC_DELAY_TIME = 35, CC_DELAY_FEEDBACK = 36, CC_DELAY_CUTOFF = 37, CC_DELAY_RESONANCE = 38, CC_DELAY_BALANCE = 39,
    CC_FILTER_CUTOFF = 41, CC_FILTER_RESONANCE = 42, CC_FILTER_TYPE = 43, CC_FILTER_DRIVE = 44,
    CC_FX_FILTER_CUTOFF = 45, CC_FX_FILTER_RESONANCE = 46, CC_FX_FILTER_TYPE = 47, CC_FX_FILTER_DRIVE = 48,
    CC_EARLY_PREDELAY = 50, CC_REVERB_PREDELAY = 52, CC_REVERB_PREFILTER = 53,
    CC_REVERB_INPUT_DIFFUSION_1 = 54, CC_REVERB_INPUT_DIFFUSION_2 = 55, CC_REVERB_DECAY_DIFFUSION = 56,
    CC_ENV_FILT_AMT = 73, CC_ENV_FILT_A = 74, CC_ENV_FILT_D = 75, CC_ENV_FILT_S = 76, CC_ENV_FILT_R = 77,
    CC_ENV_AMP_A = 108, CC_ENV_AMP_D = 109, CC_ENV_AMP_S = 110, CC_ENV_AMP_R = 111, CC_OSC_1_VOLUME = 114,
};

constexpr int kMaxDelay = 48000;  // MAX_DELAY, modules/fxlib/Fx.h:23

// ol_corelib.h:27-29: the comparison is against a double zero, the quotient is narrowed to float
inline float safediv(float num, float denom) { return denom == 0. ? 0.f : (float) (num / denom); }

// ol_corelib.h:31-44: pow is applied in double and narrowed on assignment
inline float scale(float in, float inlow, float inhigh, float outlow, float outhigh, float power) {
    const float inscale = safediv(1.f, inhigh - inlow);
    const float outdiff = outhigh - outlow;
    float value = (in - inlow) * inscale;
    if (value > 0.0) value = (float) pow((double) value, (double) power);
    else if (value < 0.0) value = (float) -pow((double) -value, (double) power);
    return (value * outdiff) + outlow;
}

// ---- daisysp::Svf, setter side ------------------------------------------------------------------
struct SvfParams {
    float sr = 48000.f, fc_max = 16000.f, res = 0.5f, pre_drive = 0.5f;
    float freq = 0.25f, damp = 0.f, drive = 0.5f;  // what Process() reads
    static float fclamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
    void recalc_damp() {
        const float a = 2.0f * (1.0f - powf(res, 0.25f)), b = fminf(2.0f, 2.0f / freq - freq * 0.5f);
        damp = fminf(a, b);
    }
    void Init(float sample_rate) {
        sr = sample_rate; res = 0.5f; drive = 0.5f; pre_drive = 0.5f; freq = 0.25f; damp = 0.f;
        fc_max = sr / 3.f;
    }
    void SetFreq(float hz) {
        const float fc = fclamp(hz, 1.0e-6, fc_max);
        freq = 2.0f * sinf(3.1415927410125732421875f * fminf(0.25f, fc / (sr * 2.0f)));
        recalc_damp();
    }
    void SetRes(float r) {
        res = fclamp(r, 0.f, 1.f);
        recalc_damp();
        drive = pre_drive * res;
    }
    void SetDrive(float d) {
        pre_drive = fclamp(d * 0.1f, 0.f, 1.f);
        drive = pre_drive * res;
    }
};

// ---- ol::fx::FilterFx (Fx.h:65-165) -------------------------------------------------------------
struct FilterParams {
    float cutoff = 20000.f, resonance = 0.f, drive = 0.f;
    int type = 0;  // LowPass, BandPass, HighPass, Notch, Peak
    SvfParams svf;
    void Update() { svf.SetFreq(cutoff); svf.SetRes(resonance); svf.SetDrive(drive); }
    void Init(float sr) { svf.Init(sr); Update(); }
    bool UpdateMidiControl(uint8_t cc, uint8_t value) {
        const float scaled = scale(value, 0, 127, 0, 1, 1);
        switch (cc) {
            case CC_FILTER_CUTOFF: cutoff = scale(value, 0, 127, 0, 20000, 1); break;
            case CC_FILTER_RESONANCE: resonance = scaled; break;
            case CC_FILTER_DRIVE: drive = scaled; break;
            case CC_FILTER_TYPE: type = (int) scale(value, 0, 127, 0, 5, 1); break;
            default: return false;
        }
        Update();
        return true;
    }
    bool UpdateHardwareControl(uint8_t cc, float value) {
        switch (cc) {
            case CC_FILTER_CUTOFF: cutoff = scale(value, 0, 1, 0, 20000, 1.02f); break;
            case CC_FILTER_RESONANCE: resonance = value; break;
            case CC_FILTER_DRIVE: drive = value; break;
            case CC_FILTER_TYPE: type = (int) scale(value, 0, 1, 0, 5, 1); break;
            default: return false;
        }
        Update();
        return true;
    }
};

// ---- ol::fx::DelayFx (Fx.h:170-268) over daisysp::DelayLine<float, 48000> -------------------------
struct DelayParams {
    float time = 0.5f, feedback = 0.5f, balance = 0.33f;
    uint32_t delay = 1;  // DelayLine::delay_
    float frac = 0.f;    // DelayLine::frac_
    FilterParams wet;
    void SetDelay(float d) {  // DelayLine::SetDelay(float)
        const int32_t i = (int32_t) d;
        frac = d - (float) i;
        delay = (size_t) i < (size_t) kMaxDelay ? (uint32_t) i : (uint32_t) (kMaxDelay - 1);
    }
    void Update() {
        SetDelay(scale(time, 0, 1, 0, kMaxDelay, 1));
        wet.Update();
    }
    void Init(float sr) {  // the rings are cleared by the bank; DelayLine::Init leaves frac_ alone and sets delay_ = 1
        delay = 1;
        wet.Init(sr);
        wet.UpdateMidiControl(CC_FILTER_CUTOFF, 64);
        wet.UpdateMidiControl(CC_FILTER_RESONANCE, 24);
        Update();
    }
    bool UpdateHardwareControl(uint8_t cc, float value) {
        switch (cc) {
            case CC_DELAY_TIME: time = value; break;
            case CC_DELAY_FEEDBACK: feedback = value; break;
            case CC_DELAY_BALANCE: balance = value; break;
            default: return false;
        }
        Update();
        return true;
    }
    bool UpdateMidiControl(uint8_t cc, uint8_t value) {
        const float scaled = scale(value, 0, 127, 0, 1, 1);
        switch (cc) {
            case CC_DELAY_TIME: time = scaled; break;
            case CC_DELAY_FEEDBACK: feedback = scaled; break;
            case CC_DELAY_BALANCE: balance = scaled; break;
            case CC_DELAY_CUTOFF: wet.UpdateMidiControl(CC_FILTER_CUTOFF, value); break;
            case CC_DELAY_RESONANCE: wet.UpdateMidiControl(CC_FILTER_RESONANCE, value); break;
            default: return false;
        }
        Update();
        return true;
    }
};

// ---- ol::fx::ReverbFx (Fx.h:272-393).  With the stub ReverbSc (Reverb.h:13-41) only `balance` is audible; the other
// members are kept (with the reference's copy-paste bug: INPUT_DIFFUSION_2 writes decay_diffusion) so a Dattorro back end
// can be driven from the same fields.
struct ReverbParams {
    float decay_time = 0.5f, cutoff = 12000.f, early_predelay = 0.1f, predelay = 0.2f, pre_cutoff = 0.5f;
    float input_diffusion1 = 0.5f, input_diffusion2 = 0.5f, decay_diffusion = 0.5f, balance = 0.1f;
    bool set(uint8_t cc, float scaled, float cutoff_scaled) {
        switch (cc) {
            case CC_REVERB_DECAY_DIFFUSION: decay_diffusion = scaled; break;
            case CC_REVERB_INPUT_DIFFUSION_1: input_diffusion1 = scaled; break;
            case CC_REVERB_INPUT_DIFFUSION_2: decay_diffusion = scaled; break;  // sic, Fx.h:323-324 and 363-364
            case CC_REVERB_CUTOFF: cutoff = cutoff_scaled; break;
            case CC_REVERB_BALANCE: balance = scaled; break;
            case CC_REVERB_PREDELAY: predelay = scaled; break;
            case CC_EARLY_PREDELAY: early_predelay = scaled; break;
            case CC_REVERB_PREFILTER: pre_cutoff = scaled; break;
            case CC_REVERB_TIME: decay_time = scaled; break;
            default: return false;
        }
        return true;
    }
    bool UpdateMidiControl(uint8_t cc, uint8_t v) { return set(cc, scale(v, 0, 127, 0, 1, 1), scale(v, 0, 127, 0, 20000, 1)); }
    bool UpdateHardwareControl(uint8_t cc, float v) { return set(cc, v, scale(v, 0, 1, 0, 20000, 1)); }
};

// ---- ol::fx::FxRack (Fx.h:397-493) ---------------------------------------------------------------
struct RackParams {
    DelayParams delay;
    ReverbParams reverb;
    FilterParams filter;
    float master_volume = 0.8f;
    void Init(float sr) {
        delay.Init(sr);
        filter.Init(sr);
        delay.Update();
        filter.Update();
    }
    void UpdateMidiControl(uint8_t cc, uint8_t value) {
        switch (cc) {
            case CC_FX_FILTER_CUTOFF: filter.UpdateMidiControl(CC_FILTER_CUTOFF, value); break;
            case CC_FX_FILTER_RESONANCE: filter.UpdateMidiControl(CC_FILTER_RESONANCE, value); break;
            case CC_FX_FILTER_DRIVE: filter.UpdateMidiControl(CC_FILTER_DRIVE, value); break;
            case CC_FX_FILTER_TYPE: filter.UpdateMidiControl(CC_FILTER_TYPE, value); break;
            default: break;
        }
        delay.UpdateMidiControl(cc, value);
        reverb.UpdateMidiControl(cc, value);
        if (cc == CC_CTL_VOLUME) master_volume = scale(value, 0, 127, 0, 1, 1);
    }
    void UpdateHardwareControl(uint8_t cc, float value) {
        switch (cc) {
            case CC_FX_FILTER_CUTOFF: filter.UpdateHardwareControl(CC_FILTER_CUTOFF, value); break;
            case CC_FX_FILTER_RESONANCE: filter.UpdateHardwareControl(CC_FILTER_RESONANCE, value); break;
            case CC_FX_FILTER_DRIVE: filter.UpdateHardwareControl(CC_FILTER_DRIVE, value); break;
            case CC_FX_FILTER_TYPE: filter.UpdateHardwareControl(CC_FILTER_TYPE, value); break;
            case CC_CTL_VOLUME: master_volume = value; break;
            default: break;
        }
        delay.UpdateHardwareControl(cc, value);
        reverb.UpdateHardwareControl(cc, value);
    }
};

// ---- daisysp::Adsr, setter side (upstream Source/Control/adsr.cpp; parity unpinned) -----------------
struct AdsrParams {
    int sr = 48000;
    float sus = 0.7f, a_shape = -1.f, a_target = 0.f, a_time = -1.f, d_time = -1.f, r_time = -1.f;
    float a_d0 = 0.f, d_d0 = 0.f, r_d0 = 0.f;
    void SetTime(float t, float *time, float *coeff) {
        if (t != *time) {
            *time = t;
            if (t > 0.f) {
                const float target = logf(1. / M_E);
                *coeff = 1.f - expf(target / (t * sr));
            } else {
                *coeff = 1.f;
            }
        }
    }
    void SetAttackTime(float t, float shape) {
        if (t != a_time || shape != a_shape) {
            a_time = t;
            a_shape = shape;
            if (t > 0.f) {
                const float target = 9.f * powf(shape, 10.f) + 0.3f * shape + 1.01f;
                a_target = target;
                const float lt = logf(1.f - (1.f / target));
                a_d0 = 1.f - expf(lt / (t * sr));
            } else {
                a_d0 = 1.f;
            }
        }
    }
    void SetDecayTime(float t) { SetTime(t, &d_time, &d_d0); }
    void SetReleaseTime(float t) { SetTime(t, &r_time, &r_d0); }
    void SetSustainLevel(float s) { sus = (s <= 0.f) ? -0.01f : (s > 1.f) ? 1.f : s; }
    void Init(float sample_rate, int block) {
        *this = AdsrParams();
        sr = (int) (sample_rate / block);
        SetAttackTime(0.1f, 0.0f);
        SetDecayTime(0.1f);
        SetReleaseTime(0.1f);
    }
};

// ---- ol::synth::SynthVoice (modules/synthlib/SynthVoice.h:16-316) with its default parts: OscillatorSoundSource
// (PolyBLEP saw, OscillatorSoundSource.h:12-38), SvfFilter (Filter.h:65-108), two DaisyAdsr (Adsr.h:35-76) and
// DaisyPortamento over the in-tree daisysp::Port (Portamento.h:13-42, 57-72).
struct VoiceParams {
    // members with the defaults of SynthVoice.h:269-315
    float filter_cutoff = 0.f, filter_resonance = 0.f, filter_drive = 0.f;
    float filter_attack = 0.f, filter_attack_shape = 1.f, filter_decay = 0.2f, filter_sustain = 0.f, filter_release = 0.f, filter_env_amount = 1.f;
    float amp_attack = 0.01f, amp_attack_shape = 1.f, amp_decay = 0.f, amp_sustain = 1.f, amp_release = 0.01f, amp_env_amount = 0.8f;
    float portamento_htime = 0.f, osc_1_mix = 0.8f;
    float sample_rate = 48000.f, freq = 0.f;
    uint8_t playing = 0;
    int gate = 0;
    bool retrigger = false;  // NoteOn: both envelopes Retrigger(hard) before the next frame
    SvfParams svf;
    AdsrParams filter_env, amp_env;
    float port_sr = 48000.f, port_coeff = 0.f;

    void SetHtime(float h) { port_coeff = expf(-1.0f / (h * port_sr)); }  // Portamento.h:34-37
    void Update() {  // SynthVoice.h:79-98 (SetFreq is commented out there)
        svf.SetRes(filter_resonance);
        svf.SetDrive(filter_drive);
        filter_env.SetAttackTime(filter_attack, filter_attack_shape);
        filter_env.SetDecayTime(filter_decay);
        filter_env.SetSustainLevel(filter_sustain);
        filter_env.SetReleaseTime(filter_release);
        amp_env.SetAttackTime(amp_attack, amp_attack_shape);
        amp_env.SetDecayTime(amp_decay);
        amp_env.SetSustainLevel(amp_sustain);
        amp_env.SetReleaseTime(amp_release);
        SetHtime(portamento_htime);
    }
    void Init(float sr) {  // SynthVoice.h:31-39: Update() is NOT called
        sample_rate = sr;
        svf.Init(sr);
        filter_env.Init(sr, 1);
        amp_env.Init(sr, 1);
        port_sr = sr;
        SetHtime(portamento_htime);
    }
    void UpdateConfig(const float *c) {  // Voice::Config field order, modules/synthlib/Voice.h:14-31
        filter_cutoff = c[0]; filter_resonance = c[1]; filter_drive = c[2]; filter_env_amount = c[3];
        filter_attack = c[4]; filter_attack_shape = c[5]; filter_decay = c[6]; filter_sustain = c[7]; filter_release = c[8];
        amp_env_amount = c[9]; amp_attack = c[10]; amp_attack_shape = c[11]; amp_decay = c[12]; amp_sustain = c[13]; amp_release = c[14];
        portamento_htime = c[15];
        Update();
    }
    bool set(uint8_t cc, float s, float porta, float cutoff, float fdecay) {
        switch (cc) {
            case CC_CTL_VOLUME: amp_env_amount = s; break;
            case CC_CTL_PORTAMENTO: portamento_htime = porta; break;
            case CC_FILTER_CUTOFF: filter_cutoff = cutoff; break;
            case CC_FILTER_RESONANCE: filter_resonance = s; break;
            case CC_FILTER_DRIVE: filter_drive = s; break;
            case CC_ENV_FILT_AMT: filter_env_amount = s; break;
            case CC_ENV_FILT_A: filter_attack = s; break;
            case CC_ENV_FILT_D: filter_decay = fdecay; break;
            case CC_ENV_FILT_S: filter_sustain = s; break;
            case CC_ENV_FILT_R: filter_release = s; break;
            case CC_ENV_AMP_A: amp_attack = s; break;
            case CC_ENV_AMP_D: amp_decay = s; break;
            case CC_ENV_AMP_S: amp_sustain = s; break;
            case CC_ENV_AMP_R: amp_release = s; break;
            case CC_OSC_1_VOLUME: osc_1_mix = s; break;
            default: return false;
        }
        Update();
        return true;
    }
    bool UpdateHardwareControl(uint8_t cc, float x) {  // SynthVoice.h:101-156
        return set(cc, x, scale(x, 0, 1, 0, 1, 4), scale(x, 0, 1, 0, 20000, 2.5f), scale(x, 0, 1, 0, 1, 3));
    }
    bool UpdateMidiControl(uint8_t cc, uint8_t v) {  // SynthVoice.h:158-229 (the printf in every case is dropped)
        return set(cc, scale(v, 0, 127, 0, 1, 1), scale(v, 0, 127, 0, 1, 4), scale(v, 0, 127, 0, 20000, 2.5f), scale(v, 0, 127, 0, 1, 3));
    }
    static float mtof(float m) { return powf(2.0f, (m - 69.0f) / 12.0f) * 440.0f; }
    void NoteOn(uint8_t note) {  // SynthVoice.h:245-251
        gate = 1; playing = note; freq = mtof(note); retrigger = true;
    }
    void NoteOff() { gate = 0; playing = 0; }  // SynthVoice.h:253-256
};

}  // namespace host
}  // namespace olb
