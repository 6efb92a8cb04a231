// TEST INFRASTRUCTURE ONLY (oracle/): nothing under ol_dsp_b200/ or include/ may include this file.
//
// daisysp.h -- stand-in for electro-smith/DaisySP, which ol_dsp pulls from
// ../.ol_dsp-deps/libs/DaisySP at `"commit": "master"` (reference dependencies.json; i.e. UNPINNED)
// and which is absent from /root/reference and from this machine (no network).
//
// What is here is a from-memory restatement of the PUBLISHED upstream algorithms of the five
// primitives the hot path touches, written so that the reference's own headers
// (modules/fxlib/Fx.h, modules/synthlib/*.h) compile UNMODIFIED against it:
//
//   daisysp::Svf          upstream Source/Filters/svf.{h,cpp}        (call sites: synthlib/Filter.h:67-107)
//   daisysp::Oscillator   upstream Source/Synthesis/oscillator.{h,cpp} (OscillatorSoundSource.h:15-36, test/fx_test.cpp:41-46)
//   daisysp::DelayLine    upstream Source/Utility/delayline.h        (fxlib/Fx.h:176-217, test/fx_test.cpp:27-28)
//   daisysp::mtof/fclamp  upstream Source/Utility/dsp.h              (SynthVoice.h:248)
//   daisysp::Adsr         upstream Source/Control/adsr.{h,cpp}       (see Control/adsr.h next to this file)
//   daisysp::LadderFilter NOT restated (constants not reproducible from memory): compiles, aborts if run.
//
// PARITY STATUS: "unpinned" at this boundary.  The only pins the reference's tests hold for these
// primitives are zero / non-zero / not-NaN checks (test/fx_test.cpp:45-54, test/synth_test.cpp:118-147),
// which tests/test_oracle_pins.py replays.  Every numeric parity claim that crosses this file reads
// "vs reference headers + DaisySP restatement (unverified against upstream)".
#pragma once
#ifndef OLB_DAISYSP_SHIM_H
#define OLB_DAISYSP_SHIM_H

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#ifndef PI_F
#define PI_F 3.1415927410125732421875f
#endif
#ifndef TWOPI_F
#define TWOPI_F (2.0f * PI_F)
#endif

namespace daisysp {

inline float fmax_(float a, float b) { return a > b ? a : b; }
inline float fmin_(float a, float b) { return a < b ? a : b; }

// upstream dsp.h: fclamp(in,min,max) = fmin(fmax(in,min),max)
inline float fclamp(float in, float mn, float mx) { return fmin_(fmax_(in, mn), mx); }

// upstream dsp.h: mtof(m) = powf(2, (m - 69) / 12) * 440
inline float mtof(float m) { return powf(2.0f, (m - 69.0f) / 12.0f) * 440.0f; }

// ---------------------------------------------------------------------------------------------
// Svf: double-sampled Chamberlin state-variable filter with cubic band saturation ("drive").
class Svf {
  public:
    void Init(float sample_rate) {
        sr_ = sample_rate;
        fc_ = 200.0f;
        res_ = 0.5f;
        drive_ = 0.5f;
        pre_drive_ = 0.5f;
        freq_ = 0.25f;
        damp_ = 0.0f;
        notch_ = low_ = high_ = band_ = peak_ = input_ = 0.0f;
        out_notch_ = out_low_ = out_high_ = out_peak_ = out_band_ = 0.0f;
        fc_max_ = sr_ / 3.f;
    }

    void Process(float in) {
        input_ = in;
        // first pass
        notch_ = input_ - damp_ * band_;
        low_ = low_ + freq_ * band_;
        high_ = notch_ - low_;
        band_ = freq_ * high_ + band_ - drive_ * band_ * band_ * band_;
        out_low_ = 0.5f * low_;
        out_high_ = 0.5f * high_;
        out_band_ = 0.5f * band_;
        out_peak_ = 0.5f * (low_ - high_);
        out_notch_ = 0.5f * notch_;
        // second pass
        notch_ = input_ - damp_ * band_;
        low_ = low_ + freq_ * band_;
        high_ = notch_ - low_;
        band_ = freq_ * high_ + band_ - drive_ * band_ * band_ * band_;
        out_low_ += 0.5f * low_;
        out_high_ += 0.5f * high_;
        out_band_ += 0.5f * band_;
        out_peak_ += 0.5f * (low_ - high_);
        out_notch_ += 0.5f * notch_;
    }

    void SetFreq(float f) {
        fc_ = fclamp(f, 1.0e-6, fc_max_);
        // fs*2 because double sampled
        freq_ = 2.0f * sinf(PI_F * fmin_(0.25f, fc_ / (sr_ * 2.0f)));
        RecalcDamp();
    }

    void SetRes(float r) {
        res_ = fclamp(r, 0.f, 1.f);
        RecalcDamp();
        drive_ = pre_drive_ * res_;
    }

    void SetDrive(float d) {
        pre_drive_ = fclamp(d * 0.1f, 0.f, 1.f);
        drive_ = pre_drive_ * res_;
    }

    float Low() { return out_low_; }
    float High() { return out_high_; }
    float Band() { return out_band_; }
    float Notch() { return out_notch_; }
    float Peak() { return out_peak_; }

  private:
    void RecalcDamp() {
        damp_ = fmin_(2.0f * (1.0f - powf(res_, 0.25f)), fmin_(2.0f, 2.0f / freq_ - freq_ * 0.5f));
    }
    float sr_, fc_, res_, drive_, freq_, damp_;
    float notch_, low_, high_, band_, peak_;
    float input_;
    float out_low_, out_high_, out_band_, out_peak_, out_notch_;
    float pre_drive_, fc_max_;
};

// ---------------------------------------------------------------------------------------------
// LadderFilter: NOT restated (see header comment).  Present so synthlib/Filter.h:35-63 compiles.
class LadderFilter {
  public:
    void Init(float) {}
    void SetFreq(float) {}
    void SetRes(float) {}
    float Process(float) {
        fprintf(stderr, "daisysp_shim: LadderFilter is not restated (upstream source unavailable)\n");
        abort();
    }
};

// ---------------------------------------------------------------------------------------------
// Oscillator: normalised phase in [0,1); waveforms the hot path uses (sine default, PolyBLEP saw).
class Oscillator {
  public:
    enum {
        WAVE_SIN,
        WAVE_TRI,
        WAVE_SAW,
        WAVE_RAMP,
        WAVE_SQUARE,
        WAVE_POLYBLEP_TRI,
        WAVE_POLYBLEP_SAW,
        WAVE_POLYBLEP_SQUARE,
        WAVE_LAST,
    };

    void Init(float sample_rate) {
        sr_ = sample_rate;
        sr_recip_ = 1.0f / sample_rate;
        freq_ = 100.0f;
        amp_ = 0.5f;
        pw_ = 0.5f;
        phase_ = 0.0f;
        phase_inc_ = CalcPhaseInc(freq_);
        waveform_ = WAVE_SIN;
        eoc_ = true;
        eor_ = true;
    }

    inline void SetFreq(const float f) {
        freq_ = f;
        phase_inc_ = CalcPhaseInc(f);
    }
    inline void SetAmp(const float a) { amp_ = a; }
    inline void SetWaveform(const uint8_t wf) { waveform_ = wf < WAVE_LAST ? wf : WAVE_SIN; }
    inline void PhaseAdd(float p) { phase_ += p; }
    inline void Reset(float p = 0.0f) { phase_ = p; }

    float Process() {
        float out, t;
        switch (waveform_) {
            case WAVE_SIN: out = sinf(phase_ * TWOPI_F); break;
            case WAVE_TRI:
                t = -1.0f + (2.0f * phase_);
                out = 2.0f * (fabsf(t) - 0.5f);
                break;
            case WAVE_SAW: out = -1.0f * (((phase_ * 2.0f)) - 1.0f); break;
            case WAVE_RAMP: out = ((phase_ * 2.0f)) - 1.0f; break;
            case WAVE_SQUARE: out = phase_ < pw_ ? (1.0f) : -1.0f; break;
            case WAVE_POLYBLEP_SAW:
                t = phase_;
                out = (2.0f * t) - 1.0f;
                out -= Polyblep(phase_inc_, t);
                out *= -1.0f;
                break;
            default: out = 0.0f; break;  // PolyBLEP tri/square are not on the hot path
        }
        phase_ += phase_inc_;
        if (phase_ > 1.0f) {
            phase_ -= 1.0f;
            eoc_ = true;
        } else {
            eoc_ = false;
        }
        eor_ = (phase_ - phase_inc_ < 0.5f && phase_ >= 0.5f);
        return out * amp_;
    }

  private:
    static float Polyblep(float phase_inc, float t) {
        float dt = phase_inc;
        if (t < dt) {
            t /= dt;
            return t + t - t * t - 1.0f;
        } else if (t > 1.0f - dt) {
            t = (t - 1.0f) / dt;
            return t * t + t + t + 1.0f;
        } else {
            return 0.0f;
        }
    }
    float CalcPhaseInc(float f) { return f * sr_recip_; }
    uint8_t waveform_;
    float amp_, freq_, pw_;
    float sr_, sr_recip_, phase_, phase_inc_;
    bool eor_, eoc_;
};

// ---------------------------------------------------------------------------------------------
// DelayLine: backwards-running write pointer, linear-interpolated read.
template <typename T, size_t max_size>
class DelayLine {
  public:
    DelayLine() {}
    ~DelayLine() {}
    void Init() { Reset(); }
    void Reset() {
        for (size_t i = 0; i < max_size; i++) line_[i] = T(0);
        write_ptr_ = 0;
        delay_ = 1;
    }
    inline void SetDelay(size_t delay) {
        frac_ = 0.0f;
        delay_ = delay < max_size ? delay : max_size - 1;
    }
    inline void SetDelay(float delay) {
        int32_t int_delay = static_cast<int32_t>(delay);
        frac_ = delay - static_cast<float>(int_delay);
        delay_ = static_cast<size_t>(int_delay) < max_size ? int_delay : max_size - 1;
    }
    inline void Write(const T sample) {
        line_[write_ptr_] = sample;
        write_ptr_ = (write_ptr_ - 1 + max_size) % max_size;
    }
    inline const T Read() const {
        T a = line_[(write_ptr_ + delay_) % max_size];
        T b = line_[(write_ptr_ + delay_ + 1) % max_size];
        return a + (b - a) * frac_;
    }

  private:
    // upstream leaves frac_ uninitialised until the first SetDelay(); zero here so the
    // reference's first Process() before Update() (test/fx_test.cpp:32-35) is defined.
    float frac_ = 0.0f;
    size_t write_ptr_;
    size_t delay_;
    T line_[max_size];
};

}  // namespace daisysp

#include "Control/adsr.h"

#endif  // OLB_DAISYSP_SHIM_H
