// TEST INFRASTRUCTURE ONLY (oracle/).  Restatement of upstream DaisySP Source/Control/adsr.{h,cpp}
// (absent from /root/reference; see ../daisysp.h for the parity status: UNPINNED).
// Included by the reference at modules/synthlib/Adsr.h:9; call sites Adsr.h:37-74.
#pragma once
#ifndef OLB_DAISYSP_SHIM_ADSR_H
#define OLB_DAISYSP_SHIM_ADSR_H

#include <cmath>
#include <cstdint>

namespace daisysp {

enum {
    ADSR_SEG_IDLE = 0,
    ADSR_SEG_ATTACK = 1,
    ADSR_SEG_DECAY = 2,
    ADSR_SEG_RELEASE = 4,
};

class Adsr {
  public:
    Adsr() {}
    ~Adsr() {}

    void Init(float sample_rate, int blockSize = 1) {
        sample_rate_ = sample_rate / blockSize;
        attackShape_ = -1.f;
        attackTarget_ = 0.0f;
        attackTime_ = -1.f;
        decayTime_ = -1.f;
        releaseTime_ = -1.f;
        sus_level_ = 0.7f;
        x_ = 0.0f;
        gate_ = false;
        mode_ = ADSR_SEG_IDLE;
        SetTime(ADSR_SEG_ATTACK, 0.1f);
        SetTime(ADSR_SEG_DECAY, 0.1f);
        SetTime(ADSR_SEG_RELEASE, 0.1f);
    }

    void Retrigger(bool hard) {
        mode_ = ADSR_SEG_ATTACK;
        if (hard) x_ = 0.f;
    }

    float Process(bool gate) {
        float out = 0.0f;
        if (gate && !gate_)  // rising edge
            mode_ = ADSR_SEG_ATTACK;
        else if (!gate && gate_)  // falling edge
            mode_ = ADSR_SEG_RELEASE;
        gate_ = gate;

        float D0(attackD0_);
        if (mode_ == ADSR_SEG_DECAY)
            D0 = decayD0_;
        else if (mode_ == ADSR_SEG_RELEASE)
            D0 = releaseD0_;

        float target = mode_ == ADSR_SEG_DECAY ? sus_level_ : -0.01f;
        switch (mode_) {
            case ADSR_SEG_IDLE: out = 0.0f; break;
            case ADSR_SEG_ATTACK:
                x_ += D0 * (attackTarget_ - x_);
                out = x_;
                if (out > 1.f) {
                    x_ = out = 1.f;
                    mode_ = ADSR_SEG_DECAY;
                }
                break;
            case ADSR_SEG_DECAY:
            case ADSR_SEG_RELEASE:
                x_ += D0 * (target - x_);
                out = x_;
                if (out < 0.0f) {
                    x_ = out = 0.f;
                    mode_ = ADSR_SEG_IDLE;
                }
            default: break;
        }
        return out;
    }

    void SetTime(int seg, float time) {
        switch (seg) {
            case ADSR_SEG_ATTACK: SetAttackTime(time, 0.0f); break;
            case ADSR_SEG_DECAY: SetTimeConstant(time, decayTime_, decayD0_); break;
            case ADSR_SEG_RELEASE: SetTimeConstant(time, releaseTime_, releaseD0_); break;
            default: return;
        }
    }

    void SetAttackTime(float timeInS, float shape = 0.0f) {
        if ((timeInS != attackTime_) || (shape != attackShape_)) {
            attackTime_ = timeInS;
            attackShape_ = shape;
            if (timeInS > 0.f) {
                float x = shape;
                float target = 9.f * powf(x, 10.f) + 0.3f * x + 1.01f;
                attackTarget_ = target;
                float logTarget = logf(1.f - (1.f / target));  // -1 for decay
                attackD0_ = 1.f - expf(logTarget / (timeInS * sample_rate_));
            } else
                attackD0_ = 1.f;  // instant change
        }
    }
    void SetDecayTime(float timeInS) { SetTimeConstant(timeInS, decayTime_, decayD0_); }
    void SetReleaseTime(float timeInS) { SetTimeConstant(timeInS, releaseTime_, releaseD0_); }

    inline void SetSustainLevel(float sus_level) {
        sus_level = (sus_level <= 0.f) ? -0.01f : (sus_level > 1.f) ? 1.f : sus_level;
        sus_level_ = sus_level;
    }

    inline uint8_t GetCurrentSegment() { return mode_; }
    inline bool IsRunning() const { return mode_ != ADSR_SEG_IDLE; }

  private:
    void SetTimeConstant(float timeInS, float &time, float &coeff) {
        if (timeInS != time) {
            time = timeInS;
            if (time > 0.f) {
                const float target = logf(1. / M_E);
                coeff = 1.f - expf(target / (time * sample_rate_));
            } else
                coeff = 1.f;  // instant change
        }
    }

    float sus_level_{0.f};
    float x_{0.f};
    float attackShape_{-1.f};
    float attackTarget_{0.0f};
    float attackTime_{-1.0f};
    float decayTime_{-1.0f};
    float releaseTime_{-1.0f};
    float attackD0_{0.f};
    float decayD0_{0.f};
    float releaseD0_{0.f};
    int sample_rate_;
    uint8_t mode_{ADSR_SEG_IDLE};
    bool gate_{false};
};

}  // namespace daisysp
#endif
