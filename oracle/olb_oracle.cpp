// TEST INFRASTRUCTURE ONLY (oracle/).  Builds oracle/libolb_oracle.so.
//
// CPU restatement of the hot path of oletizi/ol_dsp, written from the
// reference's behaviour, in flat C-style C++ (structs + free functions), one float operation per
// IEEE rounding (build flags: -O2 -ffp-contract=off -fno-fast-math).  It is the checker that
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs compare the
// CUDA bank against.  The product (ol_dsp_b200/, include/) never includes, links or loads it.
//
// Pinning (see DESIGN.md "Oracle"):
//   * Dattorro plate            -> pinned: bit-exact against libs/dattorro-verb/verb.cpp compiled unmodified
//                                  (oracle/_ref/libolref.so) and against tests/golden/dattorro_*.npz
//   * scale / FilterFx / DelayFx / ReverbFx / FxRack / SynthVoice / Polyvoice control + glue logic
//                               -> pinned against the reference's own headers run through oracle/_ref
//   * Svf / Oscillator / Adsr / DelayLine arithmetic -> PARITY UNPINNED: DaisySP (dependencies.json,
//                                  "commit": "master") is absent; restated from the published upstream
//                                  algorithm; only the reference's zero / non-zero / not-NaN tests pin it
//   * biquad / one-pole cascade -> no reference implementation exists; this file IS the
//                                  specification; parity unpinned by construction
//   * chorus                    -> spec-level restatement of modules/rnbo/patcher/mono-chorus.rnbopat + pitchshift.gendsp; unpinned
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "oracle_abi.h"

#define ORC_API extern "C" __attribute__((visibility("default")))

// ---------------------------------------------------------------------------------------------------
// corelib: ol::core::scale  (modules/corelib/ol_corelib.h:27-44)
// safediv compares the float denominator with a double zero; pow is the double overload applied to
// float arguments and the result is narrowed back to float on assignment.
static float orc_safediv(float num, float den) { return den == 0. ? 0.f : (float) (num / den); }

ORC_API float orc_scale(float in, float inlow, float inhigh, float outlow, float outhigh, float power) {
    float inscale = orc_safediv(1.f, inhigh - inlow);
    float outdiff = outhigh - outlow;
    float value = (in - inlow) * inscale;
    if (value > 0.0)
        value = (float) pow((double) value, (double) power);
    else if (value < 0.0)
        value = (float) -pow((double) -value, (double) power);
    return (value * outdiff) + outlow;
}

// CC numbers (modules/corelib/cc_map.h:7-70), only the ones the hot path reads.
enum {
    kCcPortamento = 5, kCcVolume = 7,
    kCcReverbTime = 32, kCcReverbCutoff = 33, kCcReverbBalance = 34,
    kCcDelayTime = 35, kCcDelayFeedback = 36, kCcDelayCutoff = 37, kCcDelayResonance = 38, kCcDelayBalance = 39,
    kCcFilterCutoff = 41, kCcFilterResonance = 42, kCcFilterType = 43, kCcFilterDrive = 44,
    kCcFxFilterCutoff = 45, kCcFxFilterResonance = 46, kCcFxFilterType = 47, kCcFxFilterDrive = 48,
    kCcEarlyPredelay = 50, kCcReverbPredelay = 52, kCcReverbPrefilter = 53,
    kCcReverbInDiff1 = 54, kCcReverbInDiff2 = 55, kCcReverbDecayDiff = 56,
    kCcEnvFiltAmt = 73, kCcEnvFiltA = 74, kCcEnvFiltD = 75, kCcEnvFiltS = 76, kCcEnvFiltR = 77,
    kCcEnvAmpA = 108, kCcEnvAmpD = 109, kCcEnvAmpS = 110, kCcEnvAmpR = 111, kCcOsc1Volume = 114,
};

// ---------------------------------------------------------------------------------------------------
// Counter-based white noise shared by CPU and GPU: Philox4x32-10,
// key = (seed, 0x01D5B200), counter = (frame/4 lo, frame/4 hi, channel, 0); word (frame & 3) of the
// output block gives the sample:  x = (u >> 8) * 2^-23 - 1  in [-1, 1), exactly representable.
static inline void philox_round(uint32_t c[4], const uint32_t k[2]) {
    const uint64_t p0 = (uint64_t) 0xD2511F53u * c[0];
    const uint64_t p1 = (uint64_t) 0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t) (p1 >> 32) ^ c[1] ^ k[0];
    const uint32_t n1 = (uint32_t) p1;
    const uint32_t n2 = (uint32_t) (p0 >> 32) ^ c[3] ^ k[1];
    const uint32_t n3 = (uint32_t) p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

static inline void philox4x32_10(uint32_t seed, uint32_t channel, uint64_t block, uint32_t out[4]) {
    uint32_t c[4] = {(uint32_t) block, (uint32_t) (block >> 32), channel, 0u};
    uint32_t k[2] = {seed, 0x01D5B200u};
    for (int r = 0; r < 10; r++) {
        philox_round(c, k);
        k[0] += 0x9E3779B9u;
        k[1] += 0xBB67AE85u;
    }
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}

static inline float noise_from_u32(uint32_t u) { return (float) (u >> 8) * 1.1920928955078125e-07f - 1.0f; }

static float orc_noise_sample(uint32_t seed, uint32_t channel, uint64_t frame) {
    uint32_t w[4];
    philox4x32_10(seed, channel, frame >> 2, w);
    return noise_from_u32(w[frame & 3]);
}

// out[(t * ld) + c] for t in [0, frames), c in [0, channels): frame-interleaved when ld == channels.
ORC_API void orc_noise_fill(uint32_t seed, uint32_t first_channel, uint32_t channels, uint64_t first_frame,
                            uint64_t frames, float scale, float *out, uint64_t ld_frames, uint64_t ld_channels) {
    for (uint32_t c = 0; c < channels; c++) {
        uint64_t t = 0;
        while (t < frames) {
            uint64_t f = first_frame + t;
            uint32_t w[4];
            philox4x32_10(seed, first_channel + c, f >> 2, w);
            for (uint64_t j = f & 3; j < 4 && t < frames; j++, t++)
                out[t * ld_frames + c * ld_channels] = noise_from_u32(w[j]) * scale;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// (a1) Biquad / one-pole cascade bank.  No reference class exists; the nearest reference objects are
// daisysp::Port (modules/synthlib/Portamento.h:29-32) and LowPassFilter_process
// (libs/dattorro-verb/verb.cpp:131-134), which are reproduced as dedicated section kinds.
//
// coefs: [channels][sections][5] = b0 b1 b2 a1 a2     state: [channels][sections][4] = x1 x2 y1 y2
// Section kinds (one per section index, shared by every channel of the bank):
//   0 BIQUAD_DF1   y = b0*x + (((b1*x1 + b2*x2) - a2*y2) - a1*y1)        (feedback of the newest sample enters last)
//   1 ONEPOLE_PORT z = x + c*(z - x),  c = coefs[0], z kept in y1         (Portamento.h:29-32)
//   2 ONEPOLE_LPF  o = o + (x - o)*f,  f = coefs[0], o kept in y1         (verb.cpp:131-134)
// Modes:
//   0 EXACT  every product and sum rounded separately, kinds honoured
//   1 FMA    every section runs the fused biquad recurrence
//              p = fma(b2,x2, b1*x1); p = fma(-a2,y2,p); p = fma(-a1,y1,p); y = fma(b0,x,p)
//            one-pole kinds are first rewritten as biquads:  PORT: b0 = 1-c, a1 = -c;  LPF: b0 = f, a1 = -(1-f)
enum { ORC_KIND_BIQUAD = 0, ORC_KIND_PORT = 1, ORC_KIND_LPF = 2 };
//   2 FAST   the normalised fused cascade (the throughput tier; tolerance, not bit-equal to EXACT):
//            one-pole kinds rewritten as in FMA; per section r1 = b1/b0, r2 = b2/b0 (double division, rounded to
//            float); per channel G = prod(b0) (double product, rounded to float); v = G*x, then per section
//              t = fma(r1,x1,v); t = fma(r2,x2,t); t = fma(-a2,y2,t); y = fma(-a1,y1,t)
//            so node s of the cascade carries the true signal times prod_{j>=s} b0_j (state is kept in that scaling).
//            Needs every b0 finite and non-zero and every partial product within [1e-30, 1e30]
//            (orc_biquad_fast_coefs returns 0 otherwise and callers fall back to FMA).
enum { ORC_MODE_EXACT = 0, ORC_MODE_FMA = 1, ORC_MODE_FAST = 2 };
enum { ORC_LAYOUT_FRAMES = 0 /* [T][C] */, ORC_LAYOUT_PLANAR = 1 /* [C][T] */ };

static inline float biquad_exact(const float *k, float *s, float x) {
    float p = k[1] * s[0] + k[2] * s[1];
    p = p - k[4] * s[3];
    p = p - k[3] * s[2];
    float y = k[0] * x + p;
    s[1] = s[0]; s[0] = x; s[3] = s[2]; s[2] = y;
    return y;
}

static inline float biquad_fma(const float *k, float *s, float x) {
    float p = __builtin_fmaf(k[2], s[1], k[1] * s[0]);
    p = __builtin_fmaf(-k[4], s[3], p);
    p = __builtin_fmaf(-k[3], s[2], p);
    float y = __builtin_fmaf(k[0], x, p);
    s[1] = s[0]; s[0] = x; s[3] = s[2]; s[2] = y;
    return y;
}

ORC_API void orc_biquad_fma_coefs(int kind, const float *in5, float *out5) {
    if (kind == ORC_KIND_PORT) {
        out5[0] = 1.0f - in5[0]; out5[1] = 0.f; out5[2] = 0.f; out5[3] = -in5[0]; out5[4] = 0.f;
    } else if (kind == ORC_KIND_LPF) {
        out5[0] = in5[0]; out5[1] = 0.f; out5[2] = 0.f; out5[3] = -(1.0f - in5[0]); out5[4] = 0.f;
    } else {
        memcpy(out5, in5, 5 * sizeof(float));
    }
}

// out: [sections][4] = r1 r2 -a1 -a2, *gain = G.  Returns 1 when the channel can run in FAST mode.
ORC_API int orc_biquad_fast_coefs(uint32_t sections, const uint8_t *kinds, const float *coefs, float *out4, float *gain) {
    double g = 1.0;
    int ok = 1;
    for (uint32_t s = sections; s-- > 0;) {  // partial products from the output side: the scale of node s
        float k[5];
        orc_biquad_fma_coefs(kinds ? kinds[s] : 0, coefs + 5 * s, k);
        const double b0 = k[0];
        if (!(b0 != 0.0) || !std::isfinite(b0)) { ok = 0; break; }
        g *= b0;
        if (!(fabs(g) >= 1e-30 && fabs(g) <= 1e30)) { ok = 0; break; }
        out4[4 * s + 0] = (float) ((double) k[1] / b0);
        out4[4 * s + 1] = (float) ((double) k[2] / b0);
        out4[4 * s + 2] = -k[3];
        out4[4 * s + 3] = -k[4];
    }
    *gain = (float) g;
    return ok;
}

static inline float biquad_fast(const float *k4, float *s, float x) {
    float t = __builtin_fmaf(k4[0], s[0], x);
    t = __builtin_fmaf(k4[1], s[1], t);
    t = __builtin_fmaf(k4[3], s[3], t);
    float y = __builtin_fmaf(k4[2], s[2], t);
    s[1] = s[0]; s[0] = x; s[3] = s[2]; s[2] = y;
    return y;
}

static void biquad_channel(int mode, uint32_t sections, const uint8_t *kinds, const float *coefs, float *state,
                           const float *in, float *out, uint64_t frames, uint64_t stride) {
    float kf[64 * 5];
    float gain = 1.f;
    if (mode == ORC_MODE_FAST && !orc_biquad_fast_coefs(sections, kinds, coefs, kf, &gain)) mode = ORC_MODE_FMA;
    if (mode == ORC_MODE_FMA)
        for (uint32_t s = 0; s < sections; s++) orc_biquad_fma_coefs(kinds ? kinds[s] : 0, coefs + 5 * s, kf + 5 * s);
    for (uint64_t t = 0; t < frames; t++) {
        float v = in[t * stride];
        if (mode == ORC_MODE_FAST) {
            v = gain * v;
            for (uint32_t s = 0; s < sections; s++) v = biquad_fast(kf + 4 * s, state + 4 * s, v);
            out[t * stride] = v;
            continue;
        }
        for (uint32_t s = 0; s < sections; s++) {
            const float *k = coefs + 5 * s;
            float *st = state + 4 * s;
            if (mode == ORC_MODE_FMA) {
                v = biquad_fma(kf + 5 * s, st, v);
            } else {
                int kind = kinds ? kinds[s] : ORC_KIND_BIQUAD;
                if (kind == ORC_KIND_PORT) {
                    float z = st[2];
                    z = v + k[0] * (z - v);
                    st[2] = z;
                    v = z;
                } else if (kind == ORC_KIND_LPF) {
                    float o = st[2];
                    o = o + (v - o) * k[0];
                    st[2] = o;
                    v = o;
                } else {
                    v = biquad_exact(k, st, v);
                }
            }
        }
        out[t * stride] = v;
    }
}

static int biquad_bank_mode(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs);

// state may be NULL (zero initial state, final state discarded).  sections <= 64.
ORC_API int orc_biquad_run(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs,
                           float *state, const float *in, float *out, uint64_t frames, int layout) {
    if (sections > 64) return -1;
    std::vector<float> tmp(4 * (size_t) sections);
    mode = biquad_bank_mode(mode, channels, sections, kinds, coefs);  // per BANK: one channel that cannot be normalised sends all to FMA
    for (uint32_t c = 0; c < channels; c++) {
        float *st = state ? state + (size_t) c * sections * 4 : tmp.data();
        if (!state) memset(st, 0, sizeof(float) * 4 * sections);
        const float *k = coefs + (size_t) c * sections * 5;
        if (layout == ORC_LAYOUT_FRAMES)
            biquad_channel(mode, sections, kinds, k, st, in + c, out + c, frames, channels);
        else
            biquad_channel(mode, sections, kinds, k, st, in + (size_t) c * frames, out + (size_t) c * frames, frames, 1);
    }
    return 0;
}

// The arithmetic a whole bank runs in: FAST only when EVERY channel can be normalised, else FMA.
static int biquad_bank_mode(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs) {
    if (mode != ORC_MODE_FAST) return mode;
    float k4[64 * 4], g;
    for (uint32_t c = 0; c < channels; c++)
        if (!orc_biquad_fast_coefs(sections, kinds, coefs + (size_t) c * sections * 5, k4, &g)) return ORC_MODE_FMA;
    return ORC_MODE_FAST;
}

// Node scaling of one channel: scale[n] = prod_{j >= n} b0_j in FAST (double products from the output side), 1 otherwise.
static void biquad_node_scales(int bank_mode, uint32_t sections, const uint8_t *kinds, const float *coefs, double *scale) {
    double g = 1.0;
    scale[sections] = 1.0;
    for (uint32_t s = sections; s-- > 0;) {
        float k[5];
        orc_biquad_fma_coefs(kinds ? kinds[s] : 0, coefs + 5 * s, k);
        if (bank_mode == ORC_MODE_FAST) g *= (double) k[0];
        scale[s] = bank_mode == ORC_MODE_FAST ? g : 1.0;
    }
}

// A bank keeps its FAST history in the node scaling of its coefficient table, so a coefficient (or mode) change between
// two blocks re-expresses the history in the new scaling: every history value of node n is multiplied ONCE by
// (float) (scale_new[n] / scale_old[n]) (double division, rounded to float; values whose ratio is 1 are left alone).
// state: [channels][sections][4] = x1 x2 y1 y2 (x of section s is node s, y is node s + 1), rescaled in place.
// The GPU bank does the same at the top of the first process call after the change (ol_dsp_b200/csrc/biquad_bank.cu).
ORC_API int orc_biquad_rescale_state(uint32_t channels, uint32_t sections, const uint8_t *kinds_old, const float *coefs_old, int mode_old,
                                     const uint8_t *kinds_new, const float *coefs_new, int mode_new, float *state) {
    if (sections > 64) return -1;
    const int bm_old = biquad_bank_mode(mode_old, channels, sections, kinds_old, coefs_old);
    const int bm_new = biquad_bank_mode(mode_new, channels, sections, kinds_new, coefs_new);
    double so[65], sn[65];
    for (uint32_t c = 0; c < channels; c++) {
        biquad_node_scales(bm_old, sections, kinds_old, coefs_old + (size_t) c * sections * 5, so);
        biquad_node_scales(bm_new, sections, kinds_new, coefs_new + (size_t) c * sections * 5, sn);
        float *st = state + (size_t) c * sections * 4;
        for (uint32_t s = 0; s < sections; s++) {
            if (so[s] != sn[s]) {
                const float r = (float) (sn[s] / so[s]);
                if (r != 1.0f) { st[4 * s + 0] = st[4 * s + 0] * r; st[4 * s + 1] = st[4 * s + 1] * r; }
            }
            if (so[s + 1] != sn[s + 1]) {
                const float r = (float) (sn[s + 1] / so[s + 1]);
                if (r != 1.0f) { st[4 * s + 2] = st[4 * s + 2] * r; st[4 * s + 3] = st[4 * s + 3] * r; }
            }
        }
    }
    return 0;
}

// Same cascade evaluated in double precision with double state (ground truth for the error budget).
ORC_API int orc_biquad_run_f64(uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs,
                               const float *in, double *out, uint64_t frames) {
    std::vector<double> st(4 * (size_t) sections);
    for (uint32_t c = 0; c < channels; c++) {
        std::fill(st.begin(), st.end(), 0.0);
        const float *kc = coefs + (size_t) c * sections * 5;
        for (uint64_t t = 0; t < frames; t++) {
            double v = in[t * channels + c];
            for (uint32_t s = 0; s < sections; s++) {
                const float *k = kc + 5 * s;
                double *z = &st[4 * s];
                int kind = kinds ? kinds[s] : 0;
                double y;
                if (kind == ORC_KIND_PORT) { y = v + (double) k[0] * (z[2] - v); z[2] = y; }
                else if (kind == ORC_KIND_LPF) { y = z[2] + (v - z[2]) * (double) k[0]; z[2] = y; }
                else {
                    y = (double) k[0] * v + (double) k[1] * z[0] + (double) k[2] * z[1] - (double) k[3] * z[2] - (double) k[4] * z[3];
                    z[1] = z[0]; z[0] = v; z[3] = z[2]; z[2] = y;
                }
                v = y;
            }
            out[t * channels + c] = v;
        }
    }
    return 0;
}

// RBJ-cookbook coefficient helpers (computed in double, rounded to float), normalised by a0.
// type: 0 low-pass, 1 high-pass, 2 peaking (gain_db), 3 one-pole low-pass expressed as a biquad.
ORC_API void orc_rbj(int type, double fs, double f0, double q, double gain_db, float *out5) {
    const double w0 = 2.0 * M_PI * f0 / fs, cw = cos(w0), sw = sin(w0), alpha = sw / (2.0 * q);
    double b0, b1, b2, a0, a1, a2;
    if (type == 0) { b0 = (1 - cw) / 2; b1 = 1 - cw; b2 = (1 - cw) / 2; a0 = 1 + alpha; a1 = -2 * cw; a2 = 1 - alpha; }
    else if (type == 1) { b0 = (1 + cw) / 2; b1 = -(1 + cw); b2 = (1 + cw) / 2; a0 = 1 + alpha; a1 = -2 * cw; a2 = 1 - alpha; }
    else if (type == 2) {
        const double A = pow(10.0, gain_db / 40.0);
        b0 = 1 + alpha * A; b1 = -2 * cw; b2 = 1 - alpha * A; a0 = 1 + alpha / A; a1 = -2 * cw; a2 = 1 - alpha / A;
    } else { const double p = exp(-w0); b0 = 1 - p; b1 = 0; b2 = 0; a0 = 1; a1 = -p; a2 = 0; }
    out5[0] = (float) (b0 / a0); out5[1] = (float) (b1 / a0); out5[2] = (float) (b2 / a0);
    out5[3] = (float) (a1 / a0); out5[4] = (float) (a2 / a0);
}

// ---------------------------------------------------------------------------------------------------
// DaisySP primitives, restated (PARITY UNPINNED -- see header).  Upstream file names are given for
// the day the source can be read: Source/Filters/svf.cpp, Source/Synthesis/oscillator.cpp,
// Source/Control/adsr.cpp, Source/Utility/delayline.h, Source/Utility/dsp.h.
static const float kPiF = 3.1415927410125732421875f;
static inline float fclampf(float x, float lo, float hi) { float a = x > lo ? x : lo; return a < hi ? a : hi; }
static inline float fminf_(float a, float b) { return a < b ? a : b; }

struct Svf {
    float sr, fc_max, res, pre_drive;  // setter-side
    float freq, damp, drive;           // coefficients read per sample
    float low, band;                   // recurrence state
    float o_low, o_high, o_band, o_notch, o_peak;
};

static void svf_recalc_damp(Svf *f) {
    f->damp = fminf_(2.0f * (1.0f - powf(f->res, 0.25f)), fminf_(2.0f, 2.0f / f->freq - f->freq * 0.5f));
}
static void svf_init(Svf *f, float sr) {
    memset(f, 0, sizeof(*f));
    f->sr = sr; f->res = 0.5f; f->drive = 0.5f; f->pre_drive = 0.5f; f->freq = 0.25f; f->damp = 0.f;
    f->fc_max = sr / 3.f;
}
static void svf_set_freq(Svf *f, float hz) {
    float fc = fclampf(hz, 1.0e-6, f->fc_max);
    f->freq = 2.0f * sinf(kPiF * fminf_(0.25f, fc / (f->sr * 2.0f)));
    svf_recalc_damp(f);
}
static void svf_set_res(Svf *f, float r) {
    f->res = fclampf(r, 0.f, 1.f);
    svf_recalc_damp(f);
    f->drive = f->pre_drive * f->res;
}
static void svf_set_drive(Svf *f, float d) {
    f->pre_drive = fclampf(d * 0.1f, 0.f, 1.f);
    f->drive = f->pre_drive * f->res;
}
// The fx banks' opt-in FUSED tier (olb_fx_set_mode(OLB_MODE_FMA); north_star: "fused or re-associated variants must state
// their tolerance"): the same formulas with the multiply-adds fused (one rounding instead of two), which takes the SVF's
// recurrence from 6 dependent operations per pass to 4 and its operation count from 11 to 7.  THIS FILE IS THE SPECIFICATION of
// the fused forms (fmaf = one IEEE rounding); the default everywhere is the separately rounded reference arithmetic.  The
// voice's filter is never fused.
static int g_fx_fast = 0;
ORC_API void orc_fx_set_mode(int fast) { g_fx_fast = fast ? 1 : 0; }
static inline void svf_half(Svf *f, float in, float *notch, float *high, int fast) {
    if (fast) {
        *notch = fmaf(-f->damp, f->band, in);
        f->low = fmaf(f->freq, f->band, f->low);
        *high = *notch - f->low;
        const float t = fmaf(f->freq, *high, f->band);
        const float c = (f->drive * f->band) * f->band;
        f->band = fmaf(-c, f->band, t);
        return;
    }
    *notch = in - f->damp * f->band;
    f->low = f->low + f->freq * f->band;
    *high = *notch - f->low;
    f->band = f->freq * *high + f->band - f->drive * f->band * f->band * f->band;
}
static void svf_tick(Svf *f, float in, int fast = 0) {
    float n, h;
    svf_half(f, in, &n, &h, fast);
    f->o_low = 0.5f * f->low; f->o_high = 0.5f * h; f->o_band = 0.5f * f->band;
    f->o_peak = 0.5f * (f->low - h); f->o_notch = 0.5f * n;
    svf_half(f, in, &n, &h, fast);
    if (fast) {
        f->o_low = fmaf(0.5f, f->low, f->o_low); f->o_high = fmaf(0.5f, h, f->o_high); f->o_band = fmaf(0.5f, f->band, f->o_band);
        f->o_peak = fmaf(0.5f, f->low - h, f->o_peak); f->o_notch = fmaf(0.5f, n, f->o_notch);
        return;
    }
    f->o_low += 0.5f * f->low; f->o_high += 0.5f * h; f->o_band += 0.5f * f->band;
    f->o_peak += 0.5f * (f->low - h); f->o_notch += 0.5f * n;
}

// Oscillator: phase in [0,1); sine (default) and PolyBLEP saw are the two waveforms the path uses.
struct Osc { float sr_recip, phase, inc, amp; int saw; };
static void osc_init(Osc *o, float sr) { o->sr_recip = 1.0f / sr; o->phase = 0.f; o->inc = 100.0f * o->sr_recip; o->amp = 0.5f; o->saw = 0; }
static inline void osc_set_freq(Osc *o, float f) { o->inc = f * o->sr_recip; }
static inline float polyblep(float dt, float t) {
    if (t < dt) { t /= dt; return t + t - t * t - 1.0f; }
    if (t > 1.0f - dt) { t = (t - 1.0f) / dt; return t * t + t + t + 1.0f; }
    return 0.0f;
}
static float osc_tick(Osc *o) {
    float out;
    if (o->saw) {
        float t = o->phase;
        out = (2.0f * t) - 1.0f;
        out -= polyblep(o->inc, t);
        out *= -1.0f;
    } else {
        out = sinf(o->phase * (2.0f * kPiF));
    }
    o->phase += o->inc;
    if (o->phase > 1.0f) o->phase -= 1.0f;
    return out * o->amp;
}

// Adsr: segments idle 0 / attack 1 / decay 2 / release 4; exponential approach with overshoot targets.
struct Env {
    int sr; float sus, x, a_shape, a_target, a_time, d_time, r_time, a_d0, d_d0, r_d0; int mode; int gate;
};
static void env_set_tc(Env *e, float t, float *time, float *coeff) {
    if (t != *time) {
        *time = t;
        if (t > 0.f) { const float target = logf(1. / M_E); *coeff = 1.f - expf(target / (t * e->sr)); }
        else *coeff = 1.f;
    }
}
static void env_set_attack(Env *e, float t, float shape) {
    if (t != e->a_time || shape != e->a_shape) {
        e->a_time = t; e->a_shape = shape;
        if (t > 0.f) {
            float target = 9.f * powf(shape, 10.f) + 0.3f * shape + 1.01f;
            e->a_target = target;
            float lt = logf(1.f - (1.f / target));
            e->a_d0 = 1.f - expf(lt / (t * e->sr));
        } else e->a_d0 = 1.f;
    }
}
static void env_set_sustain(Env *e, float s) { e->sus = (s <= 0.f) ? -0.01f : (s > 1.f) ? 1.f : s; }
static void env_init(Env *e, float sr, int block) {
    memset(e, 0, sizeof(*e));
    e->sr = (int) (sr / block);
    e->a_shape = -1.f; e->a_time = -1.f; e->d_time = -1.f; e->r_time = -1.f; e->sus = 0.7f;
    env_set_attack(e, 0.1f, 0.0f);
    env_set_tc(e, 0.1f, &e->d_time, &e->d_d0);
    env_set_tc(e, 0.1f, &e->r_time, &e->r_d0);
}
static void env_retrigger(Env *e, int hard) { e->mode = 1; if (hard) e->x = 0.f; }
static float env_tick(Env *e, int gate) {
    float out = 0.0f;
    if (gate && !e->gate) e->mode = 1;
    else if (!gate && e->gate) e->mode = 4;
    e->gate = gate;
    float d0 = e->a_d0;
    if (e->mode == 2) d0 = e->d_d0; else if (e->mode == 4) d0 = e->r_d0;
    float target = e->mode == 2 ? e->sus : -0.01f;
    if (e->mode == 1) {
        e->x += d0 * (e->a_target - e->x);
        out = e->x;
        if (out > 1.f) { e->x = out = 1.f; e->mode = 2; }
    } else if (e->mode == 2 || e->mode == 4) {
        e->x += d0 * (target - e->x);
        out = e->x;
        if (out < 0.0f) { e->x = out = 0.f; e->mode = 0; }
    }
    return out;
}

// DelayLine<float, 48000>: write pointer runs backwards; Read() precedes Write() in DelayFx.
enum { kMaxDelay = 48000 };  // modules/fxlib/Fx.h:23
struct Ring { float *line; size_t wp, delay; float frac; };
static void ring_init(Ring *r) { memset(r->line, 0, sizeof(float) * kMaxDelay); r->wp = 0; r->delay = 1; }
static void ring_set_delay(Ring *r, float d) {
    int32_t i = (int32_t) d;
    r->frac = d - (float) i;
    r->delay = (size_t) i < (size_t) kMaxDelay ? (size_t) i : (size_t) kMaxDelay - 1;
}
static inline float ring_read(const Ring *r) {
    float a = r->line[(r->wp + r->delay) % kMaxDelay];
    float b = r->line[(r->wp + r->delay + 1) % kMaxDelay];
    return g_fx_fast ? fmaf(b - a, r->frac, a) : a + (b - a) * r->frac;
}
static inline void ring_write(Ring *r, float s) { r->line[r->wp] = s; r->wp = (r->wp - 1 + kMaxDelay) % kMaxDelay; }

static inline float mtof(float m) { return powf(2.0f, (m - 69.0f) / 12.0f) * 440.0f; }

// ---------------------------------------------------------------------------------------------------
// (a2) FilterFx<N>  (modules/fxlib/Fx.h:65-165): ONE mono SVF whatever N is; writes channel 0 only.
struct FilterUnit { float cutoff, resonance, drive; int type; Svf svf; };
static void filter_update(FilterUnit *u) {  // Fx.h:110-114
    svf_set_freq(&u->svf, u->cutoff); svf_set_res(&u->svf, u->resonance); svf_set_drive(&u->svf, u->drive);
}
static void filter_defaults(FilterUnit *u) { u->cutoff = 20000.f; u->resonance = 0.f; u->drive = 0.f; u->type = 0; }
static void filter_init(FilterUnit *u, float sr) { svf_init(&u->svf, sr); filter_update(u); }  // Fx.h:83-86
static float filter_tick(FilterUnit *u, float in) {  // Fx.h:88-108; enum order LowPass, BandPass, HighPass, Notch, Peak
    svf_tick(&u->svf, in, g_fx_fast);
    switch (u->type) {
        case 2: return u->svf.o_high;
        case 1: return u->svf.o_band;
        case 4: return u->svf.o_peak;
        case 3: return u->svf.o_notch;
        default: return u->svf.o_low;
    }
}
static int filter_midi(FilterUnit *u, uint8_t cc, uint8_t val) {  // Fx.h:116-139
    float scaled = orc_scale(val, 0, 127, 0, 1, 1);
    switch (cc) {
        case kCcFilterCutoff: u->cutoff = orc_scale(val, 0, 127, 0, 20000, 1); break;
        case kCcFilterResonance: u->resonance = scaled; break;
        case kCcFilterDrive: u->drive = scaled; break;
        case kCcFilterType: u->type = (int) orc_scale(val, 0, 127, 0, 5, 1); break;
        default: return 0;
    }
    filter_update(u);
    return 1;
}
static int filter_hw(FilterUnit *u, uint8_t cc, float v) {  // Fx.h:141-164
    switch (cc) {
        case kCcFilterCutoff: u->cutoff = orc_scale(v, 0, 1, 0, 20000, 1.02f); break;
        case kCcFilterResonance: u->resonance = v; break;
        case kCcFilterDrive: u->drive = v; break;
        case kCcFilterType: u->type = (int) orc_scale(v, 0, 1, 0, 5, 1); break;
        default: return 0;
    }
    filter_update(u);
    return 1;
}

// (a3) DelayFx<N>  (Fx.h:170-268)
struct DelayUnit { int n; float time, feedback, balance; Ring ring[2]; FilterUnit wet; };
static void delay_update(DelayUnit *d) {  // Fx.h:210-217
    for (int i = 0; i < d->n; i++) ring_set_delay(&d->ring[i], orc_scale(d->time, 0, 1, 0, kMaxDelay, 1));
    filter_update(&d->wet);
}
static void delay_create(DelayUnit *d, int n) {
    d->n = n; d->time = 0.5f; d->feedback = 0.5f; d->balance = 0.33f;
    for (int i = 0; i < n; i++) { d->ring[i].line = (float *) malloc(sizeof(float) * kMaxDelay); d->ring[i].frac = 0.f; }
    filter_defaults(&d->wet);
}
static void delay_destroy(DelayUnit *d) { for (int i = 0; i < d->n; i++) free(d->ring[i].line); }
static void delay_init(DelayUnit *d, float sr) {  // Fx.h:183-192
    for (int i = 0; i < d->n; i++) ring_init(&d->ring[i]);
    filter_init(&d->wet, sr);
    filter_midi(&d->wet, kCcFilterCutoff, 64);
    filter_midi(&d->wet, kCcFilterResonance, 24);
    delay_update(d);
}
static void delay_tick(DelayUnit *d, const float *in, float *out) {  // Fx.h:194-208
    float buf[2];
    for (int i = 0; i < d->n; i++) {
        buf[i] = ring_read(&d->ring[i]);
        ring_write(&d->ring[i], g_fx_fast ? fmaf(d->feedback, buf[i], in[i]) : in[i] + (d->feedback * buf[i]));
    }
    buf[0] = filter_tick(&d->wet, buf[0]);  // the wet filter is mono: only channel 0 passes through it
    for (int i = 0; i < d->n; i++)
        out[i] = g_fx_fast ? fmaf(buf[i], d->balance, in[i] * (1 - d->balance)) : (buf[i] * d->balance) + (in[i] * (1 - d->balance));
}
static int delay_hw(DelayUnit *d, uint8_t cc, float v) {  // Fx.h:219-237
    switch (cc) {
        case kCcDelayTime: d->time = v; break;
        case kCcDelayFeedback: d->feedback = v; break;
        case kCcDelayBalance: d->balance = v; break;
        default: return 0;
    }
    delay_update(d);
    return 1;
}
static int delay_midi(DelayUnit *d, uint8_t cc, uint8_t val) {  // Fx.h:239-266
    float scaled = orc_scale(val, 0, 127, 0, 1, 1);
    switch (cc) {
        case kCcDelayTime: d->time = scaled; break;
        case kCcDelayFeedback: d->feedback = scaled; break;
        case kCcDelayBalance: d->balance = scaled; break;
        case kCcDelayCutoff: filter_midi(&d->wet, kCcFilterCutoff, val); break;
        case kCcDelayResonance: filter_midi(&d->wet, kCcFilterResonance, val); break;
        default: return 0;
    }
    delay_update(d);
    return 1;
}

// (a4) ReverbFx<N> around the stub ReverbSc  (Fx.h:272-393; Reverb.h:26-31,82-91).
// The setters reach only lpfreq_/feedback_ of the stub, which Process never reads; what is audible
// is `balance`.  The copy-paste bug at Fx.h:323-324/363-364 (INPUT_DIFFUSION_2 writes decay_diffusion)
// is mirrored in the bookkeeping below so a Dattorro back end can be driven from the same fields.
struct ReverbUnit {
    int n; float decay_time, cutoff, early_predelay, predelay, pre_cutoff, in_diff1, in_diff2, decay_diff, balance;
};
static void reverb_defaults(ReverbUnit *r, int n) {
    r->n = n; r->decay_time = 0.5f; r->cutoff = 12000.f; r->early_predelay = 0.1f; r->predelay = 0.2f;
    r->pre_cutoff = 0.5f; r->in_diff1 = 0.5f; r->in_diff2 = 0.5f; r->decay_diff = 0.5f; r->balance = 0.1f;
}
static void reverb_tick(ReverbUnit *r, const float *in, float *out) {
    float buf[2] = {0.f, 0.f};
    float in1 = in[0], in2 = r->n > 1 ? in[1] : in1;
    buf[0] = in1 * 0.8f;
    if (r->n > 1) buf[1] = in2 * 0.8f; else buf[0] = in2 * 0.8f;
    for (int i = 0; i < r->n; i++)
        out[i] = g_fx_fast ? fmaf(buf[i], r->balance, in[i] * (1 - r->balance)) : (buf[i] * r->balance) + (in[i] * (1 - r->balance));
}
static int reverb_set(ReverbUnit *r, uint8_t cc, float scaled, float cutoff_scaled) {
    switch (cc) {
        case kCcReverbDecayDiff: r->decay_diff = scaled; break;
        case kCcReverbInDiff1: r->in_diff1 = scaled; break;
        case kCcReverbInDiff2: r->decay_diff = scaled; break;  // sic (Fx.h:323-324)
        case kCcReverbCutoff: r->cutoff = cutoff_scaled; break;
        case kCcReverbBalance: r->balance = scaled; break;
        case kCcReverbPredelay: r->predelay = scaled; break;
        case kCcEarlyPredelay: r->early_predelay = scaled; break;
        case kCcReverbPrefilter: r->pre_cutoff = scaled; break;
        case kCcReverbTime: r->decay_time = scaled; break;
        default: return 0;
    }
    return 1;
}
static int reverb_midi(ReverbUnit *r, uint8_t cc, uint8_t val) {
    return reverb_set(r, cc, orc_scale(val, 0, 127, 0, 1, 1), orc_scale(val, 0, 127, 0, 20000, 1));
}
static int reverb_hw(ReverbUnit *r, uint8_t cc, float v) { return reverb_set(r, cc, v, orc_scale(v, 0, 1, 0, 20000, 1)); }

// (a5) FxRack<N>  (Fx.h:397-493): delay -> reverb -> filter -> * master_volume.
struct RackUnit { int n; DelayUnit delay; ReverbUnit reverb; FilterUnit filter; float master; };
static void rack_create(RackUnit *k, int n) {
    k->n = n; delay_create(&k->delay, n); reverb_defaults(&k->reverb, n); filter_defaults(&k->filter); k->master = 0.8f;
}
static void rack_init(RackUnit *k, float sr) {  // Fx.h:416-424
    delay_init(&k->delay, sr);
    filter_init(&k->filter, sr);
    delay_update(&k->delay);
    filter_update(&k->filter);
}
static void rack_tick(RackUnit *k, const float *in, float *out) {  // Fx.h:426-434
    float a[2], b[2], c[2] = {0.f, 0.f};
    delay_tick(&k->delay, in, a);
    reverb_tick(&k->reverb, a, b);
    c[0] = filter_tick(&k->filter, b[0]);  // channel 1 of buf_c is never written and stays 0
    for (int i = 0; i < k->n; i++) out[i] = c[i] * k->master;
}
static void rack_midi(RackUnit *k, uint8_t cc, uint8_t val) {  // Fx.h:443-468
    switch (cc) {
        case kCcFxFilterCutoff: filter_midi(&k->filter, kCcFilterCutoff, val); break;
        case kCcFxFilterResonance: filter_midi(&k->filter, kCcFilterResonance, val); break;
        case kCcFxFilterDrive: filter_midi(&k->filter, kCcFilterDrive, val); break;
        case kCcFxFilterType: filter_midi(&k->filter, kCcFilterType, val); break;
        default: break;
    }
    delay_midi(&k->delay, cc, val);
    reverb_midi(&k->reverb, cc, val);
    if (cc == kCcVolume) k->master = orc_scale(val, 0, 127, 0, 1, 1);
}
static void rack_hw(RackUnit *k, uint8_t cc, float v) {  // Fx.h:470-492
    switch (cc) {
        case kCcFxFilterCutoff: filter_hw(&k->filter, kCcFilterCutoff, v); break;
        case kCcFxFilterResonance: filter_hw(&k->filter, kCcFilterResonance, v); break;
        case kCcFxFilterDrive: filter_hw(&k->filter, kCcFilterDrive, v); break;
        case kCcFxFilterType: filter_hw(&k->filter, kCcFilterType, v); break;
        case kCcVolume: k->master = v; break;
        default: break;
    }
    delay_hw(&k->delay, cc, v);
    reverb_hw(&k->reverb, cc, v);
}

ORC_API int orc_filterfx_run(int n, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    if (n < 1 || n > 2) return -1;
    FilterUnit u; filter_defaults(&u); filter_init(&u, sr);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) {
            const olb_event &e = ev[next];
            if (e.kind == OLB_EV_MIDI_CC) filter_midi(&u, (uint8_t) e.a, (uint8_t) e.v);
            else if (e.kind == OLB_EV_HW_CC) filter_hw(&u, (uint8_t) e.a, e.v);
            else if (e.kind == OLB_EV_INIT) filter_init(&u, e.v);
        }
        out[t * n] = filter_tick(&u, in[t * n]);
    }
    return 0;
}

ORC_API int orc_delayfx_run(int n, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    if (n < 1 || n > 2) return -1;
    DelayUnit d; delay_create(&d, n); delay_init(&d, sr);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) {
            const olb_event &e = ev[next];
            if (e.kind == OLB_EV_MIDI_CC) delay_midi(&d, (uint8_t) e.a, (uint8_t) e.v);
            else if (e.kind == OLB_EV_HW_CC) delay_hw(&d, (uint8_t) e.a, e.v);
            else if (e.kind == OLB_EV_INIT) delay_init(&d, e.v);
        }
        delay_tick(&d, in + t * n, out + t * n);
    }
    delay_destroy(&d);
    return 0;
}

ORC_API int orc_reverbfx_run(int n, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    if (n < 1 || n > 2) return -1;
    (void) sr;
    ReverbUnit r; reverb_defaults(&r, n);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) {
            const olb_event &e = ev[next];
            if (e.kind == OLB_EV_MIDI_CC) reverb_midi(&r, (uint8_t) e.a, (uint8_t) e.v);
            else if (e.kind == OLB_EV_HW_CC) reverb_hw(&r, (uint8_t) e.a, e.v);
        }
        reverb_tick(&r, in + t * n, out + t * n);
    }
    return 0;
}

ORC_API int orc_fxrack_run(int n, float sr, const olb_event *ev, size_t n_ev, const float *in, float *out, size_t frames) {
    if (n < 1 || n > 2) return -1;
    RackUnit k; rack_create(&k, n); rack_init(&k, sr);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) {
            const olb_event &e = ev[next];
            if (e.kind == OLB_EV_MIDI_CC) rack_midi(&k, (uint8_t) e.a, (uint8_t) e.v);
            else if (e.kind == OLB_EV_HW_CC) rack_hw(&k, (uint8_t) e.a, e.v);
            else if (e.kind == OLB_EV_INIT) rack_init(&k, e.v);
        }
        rack_tick(&k, in + t * n, out + t * n);
    }
    delay_destroy(&k.delay);
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// (a7) Dattorro plate  (libs/dattorro-verb/verb.cpp:258-326; sizes and taps :173-222; state verb_structs.h:12-56)
// Restated as a table-driven network: 13 power-of-two rings in one slab, a 16-bit sample counter.
enum { DV_PRE, DV_IN0, DV_IN1, DV_IN2, DV_IN3, DV_DD1A, DV_PDA, DV_DD2A, DV_POSTA, DV_DD1B, DV_PDB, DV_DD2B, DV_POSTB, DV_RINGS };
static const uint16_t kDvDelay[DV_RINGS] = {4800, 142, 107, 379, 277, 672, 4453, 1800, 3720, 908, 4217, 2656, 3163};
// output taps (ring, delay), in the order the seven terms are summed; sign pattern + + - + - - +  (verb.cpp:302-326)
static const uint16_t kDvTapL[7][2] = {{DV_PDB, 266}, {DV_PDB, 2974}, {DV_DD2B, 1913}, {DV_POSTB, 1996}, {DV_PDA, 1990}, {DV_DD2A, 187}, {DV_POSTA, 1066}};
static const uint16_t kDvTapR[7][2] = {{DV_PDA, 353}, {DV_PDA, 3627}, {DV_DD2A, 1228}, {DV_POSTA, 2673}, {DV_PDB, 2111}, {DV_DD2B, 335}, {DV_POSTB, 121}};
static const float kDvSign[7] = {1.f, 1.f, -1.f, 1.f, -1.f, -1.f, 1.f};

struct Plate {
    float *slab; uint32_t base[DV_RINGS]; uint16_t mask[DV_RINGS]; uint16_t off[DV_RINGS];  // main-tap read offsets
    float pre_lp, damp_lp[2];
    float pre_filter, in_diff1, in_diff2, decay_diff1, damping, decay, decay_diff2;
    uint16_t t;
};
static uint32_t plate_ring_size(uint16_t delay) {  // next power of two strictly above the delay's bit length (verb.cpp:74-81)
    uint32_t bits = 0; for (uint16_t x = delay; x; x >>= 1) bits++;
    return 1u << bits;
}
ORC_API uint32_t orc_dattorro_slab_floats(void) { uint32_t n = 0; for (int i = 0; i < DV_RINGS; i++) n += plate_ring_size(kDvDelay[i]); return n; }
static double plate_clamp(float x, float lo, float hi) { if (x < lo) return lo; if (x > hi) return hi; return x; }  // returns double (verb.cpp:49-57)
static void plate_set_decay(Plate *p, float v) { p->decay = v; p->decay_diff2 = (float) plate_clamp(v + 0.15, 0.25, 0.50); }
static void plate_set_predelay(Plate *p, float v) { p->off[DV_PRE] = (uint16_t) (p->mask[DV_PRE] + 1 - (uint16_t) (v * 4800)); }
static void plate_create(Plate *p) {
    memset(p, 0, sizeof(*p));
    uint32_t n = 0;
    for (int i = 0; i < DV_RINGS; i++) {
        uint32_t sz = plate_ring_size(kDvDelay[i]);
        p->base[i] = n; p->mask[i] = (uint16_t) (sz - 1); p->off[i] = (uint16_t) (sz - kDvDelay[i]); n += sz;
    }
    p->slab = (float *) calloc(n, sizeof(float));
    plate_set_predelay(p, 0.1);
    p->pre_filter = 0.85; p->in_diff1 = 0.75; p->in_diff2 = 0.625;
    plate_set_decay(p, 0.75);
    p->decay_diff1 = 0.70; p->damping = 0.95;
}
static inline float *plate_at(Plate *p, int ring, uint32_t idx) { return p->slab + p->base[ring] + (idx & p->mask[ring]); }
static inline float plate_allpass(Plate *p, int ring, float gain, float in) {  // verb.cpp:123-128
    float delayed = *plate_at(p, ring, (uint32_t) p->t + p->off[ring]);
    in += delayed * -gain;
    *plate_at(p, ring, p->t) = in;
    return delayed + in * gain;
}
static inline float plate_delay(Plate *p, int ring, float in) {  // verb.cpp:107-110
    *plate_at(p, ring, p->t) = in;
    return *plate_at(p, ring, (uint32_t) p->t + p->off[ring]);
}
static inline float plate_lp(float *z, float f, float in) { *z += (in - *z) * f; return *z; }  // verb.cpp:131-134
static void plate_tick(Plate *p, float in) {
    if ((p->t & 0x07ff) == 0) {  // excursion: one sample every 2048, down for the first half of the 16-bit cycle
        if (p->t < (1 << 15)) { p->off[DV_DD1A]--; p->off[DV_DD1B]--; } else { p->off[DV_DD1A]++; p->off[DV_DD1B]++; }
    }
    float x = plate_delay(p, DV_PRE, in);
    x = plate_lp(&p->pre_lp, p->pre_filter, x);
    x = plate_allpass(p, DV_IN0, p->in_diff1, x);
    x = plate_allpass(p, DV_IN1, p->in_diff1, x);
    x = plate_allpass(p, DV_IN2, p->in_diff2, x);
    x = plate_allpass(p, DV_IN3, p->in_diff2, x);
    static const int dd1[2] = {DV_DD1A, DV_DD1B}, pd[2] = {DV_PDA, DV_PDB}, dd2[2] = {DV_DD2A, DV_DD2B}, post[2] = {DV_POSTA, DV_POSTB};
    for (int h = 0; h < 2; h++) {
        float y = x + *plate_at(p, post[1 - h], (uint32_t) p->t + p->off[post[1 - h]]) * p->decay;
        y = plate_allpass(p, dd1[h], -p->decay_diff1, y);
        y = plate_delay(p, pd[h], y);
        y = plate_lp(&p->damp_lp[h], p->damping, y);
        y *= p->decay;
        y = plate_allpass(p, dd2[h], p->decay_diff2, y);
        *plate_at(p, post[h], p->t) = y;
    }
    p->t++;
}
static float plate_out(Plate *p, const uint16_t taps[7][2]) {  // taps are read after t++ (verb.cpp:302-326)
    float a = 0.f;
    for (int i = 0; i < 7; i++) {
        int ring = taps[i][0];
        float v = *plate_at(p, ring, (uint32_t) p->t + (uint32_t) (p->mask[ring] + 1 - taps[i][1]));
        if (i == 0) a = v; else if (kDvSign[i] > 0) a += v; else a -= v;
    }
    return a;
}
static void plate_params(Plate *p, const float *q) {
    if (!q) return;
    if (!isnan(q[0])) plate_set_predelay(p, q[0]);
    if (!isnan(q[1])) p->pre_filter = q[1];
    if (!isnan(q[2])) p->in_diff1 = q[2];
    if (!isnan(q[3])) p->in_diff2 = q[3];
    if (!isnan(q[4])) plate_set_decay(p, q[4]);
    if (!isnan(q[5])) p->decay_diff1 = q[5];
    if (!isnan(q[6])) p->damping = q[6];
}
ORC_API int orc_dattorro_run(const float *params7, const float *in, float *out_l, float *out_r, size_t frames) {
    Plate p; plate_create(&p); plate_params(&p, params7);
    for (size_t t = 0; t < frames; t++) {
        plate_tick(&p, in[t]);
        out_l[t] = plate_out(&p, kDvTapL);
        out_r[t] = plate_out(&p, kDvTapR);
    }
    free(p.slab);
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// (a6) SynthVoice  (modules/synthlib/SynthVoice.h:16-316) and (a8) Polyvoice (Polyvoice.h:12-86)
struct VoiceUnit {
    float sr; Osc osc; Svf svf; Env fenv, aenv;
    float port_sr, port_coeff, port_z;
    float freq, cutoff, resonance, drive;
    float f_attack, f_attack_shape, f_decay, f_sustain, f_release, f_amount;
    float a_attack, a_attack_shape, a_decay, a_sustain, a_release, a_amount;
    float port_htime, osc1_mix;
    uint8_t playing; int gate;
};
static void voice_defaults(VoiceUnit *v) {  // SynthVoice.h:269-315
    memset(v, 0, sizeof(*v));
    v->port_sr = 48000.f;
    v->f_attack_shape = 1; v->f_decay = 0.2f; v->f_amount = 1;
    v->a_attack = 0.01f; v->a_attack_shape = 1; v->a_sustain = 1; v->a_release = 0.01f; v->a_amount = 0.8f;
    v->osc1_mix = 0.8f;
}
static void voice_port_htime(VoiceUnit *v, float h) { v->port_coeff = expf(-1.0f / (h * v->port_sr)); }  // Portamento.h:34-37
static void voice_update(VoiceUnit *v) {  // SynthVoice.h:79-98 (SetFreq is NOT called here)
    svf_set_res(&v->svf, v->resonance);
    svf_set_drive(&v->svf, v->drive);
    env_set_attack(&v->fenv, v->f_attack, v->f_attack_shape);
    env_set_tc(&v->fenv, v->f_decay, &v->fenv.d_time, &v->fenv.d_d0);
    env_set_sustain(&v->fenv, v->f_sustain);
    env_set_tc(&v->fenv, v->f_release, &v->fenv.r_time, &v->fenv.r_d0);
    env_set_attack(&v->aenv, v->a_attack, v->a_attack_shape);
    env_set_tc(&v->aenv, v->a_decay, &v->aenv.d_time, &v->aenv.d_d0);
    env_set_sustain(&v->aenv, v->a_sustain);
    env_set_tc(&v->aenv, v->a_release, &v->aenv.r_time, &v->aenv.r_d0);
    voice_port_htime(v, v->port_htime);
}
static void voice_init(VoiceUnit *v, float sr) {  // SynthVoice.h:31-39: Init does not call Update()
    v->sr = sr;
    osc_init(&v->osc, sr); v->osc.saw = 1;
    svf_init(&v->svf, sr);
    env_init(&v->fenv, sr, 1);
    env_init(&v->aenv, sr, 1);
    v->port_sr = sr; voice_port_htime(v, v->port_htime); v->port_z = 0.f;
}
static float voice_tick(VoiceUnit *v) {  // SynthVoice.h:41-53
    float amp = env_tick(&v->aenv, v->gate);
    amp *= v->a_amount;
    v->port_z = v->freq + v->port_coeff * (v->port_z - v->freq);
    osc_set_freq(&v->osc, v->port_z);
    float frame = osc_tick(&v->osc);
    float ff = v->cutoff + ((env_tick(&v->fenv, v->gate) * 20000) * v->f_amount);
    svf_set_freq(&v->svf, ff);
    svf_tick(&v->svf, frame);
    frame = v->svf.o_low;
    return frame * amp;
}
static void voice_config(VoiceUnit *v, const float *c) {  // Voice.h:14-31 field order; SynthVoice.h:55-77
    v->cutoff = c[0]; v->resonance = c[1]; v->drive = c[2]; v->f_amount = c[3];
    v->f_attack = c[4]; v->f_attack_shape = c[5]; v->f_decay = c[6]; v->f_sustain = c[7]; v->f_release = c[8];
    v->a_amount = c[9]; v->a_attack = c[10]; v->a_attack_shape = c[11]; v->a_decay = c[12]; v->a_sustain = c[13]; v->a_release = c[14];
    v->port_htime = c[15];
    voice_update(v);
}
static void voice_hw(VoiceUnit *v, uint8_t cc, float x) {  // SynthVoice.h:101-156
    switch (cc) {
        case kCcVolume: v->a_amount = x; break;
        case kCcPortamento: v->port_htime = orc_scale(x, 0, 1, 0, 1, 4); break;
        case kCcFilterCutoff: v->cutoff = orc_scale(x, 0, 1, 0, 20000, 2.5f); break;
        case kCcFilterResonance: v->resonance = x; break;
        case kCcFilterDrive: v->drive = x; break;
        case kCcEnvFiltAmt: v->f_amount = x; break;
        case kCcEnvFiltA: v->f_attack = x; break;
        case kCcEnvFiltD: v->f_decay = orc_scale(x, 0, 1, 0, 1, 3); break;
        case kCcEnvFiltS: v->f_sustain = x; break;
        case kCcEnvFiltR: v->f_release = x; break;
        case kCcEnvAmpA: v->a_attack = x; break;
        case kCcEnvAmpD: v->a_decay = x; break;
        case kCcEnvAmpS: v->a_sustain = x; break;
        case kCcEnvAmpR: v->a_release = x; break;
        case kCcOsc1Volume: v->osc1_mix = x; break;
        default: return;
    }
    voice_update(v);
}
static void voice_midi(VoiceUnit *v, uint8_t cc, uint8_t val) {  // SynthVoice.h:158-229
    float s = orc_scale(val, 0, 127, 0, 1, 1);
    switch (cc) {
        case kCcVolume: v->a_amount = s; break;
        case kCcPortamento: v->port_htime = orc_scale(val, 0, 127, 0, 1, 4); break;
        case kCcFilterCutoff: v->cutoff = orc_scale(val, 0, 127, 0, 20000, 2.5f); break;
        case kCcFilterResonance: v->resonance = s; break;
        case kCcFilterDrive: v->drive = s; break;
        case kCcEnvFiltAmt: v->f_amount = s; break;
        case kCcEnvFiltA: v->f_attack = s; break;
        case kCcEnvFiltD: v->f_decay = orc_scale(val, 0, 127, 0, 1, 3); break;
        case kCcEnvFiltS: v->f_sustain = s; break;
        case kCcEnvFiltR: v->f_release = s; break;
        case kCcEnvAmpA: v->a_attack = s; break;
        case kCcEnvAmpD: v->a_decay = s; break;
        case kCcEnvAmpS: v->a_sustain = s; break;
        case kCcEnvAmpR: v->a_release = s; break;
        case kCcOsc1Volume: v->osc1_mix = s; break;
        default: return;
    }
    voice_update(v);
}
static void voice_note_on(VoiceUnit *v, uint8_t note) {  // SynthVoice.h:245-251
    v->gate = 1; v->playing = note; v->freq = mtof(note);
    env_retrigger(&v->aenv, 1); env_retrigger(&v->fenv, 1);
}
static void voice_note_off(VoiceUnit *v) { v->gate = 0; v->playing = 0; }  // SynthVoice.h:253-256
static void voice_event(VoiceUnit *v, const olb_event &e) {
    switch (e.kind) {
        case OLB_EV_MIDI_CC: voice_midi(v, (uint8_t) e.a, (uint8_t) e.v); break;
        case OLB_EV_HW_CC: voice_hw(v, (uint8_t) e.a, e.v); break;
        case OLB_EV_INIT: voice_init(v, e.v); break;
        case OLB_EV_NOTE_ON: voice_note_on(v, (uint8_t) e.a); break;
        case OLB_EV_NOTE_OFF: voice_note_off(v); break;
        case OLB_EV_SET_FREQ: v->freq = e.v; break;
        case OLB_EV_GATE_ON: v->gate = 1; break;
        case OLB_EV_GATE_OFF: v->gate = 0; break;
        default: break;
    }
}

// trace (may be NULL): 4 floats per frame = oscillator output, amp envelope, filter envelope, phase after the tick.
ORC_API int orc_voice_run(float sr, const float *config16, const olb_event *ev, size_t n_ev, float *out, float *trace, size_t frames) {
    VoiceUnit v; voice_defaults(&v); voice_init(&v, sr);
    if (config16) voice_config(&v, config16);
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) voice_event(&v, ev[next]);
        if (trace) {
            // replay the tick with taps: identical arithmetic to voice_tick
            float amp_env = env_tick(&v.aenv, v.gate);
            float amp = amp_env * v.a_amount;
            v.port_z = v.freq + v.port_coeff * (v.port_z - v.freq);
            osc_set_freq(&v.osc, v.port_z);
            float frame = osc_tick(&v.osc);
            float fe = env_tick(&v.fenv, v.gate);
            float ff = v.cutoff + ((fe * 20000) * v.f_amount);
            svf_set_freq(&v.svf, ff);
            svf_tick(&v.svf, frame);
            trace[4 * t + 0] = frame; trace[4 * t + 1] = amp_env; trace[4 * t + 2] = fe; trace[4 * t + 3] = v.osc.phase;
            out[t] = v.svf.o_low * amp;
        } else {
            out[t] = voice_tick(&v);
        }
    }
    return 0;
}

ORC_API int orc_polyvoice_run(int n_voices, float sr, const float *config16, const olb_event *ev, size_t n_ev, float *out, size_t frames) {
    std::vector<VoiceUnit> vs((size_t) n_voices);
    for (auto &v : vs) { voice_defaults(&v); voice_init(&v, sr); if (config16) voice_config(&v, config16); }
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        for (; next < n_ev && ev[next].frame <= t; next++) {
            const olb_event &e = ev[next];
            if (e.voice >= 0 && e.voice < n_voices) { voice_event(&vs[e.voice], e); continue; }
            if (e.kind == OLB_EV_NOTE_ON) {  // first voice that is not playing (Polyvoice.h:35-42)
                for (auto &v : vs) if (!v.playing) { voice_note_on(&v, (uint8_t) e.a); break; }
            } else if (e.kind == OLB_EV_NOTE_OFF) {  // first voice playing that note (Polyvoice.h:44-51)
                for (auto &v : vs) if (v.playing == (uint8_t) e.a) { voice_note_off(&v); break; }
            } else if (e.kind == OLB_EV_SET_FREQ) {  // Polyvoice::SetFrequency is a nop (Polyvoice.h:76)
            } else {
                for (auto &v : vs) voice_event(&v, e);
            }
        }
        float bus = 0.f;
        for (auto &v : vs) bus += voice_tick(&v);  // index order, into a zeroed frame (Polyvoice.h:28-33)
        out[t] = bus;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// CPU baseline driver (bench.py cpu_baseline / --impl reference): one channel per std::thread at a
// time, threads = n_threads, each thread pulls channels from a shared counter and runs the scalar
// cascade (not vectorised across channels) over `frames` frames of replayable Philox noise held in
// planar host buffers.  Returns the seconds of wall time the cascades took; noise generation and
// the checksum pass are outside the timed region.
static double biquad_bench_steps(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs, uint64_t frames,
                                 uint32_t seed, int n_threads, int warmup, int steps, double *checksum) {
    // inputs first, OUTSIDE the timed region (the GPU arm's noise is resident before its clock starts too)
    std::vector<float> x((size_t) channels * frames), y((size_t) channels * frames);
    {
        std::atomic<uint32_t> nx(0);
        std::vector<std::thread> gen;
        for (int w = 0; w < n_threads; w++)
            gen.emplace_back([&]() {
                for (uint32_t c; (c = nx.fetch_add(1)) < channels;) orc_noise_fill(seed, c, 1, 0, frames, 1.0f, &x[(size_t) c * frames], 1, 0);
            });
        for (auto &th : gen) th.join();
    }
    const int bank_mode = biquad_bank_mode(mode, channels, sections, kinds, coefs);
    double total = 0.0;
    for (int it = 0; it < warmup + steps; it++) {
        std::atomic<uint32_t> next(0);
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> pool;
        for (int w = 0; w < n_threads; w++) {
            pool.emplace_back([&]() {
                std::vector<float> st(4 * (size_t) sections);
                for (uint32_t c; (c = next.fetch_add(1)) < channels;) {
                    std::fill(st.begin(), st.end(), 0.f);
                    biquad_channel(bank_mode, sections, kinds, coefs + (size_t) c * sections * 5, st.data(), &x[(size_t) c * frames],
                                   &y[(size_t) c * frames], frames, 1);
                }
            });
        }
        for (auto &th : pool) th.join();
        if (it >= warmup) total += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    if (checksum) { double s = 0; for (float v : y) s += v; *checksum = s; }
    return total;
}

ORC_API double orc_biquad_bench(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs,
                                uint64_t frames, uint32_t seed, int n_threads, double *checksum) {
    return biquad_bench_steps(mode, channels, sections, kinds, coefs, frames, seed, n_threads, 0, 1, checksum);
}

// `warmup` untimed + `steps` timed passes over ONE set of inputs; returns the seconds of the timed passes together.
ORC_API double orc_biquad_bench_steps(int mode, uint32_t channels, uint32_t sections, const uint8_t *kinds, const float *coefs,
                                      uint64_t frames, uint32_t seed, int n_threads, int warmup, int steps, double *checksum) {
    return biquad_bench_steps(mode, channels, sections, kinds, coefs, frames, seed, n_threads, warmup, steps, checksum);
}

// ================================================================================================
// (a9) Chorus.  The reference holds NO chorus code: the only statement of it is the Max/RNBO box graph
// modules/rnbo/patcher/mono-chorus.rnbopat (boxes obj-30 rate, obj-5 scale 0 1 0.01 0.5, obj-32 depth @min 0.08,
// obj-7 scale 0 1 1 12 1, obj-64 mstosamps, obj-6 cycle~, obj-38 *~, obj-36 +~ "add the depth to keep the delay
// value > 0", obj-24 delay~ @maxsize samplerate*2, obj-33 cutoff, obj-20 scale 0 1 300 15000 1, obj-34 q, obj-21 lores~,
// obj-13 mix, obj-25/29/28/27 wet*mix + dry*(1-mix)).  The Max objects' internals are not in the tree, so THIS FILE IS THE
// SPECIFICATION of the arithmetic (PARITY UNPINNED):
//   cycle~   512-point cosine table (+ guard point), linear interpolation, phase offset added to the running phase
//   delay~   write, then read `dt` samples back with linear interpolation between the two neighbours
//   lores~   y = scale*x - a1*y1 - a2*y2 with resterm = exp(q*0.125)*0.882497, fqterm = cos(2*pi*fc/sr),
//            a1 = -2*resterm*fqterm, a2 = resterm^2, scale = 1 + a1 + a2   (the published lores~ recurrence)
// Setter-time maths in double, rounded to float once; per-frame maths in float, one rounding per operation.
struct OrcChorusParams {
    float inc, d, offset, scale, a1, a2, mix;
};
ORC_API void orc_chorus_params(float sr, float rate, float depth, float cutoff, float q, float mix, float phase, float *out7) {
    const double rate_hz = 0.01 + (0.5 - 0.01) * (double) rate;                        // scale 0 1 0.01 0.5
    const double dep = depth < 0.08f ? 0.08 : (double) depth;                          // param depth @min 0.08
    const double depth_ms = 1.0 + (12.0 - 1.0) * dep;                                  // scale 0 1 1 12 1
    const double fc = 300.0 + (15000.0 - 300.0) * (double) cutoff;                     // scale 0 1 300 15000 1
    const double resterm = exp((double) q * 0.125) * 0.882497;
    const double fqterm = cos(fc * 2.0 * 3.14159265358979323846 / (double) sr);
    const double a1 = -2.0 * resterm * fqterm, a2 = resterm * resterm;
    out7[0] = (float) (rate_hz / (double) sr);
    out7[1] = (float) (depth_ms * (double) sr / 1000.0);                               // mstosamps
    out7[2] = (float) ((double) phase - floor((double) phase));                        // cycle~ phase offset, wrapped to [0,1)
    out7[3] = (float) (1.0 + a1 + a2);
    out7[4] = (float) a1;
    out7[5] = (float) a2;
    out7[6] = mix;
}
ORC_API uint32_t orc_chorus_ring_len(float sr) { return (uint32_t) ceil(2.0 * 12.0 * (double) sr / 1000.0) + 4u; }
ORC_API void orc_chorus_table(float *tab513) {
    for (int i = 0; i <= 512; i++) tab513[i] = (float) cos(2.0 * 3.14159265358979323846 * (double) i / 512.0);
}
// The gen~ pitch shifter in front of delay~ (modules/rnbo/patcher/pitchshift.gendsp; mono-chorus.rnbopat obj-8 `param pitch
// @min 0 @max 3` -> obj-3 `gen~ @file pitchshift` inlet 2 -> obj-24 delay~).  Box graph: obj-2 phasor(shift) -> (+0, +0.5) ->
// `% 1` -> two tap times `* mstosamps(param window 10)` into obj-17 `delay 44100 2`, each tap weighted by
// cos((phase - 0.5) * pi), summed.  gen~'s operator internals are not in the tree; restated here (PARITY UNPINNED) as
//   phasor   phase = wrap(phase + shift / sr); the NEW phase is the output (genlib's Phasor increments, then returns)
//   % 1      x - 1 when x >= 1 (both operands are in [0, 1.5))
//   cos      even polynomial of degree 12 in float (|argument| <= pi/2, error ~1e-7), one rounding per operation
//   delay    read before write, tap time clamped to [1, 44100], linear interpolation a + (b - a) * frac between x[n-k], x[n-k-1]
// With pitch = 0 (the default) the phase rests at 0: tap 1 = x[n-1] * cos(-pi/2) ~ -4e-8 x[n-1], tap 2 = x delayed by half
// the window (240 samples at 48 kHz) at weight 1.
// pitch2 = {phase increment shift / sr, window in samples}; state: phase; hist: the last `hist_len` inputs, hist[hist_len-1] newest.
ORC_API uint32_t orc_pitch_hist_len(float sr) {
    const double win = 10.0 * (double) sr / 1000.0;
    const double m = win < 44100.0 ? win : 44100.0;
    return (uint32_t) m + 2u;
}
ORC_API void orc_pitch_params(float sr, float pitch, float *out2) {
    const double p = pitch < 0.f ? 0.0 : (pitch > 3.f ? 3.0 : (double) pitch);  // param pitch @min 0 @max 3
    out2[0] = (float) (p / (double) sr);
    out2[1] = (float) (10.0 * (double) sr / 1000.0);                           // mstosamps(param window 10)
}
static inline float orc_cos_poly(float t) {  // cos t on [-pi/2, pi/2]
    const float u = t * t;
    float c = 2.0876756987868099e-9f;      // 1/12!
    c = c * u + -2.7557319223985888e-7f;   // -1/10!
    c = c * u + 2.4801587301587302e-5f;    // 1/8!
    c = c * u + -1.3888888888888889e-3f;   // -1/6!
    c = c * u + 4.1666666666666664e-2f;    // 1/4!
    c = c * u + -0.5f;
    c = c * u + 1.0f;
    return c;
}
static inline float orc_pitch_tap(const float *hist, uint32_t hist_len, float d) {
    d = d < 1.0f ? 1.0f : d;
    d = d > 44100.0f ? 44100.0f : d;
    const uint32_t k = (uint32_t) d;
    const float f = d - (float) k;
    const float a = hist[hist_len - k], b = hist[hist_len - k - 1];  // x[n-k], x[n-k-1]
    return a + (b - a) * f;
}
ORC_API void orc_pitch_run(const float *pitch2, uint32_t hist_len, float *phase_state, float *hist, const float *x, float *y, uint64_t frames,
                           uint64_t stride) {
    const float inc = pitch2[0], win = pitch2[1];
    float ph = *phase_state;
    for (uint64_t n = 0; n < frames; n++) {
        ph = ph + inc;
        if (ph >= 1.0f) ph = ph - 1.0f;
        float p2 = ph + 0.5f;
        if (p2 >= 1.0f) p2 = p2 - 1.0f;
        const float w1 = orc_cos_poly((ph - 0.5f) * 3.14159274f), w2 = orc_cos_poly((p2 - 0.5f) * 3.14159274f);
        const float t1 = orc_pitch_tap(hist, hist_len, ph * win), t2 = orc_pitch_tap(hist, hist_len, p2 * win);
        const float xin = x[n * stride];
        y[n * stride] = t1 * w1 + t2 * w2;
        memmove(hist, hist + 1, (hist_len - 1) * sizeof(float));
        hist[hist_len - 1] = xin;
    }
    *phase_state = ph;
}

// One instance over `frames` frames.  state: [0] phase, [1] y1, [2] y2; ring: ring_len floats; *w: write index.
// xd: what delay~ is fed (the pitch shifter's output; NULL = x itself, the patch without its gen~ stage); x stays the dry signal.
ORC_API void orc_chorus_run(const float *par7, uint32_t ring_len, float *state3, float *ring, uint32_t *w, const float *x, const float *xd,
                            float *y, uint64_t frames, uint64_t stride) {
    static float tab[513];
    static bool have = false;
    if (!have) { orc_chorus_table(tab); have = true; }
    const float inc = par7[0], d = par7[1], offset = par7[2], scale = par7[3], a1 = par7[4], a2 = par7[5], mix = par7[6];
    float phase = state3[0], y1 = state3[1], y2 = state3[2];
    uint32_t wp = *w;
    const float dmax = (float) (ring_len - 2);
    for (uint64_t n = 0; n < frames; n++) {
        const float xin = x[n * stride];
        float p = phase + offset;
        if (p >= 1.0f) p = p - 1.0f;
        const float idx = p * 512.0f;
        int i0 = (int) idx;
        if (i0 > 511) i0 = 511;
        const float f = idx - (float) i0;
        const float c0 = tab[i0], c1 = tab[i0 + 1];
        const float lfo = c0 + (c1 - c0) * f;
        float dt = lfo * d + d;
        dt = dt < 0.0f ? 0.0f : dt;
        dt = dt > dmax ? dmax : dt;
        ring[wp] = xd ? xd[n * stride] : xin;
        const uint32_t k = (uint32_t) dt;
        const float fr = dt - (float) k;
        const uint32_t ia = wp >= k ? wp - k : wp + ring_len - k;
        const uint32_t ib = ia >= 1 ? ia - 1 : ring_len - 1;
        const float a = ring[ia], b = ring[ib];
        const float wet0 = a + (b - a) * fr;
        float t = scale * wet0;
        t = t - a1 * y1;
        t = t - a2 * y2;
        y2 = y1;
        y1 = t;
        y[n * stride] = t * mix + xin * (1.0f - mix);
        phase = phase + inc;
        if (phase >= 1.0f) phase = phase - 1.0f;
        wp = wp + 1 == ring_len ? 0 : wp + 1;
    }
    state3[0] = phase; state3[1] = y1; state3[2] = y2;
    *w = wp;
}

// ------------------------------------------------------------------------------------------------
// Sample playback: ol::synth::Sample behind SampleSoundSource (modules/synthlib/Sample.cpp:9-61, SampleSoundSource.h:16-30)
// over an in-memory sample, one player.  Restated as one loop over frames:
//   playing:  read frame `cur` if it exists (SampleDataSource::Read returns 0 past the end) and advance; in Loop mode seek
//             to loop_start when nothing was read, or when loop_end is set and the new position is past it (Sample.cpp:10-22)
//   a frame that is not played is left as the caller filled it
// state5 = {playing, mode, loop_start, loop_end, cur} carries over between calls.  Events: oracle_abi.h.
ORC_API void orc_sampler_run(int channels, const float *data, uint64_t sample_frames, uint64_t *state5, const olb_event *ev, size_t n_ev,
                             float *out, size_t frames) {
    uint64_t playing = state5[0], mode = state5[1], loop_start = state5[2], loop_end = state5[3], cur = state5[4];
    size_t next = 0;
    for (size_t t = 0; t < frames; t++) {
        while (next < n_ev && ev[next].frame <= t) {
            const olb_event &e = ev[next++];
            switch (e.kind) {
                case OLB_EV_GATE_ON: cur = 0; playing = 1; break;      // Seek(0); Play()
                case OLB_EV_GATE_OFF: playing = 0; break;              // Pause()
                case OLB_EV_SAMPLE_SEEK: cur = e.a; break;
                case OLB_EV_SAMPLE_LOOP_START: loop_start = e.a; break;
                case OLB_EV_SAMPLE_LOOP_END: loop_end = e.a; break;
                case OLB_EV_SAMPLE_MODE: mode = e.a ? 1 : 0; break;
                case OLB_EV_SAMPLE_TOGGLE: playing = !playing; break;
                case OLB_EV_SAMPLE_PLAY: playing = 1; break;
                case OLB_EV_SAMPLE_PAUSE: playing = 0; break;
                default: break;
            }
        }
        if (!playing) continue;
        uint64_t read = 0;
        if (cur < sample_frames) {
            for (int c = 0; c < channels; c++) out[t * channels + c] = data[cur * channels + c];
            read = 1;
        }
        cur += read;
        if (mode == 1 && (!read || (loop_end && cur > loop_end))) cur = loop_start;
    }
    state5[0] = playing; state5[1] = mode; state5[2] = loop_start; state5[3] = loop_end; state5[4] = cur;
}
