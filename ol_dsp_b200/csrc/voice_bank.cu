// voice_bank.cu -- (a6) bank of ol::synth::SynthVoice and (a8) the Polyvoice mixdown.
//
// Reference, per output sample (modules/synthlib/SynthVoice.h:41-53):
//   amp = amp_envelope.Process(gate) * amp_env_amount               DaisyAdsr -> daisysp::Adsr      Adsr.h:48-50
//   osc.SetFreq(portamento.Process(freq_)); frame = osc.Process()   daisysp::Port Portamento.h:29-32; PolyBLEP saw
//   fc = filter_cutoff + filter_envelope.Process(gate) * 20000 * filter_env_amount
//   filter.SetFreq(fc)  (EVERY sample: sinf)                        SvfFilter -> daisysp::Svf       Filter.h:65-108
//   filter.Process(frame); frame = Low; frame *= amp
// and the bus (modules/synthlib/Polyvoice.h:28-33): *frame_out += voice frame, in voice order.
//
// One voice per thread, all state in registers for the whole block of frames.  Bit-exact pieces (every operation
// rounded separately, setter maths on the host with the host's libm): envelope stage machine and level recurrence,
// portamento one-pole, oscillator phase / wrap / PolyBLEP, SVF recurrence.  The one device transcendental is
// Svf::SetFreq's sinf per sample: CUDA's sinf differs from glibc's by <= 2 ulp, so voice OUTPUT is a tolerance tier
// (<= 1e-5 of full scale, tests/test_voice_gpu.py) while phase and envelope traces are compared bit for bit.
//
// Mixdown: each warp parks 32 frames of its 32 voices in shared memory and adds them in voice order, one partial per
// WARP per frame written coalesced 32 frames at a time; k_bus_reduce then sums the warp partials in warp order:
// deterministic run to run; differs from the reference's one sequential sum over all voices only by where the
// partial sums are cut (float re-association, tolerance tier).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "host_params.h"

namespace olb {

enum { VP_FREQ, VP_GATE, VP_RETRIG, VP_PORT_COEFF, VP_SR_RECIP, VP_CUTOFF, VP_F_AMOUNT, VP_A_AMOUNT, VP_SR2, VP_FC_MAX, VP_DAMP_A, VP_DRIVE,
       VP_A_D0, VP_A_TARGET, VP_A_DD0, VP_A_RD0, VP_A_SUS, VP_F_D0, VP_F_TARGET, VP_F_DD0, VP_F_RD0, VP_F_SUS, VP_COUNT };
enum { VS_PHASE, VS_PORT_Z, VS_LOW, VS_BAND, VS_A_X, VS_A_MODE, VS_A_GATE, VS_F_X, VS_F_MODE, VS_F_GATE, VS_COUNT };

struct EnvCoef {
    float a_d0, a_target, d_d0, r_d0, sus;
};

// daisysp::Adsr::Process (segments idle 0 / attack 1 / decay 2 / release 4), branch-free: the segment picks (d0, target)
// with selects, the one-pole step x += d0 * (target - x) is the same expression in every active segment, and the two
// segment exits (attack overshoots 1 -> decay at 1; decay / release undershoot 0 -> idle at 0) are selects as well.
// 32 voices of a warp are rarely in the same segment, so a branchy version runs every path anyway.
__device__ __forceinline__ float env_tick(float &x, int &mode, int &prev_gate, int gate, const EnvCoef &k) {
    mode = (gate && !prev_gate) ? 1 : ((!gate && prev_gate) ? 4 : mode);
    prev_gate = gate;
    const float d0 = mode == 2 ? k.d_d0 : (mode == 4 ? k.r_d0 : k.a_d0);
    const float target = mode == 1 ? k.a_target : (mode == 2 ? k.sus : -0.01f);
    const float step = __fadd_rn(x, __fmul_rn(d0, __fsub_rn(target, x)));
    const bool active = mode != 0, attack = mode == 1;
    const bool over = attack && step > 1.f, under = active && !attack && step < 0.0f;
    x = active ? (over ? 1.f : (under ? 0.f : step)) : x;
    mode = over ? 2 : (under ? 0 : mode);
    return active ? x : 0.0f;
}

__device__ __forceinline__ float polyblep(float dt, float t) {
    if (t < dt) {
        t = __fdiv_rn(t, dt);
        return __fsub_rn(__fsub_rn(__fadd_rn(t, t), __fmul_rn(t, t)), 1.0f);
    }
    if (t > __fsub_rn(1.0f, dt)) {
        t = __fdiv_rn(__fsub_rn(t, 1.0f), dt);
        return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t, t), t), t), 1.0f);
    }
    return 0.0f;
}

// sin(x) for 0 <= x <= pi/4, the only range Svf::SetFreq can ask for (its argument is pi * min(0.25, fc / (2 sr))):
// odd Taylor polynomial to x^9 in Horner form with FMAs; truncation error x^11 / 11! <= 1.8e-9, i.e. below half an ulp of
// the result, so it is as close to glibc's sinf (the oracle's) as CUDA's own sinf is, at 6 instructions instead of ~25.
__device__ __forceinline__ float sin_quarter(float x) {
    const float x2 = __fmul_rn(x, x);
    float p = __fmaf_rn(x2, 2.75573192e-6f, -1.98412698e-4f);
    p = __fmaf_rn(x2, p, 8.33333333e-3f);
    p = __fmaf_rn(x2, p, -1.66666667e-1f);
    return __fmaf_rn(__fmul_rn(x, x2), p, x);
}

// Sample-accurate events ON THE DEVICE (EVENTS = true; olb_voice_process_events): the host has applied the block's per-voice
// events to its parameter members in order and snapshotted, after each one, the addressed voice's parameter column; the
// snapshots are sorted by voice (stably: frame order within a voice).  A voice finds its own with a binary search at the
// start and, at the frame an event names, reloads its registers from the snapshot -- what the reference does when a setter
// runs between two Process calls (SynthVoice.h:101-267), at frame instead of block granularity and in ONE launch.
struct VoiceEvents {
    const uint32_t *voice;  // [count] ascending
    const uint32_t *frame;  // [count] block-relative
    const float *par;       // [count][VP_COUNT]
    uint32_t count;
};

// trace (optional): 4 floats per voice-frame = oscillator output, amp envelope, filter envelope, phase after the tick
template <bool EVENTS>
__global__ void __launch_bounds__(128) k_voice(float *out, float *partial, float *trace, uint32_t frames, uint32_t V,
                                               const float *__restrict__ par, float *retrig_flags, float *state, uint32_t partial_ld,
                                               const VoiceEvents ev) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = v < V;
    const uint32_t vv = live ? v : V - 1;  // dead lanes shadow a live voice so the warp reductions stay convergent
    auto P = [&](int p) { return par[(size_t) p * V + vv]; };
    float freq = P(VP_FREQ), port_coeff = P(VP_PORT_COEFF), sr_recip = P(VP_SR_RECIP), cutoff = P(VP_CUTOFF);
    float f_amount = P(VP_F_AMOUNT), a_amount = P(VP_A_AMOUNT), sr2 = P(VP_SR2), fc_max = P(VP_FC_MAX);
    float damp_a = P(VP_DAMP_A), drive = P(VP_DRIVE);
    int gate = (int) __float_as_uint(P(VP_GATE));
    // the NoteOn retrigger is a one-shot: read through its own (writable) pointer and cleared by the voice that consumed it, so
    // the host need not upload the table a second time to take it back
    const bool retrig = live && __float_as_uint(retrig_flags[v]) != 0;
    if (retrig) retrig_flags[v] = 0.0f;
    EnvCoef ka{P(VP_A_D0), P(VP_A_TARGET), P(VP_A_DD0), P(VP_A_RD0), P(VP_A_SUS)};
    EnvCoef kf{P(VP_F_D0), P(VP_F_TARGET), P(VP_F_DD0), P(VP_F_RD0), P(VP_F_SUS)};
    auto S = [&](int s) -> float & { return state[(size_t) s * V + vv]; };
    float phase = S(VS_PHASE), port_z = S(VS_PORT_Z), low = S(VS_LOW), band = S(VS_BAND), a_x = S(VS_A_X), f_x = S(VS_F_X);
    int a_mode = (int) __float_as_uint(S(VS_A_MODE)), a_gate = (int) __float_as_uint(S(VS_A_GATE));
    int f_mode = (int) __float_as_uint(S(VS_F_MODE)), f_gate = (int) __float_as_uint(S(VS_F_GATE));
    if (retrig) {  // NoteOn: Retrigger(hard) on both envelopes (SynthVoice.h:249-250)
        a_mode = 1; a_x = 0.f;
        f_mode = 1; f_x = 0.f;
    }
    __shared__ float ytile[4][32][33];
    // Svf::SetFreq runs every sample and its result is a tolerance tier already (device sine): its two divisions are a
    // multiplication by 1 / (2 sr) and a fast division (<= 2 ulp each)
    float inv_sr2 = __frcp_rn(sr2);
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
    const uint32_t warp_global = v >> 5;
    uint32_t ev_cur = 0, ev_next = 0xffffffffu;  // this voice's next snapshot and the frame it takes effect at
    if (EVENTS && live) {
        uint32_t lo = 0, hi = ev.count;
        while (lo < hi) {  // first snapshot with voice >= v
            const uint32_t mid = (lo + hi) >> 1;
            if (ev.voice[mid] < v) lo = mid + 1;
            else hi = mid;
        }
        ev_cur = lo;
        if (ev_cur < ev.count && ev.voice[ev_cur] == v) ev_next = ev.frame[ev_cur];
    }

    for (uint32_t n = 0; n < frames; n++) {
        if (EVENTS) {
            while (n == ev_next) {  // rare and divergent: a setter ran "between two Process calls"
                const float *q = ev.par + (size_t) ev_cur * VP_COUNT;
                freq = q[VP_FREQ]; gate = (int) __float_as_uint(q[VP_GATE]); port_coeff = q[VP_PORT_COEFF]; sr_recip = q[VP_SR_RECIP];
                cutoff = q[VP_CUTOFF]; f_amount = q[VP_F_AMOUNT]; a_amount = q[VP_A_AMOUNT]; sr2 = q[VP_SR2]; fc_max = q[VP_FC_MAX];
                damp_a = q[VP_DAMP_A]; drive = q[VP_DRIVE];
                ka = EnvCoef{q[VP_A_D0], q[VP_A_TARGET], q[VP_A_DD0], q[VP_A_RD0], q[VP_A_SUS]};
                kf = EnvCoef{q[VP_F_D0], q[VP_F_TARGET], q[VP_F_DD0], q[VP_F_RD0], q[VP_F_SUS]};
                inv_sr2 = __frcp_rn(sr2);
                if (__float_as_uint(q[VP_RETRIG]) != 0) {  // NoteOn: Retrigger(hard) on both envelopes
                    a_mode = 1; a_x = 0.f;
                    f_mode = 1; f_x = 0.f;
                }
                ev_cur++;
                ev_next = (ev_cur < ev.count && ev.voice[ev_cur] == v) ? ev.frame[ev_cur] : 0xffffffffu;
            }
        }
        const float amp_env = env_tick(a_x, a_mode, a_gate, gate, ka);
        const float amp = __fmul_rn(amp_env, a_amount);
        port_z = __fadd_rn(freq, __fmul_rn(port_coeff, __fsub_rn(port_z, freq)));  // Portamento.h:29-32
        const float inc = __fmul_rn(port_z, sr_recip);
        // PolyBLEP saw, amp 0.5 (OscillatorSoundSource.h:21-22)
        float o = __fsub_rn(__fmul_rn(2.0f, phase), 1.0f);
        o = __fsub_rn(o, polyblep(inc, phase));
        o = __fmul_rn(o, -1.0f);
        phase = __fadd_rn(phase, inc);
        if (phase > 1.0f) phase = __fsub_rn(phase, 1.0f);
        const float frame = __fmul_rn(o, 0.5f);
        const float fenv = env_tick(f_x, f_mode, f_gate, gate, kf);
        const float ff = __fadd_rn(cutoff, __fmul_rn(__fmul_rn(fenv, 20000.f), f_amount));
        // Svf::SetFreq every sample
        const float fc = fminf(fmaxf(ff, 1.0e-6f), fc_max);
        const float sfreq = __fmul_rn(2.0f, sin_quarter(__fmul_rn(3.1415927410125732421875f, fminf(0.25f, __fmul_rn(fc, inv_sr2)))));
        const float damp = fminf(damp_a, fminf(2.0f, __fsub_rn(__fdividef(2.0f, sfreq), __fmul_rn(sfreq, 0.5f))));
        // Svf::Process, two passes, low output
        float notch = __fsub_rn(frame, __fmul_rn(damp, band));
        low = __fadd_rn(low, __fmul_rn(sfreq, band));
        float high = __fsub_rn(notch, low);
        band = __fsub_rn(__fadd_rn(__fmul_rn(sfreq, high), band), __fmul_rn(__fmul_rn(__fmul_rn(drive, band), band), band));
        float o_low = __fmul_rn(0.5f, low);
        notch = __fsub_rn(frame, __fmul_rn(damp, band));
        low = __fadd_rn(low, __fmul_rn(sfreq, band));
        high = __fsub_rn(notch, low);
        band = __fsub_rn(__fadd_rn(__fmul_rn(sfreq, high), band), __fmul_rn(__fmul_rn(__fmul_rn(drive, band), band), band));
        o_low = __fadd_rn(o_low, __fmul_rn(0.5f, low));
        const float y = __fmul_rn(o_low, amp);
        if (out && live) out[(size_t) n * V + v] = y;
        if (trace && live) {
            float4 t4 = make_float4(frame, amp_env, fenv, phase);
            reinterpret_cast<float4 *>(trace)[(size_t) n * V + v] = t4;
        }
        if (partial) {
            // warp mixdown without a per-frame shuffle chain: 32 frames of the warp's 32 voices are parked in shared memory,
            // then lane f adds frame f's voices in voice order (rows are padded to 33 floats: both sweeps are conflict-free)
            ytile[wic][n & 31u][lane] = live ? y : 0.f;
            if ((n & 31u) == 31u || n + 1 == frames) {
                __syncwarp();
                const uint32_t n0 = n & ~31u;
                if (n0 + lane <= n) {
                    float s = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; j++) s = __fadd_rn(s, ytile[wic][lane][j]);
                    partial[(size_t) warp_global * partial_ld + n0 + lane] = s;
                }
                __syncwarp();
            }
        }
    }
    if (live) {
        S(VS_PHASE) = phase; S(VS_PORT_Z) = port_z; S(VS_LOW) = low; S(VS_BAND) = band; S(VS_A_X) = a_x; S(VS_F_X) = f_x;
        S(VS_A_MODE) = __uint_as_float((uint32_t) a_mode); S(VS_A_GATE) = __uint_as_float((uint32_t) a_gate);
        S(VS_F_MODE) = __uint_as_float((uint32_t) f_mode); S(VS_F_GATE) = __uint_as_float((uint32_t) f_gate);
    }
}

// bus[n * bus_stride + c] (+)= gain * sum over warps (in warp order) for every channel c < bus_channels:
// Polyvoice is mono; the caller duplicates it to stereo (modules/ol_daisy/app/synth/main.cpp:83).
__global__ void k_bus_reduce(const float *__restrict__ partial, uint32_t warps, uint32_t partial_ld, float *bus, uint32_t frames,
                             uint32_t bus_channels, int accumulate, float gain) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= frames) return;
    float s = 0.f;
    for (uint32_t w = 0; w < warps; w++) s = __fadd_rn(s, partial[(size_t) w * partial_ld + n]);
    s = __fmul_rn(s, gain);
    for (uint32_t c = 0; c < bus_channels; c++) {
        float *p = bus + (size_t) n * bus_channels + c;
        *p = accumulate ? __fadd_rn(*p, s) : s;
    }
}

}  // namespace olb

using namespace olb;

// ================================================================================================
// Host side: one host::VoiceParams per voice carries the reference object's members.
struct olb_voice {
    std::recursive_mutex mu;
    StreamOrder order;
    int device = 0;
    uint32_t V = 0;
    std::vector<host::VoiceParams> voices;
    bool dirty = false;
    uint32_t dirty_lo = 0, dirty_hi = 0;  // the voices whose parameters changed since the last upload: [lo, hi)
    bool retrig_pending = false;
    uint32_t retrig_lo = 0, retrig_hi = 0;
    void mark(uint32_t lo, uint32_t hi) {
        if (!dirty) { dirty_lo = lo; dirty_hi = hi; }
        else { dirty_lo = lo < dirty_lo ? lo : dirty_lo; dirty_hi = hi > dirty_hi ? hi : dirty_hi; }
        dirty = true;
    }
    float *d_par = nullptr, *d_state = nullptr, *d_partial = nullptr;
    unsigned char *d_ev = nullptr;  // the device-resident event snapshots of olb_voice_process_events: [E] voice, [E] frame, [E][VP_COUNT]
    size_t ev_cap = 0;
    std::vector<unsigned char> ev_stage;
    uint32_t partial_ld = 0;  // frames capacity of the warp-partial buffer
    std::vector<float> stage;
    HostStage host;
};

static float vbits_f(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// One voice's column of the device parameter table: element p goes to out[p * stride].
static void voice_column(const host::VoiceParams &u, float *out, size_t stride) {
    out[VP_FREQ * stride] = u.freq;
    out[VP_GATE * stride] = vbits_f((uint32_t) u.gate);
    out[VP_RETRIG * stride] = vbits_f(u.retrigger ? 1u : 0u);
    out[VP_PORT_COEFF * stride] = u.port_coeff;
    out[VP_SR_RECIP * stride] = 1.0f / u.sample_rate;  // Oscillator::Init
    out[VP_CUTOFF * stride] = u.filter_cutoff;
    out[VP_F_AMOUNT * stride] = u.filter_env_amount;
    out[VP_A_AMOUNT * stride] = u.amp_env_amount;
    out[VP_SR2 * stride] = u.svf.sr * 2.0f;
    out[VP_FC_MAX * stride] = u.svf.fc_max;
    out[VP_DAMP_A * stride] = 2.0f * (1.0f - powf(u.svf.res, 0.25f));
    out[VP_DRIVE * stride] = u.svf.drive;
    out[VP_A_D0 * stride] = u.amp_env.a_d0; out[VP_A_TARGET * stride] = u.amp_env.a_target; out[VP_A_DD0 * stride] = u.amp_env.d_d0;
    out[VP_A_RD0 * stride] = u.amp_env.r_d0; out[VP_A_SUS * stride] = u.amp_env.sus;
    out[VP_F_D0 * stride] = u.filter_env.a_d0; out[VP_F_TARGET * stride] = u.filter_env.a_target; out[VP_F_DD0 * stride] = u.filter_env.d_d0;
    out[VP_F_RD0 * stride] = u.filter_env.r_d0; out[VP_F_SUS * stride] = u.filter_env.sus;
}

// Uploads the columns of the voices whose parameters changed: one 2-D copy of [VP_COUNT][hi - lo] out of the [VP_COUNT][V]
// table (a NoteOn of one voice is 22 floats, not the bank's 5.8 MB).
static int voice_upload(olb_voice *h, cudaStream_t st) {
    const size_t V = h->V;
    if (h->dirty_lo >= h->dirty_hi) { h->dirty_lo = 0; h->dirty_hi = h->V; }
    const size_t lo = h->dirty_lo, hi = h->dirty_hi;
    h->stage.resize((size_t) VP_COUNT * V);
    float *s = h->stage.data();
    for (size_t i = lo; i < hi; i++) voice_column(h->voices[i], s + i, V);
    OLB_CUDA(cudaMemcpy2DAsync(h->d_par + lo, V * sizeof(float), s + lo, V * sizeof(float), (hi - lo) * sizeof(float), VP_COUNT,
                               cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaStreamSynchronize(st));  // the staging table is pageable and reused
    h->dirty = false;
    h->dirty_lo = h->dirty_hi = 0;
    return OLB_OK;
}

static int voice_init_all(olb_voice *h, float sr) {
    for (auto &u : h->voices) {
        u.Init(sr);
        u.retrigger = false;
    }
    OLB_CUDA(h->order.memset0(h->d_state, (size_t) VS_COUNT * h->V * sizeof(float)));
    h->mark(0, h->V);
    return OLB_OK;
}

static void voice_free(olb_voice *h) {
    cudaFree(h->d_par); cudaFree(h->d_state); cudaFree(h->d_partial); cudaFree(h->d_ev);
    h->host.release();
    delete h;
}

OLB_API int olb_voice_create(int device, uint32_t voices, float sample_rate, olb_voice **out) {
    OLB_ARG(out, "olb_voice_create: out is NULL");
    *out = nullptr;
    OLB_ARG(voices > 0, "olb_voice_create: voices must be > 0");
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_voice *h = new olb_voice();
    h->device = device;
    h->V = voices;
    h->voices.resize(voices);
    cudaError_t e = cudaMalloc(&h->d_par, (size_t) VP_COUNT * voices * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_state, (size_t) VS_COUNT * voices * sizeof(float));
    if (e != cudaSuccess) {
        voice_free(h);
        return fail_cuda(e, "allocating voice bank", __FILE__, __LINE__);
    }
    rc = voice_init_all(h, sample_rate);
    if (rc) {
        voice_free(h);
        return rc;
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_voice_destroy(olb_voice *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    voice_free(h);
    return OLB_OK;
}

OLB_API int olb_voice_init(olb_voice *h, float sample_rate) {
    OLB_ARG(h, "olb_voice_init: NULL handle");
    OLB_LOCK(h);
    DeviceGuard g(h->device);
    return voice_init_all(h, sample_rate);
}

#define OLB_VOICE_RANGE(fn)                                                                                              \
    OLB_ARG(h, fn ": NULL handle");                                                                                      \
    OLB_ARG((uint64_t) first + count <= h->V, fn ": range [%u, %u) exceeds %u voices", first, first + count, h->V)

OLB_API int olb_voice_update_midi(olb_voice *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value) {
    OLB_VOICE_RANGE("olb_voice_update_midi");
    OLB_LOCK(h);
    for (uint32_t i = first; i < first + count; i++) h->voices[i].UpdateMidiControl(control, value);
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_update_hw(olb_voice *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride) {
    OLB_VOICE_RANGE("olb_voice_update_hw");
    OLB_LOCK(h);
    OLB_ARG(values_host, "olb_voice_update_hw: values_host is NULL");
    for (uint32_t i = 0; i < count; i++) h->voices[first + i].UpdateHardwareControl(control, values_host[(size_t) i * stride]);
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_update_config(olb_voice *h, uint32_t first, uint32_t count, const float *config16_host) {
    OLB_VOICE_RANGE("olb_voice_update_config");
    OLB_LOCK(h);
    OLB_ARG(config16_host, "olb_voice_update_config: config is NULL");
    for (uint32_t i = first; i < first + count; i++) h->voices[i].UpdateConfig(config16_host);
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_note_on(olb_voice *h, uint32_t first, uint32_t count, const uint8_t *notes_host, uint32_t stride, uint8_t velocity) {
    OLB_VOICE_RANGE("olb_voice_note_on");
    OLB_LOCK(h);
    OLB_ARG(notes_host, "olb_voice_note_on: notes_host is NULL");
    (void) velocity;  // the reference ignores it too (SynthVoice.h:245-251)
    for (uint32_t i = 0; i < count; i++) h->voices[first + i].NoteOn(notes_host[(size_t) i * stride]);
    h->mark(first, first + count);
    if (!h->retrig_pending) { h->retrig_lo = first; h->retrig_hi = first + count; }
    else { h->retrig_lo = first < h->retrig_lo ? first : h->retrig_lo; h->retrig_hi = first + count > h->retrig_hi ? first + count : h->retrig_hi; }
    h->retrig_pending = true;
    return OLB_OK;
}

OLB_API int olb_voice_note_off(olb_voice *h, uint32_t first, uint32_t count) {
    OLB_VOICE_RANGE("olb_voice_note_off");
    OLB_LOCK(h);
    for (uint32_t i = first; i < first + count; i++) h->voices[i].NoteOff();
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_set_gate(olb_voice *h, uint32_t first, uint32_t count, int gate) {
    OLB_VOICE_RANGE("olb_voice_set_gate");
    OLB_LOCK(h);
    for (uint32_t i = first; i < first + count; i++) h->voices[i].gate = gate ? 1 : 0;
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_set_frequency(olb_voice *h, uint32_t first, uint32_t count, const float *freqs_host, uint32_t stride) {
    OLB_VOICE_RANGE("olb_voice_set_frequency");
    OLB_LOCK(h);
    OLB_ARG(freqs_host, "olb_voice_set_frequency: freqs_host is NULL");
    for (uint32_t i = 0; i < count; i++) h->voices[first + i].freq = freqs_host[(size_t) i * stride];
    h->mark(first, first + count);
    return OLB_OK;
}

OLB_API int olb_voice_playing(olb_voice *h, uint32_t first, uint32_t count, uint8_t *playing_host) {
    OLB_VOICE_RANGE("olb_voice_playing");
    OLB_LOCK(h);
    OLB_ARG(playing_host, "olb_voice_playing: playing_host is NULL");
    for (uint32_t i = 0; i < count; i++) playing_host[i] = h->voices[first + i].playing;
    return OLB_OK;
}

// The launch itself: the parameter table must be current (voice_upload) and `st` entered.  ev: the block's device-resident
// per-voice events, or count == 0.
static int voice_launch(olb_voice *h, float *out_dev, float *bus_dev, uint32_t bus_channels, int accumulate, float bus_gain, float *trace_dev,
                        uint64_t frames, cudaStream_t st, const VoiceEvents &ev) {
    const uint32_t warps = (h->V + 31) / 32;
    if (bus_dev && h->partial_ld < frames) {
        OLB_CUDA(cudaStreamSynchronize(st));
        cudaFree(h->d_partial);
        h->d_partial = nullptr;
        h->partial_ld = 0;
        const uint32_t ld = (uint32_t) ((frames + 31) / 32 * 32);
        OLB_CUDA(cudaMalloc(&h->d_partial, (size_t) warps * ld * sizeof(float)));
        h->partial_ld = ld;
    }
    if (ev.count)
        k_voice<true><<<(h->V + 127) / 128, 128, 0, st>>>(out_dev, bus_dev ? h->d_partial : nullptr, trace_dev, (uint32_t) frames, h->V, h->d_par,
                                                         h->d_par + (size_t) VP_RETRIG * h->V, h->d_state, h->partial_ld, ev);
    else
        k_voice<false><<<(h->V + 127) / 128, 128, 0, st>>>(out_dev, bus_dev ? h->d_partial : nullptr, trace_dev, (uint32_t) frames, h->V, h->d_par,
                                                          h->d_par + (size_t) VP_RETRIG * h->V, h->d_state, h->partial_ld, ev);
    OLB_LAUNCHED();
    if (bus_dev) {
        k_bus_reduce<<<(unsigned) ((frames + 127) / 128), 128, 0, st>>>(h->d_partial, warps, h->partial_ld, bus_dev, (uint32_t) frames,
                                                                       bus_channels, accumulate, bus_gain);
        OLB_LAUNCHED();
    }
    return OLB_OK;
}

// a pending NoteOn retrigger is consumed by the next launch (the kernel clears the device flags): the host clears its own
static void voice_retrig_consumed(olb_voice *h) {
    if (!h->retrig_pending) return;
    for (uint32_t i = h->retrig_lo; i < h->retrig_hi; i++) h->voices[i].retrigger = false;
    h->retrig_pending = false;
}

OLB_API int olb_voice_process(olb_voice *h, float *out_dev, float *bus_dev, uint32_t bus_channels, int accumulate, float bus_gain,
                              float *trace_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_voice_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(out_dev || bus_dev || trace_dev, "olb_voice_process: no output buffer given");
    OLB_ARG(!bus_dev || (bus_channels >= 1 && bus_channels <= 8), "olb_voice_process: bus_channels must be 1..8");
    OLB_ARG(frames < (1ull << 31), "olb_voice_process: at most 2^31-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    OLB_CUDA(h->order.enter(st));
    if (h->dirty) {
        int rc = voice_upload(h, st);
        if (rc) return rc;
    }
    int rc = voice_launch(h, out_dev, bus_dev, bus_channels, accumulate, bus_gain, trace_dev, frames, st, VoiceEvents{nullptr, nullptr, nullptr, 0});
    if (rc) return rc;
    voice_retrig_consumed(h);
    return OLB_OK;
}

// Host buffers.  out_host (optional) [frames][voices] = SynthVoice::Process of every voice; bus_host (optional)
// [frames][bus_channels] = Polyvoice::Process: with accumulate != 0 the mix is ADDED to what bus_host holds (the reference
// adds into *frame_out and the caller zeroes it: Polyvoice.h:28-33, main.cpp:79).  Chunked in time; synchronous.
OLB_API int olb_voice_process_host(olb_voice *h, float *out_host, float *bus_host, uint32_t bus_channels, int accumulate, float bus_gain,
                                   uint64_t frames) {
    OLB_ARG(h, "olb_voice_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(out_host || bus_host, "olb_voice_process_host: no output buffer given");
    OLB_ARG(!bus_host || (bus_channels >= 1 && bus_channels <= 8), "olb_voice_process_host: bus_channels must be 1..8");
    DeviceGuard g(h->device);
    const uint64_t V = h->V, chunk = host_chunk_frames(V, frames);
    int rc = OLB_OK;
    if (out_host && (rc = h->host.ensure(0, chunk * V))) return rc;
    if (bus_host && (rc = h->host.ensure(1, chunk * bus_channels))) return rc;
    if (!h->host.st && (rc = h->host.ensure(2, 1))) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        if (bus_host && accumulate)
            OLB_CUDA(cudaMemcpyAsync(h->host.d[1], bus_host + t0 * bus_channels, n * bus_channels * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_voice_process(h, out_host ? h->host.d[0] : nullptr, bus_host ? h->host.d[1] : nullptr, bus_channels, accumulate, bus_gain,
                               nullptr, n, h->host.st);
        if (rc) return rc;
        if (out_host) OLB_CUDA(cudaMemcpyAsync(out_host + t0 * V, h->host.d[0], n * V * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
        if (bus_host)
            OLB_CUDA(cudaMemcpyAsync(bus_host + t0 * bus_channels, h->host.d[1], n * bus_channels * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}

// Polyvoice::Process of a bank whose voices are sharded over the ranks of `c`: every rank runs its voices into its partial
// bus, the partial buses are summed in place over NVLink (olb_voice_bus_allreduce), and every rank gets the global mixdown
// in bus_host.  With accumulate != 0 the mix is added to what bus_host holds ON RANK 0's terms: each rank adds its own
// bus_host contents, so callers that accumulate pass the running frame on one rank and zeros on the others.
OLB_API int olb_voice_process_host_comm(olb_voice *h, olb_comm *c, float *bus_host, uint32_t bus_channels, int accumulate, float bus_gain,
                                        uint64_t frames) {
    OLB_ARG(h && c, "olb_voice_process_host_comm: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(bus_host, "olb_voice_process_host_comm: bus_host is NULL");
    OLB_ARG(bus_channels >= 1 && bus_channels <= 8, "olb_voice_process_host_comm: bus_channels must be 1..8");
    DeviceGuard g(h->device);
    const uint64_t V = h->V, chunk = host_chunk_frames(V, frames);
    int rc = h->host.ensure(1, chunk * bus_channels);
    if (rc) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        if (accumulate)
            OLB_CUDA(cudaMemcpyAsync(h->host.d[1], bus_host + t0 * bus_channels, n * bus_channels * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_voice_process(h, nullptr, h->host.d[1], bus_channels, accumulate, bus_gain, nullptr, n, h->host.st);
        if (rc) return rc;
        rc = olb_voice_bus_allreduce(c, h->host.d[1], n, bus_channels, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(bus_host + t0 * bus_channels, h->host.d[1], n * bus_channels * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}

// Sample-accurate control: `events` (sorted by frame) are applied BEFORE the frame they name, the block is cut there, and
// the pieces run back to back on the stream -- what a host does by hand when it applies MIDI between audio blocks
// (modules/juce/plugin/workout/src/PluginProcessor.cpp:24-38), at frame instead of block granularity.  Events with
// voice < 0 go through Polyvoice's allocation (first voice that is not playing / first voice playing the note:
// Polyvoice.h:35-51), events with voice >= 0 address that voice (the Voice interface, Voice.h:33-57).
OLB_API int olb_voice_process_events(olb_voice *h, const olb_voice_event *events, uint32_t n_events, float *out_dev, float *bus_dev,
                                     uint32_t bus_channels, int accumulate, float bus_gain, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_voice_process_events: NULL handle");
    OLB_LOCK(h);
    OLB_ARG(n_events == 0 || events, "olb_voice_process_events: events is NULL");
    for (uint32_t e = 0; e < n_events; e++) {
        OLB_ARG(events[e].frame <= frames && (e == 0 || events[e].frame >= events[e - 1].frame),
                "olb_voice_process_events: event %u is out of order or past the block", e);
        OLB_ARG(events[e].voice < 0 || (uint32_t) events[e].voice < h->V, "olb_voice_process_events: event %u addresses voice %d of %u", e,
                events[e].voice, h->V);
    }
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    // One event, applied to the host-side members exactly as the public setters do.  Returns the voice it addressed, -1 for an
    // event that was dropped (every voice busy / no voice plays the note), -2 for a fan-out to every voice.
    auto apply = [&](const olb_voice_event &ev, int32_t &voice_out) -> int {
        const uint8_t note = (uint8_t) ev.a;
        int32_t voice = ev.voice;
        switch (ev.kind) {
            case OLB_EV_NOTE_ON:
                if (voice < 0)
                    for (uint32_t v = 0; v < h->V && voice < 0; v++)
                        if (!h->voices[v].playing) voice = (int32_t) v;
                if (voice >= 0) h->voices[voice].NoteOn(note);  // every voice busy: dropped, as in the reference
                break;
            case OLB_EV_NOTE_OFF:
                if (voice < 0)
                    for (uint32_t v = 0; v < h->V && voice < 0; v++)
                        if (h->voices[v].playing == note) voice = (int32_t) v;
                if (voice >= 0) h->voices[voice].NoteOff();
                break;
            case OLB_EV_MIDI_CC:
                if (voice < 0) {
                    for (auto &u : h->voices) u.UpdateMidiControl((uint8_t) ev.a, (uint8_t) ev.v);
                    voice = -2;
                } else {
                    h->voices[voice].UpdateMidiControl((uint8_t) ev.a, (uint8_t) ev.v);
                }
                break;
            case OLB_EV_HW_CC:
                if (voice < 0) {
                    for (auto &u : h->voices) u.UpdateHardwareControl((uint8_t) ev.a, ev.v);
                    voice = -2;
                } else {
                    h->voices[voice].UpdateHardwareControl((uint8_t) ev.a, ev.v);
                }
                break;
            default:
                set_error("olb_voice_process_events: unknown event kind %u", ev.kind);
                return OLB_E_ARG;
        }
        voice_out = voice;
        return OLB_OK;
    };
    auto fans_out = [](const olb_voice_event &ev) { return ev.voice < 0 && (ev.kind == OLB_EV_MIDI_CC || ev.kind == OLB_EV_HW_CC); };

    // The block is cut only where an event addresses EVERY voice (a Polyvoice-wide control change: the whole table changes);
    // events that address one voice -- notes through Polyvoice's allocation, per-voice controls -- ride along with the launch.
    uint64_t pos = 0;
    uint32_t e = 0;
    while (pos < frames || e < n_events) {
        // (1) events AT the segment's first frame change the table the launch starts from
        for (; e < n_events && events[e].frame == pos; e++) {
            int32_t voice;
            int rc = apply(events[e], voice);
            if (rc) return rc;
            if (voice == -2) {
                h->mark(0, h->V);
            } else if (voice >= 0) {
                h->mark((uint32_t) voice, (uint32_t) voice + 1);
                if (events[e].kind == OLB_EV_NOTE_ON) {
                    if (!h->retrig_pending) { h->retrig_lo = (uint32_t) voice; h->retrig_hi = (uint32_t) voice + 1; }
                    else { h->retrig_lo = std::min(h->retrig_lo, (uint32_t) voice); h->retrig_hi = std::max(h->retrig_hi, (uint32_t) voice + 1); }
                    h->retrig_pending = true;
                }
            }
        }
        if (pos >= frames) break;
        uint64_t seg_end = frames;
        uint32_t e_end = e;
        for (; e_end < n_events; e_end++)
            if (fans_out(events[e_end])) {
                seg_end = events[e_end].frame;
                break;
            }
        if (e_end == n_events) seg_end = frames;
        // (2) the table as of `pos` goes to the device now: the members are about to move on
        OLB_CUDA(h->order.enter(st));
        if (h->dirty) {
            int rc = voice_upload(h, st);
            if (rc) return rc;
        }
        voice_retrig_consumed(h);  // the launch below consumes the device flags
        // (3) the segment's per-voice events: apply in order, snapshot the addressed voice's column after each
        struct Snap { uint32_t voice, frame, order; };
        std::vector<Snap> snaps;
        std::vector<float> cols;
        uint32_t lo = h->V, hi = 0;
        for (; e < e_end && events[e].frame < seg_end; e++) {  // an event AT seg_end starts the next segment (or ends the block)
            int32_t voice;
            int rc = apply(events[e], voice);
            if (rc) return rc;
            if (voice < 0) continue;  // dropped
            snaps.push_back(Snap{(uint32_t) voice, (uint32_t) (events[e].frame - pos), (uint32_t) snaps.size()});
            cols.resize(cols.size() + VP_COUNT);
            voice_column(h->voices[voice], cols.data() + cols.size() - VP_COUNT, 1);
            h->voices[voice].retrigger = false;  // carried by this snapshot alone
            lo = std::min(lo, (uint32_t) voice);
            hi = std::max(hi, (uint32_t) voice + 1);
        }
        VoiceEvents dev{nullptr, nullptr, nullptr, 0};
        if (!snaps.empty()) {
            std::stable_sort(snaps.begin(), snaps.end(), [](const Snap &a, const Snap &b) { return a.voice < b.voice; });
            const size_t E = snaps.size(), bytes = E * (2 * sizeof(uint32_t) + VP_COUNT * sizeof(float));
            h->ev_stage.resize(bytes);
            uint32_t *hv = reinterpret_cast<uint32_t *>(h->ev_stage.data()), *hf = hv + E;
            float *hp = reinterpret_cast<float *>(hf + E);
            for (size_t i = 0; i < E; i++) {
                hv[i] = snaps[i].voice;
                hf[i] = snaps[i].frame;
                memcpy(hp + i * VP_COUNT, cols.data() + (size_t) snaps[i].order * VP_COUNT, VP_COUNT * sizeof(float));
            }
            if (h->ev_cap < bytes) {
                OLB_CUDA(cudaStreamSynchronize(st));
                cudaFree(h->d_ev);
                h->d_ev = nullptr;
                h->ev_cap = 0;
                OLB_CUDA(cudaMalloc(&h->d_ev, bytes * 2));
                h->ev_cap = bytes * 2;
            }
            OLB_CUDA(cudaMemcpyAsync(h->d_ev, h->ev_stage.data(), bytes, cudaMemcpyHostToDevice, st));
            OLB_CUDA(cudaStreamSynchronize(st));  // pageable staging, reused by the next segment
            const uint32_t *dv = reinterpret_cast<const uint32_t *>(h->d_ev);
            dev = VoiceEvents{dv, dv + E, reinterpret_cast<const float *>(dv + 2 * E), (uint32_t) E};
        }
        // (4) one launch for the segment
        const uint64_t n = seg_end - pos;
        int rc = voice_launch(h, out_dev ? out_dev + pos * h->V : nullptr, bus_dev ? bus_dev + pos * bus_channels : nullptr, bus_channels, accumulate,
                              bus_gain, nullptr, n, st, dev);
        if (rc) return rc;
        if (lo < hi) h->mark(lo, hi);  // the table on the device is the segment's FIRST frame's: bring it up to date at the next call
        pos = seg_end;
    }
    return OLB_OK;
}
