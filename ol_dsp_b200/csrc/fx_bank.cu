// fx_bank.cu -- banks of the reference's fxlib units, mono instances (N = 1):
//   (a2) FilterFx<1>   modules/fxlib/Fx.h:65-165   one daisysp::Svf (two passes per sample), output selected by type
//   (a3) DelayFx<1>    modules/fxlib/Fx.h:170-268  DelayLine<float,48000> with feedback, SVF on the wet path, dry/wet
//   (a4) ReverbFx<1>   modules/fxlib/Fx.h:272-393  dry/wet around the stub ReverbSc (wet = 0.8 * in, Reverb.h:26-31)
//   (a5) FxRack<1>     modules/fxlib/Fx.h:397-493  delay -> reverb -> filter -> * master_volume
// One instance per thread, time-sequential, every float operation rounded separately in the reference's order
// (bit-exact against the reference headers built with -ffp-contract=off over the DaisySP restatement).
// All setter maths (ol::core::scale, Svf::SetFreq's sinf/powf ...) runs on the host (host_params.h).
//
// Layout in HBM
//   audio  [frames][instances] fp32
//   par    [14][instances]: main filter freq damp drive type | wet filter freq damp drive type | delay feedback balance
//          delay(bits) frac | reverb balance | master volume
//   svf    [4][instances]: main low band, wet low band
//   ring   [instances][48 000] fp32, one contiguous line per instance (delay times differ per instance, so rings are
//          walked per thread; a 32-byte sector is consumed over 8 consecutive samples out of L1/L2).  The write
//          pointer is shared by the bank (every instance is Init()ed and ticked together).
#include <stdlib.h>

#include <vector>

#include "common.cuh"
#include "host_params.h"

namespace olb {

constexpr int kRing = host::kMaxDelay;
enum { P_F_FREQ, P_F_DAMP, P_F_DRIVE, P_F_TYPE, P_W_FREQ, P_W_DAMP, P_W_DRIVE, P_W_TYPE, P_D_FEEDBACK, P_D_BALANCE, P_D_DELAY, P_D_FRAC,
       P_R_BALANCE, P_MASTER, P_COUNT };

struct SvfCoef {
    float freq, damp, drive;
    int type;
};

// daisysp::Svf::Process, two identical passes; returns the output FilterFx::Process selects (Fx.h:88-108)
__device__ __forceinline__ float svf_tick(float in, float &low, float &band, const SvfCoef &k) {
    float notch = __fsub_rn(in, __fmul_rn(k.damp, band));
    low = __fadd_rn(low, __fmul_rn(k.freq, band));
    float high = __fsub_rn(notch, low);
    band = __fsub_rn(__fadd_rn(__fmul_rn(k.freq, high), band), __fmul_rn(__fmul_rn(__fmul_rn(k.drive, band), band), band));
    float o_low = __fmul_rn(0.5f, low), o_high = __fmul_rn(0.5f, high), o_band = __fmul_rn(0.5f, band);
    float o_peak = __fmul_rn(0.5f, __fsub_rn(low, high)), o_notch = __fmul_rn(0.5f, notch);
    notch = __fsub_rn(in, __fmul_rn(k.damp, band));
    low = __fadd_rn(low, __fmul_rn(k.freq, band));
    high = __fsub_rn(notch, low);
    band = __fsub_rn(__fadd_rn(__fmul_rn(k.freq, high), band), __fmul_rn(__fmul_rn(__fmul_rn(k.drive, band), band), band));
    switch (k.type) {
        case 2: return __fadd_rn(o_high, __fmul_rn(0.5f, high));
        case 1: return __fadd_rn(o_band, __fmul_rn(0.5f, band));
        case 4: return __fadd_rn(o_peak, __fmul_rn(0.5f, __fsub_rn(low, high)));
        case 3: return __fadd_rn(o_notch, __fmul_rn(0.5f, notch));
        default: return __fadd_rn(o_low, __fmul_rn(0.5f, low));
    }
}

__device__ __forceinline__ float dry_wet(float wet, float dry, float balance) {  // Fx.h:205-207, 297-299
    return __fadd_rn(__fmul_rn(wet, balance), __fmul_rn(dry, __fsub_rn(1.0f, balance)));
}

template <int KIND>
__global__ void __launch_bounds__(64) k_fx(const float *in, float *out, uint32_t frames, uint32_t I,
                                           const float *__restrict__ par, float *svf, float *ring_all, uint32_t wp0) {
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    auto P = [&](int p) { return par[(size_t) p * I + inst]; };
    const SvfCoef kf{P(P_F_FREQ), P(P_F_DAMP), P(P_F_DRIVE), (int) __float_as_uint(P(P_F_TYPE))};
    const SvfCoef kw{P(P_W_FREQ), P(P_W_DAMP), P(P_W_DRIVE), (int) __float_as_uint(P(P_W_TYPE))};
    const float feedback = P(P_D_FEEDBACK), d_balance = P(P_D_BALANCE), frac = P(P_D_FRAC);
    const uint32_t delay = __float_as_uint(P(P_D_DELAY));
    const float r_balance = P(P_R_BALANCE), master = P(P_MASTER);
    float f_low = svf[inst], f_band = svf[(size_t) I + inst], w_low = svf[2 * (size_t) I + inst], w_band = svf[3 * (size_t) I + inst];
    float *ring = ring_all + (size_t) inst * kRing;
    uint32_t wp = wp0;

    for (uint32_t n = 0; n < frames; n++) {
        float v = in[(size_t) n * I + inst];
        if (KIND == OLB_FX_DELAY || KIND == OLB_FX_RACK) {
            // DelayLine::Read then Write (Fx.h:198-200); the write pointer runs backwards
            uint32_t ia = wp + delay;
            ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
            uint32_t ib = ia + 1;
            ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
            const float a = ring[ia], b = ring[ib];
            const float rd = __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), frac));
            ring[wp] = __fadd_rn(v, __fmul_rn(feedback, rd));
            wp = wp == 0 ? kRing - 1 : wp - 1;
            const float wet = svf_tick(rd, w_low, w_band, kw);
            v = dry_wet(wet, v, d_balance);
        }
        if (KIND == OLB_FX_REVERB || KIND == OLB_FX_RACK) v = dry_wet(__fmul_rn(v, 0.8f), v, r_balance);
        if (KIND == OLB_FX_FILTER || KIND == OLB_FX_RACK) v = svf_tick(v, f_low, f_band, kf);
        if (KIND == OLB_FX_RACK) v = __fmul_rn(v, master);
        out[(size_t) n * I + inst] = v;
    }
    svf[inst] = f_low;
    svf[(size_t) I + inst] = f_band;
    svf[2 * (size_t) I + inst] = w_low;
    svf[3 * (size_t) I + inst] = w_band;
}

// ------------------------------------------------------------------------------------------------
// k_fx_pf: the same arithmetic with (1) register double buffering -- the inputs (and, for the delay kinds, the ring
// reads) of frames n+B .. n+2B-1 are in flight while frames n .. n+B-1 run, so the HBM round trip is paid once per B
// frames instead of once per frame (in and out may alias, so the compiler will not hoist the loads itself) -- and
// (2) L active lanes per warp: a bank of a few thousand instances is only a few hundred full warps for 592 SM
// sub-partitions and each warp is one long dependent chain, so narrower warps (L = 16 or 8; a warp row is then a
// 64- or 32-byte segment, still whole sectors) give every sub-partition several chains to interleave.
// Ring reads run ahead only when the delay exceeds the look-ahead (2B + 2 samples); shorter delays take the
// direct path (read after the previous frame's write), lane by lane.
constexpr int kFxB = 8;

template <int KIND, int L>
__global__ void __launch_bounds__(64) k_fx_pf(const float *in, float *out, uint32_t frames, uint32_t I,
                                              const float *__restrict__ par, float *svf, float *ring_all, uint32_t wp0) {
    constexpr bool kHasRing = (KIND == OLB_FX_DELAY || KIND == OLB_FX_RACK);
    constexpr int B = kFxB;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const uint32_t inst = warp * L + lane;
    if (lane >= (uint32_t) L || inst >= I) return;
    auto P = [&](int p) { return par[(size_t) p * I + inst]; };
    const SvfCoef kf{P(P_F_FREQ), P(P_F_DAMP), P(P_F_DRIVE), (int) __float_as_uint(P(P_F_TYPE))};
    const SvfCoef kw{P(P_W_FREQ), P(P_W_DAMP), P(P_W_DRIVE), (int) __float_as_uint(P(P_W_TYPE))};
    const float feedback = P(P_D_FEEDBACK), d_balance = P(P_D_BALANCE), frac = P(P_D_FRAC);
    const uint32_t delay = __float_as_uint(P(P_D_DELAY));
    const float r_balance = P(P_R_BALANCE), master = P(P_MASTER);
    float f_low = svf[inst], f_band = svf[(size_t) I + inst], w_low = svf[2 * (size_t) I + inst], w_band = svf[3 * (size_t) I + inst];
    float *ring = kHasRing ? ring_all + (size_t) inst * kRing : nullptr;
    const bool far = kHasRing && delay >= (uint32_t) (2 * B + 2);
    uint32_t wp = wp0;                    // write pointer at the frame being computed
    uint32_t la = 0;                      // read index (a) of the next frame to be fetched
    float prev_a = 0.f;                   // a of the previous frame == b of this one (far path)
    if (kHasRing) {
        la = wp0 + delay;
        la = la >= (uint32_t) kRing ? la - kRing : la;
        if (far) {
            const uint32_t ib = la + 1 >= (uint32_t) kRing ? la + 1 - kRing : la + 1;
            prev_a = ring[ib];
        }
    }
    const float *src = in + inst;
    float *dst = out + inst;

    float x0[B], x1[B], a0[B], a1[B];
    auto fetch = [&](float (&x)[B], float (&a)[B], uint32_t n0) {
#pragma unroll
        for (int j = 0; j < B; j++) {
            if (n0 + j < frames) {
                x[j] = src[(size_t) (n0 + j) * I];
                if (far) {
                    a[j] = ring[la];
                    la = la == 0 ? kRing - 1 : la - 1;
                }
            }
        }
    };
    auto run = [&](const float (&x)[B], const float (&a)[B], uint32_t n0) {
#pragma unroll
        for (int j = 0; j < B; j++) {
            if (n0 + j >= frames) break;
            float v = x[j];
            if (kHasRing) {
                float ra, rb;
                if (far) {
                    ra = a[j];
                    rb = prev_a;
                    prev_a = ra;
                } else {  // DelayLine::Read straight from the ring (Fx.h:198-200)
                    uint32_t ia = wp + delay;
                    ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
                    uint32_t ib = ia + 1;
                    ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
                    ra = ring[ia];
                    rb = ring[ib];
                }
                const float rd = __fadd_rn(ra, __fmul_rn(__fsub_rn(rb, ra), frac));
                ring[wp] = __fadd_rn(v, __fmul_rn(feedback, rd));
                wp = wp == 0 ? kRing - 1 : wp - 1;
                const float wet = svf_tick(rd, w_low, w_band, kw);
                v = dry_wet(wet, v, d_balance);
            }
            if (KIND == OLB_FX_REVERB || KIND == OLB_FX_RACK) v = dry_wet(__fmul_rn(v, 0.8f), v, r_balance);
            if (KIND == OLB_FX_FILTER || KIND == OLB_FX_RACK) v = svf_tick(v, f_low, f_band, kf);
            if (KIND == OLB_FX_RACK) v = __fmul_rn(v, master);
            dst[(size_t) (n0 + j) * I] = v;
        }
    };

    fetch(x0, a0, 0);
    for (uint32_t n0 = 0; n0 < frames; n0 += 2 * B) {
        fetch(x1, a1, n0 + B);
        run(x0, a0, n0);
        fetch(x0, a0, n0 + 2 * B);
        run(x1, a1, n0 + B);
    }
    svf[inst] = f_low;
    svf[(size_t) I + inst] = f_band;
    svf[2 * (size_t) I + inst] = w_low;
    svf[3 * (size_t) I + inst] = w_band;
}

}  // namespace olb

using namespace olb;

// ================================================================================================
// Host side: one host::RackParams per instance carries the reference objects' members; the device table is rebuilt
// from it when a setter ran (setters take effect at the next Process, as Update() does in the reference).
struct olb_fx {
    int device = 0, kind = 0;
    uint32_t I = 0;
    float sample_rate = 48000.f;
    std::vector<host::RackParams> units;
    bool dirty = true;
    uint32_t wp = 0;  // DelayLine::write_ptr_, shared by the bank
    float *d_par = nullptr, *d_svf = nullptr, *d_ring = nullptr;
    std::vector<float> stage;
};

static float bits_f(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static int fx_upload(olb_fx *h, cudaStream_t st) {
    const size_t I = h->I;
    h->stage.resize((size_t) P_COUNT * I);
    float *s = h->stage.data();
    for (size_t i = 0; i < I; i++) {
        const host::RackParams &u = h->units[i];
        // a stand-alone DelayFx / FilterFx bank uses the same members as the rack's delay_ / filter1
        s[P_F_FREQ * I + i] = u.filter.svf.freq; s[P_F_DAMP * I + i] = u.filter.svf.damp; s[P_F_DRIVE * I + i] = u.filter.svf.drive;
        s[P_F_TYPE * I + i] = bits_f((uint32_t) u.filter.type);
        s[P_W_FREQ * I + i] = u.delay.wet.svf.freq; s[P_W_DAMP * I + i] = u.delay.wet.svf.damp; s[P_W_DRIVE * I + i] = u.delay.wet.svf.drive;
        s[P_W_TYPE * I + i] = bits_f((uint32_t) u.delay.wet.type);
        s[P_D_FEEDBACK * I + i] = u.delay.feedback; s[P_D_BALANCE * I + i] = u.delay.balance;
        s[P_D_DELAY * I + i] = bits_f(u.delay.delay); s[P_D_FRAC * I + i] = u.delay.frac;
        s[P_R_BALANCE * I + i] = u.reverb.balance;
        s[P_MASTER * I + i] = u.master_volume;
    }
    OLB_CUDA(cudaMemcpyAsync(h->d_par, s, h->stage.size() * sizeof(float), cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaStreamSynchronize(st));
    h->dirty = false;
    return OLB_OK;
}

static bool fx_has_ring(int kind) { return kind == OLB_FX_DELAY || kind == OLB_FX_RACK; }

// Init(sample_rate) of every instance (Fx.h:83-86, 183-192, 287-290, 416-424): state cleared, members kept.
static int fx_init(olb_fx *h, float sr) {
    h->sample_rate = sr;
    for (auto &u : h->units) {
        switch (h->kind) {
            case OLB_FX_FILTER: u.filter.Init(sr); break;
            case OLB_FX_DELAY: u.delay.Init(sr); break;
            case OLB_FX_REVERB: break;  // DaisyVerb::Init only stores the sample rate (Reverb.h:22-24)
            default: u.Init(sr); break;
        }
    }
    OLB_CUDA(cudaMemset(h->d_svf, 0, (size_t) 4 * h->I * sizeof(float)));
    if (h->d_ring) OLB_CUDA(cudaMemset(h->d_ring, 0, (size_t) h->I * kRing * sizeof(float)));
    h->wp = 0;
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_fx_create(int device, int kind, uint32_t instances, float sample_rate, olb_fx **out) {
    OLB_ARG(out, "olb_fx_create: out is NULL");
    *out = nullptr;
    OLB_ARG(kind >= OLB_FX_FILTER && kind <= OLB_FX_RACK, "olb_fx_create: unknown kind %d", kind);
    OLB_ARG(instances > 0, "olb_fx_create: instances must be > 0");
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_fx *h = new olb_fx();
    h->device = device;
    h->kind = kind;
    h->I = instances;
    h->units.resize(instances);
    cudaError_t e = cudaMalloc(&h->d_par, (size_t) P_COUNT * instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_svf, (size_t) 4 * instances * sizeof(float));
    if (e == cudaSuccess && fx_has_ring(kind)) e = cudaMalloc(&h->d_ring, (size_t) instances * kRing * sizeof(float));
    if (e != cudaSuccess) {
        cudaFree(h->d_par); cudaFree(h->d_svf); cudaFree(h->d_ring);
        delete h;
        return fail_cuda(e, "allocating fx bank", __FILE__, __LINE__);
    }
    rc = fx_init(h, sample_rate);
    if (rc) {
        cudaFree(h->d_par); cudaFree(h->d_svf); cudaFree(h->d_ring);
        delete h;
        return rc;
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_fx_destroy(olb_fx *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    cudaFree(h->d_par); cudaFree(h->d_svf); cudaFree(h->d_ring);
    delete h;
    return OLB_OK;
}

OLB_API int olb_fx_init(olb_fx *h, float sample_rate) {
    OLB_ARG(h, "olb_fx_init: NULL handle");
    DeviceGuard g(h->device);
    return fx_init(h, sample_rate);
}

OLB_API int olb_fx_update_midi(olb_fx *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value) {
    OLB_ARG(h, "olb_fx_update_midi: NULL handle");
    OLB_ARG((uint64_t) first + count <= h->I, "olb_fx_update_midi: range [%u, %u) exceeds %u instances", first, first + count, h->I);
    for (uint32_t i = first; i < first + count; i++) {
        host::RackParams &u = h->units[i];
        switch (h->kind) {
            case OLB_FX_FILTER: u.filter.UpdateMidiControl(control, value); break;
            case OLB_FX_DELAY: u.delay.UpdateMidiControl(control, value); break;
            case OLB_FX_REVERB: u.reverb.UpdateMidiControl(control, value); break;
            default: u.UpdateMidiControl(control, value); break;
        }
    }
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_fx_update_hw(olb_fx *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride) {
    OLB_ARG(h && values_host, "olb_fx_update_hw: NULL argument");
    OLB_ARG((uint64_t) first + count <= h->I, "olb_fx_update_hw: range [%u, %u) exceeds %u instances", first, first + count, h->I);
    for (uint32_t i = 0; i < count; i++) {
        host::RackParams &u = h->units[first + i];
        const float value = values_host[(size_t) i * stride];
        switch (h->kind) {
            case OLB_FX_FILTER: u.filter.UpdateHardwareControl(control, value); break;
            case OLB_FX_DELAY: u.delay.UpdateHardwareControl(control, value); break;
            case OLB_FX_REVERB: u.reverb.UpdateHardwareControl(control, value); break;
            default: u.UpdateHardwareControl(control, value); break;
        }
    }
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_fx_process(olb_fx *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_fx_process: NULL handle");
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev, "olb_fx_process: NULL buffer");
    OLB_ARG(frames < (1ull << 32), "olb_fx_process: at most 2^32-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    if (h->dirty) {
        int rc = fx_upload(h, st);
        if (rc) return rc;
    }
    const uint32_t F = (uint32_t) frames;
    // OLB_FX_KERNEL=simple selects the one-frame-at-a-time kernel; OLB_FX_LANES=32|16|8 overrides the active lanes per warp
    // (default: the widest that still leaves >= 2 warps for each of the 592 SM sub-partitions, down to 8)
    const int forced = [] {
        const char *e = getenv("OLB_FX_KERNEL");
        if (e && !strcmp(e, "simple")) return -1;
        const char *l = getenv("OLB_FX_LANES");
        return l ? atoi(l) : 0;
    }();
    if (forced < 0) {
        const unsigned grid = (h->I + 63) / 64;
        switch (h->kind) {
            case OLB_FX_FILTER: k_fx<OLB_FX_FILTER><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp); break;
            case OLB_FX_DELAY: k_fx<OLB_FX_DELAY><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp); break;
            case OLB_FX_REVERB: k_fx<OLB_FX_REVERB><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp); break;
            default: k_fx<OLB_FX_RACK><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp); break;
        }
    } else {
        const int lanes = (forced == 32 || forced == 16 || forced == 8) ? forced : (h->I >= 1184u * 32 ? 32 : (h->I >= 1184u * 16 ? 16 : 8));
        const unsigned warps = (h->I + lanes - 1) / lanes, grid = (warps + 1) / 2;
#define OLB_FX_GO(KIND, LN) k_fx_pf<KIND, LN><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp)
#define OLB_FX_LN(KIND) \
    do { if (lanes == 32) OLB_FX_GO(KIND, 32); else if (lanes == 16) OLB_FX_GO(KIND, 16); else OLB_FX_GO(KIND, 8); } while (0)
        switch (h->kind) {
            case OLB_FX_FILTER: OLB_FX_LN(OLB_FX_FILTER); break;
            case OLB_FX_DELAY: OLB_FX_LN(OLB_FX_DELAY); break;
            case OLB_FX_REVERB: OLB_FX_LN(OLB_FX_REVERB); break;
            default: OLB_FX_LN(OLB_FX_RACK); break;
        }
#undef OLB_FX_LN
#undef OLB_FX_GO
    }
    OLB_LAUNCHED();
    if (fx_has_ring(h->kind)) h->wp = (uint32_t) ((h->wp + (uint64_t) kRing - frames % kRing) % kRing);
    return OLB_OK;
}
