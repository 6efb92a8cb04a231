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

#include <type_traits>

#include "common.cuh"
#include "host_params.h"
#include "tmap.h"

namespace olb {

constexpr int kRing = host::kMaxDelay;
enum { P_F_FREQ, P_F_DAMP, P_F_DRIVE, P_F_TYPE, P_W_FREQ, P_W_DAMP, P_W_DRIVE, P_W_TYPE, P_D_FEEDBACK, P_D_BALANCE, P_D_DELAY, P_D_FRAC,
       P_R_BALANCE, P_MASTER, P_COUNT };

struct SvfCoef {
    float freq, damp, drive;
    int type;
};

// The arithmetic tier rides in the KIND template parameter: KIND | kFxFastBit = the FUSED tier of that kind
// (olb_fx_set_mode(OLB_MODE_FAST); specification: oracle/olb_oracle.cpp svf_half / ring_read / delay_tick with g_fx_fast): the
// same formulas with their multiply-adds fused -- the SVF's recurrence goes from 6 dependent operations per pass to 4, its
// operation count from 11 to 7 -- a tolerance tier (north_star: fused variants state their tolerance; measured in the tests).
// The default is the reference's separately rounded arithmetic.
constexpr int kFxFastBit = 8;
#define OLB_K(KIND) ((KIND) & 7)
#define OLB_F(KIND) (((KIND) & olb::kFxFastBit) != 0)

// daisysp::Svf::Process, two identical passes; returns the output FilterFx::Process selects (Fx.h:88-108).
// Branch-free: the per-lane filter type picks its tap with selects (a switch here puts a divergent region in every
// frame and keeps ptxas from overlapping consecutive frames); each output is 0.5 * pass1 + 0.5 * pass2 as in Svf.
__device__ __forceinline__ float svf_pick(int type, float low, float band, float high, float notch) {
    const float peak = __fsub_rn(low, high);
    float s = low;
    s = type == 1 ? band : s;
    s = type == 2 ? high : s;
    s = type == 3 ? notch : s;
    s = type == 4 ? peak : s;
    return s;
}
// one pass of the recurrence
template <bool F>
__device__ __forceinline__ void svf_half(float in, float &low, float &band, const SvfCoef &k, float &notch, float &high) {
    if (F) {
        notch = __fmaf_rn(-k.damp, band, in);
        low = __fmaf_rn(k.freq, band, low);
        high = __fsub_rn(notch, low);
        const float t = __fmaf_rn(k.freq, high, band);
        const float c = __fmul_rn(__fmul_rn(k.drive, band), band);
        band = __fmaf_rn(-c, band, t);
    } else {
        notch = __fsub_rn(in, __fmul_rn(k.damp, band));
        low = __fadd_rn(low, __fmul_rn(k.freq, band));
        high = __fsub_rn(notch, low);
        band = __fsub_rn(__fadd_rn(__fmul_rn(k.freq, high), band), __fmul_rn(__fmul_rn(__fmul_rn(k.drive, band), band), band));
    }
}
template <bool F>
__device__ __forceinline__ float svf_half_sum(float o1, float pick2) {  // 0.5 * pass 1 (already halved) + 0.5 * pass 2
    return F ? __fmaf_rn(0.5f, pick2, o1) : __fadd_rn(o1, __fmul_rn(0.5f, pick2));
}
template <bool F>
__device__ __forceinline__ float svf_tick(float in, float &low, float &band, const SvfCoef &k) {
    float notch, high;
    svf_half<F>(in, low, band, k, notch, high);
    const float o1 = __fmul_rn(0.5f, svf_pick(k.type, low, band, high, notch));
    svf_half<F>(in, low, band, k, notch, high);
    return svf_half_sum<F>(o1, svf_pick(k.type, low, band, high, notch));
}

// The same filter with the output tap fixed at compile time (T = 0..4; T < 0: per-lane selects as above).  A bank whose
// instances all use one filter type -- the usual case -- saves the 8 selects and the 2 peak subtractions of a frame.
template <int T>
__device__ __forceinline__ float svf_pick_t(int type, float low, float band, float high, float notch) {
    if (T < 0) return svf_pick(type, low, band, high, notch);
    return T == 1 ? band : T == 2 ? high : T == 3 ? notch : T == 4 ? __fsub_rn(low, high) : low;
}
template <int T, bool F>
__device__ __forceinline__ float svf_tick_t(float in, float &low, float &band, const SvfCoef &k) {
    float notch, high;
    svf_half<F>(in, low, band, k, notch, high);
    const float o1 = __fmul_rn(0.5f, svf_pick_t<T>(k.type, low, band, high, notch));
    svf_half<F>(in, low, band, k, notch, high);
    return svf_half_sum<F>(o1, svf_pick_t<T>(k.type, low, band, high, notch));
}

// Two independent SVFs advanced together, statement by statement: the same operations as two svf_tick calls, written
// interleaved because ptxas keeps close to source order and each filter alone is one dependent chain.
template <bool F>
__device__ __forceinline__ void svf_tick_pair(float inA, float &lowA, float &bandA, const SvfCoef &kA, float &outA, float inB, float &lowB,
                                              float &bandB, const SvfCoef &kB, float &outB) {
    if (F) {
        float nA, hA, nB, hB;
        svf_half<true>(inA, lowA, bandA, kA, nA, hA);
        svf_half<true>(inB, lowB, bandB, kB, nB, hB);
        const float o1A = __fmul_rn(0.5f, svf_pick(kA.type, lowA, bandA, hA, nA)), o1B = __fmul_rn(0.5f, svf_pick(kB.type, lowB, bandB, hB, nB));
        svf_half<true>(inA, lowA, bandA, kA, nA, hA);
        svf_half<true>(inB, lowB, bandB, kB, nB, hB);
        outA = __fmaf_rn(0.5f, svf_pick(kA.type, lowA, bandA, hA, nA), o1A);
        outB = __fmaf_rn(0.5f, svf_pick(kB.type, lowB, bandB, hB, nB), o1B);
        return;
    }
    float dA = __fmul_rn(kA.damp, bandA), dB = __fmul_rn(kB.damp, bandB);
    float gA = __fmul_rn(kA.freq, bandA), gB = __fmul_rn(kB.freq, bandB);
    float cA = __fmul_rn(__fmul_rn(__fmul_rn(kA.drive, bandA), bandA), bandA), cB = __fmul_rn(__fmul_rn(__fmul_rn(kB.drive, bandB), bandB), bandB);
    float notchA = __fsub_rn(inA, dA), notchB = __fsub_rn(inB, dB);
    lowA = __fadd_rn(lowA, gA); lowB = __fadd_rn(lowB, gB);
    float highA = __fsub_rn(notchA, lowA), highB = __fsub_rn(notchB, lowB);
    float tA = __fmul_rn(kA.freq, highA), tB = __fmul_rn(kB.freq, highB);
    tA = __fadd_rn(tA, bandA); tB = __fadd_rn(tB, bandB);
    bandA = __fsub_rn(tA, cA); bandB = __fsub_rn(tB, cB);
    const float o1A = __fmul_rn(0.5f, svf_pick(kA.type, lowA, bandA, highA, notchA)), o1B = __fmul_rn(0.5f, svf_pick(kB.type, lowB, bandB, highB, notchB));
    dA = __fmul_rn(kA.damp, bandA); dB = __fmul_rn(kB.damp, bandB);
    gA = __fmul_rn(kA.freq, bandA); gB = __fmul_rn(kB.freq, bandB);
    cA = __fmul_rn(__fmul_rn(__fmul_rn(kA.drive, bandA), bandA), bandA); cB = __fmul_rn(__fmul_rn(__fmul_rn(kB.drive, bandB), bandB), bandB);
    notchA = __fsub_rn(inA, dA); notchB = __fsub_rn(inB, dB);
    lowA = __fadd_rn(lowA, gA); lowB = __fadd_rn(lowB, gB);
    highA = __fsub_rn(notchA, lowA); highB = __fsub_rn(notchB, lowB);
    tA = __fmul_rn(kA.freq, highA); tB = __fmul_rn(kB.freq, highB);
    tA = __fadd_rn(tA, bandA); tB = __fadd_rn(tB, bandB);
    bandA = __fsub_rn(tA, cA); bandB = __fsub_rn(tB, cB);
    outA = __fadd_rn(o1A, __fmul_rn(0.5f, svf_pick(kA.type, lowA, bandA, highA, notchA)));
    outB = __fadd_rn(o1B, __fmul_rn(0.5f, svf_pick(kB.type, lowB, bandB, highB, notchB)));
}

template <bool F>
__device__ __forceinline__ float dry_wet(float wet, float dry, float balance) {  // Fx.h:205-207, 297-299
    const float d = __fmul_rn(dry, __fsub_rn(1.0f, balance));
    return F ? __fmaf_rn(wet, balance, d) : __fadd_rn(__fmul_rn(wet, balance), d);
}
template <bool F>
__device__ __forceinline__ float dl_read(float a, float b, float frac) {  // DelayLine::Read: a + (b - a) * frac_ (Fx.h:198)
    return F ? __fmaf_rn(__fsub_rn(b, a), frac, a) : __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), frac));
}
template <bool F>
__device__ __forceinline__ float dl_write(float x, float feedback, float rd) {  // Write(in + feedback * read) (Fx.h:199-200)
    return F ? __fmaf_rn(feedback, rd, x) : __fadd_rn(x, __fmul_rn(feedback, rd));
}

template <int KIND>
__global__ void __launch_bounds__(64) k_fx(const float *in, float *out, uint32_t frames, uint32_t I,
                                           const float *__restrict__ par, float *svf, float *ring_all, uint32_t wp0) {
    constexpr bool FM = OLB_F(KIND);
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
        if (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK) {
            // DelayLine::Read then Write (Fx.h:198-200); the write pointer runs backwards
            uint32_t ia = wp + delay;
            ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
            uint32_t ib = ia + 1;
            ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
            const float a = ring[ia], b = ring[ib];
            const float rd = dl_read<FM>(a, b, frac);
            ring[wp] = dl_write<FM>(v, feedback, rd);
            wp = wp == 0 ? kRing - 1 : wp - 1;
            const float wet = svf_tick<FM>(rd, w_low, w_band, kw);
            v = dry_wet<FM>(wet, v, d_balance);
        }
        if (OLB_K(KIND) == OLB_FX_REVERB || OLB_K(KIND) == OLB_FX_RACK) v = dry_wet<FM>(__fmul_rn(v, 0.8f), v, r_balance);
        if (OLB_K(KIND) == OLB_FX_FILTER || OLB_K(KIND) == OLB_FX_RACK) v = svf_tick<FM>(v, f_low, f_band, kf);
        if (OLB_K(KIND) == OLB_FX_RACK) v = __fmul_rn(v, master);
        out[(size_t) n * I + inst] = v;
    }
    svf[inst] = f_low;
    svf[(size_t) I + inst] = f_band;
    svf[2 * (size_t) I + inst] = w_low;
    svf[3 * (size_t) I + inst] = w_band;
}

// ------------------------------------------------------------------------------------------------
// One frame of a unit, branch-free (ring kinds: `ra`, `rb` are the two DelayLine taps already read; the value to write
// back into the ring is returned through `wr`).
template <int KIND>
__device__ __forceinline__ float fx_frame(float v, float ra, float rb, float &wr, float frac, float feedback, float d_balance,
                                          float r_balance, float master, float &w_low, float &w_band, const SvfCoef &kw,
                                          float &f_low, float &f_band, const SvfCoef &kf) {
    constexpr bool FM = OLB_F(KIND);
    if (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK) {
        const float rd = dl_read<FM>(ra, rb, frac);  // DelayLine::Read (Fx.h:198)
        wr = dl_write<FM>(v, feedback, rd);                          // Write(in + feedback * read) (Fx.h:199-200)
        const float wet = svf_tick<FM>(rd, w_low, w_band, kw);
        v = dry_wet<FM>(wet, v, d_balance);
    }
    if (OLB_K(KIND) == OLB_FX_REVERB || OLB_K(KIND) == OLB_FX_RACK) v = dry_wet<FM>(__fmul_rn(v, 0.8f), v, r_balance);
    if (OLB_K(KIND) == OLB_FX_FILTER || OLB_K(KIND) == OLB_FX_RACK) v = svf_tick<FM>(v, f_low, f_band, kf);
    if (OLB_K(KIND) == OLB_FX_RACK) v = __fmul_rn(v, master);
    return v;
}

// A batch of N frames of one instance out of shared memory, written stage by stage over register arrays so that ptxas sees
// the independent streams of one basic block: (A) every input, tap and ring write-back value of the batch, (B) the wet
// filter's recurrence, (C) the mixes, (D) the main filter's recurrence.  Only B and D are serial (12 dependent FP32 ops
// per frame each), they are independent of each other (D lags B by one frame), and A / C fill their stall slots; written
// frame by frame the compiler runs everything in sequence (the next tap load cannot move above the previous frame's
// shared-memory store) and a delay frame costs 2.3x the recurrence.  Same operations, same order per value: bit-exact.
// col: this lane's column of the audio tile (stride kFxLanes); tap: its first tap (walks down); wcol: its write-back column.
constexpr int kFxLanes = 32;
template <int KIND, int N>
__device__ __forceinline__ void fx_batch(float *col, const float *tap, float *wcol, int wstride, float &prev_a, float frac, float feedback,
                                         float d_balance, float r_balance, float master, float &w_low, float &w_band, const SvfCoef &kw,
                                         float &f_low, float &f_band, const SvfCoef &kf) {
    constexpr bool kHasRing = (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK);
    constexpr bool FM = OLB_F(KIND);
    float x[N], rd[N], v[N];
#pragma unroll
    for (int r = 0; r < N; r++) x[r] = col[r * kFxLanes];
    if (kHasRing) {
        float ra[N];
#pragma unroll
        for (int r = 0; r < N; r++) ra[r] = *(tap - r);
#pragma unroll
        for (int r = 0; r < N; r++) {
            const float rb = r == 0 ? prev_a : ra[r - 1];
            rd[r] = dl_read<FM>(ra[r], rb, frac);     // DelayLine::Read (Fx.h:198)
            wcol[r * wstride] = dl_write<FM>(x[r], feedback, rd[r]);    // Write(in + feedback * read) (Fx.h:199-200)
        }
        prev_a = ra[N - 1];
        if (OLB_K(KIND) == OLB_FX_RACK) {
            // the two recurrences side by side, the main filter one frame behind the wet filter: in source order frame r of
            // chain B sits next to frame r-1 of chain D, which is where ptxas's in-order schedule needs them to overlap
            float u = 0.f;  // main filter input of frame r-1
#pragma unroll
            for (int r = 0; r <= N; r++) {
                float wet = 0.f, y = 0.f;
                if (r > 0 && r < N) svf_tick_pair<FM>(rd[r], w_low, w_band, kw, wet, u, f_low, f_band, kf, y);
                else if (r < N) wet = svf_tick<FM>(rd[r], w_low, w_band, kw);
                else y = svf_tick<FM>(u, f_low, f_band, kf);
                if (r > 0) v[r - 1] = y;
                if (r < N) {
                    u = dry_wet<FM>(wet, x[r], d_balance);
                    u = dry_wet<FM>(__fmul_rn(u, 0.8f), u, r_balance);
                }
            }
        } else {
#pragma unroll
            for (int r = 0; r < N; r++) v[r] = svf_tick<FM>(rd[r], w_low, w_band, kw);
#pragma unroll
            for (int r = 0; r < N; r++) v[r] = dry_wet<FM>(v[r], x[r], d_balance);
        }
    } else {
#pragma unroll
        for (int r = 0; r < N; r++) v[r] = x[r];
    }
    if (OLB_K(KIND) == OLB_FX_REVERB) {
#pragma unroll
        for (int r = 0; r < N; r++) v[r] = dry_wet<FM>(__fmul_rn(v[r], 0.8f), v[r], r_balance);
    }
    if (OLB_K(KIND) == OLB_FX_FILTER) {
#pragma unroll
        for (int r = 0; r < N; r++) v[r] = svf_tick<FM>(v[r], f_low, f_band, kf);
    }
#pragma unroll
    for (int r = 0; r < N; r++) col[r * kFxLanes] = OLB_K(KIND) == OLB_FX_RACK ? __fmul_rn(v[r], master) : v[r];
}

// The ring kinds cut into pipeline stages for k_fx_split, one warp each (same operations, same order per value):
//   io_batch    DelayLine::Read + Write: taps -> interpolated read `rd` (to the rd tile) and the value written back
//   wet_batch   the wet filter, the delay's dry/wet mix and (rack) the reverb slot's mix, in place in the audio tile
//   tail_batch  (rack) the main filter and the master volume, in place
template <int N, bool FM>
__device__ __forceinline__ void io_batch(const float *col, const float *tap, float *wcol, int wstride, float *rdcol, float &prev_a, float frac,
                                         float feedback) {
    float x[N], ra[N];
#pragma unroll
    for (int r = 0; r < N; r++) x[r] = col[r * kFxLanes];
#pragma unroll
    for (int r = 0; r < N; r++) ra[r] = *(tap - r);
#pragma unroll
    for (int r = 0; r < N; r++) {
        const float rb = r == 0 ? prev_a : ra[r - 1];
        const float rd = dl_read<FM>(ra[r], rb, frac);  // DelayLine::Read (Fx.h:198)
        rdcol[r * kFxLanes] = rd;
        wcol[r * wstride] = dl_write<FM>(x[r], feedback, rd);              // Write(in + feedback * read) (Fx.h:199-200)
    }
    prev_a = ra[N - 1];
}

template <int N, bool RACK, bool FM, int T = -1>
__device__ __forceinline__ void wet_batch(float *col, const float *rdcol, float d_balance, float r_balance, float &w_low, float &w_band,
                                          const SvfCoef &kw) {
    float x[N], v[N];
#pragma unroll
    for (int r = 0; r < N; r++) x[r] = col[r * kFxLanes];
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = rdcol[r * kFxLanes];
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = svf_tick_t<T, FM>(v[r], w_low, w_band, kw);  // T < 0: the tap picked per lane
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = dry_wet<FM>(v[r], x[r], d_balance);                              // Fx.h:205-207
    if (RACK) {
#pragma unroll
        for (int r = 0; r < N; r++) v[r] = dry_wet<FM>(__fmul_rn(v[r], 0.8f), v[r], r_balance);         // Fx.h:294-300, stub reverb
    }
#pragma unroll
    for (int r = 0; r < N; r++) col[r * kFxLanes] = v[r];
}

template <int N, bool FM, int T = -1>
__device__ __forceinline__ void tail_batch(float *col, float master, float &f_low, float &f_band, const SvfCoef &kf) {
    float v[N];
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = col[r * kFxLanes];
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = svf_tick_t<T, FM>(v[r], f_low, f_band, kf);                      // Fx.h:426-434
#pragma unroll
    for (int r = 0; r < N; r++) col[r * kFxLanes] = __fmul_rn(v[r], master);
}

// ------------------------------------------------------------------------------------------------
// k_fx_pf: register double buffering -- the inputs (and, for the delay kinds, the ring reads) of frames n+B .. n+2B-1
// are in flight while frames n .. n+B-1 run (in and out may alias, so the compiler will not hoist the loads itself).
// Fallback of k_fx_tma for banks whose shape or alignment rules out a tensor map.  Ring reads run ahead only when the
// delay exceeds the look-ahead (2B + 2 samples); shorter delays take the direct path, lane by lane.
constexpr int kFxB = 8;

template <int KIND>
__global__ void __launch_bounds__(64) k_fx_pf(const float *in, float *out, uint32_t frames, uint32_t I,
                                              const float *__restrict__ par, float *svf, float *ring_all, uint32_t wp0) {
    constexpr bool kHasRing = (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK);
    constexpr bool FM = OLB_F(KIND);
    constexpr int B = kFxB;
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
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
            float ra = 0.f, rb = 0.f, wr = 0.f;
            if (kHasRing) {
                if (far) {
                    ra = a[j];
                    rb = prev_a;
                    prev_a = ra;
                } else {  // DelayLine::Read straight from the ring
                    uint32_t ia = wp + delay;
                    ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
                    uint32_t ib = ia + 1;
                    ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
                    ra = ring[ia];
                    rb = ring[ib];
                }
            }
            const float v = fx_frame<KIND>(x[j], ra, rb, wr, frac, feedback, d_balance, r_balance, master, w_low, w_band, kw, f_low, f_band, kf);
            if (kHasRing) {
                ring[wp] = wr;
                wp = wp == 0 ? kRing - 1 : wp - 1;
            }
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

// ------------------------------------------------------------------------------------------------
// k_fx_tma: the streaming kernel.  Every instance is one lane running one dependent chain, so a bank's speed is
// (instances x clock) / (cycles per frame of ONE warp) -- narrower warps buy nothing (measured: 32 / 16 / 8 active lanes
// tie, profiles/r01_run15_fx_tma_sweep.jsonl); what counts is bytes in flight and a tight, branch-free frame.
// k_fx_pf keeps only 8-16 frames x 4 B per instance in flight -- 8 192 instances are ~0.3 MB on the wire, and Little's
// law caps the bank near 0.5 TB/s.  Here ONE warp is one CTA with its own TMA pipeline (the k_bq_tma pattern):
// [TB frames][32 instances] boxes by cp.async.bulk.tensor.2d into a STAGES-deep shared-memory ring with mbarrier
// complete_tx, filtered in place (each lane walks its own column), stored by TMA; (STAGES-1) x TB frames per instance
// are in flight.  The delay kinds also stream their rings ([instances][48 000], one contiguous line per instance because
// delay times differ per instance) COALESCED: the taps of a future tile are one contiguous run per instance, copied by
// its lane as 16-byte cp.async chunks into a per-instance row of a tap tile; the tile's new samples are staged in shared
// memory and written back as float4s.  (4-byte accesses cost one instruction and 32 L1 wavefronts per frame each way and
// held the delay kinds at 1.0e11 instance-samples/s at every bank size.)  This needs every live lane's delay to exceed the look-ahead (STAGES x TB + 2 frames); any other
// warp runs the general loop, direct per-lane reads and writes.
__device__ __forceinline__ void fx_cp_async4(float *smem_dst, const float *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void fx_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void fx_cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ void fx_cp_async16(float *smem_dst, const float *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}

constexpr int kFxL = kFxLanes;  // instances per warp / CTA
constexpr int kFxBatch = 8;    // frames per fx_batch

// floats of shared memory in front of the mbarriers: audio tiles, and for the ring kinds the tap tiles + the write-back tile
template <int TB, int STAGES>
__host__ __device__ constexpr size_t fx_tma_smem_floats(bool has_ring) {
    return ((size_t) STAGES * TB * kFxL + (has_ring ? (size_t) STAGES * kFxL * (TB + 4) + (size_t) TB * kFxL : 0) + 1) & ~(size_t) 1;
}

template <int KIND, int TB, int STAGES>
__global__ void __launch_bounds__(32) k_fx_tma(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                               uint32_t frames, uint32_t I, const float *__restrict__ par, float *svf, float *ring_all,
                                               uint32_t wp0) {
    constexpr bool kHasRing = (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK);
    constexpr bool FM = OLB_F(KIND);
    constexpr int L = kFxL;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t lane = threadIdx.x, i0 = blockIdx.x * L, inst = i0 + lane;
    const bool live = inst < I;
    constexpr int RS = TB + 4;   // floats per instance row of a ring tile: TB taps + alignment slack, 16-byte multiple
    constexpr int WS = L;        // row of the write-staging tile (each lane owns a column)
    float *tiles = reinterpret_cast<float *>(smem_raw);                    // [STAGES][TB][L]
    float *rtiles = tiles + (size_t) STAGES * TB * L;                      // [STAGES][L][RS]  ring kinds only: delay-line taps
    float *wtile = rtiles + (size_t) STAGES * L * RS;                      // [TB][WS]         ring kinds only: values to write back
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + fx_tma_smem_floats<TB, STAGES>(kHasRing) * sizeof(float));
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncwarp();
    constexpr uint32_t kTileBytes = TB * L * sizeof(float);
    const uint32_t ntiles = (frames + TB - 1) / TB;

    const uint32_t ci = live ? inst : 0;  // idle lanes read instance 0's parameters and never write
    auto P = [&](int p) { return par[(size_t) p * I + ci]; };
    const SvfCoef kf{P(P_F_FREQ), P(P_F_DAMP), P(P_F_DRIVE), (int) __float_as_uint(P(P_F_TYPE))};
    const SvfCoef kw{P(P_W_FREQ), P(P_W_DAMP), P(P_W_DRIVE), (int) __float_as_uint(P(P_W_TYPE))};
    const float feedback = P(P_D_FEEDBACK), d_balance = P(P_D_BALANCE), frac = P(P_D_FRAC);
    const uint32_t delay = __float_as_uint(P(P_D_DELAY));
    const float r_balance = P(P_R_BALANCE), master = P(P_MASTER);
    float f_low = svf[ci], f_band = svf[(size_t) I + ci], w_low = svf[2 * (size_t) I + ci], w_band = svf[3 * (size_t) I + ci];
    float *ring = kHasRing ? ring_all + (size_t) ci * kRing : nullptr;
    // one filter type for the whole warp (idle lanes carry instance 0's)?  then the tap is picked at compile time
    const int type0 = __shfl_sync(0xffffffffu, kf.type, 0);
    const int uniform_type = (__all_sync(0xffffffffu, kf.type == type0 || !live) && type0 >= 0 && type0 <= 4) ? type0 : -1;
    const bool far = kHasRing && live && delay >= (uint32_t) (STAGES * TB + 2);
    // Ring traffic is coalesced only when the whole warp can run ahead (every live lane's delay exceeds the look-ahead);
    // any other warp takes the general loop: per-lane direct reads and writes, no look-ahead.
    const bool all_far = kHasRing && __all_sync(0xffffffffu, far || !live);
    uint32_t wp = wp0, la = 0, lc = 0;  // wp is the same in every lane (the bank shares DelayLine::write_ptr_)
    float prev_a = 0.f;                 // la / lc: tap index of the next frame to be fetched / computed
    if (kHasRing) {
        la = wp0 + delay;
        la = la >= (uint32_t) kRing ? la - kRing : la;
        lc = la;
        if (all_far && live) {
            const uint32_t ib = la + 1 >= (uint32_t) kRing ? la + 1 - kRing : la + 1;
            prev_a = ring[ib];
        }
    }
    // Taps of a future tile.  This lane's TB taps are one contiguous run of its ring, la - TB + 1 .. la (the read index
    // walks backwards): copied as 16-byte chunks (cp.async.cg 16 -> LDGSTS.128) into this lane's row of the tap tile, 17
    // instructions per tile instead of 64 4-byte ones.  Rings are 48 000 floats (a multiple of 4) and 16-byte aligned,
    // so a chunk never straddles the wrap.
    auto fetch_ring = [&](uint32_t tile, uint32_t stage) {  // every lane; one commit group per tile, empty or not
        if (kHasRing) {
            if (all_far && live && tile < ntiles) {
                const uint32_t rows = min((uint32_t) TB, frames - tile * TB);
                float *dst = rtiles + (size_t) stage * L * RS + lane * RS;
                const int base4 = ((int) la - (TB - 1)) & ~3;  // floor to a multiple of 4, also below zero
#pragma unroll
                for (int c = 0; c < RS / 4; c++) {
                    int idx = base4 + 4 * c;
                    if (idx <= (int) la) {
                        idx = idx < 0 ? idx + kRing : idx;
                        fx_cp_async16(dst + 4 * c, ring + idx);
                    }
                }
                la = la >= rows ? la - rows : la + kRing - rows;
            }
            fx_cp_async_commit();
        }
    };

    for (uint32_t j = 0; j < (uint32_t) (STAGES - 1); j++) {
        if (lane == 0 && j < ntiles) {
            mbar_arrive_expect_tx(&full[j], kTileBytes);
            tma_load_2d(tiles + (size_t) j * TB * L, &tm_in, (int) i0, (int) (j * TB), &full[j]);
        }
        fetch_ring(j, j);
    }

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % STAGES, ph = (i / STAGES) & 1;
        if (kHasRing) {
            fx_cp_async_wait<STAGES - 2>();  // each lane reads only the row it copied itself
        }
        mbar_wait(&full[st], ph);
        const uint32_t rows = min((uint32_t) TB, frames - i * TB);
        float *col = tiles + (size_t) st * TB * L + lane;
        if (!kHasRing) {
            if (OLB_K(KIND) == OLB_FX_FILTER && uniform_type >= 0) {  // FilterFx::Process is the filter alone (Fx.h:88-108)
                if (live) {
                    switch (uniform_type) {
#define OLB_FX_TYPE_CASE(T_)                                                                             \
    case T_:                                                                                             \
        _Pragma("unroll 4") for (uint32_t r = 0; r < rows; r++) col[r * L] = svf_tick_t<T_, FM>(col[r * L], f_low, f_band, kf); \
        break;
                        OLB_FX_TYPE_CASE(0)
                        OLB_FX_TYPE_CASE(1)
                        OLB_FX_TYPE_CASE(2)
                        OLB_FX_TYPE_CASE(3)
                        OLB_FX_TYPE_CASE(4)
#undef OLB_FX_TYPE_CASE
                    }
                }
            } else if (live) {
#pragma unroll 4
                for (uint32_t r = 0; r < rows; r++) {
                    float wr;
                    col[r * L] = fx_frame<KIND>(col[r * L], 0.f, 0.f, wr, frac, feedback, d_balance, r_balance, master, w_low, w_band, kw,
                                                f_low, f_band, kf);
                }
            }
        } else if (all_far) {
            if (live) {
                // frame r's tap a sits at virtual ring index lc - r, i.e. at row offset (lc - base4) - r of this lane's row
                const int base4 = ((int) lc - (TB - 1)) & ~3;
                const float *tap = rtiles + (size_t) st * L * RS + lane * RS + ((int) lc - base4);
                float *wcol = wtile + lane;
                uint32_t r = 0;
                for (; r + kFxBatch <= rows; r += kFxBatch)
                    fx_batch<KIND, kFxBatch>(col + r * L, tap - r, wcol + r * WS, WS, prev_a, frac, feedback, d_balance, r_balance, master, w_low,
                                             w_band, kw, f_low, f_band, kf);
                for (; r < rows; r++)
                    fx_batch<KIND, 1>(col + r * L, tap - r, wcol + r * WS, WS, prev_a, frac, feedback, d_balance, r_balance, master, w_low,
                                      w_band, kw, f_low, f_band, kf);
                lc = lc >= rows ? lc - rows : lc + kRing - rows;
            }
            // write-back: this lane's `rows` new samples are the contiguous run wp - rows + 1 .. wp of its ring (the write
            // pointer walks backwards): scalar stores up to a 16-byte boundary, then whole float4s, then the scalar rest.
            // wp is warp-uniform, so the three loops do not diverge.
            if (live) {
                const float *wcol = wtile + lane;
                uint32_t r = 0;
                auto at = [&](uint32_t rr) { return wp >= rr ? wp - rr : wp + kRing - rr; };
                for (; r < rows && (at(r) & 3u) != 3u; r++) ring[at(r)] = wcol[r * WS];
                for (; r + 4 <= rows; r += 4) {
                    const float4 q = make_float4(wcol[(r + 3) * WS], wcol[(r + 2) * WS], wcol[(r + 1) * WS], wcol[r * WS]);
                    *reinterpret_cast<float4 *>(ring + at(r) - 3) = q;
                }
                for (; r < rows; r++) ring[at(r)] = wcol[r * WS];
            }
            wp = wp >= rows ? wp - rows : wp + kRing - rows;
        } else {
            if (live) {
                uint32_t w = wp;
                for (uint32_t r = 0; r < rows; r++) {  // DelayLine::Read straight from the ring (Fx.h:198-200)
                    uint32_t ia = w + delay;
                    ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
                    uint32_t ib = ia + 1;
                    ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
                    const float ra = ring[ia], rb = ring[ib];
                    float wr;
                    col[r * L] = fx_frame<KIND>(col[r * L], ra, rb, wr, frac, feedback, d_balance, r_balance, master, w_low, w_band, kw,
                                                f_low, f_band, kf);
                    ring[w] = wr;
                    w = w == 0 ? kRing - 1 : w - 1;
                }
            }
            wp = wp >= rows ? wp - rows : wp + kRing - rows;
        }
        fence_async_smem();
        __syncwarp();
        const uint32_t nxt = i + STAGES - 1;  // goes into the stage tile i-1 was stored from
        if (lane == 0) {
            tma_store_2d(&tm_out, (int) i0, (int) (i * TB), tiles + (size_t) st * TB * L);
            tma_commit();
            if (nxt < ntiles) {
                tma_wait_read<1>();
                const uint32_t ns = nxt % STAGES;
                mbar_arrive_expect_tx(&full[ns], kTileBytes);
                tma_load_2d(tiles + (size_t) ns * TB * L, &tm_in, (int) i0, (int) (nxt * TB), &full[ns]);
            }
        }
        fetch_ring(nxt, nxt % STAGES);
    }
    if (lane == 0) tma_wait_all<0>();
    if (kHasRing) fx_cp_async_wait<0>();
    if (live) {
        svf[inst] = f_low;
        svf[(size_t) I + inst] = f_band;
        svf[2 * (size_t) I + inst] = w_low;
        svf[3 * (size_t) I + inst] = w_band;
    }
}

// ------------------------------------------------------------------------------------------------
// k_fx_split: the ring kinds (DelayFx, FxRack) as a pipeline of warps inside one CTA of 32 instances.  In k_fx_tma one warp
// issues everything a frame needs back to back: the ring traffic of the tile, the wet filter's 12-deep recurrence, the mixes
// and (rack) the main filter's recurrence -- 132 (delay) and 196 (rack) cycles per frame against the 78 of one recurrence.
// Nothing but the two filters is recurrent, and the second filter only consumes the first one's output, so the frame is
// cut at those points and each piece gets its own warp, i.e. its own scheduler slot:
//   warp 0  I/O     TMA tile of inputs + cp.async tap runs -> DelayLine::Read (to the rd tile), ring write-back
//   warp 1  wet     wet filter, dry/wet mixes, in place in the audio tile          [delay: stores the tile]
//   warp 2  tail    main filter, master volume, in place (rack only)               [rack: stores the tile]
// Tiles move down the pipeline through shared memory with one 32-arrival mbarrier per stage and hand-over (`tapd`: I/O -> wet,
// `mid`: wet -> tail); the last warp stores the tile by TMA and refills its stage, which is the only back-pressure needed
// (a stage's `full` barrier cannot complete before the tile that used it STAGES tiles ago was stored).  Shared memory per
// CTA at 64 x 4: audio 32 KB + taps 34 KB + write-back 8 KB + rd 32 KB = 106 KB, two CTAs per SM.
template <int TB, int STAGES>
__host__ __device__ constexpr size_t fx_split_smem_floats() {
    return fx_tma_smem_floats<TB, STAGES>(true) + (size_t) STAGES * TB * kFxL;
}

//
// G pipelines (32 instances each) share a CTA so that the kernel, not the block scheduler, decides which scheduler (SM
// sub-partition = warp index mod 4) each warp lands on: two co-resident 3-warp CTAs double up on sub-partitions 0-2 and
// ran 117 cycles per frame where one alone runs 82 (profiles/r01_run25_fx_split.jsonl).  G = 2: the two wet warps and
// the two tail warps get a sub-partition each, the light I/O warps share with the tails (warps 4, 5 exit at once);
// G = 4: pipeline p is warps p, p + 4, p + 8 -- all three of its stages on sub-partition p, 99 issue slots per frame each.
constexpr int kFxIo = 0, kFxWet = 1, kFxTail = 2, kFxIdle = 3;
// run(std::integral_constant<int, type>) for a warp-uniform filter type 0..4, run(<-1>) otherwise
#define OLB_FX_BY_TYPE(TYPE, RUN)                                  \
    switch (TYPE) {                                                \
        case 0: RUN(std::integral_constant<int, 0>{}); break;      \
        case 1: RUN(std::integral_constant<int, 1>{}); break;      \
        case 2: RUN(std::integral_constant<int, 2>{}); break;      \
        case 3: RUN(std::integral_constant<int, 3>{}); break;      \
        case 4: RUN(std::integral_constant<int, 4>{}); break;      \
        default: RUN(std::integral_constant<int, -1>{}); break;    \
    }
template <int KIND, int G>
__host__ __device__ constexpr int fx_split_warps() {
    return G == 1 ? (OLB_K(KIND) == OLB_FX_RACK ? 3 : 2) : G == 2 ? (OLB_K(KIND) == OLB_FX_RACK ? 8 : 4) : (OLB_K(KIND) == OLB_FX_RACK ? 12 : 8);
}
template <int TB, int STAGES>
__host__ __device__ constexpr size_t fx_split_pipe_bytes() {  // one pipeline's tiles + barriers, rounded for TMA alignment
    return (fx_split_smem_floats<TB, STAGES>() * sizeof(float) + (size_t) 3 * STAGES * sizeof(uint64_t) + 1023) & ~(size_t) 1023;
}

template <int KIND, int TB, int STAGES, int G>
__global__ void __launch_bounds__(32 * fx_split_warps<KIND, G>()) k_fx_split(const __grid_constant__ TensorMap tm_in,
                                                                              const __grid_constant__ TensorMap tm_out, uint32_t frames,
                                                                              uint32_t I, const float *__restrict__ par, float *svf,
                                                                              float *ring_all, uint32_t wp0, uint32_t type_mask) {
    static_assert(OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK, "ring kinds only");
    static_assert(G == 1 || G == 2 || G == 4, "1, 2 or 4 pipelines per CTA");
    constexpr bool kRack = OLB_K(KIND) == OLB_FX_RACK;
    constexpr bool FM = OLB_F(KIND);
    constexpr int L = kFxL, RS = TB + 4, WS = L;
    extern __shared__ __align__(1024) unsigned char smem_all[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t pipe, role;
    if (G == 1) {
        pipe = 0;
        role = warp;  // io, wet, tail
    } else if (G == 2) {
        pipe = warp & 1;
        role = kRack ? (warp < 2 ? kFxWet : warp < 4 ? kFxTail : warp < 6 ? kFxIdle : kFxIo) : (warp < 2 ? kFxWet : kFxIo);
    } else {
        pipe = warp & 3;
        role = warp >> 2;
    }
    unsigned char *smem_raw = smem_all + (size_t) pipe * fx_split_pipe_bytes<TB, STAGES>();
    const uint32_t i0 = (blockIdx.x * G + pipe) * L, inst = i0 + lane;
    const bool live = inst < I;
    float *tiles = reinterpret_cast<float *>(smem_raw);                    // [STAGES][TB][L]  audio, in place
    float *rtiles = tiles + (size_t) STAGES * TB * L;                      // [STAGES][L][RS]  delay-line taps
    float *wtile = rtiles + (size_t) STAGES * L * RS;                      // [TB][WS]         values to write back
    float *rdtiles = reinterpret_cast<float *>(smem_raw) + fx_tma_smem_floats<TB, STAGES>(true);  // [STAGES][TB][L] DelayLine::Read
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + fx_split_smem_floats<TB, STAGES>() * sizeof(float));
    uint64_t *tapd = full + STAGES, *mid = tapd + STAGES;
    if (role == kFxIo && lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&tapd[s], 32);
            mbar_init(&mid[s], 32);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (role == kFxIdle || i0 >= I) return;  // padding warps, and the pipelines past the end of the bank
    constexpr uint32_t kTileBytes = TB * L * sizeof(float);
    const uint32_t ntiles = (frames + TB - 1) / TB;
    const uint32_t ci = live ? inst : 0;  // idle lanes read instance 0's parameters and never write
    auto P = [&](int p) { return par[(size_t) p * I + ci]; };
    // the last warp of the pipeline: store tile i, then refill the stage tile i-1 was stored from
    auto store_and_refill = [&](uint32_t i, uint32_t st) {
        fence_async_smem();
        __syncwarp();
        const uint32_t nxt = i + STAGES - 1;
        if (lane == 0) {
            tma_store_2d(&tm_out, (int) i0, (int) (i * TB), tiles + (size_t) st * TB * L);
            tma_commit();
            if (nxt < ntiles) {
                tma_wait_read<1>();
                const uint32_t ns = nxt % STAGES;
                mbar_arrive_expect_tx(&full[ns], kTileBytes);
                tma_load_2d(tiles + (size_t) ns * TB * L, &tm_in, (int) i0, (int) (nxt * TB), &full[ns]);
            }
        }
    };

    if (role == kFxTail) {  // ---- tail (rack): main filter + master volume
        const SvfCoef kf{P(P_F_FREQ), P(P_F_DAMP), P(P_F_DRIVE), (int) __float_as_uint(P(P_F_TYPE))};
        const float master = P(P_MASTER);
        float f_low = svf[ci], f_band = svf[(size_t) I + ci];
        // one filter type for the whole warp (idle lanes carry instance 0's)?  then the tap is picked at compile time: 8 selects
        // and the peak subtractions of a frame less on the warps that set the pipeline's pace
        const int type0 = __shfl_sync(0xffffffffu, kf.type, 0);
        const int uniform_type = ((type_mask & 2u) && __all_sync(0xffffffffu, kf.type == type0 || !live) && type0 >= 0 && type0 <= 4) ? type0 : -1;
        auto run = [&](auto tc) {  // (the whole tile loop per type: one switch per launch, the loops as plain as without it)
            constexpr int T_ = decltype(tc)::value;
            for (uint32_t i = 0; i < ntiles; i++) {
                const uint32_t st = i % STAGES, ph = (i / STAGES) & 1;
                mbar_wait(&mid[st], ph);
                const uint32_t rows = min((uint32_t) TB, frames - i * TB);
                float *col = tiles + (size_t) st * TB * L + lane;
                if (live) {
                    uint32_t r = 0;
                    for (; r + kFxBatch <= rows; r += kFxBatch) tail_batch<kFxBatch, FM, T_>(col + r * L, master, f_low, f_band, kf);
                    for (; r < rows; r++) tail_batch<1, FM, T_>(col + r * L, master, f_low, f_band, kf);
                }
                store_and_refill(i, st);
            }
        };
        OLB_FX_BY_TYPE(uniform_type, run)
        if (lane == 0) tma_wait_all<0>();
        if (live) {
            svf[inst] = f_low;
            svf[(size_t) I + inst] = f_band;
        }
        return;
    }
    if (role == kFxWet) {  // ---- wet filter + mixes
        const SvfCoef kw{P(P_W_FREQ), P(P_W_DAMP), P(P_W_DRIVE), (int) __float_as_uint(P(P_W_TYPE))};
        const float d_balance = P(P_D_BALANCE), r_balance = P(P_R_BALANCE);
        float w_low = svf[2 * (size_t) I + ci], w_band = svf[3 * (size_t) I + ci];
        const int type0 = __shfl_sync(0xffffffffu, kw.type, 0);
        const int uniform_type = ((type_mask & 1u) && __all_sync(0xffffffffu, kw.type == type0 || !live) && type0 >= 0 && type0 <= 4) ? type0 : -1;
        if constexpr (kRack) {
            auto run = [&](auto tc) {  // the whole tile loop per type: one switch per launch
                constexpr int T_ = decltype(tc)::value;
                for (uint32_t i = 0; i < ntiles; i++) {
                    const uint32_t st = i % STAGES, ph = (i / STAGES) & 1;
                    mbar_wait(&full[st], ph);  // the inputs (async proxy) ...
                    mbar_wait(&tapd[st], ph);  // ... and warp 0's reads of the delay line
                    const uint32_t rows = min((uint32_t) TB, frames - i * TB);
                    float *col = tiles + (size_t) st * TB * L + lane;
                    const float *rdcol = rdtiles + (size_t) st * TB * L + lane;
                    if (live) {
                        uint32_t r = 0;
                        for (; r + kFxBatch <= rows; r += kFxBatch)
                            wet_batch<kFxBatch, true, FM, T_>(col + r * L, rdcol + r * L, d_balance, r_balance, w_low, w_band, kw);
                        for (; r < rows; r++) wet_batch<1, true, FM, T_>(col + r * L, rdcol + r * L, d_balance, r_balance, w_low, w_band, kw);
                    }
                    mbar_arrive(&mid[st]);
                }
            };
            OLB_FX_BY_TYPE(uniform_type, run)
        } else {
            // (the delay's two-warp pipeline runs 6 % faster with the switch INSIDE the tile loop, the rack's three-warp one 4 %
            // faster with it outside: profiles/r02_run70_fx_split_types.txt -- these warps' code is that sensitive to its shape)
            for (uint32_t i = 0; i < ntiles; i++) {
                const uint32_t st = i % STAGES, ph = (i / STAGES) & 1;
                mbar_wait(&full[st], ph);
                mbar_wait(&tapd[st], ph);
                const uint32_t rows = min((uint32_t) TB, frames - i * TB);
                float *col = tiles + (size_t) st * TB * L + lane;
                const float *rdcol = rdtiles + (size_t) st * TB * L + lane;
                if (live) {
                    auto run = [&](auto tc) {
                        constexpr int T_ = decltype(tc)::value;
                        uint32_t r = 0;
                        for (; r + kFxBatch <= rows; r += kFxBatch)
                            wet_batch<kFxBatch, false, FM, T_>(col + r * L, rdcol + r * L, d_balance, r_balance, w_low, w_band, kw);
                        for (; r < rows; r++) wet_batch<1, false, FM, T_>(col + r * L, rdcol + r * L, d_balance, r_balance, w_low, w_band, kw);
                    };
                    OLB_FX_BY_TYPE(uniform_type, run)
                }
                store_and_refill(i, st);
            }
        }
        if (!kRack && lane == 0) tma_wait_all<0>();
        if (live) {
            svf[2 * (size_t) I + inst] = w_low;
            svf[3 * (size_t) I + inst] = w_band;
        }
        return;
    }

    // ---- I/O
    const float feedback = P(P_D_FEEDBACK), frac = P(P_D_FRAC);
    const uint32_t delay = __float_as_uint(P(P_D_DELAY));
    float *ring = ring_all + (size_t) ci * kRing;
    const bool far = live && delay >= (uint32_t) (STAGES * TB + 2);
    const bool all_far = __all_sync(0xffffffffu, far || !live);  // see k_fx_tma: coalesced ring traffic needs the whole warp ahead
    uint32_t wp = wp0, la = wp0 + delay, lc;
    la = la >= (uint32_t) kRing ? la - kRing : la;
    lc = la;
    float prev_a = 0.f;
    if (all_far && live) {
        const uint32_t ib = la + 1 >= (uint32_t) kRing ? la + 1 - kRing : la + 1;
        prev_a = ring[ib];
    }
    auto fetch_ring = [&](uint32_t tile, uint32_t stage) {  // as in k_fx_tma: 16-byte runs into this lane's tap row
        if (all_far && live && tile < ntiles) {
            const uint32_t rows = min((uint32_t) TB, frames - tile * TB);
            float *dst = rtiles + (size_t) stage * L * RS + lane * RS;
            const int base4 = ((int) la - (TB - 1)) & ~3;
#pragma unroll
            for (int c = 0; c < RS / 4; c++) {
                int idx = base4 + 4 * c;
                if (idx <= (int) la) {
                    idx = idx < 0 ? idx + kRing : idx;
                    fx_cp_async16(dst + 4 * c, ring + idx);
                }
            }
            la = la >= rows ? la - rows : la + kRing - rows;
        }
        fx_cp_async_commit();
    };
    for (uint32_t j = 0; j < (uint32_t) (STAGES - 1); j++) {
        if (lane == 0 && j < ntiles) {
            mbar_arrive_expect_tx(&full[j], kTileBytes);
            tma_load_2d(tiles + (size_t) j * TB * L, &tm_in, (int) i0, (int) (j * TB), &full[j]);
        }
        fetch_ring(j, j);
    }
    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % STAGES, ph = (i / STAGES) & 1;
        fx_cp_async_wait<STAGES - 2>();  // each lane reads only the row it copied itself
        mbar_wait(&full[st], ph);
        const uint32_t rows = min((uint32_t) TB, frames - i * TB);
        const float *col = tiles + (size_t) st * TB * L + lane;
        float *rdcol = rdtiles + (size_t) st * TB * L + lane;
        if (all_far) {
            if (live) {
                const int base4 = ((int) lc - (TB - 1)) & ~3;
                const float *tap = rtiles + (size_t) st * L * RS + lane * RS + ((int) lc - base4);
                float *wcol = wtile + lane;
                uint32_t r = 0;
                for (; r + kFxBatch <= rows; r += kFxBatch) io_batch<kFxBatch, FM>(col + r * L, tap - r, wcol + r * WS, WS, rdcol + r * L, prev_a, frac, feedback);
                for (; r < rows; r++) io_batch<1, FM>(col + r * L, tap - r, wcol + r * WS, WS, rdcol + r * L, prev_a, frac, feedback);
                lc = lc >= rows ? lc - rows : lc + kRing - rows;
            }
            mbar_arrive(&tapd[st]);  // the wet warp can start while the ring is written back
            if (live) {  // write-back as in k_fx_tma: scalars to a 16-byte boundary, float4s, scalar rest
                const float *wcol = wtile + lane;
                uint32_t r = 0;
                auto at = [&](uint32_t rr) { return wp >= rr ? wp - rr : wp + kRing - rr; };
                for (; r < rows && (at(r) & 3u) != 3u; r++) ring[at(r)] = wcol[r * WS];
                for (; r + 4 <= rows; r += 4) {
                    const float4 q = make_float4(wcol[(r + 3) * WS], wcol[(r + 2) * WS], wcol[(r + 1) * WS], wcol[r * WS]);
                    *reinterpret_cast<float4 *>(ring + at(r) - 3) = q;
                }
                for (; r < rows; r++) ring[at(r)] = wcol[r * WS];
            }
        } else {
            if (live) {
                uint32_t w = wp;
                for (uint32_t r = 0; r < rows; r++) {  // DelayLine::Read straight from the ring (Fx.h:198-200)
                    uint32_t ia = w + delay;
                    ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
                    uint32_t ib = ia + 1;
                    ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
                    const float ra = ring[ia], rb = ring[ib];
                    const float rd = dl_read<FM>(ra, rb, frac);
                    rdcol[r * L] = rd;
                    ring[w] = dl_write<FM>(col[r * L], feedback, rd);
                    w = w == 0 ? kRing - 1 : w - 1;
                }
            }
            mbar_arrive(&tapd[st]);
        }
        wp = wp >= rows ? wp - rows : wp + kRing - rows;
        fetch_ring(i + STAGES - 1, (i + STAGES - 1) % STAGES);
    }
    fx_cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// k_plate_tail: the back half of the plate rack (OLB_FX_RACK_PLATE): ReverbFx's dry/wet around the Dattorro plate's left
// output (Fx.h:294-300 with the plate of libs/dattorro-verb substituted for the stub ReverbSc, the mapping the reference
// sketches in modules/fxlib/ReverbFx.cpp:11-37), then FxRack's filter and master volume (Fx.h:426-434).  One instance
// per thread, 8 frames of loads in flight.
__global__ void __launch_bounds__(64) k_plate_tail(const float *wet, const float *dry, float *out, uint32_t frames, uint32_t I,
                                                   const float *__restrict__ par, float *svf) {
    constexpr int B = kFxB;
    constexpr bool FM = false;  // the plate rack keeps the reference arithmetic
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    auto P = [&](int p) { return par[(size_t) p * I + inst]; };
    const SvfCoef kf{P(P_F_FREQ), P(P_F_DAMP), P(P_F_DRIVE), (int) __float_as_uint(P(P_F_TYPE))};
    const float r_balance = P(P_R_BALANCE), master = P(P_MASTER);
    float f_low = svf[inst], f_band = svf[(size_t) I + inst];
    for (uint32_t n0 = 0; n0 < frames; n0 += B) {
        float w[B], d[B];
#pragma unroll
        for (int j = 0; j < B; j++)
            if (n0 + j < frames) {
                w[j] = wet[(size_t) (n0 + j) * I + inst];
                d[j] = dry[(size_t) (n0 + j) * I + inst];
            }
#pragma unroll
        for (int j = 0; j < B; j++)
            if (n0 + j < frames) {
                float v = dry_wet<FM>(w[j], d[j], r_balance);
                v = svf_tick<FM>(v, f_low, f_band, kf);
                out[(size_t) (n0 + j) * I + inst] = __fmul_rn(v, master);
            }
    }
    svf[inst] = f_low;
    svf[(size_t) I + inst] = f_band;
}

}  // namespace olb

using namespace olb;

// ================================================================================================
// Host side: one host::RackParams per instance carries the reference objects' members; the device table is rebuilt
// from it when a setter ran (setters take effect at the next Process, as Update() does in the reference).
struct olb_fx {
    std::recursive_mutex mu;
    StreamOrder order;
    int device = 0, kind = 0;
    bool fast = false;  // olb_fx_set_mode: the fused arithmetic tier
    uint32_t I = 0;
    float sample_rate = 48000.f;
    std::vector<host::RackParams> units;
    bool dirty = true;
    uint32_t wp = 0;  // DelayLine::write_ptr_, shared by the bank
    float *d_par = nullptr, *d_svf = nullptr, *d_ring = nullptr;
    std::vector<float> stage;
    const char *last_kernel = "none";
    HostStage host;
    // OLB_FX_RACK_PLATE: the Dattorro bank in the reverb slot and the buffers between the three stages
    olb_dattorro *plate = nullptr;
    bool plate_dirty = true;
    float *d_mid[3] = {nullptr, nullptr, nullptr};  // delay output (dry), plate left, plate right
    size_t mid_elems = 0;
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

static bool fx_has_ring(int kind) { return kind == OLB_FX_DELAY || kind == OLB_FX_RACK || kind == OLB_FX_RACK_PLATE; }

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
    OLB_CUDA(h->order.memset0(h->d_svf, (size_t) 4 * h->I * sizeof(float)));
    if (h->d_ring) OLB_CUDA(h->order.memset0(h->d_ring, (size_t) h->I * kRing * sizeof(float)));
    h->wp = 0;
    h->dirty = true;
    if (h->plate) {  // the plate has no sample-rate dependence (verb.cpp sizes its rings in samples); re-create = clear
        int rc = olb_dattorro_init(h->plate);
        if (rc) return rc;
        h->plate_dirty = true;
    }
    return OLB_OK;
}

// k_fx_tma: one warp per CTA, [TB frames][32 instances] boxes, STAGES-deep ring (plus a second ring of the same shape for
// the delay-line reads).  OLB_FX_TMA_CFG="<TB>x<STAGES>" selects another compiled shape (tuning aid).
template <int KIND, int TB, int STAGES>
static int fx_launch_tma_cfg(olb_fx *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr bool has_ring = (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK);
    constexpr size_t smem = fx_tma_smem_floats<TB, STAGES>(has_ring) * sizeof(float) + (size_t) STAGES * sizeof(uint64_t);
    TensorMap tin, tout;
    int rc = make_tmap_frames(&tin, in, frames, h->I, kFxL, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, h->I, kFxL, TB);
    if (rc) return rc;
    auto kern = k_fx_tma<KIND, TB, STAGES>;
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    kern<<<(h->I + kFxL - 1) / kFxL, 32, smem, st>>>(tin, tout, frames, h->I, h->d_par, h->d_svf, h->d_ring, h->wp);
    return OLB_OK;
}

// k_fx_split: the ring kinds as a pipeline of two (delay) or three (rack) warps per CTA.
template <int KIND, int TB, int STAGES, int G>
static int fx_launch_split_cfg(olb_fx *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr size_t smem = (size_t) G * fx_split_pipe_bytes<TB, STAGES>();
    static_assert(smem <= 227 * 1024, "pipelines do not fit one SM");
    TensorMap tin, tout;
    int rc = make_tmap_frames(&tin, in, frames, h->I, kFxL, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, h->I, kFxL, TB);
    if (rc) return rc;
    auto kern = k_fx_split<KIND, TB, STAGES, G>;
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const uint32_t pipes = (h->I + kFxL - 1) / kFxL;
    // which warps pick the filter's output tap at compile time when their instances share a type: bit 0 wet, bit 1 tail
    // (OLB_FX_SPLIT_TYPES: tuning aid).  Default: the wet warp, the pipeline's heaviest -- 8 192 instances, of the roof:
    // delay 0.402 -> 0.435 (fused tier 0.441 -> 0.478), rack 0.386 -> 0.400 (0.464 -> 0.466); with the tail warp specialised
    // too the rack's fused tier LOSES 6 % (profiles/r02_run70_fx_split_types.txt, r02_run73_fx_split_final.txt)
    const char *et = getenv("OLB_FX_SPLIT_TYPES");
    const uint32_t type_mask = et ? (uint32_t) atoi(et) : 1u;
    kern<<<(pipes + G - 1) / G, 32 * fx_split_warps<KIND, G>(), smem, st>>>(tin, tout, frames, h->I, h->d_par, h->d_svf, h->d_ring, h->wp, type_mask);
    return OLB_OK;
}

// The ring kinds run k_fx_split unless OLB_FX_SPLIT=0 asks for the one-warp k_fx_tma (test / tuning aid).
static bool fx_use_split(int kind) {
    kind &= 7;
    if (kind != OLB_FX_DELAY && kind != OLB_FX_RACK) return false;
    const char *e = getenv("OLB_FX_SPLIT");
    return !e || atoi(e) != 0;
}

template <int KIND>
static int fx_launch_tma(olb_fx *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    int tb = 64, stages = 4;
    if (const char *e = getenv("OLB_FX_TMA_CFG")) sscanf(e, "%dx%d", &tb, &stages);
    if (fx_use_split(KIND)) {
        constexpr int SK = (OLB_K(KIND) == OLB_FX_DELAY || OLB_K(KIND) == OLB_FX_RACK) ? KIND : (OLB_FX_DELAY | (KIND & kFxFastBit));
        // pipelines per CTA: as many as it takes to keep the bank on one CTA per SM (up to 4), so that the kernel places
        // every warp on a scheduler itself; OLB_FX_SPLIT_G / OLB_FX_TMA_CFG override (test / tuning aids)
        const uint32_t pipes = (h->I + kFxL - 1) / kFxL;
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
        int g = pipes <= (uint32_t) sms ? 1 : pipes <= 2u * sms ? 2 : 4;
        if (const char *e = getenv("OLB_FX_SPLIT_G")) g = atoi(e);
        const bool tb32 = getenv("OLB_FX_TMA_CFG") ? tb == 32 : g == 4;
        if (g == 4) return fx_launch_split_cfg<SK, 32, 4, 4>(h, in, out, frames, st);  // 4 x 53 KB
        if (g == 2) return tb32 ? fx_launch_split_cfg<SK, 32, 4, 2>(h, in, out, frames, st) : fx_launch_split_cfg<SK, 64, 4, 2>(h, in, out, frames, st);
        return tb32 ? fx_launch_split_cfg<SK, 32, 4, 1>(h, in, out, frames, st) : fx_launch_split_cfg<SK, 64, 4, 1>(h, in, out, frames, st);
    }
    if (tb == 32 && stages == 4) return fx_launch_tma_cfg<KIND, 32, 4>(h, in, out, frames, st);
    if (tb == 64 && stages == 4) return fx_launch_tma_cfg<KIND, 64, 4>(h, in, out, frames, st);
    if (tb == 64 && stages == 8) return fx_launch_tma_cfg<KIND, 64, 8>(h, in, out, frames, st);
    if (tb == 32 && stages == 8) return fx_launch_tma_cfg<KIND, 32, 8>(h, in, out, frames, st);
    return fx_launch_tma_cfg<KIND, 64, 4>(h, in, out, frames, st);
}

OLB_API int olb_fx_create(int device, int kind, uint32_t instances, float sample_rate, olb_fx **out) {
    OLB_ARG(out, "olb_fx_create: out is NULL");
    *out = nullptr;
    OLB_ARG(kind >= OLB_FX_FILTER && kind <= OLB_FX_RACK_PLATE, "olb_fx_create: unknown kind %d", kind);
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
    rc = kind == OLB_FX_RACK_PLATE ? olb_dattorro_create(device, instances, &h->plate) : OLB_OK;
    if (!rc) rc = fx_init(h, sample_rate);
    if (rc) {
        olb_fx_destroy(h);
        return rc;
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_fx_destroy(olb_fx *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    cudaFree(h->d_par); cudaFree(h->d_svf); cudaFree(h->d_ring);
    for (float *p : h->d_mid) cudaFree(p);
    olb_dattorro_destroy(h->plate);
    h->host.release();
    delete h;
    return OLB_OK;
}

OLB_API int olb_fx_init(olb_fx *h, float sample_rate) {
    OLB_ARG(h, "olb_fx_init: NULL handle");
    OLB_LOCK(h);
    DeviceGuard g(h->device);
    return fx_init(h, sample_rate);
}

OLB_API int olb_fx_update_midi(olb_fx *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value) {
    OLB_ARG(h, "olb_fx_update_midi: NULL handle");
    OLB_LOCK(h);
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
    OLB_LOCK(h);
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

// The arithmetic tier of a bank (olbank.h): EXACT rounds every product and sum like the reference's -O3 build without
// contraction; FMA contracts the svf, the ring's interpolation, the feedback write and the mixes into fused operations
// (oracle: orc_fx_set_mode).  State carries over a switch -- the tiers differ in rounding only.
OLB_API int olb_fx_set_mode(olb_fx *h, int mode) {
    OLB_ARG(h, "olb_fx_set_mode: NULL handle");
    OLB_LOCK(h);
    OLB_ARG(mode == OLB_MODE_EXACT || mode == OLB_MODE_FMA, "olb_fx_set_mode: mode %d (the fx banks have OLB_MODE_EXACT and OLB_MODE_FMA)", mode);
    OLB_ARG(h->kind != OLB_FX_RACK_PLATE || mode == OLB_MODE_EXACT, "olb_fx_set_mode: the plate rack has OLB_MODE_EXACT only");
    h->fast = mode == OLB_MODE_FMA;
    return OLB_OK;
}

OLB_API const char *olb_fx_last_kernel(olb_fx *h) { return h ? h->last_kernel : "none"; }

// The plate rack's stages 2 and 3 (stage 1, the DelayFx kernel, has just written h->d_mid[0]).
static int fx_plate_stages(olb_fx *h, float *out_dev, uint32_t frames, cudaStream_t st) {
    if (h->plate_dirty) {  // ReverbFx members -> DattorroVerb setters (modules/fxlib/ReverbFx.cpp:29-37)
        const uint32_t I = h->I;
        std::vector<float> v(I);
        auto push = [&](int param, float (*get)(const host::ReverbParams &)) {
            for (uint32_t i = 0; i < I; i++) v[i] = get(h->units[i].reverb);
            return olb_dattorro_set_params(h->plate, 0, I, param, v.data(), 1);
        };
        int rc = push(OLB_PLATE_PREFILTER, [](const host::ReverbParams &r) { return r.pre_cutoff; });
        if (!rc) rc = push(OLB_PLATE_INPUT_DIFFUSION1, [](const host::ReverbParams &r) { return r.input_diffusion1; });
        if (!rc) rc = push(OLB_PLATE_INPUT_DIFFUSION2, [](const host::ReverbParams &r) { return r.input_diffusion2; });
        if (!rc) rc = push(OLB_PLATE_DECAY, [](const host::ReverbParams &r) { return r.decay_time; });
        if (!rc) rc = push(OLB_PLATE_DAMPING, [](const host::ReverbParams &r) { return host::scale(r.cutoff, 0, 20000, 0, 1, 1); });
        if (!rc) rc = push(OLB_PLATE_DECAY_DIFFUSION, [](const host::ReverbParams &r) { return r.decay_diffusion; });
        if (rc) return rc;
        h->plate_dirty = false;
    }
    int rc = olb_dattorro_process(h->plate, h->d_mid[0], h->d_mid[1], h->d_mid[2], frames, st);
    if (rc) return rc;
    k_plate_tail<<<(h->I + 63) / 64, 64, 0, st>>>(h->d_mid[1], h->d_mid[0], out_dev, frames, h->I, h->d_par, h->d_svf);
    OLB_LAUNCHED();
    return OLB_OK;
}

OLB_API int olb_fx_process(olb_fx *h, const float *in_dev, float *out_dev_arg, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_fx_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev_arg, "olb_fx_process: NULL buffer");
    OLB_ARG(frames < (1ull << 32), "olb_fx_process: at most 2^32-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    OLB_CUDA(h->order.enter(st));
    if (h->dirty) {
        h->plate_dirty = true;
        int rc = fx_upload(h, st);
        if (rc) return rc;
    }
    const uint32_t F = (uint32_t) frames;
    // the plate rack runs DelayFx -> Dattorro bank -> k_plate_tail through intermediate buffers that grow on demand
    const bool plate = h->kind == OLB_FX_RACK_PLATE;
    const int kind = plate ? OLB_FX_DELAY : h->kind;
    if (plate && h->mid_elems < frames * h->I) {
        OLB_CUDA(cudaStreamSynchronize(st));
        for (float *&p : h->d_mid) {
            cudaFree(p);
            p = nullptr;
        }
        h->mid_elems = 0;
        for (float *&p : h->d_mid) OLB_CUDA(cudaMalloc(&p, frames * h->I * sizeof(float)));
        h->mid_elems = frames * h->I;
    }
    float *out_dev = plate ? h->d_mid[0] : out_dev_arg;
    // Kernel choice.  Default: k_fx_tma (needs instances % 4 == 0 and 16-byte aligned buffers for the tensor maps), else
    // k_fx_pf.  Tuning / test aid: OLB_FX_KERNEL=simple|pf|tma.
    const char *ek = getenv("OLB_FX_KERNEL");
    const bool tma_ok = (h->I % 4 == 0) && (((uintptr_t) in_dev | (uintptr_t) out_dev) & 15) == 0;
    const int which = (ek && !strcmp(ek, "simple")) ? 0 : ((ek && !strcmp(ek, "pf")) || !tma_ok ? 1 : 2);
    const unsigned grid = (h->I + 63) / 64;
#define OLB_FX_GO(KERNEL, KIND) KERNEL<KIND><<<grid, 64, 0, st>>>(in_dev, out_dev, F, h->I, h->d_par, h->d_svf, h->d_ring, h->wp)
#define OLB_FX_KINDS(KERNEL)                                                               \
    switch (kind | (h->fast ? kFxFastBit : 0)) {                                           \
        case OLB_FX_FILTER: OLB_FX_GO(KERNEL, OLB_FX_FILTER); break;                       \
        case OLB_FX_DELAY: OLB_FX_GO(KERNEL, OLB_FX_DELAY); break;                         \
        case OLB_FX_REVERB: OLB_FX_GO(KERNEL, OLB_FX_REVERB); break;                       \
        case OLB_FX_RACK: OLB_FX_GO(KERNEL, OLB_FX_RACK); break;                           \
        case OLB_FX_FILTER | kFxFastBit: OLB_FX_GO(KERNEL, OLB_FX_FILTER | kFxFastBit); break; \
        case OLB_FX_DELAY | kFxFastBit: OLB_FX_GO(KERNEL, OLB_FX_DELAY | kFxFastBit); break;   \
        case OLB_FX_REVERB | kFxFastBit: OLB_FX_GO(KERNEL, OLB_FX_REVERB | kFxFastBit); break; \
        default: OLB_FX_GO(KERNEL, OLB_FX_RACK | kFxFastBit); break;                       \
    }
    if (which == 0) {
        OLB_FX_KINDS(k_fx);
        h->last_kernel = "k_fx";
    } else if (which == 1) {
        OLB_FX_KINDS(k_fx_pf);
        h->last_kernel = "k_fx_pf";
    } else {
        int rc = OLB_OK;
        switch (kind | (h->fast ? kFxFastBit : 0)) {
            case OLB_FX_FILTER: rc = fx_launch_tma<OLB_FX_FILTER>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_DELAY: rc = fx_launch_tma<OLB_FX_DELAY>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_REVERB: rc = fx_launch_tma<OLB_FX_REVERB>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_RACK: rc = fx_launch_tma<OLB_FX_RACK>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_FILTER | kFxFastBit: rc = fx_launch_tma<OLB_FX_FILTER | kFxFastBit>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_DELAY | kFxFastBit: rc = fx_launch_tma<OLB_FX_DELAY | kFxFastBit>(h, in_dev, out_dev, F, st); break;
            case OLB_FX_REVERB | kFxFastBit: rc = fx_launch_tma<OLB_FX_REVERB | kFxFastBit>(h, in_dev, out_dev, F, st); break;
            default: rc = fx_launch_tma<OLB_FX_RACK | kFxFastBit>(h, in_dev, out_dev, F, st); break;
        }
        if (rc) return rc;
        h->last_kernel = fx_use_split(kind) ? "k_fx_split" : "k_fx_tma";
    }
#undef OLB_FX_KINDS
#undef OLB_FX_GO
    OLB_LAUNCHED();
    if (fx_has_ring(h->kind)) h->wp = (uint32_t) ((h->wp + (uint64_t) kRing - frames % kRing) % kRing);
    if (plate) return fx_plate_stages(h, out_dev_arg, F, st);
    return OLB_OK;
}

// Host buffers [frames][instances]: chunked in time through device staging; returns when out_host is complete.
OLB_API int olb_fx_process_host(olb_fx *h, const float *in_host, float *out_host, uint64_t frames) {
    OLB_ARG(h, "olb_fx_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_host && out_host, "olb_fx_process_host: NULL buffer");
    DeviceGuard g(h->device);
    const uint64_t I = h->I, chunk = host_chunk_frames(I, frames);
    int rc = h->host.ensure(0, chunk * I);
    if (rc) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        OLB_CUDA(cudaMemcpyAsync(h->host.d[0], in_host + t0 * I, n * I * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_fx_process(h, h->host.d[0], h->host.d[0], n, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(out_host + t0 * I, h->host.d[0], n * I * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}

// ================================================================================================
// Stereo instances (CHANNEL_COUNT = 2): banks of FilterFx<2> / DelayFx<2> / ReverbFx<2> / FxRack<2>, the shapes the
// reference's production caller runs (modules/ol_daisy/app/synth/main.cpp:83-85).  What the reference's templates do with
// the second channel is reproduced as it is, not as one might wish it were:
//   FilterFx<2>::Process  filters frame_in[0] into frame_out[0]; frame_out[1] is NOT WRITTEN               (Fx.h:88-108)
//   DelayFx<2>::Process   a delay line per channel, the wet filter on channel 0 only: channel 1's wet signal is the raw
//                         DelayLine::Read                                                                   (Fx.h:194-208)
//   ReverbFx<2>::Process  the stub ReverbSc and the dry/wet on both channels                               (Fx.h:294-300)
//   FxRack<2>::Process    delay -> reverb -> filter1(buf_b, buf_c): buf_c[1] is never written and stays 0, so
//                         frame_out[1] = 0 * master_volume; channel 0 is what FxRack<1> makes of channel 0  (Fx.h:426-434)
// Channel 0 of every kind therefore IS the mono unit: it runs through the mono bank above (its TMA / warp-pipeline kernels)
// on a de-interleaved copy that is cut in time so that it stays in L2, and a merge kernel writes the stereo frames.
// Audio is [frames][instances][2] (an instance's frame is the reference's t_sample[2]).  DelayFx<2>'s channel-1 rings are
// [48 000 positions][instances] (the write pointer is shared by the bank, so a write is one coalesced row).
struct olb_fx2 {
    std::recursive_mutex mu;
    olb_fx *mono = nullptr;  // channel 0
    int device = 0, kind = 0;
    uint32_t I = 0;
    float *d_ring1 = nullptr;  // DelayFx<2>: delay_lines_[1]
    uint32_t wp1 = 0;
    float *d_tmp = nullptr;
    size_t tmp_elems = 0;
    uint32_t min_delay = 0, max_delay = 0;
    const char *last_kernel = "none";
    HostStage host;
};

namespace olb {

__global__ void __launch_bounds__(256) k_fx2_take0(const float2 *__restrict__ in, float *__restrict__ tmp, size_t n) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) tmp[i] = in[i].x;
}

// One thread per (frame, instance) of a time cut: merges channel 0 (already processed, `tmp`) with channel 1.
// OLB_K(KIND) == OLB_FX_DELAY here is the time-parallel form of channel 1's delay line, valid while no frame of the cut reads a
// ring position another frame of the cut writes (cut <= delay and delay + 1 <= 48 000 - cut for every instance; the host
// checks).  in may alias out (each thread reads its frame before writing it).
template <int KIND>
__global__ void __launch_bounds__(256) k_fx2_put(const float2 *in, const float *__restrict__ tmp, float2 *out, uint32_t frames, uint32_t I,
                                                 const float *__restrict__ par, float *ring1, uint32_t wp0) {
    constexpr bool FM = OLB_F(KIND);
    const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t) frames * I) return;
    const uint32_t n = (uint32_t) (idx / I), inst = (uint32_t) (idx - (size_t) n * I);
    auto P = [&](int p) { return par[(size_t) p * I + inst]; };
    if (OLB_K(KIND) == OLB_FX_FILTER) {
        reinterpret_cast<float *>(out)[2 * idx] = tmp[idx];  // frame_out[1] stays what the caller had there
    } else if (OLB_K(KIND) == OLB_FX_RACK) {
        out[idx] = make_float2(tmp[idx], __fmul_rn(0.0f, P(P_MASTER)));  // buf_c[1] * master_volume, buf_c[1] == 0 for ever
    } else if (OLB_K(KIND) == OLB_FX_REVERB) {
        const float2 v = in[idx];
        const float b = P(P_R_BALANCE);
        out[idx] = make_float2(dry_wet<FM>(__fmul_rn(v.x, 0.8f), v.x, b), dry_wet<FM>(__fmul_rn(v.y, 0.8f), v.y, b));
    } else {
        const float x1 = in[idx].y;
        const uint32_t delay = __float_as_uint(P(P_D_DELAY));
        const uint32_t wp = (wp0 + (uint32_t) kRing - n % (uint32_t) kRing) % (uint32_t) kRing;
        uint32_t ia = wp + delay;
        ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
        uint32_t ib = ia + 1;
        ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
        const float a = ring1[(size_t) ia * I + inst], b = ring1[(size_t) ib * I + inst];
        const float rd = dl_read<FM>(a, b, P(P_D_FRAC));           // DelayLine::Read
        ring1[(size_t) wp * I + inst] = dl_write<FM>(x1, P(P_D_FEEDBACK), rd);   // Write(in + feedback * read)
        out[idx] = make_float2(tmp[idx], dry_wet<FM>(rd, x1, P(P_D_BALANCE)));                // filter_.Process(buf, buf) left buf[1] alone
    }
}

// DelayFx<2> channel 1, one instance per thread walking time: any delay (a cut may read what it wrote).
template <bool FM>
__global__ void __launch_bounds__(64) k_fx2_delay1(const float2 *in, const float *__restrict__ tmp, float2 *out, uint32_t frames, uint32_t I,
                                                   const float *__restrict__ par, float *ring1, uint32_t wp0) {
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    auto P = [&](int p) { return par[(size_t) p * I + inst]; };
    const float feedback = P(P_D_FEEDBACK), balance = P(P_D_BALANCE), frac = P(P_D_FRAC);
    const uint32_t delay = __float_as_uint(P(P_D_DELAY));
    uint32_t wp = wp0;
    for (uint32_t n = 0; n < frames; n++) {
        const size_t idx = (size_t) n * I + inst;
        const float x1 = in[idx].y;
        uint32_t ia = wp + delay;
        ia = ia >= (uint32_t) kRing ? ia - kRing : ia;
        uint32_t ib = ia + 1;
        ib = ib >= (uint32_t) kRing ? ib - kRing : ib;
        const float a = ring1[(size_t) ia * I + inst], b = ring1[(size_t) ib * I + inst];
        const float rd = dl_read<FM>(a, b, frac);
        ring1[(size_t) wp * I + inst] = dl_write<FM>(x1, feedback, rd);
        wp = wp == 0 ? kRing - 1 : wp - 1;
        out[idx] = make_float2(tmp[idx], dry_wet<FM>(rd, x1, balance));
    }
}

}  // namespace olb

static bool fx2_kind_ok(int kind) { return kind >= OLB_FX_FILTER && kind <= OLB_FX_RACK; }

static int fx2_init(olb_fx2 *h, float sr) {
    int rc = olb_fx_init(h->mono, sr);
    if (rc) return rc;
    if (h->d_ring1) OLB_CUDA(h->mono->order.memset0(h->d_ring1, (size_t) h->I * kRing * sizeof(float)));
    h->wp1 = 0;
    return OLB_OK;
}

OLB_API int olb_fx2_create(int device, int kind, uint32_t instances, float sample_rate, olb_fx2 **out) {
    OLB_ARG(out, "olb_fx2_create: out is NULL");
    *out = nullptr;
    OLB_ARG(fx2_kind_ok(kind), "olb_fx2_create: kind %d has no stereo form in the reference", kind);
    olb_fx *mono = nullptr;
    int rc = olb_fx_create(device, kind, instances, sample_rate, &mono);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_fx2 *h = new olb_fx2();
    h->mono = mono;
    h->device = device;
    h->kind = kind;
    h->I = instances;
    if (kind == OLB_FX_DELAY) {
        cudaError_t e = cudaMalloc(&h->d_ring1, (size_t) instances * kRing * sizeof(float));
        if (e == cudaSuccess) e = cudaMemset(h->d_ring1, 0, (size_t) instances * kRing * sizeof(float));
        if (e != cudaSuccess) {
            olb_fx_destroy(mono);
            delete h;
            return fail_cuda(e, "allocating the channel-1 delay lines", __FILE__, __LINE__);
        }
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_fx2_destroy(olb_fx2 *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    h->mono->order.quiesce();
    cudaFree(h->d_ring1);
    cudaFree(h->d_tmp);
    h->host.release();
    olb_fx_destroy(h->mono);
    delete h;
    return OLB_OK;
}

OLB_API int olb_fx2_init(olb_fx2 *h, float sample_rate) {
    OLB_ARG(h, "olb_fx2_init: NULL handle");
    OLB_LOCK(h);
    DeviceGuard g(h->device);
    return fx2_init(h, sample_rate);
}

OLB_API int olb_fx2_update_midi(olb_fx2 *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value) {
    OLB_ARG(h, "olb_fx2_update_midi: NULL handle");
    OLB_LOCK(h);
    return olb_fx_update_midi(h->mono, first, count, control, value);
}

OLB_API int olb_fx2_update_hw(olb_fx2 *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride) {
    OLB_ARG(h, "olb_fx2_update_hw: NULL handle");
    OLB_LOCK(h);
    return olb_fx_update_hw(h->mono, first, count, control, values_host, stride);
}

OLB_API int olb_fx2_set_mode(olb_fx2 *h, int mode) {
    OLB_ARG(h, "olb_fx2_set_mode: NULL handle");
    OLB_LOCK(h);
    return olb_fx_set_mode(h->mono, mode);
}

OLB_API const char *olb_fx2_last_kernel(olb_fx2 *h) { return h ? h->last_kernel : "none"; }

// Channel 1 of one cut (and channel 0's interleave), in the bank's arithmetic tier.
static void fx2_put(olb_fx2 *h, const float2 *in_c, const float *tmp, float2 *out_c, uint32_t n, unsigned grid, bool par_delay, cudaStream_t st) {
    const float *par = h->mono->d_par;
#define OLB_FX2_PUT(KIND, RING, WP) k_fx2_put<KIND><<<grid, 256, 0, st>>>(in_c, tmp, out_c, n, h->I, par, RING, WP)
    if (h->mono->fast) {
        constexpr int F = kFxFastBit;
        if (h->kind == OLB_FX_REVERB) OLB_FX2_PUT(OLB_FX_REVERB | F, nullptr, 0);
        else if (h->kind == OLB_FX_FILTER) OLB_FX2_PUT(OLB_FX_FILTER | F, nullptr, 0);
        else if (h->kind == OLB_FX_RACK) OLB_FX2_PUT(OLB_FX_RACK | F, nullptr, 0);
        else if (par_delay) OLB_FX2_PUT(OLB_FX_DELAY | F, h->d_ring1, h->wp1);
        else k_fx2_delay1<true><<<(h->I + 63) / 64, 64, 0, st>>>(in_c, tmp, out_c, n, h->I, par, h->d_ring1, h->wp1);
    } else {
        if (h->kind == OLB_FX_REVERB) OLB_FX2_PUT(OLB_FX_REVERB, nullptr, 0);
        else if (h->kind == OLB_FX_FILTER) OLB_FX2_PUT(OLB_FX_FILTER, nullptr, 0);
        else if (h->kind == OLB_FX_RACK) OLB_FX2_PUT(OLB_FX_RACK, nullptr, 0);
        else if (par_delay) OLB_FX2_PUT(OLB_FX_DELAY, h->d_ring1, h->wp1);
        else k_fx2_delay1<false><<<(h->I + 63) / 64, 64, 0, st>>>(in_c, tmp, out_c, n, h->I, par, h->d_ring1, h->wp1);
    }
#undef OLB_FX2_PUT
}

OLB_API int olb_fx2_process(olb_fx2 *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_fx2_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev, "olb_fx2_process: NULL buffer");
    OLB_ARG((((uintptr_t) in_dev | (uintptr_t) out_dev) & 7) == 0, "olb_fx2_process: stereo frames must be 8-byte aligned");
    OLB_ARG(frames < (1ull << 32), "olb_fx2_process: at most 2^32-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    std::lock_guard<std::recursive_mutex> mono_lock(h->mono->mu);  // setters wait for the end of this block: one parameter set per call
    const uint64_t I = h->I;
    if (h->kind == OLB_FX_DELAY && h->mono->dirty) {
        uint32_t lo = ~0u, hi = 0;
        for (const auto &u : h->mono->units) {
            lo = u.delay.delay < lo ? u.delay.delay : lo;
            hi = u.delay.delay > hi ? u.delay.delay : hi;
        }
        h->min_delay = lo;
        h->max_delay = hi;
    }
    // time cuts: the de-interleaved channel stays in L2 between its three kernels (<= 16 MiB), whole TMA tiles
    uint64_t cut = (16ull << 20) / (I * sizeof(float));
    cut = cut < 64 ? 64 : (cut / 64) * 64;
    if (h->kind == OLB_FX_REVERB) cut = frames;  // stateless, no intermediate
    bool par_delay = false;
    if (h->kind == OLB_FX_DELAY) {
        uint64_t c = cut;
        if (h->min_delay >= 64 && c > h->min_delay) c = (h->min_delay / 64) * 64;
        par_delay = c >= 64 && c <= h->min_delay && (uint64_t) h->max_delay + 1 + c <= (uint64_t) kRing;
        if (par_delay) cut = c;
    }
    if (cut > frames) cut = frames;
    if (h->kind != OLB_FX_REVERB && h->tmp_elems < cut * I) {
        OLB_CUDA(h->mono->order.sync());
        cudaFree(h->d_tmp);
        h->d_tmp = nullptr;
        h->tmp_elems = 0;
        OLB_CUDA(cudaMalloc(&h->d_tmp, cut * I * sizeof(float)));
        h->tmp_elems = cut * I;
    }
    const float2 *in2 = reinterpret_cast<const float2 *>(in_dev);
    float2 *out2 = reinterpret_cast<float2 *>(out_dev);
    OLB_CUDA(h->mono->order.enter(st));  // the intermediate and the channel-1 rings follow the mono bank's stream order
    for (uint64_t t0 = 0; t0 < frames; t0 += cut) {
        const uint32_t n = (uint32_t) (frames - t0 < cut ? frames - t0 : cut);
        const size_t elems = (size_t) n * I;
        const unsigned grid = (unsigned) ((elems + 255) / 256);
        const float2 *in_c = in2 + t0 * I;
        float2 *out_c = out2 + t0 * I;
        if (h->kind == OLB_FX_REVERB) {
            if (h->mono->dirty) {
                int rc = fx_upload(h->mono, st);
                if (rc) return rc;
            }
            fx2_put(h, in_c, nullptr, out_c, n, grid, false, st);
            OLB_LAUNCHED();
            h->last_kernel = "k_fx2_put";
            continue;
        }
        k_fx2_take0<<<grid, 256, 0, st>>>(in_c, h->d_tmp, elems);
        OLB_LAUNCHED();
        int rc = olb_fx_process(h->mono, h->d_tmp, h->d_tmp, n, st);
        if (rc) return rc;
        fx2_put(h, in_c, h->d_tmp, out_c, n, grid, par_delay, st);
        OLB_LAUNCHED();
        if (h->kind == OLB_FX_DELAY) h->wp1 = (uint32_t) ((h->wp1 + (uint64_t) kRing - n % kRing) % kRing);
        h->last_kernel = h->kind == OLB_FX_DELAY && !par_delay ? "k_fx2_delay1" : "k_fx2_put";
    }
    return OLB_OK;
}

// Host buffers [frames][instances][2], chunked through device staging; returns when out_host is complete.
OLB_API int olb_fx2_process_host(olb_fx2 *h, const float *in_host, float *out_host, uint64_t frames) {
    OLB_ARG(h, "olb_fx2_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_host && out_host, "olb_fx2_process_host: NULL buffer");
    DeviceGuard g(h->device);
    const uint64_t W = 2ull * h->I, chunk = host_chunk_frames(W, frames);
    int rc = h->host.ensure(0, chunk * W);
    if (rc) return rc;
    // FilterFx<2> leaves frame_out[1] as the caller had it: the staging buffer starts from the caller's out frames
    const bool keep_out = h->kind == OLB_FX_FILTER && in_host != out_host;
    if (keep_out && (rc = h->host.ensure(1, chunk * W))) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        float *d_in = h->host.d[0], *d_out = keep_out ? h->host.d[1] : h->host.d[0];
        OLB_CUDA(cudaMemcpyAsync(d_in, in_host + t0 * W, n * W * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        if (keep_out) OLB_CUDA(cudaMemcpyAsync(d_out, out_host + t0 * W, n * W * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_fx2_process(h, d_in, d_out, n, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(out_host + t0 * W, d_out, n * W * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}

// ================================================================================================
// olb_fx2_process_static: olb_fx2_process WITHOUT any device allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The only scratch of the stereo path -- the de-interleaved channel-0 cut that
// k_fx2_take0 writes, the mono bank's kernel filters in place and k_fx2_put / k_fx2_delay1 merge back -- lives in a
// file-scope device array of OLB_FX2_STATIC_SCRATCH_BYTES (8 MiB) that every GPU which loads the library holds until exit.
// A cut is whole 64-frame tiles, so banks of up to OLB_FX2_STATIC_MAX_INSTANCES (32 768) instances fit; larger ones are
// rejected.  Every cut overwrites the part of the array it reads (nothing has to start at zero); ALL handles and calls on
// a device share the array, so two calls that overlap on the GPU (different streams, no ordering between them) race.
static __device__ __align__(256) float g_fx2_static_scratch[OLB_FX2_STATIC_SCRATCH_BYTES / sizeof(float)];

// In host code `g_fx2_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, asked once per device with that device current.
static std::mutex g_fx2_static_mu;
static float *g_fx2_static_ptr[64] = {nullptr};

static int fx2_static_scratch(int device, float **out) {
    OLB_ARG(device >= 0 && device < 64, "olb_fx2_process_static: device index %d out of range", device);
    std::lock_guard<std::mutex> lk(g_fx2_static_mu);
    if (!g_fx2_static_ptr[device]) {  // the caller's DeviceGuard has made `device` current
        void *p = nullptr;
        OLB_CUDA(cudaGetSymbolAddress(&p, g_fx2_static_scratch));
        g_fx2_static_ptr[device] = static_cast<float *>(p);
    }
    *out = g_fx2_static_ptr[device];
    return OLB_OK;
}

OLB_API int olb_fx2_process_static(olb_fx2 *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_fx2_process_static: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev, "olb_fx2_process_static: NULL buffer");
    OLB_ARG((((uintptr_t) in_dev | (uintptr_t) out_dev) & 7) == 0, "olb_fx2_process_static: stereo frames must be 8-byte aligned");
    OLB_ARG(frames < (1ull << 32), "olb_fx2_process_static: at most 2^32-1 frames per call");
    OLB_ARG(h->I <= OLB_FX2_STATIC_MAX_INSTANCES,
            "olb_fx2_process_static: %u instances exceed the static scratch (%u instances = one 64-frame cut in %u bytes); use olb_fx2_process",
            h->I, (unsigned) OLB_FX2_STATIC_MAX_INSTANCES, (unsigned) OLB_FX2_STATIC_SCRATCH_BYTES);
    DeviceGuard g(h->device);
    float *tmp = nullptr;
    int rc = fx2_static_scratch(h->device, &tmp);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t) stream;
    std::lock_guard<std::recursive_mutex> mono_lock(h->mono->mu);
    const uint64_t I = h->I;
    if (h->kind == OLB_FX_DELAY && h->mono->dirty) {
        uint32_t lo = ~0u, hi = 0;
        for (const auto &u : h->mono->units) {
            lo = u.delay.delay < lo ? u.delay.delay : lo;
            hi = u.delay.delay > hi ? u.delay.delay : hi;
        }
        h->min_delay = lo;
        h->max_delay = hi;
    }
    uint64_t cut = (uint64_t) OLB_FX2_STATIC_SCRATCH_BYTES / (I * sizeof(float));  // >= 64 by the instance limit
    cut = (cut / 64) * 64;
    if (h->kind == OLB_FX_REVERB) cut = frames;  // stateless, no intermediate
    bool par_delay = false;
    if (h->kind == OLB_FX_DELAY) {
        uint64_t c = cut;
        if (h->min_delay >= 64 && c > h->min_delay) c = (h->min_delay / 64) * 64;
        par_delay = c >= 64 && c <= h->min_delay && (uint64_t) h->max_delay + 1 + c <= (uint64_t) kRing;
        if (par_delay) cut = c;
    }
    if (cut > frames) cut = frames;
    const float2 *in2 = reinterpret_cast<const float2 *>(in_dev);
    float2 *out2 = reinterpret_cast<float2 *>(out_dev);
    OLB_CUDA(h->mono->order.enter(st));
    for (uint64_t t0 = 0; t0 < frames; t0 += cut) {
        const uint32_t n = (uint32_t) (frames - t0 < cut ? frames - t0 : cut);
        const size_t elems = (size_t) n * I;
        const unsigned grid = (unsigned) ((elems + 255) / 256);
        const float2 *in_c = in2 + t0 * I;
        float2 *out_c = out2 + t0 * I;
        if (h->kind == OLB_FX_REVERB) {
            if (h->mono->dirty && (rc = fx_upload(h->mono, st))) return rc;
            fx2_put(h, in_c, nullptr, out_c, n, grid, false, st);
            OLB_LAUNCHED();
            h->last_kernel = "k_fx2_put";
            continue;
        }
        k_fx2_take0<<<grid, 256, 0, st>>>(in_c, tmp, elems);
        OLB_LAUNCHED();
        rc = olb_fx_process(h->mono, tmp, tmp, n, st);  // OLB_FX_FILTER / _DELAY / _RACK: launches only, no allocation
        if (rc) return rc;
        fx2_put(h, in_c, tmp, out_c, n, grid, par_delay, st);
        OLB_LAUNCHED();
        if (h->kind == OLB_FX_DELAY) h->wp1 = (uint32_t) ((h->wp1 + (uint64_t) kRing - n % kRing) % kRing);
        h->last_kernel = h->kind == OLB_FX_DELAY && !par_delay ? "k_fx2_delay1" : "k_fx2_put";
    }
    return OLB_OK;
}
