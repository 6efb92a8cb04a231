// chorus_bank.cu -- (a9) bank of mono choruses.
//
// Reference: NO C++ exists.  The only statement of the chorus is the Max/RNBO box graph
// modules/rnbo/patcher/mono-chorus.rnbopat:  in~ -> delay~ (@maxsize samplerate*2; delay = cycle~(rate') * d + d,
// "add the depth to keep the delay value > 0") -> lores~ (cutoff', q) -> *~ mix, +~ in~ * (1 - mix) -> out~, with
// rate' = scale(rate 0..1 -> 0.01..0.5 Hz), d = mstosamps(scale(depth 0.08..1 -> 1..12 ms)),
// cutoff' = scale(cutoff 0..1 -> 300..15 000 Hz), cycle~ phase offset `phase`; parameter defaults rate .2, depth .5,
// cutoff .3, q .5, mix .5, phase 1.  Max's object internals are not in the tree: the arithmetic is specified by
// oracle/olb_oracle.cpp (orc_chorus_*), PARITY UNPINNED, and this kernel is bit-exact against that restatement.
// In front of delay~ sits the gen~ pitch shifter modules/rnbo/patcher/pitchshift.gendsp (param pitch 0..3: a phasor driving two
// cosine-windowed taps of a 10 ms window): k_pitch_* below.  stereo-chorus.rnbopat is two of these with phase 1 (left) and 0
// (right) sharing every other param: a bank of 2 x instances units, set up by the host mirrors (no kernel of its own).
//
// Layout
//   audio   [frames][instances] fp32 in HBM
//   ring    per CTA (= one warp = 32 instances) a block [R][32] fp32, R = ceil(24 ms * sr) + 4 + 8 spare (1 164 at 48 kHz, 149 KB):
//           it lives in SHARED MEMORY for the whole launch (bulk-copied in at the start and out at the end), lane-fastest,
//           so the 32 lanes' reads at 32 different delays never collide on a bank.  All instances tick together and
//           share the write index.
//   state   [3][instances]: LFO phase, lores~ y1, y2        par [7][instances]: inc d offset scale a1 a2 mix
// One CTA = 32 instances = two warps: an audio warp with its own TMA tile pipeline (the k_fx_tma pattern) and an LFO warp
// one tile ahead of it; ~216 KB of shared memory per CTA, one CTA per SM.
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"
#include "tmap.h"

namespace olb {

constexpr int kChL = 32;      // instances per warp / CTA
constexpr int kChTab = 512;   // cycle~ table points (+1 guard)
enum { CP_INC, CP_D, CP_OFFSET, CP_SCALE, CP_A1, CP_A2, CP_MIX, CP_PINC, CP_COUNT };  // CP_PINC: the pitch shifter's phasor increment

__device__ __forceinline__ void bulk_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}

struct ChorusLane {
    float inc, d, offset, scale, a1, a2, mix;
    float phase, y1, y2;
};

constexpr int kChB = 16;  // frames per batch; the ring carries kChB spare slots so a batch may write ahead of its reads

// The chorus frame has two halves that share nothing but the delay they agree on: the LFO (phase -> table lookup -> delay
// in samples: a pure function of the phase state, independent of the audio) and the audio path (delay~ write and
// interpolated read, lores~, mix).  Each half is one long per-frame chain, so they run on TWO WARPS of the CTA, the LFO
// warp one tile ahead, handing (whole samples k, fraction fr) over through shared memory.
template <int N>
__device__ __forceinline__ void chorus_lfo_batch(ChorusLane &c, uint32_t *kq, float *frq, float dmax, const float *tab, uint32_t n) {
#pragma unroll
    for (int r = 0; r < N; r++) {
        if ((uint32_t) r < n) {
            float p = __fadd_rn(c.phase, c.offset);
            p = p >= 1.0f ? __fsub_rn(p, 1.0f) : p;
            const float idx = __fmul_rn(p, 512.0f);
            int i0 = __float2int_rz(idx);
            i0 = i0 > kChTab - 1 ? kChTab - 1 : i0;
            const float f = __fsub_rn(idx, __int2float_rn(i0));
            const float c0 = tab[i0], c1 = tab[i0 + 1];
            const float lfo = __fadd_rn(c0, __fmul_rn(__fsub_rn(c1, c0), f));
            float dt = __fadd_rn(__fmul_rn(lfo, c.d), c.d);
            dt = dt < 0.0f ? 0.0f : dt;
            dt = dt > dmax ? dmax : dt;
            const uint32_t k = (uint32_t) __float2int_rz(dt);
            kq[r * kChL] = k;
            frq[r * kChL] = __fsub_rn(dt, __uint2float_rn(k));
            const float ph = __fadd_rn(c.phase, c.inc);
            c.phase = ph >= 1.0f ? __fsub_rn(ph, 1.0f) : ph;
        }
    }
}

// The phase recurrence alone over n frames: what an LFO warp runs across the tiles the other LFO warp expands.
__device__ __forceinline__ void chorus_phase_skip(ChorusLane &c, uint32_t n) {
#pragma unroll 8
    for (uint32_t r = 0; r < n; r++) {
        const float ph = __fadd_rn(c.phase, c.inc);
        c.phase = ph >= 1.0f ? __fsub_rn(ph, 1.0f) : ph;
    }
}

// A batch of up to kChB frames of the audio path, in three sweeps so that nothing but the two-op lores~ recurrence is
// serial: (1) all delay~ writes, (2) all reads (the ring has kChB slots more than the longest delay reaches, so a frame's
// taps are never overwritten by the writes of the batch's later frames; a tap at delay k < r of frame r reads what frame
// r-k just wrote, as in the frame-by-frame order), (3) interpolate, filter, mix.  `wp` is the warp-uniform write index
// of the batch's first frame, wp + n <= R.
template <int N>
__device__ __forceinline__ void chorus_audio_batch(float *col, const float *pcol, ChorusLane &c, float *ring_col, uint32_t wp, uint32_t R,
                                                   const uint32_t *kq, const float *frq, uint32_t n) {
    uint32_t k[N];
    float fr[N], x[N], a[N], b[N];
#pragma unroll
    for (int r = 0; r < N; r++) {
        if ((uint32_t) r < n) {
            k[r] = kq[r * kChL];
            fr[r] = frq[r * kChL];
            x[r] = col[r * kChL];
            ring_col[(size_t) (wp + r) * kChL] = pcol[r * kChL];  // delay~ writes (the pitch shifter's output), then reads dt samples back
        }
    }
#pragma unroll
    for (int r = 0; r < N; r++) {
        if ((uint32_t) r < n) {
            const uint32_t w = wp + r;
            const uint32_t ia = w >= k[r] ? w - k[r] : w + R - k[r];
            const uint32_t ib = ia >= 1 ? ia - 1 : R - 1;
            a[r] = ring_col[(size_t) ia * kChL];
            b[r] = ring_col[(size_t) ib * kChL];
        }
    }
#pragma unroll
    for (int r = 0; r < N; r++) {
        if ((uint32_t) r < n) {
            const float wet0 = __fadd_rn(a[r], __fmul_rn(__fsub_rn(b[r], a[r]), fr[r]));
            float t = __fmul_rn(c.scale, wet0);
            t = __fsub_rn(t, __fmul_rn(c.a1, c.y1));
            t = __fsub_rn(t, __fmul_rn(c.a2, c.y2));
            c.y2 = c.y1;
            c.y1 = t;
            col[r * kChL] = __fadd_rn(__fmul_rn(t, c.mix), __fmul_rn(x[r], __fsub_rn(1.0f, c.mix)));
        }
    }
}

__device__ __forceinline__ void mbar_arrive_all(uint64_t *bar) { mbar_arrive(bar); }

// ------------------------------------------------------------------------------------------------
// The gen~ pitch shifter in front of delay~ (modules/rnbo/patcher/pitchshift.gendsp; restated in oracle/olb_oracle.cpp:
// orc_pitch_run).  It is feed-forward -- two cosine-windowed taps of the INPUT at phasor-driven delays of at most one
// window (10 ms) -- so only the phasor is a recurrence, and that one is two operations a frame:
//   k_pitch_phase  one thread per instance walks the phase over the block and leaves it at the start of every 64-frame chunk
//   k_pitch_taps   (instances x chunks) grid: a thread re-walks its chunk's 64 phases (the same operations on the same
//                  values: identical) and reads its taps straight from the input block in HBM / L2, or from the bank's
//                  history of the previous block's last H frames for taps that reach behind the block
//   k_pitch_hist   the last H input frames of this block -> the other history buffer (before the chorus kernel may
//                  overwrite an in-place block)
// The shifted signal goes to a scratch block that k_chorus streams beside the input (delay~ is fed the shifted signal, the
// dry path the input).
constexpr int kPitchChunk = 64;

__device__ __forceinline__ float pitch_cos_poly(float t) {  // orc_cos_poly: cos t on [-pi/2, pi/2]
    const float u = __fmul_rn(t, t);
    float c = 2.0876756987868099e-9f;
    c = __fadd_rn(__fmul_rn(c, u), -2.7557319223985888e-7f);
    c = __fadd_rn(__fmul_rn(c, u), 2.4801587301587302e-5f);
    c = __fadd_rn(__fmul_rn(c, u), -1.3888888888888889e-3f);
    c = __fadd_rn(__fmul_rn(c, u), 4.1666666666666664e-2f);
    c = __fadd_rn(__fmul_rn(c, u), -0.5f);
    c = __fadd_rn(__fmul_rn(c, u), 1.0f);
    return c;
}

__global__ void __launch_bounds__(128) k_pitch_phase(uint32_t frames, uint32_t I, const float *__restrict__ pinc, float *pstate,
                                                     float *__restrict__ chunk_phase) {
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    const float inc = pinc[inst];
    float ph = pstate[inst];
    const uint32_t nchunks = (frames + kPitchChunk - 1) / kPitchChunk;
    if (inc == 0.0f) {  // pitch 0, the patch's default: the phase rests (ph + 0 wraps only from exactly 1, which it never holds)
        for (uint32_t c = 0; c < nchunks; c++) chunk_phase[(size_t) c * I + inst] = ph;
        return;
    }
    for (uint32_t c = 0; c < nchunks; c++) {
        chunk_phase[(size_t) c * I + inst] = ph;
        const uint32_t rows = min((uint32_t) kPitchChunk, frames - c * kPitchChunk);
#pragma unroll 8
        for (uint32_t r = 0; r < rows; r++) {
            ph = __fadd_rn(ph, inc);
            ph = ph >= 1.0f ? __fsub_rn(ph, 1.0f) : ph;
        }
    }
    pstate[inst] = ph;
}

// x[m] for block-relative frame m in [-H, frames): the block itself, or the history rows (row j = frame j - H)
__device__ __forceinline__ float pitch_x(const float *__restrict__ in, const float *__restrict__ hist, uint32_t H, uint32_t I, uint32_t inst, int m) {
    return m >= 0 ? in[(size_t) m * I + inst] : hist[(size_t) (H + m) * I + inst];
}
__global__ void __launch_bounds__(128) k_pitch_taps(const float *__restrict__ in, float *__restrict__ shifted, uint32_t frames, uint32_t I, uint32_t H,
                                                    const float *__restrict__ pinc, float win, const float *__restrict__ chunk_phase,
                                                    const float *__restrict__ hist) {
    const uint32_t inst = blockIdx.x * 32 + (threadIdx.x & 31), c = blockIdx.y * 4 + (threadIdx.x >> 5);
    if (c * kPitchChunk >= frames) return;  // warp-uniform
    const bool live = inst < I;
    const float inc = live ? pinc[inst] : 0.0f;
    float ph = live ? chunk_phase[(size_t) c * I + inst] : 0.0f;
    const uint32_t rows = min((uint32_t) kPitchChunk, frames - c * kPitchChunk);
    // A warp whose 32 phasors all rest (pitch 0: the patch's default) has the same phase, tap times and window weights in
    // every frame: they are computed once -- by the same operations, so the results are the general path's bit for bit --
    // and a frame is four loads, two interpolations and the weighted sum.
    if (__all_sync(0xffffffffu, inc == 0.0f)) {
        if (!live) return;
        float q = __fadd_rn(ph, 0.5f);
        q = q >= 1.0f ? __fsub_rn(q, 1.0f) : q;
        float d1 = __fmul_rn(ph, win), d2 = __fmul_rn(q, win);
        d1 = d1 < 1.0f ? 1.0f : d1; d1 = d1 > 44100.0f ? 44100.0f : d1;
        d2 = d2 < 1.0f ? 1.0f : d2; d2 = d2 > 44100.0f ? 44100.0f : d2;
        const uint32_t k1 = (uint32_t) __float2int_rz(d1), k2 = (uint32_t) __float2int_rz(d2);
        const float f1 = __fsub_rn(d1, __uint2float_rn(k1)), f2 = __fsub_rn(d2, __uint2float_rn(k2));
        const float w1 = pitch_cos_poly(__fmul_rn(__fsub_rn(ph, 0.5f), 3.14159274f)), w2 = pitch_cos_poly(__fmul_rn(__fsub_rn(q, 0.5f), 3.14159274f));
        const int n0 = (int) (c * kPitchChunk);
        // consecutive frames read consecutive rows, and a frame's second neighbour is the previous frame's first: one row
        // per tap and frame, through a pointer that walks the block (ncu on the first version: 91 warp instructions per
        // frame, two thirds of them address arithmetic for four block-or-history loads)
        const bool inblock = __all_sync(__activemask(), n0 - (int) (k1 > k2 ? k1 : k2) - 1 >= 0);
        const float *q1 = in + ((long long) n0 - (long long) k1) * (long long) I + inst, *q2 = in + ((long long) n0 - (long long) k2) * (long long) I + inst;
        float c1 = pitch_x(in, hist, H, I, inst, n0 - (int) k1 - 1), c2 = pitch_x(in, hist, H, I, inst, n0 - (int) k2 - 1);
        float *o = shifted + (size_t) n0 * I + inst;
        constexpr int B = 8;
        for (uint32_t r0 = 0; r0 < rows; r0 += B) {
            float v1[B], v2[B];
            if (inblock && r0 + B <= rows) {
#pragma unroll
                for (int j = 0; j < B; j++) {
                    v1[j] = q1[(size_t) (r0 + j) * I];
                    v2[j] = q2[(size_t) (r0 + j) * I];
                }
#pragma unroll
                for (int j = 0; j < B; j++) {
                    const float t1 = __fadd_rn(v1[j], __fmul_rn(__fsub_rn(j ? v1[j - 1] : c1, v1[j]), f1));
                    const float t2 = __fadd_rn(v2[j], __fmul_rn(__fsub_rn(j ? v2[j - 1] : c2, v2[j]), f2));
                    o[(size_t) (r0 + j) * I] = __fadd_rn(__fmul_rn(t1, w1), __fmul_rn(t2, w2));
                }
            } else {
#pragma unroll
                for (int j = 0; j < B; j++) {
                    if (r0 + j < rows) {
                        v1[j] = pitch_x(in, hist, H, I, inst, n0 + (int) r0 + j - (int) k1);
                        v2[j] = pitch_x(in, hist, H, I, inst, n0 + (int) r0 + j - (int) k2);
                    }
                }
#pragma unroll
                for (int j = 0; j < B; j++) {
                    if (r0 + j < rows) {
                        const float t1 = __fadd_rn(v1[j], __fmul_rn(__fsub_rn(j ? v1[j - 1] : c1, v1[j]), f1));
                        const float t2 = __fadd_rn(v2[j], __fmul_rn(__fsub_rn(j ? v2[j - 1] : c2, v2[j]), f2));
                        o[(size_t) (r0 + j) * I] = __fadd_rn(__fmul_rn(t1, w1), __fmul_rn(t2, w2));
                    }
                }
            }
            c1 = v1[B - 1];
            c2 = v2[B - 1];
        }
        return;
    }
    if (!live) return;
    // batches of 8 frames in three sweeps so that the 32 tap loads of a batch are in flight together: (1) the phase recurrence
    // and the tap times, (2) the loads, (3) interpolation, windows, sum
    constexpr int B = 8;
    for (uint32_t r0 = 0; r0 < rows; r0 += B) {
        float p1[B], p2[B], f1[B], f2[B], a1[B], b1[B], a2[B], b2[B];
        int m1[B], m2[B];
#pragma unroll
        for (int j = 0; j < B; j++) {
            ph = __fadd_rn(ph, inc);
            ph = ph >= 1.0f ? __fsub_rn(ph, 1.0f) : ph;
            p1[j] = ph;
            float q = __fadd_rn(ph, 0.5f);
            p2[j] = q >= 1.0f ? __fsub_rn(q, 1.0f) : q;
            float d1 = __fmul_rn(p1[j], win), d2 = __fmul_rn(p2[j], win);
            d1 = d1 < 1.0f ? 1.0f : d1; d1 = d1 > 44100.0f ? 44100.0f : d1;
            d2 = d2 < 1.0f ? 1.0f : d2; d2 = d2 > 44100.0f ? 44100.0f : d2;
            const uint32_t k1 = (uint32_t) __float2int_rz(d1), k2 = (uint32_t) __float2int_rz(d2);
            f1[j] = __fsub_rn(d1, __uint2float_rn(k1));
            f2[j] = __fsub_rn(d2, __uint2float_rn(k2));
            const int n = (int) (c * kPitchChunk + r0) + j;
            m1[j] = n - (int) k1;
            m2[j] = n - (int) k2;
        }
#pragma unroll
        for (int j = 0; j < B; j++) {
            if (r0 + j < rows) {
                a1[j] = pitch_x(in, hist, H, I, inst, m1[j]); b1[j] = pitch_x(in, hist, H, I, inst, m1[j] - 1);
                a2[j] = pitch_x(in, hist, H, I, inst, m2[j]); b2[j] = pitch_x(in, hist, H, I, inst, m2[j] - 1);
            }
        }
#pragma unroll
        for (int j = 0; j < B; j++) {
            if (r0 + j < rows) {
                const float w1 = pitch_cos_poly(__fmul_rn(__fsub_rn(p1[j], 0.5f), 3.14159274f)), w2 = pitch_cos_poly(__fmul_rn(__fsub_rn(p2[j], 0.5f), 3.14159274f));
                const float t1 = __fadd_rn(a1[j], __fmul_rn(__fsub_rn(b1[j], a1[j]), f1[j])), t2 = __fadd_rn(a2[j], __fmul_rn(__fsub_rn(b2[j], a2[j]), f2[j]));
                shifted[(size_t) (c * kPitchChunk + r0 + j) * I + inst] = __fadd_rn(__fmul_rn(t1, w1), __fmul_rn(t2, w2));
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_pitch_hist(const float *__restrict__ in, uint32_t frames, uint32_t I, uint32_t H,
                                                    const float *__restrict__ hist_old, float *__restrict__ hist_new) {
    const size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t) H * I) return;
    const uint32_t j = (uint32_t) (idx / I), inst = (uint32_t) (idx - (size_t) j * I);
    const long long m = (long long) frames - (long long) H + (long long) j;  // block-relative frame of new row j
    hist_new[idx] = m >= 0 ? in[(size_t) m * I + inst] : hist_old[(size_t) (m + (long long) H) * I + inst];
}

template <int TB, int STAGES, bool TMA, bool ring_in_smem>
__global__ void __launch_bounds__(96) k_chorus(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                               const __grid_constant__ TensorMap tm_p, const float *in, const float *pin, float *out,
                                               uint32_t frames, uint32_t I, uint32_t R,
                                               const float *__restrict__ par, float *state, float *ring_all,
                                               const float *__restrict__ tab_g, uint32_t w0, float dmax) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr int L = kChL;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, i0 = blockIdx.x * L, inst = i0 + lane;
    const bool live = inst < I;
    float *tiles = reinterpret_cast<float *>(smem_raw);          // [STAGES][TB][L]   audio tiles (TMA): input, output in place
    float *ptiles = tiles + (size_t) STAGES * TB * L;            // [2][TB][L]        pitch-shifted input tiles (read only)
    uint32_t *kq = reinterpret_cast<uint32_t *>(ptiles + (size_t) 2 * TB * L);      // [2][TB][L]  whole samples of delay
    float *frq = reinterpret_cast<float *>(kq + (size_t) 2 * TB * L);               // [2][TB][L]  fraction of delay
    // [R][L] ring: shared-memory resident for the launch when it fits (sample rates up to ~53 kHz), else used in place in HBM
    // (same code through generic pointers: the ring is lane-fastest, so its rows stay coalesced there)
    float *ring_s = frq + (size_t) 2 * TB * L;
    float *tab = ring_s + (size_t) (ring_in_smem ? R : 0) * L;   // [513] (+3 pad)
    float *ring = ring_in_smem ? ring_s : ring_all + (size_t) blockIdx.x * R * L;
    uint64_t *full = reinterpret_cast<uint64_t *>(tab + kChTab + 4);  // [STAGES] tile barriers, [STAGES] = ring barrier
    uint64_t *lfo_full = full + STAGES + 1, *lfo_empty = lfo_full + 2;  // [2] each, 32 arrivals per phase
    uint64_t *pfull = lfo_empty + 2;                                    // [2] shifted-tile barriers
    const uint32_t ring_bytes = R * L * (uint32_t) sizeof(float);
    float *ring_g = ring_all + (size_t) blockIdx.x * R * L;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s <= STAGES; s++) mbar_init(&full[s], 1);
        for (int s = 0; s < 2; s++) {
            mbar_init(&lfo_full[s], 32);
            mbar_init(&lfo_empty[s], 32);
            mbar_init(&pfull[s], 1);
        }
        mbar_fence_init();
    }
    for (uint32_t j = threadIdx.x; j <= (uint32_t) kChTab; j += blockDim.x) tab[j] = tab_g[j];
    __syncthreads();
    constexpr uint32_t kTileBytes = TB * L * sizeof(float);
    const uint32_t ntiles = (frames + TB - 1) / TB;
    const uint32_t ci = live ? inst : 0;
    ChorusLane c;
    c.inc = par[(size_t) CP_INC * I + ci]; c.d = par[(size_t) CP_D * I + ci]; c.offset = par[(size_t) CP_OFFSET * I + ci];
    c.scale = par[(size_t) CP_SCALE * I + ci]; c.a1 = par[(size_t) CP_A1 * I + ci]; c.a2 = par[(size_t) CP_A2 * I + ci];
    c.mix = par[(size_t) CP_MIX * I + ci];
    c.phase = state[ci]; c.y1 = state[(size_t) I + ci]; c.y2 = state[2 * (size_t) I + ci];

    if (warp >= 1) {
        // ---- LFO warps 1 and 2: warp 1 expands the even tiles into hand-over buffer 0, warp 2 the odd tiles into buffer 1
        // (table lookups and float <-> int conversions, ~30 instructions per frame); across the other warp's tiles each
        // runs only the 3-operation phase recurrence, so both carry the same phase sequence as the one-warp version. ----
        const uint32_t s = warp - 1;
        for (uint32_t i = 0; i < ntiles; i++) {
            const uint32_t rows = min((uint32_t) TB, frames - i * TB);
            if ((i & 1) != s) {
                if (live) chorus_phase_skip(c, rows);
                continue;
            }
            if (i >= 2) mbar_wait(&lfo_empty[s], ((i >> 1) - 1) & 1);
            uint32_t *kcol = kq + (size_t) s * TB * L + lane;
            float *fcol = frq + (size_t) s * TB * L + lane;
            if (live) {
                uint32_t r = 0;
                for (; r + kChB <= rows; r += kChB) chorus_lfo_batch<kChB>(c, kcol + r * L, fcol + r * L, dmax, tab, kChB);
                if (r < rows) chorus_lfo_batch<kChB>(c, kcol + r * L, fcol + r * L, dmax, tab, rows - r);
            }
            mbar_arrive_all(&lfo_full[s]);
        }
        if (live && warp == 1) state[inst] = c.phase;
        return;
    }

    // ---- audio warp ----
    if (lane == 0) {
        if (ring_in_smem) {
            mbar_arrive_expect_tx(&full[STAGES], ring_bytes);
            bulk_load_1d(ring, ring_g, ring_bytes, &full[STAGES]);
        }
        if (TMA) {
            for (uint32_t j = 0; j < (uint32_t) (STAGES - 1) && j < ntiles; j++) {
                mbar_arrive_expect_tx(&full[j], kTileBytes);
                tma_load_2d(tiles + (size_t) j * TB * L, &tm_in, (int) i0, (int) (j * TB), &full[j]);
            }
            for (uint32_t j = 0; j < 2 && j < ntiles; j++) {
                mbar_arrive_expect_tx(&pfull[j], kTileBytes);
                tma_load_2d(ptiles + (size_t) j * TB * L, &tm_p, (int) i0, (int) (j * TB), &pfull[j]);
            }
        }
    }
    float *ring_col = ring + lane;
    uint32_t wp = w0;
    __syncwarp();
    if (ring_in_smem) mbar_wait(&full[STAGES], 0);

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = TMA ? i % STAGES : 0, ph = (i / STAGES) & 1, ls = i & 1;
        const uint32_t rows = min((uint32_t) TB, frames - i * TB);
        float *col = tiles + (size_t) st * TB * L + lane;
        float *pcol = ptiles + (size_t) (TMA ? ls : 0) * TB * L + lane;
        const uint32_t *kcol = kq + (size_t) ls * TB * L + lane;
        const float *fcol = frq + (size_t) ls * TB * L + lane;
        if (TMA) {
            mbar_wait(&full[st], ph);
            mbar_wait(&pfull[ls], (i >> 1) & 1);
        } else if (live) {
            for (uint32_t r = 0; r < rows; r++) {
                col[r * L] = in[(size_t) (i * TB + r) * I + inst];
                pcol[r * L] = pin[(size_t) (i * TB + r) * I + inst];
            }
        }
        mbar_wait(&lfo_full[ls], (i >> 1) & 1);
        // the write index wraps at most once per tile: one or two straight segments
        for (uint32_t r0 = 0; r0 < rows;) {
            const uint32_t seg = min(rows - r0, R - wp);
            if (live) {
                uint32_t r = 0;
                for (; r + kChB <= seg; r += kChB)
                    chorus_audio_batch<kChB>(col + (r0 + r) * L, pcol + (r0 + r) * L, c, ring_col, wp + r, R, kcol + (r0 + r) * L, fcol + (r0 + r) * L, kChB);
                if (r < seg)
                    chorus_audio_batch<kChB>(col + (r0 + r) * L, pcol + (r0 + r) * L, c, ring_col, wp + r, R, kcol + (r0 + r) * L, fcol + (r0 + r) * L, seg - r);
            }
            r0 += seg;
            wp = wp + seg == R ? 0 : wp + seg;
        }
        mbar_arrive_all(&lfo_empty[ls]);
        if (TMA) {
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                if (i + 2 < ntiles) {  // this tile's shifted input has been read by every lane (__syncwarp above): refill its stage
                    mbar_arrive_expect_tx(&pfull[ls], kTileBytes);
                    tma_load_2d(ptiles + (size_t) ls * TB * L, &tm_p, (int) i0, (int) ((i + 2) * TB), &pfull[ls]);
                }
                tma_store_2d(&tm_out, (int) i0, (int) (i * TB), tiles + (size_t) st * TB * L);
                tma_commit();
                const uint32_t nxt = i + STAGES - 1;
                if (nxt < ntiles) {
                    tma_wait_read<1>();
                    const uint32_t ns = nxt % STAGES;
                    mbar_arrive_expect_tx(&full[ns], kTileBytes);
                    tma_load_2d(tiles + (size_t) ns * TB * L, &tm_in, (int) i0, (int) (nxt * TB), &full[ns]);
                }
            }
        } else if (live) {
            for (uint32_t r = 0; r < rows; r++) out[(size_t) (i * TB + r) * I + inst] = col[r * L];
        }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
        if (ring_in_smem) {
            bulk_store_1d(ring_g, ring, ring_bytes);
            tma_commit();
        }
        tma_wait_all<0>();
    }
    if (live) {
        state[(size_t) I + inst] = c.y1;
        state[2 * (size_t) I + inst] = c.y2;
    }
}

}  // namespace olb

using namespace olb;

// ================================================================================================
// Host side.  The RNBO params (0..1 domain) are kept per instance; the derived device table is rebuilt when a setter ran
// (setters take effect at the next process call).
struct olb_chorus {
    std::recursive_mutex mu;
    StreamOrder order;
    int device = 0;
    uint32_t I = 0, R = 0, w = 0, ctas = 0;
    float sample_rate = 48000.f;
    std::vector<float> raw;  // [7][I]: rate depth cutoff q mix phase pitch
    std::vector<float> stage;
    bool dirty = true;
    float *d_par = nullptr, *d_state = nullptr, *d_ring = nullptr, *d_tab = nullptr;
    // the gen~ pitch shifter: phasor phase [I], the last H input frames [2][H][I] (ping-pong), and per-call scratch that grows
    // on demand: the shifted block [cut frames][I] and the phase at every 64-frame chunk start [chunks][I]
    float *d_pstate = nullptr, *d_phist[2] = {nullptr, nullptr}, *d_shift = nullptr, *d_cphase = nullptr;
    uint32_t H = 0, hist_cur = 0;
    size_t shift_cap = 0, cphase_cap = 0;
    float win = 480.f;
    const char *last_kernel = "none";
    HostStage host;
};

// 3 stages of input / output tiles (a tile is 64 frames = ~5 700 cycles of the audio warp: two tiles ahead is plenty) + 2 of
// shifted-input tiles + 2 x 2 hand-over buffers + the ring
constexpr int kChTB = 64, kChStages = 3;
static size_t chorus_smem(uint32_t R) {  // with the ring resident; R = 0: ring left in HBM
    return (size_t) (kChStages + 2) * kChTB * kChL * sizeof(float) + (size_t) 4 * kChTB * kChL * sizeof(float) + (size_t) R * kChL * sizeof(float) +
           (kChTab + 4) * sizeof(float) + (kChStages + 1 + 6) * sizeof(uint64_t);
}
static bool chorus_ring_fits(uint32_t R) { return chorus_smem(R) <= 227u * 1024u; }
// the specification's ring length (orc_chorus_ring_len: it only enters the arithmetic through the clamp dt <= len - 2) ...
static uint32_t chorus_spec_len(float sr) { return (uint32_t) ceil(2.0 * 12.0 * (double) sr / 1000.0) + 4u; }
// ... and the bank's own: kChB spare slots so that a batch may write ahead of its reads
static uint32_t chorus_ring_len(float sr) { return chorus_spec_len(sr) + (uint32_t) kChB; }

static void chorus_derive(const olb_chorus *h, uint32_t i, float *out /* stride I */) {
    const size_t I = h->I;
    const double sr = h->sample_rate;
    const double rate = h->raw[0 * I + i], depth = h->raw[1 * I + i] < 0.08f ? 0.08 : (double) h->raw[1 * I + i];
    const double cutoff = h->raw[2 * I + i], q = h->raw[3 * I + i], phase = h->raw[5 * I + i];
    const double rate_hz = 0.01 + (0.5 - 0.01) * rate;         // [scale 0 1 0.01 0.5]
    const double depth_ms = 1.0 + (12.0 - 1.0) * depth;        // [scale 0 1 1 12 1]
    const double fc = 300.0 + (15000.0 - 300.0) * cutoff;      // [scale 0 1 300 15000 1]
    const double resterm = exp(q * 0.125) * 0.882497, fqterm = cos(fc * 2.0 * 3.14159265358979323846 / sr);
    const double a1 = -2.0 * resterm * fqterm, a2 = resterm * resterm;
    out[CP_INC * I + i] = (float) (rate_hz / sr);
    out[CP_D * I + i] = (float) (depth_ms * sr / 1000.0);      // [mstosamps]
    out[CP_OFFSET * I + i] = (float) (phase - floor(phase));
    out[CP_SCALE * I + i] = (float) (1.0 + a1 + a2);
    out[CP_A1 * I + i] = (float) a1;
    out[CP_A2 * I + i] = (float) a2;
    out[CP_MIX * I + i] = h->raw[4 * I + i];
    out[CP_PINC * I + i] = (float) ((double) h->raw[6 * I + i] / sr);  // phasor(shift): shift Hz -> phase per frame
}
static uint32_t pitch_hist_len(float sr) {  // orc_pitch_hist_len
    const double win = 10.0 * (double) sr / 1000.0;
    return (uint32_t) (win < 44100.0 ? win : 44100.0) + 2u;
}

static void chorus_free(olb_chorus *h) {
    cudaFree(h->d_par); cudaFree(h->d_state); cudaFree(h->d_ring); cudaFree(h->d_tab);
    cudaFree(h->d_pstate); cudaFree(h->d_phist[0]); cudaFree(h->d_phist[1]); cudaFree(h->d_shift); cudaFree(h->d_cphase);
    h->host.release();
    delete h;
}

static int chorus_reset(olb_chorus *h, float sr) {
    const uint32_t R = chorus_ring_len(sr);
    OLB_ARG(sr > 0.f && sr <= 768000.f, "olb_chorus: sample rate %g out of range", sr);
    if (R != h->R || !h->d_ring) {
        cudaFree(h->d_ring);
        h->d_ring = nullptr;
        OLB_CUDA(cudaMalloc(&h->d_ring, (size_t) h->ctas * R * kChL * sizeof(float)));
        h->R = R;
    }
    const uint32_t H = pitch_hist_len(sr);
    if (H != h->H || !h->d_phist[0]) {
        OLB_CUDA(h->order.sync());
        for (float *&q : h->d_phist) {
            cudaFree(q);
            q = nullptr;
            OLB_CUDA(cudaMalloc(&q, (size_t) H * h->I * sizeof(float)));
        }
        h->H = H;
    }
    h->sample_rate = sr;
    h->win = (float) (10.0 * (double) sr / 1000.0);  // mstosamps(param window 10)
    OLB_CUDA(h->order.memset0(h->d_ring, (size_t) h->ctas * R * kChL * sizeof(float)));
    OLB_CUDA(h->order.memset0(h->d_state, (size_t) 3 * h->I * sizeof(float)));
    OLB_CUDA(h->order.memset0(h->d_pstate, (size_t) h->I * sizeof(float)));
    for (float *q : h->d_phist) OLB_CUDA(h->order.memset0(q, (size_t) H * h->I * sizeof(float)));
    h->hist_cur = 0;
    h->w = 0;
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_chorus_create(int device, uint32_t instances, float sample_rate, olb_chorus **out) {
    OLB_ARG(out, "olb_chorus_create: out is NULL");
    *out = nullptr;
    OLB_ARG(instances > 0, "olb_chorus_create: instances must be > 0");
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_chorus *h = new olb_chorus();
    h->device = device;
    h->I = instances;
    h->ctas = (instances + kChL - 1) / kChL;
    h->raw.resize((size_t) 7 * instances);
    const float defaults[7] = {0.2f, 0.5f, 0.3f, 0.5f, 0.5f, 1.0f, 0.0f};  // param defaults of mono-chorus.rnbopat (pitch: none given = 0)
    for (int p = 0; p < 7; p++)
        for (uint32_t i = 0; i < instances; i++) h->raw[(size_t) p * instances + i] = defaults[p];
    float tab[kChTab + 1];
    for (int i = 0; i <= kChTab; i++) tab[i] = (float) cos(2.0 * 3.14159265358979323846 * (double) i / (double) kChTab);
    cudaError_t e = cudaMalloc(&h->d_par, (size_t) CP_COUNT * instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_state, (size_t) 3 * instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_pstate, (size_t) instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_tab, sizeof(tab));
    if (e == cudaSuccess) e = cudaMemcpy(h->d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        chorus_free(h);
        return fail_cuda(e, "allocating chorus bank", __FILE__, __LINE__);
    }
    rc = chorus_reset(h, sample_rate);
    if (rc) {
        chorus_free(h);
        return rc;
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_chorus_destroy(olb_chorus *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    chorus_free(h);
    return OLB_OK;
}

OLB_API int olb_chorus_init(olb_chorus *h, float sample_rate) {
    OLB_ARG(h, "olb_chorus_init: NULL handle");
    OLB_LOCK(h);
    DeviceGuard g(h->device);
    return chorus_reset(h, sample_rate);
}

OLB_API int olb_chorus_set_params(olb_chorus *h, uint32_t first, uint32_t count, int param, const float *values_host, uint32_t stride) {
    OLB_ARG(h && values_host, "olb_chorus_set_params: NULL argument");
    OLB_LOCK(h);
    OLB_ARG((uint64_t) first + count <= h->I, "olb_chorus_set_params: range [%u, %u) exceeds %u instances", first, first + count, h->I);
    OLB_ARG(param >= OLB_CHORUS_RATE && param <= OLB_CHORUS_PITCH, "olb_chorus_set_params: unknown parameter %d", param);
    const float vmax = param == OLB_CHORUS_PITCH ? 3.f : 1.f;  // param pitch @min 0 @max 3 (Hz of the shifter's phasor)
    for (uint32_t i = 0; i < count; i++) {
        float v = values_host[(size_t) i * stride];
        v = v < 0.f ? 0.f : (v > vmax ? vmax : v);  // RNBO params clip to @min / @max
        h->raw[(size_t) param * h->I + first + i] = v;
    }
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_chorus_set_param(olb_chorus *h, uint32_t first, uint32_t count, int param, float value) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    return olb_chorus_set_params(h, first, count, param, &value, 0);
}

OLB_API const char *olb_chorus_last_kernel(olb_chorus *h) { return h ? h->last_kernel : "none"; }

OLB_API int olb_chorus_process(olb_chorus *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_chorus_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev, "olb_chorus_process: NULL buffer");
    OLB_ARG(frames < (1ull << 32), "olb_chorus_process: at most 2^32-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    OLB_CUDA(h->order.enter(st));
    if (h->dirty) {
        h->stage.resize((size_t) CP_COUNT * h->I);
        for (uint32_t i = 0; i < h->I; i++) chorus_derive(h, i, h->stage.data());
        OLB_CUDA(cudaMemcpyAsync(h->d_par, h->stage.data(), h->stage.size() * sizeof(float), cudaMemcpyHostToDevice, st));
        OLB_CUDA(cudaStreamSynchronize(st));
        h->dirty = false;
    }
    const uint32_t in_smem = chorus_ring_fits(h->R) ? 1u : 0u;
    const size_t smem = chorus_smem(in_smem ? h->R : 0);
    const char *ek = getenv("OLB_CHORUS_KERNEL");  // "direct" forces the plain-load variant (test aid)
    const bool tma = !(ek && !strcmp(ek, "direct")) && (h->I % 4 == 0) && (((uintptr_t) in_dev | (uintptr_t) out_dev) & 15) == 0;
    // Time cuts: the pitch shifter's output is an intermediate block of its own;
    // scratch bounded (256 MiB, whole 64-frame tiles).  Long cuts matter more than L2 residency: every cut is a launch of
    // k_chorus with its ring load / store and pipeline fill (profiles/r02_run19_chorus_cuts.txt: 8 192 instances x 48 000 frames
    // 12.3 / 8.3 / 6.9 / 6.4 / 6.2 ms at cuts of 512 / 1 536 / 4 096 / 12 288 / 48 000 frames).
    const uint64_t I = h->I;
    uint64_t cut = (256ull << 20) / (I * sizeof(float));
    cut = cut < 64 ? 64 : (cut / 64) * 64;
    if (const char *ec = getenv("OLB_CHORUS_CUT")) cut = (uint64_t) atoll(ec) < 1 ? 1 : (uint64_t) atoll(ec);  // test aid: any cut length
    if (cut > frames) cut = frames;
    const uint64_t chunks = (cut + kPitchChunk - 1) / kPitchChunk;
    if (h->shift_cap < cut * I || h->cphase_cap < chunks * I) {
        OLB_CUDA(h->order.sync());
        cudaFree(h->d_shift);
        cudaFree(h->d_cphase);
        h->d_shift = h->d_cphase = nullptr;
        h->shift_cap = h->cphase_cap = 0;
        OLB_CUDA(cudaMalloc(&h->d_shift, cut * I * sizeof(float)));
        OLB_CUDA(cudaMalloc(&h->d_cphase, chunks * I * sizeof(float)));
        h->shift_cap = cut * I;
        h->cphase_cap = chunks * I;
    }
    const float dmax = (float) (chorus_spec_len(h->sample_rate) - 2);
    const float *pinc = h->d_par + (size_t) CP_PINC * I;
    for (uint64_t t0 = 0; t0 < frames; t0 += cut) {
        const uint32_t n = (uint32_t) (frames - t0 < cut ? frames - t0 : cut);
        const float *in_c = in_dev + t0 * I;
        float *out_c = out_dev + t0 * I;
        // the pitch shifter: phases, taps, then the input history for the next cut (before an in-place block is overwritten)
        const uint32_t nch = (n + kPitchChunk - 1) / kPitchChunk;
        k_pitch_phase<<<(h->I + 127) / 128, 128, 0, st>>>(n, h->I, pinc, h->d_pstate, h->d_cphase);
        OLB_LAUNCHED();
        k_pitch_taps<<<dim3(h->ctas, (nch + 3) / 4), 128, 0, st>>>(in_c, h->d_shift, n, h->I, h->H, pinc, h->win, h->d_cphase, h->d_phist[h->hist_cur]);
        OLB_LAUNCHED();
        k_pitch_hist<<<(unsigned) (((size_t) h->H * I + 255) / 256), 256, 0, st>>>(in_c, n, h->I, h->H, h->d_phist[h->hist_cur], h->d_phist[h->hist_cur ^ 1]);
        OLB_LAUNCHED();
        h->hist_cur ^= 1;
        TensorMap tin{}, tout{}, tp{};
        if (tma) {
            int rc = make_tmap_frames(&tin, in_c, n, h->I, kChL, kChTB);
            if (!rc) rc = make_tmap_frames(&tout, out_c, n, h->I, kChL, kChTB);
            if (!rc) rc = make_tmap_frames(&tp, h->d_shift, n, h->I, kChL, kChTB);
            if (rc) return rc;
            auto kern = in_smem ? k_chorus<kChTB, kChStages, true, true> : k_chorus<kChTB, kChStages, true, false>;
            OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
            kern<<<h->ctas, 96, smem, st>>>(tin, tout, tp, in_c, h->d_shift, out_c, n, h->I, h->R, h->d_par, h->d_state, h->d_ring, h->d_tab, h->w, dmax);
            h->last_kernel = "k_chorus_tma";
        } else {
            auto kern = in_smem ? k_chorus<kChTB, kChStages, false, true> : k_chorus<kChTB, kChStages, false, false>;
            OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
            kern<<<h->ctas, 96, smem, st>>>(tin, tout, tp, in_c, h->d_shift, out_c, n, h->I, h->R, h->d_par, h->d_state, h->d_ring, h->d_tab, h->w, dmax);
            h->last_kernel = "k_chorus_direct";
        }
        OLB_LAUNCHED();
        h->w = (uint32_t) ((h->w + n) % h->R);
    }
    return OLB_OK;
}

// Host buffers [frames][instances]: chunked in time through device staging; returns when out_host is complete.
OLB_API int olb_chorus_process_host(olb_chorus *h, const float *in_host, float *out_host, uint64_t frames) {
    OLB_ARG(h, "olb_chorus_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_host && out_host, "olb_chorus_process_host: NULL buffer");
    DeviceGuard g(h->device);
    const uint64_t I = h->I, chunk = host_chunk_frames(I, frames);
    int rc = h->host.ensure(0, chunk * I);
    if (rc) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        OLB_CUDA(cudaMemcpyAsync(h->host.d[0], in_host + t0 * I, n * I * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_chorus_process(h, h->host.d[0], h->host.d[0], n, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(out_host + t0 * I, h->host.d[0], n * I * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}
