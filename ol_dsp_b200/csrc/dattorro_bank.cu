// dattorro_bank.cu -- (a7) bank of Dattorro plate reverbs.
//
// Reference: libs/dattorro-verb/verb.cpp:258-326 (DattorroVerb_process / _getLeft / _getRight), ring sizes and
// taps :173-222, setters :136-170, state libs/dattorro-verb/verb_structs.h:12-56.  One reference object = one
// mono-in / stereo-out plate with 13 power-of-two ring buffers (42 368 floats), three one-pole low-passes and a
// 16-bit sample counter `t` that also drives a +-16-sample excursion of two all-pass delays every 2 048 samples.
//
// Bank layout in HBM
//   rings   [42 368 ring positions][instances]   instance fastest.  Every instance of a bank is created and ticked
//           together, so they share `t` and therefore every ring index (only the pre-delay length is a per-instance
//           parameter): a warp's 32 instances read and write 128 contiguous bytes per tap -> fully coalesced.
//   lp      [3][instances]   pre-filter and the two damping one-poles
//   par     [8][instances]   pre_filter in_diff1 in_diff2 decay_diff1 damping decay decay_diff2 pre_delay_offset
//   audio   in [frames][instances], out_l / out_r [frames][instances]
//
// Kernel: one instance per thread, time-sequential, every float operation rounded separately in the reference's
// order (bit-exact against verb.cpp compiled with -ffp-contract=off).  All 13 main-tap reads and 14 output-tap reads
// of a sample are independent of that sample's arithmetic (every delay is >= 107 samples), so they are issued
// together at the top of the iteration and the chain of filters then runs out of registers.
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace olb {

enum { DV_PRE, DV_IN0, DV_IN1, DV_IN2, DV_IN3, DV_DD1A, DV_PDA, DV_DD2A, DV_POSTA, DV_DD1B, DV_PDB, DV_DD2B, DV_POSTB, DV_RINGS };

struct PlateGeom {
    uint32_t base[DV_RINGS];  // first ring position of each ring in the slab
    uint32_t mask[DV_RINGS];
    uint32_t off[DV_RINGS];   // main-tap read offset = size - delay (DD1A / DD1B: current excursion value)
    uint32_t tap_l[7][2];     // {ring, read offset}, summed + + - + - - +  (verb.cpp:302-326)
    uint32_t tap_r[7][2];
};

static const uint16_t kDelay[DV_RINGS] = {4800, 142, 107, 379, 277, 672, 4453, 1800, 3720, 908, 4217, 2656, 3163};
static const uint16_t kTapL[7][2] = {{DV_PDB, 266}, {DV_PDB, 2974}, {DV_DD2B, 1913}, {DV_POSTB, 1996}, {DV_PDA, 1990}, {DV_DD2A, 187}, {DV_POSTA, 1066}};
static const uint16_t kTapR[7][2] = {{DV_PDA, 353}, {DV_PDA, 3627}, {DV_DD2A, 1228}, {DV_POSTA, 2673}, {DV_PDB, 2111}, {DV_DD2B, 335}, {DV_POSTB, 121}};

// buffer size = 2^(bit length of the delay)  (DelayBuffer_init, verb.cpp:67-81)
static uint32_t ring_size(uint16_t delay) {
    uint32_t bits = 0;
    for (uint16_t x = delay; x; x >>= 1) bits++;
    return 1u << bits;
}

#define RING(g, ring, idx) slab[(size_t) ((g).base[ring] + ((idx) & (g).mask[ring])) * I + inst]

__device__ __forceinline__ float allpass(float delayed, float gain, float in, float &to_write) {  // verb.cpp:123-128
    in = __fadd_rn(in, __fmul_rn(delayed, -gain));
    to_write = in;
    return __fadd_rn(delayed, __fmul_rn(in, gain));
}
__device__ __forceinline__ float lowpass(float &z, float f, float in) {  // verb.cpp:131-134
    z = __fadd_rn(z, __fmul_rn(__fsub_rn(in, z), f));
    return z;
}

__global__ void __launch_bounds__(64) k_dattorro(const float *__restrict__ in, float *__restrict__ out_l, float *__restrict__ out_r,
                                                 uint32_t frames, uint32_t I, float *slab, float *lp, const float *__restrict__ par,
                                                 const __grid_constant__ PlateGeom g0, uint32_t t0) {
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    const float pre_filter = par[0 * (size_t) I + inst], in_diff1 = par[1 * (size_t) I + inst], in_diff2 = par[2 * (size_t) I + inst];
    const float decay_diff1 = par[3 * (size_t) I + inst], damping = par[4 * (size_t) I + inst], decay = par[5 * (size_t) I + inst];
    const float decay_diff2 = par[6 * (size_t) I + inst];
    const uint32_t pre_off = __float_as_uint(par[7 * (size_t) I + inst]);
    float z_pre = lp[inst], z_a = lp[(size_t) I + inst], z_b = lp[2 * (size_t) I + inst];
    uint32_t off_a = g0.off[DV_DD1A], off_b = g0.off[DV_DD1B];
    uint32_t t = t0;  // 16-bit counter kept in 32 bits, wrapped explicitly
    const bool pre_same = ((pre_off & g0.mask[DV_PRE]) == 0);  // pre-delay of 0 samples reads what was just written

    for (uint32_t n = 0; n < frames; n++) {
        if ((t & 0x07ffu) == 0) {  // excursion (verb.cpp:262-270); offsets are uint16 in the reference
            if (t < (1u << 15)) { off_a = (off_a - 1) & 0xffffu; off_b = (off_b - 1) & 0xffffu; }
            else { off_a = (off_a + 1) & 0xffffu; off_b = (off_b + 1) & 0xffffu; }
        }
        // ---- every read of this sample (none depends on this sample's writes, see header) ----
        const float xin = in[(size_t) n * I + inst];
        const float r_pre = RING(g0, DV_PRE, t + pre_off);
        const float r_in0 = RING(g0, DV_IN0, t + g0.off[DV_IN0]), r_in1 = RING(g0, DV_IN1, t + g0.off[DV_IN1]);
        const float r_in2 = RING(g0, DV_IN2, t + g0.off[DV_IN2]), r_in3 = RING(g0, DV_IN3, t + g0.off[DV_IN3]);
        const float r_postb = RING(g0, DV_POSTB, t + g0.off[DV_POSTB]), r_posta = RING(g0, DV_POSTA, t + g0.off[DV_POSTA]);
        const float r_dd1a = RING(g0, DV_DD1A, t + off_a), r_dd1b = RING(g0, DV_DD1B, t + off_b);
        const float r_pda = RING(g0, DV_PDA, t + g0.off[DV_PDA]), r_pdb = RING(g0, DV_PDB, t + g0.off[DV_PDB]);
        const float r_dd2a = RING(g0, DV_DD2A, t + g0.off[DV_DD2A]), r_dd2b = RING(g0, DV_DD2B, t + g0.off[DV_DD2B]);
        float tl[7], tr[7];
        const uint32_t t1 = (t + 1) & 0xffffu;
#pragma unroll
        for (int i = 0; i < 7; i++) {
            tl[i] = RING(g0, g0.tap_l[i][0], t1 + g0.tap_l[i][1]);
            tr[i] = RING(g0, g0.tap_r[i][0], t1 + g0.tap_r[i][1]);
        }
        // ---- the network (verb.cpp:272-295) ----
        float w;
        RING(g0, DV_PRE, t) = xin;
        float x = pre_same ? xin : r_pre;
        x = lowpass(z_pre, pre_filter, x);
        x = allpass(r_in0, in_diff1, x, w); RING(g0, DV_IN0, t) = w;
        x = allpass(r_in1, in_diff1, x, w); RING(g0, DV_IN1, t) = w;
        x = allpass(r_in2, in_diff2, x, w); RING(g0, DV_IN2, t) = w;
        x = allpass(r_in3, in_diff2, x, w); RING(g0, DV_IN3, t) = w;
        // half 0 (cross feedback from POSTB), then half 1 (from POSTA, written 3 720 samples ago)
        float y = __fadd_rn(x, __fmul_rn(r_postb, decay));
        y = allpass(r_dd1a, -decay_diff1, y, w); RING(g0, DV_DD1A, t) = w;
        RING(g0, DV_PDA, t) = y; y = r_pda;
        y = lowpass(z_a, damping, y);
        y = __fmul_rn(y, decay);
        y = allpass(r_dd2a, decay_diff2, y, w); RING(g0, DV_DD2A, t) = w;
        RING(g0, DV_POSTA, t) = y;
        y = __fadd_rn(x, __fmul_rn(r_posta, decay));
        y = allpass(r_dd1b, -decay_diff1, y, w); RING(g0, DV_DD1B, t) = w;
        RING(g0, DV_PDB, t) = y; y = r_pdb;
        y = lowpass(z_b, damping, y);
        y = __fmul_rn(y, decay);
        y = allpass(r_dd2b, decay_diff2, y, w); RING(g0, DV_DD2B, t) = w;
        RING(g0, DV_POSTB, t) = y;
        t = t1;
        // ---- outputs, read after t++ (verb.cpp:302-326) ----
        float a = __fadd_rn(tl[0], tl[1]); a = __fsub_rn(a, tl[2]); a = __fadd_rn(a, tl[3]);
        a = __fsub_rn(a, tl[4]); a = __fsub_rn(a, tl[5]); a = __fadd_rn(a, tl[6]);
        float b = __fadd_rn(tr[0], tr[1]); b = __fsub_rn(b, tr[2]); b = __fadd_rn(b, tr[3]);
        b = __fsub_rn(b, tr[4]); b = __fsub_rn(b, tr[5]); b = __fadd_rn(b, tr[6]);
        out_l[(size_t) n * I + inst] = a;
        out_r[(size_t) n * I + inst] = b;
    }
    lp[inst] = z_pre;
    lp[(size_t) I + inst] = z_a;
    lp[2 * (size_t) I + inst] = z_b;
}

// ------------------------------------------------------------------------------------------------
// k_dattorro_pf: the same network with a D-sample deep software pipeline.  One instance per thread leaves only
// ~2 warps per SM, so the simple kernel above pays one full HBM round trip per sample.  Every ring read of sample
// n + D is independent of samples n .. n + D - 1 as long as every delay exceeds D (they are >= 107; the per-instance
// pre-delay is checked and falls back to a direct load), so each thread issues its 27 ring reads and its input for
// sample n + D as 4-byte cp.async (LDGSTS) copies into a private shared-memory slot while it computes sample n.
__device__ __forceinline__ void cp_async4(float *smem_dst, const float *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

constexpr int kPfSlots = 28;  // 13 main taps + 14 output taps + the input sample
constexpr int kPfThreads = 64;

__device__ __forceinline__ void excursion_step(uint32_t t, uint32_t &off_a, uint32_t &off_b) {  // verb.cpp:262-270
    if ((t & 0x07ffu) == 0) {
        if (t < (1u << 15)) { off_a = (off_a - 1) & 0xffffu; off_b = (off_b - 1) & 0xffffu; }
        else { off_a = (off_a + 1) & 0xffffu; off_b = (off_b + 1) & 0xffffu; }
    }
}

// L = active lanes per warp: 8 192 instances are only 256 full warps for 592 SM sub-partitions, and each warp is one
// long dependent chain; with L = 16 or 8 (a tap row is then a 64- or 32-byte segment, still whole sectors) every
// sub-partition gets several chains to interleave.
template <int D, int L>
__global__ void __launch_bounds__(kPfThreads) k_dattorro_pf(const float *__restrict__ in, float *__restrict__ out_l, float *__restrict__ out_r,
                                                            uint32_t frames, uint32_t I, float *slab, float *lp, const float *__restrict__ par,
                                                            const __grid_constant__ PlateGeom g0, uint32_t t0) {
    extern __shared__ float pf[];  // [D][kPfSlots][kAct]
    constexpr int kAct = (kPfThreads / 32) * L;  // active threads per CTA
    const uint32_t wic = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t inst = (blockIdx.x * (kPfThreads / 32) + wic) * L + lane;
    if (lane >= (uint32_t) L || inst >= I) return;
    float *mine = pf + wic * L + lane;
    auto SLOT = [&](uint32_t d, int k) -> float * { return mine + ((size_t) d * kPfSlots + k) * kAct; };
    const float pre_filter = par[0 * (size_t) I + inst], in_diff1 = par[1 * (size_t) I + inst], in_diff2 = par[2 * (size_t) I + inst];
    const float decay_diff1 = par[3 * (size_t) I + inst], damping = par[4 * (size_t) I + inst], decay = par[5 * (size_t) I + inst];
    const float decay_diff2 = par[6 * (size_t) I + inst];
    const uint32_t pre_off = __float_as_uint(par[7 * (size_t) I + inst]);
    float z_pre = lp[inst], z_a = lp[(size_t) I + inst], z_b = lp[2 * (size_t) I + inst];
    const uint32_t pre_len = (g0.mask[DV_PRE] + 1 - pre_off) & g0.mask[DV_PRE];  // pre-delay in samples
    const bool pre_same = pre_len == 0, pre_direct = pre_len <= (uint32_t) D;    // too short to fetch D samples ahead
    uint32_t t = t0, off_a = g0.off[DV_DD1A], off_b = g0.off[DV_DD1B];           // state at the sample being computed
    uint32_t ta = t0, offa_a = off_a, offb_a = off_b;                            // state at the sample being fetched

    auto fetch = [&](uint32_t n, uint32_t slot) {  // sample index n, counter value ta (before its own excursion step)
        excursion_step(ta, offa_a, offb_a);
        cp_async4(SLOT(slot, 27), in + (size_t) n * I + inst);
        cp_async4(SLOT(slot, 0), &RING(g0, DV_PRE, ta + pre_off));
        cp_async4(SLOT(slot, 1), &RING(g0, DV_IN0, ta + g0.off[DV_IN0]));
        cp_async4(SLOT(slot, 2), &RING(g0, DV_IN1, ta + g0.off[DV_IN1]));
        cp_async4(SLOT(slot, 3), &RING(g0, DV_IN2, ta + g0.off[DV_IN2]));
        cp_async4(SLOT(slot, 4), &RING(g0, DV_IN3, ta + g0.off[DV_IN3]));
        cp_async4(SLOT(slot, 5), &RING(g0, DV_POSTB, ta + g0.off[DV_POSTB]));
        cp_async4(SLOT(slot, 6), &RING(g0, DV_POSTA, ta + g0.off[DV_POSTA]));
        cp_async4(SLOT(slot, 7), &RING(g0, DV_DD1A, ta + offa_a));
        cp_async4(SLOT(slot, 8), &RING(g0, DV_DD1B, ta + offb_a));
        cp_async4(SLOT(slot, 9), &RING(g0, DV_PDA, ta + g0.off[DV_PDA]));
        cp_async4(SLOT(slot, 10), &RING(g0, DV_PDB, ta + g0.off[DV_PDB]));
        cp_async4(SLOT(slot, 11), &RING(g0, DV_DD2A, ta + g0.off[DV_DD2A]));
        cp_async4(SLOT(slot, 12), &RING(g0, DV_DD2B, ta + g0.off[DV_DD2B]));
        const uint32_t t1 = (ta + 1) & 0xffffu;
#pragma unroll
        for (int i = 0; i < 7; i++) {
            cp_async4(SLOT(slot, 13 + i), &RING(g0, g0.tap_l[i][0], t1 + g0.tap_l[i][1]));
            cp_async4(SLOT(slot, 20 + i), &RING(g0, g0.tap_r[i][0], t1 + g0.tap_r[i][1]));
        }
        ta = t1;
    };

    for (uint32_t d = 0; d < (uint32_t) D; d++) {
        if (d < frames) fetch(d, d);
        cp_async_commit();
    }
    for (uint32_t n = 0; n < frames; n++) {
        const uint32_t slot = n % D;
        cp_async_wait<D - 1>();
        excursion_step(t, off_a, off_b);
        float v[kPfSlots];
#pragma unroll
        for (int k = 0; k < kPfSlots; k++) v[k] = *SLOT(slot, k);
        if (n + D < frames) fetch(n + D, slot);
        cp_async_commit();
        const float xin = v[27];
        float r_pre = v[0];
        if (pre_direct) r_pre = pre_same ? xin : RING(g0, DV_PRE, t + pre_off);
        float w;
        RING(g0, DV_PRE, t) = xin;
        float x = lowpass(z_pre, pre_filter, r_pre);
        x = allpass(v[1], in_diff1, x, w); RING(g0, DV_IN0, t) = w;
        x = allpass(v[2], in_diff1, x, w); RING(g0, DV_IN1, t) = w;
        x = allpass(v[3], in_diff2, x, w); RING(g0, DV_IN2, t) = w;
        x = allpass(v[4], in_diff2, x, w); RING(g0, DV_IN3, t) = w;
        float y = __fadd_rn(x, __fmul_rn(v[5], decay));
        y = allpass(v[7], -decay_diff1, y, w); RING(g0, DV_DD1A, t) = w;
        RING(g0, DV_PDA, t) = y; y = v[9];
        y = lowpass(z_a, damping, y);
        y = __fmul_rn(y, decay);
        y = allpass(v[11], decay_diff2, y, w); RING(g0, DV_DD2A, t) = w;
        RING(g0, DV_POSTA, t) = y;
        y = __fadd_rn(x, __fmul_rn(v[6], decay));
        y = allpass(v[8], -decay_diff1, y, w); RING(g0, DV_DD1B, t) = w;
        RING(g0, DV_PDB, t) = y; y = v[10];
        y = lowpass(z_b, damping, y);
        y = __fmul_rn(y, decay);
        y = allpass(v[12], decay_diff2, y, w); RING(g0, DV_DD2B, t) = w;
        RING(g0, DV_POSTB, t) = y;
        t = (t + 1) & 0xffffu;
        float a = __fadd_rn(v[13], v[14]); a = __fsub_rn(a, v[15]); a = __fadd_rn(a, v[16]);
        a = __fsub_rn(a, v[17]); a = __fsub_rn(a, v[18]); a = __fadd_rn(a, v[19]);
        float b = __fadd_rn(v[20], v[21]); b = __fsub_rn(b, v[22]); b = __fadd_rn(b, v[23]);
        b = __fsub_rn(b, v[24]); b = __fsub_rn(b, v[25]); b = __fadd_rn(b, v[26]);
        out_l[(size_t) n * I + inst] = a;
        out_r[(size_t) n * I + inst] = b;
    }
    cp_async_wait<0>();
    lp[inst] = z_pre;
    lp[(size_t) I + inst] = z_a;
    lp[2 * (size_t) I + inst] = z_b;
}

// ------------------------------------------------------------------------------------------------
// k_dattorro_rp: k_dattorro_pf with the ring addressing taken out of the frame loop.  ncu on k_dattorro_pf: 531 warp
// instructions per frame, 72 % of them index and address arithmetic (LEA / IADD / IMAD / LOP / constant loads) for 41 ring
// accesses that each recompute (t + offset) & mask, + base, * instances, + lane.  Every index but the per-instance
// pre-delay tap is warp-uniform and advances by one per frame, so time is cut into BLOCKS inside which no ring index
// wraps and no excursion step falls (<= 64 frames; the smallest ring has 128 slots): per block one pointer per access
// is set up, per frame an access is one IMAD.WIDE (frame-in-block x row stride + pointer) and the memory instruction.
// Same reads, same arithmetic, same writes as k_dattorro_pf: bit-exact.
struct PlateTap {
    int ring;
    int kind;  // 0: main tap, offset g.off[ring]; 1 / 2: main tap with the excursion offset a / b; 3: output tap (reads at t + 1)
    int off;   // kind 3: index into tap_l (0..6) / tap_r (7..13)
};

template <int D>
__global__ void __launch_bounds__(kPfThreads) k_dattorro_rp(const float *__restrict__ in, float *__restrict__ out_l, float *__restrict__ out_r,
                                                            uint32_t frames, uint32_t I, float *slab, float *lp, const float *__restrict__ par,
                                                            const __grid_constant__ PlateGeom g0, uint32_t t0) {
    extern __shared__ float pf[];  // [D][kPfSlots][kPfThreads]
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    float *mine = pf + threadIdx.x;
    auto SLOT = [&](uint32_t d, int k) -> float * { return mine + ((size_t) d * kPfSlots + k) * kPfThreads; };
    const float pre_filter = par[0 * (size_t) I + inst], in_diff1 = par[1 * (size_t) I + inst], in_diff2 = par[2 * (size_t) I + inst];
    const float decay_diff1 = par[3 * (size_t) I + inst], damping = par[4 * (size_t) I + inst], decay = par[5 * (size_t) I + inst];
    const float decay_diff2 = par[6 * (size_t) I + inst];
    const uint32_t pre_off = __float_as_uint(par[7 * (size_t) I + inst]);
    float z_pre = lp[inst], z_a = lp[(size_t) I + inst], z_b = lp[2 * (size_t) I + inst];
    const uint32_t pre_len = (g0.mask[DV_PRE] + 1 - pre_off) & g0.mask[DV_PRE];  // pre-delay in samples
    const bool pre_same = pre_len == 0, pre_direct = pre_len <= (uint32_t) D;    // too short to fetch D samples ahead
    const uint32_t stride = I * (uint32_t) sizeof(float);                        // bytes from one ring position / frame to the next
    char *lane_slab = reinterpret_cast<char *>(slab + inst);
    const char *lane_in = reinterpret_cast<const char *>(in + inst);
    char *lane_l = reinterpret_cast<char *>(out_l + inst), *lane_r = reinterpret_cast<char *>(out_r + inst);
    auto at = [&](uint32_t ring, uint32_t idx) -> char * { return lane_slab + (uint64_t) (g0.base[ring] + idx) * stride; };
    auto row = [&](char *p, uint32_t r) -> float * { return reinterpret_cast<float *>(p + (uint64_t) r * stride); };

    // the 12 warp-uniform main taps in slot order 1..12, then the 14 output taps in slot order 13..26 (slot 0 = pre-delay
    // tap, per lane; slot 27 = the input sample)
    constexpr int kMain[12][2] = {{DV_IN0, 0}, {DV_IN1, 0}, {DV_IN2, 0}, {DV_IN3, 0}, {DV_POSTB, 0}, {DV_POSTA, 0}, {DV_DD1A, 1}, {DV_DD1B, 2},
                                  {DV_PDA, 0}, {DV_PDB, 0}, {DV_DD2A, 0}, {DV_DD2B, 0}};
    uint32_t t = t0, ta = t0, off_a = g0.off[DV_DD1A], off_b = g0.off[DV_DD1B];  // ta / off_*: state of the FETCH stream

    // prologue: the first D frames are fetched with plain per-frame index arithmetic
    for (uint32_t d = 0; d < (uint32_t) D; d++) {
        if (d < frames) {
            excursion_step(ta, off_a, off_b);
            cp_async4(SLOT(d, 27), in + (size_t) d * I + inst);
            cp_async4(SLOT(d, 0), row(at(DV_PRE, (ta + pre_off) & g0.mask[DV_PRE]), 0));
#pragma unroll
            for (int k = 0; k < 12; k++) {
                const uint32_t off = kMain[k][1] == 1 ? off_a : (kMain[k][1] == 2 ? off_b : g0.off[kMain[k][0]]);
                cp_async4(SLOT(d, 1 + k), row(at(kMain[k][0], (ta + off) & g0.mask[kMain[k][0]]), 0));
            }
            const uint32_t t1 = (ta + 1) & 0xffffu;
#pragma unroll
            for (int i = 0; i < 7; i++) {
                cp_async4(SLOT(d, 13 + i), row(at(g0.tap_l[i][0], (t1 + g0.tap_l[i][1]) & g0.mask[g0.tap_l[i][0]]), 0));
                cp_async4(SLOT(d, 20 + i), row(at(g0.tap_r[i][0], (t1 + g0.tap_r[i][1]) & g0.mask[g0.tap_r[i][0]]), 0));
            }
            ta = t1;
        }
        cp_async_commit();
    }

    constexpr int kTapRingL[7] = {DV_PDB, DV_PDB, DV_DD2B, DV_POSTB, DV_PDA, DV_DD2A, DV_POSTA};  // = kTapL[i][0]
    constexpr int kTapRingR[7] = {DV_PDA, DV_PDA, DV_DD2A, DV_POSTA, DV_PDB, DV_DD2B, DV_POSTB};  // = kTapR[i][0]
    uint32_t slot = 0;
    for (uint32_t n = 0; n < frames;) {
        // ---- block set-up, all warp-uniform: B frames during which no ring index wraps and no excursion step falls.
        // ta is the counter of frame n + D (the one fetched while frame n is computed).
        uint32_t B = min(frames - n, 64u);
        {
            const uint32_t e = ta & 0x7ffu;
            B = min(B, e ? 0x800u - e : 0x800u);
        }
        excursion_step(ta, off_a, off_b);  // acts only when (ta & 0x7ff) == 0, which can only be a block's first frame
        char *wp[DV_RINGS], *rp[12], *opl[7], *opr[7];
#pragma unroll
        for (int ring = 0; ring < DV_RINGS; ring++) {
            const uint32_t idx = t & g0.mask[ring];
            B = min(B, g0.mask[ring] + 1 - idx);
            wp[ring] = at(ring, idx);
        }
#pragma unroll
        for (int k = 0; k < 12; k++) {
            const int ring = kMain[k][0];
            const uint32_t off = kMain[k][1] == 1 ? off_a : (kMain[k][1] == 2 ? off_b : g0.off[ring]);
            const uint32_t idx = (ta + off) & g0.mask[ring];
            B = min(B, g0.mask[ring] + 1 - idx);
            rp[k] = at(ring, idx);
        }
#pragma unroll
        for (int i = 0; i < 7; i++) {
            const uint32_t il = (ta + 1 + g0.tap_l[i][1]) & g0.mask[kTapRingL[i]], ir = (ta + 1 + g0.tap_r[i][1]) & g0.mask[kTapRingR[i]];
            B = min(B, min(g0.mask[kTapRingL[i]] + 1 - il, g0.mask[kTapRingR[i]] + 1 - ir));
            opl[i] = at(kTapRingL[i], il);
            opr[i] = at(kTapRingR[i], ir);
        }
        // ---- the frames of the block ----
        for (uint32_t r = 0; r < B; r++) {
            cp_async_wait<D - 1>();
            float v[kPfSlots];
#pragma unroll
            for (int k = 0; k < kPfSlots; k++) v[k] = *SLOT(slot, k);
            if (n + r + D < frames) {
                cp_async4(SLOT(slot, 27), reinterpret_cast<const float *>(lane_in + (uint64_t) (n + r + D) * stride));
                cp_async4(SLOT(slot, 0), row(at(DV_PRE, (ta + r + pre_off) & g0.mask[DV_PRE]), 0));
#pragma unroll
                for (int k = 0; k < 12; k++) cp_async4(SLOT(slot, 1 + k), row(rp[k], r));
#pragma unroll
                for (int i = 0; i < 7; i++) {
                    cp_async4(SLOT(slot, 13 + i), row(opl[i], r));
                    cp_async4(SLOT(slot, 20 + i), row(opr[i], r));
                }
            }
            cp_async_commit();
            const float xin = v[27];
            float r_pre = v[0];
            if (pre_direct) r_pre = pre_same ? xin : *row(at(DV_PRE, (t + r + pre_off) & g0.mask[DV_PRE]), 0);
            float w;
            *row(wp[DV_PRE], r) = xin;
            float x = lowpass(z_pre, pre_filter, r_pre);
            x = allpass(v[1], in_diff1, x, w); *row(wp[DV_IN0], r) = w;
            x = allpass(v[2], in_diff1, x, w); *row(wp[DV_IN1], r) = w;
            x = allpass(v[3], in_diff2, x, w); *row(wp[DV_IN2], r) = w;
            x = allpass(v[4], in_diff2, x, w); *row(wp[DV_IN3], r) = w;
            float y = __fadd_rn(x, __fmul_rn(v[5], decay));
            y = allpass(v[7], -decay_diff1, y, w); *row(wp[DV_DD1A], r) = w;
            *row(wp[DV_PDA], r) = y; y = v[9];
            y = lowpass(z_a, damping, y);
            y = __fmul_rn(y, decay);
            y = allpass(v[11], decay_diff2, y, w); *row(wp[DV_DD2A], r) = w;
            *row(wp[DV_POSTA], r) = y;
            y = __fadd_rn(x, __fmul_rn(v[6], decay));
            y = allpass(v[8], -decay_diff1, y, w); *row(wp[DV_DD1B], r) = w;
            *row(wp[DV_PDB], r) = y; y = v[10];
            y = lowpass(z_b, damping, y);
            y = __fmul_rn(y, decay);
            y = allpass(v[12], decay_diff2, y, w); *row(wp[DV_DD2B], r) = w;
            *row(wp[DV_POSTB], r) = y;
            float a = __fadd_rn(v[13], v[14]); a = __fsub_rn(a, v[15]); a = __fadd_rn(a, v[16]);
            a = __fsub_rn(a, v[17]); a = __fsub_rn(a, v[18]); a = __fadd_rn(a, v[19]);
            float b = __fadd_rn(v[20], v[21]); b = __fsub_rn(b, v[22]); b = __fadd_rn(b, v[23]);
            b = __fsub_rn(b, v[24]); b = __fsub_rn(b, v[25]); b = __fadd_rn(b, v[26]);
            *row(lane_l, n + r) = a;
            *row(lane_r, n + r) = b;
            slot = slot + 1 == (uint32_t) D ? 0 : slot + 1;
        }
        n += B;
        t = (t + B) & 0xffffu;
        ta = (ta + B) & 0xffffu;
    }
    cp_async_wait<0>();
    lp[inst] = z_pre;
    lp[(size_t) I + inst] = z_a;
    lp[2 * (size_t) I + inst] = z_b;
}

// ------------------------------------------------------------------------------------------------
// k_dattorro_split: the network cut over the four warps of a CTA (32 instances).  One warp running the whole plate
// spends ~710 cycles per frame on one dependent chain (pre-filter -> four input diffusers -> tank half A -> tank half B ->
// output sums) while 8 192 instances are only 256 warps for 592 schedulers.  The pieces are coupled loosely:
//   warp 0  input    pre-delay, pre-filter, the four input diffusers            -> x, handed on through shared memory
//   warp 1  tank A   x + decay * (POSTB tap), diffuser 1A, delay, damping, diffuser 2A, delay     (verb.cpp:281-287)
//   warp 2  tank B   the same with A <-> B                                                        (verb.cpp:288-294)
//   warp 3  output   the 14 output taps and the two sums                                          (verb.cpp:302-326)
// x is the only same-frame dependency; the two halves see each other only through taps >= 3 163 frames back and the output
// taps are >= 120 frames back, so it suffices that no warp runs more than kXs tiles (of 32 frames) ahead of another and
// that the output warp starts a tile after both tank warps finished it.  Hand-over per tile with mbarriers, one arrival
// per warp (lane 0 after __syncwarp): x_full (input -> tanks), done (tanks -> output), x_empty (tanks + output -> input).
// Ring addressing is k_dattorro_rp's (blocks inside which no ring index wraps), every warp fetching only ITS taps D frames
// ahead.  Same reads, same arithmetic, same writes: bit-exact.
constexpr int kSplitSlots = 14;  // most taps any one warp fetches (the output warp)
constexpr int kXs = 4;           // x tiles in flight
constexpr int kXt = 32;          // frames per tile

template <int D>
__global__ void __launch_bounds__(128) k_dattorro_split(const float *in, float *out_l, float *out_r, uint32_t frames, uint32_t I, float *slab,
                                                        float *lp, const float *__restrict__ par, const __grid_constant__ PlateGeom g0,
                                                        uint32_t t0) {
    extern __shared__ __align__(16) float sm[];  // [D][kSplitSlots][128] prefetch slots, [kXs][kXt][32] x tiles, barriers
    float *xs = sm + (size_t) D * kSplitSlots * 128;
    uint64_t *x_full = reinterpret_cast<uint64_t *>(xs + (size_t) kXs * kXt * 32), *x_empty = x_full + kXs, *done = x_empty + kXs;
    const uint32_t lane = threadIdx.x & 31, role = threadIdx.x >> 5, inst = blockIdx.x * 32 + lane;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kXs; s++) {
            mbar_init(&x_full[s], 1);
            mbar_init(&x_empty[s], 3);
            mbar_init(&done[s], 2);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (inst >= I) return;  // lane 0 of every warp is live; arrivals are lane 0's
    float *mine = sm + threadIdx.x;
    auto SLOT = [&](uint32_t d, int k) -> float * { return mine + ((size_t) d * kSplitSlots + k) * 128; };
    const uint32_t stride = I * (uint32_t) sizeof(float);
    char *lane_slab = reinterpret_cast<char *>(slab + inst);
    auto at = [&](uint32_t ring, uint32_t idx) -> char * { return lane_slab + (uint64_t) (g0.base[ring] + idx) * stride; };
    auto row = [&](char *p, uint32_t r) -> float * { return reinterpret_cast<float *>(p + (uint64_t) r * stride); };
    constexpr int kTapRingL[7] = {DV_PDB, DV_PDB, DV_DD2B, DV_POSTB, DV_PDA, DV_DD2A, DV_POSTA};
    constexpr int kTapRingR[7] = {DV_PDA, DV_PDA, DV_DD2A, DV_POSTA, DV_PDB, DV_DD2B, DV_POSTB};
    constexpr int kMainRings[12] = {DV_IN0, DV_IN1, DV_IN2, DV_IN3, DV_POSTB, DV_POSTA, DV_DD1A, DV_DD1B, DV_PDA, DV_PDB, DV_DD2A, DV_DD2B};

    uint32_t t = t0, ta = t0, off_a = g0.off[DV_DD1A], off_b = g0.off[DV_DD1B];  // ta / off_*: state of the FETCH stream
    // frames from n on that share one set of ring pointers: no ring index of any warp wraps, no excursion step, no tile edge
    auto block_len = [&](uint32_t n) -> uint32_t {
        uint32_t B = min(frames - n, (uint32_t) kXt - (n & (kXt - 1)));
        const uint32_t e = ta & 0x7ffu;
        B = min(B, e ? 0x800u - e : 0x800u);
#pragma unroll
        for (int ring = 0; ring < DV_RINGS; ring++) B = min(B, g0.mask[ring] + 1 - (t & g0.mask[ring]));
#pragma unroll
        for (int k = 0; k < 12; k++) {
            const int ring = kMainRings[k];
            const uint32_t off = ring == DV_DD1A ? off_a : (ring == DV_DD1B ? off_b : g0.off[ring]);
            B = min(B, g0.mask[ring] + 1 - ((ta + off) & g0.mask[ring]));
        }
#pragma unroll
        for (int i = 0; i < 7; i++) {
            B = min(B, g0.mask[kTapRingL[i]] + 1 - ((ta + 1 + g0.tap_l[i][1]) & g0.mask[kTapRingL[i]]));
            B = min(B, g0.mask[kTapRingR[i]] + 1 - ((ta + 1 + g0.tap_r[i][1]) & g0.mask[kTapRingR[i]]));
        }
        return B;
    };
    auto tile_wait = [&](uint64_t *bar, uint32_t tile, bool previous_phase) {
        const uint32_t turn = tile / kXs;
        mbar_wait(&bar[tile % kXs], (previous_phase ? turn - 1 : turn) & 1);
    };
    auto tile_arrive = [&](uint64_t *bar, uint32_t tile) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar[tile % kXs]);
    };
    uint32_t slot = 0;

    if (role == 0) {  // ---------------- input section
        const float pre_filter = par[0 * (size_t) I + inst], in_diff1 = par[1 * (size_t) I + inst], in_diff2 = par[2 * (size_t) I + inst];
        const uint32_t pre_off = __float_as_uint(par[7 * (size_t) I + inst]);
        float z_pre = lp[inst];
        const uint32_t pre_len = (g0.mask[DV_PRE] + 1 - pre_off) & g0.mask[DV_PRE];
        const bool pre_same = pre_len == 0, pre_direct = pre_len <= (uint32_t) D;
        const char *lane_in = reinterpret_cast<const char *>(in + inst);
        constexpr int kRings[4] = {DV_IN0, DV_IN1, DV_IN2, DV_IN3};
        for (uint32_t d = 0; d < (uint32_t) D; d++) {
            if (d < frames) {
                excursion_step(ta, off_a, off_b);
                cp_async4(SLOT(d, 0), in + (size_t) d * I + inst);
                cp_async4(SLOT(d, 1), row(at(DV_PRE, (ta + pre_off) & g0.mask[DV_PRE]), 0));
#pragma unroll
                for (int k = 0; k < 4; k++) cp_async4(SLOT(d, 2 + k), row(at(kRings[k], (ta + g0.off[kRings[k]]) & g0.mask[kRings[k]]), 0));
                ta = (ta + 1) & 0xffffu;
            }
            cp_async_commit();
        }
        for (uint32_t n = 0; n < frames;) {
            excursion_step(ta, off_a, off_b);  // acts only when (ta & 0x7ff) == 0, which can only be a block's first frame
            const uint32_t B = block_len(n), tile = n / kXt;
            if ((n & (kXt - 1)) == 0 && tile >= (uint32_t) kXs) tile_wait(x_empty, tile, true);
            char *wp[5], *rp[4];
            wp[0] = at(DV_PRE, t & g0.mask[DV_PRE]);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                wp[1 + k] = at(kRings[k], t & g0.mask[kRings[k]]);
                rp[k] = at(kRings[k], (ta + g0.off[kRings[k]]) & g0.mask[kRings[k]]);
            }
            float *xrow = xs + ((size_t) (tile % kXs) * kXt + (n & (kXt - 1))) * 32 + lane;
            for (uint32_t r = 0; r < B; r++) {
                cp_async_wait<D - 1>();
                float v[6];
#pragma unroll
                for (int k = 0; k < 6; k++) v[k] = *SLOT(slot, k);
                if (n + r + D < frames) {
                    cp_async4(SLOT(slot, 0), reinterpret_cast<const float *>(lane_in + (uint64_t) (n + r + D) * stride));
                    cp_async4(SLOT(slot, 1), row(at(DV_PRE, (ta + r + pre_off) & g0.mask[DV_PRE]), 0));
#pragma unroll
                    for (int k = 0; k < 4; k++) cp_async4(SLOT(slot, 2 + k), row(rp[k], r));
                }
                cp_async_commit();
                const float xin = v[0];
                float r_pre = v[1];
                if (pre_direct) r_pre = pre_same ? xin : *row(at(DV_PRE, (t + r + pre_off) & g0.mask[DV_PRE]), 0);
                float w;
                *row(wp[0], r) = xin;
                float x = lowpass(z_pre, pre_filter, r_pre);
                x = allpass(v[2], in_diff1, x, w); *row(wp[1], r) = w;
                x = allpass(v[3], in_diff1, x, w); *row(wp[2], r) = w;
                x = allpass(v[4], in_diff2, x, w); *row(wp[3], r) = w;
                x = allpass(v[5], in_diff2, x, w); *row(wp[4], r) = w;
                xrow[(size_t) r * 32] = x;
                slot = slot + 1 == (uint32_t) D ? 0 : slot + 1;
            }
            n += B;
            t = (t + B) & 0xffffu;
            ta = (ta + B) & 0xffffu;
            if ((n & (kXt - 1)) == 0 || n == frames) tile_arrive(x_full, tile);
        }
        cp_async_wait<0>();
        lp[inst] = z_pre;
    } else if (role <= 2) {  // ---------------- tank half A (role 1) / B (role 2)
        const bool A = role == 1;
        const float decay_diff1 = par[3 * (size_t) I + inst], damping = par[4 * (size_t) I + inst], decay = par[5 * (size_t) I + inst];
        const float decay_diff2 = par[6 * (size_t) I + inst];
        float z = lp[(size_t) (A ? 1 : 2) * I + inst];
        const int r_cross = A ? DV_POSTB : DV_POSTA, r_dd1 = A ? DV_DD1A : DV_DD1B, r_pd = A ? DV_PDA : DV_PDB, r_dd2 = A ? DV_DD2A : DV_DD2B,
                  r_post = A ? DV_POSTA : DV_POSTB;
        auto fetch_idx = [&](int k) -> uint32_t {  // ring index of tap k of the frame whose counter is ta
            const int ring = k == 0 ? r_cross : k == 1 ? r_dd1 : k == 2 ? r_pd : r_dd2;
            const uint32_t off = k == 1 ? (A ? off_a : off_b) : g0.off[ring];
            return (ta + off) & g0.mask[ring];
        };
        auto tap_ring = [&](int k) -> int { return k == 0 ? r_cross : k == 1 ? r_dd1 : k == 2 ? r_pd : r_dd2; };
        for (uint32_t d = 0; d < (uint32_t) D; d++) {
            if (d < frames) {
                excursion_step(ta, off_a, off_b);
#pragma unroll
                for (int k = 0; k < 4; k++) cp_async4(SLOT(d, k), row(at(tap_ring(k), fetch_idx(k)), 0));
                ta = (ta + 1) & 0xffffu;
            }
            cp_async_commit();
        }
        for (uint32_t n = 0; n < frames;) {
            excursion_step(ta, off_a, off_b);  // acts only when (ta & 0x7ff) == 0, which can only be a block's first frame
            const uint32_t B = block_len(n), tile = n / kXt;
            if ((n & (kXt - 1)) == 0) tile_wait(x_full, tile, false);
            char *wp[4], *rp[4];
            wp[0] = at(r_dd1, t & g0.mask[r_dd1]);
            wp[1] = at(r_pd, t & g0.mask[r_pd]);
            wp[2] = at(r_dd2, t & g0.mask[r_dd2]);
            wp[3] = at(r_post, t & g0.mask[r_post]);
#pragma unroll
            for (int k = 0; k < 4; k++) rp[k] = at(tap_ring(k), fetch_idx(k));
            const float *xrow = xs + ((size_t) (tile % kXs) * kXt + (n & (kXt - 1))) * 32 + lane;
            for (uint32_t r = 0; r < B; r++) {
                cp_async_wait<D - 1>();
                float v[4];
#pragma unroll
                for (int k = 0; k < 4; k++) v[k] = *SLOT(slot, k);
                if (n + r + D < frames) {
#pragma unroll
                    for (int k = 0; k < 4; k++) cp_async4(SLOT(slot, k), row(rp[k], r));
                }
                cp_async_commit();
                const float x = xrow[(size_t) r * 32];
                float w;
                float y = __fadd_rn(x, __fmul_rn(v[0], decay));
                y = allpass(v[1], -decay_diff1, y, w); *row(wp[0], r) = w;
                *row(wp[1], r) = y; y = v[2];
                y = lowpass(z, damping, y);
                y = __fmul_rn(y, decay);
                y = allpass(v[3], decay_diff2, y, w); *row(wp[2], r) = w;
                *row(wp[3], r) = y;
                slot = slot + 1 == (uint32_t) D ? 0 : slot + 1;
            }
            n += B;
            t = (t + B) & 0xffffu;
            ta = (ta + B) & 0xffffu;
            if ((n & (kXt - 1)) == 0 || n == frames) {
                tile_arrive(done, tile);
                if (lane == 0) mbar_arrive(&x_empty[tile % kXs]);
            }
        }
        cp_async_wait<0>();
        lp[(size_t) (A ? 1 : 2) * I + inst] = z;
    } else {  // ---------------- output taps
        char *lane_l = reinterpret_cast<char *>(out_l + inst), *lane_r = reinterpret_cast<char *>(out_r + inst);
        for (uint32_t d = 0; d < (uint32_t) D; d++) {
            if (d < frames) {
                excursion_step(ta, off_a, off_b);
                const uint32_t t1 = (ta + 1) & 0xffffu;
#pragma unroll
                for (int i = 0; i < 7; i++) {
                    cp_async4(SLOT(d, i), row(at(kTapRingL[i], (t1 + g0.tap_l[i][1]) & g0.mask[kTapRingL[i]]), 0));
                    cp_async4(SLOT(d, 7 + i), row(at(kTapRingR[i], (t1 + g0.tap_r[i][1]) & g0.mask[kTapRingR[i]]), 0));
                }
                ta = t1;
            }
            cp_async_commit();
        }
        for (uint32_t n = 0; n < frames;) {
            excursion_step(ta, off_a, off_b);  // acts only when (ta & 0x7ff) == 0, which can only be a block's first frame
            const uint32_t B = block_len(n), tile = n / kXt;
            if ((n & (kXt - 1)) == 0) tile_wait(done, tile, false);
            char *opl[7], *opr[7];
#pragma unroll
            for (int i = 0; i < 7; i++) {
                opl[i] = at(kTapRingL[i], (ta + 1 + g0.tap_l[i][1]) & g0.mask[kTapRingL[i]]);
                opr[i] = at(kTapRingR[i], (ta + 1 + g0.tap_r[i][1]) & g0.mask[kTapRingR[i]]);
            }
            for (uint32_t r = 0; r < B; r++) {
                cp_async_wait<D - 1>();
                float v[14];
#pragma unroll
                for (int k = 0; k < 14; k++) v[k] = *SLOT(slot, k);
                if (n + r + D < frames) {
#pragma unroll
                    for (int i = 0; i < 7; i++) {
                        cp_async4(SLOT(slot, i), row(opl[i], r));
                        cp_async4(SLOT(slot, 7 + i), row(opr[i], r));
                    }
                }
                cp_async_commit();
                float a = __fadd_rn(v[0], v[1]); a = __fsub_rn(a, v[2]); a = __fadd_rn(a, v[3]);
                a = __fsub_rn(a, v[4]); a = __fsub_rn(a, v[5]); a = __fadd_rn(a, v[6]);
                float b = __fadd_rn(v[7], v[8]); b = __fsub_rn(b, v[9]); b = __fadd_rn(b, v[10]);
                b = __fsub_rn(b, v[11]); b = __fsub_rn(b, v[12]); b = __fadd_rn(b, v[13]);
                *row(lane_l, n + r) = a;
                *row(lane_r, n + r) = b;
                slot = slot + 1 == (uint32_t) D ? 0 : slot + 1;
            }
            n += B;
            t = (t + B) & 0xffffu;
            ta = (ta + B) & 0xffffu;
            if ((n & (kXt - 1)) == 0 || n == frames) tile_arrive(x_empty, tile);
        }
        cp_async_wait<0>();
    }
}

}  // namespace olb

#include "dattorro_smem.cuh"

using namespace olb;

// ================================================================================================
// Host side.  Parameter setters mirror DattorroVerb_set* (verb.cpp:136-170); the maths runs here, on the host.
struct olb_dattorro {
    std::recursive_mutex mu;
    StreamOrder order;
    int device = 0;
    uint32_t I = 0;
    PlateGeom geom{};
    uint32_t slab_positions = 0;
    uint32_t t = 0;  // the reference's uint16 sample counter, shared by every instance of the bank
    std::vector<float> par_host;  // [8][I]
    bool par_dirty = true;
    float *d_slab = nullptr, *d_lp = nullptr, *d_par = nullptr;
    const char *last_kernel = "none";
    HostStage host;
};

// banks up to this size take k_dattorro_smem for long blocks.  Measured (profiles/r02_run62_dattorro_sizes.txt, 48 000-frame
// blocks): the resident kernel runs 1.9-2.2e10 instance-samples/s from 296 instances up (two plates per SM, bounded by the
// one-pole recurrences), the HBM-ring kernel 1.8e9 at 296, 6.2e9 at 1 024, 1.48e10 at 2 560, 2.4e10 from 4 736 up: they cross
// near 3 500 instances.
constexpr uint32_t kSmemAutoMaxInstances = 3456;

static void plate_geometry(PlateGeom *g, uint32_t *positions) {
    uint32_t n = 0;
    for (int i = 0; i < DV_RINGS; i++) {
        const uint32_t sz = ring_size(kDelay[i]);
        g->base[i] = n;
        g->mask[i] = sz - 1;
        g->off[i] = sz - kDelay[i];
        n += sz;
    }
    for (int i = 0; i < 7; i++) {
        g->tap_l[i][0] = kTapL[i][0];
        g->tap_l[i][1] = g->mask[kTapL[i][0]] + 1 - kTapL[i][1];
        g->tap_r[i][0] = kTapR[i][0];
        g->tap_r[i][1] = g->mask[kTapR[i][0]] + 1 - kTapR[i][1];
    }
    *positions = n;
}

// DattorroVerb_setDecay (verb.cpp:162-165): clamp() takes t_sample arguments and returns double.
static float decay_diffusion2(float decay) {
    const float x = (float) ((double) decay + 0.15);
    const float lo = 0.25f, hi = 0.50f;
    return x < lo ? lo : (x > hi ? hi : x);
}
// DattorroVerb_setPreDelay (verb.cpp:137-139): value * MAX_PREDELAY narrowed to uint16, offset = mask + 1 - delay (uint16)
static float predelay_offset_bits(const PlateGeom &g, float value) {
    const uint16_t delay = (uint16_t) (value * 4800);
    const uint16_t off = (uint16_t) (g.mask[DV_PRE] + 1 - delay);
    const uint32_t u = off;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static void plate_defaults(olb_dattorro *h) {  // initialize(), verb.cpp:215-221
    const uint32_t I = h->I;
    h->par_host.assign((size_t) 8 * I, 0.f);
    for (uint32_t i = 0; i < I; i++) {
        h->par_host[0 * (size_t) I + i] = 0.85;
        h->par_host[1 * (size_t) I + i] = 0.75;
        h->par_host[2 * (size_t) I + i] = 0.625;
        h->par_host[3 * (size_t) I + i] = 0.70;
        h->par_host[4 * (size_t) I + i] = 0.95;
        h->par_host[5 * (size_t) I + i] = 0.75;
        h->par_host[6 * (size_t) I + i] = decay_diffusion2(0.75);
        h->par_host[7 * (size_t) I + i] = predelay_offset_bits(h->geom, 0.1);
    }
    h->par_dirty = true;
}

OLB_API int olb_dattorro_create(int device, uint32_t instances, olb_dattorro **out) {
    OLB_ARG(out, "olb_dattorro_create: out is NULL");
    *out = nullptr;
    OLB_ARG(instances > 0, "olb_dattorro_create: instances must be > 0");
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_dattorro *h = new olb_dattorro();
    h->device = device;
    h->I = instances;
    plate_geometry(&h->geom, &h->slab_positions);
    plate_defaults(h);
    const size_t slab_bytes = (size_t) h->slab_positions * instances * sizeof(float);
    cudaError_t e = cudaMalloc(&h->d_slab, slab_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_lp, (size_t) 3 * instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_par, (size_t) 8 * instances * sizeof(float));
    if (e == cudaSuccess) e = cudaMemset(h->d_slab, 0, slab_bytes);
    if (e == cudaSuccess) e = cudaMemset(h->d_lp, 0, (size_t) 3 * instances * sizeof(float));
    if (e != cudaSuccess) {
        cudaFree(h->d_slab); cudaFree(h->d_lp); cudaFree(h->d_par);
        delete h;
        return fail_cuda(e, "allocating plate bank", __FILE__, __LINE__);
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_dattorro_destroy(olb_dattorro *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    cudaFree(h->d_slab); cudaFree(h->d_lp); cudaFree(h->d_par);
    h->host.release();
    delete h;
    return OLB_OK;
}

// DattorroVerb_create()/initialize() again: rings, filters, counter and excursion cleared, parameters back to defaults.
OLB_API int olb_dattorro_init(olb_dattorro *h) {
    OLB_ARG(h, "olb_dattorro_init: NULL handle");
    OLB_LOCK(h);
    DeviceGuard g(h->device);
    OLB_CUDA(h->order.memset0(h->d_slab, (size_t) h->slab_positions * h->I * sizeof(float)));
    OLB_CUDA(h->order.memset0(h->d_lp, (size_t) 3 * h->I * sizeof(float)));
    plate_geometry(&h->geom, &h->slab_positions);
    h->t = 0;
    plate_defaults(h);
    return OLB_OK;
}

OLB_API int olb_dattorro_set_params(olb_dattorro *h, uint32_t first, uint32_t count, int param, const float *values, uint32_t stride) {
    OLB_ARG(h && values, "olb_dattorro_set_params: NULL argument");
    OLB_LOCK(h);
    OLB_ARG((uint64_t) first + count <= h->I, "olb_dattorro_set_params: range [%u, %u) exceeds %u instances", first, first + count, h->I);
    OLB_ARG(param >= OLB_PLATE_PREDELAY && param <= OLB_PLATE_DAMPING, "olb_dattorro_set_params: unknown parameter %d", param);
    const size_t I = h->I;
    for (uint32_t i = 0; i < count; i++) {
        const float v = values[(size_t) i * stride];
        const size_t c = first + i;
        switch (param) {
            case OLB_PLATE_PREDELAY: h->par_host[7 * I + c] = predelay_offset_bits(h->geom, v); break;
            case OLB_PLATE_PREFILTER: h->par_host[0 * I + c] = v; break;
            case OLB_PLATE_INPUT_DIFFUSION1: h->par_host[1 * I + c] = v; break;
            case OLB_PLATE_INPUT_DIFFUSION2: h->par_host[2 * I + c] = v; break;
            case OLB_PLATE_DECAY: h->par_host[5 * I + c] = v; h->par_host[6 * I + c] = decay_diffusion2(v); break;
            case OLB_PLATE_DECAY_DIFFUSION: h->par_host[3 * I + c] = v; break;
            default: h->par_host[4 * I + c] = v; break;
        }
    }
    h->par_dirty = true;
    return OLB_OK;
}

OLB_API const char *olb_dattorro_last_kernel(olb_dattorro *h) { return h ? h->last_kernel : "none"; }

OLB_API int olb_dattorro_set_param(olb_dattorro *h, uint32_t first, uint32_t count, int param, float value) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    return olb_dattorro_set_params(h, first, count, param, &value, 0);
}

OLB_API int olb_dattorro_process(olb_dattorro *h, const float *in_dev, float *out_l_dev, float *out_r_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_dattorro_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_l_dev && out_r_dev, "olb_dattorro_process: NULL buffer");
    OLB_ARG(frames < (1ull << 32), "olb_dattorro_process: at most 2^32-1 frames per call");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    OLB_CUDA(h->order.enter(st));
    if (h->par_dirty) {
        OLB_CUDA(cudaMemcpyAsync(h->d_par, h->par_host.data(), h->par_host.size() * sizeof(float), cudaMemcpyHostToDevice, st));
        OLB_CUDA(cudaStreamSynchronize(st));  // par_host may change right after we return
        h->par_dirty = false;
    }
    // OLB_PLATE_KERNEL=simple|pf|rp selects the one-frame-at-a-time / the index-arithmetic pipelined / the one-warp ring-pointer
    // kernel (default: k_dattorro_split);
    // OLB_PLATE_LANES=32|16|8 overrides the active lanes per warp of the pf one (default 32: narrower warps measured no faster, profiles/r01_run14_units_lane_sweep.jsonl)
    const char *ek = getenv("OLB_PLATE_KERNEL"), *el = getenv("OLB_PLATE_LANES");
    const int forced = el ? atoi(el) : 0;
    const int lanes = (forced == 32 || forced == 16 || forced == 8) ? forced : 32;
    // k_dattorro_smem (rings resident in shared memory, time-vectorised): long blocks only -- a launch gathers and scatters
    // 107 KiB of rings per plate, which a block of >= 8 192 frames amortises to < 2.3 B per instance-sample each way.
    // OLB_PLATE_KERNEL=smem forces it, any other value of the variable keeps it off.
    const bool smem_kernel = ek ? !strcmp(ek, "smem") : (frames >= 8192 && h->I <= kSmemAutoMaxInstances);
    if (smem_kernel) {
        OLB_CUDA(cudaFuncSetAttribute(k_dattorro_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kSmBytes));
        OLB_CUDA(cudaFuncSetAttribute(k_dattorro_smem, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        k_dattorro_smem<<<h->I, kSmThreads, kSmBytes, st>>>(in_dev, out_l_dev, out_r_dev, (uint32_t) frames, h->I, h->d_slab, h->d_lp, h->d_par, h->geom,
                                                   h->t);
        h->last_kernel = "k_dattorro_smem";
    } else if (ek && !strcmp(ek, "simple")) {
        h->last_kernel = "k_dattorro";
        k_dattorro<<<(h->I + 63) / 64, 64, 0, st>>>(in_dev, out_l_dev, out_r_dev, (uint32_t) frames, h->I, h->d_slab, h->d_lp, h->d_par, h->geom, h->t);
    } else if (ek && !strcmp(ek, "pf")) {
        h->last_kernel = "k_dattorro_pf";
        constexpr int D = 8;
        const unsigned warps = (h->I + lanes - 1) / lanes, grid = (warps + kPfThreads / 32 - 1) / (kPfThreads / 32);
        const size_t smem = (size_t) D * kPfSlots * (kPfThreads / 32) * lanes * sizeof(float);
#define OLB_PLATE_GO(LN)                                                                                                          \
    do {                                                                                                                          \
        OLB_CUDA(cudaFuncSetAttribute(k_dattorro_pf<D, LN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));            \
        k_dattorro_pf<D, LN><<<grid, kPfThreads, smem, st>>>(in_dev, out_l_dev, out_r_dev, (uint32_t) frames, h->I, h->d_slab, h->d_lp, \
                                                             h->d_par, h->geom, h->t);                                            \
    } while (0)
        if (lanes == 32) OLB_PLATE_GO(32); else if (lanes == 16) OLB_PLATE_GO(16); else OLB_PLATE_GO(8);
#undef OLB_PLATE_GO
    } else if (!ek || strcmp(ek, "rp")) {  // default: the network cut over the four warps of a CTA
        h->last_kernel = "k_dattorro_split";
        constexpr int D = 8;
        OLB_ARG((uint64_t) frames * h->I * sizeof(float) < (1ull << 62), "olb_dattorro_process: block too large");
        constexpr size_t smem = ((size_t) D * kSplitSlots * 128 + (size_t) kXs * kXt * 32) * sizeof(float) + 3 * kXs * sizeof(uint64_t);
        OLB_CUDA(cudaFuncSetAttribute(k_dattorro_split<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_dattorro_split<D><<<(h->I + 31) / 32, 128, smem, st>>>(in_dev, out_l_dev, out_r_dev, (uint32_t) frames, h->I, h->d_slab, h->d_lp,
                                                               h->d_par, h->geom, h->t);
    } else {  // OLB_PLATE_KERNEL=rp: one warp per 32 instances, block-wise ring pointers
        h->last_kernel = "k_dattorro_rp";
        constexpr int D = 8;
        OLB_ARG((uint64_t) frames * h->I * sizeof(float) < (1ull << 62), "olb_dattorro_process: block too large");
        constexpr size_t smem = (size_t) D * kPfSlots * kPfThreads * sizeof(float);
        OLB_CUDA(cudaFuncSetAttribute(k_dattorro_rp<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_dattorro_rp<D><<<(h->I + kPfThreads - 1) / kPfThreads, kPfThreads, smem, st>>>(in_dev, out_l_dev, out_r_dev, (uint32_t) frames, h->I,
                                                                                       h->d_slab, h->d_lp, h->d_par, h->geom, h->t);
    }
    OLB_LAUNCHED();
    // advance the shared counter and the excursion exactly as the kernel did
    uint32_t t = h->t, oa = h->geom.off[DV_DD1A], ob = h->geom.off[DV_DD1B];
    for (uint64_t n = 0; n < frames;) {
        if ((t & 0x07ffu) == 0) {
            if (t < (1u << 15)) { oa = (oa - 1) & 0xffffu; ob = (ob - 1) & 0xffffu; }
            else { oa = (oa + 1) & 0xffffu; ob = (ob + 1) & 0xffffu; }
        }
        const uint64_t to_next = 0x800u - (t & 0x07ffu);  // frames until the next excursion step
        const uint64_t step = frames - n < to_next ? frames - n : to_next;
        n += step;
        t = (uint32_t) ((t + step) & 0xffffu);
    }
    h->t = t;
    h->geom.off[DV_DD1A] = oa;
    h->geom.off[DV_DD1B] = ob;
    return OLB_OK;
}

// Host buffers, all [frames][instances]: chunked in time through device staging; returns when the outputs are complete.
OLB_API int olb_dattorro_process_host(olb_dattorro *h, const float *in_host, float *out_l_host, float *out_r_host, uint64_t frames) {
    OLB_ARG(h, "olb_dattorro_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_host && out_l_host && out_r_host, "olb_dattorro_process_host: NULL buffer");
    DeviceGuard g(h->device);
    const uint64_t I = h->I, chunk = host_chunk_frames(I, frames);
    for (int k = 0; k < 3; k++) {
        int rc = h->host.ensure(k, chunk * I);
        if (rc) return rc;
    }
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        OLB_CUDA(cudaMemcpyAsync(h->host.d[0], in_host + t0 * I, n * I * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        int rc = olb_dattorro_process(h, h->host.d[0], h->host.d[1], h->host.d[2], n, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(out_l_host + t0 * I, h->host.d[1], n * I * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
        OLB_CUDA(cudaMemcpyAsync(out_r_host + t0 * I, h->host.d[2], n * I * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}
