// dattorro_smem.cuh -- k_dattorro_smem: the Dattorro plate with its 13 rings RESIDENT IN SHARED MEMORY for the launch and
// the network vectorised over TIME (north_star (3): "per-instance ring buffers staged in
// shared memory").  Reference: libs/dattorro-verb/verb.cpp:258-326.
//
// Why: the lane-per-instance kernels keep the rings in HBM ([ring position][instance]); every one of the 27 reads and 13
// writes of a sample misses the L2 (1.4 GB of rings at 8 192 instances), 170 B of DRAM traffic per instance-sample against
// 12 algorithmic.  Every delay of the network is >= 107 samples, so inside a block of 96 frames no ring read depends on a
// ring write of the same block: the whole frame is ELEMENTWISE over the block -- except the three one-pole low-passes
// (pre-filter, two damping filters), whose inputs (the pre-delay tap and the two 4 453 / 4 217-sample taps) are known
// before the block starts.  So:
//   one CTA = one plate, its rings in shared memory with EXACT lengths (27 390 floats = 107 KiB; two CTAs per SM)
//   warp 4  ("gather")  loads the 96 input frames (three blocks ahead), runs the pre-delay ring and gathers the inputs of the
//                       three low-passes into a double-buffered table
//   warp 0  ("scan")    runs the three recurrences on lanes 0..2, 96 sequential steps (12 cycles each: sub, mul, add, all
//                       rounded as in verb.cpp:131-134) and NOTHING else: the chain is what bounds a plate, so everything
//                       that is not the chain left this warp (with the gathering in it a block took 2 120 cycles, 768 of
//                       them the chain: profiles/r02_run10_dattorro.txt)
//   warps 1-3 ("frame") lane = frame of the block: all tap reads, the four input all-passes, both tank halves, the 14 output
//                       taps, with the low-pass outputs handed over through shared memory
// gather runs up to two blocks ahead of scan, scan up to two ahead of the frame warps (mbarrier pairs per buffer); the taps
// the gatherer reads are >= 4 217 frames old, the pre-delay ring is its own.
// Per-frame operation order inside every filter is the reference's, every operation rounded separately: bit-exact.
// The one-pole chain bounds a plate at one frame per 12 cycles; an SM holds two plates.
//
// State between launches stays in the bank's ordinary format (power-of-two rings in HBM indexed by the 16-bit counter,
// shared with the other kernels): a launch gathers the last L samples of every ring at its start and scatters them back at
// its end (2 x 107 KiB per plate and launch, against 12 B x frames of audio: the kernel is used for long blocks).
#pragma once
#include <type_traits>

#include "common.cuh"

namespace olb {

constexpr int kSmB = 96;  // frames per block (every delay is >= 107): one frame per lane of the three frame warps
constexpr int kSmFrameWarps = kSmB / 32, kSmGatherWarp = 1 + kSmFrameWarps, kSmThreads = 32 * (2 + kSmFrameWarps);
static_assert(kSmB % 32 == 0 && kSmB <= 106, "a block must stay shorter than the shortest delay (107) and fill whole warps");
// exact ring lengths: the delay itself (a frame reads and then rewrites ITS OWN slot), the largest excursion for the two
// modulated all-passes (672 + 16, 908 + 16), delay + block for the pre-delay (the gather warp WRITES the block's inputs and
// then reads: the oldest of the 4 800 samples a frame may ask for must survive the block's own writes)
// (compile-time tables: every ring index below is a constant after unrolling, so the positions live in registers)
struct SmGeom {
    uint32_t len[DV_RINGS], delay[DV_RINGS], base[DV_RINGS];
    uint32_t tap_l[7][2], tap_r[7][2];
    uint32_t total;
};
constexpr SmGeom sm_geom() {
    SmGeom g{{4800 + kSmB, 142, 107, 379, 277, 688, 4453, 1800, 3720, 924, 4217, 2656, 3163},
             {4800, 142, 107, 379, 277, 672, 4453, 1800, 3720, 908, 4217, 2656, 3163},
             {},
             {{DV_PDB, 266}, {DV_PDB, 2974}, {DV_DD2B, 1913}, {DV_POSTB, 1996}, {DV_PDA, 1990}, {DV_DD2A, 187}, {DV_POSTA, 1066}},
             {{DV_PDA, 353}, {DV_PDA, 3627}, {DV_DD2A, 1228}, {DV_POSTA, 2673}, {DV_PDB, 2111}, {DV_DD2B, 335}, {DV_POSTB, 121}},
             0};
    uint32_t n = 0;
    for (int r = 0; r < DV_RINGS; r++) {
        g.base[r] = n;
        n += g.len[r];
    }
    g.total = n;
    return g;
}
// accessors instead of a namespace-scope object: device code may call constexpr functions, not read host variables
template <int R> struct SmRing {
    static constexpr uint32_t len = sm_geom().len[R], delay = sm_geom().delay[R], base = sm_geom().base[R];
};
template <int I> struct SmTap {
    static constexpr int ring_l = (int) sm_geom().tap_l[I][0], ring_r = (int) sm_geom().tap_r[I][0];
    static constexpr uint32_t back_l = sm_geom().tap_l[I][1] - 1u, back_r = sm_geom().tap_r[I][1] - 1u;  // read after t++: one sample less far back
};
constexpr int kSmRingFloats = (int) sm_geom().total;  // 27 326 + block
static_assert(kSmRingFloats == 27326 + kSmB, "ring lengths changed");
constexpr int kSmLpStride = kSmB + 1;                                                                               // padded rows
constexpr size_t kSmBytes = ((size_t) kSmRingFloats + 2 * 3 * kSmLpStride + 2 * 3 * kSmLpStride) * sizeof(float) + 8 * sizeof(uint64_t);

// a shared-memory load the compiler keeps in program order (it otherwise gathers the loads of an unrolled loop at its top)
__device__ __forceinline__ float sm_lds(const float *p) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}

struct SmRings {
    float *base;
    uint32_t pos[DV_RINGS];  // slot of the block's first frame
    __device__ __forceinline__ void init(float *b) {
        base = b;
#pragma unroll
        for (int r = 0; r < DV_RINGS; r++) pos[r] = 0;
    }
    // slot of frame `n` of the block, `back` samples in the past (0 <= n < kSmB, 0 <= back <= len); R is a compile-time ring
    template <int R>
    __device__ __forceinline__ float *at(uint32_t n, uint32_t back) const {
        constexpr uint32_t L = SmRing<R>::len;
        uint32_t i = pos[R] + n + L - back;  // < 3 * len
        if (i >= L) i -= L;
        if (i >= L) i -= L;
        return base + SmRing<R>::base + i;
    }
    template <int R>
    __device__ __forceinline__ void advance1(uint32_t frames) {
        pos[R] += frames;
        if (pos[R] >= SmRing<R>::len) pos[R] -= SmRing<R>::len;
    }
    __device__ __forceinline__ void advance(uint32_t frames);
};

// compile-time loop over the rings / taps
template <int I, int N, class F>
__device__ __forceinline__ void sm_for(F &&f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        sm_for<I + 1, N>(f);
    }
}
__device__ __forceinline__ void SmRings::advance(uint32_t frames) {
    sm_for<0, DV_RINGS>([&](auto rc) { advance1<decltype(rc)::value>(frames); });
}

__global__ void __launch_bounds__(kSmThreads, 2) k_dattorro_smem(const float *in, float *out_l, float *out_r, uint32_t frames, uint32_t I, float *slab,
                                                          float *lp, const float *__restrict__ par, const __grid_constant__ PlateGeom g0,
                                                          uint32_t t0) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm_raw);  // (first: 8-byte aligned whatever the ring lengths)
    uint64_t *full = bars, *empty = bars + 2;         // scan -> frames: the low-pass outputs of a block; frames -> scan: done with them
    uint64_t *in_full = bars + 4, *in_empty = bars + 6;  // gather -> scan: the low-pass inputs of a block; scan -> gather: taken
    float *ring_base = reinterpret_cast<float *>(sm_raw + 8 * sizeof(uint64_t));
    float *lp_in = ring_base + kSmRingFloats;        // [2][3][65] inputs of the three low-passes, double buffered
    float *lp_out = lp_in + 2 * 3 * kSmLpStride;     // [2][3][65] their outputs, double buffered
    const uint32_t inst = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SmRings R;
    R.init(ring_base);

    // ---- gather: slot s of ring r <- the sample written at time t0 - len + s (the launch's first frame goes to slot 0)
    sm_for<0, DV_RINGS>([&](auto rc) {
        constexpr int r = decltype(rc)::value;
        constexpr uint32_t L = SmRing<r>::len;
        for (uint32_t s = threadIdx.x; s < L; s += blockDim.x)  // 4-byte cp.async: all of a thread's ~285 loads in flight at once
            cp_async4(ring_base + SmRing<r>::base + s, slab + (size_t) (g0.base[r] + ((t0 + 0x10000u - L + s) & g0.mask[r])) * I + inst);
    });
    cp_async_commit();
    cp_async_wait<0>();
    if (threadIdx.x == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        mbar_init(&empty[0], kSmFrameWarps);
        mbar_init(&empty[1], kSmFrameWarps);
        mbar_init(&in_full[0], 1);
        mbar_init(&in_full[1], 1);
        mbar_init(&in_empty[0], 1);
        mbar_init(&in_empty[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

    const uint32_t nblocks = (frames + kSmB - 1) / kSmB;
    if (warp == kSmGatherWarp) {
        // ================= gather warp: input, pre-delay ring, the inputs of the three one-poles =================
        const uint32_t pre_delay = (g0.mask[DV_PRE] + 1 - __float_as_uint(par[7 * (size_t) I + inst])) & 0xffffu;  // samples, 0..4800
        // the input of a block is fetched THREE blocks ahead (one frame per lane and half block: two 4-byte loads per block,
        // each from its own DRAM sector -- the launch's input is [frames][instances] -- so their latency must stay off the path)
        struct Xin { float v[kSmFrameWarps]; };
        auto fetch = [&](uint32_t k, Xin &x) {
#pragma unroll
            for (int q = 0; q < kSmFrameWarps; q++) x.v[q] = 0.f;
            if (k < nblocks) {
                const uint32_t nvk = min((uint32_t) kSmB, frames - k * kSmB);
#pragma unroll
                for (int q = 0; q < kSmFrameWarps; q++)
                    if ((uint32_t) lane + 32 * q < nvk) x.v[q] = in[((size_t) k * kSmB + lane + 32 * q) * I + inst];
            }
        };
        Xin x0, x1, x2;
        fetch(0, x0);
        fetch(1, x1);
        fetch(2, x2);
        for (uint32_t k = 0; k < nblocks; k++) {
            const uint32_t buf = k & 1, nv = min((uint32_t) kSmB, frames - k * kSmB);
            Xin x3;
            fetch(k + 3, x3);
            // the block's input into the pre-delay ring (write, then read: verb.cpp:107-110)
#pragma unroll
            for (int q = 0; q < kSmFrameWarps; q++)
                if ((uint32_t) lane + 32 * q < nv) *R.at<DV_PRE>(lane + 32 * q, 0) = x0.v[q];
            x0 = x1; x1 = x2; x2 = x3;
            __syncwarp();
            if (k >= 2) mbar_wait(&in_empty[buf], ((k >> 1) - 1) & 1);  // the scan warp has taken block k - 2's inputs
            float *xi = lp_in + buf * 3 * kSmLpStride;
            for (uint32_t n = lane; n < nv; n += 32) {
                xi[0 * kSmLpStride + n] = *R.at<DV_PRE>(n, pre_delay);
                xi[1 * kSmLpStride + n] = *R.at<DV_PDA>(n, SmRing<DV_PDA>::delay);
                xi[2 * kSmLpStride + n] = *R.at<DV_PDB>(n, SmRing<DV_PDB>::delay);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&in_full[buf]);
            R.advance(nv);
        }
    } else if (warp == 0) {
        // ================= scan warp: the three one-pole recurrences, lanes 0..2 =================
        const float f_pre = par[0 * (size_t) I + inst], f_damp = par[4 * (size_t) I + inst];
        float z = lane < 3 ? lp[(size_t) lane * I + inst] : 0.f;
        const float f = lane == 0 ? f_pre : f_damp;
        for (uint32_t k = 0; k < nblocks; k++) {
            const uint32_t buf = k & 1, nv = min((uint32_t) kSmB, frames - k * kSmB);
            mbar_wait(&in_full[buf], (k >> 1) & 1);
            const float *xi = lp_in + (buf * 3 + (lane < 3 ? lane : 0)) * kSmLpStride;
            float *zo = lp_out + (buf * 3 + (lane < 3 ? lane : 0)) * kSmLpStride;
            if (nv == (uint32_t) kSmB) {
                // a full block: the dependent chain alone, its inputs fetched eight steps ahead in program order (a step
                // leaves three of its four issue slots free; loading the whole block first cost 460 of a block's 1 700
                // cycles, profiles/r02_run61_ncu_k_dattorro_smem.txt)
                if (k >= 2) mbar_wait(&empty[buf], ((k >> 1) - 1) & 1);  // the frame warps are done with block k - 2's outputs
                if (lane < 3) {
                    float xw[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) xw[j] = sm_lds(xi + j);
#pragma unroll
                    for (int j = 0; j < kSmB; j++) {
                        const float xj = xw[j & 7];
                        if (j + 8 < kSmB) xw[j & 7] = sm_lds(xi + j + 8);
                        z = __fadd_rn(z, __fmul_rn(__fsub_rn(xj, z), f));  // LowPassFilter_process, verb.cpp:131-134
                        zo[j] = z;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&in_empty[buf]);
            } else {
                if (k >= 2) mbar_wait(&empty[buf], ((k >> 1) - 1) & 1);
                if (lane < 3) {
                    for (uint32_t n = 0; n < nv; n++) {
                        z = __fadd_rn(z, __fmul_rn(__fsub_rn(xi[n], z), f));
                        zo[n] = z;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&in_empty[buf]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[buf]);
        }
        if (lane < 3) lp[(size_t) lane * I + inst] = z;
    } else {
        // ================= frame warps: lane = frame of the block =================
        const uint32_t n = (uint32_t) (warp - 1) * 32 + lane;
        const float in_diff1 = par[1 * (size_t) I + inst], in_diff2 = par[2 * (size_t) I + inst], decay_diff1 = par[3 * (size_t) I + inst];
        const float decay = par[5 * (size_t) I + inst], decay_diff2 = par[6 * (size_t) I + inst];
        uint32_t off_a = g0.off[DV_DD1A], off_b = g0.off[DV_DD1B];  // excursion state at the block's first frame
        uint32_t t = t0;
        for (uint32_t k = 0; k < nblocks; k++) {
            const uint32_t buf = k & 1, nv = min((uint32_t) kSmB, frames - k * kSmB);
            const bool live = n < nv;
            // the excursion of the two modulated all-passes steps when (t & 0x7ff) == 0, before the frame (verb.cpp:262-270):
            // at most one step falls into a block
            const uint32_t q = (0x800u - (t & 0x7ffu)) & 0x7ffu;  // frames until the next step
            uint32_t na = off_a, nb = off_b;
            if (q < (uint32_t) kSmB) {
                const uint32_t te = (t + q) & 0xffffu;
                if (te < (1u << 15)) { na = (off_a - 1) & 0xffffu; nb = (off_b - 1) & 0xffffu; }
                else { na = (off_a + 1) & 0xffffu; nb = (off_b + 1) & 0xffffu; }
            }
            const uint32_t my_a = n >= q ? na : off_a, my_b = n >= q ? nb : off_b;
            const uint32_t d_a = (g0.mask[DV_DD1A] + 1 - my_a) & g0.mask[DV_DD1A], d_b = (g0.mask[DV_DD1B] + 1 - my_b) & g0.mask[DV_DD1B];
            mbar_wait(&full[buf], (k >> 1) & 1);
            // ---- every ring read of the frame (all of them reach data written before this block) ----
            float r_in0 = 0, r_in1 = 0, r_in2 = 0, r_in3 = 0, r_postb = 0, r_posta = 0, r_dd1a = 0, r_dd1b = 0, r_dd2a = 0, r_dd2b = 0;
            float tl[7], tr[7], z_pre = 0, z_a = 0, z_b = 0;
            if (live) {
                r_in0 = *R.at<DV_IN0>(n, SmRing<DV_IN0>::delay); r_in1 = *R.at<DV_IN1>(n, SmRing<DV_IN1>::delay);
                r_in2 = *R.at<DV_IN2>(n, SmRing<DV_IN2>::delay); r_in3 = *R.at<DV_IN3>(n, SmRing<DV_IN3>::delay);
                r_postb = *R.at<DV_POSTB>(n, SmRing<DV_POSTB>::delay); r_posta = *R.at<DV_POSTA>(n, SmRing<DV_POSTA>::delay);
                r_dd1a = *R.at<DV_DD1A>(n, d_a); r_dd1b = *R.at<DV_DD1B>(n, d_b);
                r_dd2a = *R.at<DV_DD2A>(n, SmRing<DV_DD2A>::delay); r_dd2b = *R.at<DV_DD2B>(n, SmRing<DV_DD2B>::delay);
                sm_for<0, 7>([&](auto ic) {  // the output taps are read after t++ (verb.cpp:302-326): one sample less far back
                    constexpr int i = decltype(ic)::value;
                    tl[i] = *R.at<SmTap<i>::ring_l>(n, SmTap<i>::back_l);
                    tr[i] = *R.at<SmTap<i>::ring_r>(n, SmTap<i>::back_r);
                });
                z_pre = lp_out[(buf * 3 + 0) * kSmLpStride + n];
                z_a = lp_out[(buf * 3 + 1) * kSmLpStride + n];
                z_b = lp_out[(buf * 3 + 2) * kSmLpStride + n];
            }
            // the modulated all-passes read slots that OTHER frames of the block rewrite: all reads before any write
            asm volatile("bar.sync 1, %0;" ::"n"(32 * kSmFrameWarps) : "memory");
            if (live) {
                float w;
                float x = z_pre;
                x = allpass(r_in0, in_diff1, x, w); *R.at<DV_IN0>(n, 0) = w;
                x = allpass(r_in1, in_diff1, x, w); *R.at<DV_IN1>(n, 0) = w;
                x = allpass(r_in2, in_diff2, x, w); *R.at<DV_IN2>(n, 0) = w;
                x = allpass(r_in3, in_diff2, x, w); *R.at<DV_IN3>(n, 0) = w;
                float y = __fadd_rn(x, __fmul_rn(r_postb, decay));
                y = allpass(r_dd1a, -decay_diff1, y, w); *R.at<DV_DD1A>(n, 0) = w;
                *R.at<DV_PDA>(n, 0) = y;
                y = __fmul_rn(z_a, decay);
                y = allpass(r_dd2a, decay_diff2, y, w); *R.at<DV_DD2A>(n, 0) = w;
                *R.at<DV_POSTA>(n, 0) = y;
                y = __fadd_rn(x, __fmul_rn(r_posta, decay));
                y = allpass(r_dd1b, -decay_diff1, y, w); *R.at<DV_DD1B>(n, 0) = w;
                *R.at<DV_PDB>(n, 0) = y;
                y = __fmul_rn(z_b, decay);
                y = allpass(r_dd2b, decay_diff2, y, w); *R.at<DV_DD2B>(n, 0) = w;
                *R.at<DV_POSTB>(n, 0) = y;
                float a = __fadd_rn(tl[0], tl[1]); a = __fsub_rn(a, tl[2]); a = __fadd_rn(a, tl[3]);
                a = __fsub_rn(a, tl[4]); a = __fsub_rn(a, tl[5]); a = __fadd_rn(a, tl[6]);
                float b = __fadd_rn(tr[0], tr[1]); b = __fsub_rn(b, tr[2]); b = __fadd_rn(b, tr[3]);
                b = __fsub_rn(b, tr[4]); b = __fsub_rn(b, tr[5]); b = __fadd_rn(b, tr[6]);
                out_l[((size_t) k * kSmB + n) * I + inst] = a;
                out_r[((size_t) k * kSmB + n) * I + inst] = b;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[buf]);
            R.advance(nv);
            t = (t + nv) & 0xffffu;
            off_a = q < nv ? na : off_a;
            off_b = q < nv ? nb : off_b;
        }
    }
    __syncthreads();
    // ---- scatter: the sample of slot s was written at time t_end - len + ((s - pos) mod len)
    const uint32_t t_end = (t0 + frames) & 0xffffu;
    const uint32_t adv = frames;  // every warp advanced its own copy of the positions; recompute them here
    sm_for<0, DV_RINGS>([&](auto rc) {
        constexpr int r = decltype(rc)::value;
        constexpr uint32_t L = SmRing<r>::len;
        const uint32_t pos = adv % L;
        for (uint32_t s = threadIdx.x; s < L; s += blockDim.x) {
            const uint32_t age = (s + L - pos) % L;  // 0 = the oldest sample kept
            slab[(size_t) (g0.base[r] + ((t_end + 0x10000u - L + age) & g0.mask[r])) * I + inst] = ring_base[SmRing<r>::base + s];
        }
    });
}

}  // namespace olb
