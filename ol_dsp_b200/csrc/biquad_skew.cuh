// biquad_skew.cuh -- k_bq_skew: the skewed section-pair cascade kernel.
//
// Why: what limits the section-pipelined kernel (k_bq_sys) is not HBM but (a) the shared-memory round trip of every
// section hop (8 hops x 2 x 4 B per channel-frame = 128 B/clk/SM of shared-memory bandwidth is the wall)
// and (b) the dependent FMA chain of a warp that has its sub-partition to itself.  Here one WARP runs TWO consecutive
// sections of its 32 channels in the two halves of ONE packed fp32x2 instruction stream:
//     low half  = section 2w   at stream position m
//     high half = section 2w+1 at stream position m - 2      (its input is the low half's output of two steps ago)
// so a hop through shared memory happens once per section PAIR, a warp-frame costs 4 FFMA2 (FAST) for 2 sections x 32
// channels, the FFMA2 recurrence of one step overlaps the next one (the two-step lag keeps the low -> high hand-over off
// the critical path), and the unit of scheduling is 32 channels x 2 sections = 12 pipe-cycles: 4 096 channels put one
// warp on each of 512 SM sub-partitions, 65 536 channels are 8 192 such units (13.8 per sub-partition).
// Per-section operation order is untouched, so results are bit-identical to every other kernel of the bank (and to
// the oracle) in all three modes.
//
// Stream positions.  Each warp delays its stream by kLag = 2 positions, so with W = S/2 warps the tile that arrives as
// frames [i*TB, (i+1)*TB) leaves as frames [i*TB - 2W, (i+1)*TB - 2W): the TMA store is issued at that (possibly
// negative) row coordinate -- rows outside [0, frames) are clipped by the hardware -- and the stream is 2W positions
// longer than the block (the tail loads are out of bounds and read as zeros).  Warp w's low half carries real frames at
// positions [2w, 2w + F), its high half at [2w + 2, 2w + 2 + F); outside these a half is FROZEN (state restored after the
// step) so that at the end of the launch every node of the cascade has seen exactly frames [0, F) and the history table
// is in the bank's ordinary node format.  Only the first and last tiles take the freezing path.
#pragma once
#include "biquad_sys.cuh"

namespace olb {


template <int MODE>
__device__ __forceinline__ u64 skew_section(u64 v, u64 x1, u64 x2, u64 y1, u64 y2, const u64 (&k)[5], u64 one, u64 negzero) {
    return Lane<2>::template section<MODE>(v, x1, x2, y1, y2, k, one, negzero);
}

// Shared-memory accesses the compiler must keep in program order relative to one another (it otherwise sinks the
// prefetches down to their uses and the warp waits out the full load latency every frame).
__device__ __forceinline__ float lds_ordered(const float *p) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts_ordered(float *p, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(smem_u32(p)), "f"(v) : "memory");
}

__device__ __forceinline__ u64 sel_halves(bool lo_new, bool hi_new, u64 fresh, u64 old) {
    float fl, fh, ol, oh;
    unpack2(fresh, fl, fh);
    unpack2(old, ol, oh);
    return pack2(lo_new ? fl : ol, hi_new ? fh : oh);
}

// One tile in which both halves carry real frames throughout, fully unrolled into one basic block.  The input of step
// j + D is fetched from shared memory during step j (one load between the arithmetic of every step: back-to-back memory
// instructions of one warp issue four cycles apart, interleaved ones do not wait), so a warp that has its scheduler to
// itself still keeps the FMA pipe fed.
template <int MODE, int TB, int D, int kSkewLag, bool GAIN>
__device__ __forceinline__ void skew_tile(float *col, const u64 (&k)[5], u64 gain2, u64 &x1, u64 &x2, u64 &y1, u64 &y2,
                                          float (&p)[kSkewLag], u64 one, u64 negzero) {
    constexpr int CH = 32;
    float x[TB];
#pragma unroll
    for (int j = 0; j < D && j < TB; j++) x[j] = lds_ordered(col + j * CH);
#pragma unroll
    for (int j = 0; j < TB; j++) {
        if (j + D < TB) x[j + D] = lds_ordered(col + (j + D) * CH);
        // warp 0 of a FAST bank applies G = prod(b0) to its input: (x, p) * (G, 1) as one packed product (p * 1 is exact)
        u64 v = pack2(x[j], p[kSkewLag - 1]);
        if (GAIN) v = fmul2(v, gain2);
        const u64 y = skew_section<MODE>(v, x1, x2, y1, y2, k, one, negzero);
        x2 = x1; x1 = v; y2 = y1; y1 = y;
        float ylo, yhi;
        unpack2(y, ylo, yhi);
#pragma unroll
        for (int q = kSkewLag - 1; q > 0; q--) p[q] = p[q - 1];
        p[0] = ylo;
        sts_ordered(col + j * CH, yhi);
    }
}

// FAST tier, hand software-pipelined: the four fused operations of a step are issued for four DIFFERENT stream
// positions (A for m + 3, B for m + 2, C for m + 1, Y for m), so that every operand was produced at least three
// instructions earlier and a lone warp issues back to back; the pipeline fills at the top of a tile and drains at its end.
// Same operations on the same operands as skew_tile<OLB_MODE_FAST>: bit-identical results.
//     A = fma(r1, V[m-1], V[m])   B = fma(r2, V[m-2], A)   C = fma(-a2, Y[m-2], B)   Y = fma(-a1, Y[m-1], C)
template <int TB, int D, int kSkewLag, bool GAIN>
__device__ __forceinline__ void skew_tile_fast_unrolled(float *col, const u64 (&k)[5], float gain, u64 &x1, u64 &x2, u64 &y1, u64 &y2,
                                               float (&p)[kSkewLag], u64, u64) {
    constexpr int CH = 32, P = 3;  // pipeline depth in stream positions
    static_assert(kSkewLag >= P + 1, "the low half's output must be complete before the high half consumes it");
    float x[TB];
    u64 V[TB + 2], Y[TB + 2], A[TB], B[TB], C[TB];  // V[j + 2], Y[j + 2] = position j of this tile; [0], [1] = history
    V[1] = x1; V[0] = x2; Y[1] = y1; Y[0] = y2;
#pragma unroll
    for (int j = 0; j < D + P && j < TB; j++) {
        x[j] = lds_ordered(col + j * CH);
        if (GAIN) x[j] = __fmul_rn(gain, x[j]);  // warp 0 of a FAST bank applies G = prod(b0) to its input
    }
    // (Scaling the inputs in pairs with one packed product, away from the loads, was tried for this tile: 571 instructions
    // against 399, and with three or more warps on a scheduler -- the only banks this tile serves -- issue slots are what
    // counts: 16 384 channels 4.00e11 against 4.43e11.)
#pragma unroll
    for (int t = -P; t < TB; t++) {
        if (t + P + D < TB && t + P + D >= D + P) {
            x[t + P + D] = lds_ordered(col + (t + P + D) * CH);
            if (GAIN) x[t + P + D] = __fmul_rn(gain, x[t + P + D]);
        }
        if (t + 3 < TB) {
            const int j = t + 3;  // the high half's input: the low half's output kSkewLag positions back
            float ylo, yhi;
            if (j - kSkewLag >= 0) unpack2(Y[j - kSkewLag + 2], ylo, yhi);
            else ylo = p[kSkewLag - 1 - j];
            V[j + 2] = pack2(x[j], ylo);
            A[j] = ffma2(k[1], V[j + 1], V[j + 2]);
        }
        if (t + 2 >= 0 && t + 2 < TB) B[t + 2] = ffma2(k[2], V[t + 2], A[t + 2]);
        if (t + 1 >= 0 && t + 1 < TB) C[t + 1] = ffma2(k[4], Y[t + 1], B[t + 1]);
        if (t >= 0) {
            Y[t + 2] = ffma2(k[3], Y[t + 1], C[t]);
            float ylo, yhi;
            unpack2(Y[t + 2], ylo, yhi);
            sts_ordered(col + t * CH, yhi);
        }
    }
    x1 = V[TB + 1]; x2 = V[TB]; y1 = Y[TB + 1]; y2 = Y[TB];
#pragma unroll
    for (int q = 0; q < kSkewLag; q++) {
        float ylo, yhi;
        unpack2(Y[TB + 1 - q], ylo, yhi);
        p[q] = ylo;
    }
}

// FAST tier, rolled and modulo-scheduled (the default; kSkewLag == 4).  ptxas orders a fully unrolled tile for register
// pressure: it un-pipelines the hand-written schedule above back into one dependent A -> B -> C -> Y chain per step (four
// 4-5 cycle stalls a step, 19-20 cycles by its own stall counts whatever TB and D are: tools/sass_model.py) and sinks the
// prefetches down to their uses -- right for a scheduler with warps to switch to, wrong for a warp that has its scheduler
// to itself.  What ptxas cannot do is move work across a loop's back edge.  So the tile is a ROLLED loop of 8 positions per
// trip in which every instruction's operands were produced in an EARLIER position (the four fused operations of position t
// belong to four different frames, as above), the registers are named modulo the trip (V[8], two of everything else: no
// rotation copies), and the input of frame j is fetched 8 positions before its use into the slot frame j - 8 has just
// vacated.  Static count: 84 instructions and ~104 issue cycles per 8 positions (32 FFMA2) for either kind of warp.
// Same operations on the same operands as skew_tile<OLB_MODE_FAST>: bit-identical results.
//
// One position t (t == K mod 8):  Y_t, C_{t+1}, B_{t+2}, A_{t+3} (with V_{t+3} = (x_{t+3}, lo Y_{t-1})), store hi Y_t.
template <bool DOA, bool DOB, bool DOC, bool DOY, int K>
__device__ __forceinline__ void skew_pos(float *colt, const u64 (&k)[5], u64 (&V)[8], u64 (&Y)[2], u64 (&A)[2], u64 (&B)[2], u64 (&C)[2],
                                         float xin, float ylo_in) {
    constexpr int t = K + 8;  // keeps the modular arithmetic below positive
    if (DOY) Y[t & 1] = ffma2(k[3], Y[(t - 1) & 1], C[t & 1]);
    if (DOC) C[(t + 1) & 1] = ffma2(k[4], Y[(t - 1) & 1], B[(t + 1) & 1]);
    if (DOB) B[(t + 2) & 1] = ffma2(k[2], V[t & 7], A[(t + 2) & 1]);
    if (DOA) {
        V[(t + 3) & 7] = pack2(xin, ylo_in);
        A[(t + 3) & 1] = ffma2(k[1], V[(t + 2) & 7], V[(t + 3) & 7]);
    }
    if (DOY) {
        float ylo, yhi;
        unpack2(Y[t & 1], ylo, yhi);
        colt[0] = yhi;
    }
}

// Eight positions t .. t+7 (t == 0 mod 8).  FETCH = number of positions that still fetch an input 8 frames on (8 in the
// steady state, 5 in the last trip but one: frames beyond the tile are not read), DRAIN: the tile's last eight positions
// (the pipeline runs empty: no A after frame TB-1, and the low outputs of the last four frames go to p).
template <bool GAIN, int FETCH, bool DRAIN, int kSkewLag>
__device__ __forceinline__ void skew_block8(float *colt, const u64 (&k)[5], float gain, u64 (&V)[8], u64 (&Y)[2], u64 (&A)[2], u64 (&B)[2],
                                            u64 (&C)[2], float (&xr)[8], float (&p)[kSkewLag]) {
#define OLB_SKEW_POS(Q, DOA_, DOB_, DOC_)                                                                          \
    {                                                                                                              \
        float ylo, yhi;                                                                                            \
        unpack2(Y[((Q) + 1) & 1], ylo, yhi);                                                                       \
        if (DRAIN && (Q) >= 5) p[8 - (Q)] = ylo;                                                                   \
        float xin = xr[((Q) + 3) & 7];                                                                             \
        if (GAIN && (DOA_)) xin = __fmul_rn(gain, xin); /* warp 0 of a FAST bank applies G = prod(b0) to its input */ \
        skew_pos<DOA_, DOB_, DOC_, true, Q>(colt + (Q) * 32, k, V, Y, A, B, C, xin, ylo);                           \
        if ((Q) < FETCH) xr[((Q) + 3) & 7] = colt[((Q) + 11) * 32];                                                \
    }
    OLB_SKEW_POS(0, true, true, true)
    OLB_SKEW_POS(1, true, true, true)
    OLB_SKEW_POS(2, true, true, true)
    OLB_SKEW_POS(3, true, true, true)
    OLB_SKEW_POS(4, true, true, true)
    OLB_SKEW_POS(5, !DRAIN, true, true)
    OLB_SKEW_POS(6, !DRAIN, !DRAIN, true)
    OLB_SKEW_POS(7, !DRAIN, !DRAIN, !DRAIN)
#undef OLB_SKEW_POS
}

// nb: the rolled form runs 8 * nb >= 16 rows from `col` on (any run of rows in which both halves carry real frames: whole
// tiles in mid-stream, and what is left of the stream's first and last tiles); the unrolled form always runs TB rows.
template <int TB, int D, int kSkewLag, bool GAIN, bool ROLL, bool WHOLE>
__device__ __forceinline__ void skew_tile_fast(float *col, int nb, const u64 (&k)[5], float gain, u64 &x1, u64 &x2, u64 &y1, u64 &y2,
                                               float (&p)[kSkewLag], u64 one, u64 negzero) {
    if constexpr (!ROLL || kSkewLag != 4 || TB % 8 != 0 || TB < 24) {
        skew_tile_fast_unrolled<TB, D, kSkewLag, GAIN>(col, k, gain, x1, x2, y1, y2, p, one, negzero);
    } else {
        u64 V[8], Y[2], A[2], B[2], C[2];
        V[7] = x1; V[6] = x2; Y[1] = y1; Y[0] = y2;  // V_{-1}, V_{-2}, Y_{-1}, Y_{-2}
        A[0] = A[1] = B[0] = B[1] = C[0] = C[1] = 0ull;
        float xr[8];  // xr[j & 7] = input of frame j
#pragma unroll
        for (int q = 0; q < 8; q++) xr[q] = col[q * 32];
        // fill, positions -3, -2, -1:  A_0;  A_1 B_0;  A_2 B_1 C_0.  V_0 .. V_2 take the previous tile's low outputs from p.
        skew_pos<true, false, false, false, 5>(col, k, V, Y, A, B, C, GAIN ? __fmul_rn(gain, xr[0]) : xr[0], p[3]);
        skew_pos<true, true, false, false, 6>(col, k, V, Y, A, B, C, GAIN ? __fmul_rn(gain, xr[1]) : xr[1], p[2]);
        skew_pos<true, true, true, false, 7>(col, k, V, Y, A, B, C, GAIN ? __fmul_rn(gain, xr[2]) : xr[2], p[1]);
#pragma unroll
        for (int q = 0; q < 3; q++) xr[q] = col[(8 + q) * 32];  // frames 0 .. 2 are consumed: their slots take frames 8 .. 10
        float *colt = col;
        // steady-state blocks, two per trip (halves the copies that restore the loop's register names); WHOLE: a mid-stream
        // tile, the count known at compile time
        const int NS = (WHOLE ? TB / 8 : nb) - 2;
#pragma unroll 1
        for (int i = 0; i < NS / 2; i++, colt += 16 * 32) {
            skew_block8<GAIN, 8, false, kSkewLag>(colt, k, gain, V, Y, A, B, C, xr, p);
            skew_block8<GAIN, 8, false, kSkewLag>(colt + 8 * 32, k, gain, V, Y, A, B, C, xr, p);
        }
        if (NS & 1) {
            skew_block8<GAIN, 8, false, kSkewLag>(colt, k, gain, V, Y, A, B, C, xr, p);
            colt += 8 * 32;
        }
        skew_block8<GAIN, 5, false, kSkewLag>(colt, k, gain, V, Y, A, B, C, xr, p);
        skew_block8<GAIN, 0, true, kSkewLag>(colt + 8 * 32, k, gain, V, Y, A, B, C, xr, p);
        float yhi;
        unpack2(Y[1], p[0], yhi);  // lo Y_{TB-1}
        x1 = V[7]; x2 = V[6]; y1 = Y[1]; y2 = Y[0];
    }
}

// ROLL: the FAST tier's rolled tile (banks of at most two CTAs per SM, where a warp has its scheduler to itself or shares
// it with one other: 21.6 against 31.2 cycles per frame at 4 096 channels); with more warps per scheduler the unrolled
// tile's 7.3 instructions per step beat the rolled tile's 9.6 (16 384 channels: 72.7 against 79.1).
template <int S, int MODE, int TB, int NST, int D, int kSkewLag, bool ROLL>
__global__ void __launch_bounds__((S / 2) * 32) k_bq_skew(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                                          const __grid_constant__ TensorMap tm_out_head, uint32_t frames, uint32_t C, const float *__restrict__ coef, float *hist,
                                                          u64 one, u64 negzero) {
    constexpr int W = S / 2, CH = 32, DELAY = kSkewLag * W;  // stream positions between a frame's arrival and its departure
    static_assert(S % 2 == 0 && TB > DELAY, "bad shape");
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float *tiles = reinterpret_cast<float *>(smem_raw);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + (size_t) NST * TB * CH * sizeof(float));  // [NST][W]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t ch0 = blockIdx.x * CH;
    constexpr uint32_t kTileBytes = TB * CH * sizeof(float);
    const uint32_t total = frames + DELAY;  // stream length
    const uint32_t ntiles = (total + TB - 1) / TB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NST * W; i++) mbar_init(&bar[i], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == W - 1 && lane == 0) {
        for (uint32_t j = 0; j < (uint32_t) NST && j < ntiles; j++) {
            mbar_arrive_expect_tx(&bar[j * W], kTileBytes);
            tma_load_2d(tiles + (size_t) j * TB * CH, &tm_in, (int) ch0, (int) (j * TB), &bar[j * W]);
        }
    }

    // sections s0 (low half) and s0 + 1 (high half); history nodes s0, s0 + 1, s0 + 2
    const int s0 = warp * 2;
    const uint32_t c = ch0 + lane;
    const bool live = c < C;
    u64 k[5];
#pragma unroll
    for (int j = 0; j < 5; j++)
        k[j] = live ? pack2(coef[(size_t) (s0 * 5 + j) * C + c], coef[(size_t) ((s0 + 1) * 5 + j) * C + c]) : 0ull;
    const bool with_gain = MODE == OLB_MODE_FAST && warp == 0;
    const float gain = (with_gain && live) ? coef[c] : 1.0f;  // G = prod(b0) sits in section 0's b0 slot
    auto node = [&](int n, int i) { return live ? hist[(size_t) (n * 2 + i) * C + c] : 0.f; };
    u64 x1 = pack2(node(s0, 0), node(s0 + 1, 0)), x2 = pack2(node(s0, 1), node(s0 + 1, 1));
    u64 y1 = pack2(node(s0 + 1, 0), node(s0 + 2, 0)), y2 = pack2(node(s0 + 1, 1), node(s0 + 2, 1));
    float p[kSkewLag];  // low-half outputs of the last kSkewLag steps, on their way to the high half
#pragma unroll
    for (int q = 0; q < kSkewLag; q++) p[q] = 0.f;
    __syncthreads();  // every warp has read its nodes before any warp can finish and write one back

    const uint32_t lo0 = kSkewLag * warp, hi0 = lo0 + kSkewLag;  // first real position of each half

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % NST, par = (i / NST) & 1;
        mbar_wait(&bar[st * W + warp], par);
        float *col = tiles + (size_t) st * TB * CH + lane;
        const uint32_t m0 = i * TB;
        const uint32_t rows = min((uint32_t) TB, total - m0);
        constexpr bool kRolled = ROLL && MODE == OLB_MODE_FAST && kSkewLag == 4 && TB % 8 == 0 && TB >= 24;
        auto general = [&](uint32_t ra, uint32_t rb) {  // frame by frame, halves frozen outside their frames
            for (uint32_t r = ra; r < rb; r++) {
                const uint32_t m = m0 + r;
                const bool lo_on = m >= lo0 && m < lo0 + frames, hi_on = m >= hi0 && m < hi0 + frames;
                float xin = col[r * CH];
                if (with_gain) xin = __fmul_rn(gain, xin);
                const u64 v = pack2(xin, p[kSkewLag - 1]);
                const u64 y = skew_section<MODE>(v, x1, x2, y1, y2, k, one, negzero);
                const u64 nx2 = sel_halves(lo_on, hi_on, x1, x2), nx1 = sel_halves(lo_on, hi_on, v, x1);
                const u64 ny2 = sel_halves(lo_on, hi_on, y1, y2), ny1 = sel_halves(lo_on, hi_on, y, y1);
                x2 = nx2; x1 = nx1; y2 = ny2; y1 = ny1;
                float ylo, yhi;
                unpack2(y, ylo, yhi);
#pragma unroll
                for (int q = kSkewLag - 1; q > 0; q--) p[q] = p[q - 1];
                p[0] = ylo;
                col[r * CH] = yhi;
            }
        };
        if (m0 >= hi0 && m0 + TB <= lo0 + frames) {  // mid-stream: both halves carry real frames throughout the tile
            if (MODE == OLB_MODE_FAST) {
                if (with_gain) skew_tile_fast<TB, D, kSkewLag, true, ROLL, true>(col, TB / 8, k, gain, x1, x2, y1, y2, p, one, negzero);
                else skew_tile_fast<TB, D, kSkewLag, false, ROLL, true>(col, TB / 8, k, gain, x1, x2, y1, y2, p, one, negzero);
            } else {
                skew_tile<MODE, TB, D, kSkewLag, false>(col, k, 0ull, x1, x2, y1, y2, p, one, negzero);
            }
        } else if constexpr (!kRolled) {
            general(0, rows);  // (this branch as it was before the rolled tile learnt to run part of a tile: the unrolled
                               // tiles' code is sensitive to what surrounds it, 12 288 channels 55.2 -> 60.5 cycles per frame otherwise)
        } else {
            // the rows of this tile at which both halves carry real frames: positions [hi0, lo0 + frames); the rolled tile
            // takes them in blocks of eight, the frame-by-frame path the <= 7 + kSkewLag rows at either end
            const int lo = min(max((int) hi0 - (int) m0, 0), (int) rows), hi = min(max((int) (lo0 + frames) - (int) m0, 0), (int) rows);
            const int nb = hi > lo ? (hi - lo) / 8 : 0;
            if (nb < 2) {
                general(0, rows);
            } else {
                general(0, (uint32_t) lo);
                if (with_gain) skew_tile_fast<TB, D, kSkewLag, true, true, false>(col + lo * CH, nb, k, gain, x1, x2, y1, y2, p, one, negzero);
                else skew_tile_fast<TB, D, kSkewLag, false, true, false>(col + lo * CH, nb, k, gain, x1, x2, y1, y2, p, one, negzero);
                general((uint32_t) (lo + 8 * nb), rows);
            }
        }
        if (warp < W - 1) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[st * W + warp + 1]);
        } else {
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                // tile i leaves as frames [i*TB - DELAY, (i+1)*TB - DELAY); the first tile's leading DELAY rows are not frames
                // (its box is TB - DELAY rows high: tm_out_head), rows >= frames are clipped by the hardware
                if (i == 0) tma_store_2d(&tm_out_head, (int) ch0, 0, tiles + (size_t) DELAY * CH);
                else tma_store_2d(&tm_out, (int) ch0, (int) (i * TB) - DELAY, tiles + (size_t) st * TB * CH);
                tma_commit();
                // the stage tile i-1 was stored from is free once that store has read it: refill it with tile i-1+NST
                if (i >= 1 && i - 1 + NST < ntiles) {
                    tma_wait_read<1>();
                    const uint32_t nxt = i - 1 + NST, ns = nxt % NST;
                    mbar_arrive_expect_tx(&bar[ns * W], kTileBytes);
                    tma_load_2d(tiles + (size_t) ns * TB * CH, &tm_in, (int) ch0, (int) (nxt * TB), &bar[ns * W]);
                }
            }
        }
    }
    if (warp == W - 1 && lane == 0) tma_wait_all<0>();
    if (live) {
        // every half has now seen frames [0, F): x1.hi == y1.lo again and the table is back in node format.
        // Node s0 is the previous warp's output node (it writes it) except for warp 0, which owns the input node.
        float a, b;
        if (warp == 0) {
            unpack2(x1, a, b);
            hist[(size_t) (s0 * 2 + 0) * C + c] = a;
            unpack2(x2, a, b);
            hist[(size_t) (s0 * 2 + 1) * C + c] = a;
        }
        unpack2(y1, a, b);
        hist[(size_t) ((s0 + 1) * 2 + 0) * C + c] = a;
        hist[(size_t) ((s0 + 2) * 2 + 0) * C + c] = b;
        unpack2(y2, a, b);
        hist[(size_t) ((s0 + 1) * 2 + 1) * C + c] = a;
        hist[(size_t) ((s0 + 2) * 2 + 1) * C + c] = b;
    }
}

template <int S, int MODE, int TB, int NST, int D, int kSkewLag, bool ROLL>
static int launch_skew_cfg(uint32_t C, const float *d_coef, float *d_hist, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr int W = S / 2, CH = 32;
    constexpr size_t smem = (size_t) NST * TB * CH * sizeof(float) + (size_t) NST * W * sizeof(uint64_t);
    TensorMap tin, tout, thead;
    int rc = make_tmap_frames(&tin, in, frames, C, CH, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, C, CH, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&thead, out, frames, C, CH, TB - kSkewLag * W);
    if (rc) return rc;
    auto kern = k_bq_skew<S, MODE, TB, NST, D, kSkewLag, ROLL && MODE == OLB_MODE_FAST>;
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    const float onef = 1.0f, nzf = -0.0f;
    uint32_t ob, nb;
    memcpy(&ob, &onef, 4);
    memcpy(&nb, &nzf, 4);
    const u64 one = ((u64) ob << 32) | ob, negzero = ((u64) nb << 32) | nb;
    kern<<<(C + CH - 1) / CH, W * 32, smem, st>>>(tin, tout, thead, frames, C, d_coef, d_hist, one, negzero);
    OLB_LAUNCHED();
    return OLB_OK;
}

}  // namespace olb
