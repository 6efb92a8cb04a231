// biquad_roll.cuh -- k_bq_roll: the section-pipelined cascade with TWO CHANNELS per lane, SPW sections per warp and a rolled,
// modulo-scheduled tile (FAST tier).
//
// Why: banks of 4 737 .. ~19 000 channels -- 8 192 is one eighth of the 65 536-channel bank, the strong-scaling shard --
// give k_bq_skew's units (one warp = 32 channels x 2 sections, the two sections in the two halves of a packed stream)
// two to four warps per SM sub-partition, and a step of such a warp is 4 FFMA2 + ~6 other instructions, of which 2 are
// the shared-memory hop and 2-3 the copies that move a low-half output into the high half's input pair (ncu,
// profiles/r02_run17_ncu_k_bq_skew_8192.txt: FFMA2 40 % of the instructions, MOV 20 %).  With 64 channels to a warp the
// two halves of a packed register are two CHANNELS of the same section:
//   * the input of a step is one LDS.64 straight into the register pair the FFMA2s read, the output one STS.64;
//   * a section's output pair IS the next section's input pair: the SPW consecutive sections of one warp are chained in
//     registers, with no copy, 4 positions apart (so that the hand-over is off the critical path);
//   * a warp-frame is 4 x SPW FFMA2 + 2 memory instructions + the loop's share (ncu, profiles/r02_run44_ncu_k_bq_roll_8192.txt:
//     14.4 instructions per position with SPW = 2, FFMA2 55 %, where two k_bq_skew warps issue 20).
// 8 192 channels are 128 CTAs of 4 warps, one warp per sub-partition on 128 SMs: 37.6 cycles per frame against k_bq_skew's
// 40.2; 16 384 channels with SPW = 4 (256 CTAs of 2 warps): 63.6 against 73.5 (profiles/r02_run44_roll_sweep.txt).  What a
// lone warp cannot hide is its own issue latency: 2.5 cycles per instruction (1.0 issue + 0.9 fixed-latency wait + 0.36
// register-port dispatch stall + 0.18 shared-memory scoreboard), so the 8 FFMA2 of a position cost 21 cycles and the
// copies ptxas adds at the loop's back edge another 4.  (Below 4 737 channels 64-channel CTAs leave most SMs without work,
// and between 9 473 and 14 208 three 32-channel CTAs per SM are the finer fit: k_bq_skew keeps those banks.)
//
// Tried beside it and dropped: FOUR sections per warp as two section-pair streams of 32 channels (`k_bq_quad`: the
// hand-overs between the streams are two more copies a position; 44.4 cycles per frame at 8 192 channels).
//
// The tile is k_bq_skew's rolled form: a loop of 8 stream positions per trip in which every operand was produced in an
// earlier position, registers named modulo the trip.  Z[s] is the stream between section s-1 and section s of the warp
// (Z[0]: the warp's input, Z[SPW]: its output), 8 slots named by frame & 7; at position t section s (frame j = t - 4s):
//     Y_j = Z[s+1][j] = fma(-a1, Z[s+1][j-1], C_j)      C_{j+1} = fma(-a2, Z[s+1][j-1], B_{j+1})
//     B_{j+2} = fma(r2, Z[s][j], A_{j+2})                A_{j+3} = fma(r1, Z[s][j+2], Z[s][j+3])
// -- the four fused operations of the FAST tier (biquad_sys.cuh Lane<2>::section) for four different frames.  Row t + 8 of
// the tile is fetched into slot t & 7 of Z[0] right after its last reader (B_{t+2}, at position t); warp 0 scales it by
// G = prod(b0) in place four positions later.  Per-section operation order is untouched: bit-identical to every other
// kernel of the bank and to the oracle.
//
// Stream positions, frozen sections at the two ends of the stream, the tile ring and the TMA store coordinates are
// k_bq_skew's (biquad_skew.cuh), with a warp delaying its stream by 4 x (SPW - 1) positions.
#pragma once
#include "biquad_skew.cuh"

namespace olb {

__device__ __forceinline__ u64 lds64(const float *p) {
    u64 v;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(float *p, u64 v) {
    asm volatile("st.shared.b64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}

// SL: outputs leave for the tile SL positions after they were computed (0 or 4).  A store issued right behind the FFMA2
// that produced its operand holds that register pair until the store unit has read it, and the next writer of the pair -- a
// few instructions on, for a warp that has its scheduler to itself -- waits for that (ptxas copies the pair first, or sets a
// read barrier): 39.4 -> 37.6 cycles per frame at 8 192 channels.

// One position t == K (mod 8) of every section of the warp.  row = the tile row of position t (this lane's 8 bytes).
template <int SPW, int SL, bool GAIN, bool DOA, bool DOB, bool DOC, bool DOY, bool FETCH, bool STORE, int K>
__device__ __forceinline__ void roll_pos(float *row, const u64 (&k)[SPW][5], u64 gain2, u64 (&Z)[SPW + 1][8], u64 (&A)[SPW][2], u64 (&B)[SPW][2],
                                         u64 (&C)[SPW][2]) {
    constexpr int RS = 64;  // floats per tile row
    if (GAIN && DOA) Z[0][(K + 3) & 7] = fmul2(gain2, Z[0][(K + 3) & 7]);
#pragma unroll
    for (int s = 0; s < SPW; s++) {
        const int j = K + 32 - 4 * s;  // this section's frame at this position, mod 8
        if (DOY) Z[s + 1][j & 7] = ffma2(k[s][3], Z[s + 1][(j - 1) & 7], C[s][j & 1]);
        if (DOC) C[s][(j + 1) & 1] = ffma2(k[s][4], Z[s + 1][(j - 1) & 7], B[s][(j + 1) & 1]);
        if (DOB) B[s][(j + 2) & 1] = ffma2(k[s][2], Z[s][j & 7], A[s][(j + 2) & 1]);
        if (DOA) A[s][(j + 3) & 1] = ffma2(k[s][1], Z[s][(j + 2) & 7], Z[s][(j + 3) & 7]);
    }
    if (STORE) sts64(row - SL * RS, Z[SPW][(K + 32 - 4 * (SPW - 1) - SL) & 7]);
    if (FETCH) Z[0][K & 7] = lds64(row + 8 * RS);
}

// Eight positions t .. t + 7 (t == 0 mod 8).  NF = number of positions that still fetch the row 8 on (8 in the steady
// state, 0 in a tile's last block); DRAIN: the tile's last eight positions (no A after frame TB - 1).
// FIRST: the tile's first eight positions (the rows kRollStoreLag back belong to the tile before: nothing to store yet).
template <int SPW, int SL, bool GAIN, int NF, bool DRAIN, bool FIRST>
__device__ __forceinline__ void roll_block8(float *row, const u64 (&k)[SPW][5], u64 gain2, u64 (&Z)[SPW + 1][8], u64 (&A)[SPW][2], u64 (&B)[SPW][2],
                                            u64 (&C)[SPW][2]) {
    constexpr int RS = 64, L = SL;
    roll_pos<SPW, SL, GAIN, true, true, true, true, (0 < NF), !FIRST || (0 >= L), 0>(row + 0 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, true, true, (1 < NF), !FIRST || (1 >= L), 1>(row + 1 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, true, true, (2 < NF), !FIRST || (2 >= L), 2>(row + 2 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, true, true, (3 < NF), !FIRST || (3 >= L), 3>(row + 3 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, true, true, (4 < NF), !FIRST || (4 >= L), 4>(row + 4 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, !DRAIN, true, true, true, (5 < NF), !FIRST || (5 >= L), 5>(row + 5 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, !DRAIN, !DRAIN, true, true, (6 < NF), !FIRST || (6 >= L), 6>(row + 6 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, !DRAIN, !DRAIN, !DRAIN, true, (7 < NF), !FIRST || (7 >= L), 7>(row + 7 * RS, k, gain2, Z, A, B, C);
}

// Rotating state of a warp between tiles and on the frame-by-frame path: per section its two last inputs and outputs, and
// the outputs of the last four positions on their way to the next section (k_bq_skew's p).
template <int SPW>
struct RollState {
    u64 x1[SPW], x2[SPW], y1[SPW], y2[SPW];
    u64 p[SPW][4];  // p[s][q] = section s's output of q + 1 positions ago (the last section's go to the tile instead)
};

// 8 * nb >= 16 consecutive rows of a tile (col = the first) in which every section of the warp carries real frames.
// NB > 0: a whole tile, block count known at compile time (the mid-stream tiles: the hot loop); NB == 0: nb blocks (what is
// left of the stream's first and last tiles).
template <int SPW, int SL, int UNR, bool GAIN, int NB>
__device__ __forceinline__ void roll_tile(float *col, int nb, const u64 (&k)[SPW][5], u64 gain2, RollState<SPW> &S) {
    constexpr int RS = 64;
    u64 Z[SPW + 1][8], A[SPW][2], B[SPW][2], C[SPW][2];
    // tile-local position -1 was the last one: section s has finished frame -1 - 4s
#pragma unroll
    for (int s = 0; s < SPW; s++) {
        const int j = 32 - 1 - 4 * s;
        Z[s][j & 7] = S.x1[s];
        Z[s][(j - 1) & 7] = S.x2[s];
        Z[s + 1][j & 7] = S.y1[s];
        Z[s + 1][(j - 1) & 7] = S.y2[s];
        if (s + 1 < SPW) {
            Z[s + 1][(j - 2) & 7] = S.p[s][2];
            Z[s + 1][(j - 3) & 7] = S.p[s][3];
        }
        A[s][0] = A[s][1] = B[s][0] = B[s][1] = C[s][0] = C[s][1] = 0ull;
    }
#pragma unroll
    for (int q = 0; q < 6; q++) Z[0][q] = lds64(col + q * RS);  // slots 6 and 7 still hold frames -2 and -1
    // fill, positions -3, -2, -1:  A;  A B;  A B C  (frames -3 - 4s + 3 ... of every section)
    roll_pos<SPW, SL, GAIN, true, false, false, false, false, false, 5>(col - 3 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, false, false, true, false, 6>(col - 2 * RS, k, gain2, Z, A, B, C);
    roll_pos<SPW, SL, GAIN, true, true, true, false, true, false, 7>(col - 1 * RS, k, gain2, Z, A, B, C);
    roll_block8<SPW, SL, GAIN, 8, false, true>(col, k, gain2, Z, A, B, C);
    float *row = col + 8 * RS;
    const int NS = (NB > 0 ? NB : nb) - 2;  // steady-state blocks
    if (UNR == 2) {  // two blocks per trip: half the copies that restore the loop's register names
#pragma unroll 1
        for (int i = 0; i < NS / 2; i++, row += 16 * RS) {
            roll_block8<SPW, SL, GAIN, 8, false, false>(row, k, gain2, Z, A, B, C);
            roll_block8<SPW, SL, GAIN, 8, false, false>(row + 8 * RS, k, gain2, Z, A, B, C);
        }
        if (NS & 1) {
            roll_block8<SPW, SL, GAIN, 8, false, false>(row, k, gain2, Z, A, B, C);
            row += 8 * RS;
        }
    } else {
#pragma unroll 1
        for (int i = 0; i < NS; i++, row += 8 * RS) roll_block8<SPW, SL, GAIN, 8, false, false>(row, k, gain2, Z, A, B, C);
    }
    roll_block8<SPW, SL, GAIN, 0, true, false>(row, k, gain2, Z, A, B, C);
#pragma unroll
    for (int q = SL; q >= 1; q--)  // the outputs of the last positions, still on their way
        sts64(row + (8 - q) * RS, Z[SPW][(32 - q - 4 * (SPW - 1)) & 7]);
#pragma unroll
    for (int s = 0; s < SPW; s++) {
        const int j = 32 - 1 - 4 * s;  // the last position carries the same names as position -1
        S.x1[s] = Z[s][j & 7];
        S.x2[s] = Z[s][(j - 1) & 7];
        S.y1[s] = Z[s + 1][j & 7];
        S.y2[s] = Z[s + 1][(j - 1) & 7];
        if (s + 1 < SPW) {
#pragma unroll
            for (int q = 0; q < 4; q++) S.p[s][q] = Z[s + 1][(j - q) & 7];
        }
    }
}

template <int S, int SPW, int TB, int NST, int SL, int UNR>
__global__ void __launch_bounds__((S / SPW) * 32) k_bq_roll(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                                            const __grid_constant__ TensorMap tm_out_head, uint32_t frames, uint32_t C, const float *__restrict__ coef,
                                                            float *hist, u64 one, u64 negzero) {
    constexpr int W = S / SPW, CH = 64, LAG = 4, WLAG = LAG * (SPW - 1), DELAY = WLAG * W;
    static_assert(S % SPW == 0 && SPW >= 2 && TB > DELAY && TB % 8 == 0, "bad shape");
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

    // sections s0 .. s0 + SPW - 1 of channels c, c + 1; history nodes s0 .. s0 + SPW
    const int s0 = warp * SPW;
    const uint32_t c = ch0 + 2 * lane;
    const bool live = c < C;  // C % 4 == 0: c + 1 < C as well
    auto pair_at = [&](const float *t, size_t rowi) { return live ? *reinterpret_cast<const u64 *>(t + rowi * C + c) : 0ull; };
    u64 k[SPW][5];
#pragma unroll
    for (int s = 0; s < SPW; s++)
#pragma unroll
        for (int j = 0; j < 5; j++) k[s][j] = pair_at(coef, (size_t) ((s0 + s) * 5 + j));
    const bool with_gain = warp == 0;
    const u64 gain2 = (with_gain && live) ? pair_at(coef, 0) : one;  // G = prod(b0) sits in section 0's b0 slot
    RollState<SPW> R;
#pragma unroll
    for (int s = 0; s < SPW; s++) {
        R.x1[s] = pair_at(hist, (size_t) ((s0 + s) * 2 + 0));
        R.x2[s] = pair_at(hist, (size_t) ((s0 + s) * 2 + 1));
        R.y1[s] = pair_at(hist, (size_t) ((s0 + s + 1) * 2 + 0));
        R.y2[s] = pair_at(hist, (size_t) ((s0 + s + 1) * 2 + 1));
#pragma unroll
        for (int q = 0; q < 4; q++) R.p[s][q] = 0ull;
    }
    __syncthreads();  // every warp has read its nodes before any warp can finish and write one back

    const uint32_t base = WLAG * warp;  // first real position of the warp's first section; section s starts at base + LAG * s

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % NST, par = (i / NST) & 1;
        mbar_wait(&bar[st * W + warp], par);
        float *col = tiles + (size_t) st * TB * CH + 2 * lane;
        const uint32_t m0 = i * TB;
        const uint32_t rows = min((uint32_t) TB, total - m0);
        // frame by frame, sections frozen outside their frames: the two ends of the stream
        auto general = [&](uint32_t ra, uint32_t rb) {
            for (uint32_t r = ra; r < rb; r++) {
                const uint32_t m = m0 + r;
                u64 xin = lds64(col + r * CH);
                if (with_gain) xin = fmul2(gain2, xin);
                u64 y[SPW];
#pragma unroll
                for (int s = 0; s < SPW; s++) {
                    const bool on = m >= base + LAG * s && m < base + LAG * s + frames;
                    const u64 v = s == 0 ? xin : R.p[s - 1][3];
                    y[s] = skew_section<OLB_MODE_FAST>(v, R.x1[s], R.x2[s], R.y1[s], R.y2[s], k[s], one, negzero);
                    if (on) {
                        R.x2[s] = R.x1[s]; R.x1[s] = v; R.y2[s] = R.y1[s]; R.y1[s] = y[s];
                    }
                }
#pragma unroll
                for (int s = 0; s + 1 < SPW; s++) {
#pragma unroll
                    for (int q = 3; q > 0; q--) R.p[s][q] = R.p[s][q - 1];
                    R.p[s][0] = y[s];
                }
                sts64(col + r * CH, y[SPW - 1]);
            }
        };
        if (m0 >= base + WLAG && m0 + TB <= base + frames) {  // mid-stream: every section carries real frames throughout the tile
            if (with_gain) roll_tile<SPW, SL, UNR, true, TB / 8>(col, TB / 8, k, gain2, R);
            else roll_tile<SPW, SL, UNR, false, TB / 8>(col, TB / 8, k, gain2, R);
        } else {
            // the rows of this tile at which every section of the warp carries real frames: positions [base + WLAG, base + frames);
            // the rolled tile takes them in blocks of eight, the frame-by-frame path the <= 7 + WLAG rows at either end
            const int lo = min(max((int) (base + WLAG) - (int) m0, 0), (int) rows), hi = min(max((int) (base + frames) - (int) m0, 0), (int) rows);
            const int nb = hi > lo ? (hi - lo) / 8 : 0;
            if (nb < 2) {
                general(0, rows);
            } else {
                general(0, (uint32_t) lo);
                if (with_gain) roll_tile<SPW, SL, 1, true, 0>(col + lo * CH, nb, k, gain2, R);
                else roll_tile<SPW, SL, 1, false, 0>(col + lo * CH, nb, k, gain2, R);
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
                if (i == 0) tma_store_2d(&tm_out_head, (int) ch0, 0, tiles + (size_t) DELAY * CH);
                else tma_store_2d(&tm_out, (int) ch0, (int) (i * TB) - DELAY, tiles + (size_t) st * TB * CH);
                tma_commit();
                if (i >= 1 && i - 1 + NST < ntiles) {  // the stage tile i-1 was stored from is free once that store has read it
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
        // every section has seen frames [0, F): the table is back in node format.  Node s0 is the previous warp's output
        // node (it writes it) except for warp 0, which owns the input node.
        auto put = [&](size_t rowi, u64 v) { *reinterpret_cast<u64 *>(hist + rowi * C + c) = v; };
        if (warp == 0) {
            put((size_t) (s0 * 2 + 0), R.x1[0]);
            put((size_t) (s0 * 2 + 1), R.x2[0]);
        }
#pragma unroll
        for (int s = 0; s < SPW; s++) {
            put((size_t) ((s0 + s + 1) * 2 + 0), R.y1[s]);
            put((size_t) ((s0 + s + 1) * 2 + 1), R.y2[s]);
        }
    }
}

template <int S, int SPW, int TB, int NST, int SL, int UNR>
static int launch_roll_cfg(uint32_t C, const float *d_coef, float *d_hist, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr int W = S / SPW, CH = 64, DELAY = 4 * (SPW - 1) * W;
    constexpr size_t smem = (size_t) NST * TB * CH * sizeof(float) + (size_t) NST * W * sizeof(uint64_t);
    static_assert(smem <= 227 * 1024, "tile ring exceeds the shared memory of an SM");
    TensorMap tin, tout, thead;
    int rc = make_tmap_frames(&tin, in, frames, C, CH, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, C, CH, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&thead, out, frames, C, CH, TB - DELAY);
    if (rc) return rc;
    auto kern = k_bq_roll<S, SPW, TB, NST, SL, UNR>;
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
