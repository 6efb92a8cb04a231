// biquad_bank.cu -- (a1) biquad / one-pole cascade bank.
//
// The reference has no biquad class; the bank follows the reference's object
// conventions (Init / setters / Process) and reproduces its two one-pole forms bit-exactly:
//   daisysp::Port::Process      modules/synthlib/Portamento.h:29-32     z = x + c*(z - x)
//   LowPassFilter_process       libs/dattorro-verb/verb.cpp:131-134     o = o + (x - o)*f
//
// Data layout in HBM
//   audio    [frames][channels] fp32 (OLB_LAYOUT_FRAMES; channel fastest) or [channels][frames]
//   coef     [sections][5][channels]   b0 b1 b2 -a1 -a2   (negation is exact; channel fastest so a warp's
//                                                          prologue loads are coalesced)
//   hist     [sections+1][2][channels] the last two samples at every node of the cascade: node 0 is the
//            input, node s+1 the output of section s.  In a direct-form-1 cascade the y-history of
//            section s IS the x-history of section s+1, so 2(S+1) floats per channel replace 4S.
//
// Kernels (all: one time-sequential recurrence per channel, state and coefficients in registers)
//   k_bq_tma     the throughput kernel.  One thread = two adjacent channels in packed fp32x2 (FFMA2);
//                one warp = 64 channels with its OWN TMA pipeline: [TB frames][64 ch] tiles arrive by
//                cp.async.bulk.tensor into a per-warp shared-memory ring (mbarrier complete_tx), are
//                filtered in place and leave by TMA store.  No block-wide barrier exists in the kernel.
//   k_bq_direct  one channel per thread, plain coalesced loads, register double-buffering; handles any
//                channel count / alignment, one-pole kinds in EXACT mode, and runtime section counts.
//   k_bq_planar  [channels][frames] input: a warp transposes 32x32 tiles through shared memory.
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "tmap.h"
#include "biquad_sys.cuh"
#include "biquad_roll.cuh"

namespace olb {

constexpr int kMaxS = OLB_BIQUAD_MAX_SECTIONS;

struct Kinds {
    uint8_t k[kMaxS];
};

// ------------------------------------------------------------------------------------------------
// One cascade step, scalar.  Op order follows the oracle (oracle/olb_oracle.cpp:131-147).
template <int MODE>
__device__ __forceinline__ float section_biquad(float x, float x1, float x2, float y1, float y2, float b0, float b1, float b2,
                                                float na1, float na2) {
    if (MODE == OLB_MODE_FAST) {  // b1, b2 hold r1 = b1/b0, r2 = b2/b0; x arrives pre-scaled (b0 slot unused)
        float t = __fmaf_rn(b1, x1, x);
        t = __fmaf_rn(b2, x2, t);
        t = __fmaf_rn(na2, y2, t);
        return __fmaf_rn(na1, y1, t);
    } else if (MODE == OLB_MODE_FMA) {
        float p = __fmul_rn(b1, x1);
        p = __fmaf_rn(b2, x2, p);
        p = __fmaf_rn(na2, y2, p);
        p = __fmaf_rn(na1, y1, p);
        return __fmaf_rn(b0, x, p);
    } else {
        float p = __fadd_rn(__fmul_rn(b1, x1), __fmul_rn(b2, x2));
        p = __fadd_rn(p, __fmul_rn(na2, y2));
        p = __fadd_rn(p, __fmul_rn(na1, y1));
        return __fadd_rn(__fmul_rn(b0, x), p);
    }
}

// h[n][0] = newest, h[n][1] = the one before, for node n.  GENERIC: honour one-pole kinds (EXACT only).
template <int S, int MODE, bool GENERIC>
__device__ __forceinline__ float cascade_step(float v, float (&h)[S + 1][2], const float (&k)[S][5], const Kinds &kinds) {
    if (MODE == OLB_MODE_FAST) v = __fmul_rn(k[0][0], v);  // G = prod(b0) sits in the b0 slot of section 0
#pragma unroll
    for (int s = 0; s < S; s++) {
        float y;
        const int kind = GENERIC ? kinds.k[s] : OLB_KIND_BIQUAD;
        if (GENERIC && kind == OLB_KIND_PORT) {
            y = __fadd_rn(v, __fmul_rn(k[s][0], __fsub_rn(h[s + 1][0], v)));
        } else if (GENERIC && kind == OLB_KIND_LPF) {
            y = __fadd_rn(h[s + 1][0], __fmul_rn(__fsub_rn(v, h[s + 1][0]), k[s][0]));
        } else {
            y = section_biquad<MODE>(v, h[s][0], h[s][1], h[s + 1][0], h[s + 1][1], k[s][0], k[s][1], k[s][2], k[s][3], k[s][4]);
        }
        h[s][1] = h[s][0];
        h[s][0] = v;
        v = y;
    }
    h[S][1] = h[S][0];
    h[S][0] = v;
    return v;
}

// ------------------------------------------------------------------------------------------------
// k_bq_direct: one channel per thread, U frames of register double-buffering.
template <int S, int MODE, bool GENERIC, int U>
__global__ void __launch_bounds__(128) k_bq_direct(const float *in, float *out, uint64_t frames, uint32_t C, const float *coef,
                                                   float *hist, Kinds kinds) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float k[S][5], h[S + 1][2];
#pragma unroll
    for (int s = 0; s < S; s++)
#pragma unroll
        for (int j = 0; j < 5; j++) k[s][j] = coef[(size_t) (s * 5 + j) * C + c];
#pragma unroll
    for (int n = 0; n <= S; n++) {
        h[n][0] = hist[(size_t) (n * 2 + 0) * C + c];
        h[n][1] = hist[(size_t) (n * 2 + 1) * C + c];
    }
    const float *pi = in + c;
    float *po = out + c;
    uint64_t t = 0;
    if (frames >= (uint64_t) U) {
        float xb[U], xn[U];
#pragma unroll
        for (int i = 0; i < U; i++) xb[i] = pi[(size_t) i * C];
        pi += (size_t) U * C;
        for (; t + 2 * U <= frames; t += U) {
#pragma unroll
            for (int i = 0; i < U; i++) xn[i] = pi[(size_t) i * C];
            pi += (size_t) U * C;
#pragma unroll
            for (int i = 0; i < U; i++) po[(size_t) i * C] = cascade_step<S, MODE, GENERIC>(xb[i], h, k, kinds);
            po += (size_t) U * C;
#pragma unroll
            for (int i = 0; i < U; i++) xb[i] = xn[i];
        }
#pragma unroll
        for (int i = 0; i < U; i++) po[(size_t) i * C] = cascade_step<S, MODE, GENERIC>(xb[i], h, k, kinds);
        po += (size_t) U * C;
        t += U;
    }
    for (; t < frames; t++) {
        *po = cascade_step<S, MODE, GENERIC>(*pi, h, k, kinds);
        pi += C;
        po += C;
    }
#pragma unroll
    for (int n = 0; n <= S; n++) {
        hist[(size_t) (n * 2 + 0) * C + c] = h[n][0];
        hist[(size_t) (n * 2 + 1) * C + c] = h[n][1];
    }
}

// Runtime section count (9..16 sections, or as a last resort): state lives in local arrays.
template <int MODE>
__global__ void __launch_bounds__(128) k_bq_runtime(const float *in, float *out, uint64_t frames, uint32_t C, int S,
                                                    const float *coef, float *hist, Kinds kinds, int layout) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float k[kMaxS][5], h[kMaxS + 1][2];
    for (int s = 0; s < S; s++)
        for (int j = 0; j < 5; j++) k[s][j] = coef[(size_t) (s * 5 + j) * C + c];
    for (int n = 0; n <= S; n++) {
        h[n][0] = hist[(size_t) (n * 2 + 0) * C + c];
        h[n][1] = hist[(size_t) (n * 2 + 1) * C + c];
    }
    const size_t base = layout == OLB_LAYOUT_FRAMES ? c : (size_t) c * frames;
    const size_t stride = layout == OLB_LAYOUT_FRAMES ? C : 1;
    for (uint64_t t = 0; t < frames; t++) {
        float v = in[base + t * stride];
        if (MODE == OLB_MODE_FAST) v = __fmul_rn(k[0][0], v);
        for (int s = 0; s < S; s++) {
            float y;
            const int kind = MODE != OLB_MODE_EXACT ? OLB_KIND_BIQUAD : kinds.k[s];
            if (kind == OLB_KIND_PORT) y = __fadd_rn(v, __fmul_rn(k[s][0], __fsub_rn(h[s + 1][0], v)));
            else if (kind == OLB_KIND_LPF) y = __fadd_rn(h[s + 1][0], __fmul_rn(__fsub_rn(v, h[s + 1][0]), k[s][0]));
            else y = section_biquad<MODE>(v, h[s][0], h[s][1], h[s + 1][0], h[s + 1][1], k[s][0], k[s][1], k[s][2], k[s][3], k[s][4]);
            h[s][1] = h[s][0];
            h[s][0] = v;
            v = y;
        }
        h[S][1] = h[S][0];
        h[S][0] = v;
        out[base + t * stride] = v;
    }
    for (int n = 0; n <= S; n++) {
        hist[(size_t) (n * 2 + 0) * C + c] = h[n][0];
        hist[(size_t) (n * 2 + 1) * C + c] = h[n][1];
    }
}

// ------------------------------------------------------------------------------------------------
// k_bq_planar: [channels][frames].  One warp = 32 channels; 32x32 tiles are read with lane = frame
// (coalesced 128 B per channel row), transposed through padded shared memory, filtered with
// lane = channel, and written back the same way.
template <int S, int MODE, bool GENERIC>
__global__ void __launch_bounds__(128) k_bq_planar(const float *in, float *out, uint64_t frames, uint32_t C, const float *coef,
                                                   float *hist, Kinds kinds) {
    __shared__ float tile_all[4][32][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float(*tile)[33] = tile_all[warp];
    const uint32_t c0 = (blockIdx.x * 4 + warp) * 32;
    if (c0 >= C) return;
    const uint32_t c = c0 + lane;
    const bool live = c < C;
    const uint32_t cc = live ? c : C - 1;  // dead lanes shadow a valid channel and never store
    float k[S][5], h[S + 1][2];
#pragma unroll
    for (int s = 0; s < S; s++)
#pragma unroll
        for (int j = 0; j < 5; j++) k[s][j] = coef[(size_t) (s * 5 + j) * C + cc];
#pragma unroll
    for (int n = 0; n <= S; n++) {
        h[n][0] = hist[(size_t) (n * 2 + 0) * C + cc];
        h[n][1] = hist[(size_t) (n * 2 + 1) * C + cc];
    }
    const uint32_t nch = min(32u, C - c0);
    for (uint64_t t0 = 0; t0 < frames; t0 += 32) {
        const uint32_t nt = (uint32_t) min((uint64_t) 32, frames - t0);
        for (uint32_t r = 0; r < nch; r++)
            if ((uint32_t) lane < nt) tile[r][lane] = in[(size_t) (c0 + r) * frames + t0 + lane];
        __syncwarp();
        for (uint32_t i = 0; i < nt; i++) {
            const float y = cascade_step<S, MODE, GENERIC>(tile[lane][i], h, k, kinds);
            tile[lane][i] = y;
        }
        __syncwarp();
        for (uint32_t r = 0; r < nch; r++)
            if ((uint32_t) lane < nt) out[(size_t) (c0 + r) * frames + t0 + lane] = tile[r][lane];
        __syncwarp();
    }
    if (live) {
#pragma unroll
        for (int n = 0; n <= S; n++) {
            hist[(size_t) (n * 2 + 0) * C + c] = h[n][0];
            hist[(size_t) (n * 2 + 1) * C + c] = h[n][1];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_bq_tma: packed two-channel threads, per-warp TMA pipelines.
template <int MODE>
__device__ __forceinline__ u64 section_biquad2(u64 x, u64 x1, u64 x2, u64 y1, u64 y2, u64 b0, u64 b1, u64 b2, u64 na1, u64 na2,
                                               u64 one, u64 negzero) {
    if (MODE == OLB_MODE_FAST) {
        u64 t = ffma2(b1, x1, x);
        t = ffma2(b2, x2, t);
        t = ffma2(na2, y2, t);
        return ffma2(na1, y1, t);
    } else if (MODE == OLB_MODE_FMA) {
        u64 p = fmul2(b1, x1);
        p = ffma2(b2, x2, p);
        p = ffma2(na2, y2, p);
        p = ffma2(na1, y1, p);
        return ffma2(b0, x, p);
    } else {
        // separately rounded products and sums from FFMA2 with opaque constants (see common.cuh)
        u64 p = ffma2(ffma2(b1, x1, negzero), one, ffma2(b2, x2, negzero));
        p = ffma2(p, one, ffma2(na2, y2, negzero));
        p = ffma2(p, one, ffma2(na1, y1, negzero));
        return ffma2(ffma2(b0, x, negzero), one, p);
    }
}

template <int S, int MODE>
__device__ __forceinline__ u64 cascade_step2(u64 v, u64 (&h)[S + 1][2], const u64 (&k)[S][5], u64 one, u64 negzero) {
    if (MODE == OLB_MODE_FAST) v = fmul2(k[0][0], v);
#pragma unroll
    for (int s = 0; s < S; s++) {
        const u64 y = section_biquad2<MODE>(v, h[s][0], h[s][1], h[s + 1][0], h[s + 1][1], k[s][0], k[s][1], k[s][2], k[s][3],
                                            k[s][4], one, negzero);
        h[s][1] = h[s][0];
        h[s][0] = v;
        v = y;
    }
    h[S][1] = h[S][0];
    h[S][0] = v;
    return v;
}

constexpr int kTmaCh = 64;  // channels per warp (2 per lane)

template <int S, int MODE, int TB, int STAGES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_bq_tma(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                                       uint32_t frames, uint32_t C, const float *__restrict__ coef, float *hist,
                                                       u64 one, u64 negzero) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t group = blockIdx.x * WARPS + warp;
    const uint32_t ch0 = group * kTmaCh;
    if (ch0 >= C) return;
    float *tiles = reinterpret_cast<float *>(smem_raw) + (size_t) warp * STAGES * TB * kTmaCh;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + (size_t) WARPS * STAGES * TB * kTmaCh * sizeof(float)) + warp * STAGES;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncwarp();

    constexpr uint32_t kTileBytes = TB * kTmaCh * sizeof(float);
    const uint32_t ntiles = (frames + TB - 1) / TB;
    if (lane == 0) {
        for (uint32_t j = 0; j < (uint32_t) (STAGES - 1) && j < ntiles; j++) {
            mbar_arrive_expect_tx(&full[j], kTileBytes);
            tma_load_2d(tiles + (size_t) j * TB * kTmaCh, &tm_in, (int) ch0, (int) (j * TB), &full[j]);
        }
    }

    // coefficients and history for channels (c, c+1); lanes past the end of the bank idle on zeros
    const uint32_t c = ch0 + 2 * lane;
    const bool live = c < C;  // C is a multiple of 4 on this path, so a live lane owns two live channels
    u64 k[S][5], h[S + 1][2];
#pragma unroll
    for (int s = 0; s < S; s++)
#pragma unroll
        for (int j = 0; j < 5; j++) k[s][j] = live ? *reinterpret_cast<const u64 *>(coef + (size_t) (s * 5 + j) * C + c) : 0ull;
#pragma unroll
    for (int n = 0; n <= S; n++) {
        h[n][0] = live ? *reinterpret_cast<const u64 *>(hist + (size_t) (n * 2 + 0) * C + c) : 0ull;
        h[n][1] = live ? *reinterpret_cast<const u64 *>(hist + (size_t) (n * 2 + 1) * C + c) : 0ull;
    }

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % STAGES, par = (i / STAGES) & 1;
        mbar_wait(&full[st], par);
        u64 *col = reinterpret_cast<u64 *>(tiles + (size_t) st * TB * kTmaCh) + lane;
        const uint32_t rows = min((uint32_t) TB, frames - i * TB);
        if (rows == TB) {
#pragma unroll 4
            for (int r = 0; r < TB; r++) col[r * (kTmaCh / 2)] = cascade_step2<S, MODE>(col[r * (kTmaCh / 2)], h, k, one, negzero);
        } else {
            for (uint32_t r = 0; r < rows; r++) col[r * (kTmaCh / 2)] = cascade_step2<S, MODE>(col[r * (kTmaCh / 2)], h, k, one, negzero);
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_2d(&tm_out, (int) ch0, (int) (i * TB), tiles + (size_t) st * TB * kTmaCh);
            tma_commit();
            const uint32_t nxt = i + STAGES - 1;  // goes into the stage tile i-1 was stored from
            if (nxt < ntiles) {
                tma_wait_read<1>();
                const uint32_t ns = nxt % STAGES;
                mbar_arrive_expect_tx(&full[ns], kTileBytes);
                tma_load_2d(tiles + (size_t) ns * TB * kTmaCh, &tm_in, (int) ch0, (int) (nxt * TB), &full[ns]);
            }
        }
    }
    if (lane == 0) tma_wait_all<0>();
    if (live) {
#pragma unroll
        for (int n = 0; n <= S; n++) {
            *reinterpret_cast<u64 *>(hist + (size_t) (n * 2 + 0) * C + c) = h[n][0];
            *reinterpret_cast<u64 *>(hist + (size_t) (n * 2 + 1) * C + c) = h[n][1];
        }
    }
}

}  // namespace olb

using namespace olb;

// ================================================================================================
// Host side
//
// Threading and ordering contract (include/olbank.h): one thread at a time runs process on a handle; setters, init and
// the state calls may come from another (control) thread.  Every entry point takes the handle's mutex.  Setters touch
// HOST tables only and mark what they changed; the device tables are brought up to date at the top of the next process
// call, on that call's stream (async copies ordered before the kernel), so a parameter change is atomic at a block
// boundary and can never reach a kernel that is already queued.  Init and set_state are queued the same way.
struct olb_biquad {
    std::mutex mu;
    StreamOrder order;
    int device = 0;
    uint32_t C = 0, S = 0;
    float sample_rate = 0.f;
    int mode = OLB_MODE_EXACT;      // as set by the caller
    int eff_mode = OLB_MODE_EXACT;  // arithmetic of the device table: FAST drops to FMA while some channel cannot be normalised
    int variant = 0;  // 0 auto, 1 direct, 2 tma (whole cascade per thread), 3 systolic (sections split over warps), 4 skewed section pairs
    Kinds kinds{};
    bool all_biquad = true;
    std::vector<float> coefs_host;   // [C][S][5] as the caller set them
    std::vector<uint8_t> unnormal;   // [C] 1 = this channel cannot run the FAST arithmetic (b0 == 0, scale out of range)
    uint32_t n_unnormal = 0;
    bool unnormal_valid = false;
    std::vector<double> node_scale;  // [C][S+1] node scaling of the device table: prod_{j >= n} b0_j in FAST, 1 otherwise, 0 = no table yet
    bool hist_is_zero = true;        // the device history is all zeros (fresh / after Init): it carries any scaling
    float *d_coef = nullptr;         // [S][5][C] effective
    float *d_hist = nullptr;         // [S+1][2][C]
    float *d_ratio = nullptr;        // [S+1][C] scratch of the history rescale
    uint32_t dirty_lo = 0, dirty_hi = 0;  // channels whose device rows are stale: [lo, hi)
    bool pending_init = false;            // Init() queued: history to zero at the next process
    std::vector<float> pending_hist;      // set_state() queued: [S+1][2][C] to upload at the next process (empty = none)
    const char *last_kernel = "none";
    // host-buffer path
    float *d_stage[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // [slot][in/out]
    size_t stage_elems = 0;
    cudaStream_t s_h2d = nullptr, s_run = nullptr, s_d2h = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_run[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
};

static void mark_dirty(olb_biquad *h, uint32_t lo, uint32_t hi) {
    if (h->dirty_lo == h->dirty_hi) {
        h->dirty_lo = lo;
        h->dirty_hi = hi;
    } else {
        h->dirty_lo = lo < h->dirty_lo ? lo : h->dirty_lo;
        h->dirty_hi = hi > h->dirty_hi ? hi : h->dirty_hi;
    }
}

// One-pole kinds as biquads (the FMA and FAST tiers run the biquad recurrence everywhere).
static void as_biquad(int kind, const float *in5, float *k) {
    if (kind == OLB_KIND_PORT) {
        k[0] = 1.0f - in5[0]; k[1] = 0.f; k[2] = 0.f; k[3] = -in5[0]; k[4] = 0.f;
    } else if (kind == OLB_KIND_LPF) {
        k[0] = in5[0]; k[1] = 0.f; k[2] = 0.f; k[3] = -(1.0f - in5[0]); k[4] = 0.f;
    } else {
        memcpy(k, in5, 5 * sizeof(float));
    }
}

// FAST-tier table of one channel: out[s] = {G (s == 0 only), r1, r2, -a1, -a2}; scale[n] = prod_{j >= n} b0_j (the scaling of
// node n; scale[S] = 1).  False when the channel cannot be normalised (b0 == 0 / non-finite, or a node scale outside
// [1e-30, 1e30]).  out / scale may be NULL (test only).
static bool fast_coefs(const olb_biquad *h, const float *chan, float *out, double *scale) {
    double g = 1.0;
    if (scale) scale[h->S] = 1.0;
    for (uint32_t s = h->S; s-- > 0;) {
        float k[5];
        as_biquad(h->kinds.k[s], chan + 5 * s, k);
        const double b0 = k[0];
        if (!(b0 != 0.0) || !std::isfinite(b0)) return false;
        g *= b0;
        if (!(fabs(g) >= 1e-30 && fabs(g) <= 1e30)) return false;
        if (scale) scale[s] = g;
        if (out) {
            out[5 * s + 0] = 0.f;
            out[5 * s + 1] = (float) ((double) k[1] / b0);
            out[5 * s + 2] = (float) ((double) k[2] / b0);
            out[5 * s + 3] = -k[3];
            out[5 * s + 4] = -k[4];
        }
    }
    if (out) out[0] = (float) g;
    return true;
}

static void effective_coefs(const olb_biquad *h, const float *in5, int s, float *out5) {
    float k[5];
    if (h->eff_mode != OLB_MODE_EXACT) as_biquad(h->kinds.k[s], in5, k);
    else memcpy(k, in5, sizeof(k));
    out5[0] = k[0]; out5[1] = k[1]; out5[2] = k[2]; out5[3] = -k[3]; out5[4] = -k[4];  // negation is exact
}

// hist[n][i][first + c] *= ratio[n][c]: brings the history of `count` channels from the node scaling it was written in to
// the scaling of the new coefficient table (one rounded product per element; the oracle does the same: orc_biquad_rescale_state).
__global__ void k_bq_rescale(float *hist, const float *ratio, uint32_t first, uint32_t count, uint32_t C, uint32_t nodes) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= count) return;
    for (uint32_t n = 0; n < nodes; n++) {
        const float r = ratio[(size_t) n * count + c];
        if (r == 1.0f) continue;
        float *p = hist + (size_t) (n * 2) * C + first + c;
        p[0] = __fmul_rn(p[0], r);
        p[C] = __fmul_rn(p[C], r);
    }
}

// Brings the device tables up to date on `st` (called at the top of process, under the handle's mutex).
static int apply_pending(olb_biquad *h, cudaStream_t st) {
    const uint32_t C = h->C, S = h->S;
    const size_t hist_bytes = (size_t) (S + 1) * 2 * C * sizeof(float);
    if (h->pending_init) {
        OLB_CUDA(cudaMemsetAsync(h->d_hist, 0, hist_bytes, st));
        h->hist_is_zero = true;
        h->pending_init = false;
    }
    if (!h->pending_hist.empty()) {  // set_state: the caller speaks the scaling of the CURRENT device table (see get_state)
        OLB_CUDA(cudaMemcpyAsync(h->d_hist, h->pending_hist.data(), hist_bytes, cudaMemcpyHostToDevice, st));
        OLB_CUDA(cudaStreamSynchronize(st));  // pageable source: keep it alive until the copy has read it
        h->pending_hist.clear();
        h->pending_hist.shrink_to_fit();
        h->hist_is_zero = false;
    }
    if (h->dirty_lo == h->dirty_hi) return OLB_OK;
    // which arithmetic: FAST needs every channel of the bank normalisable
    if (h->mode == OLB_MODE_FAST) {
        const uint32_t lo = h->unnormal_valid ? h->dirty_lo : 0, hi = h->unnormal_valid ? h->dirty_hi : C;
        for (uint32_t c = lo; c < hi; c++) {
            const uint8_t bad = fast_coefs(h, &h->coefs_host[(size_t) c * S * 5], nullptr, nullptr) ? 0 : 1;
            h->n_unnormal += bad - h->unnormal[c];
            h->unnormal[c] = bad;
        }
        h->unnormal_valid = true;
    }
    const int eff = (h->mode == OLB_MODE_FAST && h->n_unnormal) ? OLB_MODE_FMA : h->mode;
    if (eff != h->eff_mode) {  // the whole table changes arithmetic
        h->eff_mode = eff;
        h->dirty_lo = 0;
        h->dirty_hi = C;
    }
    const uint32_t first = h->dirty_lo, count = h->dirty_hi - h->dirty_lo;
    std::vector<float> stage((size_t) S * 5 * count), e((size_t) S * 5), ratio;
    std::vector<double> sc(S + 1);
    bool need_rescale = false;
    for (uint32_t i = 0; i < count; i++) {
        const float *chan = &h->coefs_host[(size_t) (first + i) * S * 5];
        if (h->eff_mode == OLB_MODE_FAST) {
            fast_coefs(h, chan, e.data(), sc.data());
        } else {
            for (uint32_t s = 0; s < S; s++) effective_coefs(h, chan + 5 * s, (int) s, &e[5 * s]);
            std::fill(sc.begin(), sc.end(), 1.0);
        }
        for (uint32_t s = 0; s < S; s++)
            for (int j = 0; j < 5; j++) stage[(size_t) (s * 5 + j) * count + i] = e[5 * s + j];
        double *old = &h->node_scale[(size_t) (first + i) * (S + 1)];
        for (uint32_t n = 0; n <= S; n++) {
            if (!h->hist_is_zero && old[n] != 0.0 && old[n] != sc[n]) {
                if (ratio.empty()) ratio.assign((size_t) (S + 1) * count, 1.0f);
                ratio[(size_t) n * count + i] = (float) (sc[n] / old[n]);
                need_rescale = true;
            }
            old[n] = sc[n];
        }
    }
    if (need_rescale) {
        if (!h->d_ratio) OLB_CUDA(cudaMalloc(&h->d_ratio, (size_t) (S + 1) * C * sizeof(float)));
        OLB_CUDA(cudaMemcpyAsync(h->d_ratio, ratio.data(), ratio.size() * sizeof(float), cudaMemcpyHostToDevice, st));
        k_bq_rescale<<<(count + 127) / 128, 128, 0, st>>>(h->d_hist, h->d_ratio, first, count, C, S + 1);
        OLB_LAUNCHED();
    }
    OLB_CUDA(cudaMemcpy2DAsync(h->d_coef + first, (size_t) C * sizeof(float), stage.data(), (size_t) count * sizeof(float),
                               (size_t) count * sizeof(float), (size_t) S * 5, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaStreamSynchronize(st));  // pageable sources (stage, ratio) die with this frame
    h->dirty_lo = h->dirty_hi = 0;
    return OLB_OK;
}

OLB_API int olb_biquad_create(int device, uint32_t channels, uint32_t sections, float sample_rate, olb_biquad **out) {
    OLB_ARG(out, "olb_biquad_create: out is NULL");
    *out = nullptr;
    OLB_ARG(channels > 0, "olb_biquad_create: channels must be > 0");
    OLB_ARG(sections >= 1 && sections <= (uint32_t) kMaxS, "olb_biquad_create: sections must be in 1..%d", kMaxS);
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_biquad *h = new olb_biquad();
    h->device = device;
    h->C = channels;
    h->S = sections;
    h->sample_rate = sample_rate;
    h->coefs_host.assign((size_t) channels * sections * 5, 0.f);
    for (size_t i = 0; i < (size_t) channels * sections; i++) h->coefs_host[i * 5] = 1.0f;
    h->unnormal.assign(channels, 0);
    h->node_scale.assign((size_t) channels * (sections + 1), 0.0);
    h->dirty_lo = 0;
    h->dirty_hi = channels;
    cudaError_t e = cudaMalloc(&h->d_coef, (size_t) sections * 5 * channels * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_hist, (size_t) (sections + 1) * 2 * channels * sizeof(float));
    if (e == cudaSuccess) e = cudaMemset(h->d_hist, 0, (size_t) (sections + 1) * 2 * channels * sizeof(float));
    if (e != cudaSuccess) {
        cudaFree(h->d_coef);
        cudaFree(h->d_hist);
        delete h;
        return fail_cuda(e, "allocating bank tables", __FILE__, __LINE__);
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_biquad_destroy(olb_biquad *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    h->order.quiesce();
    cudaFree(h->d_coef);
    cudaFree(h->d_hist);
    cudaFree(h->d_ratio);
    for (int i = 0; i < 2; i++) {
        cudaFree(h->d_stage[i][0]);
        cudaFree(h->d_stage[i][1]);
        if (h->ev_in[i]) cudaEventDestroy(h->ev_in[i]);
        if (h->ev_run[i]) cudaEventDestroy(h->ev_run[i]);
        if (h->ev_out[i]) cudaEventDestroy(h->ev_out[i]);
    }
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_run) cudaStreamDestroy(h->s_run);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    h->order.release();
    delete h;
    return OLB_OK;
}

OLB_API int olb_biquad_init(olb_biquad *h, float sample_rate) {
    OLB_ARG(h, "olb_biquad_init: NULL handle");
    std::lock_guard<std::mutex> lk(h->mu);
    h->sample_rate = sample_rate;
    h->pending_init = true;
    h->pending_hist.clear();
    return OLB_OK;
}

OLB_API int olb_biquad_set_kinds(olb_biquad *h, const uint8_t *kinds, uint32_t sections) {
    OLB_ARG(h && kinds, "olb_biquad_set_kinds: NULL argument");
    OLB_ARG(sections == h->S, "olb_biquad_set_kinds: bank has %u sections, got %u", h->S, sections);
    for (uint32_t s = 0; s < sections; s++)
        OLB_ARG(kinds[s] <= OLB_KIND_LPF, "olb_biquad_set_kinds: unknown kind %u at section %u", kinds[s], s);
    std::lock_guard<std::mutex> lk(h->mu);
    bool all = true;
    for (uint32_t s = 0; s < sections; s++) {
        h->kinds.k[s] = kinds[s];
        all = all && kinds[s] == OLB_KIND_BIQUAD;
    }
    h->all_biquad = all;
    h->unnormal_valid = false;
    mark_dirty(h, 0, h->C);
    return OLB_OK;
}

OLB_API int olb_biquad_set_mode(olb_biquad *h, int mode) {
    OLB_ARG(h, "olb_biquad_set_mode: NULL handle");
    OLB_ARG(mode == OLB_MODE_EXACT || mode == OLB_MODE_FMA || mode == OLB_MODE_FAST, "olb_biquad_set_mode: unknown mode %d", mode);
    std::lock_guard<std::mutex> lk(h->mu);
    if (mode != h->mode) {
        h->unnormal_valid = false;
        mark_dirty(h, 0, h->C);
    }
    h->mode = mode;
    return OLB_OK;
}

OLB_API int olb_biquad_set_variant(olb_biquad *h, int variant) {
    OLB_ARG(h, "olb_biquad_set_variant: NULL handle");
    OLB_ARG(variant >= 0 && variant <= 4, "olb_biquad_set_variant: unknown variant %d", variant);
    std::lock_guard<std::mutex> lk(h->mu);
    h->variant = variant;
    return OLB_OK;
}

OLB_API const char *olb_biquad_last_kernel(olb_biquad *h) { return h ? h->last_kernel : "none"; }

OLB_API int olb_biquad_set_coeffs(olb_biquad *h, uint32_t first, uint32_t count, const float *coefs_host) {
    OLB_ARG(h && coefs_host, "olb_biquad_set_coeffs: NULL argument");
    OLB_ARG((uint64_t) first + count <= h->C, "olb_biquad_set_coeffs: range [%u, %u) exceeds %u channels", first, first + count, h->C);
    if (count == 0) return OLB_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    memcpy(&h->coefs_host[(size_t) first * h->S * 5], coefs_host, (size_t) count * h->S * 5 * sizeof(float));
    mark_dirty(h, first, first + count);
    return OLB_OK;
}

OLB_API int olb_biquad_design(int type, double fs, double f0, double q, double gain_db, float *c5) {
    OLB_ARG(c5, "olb_biquad_design: coefs5 is NULL");
    OLB_ARG(type >= 0 && type <= 3, "olb_biquad_design: unknown type %d", type);
    OLB_ARG(fs > 0 && f0 > 0 && f0 < fs / 2 && q > 0, "olb_biquad_design: need 0 < f0 < fs/2 and q > 0");
    const double w0 = 2.0 * M_PI * f0 / fs, cw = cos(w0), sw = sin(w0), alpha = sw / (2.0 * q);
    double b0, b1, b2, a0, a1, a2;
    switch (type) {
        case 0: b1 = 1 - cw; b0 = b2 = b1 / 2; a0 = 1 + alpha; a1 = -2 * cw; a2 = 1 - alpha; break;
        case 1: b0 = b2 = (1 + cw) / 2; b1 = -(1 + cw); a0 = 1 + alpha; a1 = -2 * cw; a2 = 1 - alpha; break;
        case 2: {
            const double A = pow(10.0, gain_db / 40.0);
            b0 = 1 + alpha * A; b1 = -2 * cw; b2 = 1 - alpha * A; a0 = 1 + alpha / A; a1 = -2 * cw; a2 = 1 - alpha / A;
            break;
        }
        default: {
            const double p = exp(-w0);
            b0 = 1 - p; b1 = 0; b2 = 0; a0 = 1; a1 = -p; a2 = 0;
        }
    }
    c5[0] = (float) (b0 / a0); c5[1] = (float) (b1 / a0); c5[2] = (float) (b2 / a0);
    c5[3] = (float) (a1 / a0); c5[4] = (float) (a2 / a0);
    return OLB_OK;
}

// ---- state in the oracle's layout [C][S][4] = x1 x2 y1 y2 ---------------------------------------
// Sections that never touch a slot (one-pole kinds in EXACT mode keep only y1) report 0 there, as a
// freshly Init()ed oracle would.  In FAST the values carry the node scaling of the table the last process call ran with.
OLB_API int olb_biquad_get_state(olb_biquad *h, float *state_host) {
    OLB_ARG(h && state_host, "olb_biquad_get_state: NULL argument");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    const uint32_t C = h->C, S = h->S;
    std::vector<float> hist((size_t) (S + 1) * 2 * C, 0.f);
    if (!h->pending_hist.empty()) {
        hist = h->pending_hist;
    } else if (!h->pending_init) {
        OLB_CUDA(h->order.sync());  // everything queued so far has written its history
        OLB_CUDA(cudaMemcpy(hist.data(), h->d_hist, hist.size() * sizeof(float), cudaMemcpyDeviceToHost));
    }
    for (uint32_t c = 0; c < C; c++)
        for (uint32_t s = 0; s < S; s++) {
            const bool full = h->eff_mode != OLB_MODE_EXACT || h->kinds.k[s] == OLB_KIND_BIQUAD;
            float *o = state_host + ((size_t) c * S + s) * 4;
            o[0] = full ? hist[(size_t) (s * 2 + 0) * C + c] : 0.f;
            o[1] = full ? hist[(size_t) (s * 2 + 1) * C + c] : 0.f;
            o[2] = hist[(size_t) ((s + 1) * 2 + 0) * C + c];
            o[3] = full ? hist[(size_t) ((s + 1) * 2 + 1) * C + c] : 0.f;
        }
    return OLB_OK;
}

OLB_API int olb_biquad_set_state(olb_biquad *h, const float *state_host) {
    OLB_ARG(h && state_host, "olb_biquad_set_state: NULL argument");
    std::lock_guard<std::mutex> lk(h->mu);
    const uint32_t C = h->C, S = h->S;
    std::vector<float> hist((size_t) (S + 1) * 2 * C, 0.f);
    auto full = [&](uint32_t s) { return h->eff_mode != OLB_MODE_EXACT || h->kinds.k[s] == OLB_KIND_BIQUAD; };
    for (uint32_t c = 0; c < C; c++) {
        const float *st = state_host + (size_t) c * S * 4;
        hist[(size_t) 0 * C + c] = st[0];
        hist[(size_t) 1 * C + c] = st[1];
        for (uint32_t s = 0; s < S; s++) {
            const float *o = st + (size_t) s * 4;
            hist[(size_t) ((s + 1) * 2 + 0) * C + c] = o[2];
            float h2 = 0.f;
            if (full(s)) h2 = o[3];
            else if (s + 1 < S && full(s + 1)) h2 = st[(size_t) (s + 1) * 4 + 1];
            hist[(size_t) ((s + 1) * 2 + 1) * C + c] = h2;
        }
    }
    h->pending_hist.swap(hist);
    h->pending_init = false;
    return OLB_OK;
}

// ---- launch -------------------------------------------------------------------------------------
// k_bq_tma runs ONE warp per CTA: the warp is the scheduling unit, so the hardware block scheduler
// spreads warp-groups of 64 channels over the 148 SMs by itself (65 536 channels = 1 024 CTAs = 6.9 per
// SM, all resident in one wave).  Shared memory per CTA = STAGES x TB x 256 B (+ barriers).
constexpr int kWarps = 1;
struct TmaCfg {
    int tb, stages;
};
// default tile shape; OLB_BQ_TMA_CFG="<TB>x<STAGES>" selects another compiled shape (tuning aid, S = 8 only)
static TmaCfg tma_cfg() {
    static TmaCfg cfg = [] {
        TmaCfg c{32, 3};
        if (const char *e = getenv("OLB_BQ_TMA_CFG")) {
            int a = 0, b = 0;
            if (sscanf(e, "%dx%d", &a, &b) == 2) c = TmaCfg{a, b};
        }
        return c;
    }();
    return cfg;
}

template <int S, int MODE, int TB, int STAGES>
static int launch_tma_cfg(olb_biquad *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr size_t smem = (size_t) kWarps * STAGES * TB * kTmaCh * sizeof(float) + (size_t) kWarps * STAGES * sizeof(uint64_t);
    TensorMap tin, tout;
    int rc = make_tmap_frames(&tin, in, frames, h->C, kTmaCh, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, h->C, kTmaCh, TB);
    if (rc) return rc;
    auto kern = k_bq_tma<S, MODE, TB, STAGES, kWarps>;
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));  // per device, cheap
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    const uint32_t groups = (h->C + kTmaCh - 1) / kTmaCh;
    const float onef = 1.0f, nzf = -0.0f;
    uint32_t ob, nb;
    memcpy(&ob, &onef, 4);
    memcpy(&nb, &nzf, 4);
    const u64 one = ((u64) ob << 32) | ob, negzero = ((u64) nb << 32) | nb;
    kern<<<(groups + kWarps - 1) / kWarps, kWarps * 32, smem, st>>>(tin, tout, frames, h->C, h->d_coef, h->d_hist, one, negzero);
    OLB_LAUNCHED();
    return OLB_OK;
}

template <int S, int MODE>
static int launch_tma(olb_biquad *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    if (S == 8) {
        const TmaCfg c = tma_cfg();
        if (c.tb == 16 && c.stages == 4) return launch_tma_cfg<S, MODE, 16, 4>(h, in, out, frames, st);
        if (c.tb == 32 && c.stages == 4) return launch_tma_cfg<S, MODE, 32, 4>(h, in, out, frames, st);
        if (c.tb == 16 && c.stages == 6) return launch_tma_cfg<S, MODE, 16, 6>(h, in, out, frames, st);
        if (c.tb == 8 && c.stages == 8) return launch_tma_cfg<S, MODE, 8, 8>(h, in, out, frames, st);
        if (c.tb == 8 && c.stages == 4) return launch_tma_cfg<S, MODE, 8, 4>(h, in, out, frames, st);
    }
    return launch_tma_cfg<S, MODE, 32, 3>(h, in, out, frames, st);
}

// ---- k_bq_sys shapes ----------------------------------------------------------------------------
// OLB_BQ_SYS_CFG="<SPW>x<TB>x<NST>x<PACK>" picks one of the compiled shapes (tuning aid).
static bool sys_cfg_env(SysCfg *c) {
    if (const char *e = getenv("OLB_BQ_SYS_CFG"))  // read per call: a tuning aid
        if (sscanf(e, "%dx%dx%dx%d", &c->spw, &c->tb, &c->nst, &c->pack) == 4) return true;
    return false;
}

// Automatic shape: packed lanes when that still gives every SM a CTA, scalar lanes below that.
static SysCfg sys_cfg_auto(uint32_t C, uint32_t S) {
    SysCfg c;
    if (sys_cfg_env(&c)) return c;
    if (C >= 64u * 148u) return SysCfg{S % 2 == 0 ? 2 : 1, 16, 7, 2};
    // few channels: one section per warp, scalar lanes, and a DEEP ring -- a CTA must keep ~100 KiB of loads in flight
    // to cover HBM latency on its own (profiles/r01_run7_small_banks.txt): 24 x 8 KiB stages with one CTA per SM,
    // 24 x 4 KiB when two CTAs share an SM
    if (S == 8) return C <= 32u * 148u ? SysCfg{1, 64, 24, 1} : SysCfg{1, 32, 24, 1};
    return SysCfg{1, 16, 12, 1};
}

static bool sys_supported(uint32_t S) { return S == 2 || S == 4 || S == 6 || S == 8; }

#define OLB_SYS_TRY(S_, SPW_, TB_, NST_, PACK_)                                                        \
    if (S == S_ && c.spw == SPW_ && c.tb == TB_ && c.nst == NST_ && c.pack == PACK_)                   \
        return launch_sys_cfg<S_, SPW_, MODE, TB_, NST_, PACK_>(h->C, h->d_coef, h->d_hist, in, out, frames, st);

template <int MODE>
static int launch_sys(olb_biquad *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    const uint32_t S = h->S;
    SysCfg c = sys_cfg_auto(h->C, S);
    if (S % c.spw) c.spw = 1;
    OLB_SYS_TRY(8, 2, 16, 7, 2)
    OLB_SYS_TRY(8, 2, 16, 6, 2)
    OLB_SYS_TRY(8, 2, 8, 12, 2)
    OLB_SYS_TRY(8, 2, 32, 6, 2)
    OLB_SYS_TRY(8, 1, 8, 12, 2)
    OLB_SYS_TRY(8, 1, 16, 12, 2)
    OLB_SYS_TRY(8, 4, 16, 6, 2)
    OLB_SYS_TRY(8, 4, 32, 3, 2)
    OLB_SYS_TRY(8, 1, 16, 12, 1)
    OLB_SYS_TRY(8, 1, 8, 16, 1)
    OLB_SYS_TRY(8, 2, 16, 8, 1)
    OLB_SYS_TRY(8, 1, 32, 24, 1)
    OLB_SYS_TRY(8, 1, 64, 24, 1)
    OLB_SYS_TRY(8, 1, 32, 48, 1)
    OLB_SYS_TRY(8, 2, 64, 24, 1)
    OLB_SYS_TRY(8, 1, 32, 24, 2)
    OLB_SYS_TRY(8, 1, 32, 12, 2)
    OLB_SYS_TRY(8, 2, 32, 24, 2)
    OLB_SYS_TRY(8, 2, 16, 12, 1)
    OLB_SYS_TRY(6, 2, 16, 7, 2)
    OLB_SYS_TRY(6, 1, 16, 12, 1)
    OLB_SYS_TRY(4, 2, 16, 7, 2)
    OLB_SYS_TRY(4, 1, 16, 12, 1)
    OLB_SYS_TRY(2, 2, 16, 7, 2)
    OLB_SYS_TRY(2, 1, 16, 12, 1)
    set_error("olb_biquad_process: no compiled systolic shape for S=%u spw=%d tb=%d nst=%d pack=%d", S, c.spw, c.tb, c.nst, c.pack);
    return OLB_E_ARG;
}

// ---- k_bq_skew shapes ---------------------------------------------------------------------------
// Tile height x ring depth x lag [r = rolled FAST tile].  A CTA is 32 channels; what changes with the bank size is how many
// CTAs share an SM.  A lone CTA must keep ~100 KiB of loads in flight to cover the HBM latency by itself
// (profiles/r01_run7_small_banks.txt), and with one or two warps per scheduler the per-tile bookkeeping (barrier wait,
// pipeline fill and drain: ~650 cycles) wants TALL tiles (profiles/r02_run13_skew_sweep.txt: 4 096 channels 31.8 / 24.9 /
// 21.6 / 20.9 cycles per frame at 48 / 96 / 192 / 256 frames).  OLB_BQ_SKEW_CFG="<TB>x<NST>x<LAG>[r]" picks a compiled shape (tuning aid).
struct SkewCfg {
    int tb, nst, lag;
    bool roll;
    int spw = 0;        // > 0: k_bq_roll, two channels per lane and spw sections per warp (FAST tier, 8 sections)
};
static SkewCfg skew_cfg_auto(uint32_t C, uint32_t frames, bool fast8) {
    SkewCfg env{};
    if (const char *e = getenv("OLB_BQ_SKEW_CFG")) {  // read per call: a tuning aid, cheap next to a launch
        char r = 0;
        int spw = 2;
        const int n = sscanf(e, "%dx%dx%d%c%d", &env.tb, &env.nst, &env.lag, &r, &spw);
        if (n >= 3) {
            env.roll = n >= 4 && r == 'r';
            env.spw = n >= 4 && r == 'c' ? spw : 0;
            return env;
        }
    }
    const uint32_t ctas = (C + 31) / 32, per_sm = (ctas + 147) / 148;
    // FAST tier, 8 sections: k_bq_roll (64-channel CTAs, two channels per lane) where it puts ONE warp on a scheduler:
    // 4 737 .. 9 472 channels with two sections per warp (8 192 channels: 37.6 cycles per frame against 40.2), 14 209 .. 18 944
    // with four (16 384: 63.6 against 73.5); in between three 32-channel CTAs per SM are the finer fit (12 288: 55.1 against
    // 66.7).  profiles/r02_run44_roll_sweep.txt.  lag: 4 = stores four positions late, 5 = and two blocks per trip.
    if (fast8 && per_sm == 2 && frames >= 512) {
        SkewCfg c = frames >= 1024 ? SkewCfg{128, 6, 4, false} : SkewCfg{64, 6, 4, false};
        c.spw = 2;
        return c;
    }
    if (fast8 && per_sm == 4 && ctas > 3 * 148 && frames >= 512) {
        SkewCfg c{64, 6, 5, false};
        c.spw = 4;
        return c;
    }
    if (!fast8) {  // EXACT / FMA tiers and other section counts: the unrolled tile (the rolled one is the FAST tier's)
        if (per_sm <= 1) return SkewCfg{48, 32, 4, false};
        if (per_sm <= 2) return SkewCfg{48, 16, 4, false};
    }
    // short blocks keep the short tiles (what is left of the stream's first and last tile beside whole 8-row blocks runs frame by frame)
    if (per_sm <= 1)  // the SM to itself: 192 KiB (256 rows is the tallest TMA box)
        return frames >= 4096 ? SkewCfg{256, 6, 4, true} : frames >= 3072 ? SkewCfg{192, 8, 4, true} : SkewCfg{48, 32, 4, true};
    if (per_sm <= 2)  // 96 KiB each
        return frames >= 2048 ? SkewCfg{128, 6, 4, true} : frames >= 1536 ? SkewCfg{96, 8, 4, true} : SkewCfg{48, 16, 4, true};
    if (per_sm <= 4) return SkewCfg{48, 8, 4, false};  // 48 KiB
    if (per_sm <= 8) return SkewCfg{24, 8, 4, false};  // 24 KiB
    return SkewCfg{24, 4, 4, false};                   // 12 KiB: up to 18 CTAs per SM
}

#define OLB_SKEW_TRY(S_, TB_, NST_, D_, LAG_, ROLL_)                                                                       \
    if constexpr (!(ROLL_) || MODE == OLB_MODE_FAST) { /* no unrolled 256-frame EXACT tiles behind the rolled shapes' names */ \
        if (S == S_ && c.tb == TB_ && c.nst == NST_ && c.lag == LAG_ && c.roll == ROLL_)                                   \
            return launch_skew_cfg<S_, MODE, TB_, NST_, D_, LAG_, ROLL_>(h->C, h->d_coef, h->d_hist, in, out, frames, st); \
    }

template <int MODE>
static int launch_skew(olb_biquad *h, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    const uint32_t S = h->S;
    SkewCfg c = skew_cfg_auto(h->C, frames, MODE == OLB_MODE_FAST && h->S == 8);
    if (MODE != OLB_MODE_FAST) c.roll = false;  // the rolled tile is the FAST tier's
    if (S != 8) c = SkewCfg{24, 8, 4, false};
    if (c.spw && MODE == OLB_MODE_FAST && S == 8) {
        h->last_kernel = "k_bq_roll";
#define OLB_ROLL_TRY(SPW_, TB_, NST_)                                                                                                                       \
    if (c.spw == SPW_ && c.tb == TB_ && c.nst == NST_) {                                                                                                    \
        if (c.lag == 0) return launch_roll_cfg<8, SPW_, TB_, NST_, 0, 1>(h->C, h->d_coef, h->d_hist, in, out, frames, st);                                  \
        if (c.lag == 1) return launch_roll_cfg<8, SPW_, TB_, NST_, 0, 2>(h->C, h->d_coef, h->d_hist, in, out, frames, st);                                  \
        if (c.lag == 5) return launch_roll_cfg<8, SPW_, TB_, NST_, 4, 2>(h->C, h->d_coef, h->d_hist, in, out, frames, st);                                  \
        return launch_roll_cfg<8, SPW_, TB_, NST_, 4, 1>(h->C, h->d_coef, h->d_hist, in, out, frames, st);                                                  \
    }
        OLB_ROLL_TRY(2, 128, 6)
        OLB_ROLL_TRY(2, 96, 8)
        OLB_ROLL_TRY(2, 64, 6)
        OLB_ROLL_TRY(2, 32, 6)
        OLB_ROLL_TRY(4, 128, 3)
        OLB_ROLL_TRY(4, 64, 6)
        OLB_ROLL_TRY(4, 32, 6)
#undef OLB_ROLL_TRY
        set_error("olb_biquad_process: no compiled k_bq_roll shape spw=%d tb=%d nst=%d", c.spw, c.tb, c.nst);
        return OLB_E_ARG;
    }
    c.spw = 0;
    OLB_SKEW_TRY(8, 256, 6, 4, 4, true)
    OLB_SKEW_TRY(8, 192, 8, 4, 4, true)
    OLB_SKEW_TRY(8, 128, 6, 4, 4, true)
    OLB_SKEW_TRY(8, 96, 8, 4, 4, true)
    OLB_SKEW_TRY(8, 96, 16, 4, 4, true)
    OLB_SKEW_TRY(8, 48, 32, 4, 4, true)
    OLB_SKEW_TRY(8, 48, 16, 4, 4, true)
    OLB_SKEW_TRY(8, 48, 8, 4, 4, true)
    OLB_SKEW_TRY(8, 24, 8, 4, 4, true)
    OLB_SKEW_TRY(8, 192, 8, 4, 4, false)
    OLB_SKEW_TRY(8, 96, 8, 4, 4, false)
    OLB_SKEW_TRY(8, 96, 16, 4, 4, false)
    OLB_SKEW_TRY(8, 48, 32, 4, 4, false)
    OLB_SKEW_TRY(8, 48, 16, 4, 4, false)
    OLB_SKEW_TRY(8, 48, 8, 4, 4, false)
    OLB_SKEW_TRY(8, 24, 8, 4, 4, false)
    OLB_SKEW_TRY(8, 24, 4, 4, 4, false)
    OLB_SKEW_TRY(8, 24, 6, 4, 4, false)
    OLB_SKEW_TRY(8, 48, 32, 4, 6, false)
    OLB_SKEW_TRY(8, 48, 16, 4, 6, false)
    OLB_SKEW_TRY(6, 24, 8, 4, 4, false)
    OLB_SKEW_TRY(4, 24, 8, 4, 4, false)
    OLB_SKEW_TRY(2, 24, 8, 4, 4, false)
    set_error("olb_biquad_process: no compiled skewed shape for S=%u tb=%d nst=%d lag=%d roll=%d", S, c.tb, c.nst, c.lag, (int) c.roll);
    return OLB_E_ARG;
}

template <int S, int MODE, bool GENERIC>
static int launch_direct(olb_biquad *h, const float *in, float *out, uint64_t frames, int layout, cudaStream_t st) {
    if (layout == OLB_LAYOUT_FRAMES) {
        k_bq_direct<S, MODE, GENERIC, 8><<<(h->C + 127) / 128, 128, 0, st>>>(in, out, frames, h->C, h->d_coef, h->d_hist, h->kinds);
    } else {
        k_bq_planar<S, MODE, GENERIC><<<(h->C + 127) / 128, 128, 0, st>>>(in, out, frames, h->C, h->d_coef, h->d_hist, h->kinds);
    }
    OLB_LAUNCHED();
    return OLB_OK;
}

template <int S>
static int dispatch_s(olb_biquad *h, const float *in, float *out, uint64_t frames, int layout, bool tma, cudaStream_t st) {
    const int m = h->eff_mode;
    if (tma) {
        return m == OLB_MODE_FAST  ? launch_tma<S, OLB_MODE_FAST>(h, in, out, (uint32_t) frames, st)
               : m == OLB_MODE_FMA ? launch_tma<S, OLB_MODE_FMA>(h, in, out, (uint32_t) frames, st)
                                   : launch_tma<S, OLB_MODE_EXACT>(h, in, out, (uint32_t) frames, st);
    }
    if (m == OLB_MODE_FAST) return launch_direct<S, OLB_MODE_FAST, false>(h, in, out, frames, layout, st);
    if (m == OLB_MODE_FMA) return launch_direct<S, OLB_MODE_FMA, false>(h, in, out, frames, layout, st);
    if (h->all_biquad) return launch_direct<S, OLB_MODE_EXACT, false>(h, in, out, frames, layout, st);
    return launch_direct<S, OLB_MODE_EXACT, true>(h, in, out, frames, layout, st);
}

static int process_dev(olb_biquad *h, const float *in, float *out, uint64_t frames, int layout, cudaStream_t st) {
    int rc0 = apply_pending(h, st);
    if (rc0) return rc0;
    h->hist_is_zero = false;  // whatever runs next writes a history in the table's scaling
    const uint32_t S = h->S;
    if (S > 8) {
        h->last_kernel = "k_bq_runtime";
        if (h->eff_mode == OLB_MODE_FAST)
            k_bq_runtime<OLB_MODE_FAST><<<(h->C + 127) / 128, 128, 0, st>>>(in, out, frames, h->C, (int) S, h->d_coef, h->d_hist, h->kinds, layout);
        else if (h->eff_mode == OLB_MODE_FMA)
            k_bq_runtime<OLB_MODE_FMA><<<(h->C + 127) / 128, 128, 0, st>>>(in, out, frames, h->C, (int) S, h->d_coef, h->d_hist, h->kinds, layout);
        else
            k_bq_runtime<OLB_MODE_EXACT><<<(h->C + 127) / 128, 128, 0, st>>>(in, out, frames, h->C, (int) S, h->d_coef, h->d_hist, h->kinds, layout);
        OLB_LAUNCHED();
        return OLB_OK;
    }
    const bool tma_ok = layout == OLB_LAYOUT_FRAMES && (h->eff_mode != OLB_MODE_EXACT || h->all_biquad) && (h->C % 4) == 0 &&
                        (((uintptr_t) in | (uintptr_t) out) & 15) == 0 && frames < (1ull << 31);
    if (h->variant >= 2 && !tma_ok) {
        set_error("olb_biquad_process: the TMA variants need FRAMES layout, channels %% 4 == 0, 16-byte aligned buffers and biquad-only kinds in EXACT mode");
        return OLB_E_ARG;
    }
    if (h->variant >= 3 && !sys_supported(S)) {
        set_error("olb_biquad_process: the systolic and skewed variants are compiled for 2, 4, 6 and 8 sections");
        return OLB_E_ARG;
    }
    if (h->variant == 4) {
        h->last_kernel = "k_bq_skew";
        const int m = h->eff_mode;
        return m == OLB_MODE_FAST  ? launch_skew<OLB_MODE_FAST>(h, in, out, (uint32_t) frames, st)
               : m == OLB_MODE_FMA ? launch_skew<OLB_MODE_FMA>(h, in, out, (uint32_t) frames, st)
                                   : launch_skew<OLB_MODE_EXACT>(h, in, out, (uint32_t) frames, st);
    }
    // automatic choice (profiles/r01_run4_sys_sweep.txt, profiles/r02_run2_sweep.txt): what limits every one of these
    // kernels is the register-operand bandwidth of an SM sub-partition, so the question is only how evenly a bank can be
    // spread over the 592 of them.  >= 32 768 channels: the whole cascade per thread (k_bq_tma: fewest instructions per
    // channel-frame); 512 .. 32 767 channels of an 8-section bank: skewed section pairs (k_bq_skew: units of 32 channels
    // x 2 sections, 1-7 per sub-partition where k_bq_tma would leave most of them idle; below 4 737 channels only for
    // blocks of 3 072 frames or more, whose tall tiles pay); other small banks: one section per warp (k_bq_sys); tiny or
    // unaligned ones run one channel per thread without TMA
    bool tma = tma_ok && (h->variant == 2 || (h->variant == 0 && h->C >= 32768u) || (h->variant == 0 && S != 8 && h->C >= 64u * 148u));
    const bool skew = !tma && tma_ok && S == 8 && h->variant == 0 && (h->C > 32u * 148u || (h->C >= 32u * 16u && frames >= 3072));
    if (skew) {
        h->last_kernel = "k_bq_skew";
        const int m = h->eff_mode;
        return m == OLB_MODE_FAST  ? launch_skew<OLB_MODE_FAST>(h, in, out, (uint32_t) frames, st)
               : m == OLB_MODE_FMA ? launch_skew<OLB_MODE_FMA>(h, in, out, (uint32_t) frames, st)
                                   : launch_skew<OLB_MODE_EXACT>(h, in, out, (uint32_t) frames, st);
    }
    bool sys = !tma && tma_ok && sys_supported(S) && (h->variant == 3 || (h->variant == 0 && h->C >= 32u * 16u));
    if (sys) {
        h->last_kernel = "k_bq_sys";
        const int m = h->eff_mode;
        return m == OLB_MODE_FAST  ? launch_sys<OLB_MODE_FAST>(h, in, out, (uint32_t) frames, st)
               : m == OLB_MODE_FMA ? launch_sys<OLB_MODE_FMA>(h, in, out, (uint32_t) frames, st)
                                   : launch_sys<OLB_MODE_EXACT>(h, in, out, (uint32_t) frames, st);
    }
    h->last_kernel = tma ? "k_bq_tma" : (layout == OLB_LAYOUT_FRAMES ? "k_bq_direct" : "k_bq_planar");
    switch (S) {
        case 1: return dispatch_s<1>(h, in, out, frames, layout, tma, st);
        case 2: return dispatch_s<2>(h, in, out, frames, layout, tma, st);
        case 3: return dispatch_s<3>(h, in, out, frames, layout, tma, st);
        case 4: return dispatch_s<4>(h, in, out, frames, layout, tma, st);
        case 5: return dispatch_s<5>(h, in, out, frames, layout, tma, st);
        case 6: return dispatch_s<6>(h, in, out, frames, layout, tma, st);
        case 7: return dispatch_s<7>(h, in, out, frames, layout, tma, st);
        default: return dispatch_s<8>(h, in, out, frames, layout, tma, st);
    }
}

OLB_API int olb_biquad_process(olb_biquad *h, const float *in_dev, float *out_dev, uint64_t frames, int layout, void *stream) {
    OLB_ARG(h, "olb_biquad_process: NULL handle");
    OLB_ARG(layout == OLB_LAYOUT_FRAMES || layout == OLB_LAYOUT_PLANAR, "olb_biquad_process: bad layout %d", layout);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_dev && out_dev, "olb_biquad_process: NULL buffer");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    OLB_CUDA(h->order.enter((cudaStream_t) stream));
    return process_dev(h, in_dev, out_dev, frames, layout, (cudaStream_t) stream);
}

// Host buffers: the block is cut in time into chunks; chunk k+1 is copied in while chunk k is
// filtered and chunk k-1 is copied out (three streams, two staging slots).  PLANAR chunks are
// strided 2-D copies.  State flows from chunk to chunk on the device.
OLB_API int olb_biquad_process_host(olb_biquad *h, const float *in_host, float *out_host, uint64_t frames, int layout) {
    OLB_ARG(h, "olb_biquad_process_host: NULL handle");
    OLB_ARG(layout == OLB_LAYOUT_FRAMES || layout == OLB_LAYOUT_PLANAR, "olb_biquad_process_host: bad layout %d", layout);
    if (frames == 0) return OLB_OK;
    OLB_ARG(in_host && out_host, "olb_biquad_process_host: NULL buffer");
    std::lock_guard<std::mutex> lk(h->mu);
    DeviceGuard g(h->device);
    const uint64_t C = h->C;
    // ~64 MiB per staging buffer, whole frames, multiple of the TMA tile height
    uint64_t chunk = (64ull << 20) / (C * sizeof(float));
    chunk = chunk < 32 ? 32 : (chunk / 32) * 32;
    if (chunk > frames) chunk = frames;
    if (!h->s_h2d) {
        OLB_CUDA(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
        OLB_CUDA(cudaStreamCreateWithFlags(&h->s_run, cudaStreamNonBlocking));
        OLB_CUDA(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) {
            OLB_CUDA(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
            OLB_CUDA(cudaEventCreateWithFlags(&h->ev_run[i], cudaEventDisableTiming));
            OLB_CUDA(cudaEventCreateWithFlags(&h->ev_out[i], cudaEventDisableTiming));
        }
    }
    if (h->stage_elems < chunk * C) {
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++) {
                cudaFree(h->d_stage[i][j]);
                h->d_stage[i][j] = nullptr;
                OLB_CUDA(cudaMalloc(&h->d_stage[i][j], chunk * C * sizeof(float)));
            }
        h->stage_elems = chunk * C;
    }
    OLB_CUDA(h->order.enter(h->s_run));
    uint64_t k = 0;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk, k++) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        const int slot = (int) (k & 1);
        float *din = h->d_stage[slot][0], *dout = h->d_stage[slot][1];
        if (k >= 2) {
            OLB_CUDA(cudaStreamWaitEvent(h->s_h2d, h->ev_run[slot], 0));  // kernel k-2 has consumed din
        }
        if (layout == OLB_LAYOUT_FRAMES)
            OLB_CUDA(cudaMemcpyAsync(din, in_host + t0 * C, n * C * sizeof(float), cudaMemcpyHostToDevice, h->s_h2d));
        else
            OLB_CUDA(cudaMemcpy2DAsync(din, n * sizeof(float), in_host + t0, frames * sizeof(float), n * sizeof(float), C,
                                       cudaMemcpyHostToDevice, h->s_h2d));
        OLB_CUDA(cudaEventRecord(h->ev_in[slot], h->s_h2d));
        OLB_CUDA(cudaStreamWaitEvent(h->s_run, h->ev_in[slot], 0));
        if (k >= 2) OLB_CUDA(cudaStreamWaitEvent(h->s_run, h->ev_out[slot], 0));  // copy-out k-2 has drained dout
        int rc = process_dev(h, din, dout, n, layout, h->s_run);
        if (rc) return rc;
        OLB_CUDA(cudaEventRecord(h->ev_run[slot], h->s_run));
        OLB_CUDA(cudaStreamWaitEvent(h->s_d2h, h->ev_run[slot], 0));
        if (layout == OLB_LAYOUT_FRAMES)
            OLB_CUDA(cudaMemcpyAsync(out_host + t0 * C, dout, n * C * sizeof(float), cudaMemcpyDeviceToHost, h->s_d2h));
        else
            OLB_CUDA(cudaMemcpy2DAsync(out_host + t0, frames * sizeof(float), dout, n * sizeof(float), n * sizeof(float), C,
                                       cudaMemcpyDeviceToHost, h->s_d2h));
        OLB_CUDA(cudaEventRecord(h->ev_out[slot], h->s_d2h));
    }
    OLB_CUDA(cudaStreamSynchronize(h->s_d2h));
    OLB_CUDA(cudaStreamSynchronize(h->s_run));
    return OLB_OK;
}
