// biquad_sys.cuh -- k_bq_sys: the section-pipelined ("systolic") cascade kernel.
//
// Why: with one whole cascade per thread the unit of scheduling is (64 channels x all sections): 65 536 channels
// are only 1 024 warps for 592 SM sub-partitions (2 on some, 1 on others -> the FP32 pipe idles 43 % of the time,
// profiles/r01_run3_ncu.md), and a 4 096-channel bank (BASELINE configs[1]) is 64 warps.  Here a CTA owns CH channels
// and its W warps own consecutive groups of SPW sections: tiles [TB frames][CH channels] arrive by TMA into a shared
// memory ring, warp 0 filters a tile IN PLACE with its sections and hands it to warp 1 through an mbarrier, ...,
// the last warp hands it to the TMA store engine and refills the stage it freed.  Work units are 2-8x finer, the
// per-thread state is a handful of registers, and per-section operation order is untouched, so results are
// bit-identical to the one-thread-per-channel kernels (and to the oracle) in both modes.
//
//   bar[stage][k]   "the tile in `stage` is ready for warp k": k = 0 completes on TMA bytes (complete_tx),
//                   k > 0 on one arrive from warp k-1.  One phase per ring revolution.
#pragma once
#include "common.cuh"
#include "tmap.h"

namespace olb {

// Packed (two channels per lane, FFMA2) and scalar lane types with identical op order.
template <int PACK>
struct Lane;
template <>
struct Lane<2> {
    typedef u64 T;
    template <int MODE>
    static __device__ __forceinline__ T gain(T g, T x) { return fmul2(g, x); }
    template <int MODE>
    static __device__ __forceinline__ T section(T x, T x1, T x2, T y1, T y2, const T (&k)[5], T one, T negzero) {
        if (MODE == OLB_MODE_FAST) {
            T t = ffma2(k[1], x1, x);
            t = ffma2(k[2], x2, t);
            t = ffma2(k[4], y2, t);
            return ffma2(k[3], y1, t);
        } else if (MODE == OLB_MODE_FMA) {
            T p = fmul2(k[1], x1);
            p = ffma2(k[2], x2, p);
            p = ffma2(k[4], y2, p);
            p = ffma2(k[3], y1, p);
            return ffma2(k[0], x, p);
        } else {
            T p = ffma2(ffma2(k[1], x1, negzero), one, ffma2(k[2], x2, negzero));
            p = ffma2(p, one, ffma2(k[4], y2, negzero));
            p = ffma2(p, one, ffma2(k[3], y1, negzero));
            return ffma2(ffma2(k[0], x, negzero), one, p);
        }
    }
};
template <>
struct Lane<1> {
    typedef float T;
    template <int MODE>
    static __device__ __forceinline__ T gain(T g, T x) { return __fmul_rn(g, x); }
    template <int MODE>
    static __device__ __forceinline__ T section(T x, T x1, T x2, T y1, T y2, const T (&k)[5], u64, u64) {
        if (MODE == OLB_MODE_FAST) {
            T t = __fmaf_rn(k[1], x1, x);
            t = __fmaf_rn(k[2], x2, t);
            t = __fmaf_rn(k[4], y2, t);
            return __fmaf_rn(k[3], y1, t);
        } else if (MODE == OLB_MODE_FMA) {
            T p = __fmul_rn(k[1], x1);
            p = __fmaf_rn(k[2], x2, p);
            p = __fmaf_rn(k[4], y2, p);
            p = __fmaf_rn(k[3], y1, p);
            return __fmaf_rn(k[0], x, p);
        } else {
            T p = __fadd_rn(__fmul_rn(k[1], x1), __fmul_rn(k[2], x2));
            p = __fadd_rn(p, __fmul_rn(k[4], y2));
            p = __fadd_rn(p, __fmul_rn(k[3], y1));
            return __fadd_rn(__fmul_rn(k[0], x), p);
        }
    }
};

template <int SPW, int MODE, int PACK>
__device__ __forceinline__ typename Lane<PACK>::T sys_step(typename Lane<PACK>::T v, typename Lane<PACK>::T (&h)[SPW + 1][2],
                                                          const typename Lane<PACK>::T (&k)[SPW][5], u64 one, u64 negzero, bool first) {
    typedef typename Lane<PACK>::T T;
    if (MODE == OLB_MODE_FAST && first) v = Lane<PACK>::template gain<MODE>(k[0][0], v);  // warp 0 applies G = prod(b0)
#pragma unroll
    for (int s = 0; s < SPW; s++) {
        const T y = Lane<PACK>::template section<MODE>(v, h[s][0], h[s][1], h[s + 1][0], h[s + 1][1], k[s], one, negzero);
        h[s][1] = h[s][0];
        h[s][0] = v;
        v = y;
    }
    h[SPW][1] = h[SPW][0];
    h[SPW][0] = v;
    return v;
}

template <int S, int SPW, int MODE, int TB, int NST, int PACK>
__global__ void __launch_bounds__((S / SPW) * 32) k_bq_sys(const __grid_constant__ TensorMap tm_in, const __grid_constant__ TensorMap tm_out,
                                                           uint32_t frames, uint32_t C, const float *__restrict__ coef, float *hist,
                                                           u64 one, u64 negzero) {
    typedef typename Lane<PACK>::T T;
    constexpr int W = S / SPW, CH = 32 * PACK, R = 8;
    static_assert(S % SPW == 0 && TB % R == 0, "bad shape");
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float *tiles = reinterpret_cast<float *>(smem_raw);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem_raw + (size_t) NST * TB * CH * sizeof(float));  // [NST][W]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t ch0 = blockIdx.x * CH;
    constexpr uint32_t kTileBytes = TB * CH * sizeof(float);
    const uint32_t ntiles = (frames + TB - 1) / TB;

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

    // this warp's sections [s0, s0+SPW) and history nodes [s0, s0+SPW]
    const int s0 = warp * SPW;
    const uint32_t c = ch0 + PACK * lane;
    const bool live = c < C;  // C % 4 == 0 on this path: a live packed lane owns two live channels
    T k[SPW][5], h[SPW + 1][2];
#pragma unroll
    for (int s = 0; s < SPW; s++)
#pragma unroll
        for (int j = 0; j < 5; j++) k[s][j] = live ? *reinterpret_cast<const T *>(coef + (size_t) ((s0 + s) * 5 + j) * C + c) : T(0);
#pragma unroll
    for (int n = 0; n <= SPW; n++) {
        h[n][0] = live ? *reinterpret_cast<const T *>(hist + (size_t) ((s0 + n) * 2 + 0) * C + c) : T(0);
        h[n][1] = live ? *reinterpret_cast<const T *>(hist + (size_t) ((s0 + n) * 2 + 1) * C + c) : T(0);
    }
    __syncthreads();  // every warp has read its nodes before any warp can finish and write a shared one back

    for (uint32_t i = 0; i < ntiles; i++) {
        const uint32_t st = i % NST, par = (i / NST) & 1;
        mbar_wait(&bar[st * W + warp], par);
        T *col = reinterpret_cast<T *>(tiles + (size_t) st * TB * CH) + lane;
        const uint32_t rows = min((uint32_t) TB, frames - i * TB);
        if (rows == TB) {
#pragma unroll
            for (int r0 = 0; r0 < TB; r0 += R) {
                T x[R];
#pragma unroll
                for (int j = 0; j < R; j++) x[j] = col[(r0 + j) * 32];
#pragma unroll
                for (int j = 0; j < R; j++) col[(r0 + j) * 32] = sys_step<SPW, MODE, PACK>(x[j], h, k, one, negzero, warp == 0);
            }
        } else {
            for (uint32_t r = 0; r < rows; r++) col[r * 32] = sys_step<SPW, MODE, PACK>(col[r * 32], h, k, one, negzero, warp == 0);
        }
        if (warp < W - 1) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[st * W + warp + 1]);
        } else {
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&tm_out, (int) ch0, (int) (i * TB), tiles + (size_t) st * TB * CH);
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
        // node s0 belongs to the previous warp (its output history) except for warp 0, which owns the input node
#pragma unroll
        for (int n = 0; n <= SPW; n++) {
            if (n == 0 && warp != 0) continue;
            *reinterpret_cast<T *>(hist + (size_t) ((s0 + n) * 2 + 0) * C + c) = h[n][0];
            *reinterpret_cast<T *>(hist + (size_t) ((s0 + n) * 2 + 1) * C + c) = h[n][1];
        }
    }
}

struct SysCfg {
    int spw, tb, nst, pack;
};

template <int S, int SPW, int MODE, int TB, int NST, int PACK>
static int launch_sys_cfg(uint32_t C, const float *d_coef, float *d_hist, const float *in, float *out, uint32_t frames, cudaStream_t st) {
    constexpr int W = S / SPW, CH = 32 * PACK;
    constexpr size_t smem = (size_t) NST * TB * CH * sizeof(float) + (size_t) NST * W * sizeof(uint64_t);
    TensorMap tin, tout;
    int rc = make_tmap_frames(&tin, in, frames, C, CH, TB);
    if (rc) return rc;
    rc = make_tmap_frames(&tout, out, frames, C, CH, TB);
    if (rc) return rc;
    auto kern = k_bq_sys<S, SPW, MODE, TB, NST, PACK>;
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    OLB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    const float onef = 1.0f, nzf = -0.0f;
    uint32_t ob, nb;
    memcpy(&ob, &onef, 4);
    memcpy(&nb, &nzf, 4);
    const u64 one = ((u64) ob << 32) | ob, negzero = ((u64) nb << 32) | nb;
    kern<<<(C + CH - 1) / CH, W * 32, smem, st>>>(tin, tout, frames, C, d_coef, d_hist, one, negzero);
    OLB_LAUNCHED();
    return OLB_OK;
}

}  // namespace olb
