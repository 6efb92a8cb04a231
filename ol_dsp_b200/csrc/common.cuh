// common.cuh -- shared host/device helpers of libolbank.so (sm_100a only; no CPU fallback).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "../../include/olbank.h"

#define OLB_API extern "C" __attribute__((visibility("default")))

namespace olb {

// ---- error plumbing: every export returns an int; the text is kept per thread -----------------
void set_error(const char *fmt, ...);
int fail_cuda(cudaError_t e, const char *what, const char *file, int line);
extern std::atomic<uint64_t> g_launches;

#define OLB_CUDA(call)                                                          \
    do {                                                                        \
        cudaError_t _e = (call);                                                \
        if (_e != cudaSuccess) return olb::fail_cuda(_e, #call, __FILE__, __LINE__); \
    } while (0)

#define OLB_ARG(cond, ...)              \
    do {                                \
        if (!(cond)) {                  \
            olb::set_error(__VA_ARGS__); \
            return OLB_E_ARG;           \
        }                               \
    } while (0)

// Counts a kernel launch and surfaces launch-time errors.
#define OLB_LAUNCHED()                                \
    do {                                              \
        olb::g_launches.fetch_add(1);                 \
        OLB_CUDA(cudaGetLastError());                 \
    } while (0)

// RAII device switch: every export may be called with any current device.
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

int require_sm100(int device);

// Every entry point of a handle takes its mutex: setters and Init may come from a control thread while the audio thread
// is inside process (the reference's hosts do exactly that, unsynchronised: workouts/workout_buddy.cpp:35-72 vs :76-100).
// Recursive, because the composite calls (process_host, process_events, the plate rack) go through the public ones.
#define OLB_LOCK(h) std::lock_guard<std::recursive_mutex> olb_lock_((h)->mu)

// ---- stream ordering of a handle's device work ------------------------------------------------------
// A bank's state lives on the device between process calls, and callers may hand every call a different stream (and
// non-blocking streams are not ordered against the legacy default stream).  enter(st) makes `st` wait for whatever the
// handle queued on the previous stream; sync() waits for everything queued so far (state read-back, destroy).
struct StreamOrder {
    cudaStream_t last = nullptr;
    bool any = false;
    cudaEvent_t ev = nullptr;
    cudaError_t enter(cudaStream_t st) {
        cudaError_t e = cudaSuccess;
        if (any && st != last) {
            if (!ev) e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventRecord(ev, last);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(st, ev, 0);
        }
        last = st;
        any = true;
        return e;
    }
    cudaStream_t stream() const { return last; }
    // control-path clear (Init): queued behind the work of the last process call, or at once when nothing was ever queued
    cudaError_t memset0(void *p, size_t bytes) const { return any ? cudaMemsetAsync(p, 0, bytes, last) : cudaMemset(p, 0, bytes); }
    cudaError_t sync() const { return any ? cudaStreamSynchronize(last) : cudaSuccess; }
    void quiesce() const {
        if (any) cudaStreamSynchronize(last);
    }
    void release() {
        if (ev) cudaEventDestroy(ev);
        ev = nullptr;
    }
};

// ---- host-buffer convenience path of the small banks (fx, plate, chorus, voices) -----------------
// Device staging that grows on demand plus one private stream: `process_host` entry points copy a time chunk in, run the
// bank's own device path on it, copy the result out, and return when the host buffer is complete (the reference's
// Process() is synchronous).  The biquad bank has its own three-stream overlapped version (biquad_bank.cu).
struct HostStage {
    float *d[3] = {nullptr, nullptr, nullptr};
    size_t cap[3] = {0, 0, 0};
    cudaStream_t st = nullptr;
    int ensure(int k, size_t elems) {
        if (!st) OLB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        if (cap[k] < elems) {
            cudaFree(d[k]);
            d[k] = nullptr;
            cap[k] = 0;
            OLB_CUDA(cudaMalloc(&d[k], elems * sizeof(float)));
            cap[k] = elems;
        }
        return OLB_OK;
    }
    void release() {
        for (int k = 0; k < 3; k++) cudaFree(d[k]);
        if (st) cudaStreamDestroy(st);
    }
};
// frames per staged chunk: ~32 MiB per buffer, whole frames, a multiple of 64 (the TMA tile height)
static inline uint64_t host_chunk_frames(uint64_t units, uint64_t frames) {
    uint64_t c = (32ull << 20) / (units * sizeof(float));
    c = c < 64 ? 64 : (c / 64) * 64;
    return c > frames ? frames : c;
}

// ---- packed fp32x2 arithmetic (sm_100 FFMA2).  One instruction, two IEEE-rounded lanes. ---------
typedef unsigned long long u64;

__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
// NOTE: ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false (probed),
// and folds fma(a, 1.0, b) / fma(a, b, -0.0) when the constants are visible.  A packed product that
// feeds only FMAs (never an add) is safe to write as mul; separately-rounded packed products and
// sums are therefore built from FFMA2 with OPAQUE constants (kernel parameters):
//     mul(a, b) = fma(a, b, NEGZERO)        add(a, b) = fma(a, ONE, b)
// both exact for every finite input (x*y + -0 == x*y including signed zeros).
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// ---- Philox4x32-10 white noise shared with the oracle (oracle/olb_oracle.cpp:65-95) -------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t seed, uint32_t channel, uint64_t block, uint32_t out[4]) {
    uint32_t c0 = (uint32_t) block, c1 = (uint32_t) (block >> 32), c2 = channel, c3 = 0u;
    uint32_t k0 = seed, k1 = 0x01D5B200u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = (uint64_t) 0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t) 0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t) p1;
        const uint32_t n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t) p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__host__ __device__ __forceinline__ float noise_from_u32(uint32_t u) {
    return (float) (u >> 8) * 1.1920928955078125e-07f - 1.0f;  // both steps exact in fp32
}

// ---- mbarrier / TMA (cp.async.bulk.tensor) wrappers ---------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 2-D tile, global -> shared, completion signalled on an mbarrier (SASS: UTMALDG)
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const void *tmap, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(smem_dst)),
        "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
// 2-D tile, shared -> global, bulk-group completion (SASS: UTMASTG)
__device__ __forceinline__ void tma_store_2d(const void *tmap, int c0, int c1, const void *smem_src) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap), "r"(c0), "r"(c1),
                 "r"(smem_u32(smem_src))
                 : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

}  // namespace olb
