// sampler_bank.cu -- sample playback bank: `instances` x ol::synth::Sample behind
// SampleSoundSource (modules/synthlib/Sample.cpp:9-61, SampleSoundSource.h:14-41) over a device-resident sample pool.
//
// The reference asks its SampleDataSource for ONE frame per Process call (Sample.cpp:11; the only implementation decodes a
// file through miniaudio, workouts/workout_buddy.h:18-42).  Here every sample lives in HBM, an instance is one lane, and a
// frame of the bank is one coalesced row `[instances * channels]`.  The state machine is the reference's:
//     if playing: read frame `cur` when it exists and advance; Loop mode: seek to loop_start when nothing was read or
//     when loop_end is set and the new position is past it (Sample.cpp:10-22); a frame that is not played is left untouched
// All of it is integer work and copies -> bit-exact.  Time is cut into tiles of kSpTile frames: a lane whose whole tile is a
// straight run (no end of sample, no loop point inside) issues kSpTile independent loads and kSpTile stores that are
// coalesced across the warp; the frame-by-frame state machine runs only in tiles where some lane of the warp is near an
// edge.  Algorithmic traffic: 4 B read + 4 B written per played sample.
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace olb {

constexpr int kSpTile = 16;
constexpr int kSpBlockChunks = 256;  // chunks (of kSpChunk = 256 frames) per scan / play launch pair: 65 536 frames
constexpr uint32_t kSpPlaying = 1u, kSpLoop = 2u, kSpSeek = 4u;

template <int CH>
struct SpVec;
template <>
struct SpVec<1> {
    typedef float T;
};
template <>
struct SpVec<2> {
    typedef float2 T;
};

// One frame of Sample::Process on the position alone (Sample.cpp:9-23): returns whether a frame is read at `c`.
__device__ __forceinline__ bool sp_step(uint32_t &c, uint32_t n, bool loop, uint32_t ls, uint32_t le) {
    const bool read = c < n;
    if (read) c++;
    if (loop && (!read || (le && c > le))) c = ls;
    return read;
}

// k_sampler_scan: where every instance stands at the start of every chunk of kSpChunk frames, so that the copy itself can
// run on the whole (instances x chunks) grid -- a position is a closed form of the elapsed frames:
//   * frames at positions below min(n, loop_end) are plain (position + 1), a run of them is one addition;
//   * any other frame of a looping instance ends at loop_start, and from loop_start the motion is periodic: the frames
//     loop_start .. loop_end when the loop end lies inside the sample, else loop_start .. n - 1 plus the one frame in which
//     Read returns nothing (Sample.cpp:13-19); whole periods are skipped with one modulo;
//   * a one-shot past its end stays where it is.
// A chunk costs a handful of iterations whatever the loop length.  Also writes the position after the block.
constexpr int kSpChunk = 256;
__global__ void __launch_bounds__(128) k_sampler_scan(uint32_t frames, uint32_t I, const uint32_t *__restrict__ len,
                                                      const uint32_t *__restrict__ loop_start, const uint32_t *__restrict__ loop_end,
                                                      const uint32_t *__restrict__ flags, const uint32_t *__restrict__ seek, uint32_t *cur,
                                                      uint32_t *chunk_start /* [chunks][I] */, int first_block) {
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    if (inst >= I) return;
    const uint32_t fl = flags[inst], n = len[inst], ls = loop_start[inst], le = loop_end[inst];
    const bool playing = fl & kSpPlaying, loop = fl & kSpLoop;
    uint32_t c = (first_block && (fl & kSpSeek)) ? seek[inst] : cur[inst];
    const uint32_t plain_end = (loop && le != 0 && le < n) ? le : n;  // frames at positions below this just advance
    uint32_t period = 1;                                              // of the motion that starts at loop_start
    if (loop && ls < n && !(le != 0 && le < ls)) period = (le != 0 && le < n) ? le - ls + 1 : n - ls + 1;
    const uint32_t chunks = (frames + kSpChunk - 1) / kSpChunk;
    for (uint32_t k = 0; k < chunks; k++) {
        chunk_start[(size_t) k * I + inst] = c;
        if (!playing) continue;
        uint32_t left = min((uint32_t) kSpChunk, frames - k * kSpChunk);
        while (left) {
            if (c < plain_end) {
                const uint32_t run = min(left, plain_end - c);
                c += run;
                left -= run;
            } else if (!loop) {
                left = 0;  // a one-shot past its end: Read returns nothing, nothing moves
            } else {
                sp_step(c, n, loop, ls, le);  // ends at loop_start
                left--;
                left %= period;
            }
        }
    }
    cur[inst] = c;
}

// k_sampler_play: one block = 128 instances x one chunk.  Tiles of kSpTile frames: a lane whose whole tile is plain issues
// kSpTile independent loads, then kSpTile stores that are coalesced across the warp; the frame-by-frame state machine runs
// only in tiles where some lane of the warp meets an edge.
template <int CH>
__global__ void __launch_bounds__(128) k_sampler_play(float *out, uint32_t frames, uint32_t I, const float *const *__restrict__ base,
                                                      const uint32_t *__restrict__ len, const uint32_t *__restrict__ loop_start,
                                                      const uint32_t *__restrict__ loop_end, const uint32_t *__restrict__ flags,
                                                      const uint32_t *__restrict__ chunk_start) {
    typedef typename SpVec<CH>::T V;
    const uint32_t inst = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = inst < I;
    const uint32_t ci = live ? inst : 0;
    const uint32_t fl = flags[ci], n = len[ci], ls = loop_start[ci], le = loop_end[ci];
    const bool playing = live && (fl & kSpPlaying), loop = fl & kSpLoop;
    if (!__any_sync(0xffffffffu, playing)) return;
    const V *src = reinterpret_cast<const V *>(base[ci]);
    V *dst = reinterpret_cast<V *>(out) + inst;
    const uint32_t chunks = (frames + kSpChunk - 1) / kSpChunk;
    for (uint32_t k = blockIdx.y; k < chunks; k += gridDim.y) {
        uint32_t c = chunk_start[(size_t) k * I + ci];
        const uint32_t f_end = min(frames, (k + 1) * (uint32_t) kSpChunk);
        for (uint32_t f0 = k * kSpChunk; f0 < f_end; f0 += kSpTile) {
            const uint32_t rows = min((uint32_t) kSpTile, f_end - f0);
            const uint64_t stop = (uint64_t) c + kSpTile;
            const bool straight = playing && rows == (uint32_t) kSpTile && stop <= n && (!loop || le == 0 || stop <= le);
            if (__all_sync(0xffffffffu, straight || !playing)) {
                if (straight) {
                    // the lane's kSpTile frames are one contiguous run of its sample: 16-byte loads from the aligned address
                    // below it (samples are 256-byte aligned and padded by 16 bytes), then the 0..3-float shift undone with
                    // selects -- a third of the L1 requests of frame-by-frame loads, which is what bounds this kernel
                    constexpr int NF = kSpTile * CH, NV = NF / 4 + 1;
                    const uint64_t p = (uint64_t) c * CH;
                    const uint32_t sh = (uint32_t) p & 3u;
                    const float4 *q = reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(src) + (p - sh));
                    float w[4 * NV];
#pragma unroll
                    for (int j = 0; j < NV; j++) {
                        const float4 t = __ldg(q + j);
                        w[4 * j] = t.x; w[4 * j + 1] = t.y; w[4 * j + 2] = t.z; w[4 * j + 3] = t.w;
                    }
                    float e[NF];
#pragma unroll
                    for (int i = 0; i < NF; i++) {
                        const float t0 = (CH == 1 && (sh & 1u)) ? w[i + 1] : w[i];
                        const float t1 = (CH == 1 && (sh & 1u)) ? w[i + 3] : w[i + 2];
                        e[i] = (sh & 2u) ? t1 : t0;
                    }
#pragma unroll
                    for (int r = 0; r < kSpTile; r++) {
                        if (CH == 1) reinterpret_cast<float *>(dst)[(size_t) (f0 + r) * I] = e[r];
                        else reinterpret_cast<float2 *>(dst)[(size_t) (f0 + r) * I] = make_float2(e[2 * r], e[2 * r + 1]);
                    }
                    c += kSpTile;
                }
            } else if (playing) {
                for (uint32_t r = 0; r < rows; r++) {  // Sample::Process, one frame (Sample.cpp:9-23)
                    const uint32_t at = c;
                    if (sp_step(c, n, loop, ls, le)) dst[(size_t) (f0 + r) * I] = __ldg(src + at);
                }
            }
        }
    }
}

}  // namespace olb

using namespace olb;

struct olb_sampler {
    std::recursive_mutex mu;
    StreamOrder order;
    int device = 0;
    uint32_t I = 0, channels = 1;
    float sample_rate = 0.f;
    struct Pooled {
        float *d = nullptr;
        uint64_t frames = 0;
    };
    std::vector<Pooled> pool;
    // host shadow of everything the host owns; `cur` is owned by the device between seeks
    std::vector<uint32_t> sample, len, ls, le, flags, seek;
    std::vector<const float *> base;
    const float **d_base = nullptr;
    uint32_t *d_u32 = nullptr;  // [6][I]: len, loop_start, loop_end, flags, seek, cur
    bool dirty = true;
    HostStage host;
    uint32_t *d_chunk = nullptr;  // [kSpBlockChunks][I] positions at the chunk starts of the block in flight
    uint32_t *dev(int k) { return d_u32 + (size_t) k * I; }
};

static int sampler_range(olb_sampler *h, uint32_t first, uint32_t count, const char *who) {
    OLB_ARG(h, "%s: NULL handle", who);
    OLB_ARG((uint64_t) first + count <= h->I, "%s: range %u+%u exceeds %u instances", who, first, count, h->I);
    return OLB_OK;
}

// positions are 32-bit on the device; anything at or past 2^32 - 64 is "past the end" of every sample the pool accepts
static uint32_t sampler_clamp(uint64_t frame_index) { return frame_index > 0xffffffbfull ? 0xffffffbfu : (uint32_t) frame_index; }

static int sampler_upload(olb_sampler *h, cudaStream_t st) {
    if (!h->dirty) return OLB_OK;
    const size_t I = h->I;
    OLB_CUDA(cudaMemcpyAsync(h->d_base, h->base.data(), I * sizeof(float *), cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaMemcpyAsync(h->dev(0), h->len.data(), I * 4, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaMemcpyAsync(h->dev(1), h->ls.data(), I * 4, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaMemcpyAsync(h->dev(2), h->le.data(), I * 4, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaMemcpyAsync(h->dev(3), h->flags.data(), I * 4, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaMemcpyAsync(h->dev(4), h->seek.data(), I * 4, cudaMemcpyHostToDevice, st));
    OLB_CUDA(cudaStreamSynchronize(st));
    h->dirty = false;
    return OLB_OK;
}

OLB_API int olb_sampler_create(int device, uint32_t instances, uint32_t channels, olb_sampler **out) {
    OLB_ARG(out, "olb_sampler_create: out is NULL");
    *out = nullptr;
    OLB_ARG(instances > 0, "olb_sampler_create: instances must be > 0");
    OLB_ARG(channels == 1 || channels == 2, "olb_sampler_create: channels must be 1 or 2, got %u", channels);
    int rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_sampler *h = new olb_sampler();
    h->device = device;
    h->I = instances;
    h->channels = channels;
    for (auto *v : {&h->sample, &h->len, &h->ls, &h->le, &h->flags, &h->seek}) v->assign(instances, 0u);
    h->base.assign(instances, nullptr);
    cudaError_t e = cudaMalloc(&h->d_base, (size_t) instances * sizeof(float *));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_u32, (size_t) 6 * instances * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(h->d_u32, 0, (size_t) 6 * instances * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&h->d_chunk, (size_t) kSpBlockChunks * instances * sizeof(uint32_t));
    if (e != cudaSuccess) {
        cudaFree(h->d_base); cudaFree(h->d_u32); cudaFree(h->d_chunk);
        delete h;
        return fail_cuda(e, "allocating sampler bank", __FILE__, __LINE__);
    }
    *out = h;
    return OLB_OK;
}

OLB_API int olb_sampler_destroy(olb_sampler *h) {
    if (!h) return OLB_OK;
    DeviceGuard g(h->device);
    for (auto &p : h->pool) cudaFree(p.d);
    cudaFree(h->d_base); cudaFree(h->d_u32); cudaFree(h->d_chunk);
    h->host.release();
    delete h;
    return OLB_OK;
}

OLB_API int olb_sampler_load_sample(olb_sampler *h, const float *frames_host, uint64_t frames, uint32_t *sample_id) {
    OLB_ARG(h && sample_id, "olb_sampler_load_sample: NULL argument");
    OLB_LOCK(h);
    OLB_ARG(frames_host || frames == 0, "olb_sampler_load_sample: NULL data");
    OLB_ARG(frames < 0xffffff00ull, "olb_sampler_load_sample: %llu frames exceed the 32-bit positions of the bank", (unsigned long long) frames);
    DeviceGuard g(h->device);
    olb_sampler::Pooled p;
    p.frames = frames;
    const size_t bytes = (size_t) frames * h->channels * sizeof(float);
    OLB_CUDA(cudaMalloc(&p.d, bytes + 16));  // 16 bytes of slack: the tile loads of k_sampler_play are 16-byte wide
    cudaError_t e = bytes ? cudaMemcpy(p.d, frames_host, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
    if (e != cudaSuccess) {
        cudaFree(p.d);
        return fail_cuda(e, "copying a sample into the pool", __FILE__, __LINE__);
    }
    h->pool.push_back(p);
    *sample_id = (uint32_t) h->pool.size() - 1;
    return OLB_OK;
}

OLB_API int olb_sampler_set_sample(olb_sampler *h, uint32_t first, uint32_t count, const uint32_t *ids_host, int per_instance) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    int rc = sampler_range(h, first, count, "olb_sampler_set_sample");
    if (rc) return rc;
    OLB_ARG(ids_host, "olb_sampler_set_sample: NULL ids");
    for (uint32_t i = 0; i < count; i++) {
        const uint32_t id = ids_host[per_instance ? i : 0];
        OLB_ARG(id < h->pool.size(), "olb_sampler_set_sample: sample %u is not in the pool (%zu loaded)", id, h->pool.size());
    }
    for (uint32_t i = 0; i < count; i++) {
        const uint32_t id = ids_host[per_instance ? i : 0], k = first + i;
        h->sample[k] = id;
        h->base[k] = h->pool[id].d;
        h->len[k] = (uint32_t) h->pool[id].frames;
        h->ls[k] = h->le[k] = 0;     // a fresh Sample object (Sample.h:41-50)
        h->flags[k] = kSpSeek;       // not playing, OneShot, position 0
        h->seek[k] = 0;
    }
    h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_sampler_init(olb_sampler *h, float sample_rate) {
    OLB_ARG(h, "olb_sampler_init: NULL handle");
    OLB_LOCK(h);
    h->sample_rate = sample_rate;  // Sample::Init stores the rate and nothing else (Sample.cpp:30-33)
    return OLB_OK;
}

#define OLB_SAMPLER_EACH(NAME, BODY)                                  \
    int rc = sampler_range(h, first, count, NAME);                    \
    if (rc) return rc;                                                \
    for (uint32_t k = first; k < first + count; k++) { BODY; }        \
    h->dirty = true;                                                  \
    return OLB_OK;

OLB_API int olb_sampler_seek(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_seek", h->seek[k] = sampler_clamp(frame_index); h->flags[k] |= kSpSeek)
}
OLB_API int olb_sampler_set_loop_start(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_set_loop_start", h->ls[k] = sampler_clamp(frame_index))
}
OLB_API int olb_sampler_set_loop_end(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_set_loop_end", h->le[k] = sampler_clamp(frame_index))
}
OLB_API int olb_sampler_set_play_mode(olb_sampler *h, uint32_t first, uint32_t count, int mode) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_ARG(mode == OLB_SAMPLE_ONESHOT || mode == OLB_SAMPLE_LOOP, "olb_sampler_set_play_mode: unknown mode %d", mode);
    OLB_SAMPLER_EACH("olb_sampler_set_play_mode", h->flags[k] = mode == OLB_SAMPLE_LOOP ? h->flags[k] | kSpLoop : h->flags[k] & ~kSpLoop)
}
OLB_API int olb_sampler_play(olb_sampler *h, uint32_t first, uint32_t count) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_play", h->flags[k] |= kSpPlaying)
}
OLB_API int olb_sampler_pause(olb_sampler *h, uint32_t first, uint32_t count) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_pause", h->flags[k] &= ~kSpPlaying)
}
OLB_API int olb_sampler_toggle_play(olb_sampler *h, uint32_t first, uint32_t count) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    OLB_SAMPLER_EACH("olb_sampler_toggle_play", h->flags[k] ^= kSpPlaying)
}
OLB_API int olb_sampler_gate(olb_sampler *h, uint32_t first, uint32_t count, int on) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    if (on) {  // SampleSoundSource::GateOn: Seek(0); Play()
        OLB_SAMPLER_EACH("olb_sampler_gate", h->seek[k] = 0; h->flags[k] |= kSpSeek | kSpPlaying)
    }
    OLB_SAMPLER_EACH("olb_sampler_gate", h->flags[k] &= ~kSpPlaying)
}

OLB_API int olb_sampler_position(olb_sampler *h, uint32_t first, uint32_t count, uint64_t *frame_index_host, uint8_t *playing_host) {
    if (!h) { olb::set_error("NULL handle"); return OLB_E_ARG; }
    OLB_LOCK(h);
    int rc = sampler_range(h, first, count, "olb_sampler_position");
    if (rc) return rc;
    DeviceGuard g(h->device);
    if (frame_index_host) {
        std::vector<uint32_t> cur(count);
        OLB_CUDA(h->order.sync());  // everything queued on the bank so far has advanced the positions
        OLB_CUDA(cudaMemcpy(cur.data(), h->dev(5) + first, (size_t) count * 4, cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < count; i++) frame_index_host[i] = (h->flags[first + i] & kSpSeek) ? h->seek[first + i] : cur[i];
    }
    if (playing_host)
        for (uint32_t i = 0; i < count; i++) playing_host[i] = (h->flags[first + i] & kSpPlaying) ? 1 : 0;
    return OLB_OK;
}

OLB_API int olb_sampler_process(olb_sampler *h, float *out_dev, uint64_t frames, void *stream) {
    OLB_ARG(h, "olb_sampler_process: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(out_dev, "olb_sampler_process: NULL buffer");
    OLB_ARG(frames <= 0x7fffffffull, "olb_sampler_process: at most 2^31 - 1 frames per call");
    OLB_ARG(((uintptr_t) out_dev & (h->channels * sizeof(float) - 1)) == 0, "olb_sampler_process: out must be %zu-byte aligned",
            h->channels * sizeof(float));
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t) stream;
    OLB_CUDA(h->order.enter(st));
    int rc = sampler_upload(h, st);
    if (rc) return rc;
    const uint32_t I = h->I, blocks = (I + 127) / 128;
    // blocks of at most kSpBlockChunks chunks: positions at the chunk starts (k_sampler_scan), then the copy on the whole
    // (instances x chunks) grid (k_sampler_play); both on the caller's stream, so consecutive blocks reuse the scratch in order
    const uint64_t block_frames = (uint64_t) kSpBlockChunks * kSpChunk, W = (uint64_t) I * h->channels;
    for (uint64_t t0 = 0; t0 < frames; t0 += block_frames) {
        const uint32_t nf = (uint32_t) (frames - t0 < block_frames ? frames - t0 : block_frames), chunks = (nf + kSpChunk - 1) / kSpChunk;
        k_sampler_scan<<<blocks, 128, 0, st>>>(nf, I, h->dev(0), h->dev(1), h->dev(2), h->dev(3), h->dev(4), h->dev(5), h->d_chunk, t0 == 0);
        OLB_LAUNCHED();
        const dim3 grid(blocks, chunks);
        if (h->channels == 1)
            k_sampler_play<1><<<grid, 128, 0, st>>>(out_dev + t0 * W, nf, I, h->d_base, h->dev(0), h->dev(1), h->dev(2), h->dev(3), h->d_chunk);
        else
            k_sampler_play<2><<<grid, 128, 0, st>>>(out_dev + t0 * W, nf, I, h->d_base, h->dev(0), h->dev(1), h->dev(2), h->dev(3), h->d_chunk);
        OLB_LAUNCHED();
    }
    // the launch consumed the pending seeks: the device owns the positions again
    bool any = false;
    for (auto &f : h->flags)
        if (f & kSpSeek) {
            f &= ~kSpSeek;
            any = true;
        }
    if (any) h->dirty = true;
    return OLB_OK;
}

OLB_API int olb_sampler_process_host(olb_sampler *h, float *out_host, uint64_t frames) {
    OLB_ARG(h, "olb_sampler_process_host: NULL handle");
    OLB_LOCK(h);
    if (frames == 0) return OLB_OK;
    OLB_ARG(out_host, "olb_sampler_process_host: NULL buffer");
    DeviceGuard g(h->device);
    const uint64_t W = (uint64_t) h->I * h->channels, chunk = host_chunk_frames(W, frames);
    int rc = h->host.ensure(0, chunk * W);
    if (rc) return rc;
    for (uint64_t t0 = 0; t0 < frames; t0 += chunk) {
        const uint64_t n = frames - t0 < chunk ? frames - t0 : chunk;
        // played INTO: frames the bank does not touch keep what the caller put there
        OLB_CUDA(cudaMemcpyAsync(h->host.d[0], out_host + t0 * W, n * W * sizeof(float), cudaMemcpyHostToDevice, h->host.st));
        rc = olb_sampler_process(h, h->host.d[0], n, h->host.st);
        if (rc) return rc;
        OLB_CUDA(cudaMemcpyAsync(out_host + t0 * W, h->host.d[0], n * W * sizeof(float), cudaMemcpyDeviceToHost, h->host.st));
    }
    OLB_CUDA(cudaStreamSynchronize(h->host.st));
    return OLB_OK;
}
