// End-to-end tool (BASELINE configs[0] on the GPU): a WAV file through the fx rack.
//   usage: wav_rack <in.wav> <out.wav> <stub|plate|decode> [frames per block]      (decode: no processing, no GPU)
// Every channel of the file is one mono rack instance (the reference's FxRack<N> processes channel 0 only, Fx.h:426-434;
// a bank processes them all).  Call order of test/fx_test.cpp / workouts: Init(sample rate of the file), defaults, Process.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ol_gpu/FxBank.hpp"
#include "ol_gpu/Wav.hpp"

template <class Rack>
static int run(const ol::gpu::WavData &w, std::vector<float> &out, uint64_t block) {
    Rack rack(w.channels);
    if (rack.Init((float) w.sample_rate) != ol::gpu::Ok) {
        fprintf(stderr, "Init failed: %s\n", rack.error().c_str());
        return 4;
    }
    for (uint64_t t = 0; t < w.frames; t += block) {  // an audio callback's blocks
        const uint64_t n = w.frames - t < block ? w.frames - t : block;
        rack.ProcessBlock(&w.samples[t * w.channels], &out[t * w.channels], n);
    }
    if (rack.status() != OLB_OK) {
        fprintf(stderr, "process failed: %s\n", rack.error().c_str());
        return 5;
    }
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s in.wav out.wav stub|plate [block]\n", argv[0]);
        return 2;
    }
    ol::gpu::WavData w;
    std::string err;
    if (!ol::gpu::ReadWav(argv[1], w, &err)) {
        fprintf(stderr, "%s: %s\n", argv[1], err.c_str());
        return 3;
    }
    const uint64_t block = argc > 4 ? (uint64_t) atoll(argv[4]) : 4096;
    std::vector<float> out(w.samples.size());
    if (!strcmp(argv[3], "decode")) {
        out = w.samples;
        return ol::gpu::WriteWavF32(argv[2], out.data(), w.frames, w.channels, w.sample_rate) ? 0 : 6;
    }
    const int rc = !strcmp(argv[3], "plate") ? run<ol::gpu::FxRackPlateBank>(w, out, block) : run<ol::gpu::FxRackBank>(w, out, block);
    if (rc) return rc;
    if (!ol::gpu::WriteWavF32(argv[2], out.data(), w.frames, w.channels, w.sample_rate)) return 6;
    printf("%u channels, %llu frames, %u Hz\n", w.channels, (unsigned long long) w.frames, w.sample_rate);
    return 0;
}
