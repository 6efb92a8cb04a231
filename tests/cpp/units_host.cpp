// C++ callers written the way a reference user writes them (Init -> setters -> Process per frame, then per block), against
// include/ol_gpu/{FxBank,DattorroBank,ChorusBank,VoiceBank}.hpp.  Built by tests/test_abi.py (link check, CPU) and run by
// tests/test_cpp_hosts_gpu.py on the GPU box, where the outputs are compared with the oracle.
//   usage: units_host <filter|delay|reverb|rack|rack2|delay2|plate|chorus|chorus2|poly|voicemap|sampler> <instances> <frames> <in.f32> <out.f32> [out2.f32]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ol_gpu/ChorusBank.hpp"
#include "ol_gpu/DattorroBank.hpp"
#include "ol_gpu/FxBank.hpp"
#include "ol_gpu/SamplerBank.hpp"
#include "ol_gpu/VoiceBank.hpp"

static bool slurp(const char *path, std::vector<float> &v) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    const size_t n = fread(v.data(), sizeof(float), v.size(), f);
    fclose(f);
    return n == v.size();
}
static void dump(const char *path, const std::vector<float> &v) {
    FILE *f = fopen(path, "wb");
    fwrite(v.data(), sizeof(float), v.size(), f);
    fclose(f);
}

// test/fx_test.cpp:25-55 shape: Init, a setter, per-frame Process in a loop -- then the rest of the block at once
template <class Bank>
static int run_fx(uint32_t I, uint32_t T, std::vector<float> &in, std::vector<float> &out, uint32_t width = 1, int mode = OLB_MODE_EXACT) {
    Bank bank(I);
    I *= width;  // floats per bank frame (stereo instances: 2 each)
    if (bank.Init(48000.f) != ol::gpu::Ok) {
        fprintf(stderr, "Init failed: %s\n", bank.error().c_str());
        return 4;
    }
    if (mode != OLB_MODE_EXACT) bank.SetMode(mode);  // the fused tier (not in the reference)
    bank.UpdateHardwareControl(35 /* CC_DELAY_TIME, cc_map.h */, 0.3f);
    bank.UpdateMidiControl(41 /* CC_FILTER_CUTOFF */, 90);
    uint32_t t = 0;
    for (; t < 3 && t < T; t++) bank.Process(&in[(size_t) t * I], &out[(size_t) t * I]);
    if (t < T) bank.ProcessBlock(&in[(size_t) t * I], &out[(size_t) t * I], T - t);
    if (bank.status() != OLB_OK) {
        fprintf(stderr, "process failed: %s\n", bank.error().c_str());
        return 5;
    }
    printf("%s\n", bank.LastKernel());
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s unit instances frames in.f32 out.f32 [out2.f32]\n", argv[0]);
        return 2;
    }
    const char *unit = argv[1];
    const uint32_t I = (uint32_t) atoi(argv[2]), T = (uint32_t) atoi(argv[3]);
    std::vector<float> in((size_t) I * T), out((size_t) I * T), out2;
    if (!strcmp(unit, "sampler")) {  // in.f32 is one sample of any length
        FILE *f = fopen(argv[4], "rb");
        if (!f) return 3;
        fseek(f, 0, SEEK_END);
        in.assign((size_t) ftell(f) / sizeof(float), 0.f);
        fclose(f);
    }
    if (strcmp(unit, "poly") && strcmp(unit, "voicemap") && !slurp(argv[4], in)) return 3;
    int rc = 0;
    if (!strcmp(unit, "filter")) rc = run_fx<ol::gpu::FilterFxBank>(I, T, in, out);
    else if (!strcmp(unit, "delay")) rc = run_fx<ol::gpu::DelayFxBank>(I, T, in, out);
    else if (!strcmp(unit, "reverb")) rc = run_fx<ol::gpu::ReverbFxBank>(I, T, in, out);
    else if (!strcmp(unit, "rack")) rc = run_fx<ol::gpu::FxRackBank>(I, T, in, out);
    else if (!strcmp(unit, "rack_fma")) rc = run_fx<ol::gpu::FxRackBank>(I, T, in, out, 1, OLB_MODE_FMA);
    else if (!strcmp(unit, "plate")) {
        out2.resize((size_t) I * T);
        ol::gpu::DattorroBank plates(I);
        if (plates.status() != OLB_OK) {
            fprintf(stderr, "create failed: %s\n", plates.error().c_str());
            return 4;
        }
        plates.SetDecay(0.8f);
        plates.SetPreDelay(0.05f);
        uint32_t t = 0;
        for (; t < 3 && t < T; t++) plates.Process(&in[(size_t) t * I], &out[(size_t) t * I], &out2[(size_t) t * I]);
        if (t < T) plates.ProcessBlock(&in[(size_t) t * I], &out[(size_t) t * I], &out2[(size_t) t * I], T - t);
        rc = plates.status() == OLB_OK ? 0 : 5;
        if (rc) fprintf(stderr, "process failed: %s\n", plates.error().c_str());
    } else if (!strcmp(unit, "rack2_fma")) {
        rc = run_fx<ol::gpu::FxRack2Bank>(I / 2, T, in, out, 2, OLB_MODE_FMA);
    } else if (!strcmp(unit, "rack2") || !strcmp(unit, "delay2")) {
        // the stereo banks: `instances` here counts FLOATS per frame (2 per stereo instance), as the buffers are sized by it
        rc = !strcmp(unit, "rack2") ? run_fx<ol::gpu::FxRack2Bank>(I / 2, T, in, out, 2) : run_fx<ol::gpu::DelayFx2Bank>(I / 2, T, in, out, 2);
    } else if (!strcmp(unit, "chorus2")) {
        ol::gpu::StereoChorusBank ch(I / 2);  // stereo-chorus.rnbopat: frames are [instances][2]
        if (ch.Init(48000.f) != ol::gpu::Ok) return 4;
        ch.SetRate(0.7f);
        ch.SetDepth(0.9f);
        ch.SetPitch(0.5f);
        uint32_t t = 0;
        for (; t < 3 && t < T; t++) ch.Process(&in[(size_t) t * I], &out[(size_t) t * I]);
        if (t < T) ch.ProcessBlock(&in[(size_t) t * I], &out[(size_t) t * I], T - t);
        rc = ch.status() == OLB_OK ? 0 : 5;
    } else if (!strcmp(unit, "chorus")) {
        ol::gpu::ChorusBank ch(I);
        if (ch.Init(48000.f) != ol::gpu::Ok) {
            fprintf(stderr, "Init failed: %s\n", ch.error().c_str());
            return 4;
        }
        ch.SetRate(0.7f);
        ch.SetDepth(0.9f);
        uint32_t t = 0;
        for (; t < 3 && t < T; t++) ch.Process(&in[(size_t) t * I], &out[(size_t) t * I]);
        if (t < T) ch.ProcessBlock(&in[(size_t) t * I], &out[(size_t) t * I], T - t);
        rc = ch.status() == OLB_OK ? 0 : 5;
        if (rc) fprintf(stderr, "process failed: %s\n", ch.error().c_str());
    } else if (!strcmp(unit, "poly")) {
        // modules/ol_daisy/app/synth/main.cpp:74-90 shape: zero the frame, Polyvoice::Process adds the voices into it
        ol::gpu::VoiceBank bank(I);
        ol::gpu::PolyvoiceBank poly(bank);
        if (poly.Init(48000.f) != ol::gpu::Ok) {
            fprintf(stderr, "Init failed: %s\n", bank.error().c_str());
            return 4;
        }
        out.assign(T, 0.f);
        poly.NoteOn(48, 100);
        poly.NoteOn(55, 100);
        uint32_t t = 0;
        for (; t < 3 && t < T; t++) poly.Process(&out[t]);
        const uint32_t half = T / 2;
        if (t < half) poly.ProcessBlock(&out[t], half - t);
        poly.NoteOff(48, 0);
        poly.NoteOn(64, 100);
        if (half < T) poly.ProcessBlock(&out[half > t ? half : t], T - (half > t ? half : t));
        rc = bank.status() == OLB_OK ? 0 : 5;
        if (rc) fprintf(stderr, "process failed: %s\n", bank.error().c_str());
        uint8_t playing[8] = {0};
        bank.Playing(playing, 0, I < 8 ? I : 8);
        for (uint32_t v = 0; v < (I < 8 ? I : 8); v++) printf("%u ", playing[v]);
        printf("\n");
    } else if (!strcmp(unit, "voicemap")) {
        // the scenario of tests/golden/make_golden.py VOICEMAP_*: pads on notes 36 / 38 / 42 / 60 / 62, channel 3 set twice
        ol::gpu::VoiceBank bank(I);
        ol::gpu::VoiceMapBank map(bank);
        if (map.Init(48000.f) != ol::gpu::Ok) {
            fprintf(stderr, "Init failed: %s\n", bank.error().c_str());
            return 4;
        }
        const int mapping[5][3] = {{0, 36, 0}, {1, 38, 1}, {2, 42, 2}, {3, 60, 3}, {3, 62, 4}};
        for (auto &m : mapping)
            if (!map.SetVoice((uint8_t) m[0], (uint8_t) m[1], (uint32_t) m[2])) return 6;
        if (map.SetVoice(5, 70, 0)) return 7;  // voice 0 already sits under note 36: refused
        struct Ev { uint32_t frame; int kind, a, v, ch; };  // kind 0 CC, 2 note on, 3 note off
        const Ev evs[] = {{0, 0, 41, 40, 1}, {0, 0, 42, 90, 3}, {0, 2, 36, 100, 0}, {3, 2, 38, 90, 0}, {400, 2, 62, 127, 0}, {401, 2, 61, 100, 0},
                          {900, 3, 36, 0, 0}, {1000, 2, 60, 80, 0}, {1200, 2, 200, 100, 0}, {1500, 0, 41, 100, 0}, {1700, 2, 36, 110, 0},
                          {2000, 3, 38, 0, 0}, {2400, 2, 42, 64, 0}, {2600, 0, 41, 20, 7}};
        out.assign(T, 0.f);
        uint32_t t = 0;
        size_t e = 0;
        const size_t n_ev = sizeof(evs) / sizeof(evs[0]);
        while (t < T) {
            for (; e < n_ev && evs[e].frame <= t; e++) {
                if (evs[e].kind == 0) map.UpdateMidiControl((uint8_t) evs[e].ch, (uint8_t) evs[e].a, (uint8_t) evs[e].v);
                else if (evs[e].kind == 2) map.NoteOn((uint8_t) evs[e].a, (uint8_t) evs[e].v);
                else map.NoteOff((uint8_t) evs[e].a, (uint8_t) evs[e].v);
            }
            const uint32_t stop = e < n_ev && evs[e].frame < T ? evs[e].frame : T;
            if (stop - t == 1) map.Process(&out[t]);
            else map.ProcessBlock(&out[t], stop - t);
            t = stop;
        }
        rc = bank.status() == OLB_OK ? 0 : 5;
        if (rc) fprintf(stderr, "process failed: %s\n", bank.error().c_str());
    } else if (!strcmp(unit, "sampler")) {
        // in.f32 is ONE mono sample of T0 = (file size) frames; every instance plays it: instance i loops [i, i + 40] from a
        // gate at frame 0, except instance 0 (one-shot from a seek to T0 - 10) -- out [T][I] pre-filled with 9
        ol::gpu::SamplerBank kit(I, 1);
        if (kit.Init(48000.f) != ol::gpu::Ok) {
            fprintf(stderr, "create failed: %s\n", kit.error().c_str());
            return 4;
        }
        const uint32_t id = kit.LoadSample(in.data(), in.size());
        kit.SetSample(id);
        for (uint32_t i = 1; i < I; i++) {
            kit.SetPlayMode(OLB_SAMPLE_LOOP, i, 1);
            kit.SetLoopStart(i, i, 1);
            kit.SetLoopEnd(i + 40, i, 1);
        }
        kit.GateOn();
        kit.Seek(in.size() - 10, 0, 1);
        out.assign((size_t) T * I, 9.f);
        uint32_t t = 0;
        for (; t < 2 && t < T; t++) kit.Process(&out[(size_t) t * I]);
        if (t < T) kit.ProcessBlock(&out[(size_t) t * I], T - t);
        rc = kit.status() == OLB_OK ? 0 : 5;
        if (rc) fprintf(stderr, "process failed: %s\n", kit.error().c_str());
        printf("%llu\n", (unsigned long long) kit.Position(0));
    } else {
        fprintf(stderr, "unknown unit %s\n", unit);
        return 2;
    }
    if (rc) return rc;
    dump(argv[5], out);
    if (!out2.empty() && argc > 6) dump(argv[6], out2);
    return 0;
}
