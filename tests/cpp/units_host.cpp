// C++ callers written the way a reference user writes them (Init -> setters -> Process per frame, then per block), against
// include/ol_gpu/{FxBank,DattorroBank,ChorusBank,VoiceBank}.hpp.  Built by tests/test_abi.py (link check, CPU) and run by
// tests/test_cpp_hosts_gpu.py on the GPU box, where the outputs are compared with the oracle.
//   usage: units_host <filter|delay|reverb|rack|plate|chorus|poly> <instances> <frames> <in.f32> <out.f32> [out2.f32]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ol_gpu/ChorusBank.hpp"
#include "ol_gpu/DattorroBank.hpp"
#include "ol_gpu/FxBank.hpp"
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
static int run_fx(uint32_t I, uint32_t T, std::vector<float> &in, std::vector<float> &out) {
    Bank bank(I);
    if (bank.Init(48000.f) != ol::gpu::Ok) {
        fprintf(stderr, "Init failed: %s\n", bank.error().c_str());
        return 4;
    }
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
    if (strcmp(unit, "poly") && !slurp(argv[4], in)) return 3;
    int rc = 0;
    if (!strcmp(unit, "filter")) rc = run_fx<ol::gpu::FilterFxBank>(I, T, in, out);
    else if (!strcmp(unit, "delay")) rc = run_fx<ol::gpu::DelayFxBank>(I, T, in, out);
    else if (!strcmp(unit, "reverb")) rc = run_fx<ol::gpu::ReverbFxBank>(I, T, in, out);
    else if (!strcmp(unit, "rack")) rc = run_fx<ol::gpu::FxRackBank>(I, T, in, out);
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
    } else {
        fprintf(stderr, "unknown unit %s\n", unit);
        return 2;
    }
    if (rc) return rc;
    dump(argv[5], out);
    if (!out2.empty() && argc > 6) dump(argv[6], out2);
    return 0;
}
