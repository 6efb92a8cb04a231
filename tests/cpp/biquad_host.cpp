// C++ caller written the way a reference user writes one (Init -> setters -> Process per frame / per block),
// against include/ol_gpu/BiquadBank.hpp.  Built by tests/test_abi.py (link check, CPU) and run by
// tests/test_biquad_gpu.py on the GPU box, where its output is compared with the oracle.
//   usage: biquad_host <channels> <frames> <in.f32> <coefs.f32> <out.f32> <mode>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ol_gpu/BiquadBank.hpp"

static bool slurp(const char *path, std::vector<float> &v) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    const size_t n = fread(v.data(), sizeof(float), v.size(), f);
    fclose(f);
    return n == v.size();
}

int main(int argc, char **argv) {
    if (argc != 7) {
        fprintf(stderr, "usage: %s channels frames in.f32 coefs.f32 out.f32 mode\n", argv[0]);
        return 2;
    }
    const uint32_t C = (uint32_t) atoi(argv[1]), T = (uint32_t) atoi(argv[2]), S = 8;
    std::vector<float> in((size_t) C * T), out((size_t) C * T), coefs((size_t) C * S * 5);
    if (!slurp(argv[3], in) || !slurp(argv[4], coefs)) return 3;
    ol::gpu::BiquadBank bank(C, S);
    if (bank.Init(48000.f) != ol::gpu::Ok) {
        fprintf(stderr, "Init failed: %s\n", bank.error().c_str());
        return 4;
    }
    bank.SetMode(atoi(argv[6]));
    bank.SetCoeffs(0, C, coefs.data());
    // first 3 frames one at a time (the reference's per-frame Process), the rest as one block
    uint32_t t = 0;
    for (; t < 3 && t < T; t++) bank.Process(&in[(size_t) t * C], &out[(size_t) t * C]);
    if (t < T) bank.ProcessBlock(&in[(size_t) t * C], &out[(size_t) t * C], T - t);
    if (bank.status() != OLB_OK) {
        fprintf(stderr, "process failed: %s\n", bank.error().c_str());
        return 5;
    }
    FILE *f = fopen(argv[5], "wb");
    fwrite(out.data(), sizeof(float), out.size(), f);
    fclose(f);
    printf("%s\n", bank.LastKernel());
    return 0;
}
