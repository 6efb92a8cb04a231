// ol_gpu/Wav.hpp -- minimal RIFF/WAVE reader and writer for the end-to-end tools.  The reference
// decodes its test material (workouts/drum_loop.wav: stereo, 24-bit PCM, 44.1 kHz) with the vendored miniaudio
// (libs/miniaudio/miniaudio.h) to interleaved f32; this reader produces the same samples for PCM 8 / 16 / 24 / 32-bit and
// IEEE float 32 (integer PCM is scaled by 1 / 2^(bits-1), 8-bit is unsigned with a 128 bias) -- own code, no dependency.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace ol {
namespace gpu {

struct WavData {
    uint32_t sample_rate = 0, channels = 0;
    uint64_t frames = 0;
    std::vector<float> samples;  // [frames][channels]
};

inline bool ReadWav(const std::string &path, WavData &out, std::string *err = nullptr) {
    auto fail = [&](const char *m) {
        if (err) *err = m;
        return false;
    };
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return fail("cannot open file");
    std::vector<uint8_t> raw;
    uint8_t buf[1 << 16];
    for (size_t n; (n = fread(buf, 1, sizeof buf, f)) > 0;) raw.insert(raw.end(), buf, buf + n);
    fclose(f);
    if (raw.size() < 12 || memcmp(&raw[0], "RIFF", 4) || memcmp(&raw[8], "WAVE", 4)) return fail("not a RIFF/WAVE file");
    auto u16 = [&](size_t o) { return (uint32_t) raw[o] | ((uint32_t) raw[o + 1] << 8); };
    auto u32 = [&](size_t o) { return u16(o) | (u16(o + 2) << 16); };
    uint32_t format = 0, bits = 0, align = 0;
    size_t data_off = 0, data_len = 0;
    for (size_t o = 12; o + 8 <= raw.size();) {
        const uint32_t len = u32(o + 4);
        if (!memcmp(&raw[o], "fmt ", 4) && o + 8 + 16 <= raw.size()) {
            format = u16(o + 8);
            out.channels = u16(o + 10);
            out.sample_rate = u32(o + 12);
            align = u16(o + 20);
            bits = u16(o + 22);
            if (format == 0xFFFE && len >= 26) format = u16(o + 8 + 24);  // WAVE_FORMAT_EXTENSIBLE: sub-format GUID's first word
        } else if (!memcmp(&raw[o], "data", 4)) {
            data_off = o + 8;
            data_len = len <= raw.size() - data_off ? len : raw.size() - data_off;
            break;
        }
        o += 8 + (size_t) len + (len & 1);
    }
    if (!out.channels || !data_off || !align) return fail("missing fmt or data chunk");
    if (!((format == 1 && (bits == 8 || bits == 16 || bits == 24 || bits == 32)) || (format == 3 && bits == 32))) return fail("unsupported sample format");
    const uint32_t bytes = bits / 8;
    out.frames = data_len / align;
    out.samples.resize(out.frames * out.channels);
    const uint8_t *p = &raw[data_off];
    for (uint64_t i = 0; i < out.frames * out.channels; i++, p += bytes) {
        float v;
        if (format == 3) {
            memcpy(&v, p, 4);
        } else if (bits == 8) {
            v = (float) ((int) p[0] - 128) * (1.0f / 128.0f);
        } else if (bits == 16) {
            v = (float) (int16_t) (p[0] | (p[1] << 8)) * (1.0f / 32768.0f);
        } else if (bits == 24) {
            const int32_t s = (int32_t) (((uint32_t) p[0] << 8) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 24)) >> 8;
            v = (float) s * (1.0f / 8388608.0f);
        } else {
            const int32_t s = (int32_t) ((uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24));
            v = (float) ((double) s * (1.0 / 2147483648.0));
        }
        out.samples[i] = v;
    }
    return true;
}

// IEEE float 32 output (format tag 3): the processed signal is written back without requantisation.
inline bool WriteWavF32(const std::string &path, const float *samples, uint64_t frames, uint32_t channels, uint32_t sample_rate) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const uint32_t data = (uint32_t) (frames * channels * 4);
    auto w32 = [&](uint32_t v) { fwrite(&v, 4, 1, f); };
    auto w16 = [&](uint16_t v) { fwrite(&v, 2, 1, f); };
    fwrite("RIFF", 1, 4, f); w32(36 + data); fwrite("WAVE", 1, 4, f);
    fwrite("fmt ", 1, 4, f); w32(16); w16(3); w16((uint16_t) channels); w32(sample_rate); w32(sample_rate * channels * 4);
    w16((uint16_t) (channels * 4)); w16(32);
    fwrite("data", 1, 4, f); w32(data);
    const bool ok = fwrite(samples, 4, frames * channels, f) == frames * channels;
    fclose(f);
    return ok;
}

}  // namespace gpu
}  // namespace ol
