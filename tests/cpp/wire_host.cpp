// Control wire format checks against include/ol_gpu/ControlWire.hpp (host-only, no GPU).
//   wire_host convert <out.bin>          every ADC value 0..4096, MIDI value 0..127 and float k/1000: uint32 records
//                                        {which, input bits, controller, adc, midi, float bits, n wire bytes} + 40 bytes holding the frame
//   wire_host parse <in.bin> <out.bin>   feeds the byte stream in ragged pieces; writes {controller, adc, midi, float bits}
#include <cstdio>
#include <cstring>
#include <vector>

#include "ol_gpu/ControlWire.hpp"

using namespace ol::gpu;

static void record(FILE *f, uint32_t which, uint32_t in_bits, const Control &c) {
    std::vector<uint8_t> wire;
    SerializeControl(c, wire);
    uint32_t fb;
    memcpy(&fb, &c.value, 4);
    const uint32_t head[7] = {which, in_bits, c.controller, c.adc, c.midi, fb, (uint32_t) wire.size()};
    fwrite(head, 4, 7, f);
    wire.resize(40, 0);
    fwrite(wire.data(), 1, 40, f);
}

int main(int argc, char **argv) {
    if (argc >= 3 && !strcmp(argv[1], "convert")) {
        FILE *f = fopen(argv[2], "wb");
        for (uint32_t a = 0; a <= 4096; a++) record(f, 2, a, Control::FromAdc((uint16_t) (a % 200), (uint16_t) a));
        for (uint32_t m = 0; m <= 127; m++) record(f, 0, m, Control::FromMidi((uint16_t) m, (uint8_t) m));
        for (uint32_t k = 0; k <= 1000; k++) {
            const float v = (float) k / 1000.0f;
            uint32_t vb;
            memcpy(&vb, &v, 4);
            record(f, 1, vb, Control::FromFloat((uint16_t) (k % 128), v));
        }
        fclose(f);
        return 0;
    }
    if (argc >= 4 && !strcmp(argv[1], "parse")) {
        std::vector<uint8_t> in;
        FILE *f = fopen(argv[2], "rb");
        if (!f) return 3;
        uint8_t buf[4096];
        for (size_t n; (n = fread(buf, 1, sizeof buf, f)) > 0;) in.insert(in.end(), buf, buf + n);
        fclose(f);
        FILE *o = fopen(argv[3], "wb");
        ControlWireParser parser;
        size_t pos = 0, step = 1;
        while (pos < in.size()) {  // ragged pieces: 1, 2, 3 ... bytes at a time, as a UART delivers them
            const size_t n = in.size() - pos < step ? in.size() - pos : step;
            parser.Feed(&in[pos], n, [&](const Control &c) {
                uint32_t fb;
                memcpy(&fb, &c.value, 4);
                const uint32_t rec[4] = {c.controller, c.adc, c.midi, fb};
                fwrite(rec, 4, 4, o);
            });
            pos += n;
            step = step % 13 + 1;
        }
        fclose(o);
        return 0;
    }
    fprintf(stderr, "usage: %s convert out.bin | parse in.bin out.bin\n", argv[0]);
    return 2;
}
