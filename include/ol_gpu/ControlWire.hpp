// ol_gpu/ControlWire.hpp -- the reference's control wire format on the host side of a bank, so
// that an existing control surface (modules/ol_teensy/app/synth/loop.h:118-134 writes these frames to a UART) can drive a
// GPU bank unchanged.  Header-only, no device code.
//
//   frame   = 5 x 0x2A, then three little-endian int64: message size (= 24, counting itself), controller, ADC value
//             (ol::io::SimpleSerializer::SerializeControl, modules/iolib/Serializer.h:113-133; int64_to_bytes,
//             modules/iolib/ol_iolib_core.h:11-19)
//   Control = one controller number with three views of its value -- 12-bit ADC (0..4096), MIDI (0..127), float (0..1) --
//             converted with ol::core::scale and truncation (modules/ctllib/Control.h:16-115)
//   parser  = the byte-at-a-time state machine of SimpleSerializer::Process (Serializer.h:72-111; disabled upstream
//             pending a UART rewrite, restated here): start bytes are counted while idle, a frame is parsed only after
//             EXACTLY five of them, the size field decides how many bytes follow.  One deviation: a stray byte outside a
//             frame is dropped (the upstream loop would emit a Control(0, 0) for it, its size check being 0 == 0).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace ol {
namespace gpu {

// ol::core::scale (modules/corelib/ol_corelib.h:27-44), restated
inline float WireScale(float in, float inlow, float inhigh, float outlow, float outhigh, float power) {
    const float d = inhigh - inlow;
    const float inscale = d == 0.f ? 0.f : 1.f / d;
    const float outdiff = outhigh - outlow;
    float value = (in - inlow) * inscale;
    if (value > 0.0f) value = std::pow(value, power);
    else if (value < 0.0f) value = -std::pow(-value, power);
    return value * outdiff + outlow;
}

struct Control {
    uint16_t controller = 0, adc = 0;
    float value = 0.f;
    uint8_t midi = 0;
    static Control FromAdc(uint16_t controller, uint16_t adc) {  // Control(CONTROLLER_TYPE, ADC_TYPE), Control.h:37-41
        Control c;
        c.controller = controller;
        c.adc = adc;
        c.midi = (uint8_t) WireScale(adc, 0, 4096, 0, 127, 1);
        c.value = WireScale(adc, 0, 4096, 0, 1, 1);
        return c;
    }
    static Control FromMidi(uint16_t controller, uint8_t midi) {  // Control(CONTROLLER_TYPE, MIDI_TYPE), Control.h:25-29
        Control c;
        c.controller = controller;
        c.midi = midi;
        c.value = WireScale(midi, 0, 127, 0, 1, 1);
        c.adc = (uint16_t) WireScale(midi, 0, 127, 0, 4096, 1);
        return c;
    }
    static Control FromFloat(uint16_t controller, float value) {  // Control(CONTROLLER_TYPE, FLOAT_TYPE), Control.h:31-35
        Control c;
        c.controller = controller;
        c.value = value;
        c.midi = (uint8_t) WireScale(value, 0, 1, 0, 127, 1);
        c.adc = (uint16_t) WireScale(value, 0, 1, 0, 4096, 1);
        return c;
    }
};

inline void SerializeControl(const Control &c, std::vector<uint8_t> &out) {
    for (int i = 0; i < 5; i++) out.push_back(0x2A);
    const int64_t fields[3] = {24, (int64_t) c.controller, (int64_t) c.adc};
    for (int64_t v : fields)
        for (int i = 0; i < 8; i++) out.push_back((uint8_t) ((v >> (i * 8)) & 0xFF));
}

class ControlWireParser {
public:
    // Feed received bytes; `on_control(const Control &)` is called for every complete frame.
    template <class F>
    void Feed(const uint8_t *bytes, size_t n, F &&on_control) {
        for (size_t i = 0; i < n; i++) {
            const uint8_t b = bytes[i];
            if (!parsing_ && b == 0x2A) {
                start_bytes_++;
                continue;
            }
            parsing_ = start_bytes_ == 5 && (parsed_ < size_ || size_ == 0);
            if (parsing_ && parsed_ < 8) {
                field_[0] |= (uint64_t) b << (8 * parsed_);
                if (++parsed_ == 8) size_ = field_[0];
            } else if (parsing_ && parsed_ < size_) {
                if (parsed_ < 16) field_[1] |= (uint64_t) b << (8 * (parsed_ - 8));
                else if (parsed_ < 24) field_[2] |= (uint64_t) b << (8 * (parsed_ - 16));
                parsed_++;
            }
            if (parsing_ && parsed_ == size_ && size_ != 0) {
                on_control(Control::FromAdc((uint16_t) field_[1], (uint16_t) field_[2]));
                Reset();
            } else if (!parsing_) {
                Reset();  // a byte that is neither a start byte nor part of a frame: drop what was collected
            }
        }
    }
    void Reset() {
        parsing_ = false;
        parsed_ = size_ = 0;
        start_bytes_ = 0;
        field_[0] = field_[1] = field_[2] = 0;
    }

private:
    bool parsing_ = false;
    uint64_t parsed_ = 0, size_ = 0, field_[3] = {0, 0, 0};
    size_t start_bytes_ = 0;
};

// The synth application's handler shape (modules/ol_daisy/app/synth/main.cpp:161-168): a received control goes to the
// bank's UpdateMidiControl with its MIDI view.
template <class Bank>
inline void ApplyControl(Bank &bank, const Control &c) {
    bank.UpdateMidiControl((uint8_t) c.controller, c.midi);
}

}  // namespace gpu
}  // namespace ol
