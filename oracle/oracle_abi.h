// TEST INFRASTRUCTURE ONLY (oracle/).  Plain-C types shared by the oracle libraries
// (oracle/_ref/libolref.so, oracle/libolb_oracle.so) and their ctypes loader (oracle/loader.py).
#ifndef OLB_ORACLE_ABI_H
#define OLB_ORACLE_ABI_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

// A control event applied BEFORE the frame with index `frame` is processed.
// Mirrors the reference's setter surface (Fx.h:116-164, SynthVoice.h:101-263, Voice.h:33-57).
enum {
    OLB_EV_MIDI_CC = 0,   // UpdateMidiControl(cc=a, value=(uint8)v)
    OLB_EV_HW_CC = 1,     // UpdateHardwareControl(cc=a, value=v)
    OLB_EV_NOTE_ON = 2,   // NoteOn(note=a, velocity=(uint8)v)
    OLB_EV_NOTE_OFF = 3,  // NoteOff(note=a, velocity=(uint8)v)
    OLB_EV_INIT = 4,      // Init(sample_rate=v)  (re-Init resets state, test/fx_test.cpp:32,42)
    OLB_EV_SET_FREQ = 5,  // SetFrequency(v)
    OLB_EV_GATE_ON = 6,
    OLB_EV_GATE_OFF = 7,
    // sample playback (Sample.h:20-38; gates go through SampleSoundSource::GateOn / GateOff, SampleSoundSource.h:21-26)
    OLB_EV_SAMPLE_SEEK = 8,        // Seek(a)
    OLB_EV_SAMPLE_LOOP_START = 9,  // SetLoopStart(a)
    OLB_EV_SAMPLE_LOOP_END = 10,   // SetLoopEnd(a)
    OLB_EV_SAMPLE_MODE = 11,       // SetPlayMode(a): 0 OneShot, 1 Loop
    OLB_EV_SAMPLE_TOGGLE = 12,     // TogglePlay()
    OLB_EV_SAMPLE_PLAY = 13,       // Play()
    OLB_EV_SAMPLE_PAUSE = 14,      // Pause()
};

typedef struct olb_event {
    uint32_t frame;
    uint32_t kind;
    uint32_t a;
    float v;
    int32_t voice;  // polyvoice drivers: >=0 addresses one voice, <0 goes through Polyvoice
} olb_event;

#ifdef __cplusplus
}
#endif
#endif
