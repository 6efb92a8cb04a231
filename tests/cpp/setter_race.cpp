// setter_race.cpp -- a control thread hammers the setters of a bank while the audio thread streams blocks through it, the
// way the reference's hosts drive their objects (MIDI thread vs audio callback with no synchronisation of their own:
// workouts/workout_buddy.cpp:35-72 vs :76-100; modules/ol_daisy/app/synth/main.cpp:161-168 vs :74-90).
// The library's contract (include/olbank.h): every entry point takes the handle's mutex, setters touch host tables only, and
// the device tables are brought up to date at the top of the next process call -- so every block is computed with ONE
// coherent parameter set (the one before or the one after a setter call, never a torn one).
//
// The control thread flips ALL channels between two coefficient sets A and B (a 0 dB and a -6 dB one-section cascade) with a
// single set_coeffs call; the audio thread feeds a constant 1.0 and checks, block by block, that all channels of a frame
// agree with each other (a torn table would give some channels A and some B inside one block).  Also run under
// -fsanitize=thread (make -C ol_dsp_b200/csrc tsan; profiles/).
//   usage: setter_race <channels> <blocks> <frames-per-block>       exit 0 = no torn block, no error code
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "ol_gpu/BiquadBank.hpp"
#include "ol_gpu/ChorusBank.hpp"
#include "ol_gpu/FxBank.hpp"
#include "ol_gpu/VoiceBank.hpp"

int main(int argc, char **argv) {
    const uint32_t C = argc > 1 ? (uint32_t) atoi(argv[1]) : 4096;
    const int blocks = argc > 2 ? atoi(argv[2]) : 400;
    const uint64_t F = argc > 3 ? (uint64_t) atoll(argv[3]) : 64;
    ol::gpu::BiquadBank bank(C, 1);
    if (bank.Init(48000.f) != ol::gpu::Ok) {
        fprintf(stderr, "Init: %s\n", bank.error().c_str());
        return 2;
    }
    bank.SetMode(OLB_MODE_FAST);  // the tier whose history follows the table's scaling: the rescale path runs under the race too
    std::vector<float> A((size_t) C * 5, 0.f), B((size_t) C * 5, 0.f);
    for (uint32_t c = 0; c < C; c++) {
        A[(size_t) c * 5] = 1.0f;   // pure gain 1
        B[(size_t) c * 5] = 0.5f;   // pure gain 0.5
    }
    bank.SetCoeffs(0, C, A.data());
    ol::gpu::FxBank<OLB_FX_FILTER> fx(256);
    if (fx.Init(48000.f) != ol::gpu::Ok) return 2;
    // the banks of round 2 under the same race: a stereo rack, a chorus (with its pitch shifter), a voice bank fed event lists
    ol::gpu::FxRack2Bank rack2(64);
    ol::gpu::ChorusBank chorus(64);
    ol::gpu::VoiceBank voices(128);
    if (rack2.Init(48000.f) != ol::gpu::Ok || chorus.Init(48000.f) != ol::gpu::Ok || voices.Init(48000.f) != ol::gpu::Ok) return 2;

    std::atomic<bool> stop{false};
    std::atomic<long> flips{0};
    int ctl_status = OLB_OK;
    std::thread control([&] {
        bool a = false;
        // the C ABI directly: the C++ mirror objects keep a last-status member that belongs to the thread that owns them
        while (!stop.load(std::memory_order_relaxed)) {
            int rc = olb_biquad_set_coeffs(bank.handle(), 0, C, a ? A.data() : B.data());
            a = !a;
            const float cutoff = 0.25f + 0.5f * (float) (flips.load() & 1);
            if (rc == OLB_OK) rc = olb_fx_update_hw(fx.handle(), 0, 256, 41 /* CC_FILTER_CUTOFF */, &cutoff, 0);
            if (rc == OLB_OK) rc = olb_fx_set_mode(fx.handle(), (flips.load() & 2) ? OLB_MODE_FMA : OLB_MODE_EXACT);  // the tier of the NEXT block
            if (rc == OLB_OK) rc = olb_fx2_update_midi(rack2.handle(), 0, 64, 35 /* CC_DELAY_TIME */, (uint8_t) (20 + (flips.load() & 63)));
            if (rc == OLB_OK) rc = olb_fx2_set_mode(rack2.handle(), (flips.load() & 4) ? OLB_MODE_FMA : OLB_MODE_EXACT);
            if (rc == OLB_OK) rc = olb_chorus_set_param(chorus.handle(), 0, 64, OLB_CHORUS_PITCH, (float) (flips.load() & 3));
            if (rc == OLB_OK) rc = olb_chorus_set_param(chorus.handle(), 0, 64, OLB_CHORUS_DEPTH, cutoff);
            if (rc == OLB_OK) rc = olb_voice_update_midi(voices.handle(), 0, 128, 41 /* CC_FILTER_CUTOFF */, (uint8_t) (40 + (flips.load() & 63)));
            if (rc != OLB_OK) ctl_status = rc;
            flips.fetch_add(1);
        }
    });

    std::vector<float> x((size_t) F * C, 1.0f), y((size_t) F * C), fx_in((size_t) F * 256, 0.1f), fx_out((size_t) F * 256);
    std::vector<float> st_in((size_t) F * 128, 0.1f), st_out((size_t) F * 128), ch_out((size_t) F * 64);
    long torn = 0, seen_a = 0, seen_b = 0;
    for (int b = 0; b < blocks; b++) {
        bank.ProcessBlock(x.data(), y.data(), F);
        fx.ProcessBlock(fx_in.data(), fx_out.data(), F);
        rack2.ProcessBlock(st_in.data(), st_out.data(), F);
        chorus.ProcessBlock(st_in.data(), ch_out.data(), F);
        {   // a NoteOn / NoteOff pair on rotating voices, sample-accurate, applied on the device
            const uint32_t vv = (uint32_t) b % 128u;
            olb_voice_event ev[2] = {{(uint32_t) (F / 4), OLB_EV_NOTE_ON, 48u + vv % 24u, 100.f, (int32_t) vv},
                                     {(uint32_t) (3 * F / 4), OLB_EV_NOTE_OFF, 48u + vv % 24u, 0.f, (int32_t) vv}};
            // (device buffers only on this entry point; none given: the voices advance, which is all the race needs)
            const int rc = olb_voice_process_events(voices.handle(), ev, 2, nullptr, nullptr, 1, 0, 1.0f, F, nullptr);
            if (rc != OLB_OK) {
                fprintf(stderr, "voices: %s\n", olb_last_error());
                stop = true;
                control.join();
                return 3;
            }
        }
        if (bank.status() != OLB_OK || fx.status() != OLB_OK || rack2.status() != OLB_OK || chorus.status() != OLB_OK) {
            fprintf(stderr, "process: %s\n", bank.error().c_str());
            stop = true;
            control.join();
            return 3;
        }
        for (uint64_t t = 0; t < F; t++) {
            const float v = y[t * C];
            if (v != 1.0f && v != 0.5f) torn++;
            for (uint32_t c = 1; c < C; c++)
                if (y[t * C + c] != v) {
                    torn++;
                    break;
                }
            if (t && v != y[0]) torn++;  // the table cannot change INSIDE a block
        }
        (y[0] == 1.0f ? seen_a : seen_b)++;
        for (size_t i = 0; i < fx_out.size(); i++)
            if (!std::isfinite(fx_out[i])) torn++;
        for (size_t i = 0; i < st_out.size(); i++)
            if (!std::isfinite(st_out[i])) torn++;
        for (size_t i = 0; i < ch_out.size(); i++)
            if (!std::isfinite(ch_out[i])) torn++;
    }
    stop = true;
    control.join();
    printf("blocks %d  setter calls %ld  blocks with A %ld / with B %ld  torn %ld  control status %d\n", blocks, flips.load(), seen_a, seen_b, torn,
           ctl_status);
    return (torn == 0 && ctl_status == OLB_OK) ? 0 : 1;
}
