/* olbank.h -- C ABI of libolbank.so: banks of independent audio-DSP units on one B200 (sm_100a).
 *
 * This is the drop-in boundary for the hot path of oletizi/ol_dsp.  The reference
 * has no FFI layer: its operator API is the duck-typed method set
 *     Init(sample_rate) / Process(frame_in, frame_out) / Update*Control(cc, value)
 * on header-only C++ classes.  Every entry point below names the reference interface it replaces
 * (file:line under the reference tree).  The C++ wrappers in include/ol_gpu/ keep the reference's
 * method names on top of these calls; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *   - every function returns 0 (OLB_OK) or a negative OLB_E_* code; olb_last_error() gives the text
 *     (the reference returns void everywhere and never throws: Fx.h:38,88,194; SoundSource.h:9-11)
 *   - audio buffers named *_dev are caller-owned DEVICE pointers on the bank's GPU; *_host are host
 *     pointers (pinned memory makes the copies asynchronous); handles own all other device memory
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream)
 *   - a bank keeps its recurrence state on the device between process calls, exactly as a reference
 *     object keeps it between Process() calls; *_init() resets it (re-Init: test/fx_test.cpp:32,42)
 *   - there is no CPU fallback: without a usable sm_100 device every create call fails
 */
#ifndef OLBANK_H
#define OLBANK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OLB_OK 0
#define OLB_E_ARG (-1)   /* bad argument (NULL handle, size 0, unsupported layout ...) */
#define OLB_E_CUDA (-2)  /* a CUDA runtime/driver call failed; see olb_last_error() */
#define OLB_E_NCCL (-3)  /* an NCCL call failed (the cross-GPU voice-bus reduce); see olb_last_error() */
#define OLB_E_STATE (-4) /* call not valid in the handle's current state */

/* Audio layouts.  A reference frame is N consecutive floats (Fx.h:38); a bank frame is one float per
 * unit, so the frame-interleaved block [frames][units] is the bank analogue of the reference's
 * "for each frame: Process(in, out)" loop (ol_daisy/app/synth/main.cpp:74-90). */
#define OLB_LAYOUT_FRAMES 0 /* [frames][units]  unit index fastest  (coalesced; the fast path) */
#define OLB_LAYOUT_PLANAR 1 /* [units][frames]  one contiguous timeline per unit */

/* Library / device ------------------------------------------------------------------------------ */
const char *olb_last_error(void);
const char *olb_version(void);
int olb_device_count(int *count);
/* Pinned host memory for the *_host entry points (cudaHostAlloc / cudaFreeHost). */
int olb_host_alloc(void **ptr, size_t bytes);
int olb_host_free(void *ptr);
/* Host placement for the *_host entry points on a multi-socket box (one process per GPU): olb_bind_thread_near pins the
 * CALLING thread to the cores of the NUMA node the GPU hangs off (/sys/bus/pci/devices/<bdf>/numa_node; a no-op when the
 * platform reports none), olb_host_alloc_near does that and then allocates pinned memory from there.  *numa_node_out
 * (optional) receives the node, -1 = unknown. */
int olb_bind_thread_near(int device, int *numa_node_out);
int olb_host_alloc_near(int device, void **ptr, size_t bytes, int *numa_node_out);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
uint64_t olb_launch_count(void);

/* Replayable white noise: Philox4x32-10, key (seed, 0x01D5B200), counter
 * (frame/4 lo, frame/4 hi, channel, 0); x = (u >> 8) * 2^-23 - 1 in [-1, 1), times `scale`.
 * Fills units [first_unit, first_unit+units) x frames [first_frame, first_frame+frames) of a buffer in
 * `layout`.  The CPU oracle replays the same stream. */
int olb_noise_fill(float *out_dev, uint32_t seed, uint32_t first_unit, uint32_t units, uint64_t first_frame,
                   uint64_t frames, float scale, int layout, void *stream);

/* (a1) Biquad / one-pole cascade bank ------------------------------------------------------------
 * No class of this name exists in the reference; the bank follows the
 * reference's object conventions and reproduces its two one-pole forms as section kinds:
 *   OLB_KIND_BIQUAD  y = b0*x + (((b1*x1 + b2*x2) - a2*y2) - a1*y1)          direct form 1
 *   OLB_KIND_PORT    z = x + c*(z - x)      daisysp::Port::Process, modules/synthlib/Portamento.h:29-32
 *   OLB_KIND_LPF     o = o + (x - o)*f      LowPassFilter_process, libs/dattorro-verb/verb.cpp:131-134
 * Modes:
 *   OLB_MODE_EXACT   every product and sum rounded separately, in the order written above
 *                    (bit-exact against the scalar oracle built with -ffp-contract=off)
 *   OLB_MODE_FMA     p = fma(b2,x2, b1*x1); p = fma(-a2,y2,p); p = fma(-a1,y1,p); y = fma(b0,x,p);
 *                    one-pole kinds are rewritten as biquads first (PORT: b0=1-c, a1=-c;
 *                    LPF: b0=f, a1=-(1-f)).  Same operations as the oracle's FMA mode, so bit-exact to it.
 *   OLB_MODE_FAST    the normalised fused cascade, 4 fused ops per section instead of 5 (the throughput tier):
 *                    r1 = b1/b0, r2 = b2/b0 per section, G = prod(b0) per channel, all computed on the host in
 *                    double and rounded to float;  v = G*x;  per section
 *                    t = fma(r1,x1,v); t = fma(r2,x2,t); t = fma(-a2,y2,t); y = fma(-a1,y1,t).
 *                    Bit-exact to the oracle's FAST mode; a tolerance tier against EXACT (DESIGN.md gives the
 *                    measured error: it is as close to the float64 truth as EXACT is).  Node s of the cascade carries
 *                    the signal times prod_{j>=s} b0_j and get/set_state speak that scaling.  A bank in which some
 *                    channel has b0 == 0 (or a partial product outside [1e-30, 1e30]) runs the FMA arithmetic. */
#define OLB_KIND_BIQUAD 0
#define OLB_KIND_PORT 1
#define OLB_KIND_LPF 2
#define OLB_MODE_EXACT 0
#define OLB_MODE_FMA 1
#define OLB_MODE_FAST 2
#define OLB_BIQUAD_MAX_SECTIONS 16

typedef struct olb_biquad olb_biquad;

/* default-construct + Init(sample_rate)  (lifecycle of every reference unit, e.g. Fx.h:83-86).
 * All coefficients start as the identity section (b0 = 1), kinds as OLB_KIND_BIQUAD, mode EXACT. */
int olb_biquad_create(int device, uint32_t channels, uint32_t sections, float sample_rate, olb_biquad **out);
int olb_biquad_destroy(olb_biquad *h);
/* Init(sample_rate): zeroes the recurrence state, keeps parameters (Fx.h:83-86 keeps cutoff etc.). */
int olb_biquad_init(olb_biquad *h, float sample_rate);
/* kinds: one uint8 per section, shared by every channel of the bank. */
int olb_biquad_set_kinds(olb_biquad *h, const uint8_t *kinds, uint32_t sections);
int olb_biquad_set_mode(olb_biquad *h, int mode);
/* Parameter setter: coefs_host is [count][sections][5] = b0 b1 b2 a1 a2 for channels
 * [first, first+count).  Takes effect at the next process call (setters in the reference take
 * effect at the next Process(): Fx.h:110-114). */
int olb_biquad_set_coeffs(olb_biquad *h, uint32_t first, uint32_t count, const float *coefs_host);
/* RBJ-cookbook designer in the reference's "setter computes coefficients on the host" style
 * (the Svf setters do the same: SURVEY.md section 2.3).  type: 0 low-pass, 1 high-pass,
 * 2 peaking (gain_db), 3 one-pole low-pass as a biquad.  Computed in double, rounded to float. */
int olb_biquad_design(int type, double sample_rate, double f0, double q, double gain_db, float *coefs5);
/* Block form of "for each frame: Process(in, out)".  in_dev == out_dev is allowed (the reference is
 * called in place: main.cpp:82-85).  frames == 1 is the per-frame call. */
int olb_biquad_process(olb_biquad *h, const float *in_dev, float *out_dev, uint64_t frames, int layout, void *stream);
/* Same with host buffers: chunks the block in time, overlaps H2D / kernel / D2H on internal
 * streams, returns when out_host is complete. */
int olb_biquad_process_host(olb_biquad *h, const float *in_host, float *out_host, uint64_t frames, int layout);
/* State as the oracle lays it out: [channels][sections][4] = x1 x2 y1 y2 (host pointers). */
int olb_biquad_get_state(olb_biquad *h, float *state_host);
int olb_biquad_set_state(olb_biquad *h, const float *state_host);
/* Diagnostics for bench.py: which kernel the last process call used ("k_bq_tma", "k_bq_roll", "k_bq_skew", "k_bq_sys",
 * "k_bq_direct", "k_bq_planar", "k_bq_runtime": DESIGN.md section 3 says which bank sizes take which), and kernel selection
 * override (0 = automatic, 1 = direct, 2 = tma, 3 = sys, 4 = skew / roll).  Every kernel computes the same bits. */
const char *olb_biquad_last_kernel(olb_biquad *h);
int olb_biquad_set_variant(olb_biquad *h, int variant);

/* (a2-a5) fxlib unit banks: FilterFx<1>, DelayFx<1>, ReverbFx<1>, FxRack<1> -----------------------------
 * One bank = `instances` mono objects of one reference class:
 *   OLB_FX_FILTER  ol::fx::FilterFx<1>  modules/fxlib/Fx.h:65-165
 *   OLB_FX_DELAY   ol::fx::DelayFx<1>   modules/fxlib/Fx.h:170-268  (+ its daisysp::DelayLine<float,48000> and wet FilterFx)
 *   OLB_FX_REVERB  ol::fx::ReverbFx<1>  modules/fxlib/Fx.h:272-393  around the stub ReverbSc (modules/fxlib/Reverb.h:13-41)
 *   OLB_FX_RACK    ol::fx::FxRack<1>    modules/fxlib/Fx.h:397-493  delay -> reverb -> filter -> master volume
 * create = construct with the reference's default members + Init(sample_rate).  The CC numbers are
 * modules/corelib/cc_map.h:7-70; values map through ol::core::scale exactly as in the reference's
 * UpdateMidiControl / UpdateHardwareControl, on the host.  Audio is [frames][instances] device memory. */
#define OLB_FX_FILTER 0
#define OLB_FX_DELAY 1
#define OLB_FX_REVERB 2
#define OLB_FX_RACK 3
/* FxRack<1> with the Dattorro plate of libs/dattorro-verb in the reverb slot instead of the stub ReverbSc -- the wiring the
 * reference sketches in modules/fxlib/ReverbFx.cpp:11-37 (commented out there): mono in -> DattorroVerb_process ->
 * getLeft as the wet signal of ReverbFx's dry/wet; PreFilter <- pre_cutoff, InputDiffusion1/2, Decay <- decay_time,
 * DecayDiffusion from the ReverbFx members.  One deviation: Damping <- cutoff / 20 000 (the sketch passes the cutoff in
 * Hz, which is outside the plate's 0..1 coefficient range). */
#define OLB_FX_RACK_PLATE 4

typedef struct olb_fx olb_fx;

int olb_fx_create(int device, int kind, uint32_t instances, float sample_rate, olb_fx **out);
int olb_fx_destroy(olb_fx *h);
/* Init(sample_rate): clears filter state and delay lines, keeps the parameter members (test/fx_test.cpp:32,42). */
int olb_fx_init(olb_fx *h, float sample_rate);
/* UpdateMidiControl(control, value 0..127) for instances [first, first+count)  (Fx.h:116-139, 239-266, 313-351, 443-468) */
int olb_fx_update_midi(olb_fx *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value);
/* UpdateHardwareControl(control, value 0..1), one value per instance at values_host[i * stride] (stride 0 = one value
 * for all)  (Fx.h:141-164, 219-237, 353-391, 470-492) */
int olb_fx_update_hw(olb_fx *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride);
/* Arithmetic tier.  OLB_MODE_EXACT (the default) rounds every product and sum separately, in the reference's order: bit-exact
 * against the reference built with -ffp-contract=off (oracle/Makefile).  OLB_MODE_FMA is the same signal flow with the
 * contractions a compiler makes under -ffp-contract=fast, written out so that they are defined:
 *     svf half-step   notch = fma(-damp, band, in); low = fma(freq, band, low); high = notch - low;
 *                     band  = fma(-(drive*band)*band, band, fma(freq, high, band))
 *     svf output      out   = fma(0.5, pick(second half), 0.5 * pick(first half))
 *     delay line      read  = fma(b - a, frac, a);  write = fma(feedback, read, in)
 *     dry/wet mixes   out   = fma(wet, balance, dry * (1 - balance))
 * bit-exact against the oracle's restatement of exactly those forms (orc_fx_set_mode), and within 1e-4 of max(1, peak) of
 * the EXACT tier over 10 s of 0.9-scale noise at five corners of the CC maps (tests/test_fx_gpu.py; measured on the device:
 * 4.3e-7 wherever the output stays below 16, 2.2e-5 with every controller at 127, where the feedback loop peaks at 4 242).
 * It is the throughput tier: FilterFx's EXACT kernel runs at the FP32 issue rate of its 100 separately rounded operations per four
 * frames.  State carries over a switch.  OLB_FX_RACK_PLATE has OLB_MODE_EXACT only; any other mode
 * value is OLB_E_ARG. */
int olb_fx_set_mode(olb_fx *h, int mode);
/* Block form of "for each frame: Process(frame_in, frame_out)"; in_dev == out_dev allowed. */
int olb_fx_process(olb_fx *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream);
/* Diagnostics: which kernel the last process call launched ("k_fx_split", "k_fx_tma", "k_fx_pf", "k_fx"). */
const char *olb_fx_last_kernel(olb_fx *h);
/* Same with HOST buffers [frames][instances] (the reference's Process takes host frames: Fx.h:88,194,294,426); chunked
 * through device staging, returns when out_host is complete.  frames == 1 is the per-frame call. */
int olb_fx_process_host(olb_fx *h, const float *in_host, float *out_host, uint64_t frames);

/* (a2-a5, CHANNEL_COUNT = 2) stereo fxlib unit banks: FilterFx<2>, DelayFx<2>, ReverbFx<2>, FxRack<2> -- the shapes the
 * reference's own caller runs (modules/ol_daisy/app/synth/main.cpp:83-85: reverb.Process / filter.Process on a stereo
 * frame).  Audio is [frames][instances][2]: an instance's frame is the reference's t_sample[2].  The second channel gets
 * exactly what the reference's templates give it:
 *   OLB_FX_FILTER  Fx.h:88-108   filters frame_in[0] into frame_out[0]; frame_out[1] is not written (out_dev keeps it)
 *   OLB_FX_DELAY   Fx.h:194-208  one delay line per channel; the wet filter runs on channel 0 only
 *   OLB_FX_REVERB  Fx.h:294-300  both channels through the stub ReverbSc and the dry/wet
 *   OLB_FX_RACK    Fx.h:426-434  channel 0 = FxRack<1> of channel 0; channel 1 = 0 * master_volume (buf_c[1] is never written)
 * Setters are per INSTANCE (one object, two channels), same CC maps as the mono banks.  OLB_FX_RACK_PLATE has no stereo
 * form.  in_dev == out_dev allowed; 8-byte aligned. */
typedef struct olb_fx2 olb_fx2;
int olb_fx2_create(int device, int kind, uint32_t instances, float sample_rate, olb_fx2 **out);
int olb_fx2_destroy(olb_fx2 *h);
int olb_fx2_init(olb_fx2 *h, float sample_rate);
int olb_fx2_update_midi(olb_fx2 *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value);
int olb_fx2_update_hw(olb_fx2 *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride);
/* Both channels in the tier (olb_fx_set_mode). */
int olb_fx2_set_mode(olb_fx2 *h, int mode);
int olb_fx2_process(olb_fx2 *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream);
/* "k_fx2_put" (merge; DelayFx<2>'s channel 1 time-parallel) or "k_fx2_delay1" (channel 1 walking time: delays shorter than a cut) */
const char *olb_fx2_last_kernel(olb_fx2 *h);
/* HOST buffers [frames][instances][2]; for OLB_FX_FILTER out_host's channel 1 is read and kept. */
int olb_fx2_process_host(olb_fx2 *h, const float *in_host, float *out_host, uint64_t frames);
/* olb_fx2_process without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same results (bit-identical) as olb_fx2_process; the stereo path's only scratch -- the
 * de-interleaved channel-0 cut -- lives in a static device array of OLB_FX2_STATIC_SCRATCH_BYTES that every GPU which loads
 * the library holds until exit.  Limits: banks of at most OLB_FX2_STATIC_MAX_INSTANCES instances (one 64-frame cut must
 * fit; OLB_E_ARG beyond); any number of frames.  The array is shared by ALL handles and calls on a device and keeps the
 * last call's contents (every cut overwrites what it reads, nothing needs clearing): calls that overlap on the GPU --
 * different streams with no ordering between them -- race; serialise them.  Calls no cudaMalloc / cudaFree. */
#define OLB_FX2_STATIC_SCRATCH_BYTES (8u << 20)
#define OLB_FX2_STATIC_MAX_INSTANCES 32768u
int olb_fx2_process_static(olb_fx2 *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream);

/* (a6) SynthVoice bank and (a8) Polyvoice mixdown -------------------------------------------------------
 * One bank = `voices` ol::synth::SynthVoice objects with the reference's default parts (PolyBLEP-saw
 * OscillatorSoundSource, SvfFilter, two DaisyAdsr, DaisyPortamento): modules/synthlib/SynthVoice.h:16-316.
 *   Init(sr)                         SynthVoice.h:31-39   -> olb_voice_create / olb_voice_init
 *   UpdateMidiControl / UpdateHardwareControl  :101-229  -> olb_voice_update_midi / olb_voice_update_hw
 *   UpdateConfig(Voice::Config&)     :55-77; Voice.h:14-31 (16 floats, that order) -> olb_voice_update_config
 *   NoteOn / NoteOff / GateOn / GateOff / SetFrequency / Playing   :231-267 -> olb_voice_note_on / _note_off / _set_gate /
 *                                                                    _set_frequency / _playing
 *   Process(frame_out) per voice     :41-53               -> olb_voice_process with out_dev [frames][voices]
 *   Polyvoice::Process: *frame_out += every voice, caller zeroes (Polyvoice.h:28-33; stereo duplicate main.cpp:83)
 *                                                         -> olb_voice_process with bus_dev [frames][bus_channels]:
 *                                                            bus (+)= bus_gain * sum of voices (accumulate != 0 adds)
 * Setters take effect at the next process call (the reference applies MIDI between audio blocks).  trace_dev
 * (optional, tests) receives 4 floats per voice-frame: oscillator output, amp envelope, filter envelope, phase. */
typedef struct olb_voice olb_voice;

int olb_voice_create(int device, uint32_t voices, float sample_rate, olb_voice **out);
int olb_voice_destroy(olb_voice *h);
int olb_voice_init(olb_voice *h, float sample_rate);
int olb_voice_update_midi(olb_voice *h, uint32_t first, uint32_t count, uint8_t control, uint8_t value);
int olb_voice_update_hw(olb_voice *h, uint32_t first, uint32_t count, uint8_t control, const float *values_host, uint32_t stride);
int olb_voice_update_config(olb_voice *h, uint32_t first, uint32_t count, const float *config16_host);
int olb_voice_note_on(olb_voice *h, uint32_t first, uint32_t count, const uint8_t *notes_host, uint32_t stride, uint8_t velocity);
int olb_voice_note_off(olb_voice *h, uint32_t first, uint32_t count);
int olb_voice_set_gate(olb_voice *h, uint32_t first, uint32_t count, int gate);
int olb_voice_set_frequency(olb_voice *h, uint32_t first, uint32_t count, const float *freqs_host, uint32_t stride);
int olb_voice_playing(olb_voice *h, uint32_t first, uint32_t count, uint8_t *playing_host);
int olb_voice_process(olb_voice *h, float *out_dev, float *bus_dev, uint32_t bus_channels, int accumulate, float bus_gain,
                      float *trace_dev, uint64_t frames, void *stream);
/* Sample-accurate control.  Events sorted by frame are applied BEFORE the frame they name.
 * Events that address ONE voice (notes; CCs with voice >= 0) are applied ON THE DEVICE: the host runs the setter maths, the
 * addressed voice's parameter column after each event is shipped with the launch, and the voice reloads it at that frame --
 * one launch however many such events the block holds.  Only a CC that fans out to every voice cuts the block.  voice < 0: Polyvoice's allocation (NoteOn -> first voice not playing, NoteOff -> first voice playing
 * that note: Polyvoice.h:35-51; CCs fan out to every voice: Polyvoice.h:53-63); voice >= 0: that voice (Voice.h:33-57). */
#define OLB_EV_MIDI_CC 0  /* UpdateMidiControl(cc = a, value = (uint8) v) */
#define OLB_EV_HW_CC 1    /* UpdateHardwareControl(cc = a, value = v) */
#define OLB_EV_NOTE_ON 2  /* NoteOn(note = a, velocity = (uint8) v) */
#define OLB_EV_NOTE_OFF 3 /* NoteOff(note = a, velocity = (uint8) v) */
typedef struct olb_voice_event {
    uint32_t frame;
    uint32_t kind;
    uint32_t a;
    float v;
    int32_t voice;
} olb_voice_event;
int olb_voice_process_events(olb_voice *h, const olb_voice_event *events, uint32_t n_events, float *out_dev, float *bus_dev,
                             uint32_t bus_channels, int accumulate, float bus_gain, uint64_t frames, void *stream);
/* Same with HOST buffers: out_host (optional) [frames][voices]; bus_host (optional) [frames][bus_channels], ADDED to when
 * accumulate != 0 (Polyvoice::Process adds into *frame_out: Polyvoice.h:28-33). */
int olb_voice_process_host(olb_voice *h, float *out_host, float *bus_host, uint32_t bus_channels, int accumulate, float bus_gain,
                           uint64_t frames);

/* (e2) Cross-GPU voice bus ------------------------------------------------------------------------------
 * Polyvoice::Process adds every voice into one frame (modules/synthlib/Polyvoice.h:28-33; the caller duplicates it to
 * stereo, modules/ol_daisy/app/synth/main.cpp:83).  With the voices sharded over the GPUs of a box -- one process per GPU,
 * no other inter-GPU traffic on the hot path -- each rank's olb_voice_process leaves a PARTIAL bus in bus_dev, and one
 * in-place sum all-reduce (NCCL over NVLink / NVSwitch) makes it the global mixdown on every rank.  The partial bus is the
 * send buffer.  Lifecycle: rank 0 calls olb_comm_unique_id and hands the OLB_COMM_ID_BYTES to the other ranks by whatever
 * means the host has; every rank then calls olb_comm_init (collective).  Failures return OLB_E_NCCL.
 *   olb_voice_bus_allreduce        on the caller's stream, after what produced the bus there
 *   olb_voice_bus_allreduce_async  on the communicator's own stream once `producer_stream` has finished what is queued on it
 *                                  now: the reduce of block k overlaps the voices of block k + 1 (written to another bus
 *                                  buffer); `slot` (0..3) names the buffer
 *   olb_comm_wait                  `stream` waits for the last reduce queued for `slot` (before reading the reduced bus, and
 *                                  before the bank writes that buffer again)
 *   olb_voice_process_host_comm    olb_voice_process_host with the bus reduced over the ranks before it is copied back:
 *                                  the block form of Polyvoice::Process for a sharded bank (collective: every rank calls it) */
#define OLB_COMM_ID_BYTES 128
typedef struct olb_comm olb_comm;
int olb_comm_unique_id(void *id_out);
int olb_comm_init(int device, int rank, int world, const void *id_bytes, olb_comm **out);
int olb_comm_destroy(olb_comm *c);
int olb_comm_info(olb_comm *c, int *rank, int *world);
int olb_voice_bus_allreduce(olb_comm *c, float *bus_dev, uint64_t frames, uint32_t channels, void *stream);
int olb_voice_bus_allreduce_async(olb_comm *c, float *bus_dev, uint64_t frames, uint32_t channels, void *producer_stream, int slot);
int olb_comm_wait(olb_comm *c, int slot, void *stream);
int olb_comm_sync(olb_comm *c);
int olb_voice_process_host_comm(olb_voice *h, olb_comm *c, float *bus_host, uint32_t bus_channels, int accumulate, float bus_gain,
                                uint64_t frames);

/* (a7) Dattorro plate reverb bank -------------------------------------------------------------------
 * Replaces libs/dattorro-verb/verb.h:5-26 for `instances` plates at once (mono in, stereo wet out):
 *   DattorroVerb_create / initialize (verb.cpp:173-222, 225-232)  -> olb_dattorro_create, olb_dattorro_init
 *   DattorroVerb_delete (verb.cpp:235-251)                         -> olb_dattorro_destroy
 *   DattorroVerb_setPreDelay .. _setDamping (verb.cpp:136-170)     -> olb_dattorro_set_param(s)
 *   DattorroVerb_process + _getLeft + _getRight (verb.cpp:258-326) -> olb_dattorro_process
 * Every instance of a bank shares the reference's 16-bit sample counter `t` (they are created and ticked together).
 * Results are bit-exact against verb.cpp compiled with -ffp-contract=off. */
#define OLB_PLATE_PREDELAY 0         /* DattorroVerb_setPreDelay: fraction of 4 800 samples */
#define OLB_PLATE_PREFILTER 1        /* DattorroVerb_setPreFilter */
#define OLB_PLATE_INPUT_DIFFUSION1 2 /* DattorroVerb_setInputDiffusion1 */
#define OLB_PLATE_INPUT_DIFFUSION2 3 /* DattorroVerb_setInputDiffusion2 */
#define OLB_PLATE_DECAY 4            /* DattorroVerb_setDecay (also derives decay diffusion 2) */
#define OLB_PLATE_DECAY_DIFFUSION 5  /* DattorroVerb_setDecayDiffusion */
#define OLB_PLATE_DAMPING 6          /* DattorroVerb_setDamping */

typedef struct olb_dattorro olb_dattorro;

int olb_dattorro_create(int device, uint32_t instances, olb_dattorro **out);
int olb_dattorro_destroy(olb_dattorro *h);
/* initialize(): rings, filters, counter cleared; parameters back to the reference defaults (verb.cpp:215-221). */
int olb_dattorro_init(olb_dattorro *h);
/* One value for instances [first, first+count) / one value per instance (values[i * stride]).  Effective at the next
 * process call. */
int olb_dattorro_set_param(olb_dattorro *h, uint32_t first, uint32_t count, int param, float value);
int olb_dattorro_set_params(olb_dattorro *h, uint32_t first, uint32_t count, int param, const float *values_host, uint32_t stride);
/* For each frame: DattorroVerb_process(in); out_l = getLeft(); out_r = getRight().  Device buffers, all
 * [frames][instances] (instance fastest). */
int olb_dattorro_process(olb_dattorro *h, const float *in_dev, float *out_l_dev, float *out_r_dev, uint64_t frames, void *stream);
/* Diagnostics: which kernel the last process call launched ("k_dattorro_smem": rings resident in shared memory, blocks of
 * >= 8 192 frames; "k_dattorro_split": rings in HBM, any block length). */
const char *olb_dattorro_last_kernel(olb_dattorro *h);
/* Same with HOST buffers (all [frames][instances]); returns when the outputs are complete. */
int olb_dattorro_process_host(olb_dattorro *h, const float *in_host, float *out_l_host, float *out_r_host, uint64_t frames);

/* (a9) Chorus bank ------------------------------------------------------------------------------------
 * The reference has no chorus class; its chorus is the Max/RNBO graph modules/rnbo/patcher/mono-chorus.rnbopat
 * (gen~ pitchshift -> LFO-modulated delay~ -> lores~ -> dry/wet).  One bank = `instances` mono choruses with that patch's params
 * (all 0..1, patch defaults): rate (scale -> 0.01..0.5 Hz), depth (@min 0.08; scale -> 1..12 ms), cutoff
 * (scale -> 300..15 000 Hz), q, mix, phase (cycle~ phase offset), pitch.  stereo-chorus.rnbopat = two of these per stereo
 * instance, phase 1 on the left and 0 on the right, all other params shared: a bank of 2 x instances units over audio
 * [frames][instances][2] (ol::gpu::StereoChorusBank / ol_dsp_b200.StereoChorusBank set it up).  Lifecycle and call shape follow the fxlib units:
 * create = construct + Init(sample_rate); init clears the delay line, filter and LFO phase and keeps the params;
 * process is the block form of "for each frame: Process(in, out)", in_dev == out_dev allowed.  Arithmetic is
 * specified by oracle/olb_oracle.cpp (orc_chorus_*): parity unpinned, bit-exact against that restatement. */
#define OLB_CHORUS_RATE 0
#define OLB_CHORUS_DEPTH 1
#define OLB_CHORUS_CUTOFF 2
#define OLB_CHORUS_Q 3
#define OLB_CHORUS_MIX 4
#define OLB_CHORUS_PHASE 5
/* param pitch @min 0 @max 3 (mono-chorus.rnbopat obj-8): the frequency in Hz of the phasor of the gen~ pitch shifter in front
 * of delay~ (modules/rnbo/patcher/pitchshift.gendsp: two cosine-windowed taps of a 10 ms window).  Not a 0..1 param: 0..3. */
#define OLB_CHORUS_PITCH 6

typedef struct olb_chorus olb_chorus;

int olb_chorus_create(int device, uint32_t instances, float sample_rate, olb_chorus **out);
int olb_chorus_destroy(olb_chorus *h);
int olb_chorus_init(olb_chorus *h, float sample_rate);
int olb_chorus_set_param(olb_chorus *h, uint32_t first, uint32_t count, int param, float value);
int olb_chorus_set_params(olb_chorus *h, uint32_t first, uint32_t count, int param, const float *values_host, uint32_t stride);
int olb_chorus_process(olb_chorus *h, const float *in_dev, float *out_dev, uint64_t frames, void *stream);
int olb_chorus_process_host(olb_chorus *h, const float *in_host, float *out_host, uint64_t frames);
const char *olb_chorus_last_kernel(olb_chorus *h);

/* ================================================================================================
 * Sample playback bank: `instances` x ol::synth::Sample behind SampleSoundSource
 * (modules/synthlib/Sample.h:16-52, Sample.cpp:9-61, SampleSoundSource.h:14-41) over a DEVICE-RESIDENT sample pool, which
 * replaces the per-frame SampleDataSource::Read (SampleDataSource.h:11-25; the reference's only implementation decodes a
 * file through miniaudio one frame at a time, workouts/workout_buddy.h:18-42).  Every instance is one Sample object: play
 * state, OneShot / Loop, loop points, position.  As in the reference, Process leaves a frame UNTOUCHED when the instance
 * is not playing or is past the end of its sample: `out` is played INTO.  Bit-exact (the path copies samples).
 * All samples of a bank have the bank's channel count (1 or 2, interleaved); out is [frames][instances * channels]. */
typedef struct olb_sampler olb_sampler;
#define OLB_SAMPLE_ONESHOT 0 /* ol::synth::SamplePlayMode (Sample.h:11-14) */
#define OLB_SAMPLE_LOOP 1
int olb_sampler_create(int device, uint32_t instances, uint32_t channels, olb_sampler **out);
int olb_sampler_destroy(olb_sampler *h);
/* Copy one sample ([frames][channels] floats, host) into the pool; *sample_id names it. */
int olb_sampler_load_sample(olb_sampler *h, const float *frames_host, uint64_t frames, uint32_t *sample_id);
/* Bind instances to samples (the Sample constructor's data source): the instance becomes a fresh Sample object
 * (not playing, OneShot, loop points 0, position 0).  ids_host: one id, or one per instance when per_instance != 0. */
int olb_sampler_set_sample(olb_sampler *h, uint32_t first, uint32_t count, const uint32_t *ids_host, int per_instance);
int olb_sampler_init(olb_sampler *h, float sample_rate);                                             /* Sample::Init */
int olb_sampler_seek(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index);          /* Sample::Seek */
int olb_sampler_set_loop_start(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index);
int olb_sampler_set_loop_end(olb_sampler *h, uint32_t first, uint32_t count, uint64_t frame_index);
int olb_sampler_set_play_mode(olb_sampler *h, uint32_t first, uint32_t count, int mode);
int olb_sampler_play(olb_sampler *h, uint32_t first, uint32_t count);                                /* Sample::Play */
int olb_sampler_pause(olb_sampler *h, uint32_t first, uint32_t count);                               /* Sample::Pause */
int olb_sampler_toggle_play(olb_sampler *h, uint32_t first, uint32_t count);                         /* Sample::TogglePlay */
/* SampleSoundSource::GateOn (Seek(0) + Play) / GateOff (Pause), SampleSoundSource.h:21-26 */
int olb_sampler_gate(olb_sampler *h, uint32_t first, uint32_t count, int on);
/* Position (next frame to be read) and play flag of a range, after all work queued so far. */
int olb_sampler_position(olb_sampler *h, uint32_t first, uint32_t count, uint64_t *frame_index_host, uint8_t *playing_host);
int olb_sampler_process(olb_sampler *h, float *out_dev, uint64_t frames, void *stream);
/* Host buffer, played INTO: copied to the device, processed, copied back. */
int olb_sampler_process_host(olb_sampler *h, float *out_host, uint64_t frames);

#ifdef __cplusplus
}
#endif
#endif /* OLBANK_H */
