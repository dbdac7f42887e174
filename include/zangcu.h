/*
 * zangcu.h — C ABI of the B200-native batch synthesis engine behind zang's module contract.
 *
 * The reference (dbandstra/zang) has no FFI; its boundary is the comptime module contract
 *   paint(self, span, outputs, temps, note_id_changed, params)            (src/modules/SineOsc.zig:24-31)
 * applied to ONE voice and ONE span at a time.  This ABI is the batched form of that contract:
 *
 *   voice  -> a column of a [frame][voice] f32 device buffer (voice is the fastest axis)
 *   self   -> row v of a "bank" (one POD state struct per voice, same fields as the Zig struct)
 *   span + note_id_changed + params -> one zcu_span record per reference paint() call of that voice;
 *            `param` indexes a caller-supplied array of the module's Params struct
 *   outputs/temps -> zcu_buf handles; every paint ACCUMULATES (+=) exactly like the reference
 *
 * One call paints every span of every voice for a whole block of mix buffers.  All calls are
 * asynchronous on the context's CUDA stream, return an int error code (0 = ok) and never abort.
 * There is NO CPU fallback: without a CUDA device every entry point that would compute returns
 * ZCU_ERR_NO_DEVICE.
 *
 * Plain C types only (pointers, sizes, PODs); usable from Zig via @cImport, from Python via ctypes.
 */
#ifndef ZANGCU_H
#define ZANGCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZCU_ABI_VERSION 1

/* ---- error codes ------------------------------------------------------------------------- */
enum {
    ZCU_OK = 0,
    ZCU_ERR_INVALID = 1,     /* bad argument (null handle, size mismatch, unknown enum) */
    ZCU_ERR_NO_DEVICE = 2,   /* no CUDA device / driver: there is no CPU fallback */
    ZCU_ERR_CUDA = 3,        /* a CUDA runtime call failed; see zcu_last_error() */
    ZCU_ERR_OOM = 4,
    ZCU_ERR_UNSUPPORTED = 5,
    ZCU_ERR_NCCL = 6
};

/* ---- module kinds (src/modules.zig:1-13) -------------------------------------------------- */
enum {
    ZCU_MOD_SINE_OSC = 0,    /* src/modules/SineOsc.zig */
    ZCU_MOD_PULSE_OSC = 1,   /* src/modules/PulseOsc.zig */
    ZCU_MOD_TRISAW_OSC = 2,  /* src/modules/TriSawOsc.zig */
    ZCU_MOD_NOISE = 3,       /* src/modules/Noise.zig */
    ZCU_MOD_ENVELOPE = 4,    /* src/modules/Envelope.zig */
    ZCU_MOD_GATE = 5,        /* src/modules/Gate.zig */
    ZCU_MOD_FILTER = 6,      /* src/modules/Filter.zig */
    ZCU_MOD_PORTAMENTO = 7,  /* src/modules/Portamento.zig */
    ZCU_MOD_CURVE = 8,       /* src/modules/Curve.zig */
    ZCU_MOD_SAMPLER = 9,     /* src/modules/Sampler.zig */
    ZCU_MOD_DECIMATOR = 10,  /* src/modules/Decimator.zig */
    ZCU_MOD_DISTORTION = 11, /* src/modules/Distortion.zig */
    ZCU_MOD_CYCLE = 12,      /* src/modules/Cycle.zig */
    /* the delay line and the echo modules built on it (next-row scope) */
    ZCU_MOD_SIMPLE_DELAY = 13,    /* examples/modules.zig:341-388 over src/zang/delay.zig:7-91 */
    ZCU_MOD_FILTERED_ECHOES = 14, /* examples/modules.zig:392-460 */
    ZCU_MOD_FM_OSCILLATOR = 15,   /* examples/example_fmsynth.zig:26-89: sine with two-sample feedback */
    ZCU_MOD_COUNT = 16
};

/* ---- opaque handles ----------------------------------------------------------------------- */
typedef struct zcu_ctx zcu_ctx;       /* device + stream (+ optional NCCL communicator) */
typedef struct zcu_buf zcu_buf;       /* [frames][voices] f32 device buffer */
typedef struct zcu_bank zcu_bank;     /* per-voice module state, one struct per voice */
typedef struct zcu_spans zcu_spans;   /* device span table of one block */
typedef struct zcu_sample zcu_sample; /* PCM sample data for the Sampler (Sampler.zig:16-21) */

/* ---- span record: one reference paint() call of one voice ---------------------------------- */
/* zang.Span (src/zang/basics.zig:3-10) + note_id_changed + params of Trigger.NewPaintReturnValue
 * (src/zang/trigger.zig:50-54).  start/end are frame indices inside the block's device buffers;
 * a span never crosses a mix-buffer boundary because the reference paints buffer by buffer. */
typedef struct zcu_span {
    uint32_t start;
    uint32_t end;   /* exclusive; start < end */
    uint32_t flags; /* ZCU_SPAN_* */
    uint32_t param; /* row of the params array passed to the paint call */
} zcu_span;
#define ZCU_SPAN_NOTE_ID_CHANGED 1u

/* ---- per-voice state PODs (field-for-field the Zig structs; also the checkpoint format) ---- */
typedef struct { float t; } zcu_sine_state;                     /* SineOsc.zig:16 */
typedef struct { uint32_t cnt; } zcu_pulse_state;               /* PulseOsc.zig:36 */
typedef struct { uint32_t cnt; float t; } zcu_trisaw_state;     /* TriSawOsc.zig:36-37 */
typedef struct { uint64_t s[4]; } zcu_noise_state;              /* Noise.zig:22 (b[7] never persists, Noise.zig:68) */
typedef struct { float t, last_value, start; } zcu_painter_state; /* painter.zig:33-36 */
typedef struct { int32_t state; zcu_painter_state painter; } zcu_envelope_state; /* Envelope.zig:15-24 */
typedef struct { float l, b; } zcu_filter_state;                /* Filter.zig:34-35 */
typedef struct { zcu_painter_state painter; } zcu_portamento_state; /* Portamento.zig:13 */
typedef struct {                                                /* Curve.zig:36-41 */
    float t;
    uint32_t current_song_note;
    int32_t current_song_note_offset;
    uint32_t next_song_note;
} zcu_curve_state;
typedef struct { float t; } zcu_sampler_state;                  /* Sampler.zig:69 */
typedef struct { float dval, dcount; } zcu_decimator_state;     /* Decimator.zig:11-12 */
typedef struct { float t; } zcu_cycle_state;                    /* Cycle.zig:14 */
/* zang.Delay (delay.zig:9-10): the index; the delay_buffer itself lives in the bank ([delay_samples][voices] f32) */
typedef struct { uint32_t index; } zcu_delay_state;
typedef struct { float t, feedback1, feedback2; } zcu_fm_oscillator_state; /* example_fmsynth.zig:37-39 */
typedef struct { uint32_t index; float l, b; } zcu_echoes_state; /* FilteredEchoes: Delay + Filter (modules.zig:403-404) */

enum { ZCU_ENV_IDLE = 0, ZCU_ENV_ATTACK = 1, ZCU_ENV_DECAY = 2, ZCU_ENV_SUSTAIN = 3, ZCU_ENV_RELEASE = 4 };

/* ---- Params PODs (one row per distinct params value; spans refer to rows) ------------------ */
/* ConstantOrBuffer (src/zang/constant_or_buffer.zig:4-7): the scalar lives in the params row and is
 * used when the corresponding buffer argument of the paint call is NULL (.constant); a non-NULL
 * zcu_buf selects the .buffer variant for every span of the call. */
enum { ZCU_CURVE_INSTANTANEOUS = 0, ZCU_CURVE_LINEAR = 1, ZCU_CURVE_SQUARED = 2, ZCU_CURVE_CUBED = 3 };
typedef struct { int32_t kind; float duration; } zcu_painter_curve;       /* painter.zig:25-30 */

typedef struct { float sample_rate, freq, phase; } zcu_sine_params;     /* SineOsc.zig:10-14 */
typedef struct { float sample_rate, freq, color; } zcu_pulse_params;    /* PulseOsc.zig:30-34 */
typedef zcu_pulse_params zcu_trisaw_params;                             /* TriSawOsc.zig:30-34 */
enum { ZCU_NOISE_WHITE = 0, ZCU_NOISE_PINK = 1 };
typedef struct { int32_t color; } zcu_noise_params;                     /* Noise.zig:18-20 */
typedef struct {                                                        /* Envelope.zig:6-13 */
    float sample_rate;
    zcu_painter_curve attack, decay, release;
    float sustain_volume;
    int32_t note_on;
} zcu_envelope_params;
typedef struct { int32_t note_on; } zcu_gate_params;                    /* Gate.zig:7-9 */
enum { ZCU_FILTER_BYPASS = 0, ZCU_FILTER_LOW_PASS = 1, ZCU_FILTER_BAND_PASS = 2,
       ZCU_FILTER_HIGH_PASS = 3, ZCU_FILTER_NOTCH = 4, ZCU_FILTER_ALL_PASS = 5 };
typedef struct { int32_t type; float cutoff, res; } zcu_filter_params;  /* Filter.zig:27-32 */
typedef struct {                                                        /* Portamento.zig:5-11 */
    float sample_rate;
    zcu_painter_curve curve;
    float goal;
    int32_t note_on, prev_note_on;
} zcu_portamento_params;
enum { ZCU_INTERP_LINEAR = 0, ZCU_INTERP_SMOOTHSTEP = 1 };
typedef struct { float value, t; } zcu_curve_node;                      /* src/zang/curve.zig:3-6 */
typedef struct {                                                        /* Curve.zig:29-33 */
    float sample_rate;
    int32_t function;
    uint32_t node_start, node_count; /* slice of the nodes array passed to zcu_paint_curve */
} zcu_curve_params;
enum { ZCU_FMT_UNSIGNED8 = 0, ZCU_FMT_SIGNED16_LSB = 1, ZCU_FMT_SIGNED24_LSB = 2, ZCU_FMT_SIGNED32_LSB = 3 };
typedef struct { float sample_rate; uint32_t channel; int32_t loop; } zcu_sampler_params; /* Sampler.zig:62-67 */
typedef struct { float sample_rate, fake_sample_rate; } zcu_decimator_params;             /* Decimator.zig:5-9 */
enum { ZCU_DIST_OVERDRIVE = 0, ZCU_DIST_CLIP = 1 };
typedef struct { int32_t type; float ingain, outgain, offset; } zcu_distortion_params;   /* Distortion.zig:13-19 */
typedef struct { float sample_rate, speed; } zcu_cycle_params;                            /* Cycle.zig:9-12 */
typedef struct { float sample_rate; uint32_t waveform; float feedback; } zcu_fm_oscillator_params; /* example_fmsynth.zig:29-35 (freq, phase = buffer args) */
typedef struct { float feedback_volume, cutoff; } zcu_echoes_params;                      /* modules.zig:396-400 (input = buffer arg) */

/* ---- library / context --------------------------------------------------------------------- */
int zcu_abi_version(void);
int zcu_device_count(void); /* 0 when no CUDA device is usable */
int zcu_ctx_create(int device, zcu_ctx** out);
/* same, but every call is enqueued on the caller's stream (a cudaStream_t of `device`, e.g. the host
 * application's audio/render stream); the stream is borrowed, not owned */
int zcu_ctx_create_on_stream(int device, void* cuda_stream, zcu_ctx** out);
void zcu_ctx_destroy(zcu_ctx* ctx);
int zcu_ctx_sync(zcu_ctx* ctx);
const char* zcu_last_error(zcu_ctx* ctx); /* never NULL */
void* zcu_ctx_stream(zcu_ctx* ctx);       /* the cudaStream_t every call is enqueued on */
uint64_t zcu_ctx_launch_count(zcu_ctx* ctx); /* kernels launched by this library on this ctx so far */

/* Diagnostic: the branch-free division by a span constant used by the buffer-frequency TriSawOsc (freq[i] / sample_rate,
 * TriSawOsc.zig:152) against the IEEE quotient, for EVERY f32 numerator.  mismatches must be 0; claimed = numerators for
 * which the fast sequence is used (the rest take the real division). */
int zcu_selftest_div_by_const(zcu_ctx* ctx, float divisor, uint64_t* mismatches, uint64_t* claimed);
/* The same for numerator / utof23(ifreq) of the buffer-frequency PulseOsc (PulseOsc.zig:138): every one of the 2^23 - 1
 * denominators the phase increment can produce. */
int zcu_selftest_div_numerator(zcu_ctx* ctx, float numerator, uint64_t* mismatches, uint64_t* claimed);

/* Arithmetic mode of the module paints (zcu_paint_*).
 * ZCU_PAINT_EXACT (default): every module walks each voice's frames in the reference's order; all state-carrying
 *   modules are bit-identical to the sequential paint() loops.
 * ZCU_PAINT_TOLERANCE: modules whose state forgets its start (Filter, src/modules/Filter.zig:124-147) are also cut
 *   into time segments; a segment starts from a warm-up over the preceding frames whose residual influence is verified
 *   on the device (homogeneous map of the window <= 1e-8), else from the previous segment's end state.  Output within
 *   1e-5 of full scale of the exact mode; every other module is unaffected. */
enum { ZCU_PAINT_EXACT = 0, ZCU_PAINT_TOLERANCE = 1 };
int zcu_ctx_set_paint_mode(zcu_ctx* ctx, int mode);
int zcu_ctx_paint_mode(const zcu_ctx* ctx);
/* Output mode of the module paints (zcu_paint_*), per context.  ZCU_OUT_ACCUMULATE (default) is the reference's contract:
 * paint() adds to what the output holds.  ZCU_OUT_OVERWRITE makes one call stand for the reference's ubiquitous pair
 * `zang.zero(span, out); module.paint(span, {out}, ...)` (examples/modules.zig:220-246): inside its spans the output is
 * 0 + what the module adds, bit for bit what the pair leaves, and the output is never read (a generator then moves 4
 * instead of 4 + 8 bytes per voice-sample).  Frames outside the spans keep their contents in both modes. */
enum { ZCU_OUT_ACCUMULATE = 0, ZCU_OUT_OVERWRITE = 1 };
int zcu_ctx_set_paint_output(zcu_ctx* ctx, int mode);
int zcu_ctx_paint_output(const zcu_ctx* ctx);
/* Noise (src/modules/Noise.zig) is painted in time segments that start from an exact jump of the xoshiro256++ state by the
 * number of samples painted before the segment.  Zig's Random.float(f32) draws a second u64 with probability 2^-41 per
 * sample; a jump cannot foresee that, and the voice's later segments of that one call are then one draw behind the
 * sequential loop.  This counter (since context creation; synchronises the stream) tells how often it happened: 0 means
 * every Noise paint so far is bit-identical to the reference loop.  Expected once per 2.2e12 noise samples. */
int zcu_ctx_noise_resync_count(zcu_ctx* ctx, uint64_t* out);
/* ZCU_PAINT_TOLERANCE diagnostics since the last call: out[0] CTAs (128 voices x time segment) that started from a warm-up
 * window, out[1] the sum and out[2] the maximum of those windows in frames, out[3] CTAs whose verification failed and which
 * waited for their predecessor segment's end state instead.  Synchronises; resets the counters. */
int zcu_ctx_paint_warm_stats(zcu_ctx*, uint32_t out[4]);
/* Host-only (no device): the same jump applied on the CPU, state <- T^n state of xoshiro256 (n < 2^24), for checking the
 * matrices against n plain steps of the generator. */
int zcu_xoshiro_jump_host(uint64_t state[4], uint32_t n);

/* device-side timing (CUDA events on the context stream) */
int zcu_timer_start(zcu_ctx* ctx);
int zcu_timer_stop_ms(zcu_ctx* ctx, float* ms); /* records, synchronises, returns elapsed */

/* ---- buffers: the batched form of the caller-owned []f32 (README.md:4-6) -------------------- */
int zcu_buf_alloc(zcu_ctx* ctx, uint32_t frames, uint32_t voices, zcu_buf** out);
int zcu_buf_wrap(zcu_ctx* ctx, void* device_ptr, uint32_t frames, uint32_t voices, zcu_buf** out);
void zcu_buf_free(zcu_buf* buf);
int zcu_buf_upload(zcu_ctx* ctx, zcu_buf* buf, const float* host);   /* host is [frames][voices] */
int zcu_buf_download(zcu_ctx* ctx, const zcu_buf* buf, float* host); /* synchronises the stream */
void* zcu_buf_device_ptr(const zcu_buf* buf);
uint32_t zcu_buf_frames(const zcu_buf* buf);
uint32_t zcu_buf_voices(const zcu_buf* buf);

/* ---- banks: `init()` of the Zig module for every voice -------------------------------------- */
/* Noise: voice v is seeded with first_seed + v, mirroring the global counter of Noise.zig:9,26. */
int zcu_bank_create(zcu_ctx* ctx, int module_kind, uint32_t voices, uint32_t first_seed, zcu_bank** out);
void zcu_bank_free(zcu_bank* bank);
int zcu_bank_reset(zcu_ctx* ctx, zcu_bank* bank, uint32_t first_seed);
size_t zcu_bank_state_size(int module_kind); /* sizeof the per-voice state POD (0 = stateless) */
int zcu_bank_get_state(zcu_ctx* ctx, const zcu_bank* bank, void* host, size_t bytes); /* checkpoint */
int zcu_bank_set_state(zcu_ctx* ctx, zcu_bank* bank, const void* host, size_t bytes); /* resume */

/* Banks of the delay modules (ZCU_MOD_SIMPLE_DELAY, ZCU_MOD_FILTERED_ECHOES): `delay_samples` is the comptime parameter
 * of zang.Delay (delay.zig:7); the bank owns one zeroed delay line per voice (delay.zig:12-17), [delay_samples][voices].
 * zcu_bank_reset zeroes it again (delay.zig:19-22).  get/set_delay_line: checkpoint / resume of the lines. */
int zcu_delay_bank_create(zcu_ctx* ctx, int module_kind, uint32_t voices, uint32_t delay_samples, zcu_bank** out);
uint32_t zcu_bank_delay_samples(const zcu_bank* bank);
int zcu_bank_get_delay_line(zcu_ctx* ctx, const zcu_bank* bank, float* host, size_t bytes);
int zcu_bank_set_delay_line(zcu_ctx* ctx, zcu_bank* bank, const float* host, size_t bytes);

/* ---- span tables ----------------------------------------------------------------------------- */
/* voice_off has voices+1 entries; spans of voice v are spans[voice_off[v] .. voice_off[v+1]), sorted
 * by start, non-overlapping, each inside [0, frames). */
int zcu_spans_upload(zcu_ctx* ctx, uint32_t voices, uint32_t frames, const uint32_t* voice_off,
                     const zcu_span* spans, zcu_spans** out);
/* every voice paints [0,buffer_frames) of each of num_buffers consecutive buffers with param row 0
 * (the un-triggered call pattern, e.g. examples/example_polyphony.zig:147-160) */
int zcu_spans_full(zcu_ctx* ctx, uint32_t voices, uint32_t num_buffers, uint32_t buffer_frames,
                   zcu_spans** out);
void zcu_spans_free(zcu_spans* spans);

/* ---- zang basics (src/zang/basics.zig:12-78) -------------------------------------------------- */
/* spans == NULL: whole buffer, fully vectorised.  spans != NULL: only inside each voice's spans. */
int zcu_zero(zcu_ctx*, const zcu_spans*, zcu_buf* dest);
int zcu_set(zcu_ctx*, const zcu_spans*, zcu_buf* dest, float a);
int zcu_copy(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* src);
int zcu_add(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b);
int zcu_add_into(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* src);
int zcu_add_scalar(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, float b);
int zcu_add_scalar_into(zcu_ctx*, const zcu_spans*, zcu_buf* dest, float a);
int zcu_multiply(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b);
int zcu_multiply_with(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a);
int zcu_multiply_scalar(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, float b);
int zcu_multiply_with_scalar(zcu_ctx*, const zcu_spans*, zcu_buf* dest, float a);
/* The same scalar ops when the reference passes a value derived from the span's params, e.g.
 * zang.set(span, temps[0], params.freq * params.ratio.constant) (examples/modules.zig:39-47): the scalar of a
 * span is scalars[span.param].  scalars is a host array of nscalars rows, copied by the call; spans != NULL. */
int zcu_set_rows(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const float* scalars, uint32_t nscalars);
int zcu_add_scalar_rows(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, const float* scalars, uint32_t nscalars);
int zcu_add_scalar_into_rows(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const float* scalars, uint32_t nscalars);
int zcu_multiply_scalar_rows(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const zcu_buf* a, const float* scalars, uint32_t nscalars);
int zcu_multiply_with_scalar_rows(zcu_ctx*, const zcu_spans*, zcu_buf* dest, const float* scalars, uint32_t nscalars);

/* ---- module paints (each replaces <Module>.paint of src/modules/<Module>.zig) ---------------- */
/* params: host array of nparams rows, copied to the device by the call.  Buffer arguments that model
 * ConstantOrBuffer may be NULL (= .constant from the params row). */
int zcu_paint_sine_osc(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_sine_params*, uint32_t nparams,
                       zcu_buf* out, const zcu_buf* freq, const zcu_buf* phase);
int zcu_paint_pulse_osc(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_pulse_params*, uint32_t nparams,
                        zcu_buf* out, const zcu_buf* freq);
int zcu_paint_trisaw_osc(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_trisaw_params*, uint32_t nparams,
                         zcu_buf* out, const zcu_buf* freq);
int zcu_paint_noise(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_noise_params*, uint32_t nparams,
                    zcu_buf* out);
int zcu_paint_envelope(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_envelope_params*, uint32_t nparams,
                       zcu_buf* out);
int zcu_paint_gate(zcu_ctx*, const zcu_spans*, const zcu_gate_params*, uint32_t nparams, zcu_buf* out);
int zcu_paint_filter(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_filter_params*, uint32_t nparams,
                     zcu_buf* out, const zcu_buf* input, const zcu_buf* cutoff, const zcu_buf* res);
int zcu_paint_portamento(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_portamento_params*,
                         uint32_t nparams, zcu_buf* out);
int zcu_paint_curve(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_curve_params*, uint32_t nparams,
                    const zcu_curve_node* nodes, uint32_t nnodes, zcu_buf* out);
int zcu_paint_sampler(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_sampler_params*, uint32_t nparams,
                      const zcu_sample* sample, zcu_buf* out);
int zcu_paint_decimator(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_decimator_params*,
                        uint32_t nparams, zcu_buf* out, const zcu_buf* input);
int zcu_paint_distortion(zcu_ctx*, const zcu_spans*, const zcu_distortion_params*, uint32_t nparams,
                         zcu_buf* out, const zcu_buf* input);
int zcu_paint_cycle(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_cycle_params*, uint32_t nparams,
                    zcu_buf* out, const zcu_buf* speed);

/* SimpleDelay.paint (examples/modules.zig:362-386): out += the input of delay_samples painted frames ago.  Exact, sequential
 * in time (one segment); the delay line is read and written in place, per voice, in paint order. */
int zcu_paint_simple_delay(zcu_ctx*, zcu_bank*, const zcu_spans*, zcu_buf* out, const zcu_buf* input);
/* FilteredEchoes.paint (examples/modules.zig:409-459): y = lowpass(delayed * feedback_volume + input), out += y, y goes
 * back into the delay line.  The reference's temps never exist: one pass, one read of input, one read-modify-write of out. */
int zcu_paint_filtered_echoes(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_echoes_params*, uint32_t nparams,
                              zcu_buf* out, const zcu_buf* input);

/* fmsynth Oscillator.paint (examples/example_fmsynth.zig:49-88): out += wave(sin((t + phase) * pi * 2 + (fb1 + fb2) * feedback)),
 * t += freq[i] / sample_rate, the last two output samples fed back.  A nonlinear recurrence: one thread per voice, sequential
 * in time.  freq is required (a slice in the reference), phase may be NULL (= 0).  waveform 0..3 as :71-76. */
int zcu_paint_fm_oscillator(zcu_ctx*, zcu_bank*, const zcu_spans*, const zcu_fm_oscillator_params*, uint32_t nparams,
                            zcu_buf* out, const zcu_buf* freq, const zcu_buf* phase);

/* Filter.cutoffFromFrequency (Filter.zig:20-23): host scalar, f32 arithmetic */
float zcu_filter_cutoff_from_frequency(float frequency, float sample_rate);

/* Sampler.Sample (Sampler.zig:16-21): data is copied to the device */
int zcu_sample_create(zcu_ctx*, const uint8_t* data, size_t len, uint32_t num_channels,
                      uint32_t sample_rate, int format, zcu_sample** out);
void zcu_sample_free(zcu_sample*);

/* ---- voice mix: what the shared outputs[0] accumulation does (example_polyphony.zig:80-91) ---- */
/* dst[f][c] += sum_v gains[v*channels + c] * src[f][v];  dst is a [frames][channels] buffer.
 * gains == NULL means 1.0 for every voice/channel. */
int zcu_mix_reduce(zcu_ctx*, const zcu_buf* src, const float* gains_host, uint32_t channels, zcu_buf* dst);

/* The reference's own accumulation order: voices paint one after the other into the shared outputs[0]
 * (examples/example_song.zig:382-396), i.e. dst[f][c] = (((dst[f][c] + g0*src[f][0]) + g1*src[f][1]) + ...) in
 * voice order, multiply and add rounded separately.  Bit-exact with the sequential host loop; use it where the
 * summation order matters (PCM output), zcu_mix_reduce where it does not. */
int zcu_mix_ordered(zcu_ctx*, const zcu_buf* src, const float* gains_host, uint32_t channels, zcu_buf* dst);

/* ---- zang.mixDown (src/zang/mixdown.zig:8-86): f32 -> s16le / s8 with clamping ---------------- */
enum { ZCU_AUDIO_SIGNED8 = 0, ZCU_AUDIO_SIGNED16_LSB = 1 };
int zcu_mix_down(zcu_ctx*, const zcu_buf* mix /* [frames][channels] */, int audio_format, float vol,
                 uint8_t* dst_host, size_t dst_len);

/* No counterpart in the reference: added at the request of the run's operators for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as zcu_mix_down, but the
 * device staging buffer for the PCM bytes is a fixed array inside the library image instead of the context's arena: the
 * call allocates nothing (no cudaMalloc / cudaMallocAsync / cudaFree).  dst_len beyond ZCU_MIX_DOWN_STATIC_MAX_BYTES is
 * rejected with ZCU_ERR_INVALID (convert in pieces, or use zcu_mix_down).  The array is shared by every call on a device and
 * keeps the last call's bytes (nothing needs clearing: a call writes every byte it copies out); calls that overlap on the
 * GPU (contexts on different streams of one device) race on it.  Every GPU that loads the library holds the array until
 * exit. */
#define ZCU_MIX_DOWN_STATIC_MAX_BYTES (4 * 1024 * 1024)
int zcu_mix_down_static(zcu_ctx*, const zcu_buf* mix /* [frames][channels] */, int audio_format, float vol,
                        uint8_t* dst_host, size_t dst_len);

/* ---- host-side event layer (bit-exact index logic of src/zang/notes.zig, src/zang/trigger.zig) - */
typedef struct zcu_impulse {  /* notes.zig:58-62 */
    uint32_t frame;    /* frame inside the current mix buffer */
    uint32_t note_id;
    uint32_t event_id;
} zcu_impulse;

/* Trigger(P) (trigger.zig:26-198).  P is represented by a param row index carried with each impulse. */
typedef struct zcu_trigger { int32_t has_note; uint32_t note_id; uint32_t param; } zcu_trigger;
void zcu_trigger_init(zcu_trigger* t);   /* trigger.zig:56-60 */
void zcu_trigger_reset(zcu_trigger* t);  /* trigger.zig:62-64 */
/* Runs Trigger.counter + the whole `while (trigger.next(&ctr))` loop for the span [start,end).
 * impulse_params[i] is the param row of impulses[i].  Writes at most max_out spans (frame indices are
 * offset by frame_base, so a block of buffers can share one table) and returns how many it produced. */
uint32_t zcu_trigger_run(zcu_trigger* t, uint32_t start, uint32_t end, const zcu_impulse* impulses,
                         const uint32_t* impulse_params, uint32_t n, uint32_t frame_base, zcu_span* out,
                         uint32_t max_out);

/* ImpulseQueue (notes.zig:72-128): capacity 32, drops when full or pushed out of order. */
typedef struct zcu_impulse_queue {
    zcu_impulse impulses[32];
    uint32_t params[32];
    uint32_t length;
    uint32_t next_event_id;
} zcu_impulse_queue;
void zcu_impulse_queue_init(zcu_impulse_queue* q);
void zcu_impulse_queue_push(zcu_impulse_queue* q, uint32_t impulse_frame, uint32_t note_id, uint32_t param);
/* returns the number of queued impulses (caller reads q->impulses / q->params) and empties the queue */
uint32_t zcu_impulse_queue_consume(zcu_impulse_queue* q, zcu_impulse* out_impulses, uint32_t* out_params);

/* IdGenerator (notes.zig:43-56) */
typedef struct zcu_id_generator { uint32_t next_id; } zcu_id_generator;
void zcu_id_generator_init(zcu_id_generator* g);
uint32_t zcu_id_generator_next(zcu_id_generator* g);

/* NoteTracker (notes.zig:138-207).  song events: time in seconds, note id, param row. */
typedef struct zcu_song_event { float t; uint32_t note_id; uint32_t param; } zcu_song_event;
typedef struct zcu_note_tracker { const zcu_song_event* song; uint32_t song_len; uint32_t next_song_event; float t; } zcu_note_tracker;
void zcu_note_tracker_init(zcu_note_tracker* nt, const zcu_song_event* song, uint32_t song_len);
void zcu_note_tracker_reset(zcu_note_tracker* nt);
/* returns the number of impulses written (at most max_out; the reference has no bound check at 32,
 * notes.zig:184-187 — events beyond max_out are consumed but dropped) */
uint32_t zcu_note_tracker_consume(zcu_note_tracker* nt, float sample_rate, uint32_t start, uint32_t end,
                                  zcu_impulse* out_impulses, uint32_t* out_params, uint32_t max_out);

/* PolyphonyDispatcher(n) (notes.zig:209-349), n <= ZCU_MAX_POLYPHONY */
#define ZCU_MAX_POLYPHONY 32
typedef struct zcu_poly_slot { int32_t used; uint32_t note_id; uint32_t event_id; int32_t note_on; } zcu_poly_slot;
typedef struct zcu_poly_dispatcher { uint32_t polyphony; zcu_poly_slot slots[ZCU_MAX_POLYPHONY]; } zcu_poly_dispatcher;
int zcu_poly_dispatcher_init(zcu_poly_dispatcher* pd, uint32_t polyphony);
void zcu_poly_dispatcher_reset(zcu_poly_dispatcher* pd);
/* note_on[i] is the note_on field of the params of impulse i.  out_* are [polyphony][32] arrays;
 * out_counts[polyphony]. */
int zcu_poly_dispatcher_dispatch(zcu_poly_dispatcher* pd, const zcu_impulse* impulses, const uint32_t* params,
                                 const int32_t* note_on, uint32_t n, zcu_impulse* out_impulses,
                                 uint32_t* out_params, uint32_t* out_counts);

/* Painter's ramp clock in closed form (painter.zig:102-106): applies `t += t_step; if (t >= 1) { t = 1; finished }`
 * up to n times with the reference's exact f32 rounding, in O(binades) instead of O(n).
 * Returns the number of adds performed (< n only when the ramp finished). */
uint32_t zcu_ramp_advance(float* t, float t_step, uint32_t n, int32_t* finished);
/* The same closed form for the un-clamped f32 accumulators: n times `t += step` (SineOsc.zig:49-62, Sampler.zig:130) and
 * n times `t += step; t -= trunc(t)` (Cycle.zig:44-46).  These are what lets a paint call start in the middle of a span. */
void zcu_accum_advance(float* t, float step, uint32_t n);
void zcu_cycle_advance(float* t, float step, uint32_t n);

/* Voice(T) of examples/example_song.zig:287-350 without the module paint: a NoteTracker feeding a
 * PolyphonyDispatcher feeding one Trigger per sub-voice. */
typedef struct zcu_poly_voice {
    zcu_note_tracker tracker;
    zcu_poly_dispatcher dispatcher;
    zcu_trigger triggers[ZCU_MAX_POLYPHONY];
} zcu_poly_voice;
int zcu_poly_voice_init(zcu_poly_voice* pv, const zcu_song_event* song, uint32_t song_len, uint32_t polyphony);
void zcu_poly_voice_reset(zcu_poly_voice* pv); /* example_song.zig:316-322 */
/* Runs Voice.paint's event logic (example_song.zig:331-347) for num_buffers consecutive mix buffers of buffer_frames
 * frames (the last one last_buffer_frames long) and writes the block's span table in the CSR layout of
 * zcu_spans_upload: sub-voice i is voice column i, frames are block relative, span.param is the song event's
 * `param`.  param_note_on[param] is the note_on flag of that params row (the dispatcher reads params.note_on,
 * notes.zig:318-323).  Returns the number of spans; if it exceeds max_spans nothing is written, *pv is left
 * untouched and the required count is returned negated. */
int64_t zcu_poly_voice_spans(zcu_poly_voice* pv, const int32_t* param_note_on, float sample_rate, uint32_t buffer_frames,
                             uint32_t num_buffers, uint32_t last_buffer_frames, uint32_t* voice_off_out, zcu_span* spans_out,
                             uint64_t max_spans);

/* The same on the DEVICE for a bank of independent polyphonic voices ("songs"): song s owns
 * events[song_off[s] .. song_off[s+1]) (chronological; event.param = a row of param_note_on, or every impulse is a note-on
 * when that is NULL), all with one polyphony.  zcu_poly_bank_spans replays NoteTracker.consume -> PolyphonyDispatcher.dispatch
 * -> one Trigger per slot for the NEXT num_buffers mix buffers of every song (the trackers keep their time, like
 * zcu_poly_voice_spans) and leaves a device span table with songs * polyphony voice columns (column = song * polyphony + slot,
 * span.param = the event's params row).  Bit-identical to zcu_poly_voice_spans per song. */
typedef struct zcu_poly_bank zcu_poly_bank;
int zcu_poly_bank_create(zcu_ctx*, uint32_t songs, uint32_t polyphony, const uint32_t* song_off_host, const zcu_song_event* events_host,
                         const int32_t* param_note_on_host, uint32_t nparams, float sample_rate, uint32_t buffer_frames, zcu_poly_bank** out);
void zcu_poly_bank_free(zcu_poly_bank*);
int zcu_poly_bank_reset(zcu_ctx*, zcu_poly_bank*);
int zcu_poly_bank_get_state(zcu_ctx*, const zcu_poly_bank*, zcu_poly_voice* state_host); /* [songs]; tracker.song is NULL */
int zcu_poly_bank_spans(zcu_ctx*, zcu_poly_bank*, uint32_t num_buffers, uint32_t last_buffer_frames, zcu_spans** out);

/* A bank of monophonic voices, each with its own ImpulseQueue + Trigger (examples/example_polyphony.zig:61-92 per voice).
 * Voice v owns events[event_off[v] .. event_off[v+1]) (absolute frames, chronological; zcu_note_event below).  Every mix
 * buffer's events are pushed into the voice's queue (capacity 32, notes.zig:108-118) and the Trigger splits the buffer.
 * Writes the span table of the block [first_frame, first_frame + num_frames) in the CSR layout of zcu_spans_upload:
 * frames block relative, span.param = GLOBAL event index.  first_frame must be a multiple of buffer_frames.  Returns the
 * span count; if it exceeds max_spans nothing is written, state is untouched and the count is returned negated. */
struct zcu_note_event;
typedef struct zcu_bank_trigger { zcu_trigger trigger; uint32_t next_event; } zcu_bank_trigger;
void zcu_bank_trigger_init(zcu_bank_trigger* state, uint32_t voices, const uint32_t* event_off);
int64_t zcu_bank_spans(zcu_bank_trigger* state, uint32_t voices, const uint32_t* event_off, const struct zcu_note_event* events,
                       uint32_t buffer_frames, uint32_t total_frames, uint32_t first_frame, uint32_t num_frames,
                       uint32_t* voice_off_out, zcu_span* spans_out, uint64_t max_spans);

/* The same on the DEVICE: the note events are uploaded once, and every block's span table is built by
 * the GPU (one thread per voice replays ImpulseQueue.push + the Trigger loop, trigger.zig:80-196, twice: count, scan,
 * write), so no span table crosses PCIe and the host does no per-block event work.  Bit-identical to zcu_bank_spans. */
typedef struct zcu_event_bank zcu_event_bank;
int zcu_event_bank_create(zcu_ctx*, uint32_t voices, const uint32_t* event_off_host, const struct zcu_note_event* events_host,
                          uint32_t buffer_frames, uint32_t total_frames, zcu_event_bank** out);
void zcu_event_bank_free(zcu_event_bank*);
int zcu_event_bank_reset(zcu_ctx*, zcu_event_bank*); /* zcu_bank_trigger_init */
int zcu_event_bank_get_state(zcu_ctx*, const zcu_event_bank*, zcu_bank_trigger* state_host); /* checkpoint */
int zcu_event_bank_set_state(zcu_ctx*, zcu_event_bank*, const zcu_bank_trigger* state_host);
/* span table of the block [first_frame, first_frame + num_frames), asynchronous on the context's stream; the Trigger state
 * advances.  The table is sized for the worst case; zcu_spans_download reads the exact count back. */
int zcu_event_bank_spans(zcu_ctx*, zcu_event_bank*, uint32_t first_frame, uint32_t num_frames, zcu_spans** out);
/* reads a device span table back (tests, checkpoints): *nspans_out = rows; rows are copied when they fit max_spans */
int zcu_spans_download(zcu_ctx*, const zcu_spans*, uint32_t* voice_off_out, zcu_span* spans_out, uint64_t max_spans, uint64_t* nspans_out);
/* Views of a Trigger table for an instrument that keeps the previous call's note_on (examples/example_portamento.zig:51
 * `defer self.prev_note_on = params.note_on`), per voice in span order, carried across blocks inside the event bank:
 *   porta: span.param = 2 * param + prev_note_on     env: span.flags = (!prev_note_on && note_on)     plain: span.param = 0 */
int zcu_spans_prev_note_on_views(zcu_ctx*, zcu_event_bank*, const zcu_spans* src, zcu_spans** porta_out, zcu_spans** env_out,
                                 zcu_spans** plain_out);

/* ---- fused instrument banks (B200-first path for the benchmark configs) ----------------------- */
/* A note event of one voice with an ABSOLUTE frame position in the whole render.  Voice v owns
 * events[event_off[v] .. event_off[v+1]), chronological.  The device replays Trigger semantics
 * per mix buffer (trigger.zig:80-196) itself, so no span table crosses PCIe. */
typedef struct zcu_note_event {
    uint32_t frame;   /* absolute frame in the render */
    uint32_t note_id;
    float freq;
    int32_t note_on;
} zcu_note_event;

typedef struct zcu_render zcu_render;

enum {
    ZCU_INSTR_NICE = 0,          /* examples/modules.zig:189-248 NiceInstrument */
    ZCU_INSTR_COUNT = 1
};

typedef struct zcu_render_config {
    int32_t instrument;       /* ZCU_INSTR_* */
    uint32_t voices;
    uint32_t buffer_frames;   /* the reference's AUDIO_BUFFER_SIZE, 1024 */
    uint32_t total_frames;    /* length of the whole render */
    uint32_t block_buffers;   /* mix buffers rendered per kernel launch (0 = default) */
    uint32_t channels;        /* 1 or 2: channels of the mix */
    float sample_rate;
    uint32_t first_seed;      /* Noise seed of voice 0 */
    int32_t mode;             /* ZCU_RENDER_* */
} zcu_render_config;
/* EXACT: every voice sequential in time; per-voice output bit-exact.
 * SCAN: one warp task per (32 voices, mix buffer); PulseOsc/Envelope start every buffer from exact closed forms, the
 *   Filter from a chained 2x2 f64 scan over per-buffer zero-state responses.  Per-voice error <= 1e-5 of full scale.
 * SEGMENTS: the time axis of a launch is cut into as many long segments as fill the GPU once; each segment is rendered
 *   sequentially, its Filter state warmed up over preceding buffers until the window's homogeneous map is <= 1e-8.
 *   Same tolerance as SCAN, about half its work; identical to EXACT when there are enough voices for one segment. */
enum { ZCU_RENDER_EXACT = 0, ZCU_RENDER_SCAN = 1, ZCU_RENDER_SEGMENTS = 2 };

/* per-voice constants: NiceInstrument.color (modules.zig:198) and the mix gains [voices][channels] */
int zcu_render_create(zcu_ctx*, const zcu_render_config* cfg, const float* voice_color_host,
                      const float* gains_host, zcu_render** out);
void zcu_render_free(zcu_render*);
/* host events -> device (H2D), resets all voice state.  ZCU_ERR_INVALID if event_off is not monotone from 0 or a voice's
 * events are not chronological inside [0, total_frames). */
int zcu_render_set_events(zcu_ctx*, zcu_render*, const uint32_t* event_off_host, const zcu_note_event* events_host);
/* renders the whole job into mix ([total_frames][channels], accumulating, like every zang paint) */
int zcu_render_run(zcu_ctx*, zcu_render*, zcu_buf* mix);
/* optional: per-voice output (before mixing) of frames [first_frame, first_frame+out->frames) for parity tests */
int zcu_render_run_voices(zcu_ctx*, zcu_render*, uint32_t first_frame, zcu_buf* out_voices);
/* the same for voices [first_voice, first_voice + out->voices) only: out is [frames][out->voices].  The window is rendered
 * with the launches and time segments of the whole job (what zcu_render_run executes), so it is the parity probe for a
 * job too large to read back whole (per-voice windows at the benchmark configuration). */
int zcu_render_run_voice_window(zcu_ctx*, zcu_render*, uint32_t first_voice, uint32_t first_frame, zcu_buf* out_voices);
uint64_t zcu_render_launches_per_run(const zcu_render*);
/* per-launch CUDA-event timing of the render kernel (the dominant kernel): after enabling, every
 * zcu_render_run records an event pair around each render-kernel launch; zcu_render_kernel_ms
 * synchronises and returns the summed duration and the number of launches since the last call. */
int zcu_render_set_profiling(zcu_render*, int enabled);
int zcu_render_kernel_ms(zcu_ctx*, zcu_render*, float* total_ms, uint32_t* launches);
/* ZCU_RENDER_SCAN only: the same for the zero-state filter pass that precedes each render-kernel launch */
int zcu_render_aux_kernel_ms(zcu_ctx*, zcu_render*, float* total_ms, uint32_t* launches);

/* ZCU_RENDER_SEGMENTS diagnostics of the last run: segments per launch, mix buffers per segment, and the sum / max
 * over all (voice group, segment) warp tasks of the warm-up length in mix buffers */
/* ZCU_RENDER_SEGMENTS: time segments per launch; 0 (default) = as many as fill the GPU's resident warps once, each at
 * least 8 mix buffers long.  Any value renders within the mode's tolerance; 1 is the exact sequential render. */
int zcu_render_set_segments(zcu_render*, uint32_t segments_per_launch);
int zcu_render_segment_info(zcu_ctx*, zcu_render*, uint32_t* segments, uint32_t* buffers_per_segment, uint32_t* warm_sum,
                            uint32_t* warm_max);
/* ZCU_RENDER_SEGMENTS, opt-in (default off): evaluate the Filter recurrence (Filter.zig:135-146) with each multiply-add
 * pair as one FMA: 10 instructions per sample instead of 16 in a kernel that is bound by instruction issue.  The Filter's
 * output then no longer has the reference's bits anywhere (each FMA skips one rounding; the recurrence is contractive, the
 * difference stays at a few 1e-7 of full scale, inside the 1e-5 this mode promises for the Filter); PulseOsc and Envelope
 * keep their exact forms.  ZCU_ERR_INVALID for another mode. */
int zcu_render_set_filter_contraction(zcu_render*, int enabled);

/* ---- fused module graphs ---------------------------------------------------------
 * The general form of zcu_render_*: an instrument's paint() body given as an instruction list and compiled at run time
 * (NVRTC, sm_100a, --fmad=false) into ONE kernel that runs the whole list sample by sample with the temps in registers.
 * The list is the reference's own: src/zangscript/codegen.zig:60-121 (`Instruction`: copy_buffer, float_to_buffer,
 * arith_*, call) as executed by src/zangscript/runtime.zig:445-740, plus the zang basics (src/zang/basics.zig:12-78) that
 * hand-written instruments use (examples/modules.zig).  Every instruction works on the span table it names; a CALL is the
 * module's paint(): it ACCUMULATES into its output, so the reference's `zero; paint` pair is two instructions (the script
 * runtime's `call`, runtime.zig:482-506, is ZERO followed by CALL).  Results equal the per-module path bit for bit
 * (tests/test_gpu_graph.py): no instruction of the reference reads any sample of a buffer but the one it is writing.
 * An instruction that reads a temp at a frame where nothing wrote it this call sees an unspecified value. */
typedef struct zcu_graph zcu_graph;

enum { ZCU_GREF_NONE = 0, ZCU_GREF_TEMP = 1, ZCU_GREF_OUTPUT = 2, ZCU_GREF_INPUT = 3 };
typedef struct zcu_graph_ref {
    int32_t kind;    /* ZCU_GREF_*: a temp of the graph, one of its outputs, or one of its external input buffers */
    uint32_t index;
} zcu_graph_ref;

enum { ZCU_GI_BASIC = 0, ZCU_GI_ARITH = 1, ZCU_GI_CALL = 2,
       /* the script runtime's `delay` block (codegen.zig:97-103, runtime.zig:704-721) in its per-sample form: DELAY_READ is
        * `zero(feedback_temp); delay.readDelayBuffer(feedback_temp)` (out = 0 + the delayed sample, delay.zig:28-56), DELAY_WRITE
        * is `delay.writeDelayBuffer(feedback_out_temp)` (in[0] goes into the same slot, the line advances, delay.zig:63-90);
        * the block's inner instructions sit between the two.  field = a ZCU_MOD_SIMPLE_DELAY bank (zcu_delay_bank_create: the
        * line and its index, i.e. zang.Delay(N)); both instructions name the same field and span table. */
       ZCU_GI_DELAY_READ = 3, ZCU_GI_DELAY_WRITE = 4 };
/* basics.zig:12-78, as zcu_zero .. zcu_multiply_with_scalar */
enum {
    ZCU_BASIC_ZERO = 0, ZCU_BASIC_SET = 1, ZCU_BASIC_COPY = 2, ZCU_BASIC_ADD = 3, ZCU_BASIC_ADD_INTO = 4, ZCU_BASIC_ADD_SCALAR = 5,
    ZCU_BASIC_ADD_SCALAR_INTO = 6, ZCU_BASIC_MULTIPLY = 7, ZCU_BASIC_MULTIPLY_WITH = 8, ZCU_BASIC_MULTIPLY_SCALAR = 9,
    ZCU_BASIC_MULTIPLY_WITH_SCALAR = 10
};
/* the script runtime's arithmetic that OVERWRITES its output (runtime.zig:528-700; add and mul are basics there) */
enum { ZCU_ARITH_SUB = 0, ZCU_ARITH_DIV = 1, ZCU_ARITH_MIN = 2, ZCU_ARITH_MAX = 3, ZCU_ARITH_NEG = 4, ZCU_ARITH_ABS = 5, ZCU_ARITH_SQRT = 6,
       /* libm per sample (CUDA sinf / cosf / powf; the reference's are Zig std's: parity within the tolerance of SineOsc) */
       ZCU_ARITH_SIN = 7, ZCU_ARITH_COS = 8, ZCU_ARITH_POW = 9 };
enum { ZCU_SCALAR_NONE = 0, ZCU_SCALAR_A = 1, ZCU_SCALAR_B = 2 };

typedef struct zcu_graph_instr {
    int32_t op;             /* ZCU_GI_* */
    int32_t sub;            /* BASIC: ZCU_BASIC_*   ARITH: ZCU_ARITH_*   CALL: ZCU_MOD_*   DELAY_*: unused */
    zcu_graph_ref out;      /* TEMP or OUTPUT */
    zcu_graph_ref in[3];    /* BASIC / ARITH: a, b.  CALL: the buffer arguments of zcu_paint_<module> in its order (SineOsc: freq,
                               phase; Filter: input, cutoff, res; ...); NONE = the scalar of the params row is used */
    float scalar;           /* BASIC: the scalar of the *_SCALAR forms and SET.  ARITH: the float operand */
    int32_t scalar_side;    /* ARITH: ZCU_SCALAR_NONE buffer op buffer, ZCU_SCALAR_A float op buffer (b = in[1]), ZCU_SCALAR_B buffer op float */
    uint32_t field;         /* CALL: which field (a module instance: bank + params rows) of the graph */
    uint32_t span_table;    /* which of the span tables given to zcu_graph_paint drives this instruction */
} zcu_graph_instr;

enum { ZCU_GRAPH_OUT_BUFFERS = 0, ZCU_GRAPH_OUT_MIX = 1 };
typedef struct zcu_graph_config {
    uint32_t voices;
    uint32_t num_temps, num_outputs, num_inputs; /* outputs, inputs <= 4 */
    uint32_t num_span_tables;                    /* <= 4 */
    uint32_t num_fields;                         /* <= 12 */
    int32_t output_mode;   /* BUFFERS: outputs are [frames][voices] buffers, accumulated (the module contract).  MIX: output 0 of
                              every voice starts each sample at 0 and is summed x gains into a [frames][channels] mix */
    uint32_t channels;     /* MIX: 1 or 2 */
} zcu_graph_config;

/* field_kinds[num_fields]: ZCU_MOD_* of every field.  Generates the kernel source, compiles and loads it. */
int zcu_graph_create(zcu_ctx*, const zcu_graph_config* cfg, const zcu_graph_instr* instrs, uint32_t num_instrs, const int32_t* field_kinds,
                     zcu_graph** out);
void zcu_graph_free(zcu_graph*);
/* the generated CUDA source of a graph (owned by the graph) */
const char* zcu_graph_source(const zcu_graph*);
/* Host-only (no device needed): generates the source of a graph and compiles it with NVRTC for sm_100a.  log (may be NULL)
 * receives the compiler log or the generator's error; *cubin_bytes the size of the machine code.  Used by the CPU tests. */
int zcu_graph_compile_check(const zcu_graph_config* cfg, const zcu_graph_instr* instrs, uint32_t num_instrs, const int32_t* field_kinds,
                            char* log, size_t log_len, size_t* cubin_bytes);
/* a field's state, params rows (`zcu_<module>_params`, row_bytes = its sizeof; kept on the device until set again) and
 * module-specific tables.  Stateless modules take bank = NULL. */
int zcu_graph_set_field(zcu_ctx*, zcu_graph*, uint32_t field, zcu_bank* bank, const void* params_host, size_t row_bytes, uint32_t nparams);
int zcu_graph_set_curve_nodes(zcu_ctx*, zcu_graph*, uint32_t field, const zcu_curve_node* nodes_host, uint32_t nnodes);
int zcu_graph_set_sample(zcu_ctx*, zcu_graph*, uint32_t field, const zcu_sample* sample);
int zcu_graph_set_gains(zcu_ctx*, zcu_graph*, const float* gains_host); /* MIX: [voices][channels] */
/* One launch paints the block: tables[num_span_tables] (all voices x frames of the buffers), outputs[num_outputs]
 * (MIX: outputs[0] is the [frames][channels] mix, accumulated), inputs[num_inputs]. */
int zcu_graph_paint(zcu_ctx*, zcu_graph*, const zcu_spans* const* tables, zcu_buf* const* outputs, const zcu_buf* const* inputs);
/* Time segments.  A graph whose modules can all skip ahead exactly from their own params (constant-frequency oscillators,
 * Envelope, Gate, Portamento, Curve, Noise, Sampler, Distortion, Cycle; Decimator through its counter) is cut along the frame axis like
 * the module paints are: every (128 voices x segment) CTA skips ahead in closed form, re-runs a short window before its
 * start silently (where a Decimator re-latches its held value: the window is sized from its params rows) and then paints.
 * Bit-identical to one segment.  zcu_graph_set_segments forces the count (0 = sized to the GPU, 1 = off);
 * zcu_graph_segment_info reports the last paint's segmentation and, in *failed, whether any segment could not establish
 * its start state since the last call (it synchronises; the painted block is then invalid). */
int zcu_graph_set_segments(zcu_graph*, uint32_t segments);
int zcu_graph_segment_info(zcu_ctx*, zcu_graph*, uint32_t* segments, uint32_t* segment_frames, uint32_t* replay_frames, uint32_t* failed);
/* A graph with a Filter (src/modules/Filter.zig) is cut into time segments only under ZCU_PAINT_TOLERANCE: a segment runs the
 * whole graph silently over a window before its start, the Filter from zero state, and paints once the window's homogeneous
 * map is below 1e-8 on every voice of the warp (the Filter has then forgotten where it started); a window that fails is
 * lengthened fourfold and the segment starts over, down to frame 0 = the loaded state, which is exact.  Voices that paint
 * nothing before the window keep their loaded state.  zcu_graph_warm_retries: (warp, segment) pairs that lengthened their
 * window since the last call (it synchronises).  Exact mode keeps such a graph in one segment. */
int zcu_graph_warm_retries(zcu_ctx*, zcu_graph*, uint32_t* retries);

/* ---- multi-GPU: sum of per-GPU partial mixes --------------------------------------------------- */
/* NCCL is loaded with dlopen("libnccl.so.2") at first use; id is the 128-byte ncclUniqueId. */
int zcu_nccl_unique_id(uint8_t id_out[128]);
int zcu_nccl_init(zcu_ctx*, int nranks, int rank, const uint8_t id[128]);
int zcu_nccl_reduce_mix(zcu_ctx*, zcu_buf* mix, int root); /* ncclReduce(sum, f32) in place on root */
int zcu_nccl_finalize(zcu_ctx*);

#ifdef __cplusplus
}
#endif
#endif /* ZANGCU_H */
