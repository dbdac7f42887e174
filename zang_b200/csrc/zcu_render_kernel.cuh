// zcu_render_kernel.cuh — the fused NiceInstrument bank kernel (device side of zcu_render.cu).
//
// examples/example_polyphony.zig:61-92 runs, per voice and per mix buffer,
//     ImpulseQueue.consume -> Trigger.next loop -> NiceInstrument.paint(span, outputs, temps, ...)
// and every voice accumulates into the shared output.  Here a CTA owns 32 voices for a whole block of
// mix buffers and runs the instrument's module chain (examples/modules.zig:212-247) as a 4-stage
// software pipeline, one warp (= one SM sub-partition) per stage, lane = voice:
//
//   warp 0  zero; PulseOsc.paint; multiplyWithScalar(0.5)          -> X tile   (PulseOsc.zig:75-114)
//   warp 1  zero; Filter.paint low_pass                      X tile -> Y tile   (Filter.zig:124-147)
//   warp 2  zero; Envelope.paint; multiply                   Y tile -> O tile   (Envelope.zig:33-109)
//   warp 3  pan gains, sum of the 32 voices                  O tile -> partial mix row in HBM
//
// Tiles are RT frames x 32 voices in shared memory, double buffered, one __syncthreads per tile.  The
// temps of the reference never exist in HBM.  Every stage replays the Trigger's span splitting itself
// (trigger.zig:80-196) with its own cursor over the voice's note events, so the only per-span data
// that is read from global memory is the note event (and the host-evaluated filter cutoff).
//
// The Filter's recurrence is an 11-operation dependent f32 chain per sample (44 cycles at 4 cycles per
// op, no FMA allowed); it is the pace setter.  Everything else is arranged to stay off that chain:
// per-sample code is branch free (selects), span/phase changes are handled between "runs" (the
// frames up to the nearest boundary of any of the warp's 32 voices, found with one warp reduction),
// and shared-memory loads are issued a group ahead of their use.
#pragma once

#include "zcu_modules.cuh"

namespace zcu {

struct NiceVoiceState {
    uint32_t cnt;                      // PulseOsc.cnt
    float l, b;                        // Filter.l/.b
    int32_t env_state;                 // Envelope.state
    float env_t, env_last, env_start;  // Envelope.painter
    int32_t has_note;                  // Trigger.note != null
    uint32_t note_id;                  // Trigger.note.id
    uint32_t cur_event;                // Trigger.note.params (row = event index)
    uint32_t next_event;               // ImpulseQueue read cursor
};

// Device copy of a note event: zcu_note_event + Filter.cutoffFromFrequency(freq*8, sample_rate), which the
// reference evaluates on the host per paint call (modules.zig:231-234).  32 bytes = two 16-byte loads.
struct __align__(16) DevEvent {
    uint32_t frame, note_id;
    float freq;
    int32_t note_on;
    float cutoff;
    uint32_t cnt0;   // ZCU_RENDER_SCAN: PulseOsc.cnt at this event's frame (k_nice_prepass)
    uint32_t flags;  // ZCU_RENDER_SCAN: EV_SUPERSEDED | EV_CHANGED (k_nice_prepass)
    uint32_t pad;
};

struct RenderArgs {
    uint32_t V, C;
    uint32_t buffer_frames, total_frames;
    uint32_t F0, F1;  // frames of this launch
    float sample_rate;
    const float* color;
    const float* gains;
    NiceVoiceState* state;
    const uint32_t* event_off;
    const DevEvent* events;  // note events with the host-evaluated filter cutoff, 32 bytes each
    float* partial;  // [nwarps][F1-F0][C] or null
    float* vout;     // [vout_frames][vout_nv] window or null
    uint32_t vout_first, vout_frames;
    uint32_t vout_v0, vout_nv;
    uint32_t debug_skip_mask;  // bit w set: stage w does no arithmetic (profiling aid; results are then wrong)
    unsigned long long* debug_cycles;  // optional [nCTA][4 warps][2]: cycles busy / cycles total (profiling aid)
};

// ---- Trigger replay, one boundary check per run (trigger.zig:80-196) ---------------------------------
// Each pipeline stage keeps its own cursor over the voice's events; all cursors take identical steps.
// The next two events are held in registers, loaded a whole note ahead of their use, so that a span
// boundary costs no memory latency.
struct TrigCursor {
    uint32_t next_event, ev_end;
    int32_t has_note;
    uint32_t note_id, cur_event;
    float cur_freq, cur_cutoff;  // params of the current note (Trigger.note.params)
    bool cur_on;
    DevEvent nx0, nx1;           // events[next_event], events[next_event + 1]; frame = UINT_MAX past the end
    uint32_t boundary, buf_end;
    bool active;

    __device__ __forceinline__ static DevEvent fetch(const DevEvent* ev, uint32_t i, uint32_t end) {
        DevEvent e;
        e.frame = 0xFFFFFFFFu;
        e.note_id = 0;
        e.freq = 0.0f;
        e.note_on = 0;
        e.cutoff = 0.0f;
        if (i < end) {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(ev + i));
            const float c = __ldg(&ev[i].cutoff);
            e.frame = a.x;
            e.note_id = a.y;
            e.freq = __uint_as_float(a.z);
            e.note_on = (int32_t)a.w;
            e.cutoff = c;
        }
        return e;
    }
    __device__ void init(const NiceVoiceState& st, uint32_t ev_end_, uint32_t F0, bool valid, const DevEvent* ev) {
        next_event = st.next_event;
        ev_end = valid ? ev_end_ : 0u;
        if (!valid) next_event = 0;
        has_note = st.has_note;
        note_id = st.note_id;
        cur_event = st.cur_event;
        cur_freq = cur_cutoff = 0.0f;
        cur_on = false;
        if (valid && has_note) {
            const DevEvent c = fetch(ev, cur_event, ev_end_);
            cur_freq = c.freq;
            cur_cutoff = c.cutoff;
            cur_on = c.note_on != 0;
        }
        nx0 = fetch(ev, next_event, ev_end);
        nx1 = fetch(ev, next_event + 1, ev_end);
        boundary = valid ? F0 : 0xFFFFFFFFu;
        buf_end = F0;
        active = false;
    }
    // at f == boundary: end the running span, consume the impulses of this frame (the last one at a frame
    // wins, trigger.zig:173-178), start the next span if a note is current.  Returns note_id_changed.
    __device__ bool step(uint32_t f, const RenderArgs& a) {
        if (f == buf_end) buf_end = min(f + a.buffer_frames, a.total_frames);  // next mix buffer
        bool changed = false;
        while (nx0.frame <= f) {
            changed = has_note ? (nx0.note_id != note_id) : true;
            // Trigger.next compares with the note of the previously RETURNED span; an impulse superseded at the
            // same frame is never returned, so only the last one of the frame is compared (and kept)
            cur_freq = nx0.freq;
            cur_cutoff = nx0.cutoff;
            cur_on = nx0.note_on != 0;
            const uint32_t id = nx0.note_id;
            cur_event = next_event;
            next_event++;
            nx0 = nx1;
            nx1 = fetch(a.events, next_event + 1, ev_end);
            if (!(nx0.frame <= f)) {
                has_note = 1;
                note_id = id;
            }
        }
        active = has_note != 0;
        boundary = min(buf_end, nx0.frame);
        return changed;
    }
};

// ---- stage 0: zero(temps[0]); PulseOsc.paint (constant frequency); multiplyWithScalar(temps[0], 0.5) ----
struct PulseStage {
    // What the stage emits is PulseOsc's value x 0.5 (multiplyWithScalar, modules.zig:226).  Halving is exact (no value of
    // the six cases is subnormal: they are sums with +-0.7 or +-0.7 itself), and fl(a/2 + b/2) = fl(a + b)/2, so the span
    // constants are kept pre-halved: g = 0.35, gdf = gain/utof23(ifreq) (the reference's gdf*2, halved), c121, c212.
    // A span whose frequency is out of range (PulseOsc.zig:82-84: paints nothing) has g = ng = 0 and ifreq = 0: the code is
    // then 0b000 or 0b011 for ever and the value +0.  None of the six values can be -0 (a cancelling sum rounds to +0), so
    // the reference's `0 + v` into the zeroed temp changes nothing.
    uint32_t cnt, ifreq, brpt;
    float gdf, col, c121, c212, g, ng;
    bool pskip, was_up;  // was_up: (cnt - ifreq) < brpt, the previous sample's is_up, carried instead of recomputed
    __device__ void init(float color, uint32_t cnt0) {
        brpt = ftou32(clamp01(color));
        col = utof23(brpt);
        cnt = cnt0;
        ifreq = 0;
        pskip = true;
        gdf = c121 = c212 = g = ng = 0.0f;
        was_up = cnt < brpt;
    }
    __device__ __forceinline__ void set_cnt(uint32_t c) {
        cnt = c;
        was_up = (uint32_t)(cnt - ifreq) < brpt;
    }
    __device__ void begin_span(float sample_rate, float freq) {
        const float gain = 0.7f;
        pskip = (freq < 0.0f || freq > sample_rate / 8.0f);  // PulseOsc.zig:82-84: paints nothing, cnt untouched
        ifreq = 0;
        g = ng = 0.0f;
        if (!pskip) {
            const float basefrq = 4294967296.0f / sample_rate;
            ifreq = (uint32_t)(basefrq * freq);
            gdf = gain / utof23(ifreq);  // the reference uses (gdf*2)*x + gain; all of it halved here
            g = gain * 0.5f;
            ng = -g;
            c121 = gdf * (col - 1.0f) + g;
            c212 = gdf * col - g;
        }
        was_up = (uint32_t)(cnt - ifreq) < brpt;  // also how PulseOsc.zig:96 seeds its history
    }
    // :98-112, then x0.5.  The reference's 2-bit history `state` always holds (previous, current) values of
    // `cnt < brpt`, so the 3-bit transition code is a pure function of cnt: bits (wrapped = cnt < ifreq, was_up, is_up).
    // The six reachable codes are selected with predicates; 0b001 and 0b110 cannot occur (they would need ifreq > 2^32).
    __device__ __forceinline__ float sample() {
        const float p = utof23(cnt);
        const bool is_up = cnt < brpt;
        const bool wrapped = cnt < ifreq;
        const float v2 = gdf * (col - p) + g;  // 0b010 up down
        const float v5 = gdf * p - g;          // 0b101 down up
        const float w1 = is_up ? (was_up ? c121 : v5) : c212;  // wrapped: 0b111, 0b101, 0b100
        const float w0 = was_up ? (is_up ? g : v2) : ng;       // not wrapped: 0b011, 0b010, 0b000
        const float v = wrapped ? w1 : w0;
        was_up = is_up;
        cnt += ifreq;  // ifreq == 0 when skipped
        return v;
    }
};

// ---- stage 1: zero(temps[1]); Filter.paint(low_pass, cutoff const, res const 0.7) ------------------------
struct FilterStage {
    float l, b, cut, res;
    __device__ void begin_span(float cutoff) {
        cut = zclamp(cutoff, 0.0f, 1.0f);       // Filter.zig:114
        res = 1.0f - zclamp(0.7f, 0.0f, 1.0f);  // modules.zig:235, Filter.zig:118
    }
    __device__ __forceinline__ float sample(float x) {  // Filter.zig:135-146
        const float dc = 3.814697265625e-6f;
        const float in = x + dc;
        l += cut * b - dc;
        b += cut * (in - b * res - l);
        l += cut * b;
        const float h = in - b * res - l;
        b += cut * h;
        // low_pass output is 0 + (l*1 + b*0 + h*0) (Filter.zig:146 into the zeroed temps[1]).  While b and h are
        // finite, b*0 and h*0 are signed zeros and the sum is l, or +0 when l is a zero of either sign.  Only the sign of
        // a zero can differ from l itself, and it cannot reach the output: the voice's value is 0 + env*y (multiply()
        // accumulates into the zeroed output, modules.zig:246), which is +0 for a product of either sign.
        return l;
    }
    // CONTRACTED: the same recurrence with every multiply-add pair as one FMA (10 instructions instead of 16).  Not the
    // reference's bits: each FMA skips one rounding, the filter is contractive, the difference stays at a few ulp of the
    // state.  Only for render modes that promise the Filter a tolerance anyway (zcu_render_set_filter_contraction).
    template <bool CONTRACTED>
    __device__ __forceinline__ float sample_t(float x) {
        if (!CONTRACTED) return sample(x);
        const float dc = 3.814697265625e-6f;
        const float in = x + dc;
        l = __fmaf_rn(cut, b, l) - dc;
        b = __fmaf_rn(cut, __fmaf_rn(-res, b, in) - l, b);
        l = __fmaf_rn(cut, b, l);
        const float h = __fmaf_rn(-res, b, in) - l;
        b = __fmaf_rn(cut, h, b);
        return l;
    }
};

// ---- stage 2: zero(temps[0]); Envelope.paint; multiply(outputs[0], temps[0], temps[1]) -------------------
struct EnvStage {
    EnvelopeCore e;
    __device__ void begin_span(float sample_rate, bool note_on, bool note_id_changed) {
        zcu_envelope_params ep;  // modules.zig:238-245
        ep.sample_rate = sample_rate;
        ep.attack.kind = ZCU_CURVE_CUBED;
        ep.attack.duration = 0.01f;
        ep.decay.kind = ZCU_CURVE_CUBED;
        ep.decay.duration = 0.1f;
        ep.release.kind = ZCU_CURVE_CUBED;
        ep.release.duration = 0.5f;
        ep.sustain_volume = 0.8f;
        ep.note_on = note_on ? 1 : 0;
        e.begin(ep, note_id_changed);
    }
    __device__ __forceinline__ bool ramping() const { return e.mode == EnvelopeCore::M_RAMP_ON || e.mode == EnvelopeCore::M_RAMP_OFF; }
    // Number of coming samples for which the ramp provably does not reach t >= 1 (painter.zig:104-106), so
    // that sample_fast() may skip the test.  f32 accumulation of t_step drifts by at most half an ulp(1) per
    // add; the margin covers a run of up to 256 samples.  0 = take the careful path for the next sample.
    __device__ __forceinline__ uint32_t safe_run() const {
        if (!ramping()) return 0xFFFFFFFFu;
        const float room = (1.0f - e.painter.t) - 3.2e-5f;
        const float k = room / e.painter.t_step - 2.0f;
        if (!(k >= 1.0f)) return 0u;  // also catches NaN / inf steps
        return k > 1.0e6f ? 1000000u : (uint32_t)k;
    }
    // Per-run invariants of the fast path (no phase change can happen inside a run, see safe_run).
    struct Run {
        // ramping: (t_step, start, goal - start); otherwise (0, flat value or 0, 0): the ramp formula start + tp*gms then
        // yields the flat value itself (tp is finite and tp*0 = +-0), so the per-sample code needs no select.  Only the
        // sign of a zero can differ from the reference's 0 + value, and that cannot reach the output (see FilterStage).
        float step, start, gms;
        bool ramp, lin, sq;
    };
    __device__ __forceinline__ Run begin_run() const {
        Run r;
        r.ramp = ramping();
        r.step = r.ramp ? e.painter.t_step : 0.0f;  // t + 0 leaves t untouched
        const float base = e.mode == EnvelopeCore::M_FLAT ? e.flat : 0.0f;
        r.start = r.ramp ? e.painter.start : base;
        r.gms = r.ramp ? e.painter.gms : 0.0f;
        r.lin = e.painter.kind == ZCU_CURVE_LINEAR;
        r.sq = e.painter.kind == ZCU_CURVE_SQUARED;
        return r;
    }
    // what the envelope adds to the zeroed temps[0] (painter.zig:103-115 without the t >= 1 test, which
    // cannot fire inside a run).  CUBED = every ramping voice of the warp uses the cubed curve.
    template <bool CUBED>
    __device__ __forceinline__ float sample_fast(const Run& r, float& lv) {
        PainterDev& p = e.painter;
        p.t = p.t + r.step;
        const float it = 1.0f - p.t;
        float tp;
        if (CUBED) {
            tp = 1.0f - it * it * it;
        } else {
            const float sq = it * it;
            const float curved = 1.0f - (r.sq ? sq : sq * it);
            tp = r.lin ? p.t : curved;
        }
        lv = r.start + tp * r.gms;
        return lv;
    }
    __device__ __forceinline__ void end_run(const Run& r, float lv, bool any) {
        if (r.ramp && any) e.painter.last_value = lv;
    }
    // the general per-sample step with the phase changes of paintOn / paintOff
    __device__ float sample_careful() {
        float acc = 0.0f;
        e.sample(acc);
        return acc;
    }
};

// 4 warps per CTA, software-pipelined over tiles of RT frames; see the file header.
template <int C, bool VOUT, int RT>
__global__ void __launch_bounds__(128, 1) k_render_nice(RenderArgs a) {
    extern __shared__ float smem[];
    float(*X)[RT][32] = reinterpret_cast<float(*)[RT][32]>(smem);
    float(*Y)[RT][32] = reinterpret_cast<float(*)[RT][32]>(smem + 2 * RT * 32);
    float(*O)[RT][33] = reinterpret_cast<float(*)[RT][33]>(smem + 4 * RT * 32);
    float(*g)[32] = reinterpret_cast<float(*)[32]>(smem + 4 * RT * 32 + 2 * RT * 33);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t v = blockIdx.x * 32 + lane;
    const bool valid = v < a.V;
    const uint32_t FULL = 0xffffffffu;
    const bool skip_math = (a.debug_skip_mask >> warp) & 1u;

    NiceVoiceState st{};
    uint32_t ev_end = 0;
    if (valid) {
        st = a.state[v];
        ev_end = a.event_off[v + 1];
    }
    TrigCursor cur;
    cur.init(st, ev_end, a.F0, valid, a.events);
    PulseStage ps;
    FilterStage fs;
    EnvStage es;
    if (warp == 0) {
        const float color = valid ? a.color[v] : 0.0f;
        ps.init(color, st.cnt);
    } else if (warp == 1) {
        fs.l = st.l;
        fs.b = st.b;
        fs.cut = fs.res = 0.0f;
    } else if (warp == 2) {
        es.e.state = st.env_state;
        es.e.painter.t = st.env_t;
        es.e.painter.last_value = st.env_last;
        es.e.painter.start = st.env_start;
        es.e.painter.t_step = 0.0f;
        es.e.painter.gms = 0.0f;
        es.e.painter.kind = ZCU_CURVE_LINEAR;
        es.e.mode = EnvelopeCore::M_NONE;
        es.e.flat = 0.0f;
        es.e.sample_rate = a.sample_rate;
        es.e.sustain_volume = 0.0f;
    } else {
        for (int c = 0; c < C; c++) g[c][lane] = valid ? a.gains[(size_t)v * C + c] : 0.0f;
    }

    const uint32_t nframes = a.F1 - a.F0;
    const uint32_t ntiles = (nframes + RT - 1) / RT;
    long long busy = 0;
    const long long t_begin = clock64();
    for (uint32_t k = 0; k < ntiles + 3; k++) {
        const int tile = (int)k - (int)warp;  // this warp's tile in this iteration
        const long long t_tile = clock64();
        if (tile >= 0 && tile < (int)ntiles) {
            const uint32_t tf = a.F0 + (uint32_t)tile * RT;
            const uint32_t n = min((uint32_t)RT, a.F1 - tf);
            const uint32_t sb = (uint32_t)tile & 1u;
            if (warp < 3) {
                uint32_t j = 0;
                while (j < n) {
                    const uint32_t f = tf + j;
                    if (f == cur.boundary) {
                        const bool changed = cur.step(f, a);
                        if (cur.active) {
                            if (warp == 0) ps.begin_span(a.sample_rate, cur.cur_freq);
                            if (warp == 1) fs.begin_span(cur.cur_cutoff);
                            if (warp == 2) es.begin_span(a.sample_rate, cur.cur_on, changed);
                        }
                    }
                    // frames until the nearest span boundary (warp 2: or possible end of an envelope ramp) of any
                    // voice in the warp: no checks inside a run
                    uint32_t dist = cur.boundary - f;
                    bool careful = false;
                    if (warp == 2 && cur.active) {
                        const uint32_t s = es.safe_run();
                        careful = (s == 0u);
                        dist = min(dist, careful ? 1u : s);
                    }
                    const uint32_t run = min(n - j, __reduce_min_sync(FULL, dist));
                    const bool fast = __all_sync(FULL, cur.active && !careful) && !skip_math;
                    if (warp == 0) {
                        float* xo = &X[sb][j][lane];
                        if (fast) {
#pragma unroll 4
                            for (uint32_t r = 0; r < run; r++) xo[r * 32] = ps.sample();
                        } else {
                            for (uint32_t r = 0; r < run; r++) xo[r * 32] = (cur.active && !skip_math) ? ps.sample() : 0.0f;
                        }
                    } else if (warp == 1) {
                        const float* xi = &X[sb][j][lane];
                        float* yo = &Y[sb][j][lane];
                        uint32_t r = 0;
                        if (fast) {
                            // groups of 4 with the next group's shared-memory loads issued before this group's math
                            float x0 = 0.f, x1 = 0.f, x2 = 0.f, x3 = 0.f;
                            if (run >= 4) {
                                x0 = xi[0];
                                x1 = xi[32];
                                x2 = xi[64];
                                x3 = xi[96];
                            }
                            for (; r + 4 <= run; r += 4) {
                                float n0 = 0.f, n1 = 0.f, n2 = 0.f, n3 = 0.f;
                                if (r + 8 <= run) {
                                    n0 = xi[(r + 4) * 32];
                                    n1 = xi[(r + 5) * 32];
                                    n2 = xi[(r + 6) * 32];
                                    n3 = xi[(r + 7) * 32];
                                }
                                const float y0 = fs.sample(x0);
                                const float y1 = fs.sample(x1);
                                const float y2 = fs.sample(x2);
                                const float y3 = fs.sample(x3);
                                yo[(r + 0) * 32] = y0;
                                yo[(r + 1) * 32] = y1;
                                yo[(r + 2) * 32] = y2;
                                yo[(r + 3) * 32] = y3;
                                x0 = n0;
                                x1 = n1;
                                x2 = n2;
                                x3 = n3;
                            }
                            for (; r < run; r++) yo[r * 32] = fs.sample(xi[r * 32]);
                        } else {
                            for (; r < run; r++) yo[r * 32] = (cur.active && !skip_math) ? fs.sample(xi[r * 32]) : 0.0f;
                        }
                    } else {
                        const float* yi = &Y[sb][j][lane];
                        float* oo = &O[sb][j][lane];
                        uint32_t r = 0;
                        if (fast) {
                            const EnvStage::Run er = es.begin_run();
                            const bool cubed = __all_sync(FULL, !er.ramp || (!er.lin && !er.sq));
                            float lv = 0.0f;
                            if (cubed) {
                                for (; r + 4 <= run; r += 4) {
                                    const float y0 = yi[(r + 0) * 32];
                                    const float y1 = yi[(r + 1) * 32];
                                    const float y2 = yi[(r + 2) * 32];
                                    const float y3 = yi[(r + 3) * 32];
                                    const float e0 = es.sample_fast<true>(er, lv);
                                    const float e1 = es.sample_fast<true>(er, lv);
                                    const float e2 = es.sample_fast<true>(er, lv);
                                    const float e3 = es.sample_fast<true>(er, lv);
                                    oo[(r + 0) * 33] = e0 * y0;
                                    oo[(r + 1) * 33] = e1 * y1;
                                    oo[(r + 2) * 33] = e2 * y2;
                                    oo[(r + 3) * 33] = e3 * y3;
                                }
                                for (; r < run; r++) oo[r * 33] = es.sample_fast<true>(er, lv) * yi[r * 32];
                            } else {
                                for (; r < run; r++) oo[r * 33] = es.sample_fast<false>(er, lv) * yi[r * 32];
                            }
                            es.end_run(er, lv, run > 0);
                        } else {
                            for (; r < run; r++) oo[r * 33] = (cur.active && !skip_math) ? es.sample_careful() * yi[r * 32] : 0.0f;
                        }
                    }
                    j += run;
                }
            } else if (VOUT) {
                for (uint32_t j = 0; j < n; j++) {
                    const uint32_t f = tf + j;
                    if (valid && f >= a.vout_first && f < a.vout_first + a.vout_frames && v - a.vout_v0 < a.vout_nv)
                        a.vout[(size_t)(f - a.vout_first) * a.vout_nv + (v - a.vout_v0)] = 0.0f + O[sb][j][lane];  // multiply() accumulates into the zeroed output: -0 becomes +0
                }
            } else {
                // lane sums frames lane, lane+32, ... over the warp's 32 voices with their pan gains
                // (the reference's voice summation order is not reproduced; tolerance 1e-5 x sum|gain|)
                for (uint32_t fr = lane; fr < n; fr += 32) {
                    float acc[C];
#pragma unroll
                    for (int c = 0; c < C; c++) acc[c] = 0.0f;
                    if (!skip_math) {
#pragma unroll 8
                        for (int vv = 0; vv < 32; vv++) {
                            const float x = O[sb][fr][vv];
#pragma unroll
                            for (int c = 0; c < C; c++) acc[c] = fmaf(g[c][vv], x, acc[c]);
                        }
                    }
                    float* dst = a.partial + ((size_t)blockIdx.x * nframes + (tf - a.F0 + fr)) * C;
                    if (C == 2)
                        *reinterpret_cast<float2*>(dst) = make_float2(acc[0], acc[1]);
                    else
                        dst[0] = acc[0];
                }
            }
        }
        busy += clock64() - t_tile;
        __syncthreads();
    }
    if (a.debug_cycles && lane == 0) {
        a.debug_cycles[((size_t)blockIdx.x * 4 + warp) * 2 + 0] += (unsigned long long)busy;
        a.debug_cycles[((size_t)blockIdx.x * 4 + warp) * 2 + 1] += (unsigned long long)(clock64() - t_begin);
    }
    if (valid) {
        NiceVoiceState* out = a.state + v;
        if (warp == 0) {
            out->cnt = ps.cnt;
            out->has_note = cur.has_note;
            out->note_id = cur.note_id;
            out->cur_event = cur.cur_event;
            out->next_event = cur.next_event;
        } else if (warp == 1) {
            out->l = fs.l;
            out->b = fs.b;
        } else if (warp == 2) {
            out->env_state = es.e.state;
            out->env_t = es.e.painter.t;
            out->env_last = es.e.painter.last_value;
            out->env_start = es.e.painter.start;
        }
    }
}

template <int RT>
constexpr size_t render_smem_bytes() {
    return (size_t)(4 * RT * 32 + 2 * RT * 33 + 2 * 32) * sizeof(float);
}

}  // namespace zcu
