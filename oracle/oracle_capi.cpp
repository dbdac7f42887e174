// oracle_capi.cpp — TEST INFRASTRUCTURE ONLY (see zang_oracle.hpp header).
//
// C entry points (ctypes) that apply the single-voice restatement in zang_oracle.hpp to the batched
// data formats of include/zangcu.h: [frame][voice] f32 buffers, per-voice span lists, POD state rows.
// For every voice and every span record the voice's column is gathered into contiguous scratch,
// the restated Module.paint is called exactly as the reference would call it, and the column is
// scattered back.  Nothing here is optimised; it is the checker and the timed CPU baseline.
#include <chrono>
#include <cstdio>
#include <thread>

#include "../include/zangcu.h"
#include "zang_oracle.hpp"
#include "bank_oracle.hpp"
#include "delay_oracle.hpp"
#include "song_oracle.hpp"

using namespace zo;

namespace {

struct Col {  // one voice's view of a [frame][voice] buffer
    float* base;
    uint32_t V, v;
    void gather(uint32_t s, uint32_t e, float* dst) const {
        if (!base) return;
        for (uint32_t f = s; f < e; f++) dst[f - s] = base[(size_t)f * V + v];
    }
    void scatter(uint32_t s, uint32_t e, const float* src) const {
        for (uint32_t f = s; f < e; f++) base[(size_t)f * V + v] = src[f - s];
    }
};

PaintCurve pc(const zcu_painter_curve& c) { return PaintCurve{(CurveKind)c.kind, c.duration}; }

}  // namespace

struct NoteOnParam {
    uint32_t row;
    bool note_on;
};

template <size_t N>
static void poly_dispatch_n(zcu_poly_slot* slots, const zcu_impulse* impulses, const uint32_t* params, const int32_t* note_on,
                            uint32_t n, zcu_impulse* out_impulses, uint32_t* out_params, uint32_t* out_counts) {
    typename Notes<NoteOnParam>::template PolyphonyDispatcher<N> pd;
    for (size_t i = 0; i < N; i++) {
        pd.slots[i].used = slots[i].used != 0;
        pd.slots[i].note_id = slots[i].note_id;
        pd.slots[i].event_id = slots[i].event_id;
        pd.slots[i].note_on = slots[i].note_on != 0;
    }
    std::vector<Impulse> imps(n);
    std::vector<NoteOnParam> ps(n);
    for (uint32_t i = 0; i < n; i++) {
        imps[i] = Impulse{impulses[i].frame, impulses[i].note_id, impulses[i].event_id};
        ps[i] = NoteOnParam{params[i], note_on[i] != 0};
    }
    typename Notes<NoteOnParam>::ImpulsesAndParamses res[N];
    pd.dispatch({imps.data(), ps.data(), n}, res);
    for (size_t i = 0; i < N; i++) {
        out_counts[i] = (uint32_t)res[i].len;
        for (size_t k = 0; k < res[i].len; k++) {
            out_impulses[i * 32 + k] = zcu_impulse{(uint32_t)res[i].impulses[k].frame, (uint32_t)res[i].impulses[k].note_id,
                                                   (uint32_t)res[i].impulses[k].event_id};
            out_params[i * 32 + k] = res[i].paramses[k].row;
        }
        slots[i].used = pd.slots[i].used;
        slots[i].note_id = (uint32_t)pd.slots[i].note_id;
        slots[i].event_id = (uint32_t)pd.slots[i].event_id;
        slots[i].note_on = pd.slots[i].note_on;
    }
}

extern "C" {

// ---- basics --------------------------------------------------------------------------------
// op: 0 zero 1 set 2 copy 3 add 4 addInto 5 addScalar 6 addScalarInto 7 multiply 8 multiplyWith
//     9 multiplyScalar 10 multiplyWithScalar
int orc_basic(int op, uint32_t V, uint32_t frames, const uint32_t* off, const zcu_span* spans, float* dest,
              const float* a, const float* b, float scalar) {
    std::vector<float> d(frames), aa(frames), bb(frames);
    for (uint32_t v = 0; v < V; v++) {
        uint32_t k0 = off ? off[v] : 0, k1 = off ? off[v + 1] : 1;
        for (uint32_t k = k0; k < k1; k++) {
            uint32_t s = off ? spans[k].start : 0, e = off ? spans[k].end : frames;
            Col cd{dest, V, v}, ca{(float*)a, V, v}, cb{(float*)b, V, v};
            cd.gather(s, e, d.data());
            ca.gather(s, e, aa.data());
            cb.gather(s, e, bb.data());
            Span sp{0, (size_t)(e - s)};
            switch (op) {
                case 0: zero(sp, d.data()); break;
                case 1: set(sp, d.data(), scalar); break;
                case 2: copy(sp, d.data(), aa.data()); break;
                case 3: add(sp, d.data(), aa.data(), bb.data()); break;
                case 4: addInto(sp, d.data(), aa.data()); break;
                case 5: addScalar(sp, d.data(), aa.data(), scalar); break;
                case 6: addScalarInto(sp, d.data(), scalar); break;
                case 7: multiply(sp, d.data(), aa.data(), bb.data()); break;
                case 8: multiplyWith(sp, d.data(), aa.data()); break;
                case 9: multiplyScalar(sp, d.data(), aa.data(), scalar); break;
                case 10: multiplyWithScalar(sp, d.data(), scalar); break;
                default: return 1;
            }
            cd.scatter(s, e, d.data());
        }
    }
    return 0;
}

// ---- module paints ---------------------------------------------------------------------------
// in0/in1/in2: optional [frame][voice] input buffers (meaning per module as in zangcu.h).
// aux: curve nodes (Curve) or PCM bytes (Sampler); aux_meta: {num_channels, sample_rate, format}.
int orc_paint(int kind, uint32_t V, uint32_t frames, const uint32_t* off, const zcu_span* spans, const void* params,
              uint32_t nparams, void* state, float* out, const float* in0, const float* in1, const float* in2,
              const void* aux, uint64_t aux_len, const int64_t* aux_meta) {
    (void)frames;
    std::vector<float> o(frames), i0(frames), i1(frames), i2(frames);
    for (uint32_t v = 0; v < V; v++) {
        Col co{out, V, v}, c0{(float*)in0, V, v}, c1{(float*)in1, V, v}, c2{(float*)in2, V, v};
        for (uint32_t k = off[v]; k < off[v + 1]; k++) {
            const zcu_span& zs = spans[k];
            if (zs.param >= nparams) return 2;
            const uint32_t s = zs.start, e = zs.end;
            const bool nic = (zs.flags & ZCU_SPAN_NOTE_ID_CHANGED) != 0;
            co.gather(s, e, o.data());
            c0.gather(s, e, i0.data());
            c1.gather(s, e, i1.data());
            c2.gather(s, e, i2.data());
            Span sp{0, (size_t)(e - s)};
            switch (kind) {
                case ZCU_MOD_SINE_OSC: {
                    auto& st = ((zcu_sine_state*)state)[v];
                    const auto& p = ((const zcu_sine_params*)params)[zs.param];
                    SineOsc m;
                    m.t = st.t;
                    m.paint(sp, o.data(), nic,
                            {p.sample_rate, in0 ? buffer(i0.data()) : constant(p.freq), in1 ? buffer(i1.data()) : constant(p.phase)});
                    st.t = m.t;
                } break;
                case ZCU_MOD_PULSE_OSC: {
                    auto& st = ((zcu_pulse_state*)state)[v];
                    const auto& p = ((const zcu_pulse_params*)params)[zs.param];
                    PulseOsc m;
                    m.cnt = st.cnt;
                    m.paint(sp, o.data(), nic, {p.sample_rate, in0 ? buffer(i0.data()) : constant(p.freq), p.color});
                    st.cnt = m.cnt;
                } break;
                case ZCU_MOD_TRISAW_OSC: {
                    auto& st = ((zcu_trisaw_state*)state)[v];
                    const auto& p = ((const zcu_trisaw_params*)params)[zs.param];
                    TriSawOsc m;
                    m.cnt = st.cnt;
                    m.t = st.t;
                    m.paint(sp, o.data(), nic, {p.sample_rate, in0 ? buffer(i0.data()) : constant(p.freq), p.color});
                    st.cnt = m.cnt;
                    st.t = m.t;
                } break;
                case ZCU_MOD_NOISE: {
                    auto& st = ((zcu_noise_state*)state)[v];
                    const auto& p = ((const zcu_noise_params*)params)[zs.param];
                    Noise m(0);
                    for (int q = 0; q < 4; q++) m.r.s[q] = st.s[q];
                    m.paint(sp, o.data(), nic, {(Noise::Color)p.color});
                    for (int q = 0; q < 4; q++) st.s[q] = m.r.s[q];
                } break;
                case ZCU_MOD_ENVELOPE: {
                    auto& st = ((zcu_envelope_state*)state)[v];
                    const auto& p = ((const zcu_envelope_params*)params)[zs.param];
                    Envelope m;
                    m.state = (Envelope::State)st.state;
                    m.painter.t = st.painter.t;
                    m.painter.last_value = st.painter.last_value;
                    m.painter.start = st.painter.start;
                    m.paint(sp, o.data(), nic,
                            {p.sample_rate, pc(p.attack), pc(p.decay), pc(p.release), p.sustain_volume, p.note_on != 0});
                    st.state = (int32_t)m.state;
                    st.painter = {m.painter.t, m.painter.last_value, m.painter.start};
                } break;
                case ZCU_MOD_GATE: {
                    const auto& p = ((const zcu_gate_params*)params)[zs.param];
                    Gate m;
                    m.paint(sp, o.data(), nic, {p.note_on != 0});
                } break;
                case ZCU_MOD_FILTER: {
                    auto& st = ((zcu_filter_state*)state)[v];
                    const auto& p = ((const zcu_filter_params*)params)[zs.param];
                    Filter m;
                    m.l = st.l;
                    m.b = st.b;
                    m.paint(sp, o.data(), nic,
                            {i0.data(), (Filter::Type)p.type, in1 ? buffer(i1.data()) : constant(p.cutoff),
                             in2 ? buffer(i2.data()) : constant(p.res)});
                    st.l = m.l;
                    st.b = m.b;
                } break;
                case ZCU_MOD_PORTAMENTO: {
                    auto& st = ((zcu_portamento_state*)state)[v];
                    const auto& p = ((const zcu_portamento_params*)params)[zs.param];
                    Portamento m;
                    m.painter.t = st.painter.t;
                    m.painter.last_value = st.painter.last_value;
                    m.painter.start = st.painter.start;
                    m.paint(sp, o.data(), nic, {p.sample_rate, pc(p.curve), p.goal, p.note_on != 0, p.prev_note_on != 0});
                    st.painter = {m.painter.t, m.painter.last_value, m.painter.start};
                } break;
                case ZCU_MOD_CURVE: {
                    auto& st = ((zcu_curve_state*)state)[v];
                    const auto& p = ((const zcu_curve_params*)params)[zs.param];
                    const zcu_curve_node* nodes = (const zcu_curve_node*)aux;
                    if ((uint64_t)p.node_start + p.node_count > aux_len) return 3;
                    Curve m;
                    m.t = st.t;
                    m.current_song_note = st.current_song_note;
                    m.current_song_note_offset = st.current_song_note_offset;
                    m.next_song_note = st.next_song_note;
                    static_assert(sizeof(CurveNode) == sizeof(zcu_curve_node), "node layout");
                    m.paint(sp, o.data(), nic,
                            {p.sample_rate, (Curve::Function)p.function, (const CurveNode*)(nodes + p.node_start), p.node_count});
                    st.t = m.t;
                    st.current_song_note = (uint32_t)m.current_song_note;
                    st.current_song_note_offset = m.current_song_note_offset;
                    st.next_song_note = (uint32_t)m.next_song_note;
                } break;
                case ZCU_MOD_SAMPLER: {
                    auto& st = ((zcu_sampler_state*)state)[v];
                    const auto& p = ((const zcu_sampler_params*)params)[zs.param];
                    Sampler m;
                    m.t = st.t;
                    Sampler::Sample smp{(size_t)aux_meta[0], (size_t)aux_meta[1], (Sampler::Format)aux_meta[2],
                                        (const uint8_t*)aux, (size_t)aux_len};
                    m.paint(sp, o.data(), nic, {p.sample_rate, smp, (size_t)p.channel, p.loop != 0});
                    st.t = m.t;
                } break;
                case ZCU_MOD_DECIMATOR: {
                    auto& st = ((zcu_decimator_state*)state)[v];
                    const auto& p = ((const zcu_decimator_params*)params)[zs.param];
                    Decimator m;
                    m.dval = st.dval;
                    m.dcount = st.dcount;
                    m.paint(sp, o.data(), nic, {p.sample_rate, i0.data(), p.fake_sample_rate});
                    st.dval = m.dval;
                    st.dcount = m.dcount;
                } break;
                case ZCU_MOD_DISTORTION: {
                    const auto& p = ((const zcu_distortion_params*)params)[zs.param];
                    Distortion m;
                    m.paint(sp, o.data(), nic, {i0.data(), (Distortion::Type)p.type, p.ingain, p.outgain, p.offset});
                } break;
                case ZCU_MOD_CYCLE: {
                    auto& st = ((zcu_cycle_state*)state)[v];
                    const auto& p = ((const zcu_cycle_params*)params)[zs.param];
                    Cycle m;
                    m.t = st.t;
                    m.paint(sp, o.data(), nic, {p.sample_rate, in0 ? buffer(i0.data()) : constant(p.speed)});
                    st.t = m.t;
                } break;
                case ZCU_MOD_FM_OSCILLATOR: {
                    auto& st = ((zcu_fm_oscillator_state*)state)[v];
                    const auto& p = ((const zcu_fm_oscillator_params*)params)[zs.param];
                    if (!in0) return 3;
                    FmOscillator m;
                    m.t = st.t;
                    m.feedback1 = st.feedback1;
                    m.feedback2 = st.feedback2;
                    m.paint(sp, o.data(), nic, {p.sample_rate, p.waveform, i0.data(), in1 ? i1.data() : nullptr, p.feedback});
                    st.t = m.t;
                    st.feedback1 = m.feedback1;
                    st.feedback2 = m.feedback2;
                } break;
                default: return 1;
            }
            co.scatter(s, e, o.data());
        }
    }
    return 0;
}

// ---- delay modules: kind = ZCU_MOD_SIMPLE_DELAY / ZCU_MOD_FILTERED_ECHOES -------------------------------------
// state: zcu_delay_state / zcu_echoes_state rows; line: [delay_samples][V] f32 (the per-voice delay_buffer, voice fastest);
// params: zcu_echoes_params rows (FilteredEchoes only).  Each span is one reference paint() call, chunked as the reference does.
int orc_paint_delay(int kind, uint32_t V, uint32_t frames, const uint32_t* off, const zcu_span* spans, const void* params,
                    uint32_t nparams, void* state, float* line, uint32_t delay_samples, float* out, const float* in0) {
    std::vector<float> o(frames), i0(frames), t0(frames), t1(frames);
    for (uint32_t v = 0; v < V; v++) {
        Col co{out, V, v}, c0{(float*)in0, V, v};
        Delay d(delay_samples);
        for (uint32_t q = 0; q < delay_samples; q++) d.delay_buffer[q] = line[(size_t)q * V + v];
        Filter flt;
        if (kind == ZCU_MOD_SIMPLE_DELAY) {
            d.delay_buffer_index = ((zcu_delay_state*)state)[v].index;
        } else {
            auto& st = ((zcu_echoes_state*)state)[v];
            d.delay_buffer_index = st.index;
            flt.l = st.l;
            flt.b = st.b;
        }
        for (uint32_t k = off[v]; k < off[v + 1]; k++) {
            const zcu_span& zs = spans[k];
            const uint32_t s = zs.start, e = zs.end;
            const bool nic = (zs.flags & ZCU_SPAN_NOTE_ID_CHANGED) != 0;
            co.gather(s, e, o.data());
            c0.gather(s, e, i0.data());
            Span sp{0, (size_t)(e - s)};
            if (kind == ZCU_MOD_SIMPLE_DELAY) {
                SimpleDelay m(1);
                m.delay = d;
                m.paint(sp, o.data(), nic, i0.data());
                d = m.delay;
            } else if (kind == ZCU_MOD_FILTERED_ECHOES) {
                if (zs.param >= nparams) return 2;
                const auto& p = ((const zcu_echoes_params*)params)[zs.param];
                FilteredEchoes m(1);
                m.delay = d;
                m.filter = flt;
                m.paint(sp, o.data(), t0.data(), t1.data(), nic, {i0.data(), p.feedback_volume, p.cutoff});
                d = m.delay;
                flt = m.filter;
            } else {
                return 1;
            }
            co.scatter(s, e, o.data());
        }
        for (uint32_t q = 0; q < delay_samples; q++) line[(size_t)q * V + v] = d.delay_buffer[q];
        if (kind == ZCU_MOD_SIMPLE_DELAY) {
            ((zcu_delay_state*)state)[v].index = (uint32_t)d.delay_buffer_index;
        } else {
            auto& st = ((zcu_echoes_state*)state)[v];
            st.index = (uint32_t)d.delay_buffer_index;
            st.l = flt.l;
            st.b = flt.b;
        }
    }
    return 0;
}

// StereoEchoes (examples/modules.zig:462-525) on V independent mono inputs: out_l/out_r/input are [frames][V]; one paint() call per
// mix buffer of buffer_frames (the un-triggered call pattern of example_delay.zig).  State starts from init() and is not returned.
int orc_stereo_echoes(uint32_t V, uint32_t frames, uint32_t buffer_frames, uint32_t main_delay, float feedback_volume, float cutoff,
                      const float* input, float* out_l, float* out_r) {
    std::vector<float> in(frames), l(frames), r(frames), t[4];
    for (auto& x : t) x.resize(frames);
    for (uint32_t v = 0; v < V; v++) {
        Col ci{(float*)input, V, v}, cl{out_l, V, v}, cr{out_r, V, v};
        ci.gather(0, frames, in.data());
        cl.gather(0, frames, l.data());
        cr.gather(0, frames, r.data());
        StereoEchoes m(main_delay);
        float* temps[4] = {t[0].data(), t[1].data(), t[2].data(), t[3].data()};
        for (uint32_t f = 0; f < frames; f += buffer_frames) {
            const Span sp{f, std::min<size_t>(frames, (size_t)f + buffer_frames)};
            m.paint(sp, l.data(), r.data(), temps, f == 0, in.data(), feedback_volume, cutoff);
        }
        cl.scatter(0, frames, l.data());
        cr.scatter(0, frames, r.data());
    }
    return 0;
}

// module init() for V voices (Noise: seed = first_seed + v, Noise.zig:26)
int orc_state_init(int kind, uint32_t V, uint32_t first_seed, void* state) {
    for (uint32_t v = 0; v < V; v++) {
        switch (kind) {
            case ZCU_MOD_SINE_OSC: ((zcu_sine_state*)state)[v] = {0.0f}; break;
            case ZCU_MOD_PULSE_OSC: ((zcu_pulse_state*)state)[v] = {0}; break;
            case ZCU_MOD_TRISAW_OSC: ((zcu_trisaw_state*)state)[v] = {0, 0.0f}; break;
            case ZCU_MOD_NOISE: {
                Noise m(first_seed + v);
                for (int q = 0; q < 4; q++) ((zcu_noise_state*)state)[v].s[q] = m.r.s[q];
            } break;
            case ZCU_MOD_ENVELOPE: ((zcu_envelope_state*)state)[v] = {ZCU_ENV_IDLE, {0.0f, 0.0f, 0.0f}}; break;
            case ZCU_MOD_FILTER: ((zcu_filter_state*)state)[v] = {0.0f, 0.0f}; break;
            case ZCU_MOD_PORTAMENTO: ((zcu_portamento_state*)state)[v] = {{0.0f, 0.0f, 0.0f}}; break;
            case ZCU_MOD_CURVE: ((zcu_curve_state*)state)[v] = {0.0f, 0, 0, 0}; break;
            case ZCU_MOD_SAMPLER: ((zcu_sampler_state*)state)[v] = {0.0f}; break;
            case ZCU_MOD_DECIMATOR: ((zcu_decimator_state*)state)[v] = {0.0f, 1.0f}; break;
            case ZCU_MOD_CYCLE: ((zcu_cycle_state*)state)[v] = {0.0f}; break;
            case ZCU_MOD_FM_OSCILLATOR: ((zcu_fm_oscillator_state*)state)[v] = {0.0f, 0.0f, 0.0f}; break;
            case ZCU_MOD_GATE:
            case ZCU_MOD_DISTORTION: break;
            default: return 1;
        }
    }
    return 0;
}

float orc_cutoff_from_frequency(float frequency, float sample_rate) { return Filter::cutoffFromFrequency(frequency, sample_rate); }

// ---- PRNG KAT hooks ----------------------------------------------------------------------------
void orc_xoshiro_seed(uint64_t seed, uint64_t s_out[4]) {
    Xoshiro256 r;
    r.seed(seed);
    for (int i = 0; i < 4; i++) s_out[i] = r.s[i];
}
void orc_xoshiro_next(uint64_t s[4], uint64_t* out, uint32_t n) {
    Xoshiro256 r;
    for (int i = 0; i < 4; i++) r.s[i] = s[i];
    for (uint32_t i = 0; i < n; i++) out[i] = r.next();
    for (int i = 0; i < 4; i++) s[i] = r.s[i];
}
void orc_xoshiro_floats(uint64_t s[4], float* out, uint32_t n) {
    Xoshiro256 r;
    for (int i = 0; i < 4; i++) r.s[i] = s[i];
    for (uint32_t i = 0; i < n; i++) out[i] = r.float32();
    for (int i = 0; i < 4; i++) s[i] = r.s[i];
}

// ---- event layer --------------------------------------------------------------------------------
struct OrcTrigger {
    int32_t has_note;
    uint32_t note_id;
    uint32_t param;
};

// Trigger(u32).counter + while(next) over one buffer span; impulses use buffer-relative frames.
uint32_t orc_trigger_run(OrcTrigger* t, uint32_t start, uint32_t end, const zcu_impulse* impulses,
                         const uint32_t* impulse_params, uint32_t n, uint32_t frame_base, zcu_span* out, uint32_t max_out) {
    std::vector<Impulse> imps(n);
    for (uint32_t i = 0; i < n; i++) imps[i] = Impulse{impulses[i].frame, impulses[i].note_id, impulses[i].event_id};
    Trigger<uint32_t> trig;
    trig.has_note = t->has_note != 0;
    trig.note = {t->note_id, t->param};
    auto ctr = trig.counter(Span{start, end}, {imps.data(), impulse_params, n});
    Trigger<uint32_t>::NewPaintReturnValue r;
    uint32_t cnt = 0;
    while (trig.next(ctr, r)) {
        if (cnt < max_out)
            out[cnt] = zcu_span{(uint32_t)r.span.start + frame_base, (uint32_t)r.span.end + frame_base,
                                r.note_id_changed ? ZCU_SPAN_NOTE_ID_CHANGED : 0u, r.params};
        cnt++;
    }
    t->has_note = trig.has_note ? 1 : 0;
    t->note_id = (uint32_t)trig.note.id;
    t->param = trig.note.params;
    return cnt;
}

int orc_poly_dispatch(uint32_t polyphony, zcu_poly_slot* slots, const zcu_impulse* impulses, const uint32_t* params,
                      const int32_t* note_on, uint32_t n, zcu_impulse* out_impulses, uint32_t* out_params, uint32_t* out_counts) {
#define ORC_PD(N) \
    case N: poly_dispatch_n<N>(slots, impulses, params, note_on, n, out_impulses, out_params, out_counts); return 0;
    switch (polyphony) {
        ORC_PD(1) ORC_PD(2) ORC_PD(3) ORC_PD(4) ORC_PD(5) ORC_PD(6) ORC_PD(7) ORC_PD(8) ORC_PD(9) ORC_PD(10) ORC_PD(11)
        ORC_PD(12) ORC_PD(13) ORC_PD(14) ORC_PD(15) ORC_PD(16) ORC_PD(32)
        default: return 1;
    }
#undef ORC_PD
}

struct OrcNoteTracker {
    uint32_t next_song_event;
    float t;
};
uint32_t orc_note_tracker_consume(OrcNoteTracker* nt, const zcu_song_event* song, uint32_t song_len, float sample_rate,
                                  uint32_t start, uint32_t end, zcu_impulse* out_impulses, uint32_t* out_params) {
    std::vector<Notes<uint32_t>::SongEvent> ev(song_len);
    for (uint32_t i = 0; i < song_len; i++) ev[i] = {song[i].param, song[i].t, song[i].note_id};
    Notes<uint32_t>::NoteTracker tr;
    tr.init(ev.data(), song_len);
    tr.next_song_event = nt->next_song_event;
    tr.t = nt->t;
    auto iap = tr.consume(sample_rate, Span{start, end});
    for (size_t i = 0; i < iap.len; i++) {
        out_impulses[i] = zcu_impulse{(uint32_t)iap.impulses[i].frame, (uint32_t)iap.impulses[i].note_id, (uint32_t)iap.impulses[i].event_id};
        out_params[i] = iap.paramses[i];
    }
    nt->next_song_event = (uint32_t)tr.next_song_event;
    nt->t = tr.t;
    return (uint32_t)iap.len;
}

void orc_impulse_queue_push(zcu_impulse_queue* q, uint32_t frame, uint32_t note_id, uint32_t param) {
    // ImpulseQueue.push (notes.zig:102-127) on the POD of zangcu.h
    Notes<uint32_t>::ImpulseQueue iq;
    iq.length = q->length;
    iq.next_event_id = q->next_event_id;
    for (uint32_t i = 0; i < q->length; i++) {
        iq.impulses_array[i] = Impulse{q->impulses[i].frame, q->impulses[i].note_id, q->impulses[i].event_id};
        iq.paramses_array[i] = q->params[i];
    }
    iq.push(frame, note_id, param);
    q->length = (uint32_t)iq.length;
    q->next_event_id = (uint32_t)iq.next_event_id;
    for (uint32_t i = 0; i < q->length; i++) {
        q->impulses[i] = zcu_impulse{(uint32_t)iq.impulses_array[i].frame, (uint32_t)iq.impulses_array[i].note_id,
                                     (uint32_t)iq.impulses_array[i].event_id};
        q->params[i] = iq.paramses_array[i];
    }
}

// ---- zang.mixDown -------------------------------------------------------------------------------
int orc_mix_down(const float* mix, uint32_t frames, uint32_t channels, int audio_format, float vol, uint8_t* dst) {
    std::vector<float> ch(frames);
    for (uint32_t c = 0; c < channels; c++) {
        for (uint32_t f = 0; f < frames; f++) ch[f] = mix[(size_t)f * channels + c];
        if (audio_format == ZCU_AUDIO_SIGNED16_LSB)
            mixDownS16LSB(dst, ch.data(), frames, channels, c, vol);
        else
            mixDownS8(dst, ch.data(), frames, channels, c, vol);
    }
    return 0;
}

// ---- voice mix (the shared-output accumulation of example_polyphony.zig:80-91, with pan gains) ----
int orc_mix_reduce(const float* src, uint32_t frames, uint32_t V, const float* gains, uint32_t channels, float* dst) {
    for (uint32_t f = 0; f < frames; f++)
        for (uint32_t v = 0; v < V; v++)
            for (uint32_t c = 0; c < channels; c++)
                dst[(size_t)f * channels + c] += (gains ? gains[(size_t)v * channels + c] : 1.0f) * src[(size_t)f * V + v];
    return 0;
}

// ---- instrument bank render: N voices of NiceInstrument driven by absolute-frame note events ------
// This is the restatement of the per-voice call structure of examples/example_polyphony.zig:61-92:
// per mix buffer, push that buffer's events into the voice's ImpulseQueue, consume, run the Trigger,
// paint the instrument on every returned span.  Per-voice output is then mixed with pan gains.
// voices_out (optional): [nframes][V] window starting at first_frame of the un-mixed voice outputs.
static void render_nice_range(uint32_t v0, uint32_t v1, uint32_t V, uint32_t total_frames, uint32_t buffer_frames,
                              float sample_rate, const uint32_t* event_off, const zcu_note_event* events, const float* color,
                              const float* gains, uint32_t channels, float* mix, float* voices_out, uint32_t first_frame,
                              uint32_t nframes) {
    std::vector<float> out(buffer_frames), t0(buffer_frames), t1(buffer_frames);
    for (uint32_t v = v0; v < v1; v++) {
        NiceInstrument instr(color[v]);
        Trigger<InstrParams> trig;
        Notes<InstrParams>::ImpulseQueue iq;
        uint32_t ei = event_off[v];
        const uint32_t ee = event_off[v + 1];
        for (uint32_t start = 0; start < total_frames; start += buffer_frames) {
            const uint32_t len = std::min(buffer_frames, total_frames - start);
            while (ei < ee && events[ei].frame < start + len) {
                iq.push(events[ei].frame - start, events[ei].note_id, InstrParams{sample_rate, events[ei].freq, events[ei].note_on != 0});
                ei++;
            }
            Span span{0, len};
            zero(span, out.data());
            auto ctr = trig.counter(span, iq.consume());
            Trigger<InstrParams>::NewPaintReturnValue r;
            float* temps[2] = {t0.data(), t1.data()};
            while (trig.next(ctr, r)) instr.paint(r.span, out.data(), temps, r.note_id_changed, r.params);
            if (mix) {
                for (uint32_t f = 0; f < len; f++)
                    for (uint32_t c = 0; c < channels; c++)
                        mix[(size_t)(start + f) * channels + c] += gains[(size_t)v * channels + c] * out[f];
            }
            if (voices_out) {
                for (uint32_t f = 0; f < len; f++) {
                    const uint32_t af = start + f;
                    if (af >= first_frame && af < first_frame + nframes) voices_out[(size_t)(af - first_frame) * V + v] = out[f];
                }
            }
        }
    }
}

// Returns wall seconds spent rendering.  With nthreads > 1 voices are partitioned over threads and
// each thread mixes into a private buffer (summed at the end, thread order).
double orc_render_nice(uint32_t V, uint32_t total_frames, uint32_t buffer_frames, float sample_rate, const uint32_t* event_off,
                       const zcu_note_event* events, const float* color, const float* gains, uint32_t channels, float* mix,
                       float* voices_out, uint32_t first_frame, uint32_t nframes, uint32_t nthreads) {
    auto t_begin = std::chrono::steady_clock::now();
    if (nthreads <= 1) {
        render_nice_range(0, V, V, total_frames, buffer_frames, sample_rate, event_off, events, color, gains, channels, mix,
                          voices_out, first_frame, nframes);
    } else {
        std::vector<std::vector<float>> partial(nthreads);
        std::vector<std::thread> th;
        for (uint32_t t = 0; t < nthreads; t++) {
            if (mix) partial[t].assign((size_t)total_frames * channels, 0.0f);
            const uint32_t a = (uint32_t)((uint64_t)V * t / nthreads), b = (uint32_t)((uint64_t)V * (t + 1) / nthreads);
            th.emplace_back(render_nice_range, a, b, V, total_frames, buffer_frames, sample_rate, event_off, events, color, gains,
                            channels, mix ? partial[t].data() : nullptr, voices_out, first_frame, nframes);
        }
        for (auto& x : th) x.join();
        if (mix)
            for (uint32_t t = 0; t < nthreads; t++)
                for (size_t i = 0; i < (size_t)total_frames * channels; i++) mix[i] += partial[t][i];
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
}

// ---- instrument banks of BASELINE.json configs[2..4] (bank_oracle.hpp) --------------------------------------------
// kind 0: SamplerVoice (voice_f = fake_sample_rate[V]; pcm = mono/multi-channel sample)   configs[3]
// kind 1: SfxVoice     (voice_i = filter type[V]; nodes = 3 curve nodes per job; seeds first_seed + v)   configs[4]
// kind 2: FmDetunedVoice (in0 = detune_in, in1 = index_in, [total_frames][V])   configs[2]
// Per voice and per mix buffer: push the buffer's events into the voice's ImpulseQueue, run the Trigger, paint.
// voices_out: [nframes][V] window of the un-mixed voice outputs starting at first_frame; mix (optional): += gains * voice.
double orc_render_bank(int kind, uint32_t V, uint32_t total_frames, uint32_t buffer_frames, float sample_rate,
                       const uint32_t* event_off, const zcu_note_event* events, const float* voice_f, const int32_t* voice_i,
                       const zcu_curve_node* nodes, const uint8_t* pcm, uint64_t pcm_len, uint32_t pcm_channels, uint32_t pcm_rate,
                       int pcm_format, const float* in0, const float* in1, uint32_t first_seed, const float* gains, uint32_t channels,
                       float* mix, float* voices_out, uint32_t first_frame, uint32_t nframes) {
    auto t_begin = std::chrono::steady_clock::now();
    std::vector<float> out(buffer_frames), a0(buffer_frames), a1(buffer_frames);
    std::vector<std::vector<float>> tb(6, std::vector<float>(buffer_frames));
    float* temps[6];
    for (int i = 0; i < 6; i++) temps[i] = tb[i].data();
    for (uint32_t v = 0; v < V; v++) {
        SamplerVoice sv;
        SfxVoice fx(first_seed + v);
        FmDetunedVoice fm;
        Trigger<InstrParams> trig;
        Notes<InstrParams>::ImpulseQueue iq;
        std::vector<CurveNode> cn(3);
        if (kind == 0) sv.fake_sample_rate = voice_f[v];
        if (kind == 1) {
            fx.type = (Filter::Type)voice_i[v];
            for (int i = 0; i < 3; i++) cn[i] = CurveNode{nodes[3 * v + i].value, nodes[3 * v + i].t};
            fx.nodes = cn.data();
        }
        const Sampler::Sample smp{pcm_channels, pcm_rate, (Sampler::Format)pcm_format, pcm, (size_t)pcm_len};
        uint32_t ei = event_off[v];
        const uint32_t ee = event_off[v + 1];
        for (uint32_t start = 0; start < total_frames; start += buffer_frames) {
            const uint32_t len = std::min(buffer_frames, total_frames - start);
            Span span{0, len};
            zero(span, out.data());
            while (ei < ee && events[ei].frame < start + len) {
                if (kind == 0)
                    sv.iq.push(events[ei].frame - start, events[ei].note_id, Sampler::Params{events[ei].freq, smp, 0, true});
                else
                    iq.push(events[ei].frame - start, events[ei].note_id, InstrParams{sample_rate, events[ei].freq, events[ei].note_on != 0});
                ei++;
            }
            if (kind == 0) {
                sv.paint(span, out.data(), temps, sample_rate);
            } else {
                if (kind == 2)
                    for (uint32_t f = 0; f < len; f++) {
                        a0[f] = in0[(size_t)(start + f) * V + v];
                        a1[f] = in1[(size_t)(start + f) * V + v];
                    }
                auto ctr = trig.counter(span, iq.consume());
                Trigger<InstrParams>::NewPaintReturnValue r;
                while (trig.next(ctr, r)) {
                    if (kind == 1)
                        fx.paint(r.span, out.data(), temps, r.note_id_changed, {r.params.sample_rate, r.params.note_on});
                    else
                        fm.paint(r.span, out.data(), temps, r.note_id_changed, r.params, a0.data(), a1.data());
                }
            }
            if (mix)
                for (uint32_t f = 0; f < len; f++)
                    for (uint32_t c = 0; c < channels; c++)
                        mix[(size_t)(start + f) * channels + c] += gains[(size_t)v * channels + c] * out[f];
            if (voices_out)
                for (uint32_t f = 0; f < len; f++) {
                    const uint32_t af = start + f;
                    if (af >= first_frame && af < first_frame + nframes) voices_out[(size_t)(af - first_frame) * V + v] = out[f];
                }
        }
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
}

// ---- the reference's own CPU workload: write_wav of the Toccata song (song_oracle.hpp) -------------
// Parsed song events of one instrument (0 Pedal, 1 RegularOrgan, 2 WeirdOrgan).  Returns the number of
// events (writes at most max_out), or -1 if the text does not parse.
int orc_song_events(const char* text, size_t len, uint32_t instrument, float* t_out, uint32_t* id_out, float* freq_out,
                    int32_t* note_on_out, uint32_t max_out) {
    Song song;
    song.parse(text, len);
    if (!song.ok || instrument >= Song::NUM_INSTRUMENTS) return -1;
    const auto& ev = song.all_notes[instrument];
    for (size_t i = 0; i < ev.size() && i < max_out; i++) {
        t_out[i] = ev[i].t;
        id_out[i] = (uint32_t)ev[i].note_id;
        freq_out[i] = ev[i].params.freq;
        note_on_out[i] = ev[i].params.note_on ? 1 : 0;
    }
    return (int)ev.size();
}

// write_wav.zig:49-95 for the first total_frames frames: pcm = s16le mono after mixDown(0.25); mix = f32 before it.
// Returns wall seconds, or a negative value if the song does not parse.
double orc_song_render(const char* text, size_t len, uint64_t total_frames, float sample_rate, uint8_t* pcm, float* mix) {
    Song song;
    song.parse(text, len);
    auto t_begin = std::chrono::steady_clock::now();
    if (!song_render(song, (size_t)total_frames, sample_rate, pcm, mix)) return -1.0;
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
}

// the same from already-parsed events (the GPU box has no /root/reference: tests ship the parsed score as a fixture).
// counts[3]; the four arrays hold the events of instrument 0, then 1, then 2.
double orc_song_render_events(const uint32_t* counts, const float* t, const uint32_t* id, const float* freq, const int32_t* note_on,
                              uint64_t total_frames, float sample_rate, uint8_t* pcm, float* mix) {
    Song song;
    size_t k = 0;
    for (size_t i = 0; i < Song::NUM_INSTRUMENTS; i++)
        for (uint32_t j = 0; j < counts[i]; j++, k++) song.all_notes[i].push_back(SongEvent{MyNoteParams{freq[k], note_on[k] != 0}, t[k], id[k]});
    song.ok = true;
    auto t_begin = std::chrono::steady_clock::now();
    if (!song_render(song, (size_t)total_frames, sample_rate, pcm, mix)) return -1.0;
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
}

uint32_t orc_hardware_threads(void) { return std::thread::hardware_concurrency(); }

}  // extern "C"
