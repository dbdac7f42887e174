// bank_oracle.hpp — TEST INFRASTRUCTURE ONLY (see zang_oracle.hpp).
//
// CPU restatement of the instrument graphs of BASELINE.json configs[2..4], one voice at a time, written the way a zang
// example writes an instrument: a struct of modules with a paint(span, outputs, temps, note_id_changed, params) body.
// The graphs are synthetic composites of the reference's examples; the device path composes the very
// same graphs from C-ABI module paints (zang_b200/banks.py).
#pragma once

#include "zang_oracle.hpp"

namespace zo {

// configs[3]: example_sampler.zig:94-118 with the Decimator of example_polyphony.zig:147-160 between sampler and distortion
struct SamplerVoice {
    Sampler sampler;
    Decimator decimator;
    Distortion distortion;
    Trigger<Sampler::Params> trigger;
    Notes<Sampler::Params>::ImpulseQueue iq;
    float fake_sample_rate = 6000.0f;

    // one mix buffer: span is the whole buffer, impulses already pushed into iq
    void paint(Span span, float* out, float** temps, float sample_rate) {
        zero(span, temps[0]);
        auto ctr = trigger.counter(span, iq.consume());
        Trigger<Sampler::Params>::NewPaintReturnValue r;
        while (trigger.next(ctr, r)) sampler.paint(r.span, temps[0], r.note_id_changed, r.params);
        multiplyWithScalar(span, temps[0], 2.5f);
        zero(span, temps[1]);
        decimator.paint(span, temps[1], false, {sample_rate, temps[0], fake_sample_rate});
        distortion.paint(span, out, false, {temps[1], Distortion::overdrive, 0.9f, 0.5f, 0.0f});
    }
};

// configs[4]: Noise -> Filter with a Curve-driven cutoff -> x Envelope (example_laser.zig:77-115, example_stereo.zig:71-98)
struct SfxVoice {
    struct Params {
        float sample_rate;
        bool note_on;
    };
    Noise noise;
    Curve curve;
    Filter flt;
    Envelope env;
    Filter::Type type = Filter::low_pass;
    const CurveNode* nodes = nullptr;  // 3 nodes
    explicit SfxVoice(uint32_t seed) : noise(seed) {}
    void paint(Span span, float* out, float** temps, bool nic, const Params& p) {
        zero(span, temps[0]);
        noise.paint(span, temps[0], nic, {Noise::white});
        zero(span, temps[1]);
        curve.paint(span, temps[1], nic, {p.sample_rate, Curve::smoothstep, nodes, 3});
        zero(span, temps[2]);
        flt.paint(span, temps[2], nic, {temps[0], type, buffer(temps[1]), constant(0.4f)});
        zero(span, temps[0]);
        env.paint(span, temps[0], nic,
                  {p.sample_rate, {CurveKind::linear, 0.005f}, {CurveKind::squared, 0.05f}, {CurveKind::cubed, 0.2f}, 0.6f, p.note_on});
        multiply(span, out, temps[2], temps[0]);
    }
};

// configs[2]: Portamento -> detuned TriSawOsc + phase-modulated SineOsc pair -> x Envelope
// (examples/modules.zig:42-75, example_detuned.zig:61-99, example_portamento.zig:20-87)
struct FmDetunedVoice {
    SineOsc modulator, carrier;
    TriSawOsc saw;
    Portamento porta;
    Envelope env;
    bool prev_note_on = false;
    void paint(Span span, float* out, float** temps, bool nic, const InstrParams& p, const float* detune_in, const float* index_in) {
        zero(span, temps[0]);
        porta.paint(span, temps[0], nic, {p.sample_rate, {CurveKind::cubed, 0.5f}, p.freq, p.note_on, prev_note_on});
        zero(span, temps[1]);
        multiply(span, temps[1], temps[0], detune_in);
        zero(span, temps[2]);
        saw.paint(span, temps[2], nic, {p.sample_rate, buffer(temps[1]), 0.0f});
        zero(span, temps[3]);
        modulator.paint(span, temps[3], nic, {p.sample_rate, buffer(temps[0]), constant(0.0f)});
        multiplyWith(span, temps[3], index_in);
        zero(span, temps[4]);
        carrier.paint(span, temps[4], nic, {p.sample_rate, buffer(temps[0]), buffer(temps[3])});
        zero(span, temps[5]);
        const bool new_note = !prev_note_on && p.note_on;  // example_portamento.zig:66
        env.paint(span, temps[5], new_note,
                  {p.sample_rate, {CurveKind::cubed, 0.025f}, {CurveKind::cubed, 0.1f}, {CurveKind::cubed, 1.0f}, 0.5f, p.note_on});
        multiply(span, out, temps[2], temps[5]);
        multiply(span, out, temps[4], temps[5]);
        prev_note_on = p.note_on;  // `defer self.prev_note_on = params.note_on` (example_portamento.zig:51)
    }
};

}  // namespace zo
