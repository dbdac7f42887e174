// zang_oracle.hpp — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement (C++17, strict IEEE f32: build with -ffp-contract=off, never -ffast-math) of the
// buffer-painting path of dbandstra/zang @30bed1e.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may build, link or call anything under oracle/.
// The product (zang_b200/, libzangcu.so) never includes this file.
//
// Pinning status:
//   * event splitting (Trigger, PolyphonyDispatcher): PINNED by the reference's own tests
//     src/zang/trigger_test.zig:38-132 and src/zang/notes_test.zig:10-119 (tests/test_event_kat.py).
//   * xoshiro256++ / SplitMix64: pinned by the Zig-std known-answer vector quoted in SURVEY.md §8(c).
//   * every paint() loop: PARITY UNPINNED by the reference (it has no test or golden audio for any
//     module); no Zig toolchain exists in this image, so the reference cannot be run.
//     The restatement follows the Zig source statement by statement; each function cites it.
//
// The reference works on one voice and one contiguous []f32 at a time; so does this file.  The
// [frame][voice] batching lives in oracle_capi.cpp.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace zo {

// ------------------------------------------------------------------------------------------
// src/zang/basics.zig
// ------------------------------------------------------------------------------------------
struct Span {  // basics.zig:3-10
    size_t start, end;
};

inline void zero(Span s, float* dest) {  // basics.zig:12-14
    for (size_t i = s.start; i < s.end; i++) dest[i] = 0.0f;
}
inline void set(Span s, float* dest, float a) {  // basics.zig:16-18
    for (size_t i = s.start; i < s.end; i++) dest[i] = a;
}
inline void copy(Span s, float* dest, const float* src) {  // basics.zig:20-22
    for (size_t i = s.start; i < s.end; i++) dest[i] = src[i];
}
inline void add(Span s, float* dest, const float* a, const float* b) {  // basics.zig:24-29
    for (size_t i = s.start; i < s.end; i++) dest[i] += a[i] + b[i];
}
inline void addInto(Span s, float* dest, const float* src) {  // basics.zig:31-36
    for (size_t i = s.start; i < s.end; i++) dest[i] += src[i];
}
inline void addScalar(Span s, float* dest, const float* a, float b) {  // basics.zig:38-43
    for (size_t i = s.start; i < s.end; i++) dest[i] += a[i] + b;
}
inline void addScalarInto(Span s, float* dest, float a) {  // basics.zig:45-50
    for (size_t i = s.start; i < s.end; i++) dest[i] += a;
}
inline void multiply(Span s, float* dest, const float* a, const float* b) {  // basics.zig:52-57
    for (size_t i = s.start; i < s.end; i++) dest[i] += a[i] * b[i];
}
inline void multiplyWith(Span s, float* dest, const float* a) {  // basics.zig:59-64
    for (size_t i = s.start; i < s.end; i++) dest[i] *= a[i];
}
inline void multiplyScalar(Span s, float* dest, const float* a, float b) {  // basics.zig:66-71
    for (size_t i = s.start; i < s.end; i++) dest[i] += a[i] * b;
}
inline void multiplyWithScalar(Span s, float* dest, float a) {  // basics.zig:73-78
    for (size_t i = s.start; i < s.end; i++) dest[i] *= a;
}

// src/zang/constant_or_buffer.zig:4-15
struct ConstantOrBuffer {
    bool is_buffer;
    float constant;
    const float* buffer;
};
inline ConstantOrBuffer constant(float x) { return {false, x, nullptr}; }
inline ConstantOrBuffer buffer(const float* b) { return {true, 0.0f, b}; }

// src/zang/curve.zig:3-6
struct CurveNode {
    float value, t;
};

// ------------------------------------------------------------------------------------------
// src/zang/painter.zig
// ------------------------------------------------------------------------------------------
struct PaintState {  // painter.zig:11-23
    float* buf;
    size_t len;
    size_t i;
    float sample_rate;
};
enum class CurveKind { instantaneous = 0, linear = 1, squared = 2, cubed = 3 };
struct PaintCurve {  // painter.zig:25-30
    CurveKind kind;
    float duration;
};

struct Painter {  // painter.zig:33-120
    float t = 0.0f, last_value = 0.0f, start = 0.0f;

    void newCurve() {  // :47-50
        start = last_value;
        t = 0.0f;
    }
    void paintFlat(PaintState& st, float value) {  // :53-58
        addScalarInto(Span{st.i, st.len}, st.buf, value);
        st.i = st.len;
    }
    bool paintToward(PaintState& st, PaintCurve curve, float goal) {  // :63-120
        if (t >= 1.0f) return true;  // :69-71
        if (curve.kind == CurveKind::instantaneous) {  // :76-80
            t = 1.0f;
            last_value = goal;
            return true;
        }
        const float duration = curve.duration;
        size_t i = st.i;
        const float t_step = 1.0f / (duration * st.sample_rate);  // :97
        bool finished = false;
        while (!finished && i < st.len) {  // :102-116
            t += t_step;
            if (t >= 1.0f) {
                t = 1.0f;
                finished = true;
            }
            const float it = 1.0f - t;
            float tp;
            switch (curve.kind) {
                case CurveKind::linear: tp = t; break;
                case CurveKind::squared: tp = 1.0f - it * it; break;
                default: tp = 1.0f - it * it * it; break;
            }
            last_value = start + tp * (goal - start);
            st.buf[i] += last_value;
            i += 1;
        }
        st.i = i;
        return finished;
    }
};

// ------------------------------------------------------------------------------------------
// Zig std.rand.DefaultPrng = Xoshiro256 (xoshiro256++) seeded through SplitMix64, and
// std.rand.Random.float(f32).  NOT in /root/reference (Zig std 0.12/0.13, README.md:2); restated
// from the published algorithms.  KAT: seed 0 -> 0x53175d61490b23df ... (tests/test_prng_kat.py).
// ------------------------------------------------------------------------------------------
struct Xoshiro256 {
    uint64_t s[4];
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    void seed(uint64_t init_s) {
        uint64_t sm = init_s;
        for (int i = 0; i < 4; i++) {
            sm += 0x9e3779b97f4a7c15ULL;
            uint64_t z = sm;
            z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
            z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
            s[i] = z ^ (z >> 31);
        }
    }
    uint64_t next() {
        const uint64_t r = rotl(s[0] + s[3], 23) + s[0];
        const uint64_t t = s[1] << 17;
        s[2] ^= s[0];
        s[3] ^= s[1];
        s[1] ^= s[2];
        s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    static int clz64(uint64_t x) { return x == 0 ? 64 : __builtin_clzll(x); }
    static int clz32(uint32_t x) { return x == 0 ? 32 : __builtin_clz(x); }
    // Random.float(f32): 23 mantissa bits from the low end, exponent from the leading-zero count.
    float float32() {
        const uint64_t r = next();
        int lz = clz64(r);
        if (lz >= 41) {
            lz = 41 + clz64(next());
            if (lz == 41 + 64) lz += clz32((uint32_t)next() | 0x7FFu);
        }
        const uint32_t mantissa = (uint32_t)r & 0x7FFFFFu;
        const uint32_t bits = ((uint32_t)(126 - lz) << 23) | mantissa;
        float f;
        std::memcpy(&f, &bits, 4);
        return f;
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/SineOsc.zig
// ------------------------------------------------------------------------------------------
static const float kPi = 3.14159265358979323846f;  // std.math.pi coerced to f32

struct SineOsc {
    struct Params {
        float sample_rate;
        ConstantOrBuffer freq, phase;
    };
    float t = 0.0f;
    static float sin2pi(float x) { return sinf(x * kPi * 2.0f); }  // SineOsc.zig:4-6
    void paint(Span span, float* out0, bool, const Params& p) {    // SineOsc.zig:24-87
        float* output = out0 + span.start;
        const size_t len = span.end - span.start;
        float tt = t;
        if (!p.freq.is_buffer) {
            const float t_step = p.freq.constant / p.sample_rate;  // :44
            if (!p.phase.is_buffer) {
                const float phase = p.phase.constant;
                for (size_t i = 0; i < len; i++) {  // :49-52
                    output[i] += sin2pi(tt + phase);
                    tt += t_step;
                }
            } else {
                const float* ph = p.phase.buffer + span.start;
                for (size_t i = 0; i < len; i++) {  // :57-60
                    output[i] += sin2pi(tt + ph[i]);
                    tt += t_step;
                }
            }
        } else {
            const float* fr = p.freq.buffer + span.start;
            const float inv_sample_rate = 1.0f / p.sample_rate;  // :66
            if (!p.phase.is_buffer) {
                const float phase = p.phase.constant;
                for (size_t i = 0; i < len; i++) {  // :71-74
                    output[i] += sin2pi(tt + phase);
                    tt += fr[i] * inv_sample_rate;
                }
            } else {
                const float* ph = p.phase.buffer + span.start;
                for (size_t i = 0; i < len; i++) {  // :79-82
                    output[i] += sin2pi(tt + ph[i]);
                    tt += fr[i] * inv_sample_rate;
                }
            }
        }
        t = tt - truncf(tt);  // :40 (defer)
    }
};

// ------------------------------------------------------------------------------------------
// helpers shared by PulseOsc.zig:12-26 and TriSawOsc.zig:8-26
// ------------------------------------------------------------------------------------------
static const float fc32bit = 4294967296.0f;
inline float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }
inline float utof23(uint32_t x) {
    const uint32_t bits = (x >> 9) | 0x3f800000u;
    float f;
    std::memcpy(&f, &bits, 4);
    return f - 1.0f;
}
inline uint32_t ftou32(float v) { return (uint32_t)(v * fc32bit * 0.99995f); }

// ------------------------------------------------------------------------------------------
// src/modules/PulseOsc.zig
// ------------------------------------------------------------------------------------------
struct PulseOsc {
    struct Params {
        float sample_rate;
        ConstantOrBuffer freq;
        float color;
    };
    uint32_t cnt = 0;

    void paint(Span span, float* out0, bool, const Params& p) {  // :44-73
        if (!p.freq.is_buffer)
            paintConstantFrequency(out0 + span.start, span.end - span.start, p.sample_rate, p.freq.constant, p.color);
        else
            paintControlledFrequency(out0 + span.start, span.end - span.start, p.sample_rate,
                                     p.freq.buffer + span.start, p.color);
    }
    void paintConstantFrequency(float* output, size_t len, float sample_rate, float freq, float color) {  // :75-114
        if (freq < 0.0f || freq > sample_rate / 8.0f) return;  // :82-84
        uint32_t c = cnt;
        const float SRfcobasefrq = fc32bit / sample_rate;
        const uint32_t ifreq = (uint32_t)(SRfcobasefrq * freq);
        const uint32_t brpt = ftou32(clamp01(color));
        const float gain = 0.7f;
        const float gdf = gain / utof23(ifreq);
        const float col = utof23(brpt);
        const float cc121 = gdf * 2.0f * (col - 1.0f) + gain;
        const float cc212 = gdf * 2.0f * col - gain;
        uint32_t state = ((uint32_t)(c - ifreq) < brpt) ? 3u : 0u;  // :96
        for (size_t i = 0; i < len; i++) {                           // :98-112
            const float pp = utof23(c);
            state = ((state << 1) | (c < brpt ? 1u : 0u)) & 3u;
            const uint32_t transition = state | ((c < ifreq ? 1u : 0u) << 2);
            float v;
            switch (transition) {
                case 3: v = gain; break;
                case 0: v = -gain; break;
                case 2: v = gdf * 2.0f * (col - pp) + gain; break;
                case 5: v = gdf * 2.0f * pp - gain; break;
                case 7: v = cc121; break;
                case 4: v = cc212; break;
                default: v = 0.0f; break;  // unreachable in the reference
            }
            output[i] += v;
            c += ifreq;
        }
        cnt = c;
    }
    void paintControlledFrequency(float* output, size_t len, float sample_rate, const float* freq, float color) {  // :116-157
        uint32_t c = cnt;
        const float SRfcobasefrq = fc32bit / sample_rate;
        const uint32_t brpt = ftou32(clamp01(color));
        const float gain = 0.7f;
        const float col = utof23(brpt);
        for (size_t i = 0; i < len; i++) {
            const float s_freq = freq[i];
            if (s_freq < 0.0f || s_freq > sample_rate / 8.0f) continue;  // :134-135
            const uint32_t ifreq = (uint32_t)(SRfcobasefrq * s_freq);
            const float gdf = gain / utof23(ifreq);
            const float cc121 = gdf * 2.0f * (col - 1.0f) + gain;
            const float cc212 = gdf * 2.0f * col - gain;
            const float pp = utof23(c);
            const uint32_t cbit = ((uint32_t)(c - ifreq) < brpt) ? 1u : 0u;
            const uint32_t state = (c < brpt ? 1u : 0u) | (cbit << 1);
            const uint32_t transition = state | ((c < ifreq ? 1u : 0u) << 2);
            float v;
            switch (transition) {
                case 3: v = gain; break;
                case 0: v = -gain; break;
                case 2: v = gdf * 2.0f * (col - pp) + gain; break;
                case 5: v = gdf * 2.0f * pp - gain; break;
                case 7: v = cc121; break;
                case 4: v = cc212; break;
                default: v = 0.0f; break;
            }
            output[i] += v;
            c += ifreq;
        }
        cnt = c;
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/TriSawOsc.zig
// ------------------------------------------------------------------------------------------
struct TriSawOsc {
    using Params = PulseOsc::Params;
    uint32_t cnt = 0;
    float t = 0.0f;
    static float sqr(float v) { return v * v; }

    void paint(Span span, float* out0, bool, const Params& p) {  // :46-75
        if (!p.freq.is_buffer)
            paintConstantFrequency(out0 + span.start, span.end - span.start, p.sample_rate, p.freq.constant, p.color);
        else
            paintControlledFrequency(out0 + span.start, span.end - span.start, p.sample_rate,
                                     p.freq.buffer + span.start, p.color);
    }
    void paintConstantFrequency(float* output, size_t len, float sample_rate, float freq, float color) {  // :77-118
        if (freq < 0.0f || freq > sample_rate / 8.0f) return;
        uint32_t c = cnt;
        const float SRfcobasefrq = fc32bit / sample_rate;
        const uint32_t ifreq = (uint32_t)(SRfcobasefrq * freq);
        const uint32_t brpt = ftou32(clamp01(color));
        const float gain = 0.7f;
        const float f = utof23(ifreq);
        const float omf = 1.0f - f;
        const float rcpf = 1.0f / f;
        const float col = utof23(brpt);
        const float c1 = gain / col;
        const float c2 = -gain / (1.0f - col);
        uint32_t state = ((uint32_t)(c - ifreq) < brpt) ? 3u : 0u;
        for (size_t i = 0; i < len; i++) {  // :102-116
            const float p = utof23(c) - col;
            state = ((state << 1) | (c < brpt ? 1u : 0u)) & 3u;
            const uint32_t s = state | ((c < ifreq ? 1u : 0u) << 2);
            float v;
            switch (s) {
                case 3: v = c1 * (p + p - f); break;
                case 0: v = c2 * (p + p - f); break;
                case 2: v = rcpf * (c2 * sqr(p) - c1 * sqr(p - f)); break;
                case 5: v = -rcpf * (gain + c2 * sqr(p + omf) - c1 * sqr(p)); break;
                case 7: v = -rcpf * (gain + c1 * omf * (p + p + omf)); break;
                case 4: v = -rcpf * (gain + c2 * omf * (p + p + omf)); break;
                default: v = 0.0f; break;
            }
            output[i] += gain + v;  // :106
            c += ifreq;
        }
        cnt = c;
    }
    void paintControlledFrequency(float* output, size_t len, float sample_rate, const float* freq, float color) {  // :120-156
        float tt = t;
        const float gain = 0.7f;
        for (size_t i = 0; i < len; i++) {
            float frac;
            if (color < 0.25f || color > 0.75f) {
                frac = (tt - floorf(tt)) * 2.0f - 1.0f;  // sawtooth :139
            } else {
                frac = tt - floorf(tt);  // triangle :142-149
                if (frac < 0.25f)
                    frac = frac * 4.0f;
                else if (frac < 0.75f)
                    frac = 1.0f - (frac - 0.25f) * 4.0f;
                else
                    frac = (frac - 0.75f) * 4.0f - 1.0f;
            }
            output[i] += gain * frac;
            tt += freq[i] / sample_rate;  // :152
        }
        t = tt - truncf(tt);  // :155
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Noise.zig
// ------------------------------------------------------------------------------------------
struct Noise {
    enum Color { white = 0, pink = 1 };
    struct Params {
        Color color;
    };
    Xoshiro256 r;
    float b[7] = {0, 0, 0, 0, 0, 0, 0};
    explicit Noise(uint32_t seed = 0) { r.seed((uint64_t)seed); }  // :25-32; caller supplies the counter value
    void paint(Span span, float* out, bool, const Params& p) {     // :34-72
        Xoshiro256 rr = r;
        if (p.color == white) {
            for (size_t i = span.start; i < span.end; i++) out[i] += rr.float32() * 2.0f - 1.0f;  // :51
        } else {
            float bb[7];
            for (int k = 0; k < 7; k++) bb[k] = b[k];
            for (size_t i = span.start; i < span.end; i++) {  // :57-67
                const float wh = rr.float32() * 2.0f - 1.0f;
                bb[0] = 0.99886f * bb[0] + wh * 0.0555179f;
                bb[1] = 0.99332f * bb[1] + wh * 0.0750759f;
                bb[2] = 0.96900f * bb[2] + wh * 0.1538520f;
                bb[3] = 0.86650f * bb[3] + wh * 0.3104856f;
                bb[4] = 0.55000f * bb[4] + wh * 0.5329522f;
                bb[5] = -0.7616f * bb[5] - wh * 0.0168980f;
                out[i] += bb[0] + bb[1] + bb[2] + bb[3] + bb[4] + bb[5] + bb[6] + wh * 0.5362f;
                bb[6] = wh * 0.115926f;
            }
            // Noise.zig:68 assigns `b = self.b` (the wrong way round): self.b is never updated.
        }
        r = rr;  // :71
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Envelope.zig
// ------------------------------------------------------------------------------------------
struct Envelope {
    struct Params {
        float sample_rate;
        PaintCurve attack, decay, release;
        float sustain_volume;
        bool note_on;
    };
    enum State { idle = 0, attack = 1, decay = 2, sustain = 3, release = 4 };
    State state = idle;
    Painter painter;

    void changeState(State ns) {  // :33-36
        state = ns;
        painter.newCurve();
    }
    void paintOn(float* buf, size_t len, const Params& p, bool new_note) {  // :38-73
        PaintState ps{buf, len, 0, p.sample_rate};
        if (new_note) changeState(attack);
        // :45 asserts state != release (undefined in ReleaseFast; generators avoid it)
        if (state == idle) changeState(attack);
        if (state == attack) {
            if (painter.paintToward(ps, p.attack, 1.0f)) {
                if (p.sustain_volume < 1.0f)
                    changeState(decay);
                else
                    changeState(sustain);
            }
        }
        if (state == decay) {
            if (painter.paintToward(ps, p.decay, p.sustain_volume)) changeState(sustain);
        }
        if (state == sustain) painter.paintFlat(ps, p.sustain_volume);
    }
    void paintOff(float* buf, size_t len, const Params& p) {  // :77-90
        if (state == idle) return;
        if (state != release) changeState(release);
        PaintState ps{buf, len, 0, p.sample_rate};
        if (painter.paintToward(ps, p.release, 0.0f)) changeState(idle);
    }
    void paint(Span span, float* out0, bool note_id_changed, const Params& p) {  // :92-109
        float* output = out0 + span.start;
        const size_t len = span.end - span.start;
        if (p.note_on)
            paintOn(output, len, p, note_id_changed);
        else
            paintOff(output, len, p);
    }
};

// src/modules/Gate.zig:15-30
struct Gate {
    struct Params {
        bool note_on;
    };
    void paint(Span span, float* out0, bool, const Params& p) {
        if (p.note_on) addScalarInto(span, out0, 1.0f);
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Filter.zig
// ------------------------------------------------------------------------------------------
inline float zclamp(float v, float lo, float hi) { return std::max(lo, std::min(v, hi)); }  // std.math.clamp

struct Filter {
    enum Type { bypass = 0, low_pass = 1, band_pass = 2, high_pass = 3, notch = 4, all_pass = 5 };
    struct Params {
        const float* input;
        Type type;
        ConstantOrBuffer cutoff, res;
    };
    float l = 0.0f, b = 0.0f;

    static float cutoffFromFrequency(float frequency, float sample_rate) {  // :20-23
        const float v = 2.0f * (1.0f - cosf(kPi * frequency / sample_rate));
        return sqrtf(zclamp(v, 0.0f, 1.0f));
    }
    void paint(Span span, float* out0, bool, const Params& p) {  // :44-72
        float* output = out0 + span.start;
        const float* input = p.input + span.start;
        const size_t len = span.end - span.start;
        paintFunction(output, input, len, p.type, !p.cutoff.is_buffer, p.cutoff.constant,
                      p.cutoff.is_buffer ? p.cutoff.buffer + span.start : nullptr, !p.res.is_buffer, p.res.constant,
                      p.res.is_buffer ? p.res.buffer + span.start : nullptr);
    }
    void paintFunction(float* output, const float* input, size_t len, Type type, bool cutoff_is_constant,
                       float cutoff_constant, const float* cutoff_buffer, bool res_is_constant, float res_constant,
                       const float* res_buffer) {  // :74-151
        const float fcdcoffset = 3.814697265625e-6f;  // :8
        float l_mul = 0.0f, b_mul = 0.0f, h_mul = 0.0f;
        switch (type) {
            case bypass:
                for (size_t i = 0; i < len; i++) output[i] += input[i];  // :91-97
                return;
            case low_pass: l_mul = 1.0f; break;
            case band_pass: b_mul = 1.0f; break;
            case high_pass: h_mul = 1.0f; break;
            case notch: l_mul = 1.0f; h_mul = 1.0f; break;
            case all_pass: l_mul = 1.0f; b_mul = 1.0f; h_mul = 1.0f; break;
        }
        float cut = 0.0f, res = 0.0f;
        if (cutoff_is_constant) cut = zclamp(cutoff_constant, 0.0f, 1.0f);
        if (res_is_constant) res = 1.0f - zclamp(res_constant, 0.0f, 1.0f);
        float ll = l, bb = b;
        for (size_t i = 0; i < len; i++) {  // :124-147
            if (!cutoff_is_constant) cut = zclamp(cutoff_buffer[i], 0.0f, 1.0f);
            if (!res_is_constant) res = 1.0f - zclamp(res_buffer[i], 0.0f, 1.0f);
            const float in = input[i] + fcdcoffset;
            ll += cut * bb - fcdcoffset;        // step 1 :138
            bb += cut * (in - bb * res - ll);   // :139
            ll += cut * bb;                     // step 2 :142
            const float h = in - bb * res - ll; // :143
            bb += cut * h;                      // :144
            output[i] += ll * l_mul + bb * b_mul + h * h_mul;
        }
        l = ll;
        b = bb;
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Portamento.zig:21-48
// ------------------------------------------------------------------------------------------
struct Portamento {
    struct Params {
        float sample_rate;
        PaintCurve curve;
        float goal;
        bool note_on, prev_note_on;
    };
    Painter painter;
    void paint(Span span, float* out0, bool note_id_changed, const Params& p) {
        float* output = out0 + span.start;
        const size_t len = span.end - span.start;
        PaintCurve curve = (p.note_on && p.prev_note_on) ? p.curve : PaintCurve{CurveKind::instantaneous, 0.0f};
        if (p.note_on && note_id_changed) painter.newCurve();
        PaintState st{output, len, 0, p.sample_rate};
        if (painter.paintToward(st, curve, p.goal)) painter.paintFlat(st, p.goal);
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Curve.zig
// ------------------------------------------------------------------------------------------
struct Curve {
    enum Function { linear = 0, smoothstep = 1 };
    struct Params {
        float sample_rate;
        Function function;
        const CurveNode* curve;
        size_t curve_len;
    };
    struct CurveSpanNode {
        int32_t frame;
        float value;
    };
    struct CurveSpan {
        size_t start, end;
        bool has_values;
        CurveSpanNode start_node, end_node;
    };
    float t = 0.0f;
    size_t current_song_note = 0;
    int32_t current_song_note_offset = 0;
    size_t next_song_note = 0;
    CurveSpanNode curve_nodes[32];

    void paint(Span span, float* out0, bool note_id_changed, const Params& p) {  // :56-128
        if (note_id_changed) {  // :66-71
            current_song_note = 0;
            current_song_note_offset = 0;
            next_song_note = 0;
            t = 0.0f;
        }
        float* out = out0 + span.start;
        const size_t out_len = span.end - span.start;
        const size_t ncn = getCurveSpanNodes(p.sample_rate, out_len, p.curve, p.curve_len);
        size_t start = 0;
        while (start < out_len) {
            const CurveSpan cs = getNextCurveSpan(curve_nodes, ncn, start, out_len);
            const int32_t paint_start = (int32_t)cs.start;
            if (cs.has_values) {
                const int32_t fstart = cs.start_node.frame;
                const int32_t fend = cs.end_node.frame;
                const float start_x = (float)(paint_start - fstart) / (float)(fend - fstart);  // :95
                const float start_value = cs.start_node.value;
                const float value_delta = cs.end_node.value - cs.start_node.value;
                const float x_step = 1.0f / (float)(fend - fstart);  // :100
                size_t i = cs.start;
                if (p.function == linear) {  // :105-113
                    float y = start_value + start_x * value_delta;
                    const float y_step = x_step * value_delta;
                    for (; i < cs.end; i++) {
                        out[i] += y;
                        y += y_step;
                    }
                } else {  // :114-122
                    float x = start_x;
                    for (; i < cs.end; i++) {
                        const float v = x * x * (3.0f - 2.0f * x) * value_delta;
                        out[i] += start_value + v;
                        x += x_step;
                    }
                }
            }
            start = cs.end;
        }
    }
    size_t getCurveSpanNodes(float sample_rate, size_t out_len, const CurveNode* nodes, size_t nnodes) {  // :130-184
        size_t count = 0;
        const float buf_time = (float)out_len / sample_rate;
        const float end_t = t + buf_time;
        if (current_song_note < next_song_note) {  // :142-148
            curve_nodes[count] = {current_song_note_offset, nodes[current_song_note].value};
            count += 1;
        }
        bool one_past = false;
        for (size_t k = next_song_note; k < nnodes; k++) {  // :151-178 (slice taken once, before the loop)
            const float note_t = nodes[k].t;
            if (note_t >= end_t) {
                if (!one_past)
                    one_past = true;
                else
                    break;
            }
            const float f = (note_t - t) / buf_time;
            const int32_t rel_frame_index = (int32_t)(f * (float)out_len);
            if (count > 0 && curve_nodes[count - 1].frame == rel_frame_index) count -= 1;
            curve_nodes[count] = {rel_frame_index, nodes[k].value};
            count += 1;
            if (!one_past) {
                current_song_note = next_song_note;
                current_song_note_offset = 0;
                next_song_note += 1;
            }
        }
        t += buf_time;                                  // :180
        current_song_note_offset -= (int32_t)out_len;   // :181
        return count;
    }
    static CurveSpan getNextCurveSpan(const CurveSpanNode* cn, size_t n, size_t dest_start_, size_t dest_end_) {  // :188-255
        const int32_t dest_start = (int32_t)dest_start_;
        const int32_t dest_end = (int32_t)dest_end_;
        for (size_t i = 0; i < n; i++) {
            const int32_t start_pos = cn[i].frame;
            if (start_pos >= dest_end) break;
            const int32_t end_pos = (i + 1 < n) ? std::min(dest_end, cn[i + 1].frame) : dest_end;
            if (end_pos <= dest_start) continue;
            const int32_t note_start_clipped = start_pos > dest_start ? start_pos : dest_start;
            if (note_start_clipped > dest_start) return {(size_t)dest_start, (size_t)note_start_clipped, false, {}, {}};
            const int32_t note_end = end_pos;
            const int32_t note_end_clipped = note_end > dest_end ? dest_end : note_end;
            if (i + 1 < n) return {(size_t)note_start_clipped, (size_t)note_end_clipped, true, cn[i], cn[i + 1]};
            return {(size_t)note_start_clipped, (size_t)note_end_clipped, false, {}, {}};
        }
        return {(size_t)dest_start, (size_t)dest_end, false, {}, {}};
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Sampler.zig
// ------------------------------------------------------------------------------------------
struct Sampler {
    enum Format { unsigned8 = 0, signed16_lsb = 1, signed24_lsb = 2, signed32_lsb = 3 };
    struct Sample {
        size_t num_channels;
        size_t sample_rate;
        Format format;
        const uint8_t* data;
        size_t data_len;
    };
    struct Params {
        float sample_rate;
        Sample sample;
        size_t channel;
        bool loop;
    };
    float t = 0.0f;

    static float decodeSigned(int byte_count, const uint8_t* slice, size_t index) {  // :23-33
        const uint8_t* sub = slice + index * byte_count;
        int64_t sval = 0;
        for (int k = 0; k < byte_count; k++) sval |= (int64_t)sub[k] << (8 * k);
        const int bits = byte_count * 8;
        if (sval & ((int64_t)1 << (bits - 1))) sval -= ((int64_t)1 << bits);  // sign extend
        const int64_t maxv = (int64_t)1 << (bits - 1);
        return (float)(int32_t)sval / (float)maxv;
    }
    static int32_t zmod(int32_t a, int32_t n) {  // @mod: floored, result in [0,n)
        int32_t r = a % n;
        return r < 0 ? r + n : r;
    }
    static float getSample(const Params& p, int32_t index1) {  // :35-58
        size_t bytes_per_sample = 1;
        switch (p.sample.format) {
            case unsigned8: bytes_per_sample = 1; break;
            case signed16_lsb: bytes_per_sample = 2; break;
            case signed24_lsb: bytes_per_sample = 3; break;
            case signed32_lsb: bytes_per_sample = 4; break;
        }
        const int32_t num_samples = (int32_t)(p.sample.data_len / bytes_per_sample / p.sample.num_channels);
        const int32_t index = p.loop ? zmod(index1, num_samples) : index1;
        if (index >= 0 && index < num_samples) {
            const size_t i = (size_t)index * p.sample.num_channels + p.channel;
            switch (p.sample.format) {
                case unsigned8: return ((float)p.sample.data[i] - 127.5f) / 127.5f;
                case signed16_lsb: return decodeSigned(2, p.sample.data, i);
                case signed24_lsb: return decodeSigned(3, p.sample.data, i);
                case signed32_lsb: return decodeSigned(4, p.sample.data, i);
            }
        }
        return 0.0f;
    }
    void paint(Span span, float* out0, bool note_id_changed, const Params& p) {  // :77-136
        if (p.channel >= p.sample.num_channels) return;
        if (note_id_changed) t = 0.0f;
        float* out = out0 + span.start;
        const size_t len = span.end - span.start;
        const float ratio = (float)p.sample.sample_rate / p.sample_rate;
        if (ratio < 0.0f && !p.loop) return;
        if (ratio > 0.9999f && ratio < 1.0001f) {  // :105-114
            const int32_t ti = (int32_t)roundf(t);
            for (size_t i = 0; i < len; i++) out[i] += getSample(p, ti + (int32_t)i);
            t += (float)len;
        } else {  // :115-131
            for (size_t i = 0; i < len; i++) {
                const int32_t t0 = (int32_t)floorf(t);
                const int32_t t1 = t0 + 1;
                const float tfrac = (float)t1 - t;
                const float s0 = getSample(p, t0);
                const float s1 = getSample(p, t1);
                const float s = s0 * (1.0f - tfrac) + s1 * tfrac;
                out[i] += s;
                t += ratio;
            }
        }
        if (t >= (float)p.sample.data_len && p.loop) t -= (float)p.sample.data_len;  // :133-135 (byte length)
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Decimator.zig:21-57
// ------------------------------------------------------------------------------------------
struct Decimator {
    struct Params {
        float sample_rate;
        const float* input;
        float fake_sample_rate;
    };
    float dval = 0.0f, dcount = 1.0f;
    void paint(Span span, float* output, bool, const Params& p) {
        if (p.fake_sample_rate >= p.sample_rate) {
            addInto(span, output, p.input);
            dval = 0.0f;
            dcount = 1.0f;
        } else if (p.fake_sample_rate > 0.0f) {
            const float ratio = p.fake_sample_rate / p.sample_rate;
            float dc = dcount, dv = dval;
            for (size_t i = span.start; i < span.end; i++) {
                dc += ratio;
                if (dc >= 1.0f) {
                    dv = p.input[i];
                    dc -= 1.0f;
                }
                output[i] += dv;
            }
            dcount = dc;
            dval = dv;
        }
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Distortion.zig:27-66
// ------------------------------------------------------------------------------------------
struct Distortion {
    enum Type { overdrive = 0, clip = 1 };
    struct Params {
        const float* input;
        Type type;
        float ingain, outgain, offset;
    };
    void paint(Span span, float* output, bool, const Params& p) {
        const float gain1 = powf(2.0f, p.ingain * 8.0f - 2.0f);  // :41 (Zig std.math.pow: parity unpinned)
        if (p.type == overdrive) {
            const float gain2 = p.outgain / atanf(gain1);
            const float offs = gain1 * p.offset;
            for (size_t i = span.start; i < span.end; i++) {
                const float a = atanf(p.input[i] * gain1 + offs);
                output[i] += gain2 * a;
            }
        } else {
            const float gain2 = p.outgain;
            const float offs = gain1 * p.offset;
            for (size_t i = span.start; i < span.end; i++) {
                const float a = p.input[i] * gain1 + offs;
                const float bq = a < -1.0f ? -1.0f : (a > 1.0f ? 1.0f : a);
                output[i] += gain2 * bq;
            }
        }
    }
};

// ------------------------------------------------------------------------------------------
// src/modules/Cycle.zig:22-59
// ------------------------------------------------------------------------------------------
struct Cycle {
    struct Params {
        float sample_rate;
        ConstantOrBuffer speed;
    };
    float t = 0.0f;
    void paint(Span span, float* out0, bool, const Params& p) {
        float tt = t;
        if (!p.speed.is_buffer) {
            const float step = p.speed.constant / p.sample_rate;
            for (size_t i = span.start; i < span.end; i++) {
                out0[i] += tt;
                tt += step;
                tt -= truncf(tt);
            }
        } else {
            const float isr = 1.0f / p.sample_rate;
            for (size_t i = span.start; i < span.end; i++) {
                out0[i] += tt;
                tt += p.speed.buffer[i] * isr;
                tt -= truncf(tt);
            }
        }
        t = tt;
    }
};

// ------------------------------------------------------------------------------------------
// src/zang/notes.zig
// ------------------------------------------------------------------------------------------
struct Impulse {  // :58-62
    size_t frame, note_id, event_id;
};
struct IdGenerator {  // :43-56
    size_t next_id = 1;
    size_t nextId() { return next_id++; }
};

template <class P>
struct Notes {
    struct ImpulsesAndParamses {  // :66-70
        const Impulse* impulses;
        const P* paramses;
        size_t len;
    };
    struct ImpulseQueue {  // :72-128
        Impulse impulses_array[32];
        P paramses_array[32];
        size_t length = 0;
        size_t next_event_id = 1;
        ImpulsesAndParamses consume() {
            ImpulsesAndParamses r{impulses_array, paramses_array, length};
            length = 0;
            return r;
        }
        void push(size_t impulse_frame, size_t note_id, const P& params) {
            if (length >= 32) return;
            if (length > 0 && impulse_frame < impulses_array[length - 1].frame) return;
            impulses_array[length] = Impulse{impulse_frame, note_id, next_event_id};
            paramses_array[length] = params;
            length += 1;
            next_event_id += 1;
        }
    };
    struct SongEvent {  // :130-134
        P params;
        float t;
        size_t note_id;
    };
    struct NoteTracker {  // :138-207
        const SongEvent* song = nullptr;
        size_t song_len = 0;
        size_t next_song_event = 0;
        float t = 0.0f;
        Impulse impulses_array[32];
        P paramses_array[32];
        void init(const SongEvent* s, size_t n) {
            song = s;
            song_len = n;
            next_song_event = 0;
            t = 0.0f;
        }
        void reset() {
            next_song_event = 0;
            t = 0.0f;
        }
        ImpulsesAndParamses consume(float sample_rate, Span span) {  // :162-206
            size_t count = 0;
            const size_t out_len = span.end - span.start;
            const float buf_time = (float)out_len / sample_rate;
            const float end_t = t + buf_time;
            const size_t first = next_song_event;  // the for loop iterates a slice taken before the loop
            for (size_t k = first; k < song_len; k++) {
                const float note_t = song[k].t;
                if (note_t < end_t) {
                    const float f = (note_t - t) / buf_time;
                    const size_t rel = std::min((size_t)(f * (float)out_len), out_len - 1);
                    next_song_event += 1;
                    if (count < 32) {  // reference has no bound check (:184-187); stay in bounds
                        impulses_array[count] = Impulse{span.start + rel, song[k].note_id, next_song_event};
                        paramses_array[count] = song[k].params;
                        count += 1;
                    }
                } else {
                    break;
                }
            }
            t = end_t;
            return {impulses_array, paramses_array, count};
        }
    };

    template <size_t polyphony>
    struct PolyphonyDispatcher {  // :209-349
        struct SlotState {
            bool used = false;
            size_t note_id = 0, event_id = 0;
            bool note_on = false;
        };
        SlotState slots[polyphony];
        Impulse impulses_array[polyphony][32];
        P paramses_array[polyphony][32];

        void reset() {
            for (auto& s : slots) s = SlotState{};
        }
        bool chooseSlot(size_t note_id, bool note_on, size_t& out) const {  // :246-306
            if (!note_on) {
                for (size_t i = 0; i < polyphony; i++)
                    if (slots[i].used && slots[i].note_id == note_id && slots[i].note_on) {
                        out = i;
                        return true;
                    }
                return false;
            }
            {
                bool have_best = false;
                size_t best = 0;
                for (size_t i = 0; i < polyphony; i++) {
                    if (slots[i].used) {
                        if (!slots[i].note_on) {
                            if (have_best) {
                                if (slots[i].event_id < slots[best].event_id) best = i;
                            } else {
                                have_best = true;
                                best = i;
                            }
                        }
                    } else {
                        out = i;
                        return true;
                    }
                }
                if (have_best) {
                    out = best;
                    return true;
                }
            }
            size_t best = 0;
            for (size_t i = 1; i < polyphony; i++)
                if (slots[i].event_id < slots[best].event_id) best = i;
            out = best;
            return true;
        }
        // P must have a note_on member
        void dispatch(ImpulsesAndParamses iap, ImpulsesAndParamses (&result)[polyphony]) {  // :310-347
            size_t counts[polyphony];
            for (auto& c : counts) c = 0;
            for (size_t i = 0; i < iap.len; i++) {
                const Impulse impulse = iap.impulses[i];
                const P params = iap.paramses[i];
                size_t slot_index;
                if (chooseSlot(impulse.note_id, params.note_on, slot_index)) {
                    slots[slot_index] = SlotState{true, impulse.note_id, impulse.event_id, params.note_on};
                    impulses_array[slot_index][counts[slot_index]] = impulse;
                    paramses_array[slot_index][counts[slot_index]] = params;
                    counts[slot_index] += 1;
                }
            }
            for (size_t i = 0; i < polyphony; i++) result[i] = {impulses_array[i], paramses_array[i], counts[i]};
        }
    };
};

// ------------------------------------------------------------------------------------------
// src/zang/trigger.zig
// ------------------------------------------------------------------------------------------
template <class P>
struct Trigger {
    struct NoteSpanNote {
        size_t id;
        P params;
    };
    struct NoteSpan {
        size_t start, end;
        bool has_note;
        NoteSpanNote note;
    };
    struct Counter {  // :43-48
        typename Notes<P>::ImpulsesAndParamses iap;
        size_t impulse_index;
        size_t start, end;
    };
    struct NewPaintReturnValue {  // :50-54
        Span span;
        P params;
        bool note_id_changed;
    };
    bool has_note = false;
    NoteSpanNote note{};

    void reset() { has_note = false; }  // :62-64
    Counter counter(Span span, typename Notes<P>::ImpulsesAndParamses iap) { return Counter{iap, 0, span.start, span.end}; }

    bool next(Counter& ctr, NewPaintReturnValue& out) {  // :80-105
        while (ctr.start < ctr.end) {
            NoteSpan ns;
            if (!carryOver(ctr, ns)) ns = getNextNoteSpan(ctr);
            ctr.start = ns.end;
            if (ns.has_note) {
                out.span = Span{ns.start, ns.end};
                out.params = ns.note.params;
                out.note_id_changed = has_note ? (ns.note.id != note.id) : true;
                has_note = true;  // defer self.note = note
                note = ns.note;
                return true;
            }
        }
        return false;
    }
    bool carryOver(const Counter& ctr, NoteSpan& out) const {  // :107-137
        if (!has_note) return false;
        if (ctr.impulse_index < ctr.iap.len) {
            const size_t next_impulse_frame = ctr.iap.impulses[ctr.impulse_index].frame;
            if (next_impulse_frame > ctr.start) {
                out = NoteSpan{ctr.start, std::min(ctr.end, next_impulse_frame), true, note};
                return true;
            }
            return false;
        }
        out = NoteSpan{ctr.start, ctr.end, true, note};
        return true;
    }
    static NoteSpan getNextNoteSpan(Counter& ctr) {  // :139-196
        const size_t base = ctr.impulse_index;  // slices taken once
        const size_t n = ctr.iap.len - base;
        for (size_t i = 0; i < n; i++) {
            const Impulse& impulse = ctr.iap.impulses[base + i];
            if (impulse.frame >= ctr.end) break;
            if (impulse.frame > ctr.start) return NoteSpan{ctr.start, impulse.frame, false, {}};
            ctr.impulse_index += 1;
            const size_t note_end_clipped = (i + 1 < n) ? std::min(ctr.end, ctr.iap.impulses[base + i + 1].frame) : ctr.end;
            if (note_end_clipped <= ctr.start) continue;
            return NoteSpan{ctr.start, note_end_clipped, true, NoteSpanNote{impulse.note_id, ctr.iap.paramses[base + i]}};
        }
        return NoteSpan{ctr.start, ctr.end, false, {}};
    }
};

// ------------------------------------------------------------------------------------------
// src/zang/mixdown.zig:28-86
// ------------------------------------------------------------------------------------------
inline void mixDownS16LSB(uint8_t* dst, const float* mix, size_t n, size_t num_channels, size_t channel_index, float vol) {
    const float mul = vol * 32767.0f;
    for (size_t i = 0; i < n; i++) {
        const float value = mix[i] * mul;
        int16_t cv;
        if (value <= -32767.0f)
            cv = -32767;
        else if (value >= 32766.0f)
            cv = 32766;
        else if (value != value)
            cv = 0;
        else
            cv = (int16_t)value;
        const size_t index = (i * num_channels + channel_index) * 2;
        dst[index + 0] = (uint8_t)(cv & 0xFF);
        dst[index + 1] = (uint8_t)((cv >> 8) & 0xFF);
    }
}
inline void mixDownS8(uint8_t* dst, const float* mix, size_t n, size_t num_channels, size_t channel_index, float vol) {
    const float mul = vol * 127.0f;
    for (size_t i = 0; i < n; i++) {
        const float value = mix[i] * mul;
        int8_t cv;
        if (value <= -127.0f)
            cv = -127;
        else if (value >= 126.0f)
            cv = 126;
        else if (value != value)
            cv = 0;
        else
            cv = (int8_t)value;
        dst[i * num_channels + channel_index] = (uint8_t)cv;
    }
}

// ------------------------------------------------------------------------------------------
// examples/modules.zig — instrument graphs (workload definitions)
// ------------------------------------------------------------------------------------------
struct PhaseModOscillator {  // modules.zig:6-77
    struct Params {
        float sample_rate, freq;
        bool relative;
        ConstantOrBuffer ratio, multiplier;
    };
    SineOsc carrier, modulator;
    void paint(Span span, float* out0, float* t0, float* t1, bool nic, const Params& p) {
        if (!p.ratio.is_buffer) {
            if (p.relative)
                set(span, t0, p.freq * p.ratio.constant);
            else
                set(span, t0, p.ratio.constant);
        } else {
            if (p.relative)
                multiplyScalar(span, t0, p.ratio.buffer, p.freq);
            else
                copy(span, t0, p.ratio.buffer);
        }
        zero(span, t1);
        modulator.paint(span, t1, nic, {p.sample_rate, buffer(t0), constant(0.0f)});
        zero(span, t0);
        if (!p.multiplier.is_buffer)
            multiplyScalar(span, t0, t1, p.multiplier.constant);
        else
            multiply(span, t0, t1, p.multiplier.buffer);
        zero(span, t1);
        carrier.paint(span, t1, nic, {p.sample_rate, constant(p.freq), buffer(t0)});
        addInto(span, out0, t1);
    }
};

struct InstrParams {  // {sample_rate, freq, note_on} shared by the example instruments
    float sample_rate, freq;
    bool note_on;
};

struct PMOscInstrument {  // modules.zig:80-128
    static constexpr int num_temps = 3;
    float release_duration;
    PhaseModOscillator osc;
    Envelope env;
    explicit PMOscInstrument(float rd = 0.4f) : release_duration(rd) {}
    void paint(Span span, float* out0, float** temps, bool nic, const InstrParams& p) {
        zero(span, temps[0]);
        osc.paint(span, temps[0], temps[1], temps[2], nic, {p.sample_rate, p.freq, true, constant(1.0f), constant(1.0f)});
        zero(span, temps[1]);
        env.paint(span, temps[1], nic,
                  {p.sample_rate, {CurveKind::cubed, 0.025f}, {CurveKind::cubed, 0.1f}, {CurveKind::cubed, release_duration}, 0.5f, p.note_on});
        multiply(span, out0, temps[0], temps[1]);
    }
};

struct NiceInstrument {  // modules.zig:189-248
    static constexpr int num_temps = 2;
    float color;
    PulseOsc osc;
    Filter flt;
    Envelope env;
    explicit NiceInstrument(float c = 0.3f) : color(c) {}
    void paint(Span span, float* out0, float** temps, bool nic, const InstrParams& p) {
        zero(span, temps[0]);
        osc.paint(span, temps[0], nic, {p.sample_rate, constant(p.freq), color});
        multiplyWithScalar(span, temps[0], 0.5f);
        zero(span, temps[1]);
        flt.paint(span, temps[1], nic,
                  {temps[0], Filter::low_pass, constant(Filter::cutoffFromFrequency(p.freq * 8.0f, p.sample_rate)), constant(0.7f)});
        zero(span, temps[0]);
        env.paint(span, temps[0], nic,
                  {p.sample_rate, {CurveKind::cubed, 0.01f}, {CurveKind::cubed, 0.1f}, {CurveKind::cubed, 0.5f}, 0.8f, p.note_on});
        multiply(span, out0, temps[0], temps[1]);
    }
};

struct HardSquareInstrument {  // modules.zig:250-289
    static constexpr int num_temps = 2;
    PulseOsc osc;
    Gate gate;
    void paint(Span span, float* out0, float** temps, bool nic, const InstrParams& p) {
        zero(span, temps[0]);
        osc.paint(span, temps[0], nic, {p.sample_rate, constant(p.freq), 0.5f});
        zero(span, temps[1]);
        gate.paint(span, temps[1], nic, {p.note_on});
        multiply(span, out0, temps[0], temps[1]);
    }
};

// src/zang-12tet.zig:3-7
inline float note12tet(int note) {
    const float semitone = powf(2.0f, (float)(1.0 / 12.0));
    return powf(semitone, (float)note);
}

struct FilteredSawtoothInstrument {  // modules.zig:130-187 (freq is ConstantOrBuffer)
    static constexpr int num_temps = 3;
    TriSawOsc osc;
    Envelope env;
    Filter flt;
    void paint(Span span, float* out0, float** temps, bool nic, float sample_rate, ConstantOrBuffer freq, bool note_on) {
        zero(span, temps[0]);
        osc.paint(span, temps[0], nic, {sample_rate, freq, 0.0f});
        multiplyWithScalar(span, temps[0], 1.5f);
        zero(span, temps[1]);
        env.paint(span, temps[1], nic,
                  {sample_rate, {CurveKind::cubed, 0.025f}, {CurveKind::cubed, 0.1f}, {CurveKind::cubed, 1.0f}, 0.5f, note_on});
        zero(span, temps[2]);
        multiply(span, temps[2], temps[0], temps[1]);
        flt.paint(span, out0, nic,
                  {temps[2], Filter::low_pass, constant(Filter::cutoffFromFrequency(440.0f * note12tet(3), sample_rate)), constant(0.7f)});
    }
};

}  // namespace zo
