// delay_oracle.hpp — TEST INFRASTRUCTURE (oracle): CPU restatement of zang's delay line and the echo modules built on it.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use anything under oracle/.
//
// Follows /root/reference/src/zang/delay.zig:7-91 (Delay), /root/reference/examples/modules.zig:341-388 (SimpleDelay),
// :392-460 (FilteredEchoes) and :462-525 (StereoEchoes), statement by statement, including the chunking of a span into
// pieces of at most `delay_samples` frames (readDelayBuffer returns how much it served; the caller loops).
// Parity unpinned: the reference holds no test or golden vector for these.
#pragma once

#include <algorithm>
#include <cstring>
#include <vector>

#include "zang_oracle.hpp"

namespace zo {

// delay.zig:7-91; delay_samples is a comptime parameter there, a run-time one here
struct Delay {
    std::vector<float> delay_buffer;
    size_t delay_buffer_index = 0;  // always < delay_samples
    explicit Delay(size_t delay_samples) : delay_buffer(delay_samples, 0.0f) {}
    size_t delay_samples() const { return delay_buffer.size(); }
    void reset() {  // :19-22
        std::fill(delay_buffer.begin(), delay_buffer.end(), 0.0f);
        delay_buffer_index = 0;
    }
    // :28-58: adds the next out_len (at most delay_samples) delayed samples into out; returns how many
    size_t readDelayBuffer(float* out, size_t out_len) {
        const size_t ds = delay_samples();
        const size_t actual = out_len > ds ? ds : out_len;
        const size_t index = delay_buffer_index;
        const size_t len = std::min(ds - index, actual);
        for (size_t i = 0; i < len; i++) out[i] += delay_buffer[index + i];
        if (len < actual) {
            const size_t b_len = actual - len;
            for (size_t i = 0; i < b_len; i++) out[len + i] += delay_buffer[i];
        }
        return actual;
    }
    // :63-90
    void writeDelayBuffer(const float* input, size_t input_len) {
        const size_t ds = delay_samples();
        const size_t index = delay_buffer_index;
        const size_t len = std::min(ds - index, input_len);
        std::memcpy(delay_buffer.data() + index, input, len * sizeof(float));
        if (len < input_len) {
            const size_t b_len = input_len - len;
            std::memcpy(delay_buffer.data(), input + len, b_len * sizeof(float));
            delay_buffer_index = b_len;
        } else {
            delay_buffer_index += len;
            if (delay_buffer_index == ds) delay_buffer_index = 0;
        }
    }
};

// examples/modules.zig:341-388
struct SimpleDelay {
    Delay delay;
    explicit SimpleDelay(size_t n) : delay(n) {}
    void paint(Span span, float* out0, bool, const float* input) {
        size_t start = span.start;
        const size_t end = span.end;
        while (start < end) {
            const size_t samples_read = delay.readDelayBuffer(out0 + start, end - start);
            delay.writeDelayBuffer(input + start, samples_read);
            start += samples_read;
        }
    }
};

// examples/modules.zig:392-460
struct FilteredEchoes {
    struct Params {
        const float* input;
        float feedback_volume, cutoff;
    };
    Delay delay;
    Filter filter;
    explicit FilteredEchoes(size_t n) : delay(n) {}
    void paint(Span span, float* output, float* temp0, float* temp1, bool note_id_changed, const Params& p) {
        size_t start = span.start;
        const size_t end = span.end;
        while (start < end) {
            zero(Span{start, end}, temp0);                                              // :430
            const size_t samples_read = delay.readDelayBuffer(temp0 + start, end - start);  // :432
            const Span span1{start, start + samples_read};
            multiplyWithScalar(span1, temp0, p.feedback_volume);                        // :437
            addInto(span1, temp0, p.input);                                             // :440
            zero(span1, temp1);                                                         // :443
            filter.paint(span1, temp1, note_id_changed, {temp0, Filter::low_pass, constant(p.cutoff), constant(0.0f)});  // :444-449
            addInto(span1, output, temp1);                                              // :452
            delay.writeDelayBuffer(temp1 + span1.start, samples_read);                  // :456
            start += samples_read;
        }
    }
};

// examples/modules.zig:462-525; MAIN_DELAY / 2 is an integer division (:463)
struct StereoEchoes {
    SimpleDelay delay0, delay1;
    FilteredEchoes echoes;
    explicit StereoEchoes(size_t main_delay) : delay0(main_delay / 2), delay1(main_delay / 2), echoes(main_delay) {}
    // outputs[2], temps[4]
    void paint(Span span, float* out_l, float* out_r, float** temps, bool nic, const float* input, float feedback_volume, float cutoff) {
        addInto(span, out_l, input);                        // :497-498
        addInto(span, out_r, input);
        zero(span, temps[0]);                               // :501
        delay0.paint(span, temps[0], nic, input);           // :502-504
        zero(span, temps[1]);                               // :506
        echoes.paint(span, temps[1], temps[2], temps[3], nic, {temps[0], feedback_volume, cutoff});  // :507-511
        addInto(span, out_l, temps[1]);                     // :513
        delay1.paint(span, out_r, nic, temps[1]);           // :514-516
    }
};

// examples/example_fmsynth.zig:26-89 (Oscillator): parity unpinned like every paint(); `sin` = glibc sinf
struct FmOscillator {
    struct Params {
        float sample_rate;
        unsigned waveform;  // u2
        const float* freq;
        const float* phase;  // may be null
        float feedback;
    };
    float t = 0.0f, feedback1 = 0.0f, feedback2 = 0.0f;
    void paint(Span span, float* out0, bool, const Params& p) {
        float* output = out0 + span.start;
        const size_t len = span.end - span.start;
        float tt = t;
        const float inv_sample_rate = 1.0f / p.sample_rate;  // :66
        for (size_t i = 0; i < len; i++) {
            const float phase = p.phase ? p.phase[span.start + i] : 0.0f;
            const float feedback = (feedback1 + feedback2) * p.feedback;  // :70
            const float pp = (tt + phase) * kPi * 2.0f + feedback;        // :72
            const float s = sinf(pp);
            float sample;
            switch (p.waveform & 3u) {  // :74-79
                case 0: sample = s; break;
                case 1: sample = s > 0.0f ? s : 0.0f; break;
                case 2: sample = fabsf(s); break;
                default: sample = sinf(pp * 2.0f) >= 0.0f ? fabsf(s) : 0.0f; break;
            }
            output[i] += sample;
            tt += p.freq[span.start + i] * inv_sample_rate;  // :83
            feedback2 = feedback1;
            feedback1 = sample;
        }
        t = tt - truncf(tt);  // :63 (defer)
    }
};

}  // namespace zo
