// song_oracle.hpp — TEST INFRASTRUCTURE ONLY (see zang_oracle.hpp header).
//
// CPU restatement of the reference's own CPU-runnable workload (BASELINE.json configs[0]):
//   examples/write_wav.zig:14-96        offline render loop, 1024-frame buffers, mixDown(vol 0.25) to s16
//   examples/example_song.zig:19-396    Toccata song: 3 polyphonic instruments (3 + 10 + 4 sub-voices)
//   examples/common/songparse1.zig      tracker-text parser
// Third-party pieces that are not in /root/reference: zig-wav (submodule, missing) is replaced by the
// canonical 44-byte PCM RIFF header; std.fmt.parseFloat(f32) is restated as strtof (both are correctly
// rounded); std.math.pow(f32) is glibc powf (unpinned).  PARITY UNPINNED: the reference
// ships no golden out.wav.
#pragma once

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "zang_oracle.hpp"

namespace zo {

struct MyNoteParams {  // example_song.zig:23-26
    float freq;
    bool note_on;
};
using SongEvent = Notes<MyNoteParams>::SongEvent;

// ---- examples/common/songparse1.zig:3-198 --------------------------------------------------------
template <size_t num_columns>
struct SongParser {
    enum NoteTag { idle, freq, off };
    struct Note {
        NoteTag tag = idle;
        float freq = 0.0f;
    };
    enum TokenTag { tok_eof, tok_word, tok_number, tok_notes, tok_error };
    struct Token {
        TokenTag tag = tok_eof;
        std::string word;
        float number = 0.0f;
        Note notes[num_columns];
        bool isWord(const char* w) const { return tag == tok_word && word == w; }
    };

    float a4;
    const char* contents;
    size_t len;
    size_t index = 0;
    size_t line_index = 0;

    bool parseNote(float& out) {  // :29-63
        if (index + 3 > len) return false;
        const char letter = contents[index], modifier = contents[index + 1], octave = contents[index + 2];
        int offset;
        if (octave >= '0' && octave <= '9')
            offset = (int)(octave - '0') * 12 - 57;
        else
            return false;
        int semitone;
        if (letter == 'C' && modifier == '-') semitone = 0;
        else if (letter == 'C' && modifier == '#') semitone = 1;
        else if (letter == 'D' && modifier == '-') semitone = 2;
        else if (letter == 'D' && modifier == '#') semitone = 3;
        else if (letter == 'E' && modifier == '-') semitone = 4;
        else if (letter == 'F' && modifier == '-') semitone = 5;
        else if (letter == 'F' && modifier == '#') semitone = 6;
        else if (letter == 'G' && modifier == '-') semitone = 7;
        else if (letter == 'G' && modifier == '#') semitone = 8;
        else if (letter == 'A' && modifier == '-') semitone = 9;
        else if (letter == 'A' && modifier == '#') semitone = 10;
        else if (letter == 'B' && modifier == '-') semitone = 11;
        else return false;
        index += 3;
        const float exp = (float)(offset + semitone) / 12.0f;
        out = a4 * powf(2.0f, exp);
        return true;
    }
    bool eat(const char* prefix) {  // :65-72
        const size_t n = strlen(prefix);
        if (index + n > len || memcmp(contents + index, prefix, n) != 0) return false;
        index += n;
        return true;
    }
    Token parseToken() {  // :74-185
        Token tok;
        while (true) {
            if (eat(" ")) {
            } else if (eat("\n")) {
                line_index += 1;
            } else if (eat("#")) {
                const void* nl = memchr(contents + index, '\n', len - index);
                if (nl) {
                    line_index += 1;
                    index = (size_t)((const char*)nl - contents) + 1;
                } else {
                    index = len;
                }
            } else {
                break;
            }
        }
        if (index >= len) return tok;  // null
        const char ch = contents[index];
        if (ch == '|') {
            index += 1;
            size_t col = 0;
            while (true) {
                if (col >= num_columns) {
                    tok.tag = tok_error;
                    return tok;
                }
                float f;
                if (parseNote(f)) {
                    tok.notes[col].tag = freq;
                    tok.notes[col].freq = f;
                } else if (eat("off")) {
                    tok.notes[col].tag = off;
                } else if (eat("   ")) {
                } else {
                    break;
                }
                if (index < len && (contents[index] == ' ' || contents[index] == '|'))
                    index += 1;
                else
                    break;
                col += 1;
            }
            if (index < len) {
                if (contents[index] == '\n') {
                    line_index += 1;
                    index += 1;
                } else {
                    tok.tag = tok_error;
                    return tok;
                }
            }
            tok.tag = tok_notes;
            return tok;
        }
        auto is_alpha = [](char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_'; };
        if (is_alpha(ch)) {
            const size_t start = index;
            index += 1;
            while (index < len && (is_alpha(contents[index]) || (contents[index] >= '0' && contents[index] <= '9'))) index += 1;
            tok.tag = tok_word;
            tok.word.assign(contents + start, index - start);
            return tok;
        }
        if (ch >= '0' && ch <= '9') {
            const size_t start = index;
            bool dot = false;
            index += 1;
            while (index < len) {
                const char c2 = contents[index];
                if (c2 == '.') {
                    if (dot) break;
                    dot = true;
                    index += 1;
                } else if (c2 >= '0' && c2 <= '9') {
                    index += 1;
                } else {
                    break;
                }
            }
            const std::string s(contents + start, index - start);
            tok.tag = tok_number;
            tok.number = strtof(s.c_str(), nullptr);  // std.fmt.parseFloat(f32): correctly rounded, as strtof
            return tok;
        }
        tok.tag = tok_error;
        return tok;
    }
};

// ---- examples/example_song.zig --------------------------------------------------------------------
struct Song {
    static constexpr size_t NUM_INSTRUMENTS = 3;
    static constexpr size_t COLUMNS_PER_VOICE[3] = {2, 8, 2};  // :89-93 (Pedal, RegularOrgan, WeirdOrgan)
    static constexpr size_t TOTAL_COLUMNS = 12;
    std::vector<SongEvent> all_notes[NUM_INSTRUMENTS];
    bool ok = false;
    size_t error_line = 0;

    // doParse, :129-264
    void parse(const char* text, size_t len) {
        const float NOTE_DURATION = 0.15f;  // :20
        SongParser<TOTAL_COLUMNS> parser{440.0f, text, len};
        struct LastNote {
            bool present = false;
            float freq = 0.0f;
            size_t id = 0;
        };
        LastNote column_last_note[TOTAL_COLUMNS];
        size_t next_id = 1;
        float t = 0.0f, rate = 1.0f, tempo = 1.0f;
        for (auto& v : all_notes) v.clear();
        ok = false;
        while (true) {
            auto token = parser.parseToken();
            if (token.tag == SongParser<TOTAL_COLUMNS>::tok_eof) break;
            if (token.tag == SongParser<TOTAL_COLUMNS>::tok_error) {
                error_line = parser.line_index + 1;
                return;
            }
            if (token.isWord("start")) {
                t = 0.0f;
                for (auto& v : all_notes) v.clear();
            } else if (token.isWord("rate") || token.isWord("tempo")) {
                const bool is_rate = token.isWord("rate");
                auto num = parser.parseToken();
                if (num.tag != SongParser<TOTAL_COLUMNS>::tok_number) {
                    error_line = parser.line_index + 1;
                    return;
                }
                (is_rate ? rate : tempo) = num.number;
            } else if (token.tag == SongParser<TOTAL_COLUMNS>::tok_notes) {
                size_t old_num[NUM_INSTRUMENTS];
                for (size_t i = 0; i < NUM_INSTRUMENTS; i++) old_num[i] = all_notes[i].size();
                for (size_t col = 0; col < TOTAL_COLUMNS; col++) {
                    size_t instrument_index = 0, first_column = 0;
                    for (size_t k = 0; k < NUM_INSTRUMENTS; k++) {
                        if (col < first_column + COLUMNS_PER_VOICE[k]) {
                            instrument_index = k;
                            break;
                        }
                        first_column += COLUMNS_PER_VOICE[k];
                    }
                    auto& notes = all_notes[instrument_index];
                    const auto& note = token.notes[col];
                    if (note.tag == SongParser<TOTAL_COLUMNS>::freq) {
                        if (column_last_note[col].present)
                            notes.push_back(SongEvent{MyNoteParams{column_last_note[col].freq, false}, t, column_last_note[col].id});
                        notes.push_back(SongEvent{MyNoteParams{note.freq, true}, t, next_id});
                        column_last_note[col] = LastNote{true, note.freq, next_id};
                        next_id += 1;
                    } else if (note.tag == SongParser<TOTAL_COLUMNS>::off) {
                        if (column_last_note[col].present) {
                            notes.push_back(SongEvent{MyNoteParams{column_last_note[col].freq, false}, t, column_last_note[col].id});
                            column_last_note[col].present = false;
                        }
                    }
                }
                t += NOTE_DURATION / (rate * tempo);  // :213
                // events of this row sorted by note id: note-offs before note-ons (:215-234; std.mem.sort is stable)
                for (size_t i = 0; i < NUM_INSTRUMENTS; i++)
                    std::stable_sort(all_notes[i].begin() + (ptrdiff_t)old_num[i], all_notes[i].end(),
                                     [](const SongEvent& a, const SongEvent& b) { return a.note_id < b.note_id; });
            } else {
                error_line = parser.line_index + 1;  // error.BadToken
                return;
            }
        }
        ok = true;
    }
};

// Voice(T), example_song.zig:287-350
template <class Module, size_t polyphony>
struct SongVoice {
    struct SubVoice {
        Module module;
        Trigger<MyNoteParams> trigger;
    };
    typename Notes<MyNoteParams>::NoteTracker tracker;
    typename Notes<MyNoteParams>::template PolyphonyDispatcher<polyphony> dispatcher;
    std::vector<SubVoice> sub_voices;
    float freq_mul;

    SongVoice(const std::vector<SongEvent>& notes, const Module& proto, float freq_mul_) : freq_mul(freq_mul_) {
        tracker.init(notes.data(), notes.size());
        for (size_t i = 0; i < polyphony; i++) sub_voices.push_back(SubVoice{proto, Trigger<MyNoteParams>{}});
    }
    void paint(float sample_rate, Span span, float* out0, float** temps) {
        auto iap = tracker.consume(sample_rate, span);
        typename Notes<MyNoteParams>::ImpulsesAndParamses poly_iap[polyphony];
        dispatcher.dispatch(iap, poly_iap);
        for (size_t i = 0; i < polyphony; i++) {
            auto ctr = sub_voices[i].trigger.counter(span, poly_iap[i]);
            typename Trigger<MyNoteParams>::NewPaintReturnValue r;
            while (sub_voices[i].trigger.next(ctr, r)) {
                // makeParams, :33-39 / :49-55 / :65-71 (Pedal plays freq * 0.5)
                const InstrParams p{sample_rate, r.params.freq * freq_mul, r.params.note_on};  // x1.0 is exact
                sub_voices[i].module.paint(r.span, out0, temps, r.note_id_changed, p);
            }
        }
    }
};

// MainModule (example_song.zig:353-396) + the render loop of write_wav.zig:49-95.
// pcm: total_frames * 2 bytes (s16le mono).  mix_f32 (optional): total_frames floats before mixDown.
inline bool song_render(const Song& song, size_t total_frames, float sample_rate, uint8_t* pcm, float* mix_f32) {
    if (!song.ok) return false;
    const size_t B = 1024;  // AUDIO_BUFFER_SIZE, :8
    SongVoice<PMOscInstrument, 3> voice0(song.all_notes[0], PMOscInstrument(0.4f), 0.5f);
    SongVoice<NiceInstrument, 10> voice1(song.all_notes[1], NiceInstrument(0.25f), 1.0f);
    SongVoice<NiceInstrument, 4> voice2(song.all_notes[2], NiceInstrument(0.1f), 1.0f);
    std::vector<float> out(B), t0(B), t1(B), t2(B);
    float* temps[3] = {t0.data(), t1.data(), t2.data()};
    size_t start = 0;
    while (start < total_frames) {
        const size_t n = std::min(B, total_frames - start);
        const Span span{0, n};
        zero(span, out.data());
        voice0.paint(sample_rate, span, out.data(), temps);
        voice1.paint(sample_rate, span, out.data(), temps);
        voice2.paint(sample_rate, span, out.data(), temps);
        if (pcm) mixDownS16LSB(pcm + start * 2, out.data(), n, 1, 0, 0.25f);
        if (mix_f32) memcpy(mix_f32 + start, out.data(), n * sizeof(float));
        start += n;
    }
    return true;
}

}  // namespace zo
