// zcu_events.cpp — host-side event layer: the index logic of src/zang/notes.zig and src/zang/trigger.zig
// on the PODs of zangcu.h.  Pure integer / f32 host code, no CUDA.  Results are bit-exact with the
// reference (pinned by its own tests, transcribed in tests/test_event_kat.py).
//
// A params value P of the reference is represented by a `param` row index that travels with the
// impulse; the caller keeps the table of rows.
#include <algorithm>
#include <cstring>
#include <vector>

#include "../../include/zangcu.h"
#include "zcu_exact_accum.cuh"

extern "C" {

// ---- Trigger (trigger.zig:26-198) ---------------------------------------------------------------------
void zcu_trigger_init(zcu_trigger* t) {
    t->has_note = 0;
    t->note_id = 0;
    t->param = 0;
}
void zcu_trigger_reset(zcu_trigger* t) { zcu_trigger_init(t); }

// What Trigger.next() produces over one buffer, stated directly:
//   * while a note is current and the next unconsumed impulse lies later, the current note plays up to
//     that impulse (carryOver, :107-137);
//   * at an impulse's frame the impulse is consumed; if another impulse has the same frame the earlier
//     one is dropped (:173-178), otherwise it becomes the current note and plays to the next impulse or
//     the end of the buffer;
//   * before the first note ever, the gap up to the first impulse is skipped (:152-158);
//   * an impulse at or past the end of the buffer ends the scan (:144-150) — the rest of the buffer is
//     a gap when nothing has played yet, and belongs to the current note otherwise.
// note_id_changed compares with the note of the previously RETURNED span (:96-100).
uint32_t zcu_trigger_run(zcu_trigger* t, uint32_t start, uint32_t end, const zcu_impulse* impulses, const uint32_t* impulse_params,
                         uint32_t n, uint32_t frame_base, zcu_span* out, uint32_t max_out) {
    uint32_t produced = 0;
    uint32_t idx = 0;   // Counter.impulse_index
    uint32_t pos = start;
    while (pos < end) {
        uint32_t span_end;
        bool has = false;
        uint32_t id = 0, param = 0;
        bool carried = false;
        if (t->has_note) {
            if (idx < n) {
                const uint32_t nf = impulses[idx].frame;
                if (nf > pos) {
                    carried = true;
                    span_end = std::min(end, nf);
                }
            } else {
                carried = true;
                span_end = end;
            }
            if (carried) {
                has = true;
                id = t->note_id;
                param = t->param;
            }
        }
        if (!carried) {
            // getNextNoteSpan: scan from idx
            span_end = end;  // "no more impulses" gap
            uint32_t i = idx;
            for (; i < n; i++) {
                const uint32_t fr = impulses[i].frame;
                if (fr >= end) break;         // gap to the end of the buffer
                if (fr > pos) {               // gap before the note begins
                    span_end = fr;
                    break;
                }
                idx++;                        // consumed (frame == pos; earlier frames are a caller bug, trigger.zig:161)
                const uint32_t note_end = (i + 1 < n) ? std::min(end, impulses[i + 1].frame) : end;
                if (note_end <= pos) continue;  // superseded by the next impulse at the same frame
                has = true;
                id = impulses[i].note_id;
                param = impulse_params[i];
                span_end = note_end;
                break;
            }
        }
        if (has) {
            const bool changed = t->has_note ? (id != t->note_id) : true;
            if (produced < max_out) out[produced] = zcu_span{pos + frame_base, span_end + frame_base, changed ? ZCU_SPAN_NOTE_ID_CHANGED : 0u, param};
            produced++;
            t->has_note = 1;
            t->note_id = id;
            t->param = param;
        }
        pos = span_end;
    }
    return produced;
}

// ---- ImpulseQueue (notes.zig:72-128) ------------------------------------------------------------------
void zcu_impulse_queue_init(zcu_impulse_queue* q) {
    q->length = 0;
    q->next_event_id = 1;
}
void zcu_impulse_queue_push(zcu_impulse_queue* q, uint32_t impulse_frame, uint32_t note_id, uint32_t param) {
    if (q->length >= 32) return;                                                      // full: dropped (:108-111)
    if (q->length > 0 && impulse_frame < q->impulses[q->length - 1].frame) return;   // out of order: dropped (:112-118)
    q->impulses[q->length] = zcu_impulse{impulse_frame, note_id, q->next_event_id};
    q->params[q->length] = param;
    q->length++;
    q->next_event_id++;
}
uint32_t zcu_impulse_queue_consume(zcu_impulse_queue* q, zcu_impulse* out_impulses, uint32_t* out_params) {
    const uint32_t n = q->length;
    if (out_impulses) std::memcpy(out_impulses, q->impulses, n * sizeof(zcu_impulse));
    if (out_params) std::memcpy(out_params, q->params, n * sizeof(uint32_t));
    q->length = 0;
    return n;
}

// ---- IdGenerator (notes.zig:43-56) ---------------------------------------------------------------------
void zcu_id_generator_init(zcu_id_generator* g) { g->next_id = 1; }
uint32_t zcu_id_generator_next(zcu_id_generator* g) { return g->next_id++; }

// ---- NoteTracker (notes.zig:138-207) -------------------------------------------------------------------
void zcu_note_tracker_init(zcu_note_tracker* nt, const zcu_song_event* song, uint32_t song_len) {
    nt->song = song;
    nt->song_len = song_len;
    nt->next_song_event = 0;
    nt->t = 0.0f;
}
void zcu_note_tracker_reset(zcu_note_tracker* nt) {
    nt->next_song_event = 0;
    nt->t = 0.0f;
}
uint32_t zcu_note_tracker_consume(zcu_note_tracker* nt, float sample_rate, uint32_t start, uint32_t end, zcu_impulse* out_impulses,
                                  uint32_t* out_params, uint32_t max_out) {
    uint32_t count = 0;
    const uint32_t out_len = end - start;
    const float flen = (float)out_len;
    const float buf_time = flen / sample_rate;
    const float end_t = nt->t + buf_time;  // f32 accumulation, as the reference
    while (nt->next_song_event < nt->song_len) {
        const zcu_song_event& ev = nt->song[nt->next_song_event];
        if (!(ev.t < end_t)) break;
        const float f = (ev.t - nt->t) / buf_time;
        uint32_t rel = (uint32_t)(long long)(f * flen);  // through 64 bits: what x86-64 does anyway, and what the device replay does
        if (rel > out_len - 1) rel = out_len - 1;
        nt->next_song_event++;
        if (count < max_out) {
            out_impulses[count] = zcu_impulse{start + rel, ev.note_id, nt->next_song_event};  // event_id is 1-based (:186-190)
            out_params[count] = ev.param;
            count++;
        }
    }
    nt->t = end_t;
    return count;
}

// ---- PolyphonyDispatcher (notes.zig:209-349) -----------------------------------------------------------
int zcu_poly_dispatcher_init(zcu_poly_dispatcher* pd, uint32_t polyphony) {
    if (!pd || polyphony == 0 || polyphony > ZCU_MAX_POLYPHONY) return ZCU_ERR_INVALID;
    pd->polyphony = polyphony;
    zcu_poly_dispatcher_reset(pd);
    return ZCU_OK;
}
void zcu_poly_dispatcher_reset(zcu_poly_dispatcher* pd) {
    for (uint32_t i = 0; i < ZCU_MAX_POLYPHONY; i++) pd->slots[i] = zcu_poly_slot{0, 0, 0, 0};
}

// slot policy (:246-306): note-off -> the slot still sounding that note id, else dropped; note-on -> the
// first never-used slot, else the released slot with the oldest event, else the sounding slot with the
// oldest event.
static int choose_slot(const zcu_poly_dispatcher* pd, uint32_t note_id, bool note_on) {
    const uint32_t n = pd->polyphony;
    if (!note_on) {
        for (uint32_t i = 0; i < n; i++)
            if (pd->slots[i].used && pd->slots[i].note_id == note_id && pd->slots[i].note_on) return (int)i;
        return -1;
    }
    int best_released = -1;
    for (uint32_t i = 0; i < n; i++) {
        if (!pd->slots[i].used) return (int)i;
        if (!pd->slots[i].note_on && (best_released < 0 || pd->slots[i].event_id < pd->slots[best_released].event_id)) best_released = (int)i;
    }
    if (best_released >= 0) return best_released;
    int best = 0;
    for (uint32_t i = 1; i < n; i++)
        if (pd->slots[i].event_id < pd->slots[best].event_id) best = (int)i;
    return best;
}

int zcu_poly_dispatcher_dispatch(zcu_poly_dispatcher* pd, const zcu_impulse* impulses, const uint32_t* params, const int32_t* note_on,
                                 uint32_t n, zcu_impulse* out_impulses, uint32_t* out_params, uint32_t* out_counts) {
    if (!pd || pd->polyphony == 0 || pd->polyphony > ZCU_MAX_POLYPHONY) return ZCU_ERR_INVALID;
    for (uint32_t i = 0; i < pd->polyphony; i++) out_counts[i] = 0;
    for (uint32_t i = 0; i < n; i++) {
        const int slot = choose_slot(pd, impulses[i].note_id, note_on[i] != 0);
        if (slot < 0) continue;
        pd->slots[slot] = zcu_poly_slot{1, impulses[i].note_id, impulses[i].event_id, note_on[i] != 0 ? 1 : 0};
        const uint32_t c = out_counts[slot];
        if (c < 32) {
            out_impulses[(size_t)slot * 32 + c] = impulses[i];
            out_params[(size_t)slot * 32 + c] = params[i];
            out_counts[slot] = c + 1;
        }
    }
    return ZCU_OK;
}

uint32_t zcu_ramp_advance(float* t, float t_step, uint32_t n, int32_t* finished) {
    bool fin = false;
    const uint32_t done = zcu::ramp_advance(*t, t_step, n, fin);
    if (finished) *finished = fin ? 1 : 0;
    return done;
}

void zcu_accum_advance(float* t, float step, uint32_t n) { zcu::accum_advance(*t, step, n); }
void zcu_cycle_advance(float* t, float step, uint32_t n) { zcu::cycle_advance(*t, step, n); }

// ---- Voice(T) event logic (examples/example_song.zig:287-350) ------------------------------------------
int zcu_poly_voice_init(zcu_poly_voice* pv, const zcu_song_event* song, uint32_t song_len, uint32_t polyphony) {
    if (!pv) return ZCU_ERR_INVALID;
    zcu_note_tracker_init(&pv->tracker, song, song_len);
    const int rc = zcu_poly_dispatcher_init(&pv->dispatcher, polyphony);
    for (uint32_t i = 0; i < ZCU_MAX_POLYPHONY; i++) zcu_trigger_init(&pv->triggers[i]);
    return rc;
}
void zcu_poly_voice_reset(zcu_poly_voice* pv) {
    zcu_note_tracker_reset(&pv->tracker);
    zcu_poly_dispatcher_reset(&pv->dispatcher);
    for (uint32_t i = 0; i < ZCU_MAX_POLYPHONY; i++) zcu_trigger_reset(&pv->triggers[i]);
}

int64_t zcu_poly_voice_spans(zcu_poly_voice* pv, const int32_t* param_note_on, float sample_rate, uint32_t buffer_frames,
                             uint32_t num_buffers, uint32_t last_buffer_frames, uint32_t* voice_off_out, zcu_span* spans_out,
                             uint64_t max_spans) {
    if (!pv || !voice_off_out || pv->dispatcher.polyphony == 0 || pv->dispatcher.polyphony > ZCU_MAX_POLYPHONY) return 0;
    const uint32_t P = pv->dispatcher.polyphony;
    zcu_poly_voice w = *pv;  // work on a copy: commit only when the table fits
    std::vector<std::vector<zcu_span>> per(P);
    zcu_impulse imps[32], poly_imps[ZCU_MAX_POLYPHONY * 32];
    uint32_t params[32], poly_params[ZCU_MAX_POLYPHONY * 32], counts[ZCU_MAX_POLYPHONY];
    int32_t on[32];
    zcu_span tmp[40];
    for (uint32_t b = 0; b < num_buffers; b++) {
        const uint32_t len = (b + 1 == num_buffers) ? last_buffer_frames : buffer_frames;
        const uint32_t base = b * buffer_frames;
        const uint32_t n = zcu_note_tracker_consume(&w.tracker, sample_rate, 0, len, imps, params, 32);
        for (uint32_t i = 0; i < n; i++) on[i] = param_note_on ? param_note_on[params[i]] : 1;
        zcu_poly_dispatcher_dispatch(&w.dispatcher, imps, params, on, n, poly_imps, poly_params, counts);
        for (uint32_t i = 0; i < P; i++) {
            const uint32_t k = zcu_trigger_run(&w.triggers[i], 0, len, poly_imps + (size_t)i * 32, poly_params + (size_t)i * 32, counts[i], base, tmp, 40);
            per[i].insert(per[i].end(), tmp, tmp + (k < 40 ? k : 40));
        }
    }
    uint64_t total = 0;
    for (uint32_t i = 0; i < P; i++) total += per[i].size();
    if (total > max_spans || (total > 0 && !spans_out)) return -(int64_t)total;
    uint32_t off = 0;
    for (uint32_t i = 0; i < P; i++) {
        voice_off_out[i] = off;
        if (!per[i].empty()) std::memcpy(spans_out + off, per[i].data(), per[i].size() * sizeof(zcu_span));
        off += (uint32_t)per[i].size();
    }
    voice_off_out[P] = off;
    *pv = w;
    return (int64_t)total;
}

// ---- a bank of monophonic voices (examples/example_polyphony.zig:61-92 per voice) -------------------------
void zcu_bank_trigger_init(zcu_bank_trigger* state, uint32_t voices, const uint32_t* event_off) {
    for (uint32_t v = 0; v < voices; v++) {
        zcu_trigger_init(&state[v].trigger);
        state[v].next_event = event_off ? event_off[v] : 0;
    }
}

int64_t zcu_bank_spans(zcu_bank_trigger* state, uint32_t voices, const uint32_t* event_off, const zcu_note_event* events,
                       uint32_t buffer_frames, uint32_t total_frames, uint32_t first_frame, uint32_t num_frames,
                       uint32_t* voice_off_out, zcu_span* spans_out, uint64_t max_spans) {
    if (!state || !event_off || !voice_off_out || buffer_frames == 0 || first_frame % buffer_frames != 0) return 0;
    std::vector<zcu_bank_trigger> w(state, state + voices);  // commit only when the table fits
    std::vector<zcu_span> all;
    std::vector<uint32_t> off(voices + 1, 0);
    zcu_impulse imps[32];
    uint32_t params[32];
    zcu_span tmp[40];
    const uint32_t end_frame = first_frame + num_frames < total_frames ? first_frame + num_frames : total_frames;
    for (uint32_t v = 0; v < voices; v++) {
        off[v] = (uint32_t)all.size();
        const uint32_t e0 = event_off[v], e1 = event_off[v + 1];
        uint32_t ei = w[v].next_event;
        for (uint32_t start = first_frame; start < end_frame; start += buffer_frames) {
            const uint32_t len = (total_frames - start < buffer_frames) ? total_frames - start : buffer_frames;
            // ImpulseQueue.push (notes.zig:102-127) of this buffer's events: capacity 32, later ones are dropped
            uint32_t n = 0;
            while (ei < e1 && events[ei].frame < start + len) {
                if (n < 32) {
                    imps[n] = zcu_impulse{events[ei].frame - start, events[ei].note_id, ei - e0 + 1};
                    params[n] = ei;
                    n++;
                }
                ei++;
            }
            const uint32_t k = zcu_trigger_run(&w[v].trigger, 0, len, imps, params, n, start - first_frame, tmp, 40);
            all.insert(all.end(), tmp, tmp + (k < 40 ? k : 40));
        }
        w[v].next_event = ei;
    }
    off[voices] = (uint32_t)all.size();
    if (all.size() > max_spans || (!all.empty() && !spans_out)) return -(int64_t)all.size();
    std::memcpy(voice_off_out, off.data(), (size_t)(voices + 1) * sizeof(uint32_t));
    if (!all.empty()) std::memcpy(spans_out, all.data(), all.size() * sizeof(zcu_span));
    std::memcpy(state, w.data(), (size_t)voices * sizeof(zcu_bank_trigger));
    return (int64_t)all.size();
}

}  // extern "C"
