// zcu_poly_dev.cu — NoteTracker + PolyphonyDispatcher + one Trigger per sub-voice ON THE DEVICE: the device
// form of zcu_poly_voice_spans (zcu_events.cpp), i.e. of Voice(T) in examples/example_song.zig:287-350 without the module
// paint.  A bank holds many independent polyphonic voices ("songs": a list of (t, note_id, params row) each, one polyphony for
// all); one thread per song replays, per mix buffer,
//     NoteTracker.consume (notes.zig:169-200: f32 time accumulate, frame = min(uint(f * len), len - 1), at most 32 impulses)
//  -> PolyphonyDispatcher.dispatch (notes.zig:246-306: note-off to the slot sounding that id, note-on to the first unused slot,
//     else the oldest released, else the oldest event)
//  -> Trigger.next loop of every slot (trigger.zig:80-196)
// twice over the block — count, exclusive scan, write — into a span table whose voice column is song * polyphony + slot.
// Written against the song list directly (not a translation of the host functions); tests/test_gpu_poly_dev.py holds it to the
// host's bits on the reference's three dispatcher tests (src/zang/notes_test.zig:10-119) and on fuzzed songs over several blocks.
#include <cub/device/device_scan.cuh>

#include <vector>

#include "zcu_internal.h"

struct zcu_poly_bank {
    zcu_ctx* ctx;
    uint32_t songs, polyphony, buffer_frames, nevents, nparams;
    float sample_rate;
    uint32_t* d_song_off;
    zcu_song_event* d_events;
    int32_t* d_note_on;  // [nparams] or null (every impulse is a note-on)
    zcu_poly_voice* d_state;  // [songs]; tracker.song is unused on the device (the list is d_events[d_song_off[s] ..])
    uint32_t* d_counts;       // [songs * polyphony + 1]
    void* d_scan_tmp;
    size_t scan_tmp_bytes;
};

namespace {

struct PolyImp {  // one impulse of the current mix buffer
    uint32_t frame, note_id, event_id, param;
};

// the Trigger loop of one slot over the impulses imp[sel[0 .. n)] of a buffer of `len` frames (same statement of
// trigger.zig:80-196 as zcu_spans_dev.cu's, over an index list)
template <bool EMIT>
__device__ uint32_t poly_trigger(zcu_trigger& t, const PolyImp* imp, const uint8_t* sel, uint32_t n, uint32_t len, uint32_t frame_base, zcu_span* out) {
    uint32_t produced = 0, idx = 0, pos = 0;
    while (pos < len) {
        uint32_t span_end = len;
        bool has = false, carried = false;
        uint32_t id = 0, param = 0;
        if (t.has_note) {
            if (idx < n) {
                const uint32_t nf = imp[sel[idx]].frame;
                if (nf > pos) {
                    carried = true;
                    span_end = min(len, nf);
                }
            } else {
                carried = true;
            }
            if (carried) {
                has = true;
                id = t.note_id;
                param = t.param;
            }
        }
        if (!carried) {
            for (uint32_t i = idx; i < n; i++) {
                const uint32_t fr = imp[sel[i]].frame;
                if (fr >= len) break;
                if (fr > pos) {
                    span_end = fr;
                    break;
                }
                idx++;
                const uint32_t note_end = (i + 1 < n) ? min(len, imp[sel[i + 1]].frame) : len;
                if (note_end <= pos) continue;
                has = true;
                id = imp[sel[i]].note_id;
                param = imp[sel[i]].param;
                span_end = note_end;
                break;
            }
        }
        if (has) {
            if (EMIT) {
                const bool changed = t.has_note ? (id != t.note_id) : true;
                out[produced] = zcu_span{pos + frame_base, span_end + frame_base, changed ? ZCU_SPAN_NOTE_ID_CHANGED : 0u, param};
            }
            produced++;
            t.has_note = 1;
            t.note_id = id;
            t.param = param;
        }
        pos = span_end;
    }
    return produced;
}

// slot policy of notes.zig:246-306
__device__ int poly_choose_slot(const zcu_poly_slot* slots, uint32_t P, uint32_t note_id, bool note_on) {
    if (!note_on) {
        for (uint32_t i = 0; i < P; i++)
            if (slots[i].used && slots[i].note_id == note_id && slots[i].note_on) return (int)i;
        return -1;
    }
    int best_released = -1;
    for (uint32_t i = 0; i < P; i++) {
        if (!slots[i].used) return (int)i;
        if (!slots[i].note_on && (best_released < 0 || slots[i].event_id < slots[best_released].event_id)) best_released = (int)i;
    }
    if (best_released >= 0) return best_released;
    int best = 0;
    for (uint32_t i = 1; i < P; i++)
        if (slots[i].event_id < slots[best].event_id) best = (int)i;
    return best;
}

template <bool EMIT>
__global__ void __launch_bounds__(64) k_poly_replay(zcu_poly_voice* __restrict__ state, const uint32_t* __restrict__ song_off,
                                                    const zcu_song_event* __restrict__ ev, const int32_t* __restrict__ note_on_rows, uint32_t S,
                                                    uint32_t P, float sample_rate, uint32_t L, uint32_t num_buffers, uint32_t last_len,
                                                    uint32_t* __restrict__ counts, const uint32_t* __restrict__ off, zcu_span* __restrict__ spans) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    zcu_poly_voice w = state[s];
    const uint32_t e0 = song_off[s], song_len = song_off[s + 1] - e0;
    uint32_t produced[ZCU_MAX_POLYPHONY];
    for (uint32_t i = 0; i < P; i++) produced[i] = 0;
    PolyImp imp[32];
    uint8_t slot_of[32], sel[32];
    for (uint32_t b = 0; b < num_buffers; b++) {
        const uint32_t len = (b + 1 == num_buffers) ? last_len : L;
        // NoteTracker.consume (notes.zig:169-200)
        const float flen = (float)len;
        const float buf_time = flen / sample_rate;
        const float end_t = w.tracker.t + buf_time;
        uint32_t n = 0;
        while (w.tracker.next_song_event < song_len) {
            const zcu_song_event e = ev[e0 + w.tracker.next_song_event];
            if (!(e.t < end_t)) break;
            const float f = (e.t - w.tracker.t) / buf_time;
            uint32_t rel = (uint32_t)(long long)(f * flen);  // as the host's conversion (through 64 bits)
            if (rel > len - 1) rel = len - 1;
            w.tracker.next_song_event++;
            if (n < 32) {
                imp[n] = PolyImp{rel, e.note_id, w.tracker.next_song_event, e.param};
                n++;
            }
        }
        w.tracker.t = end_t;
        // PolyphonyDispatcher.dispatch
        uint32_t per_slot[ZCU_MAX_POLYPHONY];
        for (uint32_t i = 0; i < P; i++) per_slot[i] = 0;
        for (uint32_t i = 0; i < n; i++) {
            const bool on = note_on_rows ? note_on_rows[imp[i].param] != 0 : true;
            const int slot = poly_choose_slot(w.dispatcher.slots, P, imp[i].note_id, on);
            slot_of[i] = 0xFF;
            if (slot < 0) continue;
            w.dispatcher.slots[slot] = zcu_poly_slot{1, imp[i].note_id, imp[i].event_id, on ? 1 : 0};
            if (per_slot[slot] < 32) {
                slot_of[i] = (uint8_t)slot;
                per_slot[slot]++;
            }
        }
        // one Trigger per slot over its impulses
        for (uint32_t k = 0; k < P; k++) {
            uint32_t m = 0;
            for (uint32_t i = 0; i < n; i++)
                if (slot_of[i] == k) sel[m++] = (uint8_t)i;
            zcu_span* out = EMIT ? spans + off[(size_t)s * P + k] + produced[k] : nullptr;
            produced[k] += poly_trigger<EMIT>(w.triggers[k], imp, sel, m, len, b * L, out);
        }
    }
    if (EMIT) {
        state[s] = w;
    } else {
        for (uint32_t k = 0; k < P; k++) counts[(size_t)s * P + k] = produced[k];
        if (s == 0) counts[(size_t)S * P] = 0;
    }
}

__global__ void k_poly_reset(zcu_poly_voice* state, uint32_t S, uint32_t P) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    zcu_poly_voice w;
    w.tracker.song = nullptr;
    w.tracker.song_len = 0;
    w.tracker.next_song_event = 0;
    w.tracker.t = 0.0f;
    w.dispatcher.polyphony = P;
    for (uint32_t i = 0; i < ZCU_MAX_POLYPHONY; i++) {
        w.dispatcher.slots[i] = zcu_poly_slot{0, 0, 0, 0};
        w.triggers[i] = zcu_trigger{0, 0, 0};
    }
    state[s] = w;
}

}  // namespace

extern "C" {

int zcu_poly_bank_create(zcu_ctx* ctx, uint32_t songs, uint32_t polyphony, const uint32_t* song_off_host, const zcu_song_event* events_host,
                         const int32_t* param_note_on_host, uint32_t nparams, float sample_rate, uint32_t buffer_frames, zcu_poly_bank** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, songs > 0 && song_off_host != nullptr, "songs and song_off required");
    ZCU_REQUIRE(ctx, polyphony >= 1 && polyphony <= ZCU_MAX_POLYPHONY, "polyphony must be 1..32");
    ZCU_REQUIRE(ctx, (uint64_t)songs * polyphony < (1ull << 31), "too many sub-voices");
    ZCU_REQUIRE(ctx, buffer_frames > 0 && sample_rate > 0.0f, "buffer_frames and sample_rate must be > 0");
    ZCU_REQUIRE(ctx, song_off_host[0] == 0, "song_off[0] must be 0");
    const uint32_t n = song_off_host[songs];
    ZCU_REQUIRE(ctx, n == 0 || events_host != nullptr, "null events");
    for (uint32_t s = 0; s < songs; s++) ZCU_REQUIRE(ctx, song_off_host[s] <= song_off_host[s + 1] && song_off_host[s + 1] <= n, "song_off must be non-decreasing");
    if (param_note_on_host)
        for (uint32_t k = 0; k < n; k++) ZCU_REQUIRE(ctx, events_host[k].param < nparams, "a song event names a params row beyond nparams");
    zcu_poly_bank* b = new zcu_poly_bank{ctx, songs, polyphony, buffer_frames, n, nparams, sample_rate, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    const size_t cols = (size_t)songs * polyphony;
    cudaError_t e = cudaMalloc(&b->d_song_off, (size_t)(songs + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_events, (size_t)(n ? n : 1) * sizeof(zcu_song_event));
    if (e == cudaSuccess && param_note_on_host) e = cudaMalloc(&b->d_note_on, (size_t)(nparams ? nparams : 1) * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_state, (size_t)songs * sizeof(zcu_poly_voice));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_counts, (cols + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) {
        cub::DeviceScan::ExclusiveSum(nullptr, b->scan_tmp_bytes, b->d_counts, b->d_counts, (int)(cols + 1), ctx->stream);
        e = cudaMalloc(&b->d_scan_tmp, b->scan_tmp_bytes ? b->scan_tmp_bytes : 16);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(b->d_song_off, song_off_host, (size_t)(songs + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(b->d_events, events_host, (size_t)n * sizeof(zcu_song_event), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && param_note_on_host && nparams)
        e = cudaMemcpyAsync(b->d_note_on, param_note_on_host, (size_t)nparams * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        k_poly_reset<<<(songs + 127) / 128, 128, 0, ctx->stream>>>(b->d_state, songs, polyphony);
        ctx->launches++;
        e = cudaStreamSynchronize(ctx->stream);  // the host arrays may be temporaries
    }
    if (e != cudaSuccess) {
        zcu_poly_bank_free(b);
        return zcu_fail(ctx, e == cudaErrorMemoryAllocation ? ZCU_ERR_OOM : ZCU_ERR_CUDA, "zcu_poly_bank_create: %s", cudaGetErrorString(e));
    }
    *out = b;
    return ZCU_OK;
}

void zcu_poly_bank_free(zcu_poly_bank* b) {
    if (!b) return;
    cudaStreamSynchronize(b->ctx->stream);
    if (b->d_song_off) cudaFree(b->d_song_off);
    if (b->d_events) cudaFree(b->d_events);
    if (b->d_note_on) cudaFree(b->d_note_on);
    if (b->d_state) cudaFree(b->d_state);
    if (b->d_counts) cudaFree(b->d_counts);
    if (b->d_scan_tmp) cudaFree(b->d_scan_tmp);
    delete b;
}

int zcu_poly_bank_reset(zcu_ctx* ctx, zcu_poly_bank* b) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, b != nullptr, "null poly bank");
    k_poly_reset<<<(b->songs + 127) / 128, 128, 0, ctx->stream>>>(b->d_state, b->songs, b->polyphony);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

int zcu_poly_bank_get_state(zcu_ctx* ctx, const zcu_poly_bank* b, zcu_poly_voice* state_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, b && state_host, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(state_host, b->d_state, (size_t)b->songs * sizeof(zcu_poly_voice), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}

int zcu_poly_bank_spans(zcu_ctx* ctx, zcu_poly_bank* b, uint32_t num_buffers, uint32_t last_buffer_frames, zcu_spans** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, b != nullptr, "null poly bank");
    ZCU_REQUIRE(ctx, num_buffers > 0 && last_buffer_frames > 0 && last_buffer_frames <= b->buffer_frames, "bad block shape");
    const uint64_t frames = (uint64_t)(num_buffers - 1) * b->buffer_frames + last_buffer_frames;
    const uint64_t cols = (uint64_t)b->songs * b->polyphony;
    const uint64_t cap = (uint64_t)b->nevents + cols * num_buffers;  // worst case: every event in this block + one carried span per slot and buffer
    ZCU_REQUIRE(ctx, frames < (1ull << 32) && cap < (1ull << 32), "span table too large");
    zcu_spans* s = new zcu_spans{ctx, (uint32_t)cols, (uint32_t)frames, (uint32_t)cap, nullptr, nullptr, b->nparams ? b->nparams - 1 : 0xFFFFFFFFu};
    cudaError_t e = cudaMalloc(&s->d_off, (size_t)(cols + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_spans, (size_t)(cap ? cap : 1) * sizeof(zcu_span));
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_OOM, "span table allocation failed: %s", cudaGetErrorString(e));
    }
    const uint32_t grid = (b->songs + 63) / 64;
    k_poly_replay<false><<<grid, 64, 0, ctx->stream>>>(b->d_state, b->d_song_off, b->d_events, b->d_note_on, b->songs, b->polyphony, b->sample_rate,
                                                       b->buffer_frames, num_buffers, last_buffer_frames, b->d_counts, nullptr, nullptr);
    cub::DeviceScan::ExclusiveSum(b->d_scan_tmp, b->scan_tmp_bytes, b->d_counts, s->d_off, (int)(cols + 1), ctx->stream);
    k_poly_replay<true><<<grid, 64, 0, ctx->stream>>>(b->d_state, b->d_song_off, b->d_events, b->d_note_on, b->songs, b->polyphony, b->sample_rate,
                                                      b->buffer_frames, num_buffers, last_buffer_frames, nullptr, s->d_off, s->d_spans);
    ctx->launches += 3;
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_CUDA, "zcu_poly_bank_spans: %s", cudaGetErrorString(e));
    }
    *out = s;
    return ZCU_OK;
}

}  // extern "C"
