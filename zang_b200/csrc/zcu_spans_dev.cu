// zcu_spans_dev.cu — span tables built ON THE DEVICE from device-resident note events: the
// device form of zcu_bank_spans (zcu_events.cpp).  Per voice and mix buffer: ImpulseQueue.push of the buffer's events
// (capacity 32, notes.zig:102-127), then the whole `while (trigger.next(&ctr))` loop (trigger.zig:80-196).  One thread per
// voice replays its Trigger over the block twice: once to count its spans, once (after an exclusive scan of the counts) to
// write them in the CSR layout of zcu_spans_upload.  Nothing but the block bounds crosses PCIe.
//
// The replay below is written against the events directly (frame / note_id read from the voice's event list, param = the
// global event index), not against the host's zcu_trigger_run; tests/test_gpu_spans_dev.py holds the two to the same bits on
// the reference's own trigger tests (src/zang/trigger_test.zig:38-132) and on fuzzed streams.
#include <cub/device/device_scan.cuh>

#include <vector>

#include "zcu_internal.h"

struct zcu_event_bank {
    zcu_ctx* ctx;
    uint32_t voices, buffer_frames, total_frames, nevents;
    uint32_t* d_event_off;
    zcu_note_event* d_events;
    zcu_bank_trigger* d_state;
    int32_t* d_prev_on;  // [voices]: note_on of the last span painted (zcu_spans_prev_note_on_views), 0 after create / reset
    uint32_t* d_counts;  // [voices + 1] scan input
    void* d_scan_tmp;
    size_t scan_tmp_bytes;
    std::vector<uint32_t> h_event_off;  // for reset
};

namespace {

// One mix buffer [start, start + len) of one voice.  EMIT = false counts, true writes.  Returns the spans produced.
template <bool EMIT>
__device__ __forceinline__ uint32_t trigger_buffer(zcu_trigger& t, const zcu_note_event* __restrict__ ev, uint32_t& ei, uint32_t e1, uint32_t start,
                                                   uint32_t len, uint32_t frame_base, zcu_span* out) {
    // the buffer's impulses: events[i0 .. i0 + n), at most 32 (later ones are dropped by the queue but still consumed)
    const uint32_t i0 = ei;
    const uint32_t lim = start + len;
    uint32_t n = 0;
    while (ei < e1 && ev[ei].frame < lim) {
        if (n < 32) n++;
        ei++;
    }
    uint32_t produced = 0;
    uint32_t idx = 0;  // Counter.impulse_index
    uint32_t pos = 0;  // frames relative to the buffer
    while (pos < len) {
        uint32_t span_end = len;
        bool has = false, carried = false;
        uint32_t id = 0, param = 0;
        if (t.has_note) {  // carryOver :107-130: the current note plays up to the next impulse
            if (idx < n) {
                const uint32_t nf = ev[i0 + idx].frame - start;
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
        if (!carried) {  // getNextNoteSpan :132-196
            for (uint32_t i = idx; i < n; i++) {
                const uint32_t fr = ev[i0 + i].frame - start;
                if (fr >= len) break;  // cannot happen here (events of this buffer only); kept for symmetry with the host
                if (fr > pos) {        // gap before the note begins
                    span_end = fr;
                    break;
                }
                idx++;
                const uint32_t note_end = (i + 1 < n) ? min(len, ev[i0 + i + 1].frame - start) : len;
                if (note_end <= pos) continue;  // superseded by the next impulse at the same frame
                has = true;
                id = ev[i0 + i].note_id;
                param = i0 + i;
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

template <bool EMIT>
__global__ void __launch_bounds__(128) k_spans_replay(zcu_bank_trigger* __restrict__ state, const uint32_t* __restrict__ event_off,
                                                      const zcu_note_event* __restrict__ ev, uint32_t V, uint32_t L, uint32_t total, uint32_t first,
                                                      uint32_t end_frame, uint32_t* __restrict__ counts, const uint32_t* __restrict__ off,
                                                      zcu_span* __restrict__ spans) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    zcu_bank_trigger w = state[v];
    const uint32_t e1 = event_off[v + 1];
    uint32_t ei = w.next_event;
    uint32_t produced = 0;
    zcu_span* out = EMIT ? spans + off[v] : nullptr;
    for (uint32_t start = first; start < end_frame; start += L) {
        const uint32_t len = min(L, total - start);
        produced += trigger_buffer<EMIT>(w.trigger, ev, ei, e1, start, len, start - first, EMIT ? out + produced : nullptr);
    }
    if (EMIT) {
        w.next_event = ei;
        state[v] = w;  // the Trigger state is carried to the next block
    } else {
        counts[v] = produced;
        if (v == 0) counts[V] = 0;
    }
}

__global__ void k_event_bank_reset(zcu_bank_trigger* state, int32_t* prev_on, const uint32_t* event_off, uint32_t V) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    prev_on[v] = 0;
    zcu_bank_trigger s;
    s.trigger.has_note = 0;
    s.trigger.note_id = 0;
    s.trigger.param = 0;
    s.next_event = event_off[v];
    state[v] = s;
}

// views of a Trigger table for an instrument that keeps `prev_note_on` (examples/example_portamento.zig:51 `defer
// self.prev_note_on = params.note_on`): per voice, in span order, prev = note_on of the previous span's event (carried
// across blocks in prev_state).  porta.param = 2*param + prev; env.flags = (!prev && on); plain.param = 0.
__global__ void __launch_bounds__(128) k_spans_prev_on_views(const uint32_t* __restrict__ off, const zcu_span* __restrict__ src,
                                                            const zcu_note_event* __restrict__ ev, int32_t* __restrict__ prev_state, uint32_t V,
                                                            zcu_span* __restrict__ porta, zcu_span* __restrict__ env, zcu_span* __restrict__ plain) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    int32_t prev = prev_state[v];
    for (uint32_t k = off[v], ke = off[v + 1]; k < ke; k++) {
        const zcu_span s = src[k];
        const int32_t on = ev[s.param].note_on;
        porta[k] = zcu_span{s.start, s.end, s.flags, 2u * s.param + (prev ? 1u : 0u)};
        env[k] = zcu_span{s.start, s.end, (!prev && on) ? 1u : 0u, s.param};
        plain[k] = zcu_span{s.start, s.end, s.flags, 0u};
        prev = on;
    }
    prev_state[v] = prev;
}

}  // namespace

extern "C" {

int zcu_event_bank_create(zcu_ctx* ctx, uint32_t voices, const uint32_t* event_off_host, const zcu_note_event* events_host, uint32_t buffer_frames,
                          uint32_t total_frames, zcu_event_bank** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, voices > 0 && event_off_host != nullptr, "voices and event_off required");
    ZCU_REQUIRE(ctx, buffer_frames > 0 && total_frames > 0, "buffer_frames and total_frames must be > 0");
    ZCU_REQUIRE(ctx, event_off_host[0] == 0, "event_off[0] must be 0");
    const uint32_t n = event_off_host[voices];
    ZCU_REQUIRE(ctx, n == 0 || events_host != nullptr, "null events");
    for (uint32_t v = 0; v < voices; v++) {
        ZCU_REQUIRE(ctx, event_off_host[v] <= event_off_host[v + 1] && event_off_host[v + 1] <= n, "event_off must be non-decreasing");
        for (uint32_t k = event_off_host[v]; k < event_off_host[v + 1]; k++) {
            ZCU_REQUIRE(ctx, events_host[k].frame < total_frames, "event frame beyond total_frames");
            ZCU_REQUIRE(ctx, k == event_off_host[v] || events_host[k - 1].frame <= events_host[k].frame, "a voice's events must be chronological");
        }
    }
    zcu_event_bank* b = new zcu_event_bank{ctx, voices, buffer_frames, total_frames, n, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, {}};
    b->h_event_off.assign(event_off_host, event_off_host + voices + 1);
    cudaError_t e = cudaMalloc(&b->d_event_off, (size_t)(voices + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_events, (size_t)(n ? n : 1) * sizeof(zcu_note_event));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_state, (size_t)voices * sizeof(zcu_bank_trigger));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_prev_on, (size_t)voices * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc(&b->d_counts, (size_t)(voices + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) {
        cub::DeviceScan::ExclusiveSum(nullptr, b->scan_tmp_bytes, b->d_counts, b->d_counts, (int)(voices + 1), ctx->stream);
        e = cudaMalloc(&b->d_scan_tmp, b->scan_tmp_bytes ? b->scan_tmp_bytes : 16);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(b->d_event_off, event_off_host, (size_t)(voices + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(b->d_events, events_host, (size_t)n * sizeof(zcu_note_event), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        k_event_bank_reset<<<(voices + 127) / 128, 128, 0, ctx->stream>>>(b->d_state, b->d_prev_on, b->d_event_off, voices);
        ctx->launches++;
        e = cudaStreamSynchronize(ctx->stream);  // the host arrays may be temporaries
    }
    if (e != cudaSuccess) {
        zcu_event_bank_free(b);
        return zcu_fail(ctx, e == cudaErrorMemoryAllocation ? ZCU_ERR_OOM : ZCU_ERR_CUDA, "zcu_event_bank_create: %s", cudaGetErrorString(e));
    }
    *out = b;
    return ZCU_OK;
}

void zcu_event_bank_free(zcu_event_bank* b) {
    if (!b) return;
    cudaStreamSynchronize(b->ctx->stream);
    if (b->d_event_off) cudaFree(b->d_event_off);
    if (b->d_events) cudaFree(b->d_events);
    if (b->d_state) cudaFree(b->d_state);
    if (b->d_prev_on) cudaFree(b->d_prev_on);
    if (b->d_counts) cudaFree(b->d_counts);
    if (b->d_scan_tmp) cudaFree(b->d_scan_tmp);
    delete b;
}

int zcu_event_bank_reset(zcu_ctx* ctx, zcu_event_bank* b) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, b != nullptr, "null event bank");
    k_event_bank_reset<<<(b->voices + 127) / 128, 128, 0, ctx->stream>>>(b->d_state, b->d_prev_on, b->d_event_off, b->voices);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

int zcu_event_bank_get_state(zcu_ctx* ctx, const zcu_event_bank* b, zcu_bank_trigger* state_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, b && state_host, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(state_host, b->d_state, (size_t)b->voices * sizeof(zcu_bank_trigger), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}

int zcu_event_bank_set_state(zcu_ctx* ctx, zcu_event_bank* b, const zcu_bank_trigger* state_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, b && state_host, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(b->d_state, state_host, (size_t)b->voices * sizeof(zcu_bank_trigger), cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}

int zcu_event_bank_spans(zcu_ctx* ctx, zcu_event_bank* b, uint32_t first_frame, uint32_t num_frames, zcu_spans** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, b != nullptr, "null event bank");
    ZCU_REQUIRE(ctx, first_frame % b->buffer_frames == 0, "first_frame must be a multiple of buffer_frames");
    ZCU_REQUIRE(ctx, num_frames > 0 && first_frame < b->total_frames, "empty block");
    const uint32_t end_frame = (uint64_t)first_frame + num_frames < b->total_frames ? first_frame + num_frames : b->total_frames;
    const uint32_t nbuf = (end_frame - first_frame + b->buffer_frames - 1) / b->buffer_frames;
    // room for the worst case (every event of the bank in this block + one carried span per voice and buffer): the table is
    // built without a host round trip, so its exact size is only known on the device (zcu_spans_count reads it back)
    const uint64_t cap = (uint64_t)b->nevents + (uint64_t)b->voices * nbuf;
    ZCU_REQUIRE(ctx, cap < (1ull << 32), "span table too large");
    zcu_spans* s = new zcu_spans{ctx, b->voices, num_frames, (uint32_t)cap, nullptr, nullptr, b->nevents ? b->nevents - 1 : 0};
    cudaError_t e = cudaMalloc(&s->d_off, (size_t)(b->voices + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_spans, (size_t)(cap ? cap : 1) * sizeof(zcu_span));
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_OOM, "span table allocation failed: %s", cudaGetErrorString(e));
    }
    const uint32_t grid = (b->voices + 127) / 128;
    k_spans_replay<false><<<grid, 128, 0, ctx->stream>>>(b->d_state, b->d_event_off, b->d_events, b->voices, b->buffer_frames, b->total_frames,
                                                         first_frame, end_frame, b->d_counts, nullptr, nullptr);
    cub::DeviceScan::ExclusiveSum(b->d_scan_tmp, b->scan_tmp_bytes, b->d_counts, s->d_off, (int)(b->voices + 1), ctx->stream);
    k_spans_replay<true><<<grid, 128, 0, ctx->stream>>>(b->d_state, b->d_event_off, b->d_events, b->voices, b->buffer_frames, b->total_frames,
                                                        first_frame, end_frame, nullptr, s->d_off, s->d_spans);
    ctx->launches += 3;
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_CUDA, "zcu_event_bank_spans: %s", cudaGetErrorString(e));
    }
    *out = s;
    return ZCU_OK;
}

int zcu_spans_download(zcu_ctx* ctx, const zcu_spans* s, uint32_t* voice_off_out, zcu_span* spans_out, uint64_t max_spans, uint64_t* nspans_out) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, s && voice_off_out && nspans_out, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(voice_off_out, s->d_off, (size_t)(s->voices + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint64_t n = voice_off_out[s->voices];
    *nspans_out = n;
    if (n > max_spans) return ZCU_OK;  // the caller sees the count and retries with room
    ZCU_REQUIRE(ctx, n == 0 || spans_out != nullptr, "null spans_out");
    if (n) ZCU_CUDA(ctx, cudaMemcpyAsync(spans_out, s->d_spans, (size_t)n * sizeof(zcu_span), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}

int zcu_spans_prev_note_on_views(zcu_ctx* ctx, zcu_event_bank* b, const zcu_spans* src, zcu_spans** porta_out, zcu_spans** env_out,
                                 zcu_spans** plain_out) {
    if (!ctx || !porta_out || !env_out || !plain_out) return ZCU_ERR_INVALID;
    *porta_out = *env_out = *plain_out = nullptr;
    ZCU_REQUIRE(ctx, b && src, "null argument");
    ZCU_REQUIRE(ctx, src->voices == b->voices, "span table / event bank voice count mismatch");
    zcu_spans* outs[3] = {nullptr, nullptr, nullptr};
    const uint32_t maxp[3] = {2u * src->max_param + 1u, src->max_param, 0u};
    for (int i = 0; i < 3; i++) {
        zcu_spans* s = new zcu_spans{ctx, src->voices, src->frames, src->nspans, nullptr, nullptr, maxp[i]};
        outs[i] = s;
        // the three views share the source's offsets: each gets its own copy so that it can be freed like any other table
        cudaError_t e = cudaMalloc(&s->d_off, (size_t)(src->voices + 1) * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMalloc(&s->d_spans, (size_t)(src->nspans ? src->nspans : 1) * sizeof(zcu_span));
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(s->d_off, src->d_off, (size_t)(src->voices + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream);
        if (e != cudaSuccess) {
            for (int j = 0; j <= i; j++) zcu_spans_free(outs[j]);
            return zcu_fail(ctx, ZCU_ERR_OOM, "span view allocation failed: %s", cudaGetErrorString(e));
        }
    }
    k_spans_prev_on_views<<<(src->voices + 127) / 128, 128, 0, ctx->stream>>>(src->d_off, src->d_spans, b->d_events, b->d_prev_on, src->voices,
                                                                              outs[0]->d_spans, outs[1]->d_spans, outs[2]->d_spans);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    *porta_out = outs[0];
    *env_out = outs[1];
    *plain_out = outs[2];
    return ZCU_OK;
}

}  // extern "C"
