// zcu_render.cu — fused instrument banks: the B200-first path for the benchmark configurations.
//
// examples/example_polyphony.zig:61-92 runs, per voice and per mix buffer,
//     ImpulseQueue.consume -> Trigger.next loop -> Instrument.paint(span, outputs, temps, ...)
// and every voice accumulates into the shared output.  Here one thread owns one voice for a whole
// block of mix buffers: it replays the Trigger's span splitting itself (trigger.zig:80-196, one
// boundary check per frame), keeps the state of every sub-module of the instrument in registers and
// evaluates the instrument's module chain sample by sample in the reference's f32 operation order.
// The temps never exist: what the reference round-trips through temps[0]/temps[1] stays in registers.
// 32 adjacent voices form a warp; their outputs go through a padded shared-memory tile so that lane j
// can sum frame j over the 32 voices with the voices' pan gains, and each warp writes one coalesced
// row of partial mix per 32 frames.  A second kernel adds the per-warp partials in fixed order
// (deterministic) into the mix.
#include <cmath>
#include <new>
#include <vector>

#include "zcu_internal.h"
#include "zcu_modules.cuh"

using namespace zcu;

struct NiceVoiceState {
    uint32_t cnt;                                // PulseOsc.cnt
    float l, b;                                  // Filter.l/.b
    int32_t env_state;                           // Envelope.state
    float env_t, env_last, env_start;            // Envelope.painter
    int32_t has_note;                            // Trigger.note != null
    uint32_t note_id;                            // Trigger.note.id
    uint32_t cur_event;                          // Trigger.note.params (row = event index)
    uint32_t next_event;                         // ImpulseQueue read cursor
};

struct zcu_render {
    zcu_ctx* ctx;
    zcu_render_config cfg;
    uint32_t block_frames;
    uint32_t nwarps;
    float* d_color;        // [V]
    float* d_gains;        // [V][C]
    NiceVoiceState* d_state;
    uint32_t* d_event_off; // [V+1]
    zcu_note_event* d_events;
    float* d_cutoff;       // per event: Filter.cutoffFromFrequency(freq*8, sr) evaluated on the host
    uint32_t nevents;
    float* d_partial;      // [nwarps][block_frames][C]
    bool profiling;
    std::vector<cudaEvent_t> ev;   // pairs around every render-kernel launch since the last query
    size_t ev_used;
};

namespace {

struct RenderArgs {
    uint32_t V, C;
    uint32_t buffer_frames, total_frames;
    uint32_t F0, F1;  // frames of this launch
    float sample_rate;
    const float* color;
    const float* gains;
    NiceVoiceState* state;
    const uint32_t* event_off;
    const zcu_note_event* events;
    const float* cutoff;
    float* partial;   // [nwarps][F1-F0][C] or null
    float* vout;      // [vout_frames][V] window or null
    uint32_t vout_first, vout_frames;
};

// NiceInstrument (examples/modules.zig:189-248) for one voice
struct NiceVoice {
    // PulseOsc (constant frequency path, PulseOsc.zig:75-114)
    uint32_t cnt, ifreq, brpt, pstate;
    float gdf2, col, cc121, cc212;
    bool pskip;
    FilterCore flt;
    EnvelopeCore env;

    __device__ void begin_span(float sample_rate, float freq, float cutoff, bool note_on, bool note_id_changed) {
        const float gain = 0.7f;
        pskip = (freq < 0.0f || freq > sample_rate / 8.0f);
        if (!pskip) {
            const float basefrq = 4294967296.0f / sample_rate;
            ifreq = (uint32_t)(basefrq * freq);
            const float gdf = gain / utof23(ifreq);
            gdf2 = gdf * 2.0f;
            cc121 = gdf2 * (col - 1.0f) + gain;
            cc212 = gdf2 * col - gain;
            pstate = ((uint32_t)(cnt - ifreq) < brpt) ? 3u : 0u;
        }
        flt.set_type(ZCU_FILTER_LOW_PASS);             // modules.zig:230
        flt.cut = zclamp(cutoff, 0.0f, 1.0f);          // Filter.zig:114
        flt.res = 1.0f - zclamp(0.7f, 0.0f, 1.0f);     // modules.zig:235, Filter.zig:118
        zcu_envelope_params ep;                        // modules.zig:238-245
        ep.sample_rate = sample_rate;
        ep.attack.kind = ZCU_CURVE_CUBED;  ep.attack.duration = 0.01f;
        ep.decay.kind = ZCU_CURVE_CUBED;   ep.decay.duration = 0.1f;
        ep.release.kind = ZCU_CURVE_CUBED; ep.release.duration = 0.5f;
        ep.sustain_volume = 0.8f;
        ep.note_on = note_on ? 1 : 0;
        env.begin(ep, note_id_changed);
    }
    __device__ __forceinline__ float sample() {
        const float gain = 0.7f;
        // zero(temps[0]); osc.paint -> temps[0]
        float t0 = 0.0f;
        if (!pskip) {
            pstate = ((pstate << 1) | (cnt < brpt ? 1u : 0u)) & 3u;
            const uint32_t transition = pstate | ((cnt < ifreq ? 1u : 0u) << 2);
            const float p = utof23(cnt);
            float v;
            switch (transition) {
                case 3: v = gain; break;
                case 0: v = -gain; break;
                case 2: v = gdf2 * (col - p) + gain; break;
                case 5: v = gdf2 * p - gain; break;
                case 7: v = cc121; break;
                case 4: v = cc212; break;
                default: v = 0.0f; break;
            }
            t0 += v;
            cnt += ifreq;
        }
        t0 *= 0.5f;                        // multiplyWithScalar(temps[0], 0.5)
        float t1 = 0.0f;                   // zero(temps[1])
        t1 += flt.step(t0);                // flt.paint -> temps[1]
        float e = 0.0f;                    // zero(temps[0])
        env.sample(e);                     // env.paint -> temps[0]
        return e * t1;                     // multiply(outputs[0], temps[0], temps[1]) adds this product
    }
};

template <int C, bool VOUT>
__global__ void __launch_bounds__(32) k_render_nice(RenderArgs a) {
    __shared__ float tile[32 * 33];
    __shared__ float g[2][32];
    const uint32_t lane = threadIdx.x;
    const uint32_t v = blockIdx.x * 32 + lane;
    const bool valid = v < a.V;

    NiceVoice nv;
    NiceVoiceState st;
    uint32_t ev_end = 0;
    if (valid) {
        st = a.state[v];
        ev_end = a.event_off[v + 1];
        const float color = a.color[v];
        nv.brpt = ftou32(clamp01(color));
        nv.col = utof23(nv.brpt);
        nv.cnt = st.cnt;
        nv.flt.l = st.l;
        nv.flt.b = st.b;
        nv.env.state = st.env_state;
        nv.env.painter.t = st.env_t;
        nv.env.painter.last_value = st.env_last;
        nv.env.painter.start = st.env_start;
        nv.env.mode = EnvelopeCore::M_NONE;
        for (int c = 0; c < C; c++) g[c][lane] = a.gains[(size_t)v * C + c];
    } else {
        st = NiceVoiceState{};
        for (int c = 0; c < C; c++) g[c][lane] = 0.0f;
    }
    __syncwarp();

    bool active = false;
    uint32_t boundary = valid ? a.F0 : 0xFFFFFFFFu;
    uint32_t buf_end = a.F0;

    for (uint32_t tf = a.F0; tf < a.F1; tf += 32) {
        const uint32_t n = min(32u, a.F1 - tf);
#pragma unroll 1
        for (uint32_t j = 0; j < n; j++) {
            const uint32_t f = tf + j;
            if (f == boundary) {
                // ---- Trigger replay (trigger.zig:80-196) at a span boundary ----
                active = false;
                if (f == buf_end) buf_end = min(f + a.buffer_frames, a.total_frames);  // next mix buffer
                bool got = false;
                uint32_t cand = 0;
                while (st.next_event < ev_end && a.events[st.next_event].frame <= f) {  // same frame: the last one wins
                    cand = st.next_event++;
                    got = true;
                }
                bool changed = false;
                if (got) {
                    const uint32_t id = a.events[cand].note_id;
                    changed = st.has_note ? (id != st.note_id) : true;
                    st.has_note = 1;
                    st.note_id = id;
                    st.cur_event = cand;
                }
                if (st.has_note) {
                    const zcu_note_event ev = a.events[st.cur_event];
                    nv.begin_span(a.sample_rate, ev.freq, a.cutoff[st.cur_event], ev.note_on != 0, changed);
                    active = true;
                }
                uint32_t nb = buf_end;
                if (st.next_event < ev_end) nb = min(nb, a.events[st.next_event].frame);
                boundary = nb;
            }
            float o = 0.0f;
            if (active) o = nv.sample();
            tile[j * 33 + lane] = o;
            if (VOUT) {
                if (valid && f >= a.vout_first && f < a.vout_first + a.vout_frames) a.vout[(size_t)(f - a.vout_first) * a.V + v] = o;
            }
        }
        if (!VOUT) {
            __syncwarp();
            if (lane < n) {
                float acc[C];
#pragma unroll
                for (int c = 0; c < C; c++) acc[c] = 0.0f;
#pragma unroll 8
                for (int vv = 0; vv < 32; vv++) {
                    const float x = tile[lane * 33 + vv];
#pragma unroll
                    for (int c = 0; c < C; c++) acc[c] = fmaf(g[c][vv], x, acc[c]);  // mix order is not pinned by the reference
                }
                float* dst = a.partial + ((size_t)blockIdx.x * (a.F1 - a.F0) + (tf - a.F0 + lane)) * C;
                if (C == 2)
                    *reinterpret_cast<float2*>(dst) = make_float2(acc[0], acc[1]);
                else
                    dst[0] = acc[0];
            }
            __syncwarp();
        }
    }
    if (valid) {
        st.cnt = nv.cnt;
        st.l = nv.flt.l;
        st.b = nv.flt.b;
        st.env_state = nv.env.state;
        st.env_t = nv.env.painter.t;
        st.env_last = nv.env.painter.last_value;
        st.env_start = nv.env.painter.start;
        a.state[v] = st;
    }
}

// mix[F0+f][c] += sum over warps (fixed order) of partial[w][f][c]
__global__ void __launch_bounds__(256) k_partial_reduce(const float* __restrict__ partial, float* __restrict__ mix, uint32_t nwarps,
                                                        size_t n /* block_frames*C */) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (uint32_t w = 0; w < nwarps; w++) s += partial[(size_t)w * n + i];
    mix[i] += s;
}

__global__ void k_state_reset(NiceVoiceState* st, const uint32_t* event_off, uint32_t V) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    NiceVoiceState s{};
    s.env_state = ZCU_ENV_IDLE;
    s.next_event = event_off[v];
    st[v] = s;
}

int render_blocks(zcu_ctx* ctx, zcu_render* r, float* mix, float* vout, uint32_t vout_first, uint32_t vout_frames, uint32_t end_frame) {
    const zcu_render_config& c = r->cfg;
    k_state_reset<<<(c.voices + 127) / 128, 128, 0, ctx->stream>>>(r->d_state, r->d_event_off, c.voices);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    for (uint32_t F0 = 0; F0 < end_frame; F0 += r->block_frames) {
        RenderArgs a;
        a.V = c.voices;
        a.C = c.channels;
        a.buffer_frames = c.buffer_frames;
        a.total_frames = c.total_frames;
        a.F0 = F0;
        a.F1 = std::min(F0 + r->block_frames, end_frame);
        a.sample_rate = c.sample_rate;
        a.color = r->d_color;
        a.gains = r->d_gains;
        a.state = r->d_state;
        a.event_off = r->d_event_off;
        a.events = r->d_events;
        a.cutoff = r->d_cutoff;
        a.partial = r->d_partial;
        a.vout = vout;
        a.vout_first = vout_first;
        a.vout_frames = vout_frames;
        if (vout) {
            if (c.channels == 2)
                k_render_nice<2, true><<<r->nwarps, 32, 0, ctx->stream>>>(a);
            else
                k_render_nice<1, true><<<r->nwarps, 32, 0, ctx->stream>>>(a);
            ctx->launches++;
        } else {
            if (r->profiling) {
                if (r->ev_used + 2 > r->ev.size()) {
                    r->ev.resize(r->ev_used + 2, nullptr);
                    ZCU_CUDA(ctx, cudaEventCreate(&r->ev[r->ev_used]));
                    ZCU_CUDA(ctx, cudaEventCreate(&r->ev[r->ev_used + 1]));
                }
                ZCU_CUDA(ctx, cudaEventRecord(r->ev[r->ev_used], ctx->stream));
            }
            if (c.channels == 2)
                k_render_nice<2, false><<<r->nwarps, 32, 0, ctx->stream>>>(a);
            else
                k_render_nice<1, false><<<r->nwarps, 32, 0, ctx->stream>>>(a);
            if (r->profiling) {
                ZCU_CUDA(ctx, cudaEventRecord(r->ev[r->ev_used + 1], ctx->stream));
                r->ev_used += 2;
            }
            const size_t n = (size_t)(a.F1 - a.F0) * c.channels;
            k_partial_reduce<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(r->d_partial, mix + (size_t)F0 * c.channels, r->nwarps, n);
            ctx->launches += 2;
        }
        ZCU_CUDA(ctx, cudaGetLastError());
    }
    return ZCU_OK;
}

}  // namespace

extern "C" {

int zcu_render_create(zcu_ctx* ctx, const zcu_render_config* cfg, const float* voice_color_host, const float* gains_host, zcu_render** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, cfg && voice_color_host && gains_host, "null argument");
    ZCU_REQUIRE(ctx, cfg->instrument == ZCU_INSTR_NICE, "unknown instrument");
    ZCU_REQUIRE(ctx, cfg->channels == 1 || cfg->channels == 2, "channels must be 1 or 2");
    ZCU_REQUIRE(ctx, cfg->voices > 0 && cfg->buffer_frames > 0 && cfg->total_frames > 0, "empty render");
    ZCU_REQUIRE(ctx, cfg->mode == ZCU_RENDER_EXACT, "only ZCU_RENDER_EXACT is implemented");
    zcu_render* r = new (std::nothrow) zcu_render();
    if (!r) return ZCU_ERR_OOM;
    r->d_color = r->d_gains = r->d_cutoff = r->d_partial = nullptr;
    r->d_state = nullptr;
    r->d_event_off = nullptr;
    r->d_events = nullptr;
    r->nevents = 0;
    r->profiling = false;
    r->ev_used = 0;
    r->ctx = ctx;
    r->cfg = *cfg;
    const uint32_t bb = cfg->block_buffers ? cfg->block_buffers : 64;
    r->block_frames = bb * cfg->buffer_frames;
    r->nwarps = (cfg->voices + 31) / 32;
    const uint32_t V = cfg->voices, C = cfg->channels;
    cudaError_t e = cudaMalloc(&r->d_color, (size_t)V * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_gains, (size_t)V * C * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_state, (size_t)V * sizeof(NiceVoiceState));
    if (e == cudaSuccess) e = cudaMalloc(&r->d_event_off, (size_t)(V + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_partial, (size_t)r->nwarps * r->block_frames * C * 4);
    if (e != cudaSuccess) {
        zcu_render_free(r);
        return zcu_fail(ctx, ZCU_ERR_OOM, "render allocation failed: %s", cudaGetErrorString(e));
    }
    ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_color, voice_color_host, (size_t)V * 4, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_gains, gains_host, (size_t)V * C * 4, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaMemsetAsync(r->d_event_off, 0, (size_t)(V + 1) * 4, ctx->stream));
    *out = r;
    return ZCU_OK;
}

void zcu_render_free(zcu_render* r) {
    if (!r) return;
    if (r->ctx) cudaStreamSynchronize(r->ctx->stream);
    cudaFree(r->d_color);
    cudaFree(r->d_gains);
    cudaFree(r->d_state);
    cudaFree(r->d_event_off);
    cudaFree(r->d_events);
    cudaFree(r->d_cutoff);
    cudaFree(r->d_partial);
    for (cudaEvent_t e : r->ev)
        if (e) cudaEventDestroy(e);
    delete r;
}

int zcu_render_set_events(zcu_ctx* ctx, zcu_render* r, const uint32_t* event_off_host, const zcu_note_event* events_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && event_off_host, "null argument");
    const uint32_t V = r->cfg.voices;
    const uint32_t n = event_off_host[V];
    ZCU_REQUIRE(ctx, n == 0 || events_host, "null events");
    if (n > r->nevents || !r->d_events) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(r->d_events);
        cudaFree(r->d_cutoff);
        r->d_events = nullptr;
        r->d_cutoff = nullptr;
        ZCU_CUDA(ctx, cudaMalloc(&r->d_events, (size_t)(n ? n : 1) * sizeof(zcu_note_event)));
        ZCU_CUDA(ctx, cudaMalloc(&r->d_cutoff, (size_t)(n ? n : 1) * sizeof(float)));
    }
    r->nevents = n;
    // per-call host scalar of modules.zig:231-234 (cos, sqrt): once per event, on the host
    std::vector<float> cutoff(n);
    for (uint32_t i = 0; i < n; i++) cutoff[i] = zcu_filter_cutoff_from_frequency(events_host[i].freq * 8.0f, r->cfg.sample_rate);
    ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_event_off, event_off_host, (size_t)(V + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (n) {
        ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_events, events_host, (size_t)n * sizeof(zcu_note_event), cudaMemcpyHostToDevice, ctx->stream));
        ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_cutoff, cutoff.data(), (size_t)n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    }
    return ZCU_OK;
}

int zcu_render_run(zcu_ctx* ctx, zcu_render* r, zcu_buf* mix) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && mix, "null argument");
    ZCU_REQUIRE(ctx, r->d_events != nullptr, "zcu_render_set_events has not been called");
    ZCU_REQUIRE(ctx, mix->frames == r->cfg.total_frames && mix->voices == r->cfg.channels, "mix must be [total_frames][channels]");
    return render_blocks(ctx, r, mix->d, nullptr, 0, 0, r->cfg.total_frames);
}

int zcu_render_run_voices(zcu_ctx* ctx, zcu_render* r, uint32_t first_frame, zcu_buf* out_voices) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && out_voices, "null argument");
    ZCU_REQUIRE(ctx, r->d_events != nullptr, "zcu_render_set_events has not been called");
    ZCU_REQUIRE(ctx, out_voices->voices == r->cfg.voices, "out_voices must be [frames][voices]");
    ZCU_REQUIRE(ctx, (uint64_t)first_frame + out_voices->frames <= r->cfg.total_frames, "window beyond the render");
    return render_blocks(ctx, r, nullptr, out_voices->d, first_frame, out_voices->frames, first_frame + out_voices->frames);
}

int zcu_render_set_profiling(zcu_render* r, int enabled) {
    if (!r) return ZCU_ERR_INVALID;
    r->profiling = enabled != 0;
    r->ev_used = 0;
    return ZCU_OK;
}

int zcu_render_kernel_ms(zcu_ctx* ctx, zcu_render* r, float* total_ms, uint32_t* launches) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && total_ms && launches, "null argument");
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double sum = 0.0;
    for (size_t i = 0; i + 1 < r->ev_used; i += 2) {
        float ms = 0.0f;
        ZCU_CUDA(ctx, cudaEventElapsedTime(&ms, r->ev[i], r->ev[i + 1]));
        sum += ms;
    }
    *total_ms = (float)sum;
    *launches = (uint32_t)(r->ev_used / 2);
    r->ev_used = 0;
    return ZCU_OK;
}

uint64_t zcu_render_launches_per_run(const zcu_render* r) {
    if (!r) return 0;
    const uint64_t nblocks = (r->cfg.total_frames + r->block_frames - 1) / r->block_frames;
    return 1 + 2 * nblocks;
}

}  // extern "C"
