// zcu_render_scan.cuh — ZCU_RENDER_SCAN: the time-parallel form of the fused NiceInstrument bank.
//
// The exact kernel (zcu_render_kernel.cuh) walks each voice's time axis sequentially, so a 4096-voice bank
// offers only 128 warps of work and the Filter's dependent f32 chain sets the pace.  Here the time axis is cut
// at the mix-buffer boundaries (chunk = one buffer of the reference's paint loop, examples/example_polyphony.zig:
// 61-92) and every (32 voices, chunk) pair is an independent warp task.  What each module needs to start in the
// middle of the render:
//
//   Trigger      the current note at a frame is the last event before it (trigger.zig:107-137 carry-over); a
//                per-(chunk, voice) cursor table is written once per job by k_nice_prepass.
//   PulseOsc     cnt is an integer accumulator (PulseOsc.zig:98-112): cnt(F) = cnt(event) + (F - frame)*ifreq,
//                exact mod 2^32.  cnt at every event is written by the prepass.
//   Envelope     the ADSR machine only changes at events and at ramp ends, and Painter's f32 ramp clock has an
//                exact closed form (zcu_exact_accum.cuh), so the state at any frame is the state at the last
//                event advanced by (F - frame) samples in O(phases x binades).  BIT-EXACT.
//   Filter       a linear recurrence in (l, b) with per-span constant coefficients (Filter.zig:135-144):
//                state_end = M * state_start + z, where z is the chunk's zero-state response (computed in strict
//                f32 by k_nice_scan_zs) and M = prod A(cut,res)^len over the chunk's spans (2x2, f64, by repeated
//                squaring).  k_nice_scan_prefix chains the chunks; k_nice_scan_render then re-runs every chunk in
//                strict f32 from its start state.  This is the only approximation: the start state of a chunk
//                differs from the sequential f32 trajectory by rounding noise (~1e-7), which the stable filter
//                damps.  Output tolerance 1e-5 of full scale (BASELINE.json north_star); measured ~2e-7.
#pragma once

#include "zcu_exact_accum.cuh"
#include "zcu_render_kernel.cuh"

namespace zcu {

enum { EV_SUPERSEDED = 1u, EV_CHANGED = 2u };

struct __align__(16) EnvSnap {  // Envelope state just before the event's span begins
    int32_t state;
    float t, last_value, start;
};

struct __align__(16) Mat2 {  // row major 2x2, f64
    double a, b, c, d;
};

struct ScanArgs {
    uint32_t V, C;
    uint32_t L;               // chunk length = buffer_frames
    uint32_t total_frames;
    uint32_t F0, F1;          // frames of this launch (F0 is chunk aligned)
    uint32_t nchunks;         // chunks of this launch
    uint32_t nvgroups;        // ceil(V / 32)
    float sample_rate;
    const float* color;
    const float* gains;
    const uint32_t* event_off;
    const DevEvent* events;
    const EnvSnap* snaps;
    const uint32_t* cursor;   // [total chunks][V]: first event with frame >= chunk start
    float2* zs;               // [nchunks][V] zero-state response of the chunk
    Mat2* M;                  // [nchunks][V] homogeneous map of the chunk
    float2* sstart;           // [nchunks][V] filter state at the chunk start
    double2* fstate;          // [V] filter state carried across launches
    float* partial;           // [nvgroups][F1-F0][C]
    float* vout;              // [vout_frames][vout_nv] window of the un-mixed voice outputs, or null
    uint32_t vout_first, vout_frames;
    uint32_t vout_v0, vout_nv;  // voices [vout_v0, vout_v0 + vout_nv) of the window
};

// NiceInstrument's envelope params (modules.zig:238-245)
__device__ __forceinline__ zcu_envelope_params nice_env_params(float sample_rate, bool note_on) {
    zcu_envelope_params ep;
    ep.sample_rate = sample_rate;
    ep.attack.kind = ZCU_CURVE_CUBED;
    ep.attack.duration = 0.01f;
    ep.decay.kind = ZCU_CURVE_CUBED;
    ep.decay.duration = 0.1f;
    ep.release.kind = ZCU_CURVE_CUBED;
    ep.release.duration = 0.5f;
    ep.sustain_volume = 0.8f;
    ep.note_on = note_on ? 1 : 0;
    return ep;
}

__device__ __forceinline__ uint32_t pulse_ifreq(float sample_rate, float freq) {  // PulseOsc.zig:82-89
    if (freq < 0.0f || freq > sample_rate / 8.0f) return 0u;
    const float basefrq = 4294967296.0f / sample_rate;
    return (uint32_t)(basefrq * freq);
}

__device__ __forceinline__ void env_reset(EnvelopeCore& e, float sample_rate) {
    e.state = ZCU_ENV_IDLE;
    e.painter.t = 0.0f;
    e.painter.last_value = 0.0f;
    e.painter.start = 0.0f;
    e.painter.t_step = 0.0f;
    e.painter.gms = 0.0f;
    e.painter.kind = ZCU_CURVE_LINEAR;
    e.mode = EnvelopeCore::M_NONE;
    e.flat = 0.0f;
    e.sample_rate = sample_rate;
    e.sustain_volume = 0.0f;
}

// ---- once per job: per-event PulseOsc phase, Envelope state and Trigger flags; per-chunk event cursors ----
struct PrepassArgs {
    uint32_t V, L, nchunks_total;
    float sample_rate;
    const uint32_t* event_off;
    DevEvent* events;
    EnvSnap* snaps;
    uint32_t* cursor;
};

// cursor[c][v] = index of the voice's first event at or after the start of mix buffer c: one warp per voice, one lane per event
__global__ void __launch_bounds__(128) k_nice_cursor(PrepassArgs a) {
    const uint32_t v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (v >= a.V) return;
    const uint32_t k0 = a.event_off[v], k1 = a.event_off[v + 1];
    for (uint32_t k = k0 + lane; k <= k1; k += 32) {  // k == k1: the buffers after the last event
        const uint32_t c_lo = k == k0 ? 0u : a.events[k - 1].frame / a.L + 1u;
        const uint32_t c_hi = k == k1 ? a.nchunks_total : min(a.nchunks_total, a.events[k].frame / a.L + 1u);
        for (uint32_t c = c_lo; c < c_hi; c++) a.cursor[(size_t)c * a.V + v] = k;
    }
}

// The envelope state at every event comes from the ramp tables of zcu_modules.cuh (RampTab): the walk over ~30 binades per
// event and the cursor loop were 3500 warp instructions per event on a one-warp-per-SM critical path (ncu, 2.4 ms per job).
__global__ void __launch_bounds__(64) k_nice_prepass(PrepassArgs a) {
    __shared__ RampTab tabs[3];
    if (threadIdx.x < 3) {
        const zcu_envelope_params ep = nice_env_params(a.sample_rate, true);
        const float dur = threadIdx.x == 0 ? ep.attack.duration : threadIdx.x == 1 ? ep.decay.duration : ep.release.duration;
        tabs[threadIdx.x].build(1.0f / (dur * a.sample_rate));  // PainterDev::try_ramp's t_step
    }
    __syncthreads();
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= a.V) return;
    const uint32_t k0 = a.event_off[v], k1 = a.event_off[v + 1];
    EnvelopeCore e;
    env_reset(e, a.sample_rate);
    uint32_t cnt = 0, ifreq = 0, note_id = 0, prev_frame = 0;
    bool has_note = false;
    uint4 raw = k0 < k1 ? *reinterpret_cast<const uint4*>(a.events + k0) : make_uint4(0, 0, 0, 0);
    for (uint32_t k = k0; k < k1; k++) {
        const uint32_t frame = raw.x, id = raw.y;
        const float freq = __uint_as_float(raw.z);
        const bool on = raw.w != 0;
        const uint4 nxt = k + 1 < k1 ? *reinterpret_cast<const uint4*>(a.events + k + 1) : make_uint4(0, 0, 0, 0);
        const uint32_t n = frame - prev_frame;
        if (has_note) {
            cnt += n * ifreq;
            env_advance_tab(e, n, tabs, 3);
        }
        const bool superseded = (k + 1 < k1) && nxt.x == frame;  // trigger.zig:173-178
        const bool changed = has_note ? (id != note_id) : true;  // trigger.zig:96-100
        a.events[k].cnt0 = cnt;
        a.events[k].flags = (superseded ? EV_SUPERSEDED : 0u) | (changed ? EV_CHANGED : 0u);
        EnvSnap s;
        s.state = e.state;
        s.t = e.painter.t;
        s.last_value = e.painter.last_value;
        s.start = e.painter.start;
        a.snaps[k] = s;
        if (!superseded) {
            has_note = true;
            note_id = id;
            ifreq = pulse_ifreq(a.sample_rate, freq);
            e.begin(nice_env_params(a.sample_rate, on), changed);
        }
        prev_frame = frame;
        raw = nxt;
    }
}

// ---- per (voice, chunk): where the Trigger stands at the chunk start ---------------------------------------
struct ChunkCursor {
    uint32_t next_event, ev_end;
    bool active;
    DevEvent cur;   // the current note's event (valid when active)
    uint32_t boundary;  // frame of events[next_event], UINT_MAX past the end

    __device__ __forceinline__ static DevEvent fetch(const DevEvent* ev, uint32_t i) {
        const uint4 x = __ldg(reinterpret_cast<const uint4*>(ev + i));
        const uint4 y = __ldg(reinterpret_cast<const uint4*>(ev + i) + 1);
        DevEvent e;
        e.frame = x.x;
        e.note_id = x.y;
        e.freq = __uint_as_float(x.z);
        e.note_on = (int32_t)x.w;
        e.cutoff = __uint_as_float(y.x);
        e.cnt0 = y.y;
        e.flags = y.z;
        e.pad = 0;
        return e;
    }
    __device__ void init(const ScanArgs& a, uint32_t v, bool valid, uint32_t chunk_global) {
        active = false;
        boundary = 0xFFFFFFFFu;
        next_event = ev_end = 0;
        cur.frame = 0;
        cur.note_id = 0;
        cur.freq = 0.0f;
        cur.note_on = 0;
        cur.cutoff = 0.0f;
        cur.cnt0 = 0;
        cur.flags = 0;
        cur.pad = 0;
        if (!valid) return;
        const uint32_t k0 = a.event_off[v];
        ev_end = a.event_off[v + 1];
        next_event = a.cursor[(size_t)chunk_global * a.V + v];
        if (next_event > k0) {
            cur = fetch(a.events, next_event - 1);
            active = true;
        }
        if (next_event < ev_end) boundary = __ldg(&a.events[next_event].frame);
    }
    // at f == boundary: consume the impulses of this frame (the last one wins).  Returns note_id_changed.
    __device__ bool step(const ScanArgs& a, uint32_t f) {
        bool changed = false;
        while (boundary == f) {
            cur = fetch(a.events, next_event);
            changed = (cur.flags & EV_CHANGED) != 0;
            next_event++;
            boundary = next_event < ev_end ? __ldg(&a.events[next_event].frame) : 0xFFFFFFFFu;
        }
        active = true;
        return changed;
    }
};

// homogeneous part of one Filter sample (Filter.zig:135-144 with in = 0 and without the DC bias), f64
__device__ inline Mat2 filter_hom(float cutf, float resf) {
    const double c = (double)cutf, r = (double)resf;
    Mat2 A;
    double l = 1.0, b = 0.0;
#pragma unroll
    for (int col = 0; col < 2; col++) {
        l = col == 0 ? 1.0 : 0.0;
        b = col == 0 ? 0.0 : 1.0;
        l = l + c * b;
        b = b + c * (0.0 - b * r - l);
        l = l + c * b;
        const double h = 0.0 - b * r - l;
        b = b + c * h;
        if (col == 0) {
            A.a = l;
            A.c = b;
        } else {
            A.b = l;
            A.d = b;
        }
    }
    return A;
}
__device__ __forceinline__ Mat2 mat2_mul(const Mat2& x, const Mat2& y) {
    Mat2 r;
    r.a = x.a * y.a + x.b * y.c;
    r.b = x.a * y.b + x.b * y.d;
    r.c = x.c * y.a + x.d * y.c;
    r.d = x.c * y.b + x.d * y.d;
    return r;
}
// M = A^n * M
__device__ inline void mat2_pow_apply(Mat2 A, uint32_t n, Mat2& M) {
    while (n) {
        if (n & 1u) M = mat2_mul(A, M);
        n >>= 1;
        if (n) A = mat2_mul(A, A);
    }
}

// warp task -> (voice group, chunk): consecutive warps take consecutive voice groups of the same chunk
__device__ __forceinline__ bool scan_task(const ScanArgs& a, uint32_t& vgroup, uint32_t& chunk) {
    const uint32_t task = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (task >= a.nvgroups * a.nchunks) return false;
    vgroup = task % a.nvgroups;
    chunk = task / a.nvgroups;
    return true;
}

// ---- pass 1: zero-state response z and homogeneous map M of every (voice, chunk) ---------------------------
__global__ void __launch_bounds__(128) k_nice_scan_zs(ScanArgs a) {
    uint32_t vgroup, chunk;
    if (!scan_task(a, vgroup, chunk)) return;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t v = vgroup * 32 + lane;
    const bool valid = v < a.V;
    const uint32_t FULL = 0xffffffffu;
    const uint32_t F = a.F0 + chunk * a.L;
    const uint32_t n = min(a.L, a.F1 - F);

    ChunkCursor cur;
    cur.init(a, v, valid, F / a.L);
    PulseStage ps;
    FilterStage fs;
    const float color = valid ? a.color[v] : 0.0f;
    ps.init(color, 0u);
    fs.l = fs.b = 0.0f;
    fs.cut = fs.res = 0.0f;
    Mat2 M{1.0, 0.0, 0.0, 1.0}, A{1.0, 0.0, 0.0, 1.0};
    uint32_t seg = 0;
    if (cur.active) {
        ps.begin_span(a.sample_rate, cur.cur.freq);
        ps.set_cnt(cur.cur.cnt0 + (F - cur.cur.frame) * ps.ifreq);
        fs.begin_span(cur.cur.cutoff);
        A = filter_hom(fs.cut, fs.res);
    }
    uint32_t j = 0;
    while (j < n) {
        const uint32_t f = F + j;
        if (f == cur.boundary) {
            if (seg) mat2_pow_apply(A, seg, M);
            seg = 0;
            cur.step(a, f);
            ps.begin_span(a.sample_rate, cur.cur.freq);
            fs.begin_span(cur.cur.cutoff);
            A = filter_hom(fs.cut, fs.res);
        }
        const uint32_t run = min(n - j, __reduce_min_sync(FULL, cur.boundary - f));
        if (__all_sync(FULL, cur.active)) {
#pragma unroll 4
            for (uint32_t r = 0; r < run; r++) fs.sample(ps.sample());
        } else if (cur.active) {
            for (uint32_t r = 0; r < run; r++) fs.sample(ps.sample());
        }
        if (cur.active) seg += run;
        j += run;
    }
    if (seg) mat2_pow_apply(A, seg, M);
    if (valid) {
        const size_t o = (size_t)chunk * a.V + v;
        a.zs[o] = make_float2(fs.l, fs.b);
        a.M[o] = M;
    }
}

// ---- pass 2: chain the chunks (per voice, f64) -------------------------------------------------------------
__global__ void __launch_bounds__(128) k_nice_scan_prefix(ScanArgs a) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= a.V) return;
    double2 s = a.fstate[v];
    constexpr int U = 8;
    for (uint32_t c0 = 0; c0 < a.nchunks; c0 += U) {
        Mat2 m[U];
        float2 z[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (c0 + u < a.nchunks) {
                const size_t o = (size_t)(c0 + u) * a.V + v;
                m[u] = a.M[o];
                z[u] = a.zs[o];
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (c0 + u < a.nchunks) {
                a.sstart[(size_t)(c0 + u) * a.V + v] = make_float2((float)s.x, (float)s.y);
                const double l = m[u].a * s.x + m[u].b * s.y + (double)z[u].x;
                const double b = m[u].c * s.x + m[u].d * s.y + (double)z[u].y;
                s.x = l;
                s.y = b;
            }
        }
    }
    a.fstate[v] = s;
}

// ---- pass 3: the render proper, every chunk from its own start state ---------------------------------------
template <int C, bool VOUT, int RT>
__global__ void __launch_bounds__(128, 4) k_nice_scan_render(ScanArgs a) {
    extern __shared__ float smem[];
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    float(*O)[33] = reinterpret_cast<float(*)[33]>(smem + (size_t)warp * RT * 33);
    float(*g)[32] = reinterpret_cast<float(*)[32]>(smem + (size_t)4 * RT * 33 + (size_t)warp * C * 32);
    uint32_t vgroup, chunk;
    if (!scan_task(a, vgroup, chunk)) return;
    const uint32_t v = vgroup * 32 + lane;
    const bool valid = v < a.V;
    const uint32_t FULL = 0xffffffffu;
    const uint32_t F = a.F0 + chunk * a.L;
    const uint32_t n = min(a.L, a.F1 - F);

    ChunkCursor cur;
    cur.init(a, v, valid, F / a.L);
    PulseStage ps;
    FilterStage fs;
    EnvStage es;
    const float color = valid ? a.color[v] : 0.0f;
    ps.init(color, 0u);
    fs.l = fs.b = 0.0f;
    fs.cut = fs.res = 0.0f;
    if (valid) {
        const float2 s = a.sstart[(size_t)chunk * a.V + v];
        fs.l = s.x;
        fs.b = s.y;
    }
    env_reset(es.e, a.sample_rate);
    if (cur.active) {
        ps.begin_span(a.sample_rate, cur.cur.freq);
        ps.set_cnt(cur.cur.cnt0 + (F - cur.cur.frame) * ps.ifreq);
        fs.begin_span(cur.cur.cutoff);
        const EnvSnap s = a.snaps[cur.next_event - 1];
        es.e.state = s.state;
        es.e.painter.t = s.t;
        es.e.painter.last_value = s.last_value;
        es.e.painter.start = s.start;
        es.begin_span(a.sample_rate, cur.cur.note_on != 0, (cur.cur.flags & EV_CHANGED) != 0);
        env_advance(es.e, F - cur.cur.frame);
    }
    if (!VOUT) {
#pragma unroll
        for (int c = 0; c < C; c++) g[c][lane] = valid ? a.gains[(size_t)v * C + c] : 0.0f;
    }
    __syncwarp();

    const uint32_t nframes = a.F1 - a.F0;
    for (uint32_t t0 = 0; t0 < n; t0 += RT) {
        const uint32_t tn = min((uint32_t)RT, n - t0);
        uint32_t j = 0;
        while (j < tn) {
            const uint32_t f = F + t0 + j;
            if (f == cur.boundary) {
                const bool changed = cur.step(a, f);
                ps.begin_span(a.sample_rate, cur.cur.freq);
                fs.begin_span(cur.cur.cutoff);
                es.begin_span(a.sample_rate, cur.cur.note_on != 0, changed);
            }
            uint32_t dist = cur.boundary - f;
            bool careful = false;
            if (cur.active) {
                const uint32_t s = es.safe_run();
                careful = (s == 0u);
                dist = min(dist, careful ? 1u : s);
            }
            const uint32_t run = min(tn - j, __reduce_min_sync(FULL, dist));
            const bool fast = __all_sync(FULL, cur.active && !careful);
            float* oo = &O[j][lane];
            if (fast) {
                const EnvStage::Run er = es.begin_run();
                const bool cubed = __all_sync(FULL, !er.ramp || (!er.lin && !er.sq));
                float lv = 0.0f;
                if (cubed) {
#pragma unroll 4
                    for (uint32_t r = 0; r < run; r++) {
                        const float y = fs.sample(ps.sample());
                        oo[r * 33] = es.sample_fast<true>(er, lv) * y;
                    }
                } else {
                    for (uint32_t r = 0; r < run; r++) {
                        const float y = fs.sample(ps.sample());
                        oo[r * 33] = es.sample_fast<false>(er, lv) * y;
                    }
                }
                es.end_run(er, lv, run > 0);
            } else {
                for (uint32_t r = 0; r < run; r++) {
                    float o = 0.0f;
                    if (cur.active) {
                        const float y = fs.sample(ps.sample());
                        o = es.sample_careful() * y;
                    }
                    oo[r * 33] = o;
                }
            }
            j += run;
        }
        __syncwarp();
        if (VOUT) {
            for (uint32_t jj = 0; jj < tn; jj++) {
                const uint32_t f = F + t0 + jj;
                if (valid && f >= a.vout_first && f < a.vout_first + a.vout_frames && v - a.vout_v0 < a.vout_nv)
                    a.vout[(size_t)(f - a.vout_first) * a.vout_nv + (v - a.vout_v0)] = 0.0f + O[jj][lane];  // multiply() accumulates into the zeroed output: -0 becomes +0
            }
        } else {
            // lane sums frames lane, lane+32, ... over the warp's 32 voices with their pan gains (the reference's
            // voice summation order is not reproduced; tolerance 1e-5 x sum|gain|)
            for (uint32_t fr = lane; fr < tn; fr += 32) {
                float acc[C];
#pragma unroll
                for (int c = 0; c < C; c++) acc[c] = 0.0f;
#pragma unroll 8
                for (int vv = 0; vv < 32; vv++) {
                    const float x = O[fr][vv];
#pragma unroll
                    for (int c = 0; c < C; c++) acc[c] = fmaf(g[c][vv], x, acc[c]);
                }
                float* dst = a.partial + ((size_t)vgroup * nframes + (F - a.F0 + t0 + fr)) * C;
                if (C == 2)
                    *reinterpret_cast<float2*>(dst) = make_float2(acc[0], acc[1]);
                else
                    dst[0] = acc[0];
            }
        }
        __syncwarp();
    }
}

template <int RT>
constexpr size_t scan_render_smem_bytes() {
    return (size_t)(4 * RT * 33 + 4 * 2 * 32) * sizeof(float);
}

}  // namespace zcu
