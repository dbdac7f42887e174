// zcu_render_seg.cuh — ZCU_RENDER_SEGMENTS: the fused NiceInstrument bank, time axis cut into a few long segments.
//
// ZCU_RENDER_SCAN pays for its parallelism with a second Filter pass over every sample (the zero-state pass).  This
// mode buys the same parallelism for ~2 % extra work: the time axis of a launch is cut into S segments of many mix
// buffers, S chosen so that (32-voice groups x S) fills the chip's resident warps exactly once, and every
// (voice group, segment) warp renders its segment sequentially in time, like the exact kernel does.  What a warp needs
// to start in the middle of the render:
//
//   Trigger / PulseOsc / Envelope   exact closed forms from the per-event tables of k_nice_prepass (zcu_render_scan.cuh)
//   Filter                          a warm-up: the warp starts W mix buffers early from (l, b) = 0 and runs PulseOsc ->
//                                   Filter only.  The error of the state at the segment start is exactly
//                                   M_window * (true state at the window start), where M_window = prod A(cut,res)^len is
//                                   the homogeneous map of the window (2x2, f64, closed form).  W is doubled per voice
//                                   until max|M_window| <= 1e-8 (or the window reaches a frame where the true state is
//                                   known: the start of the launch, or a time before the voice's first note), so the
//                                   bound holds for every cutoff/resonance, not just for well damped ones.
//
// Segment 0 of every launch starts from the carried filter state, so with S = 1 (enough voices to fill the chip)
// this mode is the exact sequential render.  Otherwise: PulseOsc and Envelope bit-exact, Filter within ~1e-7.
#pragma once

#include "zcu_render_scan.cuh"

namespace zcu {

struct SegArgs {
    ScanArgs s;            // V, C, L, total_frames, F0, F1 (launch range, F0 chunk aligned), tables, partial, vout...
    uint32_t nseg;         // segments of this launch
    uint32_t seg_chunks;   // chunks per segment
    float2* carry;         // [V] filter state at F0 (read by segment 0), at F1 (written by the last segment)
    uint32_t* warm_stats;  // optional [2]: sum and max of warm-up chunks over all warp tasks (diagnostics)
    uint32_t need_carry;   // VOUT: a later launch follows, so the last segment must run even outside the window
};

// shared-memory tile of a warp: RT frames x 32 voices, row stride 34 floats (rows 8-byte aligned, conflict-free both for the
// per-voice column writes and for the mix's 8-byte row reads)
constexpr int kSegTileStride = 34;
template <int RT>
constexpr size_t seg_render_smem_bytes() {
    return (size_t)(4 * RT * kSegTileStride + 4 * 2 * 32) * sizeof(float);
}

// Blackwell packed f32 pairs (PTX ISA 8.6, sm_100): one issue slot for two IEEE fused multiply-adds
__device__ __forceinline__ unsigned long long fma_f32x2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ float sum_f32x2(unsigned long long v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}

// M = M * A^n  (A^n acts BEFORE M in time)
__device__ inline void mat2_apply_pow_right(Mat2& M, Mat2 A, uint32_t n) {
    while (n) {
        if (n & 1u) M = mat2_mul(M, A);
        n >>= 1;
        if (n) A = mat2_mul(A, A);
    }
}

// homogeneous map of the Filter over the frames [Fw, F) of voice v; *known = the true state at Fw is exactly zero
// (no note has been triggered before Fw, so the Filter has never been painted)
__device__ inline Mat2 window_map(const ScanArgs& a, uint32_t v, uint32_t Fw, uint32_t F, bool* known) {
    Mat2 M{1.0, 0.0, 0.0, 1.0};
    const uint32_t k0 = a.event_off[v], k1 = a.event_off[v + 1];
    uint32_t k = a.cursor[(size_t)(Fw / a.L) * a.V + v];
    *known = (k == k0);
    bool have = k > k0;
    float cutoff = have ? a.events[k - 1].cutoff : 0.0f;
    uint32_t pos = Fw;
    while (pos < F) {
        const uint32_t fk = k < k1 ? a.events[k].frame : 0xFFFFFFFFu;
        const uint32_t nextf = min(fk, F);
        if (have && nextf > pos) {
            FilterStage fs;
            fs.begin_span(cutoff);
            const Mat2 A = filter_hom(fs.cut, fs.res);
            // later spans act after earlier ones: M_total = A_later^n * ... * A_first^n
            Mat2 P{1.0, 0.0, 0.0, 1.0};
            mat2_pow_apply(A, nextf - pos, P);
            M = mat2_mul(P, M);
        }
        pos = nextf;
        if (pos < F) {  // an event at pos (a superseded one contributes a zero-length span)
            cutoff = a.events[k].cutoff;
            have = true;
            k++;
        }
    }
    return M;
}

// PulseOsc -> Filter only over the frames [f0, f1): the warm-up (no envelope, no output)
template <bool FF>
__device__ __forceinline__ void seg_warm_up(const ScanArgs& a, ChunkCursor& cur, PulseStage& ps, FilterStage& fs, uint32_t f0, uint32_t f1) {
    const uint32_t FULL = 0xffffffffu;
    uint32_t f = f0;
    while (f < f1) {
        if (f == cur.boundary) {
            cur.step(a, f);
            ps.begin_span(a.sample_rate, cur.cur.freq);
            fs.begin_span(cur.cur.cutoff);
        }
        const uint32_t run = min(f1 - f, __reduce_min_sync(FULL, cur.boundary - f));
        if (__all_sync(FULL, cur.active)) {
#pragma unroll 4
            for (uint32_t r = 0; r < run; r++) fs.sample_t<FF>(ps.sample());
        } else if (cur.active) {
            for (uint32_t r = 0; r < run; r++) fs.sample_t<FF>(ps.sample());
        }
        f += run;
    }
}

// FF: the Filter's multiply-add pairs contracted to FMAs (FilterStage::sample_t<true>; zcu_render_set_filter_contraction)
template <int C, bool VOUT, int RT, bool FF>
__global__ void __launch_bounds__(128, 6) k_nice_seg_render(SegArgs g, const float2* __restrict__ carry_in, float2* __restrict__ carry_out) {
    const ScanArgs& a = g.s;
    extern __shared__ float smem[];
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    constexpr int OS = kSegTileStride;
    float(*O)[OS] = reinterpret_cast<float(*)[OS]>(smem + (size_t)warp * RT * OS);
    float(*gn)[32] = reinterpret_cast<float(*)[32]>(smem + (size_t)4 * RT * OS + (size_t)warp * C * 32);
    const uint32_t task = blockIdx.x * 4 + warp;
    if (task >= a.nvgroups * g.nseg) return;
    const uint32_t vgroup = task % a.nvgroups;
    const uint32_t seg = task / a.nvgroups;
    const uint32_t v = vgroup * 32 + lane;
    const bool valid = v < a.V;
    const uint32_t FULL = 0xffffffffu;
    const uint32_t Fs = a.F0 + seg * g.seg_chunks * a.L;
    if (Fs >= a.F1) return;
    const uint32_t Fe = (uint32_t)min((uint64_t)a.F1, (uint64_t)Fs + (uint64_t)g.seg_chunks * a.L);
    if (VOUT) {
        // A voice window is rendered with the segmentation of the whole job (the arithmetic under test is that of
        // zcu_render_run); warp tasks that cannot touch the window are dropped.
        const bool frames_hit = Fe > a.vout_first && Fs < a.vout_first + a.vout_frames;
        const bool voices_hit = vgroup * 32 + 32 > a.vout_v0 && vgroup * 32 < a.vout_v0 + a.vout_nv;
        if (!(frames_hit && voices_hit) && !(g.need_carry && Fe == a.F1)) return;
    }

    // ---- how many mix buffers of warm-up does the Filter need? (per voice, then the warp's maximum)
    uint32_t wc = 0;
    if (seg > 0 && valid) {
        const uint32_t avail = (Fs - a.F0) / a.L;  // back to the launch start, where the state is the carried one
        wc = 1;
        for (;;) {
            if (wc >= avail) {
                wc = avail;
                break;
            }
            bool known;
            const Mat2 M = window_map(a, v, Fs - wc * a.L, Fs, &known);
            const double m = fmax(fmax(fabs(M.a), fabs(M.b)), fmax(fabs(M.c), fabs(M.d)));
            if (known || m <= 1e-8) break;
            wc *= 2;
        }
    }
    wc = __reduce_max_sync(FULL, wc);
    const uint32_t Fw = Fs - wc * a.L;
    if (g.warm_stats && lane == 0) {
        atomicAdd(&g.warm_stats[0], wc);
        atomicMax(&g.warm_stats[1], wc);
    }

    ChunkCursor cur;
    cur.init(a, v, valid, Fw / a.L);
    PulseStage ps;
    FilterStage fs;
    EnvStage es;
    const float color = valid ? a.color[v] : 0.0f;
    ps.init(color, 0u);
    fs.l = fs.b = 0.0f;
    fs.cut = fs.res = 0.0f;
    if (valid && Fw == a.F0) {
        const float2 s = carry_in[v];
        fs.l = s.x;
        fs.b = s.y;
    }
    if (cur.active) {
        ps.begin_span(a.sample_rate, cur.cur.freq);
        ps.set_cnt(cur.cur.cnt0 + (Fw - cur.cur.frame) * ps.ifreq);
        fs.begin_span(cur.cur.cutoff);
    }
    if (Fw < Fs) seg_warm_up<FF>(a, cur, ps, fs, Fw, Fs);

    env_reset(es.e, a.sample_rate);
    if (cur.active) {
        const EnvSnap s = a.snaps[cur.next_event - 1];
        es.e.state = s.state;
        es.e.painter.t = s.t;
        es.e.painter.last_value = s.last_value;
        es.e.painter.start = s.start;
        es.begin_span(a.sample_rate, cur.cur.note_on != 0, (cur.cur.flags & EV_CHANGED) != 0);
        env_advance(es.e, Fs - cur.cur.frame);
    }
    if (!VOUT) {
#pragma unroll
        for (int c = 0; c < C; c++) gn[c][lane] = valid ? a.gains[(size_t)v * C + c] : 0.0f;
    }
    __syncwarp();

    const uint32_t nframes = a.F1 - a.F0;
    for (uint32_t T0 = Fs; T0 < Fe; T0 += RT) {
        const uint32_t tn = min((uint32_t)RT, Fe - T0);
        uint32_t j = 0;
        while (j < tn) {
            const uint32_t f = T0 + j;
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
                        const float y = fs.sample_t<FF>(ps.sample());
                        oo[r * OS] = es.sample_fast<true>(er, lv) * y;
                    }
                } else {
                    for (uint32_t r = 0; r < run; r++) {
                        const float y = fs.sample_t<FF>(ps.sample());
                        oo[r * OS] = es.sample_fast<false>(er, lv) * y;
                    }
                }
                es.end_run(er, lv, run > 0);
            } else {
                for (uint32_t r = 0; r < run; r++) {
                    float o = 0.0f;
                    if (cur.active) {
                        const float y = fs.sample_t<FF>(ps.sample());
                        o = es.sample_careful() * y;
                    }
                    oo[r * OS] = o;
                }
            }
            j += run;
        }
        __syncwarp();
        if (VOUT) {
            for (uint32_t jj = 0; jj < tn; jj++) {
                const uint32_t f = T0 + jj;
                if (valid && f >= a.vout_first && f < a.vout_first + a.vout_frames && v - a.vout_v0 < a.vout_nv)
                    a.vout[(size_t)(f - a.vout_first) * a.vout_nv + (v - a.vout_v0)] = 0.0f + O[jj][lane];  // multiply() accumulates into the zeroed output: -0 becomes +0
            }
        } else {
            if (C == 2 && RT == 64) {
                // Lane j sums frames j and j + 32 over the warp's 32 voices.  Voices are taken in pairs: one LDS.64 reads two
                // adjacent voices of a frame row (stride 34 floats: conflict-free for 8-byte accesses), the pan gains of four
                // voices come as one LDS.128 per channel, and each fma.rn.f32x2 does two voices of one (frame, channel), so
                // the mix costs ~1.75 issue slots per voice-sample instead of 5 (even and odd voices accumulate separately and
                // meet at the end; the voice summation order is not the reference's anyway: tolerance 1e-5 x sum|gain|).
                typedef unsigned long long u64;
                const float* rowA = &O[lane][0];
                const float* rowB = &O[lane + 32][0];
                u64 a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll 2
                for (int q = 0; q < 8; q++) {
                    const ulonglong2 g0 = *reinterpret_cast<const ulonglong2*>(&gn[0][q * 4]);
                    const ulonglong2 g1 = *reinterpret_cast<const ulonglong2*>(&gn[1][q * 4]);
                    const u64 xa0 = *reinterpret_cast<const u64*>(rowA + q * 4), xa1 = *reinterpret_cast<const u64*>(rowA + q * 4 + 2);
                    const u64 xb0 = *reinterpret_cast<const u64*>(rowB + q * 4), xb1 = *reinterpret_cast<const u64*>(rowB + q * 4 + 2);
                    a0 = fma_f32x2(g0.x, xa0, a0);
                    a1 = fma_f32x2(g1.x, xa0, a1);
                    b0 = fma_f32x2(g0.x, xb0, b0);
                    b1 = fma_f32x2(g1.x, xb0, b1);
                    a0 = fma_f32x2(g0.y, xa1, a0);
                    a1 = fma_f32x2(g1.y, xa1, a1);
                    b0 = fma_f32x2(g0.y, xb1, b0);
                    b1 = fma_f32x2(g1.y, xb1, b1);
                }
                float* dst = a.partial + ((size_t)vgroup * nframes + (T0 - a.F0 + lane)) * 2;
                if (lane < tn) *reinterpret_cast<float2*>(dst) = make_float2(sum_f32x2(a0), sum_f32x2(a1));
                if (lane + 32 < tn) *reinterpret_cast<float2*>(dst + 64) = make_float2(sum_f32x2(b0), sum_f32x2(b1));
            } else {
                for (uint32_t fr = lane; fr < tn; fr += 32) {
                    float acc[C];
#pragma unroll
                    for (int c = 0; c < C; c++) acc[c] = 0.0f;
#pragma unroll 8
                    for (int vv = 0; vv < 32; vv++) {
                        const float x = O[fr][vv];
#pragma unroll
                        for (int c = 0; c < C; c++) acc[c] = fmaf(gn[c][vv], x, acc[c]);
                    }
                    float* dst = a.partial + ((size_t)vgroup * nframes + (T0 - a.F0 + fr)) * C;
                    if (C == 2)
                        *reinterpret_cast<float2*>(dst) = make_float2(acc[0], acc[1]);
                    else
                        dst[0] = acc[0];
                }
            }
        }
        __syncwarp();
    }
    if (valid && Fe == a.F1) carry_out[v] = make_float2(fs.l, fs.b);
}

}  // namespace zcu
