// zcu_paint_kernel.cuh — the generic "lockstep" paint kernel.
//
// One thread per voice, 32 adjacent voices per warp, every warp walks the frame axis of the
// [frame][voice] buffers in lockstep, so each access is one coalesced 128-byte row.  A voice's
// spans (= the reference's paint() calls) are entered and left at the voice's own frames; module
// state lives in registers for the whole block of buffers (Filter l/b, oscillator phase, PRNG ...).
//
// Memory pipeline: the strictly sequential recurrences give each warp only one sample of work at a
// time, so HBM latency is hidden by staging tiles of T frames x 32 voices through shared memory with
// cp.async (LDGSTS), two stages deep: while tile i is consumed, tile i+1 of the output buffer and of
// every input buffer is in flight.  Each thread only ever reads the shared-memory words it copied
// itself, so no block-level barrier is needed, only cp.async.wait_group.
//
// Time parallelism: 32 voices are one warp, so a small bank would leave the chip empty.  For modules whose state has an
// exact closed-form skip-ahead (M::SKIP: integer oscillator phases, the binade-stepped f32 accumulators of SineOsc /
// Sampler / Cycle / Painter, the Envelope machine, stateless modules) the frame axis is additionally cut into segments
// (blockIdx.y): a thread first walks its voice's spans up to its segment start with M::advance() instead of painting,
// then paints its segment.  The result is bit-identical to the sequential walk; only the last segment stores the state.
#pragma once

#include <cuda_pipeline.h>

#include "zcu_modules.cuh"

namespace zcu {

template <class M>
__global__ void __launch_bounds__(32) k_paint_lockstep(PaintArgs a, uint32_t T, uint32_t seg_frames) {
    extern __shared__ float smem[];
    const uint32_t tid = threadIdx.x, nt = blockDim.x;
    const uint32_t v = blockIdx.x * nt + tid;
    const uint32_t lanes = __ballot_sync(0xffffffffu, v < a.V);  // the lanes that take part in the per-run warp reductions
    if (v >= a.V) return;

    const float* bufs[4];
    int nb = 0;
    const int i_out = nb;
    bufs[nb++] = a.out;
    int i_in0 = -1, i_in1 = -1, i_in2 = -1;
    if (M::NIN >= 1 && a.in0) { i_in0 = nb; bufs[nb++] = a.in0; }
    if (M::NIN >= 2 && a.in1) { i_in1 = nb; bufs[nb++] = a.in1; }
    if (M::NIN >= 3 && a.in2) { i_in2 = nb; bufs[nb++] = a.in2; }

    const size_t V = a.V;
    // smem layout [stage][buf][T][nt]
    auto sm = [&](uint32_t stage, int buf, uint32_t j) -> float* { return smem + ((size_t)(stage * nb + buf) * T + j) * nt + tid; };

    // this CTA's segment of the frame axis (seg_frames is a multiple of T)
    const uint32_t f_begin = M::SKIP ? blockIdx.y * seg_frames : 0u;
    const uint32_t f_end = M::SKIP ? min(a.frames, f_begin + seg_frames) : a.frames;
    if (f_begin >= f_end) return;

    auto issue = [&](uint32_t tile) {
        const uint32_t f0 = tile * T;
        const uint32_t n = min(T, f_end - f0);
        const uint32_t stage = tile & 1u;
        for (int b = 0; b < nb; b++) {
            const float* src = bufs[b] + (size_t)f0 * V + v;
            for (uint32_t j = 0; j < n; j++) __pipeline_memcpy_async(sm(stage, b, j), src + (size_t)j * V, 4);
        }
        __pipeline_commit();
    };

    M m;
    m.load(a, v);
    uint32_t k = a.off[v];
    const uint32_t ke = a.off[v + 1];
    uint32_t s_start = 0xFFFFFFFFu, s_end = 0xFFFFFFFFu, s_flags = 0, s_param = 0;
    if (k < ke) {
        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
        s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
    }
    bool active = false;
    if (M::SKIP && f_begin > 0) {
        // spans that end before the segment: state only
        while (s_end <= f_begin) {
            m.begin_span(a, s_param, s_flags, s_end - s_start);
            m.advance(a, v, s_start, s_end - s_start);
            m.end_span(a);
            k++;
            if (k < ke) {
                const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
                s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
            } else {
                s_start = s_end = 0xFFFFFFFFu;
            }
        }
        if (s_start < f_begin) {  // the span that straddles the segment start
            m.begin_span(a, s_param, s_flags, s_end - s_start);
            m.advance(a, v, s_start, f_begin - s_start);
            active = true;
        }
    }

    const uint32_t tile0 = f_begin / T;
    const uint32_t ntiles = (f_end + T - 1) / T;
    if (ntiles > tile0) issue(tile0);
    for (uint32_t tile = tile0; tile < ntiles; tile++) {
        if (tile + 1 < ntiles) {
            issue(tile + 1);
            __pipeline_wait_prior(1);
        } else {
            __pipeline_wait_prior(0);
        }
        const uint32_t f0 = tile * T;
        const uint32_t n = min(T, f_end - f0);
        const uint32_t stage = tile & 1u;
        float* optr = a.out + (size_t)f0 * V + v;
        // Frames are walked in runs: up to the nearest span boundary of any of the warp's voices (one warp reduction per
        // run), so that the per-frame code carries no span bookkeeping and can be unrolled.
        uint32_t j = 0;
        while (j < n) {
            const uint32_t f = f0 + j;
            if (active && f == s_end) {
                m.end_span(a);
                active = false;
                k++;
                if (k < ke) {
                    const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
                    s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
                } else {
                    s_start = s_end = 0xFFFFFFFFu;
                }
            }
            if (!active && f == s_start) {
                m.begin_span(a, s_param, s_flags, s_end - s_start);
                active = true;
            }
            const uint32_t next_boundary = active ? s_end : s_start;
            const uint32_t run = min(n - j, __reduce_min_sync(lanes, next_boundary - f));
            if (active) {
                const float* so = sm(stage, i_out, j);
                const float* s0 = i_in0 >= 0 ? sm(stage, i_in0, j) : so;
                const float* s1 = i_in1 >= 0 ? sm(stage, i_in1, j) : so;
                const float* s2 = i_in2 >= 0 ? sm(stage, i_in2, j) : so;
                float* op = optr + (size_t)j * V;
#pragma unroll 4
                for (uint32_t r = 0; r < run; r++) {
                    float acc = so[r * 32];
                    const float x0 = i_in0 >= 0 ? s0[r * 32] : 0.0f;
                    const float x1 = i_in1 >= 0 ? s1[r * 32] : 0.0f;
                    const float x2 = i_in2 >= 0 ? s2[r * 32] : 0.0f;
                    m.sample(acc, x0, x1, x2);
                    op[(size_t)r * V] = acc;
                }
            }
            j += run;
        }
    }
    if (f_end < a.frames) return;  // not the last segment: its state is somebody else's start
    if (active) m.end_span(a);
    m.store(a, v);
}

}  // namespace zcu
