// zcu_paint_kernel.cuh — the generic "lockstep" paint kernel.
//
// One thread per voice, 32 adjacent voices per warp, every warp walks the frame axis of the
// [frame][voice] buffers in lockstep.  A voice's spans (= the reference's paint() calls) are entered and
// left at the voice's own frames; module state lives in registers for the whole block of buffers
// (Filter l/b, oscillator phase, PRNG ...).
//
// Memory pipeline (r01c; measured with tools/ubench_rowtile.cu: 6.1-6.4 TB/s against 3.9-4.5 TB/s for one-warp CTAs
// issuing 4-byte copies): a CTA is ZCU_PAINT_WARPS warps over 32*W adjacent voices, so that a frame row of the CTA is
// W*128 contiguous bytes.  Tiles of T frames of the output buffer and of every input buffer are staged through shared
// memory with 16-byte cp.async (LDGSTS.128: one warp instruction moves a whole 512-byte row), three stages deep.  The
// module reads its operands from the tile and writes its result back INTO the tile; the finished output tile is then
// written to HBM by the whole CTA as 16-byte row stores.  HBM therefore only ever sees full-row requests in both
// directions.  Frames outside every span are written back unchanged; a tile in which no voice of the CTA painted is
// not written at all.
//
// Time parallelism: 32 voices are one warp, so a small bank would leave the chip empty.  For modules whose state has an
// exact closed-form skip-ahead (M::SKIP: integer oscillator phases, the binade-stepped f32 accumulators of SineOsc /
// Sampler / Cycle / Painter, the Envelope machine, stateless modules) the frame axis is additionally cut into segments
// (blockIdx.y): a thread first walks its voice's spans up to its segment start with M::advance() instead of painting,
// then paints its segment.  The result is bit-identical to the sequential walk; only the last segment stores the state.
#pragma once

#include <cuda_pipeline.h>

#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "zcu_modules.cuh"

namespace zcu {

template <class M, class = void>
struct WarmOf { static constexpr bool value = false; };
template <class M>
struct WarmOf<M, std::void_t<decltype(M::WARM)>> { static constexpr bool value = M::WARM; };

// M::WRITE_ONLY: the module overwrites its output without reading it (zero / set / copy): the output tile is not loaded
template <class M, class = void>
struct WriteOnlyOf { static constexpr bool value = false; };
template <class M>
struct WriteOnlyOf<M, std::void_t<decltype(M::WRITE_ONLY)>> { static constexpr bool value = M::WRITE_ONLY; };

// M::HAS_RUN: the module offers branch-free samples between its own phase changes (run_limit / run_begin / sample_run /
// run_end, zcu_modules.cuh): a run then also ends where the module says, and all 32 lanes execute the same instructions
template <class M, class = void>
struct HasRunOf { static constexpr bool value = false; };
template <class M>
struct HasRunOf<M, std::void_t<decltype(M::HAS_RUN)>> { static constexpr bool value = M::HAS_RUN; };
template <class M, bool = HasRunOf<M>::value>
struct RunCall {
    static __device__ __forceinline__ uint32_t limit(M&) { return 0xFFFFFFFFu; }
    static __device__ __forceinline__ void begin(M&) {}
    static __device__ __forceinline__ void sample(M& m, float& acc, float x0, float x1, float x2) { m.sample(acc, x0, x1, x2); }
    static __device__ __forceinline__ void end(M&) {}
};
template <class M>
struct RunCall<M, true> {
    static __device__ __forceinline__ uint32_t limit(M& m) { return m.run_limit(); }
    static __device__ __forceinline__ void begin(M& m) { m.run_begin(); }
    static __device__ __forceinline__ void sample(M& m, float& acc, float x0, float x1, float x2) { m.sample_run(acc, x0, x1, x2); }
    static __device__ __forceinline__ void end(M& m) { m.run_end(); }
};

// M::CONST_FILL: a write-only module whose stored value is one constant per call (zero / set)
template <class M, class = void>
struct ConstFillOf { static constexpr bool value = false; };
template <class M>
struct ConstFillOf<M, std::void_t<decltype(M::CONST_FILL)>> { static constexpr bool value = M::CONST_FILL; };
template <class M, bool = ConstFillOf<M>::value>
struct ConstFillCall { static __device__ __forceinline__ float value(const M&) { return 0.0f; } };
template <class M>
struct ConstFillCall<M, true> { static __device__ __forceinline__ float value(const M& m) { return m.fill_value(); } };

// m.sample_warm() for modules that have it (the call sits in a generic lambda, where a plain member call would be looked up
// for every module)
// M::WARM_EXACT: the warm-up window is computed in closed form from the spans' constants (no verification afterwards)
template <class M, class = void>
struct WarmExactOf { static constexpr bool value = false; };
template <class M>
struct WarmExactOf<M, std::void_t<decltype(M::WARM_EXACT)>> { static constexpr bool value = M::WARM_EXACT; };
template <class M, bool = WarmExactOf<M>::value>
struct WarmExactCall { static __device__ __forceinline__ uint32_t window(M&, const PaintArgs&, uint32_t, uint32_t, uint32_t, uint32_t) { return 0xFFFFFFFFu; } };
template <class M>
struct WarmExactCall<M, true> {
    static __device__ __forceinline__ uint32_t window(M& m, const PaintArgs& a, uint32_t k0, uint32_t kb, uint32_t F, uint32_t f_lo) {
        return m.warm_window_exact(a, k0, kb, F, f_lo);
    }
};

template <class M, bool = WarmOf<M>::value>
struct WarmCall { static __device__ __forceinline__ void sample(M&, float, float, float) {} };
template <class M>
struct WarmCall<M, true> { static __device__ __forceinline__ void sample(M& m, float x0, float x1, float x2) { m.sample_warm(x0, x1, x2); } };

// M::SCAN: the carried state is an integer sum over the inputs (PulseOsc's phase with a frequency buffer), so time segments
// can be chained exactly: every segment first sums its own inputs (scan tiles), takes its start state from the predecessor's
// published inclusive prefix, publishes its own, then paints
template <class M, class = void>
struct ScanOf { static constexpr bool value = false; };
template <class M>
struct ScanOf<M, std::void_t<decltype(M::SCAN)>> { static constexpr bool value = M::SCAN; };
template <class M, bool = ScanOf<M>::value>
struct ScanCall { static __device__ __forceinline__ void sample(M&, float, float, float) {} };
template <class M>
struct ScanCall<M, true> { static __device__ __forceinline__ void sample(M& m, float x0, float x1, float x2) { m.scan_sample(x0, x1, x2); } };

// M::LINE_RING: delay modules.  The k-th painted sample of a call uses line slot index0 + k, whatever frame it falls on, so the
// line can be fetched in slot order ahead of time: every tile that is issued into the cp.async ring takes up to T more slots
// of each voice's line with it, into a per-thread queue of S * T entries behind the tile ring (one more staged buffer in the
// launch geometry).  A tile paints at most T frames, so the slots it needs have landed with it (and a slot is only fetched
// when fewer than S * T are outstanding, which also keeps the fetch behind the last write to that slot as long as the line
// is longer than the queue).  Without it a delay module waits for one global-memory round trip per four frames.
template <class M, class = void>
struct LineRingOf { static constexpr bool value = false; };
template <class M>
struct LineRingOf<M, std::void_t<decltype(M::LINE_RING)>> { static constexpr bool value = M::LINE_RING; };
template <class M, bool = LineRingOf<M>::value>
struct LineRingCall {
    static __device__ __forceinline__ void bind(M&, const float*, uint32_t, uint32_t) {}
    static __device__ __forceinline__ uint32_t taken(M&) { return 0; }
    static __device__ __forceinline__ uint32_t index(M&) { return 0; }
};
template <class M>
struct LineRingCall<M, true> {
    static __device__ __forceinline__ void bind(M& m, const float* q, uint32_t stride, uint32_t L) { m.delay_line().ring_bind(q, stride, L); }
    static __device__ __forceinline__ uint32_t taken(M& m) { return m.delay_line().taken; }
    static __device__ __forceinline__ uint32_t index(M& m) { return m.delay_line().index; }
};

// M::RESETS: a call with ZCU_SPAN_NOTE_ID_CHANGED for which m.resets(a, param) holds sets the module's whole carried state,
// whatever it was before (see the skip-ahead of k_paint_lockstep)
template <class M, class = void>
struct ResetsOf { static constexpr bool value = false; };
template <class M>
struct ResetsOf<M, std::void_t<decltype(M::RESETS)>> { static constexpr bool value = M::RESETS; };

// m.advance_done() for modules whose advance() only accumulates (M::LAZY)
template <class M, class = void>
struct LazyCall { static __device__ __forceinline__ void done(M&) {} };
template <class M>
struct LazyCall<M, std::enable_if_t<M::LAZY>> { static __device__ __forceinline__ void done(M& m) { m.advance_done(); } };

// M::PREFIX: a paint() call may be split at any frame (second half entered with note_id_changed = false) without changing a
// bit, so the state at every segment start can be tabulated by one walk per voice
template <class M, class = void>
struct PrefixOf { static constexpr bool value = false; };
template <class M>
struct PrefixOf<M, std::void_t<decltype(M::PREFIX)>> { static constexpr bool value = M::PREFIX; };

// One thread per voice walks all its spans once with M::advance() and stores the module state (bank POD format) at every
// segment start s * seg_frames, s = 1 .. nseg-1, into a.prefix[s][v].
template <class M>
__global__ void __launch_bounds__(128) k_paint_prefix(PaintArgs a, uint32_t seg_frames, uint32_t nseg) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= a.V) return;
    M m;
    typename ScratchOf<M>::type m_scratch;  // a module's run-time indexed table, kept apart so that m itself stays in registers
    ScratchOf<M>::bind(m, m_scratch);
    m.load(a, v);
    PaintArgs row = a;
    uint32_t s = 1;
    uint64_t next_b = seg_frames;
    auto record = [&]() {
        row.state = (char*)a.prefix + (size_t)s * a.V * a.prefix_stride;
        m.store(row, v);
        s++;
        next_b += seg_frames;
    };
    for (uint32_t k = a.off[v], ke = a.off[v + 1]; k < ke && s < nseg; k++) {
        const uint4 sp = *reinterpret_cast<const uint4*>(a.spans + k);
        while (s < nseg && next_b <= sp.x) record();  // boundaries in the gap before this span (or exactly at its start)
        m.begin_span(a, sp.w, sp.z, sp.y - sp.x);
        uint32_t pos = sp.x;
        while (s < nseg && next_b < sp.y) {
            m.advance(a, v, pos, (uint32_t)next_b - pos);
            pos = (uint32_t)next_b;
            record();
        }
        m.advance(a, v, pos, sp.y - pos);
        m.end_span(a);
    }
    while (s < nseg) record();
}

// M::PREFIX_RING (with M::PREFIX): the carried state is an order-dependent f32 sum over an input buffer (the phase of SineOsc /
// TriSawOsc / Cycle driven by a frequency buffer): no skip-ahead exists and the per-voice walk has to read every input sample,
// but one step of it is a multiply (or an exact division) and an add where a painted sample costs a sine.
// k_paint_prefix_ring is that walk, alone on the machine before the paint launch.  One CTA per 32 voices, lane = voice, the
// input rows (128 bytes per frame) staged through a 64 KB cp.async ring in groups of 32 frames.  Only the additions are a
// dependent chain, so only they stay with the one warp that walks the frames (warp 0: span boundaries, the adds, the state
// recorded at every segment start like k_paint_prefix does); eight producer warps issue the copies and turn the landed input
// samples into increments in place, one group ahead (each with a span cursor of its own, for the span constants of its
// rows).  The module's begin_span() must leave the carried state alone (it only derives span constants): the walker does
// not call it.  The paint launch then runs its time segments in parallel from the recorded rows.
template <class M, class = void>
struct PrefixRingOf { static constexpr bool value = false; };
template <class M>
struct PrefixRingOf<M, std::void_t<decltype(M::PREFIX_RING)>> { static constexpr bool value = M::PREFIX_RING; };

constexpr uint32_t kRingRows = 512;    // rows of 32 voices in a CTA's ring
constexpr uint32_t kRingGroup = 64;    // rows per cp.async group
constexpr uint32_t kRingAhead = 1024;  // frames between the walk and its L2 prefetch
constexpr uint32_t kRingProducers = 8; // producer warps (two per scheduler: they hide each other's latencies)
constexpr uint32_t kRingThreads = 32 * (1 + kRingProducers);  // warp 0 walks

template <class M>
__global__ void __launch_bounds__(kRingThreads) k_paint_prefix_ring(PaintArgs a, uint32_t seg_frames, uint32_t row_lo, uint32_t row_hi, uint32_t vec4, uint32_t ahead) {
    extern __shared__ __align__(16) float ring[];  // [kRingRows][32]
    constexpr uint32_t G = kRingGroup, NG = kRingRows / kRingGroup;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t v0 = blockIdx.x * 32u, v = v0 + lane;
    const bool valid = v < a.V;
    const size_t V = a.V;
    // this launch walks the frames [row_lo, row_hi) x seg_frames and records the rows row_lo + 1 .. row_hi (row s = the state at
    // the start of segment s; row_lo > 0: the walk starts from row row_lo, which an earlier launch recorded)
    const uint32_t f_lo = row_lo * seg_frames, f_stop = row_hi * seg_frames;
    const int ngroups = (int)((f_stop - f_lo + G - 1) / G);
    const uint32_t ring_u32 = (uint32_t)__cvta_generic_to_shared(ring);
    const bool producer = warp != 0;
    constexpr uint32_t P = kRingProducers;
    const uint32_t pt = tid - 32u;  // producer thread
    // producers: the rows of the next group on their way into the next slot.  Groups are issued in order, so everything is a
    // running value: source pointer of this thread's first row, slot offset, prefetch pointer.
    const uint32_t chunk = vec4 ? (pt & 7u) * 4u : lane, rsub = vec4 ? pt >> 3 : pt >> 5;
    const bool col_ok = v0 + chunk < a.V;
    const uint32_t rstep = vec4 ? 4u * P : P;                      // rows between two copies of a thread
    const size_t src_rstep = (size_t)V * 4u * rstep, src_gstep = (size_t)V * 4u * G;
    const uint32_t dst_rstep = rstep * 128u;
    const char* iss_src = reinterpret_cast<const char*>(a.in0 + ((size_t)(f_lo + rsub) * V + v0 + chunk));
    // L2 prefetch, one 32-byte sector per producer thread (a prefetch brings in the sector it names, not the 128-byte line)
    const uint32_t pf_row = pt >> 2, pf_col = (pt & 3u) * 8u;
    const bool pf_ok = pf_row < G && v0 + pf_col < a.V;
    const char* iss_pf = reinterpret_cast<const char*>(a.in0 + ((size_t)(f_lo + ahead + pf_row) * V + v0 + pf_col));
    uint32_t iss_slot = 0, iss_f0 = f_lo;
    auto issue = [&]() {
        if (iss_f0 < f_stop) {
            const uint32_t n = min(G, f_stop - iss_f0);
            const char* sp = iss_src;
            uint32_t dp = ring_u32 + iss_slot * (G * 128u) + rsub * 128u + chunk * 4u;
            if (col_ok) {
                if (vec4) {  // a row = eight 16-byte chunks: 4 P rows per pass of the producer threads
#pragma unroll
                    for (uint32_t j = 0; j < G / (4u * P); j++) {
                        if (rsub + j * 4u * P < n) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dp), "l"(sp) : "memory");
                        sp += src_rstep;
                        dp += dst_rstep;
                    }
                } else {
#pragma unroll
                    for (uint32_t j = 0; j < G / P; j++) {
                        if (rsub + j * P < n) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dp), "l"(sp) : "memory");
                        sp += src_rstep;
                        dp += dst_rstep;
                    }
                }
            }
            // the rows kRingAhead frames further on are pulled into L2
            if (pf_ok && ahead != 0u && iss_f0 + ahead + pf_row < f_stop) asm volatile("prefetch.global.L2 [%0];" ::"l"(iss_pf));
            iss_src += src_gstep;
            iss_pf += src_gstep;
            iss_f0 += G;
            iss_slot = iss_slot + 1 == NG ? 0u : iss_slot + 1;
        }
        __pipeline_commit();
    };
    if (producer) {
        for (int g = 0; g + 2 < (int)NG; g++) issue();  // groups 0 .. NG-3
    }

    M m = M();
    typename ScratchOf<M>::type m_scratch;
    ScratchOf<M>::bind(m, m_scratch);
    uint32_t k = 0, ke = 0;
    PaintArgs row = a;
    if (valid) {
        if (!producer) {
            if (row_lo > 0) row.state = (char*)a.prefix + (size_t)row_lo * a.V * a.prefix_stride;
            m.load(row, v);
        }
        k = a.off[v];
        ke = a.off[v + 1];
    }
    // the span under the cursor, and the one after it already on its way (a lone warp cannot hide the fetch behind others)
    const uint4 kNoSpan = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
    uint4 cur = kNoSpan, nxt = kNoSpan;
    if (k < ke) cur = *reinterpret_cast<const uint4*>(a.spans + k);
    if (k + 1 < ke) nxt = *reinterpret_cast<const uint4*>(a.spans + k + 1);
    auto next_span = [&]() {
        k++;
        cur = nxt;
        nxt = kNoSpan;
        if (k + 1 < ke) nxt = *reinterpret_cast<const uint4*>(a.spans + k + 1);
    };
    // walker state (a launch that starts inside the block: behind the spans that ended, inside the one that straddles)
    while (cur.y <= f_lo) next_span();
    bool active = !producer && cur.x < f_lo;
    uint32_t s = row_lo + 1, next_b = f_lo + seg_frames;
    auto record = [&]() {
        row.state = (char*)a.prefix + (size_t)s * a.V * a.prefix_stride;
        if (valid) m.store(row, v);
        s++;
        next_b += seg_frames;
    };
    // producer state: the span whose constants m holds, the slot of the next group to prepare
    uint32_t k_consts = 0xFFFFFFFFu, prep_slot = 0;
    uint32_t walk_slot = 0;

    // iteration i: group i + 1 has landed (barrier), group i + NG - 1 is issued into the slot of group i - 1, the producers turn
    // group i + 1 into increments while the walker adds up group i
    for (int i = -1; i < ngroups; i++) {
        if (producer) __pipeline_wait_prior(NG - 3);
        __syncthreads();
        if (producer) {
            issue();  // group i + NG - 1
            const int g = i + 1;
            if (g < ngroups) {
                const uint32_t f0 = f_lo + (uint32_t)g * G, n = min(G, f_stop - f0);
                float* gp0 = ring + (size_t)prep_slot * G * 32u + lane;
                prep_slot = prep_slot + 1 == NG ? 0u : prep_slot + 1;
#pragma unroll 1
                for (uint32_t r0 = (warp - 1u) * (G / P), re = r0 + G / P; r0 < re && r0 < n; r0 += 8u) {
                    float* gp = gp0 + r0 * 32u;
                    const uint32_t fa = f0 + r0, nr = min(8u, n - r0);
                    while (cur.y <= fa) next_span();
                    const bool whole = cur.x <= fa && cur.y >= fa + nr;  // every row of mine inside the span under the cursor
                    const bool none = cur.x >= fa + nr;                  // ... or before it
                    if (whole && k != k_consts) {
                        m.begin_span(a, cur.w, cur.z, cur.y - cur.x);
                        k_consts = k;
                    }
                    if (nr == 8u && __all_sync(0xffffffffu, whole || none)) {
                        float xs[8], d[8];  // lanes outside every span compute values nobody reads
#pragma unroll
                        for (int q = 0; q < 8; q++) xs[q] = gp[q * 32];
                        m.scan_prep(xs, d);
#pragma unroll
                        for (int q = 0; q < 8; q++) gp[q * 32] = d[q];
                    } else {
                        for (uint32_t r = 0; r < nr; r++) {
                            const uint32_t f = fa + r;
                            while (cur.y <= f) next_span();
                            if (cur.x <= f) {
                                if (k != k_consts) {
                                    m.begin_span(a, cur.w, cur.z, cur.y - cur.x);
                                    k_consts = k;
                                }
                                gp[r * 32] = m.scan_prep1(gp[r * 32]);
                            }
                        }
                    }
                }
            }
        } else if (i >= 0) {
            const uint32_t f0 = f_lo + (uint32_t)i * G, n = min(G, f_stop - f0);
            const float* gp = ring + (size_t)walk_slot * G * 32u + lane;
            walk_slot = walk_slot + 1 == NG ? 0u : walk_slot + 1;
            uint32_t j = 0;
            while (j < n) {
                const uint32_t f = f0 + j;
                if (active && f == cur.y) {
                    m.end_span(a);
                    active = false;
                    next_span();
                }
                if (f == next_b) record();  // between the span that ended here and the one that starts here
                if (!active && f == cur.x) active = true;
                const uint32_t dist = (active ? cur.y : cur.x) - f;
                const uint32_t run = min(min(n - j, next_b - f), __reduce_min_sync(0xffffffffu, dist));
                if (active) {
                    const float* s0 = gp + j * 32u;
                    uint32_t r = 0;
                    // the common case is a run over the whole group: all the loads first, then the chain of additions
                    if (run == G) {
                        float d[G / 8][8];
#pragma unroll
                        for (int q = 0; q < (int)G; q++) d[q >> 3][q & 7] = s0[q * 32];
#pragma unroll
                        for (int b = 0; b < (int)G / 8; b++) m.scan_acc(d[b]);
                        r = G;
                    }
                    for (; r + 32 <= run; r += 32) {
                        float d[4][8];
#pragma unroll
                        for (int q = 0; q < 32; q++) d[q >> 3][q & 7] = s0[(r + q) * 32u];
#pragma unroll
                        for (int b = 0; b < 4; b++) m.scan_acc(d[b]);
                    }
                    for (; r + 8 <= run; r += 8) {
                        float d[8];
#pragma unroll
                        for (int q = 0; q < 8; q++) d[q] = s0[(r + q) * 32u];
                        m.scan_acc(d);
                    }
                    for (; r < run; r++) m.scan_add(s0[r * 32u]);
                }
                j += run;
            }
        }
    }
    if (producer) return;
    if (active && cur.y == f_stop) m.end_span(a);
    while (s <= row_hi) record();
}

// cp.async.wait_group takes an immediate: dispatch the ring depths in use
__device__ __forceinline__ void pipeline_wait_prior_rt(uint32_t n) {
    switch (n) {
        case 0: __pipeline_wait_prior(0); break;
        case 1: __pipeline_wait_prior(1); break;
        case 2: __pipeline_wait_prior(2); break;
        case 3: __pipeline_wait_prior(3); break;
        case 4: __pipeline_wait_prior(4); break;
        case 5: __pipeline_wait_prior(5); break;
        default: __pipeline_wait_prior(6); break;
    }
}

// ---- k_paint_walk: modules that cannot be cut in time, on banks too small to fill the machine ----------------------------
// A module without a skip-ahead (the delay modules, the FM oscillator, the Filter in exact mode) walks its frames in order;
// with 4096 voices that is 128 warps on 592 schedulers, each paying for its own tile copies, ring bookkeeping and write-back
// at one dependent instruction every ~3.6 cycles (ncu: SimpleDelay 62 instructions per warp and frame, 28 % issue).  Here a CTA
// is 32 voices (one CTA per SM up to 4736 voices): warp 0 walks the frames and does nothing but the module (operands from the
// tiles in shared memory, results back into the output tile, delay-line stores to global); warps 2 and 3 issue the cp.async
// copies of the output and input tiles (groups of 32 frames, eight groups in the ring) and write finished output tiles back;
// warp 1 fetches the delay line in slot order (M::LINE_RING above; idle for other modules).  One barrier per group.
constexpr uint32_t kWalkGroup = 32;    // frames per group
constexpr uint32_t kWalkGroups = 8;    // groups in the ring, at most
constexpr uint32_t kWalkThreads = 128;

// OVR (ZCU_OUT_OVERWRITE): `zero(span, out); paint(span, out)`: inside its spans the output starts from 0 (the tile is still
// staged: frames outside the spans go back unchanged)
template <class M, bool OVR = false>
__global__ void __launch_bounds__(kWalkThreads) k_paint_walk(PaintArgs a, uint32_t vec4, uint32_t RW, uint32_t NG) {
    extern __shared__ __align__(16) float smem[];
    // RW = voices of this CTA (32, 16 or 8: rows of RW floats, lanes >= RW idle), NG = groups in the ring
    constexpr uint32_t G = kWalkGroup;
    const uint32_t L = G * NG;
    __shared__ uint32_t s_taken[32];
    __shared__ uint32_t s_painted[kWalkGroups];
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t v0 = blockIdx.x * RW, v = v0 + lane;
    const bool valid = lane < RW && v < a.V;
    const size_t V = a.V;
    const uint32_t frames = a.frames;
    const int ngroups = (int)((frames + G - 1) / G);
    // staged buffers: slot 0 = output, then the inputs that are present
    int nb = 1;
    int i_in0 = -1, i_in1 = -1, i_in2 = -1;
    if (M::NIN >= 1 && a.in0) i_in0 = nb++;
    if (M::NIN >= 2 && a.in1) i_in1 = nb++;
    if (M::NIN >= 3 && a.in2) i_in2 = nb++;
    // smem: [NG][nb][G][32] tiles, then the line queue [L][32]
    const uint32_t smem_u32 = (uint32_t)__cvta_generic_to_shared(smem);
    float* lq = smem + (size_t)NG * nb * G * RW;
    const bool line_ring = LineRingOf<M>::value && a.line_ring != 0u;
    if (tid < 32) s_taken[tid] = 0;
    if (tid < NG) s_painted[tid] = 0;

    // ---- producers -----------------------------------------------------------------------------------------------------
    const uint32_t pt = tid - 64u;  // tile-copy thread 0..63 (warps 2, 3)
    const uint32_t cpr = vec4 ? RW / 4u : RW;  // copies per row
    const uint32_t chunk = (pt % cpr) * (vec4 ? 4u : 1u), rsub = pt / cpr;
    const uint32_t rstep = 64u / cpr;           // rows per pass of the 64 copy threads
    const bool col_ok = v0 + chunk < a.V;
    uint32_t iss_slot = 0, iss_f0 = 0;
    uint32_t lq_pos = 0, lq_slot = 0, lq_total = 0;
    auto copy_buf = [&](const float* buf, int b, uint32_t n) {
        const char* sp = reinterpret_cast<const char*>(buf + ((size_t)(iss_f0 + rsub) * V + v0 + chunk));
        uint32_t dp = smem_u32 + (((iss_slot * nb + b) * G + rsub) * RW + chunk) * 4u;
        const size_t sstep = V * 4u * rstep;
        const uint32_t dstep = rstep * RW * 4u;
        if (vec4) {
#pragma unroll 4
            for (uint32_t j = rsub; j < n; j += rstep) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dp), "l"(sp) : "memory");
                sp += sstep;
                dp += dstep;
            }
        } else {
#pragma unroll 4
            for (uint32_t j = rsub; j < n; j += rstep) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dp), "l"(sp) : "memory");
                sp += sstep;
                dp += dstep;
            }
        }
    };
    auto issue = [&]() {  // the next group into the next slot (groups are issued in order)
        if (iss_f0 < frames) {
            const uint32_t n = min(G, frames - iss_f0);
            if (warp >= 2 && col_ok) {
                copy_buf(a.out, 0, n);
                if (M::NIN >= 1 && i_in0 >= 0) copy_buf(a.in0, i_in0, n);
                if (M::NIN >= 2 && i_in1 >= 0) copy_buf(a.in1, i_in1, n);
                if (M::NIN >= 3 && i_in2 >= 0) copy_buf(a.in2, i_in2, n);
            }
            if constexpr (LineRingOf<M>::value) {
                if (warp == 1 && line_ring && valid) {
                    const uint32_t np = min(G, L - (lq_total - s_taken[lane]));
                    const char* sp = reinterpret_cast<const char*>(a.line + ((size_t)lq_slot * V + v));
                    uint32_t dp = smem_u32 + ((NG * nb * G + lq_pos) * RW + lane) * 4u;
                    uint32_t done = 0;
                    while (done < np) {  // straight pieces: up to the wrap of the line or of the queue
                        const uint32_t piece = min(np - done, min(a.delay_samples - lq_slot, L - lq_pos));
#pragma unroll 4
                        for (uint32_t j = 0; j < piece; j++) {
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dp), "l"(sp) : "memory");
                            sp += V * 4u;
                            dp += RW * 4u;
                        }
                        done += piece;
                        lq_slot += piece;
                        lq_pos += piece;
                        if (lq_slot == a.delay_samples) {
                            lq_slot = 0;
                            sp = reinterpret_cast<const char*>(a.line + v);
                        }
                        if (lq_pos == L) {
                            lq_pos = 0;
                            dp = smem_u32 + (NG * nb * G * RW + lane) * 4u;
                        }
                    }
                    lq_total += np;
                }
            }
            iss_f0 += G;
            iss_slot = iss_slot + 1 == NG ? 0u : iss_slot + 1;
        }
        __pipeline_commit();
    };
    auto write_back = [&](uint32_t slot, uint32_t f0) {  // finished output tile -> HBM (rows of 128 bytes)
        if (warp < 2 || !col_ok || !s_painted[slot]) return;
        const uint32_t n = min(G, frames - f0);
        const float* st = smem + ((size_t)(slot * nb) * G + rsub) * RW + chunk;
        float* dg = a.out + ((size_t)(f0 + rsub) * V + v0 + chunk);
        if (vec4) {
#pragma unroll 4
            for (uint32_t j = rsub; j < n; j += rstep) {
                *reinterpret_cast<float4*>(dg) = *reinterpret_cast<const float4*>(st);
                dg += V * rstep;
                st += rstep * RW;
            }
        } else {
#pragma unroll 4
            for (uint32_t j = rsub; j < n; j += rstep) {
                *dg = *st;
                dg += V * rstep;
                st += rstep * RW;
            }
        }
    };
    if constexpr (LineRingOf<M>::value) {
        if (warp == 1 && line_ring && valid) {  // the line is fetched from the bank's index on
            M mp;
            mp.load(a, v);
            lq_slot = LineRingCall<M>::index(mp);
        }
    }
    __syncthreads();
    const bool producer = warp != 0;
    if (producer) {
        for (uint32_t g = 0; g + 1 < NG; g++) issue();  // groups 0 .. NG-2
    }

    // ---- walker --------------------------------------------------------------------------------------------------------
    M m;
    typename ScratchOf<M>::type m_scratch;  // a module's run-time indexed table, kept apart so that m itself stays in registers
    ScratchOf<M>::bind(m, m_scratch);
    uint32_t k = 0, ke = 0;
    uint32_t s_start = 0xFFFFFFFFu, s_end = 0xFFFFFFFFu, s_flags = 0, s_param = 0;
    uint4 nxt = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);  // the span after the one under the cursor, already fetched
    bool active = false;
    if (!producer && valid) {
        m.load(a, v);
        if (line_ring) LineRingCall<M>::bind(m, lq + lane, RW, L);
        k = a.off[v];
        ke = a.off[v + 1];
        if (k < ke) {
            const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
            s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
        }
        if (k + 1 < ke) nxt = *reinterpret_cast<const uint4*>(a.spans + k + 1);
    }
    for (int i = 0; i < ngroups; i++) {
        if (producer) pipeline_wait_prior_rt(NG - 2);
        __syncthreads();  // group i has landed; group i - 1 is walked
        const uint32_t slot = (uint32_t)i % NG;
        if (producer) {
            if (i >= 1) write_back(slot == 0 ? NG - 1 : slot - 1, (uint32_t)(i - 1) * G);
            issue();  // group i + NG - 1 into the slot of group i - 1
        } else {
            const uint32_t f0 = (uint32_t)i * G, n = min(G, frames - f0);
            float* so_t = smem + (size_t)(slot * nb) * G * RW + lane;
            const float* s0_t = i_in0 >= 0 ? so_t + (size_t)i_in0 * G * RW : so_t;
            const float* s1_t = i_in1 >= 0 ? so_t + (size_t)i_in1 * G * RW : so_t;
            const float* s2_t = i_in2 >= 0 ? so_t + (size_t)i_in2 * G * RW : so_t;
            bool painted = false;
            uint32_t j = 0;
            while (j < n) {
                const uint32_t f = f0 + j;
                if (active && f == s_end) {
                    m.end_span(a);
                    active = false;
                    k++;
                    s_start = nxt.x; s_end = nxt.y; s_flags = nxt.z; s_param = nxt.w;
                    nxt = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
                    if (k + 1 < ke) nxt = *reinterpret_cast<const uint4*>(a.spans + k + 1);
                }
                if (!active && f == s_start) {
                    m.begin_span(a, s_param, s_flags, s_end - s_start);
                    active = true;
                }
                const uint32_t dist = (active ? s_end : s_start) - f;
                const uint32_t run = min(n - j, __reduce_min_sync(0xffffffffu, dist));
                if (active) {
                    painted = true;
                    float* so = so_t + j * RW;
                    const float* s0 = s0_t + j * RW;
                    const float* s1 = s1_t + j * RW;
                    const float* s2 = s2_t + j * RW;
                    uint32_t r = 0;
                    for (; r + 4 <= run; r += 4) {
                        float acc[4], x0[4], x1[4], x2[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            // (an input that is absent aliases the output tile: its value is read but no module variant
                            // that can run without the buffer looks at it)
                            acc[q] = OVR ? 0.0f : so[(r + q) * RW];
                            x0[q] = M::NIN >= 1 ? s0[(r + q) * RW] : 0.0f;
                            x1[q] = M::NIN >= 2 ? s1[(r + q) * RW] : 0.0f;
                            x2[q] = M::NIN >= 3 ? s2[(r + q) * RW] : 0.0f;
                        }
                        Sample4<M>::call(m, acc, x0, x1, x2);
#pragma unroll
                        for (int q = 0; q < 4; q++) so[(r + q) * RW] = acc[q];
                    }
                    for (; r < run; r++) {
                        float acc = OVR ? 0.0f : so[r * RW];
                        m.sample(acc, i_in0 >= 0 ? s0[r * RW] : 0.0f, i_in1 >= 0 ? s1[r * RW] : 0.0f, i_in2 >= 0 ? s2[r * RW] : 0.0f);
                        so[r * RW] = acc;
                    }
                }
                j += run;
            }
            const bool any = __any_sync(0xffffffffu, painted);
            if (lane == 0) s_painted[slot] = any ? 1u : 0u;
            if constexpr (LineRingOf<M>::value) {
                if (line_ring) s_taken[lane] = LineRingCall<M>::taken(m);
            }
        }
    }
    __syncthreads();
    if (producer) {
        if (ngroups >= 1) write_back((uint32_t)(ngroups - 1) % NG, (uint32_t)(ngroups - 1) * G);
        return;
    }
    if (!valid) return;
    if (active) m.end_span(a);
    m.store(a, v);
}

constexpr int kPaintWarps = 4;    // warps per CTA: rows of 128 voices = 512 bytes
constexpr int kPaintStages = 3;   // tiles in the cp.async ring

// OVR (ZCU_OUT_OVERWRITE): the call stands for `zero(span, out); paint(span, out)`: inside its spans the output starts from 0
// and is never read, so the output tile is not loaded (4 instead of 12 bytes per voice-sample for a generator)
template <class M, bool OVR = false>
__global__ void __launch_bounds__(32 * kPaintWarps) k_paint_lockstep(PaintArgs a, uint32_t T, uint32_t seg_frames, uint32_t vec4, uint32_t S) {
    extern __shared__ __align__(16) float smem[];
    constexpr uint32_t W = kPaintWarps, RV = 32 * W;  // S = tiles in the cp.async ring (3..8: deeper when few CTAs share an SM)
    const uint32_t tid = threadIdx.x;
    // (voice block, segment) of this CTA: blockIdx, or — in launches whose segments wait for each other — a ticket
    uint32_t bx = blockIdx.x, by = blockIdx.y + a.seg_base;
    if (a.ticket != nullptr) {
        __shared__ uint32_t s_ticket;
        if (tid == 0) s_ticket = atomicAdd(a.ticket, 1u) - a.ticket_base;
        __syncthreads();
        bx = s_ticket % gridDim.x;
        by = s_ticket / gridDim.x;
    }
    const uint32_t v0 = bx * RV;
    const uint32_t v = v0 + tid;
    const bool valid = v < a.V;
    // wait for the flag of (bx, by - 1); bounded: ~4 s, then flag the error and go on
    auto wait_predecessor = [&](uint32_t ns) {
        const volatile uint32_t* fl = a.chain_flags + (size_t)(by - 1) * gridDim.x + bx;
        uint32_t spins = 0;
        while (*fl != a.epoch) {
            __nanosleep(ns);
            if (++spins > (1u << 24)) {
                if (a.chain_err) atomicExch(a.chain_err, 1u);
                break;
            }
        }
        __threadfence();
    };

    // staged buffers: slot 0 = output, then the inputs that are present (no pointer array: it would live in local memory)
    int nb = 1;
    const int i_out = 0;
    int i_in0 = -1, i_in1 = -1, i_in2 = -1;
    if (M::NIN >= 1 && a.in0) i_in0 = nb++;
    if (M::NIN >= 2 && a.in1) i_in1 = nb++;
    if (M::NIN >= 3 && a.in2) i_in2 = nb++;

    const size_t V = a.V;
    // smem layout [stage][buf][T][RV]
    auto tile_ptr = [&](uint32_t stage, int buf) -> float* { return smem + (size_t)(stage * nb + buf) * T * RV; };

    // this CTA's segment of the frame axis (seg_frames is a multiple of T)
    constexpr bool kWarm = WarmOf<M>::value;
    constexpr bool kWriteOnly = WriteOnlyOf<M>::value || OVR;
    constexpr bool kScan = ScanOf<M>::value;
    const bool warm = kWarm && a.chain != nullptr;  // tolerance mode: segments start from a verified warm-up or the chain
    const bool scan = kScan && a.chain != nullptr;  // exact chained prefix sum: segments pass their integer state along
    const bool segmented = M::SKIP || warm || scan || (PrefixOf<M>::value && a.prefix != nullptr);
    const uint32_t f_begin = segmented ? by * seg_frames : 0u;
    const uint32_t f_end = segmented ? min(a.frames, f_begin + seg_frames) : a.frames;
    if (f_begin >= f_end) return;  // uniform over the CTA

    // cooperative row copies: a row is RV floats; thread t moves chunk c of rows r0, r0 + RPP, ...
    const uint32_t cw = vec4 ? 4u : 1u;          // floats per chunk (16-byte chunks need V % 4 == 0)
    const uint32_t cpr = RV / cw;                // chunks per row
    const uint32_t rpp = RV / cpr;               // rows per pass over the CTA's threads (4 or 1)
    const uint32_t c = (tid % cpr) * cw, r0 = tid / cpr;
    const bool c_valid = v0 + c < a.V;

    M m;
    typename ScratchOf<M>::type m_scratch;  // a module's run-time indexed table, kept apart so that m itself stays in registers
    ScratchOf<M>::bind(m, m_scratch);
    // delay modules: the line prefetch queue behind the tile ring, [S * T][RV]
    const bool line_ring = LineRingOf<M>::value && a.line_ring != 0u;
    const uint32_t lq_len = S * T;
    uint32_t lq_pos = 0, lq_slot = 0, lq_total = 0;  // next queue entry / line slot to fetch, slots fetched so far
    bool inputs_only = false;  // scan pass: the refill skips the output buffer
    // Row copies are address arithmetic more than anything else (one cp.async per 512-byte row and warp): the shared-memory
    // address is a 32-bit offset bumped by a constant, the global one a pointer bumped by four rows.
    const uint32_t smem_u32 = (uint32_t)__cvta_generic_to_shared(smem);
    const size_t src_step = (size_t)V * sizeof(float) * rpp;
    auto copy_rows = [&](const float* buf, int slot, uint32_t stage, uint32_t f0, uint32_t n) {
        const char* sp = reinterpret_cast<const char*>(buf + ((size_t)(f0 + r0) * V + v0 + c));
        uint32_t dp = smem_u32 + (((stage * nb + slot) * T + r0) * RV + c) * 4u;
        if (vec4) {
#pragma unroll 4
            for (uint32_t j = r0; j < n; j += 4) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dp), "l"(sp) : "memory");
                sp += src_step;
                dp += 4u * RV * 4u;
            }
        } else {
#pragma unroll 4
            for (uint32_t j = r0; j < n; j++) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dp), "l"(sp) : "memory");
                sp += src_step;
                dp += RV * 4u;
            }
        }
    };
    uint32_t st_iss = 0, st_cur = 0;  // ring stage of the next tile to issue / to walk (tiles of a pass are taken in order)
    auto issue = [&](uint32_t tile) {
        const uint32_t f0 = tile * T;
        const uint32_t n = min(T, f_end - f0);
        const uint32_t stage = st_iss;
        st_iss = st_iss + 1 == S ? 0u : st_iss + 1;
        if (c_valid) {
            // warm-up tiles (before the segment) only read the inputs: the output rows there belong to the previous segment
            if (!kWriteOnly && !(kWarm && f0 < f_begin) && !(kScan && inputs_only)) copy_rows(a.out, i_out, stage, f0, n);
            if (M::NIN >= 1 && i_in0 >= 0) copy_rows(a.in0, i_in0, stage, f0, n);
            if (M::NIN >= 2 && i_in1 >= 0) copy_rows(a.in1, i_in1, stage, f0, n);
            if (M::NIN >= 3 && i_in2 >= 0) copy_rows(a.in2, i_in2, stage, f0, n);
        }
        if constexpr (LineRingOf<M>::value) {
            if (line_ring && valid) {
                const uint32_t np = min(T, lq_len - (lq_total - LineRingCall<M>::taken(m)));
                const char* sp = reinterpret_cast<const char*>(a.line + ((size_t)lq_slot * V + v));
                uint32_t dp = smem_u32 + ((S * nb * T + lq_pos) * RV + tid) * 4u;
                for (uint32_t j = 0; j < np; j++) {
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dp), "l"(sp) : "memory");
                    lq_slot++;
                    sp += V * sizeof(float);
                    if (lq_slot == a.delay_samples) {
                        lq_slot = 0;
                        sp = reinterpret_cast<const char*>(a.line + v);
                    }
                    lq_pos++;
                    dp += RV * 4u;
                    if (lq_pos == lq_len) {
                        lq_pos = 0;
                        dp = smem_u32 + (S * nb * T * RV + tid) * 4u;
                    }
                }
                lq_total += np;
            }
        }
        __pipeline_commit();
    };

    uint32_t k = 0, ke = 0;
    if (valid) {
        m.load(a, v);
        k = a.off[v];
        ke = a.off[v + 1];
        if (line_ring) {
            LineRingCall<M>::bind(m, smem + (size_t)S * nb * T * RV + tid, RV, lq_len);
            lq_slot = LineRingCall<M>::index(m);
        }
    }
    uint32_t s_start = 0xFFFFFFFFu, s_end = 0xFFFFFFFFu, s_flags = 0, s_param = 0;
    if (k < ke) {
        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
        s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
    }
    const uint32_t tile0 = f_begin / T;
    const uint32_t ntiles = (f_end + T - 1) / T;
    uint32_t tile_first = tile0;
    bool active = false;
    bool pristine = false;
    // one tile of the ring: wait, refill, walk the frames in runs, write the output tile back.  Warm-up tiles (tolerance
    // mode, before the segment) and scan tiles (first pass of an exact chained prefix sum) only advance module state from the
    // inputs: nothing is read from or written to the output.
    auto do_tile = [&](uint32_t tile, auto kind_tag) {
        constexpr int KIND = decltype(kind_tag)::value;  // 0 paint, 1 warm-up tile, 2 scan tile (kScan pass 1)
        constexpr bool WT = KIND != 0;
        pipeline_wait_prior_rt(S - 2);
        __syncthreads();  // tile `tile` has landed for every thread; every thread is done with the stage refilled next
        if (tile + S - 1 < ntiles) issue(tile + S - 1); else __pipeline_commit();
        const uint32_t f0 = tile * T;
        const uint32_t n = min(T, f_end - f0);
        const uint32_t stage = st_cur;
        st_cur = st_cur + 1 == S ? 0u : st_cur + 1;
        float* so_t = tile_ptr(stage, i_out) + tid;
        const float* s0_t = i_in0 >= 0 ? tile_ptr(stage, i_in0) + tid : so_t;
        const float* s1_t = i_in1 >= 0 ? tile_ptr(stage, i_in1) + tid : so_t;
        const float* s2_t = i_in2 >= 0 ? tile_ptr(stage, i_in2) + tid : so_t;
        int painted = 0;
        uint32_t painted_mask = 0;  // kWriteOnly: the frames of this tile (T <= 16) this voice painted
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
            uint32_t dist = next_boundary - f;
            bool careful = false;  // HAS_RUN: this lane's module cannot promise even one branch-free sample
            if constexpr (HasRunOf<M>::value && !WT) {
                if (active) {
                    const uint32_t s = RunCall<M>::limit(m);
                    careful = s == 0u;
                    dist = min(dist, careful ? 1u : s);
                }
            }
            const uint32_t run = min(n - j, __reduce_min_sync(0xffffffffu, dist));
            if (active) {
                float* so = so_t + j * RV;
                const float* s0 = s0_t + j * RV;
                const float* s1 = s1_t + j * RV;
                const float* s2 = s2_t + j * RV;
                if constexpr (WT) {
#pragma unroll 4
                    for (uint32_t r = 0; r < run; r++) {
                        const float x0 = i_in0 >= 0 ? s0[r * RV] : 0.0f;
                        const float x1 = i_in1 >= 0 ? s1[r * RV] : 0.0f;
                        const float x2 = i_in2 >= 0 ? s2[r * RV] : 0.0f;
                        if constexpr (KIND == 1) WarmCall<M>::sample(m, x0, x1, x2); else ScanCall<M>::sample(m, x0, x1, x2);
                    }
                } else if constexpr (ConstFillOf<M>::value) {
                    painted_mask |= (run >= 32u ? 0xFFFFFFFFu : ((1u << run) - 1u)) << j;  // nothing else to do per frame
                } else if (HasRunOf<M>::value && !careful) {
                    // branch-free samples of a module with a run interface (no phase change of its own inside the run)
                    painted = 1;
                    if (kWriteOnly) painted_mask |= ((1u << run) - 1u) << j;
                    RunCall<M>::begin(m);
                    uint32_t r = 0;
                    for (; r + 4 <= run; r += 4) {
                        float acc[4], x0[4], x1[4], x2[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            acc[q] = kWriteOnly ? 0.0f : so[(r + q) * RV];
                            x0[q] = i_in0 >= 0 ? s0[(r + q) * RV] : 0.0f;
                            x1[q] = i_in1 >= 0 ? s1[(r + q) * RV] : 0.0f;
                            x2[q] = i_in2 >= 0 ? s2[(r + q) * RV] : 0.0f;
                        }
#pragma unroll
                        for (int q = 0; q < 4; q++) RunCall<M>::sample(m, acc[q], x0[q], x1[q], x2[q]);
#pragma unroll
                        for (int q = 0; q < 4; q++) so[(r + q) * RV] = acc[q];
                    }
                    for (; r < run; r++) {
                        float acc = kWriteOnly ? 0.0f : so[r * RV];
                        RunCall<M>::sample(m, acc, i_in0 >= 0 ? s0[r * RV] : 0.0f, i_in1 >= 0 ? s1[r * RV] : 0.0f, i_in2 >= 0 ? s2[r * RV] : 0.0f);
                        so[r * RV] = acc;
                    }
                    RunCall<M>::end(m);
                } else {
                    painted = 1;
                    if (kWriteOnly) painted_mask |= ((1u << run) - 1u) << j;
                    // batches of 4 frames: all operands are read before the first result is written, so the shared-memory
                    // loads of a batch overlap instead of queueing behind the previous sample's store (the compiler cannot
                    // know that the tiles do not alias); the module still sees the samples one by one, in order
                    uint32_t r = 0;
                    for (; r + 4 <= run; r += 4) {
                        float acc[4], x0[4], x1[4], x2[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            acc[q] = kWriteOnly ? 0.0f : so[(r + q) * RV];
                            x0[q] = i_in0 >= 0 ? s0[(r + q) * RV] : 0.0f;
                            x1[q] = i_in1 >= 0 ? s1[(r + q) * RV] : 0.0f;
                            x2[q] = i_in2 >= 0 ? s2[(r + q) * RV] : 0.0f;
                        }
                        Sample4<M>::call(m, acc, x0, x1, x2);
#pragma unroll
                        for (int q = 0; q < 4; q++) so[(r + q) * RV] = acc[q];
                    }
                    for (; r < run; r++) {
                        float acc = kWriteOnly ? 0.0f : so[r * RV];
                        const float x0 = i_in0 >= 0 ? s0[r * RV] : 0.0f;
                        const float x1 = i_in1 >= 0 ? s1[r * RV] : 0.0f;
                        const float x2 = i_in2 >= 0 ? s2[r * RV] : 0.0f;
                        m.sample(acc, x0, x1, x2);
                        so[r * RV] = acc;
                    }
                }
            }
            j += run;
        }
        if constexpr (kWriteOnly) {
            // the tile holds only what was painted: full rows when every voice of the CTA painted every frame of the tile,
            // otherwise every voice stores its own frames (one coalesced 128-byte row piece per warp and frame)
            const bool full = !valid || painted_mask == (n >= 32 ? 0xFFFFFFFFu : (1u << n) - 1u);
            if constexpr (ConstFillOf<M>::value) {
                const float cv = ConstFillCall<M>::value(m);
                if (__syncthreads_and(full)) {
                    if (c_valid) {
                        float* dstg = a.out + (size_t)f0 * V + v0 + c;
                        if (vec4) {
                            const float4 c4 = make_float4(cv, cv, cv, cv);
                            for (uint32_t jj = r0; jj < n; jj += rpp) *reinterpret_cast<float4*>(dstg + (size_t)jj * V) = c4;
                        } else {
                            for (uint32_t jj = r0; jj < n; jj += rpp) dstg[(size_t)jj * V] = cv;
                        }
                    }
                } else if (valid) {
                    float* dstv = a.out + (size_t)f0 * V + v;
                    for (uint32_t mk = painted_mask; mk != 0; mk &= mk - 1) dstv[(size_t)(__ffs(mk) - 1) * V] = cv;
                }
                return;
            }
            if (__syncthreads_and(full)) {
                if (c_valid) {
                    float* dstg = a.out + (size_t)f0 * V + v0 + c;
                    const float* st = tile_ptr(stage, i_out) + c;
                    if (vec4) {
                        for (uint32_t jj = r0; jj < n; jj += rpp) *reinterpret_cast<float4*>(dstg + (size_t)jj * V) = *reinterpret_cast<const float4*>(st + jj * RV);
                    } else {
                        for (uint32_t jj = r0; jj < n; jj += rpp) dstg[(size_t)jj * V] = st[jj * RV];
                    }
                }
            } else if (valid) {
                float* dstv = a.out + (size_t)f0 * V + v;
                for (uint32_t mk = painted_mask; mk != 0; mk &= mk - 1) {
                    const uint32_t jj = __ffs(mk) - 1;
                    dstv[(size_t)jj * V] = so_t[jj * RV];
                }
            }
            return;
        }
        // the finished output tile goes back to HBM as full rows (skipped when nobody in the CTA painted in it)
        if (__syncthreads_or(painted) && c_valid) {
            float* dstg = a.out + (size_t)f0 * V + v0 + c;
            const float* st = tile_ptr(stage, i_out) + c;
            if (vec4) {
                for (uint32_t jj = r0; jj < n; jj += rpp) *reinterpret_cast<float4*>(dstg + (size_t)jj * V) = *reinterpret_cast<const float4*>(st + jj * RV);
            } else {
                for (uint32_t jj = r0; jj < n; jj += rpp) dstg[(size_t)jj * V] = st[jj * RV];
            }
        }
    };

    if constexpr (kWarm) {
        if (warm && f_begin > 0) {
            // Tolerance mode.  The state at f_begin is approximated by running the module from zero state over the Wf frames
            // before the segment; the error is (homogeneous map of the window) x (true state at the window start), and the
            // module tracks that map (warm_ok).  Wf: the largest per-voice estimate in the CTA, from the constants of the span
            // at f_begin; no warm-up at all if that exceeds warm_cap (then the segment waits for its predecessor's end state).
            __shared__ uint32_t s_w;
            if (tid == 0) s_w = 0;
            __syncthreads();
            const uint32_t f_lo = f_begin > a.warm_cap ? f_begin - a.warm_cap : 0u;
            uint32_t kk = k;
            uint32_t p_start = s_start, p_end = s_end, p_flags = s_flags, p_param = s_param;
            uint32_t est = 0;
            pristine = s_start >= f_begin;  // nothing painted before the segment: the loaded state is the start state
            if (valid && !pristine) {
                // first span that reaches into [f_lo, ...): everything before it cannot matter
                while (p_end <= f_lo) {
                    kk++;
                    if (kk < ke) {
                        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + kk);
                        p_start = s.x; p_end = s.y; p_flags = s.z; p_param = s.w;
                    } else {
                        p_start = p_end = 0xFFFFFFFFu;
                    }
                }
                k = kk; s_start = p_start; s_end = p_end; s_flags = p_flags; s_param = p_param;
                // the span painting at f_begin (or the last one before it) provides the constants of the estimate
                uint32_t q_param = 0xFFFFFFFFu, q_flags = 0, q_len = 0, kq = kk;
                while (p_start < f_begin) {
                    q_param = p_param; q_flags = p_flags; q_len = p_end - p_start;
                    kq = kk;
                    kk++;
                    if (kk >= ke) break;
                    const uint4 s = *reinterpret_cast<const uint4*>(a.spans + kk);
                    p_start = s.x; p_end = s.y; p_flags = s.z; p_param = s.w;
                }
                est = 0xFFFFFFFFu;  // painted before, but not inside the longest window: only the chain knows the state
                if (q_param != 0xFFFFFFFFu) {
                    if constexpr (WarmExactOf<M>::value) {
                        est = WarmExactCall<M>::window(m, a, a.off[v], kq, f_begin, f_lo);
                    } else {
                        m.begin_span(a, q_param, q_flags, q_len);
                        est = m.warm_estimate(a.warm_cap);
                    }
                }
                atomicMax(&s_w, est);
            }
            __syncthreads();
            uint32_t Wf = s_w;
            if (Wf > a.warm_cap) Wf = 0;  // hopeless: chain
            Wf = min((Wf + T - 1) / T * T, f_begin / T * T);
            if (Wf == f_begin) Wf = 0;    // a window back to frame 0 is the sequential walk: chain instead
            const uint32_t f_w = f_begin - Wf;
            tile_first = f_w / T;
            if (a.warm_stats && tid == 0) {
                atomicAdd(a.warm_stats + 0, 1u);
                atomicAdd(a.warm_stats + 1, Wf);
                atomicMax(a.warm_stats + 2, Wf);
            }
            if (valid && !pristine) {
                m.warm_begin();
                while (s_end <= f_w) {
                    k++;
                    if (k < ke) {
                        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
                        s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
                    } else {
                        s_start = s_end = 0xFFFFFFFFu;
                    }
                }
                if (s_start < f_w) {  // entered in the middle: only the span's constants are needed
                    m.begin_span(a, s_param, s_flags, s_end - s_start);
                    active = true;
                }
            }
        }
    }
    if constexpr (kScan) {
        if (scan) {
            // pass 1: position the span cursor at the segment start, then sum this segment's inputs (inputs only)
            if (f_begin > 0 && valid) {
                while (s_end <= f_begin) {
                    k++;
                    if (k < ke) {
                        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
                        s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
                    } else {
                        s_start = s_end = 0xFFFFFFFFu;
                    }
                }
                if (s_start < f_begin) {
                    m.begin_span(a, s_param, s_flags, s_end - s_start);
                    active = true;
                }
            }
            const uint32_t k_0 = k, ss_0 = s_start, se_0 = s_end, sf_0 = s_flags, sp_0 = s_param;
            const bool act_0 = active;
            m.scan_begin();
            inputs_only = true;
            st_iss = st_cur = 0;
            for (uint32_t s = 0; s + 1 < S; s++) {
                if (tile0 + s < ntiles) issue(tile0 + s); else __pipeline_commit();
            }
            for (uint32_t tile = tile0; tile < ntiles; tile++) do_tile(tile, std::integral_constant<int, 2>{});
            __pipeline_wait_prior(0);
            __syncthreads();
            inputs_only = false;
            // look-back: the predecessor's inclusive prefix is this segment's start state (segment 0 keeps the loaded state)
            if (by > 0) {
                if (tid == 0) wait_predecessor(128);
                __syncthreads();
                if (valid) m.scan_load(a.chain + (size_t)(by - 1) * V + v);
            }
            if (f_end < a.frames) {
                if (valid) m.scan_store(a.chain + (size_t)by * V + v);
                __threadfence();
                __syncthreads();
                if (tid == 0) atomicExch(a.chain_flags + (size_t)by * gridDim.x + bx, a.epoch);
            }
            // pass 2 starts where pass 1 started
            k = k_0; s_start = ss_0; s_end = se_0; s_flags = sf_0; s_param = sp_0; active = act_0;
            if (active) m.begin_span(a, s_param, s_flags, s_end - s_start);
        }
    }
    // the ring is primed before the skip-ahead so that its latency hides behind it
    st_iss = st_cur = 0;
    for (uint32_t s = 0; s + 1 < S; s++) {
        if (tile_first + s < ntiles) issue(tile_first + s); else __pipeline_commit();
    }

    if (PrefixOf<M>::value && a.prefix != nullptr && f_begin > 0) {
        // the state at the segment start was tabulated by k_paint_prefix: move the span cursor, load the row, and re-enter
        // a straddling span as the second half of a split call
        if (valid) {
            while (s_end <= f_begin) {
                k++;
                if (k < ke) {
                    const uint4 s = *reinterpret_cast<const uint4*>(a.spans + k);
                    s_start = s.x; s_end = s.y; s_flags = s.z; s_param = s.w;
                } else {
                    s_start = s_end = 0xFFFFFFFFu;
                }
            }
            PaintArgs row = a;
            row.state = (char*)a.prefix + (size_t)by * a.V * a.prefix_stride;
            m.load(row, v);
            if (s_start < f_begin) {
                m.begin_span(a, s_param, s_flags & ~(uint32_t)ZCU_SPAN_NOTE_ID_CHANGED, s_end - s_start);
                active = true;
            }
        }
    } else if (M::SKIP && f_begin > 0) {
        if constexpr (ResetsOf<M>::value) {
            // A module whose whole state is reset by a call with a new note id (Sampler: t = 0) need not be walked through
            // anything before the last such call ahead of the segment: the span headers are scanned (no module work) and the
            // walk starts there.  Every lane then does its one walk from a fresh note in the same trip of the loop below,
            // instead of the warp paying such a walk at nearly every span because some lane's note starts there.
            uint32_t kr = k;
            for (uint32_t kk = k; kk < ke; kk++) {
                const uint4 sp = *reinterpret_cast<const uint4*>(a.spans + kk);
                if (sp.x >= f_begin) break;
                if ((sp.z & ZCU_SPAN_NOTE_ID_CHANGED) && m.resets(a, sp.w)) kr = kk;
            }
            if (kr != k) {
                k = kr;
                const uint4 sp = *reinterpret_cast<const uint4*>(a.spans + k);
                s_start = sp.x; s_end = sp.y; s_flags = sp.z; s_param = sp.w;
            }
        }
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
        LazyCall<M>::done(m);
    }

    if constexpr (kWarm) {
        if (warm && f_begin > 0) {
            for (uint32_t tile = tile_first; tile < tile0; tile++) do_tile(tile, std::integral_constant<int, 1>{});
            // end of the warm-up: keep the warmed state if every voice of the CTA verified its window, otherwise wait for the
            // end state the previous segment publishes (segments are handed out by ticket, so the predecessor has started:
            // no deadlock whatever the dispatch order; an undamped filter degenerates to the sequential walk)
            const bool ok = !valid || pristine || (tile_first < tile0 && m.warm_ok());
            if (!__syncthreads_and(ok)) {
                if (a.warm_stats && tid == 0) atomicAdd(a.warm_stats + 3, 1u);
                if (tid == 0) wait_predecessor(256);
                __syncthreads();
                if (valid && !pristine) m.chain_load(a.chain + (size_t)(by - 1) * V + v);
            }
        }
    }
    for (uint32_t tile = tile0; tile < ntiles; tile++) do_tile(tile, std::integral_constant<int, 0>{});

    if constexpr (kWarm) {
        if (warm && f_end < a.frames) {  // publish the end state for a successor that could not verify its warm-up
            if (valid) m.chain_store(a.chain + (size_t)by * V + v);
            __threadfence();
            __syncthreads();
            if (tid == 0) atomicExch(a.chain_flags + (size_t)by * gridDim.x + bx, a.epoch);
        }
    }
    if (f_end < a.frames) return;  // not the last segment: its state is somebody else's start
    if (!valid) return;
    if (active) m.end_span(a);
    m.store(a, v);
}

// launch geometry shared by the module and span-restricted basics launchers
struct PaintPlan {
    uint32_t T, nseg, seg_frames, vec4, grid_x, stages;
    size_t smem;
};

constexpr int kPaintMaxSmem = 200 * 1024;

inline PaintPlan zcu_paint_plan(const zcu_ctx* ctx, const PaintArgs& a, int nb, bool skippable, uint32_t max_resident = 8) {
    PaintPlan p;
    zcu_paint_segments(ctx, a.V, a.frames, nb, skippable, &p.T, &p.nseg, &p.seg_frames, max_resident);
    auto al16 = [](const void* q) { return ((uintptr_t)q & 15u) == 0; };
    p.vec4 = (a.V % 4 == 0 && al16(a.out) && al16(a.in0) && al16(a.in1) && al16(a.in2)) ? 1u : 0u;
    p.grid_x = (a.V + 32 * kPaintWarps - 1) / (32 * kPaintWarps);
    // ring depth: when the launch puts only a few CTAs on an SM (banks that cannot be cut in time), each CTA gets a deeper
    // ring of longer tiles so that enough bytes are in flight per SM to cover the HBM latency
    const uint32_t ctas = p.grid_x * p.nseg, sms = (uint32_t)ctx->sm_count;
    const uint32_t per_sm = std::min(8u, std::max(1u, (ctas + sms - 1) / sms));
    const size_t budget = (size_t)kPaintMaxSmem / per_sm;
    const size_t row = (size_t)nb * 32 * kPaintWarps * sizeof(float);
    const char* deep_env = getenv("ZCU_PAINT_DEEP");  // tuning: 0 = the fixed three-stage ring
    const bool deep = !(deep_env && atoi(deep_env) == 0);
    if (deep && !getenv("ZCU_PAINT_T") && p.T < 16 && row * 16 * kPaintStages <= budget) {
        p.T = 16;
        p.seg_frames = (p.seg_frames + 15) / 16 * 16;
        p.nseg = (a.frames + p.seg_frames - 1) / p.seg_frames;
    }
    p.stages = deep ? (uint32_t)std::min<size_t>(8, std::max<size_t>(kPaintStages, budget / (row * p.T))) : (uint32_t)kPaintStages;
    if (const char* e = getenv("ZCU_PAINT_STAGES")) p.stages = (uint32_t)std::min(8, std::max(2, atoi(e)));  // tuning
    p.smem = (size_t)p.stages * row * p.T;
    return p;
}

}  // namespace zcu
