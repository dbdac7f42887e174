// zcu_graph_rt.cuh — device-side runtime of the fused module-graph kernels that zcu_graph.cu generates and compiles with
// NVRTC (this header, zcu_modules.cuh, zcu_exact_accum.cuh and zangcu.h are embedded in libzangcu.so as text).
//
// A graph is the reference's instruction list of a script module (src/zangscript/codegen.zig:60-121, executed buffer by
// buffer by src/zangscript/runtime.zig:445-740) or, equally, the body of a hand-written instrument's paint()
// (examples/modules.zig): basics ops, direct arithmetic and module paint() calls over temps.  The reference runs every
// instruction over the whole span before the next one; no instruction reads any sample of a temp but the one it is about
// to write, so running the whole list sample by sample leaves the same bits, and the temps become registers.  One thread
// per voice, 32 adjacent voices per warp; every instruction names one of the graph's span tables (a table = the calls of
// one reference-side loop: the Trigger's spans, or whole mix buffers) and is live on the frames its table's spans cover.
#pragma once

#include "zcu_modules.cuh"

namespace zcu {

constexpr int kGraphMaxTables = 4;
constexpr int kGraphMaxFields = 12;
constexpr int kGraphMaxBufs = 4;  // external outputs / inputs each

struct GraphTable {
    const uint32_t* off;
    const zcu_span* spans;
};

struct GraphArgs {
    uint32_t V, frames;
    GraphTable tab[kGraphMaxTables];
    float* out[kGraphMaxBufs];        // ZCU_GRAPH_OUT_BUFFERS: [frames][V], accumulated like the reference's outputs
    const float* in[kGraphMaxBufs];   // external [frames][V] inputs
    // ZCU_GRAPH_OUT_MIX: output 0 of every voice x its pan gains, summed per warp into partial[warp][frames][C]
    float* partial;
    const float* gains;  // [V][C]
    uint32_t C;
    uint32_t nfields;
    // time segments (graphs whose modules all have an exact skip-ahead): blockIdx.y = segment of seg_frames frames; a segment
    // skips ahead to `replay` frames before its start, runs those frames silently (a Decimator re-latches its held value
    // there) and then paints.  err is set when a segment could not establish its start state.
    uint32_t seg_frames, replay;
    uint32_t* err;
    PaintArgs pa[kGraphMaxFields];
};

// rows of an external [frames][V] buffer this many frames ahead are pulled into L1 before they are needed
constexpr uint32_t kGraphPrefetchRows = 16;
__device__ __forceinline__ void prefetch_rows4(const float* p, size_t V) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p + V));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p + 2 * V));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p + 3 * V));
}

// cursor over one voice's spans of one table
struct SpanCursor {
    uint32_t k, ke, s_start, s_end, s_flags, s_param;
    bool active;
    __device__ __forceinline__ void fetch(const GraphTable& t) {
        if (k < ke) {
            const uint4 s = *reinterpret_cast<const uint4*>(t.spans + k);
            s_start = s.x;
            s_end = s.y;
            s_flags = s.z;
            s_param = s.w;
        } else {
            s_start = s_end = 0xFFFFFFFFu;
        }
    }
    __device__ __forceinline__ void init(const GraphTable& t, uint32_t v, bool valid) {
        k = ke = 0;
        active = false;
        if (valid) {
            k = t.off[v];
            ke = t.off[v + 1];
        }
        fetch(t);
    }
    __device__ __forceinline__ void next(const GraphTable& t) {
        active = false;
        k++;
        fetch(t);
    }
    __device__ __forceinline__ uint32_t boundary() const { return active ? s_end : s_start; }
};

// zang.Delay(N) inside a graph (src/zang/delay.zig:7-91; the script runtime's `delay` block, runtime.zig:704-721).  The
// reference serves the line in chunks of at most N frames: add the delayed samples (readDelayBuffer), run the inner
// instructions, store the new samples in the same slots (writeDelayBuffer).  A chunk only reads slots none of its own writes
// has touched, so per sample "read slot, ..., write slot, advance" leaves the same values in the same slots.
struct GraphDelayMod {
    DelayLine d;
    float* slot;
    __device__ void load(const PaintArgs& a, uint32_t v) {
        d.bind(a, v);
        d.index = ((const zcu_delay_state*)a.state)[v].index;
        slot = d.col;
    }
    __device__ void store(const PaintArgs& a, uint32_t v) { ((zcu_delay_state*)a.state)[v].index = d.index; }
    __device__ void begin_span(const PaintArgs&, uint32_t, uint32_t, uint32_t) {}
    __device__ void end_span(const PaintArgs&) {}
    __device__ __forceinline__ void read(float& acc) {  // zero(temp); readDelayBuffer(temp)
        slot = d.col + (size_t)d.index * d.V;
        acc = 0.0f + *slot;
    }
    __device__ __forceinline__ void write(float x) {  // writeDelayBuffer
        *slot = x;
        d.index = d.index + 1 == d.n ? 0u : d.index + 1;
    }
};

// ---- voice mix of a warp's tile (shared with k_nice_seg_render) ----------------------------------------------------
// tile of a warp: RT frames x 32 voices, row stride 34 floats (rows 8-byte aligned, conflict-free both for the per-voice
// column writes and for the mix's 8-byte row reads)
constexpr int kMixTileStride = 34;
constexpr int kMixTileFrames = 64;

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

// dst[frame][c] = sum over the warp's 32 voices of gn[c][voice] * O[frame][voice] for the first tn frames of a 64-frame tile.
// Lane j sums frames j and j + 32.  Voices are taken in pairs: one LDS.64 reads two adjacent voices of a frame row, the pan
// gains of four voices come as one LDS.128 per channel, and each fma.rn.f32x2 does two voices of one (frame, channel):
// ~1.75 issue slots per voice-sample (even and odd voices accumulate separately and meet at the end; the voice summation
// order is not the reference's anyway: tolerance 1e-5 x sum|gain|).  The caller brackets the call with __syncwarp().
template <int C>
__device__ __forceinline__ void mix_tile_store(const float (*O)[kMixTileStride], const float (*gn)[32], uint32_t tn, float* dst, uint32_t lane) {
    if (C == 2) {
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
        if (lane < tn) *reinterpret_cast<float2*>(dst + (size_t)lane * 2) = make_float2(sum_f32x2(a0), sum_f32x2(a1));
        if (lane + 32 < tn) *reinterpret_cast<float2*>(dst + (size_t)(lane + 32) * 2) = make_float2(sum_f32x2(b0), sum_f32x2(b1));
    } else {
        for (uint32_t fr = lane; fr < tn; fr += 32) {
            float acc = 0.0f;
#pragma unroll 8
            for (int vv = 0; vv < 32; vv++) acc = fmaf(gn[0][vv], O[fr][vv], acc);
            dst[fr] = acc;
        }
    }
}

// direct arithmetic of the script runtime (src/zangscript/runtime.zig:528-700) that is not one of the basics
__device__ __forceinline__ float zs_min(float a, float b) { return a < b ? a : b; }  // std.math.min / max: plain comparisons
__device__ __forceinline__ float zs_max(float a, float b) { return a > b ? a : b; }

}  // namespace zcu
