// zcu_basics.cu — zang's buffer primitives (src/zang/basics.zig:12-78), the voice mix, zang.mixDown
// (src/zang/mixdown.zig), bank initialisation and the no-FMA self test.
#include <algorithm>
#include <cstdlib>

#include "zcu_internal.h"
#include "zcu_paint_kernel.cuh"

using namespace zcu;

namespace {

// ---- whole-buffer elementwise kernels: float4 over the flattened [frame][voice] array ----------
template <int OP>
__device__ __forceinline__ float basic_op(float d, float a, float b, float s) {
    switch (OP) {
        case 0: return 0.0f;
        case 1: return s;
        case 2: return a;
        case 3: return d + (a + b);
        case 4: return d + a;
        case 5: return d + (a + s);
        case 6: return d + s;
        case 7: return d + a * b;
        case 8: return d * a;
        case 9: return d + a * s;
        default: return d * s;
    }
}
template <int OP>
struct OpTraits {
    static constexpr bool reads_dest = !(OP == 0 || OP == 1 || OP == 2);
    static constexpr bool reads_a = (OP == 2 || OP == 3 || OP == 4 || OP == 5 || OP == 7 || OP == 8 || OP == 9);
    static constexpr bool reads_b = (OP == 3 || OP == 7);
};

template <int OP>
__global__ void __launch_bounds__(256) k_basic_full(float* __restrict__ dest, const float* __restrict__ a,
                                                    const float* __restrict__ b, float s, size_t n4, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 d = make_float4(0, 0, 0, 0), x = d, y = d;
        if (OpTraits<OP>::reads_dest) d = reinterpret_cast<const float4*>(dest)[i];
        if (OpTraits<OP>::reads_a) x = reinterpret_cast<const float4*>(a)[i];
        if (OpTraits<OP>::reads_b) y = reinterpret_cast<const float4*>(b)[i];
        float4 r;
        r.x = basic_op<OP>(d.x, x.x, y.x, s);
        r.y = basic_op<OP>(d.y, x.y, y.y, s);
        r.z = basic_op<OP>(d.z, x.z, y.z, s);
        r.w = basic_op<OP>(d.w, x.w, y.w, s);
        reinterpret_cast<float4*>(dest)[i] = r;
    }
    // tail (n not a multiple of 4)
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float d = OpTraits<OP>::reads_dest ? dest[i] : 0.0f;
        const float x = OpTraits<OP>::reads_a ? a[i] : 0.0f;
        const float y = OpTraits<OP>::reads_b ? b[i] : 0.0f;
        dest[i] = basic_op<OP>(d, x, y, s);
    }
}

template <int OP>
int launch_full(zcu_ctx* ctx, float* dest, const float* a, const float* b, float s, size_t n) {
    const bool aligned = (((uintptr_t)dest | (uintptr_t)a | (uintptr_t)b) & 15u) == 0;
    const size_t n4 = aligned ? n / 4 : 0;
    size_t blocks = (n4 + 255) / 256;
    const size_t cap = (size_t)ctx->sm_count * 8;  // 8 resident CTAs of 256 threads per SM, grid-stride beyond
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    k_basic_full<OP><<<(unsigned)blocks, 256, 0, ctx->stream>>>(dest, a, b, s, n4, n);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

// ---- span-restricted zero / set with one scalar: a dedicated fill kernel -------------------------------------------------
// The value stored does not depend on the frame, so nothing of the generic tile pipeline is needed (no staging, no runs, no warp
// reductions: measured 3.1-4.0 TB/s through k_paint_lockstep against 6.0 TB/s for the bare store pattern,
// tools/ubench_zero_pattern.cu).  A CTA owns 128 adjacent voices and a time segment and walks it in tiles of 32 frames: every
// thread derives the 32-bit mask of the tile's frames its voice's spans cover (its own span cursor, no communication), one
// __syncthreads_and decides whether the tile is completely covered by all 128 voices — then it leaves as full 512-byte rows of
// 16-byte stores — and otherwise every voice stores its own frames.
__global__ void __launch_bounds__(128) k_fill_spans(float* __restrict__ out, const uint32_t* __restrict__ off, const zcu_span* __restrict__ spans,
                                                    uint32_t V, uint32_t frames, uint32_t seg_frames, float value, uint32_t vec4) {
    const uint32_t tid = threadIdx.x, v0 = blockIdx.x * 128u, v = v0 + tid;
    const bool valid = v < V;
    const uint32_t f_begin = blockIdx.y * seg_frames;
    if (f_begin >= frames) return;
    const uint32_t f_end = min(frames, f_begin + seg_frames);
    uint32_t k = 0, ke = 0, s_start = 0xFFFFFFFFu, s_end = 0xFFFFFFFFu;
    auto fetch = [&]() {
        if (k < ke) {
            const uint4 s = *reinterpret_cast<const uint4*>(spans + k);
            s_start = s.x;
            s_end = s.y;
        } else {
            s_start = s_end = 0xFFFFFFFFu;
        }
    };
    if (valid) {
        k = off[v];
        ke = off[v + 1];
        // first span that ends after the segment start (spans are sorted and disjoint): binary search
        uint32_t lo = k, hi = ke;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (spans[mid].end <= f_begin) lo = mid + 1; else hi = mid;
        }
        k = lo;
    }
    fetch();
    const uint32_t c = (tid & 31u) * 4u, r0 = tid >> 5;  // row copies: 32 chunks of 16 bytes per 512-byte row, 4 rows per pass
    const bool c_valid = v0 + c < V;
    const float4 v4 = make_float4(value, value, value, value);
    for (uint32_t f0 = f_begin; f0 < f_end; f0 += 32) {
        const uint32_t n = min(32u, f_end - f0);
        const uint32_t want = n >= 32 ? 0xFFFFFFFFu : (1u << n) - 1u;
        uint32_t mask = 0;
        while (s_start < f0 + n) {  // spans that intersect [f0, f0 + n)
            const uint32_t a = max(s_start, f0) - f0, b = min(s_end, f0 + n) - f0;  // a < b <= n
            mask |= (b >= 32 ? 0xFFFFFFFFu : (1u << b) - 1u) & ~((1u << a) - 1u);
            if (s_end > f0 + n) break;  // reaches into the next tile
            k++;
            fetch();
        }
        const bool full = !valid || mask == want;
        if (__syncthreads_and(full)) {
            if (c_valid) {
                float* dst = out + (size_t)f0 * V + v0 + c;
                if (vec4) {
#pragma unroll 4
                    for (uint32_t j = r0; j < n; j += 4) *reinterpret_cast<float4*>(dst + (size_t)j * V) = v4;
                } else {
                    for (uint32_t j = r0; j < n; j += 4)
                        for (uint32_t q = 0; q < 4 && v0 + c + q < V; q++) dst[(size_t)j * V + q] = value;
                }
            }
        } else if (valid) {
            float* dst = out + (size_t)f0 * V + v;
            for (uint32_t m = mask; m != 0; m &= m - 1) dst[(size_t)(__ffs(m) - 1) * V] = value;
        }
    }
}

int launch_fill_spans(zcu_ctx* ctx, const zcu_spans* sp, float* dest, float value) {
    if (sp->voices == 0 || sp->frames == 0) return ZCU_OK;
    const uint32_t grid_x = (sp->voices + 127) / 128;
    // as many time segments as fill 8 resident CTAs per SM once, each a multiple of the 32-frame tile and at least 256 frames
    uint32_t nseg = std::max(1u, 8u * (uint32_t)ctx->sm_count / grid_x);
    uint32_t seg = (sp->frames + nseg - 1) / nseg;
    seg = std::max(seg, std::min(256u, sp->frames));
    seg = (seg + 31) / 32 * 32;
    nseg = (sp->frames + seg - 1) / seg;
    const uint32_t vec4 = (sp->voices % 4 == 0 && ((uintptr_t)dest & 15u) == 0) ? 1u : 0u;
    k_fill_spans<<<dim3(grid_x, nseg), 128, 0, ctx->stream>>>(dest, sp->d_off, sp->d_spans, sp->voices, sp->frames, seg, value, vec4);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

template <int OP, bool ROWS = false>
int launch_spans(zcu_ctx* ctx, const zcu_spans* sp, float* dest, const float* a, const float* b, float s, const float* rows = nullptr,
                 uint32_t nrows = 0) {
    if (!ROWS && (OP == 0 || OP == 1) && !getenv("ZCU_FILL_GENERIC")) return launch_fill_spans(ctx, sp, dest, OP == 0 ? 0.0f : s);
    void* dscalar = nullptr;
    int rc = ROWS ? zcu_arena_upload(ctx, 0, rows, (size_t)nrows * sizeof(float), &dscalar) : zcu_arena_upload(ctx, 0, &s, sizeof(float), &dscalar);
    if (rc) return rc;
    PaintArgs pa{};
    pa.V = sp->voices;
    pa.frames = sp->frames;
    pa.off = sp->d_off;
    pa.spans = sp->d_spans;
    pa.params = dscalar;
    pa.out = dest;
    pa.in0 = OpTraits<OP>::reads_a ? a : nullptr;
    pa.in1 = OpTraits<OP>::reads_b ? b : nullptr;
    if (pa.V == 0 || pa.frames == 0) return ZCU_OK;
    int nb = 1 + (pa.in0 ? 1 : 0) + (pa.in1 ? 1 : 0);
    // the basics are pure memory traffic: four CTAs per SM, each with a deep ring of long tiles (zcu_paint_plan), move more bytes
    // than eight with three short stages (4096 voices x 65536 frames: multiply 0.83 -> 0.76 ms, multiplyWithScalar 0.45 -> 0.42)
    const PaintPlan p = zcu_paint_plan(ctx, pa, nb, true, 4u);
    if (!ctx->attr_done.count((const void*)k_paint_lockstep<BasicMod<OP, ROWS>>)) {  // per device: tracked per context
        ZCU_CUDA(ctx, cudaFuncSetAttribute(k_paint_lockstep<BasicMod<OP, ROWS>>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPaintMaxSmem));
        ctx->attr_done.insert((const void*)k_paint_lockstep<BasicMod<OP, ROWS>>);
    }
    k_paint_lockstep<BasicMod<OP, ROWS>><<<dim3(p.grid_x, p.nseg), 32 * kPaintWarps, p.smem, ctx->stream>>>(pa, p.T, p.seg_frames, p.vec4, p.stages);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

template <int OP>
int launch_op(zcu_ctx* ctx, const zcu_spans* sp, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b, float s) {
    const float* pa = a ? a->d : nullptr;
    const float* pb = b ? b->d : nullptr;
    if (sp) return launch_spans<OP>(ctx, sp, dest->d, pa, pb, s);
    return launch_full<OP>(ctx, dest->d, pa, pb, s, (size_t)dest->frames * dest->voices);
}

// ---- voice mix: dst[f][c] += sum_v gains[v][c] * src[f][v] -------------------------------------
// One CTA per frame; each thread strides over voices with coalesced float loads, then a fixed-order
// tree (warp shuffles, then one warp over the per-warp partials) so the result is deterministic.
template <int C>
__global__ void __launch_bounds__(256) k_mix_reduce(const float* __restrict__ src, const float* __restrict__ gains,
                                                    float* __restrict__ dst, uint32_t V) {
    const uint32_t f = blockIdx.x;
    const float* row = src + (size_t)f * V;
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; c++) acc[c] = 0.0f;
    for (uint32_t v = threadIdx.x; v < V; v += blockDim.x) {
        const float x = row[v];
#pragma unroll
        for (int c = 0; c < C; c++) acc[c] += (gains ? gains[(size_t)v * C + c] : 1.0f) * x;
    }
    __shared__ float part[8][C];
#pragma unroll
    for (int c = 0; c < C; c++) {
        float s = acc[c];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5][c] = s;
    }
    __syncthreads();
    if (threadIdx.x < C) {
        float s = 0.0f;
        for (int w = 0; w < 8; w++) s += part[w][threadIdx.x];
        dst[(size_t)f * C + threadIdx.x] += s;
    }
}

// ---- ordered voice mix: the reference's sequential accumulation into the shared output ----------------
// One thread per frame walks the voices in order; a CTA stages its 128 frames x V tile through shared
// memory so that the HBM reads are coalesced whatever V is.
template <int C>
__global__ void __launch_bounds__(128) k_mix_ordered(const float* __restrict__ src, const float* __restrict__ gains,
                                                     float* __restrict__ dst, uint32_t frames, uint32_t V) {
    extern __shared__ float tile[];  // [128][VT+1]
    constexpr uint32_t VT = 32;
    const uint32_t f0 = blockIdx.x * 128;
    const uint32_t nf = min(128u, frames - f0);
    const uint32_t f = f0 + threadIdx.x;
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; c++) acc[c] = (threadIdx.x < nf) ? dst[(size_t)f * C + c] : 0.0f;
    for (uint32_t v0 = 0; v0 < V; v0 += VT) {
        const uint32_t nv = min(VT, V - v0);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < nf * nv; i += 128) {
            const uint32_t r = i / nv, c = i - r * nv;
            tile[r * (VT + 1) + c] = src[(size_t)(f0 + r) * V + v0 + c];
        }
        __syncthreads();
        if (threadIdx.x < nf) {
            for (uint32_t c = 0; c < nv; c++) {
                const float x = tile[threadIdx.x * (VT + 1) + c];
#pragma unroll
                for (int ch = 0; ch < C; ch++) {
                    const float gx = gains ? gains[(size_t)(v0 + c) * C + ch] * x : x;  // rounded product, then rounded sum (-fmad=false)
                    acc[ch] = acc[ch] + gx;
                }
            }
        }
    }
    if (threadIdx.x < nf) {
#pragma unroll
        for (int c = 0; c < C; c++) dst[(size_t)f * C + c] = acc[c];
    }
}

// ---- zang.mixDown (mixdown.zig:28-86) ---------------------------------------------------------------
__global__ void k_mix_down_s16(const float* __restrict__ mix, uint8_t* __restrict__ dst, size_t n, float mul) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float value = mix[i] * mul;
    int cv;
    if (value <= -32767.0f)
        cv = -32767;
    else if (value >= 32766.0f)
        cv = 32766;
    else if (value != value)
        cv = 0;
    else
        cv = (int)value;  // @intFromFloat: truncation
    dst[i * 2 + 0] = (uint8_t)(cv & 0xFF);
    dst[i * 2 + 1] = (uint8_t)((cv >> 8) & 0xFF);
}
__global__ void k_mix_down_s8(const float* __restrict__ mix, uint8_t* __restrict__ dst, size_t n, float mul) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float value = mix[i] * mul;
    int cv;
    if (value <= -127.0f)
        cv = -127;
    else if (value >= 126.0f)
        cv = 126;
    else if (value != value)
        cv = 0;
    else
        cv = (int)value;
    dst[i] = (uint8_t)(int8_t)cv;
}

// ---- bank init: <Module>.init() for every voice ------------------------------------------------------
__global__ void k_bank_init(int kind, void* state, uint32_t V, uint32_t first_seed) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    switch (kind) {
        case ZCU_MOD_SINE_OSC: ((zcu_sine_state*)state)[v].t = 0.0f; break;
        case ZCU_MOD_PULSE_OSC: ((zcu_pulse_state*)state)[v].cnt = 0; break;
        case ZCU_MOD_TRISAW_OSC: {
            zcu_trisaw_state s;
            s.cnt = 0;
            s.t = 0.0f;
            ((zcu_trisaw_state*)state)[v] = s;
        } break;
        case ZCU_MOD_NOISE: {  // Noise.zig:25-32, seed = counter value
            XoshiroDev r;
            r.seed((uint64_t)(first_seed + v));
            zcu_noise_state s;
            s.s[0] = r.s0;
            s.s[1] = r.s1;
            s.s[2] = r.s2;
            s.s[3] = r.s3;
            ((zcu_noise_state*)state)[v] = s;
        } break;
        case ZCU_MOD_ENVELOPE: {
            zcu_envelope_state s;
            s.state = ZCU_ENV_IDLE;
            s.painter.t = s.painter.last_value = s.painter.start = 0.0f;
            ((zcu_envelope_state*)state)[v] = s;
        } break;
        case ZCU_MOD_FILTER: {
            zcu_filter_state s;
            s.l = s.b = 0.0f;
            ((zcu_filter_state*)state)[v] = s;
        } break;
        case ZCU_MOD_PORTAMENTO: {
            zcu_portamento_state s;
            s.painter.t = s.painter.last_value = s.painter.start = 0.0f;
            ((zcu_portamento_state*)state)[v] = s;
        } break;
        case ZCU_MOD_CURVE: {
            zcu_curve_state s;
            s.t = 0.0f;
            s.current_song_note = 0;
            s.current_song_note_offset = 0;
            s.next_song_note = 0;
            ((zcu_curve_state*)state)[v] = s;
        } break;
        case ZCU_MOD_SAMPLER: ((zcu_sampler_state*)state)[v].t = 0.0f; break;
        case ZCU_MOD_DECIMATOR: {
            zcu_decimator_state s;
            s.dval = 0.0f;
            s.dcount = 1.0f;
            ((zcu_decimator_state*)state)[v] = s;
        } break;
        case ZCU_MOD_CYCLE: ((zcu_cycle_state*)state)[v].t = 0.0f; break;
        case ZCU_MOD_FM_OSCILLATOR: {
            zcu_fm_oscillator_state s;
            s.t = s.feedback1 = s.feedback2 = 0.0f;
            ((zcu_fm_oscillator_state*)state)[v] = s;
        } break;
        case ZCU_MOD_SIMPLE_DELAY: ((zcu_delay_state*)state)[v].index = 0; break;  // delay.zig:12-17
        case ZCU_MOD_FILTERED_ECHOES: {
            zcu_echoes_state s;
            s.index = 0;
            s.l = s.b = 0.0f;
            ((zcu_echoes_state*)state)[v] = s;
        } break;
        default: break;
    }
}

// every f32 numerator against one divisor: div_by_const (zcu_modules.cuh) must equal the IEEE quotient wherever it claims to
__global__ void k_div_probe(float d, unsigned long long* mismatches, unsigned long long* claimed) {
    const float rd = 1.0f / d;
    unsigned long long bad = 0, ok_n = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32); i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = __uint_as_float((uint32_t)i);
        if (!(fabsf(x) <= 3.0e38f)) continue;  // NaN / inf
        bool ok = true;
        const float q = div_by_const(x, d, rd, ok);
        if (!ok) continue;
        ok_n++;
        if (__float_as_uint(q) != __float_as_uint(x / d)) bad++;
    }
    if (bad) atomicAdd(mismatches, bad);
    atomicAdd(claimed, ok_n);
}

// every denominator utof23(ifreq) can produce (k * 2^-23, k = 1 .. 2^23-1) against one numerator
__global__ void k_div_numerator_probe(float num, unsigned long long* mismatches, unsigned long long* claimed) {
    unsigned long long bad = 0, ok_n = 0;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < (1u << 23); k += gridDim.x * blockDim.x) {
        const float b = utof23(k << 9);
        if (b == 0.0f) continue;
        bool ok = true;
        const float q = div_nobranch(num, b, ok);
        if (!ok) continue;
        ok_n++;
        if (__float_as_uint(q) != __float_as_uint(num / b)) bad++;
    }
    if (bad) atomicAdd(mismatches, bad);
    atomicAdd(claimed, ok_n);
}

// a*b+c with operands chosen so that a fused multiply-add gives a different f32 than mul-then-add
__global__ void k_fma_probe(float a, float b, float c, float* out) { out[0] = a * b + c; }

}  // namespace

int zcu_launch_basic(zcu_ctx* ctx, int op, const zcu_spans* sp, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b, float s) {
    switch (op) {
        case 0: return launch_op<0>(ctx, sp, dest, a, b, s);
        case 1: return launch_op<1>(ctx, sp, dest, a, b, s);
        case 2: return launch_op<2>(ctx, sp, dest, a, b, s);
        case 3: return launch_op<3>(ctx, sp, dest, a, b, s);
        case 4: return launch_op<4>(ctx, sp, dest, a, b, s);
        case 5: return launch_op<5>(ctx, sp, dest, a, b, s);
        case 6: return launch_op<6>(ctx, sp, dest, a, b, s);
        case 7: return launch_op<7>(ctx, sp, dest, a, b, s);
        case 8: return launch_op<8>(ctx, sp, dest, a, b, s);
        case 9: return launch_op<9>(ctx, sp, dest, a, b, s);
        case 10: return launch_op<10>(ctx, sp, dest, a, b, s);
        default: return zcu_fail(ctx, ZCU_ERR_INVALID, "unknown basic op %d", op);
    }
}

int zcu_launch_basic_rows(zcu_ctx* ctx, int op, const zcu_spans* sp, zcu_buf* dest, const zcu_buf* a, const float* rows, uint32_t nrows) {
    const float* pa = a ? a->d : nullptr;
    switch (op) {
        case 1: return launch_spans<1, true>(ctx, sp, dest->d, pa, nullptr, 0.0f, rows, nrows);
        case 5: return launch_spans<5, true>(ctx, sp, dest->d, pa, nullptr, 0.0f, rows, nrows);
        case 6: return launch_spans<6, true>(ctx, sp, dest->d, pa, nullptr, 0.0f, rows, nrows);
        case 9: return launch_spans<9, true>(ctx, sp, dest->d, pa, nullptr, 0.0f, rows, nrows);
        case 10: return launch_spans<10, true>(ctx, sp, dest->d, pa, nullptr, 0.0f, rows, nrows);
        default: return zcu_fail(ctx, ZCU_ERR_INVALID, "basic op %d takes no scalar", op);
    }
}

int zcu_launch_bank_init(zcu_ctx* ctx, zcu_bank* bank, uint32_t first_seed) {
    if (bank->voices == 0 || bank->stride == 0) return ZCU_OK;
    k_bank_init<<<(bank->voices + 127) / 128, 128, 0, ctx->stream>>>(bank->kind, bank->d, bank->voices, first_seed);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    if (bank->line) ZCU_CUDA(ctx, cudaMemsetAsync(bank->line, 0, (size_t)bank->delay_samples * bank->voices * sizeof(float), ctx->stream));
    return ZCU_OK;
}

extern "C" int zcu_selftest_div_by_const(zcu_ctx* ctx, float divisor, uint64_t* mismatches, uint64_t* claimed) {
    if (!ctx || !mismatches || !claimed) return ZCU_ERR_INVALID;
    unsigned long long* d = nullptr;
    int rc = zcu_arena_reserve(ctx, 0, 2 * sizeof(unsigned long long), (void**)&d);
    if (rc) return rc;
    ZCU_CUDA(ctx, cudaMemsetAsync(d, 0, 2 * sizeof(unsigned long long), ctx->stream));
    k_div_probe<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(divisor, d, d + 1);
    ctx->launches++;
    unsigned long long h[2] = {0, 0};
    ZCU_CUDA(ctx, cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *mismatches = h[0];
    *claimed = h[1];
    return ZCU_OK;
}

extern "C" int zcu_selftest_div_numerator(zcu_ctx* ctx, float numerator, uint64_t* mismatches, uint64_t* claimed) {
    if (!ctx || !mismatches || !claimed) return ZCU_ERR_INVALID;
    unsigned long long* d = nullptr;
    int rc = zcu_arena_reserve(ctx, 0, 2 * sizeof(unsigned long long), (void**)&d);
    if (rc) return rc;
    ZCU_CUDA(ctx, cudaMemsetAsync(d, 0, 2 * sizeof(unsigned long long), ctx->stream));
    k_div_numerator_probe<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(numerator, d, d + 1);
    ctx->launches++;
    unsigned long long h[2] = {0, 0};
    ZCU_CUDA(ctx, cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *mismatches = h[0];
    *claimed = h[1];
    return ZCU_OK;
}

int zcu_selftest_no_fma(zcu_ctx* ctx) {
    // a = b = 1 + 2^-12: a*b = 1 + 2^-11 + 2^-24, a tie that rounds to 1 + 2^-11 in f32;
    // c = -(1 + 2^-11): unfused -> 0, fused -> 2^-24.
    float* d = nullptr;
    int rc = zcu_arena_reserve(ctx, 0, sizeof(float), (void**)&d);
    if (rc) return rc;
    k_fma_probe<<<1, 1, 0, ctx->stream>>>(1.0f + 0x1p-12f, 1.0f + 0x1p-12f, -(1.0f + 0x1p-11f), d);
    ctx->launches++;
    float h = 1.0f;
    ZCU_CUDA(ctx, cudaMemcpyAsync(&h, d, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h != 0.0f) return zcu_fail(ctx, ZCU_ERR_UNSUPPORTED, "library was built with FMA contraction (need -fmad=false)");
    return ZCU_OK;
}

extern "C" int zcu_mix_reduce(zcu_ctx* ctx, const zcu_buf* src, const float* gains_host, uint32_t channels, zcu_buf* dst) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, src && dst, "null buffer");
    ZCU_REQUIRE(ctx, channels == 1 || channels == 2, "channels must be 1 or 2");
    ZCU_REQUIRE(ctx, dst->voices == channels && dst->frames == src->frames, "dst must be [frames][channels]");
    float* dg = nullptr;
    if (gains_host) {
        int rc = zcu_arena_upload(ctx, 1, gains_host, (size_t)src->voices * channels * sizeof(float), (void**)&dg);
        if (rc) return rc;
    }
    if (src->frames == 0) return ZCU_OK;
    if (channels == 1)
        k_mix_reduce<1><<<src->frames, 256, 0, ctx->stream>>>(src->d, dg, dst->d, src->voices);
    else
        k_mix_reduce<2><<<src->frames, 256, 0, ctx->stream>>>(src->d, dg, dst->d, src->voices);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

extern "C" int zcu_mix_ordered(zcu_ctx* ctx, const zcu_buf* src, const float* gains_host, uint32_t channels, zcu_buf* dst) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, src && dst, "null buffer");
    ZCU_REQUIRE(ctx, channels == 1 || channels == 2, "channels must be 1 or 2");
    ZCU_REQUIRE(ctx, dst->voices == channels && dst->frames == src->frames, "dst must be [frames][channels]");
    float* dg = nullptr;
    if (gains_host) {
        int rc = zcu_arena_upload(ctx, 1, gains_host, (size_t)src->voices * channels * sizeof(float), (void**)&dg);
        if (rc) return rc;
    }
    if (src->frames == 0 || src->voices == 0) return ZCU_OK;
    const unsigned blocks = (src->frames + 127) / 128;
    const size_t smem = 128 * 33 * sizeof(float);
    if (channels == 1)
        k_mix_ordered<1><<<blocks, 128, smem, ctx->stream>>>(src->d, dg, dst->d, src->frames, src->voices);
    else
        k_mix_ordered<2><<<blocks, 128, smem, ctx->stream>>>(src->d, dg, dst->d, src->frames, src->voices);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

extern "C" int zcu_mix_down(zcu_ctx* ctx, const zcu_buf* mix, int audio_format, float vol, uint8_t* dst_host, size_t dst_len) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, mix && dst_host, "null argument");
    const size_t n = (size_t)mix->frames * mix->voices;
    const size_t bytes = audio_format == ZCU_AUDIO_SIGNED16_LSB ? n * 2 : n;
    ZCU_REQUIRE(ctx, audio_format == ZCU_AUDIO_SIGNED16_LSB || audio_format == ZCU_AUDIO_SIGNED8, "unknown audio format");
    ZCU_REQUIRE(ctx, dst_len == bytes, "dst_len must be frames*channels*bytes_per_sample (mixdown.zig:35,66)");
    if (n == 0) return ZCU_OK;
    uint8_t* d = nullptr;
    int rc = zcu_arena_reserve(ctx, 2, bytes, (void**)&d);
    if (rc) return rc;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (audio_format == ZCU_AUDIO_SIGNED16_LSB)
        k_mix_down_s16<<<blocks, 256, 0, ctx->stream>>>(mix->d, d, n, vol * 32767.0f);
    else
        k_mix_down_s8<<<blocks, 256, 0, ctx->stream>>>(mix->d, d, n, vol * 127.0f);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    ZCU_CUDA(ctx, cudaMemcpyAsync(dst_host, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
