// zcu_paint.cu — instantiation and launch of the lockstep paint kernel for every module variant.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <mutex>
#include <type_traits>
#include <vector>

#include "zcu_internal.h"
#include "zcu_paint_kernel.cuh"

using namespace zcu;

namespace {

// ZCU_PAINT_TOLERANCE: room for the end state of every (segment, voice) and a ready flag per (segment, CTA)
int ensure_chain(zcu_ctx* ctx, size_t states, size_t flags) {
    if (ctx->chain_bytes < states * sizeof(float2)) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->chain) cudaFree(ctx->chain);
        ctx->chain = nullptr;
        ctx->chain_bytes = 0;
        ZCU_CUDA(ctx, cudaMalloc(&ctx->chain, states * sizeof(float2)));
        ctx->chain_bytes = states * sizeof(float2);
    }
    if (ctx->chain_flag_count < flags) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (ctx->chain_flags) cudaFree(ctx->chain_flags);
        ctx->chain_flags = nullptr;
        ctx->chain_flag_count = 0;
        ZCU_CUDA(ctx, cudaMalloc(&ctx->chain_flags, flags * sizeof(uint32_t)));
        ZCU_CUDA(ctx, cudaMemsetAsync(ctx->chain_flags, 0, flags * sizeof(uint32_t), ctx->stream));
        ctx->chain_flag_count = flags;
        ctx->paint_epoch = 0;
    }
    if (!ctx->paint_ticket) {
        ZCU_CUDA(ctx, cudaMalloc(&ctx->paint_ticket, 8 * sizeof(uint32_t)));
        ZCU_CUDA(ctx, cudaMemsetAsync(ctx->paint_ticket, 0, 8 * sizeof(uint32_t), ctx->stream));
        ctx->paint_ticket_base = 0;
    }
    return ZCU_OK;
}

// ---- xoshiro256 jump matrices for the Noise skip-ahead ------------------------------------------------------------
// State transition of xoshiro256 (the "++" only scrambles the output), linear over GF(2).  A matrix is stored by columns:
// column j = image of the state with only bit j set, 4 x u64.
struct XMat { uint64_t c[256][4]; };
void xo_step(uint64_t s[4]) {
    const uint64_t t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = (s[3] << 45) | (s[3] >> 19);
}
void xm_apply(const XMat& m, const uint64_t x[4], uint64_t y[4]) {
    uint64_t o[4] = {0, 0, 0, 0};
    for (int j = 0; j < 256; j++)
        if ((x[j >> 6] >> (j & 63)) & 1)
            for (int q = 0; q < 4; q++) o[q] ^= m.c[j][q];
    for (int q = 0; q < 4; q++) y[q] = o[q];
}
void xm_mul(const XMat& a, const XMat& b, XMat& out) {  // out = a * b (b first)
    for (int j = 0; j < 256; j++) xm_apply(a, b.c[j], out.c[j]);
}
constexpr int kJumpLevels = 6, kJumpDigits = 15;
void build_xjump(std::vector<XMat>& tab) {
    constexpr int LEVELS = kJumpLevels, DIGITS = kJumpDigits;
    tab.resize((size_t)LEVELS * DIGITS);
    XMat p;  // T^(16^k)
    for (int j = 0; j < 256; j++) {
        uint64_t e[4] = {0, 0, 0, 0};
        e[j >> 6] = 1ull << (j & 63);
        xo_step(e);
        for (int q = 0; q < 4; q++) p.c[j][q] = e[q];
    }
    for (int k = 0; k < LEVELS; k++) {
        tab[(size_t)k * DIGITS] = p;
        for (int d = 2; d <= DIGITS; d++) xm_mul(p, tab[(size_t)k * DIGITS + d - 2], tab[(size_t)k * DIGITS + d - 1]);
        XMat next;
        xm_mul(p, tab[(size_t)k * DIGITS + DIGITS - 1], next);
        p = next;
    }
}
int ensure_xjump(zcu_ctx* ctx) {
    if (ctx->xjump) return ZCU_OK;
    std::vector<XMat> tab;
    build_xjump(tab);
    ZCU_CUDA(ctx, cudaMalloc(&ctx->xjump, tab.size() * sizeof(XMat)));
    ZCU_CUDA(ctx, cudaMemcpyAsync(ctx->xjump, tab.data(), tab.size() * sizeof(XMat), cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaMalloc((void**)&ctx->xjump_resync, sizeof(unsigned long long)));
    ZCU_CUDA(ctx, cudaMemsetAsync(ctx->xjump_resync, 0, sizeof(unsigned long long), ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // tab is a host temporary
    return ZCU_OK;
}

}  // namespace

// the jump tables, for other translation units (fused graphs with a Noise field in time segments)
int zcu_ensure_xjump(zcu_ctx* ctx) { return ensure_xjump(ctx); }

// Host-only check of the jump table (no device needed): state <- T^n state, exactly as NoiseMod::advance_done applies it.
extern "C" int zcu_xoshiro_jump_host(uint64_t state[4], uint32_t n) {
    if (!state || n >= (1u << (4 * kJumpLevels))) return ZCU_ERR_INVALID;
    static std::vector<XMat> tab;
    if (tab.empty()) build_xjump(tab);
    for (int k = 0; n != 0; k++, n >>= 4) {
        const uint32_t d = n & 15u;
        if (!d) continue;
        uint64_t y[4];
        xm_apply(tab[(size_t)k * kJumpDigits + (d - 1)], state, y);
        for (int q = 0; q < 4; q++) state[q] = y[q];
    }
    return ZCU_OK;
}

namespace {

bool overlaps(const float* a, const float* b, size_t n) { return a && b && a < b + n && b < a + n; }

// SimpleDelayParMod's second half: after the paint, line[(index0 + k) % N] = in[p_k] for the last min(N, P) painted frames p_k of
// the call (P = painted frames; the bank's index already stands at index0 + P).  One thread per (voice, chunk of 256 ordinals),
// walking the voice's spans backwards from the end of the block.
constexpr uint32_t kDelayUpdChunk = 256;
__global__ void __launch_bounds__(128) k_delay_line_update(PaintArgs a) {
    const uint32_t v = blockIdx.x * 128u + threadIdx.x;
    if (v >= a.V) return;
    const uint32_t k0 = a.off[v], ke = a.off[v + 1], n = a.delay_samples;
    const uint32_t first_back = blockIdx.y * kDelayUpdChunk;  // painted frames between the end of the call and this chunk
    // the span holding the chunk's newest frame
    uint32_t kk = ke, skip = first_back, f = 0, lo = 0;
    bool found = false;
    while (kk > k0) {
        kk--;
        const uint4 s = *reinterpret_cast<const uint4*>(a.spans + kk);
        const uint32_t len = s.y - s.x;
        if (skip < len) {
            f = s.y - 1 - skip;
            lo = s.x;
            found = true;
            break;
        }
        skip -= len;
    }
    if (!found) return;  // fewer painted frames than that
    const uint32_t index_new = ((const zcu_delay_state*)a.state)[v].index;
    uint32_t slot = index_new + n - 1 - first_back;  // first_back < n (grid), index_new < n
    slot -= slot >= n ? n : 0u;
    const uint32_t cnt = min(kDelayUpdChunk, n - first_back);
    const float* in_col = a.in0 + v;
    float* line_col = a.line + v;
    for (uint32_t i = 0; i < cnt; i++) {
        line_col[(size_t)slot * a.V] = __ldg(in_col + (size_t)f * a.V);
        slot = slot == 0 ? n - 1 : slot - 1;
        if (f == lo) {  // the previous painted frame is the last one of an earlier (non-empty) span
            bool more = false;
            while (kk > k0) {
                kk--;
                const uint4 s = *reinterpret_cast<const uint4*>(a.spans + kk);
                if (s.y > s.x) {
                    f = s.y - 1;
                    lo = s.x;
                    more = true;
                    break;
                }
            }
            if (!more) return;
        } else {
            f--;
        }
    }
}

template <class M>
int launch(zcu_ctx* ctx, PaintArgs a, int kind) {
    int nb = 1;
    if (M::NIN >= 1 && a.in0) nb++;
    if (M::NIN >= 2 && a.in1) nb++;
    if (M::NIN >= 3 && a.in2) nb++;
    // tolerance mode reads the inputs of the frames before a segment while the previous segment is still painting them:
    // only possible when no input aliases the output
    const size_t n = (size_t)a.V * a.frames;
    // delay modules: a line longer than any prefetch queue (S * T <= 8 * 16 slots) rides the cp.async ring, as one more
    // staged buffer of the launch geometry
    a.line_ring = (LineRingOf<M>::value && a.delay_samples >= 256 && !getenv("ZCU_PAINT_NO_LINE_RING")) ? 1u : 0u;
    const bool nb_line = a.line_ring != 0u;  // nb counts the line queue as a staged buffer
    if (nb_line) nb++;
    const bool warm = WarmOf<M>::value && ctx->paint_mode == ZCU_PAINT_TOLERANCE && !overlaps(a.out, a.in0, n) && !overlaps(a.out, a.in1, n) &&
                      !overlaps(a.out, a.in2, n);
    bool skip = M::SKIP;
    if (std::is_same<M, NoiseMod>::value) {
        // jumps are tabulated for fewer than 16^6 painted samples per call; longer blocks walk the frames in one segment
        skip = a.frames < (1u << 24);
        if (skip) {
            int rc = ensure_xjump(ctx);
            if (rc) return rc;
        }
        a.xjump = skip ? (const ulonglong2*)ctx->xjump : nullptr;
        a.xjump_resync = skip ? ctx->xjump_resync : nullptr;
    }
    // tolerance-mode segments pay a warm-up window each: half as many, twice as long (measured best at 4 CTAs per SM)
    // exact chained prefix sum (M::SCAN): no aliasing either, the scan pass reads inputs while earlier segments paint
    bool scan = ScanOf<M>::value && !getenv("ZCU_PAINT_NO_SCAN") && !overlaps(a.out, a.in0, n) && !overlaps(a.out, a.in1, n) &&
                      !overlaps(a.out, a.in2, n);
    uint32_t warm_resident = 4;
    if (const char* e = getenv("ZCU_PAINT_WARM_RESIDENT")) warm_resident = (uint32_t)std::min(8, std::max(1, atoi(e)));  // tuning
    // order-dependent f32 phase sums (M::PREFIX_RING): a walk of its own tabulates the state at the segment starts, if the
    // bank needs time segments to fill the machine at all
    bool ring = PrefixRingOf<M>::value && !getenv("ZCU_PAINT_NO_PREFIX") && !overlaps(a.out, a.in0, n);
    PaintPlan p = zcu_paint_plan(ctx, a, nb, skip || warm || scan || ring, warm ? warm_resident : 8u);
    if (ring && p.nseg < 12 && !getenv("ZCU_PAINT_SEGMENTS")) {  // measured: the walk pays off up to ~12000 voices (8192: 1.19 against 2.2 ms; 16384: 1.24 against 1.16)
        ring = false;
        p = zcu_paint_plan(ctx, a, nb, skip || warm || scan, 8u);
    }
    // Segments that cost a second pass over an input (the chained scan of M::SCAN reads its segment twice, a PREFIX module is
    // walked once more by k_paint_prefix) stop paying once the voices alone put ~2.5 CTAs on every SM.  Measured, 16384 frames,
    // segments / one segment: PulseOsc with a frequency buffer 32768 voices 2.22 / 2.13 ms, 49152 voices 3.26 / 2.68 ms,
    // 65536 x 8192: 2.16 / 1.67 ms; Decimator 1.40 / 1.71, 2.05 / 1.85, 1.33 / 1.23 ms.
    if ((scan || (PrefixOf<M>::value && !PrefixRingOf<M>::value)) && !warm && p.nseg > 1 && p.grid_x * 2u >= 5u * (uint32_t)ctx->sm_count &&
        !getenv("ZCU_PAINT_SEGMENTS")) {
        scan = false;
        skip = false;
        p = zcu_paint_plan(ctx, a, nb, false, 8u);
    }
    // A module that has to walk its frames in order, on a bank too small to fill the machine with 128-voice CTAs: the walker /
    // producer kernel (k_paint_walk).  One walker warp per CTA, so the bank is cut into as many CTAs as the SMs' four
    // schedulers can hold: 32, 16 or 8 voices each (narrower rows, idle lanes, but that many more walkers).  Delay modules
    // need a line longer than the prefetch queue.
    {
        const bool one_segment = !(skip || warm || scan || ring);
        const bool ovr_now = ctx->paint_output == ZCU_OUT_OVERWRITE && !WriteOnlyOf<M>::value;
        const uint32_t sms = (uint32_t)ctx->sm_count;
        uint32_t rw = 32;  // narrower CTAs (more walkers) measured no faster: see DESIGN.md §4.1
        if (const char* e = getenv("ZCU_PAINT_WALK_VOICES")) rw = (uint32_t)std::min(32, std::max(8, atoi(e))) & ~7u;  // tuning: 8, 16, 32
        if (rw == 24) rw = 16;
        const uint32_t ctas = (a.V + rw - 1) / rw;
        const bool line_ok = !LineRingOf<M>::value || a.delay_samples >= kWalkGroup * kWalkGroups + 64u;
        uint32_t max_ctas = 4 * sms;
        if (const char* e = getenv("ZCU_PAINT_WALK_MAX_CTAS")) max_ctas = (uint32_t)std::max(1, atoi(e));  // tuning
        if (one_segment && !WriteOnlyOf<M>::value && ctas <= max_ctas && line_ok && a.frames > 0 && !getenv("ZCU_PAINT_NO_WALK")) {
            a.line_ring = LineRingOf<M>::value ? 1u : 0u;
            a.chain = nullptr;
            a.prefix = nullptr;
            auto al16 = [](const void* q) { return ((uintptr_t)q & 15u) == 0; };
            const uint32_t v4 = (a.V % 4 == 0 && al16(a.out) && al16(a.in0) && al16(a.in1) && al16(a.in2)) ? 1u : 0u;
            const int nb_tiles = nb - (nb_line ? 1 : 0);  // output + inputs
            const size_t group = (size_t)(nb_tiles + (a.line_ring ? 1 : 0)) * kWalkGroup * rw * sizeof(float);
            const uint32_t per_sm = std::min(4u, (ctas + sms - 1) / sms);
            const uint32_t ng = (size_t)kWalkGroups * group * per_sm <= (size_t)kPaintMaxSmem ? kWalkGroups : kWalkGroups / 2;
            auto wk = ovr_now ? k_paint_walk<M, true> : k_paint_walk<M, false>;
            if (!ctx->attr_done.count((const void*)wk)) {
                ZCU_CUDA(ctx, cudaFuncSetAttribute(wk, cudaFuncAttributeMaxDynamicSharedMemorySize, kPaintMaxSmem));
                ctx->attr_done.insert((const void*)wk);
            }
            wk<<<ctas, kWalkThreads, ng * group, ctx->stream>>>(a, v4, rw, ng);
            ctx->launches++;
            ZCU_CUDA(ctx, cudaGetLastError());
            return ZCU_OK;
        }
    }
    a.seg_base = 0;
    a.chain = nullptr;
    a.chain_flags = nullptr;
    a.epoch = 0;
    a.warm_cap = 0;
    a.ticket = nullptr;
    a.ticket_base = 0;
    a.chain_err = nullptr;
    a.warm_stats = nullptr;
    if ((warm || scan) && p.nseg > 1) {
        int rc = ensure_chain(ctx, (size_t)p.nseg * a.V, (size_t)p.nseg * p.grid_x);
        if (rc) return rc;
        a.chain = (float2*)ctx->chain;
        a.chain_flags = ctx->chain_flags;
        a.epoch = ++ctx->paint_epoch;
        a.warm_cap = (uint32_t)std::min<uint64_t>(a.frames, 8ull * p.seg_frames);  // a window longer than this: chain instead
        a.ticket = ctx->paint_ticket;
        a.ticket_base = ctx->paint_ticket_base;
        a.chain_err = ctx->paint_ticket + 1;
        a.warm_stats = warm ? ctx->paint_ticket + 2 : nullptr;
        ctx->paint_ticket_base += p.grid_x * p.nseg;  // one ticket per CTA of this launch (mod 2^32, like the device counter)
    } else if (!skip && !ring) {
        p.nseg = 1;
        p.seg_frames = a.frames;
    }
    a.prefix = nullptr;
    a.prefix_stride = 0;
    if (PrefixOf<M>::value && (ring || !PrefixRingOf<M>::value) && p.nseg > 1 && !getenv("ZCU_PAINT_NO_PREFIX")) {
        // one sequential walk per voice tabulates the state at every segment start (instead of a re-walk per segment)
        const size_t stride = zcu_bank_state_size(kind), need = (size_t)p.nseg * a.V * stride;
        if (ctx->prefix_bytes < need) {
            ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (ctx->prefix) cudaFree(ctx->prefix);
            ctx->prefix = nullptr;
            ctx->prefix_bytes = 0;
            ZCU_CUDA(ctx, cudaMalloc(&ctx->prefix, need));
            ctx->prefix_bytes = need;
        }
        a.prefix = ctx->prefix;
        a.prefix_stride = (uint32_t)stride;
        if constexpr (!PrefixRingOf<M>::value) {
            k_paint_prefix<M><<<(a.V + 127) / 128, 128, 0, ctx->stream>>>(a, p.seg_frames, p.nseg);
            ctx->launches++;
            ZCU_CUDA(ctx, cudaGetLastError());
        }
    }
    // ZCU_OUT_OVERWRITE: `zero; paint` in one call (modules that overwrite anyway have a single form)
    const bool ovr = ctx->paint_output == ZCU_OUT_OVERWRITE && !WriteOnlyOf<M>::value;
    auto kern = ovr ? k_paint_lockstep<M, true> : k_paint_lockstep<M, false>;
    if (!ctx->attr_done.count((const void*)kern)) {  // the attribute is per device: tracked per context
        ZCU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kPaintMaxSmem));
        ctx->attr_done.insert((const void*)kern);
    }
    a.seg_base = 0;
    if constexpr (PrefixRingOf<M>::value) {
        if (a.prefix != nullptr) {
            // The phase walk runs first, on the context's stream, then all segments are painted.  (Tried, r02ay: the walk in
            // up to four pieces on a second stream, each piece's segments painted while the next piece walks.  Correct, but
            // 1.45 -> 2.2 ms in two pieces and 3.0 ms in four for SineOsc at 4096 voices: a walker is one warp per SM that lives on
            // latency, and with eight paint CTAs beside it on the SM and the paint's traffic in front of its loads it crawls.)
            const bool v4 = a.V % 4 == 0 && ((uintptr_t)a.in0 & 15u) == 0;
            auto pk = k_paint_prefix_ring<M>;
            constexpr size_t ring_bytes = kRingRows * 32 * sizeof(float);
            if (!ctx->attr_done.count((const void*)pk)) {
                ZCU_CUDA(ctx, cudaFuncSetAttribute(pk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_bytes));
                ctx->attr_done.insert((const void*)pk);
            }
            uint32_t ahead = kRingAhead;
            if (const char* e = getenv("ZCU_RING_AHEAD")) ahead = (uint32_t)std::max(0, atoi(e));  // tuning
            pk<<<(a.V + 31) / 32, kRingThreads, ring_bytes, ctx->stream>>>(a, p.seg_frames, 0u, p.nseg - 1, v4 ? 1u : 0u, ahead);
            ctx->launches++;
            ZCU_CUDA(ctx, cudaGetLastError());
        }
    }
    kern<<<dim3(p.grid_x, p.nseg), 32 * kPaintWarps, p.smem, ctx->stream>>>(a, p.T, p.seg_frames, p.vec4, p.stages);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

}  // namespace

void zcu_paint_segments(const zcu_ctx* ctx, uint32_t voices, uint32_t frames, int nbufs, bool skippable, uint32_t* T_out, uint32_t* nseg_out,
                        uint32_t* seg_frames_out, uint32_t max_resident) {
    const uint32_t rv = 32u * kPaintWarps;
    const uint32_t grid = std::max(1u, (voices + rv - 1) / rv);
    const uint32_t sms = (uint32_t)ctx->sm_count;
    // tile depth: ~24 KB of staging per CTA for one or two buffers, 24-32 KB for three or four
    uint32_t T = nbufs <= 1 ? 16u : 8u;
    if (const char* e = getenv("ZCU_PAINT_T")) T = (uint32_t)std::min(16, std::max(1, atoi(e)));  // tuning
    uint32_t nseg = 1, seg_frames = frames;
    if (skippable && frames > 0) {
        // resident CTAs per SM: 8 by registers (128 threads x <= 64 registers), or what the staging ring leaves of ~220 KB
        const uint32_t per_cta = (uint32_t)kPaintStages * (uint32_t)nbufs * T * rv * 4u + 1024u;
        const uint32_t resident = std::min(max_resident, (220u * 1024u) / per_cta) * sms;
        // one full wave: a second, partly filled wave would cost as much as the first
        nseg = std::max(1u, resident / grid);
        if (const char* e = getenv("ZCU_PAINT_SEGMENTS")) nseg = (uint32_t)std::max(1, atoi(e));  // tuning / tests
        seg_frames = (frames + nseg - 1) / nseg;
        seg_frames = std::max(seg_frames, std::min(256u, frames));
        seg_frames = (seg_frames + T - 1) / T * T;
        nseg = (frames + seg_frames - 1) / seg_frames;
    }
    *T_out = T;
    *nseg_out = nseg;
    *seg_frames_out = seg_frames;
}

int zcu_envelope_ramp_tabs(zcu_ctx* ctx, const zcu_envelope_params* rows, uint32_t nrows, const void** dev, uint32_t* count) {
    static std::mutex mu;
    static std::map<uint32_t, RampTab> cache;  // a table depends on the bits of its step alone
    *dev = nullptr;
    *count = 0;
    uint32_t steps[4];
    uint32_t ns = 0;
    for (uint32_t r = 0; r < std::min(nrows, 64u); r++) {
        const zcu_painter_curve cv[3] = {rows[r].attack, rows[r].decay, rows[r].release};
        for (const zcu_painter_curve& c : cv) {
            if (c.kind == ZCU_CURVE_INSTANTANEOUS) continue;
            volatile float prod = c.duration * rows[r].sample_rate;  // PainterDev::try_ramp, two roundings
            const float step = 1.0f / prod;
            const uint32_t sb = ea_bits(step);
            if (!(step > 0.0f) || !std::isfinite(step)) continue;
            if (std::find(steps, steps + ns, sb) != steps + ns) continue;
            if (ns < 4) steps[ns++] = sb;
        }
    }
    if (ns == 0) return ZCU_OK;
    std::vector<RampTab> host(ns);
    {
        std::lock_guard<std::mutex> lock(mu);
        for (uint32_t i = 0; i < ns; i++) {
            auto it = cache.find(steps[i]);
            if (it == cache.end()) {
                if (cache.size() > 4096) cache.clear();
                RampTab t;
                t.build(ea_float(steps[i]));
                it = cache.emplace(steps[i], t).first;
            }
            host[i] = it->second;
        }
    }
    void* d = nullptr;
    int rc = zcu_arena_upload(ctx, 3, host.data(), ns * sizeof(RampTab), &d);
    if (rc) return rc;
    *dev = d;
    *count = ns;
    return ZCU_OK;
}

int zcu_launch_paint(zcu_ctx* ctx, const PaintLaunch& pl) {
    PaintArgs a;
    a.V = pl.V;
    a.frames = pl.frames;
    a.off = pl.off;
    a.spans = pl.spans;
    a.params = pl.params;
    a.ramp_tabs = (const RampTab*)pl.ramp_tabs;
    a.n_ramp_tabs = pl.n_ramp_tabs;
    a.state = pl.state;
    a.out = pl.out;
    a.in0 = pl.in0;
    a.in1 = pl.in1;
    a.in2 = pl.in2;
    a.nodes = (const zcu_curve_node*)pl.aux;
    a.smp.data = (const uint8_t*)pl.aux;
    a.smp.len = pl.aux_len;
    a.smp.num_channels = pl.smp_channels;
    a.smp.sample_rate = pl.smp_rate;
    a.smp.format = pl.smp_format;
    zcu_sample_desc_finish(a.smp);
    a.line = pl.line;
    a.delay_samples = pl.delay_samples;
    a.xjump = nullptr;
    a.xjump_resync = nullptr;
    if (a.V == 0 || a.frames == 0) return ZCU_OK;
    const bool b0 = pl.in0 != nullptr, b1 = pl.in1 != nullptr, b2 = pl.in2 != nullptr;
    switch (pl.kind) {
        case ZCU_MOD_SINE_OSC:
            if (b0 && b1) return launch<SineMod<true, true>>(ctx, a, pl.kind);
            if (b0) return launch<SineMod<true, false>>(ctx, a, pl.kind);
            if (b1) return launch<SineMod<false, true>>(ctx, a, pl.kind);
            return launch<SineMod<false, false>>(ctx, a, pl.kind);
        case ZCU_MOD_PULSE_OSC:
            return b0 ? launch<PulseMod<true>>(ctx, a, pl.kind) : launch<PulseMod<false>>(ctx, a, pl.kind);
        case ZCU_MOD_TRISAW_OSC:
            return b0 ? launch<TriSawMod<true>>(ctx, a, pl.kind) : launch<TriSawMod<false>>(ctx, a, pl.kind);
        case ZCU_MOD_NOISE: return launch<NoiseMod>(ctx, a, pl.kind);
        case ZCU_MOD_ENVELOPE: return launch<EnvelopeMod>(ctx, a, pl.kind);
        case ZCU_MOD_GATE: return launch<GateMod>(ctx, a, pl.kind);
        case ZCU_MOD_FILTER:
            if (b1 && b2) return launch<FilterMod<true, true>>(ctx, a, pl.kind);
            if (b1) return launch<FilterMod<true, false>>(ctx, a, pl.kind);
            if (b2) return launch<FilterMod<false, true>>(ctx, a, pl.kind);
            return launch<FilterMod<false, false>>(ctx, a, pl.kind);
        case ZCU_MOD_PORTAMENTO: return launch<PortamentoMod>(ctx, a, pl.kind);
        case ZCU_MOD_CURVE: return launch<CurveMod>(ctx, a, pl.kind);
        case ZCU_MOD_SAMPLER: return launch<SamplerMod>(ctx, a, pl.kind);
        case ZCU_MOD_DECIMATOR: return launch<DecimatorMod>(ctx, a, pl.kind);
        case ZCU_MOD_DISTORTION: return launch<DistortionMod>(ctx, a, pl.kind);
        case ZCU_MOD_FM_OSCILLATOR: return b1 ? launch<FmOscMod<true>>(ctx, a, pl.kind) : launch<FmOscMod<false>>(ctx, a, pl.kind);
        case ZCU_MOD_SIMPLE_DELAY: {
            // no feedback, so the call can be cut in time (SimpleDelayParMod) when the input is not painted over by the call
            // itself; the line is brought up to date by a second, small kernel
            const size_t nel = (size_t)a.V * a.frames;
            // (measured: ahead of the sequential walk up to ~8192 voices, 1.34 against 3.2 ms at 4096; behind it from 16384 on,
            // where the walkers fill the machine and the line update is pure extra traffic)
            const bool small_bank = (a.V + 31u) / 32u <= 2u * (uint32_t)ctx->sm_count;
            const char* force = getenv("ZCU_DELAY_PARALLEL");  // tuning: 0 / 1
            const bool par = force ? atoi(force) != 0 : small_bank;
            if (par && a.in0 && a.frames > 0 && a.delay_samples > 0 && !overlaps(a.out, a.in0, nel)) {
                int rc = launch<SimpleDelayParMod>(ctx, a, pl.kind);
                if (rc) return rc;
                k_delay_line_update<<<dim3((a.V + 127) / 128, (a.delay_samples + kDelayUpdChunk - 1) / kDelayUpdChunk), 128, 0, ctx->stream>>>(a);
                ctx->launches++;
                ZCU_CUDA(ctx, cudaGetLastError());
                return ZCU_OK;
            }
            return launch<SimpleDelayMod>(ctx, a, pl.kind);
        }
        case ZCU_MOD_FILTERED_ECHOES: return launch<FilteredEchoesMod>(ctx, a, pl.kind);
        case ZCU_MOD_CYCLE: return b0 ? launch<CycleMod<true>>(ctx, a, pl.kind) : launch<CycleMod<false>>(ctx, a, pl.kind);
        default: return zcu_fail(ctx, ZCU_ERR_INVALID, "unknown module kind %d", pl.kind);
    }
}
