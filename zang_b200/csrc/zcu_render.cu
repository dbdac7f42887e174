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
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include <thread>
#include <algorithm>
#include "zcu_internal.h"
#include "zcu_render_kernel.cuh"
#include "zcu_render_scan.cuh"
#include "zcu_render_seg.cuh"

using namespace zcu;

struct zcu_render {
    zcu_ctx* ctx;
    zcu_render_config cfg;
    uint32_t block_frames;
    uint32_t nwarps;
    float* d_color;        // [V]
    float* d_gains;        // [V][C]
    NiceVoiceState* d_state;
    uint32_t* d_event_off; // [V+1]
    DevEvent* d_events;    // per event, with Filter.cutoffFromFrequency(freq*8, sr) evaluated on the host
    DevEvent* h_events[2]; // pinned staging for the H2D copy, double buffered: the host expands the next job's events while
                           // the device still renders the previous one (only a buffer's own last copy is waited for)
    uint32_t* h_off[2];    // pinned copy of the caller's event_off (nothing of the caller's is read after the call returns)
    cudaEvent_t h_done[2]; // recorded after the copies out of h_events[i] / h_off[i]
    bool h_busy[2];
    int h_flip;
    // The device tables of a job (event_off, events, snaps, cursor) exist twice, and a job's upload and pre-pass run on a
    // stream of their own: zcu_render_set_events for job i+1 proceeds while zcu_render_run of job i still executes; the run
    // waits for its set's `ready`, the next upload into a set waits for the `released` of the last run that read it.
    struct TabSet {
        uint32_t* off;
        DevEvent* events;
        EnvSnap* snaps;
        uint32_t* cursor;
        cudaEvent_t ready, released;
        bool has_ready, has_released;
    } ts[2];
    int ts_cur;
    cudaStream_t up_stream;
    uint32_t nevents, cap_events;
    float* d_partial;      // [nwarps][block_frames][C]
    uint32_t debug_skip_mask;
    unsigned long long* d_debug_cycles;  // [nwarps][4][2], only with ZCU_RENDER_DEBUG_CYCLES=1
    bool profiling;
    std::vector<cudaEvent_t> ev;   // pairs around every render-kernel launch since the last query
    size_t ev_used;
    std::vector<cudaEvent_t> ev_aux;  // ZCU_RENDER_SCAN: pairs around every k_nice_scan_zs launch
    size_t ev_aux_used;
    // ZCU_RENDER_SCAN (zcu_render_scan.cuh)
    uint32_t nchunks_total;
    EnvSnap* d_snaps;      // per event
    uint32_t* d_cursor;    // [nchunks_total][V]
    float2* d_zs;          // [block chunks][V]
    Mat2* d_M;             // [block chunks][V]
    float2* d_sstart;      // [block chunks][V]
    double2* d_fstate;     // [V]
    // ZCU_RENDER_SEGMENTS (zcu_render_seg.cuh)
    float2* d_carry[2];    // [V] filter state at a launch boundary, double buffered
    uint32_t* d_warm_stats;  // [2] sum / max of warm-up buffers over the warp tasks of the last run
    uint32_t seg_slots;    // resident warps of k_nice_seg_render on this device
    uint32_t last_nseg, last_seg_chunks;
    uint32_t forced_nseg;  // zcu_render_set_segments: segments per launch (0 = as many as fill the resident warps)
    bool filter_contraction;  // zcu_render_set_filter_contraction
    bool attr_set[3];      // cudaFuncSetAttribute done for the scan / segment / exact kernels on this render's device
};

namespace {

// un-mixed per-voice output window of a run (parity tests): frames [first, first+frames) x voices [v0, v0+nv), or d == null
struct VoutReq {
    float* d;
    uint32_t first, frames, v0, nv;
};

// mix[F0+f][c] += sum over warps (fixed order) of partial[w][f][c]
__global__ void __launch_bounds__(256) k_partial_reduce(const float* __restrict__ partial, float* __restrict__ mix, uint32_t nwarps,
                                                        size_t n /* block_frames*C */) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (uint32_t w = 0; w < nwarps; w++) s += partial[(size_t)w * n + i];
    mix[i] += s;
}

}  // namespace

// mix[i] += sum over warps of partial[w][i] (fixed order); also used by the fused module graphs (zcu_graph.cu)
void zcu_launch_partial_reduce(zcu_ctx* ctx, const float* partial, float* mix, uint32_t nwarps, size_t n) {
    if (n == 0) return;
    k_partial_reduce<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(partial, mix, nwarps, n);
    ctx->launches++;
}

namespace {

__global__ void k_state_reset(NiceVoiceState* st, const uint32_t* event_off, uint32_t V) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    NiceVoiceState s{};
    s.env_state = ZCU_ENV_IDLE;
    s.next_event = event_off[v];
    st[v] = s;
}

int record_pair_begin(zcu_ctx* ctx, std::vector<cudaEvent_t>& ev, size_t used) {
    if (used + 2 > ev.size()) {
        ev.resize(used + 2, nullptr);
        ZCU_CUDA(ctx, cudaEventCreate(&ev[used]));
        ZCU_CUDA(ctx, cudaEventCreate(&ev[used + 1]));
    }
    ZCU_CUDA(ctx, cudaEventRecord(ev[used], ctx->stream));
    return ZCU_OK;
}

// ZCU_RENDER_SCAN: per block of chunks  zero-state pass -> per-voice chain -> render -> ordered partial sum
int render_blocks_scan(zcu_ctx* ctx, zcu_render* r, float* mix, const VoutReq& vo, uint32_t end_frame) {
    const zcu_render_config& c = r->cfg;
    ZCU_CUDA(ctx, cudaMemsetAsync(r->d_fstate, 0, (size_t)c.voices * sizeof(double2), ctx->stream));
    constexpr int RT = 64;
    const size_t smem = scan_render_smem_bytes<RT>();
    bool& attr_set = r->attr_set[0];  // per render object: the attribute is per device, and contexts may live on several
    if (!attr_set) {
        ZCU_CUDA(ctx, cudaFuncSetAttribute(k_nice_scan_render<1, false, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ZCU_CUDA(ctx, cudaFuncSetAttribute(k_nice_scan_render<2, false, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ZCU_CUDA(ctx, cudaFuncSetAttribute(k_nice_scan_render<1, true, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ZCU_CUDA(ctx, cudaFuncSetAttribute(k_nice_scan_render<2, true, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    for (uint32_t F0 = 0; F0 < end_frame; F0 += r->block_frames) {
        ScanArgs a;
        a.V = c.voices;
        a.C = c.channels;
        a.L = c.buffer_frames;
        a.total_frames = c.total_frames;
        a.F0 = F0;
        a.F1 = std::min(F0 + r->block_frames, end_frame);
        a.nchunks = (a.F1 - a.F0 + a.L - 1) / a.L;
        a.nvgroups = r->nwarps;
        a.sample_rate = c.sample_rate;
        a.color = r->d_color;
        a.gains = r->d_gains;
        a.event_off = r->d_event_off;
        a.events = r->d_events;
        a.snaps = r->d_snaps;
        a.cursor = r->d_cursor;
        a.zs = r->d_zs;
        a.M = r->d_M;
        a.sstart = r->d_sstart;
        a.fstate = r->d_fstate;
        a.partial = r->d_partial;
        a.vout = vo.d;
        a.vout_first = vo.first;
        a.vout_frames = vo.frames;
        a.vout_v0 = vo.v0;
        a.vout_nv = vo.nv;
        const uint32_t tasks = a.nvgroups * a.nchunks;
        const uint32_t grid = (tasks + 3) / 4;
        const bool prof = r->profiling && !vo.d;
        if (prof) {
            int rc = record_pair_begin(ctx, r->ev_aux, r->ev_aux_used);
            if (rc != ZCU_OK) return rc;
        }
        k_nice_scan_zs<<<grid, 128, 0, ctx->stream>>>(a);
        if (prof) {
            ZCU_CUDA(ctx, cudaEventRecord(r->ev_aux[r->ev_aux_used + 1], ctx->stream));
            r->ev_aux_used += 2;
        }
        k_nice_scan_prefix<<<(c.voices + 127) / 128, 128, 0, ctx->stream>>>(a);
        if (prof) {
            int rc = record_pair_begin(ctx, r->ev, r->ev_used);
            if (rc != ZCU_OK) return rc;
        }
        if (vo.d) {
            if (c.channels == 2)
                k_nice_scan_render<2, true, RT><<<grid, 128, smem, ctx->stream>>>(a);
            else
                k_nice_scan_render<1, true, RT><<<grid, 128, smem, ctx->stream>>>(a);
            ctx->launches += 3;
        } else {
            if (c.channels == 2)
                k_nice_scan_render<2, false, RT><<<grid, 128, smem, ctx->stream>>>(a);
            else
                k_nice_scan_render<1, false, RT><<<grid, 128, smem, ctx->stream>>>(a);
            if (prof) {
                ZCU_CUDA(ctx, cudaEventRecord(r->ev[r->ev_used + 1], ctx->stream));
                r->ev_used += 2;
            }
            const size_t n = (size_t)(a.F1 - a.F0) * c.channels;
            k_partial_reduce<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(r->d_partial, mix + (size_t)F0 * c.channels, r->nwarps, n);
            ctx->launches += 4;
        }
        ZCU_CUDA(ctx, cudaGetLastError());
    }
    return ZCU_OK;
}

typedef void (*SegKernel)(SegArgs, const float2*, float2*);
template <int RT>
static SegKernel seg_kernel(int channels, bool vout, bool contracted) {
    if (contracted) {
        if (vout) return channels == 2 ? k_nice_seg_render<2, true, RT, true> : k_nice_seg_render<1, true, RT, true>;
        return channels == 2 ? k_nice_seg_render<2, false, RT, true> : k_nice_seg_render<1, false, RT, true>;
    }
    if (vout) return channels == 2 ? k_nice_seg_render<2, true, RT, false> : k_nice_seg_render<1, true, RT, false>;
    return channels == 2 ? k_nice_seg_render<2, false, RT, false> : k_nice_seg_render<1, false, RT, false>;
}

// ZCU_RENDER_SEGMENTS: one launch per block of mix buffers; (voice groups x segments) fills the resident warps once
int render_blocks_seg(zcu_ctx* ctx, zcu_render* r, float* mix, const VoutReq& vo, uint32_t end_frame) {
    const zcu_render_config& c = r->cfg;
    constexpr int RT = 64;
    const size_t smem = seg_render_smem_bytes<RT>();
    bool& attr_set = r->attr_set[1];
    if (!attr_set) {
        for (int combo = 0; combo < 8; combo++)
            ZCU_CUDA(ctx, cudaFuncSetAttribute(seg_kernel<RT>(1 + (combo & 1), (combo & 2) != 0, (combo & 4) != 0),
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    if (r->seg_slots == 0) {
        int per_sm = 0;
        ZCU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, seg_kernel<RT>(2, false, false), 128, smem));
        r->seg_slots = (uint32_t)std::max(1, per_sm) * (uint32_t)ctx->sm_count * 4u;
    }
    ZCU_CUDA(ctx, cudaMemsetAsync(r->d_carry[0], 0, (size_t)c.voices * sizeof(float2), ctx->stream));
    ZCU_CUDA(ctx, cudaMemsetAsync(r->d_warm_stats, 0, 2 * sizeof(uint32_t), ctx->stream));
    int flip = 0;
    // A voice window is rendered with the launches and segments of the whole job, so that it shows the arithmetic of
    // zcu_render_run; launches past the window are not issued, warp tasks that cannot touch it return at once.
    const uint32_t win_end = end_frame;
    if (vo.d) end_frame = c.total_frames;
    for (uint32_t F0 = 0; F0 < end_frame; F0 += r->block_frames) {
        if (vo.d && F0 >= win_end) break;
        SegArgs g;
        ScanArgs& a = g.s;
        a.V = c.voices;
        a.C = c.channels;
        a.L = c.buffer_frames;
        a.total_frames = c.total_frames;
        a.F0 = F0;
        a.F1 = std::min(F0 + r->block_frames, end_frame);
        a.nchunks = (a.F1 - a.F0 + a.L - 1) / a.L;
        a.nvgroups = r->nwarps;
        a.sample_rate = c.sample_rate;
        a.color = r->d_color;
        a.gains = r->d_gains;
        a.event_off = r->d_event_off;
        a.events = r->d_events;
        a.snaps = r->d_snaps;
        a.cursor = r->d_cursor;
        a.zs = nullptr;
        a.M = nullptr;
        a.sstart = nullptr;
        a.fstate = nullptr;
        a.partial = r->d_partial;
        a.vout = vo.d;
        a.vout_first = vo.first;
        a.vout_frames = vo.frames;
        a.vout_v0 = vo.v0;
        a.vout_nv = vo.nv;
        // segments: as many as fill the resident warps once, each at least 8 mix buffers long
        uint32_t nseg = std::max(1u, r->seg_slots / a.nvgroups);
        uint32_t min_chunks = 8;
        if (r->forced_nseg) {
            nseg = r->forced_nseg;
            min_chunks = 1;
        } else if (const char* e = getenv("ZCU_RENDER_SEGMENTS_PER_LAUNCH")) {  // tuning / tests
            nseg = std::max(1, atoi(e));
            min_chunks = 1;
        }
        uint32_t seg_chunks = std::max((a.nchunks + nseg - 1) / nseg, std::min(min_chunks, a.nchunks));
        nseg = (a.nchunks + seg_chunks - 1) / seg_chunks;
        g.nseg = nseg;
        g.seg_chunks = seg_chunks;
        g.carry = nullptr;
        g.warm_stats = r->d_warm_stats;
        g.need_carry = (vo.d && a.F1 < win_end) ? 1u : 0u;
        r->last_nseg = nseg;
        r->last_seg_chunks = seg_chunks;
        const uint32_t tasks = a.nvgroups * nseg;
        const uint32_t grid = (tasks + 3) / 4;
        const float2* cin = r->d_carry[flip];
        float2* cout = r->d_carry[flip ^ 1];
        flip ^= 1;
        const SegKernel kern = seg_kernel<RT>((int)c.channels, vo.d != nullptr, r->filter_contraction);
        if (vo.d) {
            kern<<<grid, 128, smem, ctx->stream>>>(g, cin, cout);
            ctx->launches++;
        } else {
            if (r->profiling) {
                int rc = record_pair_begin(ctx, r->ev, r->ev_used);
                if (rc != ZCU_OK) return rc;
            }
            kern<<<grid, 128, smem, ctx->stream>>>(g, cin, cout);
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

int render_blocks_mode(zcu_ctx* ctx, zcu_render* r, float* mix, const VoutReq& vo, uint32_t end_frame);
int render_blocks(zcu_ctx* ctx, zcu_render* r, float* mix, const VoutReq& vo, uint32_t end_frame) {
    zcu_render::TabSet& T = r->ts[r->ts_cur];
    if (T.has_ready) ZCU_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, T.ready, 0));  // this job's upload and pre-pass
    const int rc = render_blocks_mode(ctx, r, mix, vo, end_frame);
    if (rc != ZCU_OK) return rc;
    ZCU_CUDA(ctx, cudaEventRecord(T.released, ctx->stream));
    T.has_released = true;
    return ZCU_OK;
}
int render_blocks_mode(zcu_ctx* ctx, zcu_render* r, float* mix, const VoutReq& vo, uint32_t end_frame) {
    if (r->cfg.mode == ZCU_RENDER_SEGMENTS) return render_blocks_seg(ctx, r, mix, vo, end_frame);
    if (r->cfg.mode == ZCU_RENDER_SCAN) return render_blocks_scan(ctx, r, mix, vo, end_frame);
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
        a.partial = r->d_partial;
        a.vout = vo.d;
        a.vout_first = vo.first;
        a.vout_frames = vo.frames;
        a.vout_v0 = vo.v0;
        a.vout_nv = vo.nv;
        a.debug_skip_mask = r->debug_skip_mask;
        a.debug_cycles = r->d_debug_cycles;
        constexpr int RT = 128;
        const size_t smem = render_smem_bytes<RT>();
        bool& attr_set = r->attr_set[2];
        if (!attr_set) {
            ZCU_CUDA(ctx, cudaFuncSetAttribute(k_render_nice<1, false, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            ZCU_CUDA(ctx, cudaFuncSetAttribute(k_render_nice<2, false, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            ZCU_CUDA(ctx, cudaFuncSetAttribute(k_render_nice<1, true, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            ZCU_CUDA(ctx, cudaFuncSetAttribute(k_render_nice<2, true, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set = true;
        }
        if (vo.d) {
            if (c.channels == 2)
                k_render_nice<2, true, RT><<<r->nwarps, 128, smem, ctx->stream>>>(a);
            else
                k_render_nice<1, true, RT><<<r->nwarps, 128, smem, ctx->stream>>>(a);
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
                k_render_nice<2, false, RT><<<r->nwarps, 128, smem, ctx->stream>>>(a);
            else
                k_render_nice<1, false, RT><<<r->nwarps, 128, smem, ctx->stream>>>(a);
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
    ZCU_REQUIRE(ctx, cfg->mode == ZCU_RENDER_EXACT || cfg->mode == ZCU_RENDER_SCAN || cfg->mode == ZCU_RENDER_SEGMENTS, "unknown render mode");
    zcu_render* r = new (std::nothrow) zcu_render();
    if (!r) return ZCU_ERR_OOM;
    r->d_color = r->d_gains = r->d_partial = nullptr;
    r->h_events[0] = r->h_events[1] = nullptr;
    r->h_off[0] = r->h_off[1] = nullptr;
    for (auto& t : r->ts) t = zcu_render::TabSet{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, false, false};
    r->ts_cur = 0;
    r->up_stream = nullptr;
    r->h_done[0] = r->h_done[1] = nullptr;
    r->h_busy[0] = r->h_busy[1] = false;
    r->h_flip = 0;
    r->d_state = nullptr;
    r->d_event_off = nullptr;
    r->d_events = nullptr;
    r->nevents = 0;
    r->cap_events = 0;
    r->profiling = false;
    r->ev_used = 0;
    r->ev_aux_used = 0;
    r->nchunks_total = 0;
    r->d_snaps = nullptr;
    r->d_cursor = nullptr;
    r->d_zs = r->d_sstart = nullptr;
    r->d_M = nullptr;
    r->d_fstate = nullptr;
    r->d_carry[0] = r->d_carry[1] = nullptr;
    r->d_warm_stats = nullptr;
    r->seg_slots = 0;
    r->last_nseg = r->last_seg_chunks = 0;
    r->attr_set[0] = r->attr_set[1] = r->attr_set[2] = false;
    r->forced_nseg = 0;
    r->filter_contraction = false;
    r->debug_skip_mask = 0;
    r->d_debug_cycles = nullptr;
#ifdef ZCU_DEBUG_KNOBS  // profiling aid that produces wrong audio: only in builds that ask for it, never from a stray environment variable
    if (const char* m = getenv("ZCU_RENDER_DEBUG_SKIP_MASK")) r->debug_skip_mask = (uint32_t)atoi(m);
#endif
    r->ctx = ctx;
    r->cfg = *cfg;
    const bool scan = cfg->mode == ZCU_RENDER_SCAN;
    const bool segm = cfg->mode == ZCU_RENDER_SEGMENTS;
    uint32_t bb = cfg->block_buffers ? cfg->block_buffers : (scan ? 256 : 64);
    if (segm && !cfg->block_buffers) {
        // one launch for as much of the job as a 3 GB partial-mix buffer allows (the whole 60 s job at 4096 voices)
        const uint64_t per_buffer = (uint64_t)((cfg->voices + 31) / 32) * cfg->buffer_frames * cfg->channels * 4;
        const uint64_t fit = std::max<uint64_t>(1, 3000000000ull / per_buffer);
        const uint64_t all = (cfg->total_frames + cfg->buffer_frames - 1) / cfg->buffer_frames;
        bb = (uint32_t)std::min(fit, all);
    }
    r->block_frames = bb * cfg->buffer_frames;
    r->nwarps = (cfg->voices + 31) / 32;
    const uint32_t V = cfg->voices, C = cfg->channels;
    cudaError_t e = cudaMalloc(&r->d_color, (size_t)V * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_gains, (size_t)V * C * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_state, (size_t)V * sizeof(NiceVoiceState));
    for (auto& t : r->ts) {
        if (e == cudaSuccess) e = cudaMalloc(&t.off, (size_t)(V + 1) * 4);
        if (e == cudaSuccess) e = cudaMallocHost(&r->h_off[&t - r->ts], (size_t)(V + 1) * 4);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t.ready, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&t.released, cudaEventDisableTiming);
    }
    if (e == cudaSuccess) {  // highest priority: the small pre-pass kernels must not queue behind a render's CTAs
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        e = cudaStreamCreateWithPriority(&r->up_stream, cudaStreamNonBlocking, hi);
    }
    r->d_event_off = r->ts[0].off;
    if (e == cudaSuccess) e = cudaMalloc(&r->d_partial, (size_t)r->nwarps * r->block_frames * C * 4);
    if (segm) {
        r->nchunks_total = (cfg->total_frames + cfg->buffer_frames - 1) / cfg->buffer_frames;
        for (auto& t : r->ts)
            if (e == cudaSuccess) e = cudaMalloc(&t.cursor, (size_t)r->nchunks_total * V * 4);
        r->d_cursor = r->ts[0].cursor;
        if (e == cudaSuccess) e = cudaMalloc(&r->d_carry[0], (size_t)V * sizeof(float2));
        if (e == cudaSuccess) e = cudaMalloc(&r->d_carry[1], (size_t)V * sizeof(float2));
        if (e == cudaSuccess) e = cudaMalloc(&r->d_warm_stats, 2 * sizeof(uint32_t));
    }
    if (scan) {
        r->nchunks_total = (cfg->total_frames + cfg->buffer_frames - 1) / cfg->buffer_frames;
        const size_t bc = (size_t)bb * V;
        for (auto& t : r->ts)
            if (e == cudaSuccess) e = cudaMalloc(&t.cursor, (size_t)r->nchunks_total * V * 4);
        r->d_cursor = r->ts[0].cursor;
        if (e == cudaSuccess) e = cudaMalloc(&r->d_zs, bc * sizeof(float2));
        if (e == cudaSuccess) e = cudaMalloc(&r->d_M, bc * sizeof(Mat2));
        if (e == cudaSuccess) e = cudaMalloc(&r->d_sstart, bc * sizeof(float2));
        if (e == cudaSuccess) e = cudaMalloc(&r->d_fstate, (size_t)V * sizeof(double2));
    }
    if (e != cudaSuccess) {
        zcu_render_free(r);
        return zcu_fail(ctx, ZCU_ERR_OOM, "render allocation failed: %s", cudaGetErrorString(e));
    }
    ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_color, voice_color_host, (size_t)V * 4, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaMemcpyAsync(r->d_gains, gains_host, (size_t)V * C * 4, cudaMemcpyHostToDevice, ctx->stream));
    for (auto& t : r->ts) ZCU_CUDA(ctx, cudaMemsetAsync(t.off, 0, (size_t)(V + 1) * 4, ctx->stream));
#ifdef ZCU_DEBUG_KNOBS
    if (getenv("ZCU_RENDER_DEBUG_CYCLES")) {  // profiling aid: per-stage busy/total cycles, printed at zcu_render_free
        ZCU_CUDA(ctx, cudaMalloc(&r->d_debug_cycles, (size_t)r->nwarps * 8 * sizeof(unsigned long long)));
        ZCU_CUDA(ctx, cudaMemsetAsync(r->d_debug_cycles, 0, (size_t)r->nwarps * 8 * sizeof(unsigned long long), ctx->stream));
    }
#endif
    *out = r;
    return ZCU_OK;
}

void zcu_render_free(zcu_render* r) {
    if (!r) return;
    if (r->ctx) cudaStreamSynchronize(r->ctx->stream);
    if (r->d_debug_cycles) {
        std::vector<unsigned long long> h((size_t)r->nwarps * 8);
        cudaMemcpy(h.data(), r->d_debug_cycles, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        double busy[4] = {0, 0, 0, 0}, tot[4] = {0, 0, 0, 0};
        for (uint32_t c = 0; c < r->nwarps; c++)
            for (int w = 0; w < 4; w++) {
                busy[w] += (double)h[((size_t)c * 4 + w) * 2];
                tot[w] += (double)h[((size_t)c * 4 + w) * 2 + 1];
            }
        fprintf(stderr, "[zcu_render debug] stage busy fraction of kernel time: pulse %.3f filter %.3f env %.3f mix %.3f (cycles total/CTA %.3g)\n",
                busy[0] / tot[0], busy[1] / tot[1], busy[2] / tot[2], busy[3] / tot[3], tot[0] / r->nwarps);
        cudaFree(r->d_debug_cycles);
    }
    cudaFree(r->d_color);
    cudaFree(r->d_gains);
    cudaFree(r->d_state);
    if (r->up_stream) {
        cudaStreamSynchronize(r->up_stream);
        cudaStreamDestroy(r->up_stream);
    }
    for (int i = 0; i < 2; i++) {
        cudaFreeHost(r->h_events[i]);
        cudaFreeHost(r->h_off[i]);
        if (r->h_done[i]) cudaEventDestroy(r->h_done[i]);
        cudaFree(r->ts[i].off);
        cudaFree(r->ts[i].events);
        cudaFree(r->ts[i].snaps);
        cudaFree(r->ts[i].cursor);
        if (r->ts[i].ready) cudaEventDestroy(r->ts[i].ready);
        if (r->ts[i].released) cudaEventDestroy(r->ts[i].released);
    }
    cudaFree(r->d_partial);
    cudaFree(r->d_zs);
    cudaFree(r->d_M);
    cudaFree(r->d_sstart);
    cudaFree(r->d_fstate);
    cudaFree(r->d_carry[0]);
    cudaFree(r->d_carry[1]);
    cudaFree(r->d_warm_stats);
    for (cudaEvent_t e : r->ev)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : r->ev_aux)
        if (e) cudaEventDestroy(e);
    delete r;
}

int zcu_render_set_events(zcu_ctx* ctx, zcu_render* r, const uint32_t* event_off_host, const zcu_note_event* events_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && event_off_host, "null argument");
    const uint32_t V = r->cfg.voices;
    const uint32_t n = event_off_host[V];
    ZCU_REQUIRE(ctx, n == 0 || events_host, "null events");
    if (n > r->cap_events || !r->ts[0].events) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ZCU_CUDA(ctx, cudaStreamSynchronize(r->up_stream));
        r->d_events = nullptr;
        r->d_snaps = nullptr;
        for (int i = 0; i < 2; i++) {
            cudaFree(r->ts[i].events);
            cudaFree(r->ts[i].snaps);
            r->ts[i].events = nullptr;
            r->ts[i].snaps = nullptr;
            cudaFreeHost(r->h_events[i]);
            r->h_events[i] = nullptr;
            r->h_busy[i] = false;
            if (!r->h_done[i]) ZCU_CUDA(ctx, cudaEventCreateWithFlags(&r->h_done[i], cudaEventDisableTiming));
        }
        r->cap_events = 0;
        for (int i = 0; i < 2; i++) {
            if (r->cfg.mode != ZCU_RENDER_EXACT) ZCU_CUDA(ctx, cudaMalloc(&r->ts[i].snaps, (size_t)(n ? n : 1) * sizeof(EnvSnap)));
            ZCU_CUDA(ctx, cudaMalloc(&r->ts[i].events, (size_t)(n ? n : 1) * sizeof(DevEvent)));
            ZCU_CUDA(ctx, cudaMallocHost(&r->h_events[i], (size_t)(n ? n : 1) * sizeof(DevEvent)));
        }
        r->cap_events = n;
    }
    r->h_flip ^= 1;
    const int hb = r->h_flip;
    if (r->h_busy[hb]) ZCU_CUDA(ctx, cudaEventSynchronize(r->h_done[hb]));  // this buffer's copy of two calls ago
    r->nevents = n;
    // per-call host scalar of modules.zig:231-234 (cos, sqrt): once per event, on the host
    // (a small direct-mapped memo keyed by the bits of freq: a bank plays few distinct pitches, and the memo returns
    // the very value the function would)
    // Expanded by a few host threads (each with its own memo) straight into the pinned staging buffer: at ~1 M events per
    // job a single thread costs more than the H2D copy it feeds.
    struct Memo { uint32_t key; float cutoff; bool valid; };
    const float sr = r->cfg.sample_rate;
    const uint32_t total_frames = r->cfg.total_frames;
    DevEvent* const staged = r->h_events[hb];
    // Each thread takes a range of voices and checks the table while it expands it: offsets monotone from 0, every voice's
    // frames chronological and inside the render.  A malformed table would make the device kernels walk out of bounds.
    auto expand = [event_off_host, events_host, staged, sr, total_frames, n](uint32_t v0, uint32_t v1, int* bad) {
        std::vector<Memo> memo(1024, Memo{0u, 0.0f, false});
        for (uint32_t v = v0; v < v1; v++) {
            const uint32_t i0 = event_off_host[v], i1 = event_off_host[v + 1];
            if (i1 < i0 || i1 > n) {
                *bad = 1;
                return;
            }
            uint32_t prev = 0;
            for (uint32_t i = i0; i < i1; i++) {
                DevEvent& d = staged[i];
                d.frame = events_host[i].frame;
                if (d.frame < prev || d.frame >= total_frames) {
                    *bad = 2;
                    return;
                }
                prev = d.frame;
                d.note_id = events_host[i].note_id;
                d.freq = events_host[i].freq;
                d.note_on = events_host[i].note_on;
                uint32_t key;
                memcpy(&key, &events_host[i].freq, 4);
                Memo& m = memo[(key * 2654435761u) >> 22];
                if (!m.valid || m.key != key) {
                    m.key = key;
                    m.cutoff = zcu_filter_cutoff_from_frequency(events_host[i].freq * 8.0f, sr);
                    m.valid = true;
                }
                d.cutoff = m.cutoff;
                d.cnt0 = d.flags = d.pad = 0;
            }
        }
    };
    uint32_t nthreads = 1;
    if (n >= (1u << 16)) nthreads = std::min(4u, std::max(1u, std::thread::hardware_concurrency()));  // measured: 4 is the knee
    if (const char* e = getenv("ZCU_HOST_THREADS")) nthreads = (uint32_t)std::max(1, atoi(e));
    nthreads = std::min(nthreads, 64u);
    int bad[64] = {0};
    if (event_off_host[0] != 0) bad[0] = 1;
    if (nthreads <= 1) {
        if (!bad[0]) expand(0, V, &bad[0]);
    } else if (!bad[0]) {
        std::vector<std::thread> pool;
        for (uint32_t t = 1; t < nthreads; t++) {
            const uint32_t a = (uint32_t)((uint64_t)V * t / nthreads), b = (uint32_t)((uint64_t)V * (t + 1) / nthreads);
            if (a < b) pool.emplace_back(expand, a, b, &bad[t]);
        }
        expand(0, (uint32_t)((uint64_t)V / nthreads), &bad[0]);
        for (auto& th : pool) th.join();
    }
    for (uint32_t t = 0; t < nthreads; t++) {
        if (bad[t]) {
            r->h_flip ^= 1;  // the staging buffer was not handed to a copy
            return zcu_fail(ctx, ZCU_ERR_INVALID, bad[t] == 1 ? "event_off must start at 0 and be monotone"
                                                               : "events of a voice must be chronological with frames inside the render");
        }
    }
    // upload and pre-pass of this job on the upload stream, into the table set the current run is not reading
    const int nx = r->ts_cur ^ 1;
    zcu_render::TabSet& T = r->ts[nx];
    cudaStream_t us = r->up_stream;
    if (T.has_released) ZCU_CUDA(ctx, cudaStreamWaitEvent(us, T.released, 0));
    memcpy(r->h_off[hb], event_off_host, (size_t)(V + 1) * 4);
    ZCU_CUDA(ctx, cudaMemcpyAsync(T.off, r->h_off[hb], (size_t)(V + 1) * 4, cudaMemcpyHostToDevice, us));
    if (n) ZCU_CUDA(ctx, cudaMemcpyAsync(T.events, staged, (size_t)n * sizeof(DevEvent), cudaMemcpyHostToDevice, us));
    ZCU_CUDA(ctx, cudaEventRecord(r->h_done[hb], us));
    r->h_busy[hb] = true;
    if (r->cfg.mode != ZCU_RENDER_EXACT) {
        PrepassArgs p;
        p.V = V;
        p.L = r->cfg.buffer_frames;
        p.nchunks_total = r->nchunks_total;
        p.sample_rate = r->cfg.sample_rate;
        p.event_off = T.off;
        p.events = T.events;
        p.snaps = T.snaps;
        p.cursor = T.cursor;
        k_nice_cursor<<<(V + 3) / 4, 128, 0, us>>>(p);
        k_nice_prepass<<<(V + 63) / 64, 64, 0, us>>>(p);
        ctx->launches += 2;
        ZCU_CUDA(ctx, cudaGetLastError());
    }
    ZCU_CUDA(ctx, cudaEventRecord(T.ready, us));
    T.has_ready = true;
    r->ts_cur = nx;
    r->d_event_off = T.off;
    r->d_events = T.events;
    r->d_snaps = T.snaps;
    r->d_cursor = T.cursor;
    return ZCU_OK;
}

int zcu_render_run(zcu_ctx* ctx, zcu_render* r, zcu_buf* mix) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && mix, "null argument");
    ZCU_REQUIRE(ctx, r->d_events != nullptr, "zcu_render_set_events has not been called");
    ZCU_REQUIRE(ctx, mix->frames == r->cfg.total_frames && mix->voices == r->cfg.channels, "mix must be [total_frames][channels]");
    return render_blocks(ctx, r, mix->d, VoutReq{nullptr, 0, 0, 0, 0}, r->cfg.total_frames);
}

int zcu_render_run_voices(zcu_ctx* ctx, zcu_render* r, uint32_t first_frame, zcu_buf* out_voices) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && out_voices, "null argument");
    ZCU_REQUIRE(ctx, r->d_events != nullptr, "zcu_render_set_events has not been called");
    ZCU_REQUIRE(ctx, out_voices->voices == r->cfg.voices, "out_voices must be [frames][voices]");
    ZCU_REQUIRE(ctx, (uint64_t)first_frame + out_voices->frames <= r->cfg.total_frames, "window beyond the render");
    return render_blocks(ctx, r, nullptr, VoutReq{out_voices->d, first_frame, out_voices->frames, 0, r->cfg.voices}, first_frame + out_voices->frames);
}

int zcu_render_run_voice_window(zcu_ctx* ctx, zcu_render* r, uint32_t first_voice, uint32_t first_frame, zcu_buf* out_voices) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && out_voices, "null argument");
    ZCU_REQUIRE(ctx, r->d_events != nullptr, "zcu_render_set_events has not been called");
    ZCU_REQUIRE(ctx, out_voices->voices > 0 && (uint64_t)first_voice + out_voices->voices <= r->cfg.voices, "voice window beyond the bank");
    ZCU_REQUIRE(ctx, (uint64_t)first_frame + out_voices->frames <= r->cfg.total_frames, "window beyond the render");
    return render_blocks(ctx, r, nullptr, VoutReq{out_voices->d, first_frame, out_voices->frames, first_voice, out_voices->voices},
                         first_frame + out_voices->frames);
}

int zcu_render_set_profiling(zcu_render* r, int enabled) {
    if (!r) return ZCU_ERR_INVALID;
    r->profiling = enabled != 0;
    r->ev_used = 0;
    r->ev_aux_used = 0;
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

int zcu_render_aux_kernel_ms(zcu_ctx* ctx, zcu_render* r, float* total_ms, uint32_t* launches) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && total_ms && launches, "null argument");
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    double sum = 0.0;
    for (size_t i = 0; i + 1 < r->ev_aux_used; i += 2) {
        float ms = 0.0f;
        ZCU_CUDA(ctx, cudaEventElapsedTime(&ms, r->ev_aux[i], r->ev_aux[i + 1]));
        sum += ms;
    }
    *total_ms = (float)sum;
    *launches = (uint32_t)(r->ev_aux_used / 2);
    r->ev_aux_used = 0;
    return ZCU_OK;
}

uint64_t zcu_render_launches_per_run(const zcu_render* r) {
    if (!r) return 0;
    const uint64_t nblocks = (r->cfg.total_frames + r->block_frames - 1) / r->block_frames;
    if (r->cfg.mode == ZCU_RENDER_SEGMENTS) return 2 * nblocks;
    return r->cfg.mode == ZCU_RENDER_SCAN ? 4 * nblocks : 1 + 2 * nblocks;
}

int zcu_render_set_segments(zcu_render* r, uint32_t segments_per_launch) {
    if (!r || r->cfg.mode != ZCU_RENDER_SEGMENTS) return ZCU_ERR_INVALID;
    r->forced_nseg = segments_per_launch;
    return ZCU_OK;
}

int zcu_render_set_filter_contraction(zcu_render* r, int enabled) {
    if (!r || r->cfg.mode != ZCU_RENDER_SEGMENTS) return ZCU_ERR_INVALID;
    r->filter_contraction = enabled != 0;
    return ZCU_OK;
}

int zcu_render_segment_info(zcu_ctx* ctx, zcu_render* r, uint32_t* segments, uint32_t* buffers_per_segment, uint32_t* warm_sum, uint32_t* warm_max) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, r && r->cfg.mode == ZCU_RENDER_SEGMENTS, "not a ZCU_RENDER_SEGMENTS render");
    uint32_t h[2] = {0, 0};
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ZCU_CUDA(ctx, cudaMemcpy(h, r->d_warm_stats, sizeof(h), cudaMemcpyDeviceToHost));
    if (segments) *segments = r->last_nseg;
    if (buffers_per_segment) *buffers_per_segment = r->last_seg_chunks;
    if (warm_sum) *warm_sum = h[0];
    if (warm_max) *warm_max = h[1];
    return ZCU_OK;
}

}  // extern "C"
