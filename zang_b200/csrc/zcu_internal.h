// zcu_internal.h — handle definitions and helpers shared by the translation units of libzangcu.so.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <set>
#include <string>

#include "../../include/zangcu.h"

struct zcu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err = "";
    uint64_t launches = 0;
    // stream-ordered arena for per-call parameter tables (params rows, curve nodes, gains)
    void* arena[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t arena_bytes[4] = {0, 0, 0, 0};
    int sm_count = 148;
    // ZCU_PAINT_TOLERANCE: published segment end states + ready flags (stamped with paint_epoch) of the lockstep kernel
    int paint_mode = 0;
    int paint_output = 0;  // ZCU_OUT_ACCUMULATE / ZCU_OUT_OVERWRITE
    void* chain = nullptr;
    size_t chain_bytes = 0;
    uint32_t* chain_flags = nullptr;
    size_t chain_flag_count = 0;
    uint32_t paint_epoch = 0;
    uint32_t* paint_ticket = nullptr;  // [0] ticket counter of chained launches, [1] error flag of a timed-out wait
    uint32_t paint_ticket_base = 0;    // host shadow of the counter: tickets handed out by the launches enqueued so far
    std::set<const void*> attr_done;   // kernels whose dynamic shared-memory limit has been raised on this context's device
    void* prefix = nullptr;  // M::PREFIX modules: state rows at the segment starts
    size_t prefix_bytes = 0;
    // Noise skip-ahead: xoshiro256 jump matrices (built on first use) and the device counter of unforeseeable extra draws
    void* xjump = nullptr;
    unsigned long long* xjump_resync = nullptr;
    // NCCL (loaded lazily with dlopen)
    void* nccl_comm = nullptr;
    int nccl_rank = 0, nccl_nranks = 1;
};

struct zcu_buf {
    zcu_ctx* ctx;
    float* d;
    uint32_t frames, voices;
    bool owned;
};

struct zcu_bank {
    zcu_ctx* ctx;
    int kind;
    uint32_t voices;
    void* d;        // voices * stride bytes
    size_t stride;  // sizeof the state POD
    float* line = nullptr;        // delay modules: [delay_samples][voices]
    uint32_t delay_samples = 0;
};

struct zcu_spans {
    zcu_ctx* ctx;
    uint32_t voices, frames, nspans;
    uint32_t* d_off;    // voices+1
    zcu_span* d_spans;  // nspans
    uint32_t max_param; // largest param row referenced
};

struct zcu_sample {
    zcu_ctx* ctx;
    uint8_t* d;
    size_t len;
    uint32_t num_channels, sample_rate;
    int format;
};

// -- error plumbing -----------------------------------------------------------------------------
int zcu_fail(zcu_ctx* ctx, int code, const char* fmt, ...);

#define ZCU_CUDA(ctx, expr)                                                                          \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess)                                                                       \
            return zcu_fail((ctx), _e == cudaErrorMemoryAllocation ? ZCU_ERR_OOM : ZCU_ERR_CUDA,     \
                            "%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));    \
    } while (0)

// Every ABI entry point validates its arguments with ZCU_REQUIRE first; that is also where the calling thread is switched
// to the context's device (allocations and launches use the CURRENT device, and a host may hold contexts on several GPUs).
inline void zcu_use_device(const zcu_ctx* ctx) {
    int cur = -1;
    if (ctx && cudaGetDevice(&cur) == cudaSuccess && cur != ctx->device) cudaSetDevice(ctx->device);
}

#define ZCU_REQUIRE(ctx, cond, msg)                                                         \
    do {                                                                                    \
        zcu_use_device(ctx);                                                                \
        if (!(cond)) return zcu_fail((ctx), ZCU_ERR_INVALID, "%s: %s", __func__, (msg));    \
    } while (0)

// Tile depth T and segments of the frame axis for the lockstep paint kernel.  Modules with an exact skip-ahead are cut into
// as many segments as fill the chip's resident one-warp CTAs exactly once (each at least 256 frames, a multiple of T);
// the others get one segment.  nbufs = staged buffers (output + inputs).  ZCU_PAINT_SEGMENTS overrides (tests).
void zcu_paint_segments(const zcu_ctx* ctx, uint32_t voices, uint32_t frames, int nbufs, bool skippable, uint32_t* T, uint32_t* nseg,
                        uint32_t* seg_frames, uint32_t max_resident = 8);

// copies `bytes` of host memory into arena slot `slot` (stream ordered) and returns the device ptr
int zcu_arena_upload(zcu_ctx* ctx, int slot, const void* host, size_t bytes, void** dev_out);
int zcu_arena_reserve(zcu_ctx* ctx, int slot, size_t bytes, void** dev_out);

// implemented in zcu_paint.cu / zcu_basics.cu / zcu_render.cu
struct PaintLaunch {
    int kind;
    uint32_t V, frames;
    const uint32_t* off;
    const zcu_span* spans;
    const void* params;
    const void* ramp_tabs;  // Envelope (zcu_envelope_ramp_tabs)
    uint32_t n_ramp_tabs;
    void* state;
    float* out;
    const float* in0;
    const float* in1;
    const float* in2;
    const void* aux;      // curve nodes / PCM bytes
    uint64_t aux_len;
    uint32_t smp_channels, smp_rate;
    int smp_format;
    float* line;             // delay modules
    uint32_t delay_samples;
};
int zcu_launch_paint(zcu_ctx* ctx, const PaintLaunch& pl);
// Envelope: exact ramp-clock tables (RampTab, zcu_modules.cuh) of up to 4 distinct Painter steps named by the first rows,
// built once per step value (process-wide cache), uploaded stream-ordered; *dev stays null when no row has a ramp
int zcu_envelope_ramp_tabs(zcu_ctx* ctx, const zcu_envelope_params* rows, uint32_t nrows, const void** dev, uint32_t* count);
int zcu_launch_basic(zcu_ctx* ctx, int op, const zcu_spans* spans, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b, float scalar);
int zcu_launch_basic_rows(zcu_ctx* ctx, int op, const zcu_spans* spans, zcu_buf* dest, const zcu_buf* a, const float* rows, uint32_t nrows);
int zcu_launch_bank_init(zcu_ctx* ctx, zcu_bank* bank, uint32_t first_seed);
int zcu_selftest_no_fma(zcu_ctx* ctx);
int zcu_ensure_xjump(zcu_ctx* ctx);  // xoshiro256 jump matrices of the Noise skip-ahead (ctx->xjump, ctx->xjump_resync)
