// zcu_api.cu — contexts, buffers, banks, span tables and the exported paint/basics entry points.
#include <cmath>
#include <cstdarg>
#include <cstring>
#include <new>
#include <vector>

#include "zcu_internal.h"
#include "zcu_modules.cuh"

int zcu_fail(zcu_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

int zcu_arena_reserve(zcu_ctx* ctx, int slot, size_t bytes, void** dev_out) {
    if (bytes == 0) bytes = 16;
    if (ctx->arena_bytes[slot] < bytes) {
        // the old block may still be read by kernels already enqueued: free it stream-ordered
        if (ctx->arena[slot]) ZCU_CUDA(ctx, cudaFreeAsync(ctx->arena[slot], ctx->stream));
        ctx->arena[slot] = nullptr;
        ctx->arena_bytes[slot] = 0;
        size_t cap = bytes + bytes / 2;
        ZCU_CUDA(ctx, cudaMallocAsync(&ctx->arena[slot], cap, ctx->stream));
        ctx->arena_bytes[slot] = cap;
    }
    *dev_out = ctx->arena[slot];
    return ZCU_OK;
}

int zcu_arena_upload(zcu_ctx* ctx, int slot, const void* host, size_t bytes, void** dev_out) {
    int rc = zcu_arena_reserve(ctx, slot, bytes, dev_out);
    if (rc) return rc;
    if (bytes) ZCU_CUDA(ctx, cudaMemcpyAsync(*dev_out, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return ZCU_OK;
}

extern "C" {

int zcu_abi_version(void) { return ZCU_ABI_VERSION; }

int zcu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

static int ctx_create(int device, bool borrow, cudaStream_t borrowed, zcu_ctx** out);
int zcu_ctx_create(int device, zcu_ctx** out) { return ctx_create(device, false, nullptr, out); }
int zcu_ctx_create_on_stream(int device, void* cuda_stream, zcu_ctx** out) { return ctx_create(device, true, (cudaStream_t)cuda_stream, out); }
static int ctx_create(int device, bool borrow, cudaStream_t borrowed, zcu_ctx** out) {
    if (!out) return ZCU_ERR_INVALID;
    *out = nullptr;
    int n = zcu_device_count();
    if (n <= 0) return ZCU_ERR_NO_DEVICE;  // no CPU fallback
    if (device < 0 || device >= n) return ZCU_ERR_INVALID;
    zcu_ctx* ctx = new (std::nothrow) zcu_ctx();
    if (!ctx) return ZCU_ERR_OOM;
    ctx->device = device;
    ctx->own_stream = !borrow;
    if (borrow) ctx->stream = borrowed;
    if (cudaSetDevice(device) != cudaSuccess || (!borrow && cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        cudaGetLastError();
        delete ctx;
        return ZCU_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    int rc = zcu_selftest_no_fma(ctx);
    if (rc) {
        zcu_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return ZCU_OK;
}

void zcu_ctx_destroy(zcu_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->nccl_comm) zcu_nccl_finalize(ctx);
    for (int i = 0; i < 4; i++)
        if (ctx->arena[i]) cudaFree(ctx->arena[i]);
    if (ctx->prefix) cudaFree(ctx->prefix);
    if (ctx->xjump) cudaFree(ctx->xjump);
    if (ctx->xjump_resync) cudaFree(ctx->xjump_resync);
    if (ctx->chain) cudaFree(ctx->chain);
    if (ctx->chain_flags) cudaFree(ctx->chain_flags);
    if (ctx->paint_ticket) cudaFree(ctx->paint_ticket);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream && ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}


int zcu_ctx_sync(zcu_ctx* ctx) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->paint_ticket) {  // a chained paint launch whose wait for a predecessor segment timed out (zcu_paint_kernel.cuh)
        uint32_t err = 0;
        ZCU_CUDA(ctx, cudaMemcpy(&err, ctx->paint_ticket + 1, sizeof err, cudaMemcpyDeviceToHost));
        if (err) {
            cudaMemset(ctx->paint_ticket + 1, 0, sizeof err);
            return zcu_fail(ctx, ZCU_ERR_CUDA, "a chained paint segment timed out waiting for its predecessor; the painted block is invalid");
        }
    }
    return ZCU_OK;
}
const char* zcu_last_error(zcu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void* zcu_ctx_stream(zcu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
uint64_t zcu_ctx_launch_count(zcu_ctx* ctx) { return ctx ? ctx->launches : 0; }
int zcu_ctx_set_paint_mode(zcu_ctx* ctx, int mode) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, mode == ZCU_PAINT_EXACT || mode == ZCU_PAINT_TOLERANCE, "unknown paint mode");
    ctx->paint_mode = mode;
    return ZCU_OK;
}
int zcu_ctx_paint_mode(const zcu_ctx* ctx) { return ctx ? ctx->paint_mode : ZCU_PAINT_EXACT; }
int zcu_ctx_set_paint_output(zcu_ctx* ctx, int mode) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, mode == ZCU_OUT_ACCUMULATE || mode == ZCU_OUT_OVERWRITE, "unknown output mode");
    ctx->paint_output = mode;
    return ZCU_OK;
}
int zcu_ctx_paint_output(const zcu_ctx* ctx) { return ctx ? ctx->paint_output : ZCU_OUT_ACCUMULATE; }
int zcu_ctx_paint_warm_stats(zcu_ctx* ctx, uint32_t out[4]) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    zcu_use_device(ctx);
    out[0] = out[1] = out[2] = out[3] = 0;
    if (!ctx->paint_ticket) return ZCU_OK;
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ZCU_CUDA(ctx, cudaMemcpy(out, ctx->paint_ticket + 2, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    ZCU_CUDA(ctx, cudaMemset(ctx->paint_ticket + 2, 0, 4 * sizeof(uint32_t)));
    return ZCU_OK;
}
int zcu_ctx_noise_resync_count(zcu_ctx* ctx, uint64_t* out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = 0;
    if (!ctx->xjump_resync) return ZCU_OK;
    unsigned long long h = 0;
    ZCU_CUDA(ctx, cudaMemcpyAsync(&h, ctx->xjump_resync, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = h;
    return ZCU_OK;
}

int zcu_timer_start(zcu_ctx* ctx) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return ZCU_OK;
}
int zcu_timer_stop_ms(zcu_ctx* ctx, float* ms) {
    if (!ctx || !ms) return ZCU_ERR_INVALID;
    ZCU_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ZCU_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    ZCU_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return ZCU_OK;
}

// ---- buffers ---------------------------------------------------------------------------------------
int zcu_buf_alloc(zcu_ctx* ctx, uint32_t frames, uint32_t voices, zcu_buf** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    float* d = nullptr;
    size_t bytes = (size_t)frames * voices * sizeof(float);
    if (bytes == 0) bytes = 16;
    ZCU_CUDA(ctx, cudaSetDevice(ctx->device));
    ZCU_CUDA(ctx, cudaMalloc(&d, bytes));
    ZCU_CUDA(ctx, cudaMemsetAsync(d, 0, bytes, ctx->stream));
    *out = new zcu_buf{ctx, d, frames, voices, true};
    return ZCU_OK;
}
int zcu_buf_wrap(zcu_ctx* ctx, void* device_ptr, uint32_t frames, uint32_t voices, zcu_buf** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, device_ptr != nullptr, "null device pointer");
    ZCU_REQUIRE(ctx, ((uintptr_t)device_ptr & 3u) == 0, "device pointer must be 4-byte aligned");
    *out = new zcu_buf{ctx, (float*)device_ptr, frames, voices, false};
    return ZCU_OK;
}
void zcu_buf_free(zcu_buf* buf) {
    if (!buf) return;
    if (buf->owned && buf->d) {
        cudaStreamSynchronize(buf->ctx->stream);
        cudaFree(buf->d);
    }
    delete buf;
}
int zcu_buf_upload(zcu_ctx* ctx, zcu_buf* buf, const float* host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, buf && host, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(buf->d, host, (size_t)buf->frames * buf->voices * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    return ZCU_OK;
}
int zcu_buf_download(zcu_ctx* ctx, const zcu_buf* buf, float* host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, buf && host, "null argument");
    ZCU_CUDA(ctx, cudaMemcpyAsync(host, buf->d, (size_t)buf->frames * buf->voices * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
void* zcu_buf_device_ptr(const zcu_buf* buf) { return buf ? buf->d : nullptr; }
uint32_t zcu_buf_frames(const zcu_buf* buf) { return buf ? buf->frames : 0; }
uint32_t zcu_buf_voices(const zcu_buf* buf) { return buf ? buf->voices : 0; }

// ---- banks -------------------------------------------------------------------------------------------
size_t zcu_bank_state_size(int kind) {
    switch (kind) {
        case ZCU_MOD_SINE_OSC: return sizeof(zcu_sine_state);
        case ZCU_MOD_PULSE_OSC: return sizeof(zcu_pulse_state);
        case ZCU_MOD_TRISAW_OSC: return sizeof(zcu_trisaw_state);
        case ZCU_MOD_NOISE: return sizeof(zcu_noise_state);
        case ZCU_MOD_ENVELOPE: return sizeof(zcu_envelope_state);
        case ZCU_MOD_FILTER: return sizeof(zcu_filter_state);
        case ZCU_MOD_PORTAMENTO: return sizeof(zcu_portamento_state);
        case ZCU_MOD_CURVE: return sizeof(zcu_curve_state);
        case ZCU_MOD_SAMPLER: return sizeof(zcu_sampler_state);
        case ZCU_MOD_DECIMATOR: return sizeof(zcu_decimator_state);
        case ZCU_MOD_CYCLE: return sizeof(zcu_cycle_state);
        case ZCU_MOD_FM_OSCILLATOR: return sizeof(zcu_fm_oscillator_state);
        case ZCU_MOD_SIMPLE_DELAY: return sizeof(zcu_delay_state);
        case ZCU_MOD_FILTERED_ECHOES: return sizeof(zcu_echoes_state);
        default: return 0;  // Gate, Distortion: stateless
    }
}
int zcu_bank_create(zcu_ctx* ctx, int kind, uint32_t voices, uint32_t first_seed, zcu_bank** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, kind >= 0 && kind < ZCU_MOD_COUNT, "unknown module kind");
    ZCU_REQUIRE(ctx, kind != ZCU_MOD_SIMPLE_DELAY && kind != ZCU_MOD_FILTERED_ECHOES, "delay modules: use zcu_delay_bank_create");
    zcu_bank* b = new zcu_bank{ctx, kind, voices, nullptr, zcu_bank_state_size(kind)};
    size_t bytes = b->stride * voices;
    if (bytes) {
        cudaError_t e = cudaMalloc(&b->d, bytes);
        if (e != cudaSuccess) {
            delete b;
            return zcu_fail(ctx, ZCU_ERR_OOM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        }
    }
    int rc = zcu_launch_bank_init(ctx, b, first_seed);
    if (rc) {
        zcu_bank_free(b);
        return rc;
    }
    *out = b;
    return ZCU_OK;
}
int zcu_delay_bank_create(zcu_ctx* ctx, int kind, uint32_t voices, uint32_t delay_samples, zcu_bank** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, kind == ZCU_MOD_SIMPLE_DELAY || kind == ZCU_MOD_FILTERED_ECHOES, "not a delay module kind");
    ZCU_REQUIRE(ctx, delay_samples > 0 && delay_samples < 0x80000000u, "delay_samples must be in 1 .. 2^31-1");
    zcu_bank* b = new zcu_bank{ctx, kind, voices, nullptr, zcu_bank_state_size(kind)};
    b->delay_samples = delay_samples;
    const size_t bytes = b->stride * voices, lbytes = (size_t)delay_samples * voices * sizeof(float);
    cudaError_t e = cudaSuccess;
    if (bytes) e = cudaMalloc(&b->d, bytes);
    if (e == cudaSuccess && lbytes) e = cudaMalloc(&b->line, lbytes);
    if (e != cudaSuccess) {
        zcu_bank_free(b);
        return zcu_fail(ctx, ZCU_ERR_OOM, "cudaMalloc(%zu + %zu) failed: %s", bytes, lbytes, cudaGetErrorString(e));
    }
    int rc = zcu_launch_bank_init(ctx, b, 0);
    if (rc) {
        zcu_bank_free(b);
        return rc;
    }
    *out = b;
    return ZCU_OK;
}
uint32_t zcu_bank_delay_samples(const zcu_bank* bank) { return bank ? bank->delay_samples : 0; }
int zcu_bank_get_delay_line(zcu_ctx* ctx, const zcu_bank* bank, float* host, size_t bytes) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, bank && host, "null argument");
    ZCU_REQUIRE(ctx, bytes == (size_t)bank->delay_samples * bank->voices * sizeof(float), "size mismatch");
    if (bytes) ZCU_CUDA(ctx, cudaMemcpyAsync(host, bank->line, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
int zcu_bank_set_delay_line(zcu_ctx* ctx, zcu_bank* bank, const float* host, size_t bytes) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, bank && host, "null argument");
    ZCU_REQUIRE(ctx, bytes == (size_t)bank->delay_samples * bank->voices * sizeof(float), "size mismatch");
    if (bytes) ZCU_CUDA(ctx, cudaMemcpyAsync(bank->line, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
void zcu_bank_free(zcu_bank* bank) {
    if (!bank) return;
    if (bank->d || bank->line) cudaStreamSynchronize(bank->ctx->stream);
    if (bank->d) cudaFree(bank->d);
    if (bank->line) cudaFree(bank->line);
    delete bank;
}
int zcu_bank_reset(zcu_ctx* ctx, zcu_bank* bank, uint32_t first_seed) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, bank != nullptr, "null bank");
    return zcu_launch_bank_init(ctx, bank, first_seed);
}
int zcu_bank_get_state(zcu_ctx* ctx, const zcu_bank* bank, void* host, size_t bytes) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, bank && host, "null argument");
    ZCU_REQUIRE(ctx, bytes == bank->stride * bank->voices, "size mismatch");
    if (bytes) ZCU_CUDA(ctx, cudaMemcpyAsync(host, bank->d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
int zcu_bank_set_state(zcu_ctx* ctx, zcu_bank* bank, const void* host, size_t bytes) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, bank && host, "null argument");
    ZCU_REQUIRE(ctx, bytes == bank->stride * bank->voices, "size mismatch");
    if (bytes) ZCU_CUDA(ctx, cudaMemcpyAsync(bank->d, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return ZCU_OK;
}

// ---- span tables -------------------------------------------------------------------------------------
int zcu_spans_upload(zcu_ctx* ctx, uint32_t voices, uint32_t frames, const uint32_t* voice_off, const zcu_span* spans, zcu_spans** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, voice_off != nullptr, "null voice_off");
    ZCU_REQUIRE(ctx, voice_off[0] == 0, "voice_off[0] must be 0");
    const uint32_t n = voice_off[voices];
    ZCU_REQUIRE(ctx, n == 0 || spans != nullptr, "null spans");
    uint32_t max_param = 0;
    const char* bad = nullptr;
    for (uint32_t v = 0; v < voices && !bad; v++) {
        if (voice_off[v] > voice_off[v + 1] || voice_off[v + 1] > n) {
            bad = "voice_off must be non-decreasing";
            break;
        }
        uint32_t prev_end = 0;
        for (uint32_t k = voice_off[v]; k < voice_off[v + 1]; k++) {
            if (!(spans[k].start < spans[k].end && spans[k].end <= frames)) bad = "span outside the block or empty";
            if (spans[k].start < prev_end) bad = "spans of a voice must be sorted and disjoint";
            prev_end = spans[k].end;
            if (spans[k].param > max_param) max_param = spans[k].param;
        }
    }
    ZCU_REQUIRE(ctx, bad == nullptr, bad);
    zcu_spans* s = new zcu_spans{ctx, voices, frames, n, nullptr, nullptr, max_param};
    cudaError_t e = cudaMalloc(&s->d_off, (size_t)(voices + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_spans, (size_t)(n ? n : 1) * sizeof(zcu_span));
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_OOM, "span table allocation failed: %s", cudaGetErrorString(e));
    }
    e = cudaMemcpyAsync(s->d_off, voice_off, (size_t)(voices + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(s->d_spans, spans, (size_t)n * sizeof(zcu_span), cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
        zcu_spans_free(s);
        return zcu_fail(ctx, ZCU_ERR_CUDA, "span table upload failed: %s", cudaGetErrorString(e));
    }
    *out = s;
    return ZCU_OK;
}
int zcu_spans_full(zcu_ctx* ctx, uint32_t voices, uint32_t num_buffers, uint32_t buffer_frames, zcu_spans** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, (uint64_t)voices * num_buffers < (1ull << 32) && (uint64_t)num_buffers * buffer_frames < (1ull << 32), "table too large");
    std::vector<uint32_t> off(voices + 1);
    std::vector<zcu_span> sp((size_t)voices * num_buffers);
    for (uint32_t v = 0; v <= voices; v++) off[v] = v * num_buffers;
    for (uint32_t v = 0; v < voices; v++)
        for (uint32_t b = 0; b < num_buffers; b++) sp[(size_t)v * num_buffers + b] = zcu_span{b * buffer_frames, (b + 1) * buffer_frames, 0, 0};
    return zcu_spans_upload(ctx, voices, num_buffers * buffer_frames, off.data(), sp.data(), out);
}
void zcu_spans_free(zcu_spans* s) {
    if (!s) return;
    cudaStreamSynchronize(s->ctx->stream);
    if (s->d_off) cudaFree(s->d_off);
    if (s->d_spans) cudaFree(s->d_spans);
    delete s;
}

// ---- basics ------------------------------------------------------------------------------------------
static int basic_call(zcu_ctx* ctx, int op, const zcu_spans* sp, zcu_buf* dest, const zcu_buf* a, const zcu_buf* b, float s, bool need_a, bool need_b) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, dest != nullptr, "null dest");
    ZCU_REQUIRE(ctx, !need_a || a, "null source a");
    ZCU_REQUIRE(ctx, !need_b || b, "null source b");
    ZCU_REQUIRE(ctx, !a || (a->frames == dest->frames && a->voices == dest->voices), "shape mismatch (a)");
    ZCU_REQUIRE(ctx, !b || (b->frames == dest->frames && b->voices == dest->voices), "shape mismatch (b)");
    ZCU_REQUIRE(ctx, !sp || (sp->voices == dest->voices && sp->frames == dest->frames), "span table / buffer shape mismatch");
    return zcu_launch_basic(ctx, op, sp, dest, a, b, s);
}
int zcu_zero(zcu_ctx* c, const zcu_spans* s, zcu_buf* d) { return basic_call(c, 0, s, d, nullptr, nullptr, 0.0f, false, false); }
int zcu_set(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, float a) { return basic_call(c, 1, s, d, nullptr, nullptr, a, false, false); }
int zcu_copy(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* src) { return basic_call(c, 2, s, d, src, nullptr, 0.0f, true, false); }
int zcu_add(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, const zcu_buf* b) { return basic_call(c, 3, s, d, a, b, 0.0f, true, true); }
int zcu_add_into(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* src) { return basic_call(c, 4, s, d, src, nullptr, 0.0f, true, false); }
int zcu_add_scalar(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, float b) { return basic_call(c, 5, s, d, a, nullptr, b, true, false); }
int zcu_add_scalar_into(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, float a) { return basic_call(c, 6, s, d, nullptr, nullptr, a, false, false); }
int zcu_multiply(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, const zcu_buf* b) { return basic_call(c, 7, s, d, a, b, 0.0f, true, true); }
int zcu_multiply_with(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a) { return basic_call(c, 8, s, d, a, nullptr, 0.0f, true, false); }
int zcu_multiply_scalar(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, float b) { return basic_call(c, 9, s, d, a, nullptr, b, true, false); }
int zcu_multiply_with_scalar(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, float a) { return basic_call(c, 10, s, d, nullptr, nullptr, a, false, false); }

// per-span scalars (scalars[span.param]); the span table is mandatory
static int basic_rows_call(zcu_ctx* ctx, int op, const zcu_spans* sp, zcu_buf* dest, const zcu_buf* a, bool need_a, const float* rows, uint32_t nrows) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, dest != nullptr, "null dest");
    ZCU_REQUIRE(ctx, sp != nullptr, "per-span scalars need a span table");
    ZCU_REQUIRE(ctx, !need_a || a, "null source a");
    ZCU_REQUIRE(ctx, !a || (a->frames == dest->frames && a->voices == dest->voices), "shape mismatch (a)");
    ZCU_REQUIRE(ctx, sp->voices == dest->voices && sp->frames == dest->frames, "span table / buffer shape mismatch");
    ZCU_REQUIRE(ctx, sp->nspans == 0 || (rows && sp->max_param < nrows), "a span refers to a scalar row beyond nscalars");
    if (sp->nspans == 0) return ZCU_OK;
    return zcu_launch_basic_rows(ctx, op, sp, dest, a, rows, nrows);
}
int zcu_set_rows(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const float* r, uint32_t n) { return basic_rows_call(c, 1, s, d, nullptr, false, r, n); }
int zcu_add_scalar_rows(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, const float* r, uint32_t n) { return basic_rows_call(c, 5, s, d, a, true, r, n); }
int zcu_add_scalar_into_rows(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const float* r, uint32_t n) { return basic_rows_call(c, 6, s, d, nullptr, false, r, n); }
int zcu_multiply_scalar_rows(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const zcu_buf* a, const float* r, uint32_t n) { return basic_rows_call(c, 9, s, d, a, true, r, n); }
int zcu_multiply_with_scalar_rows(zcu_ctx* c, const zcu_spans* s, zcu_buf* d, const float* r, uint32_t n) { return basic_rows_call(c, 10, s, d, nullptr, false, r, n); }

// ---- module paints ------------------------------------------------------------------------------------
static int paint_call(zcu_ctx* ctx, int kind, zcu_bank* bank, const zcu_spans* sp, const void* params, size_t row_bytes, uint32_t nparams,
                      zcu_buf* out, const zcu_buf* in0, const zcu_buf* in1, const zcu_buf* in2, bool need_in0, const void* aux_dev,
                      uint64_t aux_len, const zcu_sample* smp) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, sp && out, "null spans or output");
    ZCU_REQUIRE(ctx, sp->voices == out->voices && sp->frames == out->frames, "span table / output shape mismatch");
    const size_t ssz = zcu_bank_state_size(kind);
    if (ssz) {
        ZCU_REQUIRE(ctx, bank != nullptr, "this module needs a bank");
        ZCU_REQUIRE(ctx, bank->kind == kind && bank->voices == out->voices, "bank kind / voice count mismatch");
    }
    ZCU_REQUIRE(ctx, params != nullptr && nparams > 0, "params required");
    ZCU_REQUIRE(ctx, sp->nspans == 0 || sp->max_param < nparams, "a span refers to a params row beyond nparams");
    ZCU_REQUIRE(ctx, !need_in0 || in0, "input buffer required");
    const zcu_buf* ins[3] = {in0, in1, in2};
    for (int i = 0; i < 3; i++)
        ZCU_REQUIRE(ctx, !ins[i] || (ins[i]->frames == out->frames && ins[i]->voices == out->voices), "input buffer shape mismatch");
    void* dparams = nullptr;
    int rc = zcu_arena_upload(ctx, 0, params, row_bytes * nparams, &dparams);
    if (rc) return rc;
    PaintLaunch pl{};
    pl.kind = kind;
    pl.V = out->voices;
    pl.frames = out->frames;
    pl.off = sp->d_off;
    pl.spans = sp->d_spans;
    pl.params = dparams;
    if (kind == ZCU_MOD_ENVELOPE && out->frames >= 8192) {  // long blocks run as time segments that skip ahead
        rc = zcu_envelope_ramp_tabs(ctx, (const zcu_envelope_params*)params, nparams, &pl.ramp_tabs, &pl.n_ramp_tabs);
        if (rc) return rc;
    }
    pl.state = bank ? bank->d : nullptr;
    pl.out = out->d;
    pl.in0 = in0 ? in0->d : nullptr;
    pl.in1 = in1 ? in1->d : nullptr;
    pl.in2 = in2 ? in2->d : nullptr;
    pl.aux = aux_dev;
    pl.aux_len = aux_len;
    pl.line = bank ? bank->line : nullptr;
    pl.delay_samples = bank ? bank->delay_samples : 0;
    if (smp) {
        pl.aux = smp->d;
        pl.aux_len = smp->len;
        pl.smp_channels = smp->num_channels;
        pl.smp_rate = smp->sample_rate;
        pl.smp_format = smp->format;
    }
    return zcu_launch_paint(ctx, pl);
}

int zcu_paint_sine_osc(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_sine_params* p, uint32_t n, zcu_buf* out, const zcu_buf* freq, const zcu_buf* phase) {
    return paint_call(c, ZCU_MOD_SINE_OSC, b, s, p, sizeof *p, n, out, freq, phase, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_pulse_osc(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_pulse_params* p, uint32_t n, zcu_buf* out, const zcu_buf* freq) {
    return paint_call(c, ZCU_MOD_PULSE_OSC, b, s, p, sizeof *p, n, out, freq, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_trisaw_osc(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_trisaw_params* p, uint32_t n, zcu_buf* out, const zcu_buf* freq) {
    return paint_call(c, ZCU_MOD_TRISAW_OSC, b, s, p, sizeof *p, n, out, freq, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_noise(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_noise_params* p, uint32_t n, zcu_buf* out) {
    return paint_call(c, ZCU_MOD_NOISE, b, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_envelope(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_envelope_params* p, uint32_t n, zcu_buf* out) {
    return paint_call(c, ZCU_MOD_ENVELOPE, b, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_gate(zcu_ctx* c, const zcu_spans* s, const zcu_gate_params* p, uint32_t n, zcu_buf* out) {
    return paint_call(c, ZCU_MOD_GATE, nullptr, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_filter(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_filter_params* p, uint32_t n, zcu_buf* out, const zcu_buf* input,
                     const zcu_buf* cutoff, const zcu_buf* res) {
    return paint_call(c, ZCU_MOD_FILTER, b, s, p, sizeof *p, n, out, input, cutoff, res, true, nullptr, 0, nullptr);
}
int zcu_paint_portamento(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_portamento_params* p, uint32_t n, zcu_buf* out) {
    return paint_call(c, ZCU_MOD_PORTAMENTO, b, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, nullptr, 0, nullptr);
}
int zcu_paint_curve(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_curve_params* p, uint32_t n, const zcu_curve_node* nodes,
                    uint32_t nnodes, zcu_buf* out) {
    if (!c) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(c, p != nullptr && n > 0, "params required");
    for (uint32_t i = 0; i < n; i++)
        ZCU_REQUIRE(c, (uint64_t)p[i].node_start + p[i].node_count <= nnodes, "curve params row refers to nodes beyond nnodes");
    void* dnodes = nullptr;
    int rc = zcu_arena_upload(c, 1, nodes, (size_t)nnodes * sizeof(zcu_curve_node), &dnodes);
    if (rc) return rc;
    return paint_call(c, ZCU_MOD_CURVE, b, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, dnodes, nnodes, nullptr);
}
int zcu_paint_sampler(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_sampler_params* p, uint32_t n, const zcu_sample* smp, zcu_buf* out) {
    if (!c) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(c, smp != nullptr, "null sample");
    return paint_call(c, ZCU_MOD_SAMPLER, b, s, p, sizeof *p, n, out, nullptr, nullptr, nullptr, false, nullptr, 0, smp);
}
int zcu_paint_decimator(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_decimator_params* p, uint32_t n, zcu_buf* out, const zcu_buf* input) {
    return paint_call(c, ZCU_MOD_DECIMATOR, b, s, p, sizeof *p, n, out, input, nullptr, nullptr, true, nullptr, 0, nullptr);
}
int zcu_paint_distortion(zcu_ctx* c, const zcu_spans* s, const zcu_distortion_params* p, uint32_t n, zcu_buf* out, const zcu_buf* input) {
    if (!c) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(c, p != nullptr && n > 0, "params required");
    // per-call host scalars of Distortion.zig:41-46 (pow, atan): evaluated once per params row on the host
    std::vector<zcu::DistortionDerived> d(n);
    for (uint32_t i = 0; i < n; i++) {
        const float gain1 = powf(2.0f, p[i].ingain * 8.0f - 2.0f);
        d[i].type = p[i].type;
        d[i].gain1 = gain1;
        d[i].gain2 = p[i].type == ZCU_DIST_OVERDRIVE ? p[i].outgain / atanf(gain1) : p[i].outgain;
        d[i].offs = gain1 * p[i].offset;
    }
    return paint_call(c, ZCU_MOD_DISTORTION, nullptr, s, d.data(), sizeof(zcu::DistortionDerived), n, out, input, nullptr, nullptr, true, nullptr, 0, nullptr);
}
int zcu_paint_cycle(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_cycle_params* p, uint32_t n, zcu_buf* out, const zcu_buf* speed) {
    return paint_call(c, ZCU_MOD_CYCLE, b, s, p, sizeof *p, n, out, speed, nullptr, nullptr, false, nullptr, 0, nullptr);
}

int zcu_paint_fm_oscillator(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_fm_oscillator_params* p, uint32_t n, zcu_buf* out,
                            const zcu_buf* freq, const zcu_buf* phase) {
    return paint_call(c, ZCU_MOD_FM_OSCILLATOR, b, s, p, sizeof *p, n, out, freq, phase, nullptr, true, nullptr, 0, nullptr);
}
int zcu_paint_simple_delay(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, zcu_buf* out, const zcu_buf* input) {
    // SimpleDelay.Params holds nothing but the input slice (modules.zig:345-347): whatever rows the spans name are empty
    const std::vector<uint32_t> rows(s ? (size_t)s->max_param + 1 : 1, 0u);
    if (c && b) ZCU_REQUIRE(c, b->line != nullptr, "bank has no delay line (zcu_delay_bank_create)");
    return paint_call(c, ZCU_MOD_SIMPLE_DELAY, b, s, rows.data(), sizeof(uint32_t), (uint32_t)rows.size(), out, input, nullptr, nullptr, true, nullptr, 0,
                      nullptr);
}
int zcu_paint_filtered_echoes(zcu_ctx* c, zcu_bank* b, const zcu_spans* s, const zcu_echoes_params* p, uint32_t n, zcu_buf* out,
                              const zcu_buf* input) {
    if (c && b) ZCU_REQUIRE(c, b->line != nullptr, "bank has no delay line (zcu_delay_bank_create)");
    return paint_call(c, ZCU_MOD_FILTERED_ECHOES, b, s, p, sizeof *p, n, out, input, nullptr, nullptr, true, nullptr, 0, nullptr);
}

float zcu_filter_cutoff_from_frequency(float frequency, float sample_rate) {  // Filter.zig:20-23
    const float pi = 3.14159265358979323846f;
    const float v = 2.0f * (1.0f - cosf(pi * frequency / sample_rate));
    return sqrtf(fmaxf(0.0f, fminf(v, 1.0f)));
}

int zcu_sample_create(zcu_ctx* ctx, const uint8_t* data, size_t len, uint32_t num_channels, uint32_t sample_rate, int format, zcu_sample** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, data != nullptr && len > 0, "empty sample");
    ZCU_REQUIRE(ctx, num_channels > 0, "num_channels must be > 0");
    ZCU_REQUIRE(ctx, format >= ZCU_FMT_UNSIGNED8 && format <= ZCU_FMT_SIGNED32_LSB, "unknown sample format");
    uint8_t* d = nullptr;
    ZCU_CUDA(ctx, cudaMalloc(&d, len));
    const cudaError_t e = cudaMemcpyAsync(d, data, len, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) {
        cudaFree(d);
        return zcu_fail(ctx, ZCU_ERR_CUDA, "sample upload failed: %s", cudaGetErrorString(e));
    }
    *out = new zcu_sample{ctx, d, len, num_channels, sample_rate, format};
    return ZCU_OK;
}
void zcu_sample_free(zcu_sample* s) {
    if (!s) return;
    cudaStreamSynchronize(s->ctx->stream);
    cudaFree(s->d);
    delete s;
}

}  // extern "C"
