// zcu_noalloc.cu — zcu_mix_down_static: zang.mixDown (src/zang/mixdown.zig:28-86) through device scratch that is part of
// the library image instead of the context's arena.
//
// zcu_mix_down converts the f32 mix to PCM bytes in a device staging buffer (arena slot 2, grown with cudaMallocAsync when a
// call needs more than any call before it) and copies the bytes to the host.  Callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same kernel arithmetic with the staging
// buffer in a file-scope __device__ array of fixed size: no cudaMalloc, cudaFree, cudaMallocAsync or any other allocating
// routine on this path.  Every GPU that loads libzangcu.so holds the ZCU_MIX_DOWN_STATIC_MAX_BYTES of the array until exit.
//
// The array is shared by every call and every context on a device and keeps the last call's bytes.  Nothing has to start at
// zero: a call writes all dst_len bytes it later copies out, and reads none it did not write.  Calls that overlap on the GPU
// (two contexts = two streams on one device) race on the array; calls on one context are stream-ordered and the entry point
// returns after its copy has completed.
#include <mutex>

#include "zcu_internal.h"

namespace {

static __device__ __align__(256) uint8_t g_mix_down_scratch[ZCU_MIX_DOWN_STATIC_MAX_BYTES];

// In host code g_mix_down_scratch is only a stand-in for the symbol: the device address is asked for once per device, with
// that device current, and cached.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t* g_scratch_ptr[kMaxDevices] = {};

int scratch_for(zcu_ctx* ctx, uint8_t** out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return zcu_fail(ctx, ZCU_ERR_INVALID, "device %d out of range", ctx->device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[ctx->device]) {
        zcu_use_device(ctx);
        void* p = nullptr;
        ZCU_CUDA(ctx, cudaGetSymbolAddress(&p, g_mix_down_scratch));
        g_scratch_ptr[ctx->device] = (uint8_t*)p;
    }
    *out = g_scratch_ptr[ctx->device];
    return ZCU_OK;
}

// same arithmetic as k_mix_down_s16 / k_mix_down_s8 (zcu_basics.cu): multiply, clamp, truncate
template <bool S16>
__global__ void k_mix_down_static(const float* __restrict__ mix, uint8_t* __restrict__ dst, size_t n, float mul) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float lo = S16 ? -32767.0f : -127.0f, hi = S16 ? 32766.0f : 126.0f;
    const float value = mix[i] * mul;
    int cv;
    if (value <= lo)
        cv = (int)lo;
    else if (value >= hi)
        cv = (int)hi;
    else if (value != value)
        cv = 0;
    else
        cv = (int)value;  // @intFromFloat: truncation
    if (S16) {
        dst[i * 2 + 0] = (uint8_t)(cv & 0xFF);
        dst[i * 2 + 1] = (uint8_t)((cv >> 8) & 0xFF);
    } else {
        dst[i] = (uint8_t)(int8_t)cv;
    }
}

}  // namespace

extern "C" int zcu_mix_down_static(zcu_ctx* ctx, const zcu_buf* mix, int audio_format, float vol, uint8_t* dst_host, size_t dst_len) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, mix && dst_host, "null argument");
    const size_t n = (size_t)mix->frames * mix->voices;
    const size_t bytes = audio_format == ZCU_AUDIO_SIGNED16_LSB ? n * 2 : n;
    ZCU_REQUIRE(ctx, audio_format == ZCU_AUDIO_SIGNED16_LSB || audio_format == ZCU_AUDIO_SIGNED8, "unknown audio format");
    ZCU_REQUIRE(ctx, dst_len == bytes, "dst_len must be frames*channels*bytes_per_sample (mixdown.zig:35,66)");
    if (bytes > (size_t)ZCU_MIX_DOWN_STATIC_MAX_BYTES)
        return zcu_fail(ctx, ZCU_ERR_INVALID, "zcu_mix_down_static: %zu bytes of PCM exceed the static scratch of %d bytes (use zcu_mix_down, or convert in pieces)",
                        bytes, (int)ZCU_MIX_DOWN_STATIC_MAX_BYTES);
    if (n == 0) return ZCU_OK;
    uint8_t* d = nullptr;
    int rc = scratch_for(ctx, &d);
    if (rc) return rc;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (audio_format == ZCU_AUDIO_SIGNED16_LSB)
        k_mix_down_static<true><<<blocks, 256, 0, ctx->stream>>>(mix->d, d, n, vol * 32767.0f);
    else
        k_mix_down_static<false><<<blocks, 256, 0, ctx->stream>>>(mix->d, d, n, vol * 127.0f);
    ctx->launches++;
    ZCU_CUDA(ctx, cudaGetLastError());
    ZCU_CUDA(ctx, cudaMemcpyAsync(dst_host, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}
