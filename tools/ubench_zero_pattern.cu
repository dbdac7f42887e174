// ubench_zero_pattern.cu — why does the span-restricted zero reach only 3.1 TB/s of writes?  Pure 16-byte stores of a constant
// into a [frames][V] f32 buffer in the lockstep kernel's tile pattern (a CTA owns 32*W adjacent voices and walks a time segment
// in tiles of T frames: row pieces of 128*W bytes at a stride of 4*V bytes), for several W, against a flat grid-stride memset.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
template <int W>
__global__ void __launch_bounds__(32 * W) k_tile(float* out, uint32_t V, uint32_t frames, uint32_t T, uint32_t seg) {
    constexpr uint32_t RV = 32 * W, CPR = RV / 4, RPP = RV / CPR;
    const uint32_t tid = threadIdx.x, c = (tid % CPR) * 4, r0 = tid / CPR;
    const uint32_t v0 = blockIdx.x * RV, f_begin = blockIdx.y * seg, f_end = min(frames, f_begin + seg);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t f0 = f_begin; f0 < f_end; f0 += T) {
        const uint32_t n = min(T, f_end - f0);
        float* dst = out + (size_t)f0 * V + v0 + c;
        for (uint32_t j = r0; j < n; j += RPP) *reinterpret_cast<float4*>(dst + (size_t)j * V) = z;
        __syncthreads();  // the real kernel synchronises per tile
    }
}
__global__ void k_flat(float4* out, size_t n4) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) out[i] = z;
}
template <class F>
float timed(F f, float* flush, size_t fn) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaMemset(flush, 1, fn);
        cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep) best = ms < best ? ms : best;
    }
    return best;
}
template <int W>
void run(float* buf, uint32_t V, uint32_t frames, float* flush, size_t fn, int sms) {
    const uint32_t gx = V / (32 * W);
    for (uint32_t per_sm : {8u * 4 / W > 0 ? 8u * 4 / W : 1u}) {
        uint32_t nseg = per_sm * sms / gx; if (nseg < 1) nseg = 1;
        uint32_t seg = (frames + nseg - 1) / nseg; seg = (seg + 15) / 16 * 16; nseg = (frames + seg - 1) / seg;
        const float ms = timed([&] { k_tile<W><<<dim3(gx, nseg), 32 * W>>>(buf, V, frames, 16, seg); }, flush, fn);
        printf("V=%u tile W=%2d (%4d-byte row pieces) nseg=%3u  %.3f ms  %.0f GB/s\n", V, W, 128 * W, nseg, ms, (double)V * frames * 4 / ms / 1e6);
    }
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const size_t fn = 512u << 20; float* flush; cudaMalloc(&flush, fn);
    for (uint32_t V : {4096u, 65536u}) {
        const uint32_t frames = V == 4096 ? 65536 : 8192;
        const size_t n = (size_t)V * frames; float* buf; cudaMalloc(&buf, n * 4);
        const float ms = timed([&] { k_flat<<<p.multiProcessorCount * 8, 256>>>((float4*)buf, n / 4); }, flush, fn);
        printf("V=%u flat grid-stride memset          %.3f ms  %.0f GB/s\n", V, ms, (double)n * 4 / ms / 1e6);
        run<4>(buf, V, frames, flush, fn, p.multiProcessorCount); run<8>(buf, V, frames, flush, fn, p.multiProcessorCount);
        run<16>(buf, V, frames, flush, fn, p.multiProcessorCount); run<32>(buf, V, frames, flush, fn, p.multiProcessorCount);
        cudaFree(buf);
    }
    return 0;
}
