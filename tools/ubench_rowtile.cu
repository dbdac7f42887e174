// ubench_rowtile.cu — memory-pipeline experiment for the lockstep paint kernel (not product code).
// Workload: out[f][v] += 1 over a [frames][V] f32 buffer, one thread per voice walking the frame axis (the access
// pattern of every paint kernel), in time segments so the chip is full.  Variants:
//   A  one-warp CTAs, 4-byte cp.async per lane and frame, direct 4-byte stores (the r01 kernel's pipeline)
//   B  W-warp CTAs over 32*W adjacent voices, 16-byte cp.async rows staged in smem, in-place update, 16-byte row stores
//   C  as B with 1-D TMA bulk copies (cp.async.bulk + mbarrier) in and bulk stores out
//   D  as B with ONE 2-D tensor-map TMA copy per tile ([T][32*W] box, cp.async.bulk.tensor.2d) in and one out (round 2)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_rowtile tools/ubench_rowtile.cu
#include <cuda.h>
#include <cuda_pipeline.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void __launch_bounds__(32) k_A(float* out, uint32_t V, uint32_t frames, uint32_t T, uint32_t seg) {
    extern __shared__ float smem[];
    const uint32_t tid = threadIdx.x, v = blockIdx.x * 32 + tid;
    const uint32_t f_begin = blockIdx.y * seg, f_end = min(frames, f_begin + seg);
    auto issue = [&](uint32_t tile) {
        const uint32_t f0 = tile * T, n = min(T, f_end - f0), st = tile & 1;
        const float* src = out + (size_t)f0 * V + v;
        for (uint32_t j = 0; j < n; j++) __pipeline_memcpy_async(smem + (st * T + j) * 32 + tid, src + (size_t)j * V, 4);
        __pipeline_commit();
    };
    const uint32_t t0 = f_begin / T, nt = (f_end + T - 1) / T;
    issue(t0);
    for (uint32_t tile = t0; tile < nt; tile++) {
        if (tile + 1 < nt) { issue(tile + 1); __pipeline_wait_prior(1); } else __pipeline_wait_prior(0);
        const uint32_t f0 = tile * T, n = min(T, f_end - f0), st = tile & 1;
        float* op = out + (size_t)f0 * V + v;
#pragma unroll 4
        for (uint32_t j = 0; j < n; j++) op[(size_t)j * V] = smem[(st * T + j) * 32 + tid] + 1.0f;
    }
}

// B: W warps, tile rows of 32*W voices; S stages.
template <int W, int S>
__global__ void __launch_bounds__(32 * W) k_B(float* out, uint32_t V, uint32_t frames, uint32_t T, uint32_t seg) {
    extern __shared__ __align__(16) float smem[];
    constexpr uint32_t RV = 32 * W;  // voices per row
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t v0 = blockIdx.x * RV;
    const uint32_t f_begin = blockIdx.y * seg, f_end = min(frames, f_begin + seg);
    // row copies: a row is RV*4 bytes = W*8 16-byte chunks; thread t copies chunk (t % (W*8)) of rows (t / (W*8)) + k*(32*W/(W*8)) = +k*4
    constexpr uint32_t CPR = W * 8;        // chunks per row
    constexpr uint32_t RPP = 32 * W / CPR;  // rows per pass = 4
    const uint32_t c = tid % CPR, r0 = tid / CPR;
    auto issue = [&](uint32_t tile) {
        const uint32_t f0 = tile * T, n = min(T, f_end - f0), st = tile % S;
        const float* src = out + (size_t)f0 * V + v0 + c * 4;
        float* dst = smem + (size_t)st * T * RV + c * 4;
        for (uint32_t j = r0; j < n; j += RPP) __pipeline_memcpy_async(dst + j * RV, src + (size_t)j * V, 16);
        __pipeline_commit();
    };
    const uint32_t t0 = f_begin / T, nt = (f_end + T - 1) / T;
    for (int s = 0; s < S - 1; s++) { if (t0 + s < nt) issue(t0 + s); else __pipeline_commit(); }
    for (uint32_t tile = t0; tile < nt; tile++) {
        __pipeline_wait_prior(S - 2);
        __syncthreads();  // tile landed for everyone; everyone finished storing tile-1 (stage to be refilled)
        if (tile + S - 1 < nt) issue(tile + S - 1); else __pipeline_commit();
        const uint32_t f0 = tile * T, n = min(T, f_end - f0), st = tile % S;
        float* sp = smem + (size_t)st * T * RV + warp * 32 + lane;
#pragma unroll 4
        for (uint32_t j = 0; j < n; j++) sp[j * RV] += 1.0f;
        __syncthreads();
        float* dstg = out + (size_t)f0 * V + v0 + c * 4;
        const float* s4 = smem + (size_t)st * T * RV + c * 4;
        for (uint32_t j = r0; j < n; j += RPP) *reinterpret_cast<float4*>(dstg + (size_t)j * V) = *reinterpret_cast<const float4*>(s4 + j * RV);
    }
}

// C: TMA 1-D bulk rows in, bulk rows out.
__device__ __forceinline__ uint32_t s2u(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s2u(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s2u(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(s2u(b)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s2u(dst)), "l"(src), "r"(bytes), "r"(s2u(b)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(s2u(src)), "r"(bytes) : "memory");
}

template <int W, int S>
__global__ void __launch_bounds__(32 * W) k_C(float* out, uint32_t V, uint32_t frames, uint32_t T, uint32_t seg) {
    extern __shared__ __align__(128) float smem[];
    __shared__ uint64_t full[S];
    constexpr uint32_t RV = 32 * W;
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t v0 = blockIdx.x * RV;
    const uint32_t f_begin = blockIdx.y * seg, f_end = min(frames, f_begin + seg);
    if (tid == 0) { for (int s = 0; s < S; s++) mbar_init(&full[s], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    // warp 0 issues loads: lane l issues rows l, l+32, ...
    auto issue = [&](uint32_t tile) {
        const uint32_t f0 = tile * T, n = min(T, f_end - f0), st = tile % S;
        if (lane == 0) mbar_expect(&full[st], n * RV * 4);
        __syncwarp();
        for (uint32_t j = lane; j < n; j += 32) bulk_g2s(smem + ((size_t)st * T + j) * RV, out + (size_t)(f0 + j) * V + v0, RV * 4, &full[st]);
    };
    const uint32_t t0 = f_begin / T, nt = (f_end + T - 1) / T;
    if (warp == 0) for (int s = 0; s < S - 1; s++) if (t0 + s < nt) issue(t0 + s);
    for (uint32_t tile = t0; tile < nt; tile++) {
        const uint32_t it = tile - t0, st = tile % S;
        // stage (tile+S-1)%S was stored from at iteration tile-1: its bulk stores must have finished reading smem
        if (warp == 0) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncthreads();
        if (warp == 0 && tile + S - 1 < nt) issue(tile + S - 1);
        mbar_wait(&full[st], (it / S) & 1);
        const uint32_t f0 = tile * T, n = min(T, f_end - f0);
        float* sp = smem + (size_t)st * T * RV + warp * 32 + lane;
#pragma unroll 4
        for (uint32_t j = 0; j < n; j++) sp[j * RV] += 1.0f;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (warp == 0) {
            for (uint32_t j = lane; j < n; j += 32) bulk_s2g(out + (size_t)(f0 + j) * V + v0, smem + ((size_t)st * T + j) * RV, RV * 4);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (warp == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// D: one 2-D TMA load and one 2-D TMA store per tile; one elected thread issues both
template <int W, int S>
__global__ void __launch_bounds__(32 * W) k_D(const CUtensorMap* __restrict__ tmap_p, uint32_t V, uint32_t frames, uint32_t T, uint32_t seg) {
    // the tensor copies need their shared-memory tile on a 128-byte boundary: no static shared memory in front of the tiles,
    // the base rounded up by hand, the mbarriers behind the tiles
    extern __shared__ float smem_raw[];
    constexpr uint32_t RV = 32 * W;
    float* smem = smem_raw + ((128u - (s2u(smem_raw) & 127u)) & 127u) / 4;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)S * T * RV);
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t v0 = blockIdx.x * RV;
    const uint32_t f_begin = blockIdx.y * seg, f_end = min(frames, f_begin + seg);
    if (tid == 0) { for (int s = 0; s < S; s++) mbar_init(&full[s], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    auto issue = [&](uint32_t tile) {
        const uint32_t st = tile % S;
        mbar_expect(&full[st], T * RV * 4);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         s2u(smem + (size_t)st * T * RV)),
                     "l"(tmap_p), "r"(v0), "r"(tile * T), "r"(s2u(&full[st]))
                     : "memory");
    };
    const uint32_t t0 = f_begin / T, nt = (f_end + T - 1) / T;
    if (tid == 0) for (int s = 0; s < S - 1; s++) if (t0 + s < nt) issue(t0 + s);
    for (uint32_t tile = t0; tile < nt; tile++) {
        const uint32_t it = tile - t0, st = tile % S;
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
        if (tid == 0 && tile + S - 1 < nt) issue(tile + S - 1);
        mbar_wait(&full[st], (it / S) & 1);
        float* sp = smem + (size_t)st * T * RV + warp * 32 + lane;
#pragma unroll 4
        for (uint32_t j = 0; j < T; j++) sp[j * RV] += 1.0f;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap_p), "r"(v0), "r"(tile * T),
                         "r"(s2u(smem + (size_t)st * T * RV))
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static CUtensorMap make_map(float* buf, uint32_t V, uint32_t frames, uint32_t T, uint32_t RV) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        cudaDriverEntryPointQueryResult q;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&fn, cudaEnableDefault, &q));
    }
    CUtensorMap m;
    const cuuint64_t dims[2] = {V, frames}, strides[1] = {(cuuint64_t)V * 4};
    const cuuint32_t box[2] = {RV, T}, estr[2] = {1, 1};
    const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", (int)r); exit(1); }
    return m;
}

static float time_it(void (*launch)(float*, uint32_t, uint32_t), float* buf, uint32_t V, uint32_t frames, float* flush, size_t flush_n) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        CK(cudaMemsetAsync(flush, 0, flush_n));
        CK(cudaEventRecord(e0));
        launch(buf, V, frames);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep) best = ms < best ? ms : best;
    }
    return best;
}

static uint32_t g_T = 32, g_nseg = 1;
template <int W, int S> static void launch_B(float* b, uint32_t V, uint32_t frames) {
    uint32_t seg = (frames + g_nseg - 1) / g_nseg; seg = (seg + g_T - 1) / g_T * g_T;
    size_t sm = (size_t)S * g_T * 32 * W * 4;
    CK(cudaFuncSetAttribute(k_B<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k_B<W, S><<<dim3(V / (32 * W), (frames + seg - 1) / seg), 32 * W, sm>>>(b, V, frames, g_T, seg);
}
template <int W, int S> static void launch_C(float* b, uint32_t V, uint32_t frames) {
    uint32_t seg = (frames + g_nseg - 1) / g_nseg; seg = (seg + g_T - 1) / g_T * g_T;
    size_t sm = (size_t)S * g_T * 32 * W * 4;
    CK(cudaFuncSetAttribute(k_C<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k_C<W, S><<<dim3(V / (32 * W), (frames + seg - 1) / seg), 32 * W, sm>>>(b, V, frames, g_T, seg);
}
template <int W, int S> static void launch_D(float* b, uint32_t V, uint32_t frames) {
    uint32_t seg = (frames + g_nseg - 1) / g_nseg; seg = (seg + g_T - 1) / g_T * g_T;
    size_t sm = (size_t)S * g_T * 32 * W * 4 + 128 + 8 * S;
    const CUtensorMap m = make_map(b, V, frames, g_T, 32 * W);
    static CUtensorMap* dm = nullptr;
    if (!dm) CK(cudaMalloc(&dm, sizeof(CUtensorMap)));
    CK(cudaMemcpy(dm, &m, sizeof m, cudaMemcpyHostToDevice));
    CK(cudaFuncSetAttribute(k_D<W, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k_D<W, S><<<dim3(V / (32 * W), (frames + seg - 1) / seg), 32 * W, sm>>>(dm, V, frames, g_T, seg);
}
static void launch_A(float* b, uint32_t V, uint32_t frames) {
    uint32_t seg = (frames + g_nseg - 1) / g_nseg; seg = (seg + g_T - 1) / g_T * g_T;
    k_A<<<dim3(V / 32, (frames + seg - 1) / seg), 32, 2 * g_T * 32 * 4>>>(b, V, frames, g_T, seg);
}

int main() {
    const size_t flush_n = 512u << 20;
    float* flush;
    CK(cudaMalloc(&flush, flush_n));
    struct Cfg { uint32_t V, frames; } cfgs[] = {{65536, 8192}, {4096, 65536}};
    for (auto cfg : cfgs) {
        const size_t n = (size_t)cfg.V * cfg.frames;
        float* buf;
        CK(cudaMalloc(&buf, n * 4));
        CK(cudaMemset(buf, 0, n * 4));
        const double gb = n * 8.0 / 1e9;
        auto report = [&](const char* name, float ms) { printf("V=%6u frames=%6u %-28s T=%3u nseg=%3u  %8.3f ms  %7.1f GB/s\n", cfg.V, cfg.frames, name, g_T, g_nseg, ms, gb / (ms * 1e-3)); fflush(stdout); };
        // correctness of each variant: count increments
        int launches = 0;
        // ROWTILE_ONESEG=1: only the 4096-voice buffer, cut into 1, 2 or 4 segments (what a launch looks like that cannot be
        // cut in time: few CTAs, all of them walking the same narrow window of rows)
        const bool oneseg = getenv("ROWTILE_ONESEG") != nullptr;
        if (oneseg && cfg.V != 4096) { CK(cudaFree(buf)); continue; }
        for (uint32_t nseg : {1u, 2u, 4u, 16u, 27u, 64u}) {
            if (oneseg != (nseg == 1 || nseg == 2 || (nseg == 4 && cfg.V == 4096))) { if (!(cfg.V == 65536 && (nseg == 1 || nseg == 4) && !oneseg)) continue; }
            if (cfg.V == 65536 && nseg > 4) continue;
            if (cfg.V == 4096 && nseg < 16 && !oneseg) continue;
            g_nseg = nseg;
            for (uint32_t T : {16u, 32u, 64u}) {
                g_T = T;
                if (T <= 64) { report("A 1warp 4B cp.async", time_it(launch_A, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4; }
                report("B W=4 S=2 16B cp.async", time_it(launch_B<4, 2>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("B W=4 S=3", time_it(launch_B<4, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("B W=8 S=3", time_it(launch_B<8, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("B W=2 S=3", time_it(launch_B<2, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("C W=4 S=3 TMA bulk", time_it(launch_C<4, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("C W=8 S=3 TMA bulk", time_it(launch_C<8, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("C W=4 S=4 TMA bulk", time_it(launch_C<4, 4>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("D W=4 S=3 TMA 2-D tile", time_it(launch_D<4, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("D W=4 S=4 TMA 2-D tile", time_it(launch_D<4, 4>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
                report("D W=8 S=3 TMA 2-D tile", time_it(launch_D<8, 3>, buf, cfg.V, cfg.frames, flush, flush_n)); launches += 4;
            }
        }
        // verify: every element must equal the number of launches
        std::vector<float> h(1 << 20);
        CK(cudaMemcpy(h.data(), buf + (n - h.size()), h.size() * 4, cudaMemcpyDeviceToHost));
        size_t bad = 0;
        for (float x : h) bad += (x != (float)launches);
        CK(cudaMemcpy(h.data(), buf, h.size() * 4, cudaMemcpyDeviceToHost));
        for (float x : h) bad += (x != (float)launches);
        printf("verify: launches=%d mismatches=%zu (first=%g)\n", launches, bad, h[0]);
        CK(cudaFree(buf));
    }
    return 0;
}
