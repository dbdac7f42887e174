// ubench_f32x2.cu — do Blackwell's packed f32x2 instructions (FADD2 / FMUL2 / FFMA2, PTX add/mul/fma.rn.f32x2) buy issue
// slots for the Filter recurrence (Filter.zig:135-146: 16 dependent strict-f32 operations per sample, no contraction)?
//   scalar : one voice per lane, separate multiply and add (__fmul_rn / __fadd_rn)
//   packed : two voices per lane, the same operations on register pairs.  ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into
//            FFMA2 even with --fmad=false, so the product is written as fma(x, y, nz) with nz = {-0.0, -0.0} read from a
//            kernel argument: x*y + (-0) == x*y bit for bit (sign of zero included), and ptxas cannot fold what it cannot see.
// Every lane checks scalar == packed bit for bit.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0,{%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1},%2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b, u64 nz) { u64 r; asm("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(nz)); return r; }

struct F { float l, b; };
__device__ __forceinline__ float filt(F& s, float x, float cut, float res) {
    const float dc = 3.814697265625e-6f;
    const float in = __fadd_rn(x, dc);
    s.l = __fadd_rn(s.l, __fsub_rn(__fmul_rn(cut, s.b), dc));
    s.b = __fadd_rn(s.b, __fmul_rn(cut, __fsub_rn(__fsub_rn(in, __fmul_rn(s.b, res)), s.l)));
    s.l = __fadd_rn(s.l, __fmul_rn(cut, s.b));
    const float h = __fsub_rn(__fsub_rn(in, __fmul_rn(s.b, res)), s.l);
    s.b = __fadd_rn(s.b, __fmul_rn(cut, h));
    return s.l;
}
__global__ void __launch_bounds__(128) k_scalar(float* out, int n, float cut0, float res) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    F s{0.f, 0.f};
    const float cut = cut0 + 1e-4f * (t & 31);
    float acc = 0.f, x = 0.35f;
#pragma unroll 4
    for (int i = 0; i < n; i++) {
        x = -x;
        acc = __fadd_rn(acc, filt(s, x, cut, res));
    }
    out[t] = acc + s.b;
}
__global__ void __launch_bounds__(128) k_packed(float* out, int n, float cut0, float res, u64 nz) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;  // lane holds voices 2t and 2t+1
    const float c0 = cut0 + 1e-4f * ((2 * t) & 31), c1 = cut0 + 1e-4f * ((2 * t + 1) & 31);
    const u64 cut = pk(c0, c1), rs = pk(res, res), dc = pk(3.814697265625e-6f, 3.814697265625e-6f);
    u64 l = pk(0.f, 0.f), b = pk(0.f, 0.f), acc = pk(0.f, 0.f);
    float x = 0.35f;
#pragma unroll 4
    for (int i = 0; i < n; i++) {
        x = -x;
        const u64 in = add2(pk(x, x), dc);
        l = add2(l, sub2(mul2(cut, b, nz), dc));
        b = add2(b, mul2(cut, sub2(sub2(in, mul2(b, rs, nz)), l), nz));
        l = add2(l, mul2(cut, b, nz));
        const u64 h = sub2(sub2(in, mul2(b, rs, nz)), l);
        b = add2(b, mul2(cut, h, nz));
        acc = add2(acc, l);
    }
    float a0, a1, b0, b1;
    upk(acc, a0, a1);
    upk(b, b0, b1);
    out[2 * t] = a0 + b0;
    out[2 * t + 1] = a1 + b1;
}
int main() {
    const int n = 1 << 16;
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const u64 nz = 0x8000000080000000ull;
    float *o1, *o2;
    for (int warps_per_smsp : {1, 2, 4, 6, 8}) {
        const int ctas = p.multiProcessorCount * warps_per_smsp;  // 128-thread CTAs: one warp per SMSP each
        const int voices = ctas * 128;
        cudaMalloc(&o1, voices * 4);
        cudaMalloc(&o2, voices * 4);
        cudaEvent_t e0, e1, e2;
        cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
        k_scalar<<<ctas, 128>>>(o1, 64, 0.2f, 0.3f);
        k_packed<<<ctas / 2, 128>>>(o2, 64, 0.2f, 0.3f, nz);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        k_scalar<<<ctas, 128>>>(o1, n, 0.2f, 0.3f);
        cudaEventRecord(e1);
        k_packed<<<ctas, 128>>>(o2, n, 0.2f, 0.3f, nz);  // same number of warps, twice the voices
        cudaEventRecord(e2);
        cudaDeviceSynchronize();
        float ms1, ms2;
        cudaEventElapsedTime(&ms1, e0, e1);
        cudaEventElapsedTime(&ms2, e1, e2);
        // correctness: voices of the packed run [0, voices) vs scalar (same cut per lane index & 31 pattern differs: compare directly)
        cudaFree(o2);
        cudaMalloc(&o2, 2 * voices * 4);
        k_packed<<<ctas, 128>>>(o2, 4096, 0.2f, 0.3f, nz);
        k_scalar<<<ctas, 128>>>(o1, 4096, 0.2f, 0.3f);
        float *h1 = new float[voices], *h2 = new float[2 * voices];
        cudaMemcpy(h1, o1, voices * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(h2, o2, 2 * voices * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int v = 0; v < 32; v++) {  // voice v of the scalar run has cut index v & 31; packed voice v likewise
            uint32_t a, b;
            memcpy(&a, &h1[v], 4);
            memcpy(&b, &h2[v], 4);
            bad += a != b;
        }
        const double vs1 = (double)voices * n, vs2 = 2.0 * voices * n;
        printf("warps/SMSP %d: scalar %.3f ms (%.3e voice-samples/s, %.2f cyc/sample/warp @1.965GHz)  packed %.3f ms (%.3e voice-samples/s)  speedup x%.2f  mismatches %d\n",
               warps_per_smsp, ms1, vs1 / ms1 * 1e3, ms1 * 1e-3 * 1.965e9 / n / warps_per_smsp, ms2, vs2 / ms2 * 1e3, (vs2 / ms2) / (vs1 / ms1), bad);
        delete[] h1; delete[] h2;
        cudaFree(o1); cudaFree(o2);
    }
    return 0;
}
