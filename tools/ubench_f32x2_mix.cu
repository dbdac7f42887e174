// ubench_f32x2_mix.cu — does a packed f32x2 instruction (FFMA2/FADD2) block the warp scheduler's issue port for its two pipe
// cycles, or can ALU-pipe instructions of other warps issue in its shadow?  Three kernels over the same FP work (the Filter
// recurrence of two voices) plus K independent integer ops per sample:
//   scalar : 2 x 16 scalar FP ops + K ALU ops      packed : 16 packed FP ops + K ALU ops      alu : K ALU ops alone
// If packed ~ max(2*16, K) cycles per sample the shadow is usable (fewer issue slots help a mixed kernel); if ~ 2*16 + K it is not.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0,{%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1},%2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b, u64 nz) { u64 r; asm("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(nz)); return r; }
template <int K>
__device__ __forceinline__ void alu(uint32_t (&z)[4], uint32_t c) {
#pragma unroll
    for (int k = 0; k < K; k++) z[k & 3] = (z[k & 3] ^ c) + (z[(k + 1) & 3] >> 3);  // LOP3 + SHF/IADD-ish, 4 independent chains
}
template <int K, int MODE>
__global__ void __launch_bounds__(128) k(float* out, int n, float cut0, float res, u64 nz, uint32_t c) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t z[4] = {(uint32_t)t, 2u * t, 3u * t, 5u * t};
    const float dcf = 3.814697265625e-6f;
    float x = 0.35f, r0 = 0.f, r1 = 0.f;
    if (MODE == 0) {
        float l0 = 0, b0 = 0, l1 = 0, b1 = 0;
        const float c0 = cut0 + 1e-4f * (t & 31), c1 = c0 + 1e-5f;
#pragma unroll 4
        for (int i = 0; i < n; i++) {
            x = -x;
            const float in = __fadd_rn(x, dcf);
            l0 = __fadd_rn(l0, __fsub_rn(__fmul_rn(c0, b0), dcf)); l1 = __fadd_rn(l1, __fsub_rn(__fmul_rn(c1, b1), dcf));
            b0 = __fadd_rn(b0, __fmul_rn(c0, __fsub_rn(__fsub_rn(in, __fmul_rn(b0, res)), l0))); b1 = __fadd_rn(b1, __fmul_rn(c1, __fsub_rn(__fsub_rn(in, __fmul_rn(b1, res)), l1)));
            l0 = __fadd_rn(l0, __fmul_rn(c0, b0)); l1 = __fadd_rn(l1, __fmul_rn(c1, b1));
            const float h0 = __fsub_rn(__fsub_rn(in, __fmul_rn(b0, res)), l0), h1 = __fsub_rn(__fsub_rn(in, __fmul_rn(b1, res)), l1);
            b0 = __fadd_rn(b0, __fmul_rn(c0, h0)); b1 = __fadd_rn(b1, __fmul_rn(c1, h1));
            alu<K>(z, c);
        }
        r0 = l0 + b0; r1 = l1 + b1;
    } else if (MODE == 1) {
        const float c0 = cut0 + 1e-4f * (t & 31), c1 = c0 + 1e-5f;
        const u64 cut = pk(c0, c1), rs = pk(res, res), dc = pk(dcf, dcf);
        u64 l = pk(0.f, 0.f), b = pk(0.f, 0.f);
#pragma unroll 4
        for (int i = 0; i < n; i++) {
            x = -x;
            const u64 in = add2(pk(x, x), dc);
            l = add2(l, sub2(mul2(cut, b, nz), dc));
            b = add2(b, mul2(cut, sub2(sub2(in, mul2(b, rs, nz)), l), nz));
            l = add2(l, mul2(cut, b, nz));
            const u64 h = sub2(sub2(in, mul2(b, rs, nz)), l);
            b = add2(b, mul2(cut, h, nz));
            alu<K>(z, c);
        }
        float a0, a1, b0, b1; upk(l, a0, a1); upk(b, b0, b1); r0 = a0 + b0; r1 = a1 + b1;
    } else {
#pragma unroll 4
        for (int i = 0; i < n; i++) alu<K>(z, c);
    }
    out[t] = r0 + r1 + (float)(z[0] ^ z[1] ^ z[2] ^ z[3]);
}
template <int K, int MODE>
float run(float* o, int ctas, int n) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<K, MODE><<<ctas, 128>>>(o, 64, 0.2f, 0.3f, 0x8000000080000000ull, 0x9e3779b9u);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<K, MODE><<<ctas, 128>>>(o, n, 0.2f, 0.3f, 0x8000000080000000ull, 0x9e3779b9u);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
template <int K>
void row(float* o, int sms, int wps, int n) {
    const int ctas = sms * wps;
    const float s = run<K, 0>(o, ctas, n), p = run<K, 1>(o, ctas, n), a = run<K, 2>(o, ctas, n);
    const double cyc = 1.965e9 * 1e-3 / n / wps;  // cycles per sample per warp slot of a scheduler
    printf("K=%2d ALU ops/sample, %d warps/SMSP: scalar(32 FP) %.1f cyc  packed(16 FP2) %.1f cyc  alu alone %.1f cyc   [scalar issue slots %d, packed %d]\n",
           K, wps, s * cyc, p * cyc, a * cyc, 32 + 2 * K + 2, 16 + 2 * K + 2);
}
int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    float* o; cudaMalloc(&o, (size_t)pr.multiProcessorCount * 8 * 128 * 4);
    const int n = 1 << 15;
    for (int wps : {4, 6}) { row<0>(o, pr.multiProcessorCount, wps, n); row<8>(o, pr.multiProcessorCount, wps, n); row<16>(o, pr.multiProcessorCount, wps, n); row<24>(o, pr.multiProcessorCount, wps, n); }
    return 0;
}
