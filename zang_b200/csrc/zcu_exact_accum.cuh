// zcu_exact_accum.cuh — exact closed form of the reference's f32 accumulators.
//
// painter.zig:102-106 runs `t += t_step; if (t >= 1) { t = 1; finished }` once per sample in f32.  A
// closed form `n * t_step` drifts from that by up to ~1e-4, but repeated f32 addition of
// a constant IS piecewise closed-form: while t stays in one binade [2^e, 2^(e+1)) with U = ulp(t), every add
// moves the 24-bit significand by the same integer d = round-to-nearest-even(t_step / U).  So a ramp is
// at most ~25 (binade, d) segments; inside a segment n adds are one integer multiply-add.  The add that
// leaves a binade (and everything unusual: zero, subnormals, ties on an odd significand, non-positive or
// non-finite steps) is done as one real f32 add, so the result is the reference's bit pattern by construction.
//
// Host + device: the same code is unit-tested on the CPU against the plain loop (tests/test_exact_accum.py).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define ZCU_HD __host__ __device__ __forceinline__
#else
#define ZCU_HD inline
#endif

namespace zcu {

ZCU_HD uint32_t ea_bits(float x) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(x);
#else
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
#endif
}
ZCU_HD float ea_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float x;
    memcpy(&x, &u, 4);
    return x;
#endif
}
// one real f32 add that the compiler may not contract or reassociate
ZCU_HD float ea_add(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    volatile float r = a + b;
    return r;
#endif
}

// Applies `t += step` up to n times with the reference's exact f32 rounding, in O(binades) instead of O(n).
//   STOP   stop after the add that makes t >= 1 (*finished = true); CLAMP additionally sets t = 1 then
//          (painter.zig:102-106).  Without STOP every one of the n adds is performed (SineOsc.zig:49-62, Sampler.zig:130).
// Returns the number of adds performed.
//   ONE    return after the first piece (one in-binade jump or one real add): the table builder of the bank render's
//          pre-pass walks a ramp piece by piece with it (zcu_render_scan.cuh, RampTab).
template <bool STOP, bool CLAMP, bool ONE = false>
ZCU_HD uint32_t accum_core(float& t, float step, uint32_t n, bool& finished) {
    finished = false;
    uint32_t done = 0;
    const uint32_t sb = ea_bits(step);
    const uint32_t es = (sb >> 23) & 0xFFu;
    const bool fast_ok = (sb >> 31) == 0 && es != 0 && es != 0xFFu;  // positive, normal, finite step
    const uint32_t ms = (sb & 0x7FFFFFu) | 0x800000u;
    while (done < n) {
        const uint32_t tb = ea_bits(t);
        const uint32_t et = (tb >> 23) & 0xFFu;
        uint32_t d = 0, k = 0;
        bool jump = false;
        // t positive and normal (below 1 when the clock stops there), step's exponent below t's
        if (fast_ok && (tb >> 31) == 0 && et != 0 && et < (STOP ? 127u : 0xFFu) && es < et) {
            const uint32_t shift = et - es;  // >= 1
            const uint32_t m = (tb & 0x7FFFFFu) | 0x800000u;
            bool settled = true;
            if (shift > 24) {
                d = 0;
            } else {
                const uint32_t q = ms >> shift;
                const uint32_t rem = ms & ((1u << shift) - 1u);
                const uint32_t half = 1u << (shift - 1);
                if (rem > half)
                    d = q + 1;
                else if (rem < half)
                    d = q;
                else if (m & 1u)
                    settled = false;  // tie on an odd significand: one real add makes it even
                else
                    d = (q & 1u) ? q + 1 : q;
            }
            if (settled) {
                if (d == 0) return n;  // step is below half an ulp: t never moves again
                // adds that stay inside the binade: min(room / d, n - done), without the division when the rest of the run fits
                const uint32_t room = 0xFFFFFFu - m, want = n - done;
                k = (unsigned long long)want * d <= room ? want : room / d;
                if (k > 0) {
                    t = ea_float((et << 23) | ((m + k * d) & 0x7FFFFFu));
                    done += k;
                    jump = true;
                }
            }
        }
        // no jump: one real add.  After a jump the next add usually leaves the binade: it is done here as a real add too
        // (a real add is always right), so that a binade costs one trip round this loop instead of two
        if (!jump || (!ONE && done < n)) {
            t = ea_add(t, step);
            done++;
            if (STOP && t >= 1.0f) {
                if (CLAMP) t = 1.0f;
                finished = true;
                return done;
            }
        }
        if (ONE) return done;
    }
    return done;
}

// Painter's ramp clock: `t += step; if (t >= 1) { t = 1; finished }` up to n times, stopping after the add that finishes.
ZCU_HD uint32_t ramp_advance(float& t, float step, uint32_t n, bool& finished) { return accum_core<true, true>(t, step, n, finished); }

// the first piece of the same walk: up to n adds inside t's binade in one jump, or one real add
ZCU_HD uint32_t ramp_advance_piece(float& t, float step, uint32_t n, bool& finished) { return accum_core<true, true, true>(t, step, n, finished); }

// n plain adds (SineOsc's and Sampler's phase between the wraps at the end of a paint call)
ZCU_HD void accum_advance(float& t, float step, uint32_t n) {
    bool f;
    accum_core<false, false>(t, step, n, f);
}

// Cycle.zig:44-46: `t += step; t -= trunc(t)` n times.  Between wraps it is a plain accumulator below 1.
ZCU_HD void cycle_advance(float& t, float step, uint32_t n) {
    if (!(step > 0.0f) || !(step < 0.00390625f) || !(t >= 0.0f) || !(t < 1.0f)) {  // fast wraps (or odd values): the plain loop
        for (uint32_t i = 0; i < n; i++) {
            t = ea_add(t, step);
            t = t - truncf(t);
        }
        return;
    }
    while (n > 0) {
        bool wrapped;
        const uint32_t k = accum_core<true, false>(t, step, n, wrapped);
        n -= k;
        if (wrapped) t = t - truncf(t);
    }
}

}  // namespace zcu
