"""zcu_ramp_advance (zang_b200/csrc/zcu_exact_accum.cuh): the closed form of Painter's f32 ramp clock
(painter.zig:102-106) must give the plain loop's bits for every (t0, step, n).  SURVEY.md App. B."""
import ctypes as C

import numpy as np

from zang_b200._lib import lib


def loop(t, step, n):
    t, step = np.float32(t), np.float32(step)
    with np.errstate(all="ignore"):
        for i in range(n):
            t = np.float32(t + step)
            if t >= np.float32(1.0):
                return np.float32(1.0), i + 1, True
    return t, n, False


def closed(t, step, n):
    tt, fin = C.c_float(t), C.c_int32(0)
    done = lib.zcu_ramp_advance(C.byref(tt), float(step), n, C.byref(fin))
    return np.float32(tt.value), done, bool(fin.value)


def check(t0, step, n):
    a, b = loop(t0, step, n), closed(t0, step, n)
    assert a[1:] == b[1:] and np.float32(a[0]).view(np.uint32) == np.float32(b[0]).view(np.uint32), (t0, step, n, a, b)


def test_envelope_steps_every_prefix():
    # the t_steps of the example instruments (modules.zig:100-103,238-243) at 48 kHz and 44.1 kHz
    for sr in (48000.0, 44100.0):
        for dur in (0.01, 0.025, 0.1, 0.4, 0.5, 1.0):
            step = np.float32(1.0) / np.float32(np.float32(dur) * np.float32(sr))
            total = loop(0.0, step, 10 ** 6)[1]
            for n in list(range(0, 40)) + [total - 2, total - 1, total, total + 5, total // 2, total // 3, 777]:
                check(0.0, step, max(0, n))


def test_random_starts_and_steps():
    rng = np.random.default_rng(5)
    for _ in range(400):
        t0 = np.float32(rng.random() * rng.choice([1e-6, 1e-3, 0.1, 0.999]))
        step = np.float32(rng.random() * rng.choice([1e-7, 1e-5, 1e-3, 0.3]))
        check(t0, step, int(rng.integers(0, 6000)))


def test_ties_stuck_and_odd_inputs():
    for t0, step, n in [
        (0.5, 2.0 ** -25, 100),               # exact tie on an even significand: never moves
        (0.5 + 2.0 ** -24, 2.0 ** -25, 100),  # tie on an odd significand: one add, then stuck
        (0.25, 3 * 2.0 ** -26, 5000),         # tie with odd q
        (0.75, 2.0 ** -30, 1000),             # below half an ulp: stuck
        (0.0, 2.0 ** -10, 2000),              # power-of-two step, finishes exactly at 1
        (0.0, 1.3, 3), (0.0, 1.0, 3), (0.999, 0.5, 3),
        (0.3, 0.0, 50), (0.3, -0.01, 50), (0.0, 1e-42, 50), (1e-40, 1e-3, 50),
        (0.0, float("inf"), 4), (-0.5, 0.01, 200), (1.5, 0.01, 5),
    ]:
        check(t0, step, n)
    t, done, fin = closed(0.3, float("nan"), 7)
    assert np.isnan(t) and done == 7 and not fin


# ---- the un-clamped accumulators (what lets a paint call start in the middle of a span) ------------------------------
def accum_loop(t, step, n):
    t, step = np.float32(t), np.float32(step)
    with np.errstate(all="ignore"):
        for _ in range(n):
            t = np.float32(t + step)
    return t


def cycle_loop(t, step, n):
    t, step = np.float32(t), np.float32(step)
    with np.errstate(all="ignore"):
        for _ in range(n):
            t = np.float32(t + step)
            t = np.float32(t - np.trunc(t))
    return t


def same_bits(a, b):
    return np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32) or (np.isnan(a) and np.isnan(b))


def test_accum_advance_matches_the_plain_loop():
    """SineOsc.zig:49-62 (t += freq/sr, up to ~100 cycles inside a span) and Sampler.zig:130 (t += ratio, t up to 3e5)."""
    rng = np.random.default_rng(9)
    cases = [(0.0, 440.0 / 48000.0, 1024), (0.73, 3951.0 / 48000.0, 1024), (0.0, 0.76923, 5000), (262000.0, 0.76923, 3000),
             (3.0e5, 1.3, 4000), (0.0, 2.0 ** -10, 5000), (0.5, 3 * 2.0 ** -25, 5000), (1.0, 1.0, 100), (0.0, 0.0, 10),
             (5.0, -0.3, 200), (-2.0, 0.25, 100), (0.0, 1e-42, 30), (1e30, 1.0, 10), (0.0, float("inf"), 3)]
    for _ in range(300):
        cases.append((float(rng.random() * rng.choice([1.0, 50.0, 3e5])), float(rng.random() * rng.choice([1e-3, 0.1, 1.5])), int(rng.integers(0, 5000))))
    for t0, step, n in cases:
        tt = C.c_float(t0)
        lib.zcu_accum_advance(C.byref(tt), float(step), n)
        assert same_bits(tt.value, accum_loop(t0, step, n)), (t0, step, n)


def test_cycle_advance_matches_the_plain_loop():
    """Cycle.zig:44-46: slow LFO speeds take the closed form between wraps, fast ones the plain loop."""
    rng = np.random.default_rng(10)
    cases = [(0.0, 3.0 / 48000.0, 48000), (0.999, 3.0 / 48000.0, 100), (0.0, 0.0625, 3000), (0.0, 2.0 ** -9, 4000),
             (0.5, 0.0, 10), (0.2, -0.001, 50), (1.7, 0.001, 50)]
    for _ in range(150):
        cases.append((float(rng.random()), float(rng.random() * rng.choice([1e-5, 1e-3, 3.9e-3, 0.05])), int(rng.integers(0, 20000))))
    for t0, step, n in cases:
        tt = C.c_float(t0)
        lib.zcu_cycle_advance(C.byref(tt), float(step), n)
        assert same_bits(tt.value, cycle_loop(t0, step, n)), (t0, step, n)
