"""Known-answer tests for the arithmetic that lives OUTSIDE /root/reference: Zig std's DefaultPrng
(xoshiro256++ seeded through SplitMix64) and Random.float(f32), used by src/modules/Noise.zig:22-29,51,58.

The reference pins none of this with a test of its own.  Anchors used here:
  * xoshiro256++ seeded with SplitMix64(0): the Zig std test vector quoted in SURVEY.md §8(c)
  * SplitMix64(seed = 1234567) first outputs: the reference vector of the public-domain splitmix64.c
  * the float conversion: structural properties (range, mantissa = low 23 bits, exponent from the leading zeros)
"""
import numpy as np

import oracle
from zang_b200 import _lib as L

ptr = L.ptr

XOSHIRO_SEED0 = [0x53175d61490b23df, 0x61da6f3dc380d507, 0x5c0fdf91ec9a7bfc, 0x02eebf8c3bbe5e1a, 0x7eca04ebaf4a5eea, 0x0543c37757f08d9a]


def splitmix(seed, n):
    out, s, M = [], seed, (1 << 64) - 1
    for _ in range(n):
        s = (s + 0x9e3779b97f4a7c15) & M
        z = s
        z = ((z ^ (z >> 30)) * 0xbf58476d1ce4e5b9) & M
        z = ((z ^ (z >> 27)) * 0x94d049bb133111eb) & M
        out.append(z ^ (z >> 31))
    return out


def test_splitmix64_reference_vector():
    # splitmix64.c (Vigna), seed 1234567
    assert splitmix(1234567, 5) == [6457827717110365317, 3203168211198807973, 9817491932198370423, 4593380528125082431, 16408922859458223821]


def test_xoshiro_seed_is_splitmix():
    s = np.zeros(4, dtype=np.uint64)
    for seed in (0, 1, 5, 4095, 2 ** 32 - 1):
        oracle.orc.orc_xoshiro_seed(seed, ptr(s))
        assert [int(x) for x in s] == splitmix(seed, 4)


def test_xoshiro256pp_zig_std_vector():
    s = np.zeros(4, dtype=np.uint64)
    oracle.orc.orc_xoshiro_seed(0, ptr(s))
    out = np.zeros(6, dtype=np.uint64)
    oracle.orc.orc_xoshiro_next(ptr(s), ptr(out), 6)
    assert [int(x) for x in out] == XOSHIRO_SEED0


def py_float32(r):
    lz = 64 - r.bit_length()
    assert lz < 41  # the >= 41 branch has probability 2^-41 per draw
    bits = ((126 - lz) << 23) | (r & 0x7FFFFF)
    return np.array([bits], dtype=np.uint32).view(np.float32)[0]


def test_float_conversion_structure():
    s = np.zeros(4, dtype=np.uint64)
    oracle.orc.orc_xoshiro_seed(7, ptr(s))
    s2 = s.copy()
    raw = np.zeros(4096, dtype=np.uint64)
    oracle.orc.orc_xoshiro_next(ptr(s), ptr(raw), raw.size)
    fl = np.zeros(4096, dtype=np.float32)
    oracle.orc.orc_xoshiro_floats(ptr(s2), ptr(fl), fl.size)
    assert np.array_equal(s, s2)  # one u64 per float
    assert fl.min() >= 0.0 and fl.max() < 1.0
    want = np.array([py_float32(int(r)) for r in raw], dtype=np.float32)
    assert np.array_equal(fl.view(np.uint32), want.view(np.uint32))
    assert 0.45 < fl.mean() < 0.55


def test_noise_seed_counter():
    # Noise.zig:26: the k-th Noise.init() in program order gets seed k; bank voice v gets first_seed + v
    st = oracle.state_init(L.MOD_NOISE, 5, first_seed=10)
    for v in range(5):
        assert [int(x) for x in st["s"][v]] == splitmix(10 + v, 4)


def test_jump_matrices_equal_plain_stepping():
    """The Noise skip-ahead multiplies the xoshiro256 state by T^n (hex-digit matrices built in libzangcu, zcu_paint.cu); the
    host copy of that product must equal n plain next() calls of the oracle's generator, for n across all digit positions."""
    import ctypes as C
    from zang_b200._lib import lib
    rng = np.random.default_rng(2024)
    for n in [1, 2, 15, 16, 17, 255, 256, 4095, 4096, 65535, 65536, 99991, (1 << 20) + 12345] + [int(x) for x in rng.integers(1, 1 << 21, size=6)]:
        s = np.zeros(4, dtype=np.uint64)
        oracle.orc.orc_xoshiro_seed(int(rng.integers(0, 1 << 32)), L.ptr(s))
        jumped = s.copy()
        assert lib.zcu_xoshiro_jump_host(L.ptr(jumped), n) == 0
        out = np.zeros(n, dtype=np.uint64)
        oracle.orc.orc_xoshiro_next(L.ptr(s), L.ptr(out), n)  # advances s by n draws
        assert np.array_equal(jumped, s), n
    bad = np.zeros(4, dtype=np.uint64)
    assert lib.zcu_xoshiro_jump_host(L.ptr(bad), 1 << 24) != 0  # beyond the table


M64 = (1 << 64) - 1


def _rotl(x, k):
    return ((x << k) | (x >> (64 - k))) & M64


def py_next(s):
    s0, s1, s2, s3 = s
    r = (_rotl((s0 + s3) & M64, 23) + s0) & M64
    t = (s1 << 17) & M64
    s2 ^= s0
    s3 ^= s1
    s1 ^= s2
    s0 ^= s3
    s2 ^= t
    s3 = _rotl(s3, 45)
    return r, [s0, s1, s2, s3]


def py_prev(s):
    """the state before the last step: the form XoshiroDev::prev uses on the device (zcu_modules.cuh) when a batch of four
    draws has to be taken again one sample at a time"""
    s0, s1, s2, s3 = s
    s3a = _rotl(s3, 19)
    c = s1 ^ s2
    s1o = (c ^ (c << 17) ^ (c << 34) ^ (c << 51)) & M64
    s2a = s2 ^ ((s1o << 17) & M64)
    s0o = s0 ^ s3a
    return [s0o, s1o, s2a ^ s0o, s3a ^ s1o]


def test_xoshiro_step_and_its_inverse():
    rng = np.random.default_rng(5)
    for _ in range(50):
        s = [int(x) for x in rng.integers(0, 1 << 63, size=4, dtype=np.uint64)]
        arr = np.array(s, dtype=np.uint64)
        out = np.zeros(3, dtype=np.uint64)
        oracle.orc.orc_xoshiro_next(ptr(arr), ptr(out), 3)
        st, draws = s, []
        for _ in range(3):
            r, st = py_next(st)
            draws.append(r)
        assert draws == [int(x) for x in out] and st == [int(x) for x in arr]
        for _ in range(3):
            st = py_prev(st)
        assert st == s


def py_float_full(state):
    """Random.float(f32) with its refill branches; returns (bits, state after, draws taken)"""
    clz = lambda x, w: w - x.bit_length()
    r, state = py_next(state)
    lz, taken = clz(r, 64), 1
    if lz >= 41:
        r2, state = py_next(state)
        lz, taken = 41 + clz(r2, 64), 2
        if lz == 41 + 64:
            r3, state = py_next(state)
            lz += clz((r3 & 0xFFFFFFFF) | 0x7FF, 32)
            taken = 3
    return ((126 - lz) << 23) | (r & 0x7FFFFF), state, taken


def test_float_refill_branches():
    """Draws the float conversion refills on: an empty upper word without a refill (lz 32..40), a draw below 2^23 (one more
    draw decides the exponent) and two zero draws in a row (a third draw).  States are built so that the first draw is such a
    value; the oracle's conversion must take as many draws as the restatement and give its bits."""
    rng = np.random.default_rng(6)
    seen = set()
    for variant in range(3):
        for _ in range(5):
            s1, s2 = (int(x) for x in rng.integers(1, 1 << 63, size=2, dtype=np.uint64))
            s = [[0, s1, s2, 3 + int(rng.integers(0, 500))], [0, s1, s2, 0], [0, 0, s2, 0]][variant]
            bits, after, taken = py_float_full(s)
            seen.add(taken)
            arr = np.array(s, dtype=np.uint64)
            fl = np.zeros(1, dtype=np.float32)
            oracle.orc.orc_xoshiro_floats(ptr(arr), ptr(fl), 1)
            assert int(fl.view(np.uint32)[0]) == bits, (variant, hex(bits))
            assert [int(x) for x in arr] == after, (variant, taken)
            assert 0.0 <= float(fl[0]) < 1.0
    assert seen == {1, 2, 3}
