"""Shared generators for the parity tests."""
import numpy as np

from zang_b200 import _lib as L


def random_spans(rng, voices, num_buffers, buffer_frames, nrows, max_cuts=3, p_empty=0.15, p_gap=0.3):
    """Per voice and per mix buffer: a few disjoint spans with random gaps, flags and params rows.
    Spans never cross a buffer boundary (the reference paints buffer by buffer)."""
    off = [0]
    rows = []
    for v in range(voices):
        n = 0
        for b in range(num_buffers):
            base = b * buffer_frames
            if rng.random() < p_empty:
                continue
            k = int(rng.integers(0, max_cuts + 1))
            cuts = sorted(set(int(x) for x in rng.integers(1, buffer_frames, size=k))) if buffer_frames > 1 else []
            edges = [0] + cuts + [buffer_frames]
            if rng.random() < 0.3 and len(edges) > 2:
                edges[1] = 1  # a one-frame span at the start
                edges = sorted(set(edges))
            for s, e in zip(edges[:-1], edges[1:]):
                if rng.random() < p_gap:
                    continue
                rows.append((base + s, base + e, int(rng.integers(0, 2)), int(rng.integers(0, nrows))))
                n += 1
        off.append(off[-1] + n)
    spans = np.array(rows, dtype=L.span_dtype) if rows else np.zeros(0, dtype=L.span_dtype)
    return np.array(off, dtype=np.uint32), spans


def full_spans(voices, num_buffers, buffer_frames, row=0, flag_first=True):
    off = np.arange(voices + 1, dtype=np.uint32) * num_buffers
    spans = np.zeros(voices * num_buffers, dtype=L.span_dtype)
    for v in range(voices):
        for b in range(num_buffers):
            spans[v * num_buffers + b] = (b * buffer_frames, (b + 1) * buffer_frames, 1 if (flag_first and b == 0) else 0, row)
    return off, spans


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_bit_exact(got, want, what=""):
    g, w = bits(got), bits(want)
    if not np.array_equal(g, w):
        bad = np.argwhere(g != w)
        i = tuple(bad[0])
        raise AssertionError(f"{what}: {bad.shape[0]} of {g.size} values differ; first at {i}: got {got[i]!r} want {want[i]!r} "
                             f"(max abs diff {np.nanmax(np.abs(got.astype(np.float64) - want.astype(np.float64)))})")


def assert_close(got, want, tol, what=""):
    d = np.abs(got.astype(np.float64) - want.astype(np.float64))
    m = float(d.max()) if d.size else 0.0
    assert m <= tol, f"{what}: max abs error {m} > {tol}"
    return m


def state_equal(a, b):
    if a is None and b is None:
        return True
    return a.tobytes() == b.tobytes()
