"""NoteTracker + PolyphonyDispatcher + per-slot Trigger replayed on the device (zcu_poly_bank_*) against the
host replay of the same song (zcu_poly_voice_spans, itself held to the reference's tests and the oracle in
tests/test_event_kat.py / test_song_cpu.py): the reference's three dispatcher tests (src/zang/notes_test.zig:10-119) as
one-buffer songs, then fuzzed songs (note-offs for sounding and for already stolen notes, more notes than slots, several notes
on one frame, more than 32 notes in a buffer) over several blocks with the tracker time, the slots and the Triggers carried.
Bit-exact: index work plus the tracker's f32 time arithmetic."""
import ctypes as C

import numpy as np
import pytest

from zang_b200 import _lib as L
from zang_b200._lib import lib, ptr
from zang_b200.events import DevicePolyBank

pytestmark = pytest.mark.gpu


def _host_spans(song, polyphony, note_on, sr, bf, nbuf, last, pod=None):
    se = np.zeros(len(song), dtype=L.song_event_dtype)
    for i, e in enumerate(song):
        se[i] = e
    if pod is None:
        pod = L.PolyVoicePOD()
        assert lib.zcu_poly_voice_init(C.byref(pod), ptr(se), se.size, polyphony) == L.OK
    pod.tracker.song = se.ctypes.data
    off = np.zeros(polyphony + 1, dtype=np.uint32)
    cap = len(song) + polyphony * nbuf + 8
    spans = np.zeros(cap, dtype=L.span_dtype)
    n = lib.zcu_poly_voice_spans(C.byref(pod), ptr(note_on), sr, bf, nbuf, last, ptr(off), ptr(spans), cap)
    assert n >= 0
    return off, spans[:n].copy(), pod, se


def _notes_of_slot(off, spans, song_rows, slot):
    return [song_rows[int(s["param"])] for s in spans[off[slot]:off[slot + 1]]]


@pytest.mark.parametrize("impulses,ons,want", [
    ([(100, 1), (200, 2), (300, 3), (400, 4), (500, 5)], [1, 1, 1, 1, 1], [[1, 4], [2, 5], [3]]),                       # notes_test.zig:10-41
    ([(100, 1), (200, 1), (300, 2), (400, 2), (500, 3)], [1, 0, 1, 0, 1], [[1, 1], [2, 2], [3]]),                       # :43-74
    ([(100, 1), (200, 2), (300, 3), (400, 3), (500, 2), (600, 1), (700, 4)], [1, 1, 1, 0, 0, 0, 1], [[1, 1], [2, 2], [3, 3, 4]]),  # :76-119
])
def test_reference_dispatcher_cases_on_the_device(ctx, impulses, ons, want):
    sr, bf = 1024.0, 1024  # one buffer = one second: event time = frame / 1024 lands on that frame exactly
    song = [(f / 1024.0, nid, i) for i, (f, nid) in enumerate(impulses)]
    bank = DevicePolyBank(ctx, [song], 3, ons, sr, bf)
    sp = bank.spans(1)
    off, spans = sp.download()
    sp.free()
    ids = [nid for _, nid in impulses]
    for slot in range(3):
        got = [ids[int(s["param"])] for s in spans[off[slot]:off[slot + 1]]]
        assert got == want[slot], (slot, got)
        first = int(spans[off[slot]]["start"])
        assert first == [f for f, nid in impulses if nid == want[slot][0]][0], "the slot's first span starts at its first note's frame"
    bank.free()


def _fuzz_song(rng, seconds, dense=False):
    t, nid, rows, song = 0.0, 0, [], []
    sounding = []
    while True:
        t += float(rng.exponential(0.004 if dense else 0.05))
        if t >= seconds:
            break
        if sounding and rng.random() < 0.45:
            k = sounding.pop(int(rng.integers(0, len(sounding))))
            song.append((np.float32(t), k, len(rows)))
            rows.append(0)
        else:
            nid += 1
            sounding.append(nid)
            song.append((np.float32(t), nid, len(rows)))
            rows.append(1)
        if rng.random() < 0.15 and song:  # another note on the very same instant
            nid += 1
            sounding.append(nid)
            song.append((song[-1][0], nid, len(rows)))
            rows.append(1)
    return song, rows


@pytest.mark.parametrize("polyphony,songs,dense", [(3, 40, False), (10, 70, False), (4, 9, True), (1, 33, False), (32, 5, True)])
def test_device_poly_bank_equals_the_host_replay(ctx, polyphony, songs, dense):
    rng = np.random.default_rng(polyphony * 100 + songs)
    sr, bf = 48000.0, 1024
    blocks = [(7, 1024), (5, 300), (9, 1024)]  # (buffers, frames of the last buffer)
    seconds = sum((n - 1) * bf + last for n, last in blocks) / sr
    all_songs, all_rows, base = [], [], 0
    for s in range(songs):
        song, rows = _fuzz_song(rng, seconds * 1.1, dense)
        all_songs.append([(t, nid, row + base) for t, nid, row in song])
        all_rows += rows
        base += len(rows)
    note_on = np.array(all_rows if all_rows else [1], dtype=np.int32)
    bank = DevicePolyBank(ctx, all_songs, polyphony, note_on, sr, bf)
    pods = [None] * songs
    keep = [None] * songs
    total = 0
    for nbuf, last in blocks:
        sp = bank.spans(nbuf, last)
        doff, dspans = sp.download()
        sp.free()
        for s in range(songs):
            hoff, hspans, pods[s], keep[s] = _host_spans(all_songs[s], polyphony, note_on, sr, bf, nbuf, last, pods[s])
            for k in range(polyphony):
                col = s * polyphony + k
                got = dspans[doff[col]:doff[col + 1]]
                want = hspans[hoff[k]:hoff[k + 1]]
                assert got.tobytes() == want.tobytes(), f"song {s} slot {k}: spans differ in a block of {nbuf} buffers"
            total += hspans.size
        st = bank.state()
        for s in range(songs):
            assert st[s].tracker.next_song_event == pods[s].tracker.next_song_event and st[s].tracker.t == pods[s].tracker.t, f"tracker state of song {s}"
            assert bytes(st[s].dispatcher) == bytes(pods[s].dispatcher), f"dispatcher slots of song {s}"
            assert bytes(st[s].triggers)[:12 * polyphony] == bytes(pods[s].triggers)[:12 * polyphony], f"trigger state of song {s}"
    assert total > songs * polyphony
    bank.free()
