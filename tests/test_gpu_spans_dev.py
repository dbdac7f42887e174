"""Span tables built on the device (zcu_event_bank_*) against the host replay of the same bank
(zcu_bank_spans, itself held to the reference's tests and the oracle in tests/test_event_kat.py): the reference's five
Trigger tests (src/zang/trigger_test.zig:38-132) as one-buffer banks, then fuzzed streams over several blocks with carried
Trigger state (same-frame impulses, more than 32 impulses in a buffer, silent voices, a partial last buffer), and the
prev_note_on views of the configs[2] instrument.  Bit-exact: this is index work."""
import numpy as np
import pytest

from zang_b200 import _lib as L
from zang_b200 import banks, synth
from zang_b200.events import BankTrigger, DeviceBankTrigger

pytestmark = pytest.mark.gpu


def _events(rows):
    ev = np.zeros(len(rows), dtype=L.note_event_dtype)
    for i, (frame, note_id, freq) in enumerate(rows):
        ev[i] = (frame, note_id, freq, 1)
    return ev


def _run_device(ctx, buffers, total):
    """buffers: list of per-buffer impulse lists [(frame_in_buffer, note_id, freq)], one voice.  Returns per buffer the list
    the reference's `while (trigger.next(&ctr))` loop yields: (start, end, freq, note_id_changed)."""
    rows = [(b * 1024 + f, nid, fr) for b, lst in enumerate(buffers) for (f, nid, fr) in lst]
    ev = _events(rows)
    off = np.array([0, ev.size], dtype=np.uint32)
    dt = DeviceBankTrigger(ctx, off, ev, total, 1024)
    out = []
    for b in range(len(buffers)):
        sp = dt.spans(b * 1024, 1024)
        voff, spans = sp.download()
        sp.free()
        out.append([(int(s["start"]), int(s["end"]), float(ev[int(s["param"])]["freq"]), bool(s["flags"] & 1)) for s in spans[voff[0]:voff[1]]])
    dt.free()
    return out


def test_trigger_no_notes(ctx):  # trigger_test.zig:38-45
    assert _run_device(ctx, [[]], 1024) == [[]]


def test_trigger_first_note_at_frame_0(ctx):  # :47-60
    assert _run_device(ctx, [[(0, 1, 440.0)]], 1024) == [[(0, 1024, 440.0, True)]]


def test_trigger_first_note_after_frame_0(ctx):  # :62-75
    assert _run_device(ctx, [[(500, 1, 440.0)]], 1024) == [[(500, 1024, 440.0, True)]]


def test_trigger_carryover(ctx):  # :77-115
    got = _run_device(ctx, [[(0, 1, 440.0), (200, 2, 220.0)], [(500, 3, 330.0), (600, 3, 660.0)], []], 3 * 1024)
    assert got == [[(0, 200, 440.0, True), (200, 1024, 220.0, True)],
                   [(0, 500, 220.0, False), (500, 600, 330.0, True), (600, 1024, 660.0, False)],
                   [(0, 1024, 660.0, False)]]


def test_trigger_two_notes_same_frame(ctx):  # :117-132
    assert _run_device(ctx, [[(200, 1, 440.0), (200, 2, 220.0)]], 1024) == [[(200, 1024, 220.0, True)]]


def _fuzz_bank(rng, voices, total, dense_voice=None):
    off, rows = [0], []
    for v in range(voices):
        n = int(rng.integers(0, 40)) if v % 7 else 0  # some voices never play
        frames = np.sort(rng.integers(0, total, size=n))
        if n > 3:
            frames[2] = frames[1]  # two impulses on one frame
        if v == dense_voice:  # more than 32 impulses inside one mix buffer: the queue drops the rest (notes.zig:108-111)
            frames = np.sort(np.concatenate([frames, rng.integers(2048, 3072, size=50)]))
        nid = 1
        for f in frames:
            if rng.random() < 0.6:
                nid += 1  # a new note; otherwise the same id again (note off / retrigger of the same note)
            rows.append((int(f), nid, float(rng.uniform(50, 2000)), int(rng.random() < 0.6)))
        off.append(len(rows))
    ev = np.zeros(len(rows), dtype=L.note_event_dtype)
    for i, r in enumerate(rows):
        ev[i] = r
    return np.array(off, dtype=np.uint32), ev


@pytest.mark.parametrize("voices,total,block", [(70, 9 * 1024 + 300, 3 * 1024), (300, 16 * 1024, 16 * 1024), (33, 5 * 1024 + 1, 1024)])
def test_device_span_tables_equal_the_host_replay(ctx, voices, total, block):
    rng = np.random.default_rng(voices)
    off, ev = _fuzz_bank(rng, voices, total, dense_voice=5)
    host = BankTrigger(off, ev, total, 1024)
    dev = DeviceBankTrigger(ctx, off, ev, total, 1024)
    nspans = 0
    for first in range(0, total, block):
        n = min(block, total - first)
        hoff, hspans = host.spans(first, n)
        sp = dev.spans(first, n)
        doff, dspans = sp.download()
        sp.free()
        assert np.array_equal(hoff, doff), f"block at {first}: span offsets differ"
        assert hspans.tobytes() == dspans.tobytes(), f"block at {first}: spans differ"
        assert host.state.tobytes() == dev.state().tobytes(), f"block at {first}: carried Trigger state differs"
        nspans += hspans.size
    assert nspans > voices
    # reset: the same table again
    dev.reset()
    host2 = BankTrigger(off, ev, total, 1024)
    sp = dev.spans(0, min(block, total))
    assert host2.spans(0, min(block, total))[1].tobytes() == sp.download()[1].tobytes()
    sp.free()
    dev.free()


def test_prev_note_on_views_equal_the_host_tables(ctx):
    sr, voices, total, block = 48000.0, 50, 9 * 1024 + 50, 4 * 1024
    wl = synth.fm_bank(voices, total, sr, seed=21, mean_note=0.04)
    bank = banks.FmDetunedBank(ctx, wl, sr, total)
    dev = DeviceBankTrigger(ctx, wl["event_off"], wl["events"], total, 1024)
    for first in range(0, total, block):
        n = min(block, total - first)
        off, _, a_porta, a_env, a_plain = bank.span_tables(first, n)
        src = dev.spans(first, n)
        views = dev.prev_note_on_views(src)
        for name, want, got in zip(("porta", "env", "plain"), (a_porta, a_env, a_plain), views):
            goff, gspans = got.download()
            assert np.array_equal(goff, off) and gspans.tobytes() == want.tobytes(), f"{name} view differs in the block at {first}"
            got.free()
        src.free()
    dev.free()
