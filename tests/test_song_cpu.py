"""Host side of the reference's own CPU workload (BASELINE.json configs[0]: write_wav of the Toccata song,
examples/write_wav.zig + examples/example_song.zig): the product's score parser and Voice(T) event logic against
the oracle's restatement, and the oracle against its golden hashes.  No GPU needed."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from zang_b200 import _lib as L
from zang_b200 import song
from zang_b200._lib import lib, ptr

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SONG = "/root/reference/examples/example_song.txt"

# a score in the reference's tracker syntax (songparse1.zig) written for this test: comments, rate/tempo words,
# 12 columns in groups of 2 | 8 | 2, note-offs, same-column retriggers, idle rows, a `start` word that discards
# what came before it
MINI_SONG = """# warm-up that `start` throws away
|C-4
|off
start
rate 2.0
|D-2 D-3|A-4     E-5             F#3 G-4|C#5
|       |    B-4
tempo 0.5
|off    |A-4 off                        |    A#5
|
tempo 1.0   |D-3    |off     off             off off|off off
rate 1.5
|E-2    |C-4 E-4 G-4 C-5 E-5 G-5 C-6 E-6|G#4 B-4
|off    |off off off off off off off off|off off
"""


def load_fixture():
    z = np.load(os.path.join(HERE, "golden", "toccata_events.npz"))
    return [z["pedal"], z["regular_organ"], z["weird_organ"]]


def same_events(a, b):
    return all(x.tobytes() == y.tobytes() for x, y in zip(a, b)) and len(a) == len(b)


def test_parser_matches_oracle_on_mini_song():
    got, want = song.parse_song(MINI_SONG), oracle.song_parse(MINI_SONG)
    assert [n.size for n in want] == [7, 28, 8]
    assert same_events(got, want)
    # `start` dropped the warm-up rows but ids keep counting (example_song.zig:147-154)
    assert want[1]["note_id"].min() > 1 and want[0]["t"][0] == 0.0


def test_parser_rejects_bad_rows():
    with pytest.raises(song.SongSyntaxError):
        song.parse_song("|C-4 D-4 E-4 F-4 G-4 A-4 B-4 C-5 D-5 E-5 F-5 G-5 A-5\n")  # 13 columns
    with pytest.raises(song.SongSyntaxError):
        song.parse_song("rate fast\n")
    with pytest.raises(song.SongSyntaxError):
        song.parse_song("|C-4 x\n")
    assert oracle.orc.orc_song_events(b"|C-4 x\n", 7, 0, None, None, None, None, 0) == -1


@pytest.mark.skipif(not os.path.exists(REF_SONG), reason="the reference tree is not mounted on this machine")
def test_parser_matches_oracle_on_the_reference_score_and_fixture_is_current():
    text = open(REF_SONG, "rb").read()
    want = oracle.song_parse(text)
    assert same_events(song.parse_song(text), want)
    assert same_events(load_fixture(), want), "tests/golden/toccata_events.npz is stale: run tests/golden/make_song_golden.py"


def test_oracle_song_golden_prefix():
    gold = json.load(open(os.path.join(HERE, "golden", "song_golden.json")))
    notes = load_fixture()
    assert [int(n.size) for n in notes] == gold["events"]
    n = 600 * 1024
    pcm, mix, _ = oracle.song_render(notes, n)
    assert hashlib.sha256(pcm.tobytes()).hexdigest() == gold["pcm_sha256"][str(n)]
    assert hashlib.sha256(pcm[:64 * 1024 * 2].tobytes()).hexdigest() == gold["pcm_sha256"][str(64 * 1024)]
    assert np.abs(mix).max() > 0.5


def oracle_voice_spans(notes, polyphony, num_buffers, buffer_frames, last_frames, sr=48000.0):
    """Voice.paint's event logic through the oracle's NoteTracker / PolyphonyDispatcher / Trigger restatements."""
    se = np.zeros(notes.size, dtype=L.song_event_dtype)
    se["t"], se["note_id"], se["param"] = notes["t"], notes["note_id"], np.arange(notes.size)
    tracker = (C.c_uint32 * 2)()  # OrcNoteTracker {next_song_event, t}
    slots = np.zeros(polyphony, dtype=L.poly_slot_dtype)
    trig = [(C.c_uint32 * 3)() for _ in range(polyphony)]  # {has_note, note_id, param}
    per = [[] for _ in range(polyphony)]
    for b in range(num_buffers):
        n_frames = last_frames if b + 1 == num_buffers else buffer_frames
        imps = np.zeros(64, dtype=L.impulse_dtype)
        prm = np.zeros(64, dtype=np.uint32)
        n = oracle.orc.orc_note_tracker_consume(tracker, ptr(se), se.size, sr, 0, n_frames, ptr(imps), ptr(prm))
        on = np.ascontiguousarray(notes["note_on"][prm[:n]], dtype=np.int32)
        oi = np.zeros(polyphony * 32, dtype=L.impulse_dtype)
        op = np.zeros(polyphony * 32, dtype=np.uint32)
        oc = np.zeros(polyphony, dtype=np.uint32)
        assert oracle.orc.orc_poly_dispatch(polyphony, ptr(slots), ptr(imps), ptr(prm), ptr(on), n, ptr(oi), ptr(op), ptr(oc)) == 0
        for i in range(polyphony):
            out = np.zeros(40, dtype=L.span_dtype)
            k = oracle.orc.orc_trigger_run(trig[i], 0, n_frames, ptr(oi[i * 32:]), ptr(op[i * 32:]), int(oc[i]), b * buffer_frames, ptr(out), 40)
            per[i].extend(out[:k].tolist())
    off = np.cumsum([0] + [len(p) for p in per]).astype(np.uint32)
    spans = np.array([s for p in per for s in p], dtype=L.span_dtype)
    return off, spans


@pytest.mark.parametrize("which,polyphony", [(0, 3), (1, 10), (2, 4)])
def test_poly_voice_spans_match_oracle(which, polyphony):
    notes = load_fixture()[which]
    if which == 2:  # the weird organ only plays around 178-222 s: shift it to the start
        notes = notes.copy()
        notes["t"] -= np.float32(177.0)
    nb, B, last = 700, 1024, 777
    want_off, want = oracle_voice_spans(notes, polyphony, nb, B, last)
    assert want.size > 100

    class FakeBank:
        Params = np.dtype([("sample_rate", "<f4"), ("freq", "<f4"), ("note_on", "<i4")])
        num_temps = 0
    pv = song.PolyVoice(None, FakeBank, notes, polyphony)
    # two consecutive blocks must equal one long block (tracker / dispatcher / trigger state carried)
    off1, sp1 = pv.span_table(300, B, B)
    off2, sp2 = pv.span_table(nb - 300, B, last)
    sp2 = sp2.copy()
    sp2["start"] += 300 * B
    sp2["end"] += 300 * B
    got = np.concatenate([np.concatenate([sp1[off1[i]:off1[i + 1]], sp2[off2[i]:off2[i + 1]]]) for i in range(polyphony)])
    assert got.tobytes() == want.tobytes()
    assert np.array_equal(off1 + off2, want_off)


def test_poly_voice_spans_capacity_is_transactional():
    notes = load_fixture()[1]
    se = np.zeros(notes.size, dtype=L.song_event_dtype)
    se["t"], se["note_id"], se["param"] = notes["t"], notes["note_id"], np.arange(notes.size)
    on = np.ascontiguousarray(notes["note_on"], dtype=np.int32)
    pod = L.PolyVoicePOD()
    assert lib.zcu_poly_voice_init(C.byref(pod), ptr(se), se.size, 10) == L.OK
    before = bytes(pod)
    off = np.zeros(11, dtype=np.uint32)
    small = np.zeros(4, dtype=L.span_dtype)
    n = lib.zcu_poly_voice_spans(C.byref(pod), ptr(on), 48000.0, 1024, 200, 1024, ptr(off), ptr(small), 4)
    assert n < 0 and bytes(pod) == before  # nothing consumed
    big = np.zeros(-n, dtype=L.span_dtype)
    assert lib.zcu_poly_voice_spans(C.byref(pod), ptr(on), 48000.0, 1024, 200, 1024, ptr(off), ptr(big), -n) == -n
    assert off[10] == -n and bytes(pod) != before
    assert lib.zcu_poly_voice_init(C.byref(pod), ptr(se), se.size, 33) == L.ERR_INVALID


def test_wav_header():
    h = song.wav_header(1, 48000, 2, 1000)
    assert len(h) == 44 and h[:4] == b"RIFF" and h[8:16] == b"WAVEfmt " and h[36:40] == b"data"
    assert int.from_bytes(h[4:8], "little") == 1036 and int.from_bytes(h[40:44], "little") == 1000
    assert int.from_bytes(h[24:28], "little") == 48000 and int.from_bytes(h[34:36], "little") == 16
