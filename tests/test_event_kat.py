"""Known-answer tests of the event layer: the reference's own 8 tests, transcribed.

  src/zang/trigger_test.zig:38-132  (5 tests over Trigger(f32), span (0,1024))
  src/zang/notes_test.zig:10-119    (3 tests over PolyphonyDispatcher(3))

Each runs against BOTH the oracle's line-by-line restatement (pinning the oracle) and the product's
host event layer in libzangcu.so (zcu_trigger_run / zcu_poly_dispatcher_dispatch), then the two are
fuzzed against each other bit for bit.
"""
import ctypes as C

import numpy as np
import pytest

import oracle
from zang_b200 import _lib as L
from zang_b200.events import ImpulseQueue, NoteTracker, PolyphonyDispatcher, Trigger
from zang_b200._lib import lib, ptr

SPAN = (0, 1024)


class OracleTrigger:
    """same face as zang_b200.events.Trigger, on oracle/zang_oracle.hpp's Trigger<u32>"""

    def __init__(self):
        self._t = L.TriggerPOD(0, 0, 0)
        self._cur = None

    def run(self, span, impulses, params, frame_base=0):
        imp = np.ascontiguousarray(impulses, dtype=L.impulse_dtype).reshape(-1)
        n = imp.size
        table = [self._cur] + list(params)
        self._t.param = 0
        rows = np.arange(1, n + 1, dtype=np.uint32)
        out = np.zeros(n + 2, dtype=L.span_dtype)
        cnt = oracle.orc.orc_trigger_run(C.byref(self._t), span[0], span[1], ptr(imp), ptr(rows), n, frame_base, ptr(out), out.size)
        res = [(int(s["start"]), int(s["end"]), table[int(s["param"])], bool(s["flags"] & 1)) for s in out[:cnt]]
        if self._t.has_note:
            self._cur = table[self._t.param]
        return res


def imps(*rows):
    a = np.zeros(len(rows), dtype=L.impulse_dtype)
    for i, r in enumerate(rows):
        a[i] = r
    return a


@pytest.fixture(params=["oracle", "libzangcu"])
def trigger(request):
    return OracleTrigger() if request.param == "oracle" else Trigger()


def test_trigger_no_notes(trigger):  # trigger_test.zig:38-45
    assert trigger.run(SPAN, imps(), []) == []


def test_trigger_first_note_at_frame_0(trigger):  # :47-60
    assert trigger.run(SPAN, imps((0, 1, 1)), [440.0]) == [(0, 1024, 440.0, True)]


def test_trigger_first_note_after_frame_0(trigger):  # :62-75
    assert trigger.run(SPAN, imps((500, 1, 1)), [440.0]) == [(500, 1024, 440.0, True)]


def test_trigger_carryover(trigger):  # :77-115
    assert trigger.run(SPAN, imps((0, 1, 1), (200, 2, 2)), [440.0, 220.0]) == [(0, 200, 440.0, True), (200, 1024, 220.0, True)]
    assert trigger.run(SPAN, imps((500, 3, 1), (600, 3, 2)), [330.0, 660.0]) == [
        (0, 500, 220.0, False), (500, 600, 330.0, True), (600, 1024, 660.0, False)]
    assert trigger.run(SPAN, imps(), []) == [(0, 1024, 660.0, False)]


def test_trigger_two_notes_same_frame(trigger):  # :117-132
    assert trigger.run(SPAN, imps((200, 1, 1), (200, 2, 2)), [440.0, 220.0]) == [(200, 1024, 220.0, True)]


class OracleDispatcher:
    def __init__(self, polyphony):
        self.polyphony = polyphony
        self.slots = np.zeros(polyphony, dtype=L.poly_slot_dtype)

    def dispatch(self, impulses, params):
        imp = np.ascontiguousarray(impulses, dtype=L.impulse_dtype).reshape(-1)
        n = imp.size
        rows = np.arange(n, dtype=np.uint32)
        on = np.array([1 if p["note_on"] else 0 for p in params], dtype=np.int32)
        oi = np.zeros((self.polyphony, 32), dtype=L.impulse_dtype)
        op = np.zeros((self.polyphony, 32), dtype=np.uint32)
        oc = np.zeros(self.polyphony, dtype=np.uint32)
        rc = oracle.orc.orc_poly_dispatch(self.polyphony, ptr(self.slots), ptr(imp), ptr(rows), ptr(on), n, ptr(oi), ptr(op), ptr(oc))
        assert rc == 0
        return [(oi[s, :oc[s]].copy(), [params[r] for r in op[s, :oc[s]]]) for s in range(self.polyphony)]


@pytest.fixture(params=["oracle", "libzangcu"])
def make_pd(request):
    return OracleDispatcher if request.param == "oracle" else PolyphonyDispatcher


def on(b):
    return {"note_on": b}


def test_dispatcher_5_note_ons_3_slots(make_pd):  # notes_test.zig:10-41
    pd = make_pd(3)
    r = pd.dispatch(imps((100, 1, 1), (200, 2, 2), (300, 3, 3), (400, 4, 4), (500, 5, 5)), [on(True)] * 5)
    assert [int(x) for x in r[0][0]["note_id"]] == [1, 4]
    assert [int(x) for x in r[1][0]["note_id"]] == [2, 5]
    assert [int(x) for x in r[2][0]["note_id"]] == [3]


def test_dispatcher_single_note_on_and_off(make_pd):  # :43-74
    pd = make_pd(3)
    r = pd.dispatch(imps((100, 1, 1), (200, 1, 2), (300, 2, 3), (400, 2, 4), (500, 3, 5)),
                    [on(True), on(False), on(True), on(False), on(True)])
    assert [int(x) for x in r[0][0]["note_id"]] == [1, 1]
    assert [int(x) for x in r[1][0]["note_id"]] == [2, 2]
    assert [int(x) for x in r[2][0]["note_id"]] == [3]


def test_dispatcher_reuse_least_recently_released(make_pd):  # :76-119
    pd = make_pd(3)
    r = pd.dispatch(imps((100, 1, 1), (200, 2, 2), (300, 3, 3), (400, 3, 4), (500, 2, 5), (600, 1, 6), (700, 4, 7)),
                    [on(True), on(True), on(True), on(False), on(False), on(False), on(True)])
    assert [int(x) for x in r[0][0]["note_id"]] == [1, 1]
    assert [int(x) for x in r[1][0]["note_id"]] == [2, 2]
    assert [int(x) for x in r[2][0]["note_id"]] == [3, 3, 4]


# ---- fuzz: product event layer == oracle restatement, bit for bit ------------------------------------
def random_impulses(rng, n, lo=0, hi=1024, dup_prob=0.3, ids=6):
    frames = np.sort(rng.integers(lo, hi, size=n))
    for i in range(1, n):
        if rng.random() < dup_prob:
            frames[i] = frames[i - 1]
    a = np.zeros(n, dtype=L.impulse_dtype)
    a["frame"] = frames
    a["note_id"] = rng.integers(1, ids + 1, size=n)
    a["event_id"] = np.arange(1, n + 1)
    return a


def test_trigger_fuzz_against_oracle():
    rng = np.random.default_rng(1)
    for trial in range(300):
        a, b = OracleTrigger(), Trigger()
        for buf in range(6):
            n = int(rng.integers(0, 9))
            length = int(rng.choice([1024, 1024, 1000, 7, 1]))
            imp = random_impulses(rng, n, 0, length)
            params = [float(x) for x in rng.integers(100, 1000, size=n)]
            assert a.run((0, length), imp, params, frame_base=buf * 1024) == b.run((0, length), imp, params, frame_base=buf * 1024), (trial, buf)


def test_trigger_nonzero_span_start():
    # Trigger.counter takes any span (trigger.zig:66-78); impulses are absolute inside it
    rng = np.random.default_rng(2)
    for trial in range(100):
        a, b = OracleTrigger(), Trigger()
        for buf in range(3):
            n = int(rng.integers(0, 6))
            imp = random_impulses(rng, n, 100, 900)
            params = list(range(n))
            assert a.run((100, 900), imp, params) == b.run((100, 900), imp, params)


def test_dispatcher_fuzz_against_oracle():
    rng = np.random.default_rng(3)
    for trial in range(200):
        poly = int(rng.integers(1, 11))
        a, b = OracleDispatcher(poly), PolyphonyDispatcher(poly)
        eid = 1
        for buf in range(5):
            n = int(rng.integers(0, 12))
            imp = random_impulses(rng, n, ids=8, dup_prob=0.1)
            imp["event_id"] = np.arange(eid, eid + n)
            eid += n
            params = [on(bool(x)) for x in rng.integers(0, 2, size=n)]
            ra, rb = a.dispatch(imp, params), b.dispatch(imp, params)
            for s in range(poly):
                assert np.array_equal(ra[s][0], rb[s][0]) and ra[s][1] == rb[s][1], (trial, buf, s)


def test_note_tracker_against_oracle():
    rng = np.random.default_rng(4)
    for trial in range(50):
        n = int(rng.integers(1, 60))
        ts = np.sort(rng.uniform(0.0, 0.5, size=n)).astype(np.float32)
        song = [(float(ts[i]), int(rng.integers(1, 9)), i) for i in range(n)]
        nt = NoteTracker(song)
        osong = np.zeros(n, dtype=L.song_event_dtype)
        for i, (t, nid, p) in enumerate(song):
            osong[i] = (t, nid, p)
        ost = np.zeros(1, dtype=np.dtype([("next", "<u4"), ("t", "<f4")]))
        for buf in range(30):
            length = 1024 if buf % 5 else 333
            imp, params = nt.consume(48000.0, (0, length))
            oi = np.zeros(64, dtype=L.impulse_dtype)
            op = np.zeros(64, dtype=np.uint32)
            cnt = oracle.orc.orc_note_tracker_consume(ptr(ost), ptr(osong), n, 48000.0, 0, length, ptr(oi), ptr(op))
            assert cnt == imp.size and np.array_equal(oi[:cnt], imp) and list(op[:cnt]) == params


def test_note_tracker_frame_formula():
    # notes.zig:169-183: frame = start + min(int((t - self.t)/buf_time * len), len-1), event ids 1-based
    nt = NoteTracker([(0.0, 1, "a"), (0.01, 2, "b"), (0.0213, 3, "c"), (0.03, 4, "d")])
    imp, params = nt.consume(48000.0, (0, 1024))
    bt = np.float32(1024) / np.float32(48000)
    exp = [min(int((np.float32(t) - np.float32(0)) / bt * np.float32(1024)), 1023) for t in (0.0, 0.01, 0.0213)]
    assert [int(x) for x in imp["frame"]] == exp and params == ["a", "b", "c"]
    assert [int(x) for x in imp["event_id"]] == [1, 2, 3]
    imp, params = nt.consume(48000.0, (0, 1024))
    assert params == ["d"]


def test_impulse_queue_semantics():  # notes.zig:102-127
    q = ImpulseQueue()
    q.push(10, 1, "a")
    q.push(5, 2, "dropped: out of order")
    q.push(10, 3, "b")
    imp, params = q.consume()
    assert [tuple(int(x) for x in r) for r in imp] == [(10, 1, 1), (10, 3, 2)] and params == ["a", "b"]
    for i in range(40):
        q.push(i, i + 1, i)
    imp, params = q.consume()
    assert imp.size == 32 and params == list(range(32))  # capacity 32, the rest dropped
    assert [int(x) for x in imp["event_id"]] == list(range(3, 35))  # next_event_id keeps counting (:126)
    assert q.consume()[0].size == 0
    # product POD vs oracle restatement on the same pushes
    a, b = L.ImpulseQueuePOD(), L.ImpulseQueuePOD()
    lib.zcu_impulse_queue_init(C.byref(a))
    lib.zcu_impulse_queue_init(C.byref(b))
    rng = np.random.default_rng(5)
    for i in range(50):
        fr, nid = int(rng.integers(0, 1024)), int(rng.integers(1, 5))
        lib.zcu_impulse_queue_push(C.byref(a), fr, nid, i)
        oracle.orc.orc_impulse_queue_push(C.byref(b), fr, nid, i)
        assert a.length == b.length and a.next_event_id == b.next_event_id
        for k in range(a.length):
            assert (a.impulses[k].frame, a.impulses[k].note_id, a.impulses[k].event_id, a.params[k]) == (
                b.impulses[k].frame, b.impulses[k].note_id, b.impulses[k].event_id, b.params[k])


def test_bank_spans_match_the_per_buffer_trigger_replay():
    """zcu_bank_spans (one C call per block for a whole bank) against the per-voice, per-buffer zcu_trigger_run replay,
    block by block with carried state, including same-frame events and more than 32 events in one buffer."""
    from zang_b200 import synth
    from zang_b200.events import BankTrigger
    total, bf = 9 * 1024 + 77, 1024
    wl = synth.fm_bank(23, total, 48000.0, seed=4, mean_note=0.002)   # ~10 events per buffer
    ev, eo = wl["events"], wl["event_off"]
    ev["frame"][eo[3] + 1] = ev["frame"][eo[3]]  # two impulses on one frame: the second wins (trigger_test.zig:117-132)
    # voice 0: 40 events inside mix buffer 2 -> the ImpulseQueue keeps the first 32 (notes.zig:108-118)
    first_in_b2 = int(eo[0] + np.searchsorted(ev["frame"][eo[0]:eo[1]], 2 * bf))
    assert first_in_b2 + 40 < eo[1]
    ev["frame"][first_in_b2:first_in_b2 + 40] = 2 * bf + 5 + 20 * np.arange(40)
    later = slice(first_in_b2 + 40, int(eo[1]))
    ev["frame"][later] = np.maximum(ev["frame"][later], 3 * bf + 1)
    ev["frame"][later] = np.maximum.accumulate(ev["frame"][later])
    assert np.diff(eo).max() > 100
    bt = BankTrigger(eo, ev, total, bf)
    for first, n in ((0, 4 * bf), (4 * bf, 2 * bf), (6 * bf, total - 6 * bf)):
        off, spans = bt.spans(first, n)
        off2, spans2 = synth.spans_from_events(eo, ev, total, bf, first_frame=first, num_frames=n)
        for v in range(1, eo.size - 1):  # spans_from_events does not model the queue capacity: voice 0 is checked below
            assert np.array_equal(spans[off[v]:off[v + 1]], spans2[off2[v]:off2[v + 1]]), v
        if first == 0:
            s0 = spans[off[0]:off[1]]
            in_b2 = s0[(s0["start"] >= 2 * bf) & (s0["start"] < 3 * bf)]
            assert in_b2["param"].max() == first_in_b2 + 31  # events 33..40 of that buffer were dropped
            assert in_b2["end"].max() == 3 * bf
