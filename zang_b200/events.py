"""Python face of the host-side event layer exported by libzangcu.so (src/zang/notes.zig, src/zang/trigger.zig).

All logic lives in zang_b200/csrc/zcu_events.cpp; these classes only marshal arguments so the tests can
be written like the reference's own (src/zang/trigger_test.zig, src/zang/notes_test.zig).
"""
import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import lib, ptr


class IdGenerator:
    def __init__(self):
        self._g = L.IdGeneratorPOD()
        lib.zcu_id_generator_init(C.byref(self._g))

    def nextId(self):
        return int(lib.zcu_id_generator_next(C.byref(self._g)))


class ImpulseQueue:
    """Notes(P).ImpulseQueue: push(impulse_frame, note_id, params) / consume()."""

    def __init__(self):
        self._q = L.ImpulseQueuePOD()
        lib.zcu_impulse_queue_init(C.byref(self._q))
        self._rows = []  # params values by row

    def push(self, impulse_frame, note_id, params):
        before = self._q.length
        lib.zcu_impulse_queue_push(C.byref(self._q), int(impulse_frame), int(note_id), len(self._rows))
        if self._q.length != before:
            self._rows.append(params)

    def consume(self):
        imp = np.zeros(32, dtype=L.impulse_dtype)
        rows = np.zeros(32, dtype=np.uint32)
        n = lib.zcu_impulse_queue_consume(C.byref(self._q), ptr(imp), ptr(rows))
        params = [self._rows[r] for r in rows[:n]]
        self._rows = []
        return imp[:n].copy(), params


class Trigger:
    """Trigger(P): `run(span, impulses, params)` returns the list the reference's
    `while (trigger.next(&ctr))` loop would yield: (start, end, params, note_id_changed)."""

    def __init__(self):
        self._t = L.TriggerPOD()
        lib.zcu_trigger_init(C.byref(self._t))
        self._cur = None  # params value of the carried note

    def reset(self):
        lib.zcu_trigger_reset(C.byref(self._t))
        self._cur = None

    def run(self, span, impulses, params, frame_base=0):
        start, end = span
        imp = np.ascontiguousarray(impulses, dtype=L.impulse_dtype).reshape(-1)
        n = imp.size
        # row 0 = the carried note's params, rows 1.. = this call's impulses
        table = [self._cur] + list(params)
        self._t.param = 0
        rows = np.arange(1, n + 1, dtype=np.uint32)
        out = np.zeros(n + 2, dtype=L.span_dtype)
        cnt = lib.zcu_trigger_run(C.byref(self._t), start, end, ptr(imp), ptr(rows), n, frame_base, ptr(out), out.size)
        assert cnt <= out.size
        res = [(int(s["start"]), int(s["end"]), table[int(s["param"])], bool(s["flags"] & L.SPAN_NOTE_ID_CHANGED)) for s in out[:cnt]]
        if self._t.has_note:
            self._cur = table[self._t.param]
        return res


class NoteTracker:
    """Notes(P).NoteTracker over a song = list of (t, note_id, params)."""

    def __init__(self, song):
        self._params = [e[2] for e in song]
        self._song = np.zeros(len(song), dtype=L.song_event_dtype)
        for i, (t, nid, _) in enumerate(song):
            self._song[i] = (t, nid, i)
        self._nt = L.NoteTrackerPOD()
        lib.zcu_note_tracker_init(C.byref(self._nt), ptr(self._song), len(song))

    def reset(self):
        lib.zcu_note_tracker_reset(C.byref(self._nt))

    def consume(self, sample_rate, span):
        imp = np.zeros(32, dtype=L.impulse_dtype)
        rows = np.zeros(32, dtype=np.uint32)
        n = lib.zcu_note_tracker_consume(C.byref(self._nt), float(sample_rate), span[0], span[1], ptr(imp), ptr(rows), 32)
        return imp[:n].copy(), [self._params[r] for r in rows[:n]]


class PolyphonyDispatcher:
    """Notes(P).PolyphonyDispatcher(polyphony); params values must expose note_on (dict key or attribute)."""

    def __init__(self, polyphony):
        self.polyphony = polyphony
        self._pd = L.PolyDispatcherPOD()
        rc = lib.zcu_poly_dispatcher_init(C.byref(self._pd), polyphony)
        if rc != L.OK:
            raise ValueError("polyphony must be 1..32")

    def reset(self):
        lib.zcu_poly_dispatcher_reset(C.byref(self._pd))

    def dispatch(self, impulses, params):
        imp = np.ascontiguousarray(impulses, dtype=L.impulse_dtype).reshape(-1)
        n = imp.size
        rows = np.arange(n, dtype=np.uint32)
        on = np.array([1 if (p["note_on"] if isinstance(p, dict) else p.note_on) else 0 for p in params], dtype=np.int32)
        oi = np.zeros((self.polyphony, 32), dtype=L.impulse_dtype)
        op = np.zeros((self.polyphony, 32), dtype=np.uint32)
        oc = np.zeros(self.polyphony, dtype=np.uint32)
        rc = lib.zcu_poly_dispatcher_dispatch(C.byref(self._pd), ptr(imp), ptr(rows), ptr(on), n, ptr(oi), ptr(op), ptr(oc))
        assert rc == L.OK
        return [(oi[s, :oc[s]].copy(), [params[r] for r in op[s, :oc[s]]]) for s in range(self.polyphony)]


class BankTrigger:
    """One ImpulseQueue + Trigger per voice of a bank (examples/example_polyphony.zig:61-92 per voice), replayed on the
    host by libzangcu (zcu_bank_spans).  Events carry absolute frames; span.param is the global event index."""

    def __init__(self, event_off, events, total_frames, buffer_frames=1024):
        self.event_off = np.ascontiguousarray(event_off, dtype=np.uint32)
        self.events = np.ascontiguousarray(events, dtype=L.note_event_dtype)
        self.voices = self.event_off.size - 1
        self.total_frames, self.buffer_frames = int(total_frames), int(buffer_frames)
        self.state = np.zeros(self.voices, dtype=L.bank_trigger_dtype)
        lib.zcu_bank_trigger_init(ptr(self.state), self.voices, ptr(self.event_off))

    def spans(self, first_frame, num_frames):
        """(voice_off[V+1], spans) of the block; Trigger state is carried to the next call."""
        nbuf = (num_frames + self.buffer_frames - 1) // self.buffer_frames
        cap = self.voices * nbuf + int(self.events.size) + 16
        off = np.zeros(self.voices + 1, dtype=np.uint32)
        spans = np.zeros(cap, dtype=L.span_dtype)
        n = lib.zcu_bank_spans(ptr(self.state), self.voices, ptr(self.event_off), ptr(self.events), self.buffer_frames, self.total_frames,
                               int(first_frame), int(num_frames), ptr(off), ptr(spans), cap)
        if n < 0:
            raise RuntimeError(f"span table needs {-n} rows, {cap} provided")
        return off, spans[:n].copy()


def _err(ctx):
    m = lib.zcu_last_error(ctx._h)
    return m.decode() if m else ""


class DeviceSpans:
    """A span table that was built on the device (same handle type as engine.Spans)."""

    def __init__(self, ctx, handle, voices, frames):
        self.ctx, self._h, self.voices, self.frames = ctx, handle, voices, frames

    def download(self):
        """(voice_off[V+1], spans) read back from the device."""
        off = np.zeros(self.voices + 1, dtype=np.uint32)
        n = C.c_uint64()
        rc = lib.zcu_spans_download(self.ctx._h, self._h, ptr(off), None, 0, C.byref(n))
        assert rc == L.OK, _err(self.ctx)
        spans = np.zeros(max(int(n.value), 1), dtype=L.span_dtype)
        rc = lib.zcu_spans_download(self.ctx._h, self._h, ptr(off), ptr(spans), int(n.value), C.byref(n))
        assert rc == L.OK, _err(self.ctx)
        return off, spans[:int(n.value)].copy()

    def free(self):
        if self._h:
            lib.zcu_spans_free(self._h)
            self._h = None


class DeviceBankTrigger:
    """BankTrigger with the events resident on the GPU and the Trigger replayed there (zcu_event_bank_*): `spans()` returns a
    device span table without any per-block host work."""

    def __init__(self, ctx, event_off, events, total_frames, buffer_frames=1024):
        self.ctx = ctx
        off = np.ascontiguousarray(event_off, dtype=np.uint32)
        ev = np.ascontiguousarray(events, dtype=L.note_event_dtype)
        self.voices = off.size - 1
        self.total_frames, self.buffer_frames = int(total_frames), int(buffer_frames)
        h = C.c_void_p()
        rc = lib.zcu_event_bank_create(ctx._h, self.voices, ptr(off), ptr(ev), self.buffer_frames, self.total_frames, C.byref(h))
        if rc != L.OK:
            raise RuntimeError(f"zcu_event_bank_create: {_err(ctx)}")
        self._h = h

    def reset(self):
        assert lib.zcu_event_bank_reset(self.ctx._h, self._h) == L.OK

    def spans(self, first_frame, num_frames):
        h = C.c_void_p()
        rc = lib.zcu_event_bank_spans(self.ctx._h, self._h, int(first_frame), int(num_frames), C.byref(h))
        if rc != L.OK:
            raise RuntimeError(f"zcu_event_bank_spans: {_err(self.ctx)}")
        return DeviceSpans(self.ctx, h, self.voices, int(num_frames))

    def prev_note_on_views(self, src):
        """(porta, env, plain) views of a Trigger table for an instrument that keeps prev_note_on (example_portamento.zig:51)."""
        hs = [C.c_void_p() for _ in range(3)]
        rc = lib.zcu_spans_prev_note_on_views(self.ctx._h, self._h, src._h, *[C.byref(h) for h in hs])
        if rc != L.OK:
            raise RuntimeError(f"zcu_spans_prev_note_on_views: {_err(self.ctx)}")
        return tuple(DeviceSpans(self.ctx, h, self.voices, src.frames) for h in hs)

    def state(self):
        st = np.zeros(self.voices, dtype=L.bank_trigger_dtype)
        assert lib.zcu_event_bank_get_state(self.ctx._h, self._h, ptr(st)) == L.OK
        return st

    def free(self):
        if self._h:
            lib.zcu_event_bank_free(self._h)
            self._h = None


class DevicePolyBank:
    """Many independent polyphonic voices (NoteTracker -> PolyphonyDispatcher -> Trigger per slot) replayed on the GPU
    (zcu_poly_bank_*): songs = list of [(t, note_id, params_row)], one polyphony for all."""

    def __init__(self, ctx, songs, polyphony, param_note_on, sample_rate, buffer_frames=1024):
        self.ctx, self.songs, self.polyphony, self.buffer_frames = ctx, len(songs), int(polyphony), int(buffer_frames)
        off = np.zeros(len(songs) + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(s) for s in songs])
        ev = np.zeros(int(off[-1]), dtype=L.song_event_dtype)
        k = 0
        for s in songs:
            for t, nid, row in s:
                ev[k] = (t, nid, row)
                k += 1
        on = None if param_note_on is None else np.ascontiguousarray(param_note_on, dtype=np.int32)
        h = C.c_void_p()
        rc = lib.zcu_poly_bank_create(ctx._h, self.songs, self.polyphony, ptr(off), ptr(ev), ptr(on), 0 if on is None else on.size,
                                      float(sample_rate), self.buffer_frames, C.byref(h))
        if rc != L.OK:
            raise RuntimeError(f"zcu_poly_bank_create: {_err(ctx)}")
        self._h = h

    def spans(self, num_buffers, last_buffer_frames=None):
        last = self.buffer_frames if last_buffer_frames is None else int(last_buffer_frames)
        h = C.c_void_p()
        rc = lib.zcu_poly_bank_spans(self.ctx._h, self._h, int(num_buffers), last, C.byref(h))
        if rc != L.OK:
            raise RuntimeError(f"zcu_poly_bank_spans: {_err(self.ctx)}")
        return DeviceSpans(self.ctx, h, self.songs * self.polyphony, (int(num_buffers) - 1) * self.buffer_frames + last)

    def reset(self):
        assert lib.zcu_poly_bank_reset(self.ctx._h, self._h) == L.OK

    def state(self):
        st = (L.PolyVoicePOD * self.songs)()
        assert lib.zcu_poly_bank_get_state(self.ctx._h, self._h, st) == L.OK
        return st

    def free(self):
        if self._h:
            lib.zcu_poly_bank_free(self._h)
            self._h = None
