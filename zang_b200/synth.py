"""Synthetic workloads for the benchmark configurations.  Deterministic from one seed.

polyphony bank (BASELINE.json configs[1]): V voices of NiceInstrument, each driven by its own note
stream: alternating note-on / note-off with exponentially distributed durations (mean 0.4 s on,
0.2 s off), a fresh note id and a fresh MIDI-like pitch for every note-on, frames anywhere in the mix
buffer.  Every voice has its own pulse colour and pan position.
"""
import numpy as np

from . import _lib as L

SEED = 0x5EED


def polyphony_bank(voices, total_frames, sample_rate=48000.0, seed=SEED, mean_on=0.4, mean_off=0.2):
    """Returns dict(event_off[V+1] u32, events[note_event_dtype], color[V] f32, gains[V][2] f32)."""
    rng = np.random.default_rng(seed)
    total_t = total_frames / sample_rate
    k = int(total_t / (mean_on + mean_off) * 1.6) + 16
    on_d = rng.exponential(mean_on, size=(voices, k))
    off_d = rng.exponential(mean_off, size=(voices, k))
    first = rng.uniform(0.0, mean_off, size=(voices, 1))
    # event times: on_0 = first, off_0 = on_0 + on_d0, on_1 = off_0 + off_d0, ...
    seg = np.stack([on_d, off_d], axis=2).reshape(voices, 2 * k)
    t = first + np.concatenate([np.zeros((voices, 1)), np.cumsum(seg, axis=1)[:, :-1]], axis=1)  # [V, 2k]
    frames = np.floor(t * sample_rate).astype(np.int64)
    is_on = np.tile(np.array([1, 0], dtype=np.int32), k)[None, :].repeat(voices, axis=0)
    note_num = rng.integers(36, 97, size=(voices, k))
    freq_note = (np.float32(440.0) * np.power(np.float32(2.0), ((note_num - 69).astype(np.float32) / np.float32(12.0)))).astype(np.float32)
    freq = np.repeat(freq_note, 2, axis=1)
    note_id = np.repeat(np.arange(1, k + 1, dtype=np.uint32)[None, :], 2, axis=1).reshape(1, 2 * k).repeat(voices, axis=0)
    keep = frames < total_frames
    counts = keep.sum(axis=1)
    event_off = np.zeros(voices + 1, dtype=np.uint32)
    np.cumsum(counts, out=event_off[1:])
    ev = np.zeros(int(event_off[-1]), dtype=L.note_event_dtype)
    ev["frame"] = frames[keep].astype(np.uint32)
    ev["note_id"] = note_id[keep]
    ev["freq"] = freq[keep]
    ev["note_on"] = is_on[keep]
    color = rng.uniform(0.1, 0.9, size=voices).astype(np.float32)
    pan = rng.uniform(0.0, 1.0, size=voices).astype(np.float32)
    g = np.float32(1.0 / np.sqrt(voices))
    gains = np.stack([(np.float32(1.0) - pan) * g, pan * g], axis=1).astype(np.float32)
    return {"event_off": event_off, "events": ev, "color": color, "gains": gains}


def shard_voices(wl, rank, world):
    """Contiguous voice range of `rank`: voices are independent, no data-path exchange."""
    V = wl["color"].shape[0]
    a, b = V * rank // world, V * (rank + 1) // world
    off = wl["event_off"]
    ev = wl["events"][off[a]:off[b]]
    return {"event_off": (off[a:b + 1] - off[a]).astype(np.uint32), "events": ev, "color": wl["color"][a:b], "gains": wl["gains"][a:b],
            "voice_range": (a, b)}


def spans_from_events(event_off, events, total_frames, buffer_frames=1024, first_frame=0, num_frames=None):
    """Host Trigger replay (through libzangcu's zcu_trigger_run) of absolute-frame note events into the span
    table of a block [first_frame, first_frame+num_frames), plus the instrument params rows
    (one row per event).  Used by the tests of the un-fused path."""
    import ctypes as C
    from ._lib import lib, ptr
    V = event_off.size - 1
    if num_frames is None:
        num_frames = total_frames - first_frame
    spans_all, off = [], [0]
    out = np.zeros(64, dtype=L.span_dtype)
    for v in range(V):
        trig = L.TriggerPOD()
        lib.zcu_trigger_init(C.byref(trig))
        e0, e1 = int(event_off[v]), int(event_off[v + 1])
        ei = e0
        mine = []
        for start in range(0, first_frame + num_frames, buffer_frames):
            ln = min(buffer_frames, total_frames - start)
            ej = ei
            while ej < e1 and events["frame"][ej] < start + ln:
                ej += 1
            n = ej - ei
            imp = np.zeros(n, dtype=L.impulse_dtype)
            imp["frame"] = events["frame"][ei:ej] - start
            imp["note_id"] = events["note_id"][ei:ej]
            imp["event_id"] = np.arange(ei - e0 + 1, ej - e0 + 1)
            rows = np.arange(ei, ej, dtype=np.uint32)  # params row = global event index
            cnt = lib.zcu_trigger_run(C.byref(trig), 0, ln, ptr(imp), ptr(rows), n, start, ptr(out), out.size)
            assert cnt <= out.size
            if start >= first_frame:
                s = out[:cnt].copy()
                s["start"] -= first_frame
                s["end"] -= first_frame
                mine.append(s)
            ei = ej
        if mine:
            m = np.concatenate(mine)
            spans_all.append(m)
            off.append(off[-1] + m.size)
        else:
            off.append(off[-1])
    spans = np.concatenate(spans_all) if spans_all else np.zeros(0, dtype=L.span_dtype)
    return np.array(off, dtype=np.uint32), spans
