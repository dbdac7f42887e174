"""Synthetic workloads for the benchmark configurations.  Deterministic from one seed.

polyphony bank (BASELINE.json configs[1]): V voices of NiceInstrument, each driven by its own note
stream: alternating note-on / note-off with exponentially distributed durations (mean 0.4 s on,
0.2 s off), a fresh note id and a fresh MIDI-like pitch for every note-on, frames anywhere in the mix
buffer.  Every voice has its own pulse colour and pan position.
"""
import numpy as np

from . import abi as L  # dtypes only: generating a workload must not need the CUDA library

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


def window_bank(wl, f0, f1):
    """The events of a bank inside the frame window [f0, f1), re-based to frame 0, plus every voice's last event before f0 placed
    at frame 0 (the note it is holding when the window opens): a short job whose voices are all in the state a long render has
    in its middle.  Used to measure the module chain on a steady-state block instead of the job's first seconds, in which the
    voices have not started yet."""
    off, ev = wl["event_off"], wl["events"]
    V = off.size - 1
    parts, counts = [], np.zeros(V, dtype=np.int64)
    for v in range(V):
        e = ev[off[v]:off[v + 1]]
        a = int(np.searchsorted(e["frame"], f0, side="left"))
        b = int(np.searchsorted(e["frame"], f1, side="left"))
        w = e[max(a - 1, 0):b].copy() if a > 0 else e[a:b].copy()
        fr = w["frame"].astype(np.int64) - f0
        fr[fr < 0] = 0
        w["frame"] = fr.astype(np.uint32)
        parts.append(w)
        counts[v] = w.size
    out = dict(wl)
    out["event_off"] = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    out["events"] = np.concatenate(parts) if parts else ev[:0]
    return out


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


# ---- BASELINE.json configs[2..4]: synthetic workloads for the other instrument banks ---------------
def _splitmix_rng(seed):
    return np.random.default_rng(seed)


def pcm_loop(frames=192437, seed=SEED):
    """A mono 16-bit loop standing in for examples/drumloop.wav (mono s16 44.1 kHz, 192 437 frames; the asset itself is
    not shipped): decaying noise bursts on a 4-beat grid plus a low sine, deterministic."""
    rng = _splitmix_rng(seed ^ 0xD2)
    t = np.arange(frames, dtype=np.float64)
    beat = frames / 8.0
    env = np.exp(-((t % beat) / (0.12 * beat)))
    x = 0.6 * env * rng.uniform(-1.0, 1.0, size=frames) + 0.3 * np.sin(2 * np.pi * 55.0 * t / 44100.0)
    return np.clip(np.round(x * 32767.0), -32768, 32767).astype("<i2").view(np.uint8).copy()


DRUMLOOP_MD5 = "29dec84da4c4abc7d54416c1d8a3bf5f"  # /root/reference/examples/drumloop.wav


def drumloop_pcm():
    """PCM bytes of the reference's own Sampler asset (examples/example_sampler.zig:66 embeds examples/drumloop.wav: mono
    s16 44.1 kHz, 384 874 data bytes after a 44-byte RIFF header, parsed as example_sampler.zig:26-45 does), shipped as the
    fixture tests/golden/drumloop.wav.  Returns (uint8 array, name); falls back to the synthetic stand-in when the fixture
    is absent, and says so in the name."""
    import hashlib
    import os
    import struct
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "drumloop.wav")
    try:
        raw = open(path, "rb").read()
    except OSError:
        return pcm_loop(), "synthetic stand-in for drumloop.wav"
    assert hashlib.md5(raw).hexdigest() == DRUMLOOP_MD5, "tests/golden/drumloop.wav is not the reference's asset"
    assert raw[:4] == b"RIFF" and raw[8:12] == b"WAVE" and raw[12:16] == b"fmt " and raw[36:40] == b"data"
    fmt, channels, rate, _, _, bits = struct.unpack("<HHIIHH", raw[20:36])
    (nbytes,) = struct.unpack("<I", raw[40:44])
    assert (fmt, channels, rate, bits, nbytes) == (1, 1, 44100, 16, 384874)
    return np.frombuffer(raw[44:44 + nbytes], dtype=np.uint8).copy(), "examples/drumloop.wav"


def sampler_bank(voices, total_frames, sample_rate=44100.0, seed=SEED, mean_gap=0.5):
    """configs[3]: every voice retriggers the loop at random times with a random playback speed
    (example_sampler.zig:121-130: params.sample_rate = AUDIO_SAMPLE_RATE * (0.5 + rand)); a per-voice Decimator rate
    from example_polyphony.zig:151-158 and a pan position.  events.freq carries Sampler.Params.sample_rate."""
    rng = _splitmix_rng(seed ^ 0xC4)
    k = int(total_frames / sample_rate / mean_gap * 1.8) + 8
    gaps = rng.exponential(mean_gap, size=(voices, k))
    gaps[:, 0] = rng.uniform(0.0, 0.05, size=voices)
    frames = np.floor(np.cumsum(gaps, axis=1) * sample_rate).astype(np.int64)
    speed = (np.float32(sample_rate) * (np.float32(0.5) + rng.random(size=(voices, k), dtype=np.float32))).astype(np.float32)
    keep = frames < total_frames
    counts = keep.sum(axis=1)
    event_off = np.zeros(voices + 1, dtype=np.uint32)
    np.cumsum(counts, out=event_off[1:])
    ev = np.zeros(int(event_off[-1]), dtype=L.note_event_dtype)
    ev["frame"] = frames[keep].astype(np.uint32)
    ev["note_id"] = np.tile(np.arange(1, k + 1, dtype=np.uint32), (voices, 1))[keep]
    ev["freq"] = speed[keep]
    ev["note_on"] = 1
    fake = rng.choice(np.array([6000.0, 5000.0, 4000.0, 3000.0, 2000.0, 1000.0], dtype=np.float32), size=voices).astype(np.float32)
    pan = rng.uniform(0.0, 1.0, size=voices).astype(np.float32)
    g = np.float32(1.0 / np.sqrt(voices))
    gains = np.stack([(np.float32(1.0) - pan) * g, pan * g], axis=1).astype(np.float32)
    return {"event_off": event_off, "events": ev, "fake_sample_rate": fake, "gains": gains}


def sfx_bank(jobs, job_frames=24576, seed=SEED):
    """configs[4]: one short sound effect per job (Noise -> Filter with a Curve-driven cutoff -> x Envelope, in the style of
    example_laser.zig / example_stereo.zig).  Note-on at frame 0, note-off at 60 % of the job, same note id."""
    rng = _splitmix_rng(seed ^ 0xC5)
    event_off = (np.arange(jobs + 1, dtype=np.uint32) * 2)
    ev = np.zeros(2 * jobs, dtype=L.note_event_dtype)
    ev["frame"][0::2] = 0
    ev["frame"][1::2] = int(job_frames * 0.6)
    ev["note_id"] = 1
    ev["note_on"][0::2] = 1
    ftype = rng.choice(np.array([L.FILTER_LOW_PASS, L.FILTER_BAND_PASS, L.FILTER_HIGH_PASS, L.FILTER_NOTCH], dtype=np.int32), size=jobs).astype(np.int32)
    # 3-node cutoff curve per job: values are filter cutoff coefficients, times in seconds
    nodes = np.zeros((jobs, 3), dtype=L.curve_node_dtype)
    nodes["value"] = rng.uniform(0.05, 0.8, size=(jobs, 3)).astype(np.float32)
    nodes["t"][:, 1] = rng.uniform(0.03, 0.15, size=jobs).astype(np.float32)
    nodes["t"][:, 2] = nodes["t"][:, 1] + rng.uniform(0.05, 0.25, size=jobs).astype(np.float32)
    pan = rng.uniform(0.0, 1.0, size=jobs).astype(np.float32)
    g = np.float32(1.0 / np.sqrt(jobs))
    gains = np.stack([(np.float32(1.0) - pan) * g, pan * g], axis=1).astype(np.float32)
    return {"event_off": event_off, "events": ev, "filter_type": ftype, "nodes": nodes.reshape(-1), "gains": gains}


def fm_bank(voices, total_frames, sample_rate=48000.0, seed=SEED, mean_note=0.25, p_off=0.3):
    """configs[2]: legato note streams (a new note-on usually arrives while the previous note is still held, so the
    Portamento really glides; with probability p_off the key is released first), MIDI-like pitches."""
    rng = _splitmix_rng(seed ^ 0xC3)
    k = int(total_frames / sample_rate / mean_note * 1.8) + 8
    gaps = rng.exponential(mean_note, size=(voices, k))
    t_on = np.cumsum(gaps, axis=1)
    release = rng.random(size=(voices, k)) < p_off
    t_off = t_on + gaps * rng.uniform(0.3, 0.8, size=(voices, k))  # before the next note-on... of the NEXT gap
    t_off = np.minimum(t_off, np.concatenate([t_on[:, 1:], t_on[:, -1:] + 1.0], axis=1) - 1.0 / sample_rate)
    note_num = rng.integers(36, 85, size=(voices, k))
    freq = (np.float32(440.0) * np.power(np.float32(2.0), ((note_num - 69).astype(np.float32) / np.float32(12.0)))).astype(np.float32)
    ids = np.tile(np.arange(1, k + 1, dtype=np.uint32), (voices, 1))
    # interleave on / optional off
    tt = np.stack([t_on, t_off], axis=2).reshape(voices, 2 * k)
    on = np.tile(np.array([1, 0], dtype=np.int32), k)[None, :].repeat(voices, axis=0)
    valid = np.stack([np.ones_like(release), release], axis=2).reshape(voices, 2 * k)
    fr = np.floor(tt * sample_rate).astype(np.int64)
    keep = valid & (fr < total_frames) & (fr >= 0)
    counts = keep.sum(axis=1)
    event_off = np.zeros(voices + 1, dtype=np.uint32)
    np.cumsum(counts, out=event_off[1:])
    ev = np.zeros(int(event_off[-1]), dtype=L.note_event_dtype)
    ev["frame"] = fr[keep].astype(np.uint32)
    ev["note_id"] = np.repeat(ids, 2, axis=1)[keep]
    ev["freq"] = np.repeat(freq, 2, axis=1)[keep]
    ev["note_on"] = on[keep]
    pan = rng.uniform(0.0, 1.0, size=voices).astype(np.float32)
    g = np.float32(1.0 / np.sqrt(voices))
    gains = np.stack([(np.float32(1.0) - pan) * g, pan * g], axis=1).astype(np.float32)
    return {"event_off": event_off, "events": ev, "gains": gains}


def fm_mod_inputs(voices, frames, first_frame=0, seed=SEED):
    """The per-sample modulation inputs of configs[2] for a block: detune in [0.98, 1.02] (a slow per-voice wobble) and
    FM index in [0, 2], [frames][voices] f32.  Pure functions of (frame, voice) so that any block can be generated alone."""
    f = (np.arange(first_frame, first_frame + frames, dtype=np.float64))[:, None]
    v = np.arange(voices, dtype=np.float64)[None, :]
    detune = 1.0 + 0.02 * np.sin(2 * np.pi * (0.37 + 0.013 * (v % 17)) * f / 48000.0 + 0.7 * v)
    index = 1.0 + np.sin(2 * np.pi * (0.11 + 0.007 * (v % 23)) * f / 48000.0 + 1.3 * v)
    return detune.astype(np.float32), index.astype(np.float32)
