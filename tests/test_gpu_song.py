"""BASELINE.json configs[0] — the reference's own CPU-runnable workload on the device path: write_wav
(examples/write_wav.zig) of the Toccata song (examples/example_song.zig), 17 sub-voices of PMOscInstrument /
NiceInstrument behind NoteTracker -> PolyphonyDispatcher -> Trigger, summed in the reference's order and
mixed down to 16-bit PCM.

Every per-sample operation of this graph is strict f32 except SineOsc's sin (CUDA sinf vs the oracle's glibc sinf,
<= 2e-6 abs, tests/test_gpu_modules.py), so the f32 mix is held to 17 voices x 2e-6 and the PCM to +-1 LSB; the
NiceInstrument-only stretch of the song (the pedal enters at 9.75 s) is BIT-EXACT in f32 and in PCM."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from helpers import assert_bit_exact
from zang_b200 import song

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def load_fixture():
    z = np.load(os.path.join(HERE, "golden", "toccata_events.npz"))
    return [z["pedal"], z["regular_organ"], z["weird_organ"]]


def pcm_i16(pcm):
    return pcm.view("<i2").astype(np.int32)


def test_song_opening_bit_exact(ctx):
    # before the pedal's first note (t = 9.75 s) only NiceInstrument voices sound: no transcendental per sample
    notes = load_fixture()
    n = 400 * 1024 + 123  # 8.5 s, ends inside a mix buffer
    want_pcm, want_mix, _ = oracle.song_render(notes, n)
    got_pcm, got_mix = song.render_song(ctx, notes, n, block_buffers=96, want_mix=True)
    assert np.abs(want_mix).max() > 0.5
    assert_bit_exact(got_mix, want_mix, "song mix (NiceInstrument stretch)")
    assert got_pcm.tobytes() == want_pcm.tobytes()
    gold = json.load(open(os.path.join(HERE, "golden", "song_golden.json")))
    assert hashlib.sha256(got_pcm[:64 * 1024 * 2].tobytes()).hexdigest() == gold["pcm_sha256"][str(64 * 1024)]


def test_song_dispatched_on_the_device(ctx):
    """configs[0] with the whole event path on the GPU: NoteTracker -> PolyphonyDispatcher -> Trigger of the three voices
    (polyphony 3 / 10 / 4, example_song.zig:28-74) replayed by zcu_poly_bank_*; the PCM must be the host-dispatched render's,
    byte for byte, over 16 s that include the pedal's entry, with a block size that does not divide anything."""
    notes = load_fixture()
    n = 750 * 1024 + 77
    host_pcm, host_mix = song.render_song(ctx, notes, n, block_buffers=96, want_mix=True)
    dev_pcm, dev_mix = song.render_song(ctx, notes, n, block_buffers=61, want_mix=True, device_events=True)
    assert np.abs(host_mix).max() > 0.5
    assert_bit_exact(dev_mix, host_mix, "song mix, device-side dispatch vs host-side dispatch")
    assert dev_pcm.tobytes() == host_pcm.tobytes()


def test_song_with_pedal_within_tolerance(ctx):
    notes = load_fixture()
    n = 1500 * 1024  # 32 s: pedal (PMOscInstrument, two SineOscs) + regular organ
    want_pcm, want_mix, _ = oracle.song_render(notes, n)
    got_pcm, got_mix = song.render_song(ctx, notes, n, block_buffers=512, want_mix=True)
    err = float(np.abs(got_mix.astype(np.float64) - want_mix).max())
    assert err <= 17 * 2e-6, err
    d = np.abs(pcm_i16(got_pcm) - pcm_i16(want_pcm))
    assert d.max() <= 1, d.max()
    assert (d != 0).mean() < 1e-3
    # block size must not matter: the same bits with a different block partition
    got_pcm2, got_mix2 = song.render_song(ctx, notes, 300 * 1024, block_buffers=7, want_mix=True)
    assert_bit_exact(got_mix2, got_mix[:300 * 1024], "block partition independence")
    assert got_pcm2.tobytes() == got_pcm[:300 * 1024 * 2].tobytes()


def test_song_weird_organ_section(ctx):
    # the third instrument only plays at 178-222 s; move it to the start so that all three instruments overlap
    notes = load_fixture()
    notes[2] = notes[2].copy()
    notes[2]["t"] -= np.float32(170.0)
    n = 900 * 1024
    want_pcm, want_mix, _ = oracle.song_render(notes, n)
    got_pcm, got_mix = song.render_song(ctx, notes, n, block_buffers=300, want_mix=True)
    assert float(np.abs(got_mix.astype(np.float64) - want_mix).max()) <= 17 * 2e-6
    assert np.abs(pcm_i16(got_pcm) - pcm_i16(want_pcm)).max() <= 1
