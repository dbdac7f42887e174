"""Renders the whole 385 s Toccata (BASELINE.json configs[0]) on cuda:0 through the per-module C ABI, times it, and
compares the PCM with the oracle's.  Usage: python tools/song_full.py [out.wav]"""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # checker only  # noqa: E402
from zang_b200 import engine as E  # noqa: E402
from zang_b200 import song  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "toccata_events.npz"))
notes = [z["pedal"], z["regular_organ"], z["weird_organ"]]
gold = json.load(open(os.path.join(ROOT, "tests", "golden", "song_golden.json")))
total = gold["total_frames"]
ctx = E.Context(0)
song.render_song(ctx, notes, 64 * 1024)  # warm-up
t0 = time.time()
l0 = ctx.launch_count
pcm, _ = song.render_song(ctx, notes, total, block_buffers=1024)
gpu_s = time.time() - t0
launches = ctx.launch_count - l0
want, _, cpu_s = oracle.song_render(notes, total, want_mix=False)
assert hashlib.sha256(want.tobytes()).hexdigest() == gold["pcm_sha256"][str(total)], "oracle drifted from its golden hash"
d = np.abs(pcm.view("<i2").astype(np.int32) - want.view("<i2").astype(np.int32))
res = {"workload": "write_wav Toccata, 385 s @48 kHz, 17 sub-voices (configs[0])", "frames": total, "gpu_wall_s": round(gpu_s, 3),
       "oracle_1_thread_s": round(cpu_s, 3), "gpu_launches": launches, "pcm_max_abs_diff_lsb": int(d.max()),
       "pcm_samples_differing": int((d != 0).sum()), "pcm_sha256_gpu": hashlib.sha256(pcm.tobytes()).hexdigest(),
       "pcm_sha256_oracle": gold["pcm_sha256"][str(total)]}
print(json.dumps(res))
if len(sys.argv) > 1:
    with open(sys.argv[1], "wb") as f:
        f.write(song.wav_header(1, song.AUDIO_SAMPLE_RATE, 2, pcm.size))
        f.write(pcm.tobytes())
