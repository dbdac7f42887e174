"""Regenerates the Toccata fixtures (BASELINE.json configs[0], the reference's own CPU-runnable workload).

  tests/golden/toccata_events.npz   the score examples/example_song.txt of the reference after doParse
                                    (example_song.zig:129-264): per instrument (t, note_id, freq, note_on).
                                    /root/reference does not exist on the GPU box, so the parsed events travel
                                    instead of the text.  Produced by the ORACLE's parser from the reference's
                                    file; tests/test_song_cpu.py checks the product's parser against the same
                                    text whenever /root/reference is present.
  tests/golden/song_golden.json     sha256 of the oracle's PCM for the whole 385 s render and for prefixes.

The reference ships no golden out.wav and cannot be run here (Zig): these pin the oracle to itself.
Run:  python tests/golden/make_song_golden.py   (needs /root/reference)
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402

SONG = "/root/reference/examples/example_song.txt"
TOTAL = 385 * 48000  # write_wav.zig:7


def main():
    text = open(SONG, "rb").read()
    notes = oracle.song_parse(text)
    np.savez_compressed(os.path.join(HERE, "toccata_events.npz"), pedal=notes[0], regular_organ=notes[1], weird_organ=notes[2])
    pcm, mix, secs = oracle.song_render(notes, TOTAL)
    gold = {
        "source": "examples/example_song.txt (md5 %s) through oracle/song_oracle.hpp" % hashlib.md5(text).hexdigest(),
        "events": [int(n.size) for n in notes],
        "total_frames": TOTAL,
        "pcm_sha256": {str(n): hashlib.sha256(pcm[:n * 2].tobytes()).hexdigest() for n in (64 * 1024, 600 * 1024, 4000 * 1024 + 321, TOTAL)},
        "mix_abs_max": float(np.abs(mix).max()),
        "oracle_seconds_1_thread": round(secs, 2),
    }
    json.dump(gold, open(os.path.join(HERE, "song_golden.json"), "w"), indent=1)
    print(gold)


if __name__ == "__main__":
    main()
