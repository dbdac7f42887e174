"""The reference's own CPU-runnable workload on the device path (BASELINE.json configs[0]):
examples/write_wav.zig rendering examples/example_song.zig (Bach's Toccata, 3 polyphonic instruments,
3 + 10 + 4 sub-voices) into 16-bit PCM.

Host side (this file + the C event layer of libzangcu.so): the tracker-text parser
(examples/common/songparse1.zig), doParse (example_song.zig:129-264) and, per block of mix buffers,
Voice.paint's event logic (NoteTracker -> PolyphonyDispatcher -> Trigger, zcu_poly_voice_spans).
Device side: every module paint of PMOscInstrument / NiceInstrument (examples/modules.zig) through the
per-module C ABI on [frame][sub-voice] buffers, the ordered voice sum (the reference accumulates the
sub-voices one after the other into outputs[0]) and zang.mixDown.  There is no CPU fallback.
"""
import ctypes as C
import ctypes.util
import struct

import numpy as np

from . import _lib as L
from . import engine as E
from ._lib import lib, ptr

AUDIO_SAMPLE_RATE = 48000      # example_song.zig:7
AUDIO_BUFFER_SIZE = 1024       # example_song.zig:8
NUM_SECONDS = 6 * 60 + 25      # write_wav.zig:7
A4 = 440.0                     # example_song.zig:19
NOTE_DURATION = np.float32(0.15)  # example_song.zig:20
COLUMNS_PER_VOICE = (2, 8, 2)  # Pedal, RegularOrgan, WeirdOrgan (example_song.zig:41,57,73)
TOTAL_COLUMNS = sum(COLUMNS_PER_VOICE)

_libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.powf.restype = C.c_float
_libm.powf.argtypes = [C.c_float, C.c_float]

_SEMITONES = {"C-": 0, "C#": 1, "D-": 2, "D#": 3, "E-": 4, "F-": 5, "F#": 6, "G-": 7, "G#": 8, "A-": 9, "A#": 10, "B-": 11}

song_note_dtype = np.dtype([("t", "<f4"), ("note_id", "<u4"), ("freq", "<f4"), ("note_on", "<i4")])


class SongSyntaxError(ValueError):
    pass


class _Parser:
    """songparse1.zig:3-198.  Tokens: ('word', str) | ('number', f32) | ('notes', [None | 'off' | f32] * columns)."""

    def __init__(self, contents, num_columns, a4=A4):
        self.s, self.i, self.line, self.ncol, self.a4 = contents, 0, 0, num_columns, np.float32(a4)

    def _note(self):  # :29-63
        s, i = self.s, self.i
        if i + 3 > len(s):
            return None
        octave = s[i + 2]
        if not ("0" <= octave <= "9"):
            return None
        semitone = _SEMITONES.get(s[i:i + 2])
        if semitone is None:
            return None
        self.i += 3
        exp = np.float32(np.float32((ord(octave) - 48) * 12 - 57 + semitone) / np.float32(12.0))
        return np.float32(self.a4 * np.float32(_libm.powf(2.0, float(exp))))  # std.math.pow(f32, 2.0, exp): host scalar, libm here

    def _eat(self, prefix):
        if self.s.startswith(prefix, self.i):
            self.i += len(prefix)
            return True
        return False

    def token(self):  # :74-185
        s = self.s
        while True:
            if self._eat(" "):
                pass
            elif self._eat("\n"):
                self.line += 1
            elif self._eat("#"):
                pos = s.find("\n", self.i)
                if pos >= 0:
                    self.line += 1
                    self.i = pos + 1
                else:
                    self.i = len(s)
            else:
                break
        if self.i >= len(s):
            return None
        ch = s[self.i]
        if ch == "|":
            self.i += 1
            notes = [None] * self.ncol
            col = 0
            while True:
                if col >= self.ncol:
                    raise SongSyntaxError(f"too many columns on line {self.line + 1}")
                f = self._note()
                if f is not None:
                    notes[col] = f
                elif self._eat("off"):
                    notes[col] = "off"
                elif self._eat("   "):
                    pass
                else:
                    break
                if self.i < len(s) and s[self.i] in " |":
                    self.i += 1
                else:
                    break
                col += 1
            if self.i < len(s):
                if s[self.i] == "\n":
                    self.line += 1
                    self.i += 1
                else:
                    raise SongSyntaxError(f"syntax error on line {self.line + 1}")
            return ("notes", notes)
        if ch.isascii() and (ch.isalpha() or ch == "_"):
            start = self.i
            self.i += 1
            while self.i < len(s) and s[self.i].isascii() and (s[self.i].isalnum() or s[self.i] == "_"):
                self.i += 1
            return ("word", s[start:self.i])
        if "0" <= ch <= "9":
            start = self.i
            dot = False
            self.i += 1
            while self.i < len(s):
                c2 = s[self.i]
                if c2 == ".":
                    if dot:
                        break
                    dot = True
                    self.i += 1
                elif "0" <= c2 <= "9":
                    self.i += 1
                else:
                    break
            return ("number", np.float32(s[start:self.i]))  # std.fmt.parseFloat(f32): correctly rounded
        raise SongSyntaxError(f"syntax error on line {self.line + 1}")


def parse_song(text):
    """doParse (example_song.zig:129-264): returns one chronological event array (song_note_dtype) per instrument."""
    if isinstance(text, bytes):
        text = text.decode("latin-1")
    parser = _Parser(text, TOTAL_COLUMNS)
    col_instrument = [k for k, n in enumerate(COLUMNS_PER_VOICE) for _ in range(n)]
    notes = [[] for _ in COLUMNS_PER_VOICE]
    last = [None] * TOTAL_COLUMNS  # (freq, id) of the note sounding in the column
    next_id = 1
    t, rate, tempo = np.float32(0.0), np.float32(1.0), np.float32(1.0)
    while True:
        tok = parser.token()
        if tok is None:
            break
        kind, val = tok
        if kind == "word" and val == "start":
            t = np.float32(0.0)
            notes = [[] for _ in COLUMNS_PER_VOICE]
        elif kind == "word" and val in ("rate", "tempo"):
            num = parser.token()
            if num is None or num[0] != "number":
                raise SongSyntaxError(f"expected a number on line {parser.line + 1}")
            if val == "rate":
                rate = num[1]
            else:
                tempo = num[1]
        elif kind == "notes":
            old = [len(n) for n in notes]
            for col, note in enumerate(val):
                dst = notes[col_instrument[col]]
                if note is None:
                    continue
                if isinstance(note, str):  # off
                    if last[col] is not None:
                        dst.append((t, last[col][1], last[col][0], 0))
                        last[col] = None
                else:
                    if last[col] is not None:  # note-off of the previous note in this column
                        dst.append((t, last[col][1], last[col][0], 0))
                    dst.append((t, next_id, note, 1))
                    last[col] = (note, next_id)
                    next_id += 1
            t = np.float32(t + np.float32(NOTE_DURATION / np.float32(rate * tempo)))  # :213
            for k, n in enumerate(notes):  # this row's events by note id: note-offs before note-ons (:215-234)
                n[old[k]:] = sorted(n[old[k]:], key=lambda e: e[1])
        else:
            raise SongSyntaxError(f"bad token on line {parser.line + 1}")
    return [np.array(n, dtype=song_note_dtype) if n else np.zeros(0, dtype=song_note_dtype) for n in notes]


class PolyVoice:
    """Voice(T) (example_song.zig:287-350): NoteTracker + PolyphonyDispatcher + one Trigger per sub-voice on
    the host, the instrument bank with `polyphony` voice columns on the device."""

    def __init__(self, ctx, bank, notes, polyphony, freq_mul=1.0, sample_rate=AUDIO_SAMPLE_RATE, device_events=False):
        self.ctx, self.bank, self.polyphony, self.sample_rate = ctx, bank, polyphony, float(sample_rate)
        self.dev = None
        self.song = np.zeros(notes.size, dtype=L.song_event_dtype)
        self.song["t"], self.song["note_id"], self.song["param"] = notes["t"], notes["note_id"], np.arange(notes.size)
        self.note_on = np.ascontiguousarray(notes["note_on"], dtype=np.int32)
        # T.makeParams (:33-39, :49-55, :65-71): one instrument params row per song event
        self.params = np.zeros(notes.size, dtype=bank.Params)
        self.params["sample_rate"] = sample_rate
        self.params["freq"] = notes["freq"] * np.float32(freq_mul)
        self.params["note_on"] = notes["note_on"]
        self.pod = L.PolyVoicePOD()
        rc = lib.zcu_poly_voice_init(C.byref(self.pod), ptr(self.song), self.song.size, polyphony)
        if rc != L.OK:
            raise E.ZangCuError(f"zcu_poly_voice_init failed ({rc})")
        self.bufs = None
        if device_events:  # the same event logic replayed on the GPU (zcu_poly_bank_*): no span table crosses PCIe
            from .events import DevicePolyBank
            song = list(zip(self.song["t"].tolist(), self.song["note_id"].tolist(), self.song["param"].tolist()))
            self.dev = DevicePolyBank(ctx, [[(np.float32(t), nid, row) for t, nid, row in song]], polyphony, self.note_on, sample_rate, AUDIO_BUFFER_SIZE)

    def _buffers(self, frames):
        if self.bufs is None or self.bufs[0].frames != frames:
            for b in self.bufs or ():
                b.free()
            self.bufs = [self.ctx.buf(frames, self.polyphony) for _ in range(1 + self.bank.num_temps)]
        return self.bufs

    def span_table(self, num_buffers, buffer_frames, last_frames):
        cap = max(64, num_buffers * self.polyphony * 2)
        while True:
            off = np.zeros(self.polyphony + 1, dtype=np.uint32)
            spans = np.zeros(cap, dtype=L.span_dtype)
            n = lib.zcu_poly_voice_spans(C.byref(self.pod), ptr(self.note_on), self.sample_rate, buffer_frames, num_buffers, last_frames,
                                         ptr(off), ptr(spans), cap)
            if n >= 0:
                return off, spans[:n]
            cap = -n

    def paint(self, num_buffers, buffer_frames, last_frames, mix):
        """Paints the block's spans of every sub-voice and adds the sub-voices, in order, into mix ([frames][1])."""
        frames = (num_buffers - 1) * buffer_frames + last_frames
        if self.dev is not None:
            assert buffer_frames == self.dev.buffer_frames
            out, *temps = self._buffers(frames)
            sp = self.dev.spans(num_buffers, last_frames)
            self.ctx.zero(None, out)
            self.bank.paint(sp, [out], temps, self.params)
            self.ctx.mix_ordered(out, None, 1, mix)
            sp.free()
            return -1
        off, spans = self.span_table(num_buffers, buffer_frames, last_frames)
        if spans.size == 0:
            return 0
        out, *temps = self._buffers(frames)
        sp = self.ctx.spans(self.polyphony, frames, off, spans)
        self.ctx.zero(None, out)
        self.bank.paint(sp, [out], temps, self.params)
        self.ctx.mix_ordered(out, None, 1, mix)
        sp.free()
        return int(spans.size)

    def free(self):
        for b in self.bufs or ():
            b.free()
        self.bufs = None
        if self.dev is not None:
            self.dev.free()
            self.dev = None


def wav_header(num_channels, sample_rate, bytes_per_sample, data_bytes):
    """Canonical 44-byte PCM RIFF header (zig-wav's writeHeader + patchHeader; the submodule is not in the tree)."""
    return struct.pack("<4sI4s4sIHHIIHH4sI", b"RIFF", 36 + data_bytes, b"WAVE", b"fmt ", 16, 1, num_channels, sample_rate,
                       sample_rate * num_channels * bytes_per_sample, num_channels * bytes_per_sample, 8 * bytes_per_sample, b"data", data_bytes)


def render_song(ctx, notes, total_frames=NUM_SECONDS * AUDIO_SAMPLE_RATE, block_buffers=512, want_mix=False, device_events=False):
    """write_wav.zig:49-95 for the first total_frames frames.  `notes` is the song text or the three event arrays
    parse_song returns.  Returns (pcm bytes as uint8 array, f32 mix or None)."""
    if isinstance(notes, (str, bytes)):
        notes = parse_song(notes)
    sr, B = AUDIO_SAMPLE_RATE, AUDIO_BUFFER_SIZE
    voices = [  # example_song.zig:28-74
        PolyVoice(ctx, E.PMOscInstrumentBank(ctx, 3, 0.4), notes[0], 3, freq_mul=0.5, sample_rate=sr, device_events=device_events),
        PolyVoice(ctx, E.NiceInstrumentBank(ctx, 10, 0.25), notes[1], 10, sample_rate=sr, device_events=device_events),
        PolyVoice(ctx, E.NiceInstrumentBank(ctx, 4, 0.1), notes[2], 4, sample_rate=sr, device_events=device_events),
    ]
    pcm = np.zeros(total_frames * 2, dtype=np.uint8)
    mix_all = np.zeros(total_frames, dtype=np.float32) if want_mix else None
    start = 0
    mix = None
    while start < total_frames:
        frames = min(block_buffers * B, total_frames - start)
        nb = (frames + B - 1) // B
        last = frames - (nb - 1) * B
        if mix is None or mix.frames != frames:
            if mix is not None:
                mix.free()
            mix = ctx.buf(frames, 1)
        ctx.zero(None, mix)  # write_wav.zig:63-64
        for v in voices:     # MainModule.paint, example_song.zig:382-396
            v.paint(nb, B, last, mix)
        pcm[start * 2:(start + frames) * 2] = ctx.mixDown(mix, L.AUDIO_SIGNED16_LSB, 0.25)  # write_wav.zig:71-78
        if want_mix:
            mix_all[start:start + frames] = mix.download()[:, 0]
        start += frames
    if mix is not None:
        mix.free()
    for v in voices:
        v.free()
    return pcm, mix_all


def write_wav(ctx, notes, path, total_frames=NUM_SECONDS * AUDIO_SAMPLE_RATE, block_buffers=512):
    pcm, _ = render_song(ctx, notes, total_frames, block_buffers)
    with open(path, "wb") as f:
        f.write(wav_header(1, AUDIO_SAMPLE_RATE, 2, pcm.size))
        f.write(pcm.tobytes())
    return pcm
