"""Host-side mirror of zang's module contract on top of the C ABI (include/zangcu.h).

The reference paints ONE voice:  Module.paint(span, outputs, temps, note_id_changed, params).
A bank paints EVERY voice of a block:  bank.paint(spans, outputs, temps, params) where `spans` is the
table of all (span, note_id_changed, params-row) records and outputs/temps are [frame][voice] device
buffers.  Names, argument meaning and accumulate (+=) semantics follow src/modules/*.zig.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from ._lib import lib, ptr


class ZangCuError(RuntimeError):
    pass


def _check(ctx, rc):
    if rc != L.OK:
        msg = lib.zcu_last_error(ctx._h if ctx is not None else None)
        names = {1: "INVALID", 2: "NO_DEVICE", 3: "CUDA", 4: "OOM", 5: "UNSUPPORTED", 6: "NCCL"}
        raise ZangCuError(f"zangcu error {rc} ({names.get(rc, '?')}): {msg.decode() if msg else ''}")


class Context:
    def __init__(self, device=0, stream=None):
        h = C.c_void_p()
        if stream is None:
            rc = lib.zcu_ctx_create(int(device), C.byref(h))
        else:
            rc = lib.zcu_ctx_create_on_stream(int(device), C.c_void_p(stream), C.byref(h))
        if rc == L.ERR_NO_DEVICE:
            raise ZangCuError("no CUDA device: libzangcu has no CPU fallback")
        if rc != L.OK:
            raise ZangCuError(f"zcu_ctx_create failed with code {rc}")
        self._h = h
        self.device = device

    def close(self):
        if self._h:
            lib.zcu_ctx_destroy(self._h)
            self._h = None

    def sync(self):
        _check(self, lib.zcu_ctx_sync(self._h))

    @property
    def stream(self):
        return lib.zcu_ctx_stream(self._h)

    @property
    def launch_count(self):
        return int(lib.zcu_ctx_launch_count(self._h))

    def set_paint_mode(self, mode):
        """L.PAINT_EXACT (default, bit-exact) or L.PAINT_TOLERANCE (Filter cut into verified time segments)."""
        _check(self, lib.zcu_ctx_set_paint_mode(self._h, int(mode)))

    def set_paint_output(self, mode):
        """OUT_ACCUMULATE (the reference's += contract) or OUT_OVERWRITE (`zero; paint` in one call)."""
        _check(self, lib.zcu_ctx_set_paint_output(self._h, int(mode)))

    def paint_output(self):
        return int(lib.zcu_ctx_paint_output(self._h))

    def paint_mode(self):
        return int(lib.zcu_ctx_paint_mode(self._h))

    def noise_resync_count(self):
        n = C.c_uint64()
        _check(self, lib.zcu_ctx_noise_resync_count(self._h, C.byref(n)))
        return int(n.value)

    def timer_start(self):
        _check(self, lib.zcu_timer_start(self._h))

    def timer_stop_ms(self):
        ms = C.c_float()
        _check(self, lib.zcu_timer_stop_ms(self._h, C.byref(ms)))
        return ms.value

    # ---- buffers -------------------------------------------------------------------------------
    def buf(self, frames, voices, data=None):
        b = Buffer(self, frames, voices)
        if data is not None:
            b.upload(data)
        return b

    def wrap(self, device_ptr, frames, voices):
        return Buffer(self, frames, voices, device_ptr=device_ptr)

    def bank(self, kind, voices, first_seed=0):
        return Bank(self, kind, voices, first_seed)

    def spans(self, voices, frames, voice_off, spans):
        return Spans(self, voices, frames, voice_off, spans)

    def spans_full(self, voices, num_buffers, buffer_frames):
        h = C.c_void_p()
        _check(self, lib.zcu_spans_full(self._h, voices, num_buffers, buffer_frames, C.byref(h)))
        s = Spans.__new__(Spans)
        s.ctx, s._h, s.voices, s.frames = self, h, voices, num_buffers * buffer_frames
        return s

    def sample(self, data, num_channels, sample_rate, fmt):
        return Sample(self, data, num_channels, sample_rate, fmt)

    # ---- zang basics (src/zang/basics.zig) --------------------------------------------------------
    def zero(self, spans, dest):
        _check(self, lib.zcu_zero(self._h, _h(spans), dest._h))

    def set(self, spans, dest, a):
        _check(self, lib.zcu_set(self._h, _h(spans), dest._h, float(a)))

    def copy(self, spans, dest, src):
        _check(self, lib.zcu_copy(self._h, _h(spans), dest._h, src._h))

    def add(self, spans, dest, a, b):
        _check(self, lib.zcu_add(self._h, _h(spans), dest._h, a._h, b._h))

    def addInto(self, spans, dest, src):
        _check(self, lib.zcu_add_into(self._h, _h(spans), dest._h, src._h))

    def addScalar(self, spans, dest, a, b):
        _check(self, lib.zcu_add_scalar(self._h, _h(spans), dest._h, a._h, float(b)))

    def addScalarInto(self, spans, dest, a):
        _check(self, lib.zcu_add_scalar_into(self._h, _h(spans), dest._h, float(a)))

    def multiply(self, spans, dest, a, b):
        _check(self, lib.zcu_multiply(self._h, _h(spans), dest._h, a._h, b._h))

    def multiplyWith(self, spans, dest, a):
        _check(self, lib.zcu_multiply_with(self._h, _h(spans), dest._h, a._h))

    def multiplyScalar(self, spans, dest, a, b):
        _check(self, lib.zcu_multiply_scalar(self._h, _h(spans), dest._h, a._h, float(b)))

    def multiplyWithScalar(self, spans, dest, a):
        _check(self, lib.zcu_multiply_with_scalar(self._h, _h(spans), dest._h, float(a)))

    # the same with a per-span scalar (scalars[span.param]), e.g. zang.set(span, t, params.freq * ratio)
    def setRows(self, spans, dest, scalars):
        r = np.ascontiguousarray(scalars, dtype=np.float32).reshape(-1)
        _check(self, lib.zcu_set_rows(self._h, spans._h, dest._h, ptr(r), r.size))

    def addScalarRows(self, spans, dest, a, scalars):
        r = np.ascontiguousarray(scalars, dtype=np.float32).reshape(-1)
        _check(self, lib.zcu_add_scalar_rows(self._h, spans._h, dest._h, a._h, ptr(r), r.size))

    def addScalarIntoRows(self, spans, dest, scalars):
        r = np.ascontiguousarray(scalars, dtype=np.float32).reshape(-1)
        _check(self, lib.zcu_add_scalar_into_rows(self._h, spans._h, dest._h, ptr(r), r.size))

    def multiplyScalarRows(self, spans, dest, a, scalars):
        r = np.ascontiguousarray(scalars, dtype=np.float32).reshape(-1)
        _check(self, lib.zcu_multiply_scalar_rows(self._h, spans._h, dest._h, a._h, ptr(r), r.size))

    def multiplyWithScalarRows(self, spans, dest, scalars):
        r = np.ascontiguousarray(scalars, dtype=np.float32).reshape(-1)
        _check(self, lib.zcu_multiply_with_scalar_rows(self._h, spans._h, dest._h, ptr(r), r.size))

    # ---- mix ------------------------------------------------------------------------------------
    def mix_reduce(self, src, gains, channels, dst):
        g = None if gains is None else np.ascontiguousarray(gains, dtype=np.float32)
        _check(self, lib.zcu_mix_reduce(self._h, src._h, ptr(g), channels, dst._h))

    def mix_ordered(self, src, gains, channels, dst):
        """the reference's sequential voice accumulation order (bit-exact)"""
        g = None if gains is None else np.ascontiguousarray(gains, dtype=np.float32)
        _check(self, lib.zcu_mix_ordered(self._h, src._h, ptr(g), channels, dst._h))

    def mixDown(self, mix, audio_format, vol):
        n = mix.frames * mix.voices * (2 if audio_format == L.AUDIO_SIGNED16_LSB else 1)
        out = np.zeros(n, dtype=np.uint8)
        _check(self, lib.zcu_mix_down(self._h, mix._h, audio_format, float(vol), ptr(out), n))
        return out


def _h(x):
    return None if x is None else x._h


class Buffer:
    def __init__(self, ctx, frames, voices, device_ptr=None):
        self.ctx, self.frames, self.voices = ctx, int(frames), int(voices)
        h = C.c_void_p()
        if device_ptr is None:
            _check(ctx, lib.zcu_buf_alloc(ctx._h, self.frames, self.voices, C.byref(h)))
        else:
            _check(ctx, lib.zcu_buf_wrap(ctx._h, C.c_void_p(device_ptr), self.frames, self.voices, C.byref(h)))
        self._h = h

    def upload(self, data):
        a = np.ascontiguousarray(data, dtype=np.float32)
        assert a.shape == (self.frames, self.voices), (a.shape, self.frames, self.voices)
        _check(self.ctx, lib.zcu_buf_upload(self.ctx._h, self._h, ptr(a)))
        self.ctx.sync()  # `a` may be a temporary

    def download(self):
        out = np.empty((self.frames, self.voices), dtype=np.float32)
        _check(self.ctx, lib.zcu_buf_download(self.ctx._h, self._h, ptr(out)))
        return out

    @property
    def device_ptr(self):
        return lib.zcu_buf_device_ptr(self._h)

    def free(self):
        if self._h:
            lib.zcu_buf_free(self._h)
            self._h = None


class Bank:
    """<Module>.init() for every voice; row v is the `self` of voice v."""

    def __init__(self, ctx, kind, voices, first_seed=0):
        self.ctx, self.kind, self.voices = ctx, kind, int(voices)
        h = C.c_void_p()
        _check(ctx, lib.zcu_bank_create(ctx._h, kind, self.voices, first_seed, C.byref(h)))
        self._h = h

    def get_state(self):
        dt = L.state_dtypes[self.kind]
        if dt is None:
            return None
        out = np.zeros(self.voices, dtype=dt)
        _check(self.ctx, lib.zcu_bank_get_state(self.ctx._h, self._h, ptr(out), out.nbytes))
        return out

    def set_state(self, st):
        st = np.ascontiguousarray(st, dtype=L.state_dtypes[self.kind])
        _check(self.ctx, lib.zcu_bank_set_state(self.ctx._h, self._h, ptr(st), st.nbytes))
        self.ctx.sync()

    def reset(self, first_seed=0):
        _check(self.ctx, lib.zcu_bank_reset(self.ctx._h, self._h, first_seed))

    def free(self):
        if self._h:
            lib.zcu_bank_free(self._h)
            self._h = None


class Spans:
    def __init__(self, ctx, voices, frames, voice_off, spans):
        self.ctx, self.voices, self.frames = ctx, int(voices), int(frames)
        off = np.ascontiguousarray(voice_off, dtype=np.uint32)
        sp = np.ascontiguousarray(spans, dtype=L.span_dtype)
        assert off.shape == (voices + 1,)
        h = C.c_void_p()
        _check(ctx, lib.zcu_spans_upload(ctx._h, self.voices, self.frames, ptr(off), ptr(sp), C.byref(h)))
        ctx.sync()
        self._h = h

    def free(self):
        if self._h:
            lib.zcu_spans_free(self._h)
            self._h = None


class Sample:
    """Sampler.Sample (Sampler.zig:16-21)"""

    def __init__(self, ctx, data, num_channels, sample_rate, fmt):
        self.ctx = ctx
        d = np.ascontiguousarray(np.frombuffer(bytes(data), dtype=np.uint8))
        h = C.c_void_p()
        _check(ctx, lib.zcu_sample_create(ctx._h, ptr(d), d.size, num_channels, sample_rate, fmt, C.byref(h)))
        ctx.sync()
        self._h = h
        self.data, self.num_channels, self.sample_rate, self.format = d, num_channels, sample_rate, fmt

    def free(self):
        if self._h:
            lib.zcu_sample_free(self._h)
            self._h = None


def _params(kind, params):
    p = np.ascontiguousarray(params, dtype=L.param_dtypes[kind])
    return p.reshape(-1)


class _Module:
    """Shared shape of the module contract: num_outputs / num_temps / Params / paint."""
    kind = None
    num_outputs = 1
    num_temps = 0

    def __init__(self, ctx, voices, first_seed=0):
        self.ctx = ctx
        self.voices = voices
        self.Params = L.param_dtypes[self.kind]
        self.bank = Bank(ctx, self.kind, voices, first_seed) if L.state_dtypes[self.kind] is not None else None

    def _bh(self):
        return self.bank._h if self.bank is not None else None


class SineOsc(_Module):
    kind = L.MOD_SINE_OSC

    def paint(self, spans, outputs, temps, params, freq=None, phase=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_sine_osc(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(freq), _h(phase)))


class PulseOsc(_Module):
    kind = L.MOD_PULSE_OSC

    def paint(self, spans, outputs, temps, params, freq=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_pulse_osc(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(freq)))


class TriSawOsc(_Module):
    kind = L.MOD_TRISAW_OSC

    def paint(self, spans, outputs, temps, params, freq=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_trisaw_osc(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(freq)))


class Noise(_Module):
    kind = L.MOD_NOISE

    def paint(self, spans, outputs, temps, params):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_noise(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h))


class Envelope(_Module):
    kind = L.MOD_ENVELOPE

    def paint(self, spans, outputs, temps, params):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_envelope(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h))


class Gate(_Module):
    kind = L.MOD_GATE

    def paint(self, spans, outputs, temps, params):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_gate(self.ctx._h, spans._h, ptr(p), p.size, outputs[0]._h))


class Filter(_Module):
    kind = L.MOD_FILTER

    @staticmethod
    def cutoffFromFrequency(frequency, sample_rate):
        return float(lib.zcu_filter_cutoff_from_frequency(float(frequency), float(sample_rate)))

    def paint(self, spans, outputs, temps, params, input, cutoff=None, res=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_filter(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(input), _h(cutoff), _h(res)))


class Portamento(_Module):
    kind = L.MOD_PORTAMENTO

    def paint(self, spans, outputs, temps, params):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_portamento(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h))


class Curve(_Module):
    kind = L.MOD_CURVE

    def paint(self, spans, outputs, temps, params, nodes):
        p = _params(self.kind, params)
        n = np.ascontiguousarray(nodes, dtype=L.curve_node_dtype).reshape(-1)
        _check(self.ctx, lib.zcu_paint_curve(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, ptr(n), n.size, outputs[0]._h))


class Sampler(_Module):
    kind = L.MOD_SAMPLER

    def paint(self, spans, outputs, temps, params, sample):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_sampler(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, sample._h, outputs[0]._h))


class Decimator(_Module):
    kind = L.MOD_DECIMATOR

    def paint(self, spans, outputs, temps, params, input):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_decimator(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(input)))


class Distortion(_Module):
    kind = L.MOD_DISTORTION

    def paint(self, spans, outputs, temps, params, input):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_distortion(self.ctx._h, spans._h, ptr(p), p.size, outputs[0]._h, _h(input)))


class Cycle(_Module):
    kind = L.MOD_CYCLE

    def paint(self, spans, outputs, temps, params, speed=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_cycle(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(speed)))


class FmOscillator(_Module):
    """examples/example_fmsynth.zig:26-89: Params = {sample_rate, waveform, freq (buffer), phase (buffer or None), feedback}."""
    kind = L.MOD_FM_OSCILLATOR

    def paint(self, spans, outputs, temps, params, freq, phase=None):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_fm_oscillator(self.ctx._h, self._bh(), spans._h, ptr(p), p.size, outputs[0]._h, _h(freq), _h(phase)))


class DelayBank(Bank):
    """zang.Delay(delay_samples).init() for every voice (src/zang/delay.zig:7-17): index POD + a zeroed delay line per voice."""

    def __init__(self, ctx, kind, voices, delay_samples):
        self.ctx, self.kind, self.voices, self.delay_samples = ctx, kind, int(voices), int(delay_samples)
        h = C.c_void_p()
        _check(ctx, lib.zcu_delay_bank_create(ctx._h, kind, self.voices, self.delay_samples, C.byref(h)))
        self._h = h

    def get_delay_line(self):
        out = np.zeros((self.delay_samples, self.voices), dtype=np.float32)
        _check(self.ctx, lib.zcu_bank_get_delay_line(self.ctx._h, self._h, ptr(out), out.nbytes))
        return out

    def set_delay_line(self, line):
        line = np.ascontiguousarray(line, dtype=np.float32)
        _check(self.ctx, lib.zcu_bank_set_delay_line(self.ctx._h, self._h, ptr(line), line.nbytes))


class SimpleDelay:
    """examples/modules.zig:341-388: Params = {input}; outputs[0] += input delayed by delay_samples painted frames."""
    kind = L.MOD_SIMPLE_DELAY
    num_outputs = 1
    num_temps = 0

    def __init__(self, ctx, voices, delay_samples):
        self.ctx, self.voices = ctx, voices
        self.bank = DelayBank(ctx, self.kind, voices, delay_samples)

    def paint(self, spans, outputs, temps, input):
        _check(self.ctx, lib.zcu_paint_simple_delay(self.ctx._h, self.bank._h, spans._h, outputs[0]._h, _h(input)))


class FilteredEchoes:
    """examples/modules.zig:392-460: Params = {input, feedback_volume, cutoff}.  The reference's two temps are not needed."""
    kind = L.MOD_FILTERED_ECHOES
    num_outputs = 1
    num_temps = 0
    Params = L.param_dtypes[L.MOD_FILTERED_ECHOES]

    def __init__(self, ctx, voices, delay_samples):
        self.ctx, self.voices = ctx, voices
        self.bank = DelayBank(ctx, self.kind, voices, delay_samples)

    def paint(self, spans, outputs, temps, params, input):
        p = _params(self.kind, params)
        _check(self.ctx, lib.zcu_paint_filtered_echoes(self.ctx._h, self.bank._h, spans._h, ptr(p), p.size, outputs[0]._h, _h(input)))


class StereoEchoes:
    """examples/modules.zig:462-525 composed from the C-ABI calls, line by line: outputs = [left, right], temps = 2 buffers
    (the reference needs 4: two of them are FilteredEchoes' internal temps, which the device module keeps in registers)."""
    num_outputs = 2
    num_temps = 2
    Params = L.param_dtypes[L.MOD_FILTERED_ECHOES]

    def __init__(self, ctx, voices, main_delay):
        self.ctx = ctx
        half = int(main_delay) // 2  # :463
        self.delay0 = SimpleDelay(ctx, voices, half)
        self.delay1 = SimpleDelay(ctx, voices, half)
        self.echoes = FilteredEchoes(ctx, voices, int(main_delay))

    def paint(self, spans, outputs, temps, params, input):
        ctx = self.ctx
        ctx.addInto(spans, outputs[0], input)                                   # :497
        ctx.addInto(spans, outputs[1], input)                                   # :498
        ctx.zero(spans, temps[0])                                               # :501
        self.delay0.paint(spans, [temps[0]], [], input)                         # :502-504
        ctx.zero(spans, temps[1])                                               # :506
        self.echoes.paint(spans, [temps[1]], [], params, input=temps[0])        # :507-511
        ctx.addInto(spans, outputs[0], temps[1])                                # :513
        self.delay1.paint(spans, [outputs[1]], [], temps[1])                    # :514-516


MODULE_CLASSES = {c.kind: c for c in (SineOsc, PulseOsc, TriSawOsc, Noise, Envelope, Gate, Filter, Portamento, Curve, Sampler,
                                      Decimator, Distortion, Cycle)}


class NiceInstrumentBank:
    """examples/modules.zig:189-248 composed from the module paints (the un-fused, drop-in path).

    instrument params rows: (sample_rate, freq, note_on); `color` is per bank as in NiceInstrument.init.
    """
    num_outputs = 1
    num_temps = 2
    Params = np.dtype([("sample_rate", "<f4"), ("freq", "<f4"), ("note_on", "<i4")])

    def __init__(self, ctx, voices, color):
        self.ctx, self.color = ctx, float(color)
        self.osc = PulseOsc(ctx, voices)
        self.flt = Filter(ctx, voices)
        self.env = Envelope(ctx, voices)
        self._derived = None

    def derive(self, params):
        """Module params rows of every instrument params row (the per-call scalars of modules.zig:221-245).
        Cached for the params array last seen: a song passes the same table block after block."""
        if self._derived is not None and self._derived[0] is params:
            return self._derived[1]
        src = params
        params = np.ascontiguousarray(params, dtype=self.Params).reshape(-1)
        n = params.size
        pp = np.zeros(n, dtype=L.param_dtypes[L.MOD_PULSE_OSC])
        pp["sample_rate"], pp["freq"], pp["color"] = params["sample_rate"], params["freq"], self.color
        fp = np.zeros(n, dtype=L.param_dtypes[L.MOD_FILTER])
        fp["type"] = L.FILTER_LOW_PASS
        fp["cutoff"] = [Filter.cutoffFromFrequency(np.float32(f) * np.float32(8.0), sr) for f, sr in zip(params["freq"], params["sample_rate"])]
        fp["res"] = 0.7
        ep = np.zeros(n, dtype=L.param_dtypes[L.MOD_ENVELOPE])
        ep["sample_rate"] = params["sample_rate"]
        ep["attack"] = (L.CURVE_CUBED, 0.01)
        ep["decay"] = (L.CURVE_CUBED, 0.1)
        ep["release"] = (L.CURVE_CUBED, 0.5)
        ep["sustain_volume"] = 0.8
        ep["note_on"] = params["note_on"]
        self._derived = (src, (pp, fp, ep))
        return pp, fp, ep

    def paint(self, spans, outputs, temps, params, overwrite=False):
        """overwrite: every `zero; paint` pair of the body as ONE call in ZCU_OUT_OVERWRITE mode (same bits, 5 calls
        instead of 8, no zero passes and no reads of the temps by the generators)."""
        ctx = self.ctx
        pp, fp, ep = self.derive(params)
        if overwrite:
            ctx.set_paint_output(L.OUT_OVERWRITE)
            try:
                self.osc.paint(spans, [temps[0]], [], pp)
                ctx.multiplyWithScalar(spans, temps[0], 0.5)
                self.flt.paint(spans, [temps[1]], [], fp, input=temps[0])
                self.env.paint(spans, [temps[0]], [], ep)
            finally:
                ctx.set_paint_output(L.OUT_ACCUMULATE)
            ctx.multiply(spans, outputs[0], temps[0], temps[1])
            return
        ctx.zero(spans, temps[0])
        self.osc.paint(spans, [temps[0]], [], pp)
        ctx.multiplyWithScalar(spans, temps[0], 0.5)
        ctx.zero(spans, temps[1])
        self.flt.paint(spans, [temps[1]], [], fp, input=temps[0])
        ctx.zero(spans, temps[0])
        self.env.paint(spans, [temps[0]], [], ep)
        ctx.multiply(spans, outputs[0], temps[0], temps[1])


class PhaseModOscillatorBank:
    """examples/modules.zig:6-77 PhaseModOscillator with constant ratio and multiplier (the variants the
    example instruments use), composed from the module paints.  params rows: (sample_rate, freq)."""
    num_outputs = 1
    num_temps = 2

    def __init__(self, ctx, voices, relative=True, ratio=1.0, multiplier=1.0):
        self.ctx, self.relative, self.ratio, self.multiplier = ctx, bool(relative), np.float32(ratio), np.float32(multiplier)
        self.carrier = SineOsc(ctx, voices)
        self.modulator = SineOsc(ctx, voices)

    def derive(self, sample_rate, freq):
        n = freq.size
        mod_freq = (freq.astype(np.float32) * self.ratio) if self.relative else np.full(n, self.ratio, dtype=np.float32)  # :41-47
        mp = np.zeros(n, dtype=L.param_dtypes[L.MOD_SINE_OSC])
        mp["sample_rate"], mp["phase"] = sample_rate, 0.0  # freq comes from the buffer (:56-60)
        cp = np.zeros(n, dtype=L.param_dtypes[L.MOD_SINE_OSC])
        cp["sample_rate"], cp["freq"] = sample_rate, freq  # phase comes from the buffer (:69-73)
        return mod_freq, mp, cp

    def paint(self, spans, outputs, temps, derived):
        ctx = self.ctx
        mod_freq, mp, cp = derived
        ctx.setRows(spans, temps[0], mod_freq)
        ctx.zero(spans, temps[1])
        self.modulator.paint(spans, [temps[1]], [], mp, freq=temps[0])
        ctx.zero(spans, temps[0])
        ctx.multiplyScalar(spans, temps[0], temps[1], float(self.multiplier))
        ctx.zero(spans, temps[1])
        self.carrier.paint(spans, [temps[1]], [], cp, phase=temps[0])
        ctx.addInto(spans, outputs[0], temps[1])


class PMOscInstrumentBank:
    """examples/modules.zig:80-128 PMOscInstrument composed from the module paints.
    instrument params rows: (sample_rate, freq, note_on); release_duration is per bank as in init()."""
    num_outputs = 1
    num_temps = 3
    Params = NiceInstrumentBank.Params

    def __init__(self, ctx, voices, release_duration):
        self.ctx, self.release_duration = ctx, float(release_duration)
        self.osc = PhaseModOscillatorBank(ctx, voices, relative=True, ratio=1.0, multiplier=1.0)
        self.env = Envelope(ctx, voices)
        self._derived = None

    def derive(self, params):
        if self._derived is not None and self._derived[0] is params:
            return self._derived[1]
        src = params
        params = np.ascontiguousarray(params, dtype=self.Params).reshape(-1)
        n = params.size
        od = self.osc.derive(params["sample_rate"], params["freq"])
        ep = np.zeros(n, dtype=L.param_dtypes[L.MOD_ENVELOPE])
        ep["sample_rate"] = params["sample_rate"]
        ep["attack"] = (L.CURVE_CUBED, 0.025)
        ep["decay"] = (L.CURVE_CUBED, 0.1)
        ep["release"] = (L.CURVE_CUBED, self.release_duration)
        ep["sustain_volume"] = 0.5
        ep["note_on"] = params["note_on"]
        self._derived = (src, (od, ep))
        return od, ep

    def paint(self, spans, outputs, temps, params):
        ctx = self.ctx
        od, ep = self.derive(params)
        ctx.zero(spans, temps[0])
        self.osc.paint(spans, [temps[0]], [temps[1], temps[2]], od)
        ctx.zero(spans, temps[1])
        self.env.paint(spans, [temps[1]], [], ep)
        ctx.multiply(spans, outputs[0], temps[0], temps[1])


class Render:
    """Fused instrument bank (zcu_render_*): whole-job render of V voices into a [frames][channels] mix."""

    def __init__(self, ctx, voices, total_frames, color, gains, channels=2, sample_rate=48000.0, buffer_frames=1024, block_buffers=0,
                 instrument=L.INSTR_NICE, mode=L.RENDER_EXACT):
        self.ctx = ctx
        cfg = L.RenderConfig(instrument, voices, buffer_frames, total_frames, block_buffers, channels, sample_rate, 0, mode)
        self.cfg = cfg
        color = np.ascontiguousarray(color, dtype=np.float32)
        gains = np.ascontiguousarray(gains, dtype=np.float32)
        assert color.shape == (voices,) and gains.shape == (voices, channels)
        h = C.c_void_p()
        _check(ctx, lib.zcu_render_create(ctx._h, C.byref(cfg), ptr(color), ptr(gains), C.byref(h)))
        ctx.sync()
        self._h = h
        self.voices, self.total_frames, self.channels = voices, total_frames, channels

    def set_events(self, event_off, events, sync=True):
        off = np.ascontiguousarray(event_off, dtype=np.uint32)
        ev = np.ascontiguousarray(events, dtype=L.note_event_dtype)
        _check(self.ctx, lib.zcu_render_set_events(self.ctx._h, self._h, ptr(off), ptr(ev)))
        if sync:  # the host arrays may be temporaries: the copies are stream ordered, wait for them
            self.ctx.sync()

    def run(self, mix):
        _check(self.ctx, lib.zcu_render_run(self.ctx._h, self._h, mix._h))

    def run_voices(self, first_frame, out_voices):
        _check(self.ctx, lib.zcu_render_run_voices(self.ctx._h, self._h, first_frame, out_voices._h))

    def run_voice_window(self, first_voice, first_frame, out_voices):
        """Un-mixed output of voices [first_voice, first_voice + out.voices) over [first_frame, first_frame + out.frames),
        rendered with the launches and time segments of the whole job."""
        _check(self.ctx, lib.zcu_render_run_voice_window(self.ctx._h, self._h, int(first_voice), int(first_frame), out_voices._h))

    def set_profiling(self, enabled=True):
        _check(self.ctx, lib.zcu_render_set_profiling(self._h, 1 if enabled else 0))

    def kernel_ms(self):
        """(summed render-kernel milliseconds, launches) since the last call; synchronises."""
        ms, n = C.c_float(), C.c_uint32()
        _check(self.ctx, lib.zcu_render_kernel_ms(self.ctx._h, self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def aux_kernel_ms(self):
        """ZCU_RENDER_SCAN: (summed milliseconds, launches) of the zero-state filter pass since the last call."""
        ms, n = C.c_float(), C.c_uint32()
        _check(self.ctx, lib.zcu_render_aux_kernel_ms(self.ctx._h, self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def set_segments(self, segments_per_launch):
        """ZCU_RENDER_SEGMENTS: force the number of time segments per launch (0 = sized to the GPU)."""
        rc = lib.zcu_render_set_segments(self._h, int(segments_per_launch))
        if rc != L.OK:
            raise ZangCuError("zcu_render_set_segments: not a ZCU_RENDER_SEGMENTS render")

    def set_filter_contraction(self, enabled):
        """ZCU_RENDER_SEGMENTS, opt-in: the Filter recurrence with FMAs (faster; Filter output inside the mode's 1e-5, no longer
        the reference's bits)."""
        if lib.zcu_render_set_filter_contraction(self._h, 1 if enabled else 0) != L.OK:
            raise ZangCuError("zcu_render_set_filter_contraction: not a ZCU_RENDER_SEGMENTS render")

    def segment_info(self):
        """ZCU_RENDER_SEGMENTS: dict(segments, buffers_per_segment, warm_sum, warm_max) of the last run."""
        v = [C.c_uint32() for _ in range(4)]
        _check(self.ctx, lib.zcu_render_segment_info(self.ctx._h, self._h, *[C.byref(x) for x in v]))
        return dict(zip(("segments", "buffers_per_segment", "warm_sum", "warm_max"), (x.value for x in v)))

    @property
    def launches_per_run(self):
        return int(lib.zcu_render_launches_per_run(self._h))

    def free(self):
        if self._h:
            lib.zcu_render_free(self._h)
            self._h = None


# ---- fused module graphs (zcu_graph_*) -------------------------------------------------------------------------------
def T(i):
    """temp i of a graph"""
    return (L.GREF_TEMP, int(i))


def OUT(i=0):
    """output i of a graph"""
    return (L.GREF_OUTPUT, int(i))


def IN(i):
    """external input buffer i of a graph"""
    return (L.GREF_INPUT, int(i))


class GraphBuilder:
    """Instruction list of a fused graph: the body of an instrument's paint() (examples/modules.zig) or a script module's
    instructions (src/zangscript/codegen.zig:108-121), one method per instruction."""

    def __init__(self):
        self.instrs = []
        self.field_kinds = []

    def _add(self, op, sub, out, ins=(), scalar=0.0, scalar_side=L.SCALAR_NONE, field=0, table=0):
        x = L.GraphInstr()
        x.op, x.sub = op, sub
        x.out = L.GraphRef(*out)
        for q in range(3):
            r = ins[q] if q < len(ins) and ins[q] is not None else (L.GREF_NONE, 0)
            x.in_[q] = L.GraphRef(*r)
        x.scalar, x.scalar_side, x.field, x.span_table = float(scalar), scalar_side, field, table
        self.instrs.append(x)
        return self

    # zang basics (basics.zig:12-78)
    def zero(self, d, table=0): return self._add(L.GI_BASIC, L.BASIC_ZERO, d, table=table)
    def set(self, d, s, table=0): return self._add(L.GI_BASIC, L.BASIC_SET, d, scalar=s, table=table)
    def copy(self, d, a, table=0): return self._add(L.GI_BASIC, L.BASIC_COPY, d, (a,), table=table)
    def add(self, d, a, b, table=0): return self._add(L.GI_BASIC, L.BASIC_ADD, d, (a, b), table=table)
    def addInto(self, d, a, table=0): return self._add(L.GI_BASIC, L.BASIC_ADD_INTO, d, (a,), table=table)
    def addScalar(self, d, a, s, table=0): return self._add(L.GI_BASIC, L.BASIC_ADD_SCALAR, d, (a,), scalar=s, table=table)
    def addScalarInto(self, d, s, table=0): return self._add(L.GI_BASIC, L.BASIC_ADD_SCALAR_INTO, d, scalar=s, table=table)
    def multiply(self, d, a, b, table=0): return self._add(L.GI_BASIC, L.BASIC_MULTIPLY, d, (a, b), table=table)
    def multiplyWith(self, d, a, table=0): return self._add(L.GI_BASIC, L.BASIC_MULTIPLY_WITH, d, (a,), table=table)
    def multiplyScalar(self, d, a, s, table=0): return self._add(L.GI_BASIC, L.BASIC_MULTIPLY_SCALAR, d, (a,), scalar=s, table=table)
    def multiplyWithScalar(self, d, s, table=0): return self._add(L.GI_BASIC, L.BASIC_MULTIPLY_WITH_SCALAR, d, scalar=s, table=table)

    # the script runtime's overwriting arithmetic (runtime.zig:528-700); a or b may be a Python float
    def arith(self, sub, d, a, b=None, table=0):
        if isinstance(a, (int, float)):
            return self._add(L.GI_ARITH, sub, d, (None, b), scalar=a, scalar_side=L.SCALAR_A, table=table)
        if isinstance(b, (int, float)):
            return self._add(L.GI_ARITH, sub, d, (a, None), scalar=b, scalar_side=L.SCALAR_B, table=table)
        return self._add(L.GI_ARITH, sub, d, (a, b), table=table)

    def call(self, kind, d, ins=(), table=0):
        """module.paint(span, {d}, ...): accumulates into d.  Returns the new field's index."""
        field = len(self.field_kinds)
        self.field_kinds.append(int(kind))
        self._add(L.GI_CALL, int(kind), d, ins, field=field, table=table)
        return field

    def delay_field(self):
        """a zang.Delay(N) of the graph (bank: DelayBank of kind MOD_SIMPLE_DELAY); returns the field index"""
        self.field_kinds.append(int(L.MOD_SIMPLE_DELAY))
        return len(self.field_kinds) - 1

    def delay_read(self, field, d, table=0):
        """zero(d); delay.readDelayBuffer(d)  (runtime.zig:711-713)"""
        return self._add(L.GI_DELAY_READ, 0, d, field=field, table=table)

    def delay_write(self, field, a, table=0):
        """delay.writeDelayBuffer(a)  (runtime.zig:718)"""
        return self._add(L.GI_DELAY_WRITE, 0, a, (a,), field=field, table=table)

    def arrays(self):
        arr = (L.GraphInstr * len(self.instrs))(*self.instrs)
        kinds = (C.c_int32 * max(1, len(self.field_kinds)))(*self.field_kinds)
        return arr, kinds


def graph_config(voices, builder, num_temps, num_outputs=1, num_inputs=0, num_span_tables=1, output_mode=L.GRAPH_OUT_BUFFERS, channels=2):
    return L.GraphConfig(int(voices), int(num_temps), int(num_outputs), int(num_inputs), int(num_span_tables), len(builder.field_kinds),
                         int(output_mode), int(channels))


def graph_compile_check(cfg, builder):
    """Host-only: generate + NVRTC-compile for sm_100a.  Returns (rc, log, cubin_bytes)."""
    arr, kinds = builder.arrays()
    log = C.create_string_buffer(1 << 16)
    n = C.c_size_t()
    rc = lib.zcu_graph_compile_check(C.byref(cfg), arr, len(builder.instrs), kinds, log, len(log), C.byref(n))
    return rc, log.value.decode(errors="replace"), n.value


class Graph:
    """One fused kernel for a whole instruction list (zcu_graph_create / zcu_graph_paint)."""

    def __init__(self, ctx, cfg, builder):
        self.ctx, self.cfg = ctx, cfg
        arr, kinds = builder.arrays()
        h = C.c_void_p()
        _check(ctx, lib.zcu_graph_create(ctx._h, C.byref(cfg), arr, len(builder.instrs), kinds, C.byref(h)))
        self._h = h
        self.field_kinds = list(builder.field_kinds)
        self._keep = []

    @property
    def source(self):
        return lib.zcu_graph_source(self._h).decode()

    def set_field(self, field, bank, params):
        kind = self.field_kinds[field]
        if params is None:
            _check(self.ctx, lib.zcu_graph_set_field(self.ctx._h, self._h, field, _h(bank), None, 0, 0))
            return
        p = _params(kind, params)
        _check(self.ctx, lib.zcu_graph_set_field(self.ctx._h, self._h, field, _h(bank), ptr(p), p.dtype.itemsize, p.size))

    def set_curve_nodes(self, field, nodes):
        n = np.ascontiguousarray(nodes, dtype=L.curve_node_dtype).reshape(-1)
        _check(self.ctx, lib.zcu_graph_set_curve_nodes(self.ctx._h, self._h, field, ptr(n), n.size))

    def set_sample(self, field, sample):
        self._keep.append(sample)
        _check(self.ctx, lib.zcu_graph_set_sample(self.ctx._h, self._h, field, sample._h))

    def set_gains(self, gains):
        g = np.ascontiguousarray(gains, dtype=np.float32)
        assert g.shape == (self.cfg.voices, self.cfg.channels)
        _check(self.ctx, lib.zcu_graph_set_gains(self.ctx._h, self._h, ptr(g)))

    def paint(self, tables, outputs, inputs=()):
        tabs = (C.c_void_p * len(tables))(*[t._h for t in tables])
        outs = (C.c_void_p * len(outputs))(*[o._h for o in outputs])
        ins = (C.c_void_p * max(1, len(inputs)))(*[i._h for i in inputs])
        _check(self.ctx, lib.zcu_graph_paint(self.ctx._h, self._h, tabs, outs, ins if len(inputs) else None))

    def set_segments(self, n):
        """time segments of the next paints: 0 = sized to the GPU (default), 1 = off"""
        assert lib.zcu_graph_set_segments(self._h, int(n)) == L.OK

    def segment_info(self):
        """dict(segments, segment_frames, replay_frames, failed) of the last paint; synchronises"""
        v = [C.c_uint32() for _ in range(4)]
        _check(self.ctx, lib.zcu_graph_segment_info(self.ctx._h, self._h, *[C.byref(x) for x in v]))
        return dict(zip(("segments", "segment_frames", "replay_frames", "failed"), (x.value for x in v)))

    def warm_retries(self):
        """(warp, segment) pairs whose Filter warm-up window had to be lengthened since the last call; synchronises"""
        n = C.c_uint32()
        _check(self.ctx, lib.zcu_graph_warm_retries(self.ctx._h, self._h, C.byref(n)))
        return int(n.value)

    def free(self):
        if self._h:
            lib.zcu_graph_free(self._h)
            self._h = None
