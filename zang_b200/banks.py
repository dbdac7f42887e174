"""Instrument graphs of BASELINE.json configs[2..4] composed from the module paints (the drop-in, un-fused path):
every line of a reference-style `paint()` body becomes one C-ABI call over all voices of the block.

The graphs are synthetic composites of the reference's examples; tests/test_gpu_banks.py compares them
with a per-voice CPU restatement of the same graphs.

  FmDetunedBank   configs[2]: Portamento -> detuned TriSawOsc + phase-modulated SineOsc pair -> x Envelope
                  (examples/modules.zig:42-75, example_detuned.zig:61-99, example_portamento.zig:53-85)
  SamplerVoiceBank configs[3]: Sampler (Trigger spans) x2.5 -> Decimator -> Distortion overdrive
                  (example_sampler.zig:94-118, example_polyphony.zig:147-160)
  SfxBank         configs[4]: Noise -> Filter with a Curve-driven cutoff -> x Envelope
                  (example_laser.zig:77-115, example_stereo.zig:71-98)
"""
import numpy as np

from . import _lib as L
from . import engine as E
from .events import BankTrigger, DeviceBankTrigger


def full_span_table(voices, num_frames, buffer_frames, rows):
    """One span per voice per mix buffer covering the whole buffer (the un-triggered call pattern,
    example_polyphony.zig:147-160), span.param = rows[v]."""
    starts = np.arange(0, num_frames, buffer_frames, dtype=np.uint32)
    ends = np.minimum(starts + buffer_frames, num_frames).astype(np.uint32)
    nb = starts.size
    spans = np.zeros(voices * nb, dtype=L.span_dtype)
    spans["start"] = np.tile(starts, voices)
    spans["end"] = np.tile(ends, voices)
    spans["param"] = np.repeat(np.asarray(rows, dtype=np.uint32), nb)
    off = (np.arange(voices + 1, dtype=np.uint32) * nb).astype(np.uint32)
    return off, spans


class _Kept:
    """a span table owned by its bank: callers that free their prepared tables leave it alone"""

    def __init__(self, spans):
        self.spans, self._h, self.voices, self.frames = spans, spans._h, spans.voices, spans.frames

    def free(self):
        pass


class _DeviceSpansMixin:
    """use_device_spans(): the bank's Trigger is replayed on the GPU from device-resident events (zcu_event_bank_*), so a
    block's span tables cost the host nothing; the default (host replay in libzangcu + upload) stays available."""
    dev_trigger = None

    def use_device_spans(self, wl):
        self._full_tables = {}
        self.dev_trigger = DeviceBankTrigger(self.ctx, wl["event_off"], wl["events"], self.total_frames, self.buffer_frames)
        return self


class SamplerVoiceBank(_DeviceSpansMixin):
    """paint body per voice (per mix buffer):
        zero(span, t0); while (trigger.next) sampler.paint(result.span, {t0}, ..., result.params)
        multiplyWithScalar(span, t0, 2.5)                                   (example_sampler.zig:94-108)
        zero(span, t1); decimator.paint(span, {t1}, .{input = t0, fake_sample_rate})   (example_polyphony.zig:147-160)
        distortion.paint(span, {out}, .{input = t1, overdrive, 0.9, 0.5, 0.0})         (example_sampler.zig:109-116)
    """
    num_outputs, num_temps = 1, 2

    def __init__(self, ctx, wl, sample, sample_rate, total_frames, buffer_frames=1024):
        self.ctx, self.sample = ctx, sample
        self.voices = wl["fake_sample_rate"].size
        self.sample_rate, self.total_frames, self.buffer_frames = float(sample_rate), total_frames, buffer_frames
        self.trigger = BankTrigger(wl["event_off"], wl["events"], total_frames, buffer_frames)
        self.sampler = E.Sampler(ctx, self.voices)
        self.decimator = E.Decimator(ctx, self.voices)
        self.distortion = E.Distortion(ctx, self.voices)
        ev = wl["events"]
        self.sampler_rows = np.zeros(max(ev.size, 1), dtype=L.param_dtypes[L.MOD_SAMPLER])
        self.sampler_rows["sample_rate"][:ev.size] = ev["freq"]
        self.sampler_rows["channel"] = 0
        self.sampler_rows["loop"] = 1
        self.dec_rows = np.zeros(self.voices, dtype=L.param_dtypes[L.MOD_DECIMATOR])
        self.dec_rows["sample_rate"] = sample_rate
        self.dec_rows["fake_sample_rate"] = wl["fake_sample_rate"]
        self.dist_rows = np.zeros(1, dtype=L.param_dtypes[L.MOD_DISTORTION])
        self.dist_rows[0] = (L.DIST_OVERDRIVE, 0.9, 0.5, 0.0)

    def prepare_block(self, first_frame, num_frames):
        """Host side of a block: Trigger replay and upload of the span tables (the reference's per-buffer event logic)."""
        ctx, V = self.ctx, self.voices
        if self.dev_trigger is not None:
            trig = self.dev_trigger.spans(first_frame, num_frames)
        else:
            trig = ctx.spans(V, num_frames, *self.trigger.spans(first_frame, num_frames))
        if self.dev_trigger is None:
            return (trig,
                    ctx.spans(V, num_frames, *full_span_table(V, num_frames, self.buffer_frames, np.arange(V))),
                    ctx.spans(V, num_frames, *full_span_table(V, num_frames, self.buffer_frames, np.zeros(V))))
        # the whole-buffer tables depend on the block length alone: built once, kept on the device
        if num_frames not in self._full_tables:
            self._full_tables[num_frames] = tuple(_Kept(ctx.spans(V, num_frames, *full_span_table(V, num_frames, self.buffer_frames, rows)))
                                                  for rows in (np.arange(V), np.zeros(V)))
        return (trig,) + self._full_tables[num_frames]

    def paint_block(self, first_frame, num_frames, out, temps, prepared=None):
        ctx = self.ctx
        t0, t1 = temps
        sp_trig, sp_voice, sp_row0 = prepared or self.prepare_block(first_frame, num_frames)
        ctx.zero(None, t0)
        self.sampler.paint(sp_trig, [t0], [], self.sampler_rows, self.sample)
        ctx.multiplyWithScalar(None, t0, 2.5)
        ctx.zero(None, t1)
        self.decimator.paint(sp_voice, [t1], [], self.dec_rows, input=t0)
        self.distortion.paint(sp_row0, [out], [], self.dist_rows, input=t1)
        if prepared is None:
            for s in (sp_trig, sp_voice, sp_row0):
                s.free()


    # ---- the same body as ONE fused kernel (zcu_graph_*): temps in registers ----
    @staticmethod
    def graph_list(voices, output_mode=L.GRAPH_OUT_BUFFERS, channels=2):
        """(builder, config, field indices) of the body above; host only."""
        b = E.GraphBuilder()
        T, OUT = E.T, E.OUT
        TRIG, VOICE, ROW0 = 0, 1, 2  # span tables: the Trigger's spans, whole buffers with row = voice, whole buffers with row 0
        b.zero(T(0), table=VOICE)
        f_smp = b.call(L.MOD_SAMPLER, T(0), table=TRIG)
        b.multiplyWithScalar(T(0), 2.5, table=VOICE)
        b.zero(T(1), table=VOICE)
        f_dec = b.call(L.MOD_DECIMATOR, T(1), (T(0),), table=VOICE)
        f_dist = b.call(L.MOD_DISTORTION, OUT(0), (T(1),), table=ROW0)
        cfg = E.graph_config(voices, b, 2, num_span_tables=3, output_mode=output_mode, channels=channels)
        return b, cfg, (f_smp, f_dec, f_dist)

    def build_graph(self, output_mode=L.GRAPH_OUT_BUFFERS, channels=2, gains=None):
        b, cfg, (f_smp, f_dec, f_dist) = self.graph_list(self.voices, output_mode, channels)
        g = E.Graph(self.ctx, cfg, b)
        g.set_field(f_smp, self.sampler.bank, self.sampler_rows)
        g.set_sample(f_smp, self.sample)
        g.set_field(f_dec, self.decimator.bank, self.dec_rows)
        g.set_field(f_dist, None, self.dist_rows)
        if gains is not None:
            g.set_gains(gains)
        self.graph = g
        return g

    def paint_block_fused(self, first_frame, num_frames, out, prepared=None):
        tabs = prepared or self.prepare_block(first_frame, num_frames)
        self.graph.paint(tabs, [out])
        if prepared is None:
            for s in tabs:
                s.free()


class SfxBank(_DeviceSpansMixin):
    """paint body per job (per Trigger span):
        zero(t0); noise.paint(span, {t0}, .{ .color = .white })
        zero(t1); curve.paint(span, {t1}, note_id_changed, .{ .function = .smoothstep, .curve = job's 3 nodes })
        zero(t2); filter.paint(span, {t2}, .{ .input = t0, .type = job's, .cutoff = buffer(t1), .res = constant(0.4) })
        zero(t0); env.paint(span, {t0}, note_id_changed, .{ linear 0.005, squared 0.05, cubed 0.2, sustain 0.6, note_on })
        multiply(span, out, t2, t0)
    """
    num_outputs, num_temps = 1, 3

    def __init__(self, ctx, wl, sample_rate, total_frames, buffer_frames=1024, first_seed=0):
        self.ctx = ctx
        self.voices = V = wl["filter_type"].size
        self.sample_rate, self.total_frames, self.buffer_frames = float(sample_rate), total_frames, buffer_frames
        self.trigger = BankTrigger(wl["event_off"], wl["events"], total_frames, buffer_frames)
        self.noise = E.Noise(ctx, V, first_seed=first_seed)
        self.curve = E.Curve(ctx, V)
        self.filter = E.Filter(ctx, V)
        self.env = E.Envelope(ctx, V)
        ev, eo = wl["events"], wl["event_off"]
        n = max(ev.size, 1)
        voice_of_event = np.repeat(np.arange(V, dtype=np.uint32), np.diff(eo).astype(np.int64))
        self.nodes = np.ascontiguousarray(wl["nodes"], dtype=L.curve_node_dtype)
        self.noise_rows = np.zeros(n, dtype=L.param_dtypes[L.MOD_NOISE])
        self.noise_rows["color"] = L.NOISE_WHITE
        self.curve_rows = np.zeros(n, dtype=L.param_dtypes[L.MOD_CURVE])
        self.curve_rows["sample_rate"] = sample_rate
        self.curve_rows["function"] = L.INTERP_SMOOTHSTEP
        self.curve_rows["node_start"][:ev.size] = 3 * voice_of_event
        self.curve_rows["node_count"] = 3
        self.filter_rows = np.zeros(n, dtype=L.param_dtypes[L.MOD_FILTER])
        self.filter_rows["type"][:ev.size] = wl["filter_type"][voice_of_event]
        self.filter_rows["res"] = 0.4
        self.env_rows = np.zeros(n, dtype=L.param_dtypes[L.MOD_ENVELOPE])
        self.env_rows["sample_rate"] = sample_rate
        self.env_rows["attack"] = (L.CURVE_LINEAR, 0.005)
        self.env_rows["decay"] = (L.CURVE_SQUARED, 0.05)
        self.env_rows["release"] = (L.CURVE_CUBED, 0.2)
        self.env_rows["sustain_volume"] = 0.6
        self.env_rows["note_on"][:ev.size] = ev["note_on"]

    def prepare_block(self, first_frame, num_frames):
        if self.dev_trigger is not None:
            return (self.dev_trigger.spans(first_frame, num_frames),)
        off, spans = self.trigger.spans(first_frame, num_frames)
        return (self.ctx.spans(self.voices, num_frames, off, spans),)

    def paint_block(self, first_frame, num_frames, out, temps, prepared=None):
        ctx = self.ctx
        t0, t1, t2 = temps
        (sp,) = prepared or self.prepare_block(first_frame, num_frames)
        ctx.zero(sp, t0)
        self.noise.paint(sp, [t0], [], self.noise_rows)
        ctx.zero(sp, t1)
        self.curve.paint(sp, [t1], [], self.curve_rows, self.nodes)
        ctx.zero(sp, t2)
        self.filter.paint(sp, [t2], [], self.filter_rows, input=t0, cutoff=t1)
        ctx.zero(sp, t0)
        self.env.paint(sp, [t0], [], self.env_rows)
        ctx.multiply(sp, out, t2, t0)
        if prepared is None:
            sp.free()


    # ---- the same body as ONE fused kernel (zcu_graph_*): temps in registers ----
    @staticmethod
    def graph_list(voices, output_mode=L.GRAPH_OUT_BUFFERS, channels=2):
        """(builder, config, field indices) of the body above; host only."""
        b = E.GraphBuilder()
        T, OUT = E.T, E.OUT
        b.zero(T(0))
        f_noise = b.call(L.MOD_NOISE, T(0))
        b.zero(T(1))
        f_curve = b.call(L.MOD_CURVE, T(1))
        b.zero(T(2))
        f_filter = b.call(L.MOD_FILTER, T(2), (T(0), T(1)))  # input, cutoff buffer; res from the params row
        b.zero(T(0))
        f_env = b.call(L.MOD_ENVELOPE, T(0))
        b.multiply(OUT(0), T(2), T(0))
        cfg = E.graph_config(voices, b, 3, output_mode=output_mode, channels=channels)
        return b, cfg, (f_noise, f_curve, f_filter, f_env)

    def build_graph(self, output_mode=L.GRAPH_OUT_BUFFERS, channels=2, gains=None):
        b, cfg, (f_noise, f_curve, f_filter, f_env) = self.graph_list(self.voices, output_mode, channels)
        g = E.Graph(self.ctx, cfg, b)
        g.set_field(f_noise, self.noise.bank, self.noise_rows)
        g.set_field(f_curve, self.curve.bank, self.curve_rows)
        g.set_curve_nodes(f_curve, self.nodes)
        g.set_field(f_filter, self.filter.bank, self.filter_rows)
        g.set_field(f_env, self.env.bank, self.env_rows)
        if gains is not None:
            g.set_gains(gains)
        self.graph = g
        return g

    def paint_block_fused(self, first_frame, num_frames, out, prepared=None):
        tabs = prepared or self.prepare_block(first_frame, num_frames)
        self.graph.paint(tabs, [out])
        if prepared is None:
            for s in tabs:
                s.free()


class FmDetunedBank(_DeviceSpansMixin):
    """paint body per voice (per Trigger span), params = {sample_rate, freq, note_on}, two per-sample inputs:
        defer self.prev_note_on = params.note_on                                        (example_portamento.zig:51)
        zero(t0); porta.paint(span, {t0}, note_id_changed, .{ .curve = .{ .cubed = 0.5 }, .goal = freq, note_on, prev_note_on })
        zero(t1); multiply(span, t1, t0, detune_in)                                      (example_detuned.zig: freq x warble)
        zero(t2); trisaw.paint(span, {t2}, .{ .freq = buffer(t1), .color = 0.0 })
        zero(t3); modulator.paint(span, {t3}, .{ .freq = buffer(t0), .phase = constant(0) })   (modules.zig:42-60)
        multiplyWith(span, t3, index_in)
        zero(t4); carrier.paint(span, {t4}, .{ .freq = buffer(t0), .phase = buffer(t3) })      (modules.zig:62-75)
        zero(t5); env.paint(span, {t5}, !prev_note_on and note_on, .{ cubed 0.025, cubed 0.1, cubed 1.0, sustain 0.5, note_on })
        multiply(span, out, t2, t5); multiply(span, out, t4, t5)
    """
    num_outputs, num_temps = 1, 6

    def __init__(self, ctx, wl, sample_rate, total_frames, buffer_frames=1024):
        self.ctx = ctx
        self.voices = V = wl["event_off"].size - 1
        self.sample_rate, self.total_frames, self.buffer_frames = float(sample_rate), total_frames, buffer_frames
        self.trigger = BankTrigger(wl["event_off"], wl["events"], total_frames, buffer_frames)
        self.porta = E.Portamento(ctx, V)
        self.trisaw = E.TriSawOsc(ctx, V)
        self.modulator = E.SineOsc(ctx, V)
        self.carrier = E.SineOsc(ctx, V)
        self.env = E.Envelope(ctx, V)
        self.prev_note_on = np.zeros(V, dtype=np.int32)  # Instrument.prev_note_on, carried across blocks
        ev = wl["events"]
        n = max(ev.size, 1)
        self.event_on = np.zeros(n, dtype=np.int32)
        self.event_on[:ev.size] = ev["note_on"]
        # Portamento rows: 2 per event, row 2e + prev_note_on
        self.porta_rows = np.zeros(2 * n, dtype=L.param_dtypes[L.MOD_PORTAMENTO])
        self.porta_rows["sample_rate"] = sample_rate
        self.porta_rows["curve"] = (L.CURVE_CUBED, 0.5)
        self.porta_rows["goal"][0:2 * ev.size:2] = ev["freq"]
        self.porta_rows["goal"][1:2 * ev.size:2] = ev["freq"]
        self.porta_rows["note_on"][0:2 * ev.size:2] = ev["note_on"]
        self.porta_rows["note_on"][1:2 * ev.size:2] = ev["note_on"]
        self.porta_rows["prev_note_on"][1::2] = 1
        self.env_rows = np.zeros(n, dtype=L.param_dtypes[L.MOD_ENVELOPE])
        self.env_rows["sample_rate"] = sample_rate
        self.env_rows["attack"] = (L.CURVE_CUBED, 0.025)
        self.env_rows["decay"] = (L.CURVE_CUBED, 0.1)
        self.env_rows["release"] = (L.CURVE_CUBED, 1.0)
        self.env_rows["sustain_volume"] = 0.5
        self.env_rows["note_on"][:ev.size] = ev["note_on"]
        self.trisaw_rows = np.zeros(1, dtype=L.param_dtypes[L.MOD_TRISAW_OSC])
        self.trisaw_rows[0] = (sample_rate, 0.0, 0.0)
        self.sine_rows = np.zeros(1, dtype=L.param_dtypes[L.MOD_SINE_OSC])
        self.sine_rows[0] = (sample_rate, 0.0, 0.0)

    def span_tables(self, first_frame, num_frames):
        """The Trigger's span table and the two tables derived from the instrument's own host state (prev_note_on)."""
        off, spans = self.trigger.spans(first_frame, num_frames)
        on = self.event_on[spans["param"]] if spans.size else np.zeros(0, np.int32)
        prev = np.zeros(spans.size, dtype=np.int32)
        if spans.size:
            prev[1:] = on[:-1]
            counts = np.diff(off).astype(np.int64)
            has = counts > 0
            prev[off[:-1][has]] = self.prev_note_on[has]   # first span of each voice in this block: carried state
            last = (off[1:][has] - 1).astype(np.int64)
            self.prev_note_on[has] = on[last]
        sp_porta = spans.copy()
        sp_porta["param"] = 2 * spans["param"] + prev.astype(np.uint32)
        sp_env = spans.copy()
        sp_env["flags"] = ((prev == 0) & (on != 0)).astype(np.uint32)
        sp_plain = spans.copy()
        sp_plain["param"] = 0
        return off, spans, sp_porta, sp_env, sp_plain

    def prepare_block(self, first_frame, num_frames):
        if self.dev_trigger is not None:  # Trigger replay and the prev_note_on bookkeeping both on the device
            src = self.dev_trigger.spans(first_frame, num_frames)
            views = self.dev_trigger.prev_note_on_views(src)
            src.free()
            return views
        off, _, a_porta, a_env, a_plain = self.span_tables(first_frame, num_frames)
        return tuple(self.ctx.spans(self.voices, num_frames, off, a) for a in (a_porta, a_env, a_plain))

    def paint_block(self, first_frame, num_frames, out, temps, detune_in, index_in, prepared=None):
        ctx = self.ctx
        t0, t1, t2, t3, t4, t5 = temps
        sp_porta, sp_env, sp = prepared or self.prepare_block(first_frame, num_frames)
        ctx.zero(sp, t0)
        self.porta.paint(sp_porta, [t0], [], self.porta_rows)
        ctx.zero(sp, t1)
        ctx.multiply(sp, t1, t0, detune_in)
        ctx.zero(sp, t2)
        self.trisaw.paint(sp, [t2], [], self.trisaw_rows, freq=t1)
        ctx.zero(sp, t3)
        self.modulator.paint(sp, [t3], [], self.sine_rows, freq=t0)
        ctx.multiplyWith(sp, t3, index_in)
        ctx.zero(sp, t4)
        self.carrier.paint(sp, [t4], [], self.sine_rows, freq=t0, phase=t3)
        ctx.zero(sp, t5)
        self.env.paint(sp_env, [t5], [], self.env_rows)
        ctx.multiply(sp, out, t2, t5)
        ctx.multiply(sp, out, t4, t5)
        if prepared is None:
            for s in (sp_porta, sp_env, sp):
                s.free()

    # ---- the same body as ONE fused kernel (zcu_graph_*): temps in registers ----
    @staticmethod
    def graph_list(voices, output_mode=L.GRAPH_OUT_BUFFERS, channels=2):
        """(builder, config, field indices) of the body above; host only."""
        b = E.GraphBuilder()
        T, OUT, IN = E.T, E.OUT, E.IN
        PORTA, ENV, PLAIN = 0, 1, 2  # the three span tables of prepare_block (same spans; rows / flags differ)
        b.zero(T(0), table=PLAIN)
        f_porta = b.call(L.MOD_PORTAMENTO, T(0), table=PORTA)
        b.zero(T(1), table=PLAIN)
        b.multiply(T(1), T(0), IN(0), table=PLAIN)
        b.zero(T(2), table=PLAIN)
        f_trisaw = b.call(L.MOD_TRISAW_OSC, T(2), (T(1),), table=PLAIN)
        b.zero(T(3), table=PLAIN)
        f_mod = b.call(L.MOD_SINE_OSC, T(3), (T(0),), table=PLAIN)
        b.multiplyWith(T(3), IN(1), table=PLAIN)
        b.zero(T(4), table=PLAIN)
        f_car = b.call(L.MOD_SINE_OSC, T(4), (T(0), T(3)), table=PLAIN)
        b.zero(T(5), table=PLAIN)
        f_env = b.call(L.MOD_ENVELOPE, T(5), table=ENV)
        b.multiply(OUT(0), T(2), T(5), table=PLAIN)
        b.multiply(OUT(0), T(4), T(5), table=PLAIN)
        cfg = E.graph_config(voices, b, 6, num_inputs=2, num_span_tables=3, output_mode=output_mode, channels=channels)
        return b, cfg, (f_porta, f_trisaw, f_mod, f_car, f_env)

    def build_graph(self, output_mode=L.GRAPH_OUT_BUFFERS, channels=2, gains=None):
        b, cfg, (f_porta, f_trisaw, f_mod, f_car, f_env) = self.graph_list(self.voices, output_mode, channels)
        g = E.Graph(self.ctx, cfg, b)
        g.set_field(f_porta, self.porta.bank, self.porta_rows)
        g.set_field(f_trisaw, self.trisaw.bank, self.trisaw_rows)
        g.set_field(f_mod, self.modulator.bank, self.sine_rows)
        g.set_field(f_car, self.carrier.bank, self.sine_rows)
        g.set_field(f_env, self.env.bank, self.env_rows)
        if gains is not None:
            g.set_gains(gains)
        self.graph = g
        return g

    def paint_block_fused(self, first_frame, num_frames, out, detune_in, index_in, prepared=None):
        tabs = prepared or self.prepare_block(first_frame, num_frames)
        self.graph.paint(tabs, [out], [detune_in, index_in])
        if prepared is None:
            for s in tabs:
                s.free()
