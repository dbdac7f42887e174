// zangcu.zig — the Zig side of the boundary: module-contract-shaped wrappers over include/zangcu.h.
//
// SOURCE ONLY: there is no Zig toolchain in the build image or on the GPU box, so this file has never been
// compiled.  It shows the binding a zang maintainer would add (build.zig: exe.addIncludePath("include");
// exe.linkSystemLibrary("zangcu"); exe.linkLibC();).  The Python ctypes harness in zang_b200/ is the
// binding that actually runs in this repository's tests.
//
// A zang module paints ONE voice:   paint(self, span, outputs, temps, note_id_changed, params)
// A bank paints EVERY voice of a block of mix buffers: the per-voice (span, note_id_changed, params) triples
// that Trigger.next() yields are collected into a span table, then one call paints them all on the GPU.
const std = @import("std");
pub const c = @cImport({
    @cInclude("zangcu.h");
});

pub const Error = error{ Invalid, NoDevice, Cuda, OutOfMemory, Unsupported, Nccl };

fn check(rc: c_int) Error!void {
    return switch (rc) {
        c.ZCU_OK => {},
        c.ZCU_ERR_INVALID => Error.Invalid,
        c.ZCU_ERR_NO_DEVICE => Error.NoDevice, // there is no CPU fallback
        c.ZCU_ERR_CUDA => Error.Cuda,
        c.ZCU_ERR_OOM => Error.OutOfMemory,
        c.ZCU_ERR_UNSUPPORTED => Error.Unsupported,
        else => Error.Nccl,
    };
}

pub const Context = struct {
    h: *c.zcu_ctx,

    pub fn init(device: c_int) Error!Context {
        var h: ?*c.zcu_ctx = null;
        try check(c.zcu_ctx_create(device, &h));
        return .{ .h = h.? };
    }
    pub fn deinit(self: Context) void {
        c.zcu_ctx_destroy(self.h);
    }
    pub fn sync(self: Context) Error!void {
        try check(c.zcu_ctx_sync(self.h));
    }
    /// ZCU_PAINT_EXACT (default) or ZCU_PAINT_TOLERANCE (Filter.paint in device-verified time segments, <= 1e-5)
    pub fn setPaintMode(self: Context, mode: c_int) Error!void {
        try check(c.zcu_ctx_set_paint_mode(self.h, mode));
    }
    /// 0 = every time-segmented Noise paint so far is bit-identical to the sequential loop
    pub fn noiseResyncCount(self: Context) Error!u64 {
        var n: u64 = 0;
        try check(c.zcu_ctx_noise_resync_count(self.h, &n));
        return n;
    }
};

/// zang.Delay(DELAY_SAMPLES) for every voice (src/zang/delay.zig:7-17) + SimpleDelay.paint (examples/modules.zig:341-388).
/// The comptime delay length becomes a run-time argument of the bank; the module contract is unchanged.
pub fn SimpleDelayBank(comptime DELAY_SAMPLES: u32) type {
    return struct {
        pub const num_outputs = 1;
        pub const num_temps = 0;
        ctx: Context,
        bank: *c.zcu_bank,

        pub fn init(ctx: Context, voices: u32) Error!@This() {
            var b: ?*c.zcu_bank = null;
            try check(c.zcu_delay_bank_create(ctx.h, c.ZCU_MOD_SIMPLE_DELAY, voices, DELAY_SAMPLES, &b));
            return .{ .ctx = ctx, .bank = b.? };
        }
        pub fn reset(self: *@This()) Error!void {
            try check(c.zcu_bank_reset(self.ctx.h, self.bank, 0));
        }
        pub fn paint(self: *@This(), spans: *const c.zcu_spans, outputs: [1]Buffer, temps: [0]Buffer, input: Buffer) Error!void {
            _ = temps;
            try check(c.zcu_paint_simple_delay(self.ctx.h, self.bank, spans, outputs[0].h, input.h));
        }
    };
}

/// FilteredEchoes(DELAY_SAMPLES) (examples/modules.zig:392-460): Params{input, feedback_volume, cutoff}; the reference's two
/// temps are not needed (the device module keeps them in registers).
pub fn FilteredEchoesBank(comptime DELAY_SAMPLES: u32) type {
    return struct {
        pub const num_outputs = 1;
        pub const num_temps = 0;
        pub const Params = c.zcu_echoes_params;
        ctx: Context,
        bank: *c.zcu_bank,

        pub fn init(ctx: Context, voices: u32) Error!@This() {
            var b: ?*c.zcu_bank = null;
            try check(c.zcu_delay_bank_create(ctx.h, c.ZCU_MOD_FILTERED_ECHOES, voices, DELAY_SAMPLES, &b));
            return .{ .ctx = ctx, .bank = b.? };
        }
        pub fn paint(self: *@This(), spans: *const c.zcu_spans, outputs: [1]Buffer, temps: [0]Buffer, params: []const Params, input: Buffer) Error!void {
            _ = temps;
            try check(c.zcu_paint_filtered_echoes(self.ctx.h, self.bank, spans, params.ptr, @intCast(params.len), outputs[0].h, input.h));
        }
    };
}

/// The fmsynth example's feedback Oscillator (examples/example_fmsynth.zig:26-89): Params{sample_rate, waveform, freq, phase,
/// feedback}; freq is a buffer argument (a slice in the reference), phase null stands for the reference's zero phase.
pub const FmOscillatorBank = struct {
    pub const num_outputs = 1;
    pub const num_temps = 0;
    pub const Params = c.zcu_fm_oscillator_params;
    ctx: Context,
    bank: *c.zcu_bank,

    pub fn init(ctx: Context, voices: u32) Error!FmOscillatorBank {
        var b: ?*c.zcu_bank = null;
        try check(c.zcu_bank_create(ctx.h, c.ZCU_MOD_FM_OSCILLATOR, voices, 0, &b));
        return .{ .ctx = ctx, .bank = b.? };
    }
    pub fn paint(self: *FmOscillatorBank, spans: *const c.zcu_spans, outputs: [1]Buffer, temps: [0]Buffer, rows: []const Params, freq: Buffer, phase: ?Buffer) Error!void {
        _ = temps;
        try check(c.zcu_paint_fm_oscillator(self.ctx.h, self.bank, spans, rows.ptr, @intCast(rows.len), outputs[0].h, freq.h, if (phase) |p| p.h else null));
    }
};

/// [frames][voices] f32 device buffer: the batched form of the caller-owned []f32
pub const Buffer = struct {
    h: *c.zcu_buf,
    pub fn init(ctx: Context, frames: u32, voices: u32) Error!Buffer {
        var h: ?*c.zcu_buf = null;
        try check(c.zcu_buf_alloc(ctx.h, frames, voices, &h));
        return .{ .h = h.? };
    }
    pub fn deinit(self: Buffer) void {
        c.zcu_buf_free(self.h);
    }
};

/// Collects what `while (trigger.next(&ctr)) |result|` yields for every voice of a block.
pub const SpanTableBuilder = struct {
    voice_off: std.ArrayList(u32),
    spans: std.ArrayList(c.zcu_span),

    pub fn init(allocator: std.mem.Allocator) SpanTableBuilder {
        return .{ .voice_off = std.ArrayList(u32).init(allocator), .spans = std.ArrayList(c.zcu_span).init(allocator) };
    }
    pub fn beginVoice(self: *SpanTableBuilder) !void {
        try self.voice_off.append(@intCast(self.spans.items.len));
    }
    /// frame_base = index of the mix buffer inside the block times its length
    pub fn add(self: *SpanTableBuilder, frame_base: u32, start: usize, end: usize, note_id_changed: bool, params_row: u32) !void {
        try self.spans.append(.{
            .start = frame_base + @as(u32, @intCast(start)),
            .end = frame_base + @as(u32, @intCast(end)),
            .flags = if (note_id_changed) c.ZCU_SPAN_NOTE_ID_CHANGED else 0,
            .param = params_row,
        });
    }
    pub fn upload(self: *SpanTableBuilder, ctx: Context, voices: u32, frames: u32) Error!*c.zcu_spans {
        self.voice_off.append(@intCast(self.spans.items.len)) catch return Error.OutOfMemory;
        var h: ?*c.zcu_spans = null;
        try check(c.zcu_spans_upload(ctx.h, voices, frames, self.voice_off.items.ptr, self.spans.items.ptr, &h));
        return h.?;
    }
};

/// mod.PulseOsc for `voices` voices: same contract, one more axis.
pub const PulseOscBank = struct {
    pub const num_outputs = 1;
    pub const num_temps = 0;
    pub const Params = c.zcu_pulse_params; // { sample_rate, freq, color }

    ctx: Context,
    bank: *c.zcu_bank,

    pub fn init(ctx: Context, voices: u32) Error!PulseOscBank {
        var b: ?*c.zcu_bank = null;
        try check(c.zcu_bank_create(ctx.h, c.ZCU_MOD_PULSE_OSC, voices, 0, &b));
        return .{ .ctx = ctx, .bank = b.? };
    }
    /// `freq` null = zang.constant(params.freq), non-null = zang.buffer(freq)
    pub fn paint(
        self: *PulseOscBank,
        spans: *const c.zcu_spans,
        outputs: [num_outputs]Buffer,
        temps: [num_temps]Buffer,
        params: []const Params,
        freq: ?Buffer,
    ) Error!void {
        _ = temps;
        try check(c.zcu_paint_pulse_osc(self.ctx.h, self.bank, spans, params.ptr, @intCast(params.len), outputs[0].h, if (freq) |f| f.h else null));
    }
};

/// mod.Filter
pub const FilterBank = struct {
    pub const num_outputs = 1;
    pub const num_temps = 0;
    pub const Params = c.zcu_filter_params; // { type, cutoff, res }; input/cutoff/res buffers are call arguments

    ctx: Context,
    bank: *c.zcu_bank,

    pub fn init(ctx: Context, voices: u32) Error!FilterBank {
        var b: ?*c.zcu_bank = null;
        try check(c.zcu_bank_create(ctx.h, c.ZCU_MOD_FILTER, voices, 0, &b));
        return .{ .ctx = ctx, .bank = b.? };
    }
    pub fn cutoffFromFrequency(frequency: f32, sample_rate: f32) f32 {
        return c.zcu_filter_cutoff_from_frequency(frequency, sample_rate);
    }
    pub fn paint(self: *FilterBank, spans: *const c.zcu_spans, outputs: [1]Buffer, temps: [0]Buffer, params: []const Params, input: Buffer, cutoff: ?Buffer, res: ?Buffer) Error!void {
        _ = temps;
        try check(c.zcu_paint_filter(self.ctx.h, self.bank, spans, params.ptr, @intCast(params.len), outputs[0].h, input.h, if (cutoff) |b| b.h else null, if (res) |b| b.h else null));
    }
};

/// mod.Envelope
pub const EnvelopeBank = struct {
    pub const num_outputs = 1;
    pub const num_temps = 0;
    pub const Params = c.zcu_envelope_params;

    ctx: Context,
    bank: *c.zcu_bank,

    pub fn init(ctx: Context, voices: u32) Error!EnvelopeBank {
        var b: ?*c.zcu_bank = null;
        try check(c.zcu_bank_create(ctx.h, c.ZCU_MOD_ENVELOPE, voices, 0, &b));
        return .{ .ctx = ctx, .bank = b.? };
    }
    pub fn paint(self: *EnvelopeBank, spans: *const c.zcu_spans, outputs: [1]Buffer, temps: [0]Buffer, params: []const Params) Error!void {
        _ = temps;
        try check(c.zcu_paint_envelope(self.ctx.h, self.bank, spans, params.ptr, @intCast(params.len), outputs[0].h));
    }
};

/// examples/modules.zig NiceInstrument for a bank of voices, composed exactly like the original.
pub const NiceInstrumentBank = struct {
    pub const num_outputs = 1;
    pub const num_temps = 2;
    pub const Params = struct { sample_rate: f32, freq: f32, note_on: bool };

    ctx: Context,
    color: f32,
    osc: PulseOscBank,
    flt: FilterBank,
    env: EnvelopeBank,

    pub fn paint(
        self: *NiceInstrumentBank,
        allocator: std.mem.Allocator,
        spans: *const c.zcu_spans,
        outputs: [num_outputs]Buffer,
        temps: [num_temps]Buffer,
        params: []const Params, // one row per note event; spans refer to rows
    ) !void {
        const h = self.ctx.h;
        var pp = try allocator.alloc(PulseOscBank.Params, params.len);
        defer allocator.free(pp);
        var fp = try allocator.alloc(FilterBank.Params, params.len);
        defer allocator.free(fp);
        var ep = try allocator.alloc(EnvelopeBank.Params, params.len);
        defer allocator.free(ep);
        for (params, 0..) |p, i| {
            pp[i] = .{ .sample_rate = p.sample_rate, .freq = p.freq, .color = self.color };
            fp[i] = .{ .type = c.ZCU_FILTER_LOW_PASS, .cutoff = FilterBank.cutoffFromFrequency(p.freq * 8.0, p.sample_rate), .res = 0.7 };
            ep[i] = .{
                .sample_rate = p.sample_rate,
                .attack = .{ .kind = c.ZCU_CURVE_CUBED, .duration = 0.01 },
                .decay = .{ .kind = c.ZCU_CURVE_CUBED, .duration = 0.1 },
                .release = .{ .kind = c.ZCU_CURVE_CUBED, .duration = 0.5 },
                .sustain_volume = 0.8,
                .note_on = @intFromBool(p.note_on),
            };
        }
        try check(c.zcu_zero(h, spans, temps[0].h));
        try self.osc.paint(spans, .{temps[0]}, .{}, pp, null);
        try check(c.zcu_multiply_with_scalar(h, spans, temps[0].h, 0.5));
        try check(c.zcu_zero(h, spans, temps[1].h));
        try self.flt.paint(spans, .{temps[1]}, .{}, fp, temps[0], null, null);
        try check(c.zcu_zero(h, spans, temps[0].h));
        try self.env.paint(spans, .{temps[0]}, .{}, ep);
        try check(c.zcu_multiply(h, spans, outputs[0].h, temps[0].h, temps[1].h));
    }
};

/// zang.Trigger(P) with P = a params-row index; the host keeps the rows.
pub const Trigger = struct {
    t: c.zcu_trigger,
    pub fn init() Trigger {
        var t: c.zcu_trigger = undefined;
        c.zcu_trigger_init(&t);
        return .{ .t = t };
    }
    pub fn reset(self: *Trigger) void {
        c.zcu_trigger_reset(&self.t);
    }
    /// the whole `while (trigger.next(&ctr))` loop of one mix buffer, appended to `out`
    pub fn run(self: *Trigger, start: u32, end: u32, impulses: []const c.zcu_impulse, rows: []const u32, frame_base: u32, out: []c.zcu_span) u32 {
        return c.zcu_trigger_run(&self.t, start, end, impulses.ptr, rows.ptr, @intCast(impulses.len), frame_base, out.ptr, @intCast(out.len));
    }
};

// ---- round 2: device-side Trigger, fused graphs, overwrite mode (INTEGRATION.md §6) -------------------------------------

/// ImpulseQueue + Trigger of every voice of a bank replayed on the GPU (examples/example_polyphony.zig:61-92 per voice).
pub const EventBank = struct {
    ctx: Context,
    h: *c.zcu_event_bank,

    pub fn init(ctx: Context, event_off: []const u32, events: []const c.zcu_note_event, buffer_frames: u32, total_frames: u32) !EventBank {
        var h: ?*c.zcu_event_bank = null;
        try check(c.zcu_event_bank_create(ctx.h, @intCast(event_off.len - 1), event_off.ptr, events.ptr, buffer_frames, total_frames, &h));
        return .{ .ctx = ctx, .h = h.? };
    }
    pub fn deinit(self: *EventBank) void {
        c.zcu_event_bank_free(self.h);
    }
    /// span table of the block, built on the device; the Trigger state is carried to the next call
    pub fn spans(self: *EventBank, first_frame: u32, num_frames: u32) !*c.zcu_spans {
        var s: ?*c.zcu_spans = null;
        try check(c.zcu_event_bank_spans(self.ctx.h, self.h, first_frame, num_frames, &s));
        return s.?;
    }
};

/// `zang.zero(span, out); module.paint(span, .{out}, ...)` as ONE call for everything painted inside `body`.
pub fn overwriting(ctx: Context, body: anytype) !void {
    try check(c.zcu_ctx_set_paint_output(ctx.h, c.ZCU_OUT_OVERWRITE));
    defer _ = c.zcu_ctx_set_paint_output(ctx.h, c.ZCU_OUT_ACCUMULATE);
    try body.run();
}

/// One fused kernel for an instrument body / a script module's instruction list (src/zangscript/codegen.zig:60-121).
pub const Graph = struct {
    ctx: Context,
    h: *c.zcu_graph,

    pub fn init(ctx: Context, cfg: c.zcu_graph_config, instrs: []const c.zcu_graph_instr, field_kinds: []const i32) !Graph {
        var h: ?*c.zcu_graph = null;
        try check(c.zcu_graph_create(ctx.h, &cfg, instrs.ptr, @intCast(instrs.len), field_kinds.ptr, &h));
        return .{ .ctx = ctx, .h = h.? };
    }
    pub fn deinit(self: *Graph) void {
        c.zcu_graph_free(self.h);
    }
    /// rows: a slice of the field's `zcu_<module>_params`
    pub fn setField(self: *Graph, field: u32, bank: ?*c.zcu_bank, rows: anytype) !void {
        try check(c.zcu_graph_set_field(self.ctx.h, self.h, field, bank, rows.ptr, @sizeOf(@TypeOf(rows[0])), @intCast(rows.len)));
    }
    pub fn paint(self: *Graph, tables: []const ?*const c.zcu_spans, outputs: []const ?*c.zcu_buf, inputs: []const ?*const c.zcu_buf) !void {
        try check(c.zcu_graph_paint(self.ctx.h, self.h, tables.ptr, outputs.ptr, if (inputs.len > 0) inputs.ptr else null));
    }
    /// 0 = sized to the GPU, 1 = off.  A graph with a Filter is cut in time only after
    /// zcu_ctx_set_paint_mode(ZCU_PAINT_TOLERANCE) (and, if a buffer drives its cutoff or resonance, only when forced here).
    pub fn setSegments(self: *Graph, segments: u32) !void {
        try check(c.zcu_graph_set_segments(self.h, segments));
    }
    /// (warp, segment) pairs whose Filter warm-up window had to be lengthened since the last call; synchronises
    pub fn warmRetries(self: *Graph) !u32 {
        var n: u32 = 0;
        try check(c.zcu_graph_warm_retries(self.ctx.h, self.h, &n));
        return n;
    }
};

/// zang.mixDown (src/zang/mixdown.zig:8-86) into `dst` (s16le or s8), through device scratch that is part of the library
/// image: the call allocates nothing (hosts that must not call cudaMalloc).  dst.len <= ZCU_MIX_DOWN_STATIC_MAX_BYTES.
/// No counterpart in the reference; added at the request of the run's operators.
pub fn mixDownStatic(ctx: Context, mix: *const c.zcu_buf, audio_format: c_int, vol: f32, dst: []u8) !void {
    try check(c.zcu_mix_down_static(ctx.h, mix, audio_format, vol, dst.ptr, dst.len));
}
