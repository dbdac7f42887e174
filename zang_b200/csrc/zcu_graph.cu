// zcu_graph.cu — fused module graphs: an instrument's instruction list (include/zangcu.h, "fused module graphs";
// reference: src/zangscript/codegen.zig:60-121, src/zangscript/runtime.zig:445-740, examples/modules.zig) is turned into
// CUDA source for ONE kernel over zcu_graph_rt.cuh / zcu_modules.cuh, compiled with NVRTC for sm_100a (--fmad=false, the
// same arithmetic contract as the rest of the library) and loaded through the runtime's library API.
//
// Shape of a generated kernel (one thread per voice, 32 adjacent voices per warp, 4 warps per CTA):
//   cursors over the voice's spans of every table; module objects with their state in registers; frames walked in
//   warp-uniform RUNS (up to the nearest span boundary of any table of any of the warp's voices); inside a run the whole
//   instruction list is evaluated per sample with the temps in registers.  Two copies of the per-sample body: one for runs
//   in which every table is live on every lane (no predicates, unrolled x4), one predicated per table.
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "zcu_graph_rt.cuh"
#include "zcu_internal.h"

using namespace zcu;

namespace {

// ---- embedded device headers (build/zcu_embedded_headers.inc: the files as string literals) ------------------------
#include "build/zcu_embedded_headers.inc"

const char* kStubInts =
    "#pragma once\n"
    "typedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\n"
    "typedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n"
    "typedef unsigned long size_t; typedef unsigned long uintptr_t;\n";
const char* kStubEmpty = "#pragma once\n";

// ---- NVRTC, loaded lazily -------------------------------------------------------------------------------------------
struct Nvrtc {
    void* lib = nullptr;
    int (*CreateProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*DestroyProgram)(void**) = nullptr;
    int (*CompileProgram)(void*, int, const char* const*) = nullptr;
    int (*GetProgramLogSize)(void*, size_t*) = nullptr;
    int (*GetProgramLog)(void*, char*) = nullptr;
    int (*GetCUBINSize)(void*, size_t*) = nullptr;
    int (*GetCUBIN)(void*, char*) = nullptr;
    std::string err;
    bool load() {
        if (lib) return true;
        const char* names[] = {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so"};
        for (const char* n : names) {
            lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (lib) break;
        }
        if (!lib) {
            const char* e = dlerror();
            err = std::string("dlopen(libnvrtc) failed: ") + (e ? e : "unknown");
            return false;
        }
        bool ok = true;
        auto sym = [&](const char* n) {
            void* p = dlsym(lib, n);
            if (!p) {
                ok = false;
                err = std::string("libnvrtc lacks ") + n;
            }
            return p;
        };
        CreateProgram = (decltype(CreateProgram))sym("nvrtcCreateProgram");
        DestroyProgram = (decltype(DestroyProgram))sym("nvrtcDestroyProgram");
        CompileProgram = (decltype(CompileProgram))sym("nvrtcCompileProgram");
        GetProgramLogSize = (decltype(GetProgramLogSize))sym("nvrtcGetProgramLogSize");
        GetProgramLog = (decltype(GetProgramLog))sym("nvrtcGetProgramLog");
        GetCUBINSize = (decltype(GetCUBINSize))sym("nvrtcGetCUBINSize");
        GetCUBIN = (decltype(GetCUBIN))sym("nvrtcGetCUBIN");
        if (!ok) {
            dlclose(lib);
            lib = nullptr;
        }
        return ok;
    }
};
Nvrtc g_nvrtc;
std::mutex g_nvrtc_mu;

bool nvrtc_compile(const std::string& src, std::vector<char>& cubin, std::string& log) {
    std::lock_guard<std::mutex> lock(g_nvrtc_mu);
    if (!g_nvrtc.load()) {
        log = g_nvrtc.err;
        return false;
    }
    const char* hn[] = {"zcu_graph_rt.cuh", "zcu_modules.cuh", "zcu_exact_accum.cuh", "../../include/zangcu.h", "cstdint", "stdint.h", "stddef.h",
                        "cmath",            "cstring"};
    const char* hs[] = {kEmbedded_zcu_graph_rt, kEmbedded_zcu_modules, kEmbedded_zcu_exact_accum, kEmbedded_zangcu_h, kStubInts, kStubInts, kStubInts,
                        kStubEmpty,             kStubEmpty};
    void* prog = nullptr;
    if (g_nvrtc.CreateProgram(&prog, src.c_str(), "zcu_graph_kernel.cu", 9, hs, hn) != 0) {
        log = "nvrtcCreateProgram failed";
        return false;
    }
    // the library's arithmetic contract (Makefile): no contraction, IEEE division and square root, denormals kept
    const char* opts[] = {"--gpu-architecture=sm_100a", "--fmad=false", "--prec-div=true", "--prec-sqrt=true", "--ftz=false", "-std=c++17",
                          "-lineinfo",                  "-default-device"};
    const int rc = g_nvrtc.CompileProgram(prog, 8, opts);
    size_t ls = 0;
    g_nvrtc.GetProgramLogSize(prog, &ls);
    if (ls > 1) {
        std::vector<char> l(ls + 1, 0);
        g_nvrtc.GetProgramLog(prog, l.data());
        log = l.data();
    }
    bool ok = rc == 0;
    if (ok) {
        size_t cs = 0;
        ok = g_nvrtc.GetCUBINSize(prog, &cs) == 0 && cs > 0;
        if (ok) {
            cubin.resize(cs);
            ok = g_nvrtc.GetCUBIN(prog, cubin.data()) == 0;
        }
        if (!ok) log += "\nnvrtcGetCUBIN failed";
    }
    g_nvrtc.DestroyProgram(&prog);
    if (const char* dir = getenv("ZCU_GRAPH_DUMP")) {  // inspection: the generated source and machine code of the last graph
        const std::string base = std::string(dir) + "/zcu_graph_kernel";
        if (FILE* f = fopen((base + ".cu").c_str(), "w")) {
            fwrite(src.data(), 1, src.size(), f);
            fclose(f);
        }
        if (ok)
            if (FILE* f = fopen((base + ".cubin").c_str(), "wb")) {
                fwrite(cubin.data(), 1, cubin.size(), f);
                fclose(f);
            }
    }
    return ok;
}

// ---- code generation -----------------------------------------------------------------------------------------------
void appendf(std::string& s, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    s += buf;
}

std::string float_lit(float x) {  // exact: the bits
    uint32_t u;
    memcpy(&u, &x, 4);
    char b[64];
    snprintf(b, sizeof b, "__uint_as_float(0x%08xu)", u);
    return b;
}

struct FieldInfo {
    int kind = -1;
    std::string type;  // device module type
    uint32_t table = 0;
    bool used = false;
    bool has_run = false;  // the module offers run_limit / run_begin / sample_run / run_end (zcu_modules.cuh: HAS_RUN)
    bool skippable = false;  // exact skip-ahead from its own params alone (advance()), or a Decimator (counter + replay window)
    bool warmable = false;   // a Filter: a time segment may start it from a verified warm-up window (ZCU_PAINT_TOLERANCE)
    bool is_delay = false;   // a zang.Delay(N) served by DELAY_READ / DELAY_WRITE
    int delay_reads = 0, delay_writes = 0;
};

bool module_type(int kind, const zcu_graph_instr& in, std::string& type, std::string& err) {
    const bool b0 = in.in[0].kind != ZCU_GREF_NONE, b1 = in.in[1].kind != ZCU_GREF_NONE, b2 = in.in[2].kind != ZCU_GREF_NONE;
    auto tf = [](bool b) { return b ? "true" : "false"; };
    auto need0 = [&]() {
        if (!b0) err = "this module needs its input buffer (in[0])";
        return b0;
    };
    char t[96];
    switch (kind) {
        case ZCU_MOD_SINE_OSC: snprintf(t, sizeof t, "SineMod<%s, %s>", tf(b0), tf(b1)); break;
        case ZCU_MOD_PULSE_OSC: snprintf(t, sizeof t, "PulseMod<%s>", tf(b0)); break;
        case ZCU_MOD_TRISAW_OSC: snprintf(t, sizeof t, "TriSawMod<%s>", tf(b0)); break;
        case ZCU_MOD_NOISE: snprintf(t, sizeof t, "NoiseMod"); break;
        case ZCU_MOD_ENVELOPE: snprintf(t, sizeof t, "EnvelopeMod"); break;
        case ZCU_MOD_GATE: snprintf(t, sizeof t, "GateMod"); break;
        case ZCU_MOD_FILTER:
            if (!need0()) return false;
            snprintf(t, sizeof t, "FilterMod<%s, %s>", tf(b1), tf(b2));
            break;
        case ZCU_MOD_PORTAMENTO: snprintf(t, sizeof t, "PortamentoMod"); break;
        case ZCU_MOD_CURVE: snprintf(t, sizeof t, "CurveMod"); break;
        case ZCU_MOD_SAMPLER: snprintf(t, sizeof t, "SamplerMod"); break;
        case ZCU_MOD_DECIMATOR:
            if (!need0()) return false;
            snprintf(t, sizeof t, "DecimatorMod");
            break;
        case ZCU_MOD_DISTORTION:
            if (!need0()) return false;
            snprintf(t, sizeof t, "DistortionMod");
            break;
        case ZCU_MOD_CYCLE: snprintf(t, sizeof t, "CycleMod<%s>", tf(b0)); break;
        case ZCU_MOD_FM_OSCILLATOR:
            if (!need0()) return false;
            snprintf(t, sizeof t, "FmOscMod<%s>", tf(b1));
            break;
        case ZCU_MOD_SIMPLE_DELAY:
            if (!need0()) return false;
            snprintf(t, sizeof t, "SimpleDelayMod");
            break;
        case ZCU_MOD_FILTERED_ECHOES:
            if (!need0()) return false;
            snprintf(t, sizeof t, "FilteredEchoesMod");
            break;
        default: err = "unknown module kind"; return false;
    }
    type = t;
    return true;
}

bool codegen(const zcu_graph_config& c, const zcu_graph_instr* ins, uint32_t n, const int32_t* field_kinds, std::string& src, std::string& err,
             bool* segmentable_out = nullptr, bool* needs_tolerance_out = nullptr, std::vector<int>* filter_buffered_out = nullptr) {
    auto fail = [&](uint32_t i, const char* what) {
        char b[256];
        snprintf(b, sizeof b, "instruction %u: %s", i, what);
        err = b;
        return false;
    };
    if (c.voices == 0) return err = "voices must be > 0", false;
    if (c.num_outputs == 0 || c.num_outputs > (uint32_t)kGraphMaxBufs || c.num_inputs > (uint32_t)kGraphMaxBufs)
        return err = "1..4 outputs and 0..4 inputs", false;
    if (c.num_span_tables == 0 || c.num_span_tables > (uint32_t)kGraphMaxTables) return err = "1..4 span tables", false;
    if (c.num_fields > (uint32_t)kGraphMaxFields) return err = "at most 12 fields", false;
    if (c.num_temps > 32) return err = "at most 32 temps", false;
    if (c.num_fields && !field_kinds) return err = "null field_kinds", false;
    const bool mix = c.output_mode == ZCU_GRAPH_OUT_MIX;
    if (c.output_mode != ZCU_GRAPH_OUT_BUFFERS && !mix) return err = "unknown output_mode", false;
    if (mix && (c.channels < 1 || c.channels > 2 || c.num_outputs != 1)) return err = "MIX: one output, 1 or 2 channels", false;
    if (n == 0 || !ins) return err = "empty instruction list", false;

    const uint32_t K = c.num_span_tables;
    std::vector<FieldInfo> fields(c.num_fields);
    std::vector<uint32_t> out_tables(c.num_outputs, 0);  // bit k: an instruction on table k writes this output
    std::vector<bool> in_used(c.num_inputs, false);
    auto ref_ok = [&](const zcu_graph_ref& r, bool dest) {
        switch (r.kind) {
            case ZCU_GREF_TEMP: return r.index < c.num_temps;
            case ZCU_GREF_OUTPUT: return r.index < c.num_outputs;
            case ZCU_GREF_INPUT: return !dest && r.index < c.num_inputs;
            default: return false;
        }
    };
    for (uint32_t i = 0; i < n; i++) {
        const zcu_graph_instr& x = ins[i];
        if (x.span_table >= K) return fail(i, "span_table out of range");
        if (x.op != ZCU_GI_DELAY_WRITE) {  // (DELAY_WRITE has no destination buffer)
            if (!ref_ok(x.out, true)) return fail(i, "out must be a temp or an output of the graph");
            if (x.out.kind == ZCU_GREF_OUTPUT) out_tables[x.out.index] |= 1u << x.span_table;
        }
        for (int q = 0; q < 3; q++) {
            if (x.in[q].kind == ZCU_GREF_NONE) continue;
            if (!ref_ok(x.in[q], false)) return fail(i, "operand reference out of range");
            if (x.in[q].kind == ZCU_GREF_INPUT) in_used[x.in[q].index] = true;
        }
        if (x.op == ZCU_GI_CALL) {
            if (x.field >= c.num_fields) return fail(i, "field out of range");
            FieldInfo& f = fields[x.field];
            if (f.used) return fail(i, "a field is one module instance: it can be called once per graph");
            if (field_kinds[x.field] != x.sub) return fail(i, "CALL.sub differs from the field's module kind");
            std::string e2;
            if (!module_type(x.sub, x, f.type, e2)) return fail(i, e2.c_str());
            f.kind = x.sub;
            f.table = x.span_table;
            f.used = true;
            f.has_run = x.sub == ZCU_MOD_ENVELOPE || x.sub == ZCU_MOD_PORTAMENTO || x.sub == ZCU_MOD_CURVE;
            {
                const bool b0 = x.in[0].kind != ZCU_GREF_NONE, b1 = x.in[1].kind != ZCU_GREF_NONE;
                switch (x.sub) {
                    case ZCU_MOD_SINE_OSC: f.skippable = !b0 && !b1; break;  // buffer-driven phases are sequential by nature
                    case ZCU_MOD_PULSE_OSC:
                    case ZCU_MOD_TRISAW_OSC:
                    case ZCU_MOD_CYCLE: f.skippable = !b0; break;
                    case ZCU_MOD_ENVELOPE:
                    case ZCU_MOD_GATE:
                    case ZCU_MOD_PORTAMENTO:
                    case ZCU_MOD_SAMPLER:
                    case ZCU_MOD_DECIMATOR:
                    case ZCU_MOD_NOISE:  // exact GF(2) jump (zcu_modules.cuh), applied once after the walk
                    case ZCU_MOD_CURVE:
                    case ZCU_MOD_DISTORTION: f.skippable = true; break;
                    default: f.skippable = false; break;  // Filter, delays, FM oscillator: one segment
                }
            }
        } else if (x.op == ZCU_GI_DELAY_READ || x.op == ZCU_GI_DELAY_WRITE) {
            if (x.field >= c.num_fields) return fail(i, "field out of range");
            FieldInfo& f = fields[x.field];
            if (field_kinds[x.field] != ZCU_MOD_SIMPLE_DELAY) return fail(i, "a delay instruction needs a ZCU_MOD_SIMPLE_DELAY field (zang.Delay)");
            if (f.used && !f.is_delay) return fail(i, "this field is a module instance, not a delay line");
            if (f.used && f.table != x.span_table) return fail(i, "DELAY_READ and DELAY_WRITE of a field must name the same span table");
            if (x.op == ZCU_GI_DELAY_READ) {
                if (f.delay_reads != f.delay_writes) return fail(i, "DELAY_READ twice without a DELAY_WRITE in between");
                f.delay_reads++;
            } else {
                if (f.delay_reads != f.delay_writes + 1) return fail(i, "DELAY_WRITE without a preceding DELAY_READ");
                if (x.in[0].kind == ZCU_GREF_NONE) return fail(i, "DELAY_WRITE needs the buffer to store (in[0])");
                f.delay_writes++;
            }
            if (f.delay_reads > 1) return fail(i, "a delay line is served once per graph");
            f.kind = ZCU_MOD_SIMPLE_DELAY;
            f.type = "GraphDelayMod";
            f.table = x.span_table;
            f.used = true;
            f.is_delay = true;
        } else if (x.op == ZCU_GI_BASIC) {
            if (x.sub < 0 || x.sub > 10) return fail(i, "unknown basic op");
            const bool na = x.sub == 2 || x.sub == 3 || x.sub == 4 || x.sub == 5 || x.sub == 7 || x.sub == 8 || x.sub == 9;
            const bool nb = x.sub == 3 || x.sub == 7;
            if (na && x.in[0].kind == ZCU_GREF_NONE) return fail(i, "this basic op needs operand a");
            if (nb && x.in[1].kind == ZCU_GREF_NONE) return fail(i, "this basic op needs operand b");
        } else if (x.op == ZCU_GI_ARITH) {
            if (x.sub < 0 || x.sub > ZCU_ARITH_POW) return fail(i, "unknown arith op");
            const bool unary = x.sub >= ZCU_ARITH_NEG && x.sub <= ZCU_ARITH_COS;
            if (unary) {
                if (x.in[0].kind == ZCU_GREF_NONE) return fail(i, "unary arith needs operand a");
            } else {
                if (x.scalar_side != ZCU_SCALAR_A && x.in[0].kind == ZCU_GREF_NONE) return fail(i, "arith needs operand a");
                if (x.scalar_side != ZCU_SCALAR_B && x.in[1].kind == ZCU_GREF_NONE) return fail(i, "arith needs operand b");
            }
        } else {
            return fail(i, "unknown instruction kind");
        }
    }
    for (uint32_t f = 0; f < c.num_fields; f++) {
        if (!fields[f].used) return err = "a field is never called", false;
        if (fields[f].is_delay && fields[f].delay_reads != fields[f].delay_writes) return err = "a DELAY_READ has no DELAY_WRITE", false;
    }

    auto ref = [&](const zcu_graph_ref& r) -> std::string {
        char b[32];
        switch (r.kind) {
            case ZCU_GREF_TEMP: snprintf(b, sizeof b, "t%u", r.index); break;
            case ZCU_GREF_OUTPUT: snprintf(b, sizeof b, "o%u", r.index); break;
            case ZCU_GREF_INPUT: snprintf(b, sizeof b, "x%u", r.index); break;
            default: snprintf(b, sizeof b, "0.0f"); break;
        }
        return b;
    };
    // one instruction on one sample: d, a, b, cc are the C expressions of its operands
    auto statement = [&](const zcu_graph_instr& x, const std::string& d, const std::string& a, const std::string& b, const std::string& cc,
                         bool run_form, bool silent = false) -> std::string {
        const std::string sc = float_lit(x.scalar);
        std::string stmt;
        if (x.op == ZCU_GI_CALL) {
            char t[200];
            snprintf(t, sizeof t, "m%u.%s(%s, %s, %s, %s);", x.field,
                     (silent && fields[x.field].warmable) ? "sample_silent" : ((run_form && fields[x.field].has_run) ? "sample_run" : "sample"), d.c_str(),
                     a.c_str(), b.c_str(), cc.c_str());
            stmt = t;
        } else if (x.op == ZCU_GI_DELAY_READ) {
            char t[96];
            snprintf(t, sizeof t, "m%u.read(%s);", x.field, d.c_str());
            stmt = t;
        } else if (x.op == ZCU_GI_DELAY_WRITE) {
            char t[96];
            snprintf(t, sizeof t, "m%u.write(%s);", x.field, a.c_str());
            stmt = t;
        } else if (x.op == ZCU_GI_BASIC) {
            switch (x.sub) {  // basics.zig:12-78, the expression forms of BasicMod (zcu_modules.cuh)
                case 0: stmt = d + " = 0.0f;"; break;
                case 1: stmt = d + " = " + sc + ";"; break;
                case 2: stmt = d + " = " + a + ";"; break;
                case 3: stmt = d + " += " + a + " + " + b + ";"; break;
                case 4: stmt = d + " += " + a + ";"; break;
                case 5: stmt = d + " += " + a + " + " + sc + ";"; break;
                case 6: stmt = d + " += " + sc + ";"; break;
                case 7: stmt = d + " += " + a + " * " + b + ";"; break;
                case 8: stmt = d + " *= " + a + ";"; break;
                case 9: stmt = d + " += " + a + " * " + sc + ";"; break;
                case 10: stmt = d + " *= " + sc + ";"; break;
            }
        } else {
            const std::string A = x.scalar_side == ZCU_SCALAR_A ? sc : a, B = x.scalar_side == ZCU_SCALAR_B ? sc : b;
            switch (x.sub) {  // runtime.zig:528-700: these overwrite
                case ZCU_ARITH_SUB: stmt = d + " = " + A + " - " + B + ";"; break;
                case ZCU_ARITH_DIV: stmt = d + " = " + A + " / " + B + ";"; break;
                case ZCU_ARITH_MIN: stmt = d + " = zs_min(" + A + ", " + B + ");"; break;
                case ZCU_ARITH_MAX: stmt = d + " = zs_max(" + A + ", " + B + ");"; break;
                case ZCU_ARITH_NEG: stmt = d + " = -" + a + ";"; break;
                case ZCU_ARITH_ABS: stmt = d + " = fabsf(" + a + ");"; break;
                case ZCU_ARITH_SQRT: stmt = d + " = sqrtf(" + a + ");"; break;
                case ZCU_ARITH_SIN: stmt = d + " = sinf(" + a + ");"; break;
                case ZCU_ARITH_COS: stmt = d + " = cosf(" + a + ");"; break;
                case ZCU_ARITH_POW: stmt = d + " = powf(" + A + ", " + B + ");"; break;
            }
        }
        return stmt;
    };
    // Four consecutive samples at once, instruction by instruction (all tables live on every lane): the reference's order at a
    // granularity of four frames instead of a whole span.  Every temp is a register quad; modules with a four-sample form
    // (sample4: branch-free sines, exact divisions and PCM taps arranged so that the four evaluations overlap) use it.  The
    // operands of the four frames are loaded before the first instruction, the results stored after the last.
    // (decided below, once it is known whether the graph is cut in time: body4 is only called after that)
    bool pipelined = (c.voices + 31u) / 32u < 2u * 592u && !getenv("ZCU_GRAPH_NO_PIPELINE");
    auto body4 = [&](const char* ind) {
        std::string s;
        auto arr = [&](const zcu_graph_ref& r) -> std::string {
            char b[32];
            switch (r.kind) {
                case ZCU_GREF_TEMP: snprintf(b, sizeof b, "t%u_", r.index); break;
                case ZCU_GREF_OUTPUT: snprintf(b, sizeof b, "o%u_", r.index); break;
                case ZCU_GREF_INPUT: snprintf(b, sizeof b, "x%u_", r.index); break;
                default: snprintf(b, sizeof b, "z4_"); break;
            }
            return b;
        };
        auto elem = [&](const zcu_graph_ref& r) -> std::string { return r.kind == ZCU_GREF_NONE ? std::string("0.0f") : arr(r) + "[q]"; };
        appendf(s, "%sconst float z4_[4] = {0.0f, 0.0f, 0.0f, 0.0f};\n%s(void)z4_;\n", ind, ind);
        // Small banks (fewer than about two warps per scheduler): external operands are software-pipelined one batch deep: this batch takes what the previous one loaded (xn / on), and
        // the loads of the next batch are issued before this batch's arithmetic (a bank that puts one warp on a scheduler has
        // nothing else to hide the HBM latency behind; ncu on the FmDetuned graph: 2.5 long-scoreboard stalls per issue)
        // (large banks have the warps to hide it and keep the registers: plain loads)
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i] && !pipelined)
                appendf(s, "%sfloat x%u_[4];\n#pragma unroll\n%sfor (int q = 0; q < 4; q++) x%u_[q] = __ldg(g.in[%u] + base + (size_t)(r + q) * g.V);\n", ind, i,
                        ind, i, i);
        for (uint32_t o = 0; o < c.num_outputs; o++)
            if (!mix && !pipelined)
                appendf(s, "%sfloat o%u_[4];\n#pragma unroll\n%sfor (int q = 0; q < 4; q++) o%u_[q] = g.out[%u][base + (size_t)(r + q) * g.V];\n", ind, o, ind, o, o);
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i] && pipelined) {
                appendf(s, "%sfloat x%u_[4];\n#pragma unroll\n%sfor (int q = 0; q < 4; q++) x%u_[q] = xn%u_[q];\n", ind, i, ind, i, i);
                appendf(s, "%sif (f + r + 8 <= g.frames) {\n#pragma unroll\n%s    for (int q = 0; q < 4; q++) xn%u_[q] = __ldg(g.in[%u] + base + (size_t)(r + 4 + q) * g.V);\n%s}\n",
                        ind, ind, i, i, ind);
            }
        for (uint32_t o = 0; o < c.num_outputs; o++) {
            if (mix) {
                appendf(s, "%sfloat o%u_[4] = {0.0f, 0.0f, 0.0f, 0.0f};\n", ind, o);
            } else if (pipelined) {
                appendf(s, "%sfloat o%u_[4];\n#pragma unroll\n%sfor (int q = 0; q < 4; q++) o%u_[q] = on%u_[q];\n", ind, o, ind, o, o);
                appendf(s, "%sif (f + r + 8 <= g.frames) {\n#pragma unroll\n%s    for (int q = 0; q < 4; q++) on%u_[q] = g.out[%u][base + (size_t)(r + 4 + q) * g.V];\n%s}\n",
                        ind, ind, o, o, ind);
            }
        }
        if (pipelined) appendf(s, "%sxn_f = f + r + 8 <= g.frames ? f + r + 4 : 0xFFFFFFFFu;\n", ind);
        for (uint32_t t = 0; t < c.num_temps; t++) appendf(s, "%sfloat t%u_[4] = {t%u, t%u, t%u, t%u};\n", ind, t, t, t, t, t);
        for (uint32_t i = 0; i < n; i++) {
            const zcu_graph_instr& x = ins[i];
            if (x.op == ZCU_GI_CALL && !fields[x.field].has_run) {
                // inputs that alias the output are copied first (a four-sample form may write before it has read everything)
                std::string a[3];
                for (int q = 0; q < 3; q++) {
                    a[q] = arr(x.in[q]);
                    if (x.in[q].kind != ZCU_GREF_NONE && x.in[q].kind == x.out.kind && x.in[q].index == x.out.index) {
                        appendf(s, "%sconst float c%u_%d_[4] = {%s[0], %s[1], %s[2], %s[3]};\n", ind, i, q, a[q].c_str(), a[q].c_str(), a[q].c_str(), a[q].c_str());
                        char b[32];
                        snprintf(b, sizeof b, "c%u_%d_", i, q);
                        a[q] = b;
                    }
                }
                appendf(s, "%sSample4<decltype(m%u)>::call(m%u, %s, %s, %s, %s);\n", ind, x.field, x.field, arr(x.out).c_str(), a[0].c_str(), a[1].c_str(),
                        a[2].c_str());
            } else {
                const std::string stmt = statement(x, arr(x.out) + "[q]", elem(x.in[0]), elem(x.in[1]), elem(x.in[2]), true);
                appendf(s, "#pragma unroll\n%sfor (int q = 0; q < 4; q++) { %s }\n", ind, stmt.c_str());
            }
        }
        for (uint32_t o = 0; o < c.num_outputs; o++) {
            if (mix)
                appendf(s, "#pragma unroll\n%sfor (int q = 0; q < 4; q++) O[j + r + q][lane] = o%u_[q];\n", ind, o);
            else
                appendf(s, "#pragma unroll\n%sfor (int q = 0; q < 4; q++) g.out[%u][base + (size_t)(r + q) * g.V] = o%u_[q];\n", ind, o, o);
        }
        for (uint32_t t = 0; t < c.num_temps; t++) appendf(s, "%st%u = t%u_[3];\n", ind, t, t);
        return s;
    };
    // the per-sample body; live[k] = expression telling whether table k is live on this lane ("true" in the all-live copy)
    auto body = [&](bool all_live, const char* ind, const char* row, bool emit = true, bool silent = false) {
        std::string s;
        auto live = [&](uint32_t k) -> std::string {
            if (all_live) return "true";
            char b[16];
            snprintf(b, sizeof b, "a%u", k);
            return b;
        };
        auto any_live = [&](uint32_t mask) -> std::string {
            if (all_live) return "true";
            std::string e;
            for (uint32_t k = 0; k < K; k++)
                if (mask >> k & 1u) e += (e.empty() ? "" : " || ") + live(k);
            return e.empty() ? "false" : e;
        };
        if ((!mix && emit) || c.num_inputs) appendf(s, "%sconst size_t idx = base + (size_t)%s * g.V;\n", ind, row);
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i]) appendf(s, "%sconst float x%u = valid ? __ldg(g.in[%u] + idx) : 0.0f;\n", ind, i, i);
        for (uint32_t o = 0; o < c.num_outputs; o++) {
            if (mix || !emit)
                appendf(s, "%sfloat o%u = 0.0f;\n", ind, o);
            else
                appendf(s, "%sfloat o%u = (%s) ? g.out[%u][idx] : 0.0f;\n", ind, o, any_live(out_tables[o]).c_str(), o);
        }
        for (uint32_t i = 0; i < n; i++) {
            const zcu_graph_instr& x = ins[i];
            const std::string stmt = statement(x, ref(x.out), ref(x.in[0]), ref(x.in[1]), ref(x.in[2]), all_live, silent);
            if (all_live)
                appendf(s, "%s%s\n", ind, stmt.c_str());
            else
                appendf(s, "%sif (a%u) { %s }\n", ind, x.span_table, stmt.c_str());
        }
        for (uint32_t o = 0; o < c.num_outputs; o++) {
            if (!emit)
                appendf(s, "%s(void)o%u;\n", ind, o);
            else if (mix)
                appendf(s, "%sO[j + %s][lane] = o%u;\n", ind, row, o);
            else
                appendf(s, "%sif (%s) g.out[%u][idx] = o%u;\n", ind, any_live(out_tables[o]).c_str(), o, o);
        }
        return s;
    };

    // time segments: every module must be able to skip ahead exactly from its own params (FieldInfo::skippable)
    // ... or be a Filter, which a segment starts from a warm-up over the frames before it (tolerance mode only: the launcher
    // keeps such a graph in one segment otherwise)
    bool segmentable = c.num_fields > 0, warm_any = false;
    for (uint32_t f = 0; f < c.num_fields; f++) {
        fields[f].warmable = fields[f].kind == ZCU_MOD_FILTER;
        segmentable = segmentable && (fields[f].skippable || fields[f].warmable);
        warm_any = warm_any || fields[f].warmable;
    }
    warm_any = warm_any && segmentable;
    // a graph that is cut in time exactly fills the machine with its segments whatever the voice count: plain operand loads
    // (measured on the SamplerVoice graph, 2048 voices x 46 segments: 1.74 ms plain, 1.87 ms pipelined)
    if (segmentable && !warm_any) pipelined = false;
    if (segmentable_out) *segmentable_out = segmentable;
    if (needs_tolerance_out) *needs_tolerance_out = warm_any;
    if (filter_buffered_out) {  // per field: a Filter whose cutoff or resonance comes from a buffer (its decay is not known from the params)
        filter_buffered_out->assign(c.num_fields, 0);
        for (uint32_t f = 0; f < c.num_fields; f++)
            (*filter_buffered_out)[f] = fields[f].warmable && fields[f].type != "FilterMod<false, false>" ? 1 : 0;
    }

    src.clear();
    src += "// generated by libzangcu (zcu_graph.cu) from a zcu_graph_instr list\n#include \"zcu_graph_rt.cuh\"\nusing namespace zcu;\n\n";
    src += "extern \"C\" __global__ void __launch_bounds__(128) zcu_graph_kernel(const GraphArgs g) {\n";
    src += "    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;\n";
    src += "    const uint32_t v = blockIdx.x * 128u + threadIdx.x;\n";
    src += "    const bool valid = v < g.V;\n";
    src += "    const uint32_t FULL = 0xffffffffu;\n";
    src += "    (void)lane; (void)warp;\n";
    if (mix) {
        appendf(src, "    constexpr int C = %u;\n", c.channels);
        src += "    extern __shared__ float smem[];\n";
        src += "    float(*O)[kMixTileStride] = reinterpret_cast<float(*)[kMixTileStride]>(smem + (size_t)warp * kMixTileFrames * kMixTileStride);\n";
        src += "    float(*gn)[32] = reinterpret_cast<float(*)[32]>(smem + (size_t)4 * kMixTileFrames * kMixTileStride + (size_t)warp * 2 * 32);\n";
        src += "    for (int c = 0; c < C; c++) gn[c][lane] = valid ? g.gains[(size_t)v * C + c] : 0.0f;\n";
        src += "    __syncwarp();\n";
        src += "    const uint32_t gwarp = blockIdx.x * 4u + warp;\n";
    }
    if (warm_any) {
        // a Filter's warm-up window that fails its check is lengthened and the segment starts over (W == F_begin: from the
        // loaded state, which is exact)
        src += "    uint32_t W = g.replay;\nretry:;\n";
    }
    for (uint32_t k = 0; k < K; k++) appendf(src, "    SpanCursor c%u;\n    c%u.init(g.tab[%u], v, valid);\n", k, k, k);
    for (uint32_t f = 0; f < c.num_fields; f++) {
        appendf(src, "    %s m%u;\n", fields[f].type.c_str(), f);
        appendf(src, "    ScratchOf<decltype(m%u)>::type s%u_;\n    ScratchOf<decltype(m%u)>::bind(m%u, s%u_);\n", f, f, f, f, f);
    }
    src += "    if (valid) {\n";
    for (uint32_t f = 0; f < c.num_fields; f++) appendf(src, "        m%u.load(g.pa[%u], v);\n", f, f);
    src += "    }\n";
    for (uint32_t t = 0; t < c.num_temps; t++) appendf(src, "    float t%u = 0.0f;\n", t);
    if (segmentable) {
        // this CTA's segment of the frame axis; the modules skip ahead to F_play (closed forms), the frames [F_play, F_begin)
        // are then run silently (a Decimator re-latches its held value there), [F_begin, F_end) are painted
        src += "    const uint32_t F_begin = blockIdx.y * g.seg_frames;\n";
        src += "    if (F_begin >= g.frames) return;\n";
        src += "    const uint32_t F_end = min(g.frames, F_begin + g.seg_frames);\n";
        src += warm_any ? "    const uint32_t F_play = F_begin > W ? F_begin - W : 0u;\n" : "    const uint32_t F_play = F_begin > g.replay ? F_begin - g.replay : 0u;\n";
        src += "    if (F_play > 0 && valid) {\n";
        for (uint32_t k = 0; k < K; k++) {
            // a table whose modules are all reset by a call with a new note id (the Sampler: t = 0) is walked from the voice's
            // last such call before the segment only (zcu_paint_kernel.cuh, M::RESETS): the span headers are scanned first
            bool resets = false, others = false;
            for (uint32_t f = 0; f < c.num_fields; f++)
                if (fields[f].table == k) (fields[f].kind == ZCU_MOD_SAMPLER ? resets : others) = true;
            if (resets && !others) {
                appendf(src, "        {\n            uint32_t kr = c%u.k;\n            for (uint32_t kk = c%u.k; kk < c%u.ke; kk++) {\n"
                             "                const uint4 sp = *reinterpret_cast<const uint4*>(g.tab[%u].spans + kk);\n"
                             "                if (sp.x >= F_play) break;\n                if ((sp.z & ZCU_SPAN_NOTE_ID_CHANGED)", k, k, k, k);
                for (uint32_t f = 0; f < c.num_fields; f++)
                    if (fields[f].table == k) appendf(src, " && m%u.resets(g.pa[%u], sp.w)", f, f);
                appendf(src, ") kr = kk;\n            }\n            if (kr != c%u.k) {\n                c%u.k = kr;\n                c%u.fetch(g.tab[%u]);\n            }\n        }\n",
                        k, k, k, k);
            }
            appendf(src, "        while (c%u.s_end <= F_play) {\n", k);
            for (uint32_t f = 0; f < c.num_fields; f++)
                if (fields[f].table == k)
                    appendf(src, "            m%u.begin_span(g.pa[%u], c%u.s_param, c%u.s_flags, c%u.s_end - c%u.s_start);\n"
                                 "            m%u.advance(g.pa[%u], v, c%u.s_start, c%u.s_end - c%u.s_start);\n            m%u.end_span(g.pa[%u]);\n",
                            f, f, k, k, k, k, f, f, k, k, k, f, f);
            appendf(src, "            c%u.next(g.tab[%u]);\n        }\n", k, k);
            appendf(src, "        if (c%u.s_start < F_play) {\n", k);
            for (uint32_t f = 0; f < c.num_fields; f++)
                if (fields[f].table == k)
                    appendf(src, "            m%u.begin_span(g.pa[%u], c%u.s_param, c%u.s_flags, c%u.s_end - c%u.s_start);\n"
                                 "            m%u.advance(g.pa[%u], v, c%u.s_start, F_play - c%u.s_start);\n",
                            f, f, k, k, k, k, f, f, k, k);
            appendf(src, "            c%u.active = true;\n        }\n", k);
        }
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].kind == ZCU_MOD_NOISE) appendf(src, "        m%u.advance_done();  // the draws skipped so far, as one jump\n", f);
        src += "    }\n";
        if (warm_any) {
            src += "    if (F_play > 0) {\n";
            for (uint32_t f = 0; f < c.num_fields; f++)
                if (fields[f].warmable)
                    appendf(src, "        m%u.warm_begin_graph(!valid || (c%u.k == g.tab[%u].off[v] && !c%u.active));\n", f, fields[f].table, fields[f].table,
                            fields[f].table);
            src += "    }\n";
        }
    } else {
        src += "    const uint32_t F_begin = 0, F_end = g.frames, F_play = 0;\n    (void)F_begin;\n";
    }
    if (mix) src += "    uint32_t j = 0, T0 = F_begin;\n";
    src += "    uint32_t f = F_play;\n";
    {  // external operands loaded ahead of their batch (body4): they outlive a run, so that the next run can start on them
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i]) appendf(src, "    float xn%u_[4] = {0.0f, 0.0f, 0.0f, 0.0f};\n", i);
        if (!mix)
            for (uint32_t o = 0; o < c.num_outputs; o++) appendf(src, "    float on%u_[4] = {0.0f, 0.0f, 0.0f, 0.0f};\n", o);
        src += "    uint32_t xn_f = 0xFFFFFFFFu;  // the frame whose four rows xn / on hold\n    (void)xn_f;\n";
    }
    src += "    while (f < F_end) {\n";
    if (warm_any) {
        src += "        if (f == F_begin && F_play > 0) {  // every Filter must have forgotten the start of its warm-up window\n            bool ok = true;\n";
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].warmable) appendf(src, "            ok = ok && m%u.warm_ok_graph();\n            m%u.track_map = false;\n", f, f);
        src += "            if (!__all_sync(FULL, ok)) {\n                if (lane == 0 && g.err) atomicAdd(g.err + 1, 1u);\n                W = W >= F_begin / 4u ? F_begin : W * 4u;\n                goto retry;\n            }\n        }\n";
    }
    if (segmentable) {
        bool any_dec = false;
        for (uint32_t f = 0; f < c.num_fields; f++) any_dec = any_dec || fields[f].kind == ZCU_MOD_DECIMATOR;
        if (any_dec) {
            src += "        if (f == F_begin && F_begin > 0 && valid) {  // the replay window must have re-latched every Decimator\n";
            for (uint32_t f = 0; f < c.num_fields; f++)
                if (fields[f].kind == ZCU_MOD_DECIMATOR) appendf(src, "            if (m%u.dval_unknown && g.err) atomicExch(g.err, 1u);\n", f);
            src += "        }\n";
        }
    }
    for (uint32_t k = 0; k < K; k++) {
        appendf(src, "        if (c%u.active && f == c%u.s_end) {\n", k, k);
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].table == k) appendf(src, "            m%u.end_span(g.pa[%u]);\n", f, f);
        appendf(src, "            c%u.next(g.tab[%u]);\n        }\n", k, k);
        appendf(src, "        if (!c%u.active && f == c%u.s_start) {\n", k, k);
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].table == k)
                appendf(src, "            m%u.begin_span(g.pa[%u], c%u.s_param, c%u.s_flags, c%u.s_end - c%u.s_start);\n", f, f, k, k, k, k);
        appendf(src, "            c%u.active = true;\n        }\n", k);
    }
    src += "        uint32_t dist = c0.boundary() - f;\n";
    for (uint32_t k = 1; k < K; k++) appendf(src, "        dist = min(dist, c%u.boundary() - f);\n", k);
    src += "        const bool silent = f < F_begin;  // warp-uniform\n";
    src += "        const uint32_t limit = silent ? F_begin : F_end;\n";
    if (mix) src += "        if (!silent) dist = min(dist, (uint32_t)kMixTileFrames - j);\n";
    // modules with a run interface: no phase change of theirs inside a run (painter_safe_run's margin covers 256 samples);
    // a module that cannot promise a single sample sends the whole warp through the general per-sample path for one frame
    bool any_run = false;
    for (uint32_t f = 0; f < c.num_fields; f++) any_run = any_run || fields[f].has_run;
    src += "        bool careful = false;\n";
    if (any_run) {
        src += "        dist = min(dist, 256u);\n";
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].has_run)
                appendf(src, "        if (c%u.active) {\n            const uint32_t s = m%u.run_limit();\n            careful = careful || s == 0u;\n"
                             "            dist = min(dist, s == 0u ? 1u : s);\n        }\n", fields[f].table, f);
    }
    src += "        const uint32_t run = min(limit - f, __reduce_min_sync(FULL, dist));\n";
    std::string all = "", any = "";
    for (uint32_t k = 0; k < K; k++) {
        appendf(src, "        const bool a%u = c%u.active;\n", k, k);
        char b[16];
        snprintf(b, sizeof b, "a%u", k);
        all += (k ? std::string(" && ") : std::string("")) + b;
        any += (k ? std::string(" || ") : std::string("")) + b;
    }
    src += "        const size_t base = (size_t)f * g.V + v;\n";
    // all tables live on every lane: batches of four samples; the rows of every external buffer 16 frames ahead are pulled
    // into L1 first (banks too small to fill the chip run one warp per scheduler: nothing else would hide the HBM latency)
    // ... or live on some lanes and dead on all the others (voices that have not started yet, the tail of a partial warp):
    // the dead lanes sit the run out
    appendf(src, "        const bool lane_live = %s;\n        const bool lane_idle = !(%s);\n", all.c_str(), any.c_str());
    if (segmentable) {
        src += "        if (silent) {\n            for (uint32_t r = 0; r < run; r++) {\n";
        src += body(false, "                ", "r", false, true);
        src += "            }\n        } else\n";
    }
    src += "        if (__all_sync(FULL, (lane_live && !careful) || lane_idle) && __any_sync(FULL, lane_live)) {\n";
    src += "          if (lane_live) {\n";
    for (uint32_t f = 0; f < c.num_fields; f++)
        if (fields[f].has_run) appendf(src, "            m%u.run_begin();\n", f);
    bool has_delay = false;  // a delay line is read and written slot by slot: no four-sample batches
    for (uint32_t f = 0; f < c.num_fields; f++) has_delay = has_delay || fields[f].is_delay;
    src += "            uint32_t r = 0;\n";
    {  // the first batch's external operands (a graph with a delay line has no batches: the arrays only have to exist)
        const char* first = (has_delay || !pipelined) ? "false" : "run >= 4 && xn_f != f";
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i])
                appendf(src, "            if (%s) {\n#pragma unroll\n                for (int q = 0; q < 4; q++) xn%u_[q] = "
                             "__ldg(g.in[%u] + base + (size_t)q * g.V);\n            }\n", first, i, i);
        if (!mix)
            for (uint32_t o = 0; o < c.num_outputs; o++)
                appendf(src, "            if (%s) {\n#pragma unroll\n                for (int q = 0; q < 4; q++) on%u_[q] = "
                             "g.out[%u][base + (size_t)q * g.V];\n            }\n", first, o, o);
    }
    src += has_delay ? "            for (; false;) {\n" : "            for (; r + 4 <= run; r += 4) {\n";
    {
        std::string pf;
        for (uint32_t i = 0; i < c.num_inputs; i++)
            if (in_used[i]) appendf(pf, "                    prefetch_rows4(g.in[%u] + pidx, g.V);\n", i);
        if (!mix)
            for (uint32_t o = 0; o < c.num_outputs; o++) appendf(pf, "                    prefetch_rows4(g.out[%u] + pidx, g.V);\n", o);
        if (!pf.empty()) {
            src += "                if (valid && f + r + kGraphPrefetchRows + 4 <= g.frames) {\n";
            src += "                    const size_t pidx = base + (size_t)(r + kGraphPrefetchRows) * g.V;\n";
            src += pf;
            src += "                }\n";
        }
    }
    src += "                {\n";
    src += body4("                    ");
    src += "                }\n            }\n            for (; r < run; r++) {\n";
    src += body(true, "                ", "r");
    src += "            }\n";
    for (uint32_t f = 0; f < c.num_fields; f++)
        if (fields[f].has_run) appendf(src, "            m%u.run_end();\n", f);
    if (mix)
        src += "          } else {\n            for (uint32_t r = 0; r < run; r++) O[j + r][lane] = 0.0f;\n          }\n";
    else
        src += "          }\n";
    if (mix)
        src += "        } else {\n            for (uint32_t r = 0; r < run; r++) {\n";
    else
        appendf(src, "        } else if (%s) {\n            for (uint32_t r = 0; r < run; r++) {\n", any.c_str());
    src += body(false, "                ", "r");
    src += "            }\n        }\n";
    src += "        f += run;\n";
    if (mix) {
        src += "        if (!silent) j += run;\n";
        src += "        if (j == (uint32_t)kMixTileFrames || (f == F_end && j > 0)) {\n";
        src += "            __syncwarp();\n";
        src += "            mix_tile_store<C>(O, gn, j, g.partial + ((size_t)gwarp * g.frames + T0) * C, lane);\n";
        src += "            __syncwarp();\n";
        src += "            T0 += j;\n            j = 0;\n        }\n";
    }
    src += "    }\n";
    src += "    if (!valid || F_end < g.frames) return;  // only the last segment carries the state out\n";
    for (uint32_t k = 0; k < K; k++) {
        appendf(src, "    if (c%u.active) {\n", k);
        for (uint32_t f = 0; f < c.num_fields; f++)
            if (fields[f].table == k) appendf(src, "        m%u.end_span(g.pa[%u]);\n", f, f);
        src += "    }\n";
    }
    for (uint32_t f = 0; f < c.num_fields; f++) appendf(src, "    m%u.store(g.pa[%u], v);\n", f, f);
    src += "}\n";
    return true;
}

size_t params_row_bytes(int kind) {
    switch (kind) {
        case ZCU_MOD_SINE_OSC: return sizeof(zcu_sine_params);
        case ZCU_MOD_PULSE_OSC: return sizeof(zcu_pulse_params);
        case ZCU_MOD_TRISAW_OSC: return sizeof(zcu_trisaw_params);
        case ZCU_MOD_NOISE: return sizeof(zcu_noise_params);
        case ZCU_MOD_ENVELOPE: return sizeof(zcu_envelope_params);
        case ZCU_MOD_GATE: return sizeof(zcu_gate_params);
        case ZCU_MOD_FILTER: return sizeof(zcu_filter_params);
        case ZCU_MOD_PORTAMENTO: return sizeof(zcu_portamento_params);
        case ZCU_MOD_CURVE: return sizeof(zcu_curve_params);
        case ZCU_MOD_SAMPLER: return sizeof(zcu_sampler_params);
        case ZCU_MOD_DECIMATOR: return sizeof(zcu_decimator_params);
        case ZCU_MOD_DISTORTION: return sizeof(zcu_distortion_params);
        case ZCU_MOD_CYCLE: return sizeof(zcu_cycle_params);
        case ZCU_MOD_FM_OSCILLATOR: return sizeof(zcu_fm_oscillator_params);
        case ZCU_MOD_FILTERED_ECHOES: return sizeof(zcu_echoes_params);
        case ZCU_MOD_SIMPLE_DELAY: return sizeof(uint32_t);
        default: return 0;
    }
}

}  // namespace

struct zcu_graph {
    zcu_ctx* ctx = nullptr;
    zcu_graph_config cfg{};
    std::vector<int32_t> kinds;
    std::string source;
    cudaLibrary_t lib = nullptr;
    cudaKernel_t kernel = nullptr;
    size_t smem = 0;
    struct Field {
        zcu_bank* bank = nullptr;
        void* d_params = nullptr;
        size_t params_bytes = 0;
        uint32_t nparams = 0;
        void* d_nodes = nullptr;
        uint32_t nnodes = 0;
        const zcu_sample* sample = nullptr;
        uint32_t replay = 0;     // frames a segment must re-run before its start to re-establish this module's state
        bool seg_ok = true;      // false: these params rule time segments out
        bool guess = false;      // a Filter with buffer-driven coefficients: the window is a guess (segments only when forced)
    };
    std::vector<Field> fields;
    float* d_gains = nullptr;
    float* d_partial = nullptr;
    size_t partial_bytes = 0;
    // time segments
    bool segmentable = false;     // every module can skip ahead exactly, or is a Filter (codegen)
    bool needs_tolerance = false; // ... the latter: time segments only under ZCU_PAINT_TOLERANCE
    uint32_t forced_nseg = 0;     // zcu_graph_set_segments: 0 = sized to the GPU
    int resident_per_sm = 0;      // occupancy of the compiled kernel
    uint32_t last_nseg = 1, last_seg_frames = 0, last_replay = 0;
    uint32_t* d_err = nullptr;
};

// Frames after which the Filter's homogeneous map (FilterMod::hom_step, zcu_modules.cuh) is below 1e-8 in every entry, for
// constant (cutoff, res) as FilterMod::begin_span clamps them; + 12 % + 8 as in warm_estimate; 0xFFFFFFFF if > 2^16.
uint32_t filter_forget_window(float cutoff, float res_param) {
    const float cut = std::min(std::max(cutoff, 0.0f), 1.0f), res = 1.0f - std::min(std::max(res_param, 0.0f), 1.0f);
    auto hom = [&](float& l, float& b) {
        l += cut * b;
        b += cut * (0.0f - b * res - l);
        l += cut * b;
        const float h = 0.0f - b * res - l;
        b += cut * h;
    };
    auto small = [](const double* m) { return std::max(std::max(std::fabs(m[0]), std::fabs(m[1])), std::max(std::fabs(m[2]), std::fabs(m[3]))) <= 1e-8; };
    float c1l = 1.0f, c1b = 0.0f, c2l = 0.0f, c2b = 1.0f;
    hom(c1l, c1b);
    hom(c2l, c2b);
    double p[17][4];
    double a[4] = {c1l, c2l, c1b, c2b};  // row-major 2x2: columns are the images of (1,0) and (0,1)
    int k = 0;
    for (;; k++) {
        if (small(a)) break;
        if (k == 16) return 0xFFFFFFFFu;
        for (int i = 0; i < 4; i++) p[k][i] = a[i];
        const double b[4] = {a[0] * a[0] + a[1] * a[2], a[0] * a[1] + a[1] * a[3], a[2] * a[0] + a[3] * a[2], a[2] * a[1] + a[3] * a[3]};
        for (int i = 0; i < 4; i++) a[i] = b[i];
    }
    double q[4] = {1.0, 0.0, 0.0, 1.0};
    uint32_t n = 0;
    for (int j = k - 1; j >= 0; j--) {
        const double b[4] = {q[0] * p[j][0] + q[1] * p[j][2], q[0] * p[j][1] + q[1] * p[j][3], q[2] * p[j][0] + q[3] * p[j][2], q[2] * p[j][1] + q[3] * p[j][3]};
        if (!small(b)) {
            for (int i = 0; i < 4; i++) q[i] = b[i];
            n += 1u << j;
        }
    }
    return n + 1 + (n >> 3) + 8;
}

void zcu_launch_partial_reduce(zcu_ctx* ctx, const float* partial, float* mix, uint32_t nwarps, size_t n);  // zcu_render.cu

extern "C" {

int zcu_graph_compile_check(const zcu_graph_config* cfg, const zcu_graph_instr* instrs, uint32_t num_instrs, const int32_t* field_kinds, char* log,
                            size_t log_len, size_t* cubin_bytes) {
    if (cubin_bytes) *cubin_bytes = 0;
    if (log && log_len) log[0] = 0;
    if (!cfg) return ZCU_ERR_INVALID;
    std::string src, err;
    auto put = [&](const std::string& s) {
        if (log && log_len) snprintf(log, log_len, "%s", s.c_str());
    };
    if (!codegen(*cfg, instrs, num_instrs, field_kinds, src, err)) {
        put(err);
        return ZCU_ERR_INVALID;
    }
    std::vector<char> cubin;
    if (!nvrtc_compile(src, cubin, err)) {
        put(err);
        return ZCU_ERR_UNSUPPORTED;
    }
    put(err);
    if (cubin_bytes) *cubin_bytes = cubin.size();
    return ZCU_OK;
}

int zcu_graph_create(zcu_ctx* ctx, const zcu_graph_config* cfg, const zcu_graph_instr* instrs, uint32_t num_instrs, const int32_t* field_kinds,
                     zcu_graph** out) {
    if (!ctx || !out) return ZCU_ERR_INVALID;
    *out = nullptr;
    ZCU_REQUIRE(ctx, cfg != nullptr, "null config");
    std::string src, err;
    bool segmentable = false, needs_tolerance = false;
    std::vector<int> filter_buffered;
    if (!codegen(*cfg, instrs, num_instrs, field_kinds, src, err, &segmentable, &needs_tolerance, &filter_buffered))
        return zcu_fail(ctx, ZCU_ERR_INVALID, "zcu_graph_create: %s", err.c_str());
    std::vector<char> cubin;
    if (!nvrtc_compile(src, cubin, err)) return zcu_fail(ctx, ZCU_ERR_UNSUPPORTED, "zcu_graph_create: NVRTC: %s", err.c_str());
    zcu_graph* g = new zcu_graph;
    g->ctx = ctx;
    g->cfg = *cfg;
    g->kinds.assign(field_kinds, field_kinds + cfg->num_fields);
    g->source = src;
    g->segmentable = segmentable;
    g->needs_tolerance = needs_tolerance;
    g->fields.resize(cfg->num_fields);
    for (uint32_t f = 0; f < cfg->num_fields; f++) g->fields[f].guess = filter_buffered[f] != 0;
    cudaError_t e = cudaLibraryLoadData(&g->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e == cudaSuccess) e = cudaLibraryGetKernel(&g->kernel, g->lib, "zcu_graph_kernel");
    if (e != cudaSuccess) {
        zcu_graph_free(g);
        return zcu_fail(ctx, ZCU_ERR_CUDA, "zcu_graph_create: loading the compiled kernel failed: %s", cudaGetErrorString(e));
    }
    if (cfg->output_mode == ZCU_GRAPH_OUT_MIX) {
        g->smem = (size_t)(4 * kMixTileFrames * kMixTileStride + 4 * 2 * 32) * sizeof(float);
        e = cudaFuncSetAttribute((const void*)g->kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g->smem);
        if (e != cudaSuccess) {
            zcu_graph_free(g);
            return zcu_fail(ctx, ZCU_ERR_CUDA, "zcu_graph_create: shared-memory attribute: %s", cudaGetErrorString(e));
        }
    }
    if (segmentable) {
        // [0] a segment could not establish its start state, [1] warm-up windows that had to be lengthened
        if (cudaMalloc((void**)&g->d_err, 2 * sizeof(uint32_t)) != cudaSuccess || cudaMemsetAsync(g->d_err, 0, 2 * sizeof(uint32_t), ctx->stream) != cudaSuccess) {
            zcu_graph_free(g);
            return zcu_fail(ctx, ZCU_ERR_OOM, "zcu_graph_create: status word allocation failed");
        }
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g->resident_per_sm, (const void*)g->kernel, 128, g->smem) != cudaSuccess || g->resident_per_sm < 1) {
            cudaGetLastError();
            g->resident_per_sm = 4;
        }
    }
    *out = g;
    return ZCU_OK;
}

int zcu_graph_set_segments(zcu_graph* g, uint32_t segments) {
    if (!g) return ZCU_ERR_INVALID;
    g->forced_nseg = segments;
    return ZCU_OK;
}

int zcu_graph_segment_info(zcu_ctx* ctx, zcu_graph* g, uint32_t* segments, uint32_t* segment_frames, uint32_t* replay_frames, uint32_t* failed) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr, "null graph");
    if (segments) *segments = g->last_nseg;
    if (segment_frames) *segment_frames = g->last_seg_frames;
    if (replay_frames) *replay_frames = g->last_replay;
    if (failed) {
        *failed = 0;
        if (g->d_err) {
            ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            ZCU_CUDA(ctx, cudaMemcpy(failed, g->d_err, sizeof(uint32_t), cudaMemcpyDeviceToHost));
            ZCU_CUDA(ctx, cudaMemset(g->d_err, 0, sizeof(uint32_t)));
        }
    }
    return ZCU_OK;
}

int zcu_graph_warm_retries(zcu_ctx* ctx, zcu_graph* g, uint32_t* retries) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && retries != nullptr, "null argument");
    *retries = 0;
    if (g->d_err) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ZCU_CUDA(ctx, cudaMemcpy(retries, g->d_err + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost));
        ZCU_CUDA(ctx, cudaMemset(g->d_err + 1, 0, sizeof(uint32_t)));
    }
    return ZCU_OK;
}

void zcu_graph_free(zcu_graph* g) {
    if (!g) return;
    if (g->ctx) cudaStreamSynchronize(g->ctx->stream);
    for (auto& f : g->fields) {
        if (f.d_params) cudaFree(f.d_params);
        if (f.d_nodes) cudaFree(f.d_nodes);
    }
    if (g->d_gains) cudaFree(g->d_gains);
    if (g->d_partial) cudaFree(g->d_partial);
    if (g->d_err) cudaFree(g->d_err);
    if (g->lib) cudaLibraryUnload(g->lib);
    delete g;
}

const char* zcu_graph_source(const zcu_graph* g) { return g ? g->source.c_str() : ""; }

int zcu_graph_set_field(zcu_ctx* ctx, zcu_graph* g, uint32_t field, zcu_bank* bank, const void* params_host, size_t row_bytes, uint32_t nparams) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && field < g->cfg.num_fields, "null graph or field out of range");
    const int kind = g->kinds[field];
    const size_t ssz = zcu_bank_state_size(kind);
    if (ssz) {
        ZCU_REQUIRE(ctx, bank != nullptr, "this module needs a bank");
        ZCU_REQUIRE(ctx, bank->kind == kind && bank->voices == g->cfg.voices, "bank kind / voice count mismatch");
        if (kind == ZCU_MOD_SIMPLE_DELAY || kind == ZCU_MOD_FILTERED_ECHOES) ZCU_REQUIRE(ctx, bank->line != nullptr, "bank has no delay line");
    }
    zcu_graph::Field& f = g->fields[field];
    f.bank = bank;
    if (kind == ZCU_MOD_SIMPLE_DELAY && !params_host) {  // SimpleDelay.Params holds nothing but the input slice
        static const uint32_t zero_row = 0;
        params_host = &zero_row;
        row_bytes = sizeof(uint32_t);
        nparams = 1;
    }
    ZCU_REQUIRE(ctx, params_host != nullptr && nparams > 0, "params required");
    ZCU_REQUIRE(ctx, row_bytes == params_row_bytes(kind), "row_bytes is not the sizeof of this module's params struct");
    f.replay = 0;
    f.seg_ok = true;
    if (kind == ZCU_MOD_FILTER && !f.guess) {
        // constant coefficients per span: the window after which a segment's warm-up has forgotten its start follows from the
        // rows (the largest over the rows; a row that needs more than 16384 frames rules segments out)
        const zcu_filter_params* p = (const zcu_filter_params*)params_host;
        for (uint32_t i = 0; i < nparams; i++) {
            if (p[i].type == ZCU_FILTER_BYPASS) continue;
            const uint32_t w = filter_forget_window(p[i].cutoff, p[i].res);
            if (w > 16384u)
                f.seg_ok = false;
            else
                f.replay = std::max(f.replay, w);
        }
    }
    if (kind == ZCU_MOD_DECIMATOR) {
        // a time segment re-latches the held value by running the frames before its start again: the window must hold a
        // latch for every row (Decimator.zig:40-55: one latch per 1/ratio samples); a row that paints nothing (fake rate <= 0)
        // leaves a value of unbounded age behind, so it rules segments out
        const zcu_decimator_params* p = (const zcu_decimator_params*)params_host;
        for (uint32_t i = 0; i < nparams; i++) {
            if (!(p[i].fake_sample_rate > 0.0f)) {
                f.seg_ok = false;
            } else if (p[i].fake_sample_rate < p[i].sample_rate) {
                const float ratio = p[i].fake_sample_rate / p[i].sample_rate;
                const double need = std::ceil(1.0 / (double)ratio) + 2.0;
                if (need > 65536.0)
                    f.seg_ok = false;
                else
                    f.replay = std::max(f.replay, (uint32_t)need);
            }
        }
    }
    std::vector<DistortionDerived> derived;
    const void* src = params_host;
    size_t bytes = row_bytes * nparams;
    if (kind == ZCU_MOD_DISTORTION) {  // per-call host scalars of Distortion.zig:41-46 (pow, atan), as zcu_paint_distortion
        const zcu_distortion_params* p = (const zcu_distortion_params*)params_host;
        derived.resize(nparams);
        for (uint32_t i = 0; i < nparams; i++) {
            const float gain1 = powf(2.0f, p[i].ingain * 8.0f - 2.0f);
            derived[i].type = p[i].type;
            derived[i].gain1 = gain1;
            derived[i].gain2 = p[i].type == ZCU_DIST_OVERDRIVE ? p[i].outgain / atanf(gain1) : p[i].outgain;
            derived[i].offs = gain1 * p[i].offset;
        }
        src = derived.data();
        bytes = derived.size() * sizeof(DistortionDerived);
    }
    if (f.params_bytes < bytes) {
        ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (f.d_params) cudaFree(f.d_params);
        f.d_params = nullptr;
        f.params_bytes = 0;
        ZCU_CUDA(ctx, cudaMalloc(&f.d_params, bytes));
        f.params_bytes = bytes;
    }
    ZCU_CUDA(ctx, cudaMemcpyAsync(f.d_params, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // params_host / derived may be temporaries
    f.nparams = nparams;
    return ZCU_OK;
}

int zcu_graph_set_curve_nodes(zcu_ctx* ctx, zcu_graph* g, uint32_t field, const zcu_curve_node* nodes_host, uint32_t nnodes) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && field < g->cfg.num_fields && g->kinds[field] == ZCU_MOD_CURVE, "not a Curve field");
    ZCU_REQUIRE(ctx, nodes_host != nullptr && nnodes > 0, "nodes required");
    zcu_graph::Field& f = g->fields[field];
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (f.d_nodes) cudaFree(f.d_nodes);
    f.d_nodes = nullptr;
    ZCU_CUDA(ctx, cudaMalloc(&f.d_nodes, (size_t)nnodes * sizeof(zcu_curve_node)));
    ZCU_CUDA(ctx, cudaMemcpyAsync(f.d_nodes, nodes_host, (size_t)nnodes * sizeof(zcu_curve_node), cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    f.nnodes = nnodes;
    return ZCU_OK;
}

int zcu_graph_set_sample(zcu_ctx* ctx, zcu_graph* g, uint32_t field, const zcu_sample* sample) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && field < g->cfg.num_fields && g->kinds[field] == ZCU_MOD_SAMPLER, "not a Sampler field");
    ZCU_REQUIRE(ctx, sample != nullptr, "null sample");
    g->fields[field].sample = sample;
    return ZCU_OK;
}

int zcu_graph_set_gains(zcu_ctx* ctx, zcu_graph* g, const float* gains_host) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && g->cfg.output_mode == ZCU_GRAPH_OUT_MIX, "not a MIX graph");
    ZCU_REQUIRE(ctx, gains_host != nullptr, "null gains");
    const size_t bytes = (size_t)g->cfg.voices * g->cfg.channels * sizeof(float);
    if (!g->d_gains) ZCU_CUDA(ctx, cudaMalloc((void**)&g->d_gains, bytes));
    ZCU_CUDA(ctx, cudaMemcpyAsync(g->d_gains, gains_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZCU_OK;
}

int zcu_graph_paint(zcu_ctx* ctx, zcu_graph* g, const zcu_spans* const* tables, zcu_buf* const* outputs, const zcu_buf* const* inputs) {
    if (!ctx) return ZCU_ERR_INVALID;
    ZCU_REQUIRE(ctx, g != nullptr && tables != nullptr && outputs != nullptr, "null argument");
    const zcu_graph_config& c = g->cfg;
    const bool mix = c.output_mode == ZCU_GRAPH_OUT_MIX;
    ZCU_REQUIRE(ctx, tables[0] != nullptr, "null span table");
    const uint32_t frames = tables[0]->frames;
    GraphArgs a;
    memset(&a, 0, sizeof a);
    a.V = c.voices;
    a.frames = frames;
    for (uint32_t k = 0; k < c.num_span_tables; k++) {
        ZCU_REQUIRE(ctx, tables[k] != nullptr, "null span table");
        ZCU_REQUIRE(ctx, tables[k]->voices == c.voices && tables[k]->frames == frames, "span tables must share the block's shape");
        a.tab[k].off = tables[k]->d_off;
        a.tab[k].spans = tables[k]->d_spans;
    }
    for (uint32_t o = 0; o < c.num_outputs; o++) {
        ZCU_REQUIRE(ctx, outputs[o] != nullptr, "null output");
        if (mix)
            ZCU_REQUIRE(ctx, outputs[o]->frames == frames && outputs[o]->voices == c.channels, "MIX output must be [frames][channels]");
        else
            ZCU_REQUIRE(ctx, outputs[o]->frames == frames && outputs[o]->voices == c.voices, "output shape mismatch");
        a.out[o] = outputs[o]->d;
    }
    ZCU_REQUIRE(ctx, c.num_inputs == 0 || inputs != nullptr, "null inputs");
    for (uint32_t i = 0; i < c.num_inputs; i++) {
        ZCU_REQUIRE(ctx, inputs[i] != nullptr && inputs[i]->frames == frames && inputs[i]->voices == c.voices, "input shape mismatch");
        a.in[i] = inputs[i]->d;
    }
    a.nfields = c.num_fields;
    for (uint32_t f = 0; f < c.num_fields; f++) {
        const zcu_graph::Field& fd = g->fields[f];
        const int kind = g->kinds[f];
        ZCU_REQUIRE(ctx, fd.d_params != nullptr, "a field has no params (zcu_graph_set_field)");
        // every span of the field's table must name an existing params row: checked against the largest row of ANY table
        // (the generator knows which table drives the field; the bound is kept simple and conservative here)
        PaintArgs& p = a.pa[f];
        p.V = c.voices;
        p.frames = frames;
        p.params = fd.d_params;
        p.state = fd.bank ? fd.bank->d : nullptr;
        p.line = fd.bank ? fd.bank->line : nullptr;
        p.delay_samples = fd.bank ? fd.bank->delay_samples : 0;
        p.line_ring = 0;
        if (kind == ZCU_MOD_CURVE) {
            ZCU_REQUIRE(ctx, fd.d_nodes != nullptr, "a Curve field has no nodes (zcu_graph_set_curve_nodes)");
            p.nodes = (const zcu_curve_node*)fd.d_nodes;
        }
        if (kind == ZCU_MOD_SAMPLER) {
            ZCU_REQUIRE(ctx, fd.sample != nullptr, "a Sampler field has no sample (zcu_graph_set_sample)");
            p.smp.data = fd.sample->d;
            p.smp.len = fd.sample->len;
            p.smp.num_channels = fd.sample->num_channels;
            p.smp.sample_rate = fd.sample->sample_rate;
            p.smp.format = fd.sample->format;
            zcu_sample_desc_finish(p.smp);
        }
    }
    if (frames == 0) return ZCU_OK;
    const uint32_t grid = (c.voices + 127) / 128;
    if (mix) {
        ZCU_REQUIRE(ctx, g->d_gains != nullptr, "MIX graph has no gains (zcu_graph_set_gains)");
        const size_t need = (size_t)grid * 4 * frames * c.channels * sizeof(float);
        if (g->partial_bytes < need) {
            ZCU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (g->d_partial) cudaFree(g->d_partial);
            g->d_partial = nullptr;
            g->partial_bytes = 0;
            ZCU_CUDA(ctx, cudaMalloc((void**)&g->d_partial, need));
            g->partial_bytes = need;
        }
        a.partial = g->d_partial;
        a.gains = g->d_gains;
        a.C = c.channels;
        a.out[0] = nullptr;
    }
    // time segments: as many as fill the resident CTAs once (each at least 512 frames, a multiple of the mix tile)
    uint32_t nseg = 1, seg_frames = frames, replay = 0;
    if (g->segmentable) {
        bool ok = true;
        for (const auto& fd : g->fields) {
            ok = ok && fd.seg_ok;
            replay = std::max(replay, fd.replay);
        }
        if (g->needs_tolerance) {
            // a Filter starts a segment from a warm-up window: only in tolerance mode; the window is a first guess that the
            // kernel verifies (and lengthens where it was too short)
            ok = ok && ctx->paint_mode == ZCU_PAINT_TOLERANCE;
            for (const auto& fd : g->fields) {
                if (!fd.guess) continue;
                // coefficients from a buffer: nothing is known about the decay (a sweep towards a low cutoff forgets slowly and
                // the retries then cost more than the segments bring): segments only on request, from a first guess
                ok = ok && g->forced_nseg != 0;
                uint32_t w = 1024;
                if (const char* e = getenv("ZCU_GRAPH_WARM")) w = (uint32_t)std::max(1, atoi(e));  // tuning
                replay = std::max(replay, w);
            }
        }
        bool has_noise = false;
        for (uint32_t f = 0; f < c.num_fields; f++) has_noise = has_noise || g->kinds[f] == ZCU_MOD_NOISE;
        if (has_noise) {  // jumps are tabulated for fewer than 16^6 draws per call
            ok = ok && frames < (1u << 24);
            if (ok) {
                int rc = zcu_ensure_xjump(ctx);
                if (rc) return rc;
                for (uint32_t f = 0; f < c.num_fields; f++)
                    if (g->kinds[f] == ZCU_MOD_NOISE) {
                        a.pa[f].xjump = (const ulonglong2*)ctx->xjump;
                        a.pa[f].xjump_resync = ctx->xjump_resync;
                    }
            }
        }
        if (ok) {
            nseg = g->forced_nseg ? g->forced_nseg : std::max(1u, (uint32_t)g->resident_per_sm * (uint32_t)ctx->sm_count / grid);
            seg_frames = (frames + nseg - 1) / nseg;
            if (!g->forced_nseg) seg_frames = std::max(seg_frames, std::min(frames, 512u));
            seg_frames = std::max(seg_frames, replay + 1);
            if (g->needs_tolerance && !g->forced_nseg) seg_frames = std::max(seg_frames, 2 * replay);  // warm-up at most a third of a segment's work
            seg_frames = (seg_frames + kMixTileFrames - 1) / kMixTileFrames * kMixTileFrames;
            nseg = (frames + seg_frames - 1) / seg_frames;
        }
    }
    a.seg_frames = seg_frames;
    a.replay = replay;
    a.err = g->d_err;
    g->last_nseg = nseg;
    g->last_seg_frames = seg_frames;
    g->last_replay = replay;
    void* args[] = {&a};
    ZCU_CUDA(ctx, cudaLaunchKernel((const void*)g->kernel, dim3(grid, nseg), dim3(128), args, g->smem, ctx->stream));
    ctx->launches++;
    if (mix) zcu_launch_partial_reduce(ctx, g->d_partial, outputs[0]->d, grid * 4, (size_t)frames * c.channels);
    ZCU_CUDA(ctx, cudaGetLastError());
    return ZCU_OK;
}

}  // extern "C"
