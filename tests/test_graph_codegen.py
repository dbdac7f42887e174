"""Host side of the fused module graphs (no GPU): the generator accepts the three bank bodies of zang_b200/banks.py and NVRTC
compiles them for sm_100a inside libzangcu (embedded headers, --fmad=false); malformed lists are rejected with a message."""
import ctypes as C

import pytest

from zang_b200 import _lib as L
from zang_b200 import banks, engine as E


@pytest.mark.parametrize("cls,mode", [(banks.SfxBank, L.GRAPH_OUT_BUFFERS), (banks.SfxBank, L.GRAPH_OUT_MIX),
                                      (banks.SamplerVoiceBank, L.GRAPH_OUT_BUFFERS), (banks.FmDetunedBank, L.GRAPH_OUT_BUFFERS)])
def test_bank_graphs_compile_for_sm100a(cls, mode):
    b, cfg, fields = cls.graph_list(4096, mode, 2)
    rc, log, cubin_bytes = E.graph_compile_check(cfg, b)
    assert rc == L.OK, log
    assert cubin_bytes > 10000
    assert "error" not in log.lower()
    assert len(fields) == cfg.num_fields


def test_generator_rejects_malformed_lists():
    def check(b, cfg, needle):
        rc, log, n = E.graph_compile_check(cfg, b)
        assert rc == L.ERR_INVALID and n == 0 and needle in log, (rc, log)

    b = E.GraphBuilder()
    b.call(L.MOD_FILTER, E.T(0))
    check(b, E.graph_config(8, b, 1), "needs its input buffer")
    b = E.GraphBuilder()
    b.zero(E.T(2))
    check(b, E.graph_config(8, b, 2), "out must be a temp or an output")
    b = E.GraphBuilder()
    b.zero(E.T(0), table=1)
    check(b, E.graph_config(8, b, 1), "span_table out of range")
    b = E.GraphBuilder()
    b.copy(E.IN(0), E.T(0))
    check(b, E.graph_config(8, b, 1, num_inputs=1), "out must be a temp or an output")
    b = E.GraphBuilder()
    b.multiply(E.OUT(0), E.T(0), E.IN(3))
    check(b, E.graph_config(8, b, 1, num_inputs=1), "operand reference out of range")
    b = E.GraphBuilder()
    b.zero(E.OUT(0))
    cfg = E.graph_config(8, b, 0, output_mode=L.GRAPH_OUT_MIX, channels=3)
    check(b, cfg, "MIX")


def test_delay_block_instructions_are_checked():
    def check(b, cfg, needle):
        rc, log, n = E.graph_compile_check(cfg, b)
        assert rc == L.ERR_INVALID and needle in log, (rc, log)

    b = E.GraphBuilder()
    fd = b.delay_field()
    b.delay_write(fd, E.T(0))
    check(b, E.graph_config(8, b, 1), "without a preceding DELAY_READ")
    b = E.GraphBuilder()
    fd = b.delay_field()
    b.delay_read(fd, E.T(0))
    check(b, E.graph_config(8, b, 1), "has no DELAY_WRITE")
    b = E.GraphBuilder()
    f = b.call(L.MOD_GATE, E.T(0))
    b.delay_read(f, E.T(0))
    check(b, E.graph_config(8, b, 1), "needs a ZCU_MOD_SIMPLE_DELAY field")
    b = E.GraphBuilder()
    fd = b.delay_field()
    b.delay_read(fd, E.T(0))
    b.arith(L.ARITH_POW, E.T(0), 2.0, E.T(0))
    b.arith(L.ARITH_SIN, E.T(0), E.T(0))
    b.delay_write(fd, E.T(0))
    b.addInto(E.OUT(0), E.T(0))
    rc, log, n = E.graph_compile_check(E.graph_config(8, b, 1), b)
    assert rc == L.OK and n > 10000, log


def test_struct_sizes_match_the_header():
    assert C.sizeof(L.GraphRef) == 8 and C.sizeof(L.GraphInstr) == 56 and C.sizeof(L.GraphConfig) == 32


def test_curve_graph_keeps_only_the_node_list_in_local_memory(tmp_path, monkeypatch):
    """A module object that holds a run-time indexed array lives in local memory as a whole (measured on Curve: every
    `pos == sub_end` a local load).  CurveMod's node list is therefore declared beside the module (ScratchOf<M>,
    zcu_modules.cuh); the compiled graph's stack frame must be the list (32 nodes x 8 bytes) and nothing else."""
    import re
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    monkeypatch.setenv("ZCU_GRAPH_DUMP", str(tmp_path))
    b = E.GraphBuilder()
    b.zero(E.T(0))
    b.call(L.MOD_CURVE, E.T(0), ())
    b.addInto(E.OUT(0), E.T(0))
    cfg = E.graph_config(4096, b, 1)
    rc, log, n = E.graph_compile_check(cfg, b)
    assert rc == L.OK, log
    src = (tmp_path / "zcu_graph_kernel.cu").read_text()
    assert "ScratchOf<decltype(m0)>::bind(m0, s0_)" in src
    res = subprocess.run(["cuobjdump", "-res-usage", str(tmp_path / "zcu_graph_kernel.cubin")], capture_output=True, text=True).stdout
    m = re.search(r"STACK:(\d+)", res)
    assert m, res
    assert int(m.group(1)) <= 256 + 32, res
