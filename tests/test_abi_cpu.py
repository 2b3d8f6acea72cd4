"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol the
header declares, and refuses to run without a GPU (no silent CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from monitor_systam_flink_quickstart_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "fmq_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fmq_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    names = _declared_symbols()
    assert len(names) >= 20, names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/fmq_b200.h but not exported"
    assert lib.fmq_abi_version() == 1
    assert b"sm_100a" in lib.fmq_build_info()


def test_config_struct_layout_matches_header():
    """sizeof(fmq_config) as the C compiler sees it == ctypes mirror."""
    import subprocess, tempfile
    src = '#include "fmq_b200.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu", sizeof(fmq_config), sizeof(fmq_counters), sizeof(fmq_fired_cols), sizeof(fmq_packed_columns));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c"); open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", exe, c])
        a, b, f, p = map(int, subprocess.check_output([exe]).split())
    assert a == C.sizeof(capi.FmqConfig) and b == C.sizeof(capi.FmqCounters) and f == C.sizeof(capi.FmqFiredCols)
    assert p == C.sizeof(capi.FmqPackedColumns)


def test_bad_arguments_are_rejected_before_touching_a_device():
    lib = capi.load_library()
    h = C.c_void_p()
    cfg = capi.make_config(window_enabled=1, window_size_ms=10, window_slide_ms=3)
    cfg.struct_size = 4
    assert lib.fmq_create(C.byref(cfg), C.byref(h)) == 1
    assert b"struct_size" in lib.fmq_last_error(None)
    cfg = capi.make_config(window_enabled=1, window_size_ms=10, window_slide_ms=3)
    assert lib.fmq_create(C.byref(cfg), C.byref(h)) == 1          # size % slide != 0
    cfg = capi.make_config(window_enabled=1, window_size_ms=10, window_slide_ms=10, max_out_of_orderness_ms=-1)
    assert lib.fmq_create(C.byref(cfg), C.byref(h)) == 1          # chapter3/README.md:356-359
    # entry points that take a handle answer FMQ_ERR_BAD_ARG to a null one (no device is touched)
    pc = capi.FmqPackedColumns(C.sizeof(capi.FmqPackedColumns), 4, 4, 8, 0, None, None, None)
    assert lib.fmq_push_packed(None, C.byref(pc), 0, 0, capi.FMQ_MEM_HOST) == 1
    assert lib.fmq_stage_host(None, C.byref(pc), 1) == 1
    assert lib.fmq_push_local_packed(None, C.byref(pc), 0, capi.FMQ_MEM_HOST) == 1
    assert lib.fmq_combiner_overlap(None, 1) == 1 and lib.fmq_fire_stream(None) is None


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.FmqError) as e:
        capi.Engine(window_enabled=1, window_size_ms=1000, window_slide_ms=1000)
    assert e.value.status == 7 and "no CPU fallback" in str(e.value)


def test_routing_functions_host_side():
    """fmq_flink_key_group / fmq_flink_subtask reproduce golden G8; fmq_owner_of is a pure function."""
    lib = capi.load_library()
    for name, kg, sub in [("10.8.22.1", 78, 2), ("10.8.22.2", 90, 2), ("www.163.com", 58, 1)]:
        h = lib.fmq_java_string_hash(name.encode(), len(name))
        assert lib.fmq_flink_key_group(h, 128) == kg
        assert lib.fmq_flink_subtask(kg, 4, 128) == sub
    owners = [lib.fmq_owner_of(k, 8) for k in range(1000)]
    assert set(owners) == set(range(8)) and owners == [lib.fmq_owner_of(k, 8) for k in range(1000)]
