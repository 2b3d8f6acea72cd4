"""Host logic above the C ABI, on CPU: the Python operator mirror (buffering, key dictionary,
watermark hand-off, drain order) driven with the CPU oracle injected as the engine — a test double;
the product default is the CUDA engine."""
import json
import os

import numpy as np

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.operator import GpuWindowAggOperator
from monitor_systam_flink_quickstart_b200.text import parse_cpu_line, parse_flow_line, iso_local_to_epoch_ms
from oracle_lib import OracleEngine

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_sessions.json"), encoding="utf-8"))
ORACLE = lambda **kw: OracleEngine(**kw)  # noqa: E731


def test_text_parsing_matches_reference_map_functions():
    assert parse_cpu_line("1563452056 10.8.22.1 cpu0 80.5") == (1563452056, "10.8.22.1", "cpu0", 80.5)      # Main.java:21-25
    assert parse_flow_line("2019-08-28T10:00:00 www.163.com 10000") == (1566957600000, "www.163.com", 10000)   # …EventTime.java:33-44
    assert iso_local_to_epoch_ms("2038-01-19T11:14:08") == -2147483648 * 1000                                  # the (int) cast wraps


def test_operator_replays_g7_like_a_print_sink():
    out = []
    op = GpuWindowAggOperator(lambda kind, row: out.append((kind, row)), batch_capacity=4, engine_factory=ORACLE,
                              value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, emit_fired=0,
                              window_size_ms=300000, window_slide_ms=5000, max_out_of_orderness_ms=60000,
                              walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                              walert_div1=1024.0, walert_div2=1024.0, walert_threshold=100.0).open()
    for line in G["G7_event_time"]["input"]:
        t, ch, flow = parse_flow_line(line)
        op.process_element(t, ch, flow)
        op.process_watermark()                       # auto watermark after every record
    assert op.metrics()["late_dropped"] == 1
    printed = [f"({row[0]},{row[2]!r})" for kind, row in out if kind == "window_alert"]
    for want in G["G7_event_time"]["output"]:
        assert want.split("> ")[1] in printed
    assert len(printed) == 60
    # record timestamps are window_end - 1 and come out in timer order
    stamps = [row[1] for kind, row in out if kind == "window_alert"]
    assert stamps == sorted(stamps) and stamps[0] == 1566957605000 - 1
    op.close()
    assert len([1 for kind, _ in out if kind == "window_alert"]) == 60 + 72


def test_operator_buffer_full_pushes_and_rolling_max():
    out = []
    op = GpuWindowAggOperator(lambda kind, row: out.append((kind, row)), batch_capacity=2, engine_factory=ORACLE, rolling_max_enabled=1).open()
    for line in G["G3_rolling_max"]["input"]:
        t, host, _, usage = parse_cpu_line(line)
        op.process_element(t * 1000, host, usage)     # third element forces a push of the first two
    op.close()
    assert [row for kind, row in out if kind == "rolling_max"] == [("10.8.22.1", 80.5), ("10.8.22.1", 80.5), ("10.8.22.1", 99.9)]
