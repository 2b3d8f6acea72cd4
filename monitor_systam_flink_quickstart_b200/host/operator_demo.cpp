// Replays the reference's event-time session (chapter3/README.md:284-297) through the C++ operator
// mirror and the C ABI, one watermark per record, and prints what a Flink `print()` sink would.
// Exit code 0 iff the two README doubles appear and exactly one record was dropped as late.
#include <cstdio>
#include <cmath>
#include <set>
#include "gpu_window_operator.hpp"

int main() {
    fmq_config cfg; std::memset(&cfg, 0, sizeof(cfg));
    cfg.value_type = FMQ_VAL_I64; cfg.window_enabled = 1; cfg.agg_mask = FMQ_AGG_COUNT | FMQ_AGG_SUM; cfg.emit_fired = 0;
    cfg.window_size_ms = 300000; cfg.window_slide_ms = 5000; cfg.max_out_of_orderness_ms = 60000;
    cfg.walert_source = FMQ_ALERT_SUM; cfg.walert_op = FMQ_CMP_LT; cfg.walert_mul = 8.0; cfg.walert_div0 = 60; cfg.walert_div1 = 1024; cfg.walert_div2 = 1024;
    cfg.walert_threshold = 100.0; cfg.table_capacity = 1024; cfg.pane_ring = 128;
    std::set<double> values; size_t alerts = 0;
    try {
        fmq_host::GpuWindowAggOperator op(cfg, 1024, [&](const fmq_host::WindowResult*, const fmq_host::WindowAlert* a) {
            if (a) { ++alerts; if (values.insert(a->value).second) std::printf("(%s,%.17g)\n", a->key.c_str(), a->value); }
        });
        op.open();
        const int64_t t0 = 1566957600000LL;                      // 2019-08-28T10:00:00+08:00
        const int64_t ts[5] = {t0, t0 + 60000, t0 + 120000, t0 - 3540000, t0 + 360000};
        const int64_t flow[5] = {10000, 100, 100, 100, 100};
        for (int i = 0; i < 5; ++i) { op.processElement<int64_t>(ts[i], "www.163.com", flow[i]); op.processWatermark(FMQ_WM_AUTO); }
        const uint64_t late = op.numLateRecordsDropped();
        op.close();
        const bool ok = values.count(0.0012715657552083333) && values.count(0.0012969970703125) && late == 1 && alerts == 60 + 72;
        std::printf("alerts=%zu late_dropped=%llu %s\n", alerts, (unsigned long long)late, ok ? "OK" : "MISMATCH");
        return ok ? 0 : 1;
    } catch (const std::exception& e) { std::printf("error: %s\n", e.what()); return 2; }
}
