// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// CPU oracle: a literal, single-threaded restatement of the Flink 1.8.0 operator semantics that
// the reference jobs run on (the reference itself only contributes lambdas; the arithmetic lives
// in the un-vendored dependency org.apache.flink:flink-streaming-java_2.11:1.8.0, pom.xml:19,43-55,
// whose source is NOT on disk — restated from SURVEY.md §8a / Appendix A).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (libfmq_b200.so) never links or calls it.
//
// PARITY PINNING: the reference has no automated tests; the only pins are the README sessions
// G1-G8, reproduced by tests/test_oracle_golden.py.  For everything those sessions
// do not cover (late-drop schedule, sliding fan-out, consecutive-breach rule) parity is UNPINNED
// by the reference and rests on this restatement plus the independent Python model
// (oracle/naive_model.py).
//
// Per-record call chain restated here:
//   WindowOperator.processElement  -> assignWindows -> isWindowLate -> state.add -> register timer
//   WindowOperator.processWatermark-> InternalTimerServiceImpl.advanceWatermark -> onEventTime
//   StreamFilter.processElement    (chapter1/.../Main.java:27-32)
//   StreamGroupedReduce.processElement + ComparableAggregator (chapter2/.../ComputeCpuMax.java:26)
#include <cstdint>
#include <cstring>
#include <cmath>
#include <map>
#include <set>
#include <tuple>
#include <vector>
#include <string>
#include <unordered_map>
#include <algorithm>
#include <limits>
#include "../include/fmq_b200.h"

namespace {

template <typename V> struct Acc { int64_t count; V sum; V mn; V mx; };

// Math.max / Math.min of java.lang.Math for doubles: NaN propagates, -0.0 < +0.0.
inline double java_max(double a, double b) {
    if (a != a) return a;
    if (a == 0.0 && b == 0.0) { return std::signbit(a) ? b : a; }
    return (a >= b) ? a : b;   // b NaN -> (a>=b) false -> b
}
inline double java_min(double a, double b) {
    if (a != a) return a;
    if (a == 0.0 && b == 0.0) { return std::signbit(a) ? a : b; }
    return (a <= b) ? a : b;
}
inline int64_t java_max(int64_t a, int64_t b) { return a >= b ? a : b; }
inline int64_t java_min(int64_t a, int64_t b) { return a <= b ? a : b; }

// Double.compareTo / Long.compareTo (ComparableAggregator uses Comparable#compareTo).
inline int java_compare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    auto bits = [](double d) -> int64_t {
        if (d != d) return 0x7ff8000000000000LL;           // doubleToLongBits canonical NaN
        int64_t r; std::memcpy(&r, &d, 8); return r; };
    int64_t x = bits(a), y = bits(b);
    return x == y ? 0 : (x < y ? -1 : 1);
}
inline int java_compare(int64_t a, int64_t b) { return a < b ? -1 : (a > b ? 1 : 0); }

inline bool cmp(double v, int op, double thr) {
    switch (op) { case FMQ_CMP_GT: return v > thr; case FMQ_CMP_LT: return v < thr;
                  case FMQ_CMP_GE: return v >= thr; default: return v <= thr; }
}
inline double one_if_zero(double d) { return d == 0.0 ? 1.0 : d; }

struct FiredRow { uint64_t key; int64_t wend; int64_t count; uint64_t sum, mn, mx; };
struct WAlert   { uint64_t key; int64_t wend; double value; };
struct EvRow    { int64_t ts; uint64_t key; uint64_t val; };

template <typename V> inline uint64_t bits_of(V v) { uint64_t r; std::memcpy(&r, &v, 8); return r; }

struct OracleBase {
    fmq_config cfg{};
    std::string err;
    virtual ~OracleBase() {}
    virtual int push(const int64_t* ts, const uint64_t* key, const void* val, size_t n, int64_t wm_after) = 0;
    virtual int flush(int64_t wm) = 0;
    std::vector<FiredRow> fired; std::vector<WAlert> walerts, medians; std::vector<EvRow> ealerts, calerts, late;
    std::vector<uint64_t> rolling;   // rolling max output of the last batch
    fmq_counters ctr{};
    virtual int key_state(uint64_t key, int32_t* consec, uint64_t* rolling_bits) = 0;
};

template <typename V>
struct Oracle : OracleBase {
    using WKey = std::pair<uint64_t, int64_t>;                 // (key, window start) = (key, namespace)
    std::map<WKey, Acc<V>> wstate;                             // Heap{Reducing,Aggregating}State
    std::map<WKey, std::vector<V>> wlist;                      // HeapListState: ComputeCpuMiddle.java:34 (process(...))
    std::set<std::tuple<int64_t, uint64_t, int64_t>> timers;   // (timestamp, key, window start)
    std::set<std::tuple<int64_t, uint64_t, int64_t>> cleanup;  // allowedLateness > 0: (maxTimestamp + lateness, key, window start)
    std::set<std::pair<int64_t, uint64_t>> refire;             // (window start, key) that took a record after having fired, this batch
    int64_t current_wm = INT64_MIN;                            // InternalTimerServiceImpl.currentWatermark
    // BoundedOutOfOrdernessTimestampExtractor (chapter3/README.md:342-397)
    int64_t current_max_ts; int64_t last_emitted_wm = INT64_MIN;
    struct KState { bool has_max = false; V mx{}; int32_t consec = 0; };
    std::unordered_map<uint64_t, KState> kstate;               // HeapValueState per key

    explicit Oracle(const fmq_config& c) { cfg = c; current_max_ts = INT64_MIN + cfg.max_out_of_orderness_ms; ctr.watermark = INT64_MIN; }

    // TimeWindow.getWindowStartWithOffset (Appendix A.2), Java % == C++ % (truncating)
    static int64_t window_start(int64_t ts, int64_t off, int64_t w) { return ts - (ts - off + w) % w; }

    void process_element(int64_t ts, uint64_t key, V val) {
        ctr.events_in++;
        // --- chapter1 Main.java:31  value.f2 > 90 ---
        if (cfg.efilter_enabled && cmp((double)val, cfg.efilter_op, cfg.efilter_threshold)) {
            ealerts.push_back({ts, key, bits_of(val)}); ctr.event_alerts++;
        }
        // --- keyed ValueState operators, per-key arrival order, no event-time logic ---
        if (cfg.rolling_max_enabled || cfg.consec_n > 0) {
            KState& st = kstate[key];
            if (cfg.rolling_max_enabled) {
                // StreamGroupedReduce.processElement + ComparableAggregator.reduce (Appendix A.9)
                if (!st.has_max) { st.has_max = true; st.mx = val; }
                else if (!(java_compare(st.mx, val) > 0)) st.mx = val;
                rolling.push_back(bits_of(st.mx));
            }
            if (cfg.consec_n > 0) {
                // SURVEY.md §8 A7-ii (spec fixed there; absent from the reference)
                bool over = cmp((double)val, cfg.consec_op, cfg.consec_threshold);
                int32_t c = over ? st.consec + 1 : 0;
                bool fire = cfg.consec_reset_on_fire ? (c == cfg.consec_n) : (c >= cfg.consec_n);
                if (fire) { calerts.push_back({ts, key, bits_of(val)}); ctr.consec_alerts++; if (cfg.consec_reset_on_fire) c = 0; }
                st.consec = c;
            }
        }
        if (!cfg.window_enabled) return;
        // --- max-timestamp tracking of the watermark assigner (chapter3/README.md:389-396) ---
        if (ts > current_max_ts) current_max_ts = ts;
        // --- WindowOperator.processElement, non-merging branch (Appendix A.3, A.4) ---
        const int64_t size = cfg.window_size_ms, slide = cfg.window_slide_ms, off = cfg.window_offset_ms;
        const int64_t lateness = cfg.allowed_lateness_ms;
        bool skipped = true;
        int64_t last_start = window_start(ts, off, slide);
        for (int64_t s = last_start; s > ts - size; s -= slide) {
            int64_t max_ts = s + size - 1;                       // TimeWindow.maxTimestamp()
            if (max_ts + lateness <= current_wm) continue;       // isWindowLate: cleanupTime(window) <= currentWatermark
            skipped = false;
            auto it = wstate.find({key, s});
            if (it == wstate.end()) {
                wstate[{key, s}] = Acc<V>{1, val, val, val};     // first element stored as-is (A.7)
            } else {
                Acc<V>& a = it->second;
                a.count += 1; a.sum = a.sum + val;               // arrival order; long wraps (two's complement)
                a.mn = java_min(a.mn, val); a.mx = java_max(a.mx, val);
            }
            if (cfg.median_log_capacity > 0) wlist[{key, s}].push_back(val);
            // EventTimeTrigger.onElement: a window whose maxTimestamp has passed FIREs at once (allowedLateness > 0
            // only: otherwise such a window is late).  Here outputs exist at micro-batch boundaries, so the window
            // re-fires once at the end of this batch (see fmq_config.allowed_lateness_ms).
            if (max_ts <= current_wm) refire.insert({s, key});
            else timers.insert({max_ts, key, s});                // registerEventTimeTimer(maxTimestamp)
            if (lateness > 0) cleanup.insert({max_ts + lateness, key, s});   // registerCleanupTimer
        }
        if (skipped && ts + lateness <= current_wm) {                   // numLateRecordsDropped / sideOutput(lateDataOutputTag, element)
            ctr.late_dropped++;
            if (cfg.late_capacity > 0) late.push_back({ts, key, bits_of(val)});
        }
    }

    double alert_value(const Acc<V>& a) const {
        double src;
        switch (cfg.walert_source) {
            case FMQ_ALERT_SUM:   src = (double)a.sum; break;
            case FMQ_ALERT_AVG:   return a.count == 0 ? 0.0 : (double)a.sum / (double)a.count;  // ComputeCpuAvg.java:49
            case FMQ_ALERT_MIN:   src = (double)a.mn; break;
            case FMQ_ALERT_MAX:   src = (double)a.mx; break;
            default:              src = (double)a.count; break;
        }
        // BandwidthMonitorWithEventTime.java:51   f2 * 8.0 / 60 / 1024 / 1024  (left to right)
        return src * one_if_zero(cfg.walert_mul) / one_if_zero(cfg.walert_div0) / one_if_zero(cfg.walert_div1) / one_if_zero(cfg.walert_div2);
    }

    // emitWindowContents for one (key, window): the fired row and the alert evaluated on it
    void emit_window(uint64_t key, int64_t s, const Acc<V>& a) {
        const int64_t wend = s + cfg.window_size_ms;
        ctr.windows_fired++;
        if (cfg.emit_fired) fired.push_back({key, wend, a.count, bits_of(a.sum), bits_of(a.mn), bits_of(a.mx)});
        if (cfg.walert_source != FMQ_ALERT_NONE) {
            double v = alert_value(a);
            if (cmp(v, cfg.walert_op, cfg.walert_threshold)) { walerts.push_back({key, wend, v}); ctr.window_alerts++; }
        }
    }
    // windows that had fired already and took records in this batch: one more firing each, before the next watermark
    void refire_pending() {
        for (const auto& r : refire) {
            auto it = wstate.find({r.second, r.first});
            if (it != wstate.end()) emit_window(r.second, r.first, it->second);
        }
        refire.clear();
    }

    void advance_watermark(int64_t wm) {
        if (wm <= current_wm) return;
        current_wm = wm; ctr.watermark = wm;
        // InternalTimerServiceImpl.advanceWatermark (Appendix A.5)
        while (!timers.empty() && std::get<0>(*timers.begin()) <= wm) {
            auto t = *timers.begin(); timers.erase(timers.begin());
            int64_t tts = std::get<0>(t); uint64_t key = std::get<1>(t); int64_t s = std::get<2>(t);
            int64_t wend = s + cfg.window_size_ms;
            if (tts != wend - 1) continue;                       // EventTimeTrigger.onEventTime
            auto it = wstate.find({key, s});
            if (it == wstate.end()) continue;
            emit_window(key, s, it->second);
            if (cfg.median_log_capacity > 0) {
                // ComputeCpuMiddle.java:36-47: copy, Collections.sort (Double.compareTo), middle element(s)
                std::vector<V>& vals = wlist[{key, s}];
                std::sort(vals.begin(), vals.end(), [](V x, V y) { return java_compare(x, y) < 0; });
                const size_t n = vals.size();
                double med = n == 0 ? 0.0 : (n % 2 != 0 ? (double)vals[n / 2] : ((double)vals[n / 2] + (double)vals[n / 2 - 1]) / 2);
                medians.push_back({key, wend, med});
                wlist.erase({key, s});
            }
            if (cfg.allowed_lateness_ms == 0) wstate.erase(it);  // clearAllState: cleanup time == fire time
        }
        // cleanup timers (maxTimestamp + allowedLateness): the window's state goes, later records for it are late
        while (!cleanup.empty() && std::get<0>(*cleanup.begin()) <= wm) {
            auto t = *cleanup.begin(); cleanup.erase(cleanup.begin());
            wstate.erase({std::get<1>(t), std::get<2>(t)});
        }
    }

    int64_t auto_watermark() {                                   // getCurrentWatermark (README.md:379-386)
        int64_t potential = current_max_ts - cfg.max_out_of_orderness_ms;
        if (potential >= last_emitted_wm) last_emitted_wm = potential;
        return last_emitted_wm;
    }

    int push(const int64_t* ts, const uint64_t* key, const void* val, size_t n, int64_t wm_after) override {
        const V* v = static_cast<const V*>(val);
        rolling.clear();
        for (size_t i = 0; i < n; i++) {
            if (ts[i] < 0 || key[i] == UINT64_MAX) { err = "negative timestamp or reserved key"; return FMQ_ERR_BAD_TS; }
            process_element(ts[i], key[i], v[i]);
        }
        // FMQ_WM_KEEP: no Watermark element follows these records (WindowOperator#processElement never moves time)
        if (cfg.window_enabled) refire_pending();
        if (cfg.window_enabled && wm_after != FMQ_WM_KEEP) advance_watermark(wm_after == FMQ_WM_AUTO ? auto_watermark() : wm_after);
        ctr.keys = kstate.size();
        return FMQ_OK;
    }
    int flush(int64_t wm) override { if (cfg.window_enabled) advance_watermark(wm); return FMQ_OK; }
    int key_state(uint64_t key, int32_t* consec, uint64_t* rolling_bits) override {
        auto it = kstate.find(key); if (it == kstate.end()) return 0;
        *consec = it->second.consec; *rolling_bits = it->second.has_max ? bits_of(it->second.mx) : 0; return 1;
    }
};

template <typename Row, typename F>
size_t drain(std::vector<Row>& v, size_t cap, F emit) {
    size_t n = std::min(cap, v.size());
    for (size_t i = 0; i < n; i++) emit(i, v[i]);
    v.erase(v.begin(), v.begin() + n);
    return n;
}
}  // namespace

extern "C" {

int32_t orc_create(const fmq_config* cfg, void** out) {
    if (!cfg || !out || cfg->struct_size != sizeof(fmq_config)) return FMQ_ERR_BAD_ARG;
    if (cfg->window_enabled) {
        if (cfg->window_size_ms <= 0 || cfg->window_slide_ms <= 0) return FMQ_ERR_BAD_ARG;
        if (cfg->window_offset_ms < 0 || cfg->window_offset_ms >= cfg->window_slide_ms) return FMQ_ERR_BAD_ARG;
        if (cfg->max_out_of_orderness_ms < 0) return FMQ_ERR_BAD_ARG;   // README.md:356-359
    }
    if (cfg->value_type == FMQ_VAL_F64) *out = new Oracle<double>(*cfg);
    else if (cfg->value_type == FMQ_VAL_I64) *out = new Oracle<int64_t>(*cfg);
    else return FMQ_ERR_BAD_ARG;
    return FMQ_OK;
}
int32_t orc_push_batch(void* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int64_t wm_after) {
    return static_cast<OracleBase*>(h)->push(ts, key, val, n, wm_after);
}
int32_t orc_flush(void* h, int64_t wm) { return static_cast<OracleBase*>(h)->flush(wm); }
int32_t orc_poll_fired(void* h, const fmq_fired_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    auto* b = static_cast<OracleBase*>(h);
    *n_out = drain(b->fired, cap, [&](size_t i, const FiredRow& r) {
        o->key[i] = r.key; o->window_end[i] = r.wend; o->count[i] = r.count;
        ((uint64_t*)o->sum)[i] = r.sum; ((uint64_t*)o->min)[i] = r.mn; ((uint64_t*)o->max)[i] = r.mx; });
    if (n_left) *n_left = b->fired.size();
    return FMQ_OK;
}
int32_t orc_poll_window_alerts(void* h, const fmq_walert_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    auto* b = static_cast<OracleBase*>(h);
    *n_out = drain(b->walerts, cap, [&](size_t i, const WAlert& r) { o->key[i] = r.key; o->window_end[i] = r.wend; o->value[i] = r.value; });
    if (n_left) *n_left = b->walerts.size();
    return FMQ_OK;
}
int32_t orc_poll_medians(void* h, const fmq_walert_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    auto* b = static_cast<OracleBase*>(h);
    *n_out = drain(b->medians, cap, [&](size_t i, const WAlert& r) { o->key[i] = r.key; o->window_end[i] = r.wend; o->value[i] = r.value; });
    if (n_left) *n_left = b->medians.size();
    return FMQ_OK;
}
static int32_t poll_ev(std::vector<EvRow>& v, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    *n_out = drain(v, cap, [&](size_t i, const EvRow& r) { o->ts[i] = r.ts; o->key[i] = r.key; ((uint64_t*)o->val)[i] = r.val; });
    if (n_left) *n_left = v.size();
    return FMQ_OK;
}
int32_t orc_poll_event_alerts(void* h, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    return poll_ev(static_cast<OracleBase*>(h)->ealerts, o, cap, n_out, n_left);
}
int32_t orc_poll_late(void* h, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    return poll_ev(static_cast<OracleBase*>(h)->late, o, cap, n_out, n_left);
}
int32_t orc_poll_consec_alerts(void* h, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    return poll_ev(static_cast<OracleBase*>(h)->calerts, o, cap, n_out, n_left);
}
int32_t orc_poll_rolling_max(void* h, void* out, size_t cap, size_t* n_out) {
    auto* b = static_cast<OracleBase*>(h);
    size_t n = std::min(cap, b->rolling.size());
    std::memcpy(out, b->rolling.data(), n * 8); *n_out = n; return FMQ_OK;
}
int32_t orc_get_counters(void* h, fmq_counters* out) { *out = static_cast<OracleBase*>(h)->ctr; return FMQ_OK; }
int32_t orc_get_key_state(void* h, const uint64_t* keys, size_t n, int32_t* found, int32_t* consec, void* rolling) {
    auto* b = static_cast<OracleBase*>(h);
    for (size_t i = 0; i < n; i++) { consec[i] = 0; ((uint64_t*)rolling)[i] = 0; found[i] = b->key_state(keys[i], &consec[i], &((uint64_t*)rolling)[i]); }
    return FMQ_OK;
}
int32_t orc_destroy(void* h) { delete static_cast<OracleBase*>(h); return FMQ_OK; }

// ---- keyBy routing: Flink 1.8 MathUtils.murmurHash / KeyGroupRangeAssignment (Appendix A.10) ----
int32_t orc_java_string_hash(const char* s, size_t len) {
    // String.hashCode over UTF-16 code units; ASCII-only inputs here (hosts, domains)
    uint32_t h = 0; for (size_t i = 0; i < len; i++) h = 31u * h + (uint8_t)s[i]; return (int32_t)h;
}
int32_t orc_murmur_hash(int32_t code_in) {
    uint32_t code = (uint32_t)code_in;
    code *= 0xcc9e2d51u; code = (code << 15) | (code >> 17); code *= 0x1b873593u;
    code = (code << 13) | (code >> 19); code = code * 5u + 0xe6546b64u;
    code ^= 4u;
    code ^= code >> 16; code *= 0x85ebca6bu; code ^= code >> 13; code *= 0xc2b2ae35u; code ^= code >> 16;  // fmix32
    int32_t s = (int32_t)code;
    if (s >= 0) return s;
    if (s != INT32_MIN) return -s;
    return 0;
}
int32_t orc_key_group(int32_t key_hash, int32_t max_parallelism) { return orc_murmur_hash(key_hash) % max_parallelism; }
int32_t orc_subtask(int32_t key_group, int32_t parallelism, int32_t max_parallelism) { return key_group * parallelism / max_parallelism; }

}  // extern "C"
