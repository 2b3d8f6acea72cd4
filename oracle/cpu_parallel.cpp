// TEST / BASELINE INFRASTRUCTURE — NOT PRODUCT CODE.
// Key-partitioned multi-thread CPU restatement of the same Flink semantics as flink_literal.cpp,
// organised the way a Flink job with parallelism p runs it: a hash exchange
// (keyBy, BandwidthMonitorWithEventTime.java:45) in front of p independent keyed operator
// instances, each doing per-record / per-window work (assignWindows fan-out, late test, keyed
// accumulator update) and firing on the watermark.  It is the timed CPU baseline of bench.py
// ("C++ restatement of the reference semantics", kind "port" — NOT Flink, which cannot run on
// this box: no JVM).  tests/test_oracle_cross.py checks it row-for-row against the literal oracle.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <map>
#include <thread>
#include <vector>
#include <algorithm>
#include "../include/fmq_b200.h"

namespace {

inline uint64_t mix64(uint64_t x) { x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31; return x; }

template <typename T> struct FlatMap {            // open addressing, linear probing, no deletes
    std::vector<uint64_t> keys; std::vector<T> vals; size_t used = 0, mask = 0;
    FlatMap() { rehash(1024); }
    void rehash(size_t cap) {
        std::vector<uint64_t> ok; std::vector<T> ov; ok.swap(keys); ov.swap(vals);
        keys.assign(cap, UINT64_MAX); vals.assign(cap, T{}); mask = cap - 1; used = 0;
        for (size_t i = 0; i < ok.size(); i++) if (ok[i] != UINT64_MAX) { bool f; *get(ok[i], &f) = ov[i]; }
    }
    const T* find(uint64_t k) const {
        size_t h = mix64(k) & mask;
        while (true) {
            if (keys[h] == k) return &vals[h];
            if (keys[h] == UINT64_MAX) return nullptr;
            h = (h + 1) & mask;
        }
    }
    T* get(uint64_t k, bool* fresh) {
        if ((used + 1) * 10 > keys.size() * 6) rehash(keys.size() * 2);
        size_t h = mix64(k) & mask;
        while (true) {
            if (keys[h] == k) { *fresh = false; return &vals[h]; }
            if (keys[h] == UINT64_MAX) { keys[h] = k; used++; *fresh = true; return &vals[h]; }
            h = (h + 1) & mask;
        }
    }
};

template <typename V> struct Acc { int64_t count; V sum, mn, mx; };
inline double jmax(double a, double b) { if (a != a) return a; if (a == 0.0 && b == 0.0) return std::signbit(a) ? b : a; return a >= b ? a : b; }
inline double jmin(double a, double b) { if (a != a) return a; if (a == 0.0 && b == 0.0) return std::signbit(a) ? a : b; return a <= b ? a : b; }
inline int64_t jmax(int64_t a, int64_t b) { return a >= b ? a : b; }
inline int64_t jmin(int64_t a, int64_t b) { return a <= b ? a : b; }
inline int jcmp(double a, double b) {
    if (a < b) return -1; if (a > b) return 1;
    auto bits = [](double d) -> int64_t { if (d != d) return 0x7ff8000000000000LL; int64_t r; std::memcpy(&r, &d, 8); return r; };
    int64_t x = bits(a), y = bits(b); return x == y ? 0 : (x < y ? -1 : 1);
}
inline int jcmp(int64_t a, int64_t b) { return a < b ? -1 : (a > b ? 1 : 0); }
inline bool cmp(double v, int op, double t) { switch (op) { case 0: return v > t; case 1: return v < t; case 2: return v >= t; default: return v <= t; } }
inline double oz(double d) { return d == 0.0 ? 1.0 : d; }
template <typename V> inline uint64_t bits_of(V v) { uint64_t r; std::memcpy(&r, &v, 8); return r; }

struct FiredRow { uint64_t key; int64_t wend; int64_t count; uint64_t sum, mn, mx; };
struct WAlert { uint64_t key; int64_t wend; double value; };
struct EvRow { int64_t ts; uint64_t key; uint64_t val; };

struct ParBase {
    virtual ~ParBase() {}
    virtual int push(const int64_t*, const uint64_t*, const void*, size_t, int64_t) = 0;
    virtual int flush(int64_t) = 0;
    virtual void counters(fmq_counters*) = 0;
    virtual int key_state(uint64_t key, int32_t* consec, uint64_t* rolling_bits) = 0;
    std::vector<FiredRow> fired; std::vector<WAlert> walerts; std::vector<EvRow> ealerts, calerts;
};

template <typename V>
struct Par : ParBase {
    fmq_config cfg; int T;
    struct KState { bool has = false; V mx{}; int32_t c = 0; };
    struct Sub {                                               // one keyed operator instance (subtask)
        std::map<int64_t, FlatMap<Acc<V>>> windows;            // window start -> (key -> accumulator)
        FlatMap<KState> kstate;
        std::vector<FiredRow> fired; std::vector<WAlert> walerts; std::vector<EvRow> ealerts, calerts;
        uint64_t late = 0, nfired = 0;
    };
    std::vector<Sub> subs;
    std::vector<std::vector<std::vector<uint32_t>>> buckets;   // [source slice][owner] -> indices
    int64_t wm = INT64_MIN, cur_max_ts, last_wm = INT64_MIN; uint64_t events_in = 0;

    Par(const fmq_config& c, int threads) : cfg(c), T(threads), subs(threads), buckets(threads, std::vector<std::vector<uint32_t>>(threads)) {
        cur_max_ts = INT64_MIN + cfg.max_out_of_orderness_ms;
    }
    template <typename F> void parallel(F f) {
        std::vector<std::thread> th; for (int t = 1; t < T; t++) th.emplace_back(f, t);
        f(0); for (auto& x : th) x.join();
    }
    void element(Sub& s, int64_t ts, uint64_t key, V val) {
        if (cfg.rolling_max_enabled || cfg.consec_n > 0) {
            bool fresh; KState* st = s.kstate.get(key, &fresh);
            if (cfg.rolling_max_enabled) { if (!st->has) { st->has = true; st->mx = val; } else if (!(jcmp(st->mx, val) > 0)) st->mx = val; }
            if (cfg.consec_n > 0) {
                bool over = cmp((double)val, cfg.consec_op, cfg.consec_threshold);
                int32_t c = over ? st->c + 1 : 0;
                bool fire = cfg.consec_reset_on_fire ? (c == cfg.consec_n) : (c >= cfg.consec_n);
                if (fire) { s.calerts.push_back({ts, key, bits_of(val)}); if (cfg.consec_reset_on_fire) c = 0; }
                st->c = c;
            }
        }
        if (!cfg.window_enabled) return;
        const int64_t size = cfg.window_size_ms, slide = cfg.window_slide_ms, off = cfg.window_offset_ms;
        bool skipped = true;
        for (int64_t st = ts - (ts - off + slide) % slide; st > ts - size; st -= slide) {   // assignWindows
            if (st + size - 1 <= wm) continue;                                                // isWindowLate
            skipped = false;
            bool fresh; Acc<V>* a = s.windows[st].get(key, &fresh);
            if (fresh) *a = Acc<V>{1, val, val, val};
            else { a->count++; a->sum = a->sum + val; a->mn = jmin(a->mn, val); a->mx = jmax(a->mx, val); }
        }
        if (skipped && ts <= wm) s.late++;
    }
    double alert_value(const Acc<V>& a) const {
        double src;
        switch (cfg.walert_source) {
            case FMQ_ALERT_SUM: src = (double)a.sum; break;
            case FMQ_ALERT_AVG: return a.count == 0 ? 0.0 : (double)a.sum / (double)a.count;
            case FMQ_ALERT_MIN: src = (double)a.mn; break;
            case FMQ_ALERT_MAX: src = (double)a.mx; break;
            default: src = (double)a.count; break;
        }
        return src * oz(cfg.walert_mul) / oz(cfg.walert_div0) / oz(cfg.walert_div1) / oz(cfg.walert_div2);
    }
    void fire(Sub& s, int64_t w) {
        while (!s.windows.empty() && s.windows.begin()->first + cfg.window_size_ms - 1 <= w) {
            int64_t wend = s.windows.begin()->first + cfg.window_size_ms;
            auto& m = s.windows.begin()->second;
            for (size_t i = 0; i < m.keys.size(); i++) if (m.keys[i] != UINT64_MAX) {
                const Acc<V>& a = m.vals[i]; s.nfired++;
                if (cfg.emit_fired) s.fired.push_back({m.keys[i], wend, a.count, bits_of(a.sum), bits_of(a.mn), bits_of(a.mx)});
                if (cfg.walert_source != FMQ_ALERT_NONE) { double v = alert_value(a); if (cmp(v, cfg.walert_op, cfg.walert_threshold)) s.walerts.push_back({m.keys[i], wend, v}); }
            }
            s.windows.erase(s.windows.begin());
        }
    }
    void advance(int64_t w) {
        if (w <= wm) return; wm = w;
        parallel([&](int t) { fire(subs[t], w); });
    }
    void collect() {
        for (auto& s : subs) {
            fired.insert(fired.end(), s.fired.begin(), s.fired.end()); s.fired.clear();
            walerts.insert(walerts.end(), s.walerts.begin(), s.walerts.end()); s.walerts.clear();
            ealerts.insert(ealerts.end(), s.ealerts.begin(), s.ealerts.end()); s.ealerts.clear();
            calerts.insert(calerts.end(), s.calerts.begin(), s.calerts.end()); s.calerts.clear();
        }
    }
    int push(const int64_t* ts, const uint64_t* key, const void* valp, size_t n, int64_t wm_after) override {
        const V* val = static_cast<const V*>(valp);
        events_in += n;
        std::vector<int64_t> lmax(T, INT64_MIN);
        // phase 1: the rebalance + keyBy exchange — each source slice routes its records to owners
        parallel([&](int t) {
            size_t lo = n * t / T, hi = n * (t + 1) / T; int64_t m = INT64_MIN;
            for (auto& b : buckets[t]) b.clear();
            for (size_t i = lo; i < hi; i++) {
                if (ts[i] > m) m = ts[i];
                if (cfg.efilter_enabled && cmp((double)val[i], cfg.efilter_op, cfg.efilter_threshold)) subs[t].ealerts.push_back({ts[i], key[i], bits_of(val[i])});
                buckets[t][mix64(key[i]) % T].push_back((uint32_t)i);
            }
            lmax[t] = m;
        });
        // phase 2: each keyed subtask consumes its input channels in source order
        parallel([&](int o) {
            for (int t = 0; t < T; t++) for (uint32_t i : buckets[t][o]) element(subs[o], ts[i], key[i], val[i]);
        });
        if (cfg.window_enabled) {
            for (int t = 0; t < T; t++) if (lmax[t] > cur_max_ts) cur_max_ts = lmax[t];
            int64_t w = wm_after;
            if (wm_after == FMQ_WM_AUTO) { int64_t p = cur_max_ts - cfg.max_out_of_orderness_ms; if (p >= last_wm) last_wm = p; w = last_wm; }
            if (wm_after != FMQ_WM_KEEP) advance(w);
        }
        collect();
        return FMQ_OK;
    }
    int flush(int64_t w) override { if (cfg.window_enabled) advance(w); collect(); return FMQ_OK; }
    int key_state(uint64_t key, int32_t* consec, uint64_t* rolling_bits) override {
        const KState* st = subs[mix64(key) % T].kstate.find(key);
        if (!st) return 0;
        *consec = st->c; *rolling_bits = st->has ? bits_of(st->mx) : 0; return 1;
    }
    void counters(fmq_counters* c) override {
        std::memset(c, 0, sizeof(*c)); c->events_in = events_in; c->watermark = wm;
        for (auto& s : subs) { c->late_dropped += s.late; c->windows_fired += s.nfired; c->keys += s.kstate.used; }
    }
};

template <typename Row, typename F> size_t drain(std::vector<Row>& v, size_t cap, F emit) {
    size_t n = std::min(cap, v.size()); for (size_t i = 0; i < n; i++) emit(i, v[i]); v.erase(v.begin(), v.begin() + n); return n;
}
}  // namespace

extern "C" {
int32_t orcp_create(const fmq_config* cfg, int32_t threads, void** out) {
    if (!cfg || !out || cfg->struct_size != sizeof(fmq_config) || threads < 1) return FMQ_ERR_BAD_ARG;
    if (cfg->allowed_lateness_ms != 0) return FMQ_ERR_BAD_ARG;      // restated in flink_literal.cpp only
    if (cfg->value_type == FMQ_VAL_F64) *out = new Par<double>(*cfg, threads); else *out = new Par<int64_t>(*cfg, threads);
    return FMQ_OK;
}
int32_t orcp_push_batch(void* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int64_t wm) { return static_cast<ParBase*>(h)->push(ts, key, val, n, wm); }
int32_t orcp_flush(void* h, int64_t wm) { return static_cast<ParBase*>(h)->flush(wm); }
int32_t orcp_get_counters(void* h, fmq_counters* c) { static_cast<ParBase*>(h)->counters(c); return FMQ_OK; }
int32_t orcp_poll_fired(void* h, const fmq_fired_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    auto* b = static_cast<ParBase*>(h);
    *n_out = drain(b->fired, cap, [&](size_t i, const FiredRow& r) { o->key[i] = r.key; o->window_end[i] = r.wend; o->count[i] = r.count; ((uint64_t*)o->sum)[i] = r.sum; ((uint64_t*)o->min)[i] = r.mn; ((uint64_t*)o->max)[i] = r.mx; });
    if (n_left) *n_left = b->fired.size(); return FMQ_OK;
}
int32_t orcp_poll_window_alerts(void* h, const fmq_walert_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    auto* b = static_cast<ParBase*>(h);
    *n_out = drain(b->walerts, cap, [&](size_t i, const WAlert& r) { o->key[i] = r.key; o->window_end[i] = r.wend; o->value[i] = r.value; });
    if (n_left) *n_left = b->walerts.size(); return FMQ_OK;
}
static int32_t poll_ev(std::vector<EvRow>& v, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) {
    *n_out = drain(v, cap, [&](size_t i, const EvRow& r) { o->ts[i] = r.ts; o->key[i] = r.key; ((uint64_t*)o->val)[i] = r.val; });
    if (n_left) *n_left = v.size(); return FMQ_OK;
}
int32_t orcp_poll_event_alerts(void* h, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) { return poll_ev(static_cast<ParBase*>(h)->ealerts, o, cap, n_out, n_left); }
int32_t orcp_poll_consec_alerts(void* h, const fmq_event_cols* o, size_t cap, size_t* n_out, size_t* n_left) { return poll_ev(static_cast<ParBase*>(h)->calerts, o, cap, n_out, n_left); }
int32_t orcp_discard_outputs(void* h) { auto* b = static_cast<ParBase*>(h); b->fired.clear(); b->walerts.clear(); b->ealerts.clear(); b->calerts.clear(); return FMQ_OK; }
int32_t orcp_get_key_state(void* h, const uint64_t* keys, size_t n, int32_t* found, int32_t* consec, void* rolling) {
    auto* b = static_cast<ParBase*>(h);
    for (size_t i = 0; i < n; i++) { consec[i] = 0; ((uint64_t*)rolling)[i] = 0; found[i] = b->key_state(keys[i], &consec[i], &((uint64_t*)rolling)[i]); }
    return FMQ_OK;
}
int32_t orcp_destroy(void* h) { delete static_cast<ParBase*>(h); return FMQ_OK; }
}
