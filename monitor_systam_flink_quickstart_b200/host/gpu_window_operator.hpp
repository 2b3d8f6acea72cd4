// C++ mirror of the Flink operator the reference's DSL chain instantiates, written against the C ABI
// only (include/fmq_b200.h) — this is the host code a JVM shim would contain, in the language this
// image can compile.  Same hooks as OneInputStreamOperator [Flink 1.8.0]:
//   open() / processElement(ts, key, value) / processWatermark(wm) / close()
// replacing keyBy(1).timeWindow(..).reduce(..).map(..).filter(..)
// (chapter3/.../BandwidthMonitorWithEventTime.java:45-55).
#pragma once
#include <cstdint>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>
#include "../../include/fmq_b200.h"

namespace fmq_host {

struct WindowResult { std::string key; int64_t timestamp; int64_t count; uint64_t sum_bits, min_bits, max_bits; };
struct WindowAlert  { std::string key; int64_t timestamp; double value; };

class GpuWindowAggOperator {
public:
    using Collector = std::function<void(const WindowResult*, const WindowAlert*)>;

    GpuWindowAggOperator(const fmq_config& cfg, size_t batch_capacity, Collector out)
        : cfg_(cfg), cap_(batch_capacity), out_(std::move(out)) {
        cfg_.struct_size = sizeof(fmq_config);
        if (cfg_.max_batch < cap_) cfg_.max_batch = cap_;
    }
    ~GpuWindowAggOperator() { if (h_) fmq_destroy(h_); }

    void open() {                                                       // OneInputStreamOperator#open
        check(fmq_create(&cfg_, &h_), nullptr);
        ts_.resize(cap_); key_.resize(cap_); val_.resize(cap_); n_ = 0;
    }
    template <typename V>
    void processElement(int64_t ts, const std::string& key, V value) {  // #processElement
        static_assert(sizeof(V) == 8, "values are 8-byte (double or int64_t)");
        if (n_ == cap_) push(FMQ_WM_KEEP);                               // buffer full between two upstream watermarks: time does not move
        ts_[n_] = ts; key_[n_] = encode(key); std::memcpy(&val_[n_], &value, 8); ++n_;
    }
    void processWatermark(int64_t wm) { push(wm); }                     // #processWatermark
    void close() {                                                      // #close: Watermark(Long.MAX_VALUE)
        if (n_) push(FMQ_WM_KEEP);
        check(fmq_flush(h_, FMQ_WM_FINAL), h_); drain();
        fmq_destroy(h_); h_ = nullptr;
    }
    uint64_t numLateRecordsDropped() { fmq_counters c; check(fmq_get_counters(h_, &c), h_); return c.late_dropped; }

private:
    uint64_t encode(const std::string& k) {                             // String key -> dense id
        auto it = dict_.find(k);
        if (it != dict_.end()) return it->second;
        uint64_t id = names_.size(); dict_.emplace(k, id); names_.push_back(k); return id;
    }
    void push(int64_t wm) {
        check(fmq_push_batch(h_, ts_.data(), key_.data(), val_.data(), n_, wm, FMQ_MEM_HOST), h_);
        n_ = 0; drain();
    }
    void drain() {
        const size_t C = 4096;
        std::vector<uint64_t> k(C), s(C), mn(C), mx(C); std::vector<int64_t> we(C), cnt(C); std::vector<double> v(C);
        size_t got = 0, left = 0;
        if (cfg_.window_enabled && cfg_.emit_fired) do {
            fmq_fired_cols cols{k.data(), we.data(), cnt.data(), s.data(), mn.data(), mx.data()};
            check(fmq_poll_fired(h_, &cols, C, FMQ_MEM_HOST, &got, &left), h_);
            for (size_t i = 0; i < got; ++i) { WindowResult r{names_.at(k[i]), we[i] - 1, cnt[i], s[i], mn[i], mx[i]}; out_(&r, nullptr); }
        } while (left);
        if (cfg_.window_enabled && cfg_.walert_source != FMQ_ALERT_NONE) do {
            fmq_walert_cols cols{k.data(), we.data(), v.data()};
            check(fmq_poll_window_alerts(h_, &cols, C, FMQ_MEM_HOST, &got, &left), h_);
            for (size_t i = 0; i < got; ++i) { WindowAlert a{names_.at(k[i]), we[i] - 1, v[i]}; out_(nullptr, &a); }
        } while (left);
    }
    static void check(int32_t st, fmq_handle* h) {
        if (st != FMQ_OK) throw std::runtime_error(std::string("fmq status ") + std::to_string(st) + ": " + fmq_last_error(h));
    }
    fmq_config cfg_; size_t cap_; Collector out_; fmq_handle* h_ = nullptr;
    std::vector<int64_t> ts_; std::vector<uint64_t> key_, val_; size_t n_ = 0;
    std::unordered_map<std::string, uint64_t> dict_; std::vector<std::string> names_;
};

}  // namespace fmq_host
