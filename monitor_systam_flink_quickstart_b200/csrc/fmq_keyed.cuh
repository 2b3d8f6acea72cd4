// Keyed ValueState operators in per-key ARRIVAL order:
//   (i)  rolling max  = keyBy(0).max(2), chapter2/.../ComputeCpuMax.java:26
//        (StreamGroupedReduce + ComparableAggregator: one output per input, Double.compareTo order)
//   (ii) N consecutive over-threshold readings (BASELINE.json north_star; spec in SURVEY §8 A7-ii)
// Pipeline per batch: resolve key -> directory slot, stable LSD radix sort of (slot, arrival idx),
// then every run of equal slots is walked in order, seeded from / written back to the per-key state.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

// per-key state, indexed by directory slot
//   kflags[slot] : bit 31 = rolling max present, bits 0..30 = consecutive-breach counter
//   kmax[slot]   : rolling max value bits
constexpr u32 KF_HAS_MAX = 0x80000000u;
constexpr u32 KF_SEEN = 0x40000000u;       // dense-id mode: this id has been seen by a keyed operator
constexpr u32 KF_COUNT_MASK = 0x3FFFFFFFu;

__global__ void __launch_bounds__(256) resolve_slots_kernel(const u64* __restrict__ key, size_t n, u64* dkeys, u64 dcap, int dense,
                                                            DevState* st, u32* __restrict__ slot_out, u32* __restrict__ idx_out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const u64 k = ldg_stream(key + i);
        u32 s = SLOT_INVALID;
        if (k == KEY_EMPTY || (dense && k >= dcap)) atomicOr(&st->err_flags, ERRB_BAD_INPUT);
        else s = dense ? (u32)k : dir_find_or_insert(dkeys, dcap, k, st);
        slot_out[i] = s; idx_out[i] = (u32)i;
    }
}

struct KeyedArgs {
    const u32* slot_sorted; const u32* idx_sorted; size_t n;
    const i64* ts; const u64* key; const void* val;
    DevState* st; u32* kflags; u64* kmax;
    int rolling_enabled; u64* rolling_out;          // [n], input order
    int consec_n, consec_op, consec_reset; double consec_thr;
    i64* ca_ts; u64* ca_key; u64* ca_val; u64 ca_cap;
};

// Thread t owns the run of equal slots that STARTS at sorted position t (if any) and walks it in
// arrival order.  Runs are short for monitoring streams (events-per-key-per-batch); heavy-hitter
// keys serialise on one thread — the associative-scan formulation is the planned replacement.
template <typename V>
__global__ void __launch_bounds__(256) keyed_apply_kernel(const __grid_constant__ KeyedArgs a) {
    const V* __restrict__ val = static_cast<const V*>(a.val);
    u64 nalert = 0;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < a.n; t += (size_t)gridDim.x * blockDim.x) {
        const u32 s = a.slot_sorted[t];
        if (s == SLOT_INVALID) continue;
        if (t > 0 && a.slot_sorted[t - 1] == s) continue;          // not a run head
        u32 fl = a.kflags[s];
        bool has = fl & KF_HAS_MAX; int c = (int)(fl & KF_COUNT_MASK);
        u64 mbits = a.kmax[s];
        for (size_t j = t; j < a.n && a.slot_sorted[j] == s; ++j) {
            const u32 i = a.idx_sorted[j];
            const V v = val[i];
            u64 vb; memcpy(&vb, &v, 8);
            if (a.rolling_enabled) {
                // ComparableAggregator.reduce: keep cur iff cur.compareTo(v) > 0, else take v
                if (!has || !(Codec<V>::cmp_key(Codec<V>::from_bits(mbits)) > Codec<V>::cmp_key(v))) { mbits = vb; has = true; }
                a.rolling_out[i] = mbits;
            }
            if (a.consec_n > 0) {
                const bool over = cmp_op((double)v, a.consec_op, a.consec_thr);
                c = over ? c + 1 : 0;
                const bool fire = a.consec_reset ? (c == a.consec_n) : (c >= a.consec_n);
                if (fire) {
                    const u64 row = atomicAdd(&a.st->n_calert, 1ull);
                    if (row < a.ca_cap) { a.ca_ts[row] = a.ts[i]; a.ca_key[row] = a.key[i]; a.ca_val[row] = vb; }
                    else atomicOr(&a.st->err_flags, ERRB_OUTPUT);
                    nalert++;
                    if (a.consec_reset) c = 0;
                }
            }
        }
        a.kflags[s] = (has ? KF_HAS_MAX : 0u) | KF_SEEN | ((u32)c & KF_COUNT_MASK);
        a.kmax[s] = mbits;
    }
    nalert = warp_sum_u64(nalert);
    if ((threadIdx.x & 31) == 0 && nalert) atomicAdd(&a.st->consec_alerts, nalert);
}

__global__ void key_state_lookup_kernel(const u64* keys, size_t n, const u64* dkeys, u64 dcap, int dense, const u32* kflags, const u64* kmax,
                                        int* found, int* consec, u64* rolling) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 s = dense ? (keys[i] < dcap ? (u32)keys[i] : SLOT_INVALID) : dir_find(dkeys, dcap, keys[i]);
    if (s == SLOT_INVALID || (dense && !(kflags[s] & KF_SEEN))) { found[i] = 0; consec[i] = 0; rolling[i] = 0; return; }
    const u32 fl = kflags[s];
    found[i] = 1; consec[i] = (int)(fl & KF_COUNT_MASK); rolling[i] = (fl & KF_HAS_MAX) ? kmax[s] : 0;
}

}  // namespace fmq
