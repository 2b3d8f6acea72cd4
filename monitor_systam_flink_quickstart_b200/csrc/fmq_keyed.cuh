// Keyed ValueState operators in per-key ARRIVAL order:
//   (i)  rolling max  = keyBy(0).max(2), chapter2/.../ComputeCpuMax.java:26
//        (StreamGroupedReduce + ComparableAggregator: one output per input, Double.compareTo order)
//   (ii) N consecutive over-threshold readings (BASELINE.json north_star; spec in SURVEY §8 A7-ii)
// Pipeline per batch: resolve key -> directory slot, stable LSD radix sort of (slot, arrival idx),
// then every run of equal slots is walked in order, seeded from / written back to the per-key state.
#pragma once
#include "fmq_device.cuh"
#include "fmq_sort.cuh"

namespace fmq {

// per-key state, indexed by directory slot
//   kflags[slot] : bit 31 = rolling max present, bits 0..30 = consecutive-breach counter
//   kmax[slot]   : rolling max value bits
constexpr u32 KF_HAS_MAX = 0x80000000u;
constexpr u32 KF_SEEN = 0x40000000u;       // dense-id mode: this id has been seen by a keyed operator
constexpr u32 KF_COUNT_MASK = 0x3FFFFFFFu;

// ---- step 1: key -> (slot << 32 | arrival index), one packed 8-byte item per event ----------------
// When only the consecutive-readings rule runs (no rolling max), the one thing the rule needs from a
// record, "is the value over the threshold", rides along as bit 31 of the item (KI_OVER): the scan then
// streams the sorted items and never gathers values (ncu: the gathers were 1.15 GB of DRAM traffic per
// 2^24 records, 0.31 ms; profiles/r01_ncu_keyed_chain_summary.txt).  max_batch <= 2^31 in that mode.
constexpr u32 KI_OVER = 0x80000000u;
// ---- step 2: stable LSD radix sort of the packed items by slot, BITS (<= 11) bits per pass -------
// Same scheme as the owner split (a warp owns a contiguous chunk, counts[digit][chunk] scanned
// digit-major), tuned for this job: 8-byte items, up to 2048 bins so that 22 slot bits need two
// passes, and four rounds of loads issued ahead of the ranking so that a warp is not one
// dependent load per 32 items.
constexpr u32 PS_CHUNK = 4096;
constexpr int PS_AHEAD = 4;
__host__ __device__ inline u32 ps_num_chunks(size_t n) { return (u32)((n + PS_CHUNK - 1) / PS_CHUNK); }
// digit width of a sort over `key_bits` bits: as few passes as 11-bit digits allow, bits spread evenly
__host__ __device__ inline u32 ps_pass_bits(u32 key_bits) { const u32 passes = (key_bits + 10) / 11; const u32 b = (key_bits + passes - 1) / passes; return b < 6 ? 6 : b; }

// KeyFn: u64 operator()(u64 item) -> the sort key of an item; a pass ranks by (key >> shift) & (2^BITS - 1)
struct SlotKey {            // item = (slot << 32 | index): sort by slot
    __device__ __forceinline__ u64 operator()(u64 item) const { return item >> 32; }
};

// Counting half of a radix pass: every warp counts the digits of one chunk of PS_CHUNK items into its own bins
// (counting needs no order: one shared-memory atomic per item), counts[digit][chunk].  `SCATTER` is kept in the
// signature for the call sites; only the counting form exists (the scatter is ps_scatter_tile_kernel).
template <int BITS, bool SCATTER, typename KeyFn>
__global__ void __launch_bounds__(256) ps_pass_kernel(const u64* __restrict__ in, u64* __restrict__ /*out*/, size_t n, u32 nchunks, u32 shift,
                                                      u32* __restrict__ counts, KeyFn keyfn) {
    static_assert(!SCATTER, "the scatter half is ps_scatter_tile_kernel");
    extern __shared__ u32 ps_smem[];
    constexpr u32 NB = 1u << BITS;
    const u32 wl = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u32* bins = ps_smem + wl * NB;
    const u32 chunk = blockIdx.x * 8 + wl;
    if (chunk >= nchunks) return;
    for (u32 d = lane; d < NB; d += 32) bins[d] = 0u;
    __syncwarp();
    const size_t lo = (size_t)chunk * PS_CHUNK;
    const size_t hi = (lo + PS_CHUNK < n) ? lo + PS_CHUNK : n;
    for (size_t i0 = lo; i0 < hi; i0 += 32 * PS_AHEAD) {
        u64 it[PS_AHEAD];
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r) { const size_t i = i0 + r * 32 + lane; it[r] = i < hi ? in[i] : ~0ull; }
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r)
            if (i0 + r * 32 + lane < hi) atomicAdd(&bins[(u32)(keyfn(it[r]) >> shift) & (NB - 1)], 1u);
    }
    __syncwarp();
    for (u32 d = lane; d < NB; d += 32) counts[(size_t)d * nchunks + chunk] = bins[d];
}

// Step 1, fused with the counting half of the first radix pass: the warp that resolves a chunk of keys also
// counts the first digit of the items it writes (saves one read of the packed items: 35 us per 2^24).
template <typename V, int BITS>
__global__ void __launch_bounds__(256) resolve_hist_kernel(const u64* __restrict__ key, const V* __restrict__ val, size_t n, u32 nchunks, u64* dkeys, u64 dcap, int dense,
                                                           int pack_over, int op, double thr, DevState* st, u64* __restrict__ packed, u32* __restrict__ counts) {
    extern __shared__ u32 ps_smem[];
    constexpr u32 NB = 1u << BITS;
    const u32 wl = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u32* bins = ps_smem + wl * NB;
    const u32 chunk = blockIdx.x * 8 + wl;
    if (chunk >= nchunks) return;
    for (u32 d = lane; d < NB; d += 32) bins[d] = 0u;
    __syncwarp();
    const size_t lo = (size_t)chunk * PS_CHUNK;
    const size_t hi = (lo + PS_CHUNK < n) ? lo + PS_CHUNK : n;
    for (size_t i0 = lo; i0 < hi; i0 += 32 * PS_AHEAD) {
        u64 kk[PS_AHEAD], vv[PS_AHEAD];
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r) {
            const size_t i = i0 + r * 32 + lane;
            kk[r] = i < hi ? ldg_stream(key + i) : 0; vv[r] = (pack_over && i < hi) ? ldg_stream(val + i) : 0;
        }
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r) {
            const size_t i = i0 + r * 32 + lane;
            if (i >= hi) continue;
            const u64 k = kk[r];
            u32 s = SLOT_INVALID;
            if (k == KEY_EMPTY || (dense && k >= dcap)) atomicOr(&st->err_flags, ERRB_BAD_INPUT);
            else s = dense ? (u32)k : dir_find_or_insert(dkeys, dcap, k, st);
            u32 low = (u32)i;
            if (pack_over) { V v; memcpy(&v, &vv[r], 8); if (cmp_op((double)v, op, thr)) low |= KI_OVER; }
            packed[i] = ((u64)s << 32) | low;
            atomicAdd(&bins[s & (NB - 1)], 1u);
        }
    }
    __syncwarp();
    for (u32 d = lane; d < NB; d += 32) counts[(size_t)d * nchunks + chunk] = bins[d];
}

// Scatter half of a radix pass, one BLOCK per chunk of PS_CHUNK items: the chunk is ranked by digit in shared memory
// (stable: warp w ranks its 512 consecutive items round by round, a cross-warp scan per digit orders the warps) and then
// copied out position by position, so consecutive threads write consecutive addresses of a digit's run (lone 8-byte
// stores were 782 MB of DRAM traffic for 268 MB of payload).
// The chunk arrives as ONE 32 KB bulk copy (TMA) into shared memory: with plain loads ptxas, short of registers, issued
// each round's load two rounds before its use, and every one of the 16 ranking rounds waited on DRAM (ncu: long
// scoreboard in every round, 59 % of all samples in the ranking phase).  Items are never held in registers: a round
// reads its item from shared memory, and the sorted order is a 16-bit source index per position.
constexpr int PS_IPT = PS_CHUNK / 256;          // items per thread
#ifndef FMQ_PS_MINB
#define FMQ_PS_MINB 4
#endif
constexpr size_t ps_scatter_smem(u32 nb) { return (size_t)PS_CHUNK * 8 + (size_t)nb * 8 + (size_t)8 * nb * 2 + (size_t)PS_CHUNK * 2; }

template <int BITS, typename KeyFn>
__global__ void __launch_bounds__(256, FMQ_PS_MINB) ps_scatter_tile_kernel(const u64* __restrict__ in, u64* __restrict__ out, size_t n, u32 nchunks, u32 shift,
                                                              const u32* __restrict__ offsets, KeyFn keyfn) {
    constexpr u32 NB = 1u << BITS;
    extern __shared__ __align__(128) unsigned char ps_tile_smem[];
    u64* in_s = reinterpret_cast<u64*>(ps_tile_smem);                            // [PS_CHUNK] the chunk as it arrived
    u32* goff = reinterpret_cast<u32*>(in_s + PS_CHUNK);                         // [NB] global offset of (digit, this chunk)
    u32* dbase = goff + NB;                                                      // [NB] start of the digit inside the sorted chunk
    unsigned short* wcnt = reinterpret_cast<unsigned short*>(dbase + NB);        // [8][NB] per-warp counts, then exclusive prefixes
    unsigned short* src_s = wcnt + 8 * NB;                                       // [PS_CHUNK] sorted position -> index in the chunk
    __shared__ u32 scan_ws[8];
    __shared__ __align__(8) u64 bar;
    const u32 tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const u32 chunk = blockIdx.x;
    const size_t lo = (size_t)chunk * PS_CHUNK;
    const size_t hi = (lo + PS_CHUNK < n) ? lo + PS_CHUNK : n;
    const u32 cnt = (u32)(hi - lo);
    const bool bulk = cnt == PS_CHUNK && ((uintptr_t)in & 15) == 0;
    if (tid == 0 && bulk) {
        mbar_init(&bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        u64 policy;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
        mbar_expect_tx(&bar, PS_CHUNK * 8u);
        tma_load_1d(in_s, in + lo, PS_CHUNK * 8u, &bar, policy);
    }
    if (!bulk) for (u32 i = tid; i < PS_CHUNK; i += 256) in_s[i] = i < cnt ? in[lo + i] : ~0ull;     // ragged last chunk
    for (u32 i = tid; i < 8 * NB; i += 256) wcnt[i] = 0;
    for (u32 d = tid; d < NB; d += 256) goff[d] = offsets[(size_t)d * nchunks + chunk];
    __syncthreads();
    if (bulk) mbar_wait(&bar, 0);
    // phase 1: rank inside the warp's 512 consecutive items, in order
    unsigned short rk[PS_IPT];
    const u32 wlo = w * (PS_CHUNK / 8);
#pragma unroll
    for (int r = 0; r < PS_IPT; ++r) {
        const u32 i = wlo + r * 32 + lane;
        const bool act = i < cnt;
        const unsigned am = __ballot_sync(0xffffffffu, act);
        const u32 d = (u32)(keyfn(in_s[i]) >> shift) & (NB - 1);                // (every lane takes part in the votes and the shuffle)
        const unsigned m = match_digit<BITS>(am, d);
        const int leader = (__ffs(m) - 1) & 31;
        u32 base = 0;
        if (act && (int)lane == leader) { base = wcnt[w * NB + d]; wcnt[w * NB + d] = (unsigned short)(base + __popc(m)); }
        base = __shfl_sync(0xffffffffu, base, leader);
        rk[r] = (unsigned short)(base + __popc(m & ((1u << lane) - 1)));
        __syncwarp();
    }
    __syncthreads();
    // phase 2: per digit, exclusive prefix over the 8 warps; then exclusive prefix over the digits
    u32 tot_local = 0;
    constexpr u32 DPT = NB / 256 ? NB / 256 : 1;                                 // digits per thread (contiguous)
    u32 dtot[DPT];
#pragma unroll
    for (u32 q = 0; q < DPT; ++q) {
        const u32 d = tid * DPT + q;
        u32 run = 0;
        if (d < NB) for (int ww = 0; ww < 8; ++ww) { const u32 c = wcnt[ww * NB + d]; wcnt[ww * NB + d] = (unsigned short)run; run += c; }
        dtot[q] = run; tot_local += run;
    }
    u32 inc = tot_local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += x; }
    if (lane == 31) scan_ws[w] = inc;
    __syncthreads();
    u32 wb = 0;
    for (u32 i = 0; i < w; ++i) wb += scan_ws[i];
    u32 run = wb + inc - tot_local;
#pragma unroll
    for (u32 q = 0; q < DPT; ++q) { const u32 d = tid * DPT + q; if (d < NB) dbase[d] = run; run += dtot[q]; }
    __syncthreads();
    // phase 3: the sorted position of every item
#pragma unroll
    for (int r = 0; r < PS_IPT; ++r) {
        const u32 i = wlo + r * 32 + lane;
        if (i < cnt) {
            const u32 d = (u32)(keyfn(in_s[i]) >> shift) & (NB - 1);
            src_s[dbase[d] + wcnt[w * NB + d] + rk[r]] = (unsigned short)i;
        }
    }
    __syncthreads();
    // phase 4: copy out; consecutive threads -> consecutive positions -> consecutive addresses inside a run
    for (u32 pos = tid; pos < cnt; pos += 256) {
        const u64 item = in_s[src_s[pos]];
        const u32 d = (u32)(keyfn(item) >> shift) & (NB - 1);
        out[(size_t)goff[d] + (pos - dbase[d])] = item;
    }
}

// ---- step 3: apply the operators along every key's run, in arrival order ---------------------------
// The sorted array is cut into ranges of KA_RANGE positions, one warp each.  A record's new state is a
// FUNCTION of the key's state before it:
//   consecutive-breach counter:  over -> c+1 (alert when it reaches N, then 0 if reset-on-fire), else -> 0
//       composite of a sub-run = (hr, k):  c -> hr ? k : c + k            (taken mod N at the end in reset mode)
//   rolling max (ComparableAggregator: keep cur iff cur.compareTo(v) > 0, else take v)
//       composite = (mc, m):  cur -> mc ? m : max(cur, m), ties to the later record
// and composition is associative, so a run is scanned in parallel: 32 positions per round with a
// segmented warp scan, rounds chained through a carried function.  Three kernels:
//   A  keyed_scan_kernel   every range: runs whose head lies inside the range are seeded from the key's
//                          stored state and finished (outputs, alerts, state write-back at the tail);
//                          the leading part of a range that continues a run from an earlier range is
//                          only summarised.  Each range also records its end function E (a constant if
//                          its last run has its head inside, the composite of the whole range otherwise).
//   B  keyed_carry_kernel  one block: exclusive scan of the E's = state entering every range.
//   C  keyed_lead_kernel   ranges that start inside a run: redo that leading part with the carried state.
// A heavy-hitter key (Zipf: one key = 18 % of a batch, a run of 3 M records) is thereby spread over
// thousands of warps instead of being one warp's job (measured 215 ms -> see profiles/).
struct KeyedArgs {
    const u64* sorted; size_t n;                    // (slot << 32 | arrival index), sorted by slot, stable
    const i64* ts; const u64* key; const void* val;
    DevState* st; u32* kflags; u64* kmax;
    int rolling_enabled; u64* rolling_out;          // [n], input order
    int consec_n, consec_op, consec_reset; double consec_thr;
    i64* ca_ts; u64* ca_key; u64* ca_val; u64 ca_cap;
    // per-range summaries / carries
    u32* r_lead; u32* r_fn_k; u32* r_fn_flags; u64* r_fn_mkey; u64* r_fn_mbits;     // E: flags bit0 = hr, bit1 = mc, bit2 = has max
    u32* r_in_c; u64* r_in_mkey; u64* r_in_mbits;                                     // state entering the range
};

constexpr u32 KA_RANGE = 2048;                      // sorted positions per warp
constexpr u32 KFN_HR = 1u, KFN_MC = 2u, KFN_HASM = 4u;

struct KCarry { bool hr; u32 k; bool mc; bool hasm; u64 mkey, mbits; };

// Scans sorted positions [lo, hi) with one warp.  `c` = function carried into the first position (identity
// for kernel A, the constant entering state for kernel C).  Outputs are produced for the positions whose
// state is fully known (c.hr && c.mc or a run head at/before them).  Returns the carried function after
// `hi` and the number of positions before the first run head (`lead`).
template <typename V>
__device__ __forceinline__ void keyed_scan_range(const KeyedArgs& a, size_t lo, size_t hi, KCarry& c, u32& lead, u64& nalert) {
    const bool rolling = a.rolling_enabled;
    const V* __restrict__ val = static_cast<const V*>(a.val);
    const u32 lane = threadIdx.x & 31;
    const u32 N = (u32)a.consec_n;
    u32 c_slot = lo > 0 ? (u32)(a.sorted[lo - 1] >> 32) : SLOT_INVALID;
    bool seen_head = false; lead = (u32)(hi - lo);
    // Software pipeline, two deep: items are requested two rounds ahead; the value (random gather) and the
    // stored state at run heads (another gather), which need the item, one round ahead.  Items are read up
    // to a.n, not hi: the first item after the range tells the last lane whether its run ends here.
    struct Stage { V v; bool act, head; u32 fl; u64 smax; };
    auto load_item = [&](size_t pos) { const size_t p = pos + lane; return p < a.n ? a.sorted[p] : ~0ull; };
    auto classify = [&](u64 item, size_t pos, u32 prev_last_slot) {
        Stage s;
        const size_t p = pos + lane;
        const u32 slot = (u32)(item >> 32);
        s.act = p < hi && slot != SLOT_INVALID;
        const u32 prev_slot = __shfl_up_sync(0xffffffffu, slot, 1);
        s.head = s.act && (lane == 0 ? (slot != prev_last_slot || p == 0) : (slot != prev_slot));
        s.v = s.act ? val[(u32)item] : V(0);
        s.fl = 0; s.smax = 0;
        if (s.head) { s.fl = a.kflags[slot]; s.smax = a.kmax[slot]; }
        return s;
    };
    u64 it_cur = load_item(lo), it_nxt = load_item(lo + 32);
    Stage cur = classify(it_cur, lo, c_slot);
    for (size_t pos = lo; pos < hi; pos += 32) {
        const u64 it_nn = load_item(pos + 64);                       // in flight for two rounds
        const u64 item = it_cur;
        const u32 slot = (u32)(item >> 32), idx = (u32)item;
        const bool act = cur.act, head = cur.head;
        const unsigned am0 = __ballot_sync(0xffffffffu, act);
        const u32 last_slot = am0 ? __shfl_sync(0xffffffffu, slot, 31 - __clz(am0)) : c_slot;
        const Stage nxt = classify(it_nxt, pos + 32, last_slot);      // its gathers are in flight for one round
        const u32 first_nxt = __shfl_sync(0xffffffffu, (u32)(it_nxt >> 32), 0);
        const unsigned hm = __ballot_sync(0xffffffffu, head);
        if (hm && !seen_head) { seen_head = true; lead = (u32)(pos - lo) + (u32)(__ffs(hm) - 1); }
        const V v = cur.v;
        u64 vb; memcpy(&vb, &v, 8);
        const u32 fl = cur.fl; const u64 smax = cur.smax;
        const bool over = act && N > 0 && cmp_op((double)v, a.consec_op, a.consec_thr);
        // ---- elements ----
        bool hr = false; u32 k = 0;
        if (N > 0 && act) { hr = !over; k = over ? 1u : 0u; if (head) { k = over ? (fl & KF_COUNT_MASK) + 1 : 0; hr = true; } }
        u64 mkey = act ? Codec<V>::cmp_key(v) : 0, mbits = vb; bool seg = head;        // seg = "a head lies at or before me in this round"
        if (rolling && head && (fl & KF_HAS_MAX)) {
            const u64 sk = Codec<V>::cmp_key(Codec<V>::from_bits(smax));
            if (sk > mkey) { mkey = sk; mbits = smax; }
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const bool hr_p = __shfl_up_sync(0xffffffffu, (int)hr, o) != 0; const u32 k_p = __shfl_up_sync(0xffffffffu, k, o);
            const u64 mk_p = __shfl_up_sync(0xffffffffu, mkey, o), mb_p = __shfl_up_sync(0xffffffffu, mbits, o);
            const bool seg_p = __shfl_up_sync(0xffffffffu, (int)seg, o) != 0;
            if ((int)lane >= o && !seg) {                          // my sub-run does not start with a head: prepend the left neighbour's
                if (!hr) { k += k_p; hr = hr_p; }
                if (mk_p > mkey) { mkey = mk_p; mbits = mb_p; }
                seg = seg_p;
            }
        }
        // prepend the function carried from the previous round for lanes whose run began before this round
        bool known_c = seg, known_m = seg;                          // is my state a known value (not a function of an unknown)?
        if (act && !seg) {
            if (!hr) { k += c.k; hr = c.hr; }
            if (c.hasm && c.mkey > mkey) { mkey = c.mkey; mbits = c.mbits; }
            known_c = hr; known_m = c.mc;
        }
        // Outputs are produced only for runs whose head this warp has seen (or, in kernel C, whose entering
        // state was handed in): known_m.  known_c may hold earlier (an under-threshold record resets the
        // counter whatever preceded it) and is what the carried function needs.
        u32 state = k; bool alert = false;
        if (N > 0 && act && known_c) {
            if (a.consec_reset) { state = k % N; alert = over && state == 0; }         // the counter just reached N (and went back to 0)
            else { if (state > KF_COUNT_MASK) state = KF_COUNT_MASK; alert = state >= N; }
        }
        const bool emit = act && known_m;
        if (N > 0) {
            const u64 row = warp_append(alert && emit, &a.st->n_calert, a.ca_cap, a.st);   // one reservation per warp round
            if (row != ~0ull) { a.ca_ts[row] = a.ts[idx]; a.ca_key[row] = a.key[idx]; a.ca_val[row] = vb; }
            nalert += (alert && emit) ? 1 : 0;
        }
        if (rolling && emit) a.rolling_out[idx] = mbits;
        // run tails write the key's state back
        const u32 next_slot = __shfl_down_sync(0xffffffffu, slot, 1);
        const bool tail = emit && (lane == 31 ? first_nxt : next_slot) != slot;
        if (tail) {
            a.kflags[slot] = (rolling ? KF_HAS_MAX : 0u) | KF_SEEN | (state & KF_COUNT_MASK);
            if (rolling) a.kmax[slot] = mbits;
        }
        // carry = function/state after the last active lane of this round
        const unsigned am = __ballot_sync(0xffffffffu, act);
        if (am) {
            const int last = 31 - __clz(am);
            const bool l_seg = __shfl_sync(0xffffffffu, (int)seg, last) != 0;
            c.hr = __shfl_sync(0xffffffffu, (int)(hr), last) != 0 || l_seg;
            c.k = __shfl_sync(0xffffffffu, known_c ? state : k, last);
            c.mc = l_seg || c.mc; c.hasm = rolling;
            c.mkey = __shfl_sync(0xffffffffu, mkey, last); c.mbits = __shfl_sync(0xffffffffu, mbits, last);
            c_slot = __shfl_sync(0xffffffffu, slot, last);
        }
        it_cur = it_nxt; it_nxt = it_nn; cur = nxt;
    }
}

// The consecutive-readings rule alone (items carry KI_OVER): the same contract as keyed_scan_range, but a
// round of 32 records needs no scan at all.  With Z = lanes whose reading is NOT over the threshold and
// H = run heads (both as ballots), lane i's counter is
//     i - z                      if the last not-over lane z <= i lies at or after the last head h <= i
//     stored(h) + (i - h + 1)    if the last head lies after it (lanes h..i are all over)
//     carried + (i + 1)          if the round has neither up to lane i
// i.e. two ballots, two count-leading-zeros and one shuffle instead of five scan steps, and no value is
// read except for the rare alert rows.
template <typename V>
__device__ __forceinline__ void consec_scan_range(const KeyedArgs& a, size_t lo, size_t hi, KCarry& c, u32& lead, u64& nalert) {
    const V* __restrict__ val = static_cast<const V*>(a.val);
    const u32 lane = threadIdx.x & 31;
    const u32 N = (u32)a.consec_n;
    const u32 le = 0xFFFFFFFFu >> (31 - lane);                        // lanes <= me
    u32 c_slot = lo > 0 ? (u32)(a.sorted[lo - 1] >> 32) : SLOT_INVALID;
    bool seen_head = false; lead = (u32)(hi - lo);
    // Software pipeline, two deep: items are requested two rounds ahead, the stored state of run heads
    // (a gather that needs the item) one round ahead.  Items are read up to a.n, not hi: the first item
    // after the range tells the last lane whether its run ends here.
    auto load_item = [&](size_t pos) { const size_t p = pos + lane; return p < a.n ? a.sorted[p] : ~0ull; };
    auto classify = [&](u64 item, size_t pos, u32 prev_last_slot, bool& act, bool& head, u32& fl) {
        const size_t p = pos + lane;
        const u32 slot = (u32)(item >> 32);
        act = p < hi && slot != SLOT_INVALID;
        const u32 prev_slot = __shfl_up_sync(0xffffffffu, slot, 1);
        head = act && (lane == 0 ? (slot != prev_last_slot || p == 0) : (slot != prev_slot));
        fl = head ? a.kflags[slot] : 0u;
    };
    u64 it_cur = load_item(lo), it_nxt = load_item(lo + 32);
    bool act, head; u32 fl;
    classify(it_cur, lo, c_slot, act, head, fl);
    for (size_t pos = lo; pos < hi; pos += 32) {
        const u64 it_nn = load_item(pos + 64);                       // in flight for two rounds
        const u32 slot = (u32)(it_cur >> 32), idx = (u32)it_cur & ~KI_OVER;
        const bool over = act && ((u32)it_cur & KI_OVER);
        const unsigned am = __ballot_sync(0xffffffffu, act);         // a prefix of the lanes: invalid items sort last
        const u32 last_slot = am ? __shfl_sync(0xffffffffu, slot, 31 - __clz(am)) : c_slot;
        bool act_n, head_n; u32 fl_n;
        classify(it_nxt, pos + 32, last_slot, act_n, head_n, fl_n);  // its state gather is in flight for one round
        const unsigned hm = __ballot_sync(0xffffffffu, head);
        const unsigned zm = __ballot_sync(0xffffffffu, act && !over);
        if (hm && !seen_head) { seen_head = true; lead = (u32)(pos - lo) + (u32)(__ffs(hm) - 1); }
        const int z = 31 - __clz((int)(zm & le)), h = 31 - __clz((int)(hm & le));          // -1 if none
        const u32 stored = __shfl_sync(0xffffffffu, fl & KF_COUNT_MASK, h < 0 ? 0 : h);
        u32 k; bool known_c;
        if (z >= 0 && z >= h) { k = lane - (u32)z; known_c = true; }
        else if (h >= 0)      { k = stored + (lane - (u32)h + 1); known_c = true; }
        else                  { k = c.k + lane + 1; known_c = c.hr; }
        const bool known_m = h >= 0 || c.mc;                          // a head of my run has been seen by this warp (or its state was handed in)
        u32 state = k; bool alert = false;
        if (act && known_c) {
            if (a.consec_reset) { state = k % N; alert = over && state == 0; }             // the counter just reached N (and went back to 0)
            else { if (state > KF_COUNT_MASK) state = KF_COUNT_MASK; alert = state >= N; }
        }
        const bool emit = act && known_m;
        const u64 row = warp_append(alert && emit, &a.st->n_calert, a.ca_cap, a.st);       // one reservation per warp round
        if (row != ~0ull) { a.ca_ts[row] = a.ts[idx]; a.ca_key[row] = a.key[idx]; memcpy(&a.ca_val[row], &val[idx], 8); }
        nalert += (alert && emit) ? 1 : 0;
        // run tails write the key's state back
        const u32 down_slot = __shfl_down_sync(0xffffffffu, slot, 1);
        const u32 first_nxt = __shfl_sync(0xffffffffu, (u32)(it_nxt >> 32), 0);
        const bool tail = emit && (lane == 31 ? first_nxt : down_slot) != slot;
        if (tail) a.kflags[slot] = KF_SEEN | (state & KF_COUNT_MASK);
        // carry = function/state after the last active lane of this round
        if (am) {
            const int last = 31 - __clz(am);
            c.hr = __shfl_sync(0xffffffffu, (int)known_c, last) != 0;
            c.k = __shfl_sync(0xffffffffu, known_c ? state : k, last);
            c.mc = c.mc || (hm & (0xFFFFFFFFu >> (31 - last))) != 0;
            c_slot = __shfl_sync(0xffffffffu, slot, last);
        }
        it_cur = it_nxt; it_nxt = it_nn; act = act_n; head = head_n; fl = fl_n;
    }
    c.hasm = false; c.mkey = 0; c.mbits = 0;
}

template <typename V, bool GATHER>
__global__ void __launch_bounds__(256) keyed_scan_kernel(const __grid_constant__ KeyedArgs a) {
    const u32 lane = threadIdx.x & 31;
    const size_t range = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t lo = range * KA_RANGE;
    if (lo >= a.n) return;
    const size_t hi = (lo + KA_RANGE < a.n) ? lo + KA_RANGE : a.n;
    KCarry c{false, 0u, false, false, 0ull, 0ull};                 // identity: nothing known about what precedes
    u32 lead; u64 nalert = 0;
    if constexpr (GATHER) keyed_scan_range<V>(a, lo, hi, c, lead, nalert); else consec_scan_range<V>(a, lo, hi, c, lead, nalert);
    if (lane == 0) {
        a.r_lead[range] = lead;
        a.r_fn_k[range] = c.k; a.r_fn_flags[range] = (c.hr ? KFN_HR : 0u) | (c.mc ? KFN_MC : 0u) | (c.hasm ? KFN_HASM : 0u);
        a.r_fn_mkey[range] = c.mkey; a.r_fn_mbits[range] = c.mbits;
    }
    nalert = warp_sum_u64(nalert);
    if (lane == 0 && nalert) atomicAdd(&a.st->consec_alerts, nalert);
}

// state entering range r = E[r-1] o ... o E[0]; one block, each thread owns a contiguous chunk of ranges.
// Composition of the range functions is associative: the 1024 chunk functions are scanned with warp
// shuffles (a first version let thread 0 walk them: 65 us per 8192 ranges, 9 % of the whole keyed chain).
struct KFn { u32 k, fl; u64 mk, mb; };
__device__ __forceinline__ void kfn_then(KFn& f, const KFn& g, u32 N, int reset) {                       // f := "f, then g"
    if (g.fl & KFN_HR) { f.k = g.k; f.fl |= KFN_HR; } else { f.k += g.k; if (reset && N) f.k %= N; }
    if (g.fl & KFN_MC) { f.mk = g.mk; f.mb = g.mb; f.fl |= KFN_MC | KFN_HASM; }
    else if (g.fl & KFN_HASM) { if (!(f.fl & KFN_HASM) || !(f.mk > g.mk)) { f.mk = g.mk; f.mb = g.mb; } f.fl |= KFN_HASM; }
}
__device__ __forceinline__ KFn kfn_shfl_up(const KFn& f, int o) {
    KFn r; r.k = __shfl_up_sync(0xffffffffu, f.k, o); r.fl = __shfl_up_sync(0xffffffffu, f.fl, o);
    r.mk = __shfl_up_sync(0xffffffffu, f.mk, o); r.mb = __shfl_up_sync(0xffffffffu, f.mb, o);
    return r;
}
__global__ void __launch_bounds__(1024) keyed_carry_kernel(const KeyedArgs a, u32 nranges, u32 N, int reset) {
    __shared__ KFn s_w[32];
    const u32 t = threadIdx.x, lane = t & 31, w = t >> 5;
    const u32 per = (nranges + 1023) / 1024;
    const u32 r0 = t * per < nranges ? t * per : nranges, r1 = (r0 + per < nranges) ? r0 + per : nranges;
    KFn mine{0, 0, 0, 0};                                           // identity
    for (u32 r = r0; r < r1; ++r) kfn_then(mine, KFn{a.r_fn_k[r], a.r_fn_flags[r], a.r_fn_mkey[r], a.r_fn_mbits[r]}, N, reset);
    // inclusive scan inside the warp
    KFn inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        KFn left = kfn_shfl_up(inc, o);
        if ((int)lane >= o) { kfn_then(left, inc, N, reset); inc = left; }
    }
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    if (w == 0) {                                                   // exclusive scan of the 32 warp totals
        KFn tot = s_w[lane], x = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            KFn left = kfn_shfl_up(x, o);
            if ((int)lane >= o) { kfn_then(left, x, N, reset); x = left; }
        }
        KFn ex = kfn_shfl_up(x, 1);
        if (lane == 0) ex = KFn{0, 0, 0, 0};
        s_w[lane] = ex;
    }
    __syncthreads();
    KFn f = s_w[w];                                                 // everything before my warp ...
    KFn prev = kfn_shfl_up(inc, 1);
    if (lane > 0) kfn_then(f, prev, N, reset);                      // ... then the lanes before me
    for (u32 r = r0; r < r1; ++r) {
        a.r_in_c[r] = f.k; a.r_in_mkey[r] = f.mk; a.r_in_mbits[r] = f.mb;
        kfn_then(f, KFn{a.r_fn_k[r], a.r_fn_flags[r], a.r_fn_mkey[r], a.r_fn_mbits[r]}, N, reset);
    }
}

template <typename V, bool GATHER>
__global__ void __launch_bounds__(256) keyed_lead_kernel(const __grid_constant__ KeyedArgs a) {
    const u32 lane = threadIdx.x & 31;
    const size_t range = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t lo = range * KA_RANGE;
    if (lo >= a.n) return;
    const u32 lead_len = a.r_lead[range];
    if (lead_len == 0) return;
    KCarry c{true, a.r_in_c[range], true, a.rolling_enabled != 0, a.r_in_mkey[range], a.r_in_mbits[range]};   // the entering state is a constant
    u32 lead; u64 nalert = 0;
    if constexpr (GATHER) keyed_scan_range<V>(a, lo, lo + lead_len, c, lead, nalert); else consec_scan_range<V>(a, lo, lo + lead_len, c, lead, nalert);
    nalert = warp_sum_u64(nalert);
    if (lane == 0 && nalert) atomicAdd(&a.st->consec_alerts, nalert);
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
