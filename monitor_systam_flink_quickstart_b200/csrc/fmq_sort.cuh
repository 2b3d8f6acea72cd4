// Stable multi-split ("one radix pass") used three ways:
//   * keyBy exchange (A4): split a batch into `world` owner buckets before the all-to-all;
//   * keyed ValueState path (A7): LSD radix sort of (key slot -> arrival index) so that every
//     key's records are contiguous AND in arrival order;
// A block owns a contiguous chunk of the input: counts[digit][chunk] is scanned digit-major, and the block
// then sorts its chunk in shared memory and copies every digit's run out in one piece.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

constexpr int RS_DIGITS = 256;                  // largest world an owner split supports
constexpr u32 XS_CHUNK = 2048;                  // events per block
constexpr int XS_THREADS = 256;                 // counting kernel, and the scatter when it shares the SMs with the ingest (keyBy exchange)
constexpr int XS_WARPS = XS_THREADS / 32;
constexpr int XS_IPT = XS_CHUNK / XS_THREADS;   // 8 events per thread
constexpr int XS_THREADS_WIDE = 512;            // the scatter by itself: 4 events per thread, all three columns requested at once

__host__ __device__ inline u32 xs_num_chunks(size_t n) { return (u32)((n + XS_CHUNK - 1) / XS_CHUNK); }

// keyBy routing on the GPU (A4): hashed keys -> mix64(key) % world (mode 0); dense ids -> id % world (mode 1).
// Mode 2 is not routing but locality: `world` equal ranges of the hash, i.e. of the key directory's home buckets
// (dir_home_bucket = mulhi(hash, buckets) is monotone in the hash) and so of the accumulator slots behind them.
constexpr u32 SPLIT_HASHED = 0, SPLIT_DENSE = 1, SPLIT_HASH_RANGE = 2;
// x % w for w < 2^16 without a 64-bit division (a run-time 64-bit divisor costs ~100 instructions per event: the
// counting kernel ran at 88 us per 2^24 keys for a 21 us read): a mask when w is a power of two, else three 32-bit
// remainders — x = hi * 2^32 + lo, so x % w = ((hi % w) * (2^32 % w) + lo % w) % w, every product below 2^32.
__device__ __forceinline__ u32 mod_small(u64 x, u32 w) {
    if ((w & (w - 1u)) == 0u) return (u32)x & (w - 1u);
    const u32 c = (0xFFFFFFFFu % w + 1u) % w;                             // 2^32 % w
    return (((u32)(x >> 32) % w) * c + (u32)x % w) % w;
}
__device__ __forceinline__ u64 div_small(u64 x, u32 w) {                  // dense id -> the owner's local id
    if ((w & (w - 1u)) == 0u) return x >> (31 - __clz((int)w));
    return (x >> 32) ? x / w : (u64)((u32)x / w);
}
__device__ __forceinline__ u32 owner_digit(u64 key, u32 world, u32 mode) {
    if (mode == SPLIT_HASH_RANGE) return (u32)__umul64hi(mix64(key), (u64)world);
    return mod_small(mode == SPLIT_DENSE ? key : mix64(key), world);
}

// Lanes of `am` holding the same BITS-bit digit, from BITS ballots.  MATCH.ANY does this in one
// instruction but was measured at ~150 cycles per warp instruction per SM on B200; votes are cheap.
// ALL 32 lanes call this, converged (lanes outside `am` pass any digit; their result is meaningless): a vote or
// shuffle under a run-time mask inside a divergent branch compiles to WARPSYNC + MATCH.ANY on the mask itself —
// the slow instruction again, 16 times per chunk (found in the SASS of the scatter kernels).
template <int BITS>
__device__ __forceinline__ unsigned match_digit(unsigned am, u32 d) {
    unsigned m = am;
#pragma unroll
    for (int b = 0; b < BITS; ++b) {
        const unsigned v = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        m &= ((d >> b) & 1u) ? v : ~v;
    }
    return m;
}

// Counting half of the owner split: one block per chunk of XS_CHUNK events, counts[digit][chunk].
// Every one of the 2^BITS x nchunks counters is STORED (nothing has to start at zero).
template <int BITS>
__global__ void __launch_bounds__(XS_THREADS) xs_hist_kernel(const u64* __restrict__ key, size_t n, u32 nchunks, u32 world, u32 dense, u32* __restrict__ counts) {
    constexpr u32 NB = 1u << BITS;
    __shared__ u32 cnt[NB];
    const u32 tid = threadIdx.x;
    const u32 chunk = blockIdx.x;
    for (u32 d = tid; d < NB; d += XS_THREADS) cnt[d] = 0;
    __syncthreads();
    const size_t lo = (size_t)chunk * XS_CHUNK;
    const size_t hi = (lo + XS_CHUNK < n) ? lo + XS_CHUNK : n;
    u64 kk[XS_IPT];
#pragma unroll
    for (int r = 0; r < XS_IPT; ++r) { const size_t i = lo + (size_t)r * XS_THREADS + tid; kk[r] = i < hi ? ldg_stream(key + i) : 0; }
    if constexpr (BITS == 3) {
        // 8 owners (one NVSwitch node): every thread counts its own 8 events in eight 16-bit fields of two registers, the
        // warp adds them up with shuffles (32 x 8 = 256 fits a field) and lane 0 makes the 8 shared-memory additions
        u64 c0 = 0, c1 = 0;
#pragma unroll
        for (int r = 0; r < XS_IPT; ++r) {
            if (lo + (size_t)r * XS_THREADS + tid < hi) {
                const u32 d = owner_digit(kk[r], world, dense);
                const u64 one = 1ull << ((d & 3u) * 16u);
                if (d & 4u) c1 += one; else c0 += one;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { c0 += __shfl_xor_sync(0xffffffffu, c0, o); c1 += __shfl_xor_sync(0xffffffffu, c1, o); }
        if ((tid & 31) == 0) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const u32 x0 = (u32)(c0 >> (16 * b)) & 0xFFFFu, x1 = (u32)(c1 >> (16 * b)) & 0xFFFFu;
                if (x0) atomicAdd(&cnt[b], x0);
                if (x1) atomicAdd(&cnt[4 + b], x1);
            }
        }
    } else
#pragma unroll
    for (int r = 0; r < XS_IPT; ++r) {
        const bool act = lo + (size_t)r * XS_THREADS + tid < hi;
        const unsigned am = __ballot_sync(0xffffffffu, act);
        const u32 d = act ? owner_digit(kk[r], world, dense) : 0u;
        const unsigned m = match_digit<BITS>(am, d);
        if (act && (int)(tid & 31) == __ffs(m) - 1) atomicAdd(&cnt[d], (u32)__popc(m));   // one shared atomic per distinct digit of the warp round
    }
    __syncthreads();
    for (u32 d = tid; d < NB; d += XS_THREADS) counts[(size_t)d * nchunks + chunk] = cnt[d];
}

// Scatter half: the block sorts its chunk by owner in shared memory (stable: warp w ranks its 256 consecutive
// events round by round, a cross-warp prefix per owner orders the warps) and copies every owner's run out with
// consecutive threads writing consecutive addresses — 2 KB runs per column at world = 8 instead of lone 8-byte
// stores.  The destination of owner d is either one shared set of output columns (plain partition: the buckets lie
// one after the other) or that owner's OWN columns (keyBy exchange: dst[d] is a segment of rank d's receive
// buffer, local or mapped over NVLink — the scatter IS the transfer).  Dense ids become the owner's local id
// (id / world).  The largest timestamp of the batch is reduced on the way (one atomic per block).
struct SplitDst {
    u64* ts[FMQ_MAX_PEERS]; size_t col_stride;     // per-owner mode: key column = ts + col_stride, value column = ts + 2 * col_stride
    u64* out_ts; u64* out_key; u64* out_val;       // single-output mode
};
struct SplitArgs {
    const u64* ts; const u64* key; const u64* val; size_t n; u32 nchunks; u32 world; u32 dense;
    const u32* offsets;                            // counts[digit][chunk] after the exclusive scan (digit-major); unused when `fill` is set
    SplitDst dst; i64* max_ts;                     // max_ts may be null
    // Unordered mode (per-owner destinations only): no counting pass and no scan; a block reserves the place of each of
    // its runs with one atomic on fill[owner].  Runs of different blocks then land in completion order, which is
    // all a commutative aggregate needs; jobs with order-dependent keyed state use the stable two-pass form.
    unsigned long long* fill;
};
// (Tried: the three columns as TMA bulk copies into shared memory and a 16-bit source index per sorted position, as in
// ps_scatter_tile_kernel.  By itself the kernel did not change — 175 us per 2^24 events either way once the owner digit
// lost its 64-bit division — and beside the ingest kernel (keyBy exchange, N = 2, same box) the step went from 0.533 to
// 0.576 ms: 512 threads polling an mbarrier take issue slots from the ingest warps.  Kept: per-thread loads.)
constexpr size_t xs_scatter_smem(u32 nb, int threads) { return (size_t)3 * XS_CHUNK * 8 + XS_CHUNK * 2 + (size_t)(threads / 32) * nb * 2 + nb * 4 * 2; }

template <int BITS, bool PER_OWNER, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? 4 : 2) xs_scatter_kernel(const __grid_constant__ SplitArgs a) {
    constexpr int WARPS = THREADS / 32, IPT = (int)XS_CHUNK / THREADS;
    constexpr bool EARLY = THREADS > 256;         // enough registers to hold all three columns through the ranking
    constexpr u32 NB = 1u << BITS;
    extern __shared__ __align__(16) unsigned char xs_smem[];
    u64* ev_s = reinterpret_cast<u64*>(xs_smem);                                  // [3][XS_CHUNK]
    unsigned short* dig_s = reinterpret_cast<unsigned short*>(ev_s + 3 * XS_CHUNK);   // [XS_CHUNK] owner of every sorted position
    unsigned short* wcnt = dig_s + XS_CHUNK;                                      // [WARPS][NB] per-warp counts, then exclusive prefixes
    u32* dbase = reinterpret_cast<u32*>(wcnt + WARPS * NB);                    // [NB] start of the owner's run inside the sorted chunk
    u32* goff = dbase + NB;                                                       // [NB] where that run goes in the owner's destination
    __shared__ u32 scan_ws[WARPS];
    __shared__ i64 max_ws[WARPS];
    const u32 tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    i64 tmax = I64_MIN_;
    // persistent over chunks: the grid may be much smaller than the number of chunks (the keyBy exchange runs this kernel
    // beside the ingest of the previous batch and leaves it most of every SM)
    for (u32 chunk = blockIdx.x; chunk < a.nchunks; chunk += gridDim.x) {
    const size_t lo = (size_t)chunk * XS_CHUNK;
    const size_t hi = (lo + XS_CHUNK < a.n) ? lo + XS_CHUNK : a.n;
    const bool unordered = PER_OWNER && a.fill != nullptr;
    for (u32 i = tid; i < WARPS * NB; i += THREADS) wcnt[i] = 0;
    if (!unordered) for (u32 d = tid; d < NB; d += THREADS) {
        const u32 o = a.offsets[(size_t)d * a.nchunks + chunk];
        goff[d] = PER_OWNER ? o - a.offsets[(size_t)d * a.nchunks] : o;           // per-owner destinations start at 0
    }
    __syncthreads();
    // phase 1: rank inside the warp's 256 consecutive events, in arrival order
    // all three columns of the thread's events are requested at once (one memory round trip per chunk, not two)
    u64 kk[IPT], tt[EARLY ? IPT : 1], vv[EARLY ? IPT : 1]; unsigned short dg[IPT], rk[IPT];
    const size_t wlo = lo + (size_t)w * (XS_CHUNK / WARPS);
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        const size_t i = wlo + r * 32 + lane; const bool in = i < hi;
        kk[r] = in ? ldg_stream(a.key + i) : 0;
        if constexpr (EARLY) { tt[r] = in ? ldg_stream(a.ts + i) : 0; vv[r] = in ? ldg_stream(a.val + i) : 0; }
    }
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        const bool act = wlo + r * 32 + lane < hi;
        const unsigned am = __ballot_sync(0xffffffffu, act);
        const u32 d = act ? owner_digit(kk[r], a.world, a.dense) : 0u;           // (every lane takes part in the votes and the shuffle)
        const unsigned m = match_digit<BITS>(am, d);
        const int leader = (__ffs(m) - 1) & 31;
        u32 base = 0;
        if (act && (int)lane == leader) { base = wcnt[w * NB + d]; wcnt[w * NB + d] = (unsigned short)(base + __popc(m)); }
        base = __shfl_sync(0xffffffffu, base, leader);
        dg[r] = (unsigned short)d; rk[r] = act ? (unsigned short)(base + __popc(m & ((1u << lane) - 1))) : (unsigned short)0;
        __syncwarp();
    }
    __syncthreads();
    // phase 2: per owner, exclusive prefix over the warps; then exclusive prefix over the owners
    constexpr u32 DPT = NB / THREADS ? NB / THREADS : 1;
    u32 tot_local = 0, dtot[DPT];
#pragma unroll
    for (u32 q = 0; q < DPT; ++q) {
        const u32 d = tid * DPT + q;
        u32 run = 0;
        if (d < NB) for (int ww = 0; ww < WARPS; ++ww) { const u32 c = wcnt[ww * NB + d]; wcnt[ww * NB + d] = (unsigned short)run; run += c; }
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
    for (u32 q = 0; q < DPT; ++q) {
        const u32 d = tid * DPT + q;
        if (d < NB) {
            dbase[d] = run;
            if (unordered) goff[d] = dtot[q] ? (u32)atomicAdd(&a.fill[d], (unsigned long long)dtot[q]) : 0u;   // this block's run of owner d
        }
        run += dtot[q];
    }
    __syncthreads();
    // phase 3: the three columns of every event go to its sorted position
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        const size_t i = wlo + r * 32 + lane;
        if (i < hi) {
            const u32 pos = dbase[dg[r]] + wcnt[w * NB + dg[r]] + rk[r];
            u64 t, v; const u64 k = kk[r];
            if constexpr (EARLY) { t = tt[r]; v = vv[r]; } else { t = ldg_stream(a.ts + i); v = ldg_stream(a.val + i); }
            tmax = (i64)t > tmax ? (i64)t : tmax;
            ev_s[pos] = t;
            ev_s[XS_CHUNK + pos] = (a.dense == SPLIT_DENSE && k != KEY_EMPTY) ? div_small(k, a.world) : k;
            ev_s[2 * XS_CHUNK + pos] = v;
            dig_s[pos] = dg[r];
        }
    }
    __syncthreads();
    // phase 4: copy out; consecutive threads -> consecutive positions -> consecutive addresses inside a run
    const u32 cnt = (u32)(hi - lo);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        for (u32 pos = tid; pos < cnt; pos += THREADS) {
            const u32 d = dig_s[pos];
            const size_t o = (size_t)goff[d] + (pos - dbase[d]);
            u64* dstc = PER_OWNER ? a.dst.ts[d] + (size_t)c * a.dst.col_stride : (c == 0 ? a.dst.out_ts : (c == 1 ? a.dst.out_key : a.dst.out_val));
            dstc[o] = ev_s[c * XS_CHUNK + pos];
        }
    }
    __syncthreads();                               // the staging area is reused by the next chunk
    }
    if (a.max_ts) {                                // one atomic per block for the batch's largest timestamp
        tmax = warp_max_i64(tmax); if (lane == 0) max_ws[w] = tmax;
        __syncthreads();
        if (tid == 0) {
            i64 m = max_ws[0];
            for (int i = 1; i < WARPS; ++i) m = max_ws[i] > m ? max_ws[i] : m;
            if (m != I64_MIN_) atomicMax(reinterpret_cast<long long*>(a.max_ts), (long long)m);
        }
    }
    // stores to a peer's memory must have landed before a later kernel of this stream raises that peer's flag
    if (PER_OWNER) __threadfence_system();
}

// ---- exclusive scan over u32 (three small kernels; the counter matrix is tiny next to the data) ----
constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_IPT = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;

__device__ __forceinline__ u32 block_exclusive_scan(u32 v, u32* total) {
    __shared__ u32 wsum[32];
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = v;
    for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += t; }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        u32 s = wsum[lane];
        for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, s, o); if ((int)lane >= o) s += t; }
        wsum[lane] = s;
    }
    __syncthreads();
    const u32 wbase = w ? wsum[w - 1] : 0;
    if (total) *total = wsum[31];
    __syncthreads();
    return wbase + inc - v;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums(const u32* __restrict__ in, size_t n, u32* __restrict__ tile_sums) {
    const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_IPT;
    u32 s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; ++k) if (base + k < n) s += in[base + k];
    u32 total; block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
// single block: in-place exclusive scan of up to SCAN_TILE * k tile sums
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_offsets(u32* __restrict__ tile_sums, u32 ntiles) {
    __shared__ u32 carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (u32 b = 0; b < ntiles; b += SCAN_THREADS) {
        const u32 i = b + threadIdx.x;
        const u32 v = i < ntiles ? tile_sums[i] : 0;
        u32 total; const u32 ex = block_exclusive_scan(v, &total);
        const u32 carry = carry_s;
        if (i < ntiles) tile_sums[i] = ex + carry;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply(u32* __restrict__ data, size_t n, const u32* __restrict__ tile_offsets) {
    const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_IPT;
    u32 v[SCAN_IPT]; u32 s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; ++k) { v[k] = (base + k < n) ? data[base + k] : 0; s += v[k]; }
    u32 ex = block_exclusive_scan(s, nullptr) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_IPT; ++k) { if (base + k < n) data[base + k] = ex; ex += v[k]; }
}

}  // namespace fmq
