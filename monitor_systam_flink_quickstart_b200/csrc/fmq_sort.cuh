// Stable multi-split ("one radix pass") used three ways:
//   * keyBy exchange (A4): split a batch into `world` owner buckets before the all-to-all;
//   * keyed ValueState path (A7): LSD radix sort of (key slot -> arrival index) so that every
//     key's records are contiguous AND in arrival order;
// A warp owns a contiguous chunk of the input, so stability needs no cross-warp coordination:
// counts[digit][chunk] is scanned digit-major and each warp then walks its chunk in order.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

constexpr int RS_THREADS = 256;                 // 8 warps per block
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_DIGITS = 256;
constexpr u32 RS_CHUNK = 2048;                  // items per warp chunk

__host__ __device__ inline u32 rs_num_chunks(size_t n) { return (u32)((n + RS_CHUNK - 1) / RS_CHUNK); }

// DigitFn: u32 operator()(size_t i) -> [0, 256)
template <typename DigitFn>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(size_t n, u32 nchunks, DigitFn digit, u32* __restrict__ counts) {
    __shared__ u32 cnt[RS_WARPS][RS_DIGITS];
    const u32 wl = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const u32 chunk = blockIdx.x * RS_WARPS + wl;
    for (int d = lane; d < RS_DIGITS; d += 32) cnt[wl][d] = 0;
    __syncwarp();
    if (chunk < nchunks) {
        const size_t lo = (size_t)chunk * RS_CHUNK;
        const size_t hi = (lo + RS_CHUNK < n) ? lo + RS_CHUNK : n;
        for (size_t i0 = lo; i0 < hi; i0 += 32) {
            const size_t i = i0 + lane;
            const bool act = i < hi;
            const unsigned am = __ballot_sync(0xffffffffu, act);
            if (act) {
                const u32 d = digit(i);
                const unsigned m = __match_any_sync(am, d);
                if ((int)lane == __ffs(m) - 1) cnt[wl][d] += __popc(m);
            }
            __syncwarp();
        }
        for (int d = lane; d < RS_DIGITS; d += 32) counts[(size_t)d * nchunks + chunk] = cnt[wl][d];
    }
}

// Emit: void operator()(size_t src, size_t dst) — moves item src to position dst.
template <typename DigitFn, typename Emit>
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(size_t n, u32 nchunks, DigitFn digit, Emit emit, const u32* __restrict__ offsets) {
    __shared__ u32 off[RS_WARPS][RS_DIGITS];
    const u32 wl = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const u32 chunk = blockIdx.x * RS_WARPS + wl;
    if (chunk >= nchunks) return;
    for (int d = lane; d < RS_DIGITS; d += 32) off[wl][d] = offsets[(size_t)d * nchunks + chunk];
    __syncwarp();
    const size_t lo = (size_t)chunk * RS_CHUNK;
    const size_t hi = (lo + RS_CHUNK < n) ? lo + RS_CHUNK : n;
    for (size_t i0 = lo; i0 < hi; i0 += 32) {
        const size_t i = i0 + lane;
        const bool act = i < hi;
        const unsigned am = __ballot_sync(0xffffffffu, act);
        if (act) {
            const u32 d = digit(i);
            const unsigned m = __match_any_sync(am, d);
            const int leader = __ffs(m) - 1;
            u32 base = 0;
            if ((int)lane == leader) { base = off[wl][d]; off[wl][d] = base + __popc(m); }
            base = __shfl_sync(m, base, leader);
            emit(i, (size_t)base + __popc(m & ((1u << lane) - 1)));
        }
        __syncwarp();
    }
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

// ---- digit / emit functors -----------------------------------------------------------------------
struct OwnerDigit {      // keyBy routing on the GPU (A4): hashed keys -> mix64(key) % world; dense ids -> id % world
    const u64* key; u32 world; u32 dense;
    __device__ __forceinline__ u32 operator()(size_t i) const { return (u32)((dense ? key[i] : mix64(key[i])) % world); }
};
struct MoveEvent3 {      // moves one 24 B event (three 8 B columns); dense ids become the owner's local id (id / world)
    const u64* ts; const u64* key; const u64* val; u64* ts_o; u64* key_o; u64* val_o; u32 world; u32 dense;
    __device__ __forceinline__ void operator()(size_t s, size_t d) const {
        const u64 k = key[s];
        ts_o[d] = ts[s]; key_o[d] = (dense && k != KEY_EMPTY) ? k / world : k; val_o[d] = val[s];
    }
};
struct SlotDigit {       // one 8-bit digit of the key-directory slot
    const u32* slot; u32 shift;
    __device__ __forceinline__ u32 operator()(size_t i) const { return (slot[i] >> shift) & 0xFFu; }
};
struct MovePair {        // (slot, arrival index) pair
    const u32* slot; const u32* idx; u32* slot_o; u32* idx_o;
    __device__ __forceinline__ void operator()(size_t s, size_t d) const { slot_o[d] = slot[s]; idx_o[d] = idx[s]; }
};

}  // namespace fmq
