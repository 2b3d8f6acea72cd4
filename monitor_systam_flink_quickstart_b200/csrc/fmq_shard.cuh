// keyBy exchange between the GPUs of one NVSwitch node, entirely on the device.
//
// The reference's structure is a hash exchange in front of owner-held keyed state
// (chapter3/.../BandwidthMonitorWithEventTime.java:45 `.keyBy(1)` -> :46-47 window + reduce on the owning
// subtask).  Here every rank owns a receive buffer that all peers have mapped (CUDA IPC, or plain pointers inside
// one process): 2 parities x `world` segments, one per source rank, three 8-byte columns each.  Per micro-batch
//   SEND  (send stream)  wait until every peer has consumed the batch that last used this parity (ack flags) ->
//                        count owners per chunk -> scan -> xs_scatter_kernel writes every owner's run STRAIGHT INTO
//                        that owner's segment (local stores or NVLink peer stores: the scatter is the transfer) ->
//                        publish: per peer {count, largest timestamp of my slice}, fence, then the batch number;
//   RECV  (main stream)  gather: spin on the `world` batch numbers in MY memory, read the counts and the largest
//                        timestamp of the whole batch (one watermark for the job: the reference has ONE assigner in
//                        front of the keyBy, BandwidthMonitorWithEventTime.java:30-45) -> ingest / keyed chain / plan /
//                        fire read the segments in source-rank order with device-side counts -> ack to every source.
// No collective library, no host round trip: the host only queues kernels.  Spins give up after a timeout and raise
// a sticky error instead of hanging the GPU.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

constexpr u32 ERRB_PEER = 16u;                       // a peer did not publish / acknowledge in time

struct XPub { u64 seq; u64 count; i64 max_ts; u64 pad; };              // written by source rank s into pub[parity][s] of the receiver
struct XCtl {
    XPub pub[2][FMQ_MAX_PEERS];
    u64 ack[2][FMQ_MAX_PEERS];                       // ack[parity][d]: receiver d has consumed my batch `value` of that parity
    // combiner mode: batch scalars {seq, max ts, -(min pane), max pane} of every rank, and "done reading" flags
    i64 cmb[2][FMQ_MAX_PEERS][4];
    u64 done[2][FMQ_MAX_PEERS];
};
constexpr size_t XCTL_BYTES = 4096;
static_assert(sizeof(XCtl) <= XCTL_BYTES, "control block");

__device__ __forceinline__ u64 ld_acquire_sys(const u64* p) {
    u64 v; asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_release_sys(u64* p, u64 v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 global_timer_ns() { u64 t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

// spins until *p >= want (flags only ever grow); false on timeout
__device__ __forceinline__ bool spin_until_ge(const u64* p, u64 want, u64 timeout_ns) {
    if (ld_acquire_sys(p) >= want) return true;
    const u64 t0 = global_timer_ns();
    for (;;) {
        for (int i = 0; i < 64; ++i) { if (ld_acquire_sys(p) >= want) return true; __nanosleep(200); }
        if (global_timer_ns() - t0 > timeout_ns) return false;
    }
}

// SEND step 0: every peer has taken the batch that last used this parity out of its segment of mine
__global__ void xs_wait_acks_kernel(const XCtl* ctl, u32 world, u32 parity, u64 want, DevState* st, u64 timeout_ns) {
    const u32 d = threadIdx.x;
    if (d < world && want > 0 && !spin_until_ge(&ctl->ack[parity][d], want, timeout_ns)) atomicOr(&st->err_flags, ERRB_PEER);
}

// SEND last step: tell every owner how many events of mine lie in its segment, and my slice's largest timestamp
struct XPeers { XCtl* ctl[FMQ_MAX_PEERS]; };
__global__ void xs_publish_kernel(const u32* __restrict__ offsets, u32 nchunks, u64 n, unsigned long long* fill, u32 world, u32 rank, u32 parity, u64 seq, i64* max_ts, XPeers peers) {
    const u32 d = threadIdx.x;
    const i64 m = *max_ts;
    __syncwarp();
    if (d < world) {
        u64 cnt;
        if (fill) { cnt = fill[d]; fill[d] = 0; }                      // unordered scatter: the owners' fill counters are the counts
        else { const u64 lo = n ? offsets[(size_t)d * nchunks] : 0, hi = n ? (d + 1 < world ? offsets[(size_t)(d + 1) * nchunks] : n) : 0; cnt = hi - lo; }
        XPub* p = &peers.ctl[d]->pub[parity][rank];
        p->count = cnt; p->max_ts = m;
        __threadfence_system();                      // the record, and the scatter kernel's stores before it in stream order, before the flag
        st_release_sys(&p->seq, seq);
    }
    __syncwarp();
    if (d == 0) *max_ts = I64_MIN_;                  // re-armed for the next batch
}

// RECV first step: wait for all `world` sources, hand their counts to the kernels that follow
__global__ void xs_gather_kernel(const XCtl* ctl, u32 world, u32 parity, u64 seq, u64* seg_n, u64* n_total, DevState* st, u64 timeout_ns) {
    const u32 s = threadIdx.x;
    u64 cnt = 0; i64 m = I64_MIN_;
    if (s < world) {
        if (spin_until_ge(&ctl->pub[parity][s].seq, seq, timeout_ns)) { cnt = ctl->pub[parity][s].count; m = ctl->pub[parity][s].max_ts; }
        else atomicOr(&st->err_flags, ERRB_PEER);
        seg_n[s] = cnt;
    }
    m = warp_max_i64(m); cnt = warp_sum_u64(cnt);
    if (s == 0) {
        if (m > st->batch_max_ts) st->batch_max_ts = m;          // A1: the max over EVERY rank's slice, not only over what this rank received
        *n_total = cnt;
    }
}

// RECV last step: my segments of this parity are free again
__global__ void xs_ack_kernel(XPeers peers, u32 world, u32 rank, u32 parity, u64 seq) {
    const u32 s = threadIdx.x;
    if (s < world) { __threadfence_system(); st_release_sys(&peers.ctl[s]->ack[parity][rank], seq); }
}

// keyed ValueState operators want one contiguous batch: the segments one after the other, in source-rank order
// (= the order in which a Flink subtask drains its input channels in the restatement, oracle/cpu_parallel.cpp)
__global__ void __launch_bounds__(256) xs_compact_kernel(const u64* __restrict__ seg_base, size_t seg_stride, size_t col_stride, const u64* __restrict__ seg_n, u32 world,
                                                         u64* __restrict__ ts, u64* __restrict__ key, u64* __restrict__ val, size_t cap) {
    size_t start = 0;
    for (u32 s = 0; s < world; ++s) {
        const size_t n = seg_n[s];
        const u64* src = seg_base + (size_t)s * seg_stride;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n && start + i < cap; i += (size_t)gridDim.x * blockDim.x) {
            ts[start + i] = src[i]; key[start + i] = src[col_stride + i]; val[start + i] = src[2 * col_stride + i];
        }
        start += n;
    }
}

// ---- combiner mode without a collective library: the 32-byte all-reduce and the "done reading" barrier ----
__global__ void cmb_publish_kernel(const DevState* st, XPeers peers, u32 world, u32 rank, u32 parity, u64 seq) {
    const u32 d = threadIdx.x;
    if (d >= world) return;
    i64 lo = st->min_pane;
    if (st->anchored && lo < st->origin_pane) lo = st->origin_pane;      // panes before the origin have expired (see pack_scalars_kernel)
    const bool has = lo <= st->max_pane;
    i64* x = peers.ctl[d]->cmb[parity][rank];
    x[1] = st->batch_max_ts; x[2] = has ? -lo : I64_MIN_; x[3] = has ? st->max_pane : I64_MIN_;
    __threadfence_system();
    st_release_sys(reinterpret_cast<u64*>(&x[0]), seq);
}
// max over all ranks, into this rank's device state (what unpack_scalars_kernel did after the NCCL all-reduce)
__global__ void cmb_reduce_kernel(DevState* st, const XCtl* ctl, u32 world, u32 parity, u64 seq, u64 timeout_ns) {
    const u32 s = threadIdx.x;
    i64 a = I64_MIN_, b = I64_MIN_, c = I64_MIN_;
    if (s < world) {
        const i64* x = ctl->cmb[parity][s];
        if (spin_until_ge(reinterpret_cast<const u64*>(&x[0]), seq, timeout_ns)) { a = x[1]; b = x[2]; c = x[3]; }
        else atomicOr(&st->err_flags, ERRB_PEER);
    }
    a = warp_max_i64(a); b = warp_max_i64(b); c = warp_max_i64(c);
    if (s == 0) { st->batch_max_ts = a; if (b != I64_MIN_) { st->min_pane = -b; st->max_pane = c; } }
}
__global__ void cmb_done_kernel(XPeers peers, u32 world, u32 rank, u32 parity, u64 seq) {
    const u32 d = threadIdx.x;
    if (d < world) { __threadfence_system(); st_release_sys(&peers.ctl[d]->done[parity][rank], seq); }
}
__global__ void cmb_wait_done_kernel(const XCtl* ctl, u32 world, u32 parity, u64 seq, DevState* st, u64 timeout_ns) {
    const u32 s = threadIdx.x;
    if (s < world && !spin_until_ge(&ctl->done[parity][s], seq, timeout_ns)) atomicOr(&st->err_flags, ERRB_PEER);
}

}  // namespace fmq
