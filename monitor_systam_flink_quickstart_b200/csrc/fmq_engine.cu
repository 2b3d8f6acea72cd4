// libfmq_b200: host side of the C ABI declared in include/fmq_b200.h.
// One handle = one operator subtask = one CUDA stream + all device state.  No CPU fallback:
// without a usable device fmq_create fails with FMQ_ERR_NO_DEVICE.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <new>
#include <mutex>
#include "../../include/fmq_b200.h"
#include "fmq_device.cuh"
#include "fmq_window.cuh"
#include "fmq_sort.cuh"
#include "fmq_keyed.cuh"
#include "fmq_text.cuh"
#include "fmq_median.cuh"
#include "fmq_shard.cuh"

using namespace fmq;

#define FMQ_STR2(x) #x
#define FMQ_STR(x) FMQ_STR2(x)

static thread_local std::string g_create_err;

// the columns of one micro-batch as the caller hands them over
struct InCols { const void* ts; const void* key; const void* val; int ts_w, key_w, val_w; i64 ts_base; };

struct fmq_handle {
    fmq_config cfg;
    int device = 0, num_sms = 148;
    cudaStream_t stream = nullptr;
    // per-push stage timing: ring of event quadruples {start, after ingest, after keyed, end}
    static constexpr int TIMING_RING = 64;
    cudaEvent_t tev[TIMING_RING][4] = {};
    uint64_t pushes = 0;
    std::string err;
    int sticky = 0;                      // sticky fmq_status once the device state is invalid
    bool anchored_known = false;         // host knows a finite watermark is in force (device state is the authority)
    // geometry
    WindowGeom g{};
    u64 capacity = 0; u32 ring = 0; u32 slot_bits = 0;
    size_t max_batch = 0;
    // device state
    DevState* d_st = nullptr; DevState* h_st = nullptr;   // h_st: pinned mirror
    u32* h_hint = nullptr;                                // pinned word the plan kernel writes: skewed keys seen (picks the ingest variant)
    u64* d_keys = nullptr; Acc32* d_acc = nullptr; u32* d_kflags = nullptr; u64* d_kmax = nullptr;
    u32* d_dirty = nullptr;              // allowedLateness: one bit per (ring pane, key slot)
    // staging for host-resident inputs
    i64* d_in_ts = nullptr; u64* d_in_key = nullptr; u64* d_in_val = nullptr;
    // host pushes: two staging sets filled on a copy stream, so that the copy-in of batch k+1 runs
    // while the kernels of batch k do (allocated by the first host push)
    static constexpr int HOST_SETS = 3;          // one being consumed + up to two staged ahead (fmq_stage_host)
    i64* hs_ts[HOST_SETS] = {}; u64* hs_key[HOST_SETS] = {}; u64* hs_val[HOST_SETS] = {};
    u32* hs_narrow[HOST_SETS] = {};              // 3 x max_batch 4-byte columns of fmq_push_packed
    cudaStream_t copy_stream = nullptr; cudaEvent_t ev_free[HOST_SETS] = {}, ev_piece[HOST_SETS][4] = {}; uint64_t host_pushes = 0;
    struct Staged { InCols c{}; size_t n = 0, chunk = 0; int set = 0, pieces = 0; };
    Staged staged[2]; int n_staged = 0;          // FIFO of batches whose copy-in has been queued but not consumed
    // outputs
    u64 *f_key = nullptr, *f_sum = nullptr, *f_min = nullptr, *f_max = nullptr; i64 *f_wend = nullptr, *f_count = nullptr; u64 f_cap = 0;
    u64* wa_key = nullptr; i64* wa_wend = nullptr; double* wa_val = nullptr; u64 wa_cap = 0;
    i64* ea_ts = nullptr; u64 *ea_key = nullptr, *ea_val = nullptr; u64 ea_cap = 0;
    i64* ca_ts = nullptr; u64 *ca_key = nullptr, *ca_val = nullptr; u64 ca_cap = 0;
    u64* d_rolling = nullptr; size_t rolling_n = 0;
    size_t cur_fired = 0, cur_walert = 0, cur_ealert = 0, cur_calert = 0;   // host read cursors
    // keyed-path / partition scratch
    u64* d_pack[2] = {nullptr, nullptr};   // (slot << 32 | arrival index), ping-pong
    u32 *kr_lead = nullptr, *kr_k = nullptr, *kr_fl = nullptr, *kr_in_c = nullptr; u64 *kr_mk = nullptr, *kr_mb = nullptr, *kr_in_mk = nullptr, *kr_in_mb = nullptr;
    u32* d_counts = nullptr; u32* d_tiles = nullptr; size_t counts_len = 0, tiles_len = 0;
    uint64_t launches = 0;
    // combiner mode (fmq_attach_peers)
    int world = 1, rank = 0;
    const Acc32* peer_acc[FMQ_MAX_PEERS] = {};
    i64* d_xchg = nullptr;               // 4 x int64 exchange buffer for the all-reduce of the batch scalars
    // overlapped combiner mode (fmq_combiner_overlap): fire / retire run on fire_stream while the next batch is ingested
    bool overlap = false; cudaStream_t fire_stream = nullptr; cudaEvent_t ev_plan = nullptr, ev_bdone = nullptr;
    size_t pending_n = 0;                // events ingested by fmq_push_local and not yet accounted by fmq_fire_global
    // peer-pull keyBy exchange: two parities x three columns of owner-bucketed events, mapped by every peer
    char* d_xsend = nullptr; const char* peer_xsend[FMQ_MAX_PEERS] = {}; int x_world = 0, x_rank = 0;
    // keyBy exchange on the device (fmq_shard_*): receive buffer all peers map, control block at its start
    char* d_xbuf = nullptr; size_t xs_seg_cap = 0; int xs_world = 0, xs_rank = -1; bool xs_attached = false;
    XCtl* xs_ctl[FMQ_MAX_PEERS] = {}; char* xs_base[FMQ_MAX_PEERS] = {}; void* xs_opened[FMQ_MAX_PEERS] = {};
    cudaStream_t send_stream = nullptr; u32* d_xcounts = nullptr; u32* d_xtiles = nullptr; i64* d_send_max = nullptr;
    unsigned long long* d_xfill = nullptr;   // [FMQ_MAX_PEERS] events placed so far in each owner's segment (unordered scatter)
    u64* d_seg_n = nullptr;              // [FMQ_MAX_PEERS] counts of the batch being received + [1] their sum
    u64* h_ntotal = nullptr;             // pinned: the sum read back for the keyed chain (which sizes its launches on the host)
    u64* xs_stage = nullptr;             // 3 x seg_cap: host columns of a send are copied here first
    uint64_t xs_send_seq = 0, xs_recv_seq = 0;
    uint64_t cmb_seq = 0; bool cmb_flags = false;   // combiner mode over the same control block instead of a collective library
    // second set of fired / window-alert / event-alert columns + a copy-out stream, for overlapped polls
    void* alt_out[12] = {}; cudaStream_t out_stream = nullptr; cudaEvent_t ev_out = nullptr; bool out_pending = false;
    i64* lt_ts = nullptr; u64 *lt_key = nullptr, *lt_val = nullptr; u64 lt_cap = 0; size_t cur_late = 0;
    // median path (N3)
    i64* ml_pane[2] = {nullptr, nullptr}; u32* ml_slot[2] = {nullptr, nullptr}; u64* ml_val[2] = {nullptr, nullptr}; int ml_cur = 0; u64 ml_cap = 0;
    u64* d_loc[3] = {nullptr, nullptr, nullptr}; u32 loc_parts = 0;     // hash-range partition in front of the ingest (tables much larger than the L2)
    u64* mw_items[2] = {nullptr, nullptr}; u64* mw_vals = nullptr; u64* mw_big = nullptr; u64 mw_big_cap = 0;
    u64* md_key = nullptr; i64* md_wend = nullptr; double* md_val = nullptr; u64 md_cap = 0; size_t cur_median = 0;
    // text ingest
    char* d_text = nullptr; u64* d_line_start = nullptr; u32* d_txt_status = nullptr; unsigned long long* d_txt_bad = nullptr;
    // device dictionary (text_dict_capacity > 0): hash table, names, where every line's key field lies
    u64* dt_hash = nullptr; u32* dt_id = nullptr; u64 dt_mask = 0; u64* dn_off = nullptr; u32* dn_len = nullptr; char* dn_arena = nullptr; u64 dn_arena_cap = 0;
    unsigned long long* dt_counters = nullptr; u64* d_key_off = nullptr; u32* d_key_len = nullptr;
};

#define CU(h, call)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            (h)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                            \
            return FMQ_ERR_CUDA;                                                                      \
        }                                                                                             \
    } while (0)

static int set_err(fmq_handle* h, int st, const char* msg) { if (h) h->err = msg; else g_create_err = msg; return st; }

template <typename T> static cudaError_t dalloc(T** p, size_t count) { return cudaMalloc((void**)p, count * sizeof(T)); }

static bool keyed_enabled(const fmq_config& c) { return c.rolling_max_enabled || c.consec_n > 0; }

extern "C" {

int32_t fmq_abi_version(void) { return FMQ_ABI_VERSION; }
const char* fmq_build_info(void) { return "libfmq_b200 sm_100a (CUDA " FMQ_STR(CUDART_VERSION) ") hashed key directory + pane ring, no CPU fallback"; }

const char* fmq_last_error(fmq_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

static void free_all(fmq_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    void* ptrs[] = {h->dt_hash, h->dt_id, h->dn_off, h->dn_len, h->dn_arena, h->dt_counters, h->d_key_off, h->d_key_len, h->d_dirty, h->d_st, h->d_keys, h->d_acc, h->d_kflags, h->d_kmax, h->d_in_ts, h->d_in_key, h->d_in_val,
                    h->f_key, h->f_sum, h->f_min, h->f_max, h->f_wend, h->f_count, h->wa_key, h->wa_wend, h->wa_val,
                    h->ea_ts, h->ea_key, h->ea_val, h->ca_ts, h->ca_key, h->ca_val, h->d_rolling,
                    h->d_pack[0], h->d_pack[1], h->kr_lead, h->kr_k, h->kr_fl, h->kr_in_c, h->kr_mk, h->kr_mb, h->kr_in_mk, h->kr_in_mb, h->d_counts, h->d_tiles, h->d_xchg, h->d_xsend, h->d_text, h->d_line_start, h->d_txt_status, h->d_txt_bad,
                    h->ml_pane[0], h->ml_pane[1], h->ml_slot[0], h->ml_slot[1], h->ml_val[0], h->ml_val[1], h->mw_items[0], h->mw_items[1], h->mw_vals, h->mw_big, h->d_loc[0], h->d_loc[1], h->d_loc[2],
                    h->md_key, h->md_wend, h->md_val, h->lt_ts, h->lt_key, h->lt_val};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (void* p : h->alt_out) if (p) cudaFree(p);
    for (void* p : h->xs_opened) if (p) cudaIpcCloseMemHandle(p);
    { void* xs[] = {h->d_xbuf, h->d_xcounts, h->d_xtiles, h->d_send_max, h->d_seg_n, h->xs_stage, h->d_xfill}; for (void* p : xs) if (p) cudaFree(p); }
    if (h->h_ntotal) cudaFreeHost(h->h_ntotal);
    if (h->send_stream) { cudaStreamSynchronize(h->send_stream); cudaStreamDestroy(h->send_stream); }
    for (int i = 0; i < fmq_handle::HOST_SETS; ++i) {
        void* q[] = {h->hs_ts[i], h->hs_key[i], h->hs_val[i], h->hs_narrow[i]};
        for (void* p : q) if (p) cudaFree(p);
        if (h->ev_free[i]) cudaEventDestroy(h->ev_free[i]);
    }
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    for (auto& q : h->ev_piece) for (cudaEvent_t e : q) if (e) cudaEventDestroy(e);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->out_stream) cudaStreamDestroy(h->out_stream);
    if (h->fire_stream) cudaStreamDestroy(h->fire_stream);
    if (h->ev_plan) cudaEventDestroy(h->ev_plan);
    if (h->ev_bdone) cudaEventDestroy(h->ev_bdone);
    if (h->ev_out) cudaEventDestroy(h->ev_out);
    if (h->h_st) cudaFreeHost(h->h_st);
    if (h->h_hint) cudaFreeHost(h->h_hint);
    for (auto& q : h->tev) for (cudaEvent_t e : q) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

static size_t locality_min_mb();
int32_t fmq_create(const fmq_config* cfg, fmq_handle** out) {
    if (!cfg || !out) return set_err(nullptr, FMQ_ERR_BAD_ARG, "null argument");
    *out = nullptr;
    if (cfg->struct_size != sizeof(fmq_config)) return set_err(nullptr, FMQ_ERR_BAD_ARG, "fmq_config.struct_size mismatch");
    if (cfg->value_type != FMQ_VAL_F64 && cfg->value_type != FMQ_VAL_I64) return set_err(nullptr, FMQ_ERR_BAD_ARG, "bad value_type");
    if (cfg->window_enabled) {
        if (cfg->window_size_ms <= 0 || cfg->window_slide_ms <= 0) return set_err(nullptr, FMQ_ERR_BAD_ARG, "window size/slide must be positive");
        if (cfg->window_size_ms % cfg->window_slide_ms != 0) return set_err(nullptr, FMQ_ERR_BAD_ARG, "window_size_ms must be a multiple of window_slide_ms (pane pre-aggregation)");
        if (cfg->window_offset_ms < 0 || cfg->window_offset_ms >= cfg->window_slide_ms) return set_err(nullptr, FMQ_ERR_BAD_ARG, "need 0 <= window_offset_ms < window_slide_ms");
        if (cfg->max_out_of_orderness_ms < 0) return set_err(nullptr, FMQ_ERR_BAD_ARG, "max_out_of_orderness_ms cannot be negative");   // chapter3/README.md:356-359
    }
    if (cfg->consec_n < 0) return set_err(nullptr, FMQ_ERR_BAD_ARG, "consec_n cannot be negative");
    if (cfg->consec_n > 0x3FFFFFFF) return set_err(nullptr, FMQ_ERR_BAD_ARG, "consec_n too large");
    if (cfg->key_mode != FMQ_KEYS_HASHED && cfg->key_mode != FMQ_KEYS_DENSE) return set_err(nullptr, FMQ_ERR_BAD_ARG, "bad key_mode");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return set_err(nullptr, FMQ_ERR_NO_DEVICE, "no CUDA device visible (this library has no CPU fallback)"); }
    if (cfg->device < 0 || cfg->device >= ndev) return set_err(nullptr, FMQ_ERR_NO_DEVICE, "device ordinal out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess) return set_err(nullptr, FMQ_ERR_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major < 10) return set_err(nullptr, FMQ_ERR_NO_DEVICE, "kernels are built for sm_100a only");

    fmq_handle* h = new (std::nothrow) fmq_handle();
    if (!h) return set_err(nullptr, FMQ_ERR_BAD_ARG, "out of host memory");
    h->cfg = *cfg; h->device = cfg->device; h->num_sms = prop.multiProcessorCount;
    auto fail = [&](int st, const std::string& m) { g_create_err = m; free_all(h); return st; };
#define CUC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(FMQ_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
    CUC(cudaSetDevice(h->device));
    CUC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    for (auto& q : h->tev) for (cudaEvent_t& e : q) CUC(cudaEventCreate(&e));

    h->capacity = cfg->table_capacity ? cfg->table_capacity : (1u << 20);
    if (h->capacity < 32) h->capacity = 32;
    h->capacity = (h->capacity + 3) & ~3ull;            // whole 4-key buckets
    if (h->capacity >= 0xFFFFFFF0ull) return fail(FMQ_ERR_BAD_ARG, "table_capacity must be < 2^32");
    u32 ring = cfg->pane_ring ? cfg->pane_ring : 8; u32 r2 = 1; while (r2 < ring) r2 <<= 1; h->ring = r2;
    h->slot_bits = 1; while ((1ull << h->slot_bits) < h->capacity) h->slot_bits++;
    h->max_batch = cfg->max_batch ? (size_t)cfg->max_batch : ((size_t)1 << 22);
    if (h->max_batch >= 0xFFFFFFF0ull) return fail(FMQ_ERR_BAD_ARG, "max_batch must be < 2^32");
    if (cfg->consec_n > 0 && h->max_batch > 0x80000000ull) return fail(FMQ_ERR_BAD_ARG, "max_batch must be <= 2^31 with the consecutive-readings rule");
    h->g.size = cfg->window_size_ms; h->g.slide = cfg->window_slide_ms; h->g.off = cfg->window_offset_ms;
    h->g.ring_mask = h->ring - 1; h->g.agg_mask = cfg->agg_mask | FMQ_AGG_COUNT; h->g.capacity = h->capacity;
    h->g.slide32 = (cfg->window_slide_ms > 0 && cfg->window_slide_ms < 0x100000000ll) ? (u32)cfg->window_slide_ms : 0;
    h->g.panes_per_window = cfg->window_enabled ? (u32)(cfg->window_size_ms / cfg->window_slide_ms) : 0;
    if (cfg->window_enabled && h->g.panes_per_window >= h->ring) return fail(FMQ_ERR_BAD_ARG, "pane_ring must exceed window_size_ms / window_slide_ms");
    if (cfg->allowed_lateness_ms < 0 || cfg->allowed_lateness_ms > (1ll << 40)) return fail(FMQ_ERR_BAD_ARG, "allowed_lateness_ms must be in [0, 2^40]");
    if (cfg->allowed_lateness_ms > 0) {
        if (!cfg->window_enabled) return fail(FMQ_ERR_BAD_ARG, "allowed_lateness_ms needs window_enabled");
        if (cfg->median_log_capacity > 0) return fail(FMQ_ERR_BAD_ARG, "allowed_lateness_ms with the window median is not supported");
        if ((uint64_t)(cfg->window_size_ms + cfg->allowed_lateness_ms) / (uint64_t)cfg->window_slide_ms + 1 >= h->ring)
            return fail(FMQ_ERR_BAD_ARG, "pane_ring must exceed (window_size_ms + allowed_lateness_ms) / window_slide_ms + 1");
        h->g.lateness = cfg->allowed_lateness_ms;
        h->g.dirty_words = (h->capacity + 31) / 32;
    }

    CUC(dalloc(&h->d_st, 1)); CUC(cudaMallocHost((void**)&h->h_st, sizeof(DevState)));
    CUC(cudaMallocHost((void**)&h->h_hint, 64)); memset(h->h_hint, 0, 64);
    DevState init; memset(&init, 0, sizeof(init));
    init.wm = I64_MIN_; init.cur_max_ts = I64_MIN_ + cfg->max_out_of_orderness_ms; init.last_emitted_wm = I64_MIN_; init.batch_max_ts = I64_MIN_;
    init.min_pane = I64_MAX_; init.max_pane = I64_MIN_; init.expire_lo = 1; init.expire_hi = 0;
    CUC(cudaMemcpy(h->d_st, &init, sizeof(init), cudaMemcpyHostToDevice));
    CUC(dalloc(&h->d_keys, h->capacity)); CUC(cudaMemset(h->d_keys, 0xFF, h->capacity * 8));
    if (cfg->window_enabled) { CUC(dalloc(&h->d_acc, (size_t)h->ring * h->capacity)); CUC(cudaMemset(h->d_acc, 0, (size_t)h->ring * h->capacity * sizeof(Acc32))); }
    if (h->g.lateness > 0) {
        CUC(dalloc(&h->d_dirty, (size_t)h->ring * h->g.dirty_words)); CUC(cudaMemset(h->d_dirty, 0, (size_t)h->ring * h->g.dirty_words * 4));
        h->g.dirty = h->d_dirty;
    }
    CUC(dalloc(&h->d_in_ts, h->max_batch)); CUC(dalloc(&h->d_in_key, h->max_batch)); CUC(dalloc(&h->d_in_val, h->max_batch));
    if (cfg->window_enabled && cfg->emit_fired) {
        h->f_cap = cfg->fired_capacity ? cfg->fired_capacity : (2 * h->capacity > (1u << 20) ? 2 * h->capacity : (1u << 20));
        CUC(dalloc(&h->f_key, h->f_cap)); CUC(dalloc(&h->f_wend, h->f_cap)); CUC(dalloc(&h->f_count, h->f_cap));
        CUC(dalloc(&h->f_sum, h->f_cap)); CUC(dalloc(&h->f_min, h->f_cap)); CUC(dalloc(&h->f_max, h->f_cap));
    }
    if (cfg->window_enabled && cfg->walert_source != FMQ_ALERT_NONE) {
        h->wa_cap = cfg->walert_capacity ? cfg->walert_capacity : (2 * h->capacity > (1u << 20) ? 2 * h->capacity : (1u << 20));
        CUC(dalloc(&h->wa_key, h->wa_cap)); CUC(dalloc(&h->wa_wend, h->wa_cap)); CUC(dalloc(&h->wa_val, h->wa_cap));
    }
    if (cfg->efilter_enabled) {
        h->ea_cap = cfg->ealert_capacity ? cfg->ealert_capacity : h->max_batch;
        CUC(dalloc(&h->ea_ts, h->ea_cap)); CUC(dalloc(&h->ea_key, h->ea_cap)); CUC(dalloc(&h->ea_val, h->ea_cap));
    }
    if (keyed_enabled(*cfg)) {
        CUC(dalloc(&h->d_kflags, h->capacity)); CUC(cudaMemset(h->d_kflags, 0, h->capacity * 4));
        CUC(dalloc(&h->d_kmax, h->capacity)); CUC(cudaMemset(h->d_kmax, 0, h->capacity * 8));
        for (int k = 0; k < 2; ++k) CUC(dalloc(&h->d_pack[k], h->max_batch));
        const size_t nr = (h->max_batch + KA_RANGE - 1) / KA_RANGE + 1;
        CUC(dalloc(&h->kr_lead, nr)); CUC(dalloc(&h->kr_k, nr)); CUC(dalloc(&h->kr_fl, nr)); CUC(dalloc(&h->kr_in_c, nr));
        CUC(dalloc(&h->kr_mk, nr)); CUC(dalloc(&h->kr_mb, nr)); CUC(dalloc(&h->kr_in_mk, nr)); CUC(dalloc(&h->kr_in_mb, nr));
        if (cfg->rolling_max_enabled) CUC(dalloc(&h->d_rolling, h->max_batch));
        if (cfg->consec_n > 0) {
            h->ca_cap = cfg->calert_capacity ? cfg->calert_capacity : h->max_batch;
            CUC(dalloc(&h->ca_ts, h->ca_cap)); CUC(dalloc(&h->ca_key, h->ca_cap)); CUC(dalloc(&h->ca_val, h->ca_cap));
        }
    }
    // radix counter matrix (keyed path and owner partition)
    h->counts_len = (size_t)RS_DIGITS * xs_num_chunks(h->max_batch);
    if (keyed_enabled(*cfg)) { const size_t k = (size_t)2048 * ps_num_chunks(h->max_batch); if (k > h->counts_len) h->counts_len = k; }
    if (cfg->median_log_capacity > 0) { const size_t k = (size_t)2048 * ps_num_chunks(cfg->median_log_capacity); if (k > h->counts_len) h->counts_len = k; }
    h->tiles_len = (h->counts_len + SCAN_TILE - 1) / SCAN_TILE;
    CUC(dalloc(&h->d_counts, h->counts_len)); CUC(dalloc(&h->d_tiles, h->tiles_len));
    CUC(dalloc(&h->d_xchg, 4)); CUC(cudaMemset(h->d_xchg, 0, 32));
    // Hashed keys over a table whose panes are much larger than the L2 (126 MB): every accumulator request would go to DRAM
    // at random.  Such handles split each batch into ranges of the hash first (= ranges of directory buckets and of
    // accumulator slots: ~24 MB of cells per range and pane), so that the ingest kernel sweeps the table range by range.
    // Measured (profiles/r02_notes.md): configs[3], 480 MB per pane: 1.57 -> 1.35 ms per 2^24 events; a 128 MB pane under
    // Zipf keys (configs[4]) loses more to the split than it gains (1.80 -> 1.96), hence the 256 MB default threshold.
    if (cfg->window_enabled && cfg->key_mode == FMQ_KEYS_HASHED) {
        const size_t pane_mb = (size_t)h->capacity * sizeof(Acc32) >> 20;
        if (pane_mb >= locality_min_mb()) {
            u32 p = 2; while (p < 32 && (size_t)p * 24 < pane_mb) p <<= 1;
            h->loc_parts = p;
            for (int k = 0; k < 3; ++k) CUC(dalloc(&h->d_loc[k], h->max_batch));
        }
    }
    if (cfg->overlapped_polls) {
        CUC(cudaStreamCreateWithFlags(&h->out_stream, cudaStreamNonBlocking)); CUC(cudaEventCreateWithFlags(&h->ev_out, cudaEventDisableTiming));
        const size_t caps[12] = {h->f_cap, h->f_cap, h->f_cap, h->f_cap, h->f_cap, h->f_cap, h->wa_cap, h->wa_cap, h->wa_cap, h->ea_cap, h->ea_cap, h->ea_cap};
        for (int i = 0; i < 12; ++i) if (caps[i]) CUC(cudaMalloc(&h->alt_out[i], caps[i] * 8));
    }
    if (cfg->late_capacity > 0 && cfg->window_enabled) {
        h->lt_cap = cfg->late_capacity;
        CUC(dalloc(&h->lt_ts, h->lt_cap)); CUC(dalloc(&h->lt_key, h->lt_cap)); CUC(dalloc(&h->lt_val, h->lt_cap));
    }
    if (cfg->median_log_capacity > 0) {
        if (!cfg->window_enabled || cfg->value_type != FMQ_VAL_F64) return fail(FMQ_ERR_BAD_ARG, "the window median needs window_enabled and f64 values");
        if (cfg->median_log_capacity >= 0xFFFFFFF0ull) return fail(FMQ_ERR_BAD_ARG, "median_log_capacity must be < 2^32");
        h->ml_cap = cfg->median_log_capacity;
        for (int k = 0; k < 2; ++k) { CUC(dalloc(&h->ml_pane[k], h->ml_cap)); CUC(dalloc(&h->ml_slot[k], h->ml_cap)); CUC(dalloc(&h->ml_val[k], h->ml_cap)); CUC(dalloc(&h->mw_items[k], h->ml_cap)); }
        CUC(dalloc(&h->mw_vals, h->ml_cap));
        h->mw_big_cap = h->ml_cap / (MED_SMALL + 1) + 1;                    // runs longer than MED_SMALL values: (start, length) pairs behind a counter
        CUC(dalloc(&h->mw_big, 2 + 2 * h->mw_big_cap));
        h->md_cap = cfg->median_capacity ? cfg->median_capacity : (2 * h->capacity > (1u << 20) ? 2 * h->capacity : (1u << 20));
        CUC(dalloc(&h->md_key, h->md_cap)); CUC(dalloc(&h->md_wend, h->md_cap)); CUC(dalloc(&h->md_val, h->md_cap));
    }
    if (cfg->text_max_bytes > 0) {
        CUC(dalloc(&h->d_text, cfg->text_max_bytes + 64)); CUC(dalloc(&h->d_line_start, h->max_batch + 2));
        CUC(dalloc(&h->d_txt_status, h->max_batch)); CUC(dalloc(&h->d_txt_bad, 1));
        if (cfg->text_dict_capacity > 0) {
            if (cfg->text_dict_capacity >= 0xFFFFFFF0ull) return fail(FMQ_ERR_BAD_ARG, "text_dict_capacity must be < 2^32");
            if (cfg->key_mode == FMQ_KEYS_DENSE && cfg->text_dict_capacity > h->capacity) return fail(FMQ_ERR_BAD_ARG, "text_dict_capacity cannot exceed table_capacity of a FMQ_KEYS_DENSE handle (ids index the table)");
            u64 slots = 64; while (slots < 2 * cfg->text_dict_capacity) slots <<= 1;
            h->dt_mask = slots - 1;
            h->dn_arena_cap = cfg->text_dict_arena_bytes ? cfg->text_dict_arena_bytes : 32 * cfg->text_dict_capacity;
            CUC(dalloc(&h->dt_hash, slots)); CUC(cudaMemset(h->dt_hash, 0xFF, slots * 8));
            CUC(dalloc(&h->dt_id, slots)); CUC(cudaMemset(h->dt_id, 0xFF, slots * 4));
            CUC(dalloc(&h->dn_off, cfg->text_dict_capacity)); CUC(dalloc(&h->dn_len, cfg->text_dict_capacity));
            CUC(dalloc(&h->dn_arena, h->dn_arena_cap)); CUC(dalloc(&h->dt_counters, 4)); CUC(cudaMemset(h->dt_counters, 0, 32));
            CUC(dalloc(&h->d_key_off, h->max_batch)); CUC(dalloc(&h->d_key_len, h->max_batch));
        }
        const size_t nb = (cfg->text_max_bytes + TXT_CHUNK - 1) / TXT_CHUNK + 1;
        if (nb > h->counts_len) return fail(FMQ_ERR_BAD_ARG, "text_max_bytes too large for this max_batch (raise max_batch)");
    }
    h->peer_acc[0] = h->d_acc;
    CUC(cudaDeviceSynchronize());
#undef CUC
    *out = h;
    return FMQ_OK;
}

int32_t fmq_destroy(fmq_handle* h) {
    if (!h) return FMQ_ERR_BAD_ARG;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    free_all(h);
    return FMQ_OK;
}

}  // extern "C"

// ---- internal helpers ----------------------------------------------------------------------------
static int launch_check(fmq_handle* h, const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { h->err = std::string(what) + ": " + cudaGetErrorString(e); return FMQ_ERR_CUDA; }
    h->launches++;
    return FMQ_OK;
}
#define LAUNCHED(h, what) do { int s_ = launch_check(h, what); if (s_) return s_; } while (0)

static size_t locality_min_mb() {              // FMQ_LOCALITY_MIN_MB=0 switches the hash-range split off
    static const size_t v = [] { const char* e = getenv("FMQ_LOCALITY_MIN_MB"); const long m = e ? atol(e) : 256; return m <= 0 ? (size_t)1 << 40 : (size_t)m; }();
    return v;
}
// exclusive scan of `len` u32 counters in place (three small kernels); `tiles` holds one partial sum per 4096 counters
static cudaError_t scan_counts_on(cudaStream_t stream, u32* counts, u32* tiles, size_t len) {
    const u32 ntiles = (u32)((len + SCAN_TILE - 1) / SCAN_TILE);
    scan_tile_sums<<<ntiles, SCAN_THREADS, 0, stream>>>(counts, len, tiles);
    scan_tile_offsets<<<1, SCAN_THREADS, 0, stream>>>(tiles, ntiles);
    scan_apply<<<ntiles, SCAN_THREADS, 0, stream>>>(counts, len, tiles);
    return cudaGetLastError();
}
static int exclusive_scan_counts(fmq_handle* h, size_t len) {
    cudaError_t e = scan_counts_on(h->stream, h->d_counts, h->d_tiles, len);
    if (e != cudaSuccess) { h->err = std::string("scan kernels: ") + cudaGetErrorString(e); return FMQ_ERR_CUDA; }
    h->launches += 3;
    return FMQ_OK;
}

// Stable owner split of a device-resident batch (A4): count -> scan -> block-staged scatter, 5 launches on `stream`.
// 3-bit digits while world <= 8 (one NVSwitch node), 8-bit digits up to 256.  After it, counts[d * nchunks] is the
// start of owner d's bucket (single-output mode) or what precedes owner d's events in digit-major order.
struct SplitScratch { u32* counts; u32* tiles; };
static cudaError_t owner_split(cudaStream_t stream, const SplitScratch& sc, const u64* ts, const u64* key, const u64* val, size_t n, u32 world, u32 dense,
                               const SplitDst& dst, bool per_owner, i64* max_ts, unsigned long long* fill = nullptr, u32 max_blocks = 0) {
    const u32 nchunks = xs_num_chunks(n);
    SplitArgs a{ts, key, val, n, nchunks, world, dense, sc.counts, dst, max_ts, per_owner ? fill : nullptr};
    const u32 grid = (max_blocks && max_blocks < nchunks) ? max_blocks : nchunks;      // the scatter is persistent over chunks
    auto run = [&](auto hist, auto scat_single, auto scat_owner, u32 nb) -> cudaError_t {
        cudaError_t e;
        if (!a.fill) {                                                                // stable form: count, scan, then scatter
            hist<<<nchunks, XS_THREADS, 0, stream>>>(key, n, nchunks, world, dense, sc.counts);
            e = scan_counts_on(stream, sc.counts, sc.tiles, (size_t)nb * nchunks); if (e != cudaSuccess) return e;
        }
        // beside the ingest of the previous batch (keyBy exchange): 256-thread blocks; by itself: 512-thread blocks
        if (per_owner) {
            const size_t smem = xs_scatter_smem(nb, XS_THREADS); e = cudaFuncSetAttribute(scat_owner, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(scat_owner, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared); if (e != cudaSuccess) return e;   // it shares the SMs with the ingest kernel
            scat_owner<<<grid, XS_THREADS, smem, stream>>>(a);
        }
        else           { const size_t smem = xs_scatter_smem(nb, XS_THREADS_WIDE); e = cudaFuncSetAttribute(scat_single, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); if (e != cudaSuccess) return e; scat_single<<<grid, XS_THREADS_WIDE, smem, stream>>>(a); }
        return cudaGetLastError();
    };
    if (world <= 8) return run(xs_hist_kernel<3>, xs_scatter_kernel<3, false, XS_THREADS_WIDE>, xs_scatter_kernel<3, true, XS_THREADS>, 8u);
    if (world <= 32 && !per_owner) return run(xs_hist_kernel<5>, xs_scatter_kernel<5, false, XS_THREADS_WIDE>, xs_scatter_kernel<5, false, XS_THREADS_WIDE>, 32u);
    return run(xs_hist_kernel<8>, xs_scatter_kernel<8, false, XS_THREADS_WIDE>, xs_scatter_kernel<8, false, XS_THREADS_WIDE>, 256u);
}

static int status_from_flags(fmq_handle* h, u32 flags) {
    if (!flags) return FMQ_OK;
    int st; const char* m;
    if (flags & ERRB_TABLE_FULL) { st = FMQ_ERR_TABLE_FULL; m = "key directory full: raise fmq_config.table_capacity"; }
    else if (flags & ERRB_RING) { st = FMQ_ERR_RING_OVERFLOW; m = "open event-time span exceeds pane_ring panes: raise fmq_config.pane_ring or push smaller batches"; }
    else if (flags & ERRB_PEER) { st = FMQ_ERR_PEER_TIMEOUT; m = "a peer rank did not publish or acknowledge a batch in time (keyBy exchange / combiner flags)"; }
    else if (flags & ERRB_OUTPUT) { st = FMQ_ERR_OUTPUT_OVERFLOW; m = "an output buffer overflowed before it was polled: poll more often or raise its capacity"; }
    else { st = FMQ_ERR_BAD_TS; m = "negative timestamp or reserved key (0xFFFFFFFFFFFFFFFF) in input"; }
    h->err = m; h->sticky = st;
    return st;
}

// synchronise the stream and refresh the pinned mirror of DevState
static int sync_state(fmq_handle* h) {
    CU(h, cudaSetDevice(h->device));
    if (h->fire_stream) CU(h, cudaStreamSynchronize(h->fire_stream));
    CU(h, cudaMemcpyAsync(h->h_st, h->d_st, sizeof(DevState), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->h_st->anchored) h->anchored_known = true;
    return status_from_flags(h, h->h_st->err_flags);
}

// Stable LSD radix sort of 8-byte items between buf[0] and buf[1] by the low `key_bits` of keyfn(item):
// as few passes as 11-bit digits allow, bits spread evenly over them.  *cur = index of the sorted buffer.
template <typename KeyFn>
static int sort_items(fmq_handle* h, u64* const* buf, size_t n, u32 key_bits, KeyFn keyfn, int* cur, bool first_counts_done = false) {
    const u32 passes = (key_bits + 10) / 11;
    const u32 bits = ps_pass_bits(key_bits);
    const u32 nchunks = ps_num_chunks(n), nblocks = (nchunks + 7) / 8;
    if ((size_t)(1u << bits) * nchunks > h->counts_len) return set_err(h, FMQ_ERR_BAD_ARG, "internal: radix counter matrix too small");
    for (u32 pass = 0; pass < passes; ++pass) {
        const u32 shift = pass * bits;
        const size_t smem = (size_t)8 * (1u << bits) * 4;
        auto launch = [&](auto hist, auto scat) -> int {
            CU(h, cudaFuncSetAttribute(hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            if (!(first_counts_done && pass == 0)) {                  // the producer of the items has already counted their first digit
                hist<<<nblocks, 256, smem, h->stream>>>(buf[*cur], nullptr, n, nchunks, shift, h->d_counts, keyfn); LAUNCHED(h, "ps_pass_kernel(hist)");
            }
            int s_ = exclusive_scan_counts(h, (size_t)(1u << bits) * nchunks); if (s_) return s_;
            const size_t tsmem = ps_scatter_smem(1u << bits);
            CU(h, cudaFuncSetAttribute(scat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem));
            scat<<<nchunks, 256, tsmem, h->stream>>>(buf[*cur], buf[*cur ^ 1], n, nchunks, shift, h->d_counts, keyfn); LAUNCHED(h, "ps_scatter_tile_kernel");
            return FMQ_OK;
        };
        int s_ = FMQ_OK;
        switch (bits) {
            case 6: s_ = launch(ps_pass_kernel<6, false, KeyFn>, ps_scatter_tile_kernel<6, KeyFn>); break;
            case 7: s_ = launch(ps_pass_kernel<7, false, KeyFn>, ps_scatter_tile_kernel<7, KeyFn>); break;
            case 8: s_ = launch(ps_pass_kernel<8, false, KeyFn>, ps_scatter_tile_kernel<8, KeyFn>); break;
            case 9: s_ = launch(ps_pass_kernel<9, false, KeyFn>, ps_scatter_tile_kernel<9, KeyFn>); break;
            case 10: s_ = launch(ps_pass_kernel<10, false, KeyFn>, ps_scatter_tile_kernel<10, KeyFn>); break;
            default: s_ = launch(ps_pass_kernel<11, false, KeyFn>, ps_scatter_tile_kernel<11, KeyFn>); break;
        }
        if (s_) return s_;
        *cur ^= 1;
    }
    return FMQ_OK;
}

static size_t keyed_sub_batch() {
    static const size_t v = [] { const char* e = getenv("FMQ_KEYED_SUB_LOG2"); const int l = e ? atoi(e) : 23; return (size_t)1 << (l < 12 ? 12 : (l > 31 ? 31 : l)); }();
    return v;
}

// keyed ValueState operators over one (sub-)batch: resolve -> stable sort by slot -> scan / carry / lead
template <typename V>
static int stage_keyed(fmq_handle* h, const i64* ts, const u64* key, const V* val, size_t n, size_t out_off) {
    const fmq_config& c = h->cfg;
    {
        // consecutive rule alone: the over-threshold bit travels inside the sorted items, no value gathers
        const bool pack_over = c.consec_n > 0 && !c.rolling_max_enabled;
        const u32 nchunks = ps_num_chunks(n), bits = ps_pass_bits(h->slot_bits);
        const size_t smem = (size_t)8 * (1u << bits) * 4;
        auto resolve = [&](auto kern) -> int {
            CU(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            kern<<<(nchunks + 7) / 8, 256, smem, h->stream>>>(key, val, n, nchunks, h->d_keys, h->capacity, c.key_mode == FMQ_KEYS_DENSE,
                                                             pack_over, c.consec_op, c.consec_threshold, h->d_st, h->d_pack[0], h->d_counts);
            LAUNCHED(h, "resolve_hist_kernel");
            return FMQ_OK;
        };
        int rs = FMQ_OK;
        switch (bits) {
            case 6: rs = resolve(resolve_hist_kernel<V, 6>); break;
            case 7: rs = resolve(resolve_hist_kernel<V, 7>); break;
            case 8: rs = resolve(resolve_hist_kernel<V, 8>); break;
            case 9: rs = resolve(resolve_hist_kernel<V, 9>); break;
            case 10: rs = resolve(resolve_hist_kernel<V, 10>); break;
            default: rs = resolve(resolve_hist_kernel<V, 11>); break;
        }
        if (rs) return rs;
        int cur = 0;
        { int s_ = sort_items(h, h->d_pack, n, h->slot_bits, SlotKey{}, &cur, true); if (s_) return s_; }
        KeyedArgs k{};
        k.sorted = h->d_pack[cur]; k.n = n; k.ts = ts; k.key = key; k.val = val; k.st = h->d_st;
        k.kflags = h->d_kflags; k.kmax = h->d_kmax; k.rolling_enabled = c.rolling_max_enabled; k.rolling_out = h->d_rolling ? h->d_rolling + out_off : nullptr;
        k.consec_n = c.consec_n; k.consec_op = c.consec_op; k.consec_reset = c.consec_reset_on_fire; k.consec_thr = c.consec_threshold;
        k.ca_ts = h->ca_ts; k.ca_key = h->ca_key; k.ca_val = h->ca_val; k.ca_cap = h->ca_cap;
        k.r_lead = h->kr_lead; k.r_fn_k = h->kr_k; k.r_fn_flags = h->kr_fl; k.r_fn_mkey = h->kr_mk; k.r_fn_mbits = h->kr_mb;
        k.r_in_c = h->kr_in_c; k.r_in_mkey = h->kr_in_mk; k.r_in_mbits = h->kr_in_mb;
        const size_t nranges = (n + KA_RANGE - 1) / KA_RANGE;
        const unsigned sgrid = (unsigned)((nranges + 7) / 8);
        if (pack_over) keyed_scan_kernel<V, false><<<sgrid, 256, 0, h->stream>>>(k); else keyed_scan_kernel<V, true><<<sgrid, 256, 0, h->stream>>>(k);
        LAUNCHED(h, "keyed_scan_kernel");
        keyed_carry_kernel<<<1, 1024, 0, h->stream>>>(k, (u32)nranges, (u32)c.consec_n, c.consec_reset_on_fire); LAUNCHED(h, "keyed_carry_kernel");
        if (pack_over) keyed_lead_kernel<V, false><<<sgrid, 256, 0, h->stream>>>(k); else keyed_lead_kernel<V, true><<<sgrid, 256, 0, h->stream>>>(k);
        LAUNCHED(h, "keyed_lead_kernel");
    }
    return FMQ_OK;
}

// stage 1 of a push: everything that touches only this handle's own events
static u64 peer_timeout_ns();
static u32 xs_ingest_blocks();
static u32 cmb_ingest_blocks();
struct SegIn { u32 nseg = 0; size_t stride = 0; const u64* seg_n = nullptr; };   // segmented input with device-side counts (keyBy exchange)
template <typename V>
static int stage_ingest(fmq_handle* h, const i64* ts, const u64* key, const void* val, size_t n, const SegIn& seg = SegIn()) {
    const fmq_config& c = h->cfg;
    cudaEvent_t* tev = h->tev[h->pushes % fmq_handle::TIMING_RING];
    if (n > 0 && (c.window_enabled || c.efilter_enabled)) {
        IngestArgs a{};
        a.nseg = seg.nseg; a.seg_stride = seg.stride; a.seg_n = seg.seg_n;
        a.ts = ts; a.key = key; a.val = val; a.n = n; a.st = h->d_st; a.dkeys = h->d_keys; a.acc = h->d_acc; a.g = h->g;
        if (h->loc_parts && seg.nseg == 0 && n >= ((size_t)1 << 20)) {      // (small batches touch little of the table anyway)
            SplitDst dst{}; dst.out_ts = h->d_loc[0]; dst.out_key = h->d_loc[1]; dst.out_val = h->d_loc[2];
            CU(h, owner_split(h->stream, SplitScratch{h->d_counts, h->d_tiles}, reinterpret_cast<const u64*>(ts), key, static_cast<const u64*>(val), n, h->loc_parts, SPLIT_HASH_RANGE, dst, false, nullptr));
            h->launches += 5;
            a.ts = reinterpret_cast<const i64*>(h->d_loc[0]); a.key = h->d_loc[1]; a.val = h->d_loc[2];
        }
        a.window_enabled = c.window_enabled; a.vec32 = ((((uintptr_t)ts | (uintptr_t)key | (uintptr_t)val) & 31) == 0);
        a.efilter_enabled = c.efilter_enabled; a.efilter_op = c.efilter_op; a.efilter_thr = c.efilter_threshold;
        a.ea_ts = h->ea_ts; a.ea_key = h->ea_key; a.ea_val = h->ea_val; a.ea_cap = h->ea_cap;
        a.lt_ts = h->lt_ts; a.lt_key = h->lt_key; a.lt_val = h->lt_val; a.lt_cap = h->lt_cap;
        a.mlog_enabled = h->ml_cap > 0; a.ml_pane = h->ml_pane[h->ml_cur]; a.ml_slot = h->ml_slot[h->ml_cur]; a.ml_val = h->ml_val[h->ml_cur]; a.ml_cap = h->ml_cap;
        // The kernels pick themselves on the device-resident stream position (no host round trip):
        // the fast kernel returns at once unless a watermark anchors the pane origin (filter-only jobs
        // need none), the generic one unless it does not.  Launches known to be no-ops are skipped.
        const bool fast_geom = !c.window_enabled || (h->g.slide32 != 0 && (u64)h->ring * h->g.slide32 < 0x100000000ull &&
                                                     (u64)h->ring * h->capacity < 0x80000000ull);
        const bool dense = c.key_mode == FMQ_KEYS_DENSE;
        const bool extras = h->ml_cap > 0 || h->lt_cap > 0;
        if (fast_geom) {
            // persistent grid: every warp streams its own ring of 128-event tiles
            size_t blocks = (n * (seg.nseg ? seg.nseg : 1) + (size_t)IW_TILE * IW_WARPS - 1) / ((size_t)IW_TILE * IW_WARPS);
            // a rank of the keyBy exchange leaves room on every SM for the scatter of the next batch (one block of ~53 KB)
            // (overlapped combiner: the fire kernel of the previous batch runs beside this ingest; FMQ_CMB_INGEST_BLOCKS leaves it room)
            const size_t maxb = (size_t)h->num_sms * ((h->xs_attached && !h->cmb_flags) ? xs_ingest_blocks() : (h->overlap ? cmb_ingest_blocks() : FMQ_IW_BLOCKS));
            if (blocks > maxb) blocks = maxb;
            const bool comb = c.window_enabled && h->h_hint && *(volatile u32*)h->h_hint != 0;     // skew seen by an earlier batch's sample
            auto pick = [&](auto lean, auto combining) { return comb ? combining : lean; };
            auto kfn = dense ? (extras ? pick(ingest_ring_kernel<V, true, true, false>, ingest_ring_kernel<V, true, true, true>)
                                       : pick(ingest_ring_kernel<V, true, false, false>, ingest_ring_kernel<V, true, false, true>))
                             : (extras ? pick(ingest_ring_kernel<V, false, true, false>, ingest_ring_kernel<V, false, true, true>)
                                       : pick(ingest_ring_kernel<V, false, false, false>, ingest_ring_kernel<V, false, false, true>));
            const int smem = iw_smem_bytes(extras);
            CU(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            if (h->xs_attached) CU(h, cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));   // beside the scatter of the next batch
            kfn<<<(unsigned)blocks, IW_THREADS, smem, h->stream>>>(a);
            LAUNCHED(h, "ingest_ring_kernel");
        }
        const size_t tile = (size_t)INGEST_THREADS * INGEST_EPT;
        size_t blocks = (n * (seg.nseg ? seg.nseg : 1) + tile - 1) / tile; const size_t maxb = (size_t)h->num_sms * FMQ_INGEST_MIN_BLOCKS;   // persistent: resident blocks only
        if (blocks > maxb) blocks = maxb;
        if (c.window_enabled && (!fast_geom || !h->anchored_known)) {
            IngestArgs g = a;
            if (dense) ingest_generic_kernel<V, true><<<(unsigned)blocks, INGEST_THREADS, 0, h->stream>>>(g);
            else       ingest_generic_kernel<V, false><<<(unsigned)blocks, INGEST_THREADS, 0, h->stream>>>(g);
            LAUNCHED(h, "ingest_generic_kernel");
        }
    }
    CU(h, cudaEventRecord(tev[1], h->stream));
    if (n > 0 && keyed_enabled(c) && seg.nseg > 1) return set_err(h, FMQ_ERR_BAD_ARG, "internal: the keyed chain needs one contiguous batch");
    if (n > 0 && keyed_enabled(c)) {
        // Rolling max gathers every value by arrival index and scatters every output back to it: random
        // 8-byte accesses.  Sub-batches whose value / output columns are L2-sized (2^23 records = 64 MB each)
        // keep them mostly there (2^23 measured best: 1.39 -> 1.18 ms; 2^21 and below lose to per-launch
        // costs); per-key state carries over exactly as it does from one push to the next.
        const size_t sub = c.rolling_max_enabled ? keyed_sub_batch() : n;
        for (size_t off = 0; off < n; off += sub) {
            const size_t m = n - off < sub ? n - off : sub;
            int s_ = stage_keyed<V>(h, ts + off, key + off, static_cast<const V*>(val) + off, m, off); if (s_) return s_;
        }
        h->rolling_n = c.rolling_max_enabled ? n : 0;
    } else if (keyed_enabled(c)) {
        h->rolling_n = 0;
    }
    CU(h, cudaEventRecord(tev[2], h->stream));
    return FMQ_OK;
}

__global__ void set_mlog_kernel(DevState* st) { st->n_mlog = st->n_mwork; }

// N3: the records of every window the plan says is firing are selected from the value log — all windows of the
// fire in one pass (in groups only when their (record, window) pairs would not fit the work buffer) — grouped by
// (window, key) with two or three radix passes; each run's middle value(s) are then found where the run lies
// (csrc/fmq_median.cuh: one thread for a short run, a block's radix select for a long one).
// One read-back of the plan per push sizes the launches; nothing is synchronised per window.
static int sync_state(fmq_handle* h);
static int fire_medians(fmq_handle* h) {
    int s = sync_state(h); if (s) return s;
    const DevState d = *h->h_st;
    const i64 size = h->g.size, slide = h->g.slide, off = h->g.off, k = size / slide;
    const size_t off_mwork = offsetof(DevState, n_mwork);
    const unsigned grid = (unsigned)h->num_sms * 4;
    if (d.fire_count > 0 && d.n_mlog > 0) {
        // windows per group: every record lies in at most min(G, k) of G consecutive windows
        i64 G = d.fire_count;
        const u64 per = (u64)(G < k ? G : k);
        if (d.n_mlog * per > h->ml_cap) { G = (i64)(h->ml_cap / d.n_mlog); if (G > k) G = k; if (G < 1) G = 1; }
        u32 jbits = 1; while ((1ll << jbits) <= G) ++jbits;                  // 2^jbits > G: a real group key is never all ones
        if (jbits + h->slot_bits > 32) return set_err(h, FMQ_ERR_BAD_ARG, "window median: too many windows fire at once for this table size (advance the watermark in smaller steps)");
        for (i64 j0 = 0; j0 < d.fire_count; j0 += G) {
            const i64 nj = d.fire_count - j0 < G ? d.fire_count - j0 : G;
            u64 bound = d.n_mlog * (u64)(nj < k ? nj : k); if (bound > h->ml_cap) bound = h->ml_cap;
            CU(h, cudaMemsetAsync((char*)h->d_st + off_mwork, 0, 8, h->stream));
            CU(h, cudaMemsetAsync(h->mw_items[0], 0xFF, bound * 8, h->stream));
            median_select_all_kernel<<<grid, MED_THREADS, 0, h->stream>>>(h->ml_pane[h->ml_cur], h->ml_slot[h->ml_cur], h->ml_val[h->ml_cur], h->d_st, size, slide, off, j0, nj,
                                                                         h->slot_bits, h->mw_items[0], h->mw_vals, bound); LAUNCHED(h, "median_select_all_kernel");
            int cur = 0;
            s = sort_items(h, h->mw_items, (size_t)bound, h->slot_bits + jbits, SlotKey{}, &cur); if (s) return s;          // groups by (window, key)
            CU(h, cudaMemsetAsync(h->mw_big, 0, 8, h->stream));
            MedianPickArgs m{h->mw_items[cur], h->mw_items[cur ^ 1], h->mw_vals, bound, h->d_keys, h->cfg.key_mode == FMQ_KEYS_DENSE, d.fire_first_end + j0 * slide, h->d_st,
                             h->md_key, h->md_wend, h->md_val, h->md_cap, h->slot_bits, slide, h->mw_big, h->mw_big_cap};
            median_pick_kernel<<<(unsigned)h->num_sms * 8, MED_THREADS, 0, h->stream>>>(m); LAUNCHED(h, "median_pick_kernel");   // runs of up to MED_SMALL values
            median_long_kernel<<<grid, MED_THREADS, 0, h->stream>>>(m); LAUNCHED(h, "median_long_kernel");      // the others (returns at once if there are none)
        }
    }
    if (d.expire_lo <= d.expire_hi && d.n_mlog > 0) {
        CU(h, cudaMemsetAsync((char*)h->d_st + off_mwork, 0, 8, h->stream));
        const int c = h->ml_cur;
        median_compact_kernel<<<grid, MED_THREADS, 0, h->stream>>>(h->ml_pane[c], h->ml_slot[c], h->ml_val[c], d.n_mlog, h->d_st, d.expire_hi,
                                                                  h->ml_pane[c ^ 1], h->ml_slot[c ^ 1], h->ml_val[c ^ 1], h->ml_cap); LAUNCHED(h, "median_compact_kernel");
        set_mlog_kernel<<<1, 1, 0, h->stream>>>(h->d_st); LAUNCHED(h, "set_mlog_kernel");
        h->ml_cur = c ^ 1;
    }
    return FMQ_OK;
}

// stage 2 of a push: watermark, fire, cleanup.  In combiner mode (`retire_now` false) the cleanup is
// a separate step (fmq_retire) because peers may still be reading this rank's table.
template <typename V>
static int stage_fire(fmq_handle* h, size_t n, int64_t wm_after, bool retire_now, const SegIn& seg = SegIn(), size_t seg_cap = 0) {
    const fmq_config& c = h->cfg;
    PlanArgs p{};
    p.seg_n = seg.seg_n; p.nseg = seg.nseg; p.seg_cap = seg_cap;
    p.st = h->d_st; p.g = h->g; p.bound = c.max_out_of_orderness_ms; p.watermark_after = wm_after; p.n_events = (u64)n; p.window_enabled = c.window_enabled;
    p.is_auto = wm_after == FMQ_WM_AUTO; p.is_keep = wm_after == FMQ_WM_KEEP; p.overlap = h->overlap ? 1 : 0;
    p.combine_hint = h->h_hint;
    auto fire_args = [&](bool refire) {
        FireArgs f{};
        f.st = h->d_st; f.dkeys = h->d_keys; f.acc = h->d_acc; f.g = h->g; f.dense_keys = c.key_mode == FMQ_KEYS_DENSE; f.emit_fired = c.emit_fired;
        f.walert_source = c.walert_source; f.walert_op = c.walert_op;
        f.mul = c.walert_mul == 0.0 ? 1.0 : c.walert_mul; f.d0 = c.walert_div0 == 0.0 ? 1.0 : c.walert_div0;
        f.d1 = c.walert_div1 == 0.0 ? 1.0 : c.walert_div1; f.d2 = c.walert_div2 == 0.0 ? 1.0 : c.walert_div2; f.thr = c.walert_threshold;
        f.f_key = h->f_key; f.f_wend = h->f_wend; f.f_count = h->f_count; f.f_sum = h->f_sum; f.f_min = h->f_min; f.f_max = h->f_max; f.f_cap = h->f_cap;
        f.wa_key = h->wa_key; f.wa_wend = h->wa_wend; f.wa_val = h->wa_val; f.wa_cap = h->wa_cap;
        f.world = h->world; f.clear_in_fire = (retire_now && !refire) ? 1 : 0; f.refire = refire ? 1 : 0;
        const u64 per = (h->capacity + h->world - 1) / h->world;            // this rank fires a contiguous slot range
        f.slot_lo = per * h->rank < h->capacity ? per * h->rank : h->capacity;
        f.slot_hi = f.slot_lo + per < h->capacity ? f.slot_lo + per : h->capacity;
        for (int r = 0; r < h->world; ++r) f.peer_acc[r] = h->peer_acc[r];
        return f;
    };
    auto fire_blocks = [&](const FireArgs& f) {
        size_t blocks = (f.slot_hi - f.slot_lo + FIRE_THREADS * FIRE_SPT - 1) / (FIRE_THREADS * FIRE_SPT); const size_t maxb = (size_t)h->num_sms * FMQ_FIRE_GRID;
        if (blocks > maxb) blocks = maxb;
        return (unsigned)(blocks == 0 ? 1 : blocks);
    };
    if (c.window_enabled && h->g.lateness > 0) {
        // allowedLateness: windows that had fired before this batch and took records in it fire again, under the watermark
        // still in force, before the plan moves it (the kernel returns at once if the ingest marked nothing)
        const FireArgs f = fire_args(true);
        fire_kernel<V, false><<<fire_blocks(f), FIRE_THREADS, 0, h->stream>>>(f); LAUNCHED(h, "fire_kernel(refire)");
    }
    plan_kernel<<<1, 32, 0, h->stream>>>(p); LAUNCHED(h, "plan_kernel");
    cudaStream_t fs = h->stream;
    if (h->overlap) {                                          // everything from the fire kernel on runs beside the next ingest
        CU(h, cudaEventRecord(h->ev_plan, h->stream));
        CU(h, cudaStreamWaitEvent(h->fire_stream, h->ev_plan, 0));
        fs = h->fire_stream;
    }
    if (c.window_enabled && wm_after != FMQ_WM_AUTO && wm_after != FMQ_WM_KEEP &&
        wm_after >= INT64_MIN + c.window_size_ms + c.window_offset_ms + c.window_slide_ms + 2) h->anchored_known = true;
    if (c.window_enabled && h->ml_cap > 0) { int s_ = fire_medians(h); if (s_) return s_; }
    if (c.window_enabled) {
        const FireArgs f = fire_args(false);
        const unsigned blocks = fire_blocks(f);
        if (h->world > 1) fire_kernel<V, true><<<(unsigned)blocks, FIRE_THREADS, 0, fs>>>(f);
        else              fire_kernel<V, false><<<(unsigned)blocks, FIRE_THREADS, 0, fs>>>(f);
        LAUNCHED(h, "fire_kernel");
        if (retire_now) { commit_kernel<<<1, 32, 0, h->stream>>>(h->d_st); LAUNCHED(h, "commit_kernel"); }
    }
    return FMQ_OK;
}

template <typename V>
static int run_chain(fmq_handle* h, const i64* ts, const u64* key, const void* val, size_t n, int64_t wm_after) {
    int s = stage_ingest<V>(h, ts, key, val, n); if (s) return s;
    return stage_fire<V>(h, n, wm_after, true);
}

// ---- the columns of one micro-batch as the caller hands them over -----------------------------------

// widens the 4-byte columns of fmq_push_packed; a null source leaves that column alone
__global__ void __launch_bounds__(256) widen_columns_kernel(const u32* __restrict__ ts32, i64 ts_base, const u32* __restrict__ key32, const int* __restrict__ val32,
                                                            i64* __restrict__ ts, u64* __restrict__ key, u64* __restrict__ val, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if (ts32) ts[i] = ts_base + (i64)ts32[i];
        if (key32) key[i] = (u64)key32[i];
        if (val32) val[i] = (u64)(i64)val32[i];
    }
}

static int ensure_host_staging(fmq_handle* h, bool narrow) {
    if (!h->copy_stream) {
        CU(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (cudaEvent_t& e : h->ev_free) CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& q : h->ev_piece) for (cudaEvent_t& e : q) CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    const size_t mb = h->max_batch ? h->max_batch : 1;
    for (int i = 0; i < fmq_handle::HOST_SETS; ++i) {
        if (!h->hs_ts[i]) { CU(h, dalloc(&h->hs_ts[i], mb)); CU(h, dalloc(&h->hs_key[i], mb)); CU(h, dalloc(&h->hs_val[i], mb)); }
        if (narrow && !h->hs_narrow[i]) CU(h, dalloc(&h->hs_narrow[i], 3 * mb));
    }
    return FMQ_OK;
}

// Host columns are copied in on the copy stream in up to HOST_CHUNKS pieces; the ingest of a piece
// runs while the next one is in flight, so that only the last piece's kernels and the fire chain are
// left when the copy ends.  Ingest does not depend on how a micro-batch is cut (the pane origin and
// the watermark move in the plan / commit kernels only).
constexpr int HOST_CHUNKS = 4;
constexpr size_t HOST_CHUNK_MIN = 1u << 18;

// queues the copy-in of one host batch on the copy stream, into the staging set whose turn it is
// (`ahead`: called from fmq_stage_host — the copy runs under the previous batch's kernels anyway, so it goes
// over as one piece: fewer, larger copies measured 1.6 % faster end to end)
static int queue_host_copy(fmq_handle* h, const InCols& c, size_t n, bool ahead) {
    const bool narrow = c.ts_w == 4 || c.key_w == 4 || c.val_w == 4;
    int s_ = ensure_host_staging(h, narrow); if (s_) return s_;
    fmq_handle::Staged& g = h->staged[h->n_staged];
    g.c = c; g.n = n; g.set = (int)(h->host_pushes++ % fmq_handle::HOST_SETS); g.pieces = 0;
    const size_t mb = h->max_batch;
    u32* nar = h->hs_narrow[g.set];
    size_t chunk = (n + HOST_CHUNKS - 1) / HOST_CHUNKS;
    if (chunk < HOST_CHUNK_MIN) chunk = HOST_CHUNK_MIN;
    if (h->cfg.rolling_max_enabled || ahead) chunk = n;                 // rolling max: its output column is indexed by position in the push
    g.chunk = (chunk + 3) & ~(size_t)3;                                 // keeps every piece 32-byte aligned
    // the kernels of the push that used this set last: wait for them, on the copy stream only
    CU(h, cudaStreamWaitEvent(h->copy_stream, h->ev_free[g.set], 0));
    for (size_t off = 0; off < n; off += g.chunk, ++g.pieces) {
        const size_t m = n - off < g.chunk ? n - off : g.chunk;
        CU(h, cudaMemcpyAsync(c.ts_w == 8 ? (void*)(h->hs_ts[g.set] + off) : (void*)(nar + off), (const char*)c.ts + off * c.ts_w, m * c.ts_w, cudaMemcpyHostToDevice, h->copy_stream));
        CU(h, cudaMemcpyAsync(c.key_w == 8 ? (void*)(h->hs_key[g.set] + off) : (void*)(nar + mb + off), (const char*)c.key + off * c.key_w, m * c.key_w, cudaMemcpyHostToDevice, h->copy_stream));
        CU(h, cudaMemcpyAsync(c.val_w == 8 ? (void*)(h->hs_val[g.set] + off) : (void*)(nar + 2 * mb + off), (const char*)c.val + off * c.val_w, m * c.val_w, cudaMemcpyHostToDevice, h->copy_stream));
        CU(h, cudaEventRecord(h->ev_piece[g.set][g.pieces], h->copy_stream));
    }
    h->n_staged++;
    return FMQ_OK;
}

static bool same_cols(const InCols& a, const InCols& b) {
    return a.ts == b.ts && a.key == b.key && a.val == b.val && a.ts_w == b.ts_w && a.key_w == b.key_w && a.val_w == b.val_w && a.ts_base == b.ts_base;
}

// Stage 1 for columns as the caller hands them over: copy-in (host), widening (narrow columns) and
// ingest, all queued without waiting.  *wait_copy = the event after which host columns are free again.
template <typename V>
static int ingest_cols(fmq_handle* h, const InCols& c, size_t n, int32_t mem, cudaEvent_t* wait_copy) {
    const bool narrow = c.ts_w == 4 || c.key_w == 4 || c.val_w == 4;
    const unsigned wgrid = (unsigned)h->num_sms * 8;
    *wait_copy = nullptr;
    if (mem == FMQ_MEM_HOST && n > 0) {
        if (h->n_staged > 0 && !(h->staged[0].n == n && same_cols(h->staged[0].c, c)))
            return set_err(h, FMQ_ERR_BAD_ARG, "batches were staged with fmq_stage_host: the next host push must pass the columns and n of the oldest one");
        if (h->n_staged == 0) { int s_ = queue_host_copy(h, c, n, false); if (s_) return s_; }
        const fmq_handle::Staged g = h->staged[0];                   // consume the oldest staged batch
        h->staged[0] = h->staged[1]; h->n_staged--;
        const size_t mb = h->max_batch;
        const u32* nar = h->hs_narrow[g.set];
        i64* sts = h->hs_ts[g.set]; u64* skey = h->hs_key[g.set]; u64* sval = h->hs_val[g.set];
        int piece = 0;
        for (size_t off = 0; off < n; off += g.chunk, ++piece) {
            const size_t m = n - off < g.chunk ? n - off : g.chunk;
            CU(h, cudaStreamWaitEvent(h->stream, h->ev_piece[g.set][piece], 0));
            if (narrow) {
                widen_columns_kernel<<<wgrid, 256, 0, h->stream>>>(c.ts_w == 4 ? nar + off : nullptr, c.ts_base, c.key_w == 4 ? nar + mb + off : nullptr,
                                                                   c.val_w == 4 ? (const int*)(nar + 2 * mb + off) : nullptr, sts + off, skey + off, sval + off, m);
                LAUNCHED(h, "widen_columns_kernel");
            }
            int s_ = stage_ingest<V>(h, sts + off, skey + off, sval + off, m); if (s_) return s_;
        }
        CU(h, cudaEventRecord(h->ev_free[g.set], h->stream));              // nothing after ingest reads the staging set
        *wait_copy = h->ev_piece[g.set][piece - 1];
        return FMQ_OK;
    }
    const i64* dts = (const i64*)c.ts; const u64* dkey = (const u64*)c.key; const void* dval = c.val;
    if (narrow && n > 0) {
        // device-resident narrow columns: widen into the handle's staging columns on its own stream
        widen_columns_kernel<<<wgrid, 256, 0, h->stream>>>(c.ts_w == 4 ? (const u32*)c.ts : nullptr, c.ts_base, c.key_w == 4 ? (const u32*)c.key : nullptr,
                                                           c.val_w == 4 ? (const int*)c.val : nullptr, h->d_in_ts, h->d_in_key, h->d_in_val, n);
        LAUNCHED(h, "widen_columns_kernel");
        if (c.ts_w == 4) dts = h->d_in_ts;
        if (c.key_w == 4) dkey = h->d_in_key;
        if (c.val_w == 4) dval = h->d_in_val;
    }
    return stage_ingest<V>(h, dts, dkey, dval, n);
}

template <typename V>
static int push_cols_t(fmq_handle* h, const InCols& c, size_t n, int64_t wm_after, int32_t mem) {
    cudaEvent_t* tev = h->tev[h->pushes % fmq_handle::TIMING_RING];
    CU(h, cudaEventRecord(tev[0], h->stream));
    cudaEvent_t wait_copy = nullptr;
    int s = ingest_cols<V>(h, c, n, mem, &wait_copy); if (s) return s;
    s = stage_fire<V>(h, n, wm_after, true); if (s) return s;
    CU(h, cudaEventRecord(tev[3], h->stream));
    h->pushes++;
    // the caller owns host columns only until this call returns: wait for the last copy
    if (wait_copy) CU(h, cudaEventSynchronize(wait_copy));
    return FMQ_OK;
}

// argument checks shared by every call that takes input columns
static int check_cols(fmq_handle* h, const InCols& c, size_t n, int32_t mem) {
    if (h->sticky) return FMQ_ERR_POISONED;
    if (n > 0 && (!c.ts || !c.key || !c.val)) return set_err(h, FMQ_ERR_BAD_ARG, "null column pointer");
    if (n > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "batch larger than fmq_config.max_batch");
    if (mem != FMQ_MEM_HOST && mem != FMQ_MEM_DEVICE) return set_err(h, FMQ_ERR_BAD_ARG, "bad mem");
    if (mem == FMQ_MEM_DEVICE && n > 0) {
        const uintptr_t wide = (c.ts_w == 8 ? (uintptr_t)c.ts : 0) | (c.key_w == 8 ? (uintptr_t)c.key : 0) | (c.val_w == 8 ? (uintptr_t)c.val : 0);
        const uintptr_t thin = (c.ts_w == 4 ? (uintptr_t)c.ts : 0) | (c.key_w == 4 ? (uintptr_t)c.key : 0) | (c.val_w == 4 ? (uintptr_t)c.val : 0);
        if ((wide & 7) || (thin & 3)) return set_err(h, FMQ_ERR_BAD_ARG, "device columns must be aligned to their element size (32-byte alignment of 8-byte columns enables 256-bit loads)");
    }
    return FMQ_OK;
}

static int push_cols(fmq_handle* h, const InCols& c, size_t n, int64_t wm_after, int32_t mem) {
    int s = check_cols(h, c, n, mem); if (s) return s;
    if (h->world > 1 || h->overlap) return set_err(h, FMQ_ERR_BAD_ARG, "peers attached or overlapped schedule: use fmq_push_local / fmq_fire_global / fmq_retire");
    CU(h, cudaSetDevice(h->device));
    return h->cfg.value_type == FMQ_VAL_F64 ? push_cols_t<double>(h, c, n, wm_after, mem) : push_cols_t<i64>(h, c, n, wm_after, mem);
}

static u32 xs_ingest_blocks();
static XPeers xs_peers(const fmq_handle* h);

// combiner mode, stage 1 (fmq_push_local / fmq_push_local_packed)
static int push_local_cols(fmq_handle* h, const InCols& c, size_t n, int32_t mem) {
    int s = check_cols(h, c, n, mem); if (s) return s;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaEventRecord(h->tev[h->pushes % fmq_handle::TIMING_RING][0], h->stream));
    cudaEvent_t wait_copy = nullptr;
    s = h->cfg.value_type == FMQ_VAL_F64 ? ingest_cols<double>(h, c, n, mem, &wait_copy) : ingest_cols<i64>(h, c, n, mem, &wait_copy);
    if (s) return s;
    if (h->cmb_flags) {
        // this rank's batch scalars go straight into every peer's control block, behind the batch number
        const u64 seq = ++h->cmb_seq;
        cmb_publish_kernel<<<1, 32, 0, h->stream>>>(h->d_st, xs_peers(h), (u32)h->world, (u32)h->rank, (u32)(seq & 1), seq); LAUNCHED(h, "cmb_publish_kernel");
    } else {
        pack_scalars_kernel<<<1, 32, 0, h->stream>>>(h->d_st, h->d_xchg); LAUNCHED(h, "pack_scalars_kernel");
    }
    h->pending_n += n;
    if (wait_copy) CU(h, cudaEventSynchronize(wait_copy));
    return FMQ_OK;
}

extern "C" {

int32_t fmq_push_batch(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int64_t wm_after, int32_t mem) {
    if (!h) return FMQ_ERR_BAD_ARG;
    const InCols c{ts, key, val, 8, 8, 8, 0};
    return push_cols(h, c, n, wm_after, mem);
}

static int check_packed(fmq_handle* h, const fmq_packed_columns* p) {
    if (!p || p->struct_size != sizeof(fmq_packed_columns)) return set_err(h, FMQ_ERR_BAD_ARG, "fmq_packed_columns.struct_size mismatch");
    if ((p->ts_width != 4 && p->ts_width != 8) || (p->key_width != 4 && p->key_width != 8) || (p->val_width != 4 && p->val_width != 8))
        return set_err(h, FMQ_ERR_BAD_ARG, "column widths must be 4 or 8");
    if (p->val_width == 4 && h->cfg.value_type != FMQ_VAL_I64) return set_err(h, FMQ_ERR_BAD_ARG, "4-byte values are int32 and need an FMQ_VAL_I64 handle");
    if (p->ts_width == 4 && p->ts_base > INT64_MAX - 0xFFFFFFFFll) return set_err(h, FMQ_ERR_BAD_ARG, "ts_base + 2^32 overflows int64");
    return FMQ_OK;
}

int32_t fmq_stage_host(fmq_handle* h, const fmq_packed_columns* p, size_t n) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    int s = check_packed(h, p); if (s) return s;
    if (n == 0 || !p->ts || !p->key || !p->val) return set_err(h, FMQ_ERR_BAD_ARG, "fmq_stage_host needs a non-empty batch");
    if (n > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "batch larger than fmq_config.max_batch");
    if (h->n_staged >= 2) return set_err(h, FMQ_ERR_BAD_ARG, "two batches are already staged: push the oldest first");
    CU(h, cudaSetDevice(h->device));
    const InCols c{p->ts, p->key, p->val, (int)p->ts_width, (int)p->key_width, (int)p->val_width, p->ts_width == 4 ? p->ts_base : 0};
    return queue_host_copy(h, c, n, true);
}

int32_t fmq_push_packed(fmq_handle* h, const fmq_packed_columns* p, size_t n, int64_t wm_after, int32_t mem) {
    if (!h) return FMQ_ERR_BAD_ARG;
    int s = check_packed(h, p); if (s) return s;
    const InCols c{p->ts, p->key, p->val, (int)p->ts_width, (int)p->key_width, (int)p->val_width, p->ts_width == 4 ? p->ts_base : 0};
    return push_cols(h, c, n, wm_after, mem);
}

int32_t fmq_flush(fmq_handle* h, int64_t watermark) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (watermark == FMQ_WM_AUTO || watermark == FMQ_WM_KEEP) return set_err(h, FMQ_ERR_BAD_ARG, "flush needs an explicit watermark");
    if (h->world > 1 || h->overlap) return set_err(h, FMQ_ERR_BAD_ARG, "peers attached or overlapped schedule: use fmq_fire_global / fmq_retire");
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaEventRecord(h->tev[h->pushes % fmq_handle::TIMING_RING][0], h->stream));
    int s = (h->cfg.value_type == FMQ_VAL_F64) ? run_chain<double>(h, nullptr, nullptr, nullptr, 0, watermark)
                                               : run_chain<i64>(h, nullptr, nullptr, nullptr, 0, watermark);
    if (s) return s;
    CU(h, cudaEventRecord(h->tev[h->pushes % fmq_handle::TIMING_RING][3], h->stream));
    h->pushes++;
    return FMQ_OK;
}

// copies rows [cur, cur+take) of `ncols` 8-byte columns to the caller; resets the device cursor
// when everything has been drained
static int drain_cols(fmq_handle* h, void* const* dcols, void* const* ocols, int ncols, u64 DevState::*counter, size_t* cursor,
                      size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!n_out) return set_err(h, FMQ_ERR_BAD_ARG, "n_out is null");
    if (h->sticky) return FMQ_ERR_POISONED;
    int s = sync_state(h); if (s) return s;
    const size_t total = (size_t)(h->h_st->*counter);
    const size_t avail = total > *cursor ? total - *cursor : 0;
    const size_t take = avail < cap ? avail : cap;
    const cudaMemcpyKind kind = mem == FMQ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (take > 0) {
        for (int c = 0; c < ncols; ++c) {
            if (!ocols[c] || !dcols[c]) continue;
            CU(h, cudaMemcpyAsync(ocols[c], (const char*)dcols[c] + *cursor * 8, take * 8, kind, h->stream));
        }
    }
    *cursor += take;
    if (n_left) *n_left = avail - take;
    if (avail - take == 0 && total > 0) {
        const size_t off = (size_t)((char*)&(h->d_st->*counter) - (char*)h->d_st);
        CU(h, cudaMemsetAsync((char*)h->d_st + off, 0, 8, h->stream));
        *cursor = 0;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    *n_out = take;
    return FMQ_OK;
}

int32_t fmq_poll_fired(fmq_handle* h, const fmq_fired_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    void* d[6] = {h->f_key, h->f_wend, h->f_count, h->f_sum, h->f_min, h->f_max};
    void* c[6] = {o->key, o->window_end, o->count, o->sum, o->min, o->max};
    return drain_cols(h, d, c, 6, &DevState::n_fired, &h->cur_fired, cap, mem, n_out, n_left);
}
int32_t fmq_poll_window_alerts(fmq_handle* h, const fmq_walert_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    void* d[3] = {h->wa_key, h->wa_wend, h->wa_val};
    void* c[3] = {o->key, o->window_end, o->value};
    return drain_cols(h, d, c, 3, &DevState::n_walert, &h->cur_walert, cap, mem, n_out, n_left);
}
// ---- overlapped polls --------------------------------------------------------------------------------
// fmq_poll_all_begin waits for the handle's kernels, reads the row counts, queues the device->host
// copies of the fired / window-alert / event-alert columns on a second stream and switches the
// kernels to a second set of output columns, so the caller can already push the next micro-batch
// (its host->device copy runs on the other PCIe direction at the same time).  fmq_poll_all_end waits
// for the copies.  Host buffers should be pinned.
int32_t fmq_poll_all_begin(fmq_handle* h, const fmq_fired_cols* f, size_t f_cap, const fmq_walert_cols* wa, size_t wa_cap,
                           const fmq_event_cols* ea, size_t ea_cap, size_t* counts3) {
    if (!h || !counts3) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (!h->out_stream) return set_err(h, FMQ_ERR_BAD_ARG, "overlapped polls need fmq_config.overlapped_polls = 1");
    if (h->out_pending) return set_err(h, FMQ_ERR_BAD_ARG, "fmq_poll_all_end has not been called for the previous fmq_poll_all_begin");
    if (h->cur_fired || h->cur_walert || h->cur_ealert) return set_err(h, FMQ_ERR_BAD_ARG, "finish the partial fmq_poll_* first");
    int s = sync_state(h); if (s) return s;
    const DevState& d = *h->h_st;
    const size_t nf = d.n_fired, nw = d.n_walert, ne = d.n_ealert;
    if ((nf && (!f || nf > f_cap)) || (nw && (!wa || nw > wa_cap)) || (ne && (!ea || ne > ea_cap))) return set_err(h, FMQ_ERR_BAD_ARG, "caller buffers smaller than the rows waiting");
    CU(h, cudaEventRecord(h->ev_out, h->stream));
    CU(h, cudaStreamWaitEvent(h->out_stream, h->ev_out, 0));
    void* cur[12] = {h->f_key, h->f_wend, h->f_count, h->f_sum, h->f_min, h->f_max, h->wa_key, h->wa_wend, h->wa_val, h->ea_ts, h->ea_key, h->ea_val};
    void* dst[12] = {f ? f->key : nullptr, f ? f->window_end : nullptr, f ? f->count : nullptr, f ? f->sum : nullptr, f ? f->min : nullptr, f ? f->max : nullptr,
                     wa ? wa->key : nullptr, wa ? wa->window_end : nullptr, wa ? wa->value : nullptr, ea ? ea->ts : nullptr, ea ? ea->key : nullptr, ea ? ea->val : nullptr};
    const size_t cnt[12] = {nf, nf, nf, nf, nf, nf, nw, nw, nw, ne, ne, ne};
    for (int i = 0; i < 12; ++i) if (cnt[i] && dst[i] && cur[i]) CU(h, cudaMemcpyAsync(dst[i], cur[i], cnt[i] * 8, cudaMemcpyDeviceToHost, h->out_stream));
    // the kernels of the next push append into the other set, from row 0
    CU(h, cudaMemsetAsync((char*)h->d_st + offsetof(DevState, n_fired), 0, 8, h->stream));
    CU(h, cudaMemsetAsync((char*)h->d_st + offsetof(DevState, n_walert), 0, 8, h->stream));
    CU(h, cudaMemsetAsync((char*)h->d_st + offsetof(DevState, n_ealert), 0, 8, h->stream));
    void** slots[12] = {(void**)&h->f_key, (void**)&h->f_wend, (void**)&h->f_count, (void**)&h->f_sum, (void**)&h->f_min, (void**)&h->f_max,
                        (void**)&h->wa_key, (void**)&h->wa_wend, (void**)&h->wa_val, (void**)&h->ea_ts, (void**)&h->ea_key, (void**)&h->ea_val};
    for (int i = 0; i < 12; ++i) { void* t = *slots[i]; *slots[i] = h->alt_out[i]; h->alt_out[i] = t; }
    counts3[0] = nf; counts3[1] = nw; counts3[2] = ne;
    h->out_pending = true;
    return FMQ_OK;
}
int32_t fmq_poll_all_end(fmq_handle* h) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (!h->out_pending) return FMQ_OK;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->out_stream));
    h->out_pending = false;
    return FMQ_OK;
}

int32_t fmq_poll_late(fmq_handle* h, const fmq_event_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    if (!h->lt_cap) return set_err(h, FMQ_ERR_BAD_ARG, "the late side output is not enabled (fmq_config.late_capacity)");
    void* d[3] = {h->lt_ts, h->lt_key, h->lt_val};
    void* c[3] = {o->ts, o->key, o->val};
    return drain_cols(h, d, c, 3, &DevState::n_late, &h->cur_late, cap, mem, n_out, n_left);
}
int32_t fmq_poll_medians(fmq_handle* h, const fmq_walert_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    if (!h->ml_cap) return set_err(h, FMQ_ERR_BAD_ARG, "the window median is not enabled (fmq_config.median_log_capacity)");
    void* d[3] = {h->md_key, h->md_wend, h->md_val};
    void* c[3] = {o->key, o->window_end, o->value};
    return drain_cols(h, d, c, 3, &DevState::n_median, &h->cur_median, cap, mem, n_out, n_left);
}
int32_t fmq_poll_event_alerts(fmq_handle* h, const fmq_event_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    void* d[3] = {h->ea_ts, h->ea_key, h->ea_val};
    void* c[3] = {o->ts, o->key, o->val};
    return drain_cols(h, d, c, 3, &DevState::n_ealert, &h->cur_ealert, cap, mem, n_out, n_left);
}
int32_t fmq_poll_consec_alerts(fmq_handle* h, const fmq_event_cols* o, size_t cap, int32_t mem, size_t* n_out, size_t* n_left) {
    if (!h || !o) return FMQ_ERR_BAD_ARG;
    void* d[3] = {h->ca_ts, h->ca_key, h->ca_val};
    void* c[3] = {o->ts, o->key, o->val};
    return drain_cols(h, d, c, 3, &DevState::n_calert, &h->cur_calert, cap, mem, n_out, n_left);
}
int32_t fmq_poll_rolling_max(fmq_handle* h, void* out, size_t cap, int32_t mem, size_t* n_out) {
    if (!h || !n_out) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    int s = sync_state(h); if (s) return s;
    const size_t take = h->rolling_n < cap ? h->rolling_n : cap;
    if (take > 0) {
        if (!out) return set_err(h, FMQ_ERR_BAD_ARG, "out is null");
        CU(h, cudaMemcpyAsync(out, h->d_rolling, take * 8, mem == FMQ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
    }
    *n_out = take;
    return FMQ_OK;
}

int32_t fmq_get_counters(fmq_handle* h, fmq_counters* out) {
    if (!h || !out) return FMQ_ERR_BAD_ARG;
    int s = sync_state(h);
    const DevState& d = *h->h_st;
    out->events_in = d.events_in; out->late_dropped = d.late_dropped; out->windows_fired = d.windows_fired;
    out->window_alerts = d.window_alerts; out->event_alerts = d.event_alerts; out->consec_alerts = d.consec_alerts;
    out->keys = d.keys; out->watermark = d.wm; out->gpu_launches = h->launches;
    return s;
}

int32_t fmq_get_key_state(fmq_handle* h, const uint64_t* keys, size_t n, int32_t* found, int32_t* consec, void* rolling) {
    if (!h || (n && (!keys || !found || !consec || !rolling))) return FMQ_ERR_BAD_ARG;
    if (!keyed_enabled(h->cfg)) return set_err(h, FMQ_ERR_BAD_ARG, "no keyed state operator is enabled");
    if (n == 0) return FMQ_OK;
    CU(h, cudaSetDevice(h->device));
    // one temporary buffer, released on every path out of this function
    struct Tmp { void* p = nullptr; ~Tmp() { if (p) cudaFree(p); } } tmp;
    CU(h, cudaMalloc(&tmp.p, n * 24));
    u64* dk = static_cast<u64*>(tmp.p); u64* dr = dk + n; int* df = reinterpret_cast<int*>(dr + n); int* dc = df + n;
    CU(h, cudaMemcpyAsync(dk, keys, n * 8, cudaMemcpyHostToDevice, h->stream));
    key_state_lookup_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(dk, n, h->d_keys, h->capacity, h->cfg.key_mode == FMQ_KEYS_DENSE, h->d_kflags, h->d_kmax, df, dc, dr);
    LAUNCHED(h, "key_state_lookup_kernel");
    CU(h, cudaMemcpyAsync(found, df, n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(consec, dc, n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(rolling, dr, n * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return FMQ_OK;
}

// ---- combiner mode: multi-GPU without an event shuffle ------------------------------------------
void* fmq_table_ptr(fmq_handle* h) { return h ? (void*)h->d_acc : nullptr; }
void* fmq_exchange_ptr(fmq_handle* h) { return h ? (void*)h->d_xchg : nullptr; }

int32_t fmq_ipc_export(fmq_handle* h, void* handle64) {
    if (!h || !handle64 || !h->d_acc) return h ? set_err(h, FMQ_ERR_BAD_ARG, "no accumulator table to export") : FMQ_ERR_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CU(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t ih;
    CU(h, cudaIpcGetMemHandle(&ih, h->d_acc));
    memcpy(handle64, &ih, 64);
    return FMQ_OK;
}
int32_t fmq_ipc_open(fmq_handle* h, const void* handle64, void** ptr) {
    if (!h || !handle64 || !ptr) return FMQ_ERR_BAD_ARG;
    CU(h, cudaSetDevice(h->device));
    cudaIpcMemHandle_t ih; memcpy(&ih, handle64, 64);
    CU(h, cudaIpcOpenMemHandle(ptr, ih, cudaIpcMemLazyEnablePeerAccess));
    return FMQ_OK;
}
int32_t fmq_attach_peers(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_tables) {
    if (!h || !peer_tables) return FMQ_ERR_BAD_ARG;
    if (world < 1 || world > FMQ_MAX_PEERS || rank < 0 || rank >= world) return set_err(h, FMQ_ERR_BAD_ARG, "need 1 <= world <= 8 and 0 <= rank < world");
    if (!h->cfg.window_enabled) return set_err(h, FMQ_ERR_BAD_ARG, "combiner mode needs window aggregation");
    if (h->cfg.key_mode != FMQ_KEYS_DENSE) return set_err(h, FMQ_ERR_BAD_ARG, "combiner mode needs FMQ_KEYS_DENSE (slots must mean the same key on every rank)");
    if (h->ml_cap) return set_err(h, FMQ_ERR_BAD_ARG, "the window median needs every value of a key on one rank: route it through the all-to-all exchange instead");
    if (keyed_enabled(h->cfg)) return set_err(h, FMQ_ERR_BAD_ARG, "keyed ValueState operators depend on per-key order: route them through the all-to-all exchange instead");
    if (h->g.lateness > 0) return set_err(h, FMQ_ERR_BAD_ARG, "allowed_lateness_ms is not supported in combiner mode: route the job through the keyBy exchange instead");
    for (int r = 0; r < world; ++r) {
        if (r != rank && !peer_tables[r]) return set_err(h, FMQ_ERR_BAD_ARG, "null peer table");
        h->peer_acc[r] = (r == rank) ? h->d_acc : static_cast<const Acc32*>(peer_tables[r]);
    }
    h->world = world; h->rank = rank;
    return FMQ_OK;
}
int32_t fmq_push_local(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int32_t mem) {
    if (!h) return FMQ_ERR_BAD_ARG;
    const InCols c{ts, key, val, 8, 8, 8, 0};
    return push_local_cols(h, c, n, mem);
}
int32_t fmq_push_local_packed(fmq_handle* h, const fmq_packed_columns* p, size_t n, int32_t mem) {
    if (!h) return FMQ_ERR_BAD_ARG;
    int s = check_packed(h, p); if (s) return s;
    const InCols c{p->ts, p->key, p->val, (int)p->ts_width, (int)p->key_width, (int)p->val_width, p->ts_width == 4 ? p->ts_base : 0};
    return push_local_cols(h, c, n, mem);
}
int32_t fmq_fire_global(fmq_handle* h, int64_t watermark_after) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    CU(h, cudaSetDevice(h->device));
    if (h->overlap) CU(h, cudaStreamWaitEvent(h->stream, h->ev_bdone, 0));     // the previous batch's retire / commit
    if (h->cmb_flags) {
        if (h->cmb_seq == 0) return set_err(h, FMQ_ERR_BAD_ARG, "with device flags every fmq_fire_global follows an fmq_push_local (n = 0 for a flush)");
        // waits, on the device, until every rank has published this batch's scalars (= "every rank has ingested") and takes their maximum
        cmb_reduce_kernel<<<1, 32, 0, h->stream>>>(h->d_st, h->xs_ctl[h->rank], (u32)h->world, (u32)(h->cmb_seq & 1), h->cmb_seq, peer_timeout_ns()); LAUNCHED(h, "cmb_reduce_kernel");
    } else {
        unpack_scalars_kernel<<<1, 32, 0, h->stream>>>(h->d_st, h->d_xchg); LAUNCHED(h, "unpack_scalars_kernel");
    }
    int s = (h->cfg.value_type == FMQ_VAL_F64) ? stage_fire<double>(h, h->pending_n, watermark_after, false)
                                               : stage_fire<i64>(h, h->pending_n, watermark_after, false);
    if (s) return s;
    if (h->cmb_flags) {                                            // "this rank has finished reading its peers' tables", after the fire kernel
        cudaStream_t fs = h->overlap ? h->fire_stream : h->stream;
        cmb_done_kernel<<<1, 32, 0, fs>>>(xs_peers(h), (u32)h->world, (u32)h->rank, (u32)(h->cmb_seq & 1), h->cmb_seq); LAUNCHED(h, "cmb_done_kernel");
    }
    h->pending_n = 0;
    return FMQ_OK;
}
int32_t fmq_retire(fmq_handle* h) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    CU(h, cudaSetDevice(h->device));
    cudaStream_t fs = h->overlap ? h->fire_stream : h->stream;
    if (h->cmb_flags) {                                            // every rank has finished reading this rank's table
        cmb_wait_done_kernel<<<1, 32, 0, fs>>>(h->xs_ctl[h->rank], (u32)h->world, (u32)(h->cmb_seq & 1), h->cmb_seq, h->d_st, peer_timeout_ns()); LAUNCHED(h, "cmb_wait_done_kernel");
    }
    retire_kernel<<<(unsigned)h->num_sms * 4, 256, 0, fs>>>(h->d_st, h->d_acc, h->capacity, h->g.ring_mask); LAUNCHED(h, "retire_kernel");
    if (h->overlap) {
        commit_overlapped_kernel<<<1, 32, 0, fs>>>(h->d_st); LAUNCHED(h, "commit_overlapped_kernel");
        CU(h, cudaEventRecord(h->ev_bdone, fs));
    } else {
        commit_kernel<<<1, 32, 0, fs>>>(h->d_st); LAUNCHED(h, "commit_kernel");
        // pre-arm the exchange buffer so that a fire with no preceding local push contributes nothing
        if (!h->cmb_flags) { pack_scalars_kernel<<<1, 32, 0, fs>>>(h->d_st, h->d_xchg); LAUNCHED(h, "pack_scalars_kernel"); }
    }
    CU(h, cudaEventRecord(h->tev[h->pushes % fmq_handle::TIMING_RING][3], fs));
    h->pushes++;
    return FMQ_OK;
}
int32_t fmq_combiner_overlap(fmq_handle* h, int32_t enable) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (enable && h->g.size != h->g.slide) return set_err(h, FMQ_ERR_BAD_ARG, "only tumbling windows can fire beside the next ingest (a sliding window's panes are still being updated)");
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->fire_stream) CU(h, cudaStreamSynchronize(h->fire_stream));
    if (enable && !h->fire_stream) {
        CU(h, cudaStreamCreateWithFlags(&h->fire_stream, cudaStreamNonBlocking));
        CU(h, cudaEventCreateWithFlags(&h->ev_plan, cudaEventDisableTiming));
        CU(h, cudaEventCreateWithFlags(&h->ev_bdone, cudaEventDisableTiming));
    }
    h->overlap = enable != 0;
    return FMQ_OK;
}
void* fmq_fire_stream(fmq_handle* h) { return h ? (void*)(h->overlap ? h->fire_stream : h->stream) : nullptr; }
int32_t fmq_combiner_use_flags(fmq_handle* h, int32_t enable) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (enable) {
        if (h->world < 2) return set_err(h, FMQ_ERR_BAD_ARG, "fmq_combiner_use_flags follows fmq_attach_peers");
        if (!h->xs_attached || h->xs_world != h->world || h->xs_rank != h->rank)
            return set_err(h, FMQ_ERR_BAD_ARG, "the control blocks of all ranks must be attached first (fmq_shard_alloc + fmq_shard_attach[_ipc], same rank and world)");
    }
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->fire_stream) CU(h, cudaStreamSynchronize(h->fire_stream));
    h->cmb_flags = enable != 0;
    return FMQ_OK;
}

// ---- text ingest ------------------------------------------------------------------------------------
uint64_t fmq_text_key_hash(const char* s, size_t len) {
    u64 hsh = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < len; ++i) { hsh ^= (unsigned char)s[i]; hsh *= 0x100000001b3ull; }
    hsh = mix64(hsh);
    return hsh == KEY_EMPTY ? 0 : hsh;
}
// parses into the handle's staging columns (d_in_ts / d_in_key / d_in_val) and d_txt_status
static int parse_text_to_staging(fmq_handle* h, const char* text, size_t bytes, int32_t text_mem, int32_t format, int32_t utc, size_t* n_lines, size_t* n_bad) {
    if (!h->d_text) return set_err(h, FMQ_ERR_BAD_ARG, "text ingest needs fmq_config.text_max_bytes > 0");
    if (bytes > h->cfg.text_max_bytes) return set_err(h, FMQ_ERR_BAD_ARG, "text larger than fmq_config.text_max_bytes");
    if (format != FMQ_TEXT_CPU && format != FMQ_TEXT_FLOW) return set_err(h, FMQ_ERR_BAD_ARG, "bad text format");
    if ((format == FMQ_TEXT_CPU) != (h->cfg.value_type == FMQ_VAL_F64)) return set_err(h, FMQ_ERR_BAD_ARG, "FMQ_TEXT_CPU yields f64 values, FMQ_TEXT_FLOW int64: value_type does not match");
    *n_lines = 0; *n_bad = 0;
    if (bytes == 0) return FMQ_OK;
    if (!text) return set_err(h, FMQ_ERR_BAD_ARG, "null text");
    CU(h, cudaSetDevice(h->device));
    const char* d = text;
    if (text_mem == FMQ_MEM_HOST) { CU(h, cudaMemcpyAsync(h->d_text, text, bytes, cudaMemcpyHostToDevice, h->stream)); d = h->d_text; }
    const u32 nblocks = (u32)((bytes + TXT_CHUNK - 1) / TXT_CHUNK);
    txt_count_kernel<<<nblocks, TXT_THREADS, 0, h->stream>>>(d, bytes, h->d_counts); LAUNCHED(h, "txt_count_kernel");
    CU(h, cudaMemsetAsync(h->d_counts + nblocks, 0, 4, h->stream));
    int s = exclusive_scan_counts(h, (size_t)nblocks + 1); if (s) return s;
    u32 newlines = 0; char last = 0;
    CU(h, cudaMemcpyAsync(&newlines, h->d_counts + nblocks, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(&last, d + bytes - 1, 1, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    const size_t lines = (size_t)newlines + (last != '\n' ? 1 : 0);
    if (lines > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "more lines than fmq_config.max_batch");
    txt_mark_kernel<<<nblocks, TXT_THREADS, 0, h->stream>>>(d, bytes, h->d_counts, h->d_line_start, h->max_batch + 2); LAUNCHED(h, "txt_mark_kernel");
    if (last != '\n') { const u64 end = bytes; CU(h, cudaMemcpyAsync(h->d_line_start + lines, &end, 8, cudaMemcpyHostToDevice, h->stream)); }
    CU(h, cudaMemsetAsync(h->d_txt_bad, 0, 8, h->stream));
    TextArgs a{d, bytes, h->d_line_start, lines, format, utc, h->d_in_ts, h->d_in_key, h->d_in_val, h->d_txt_status, h->d_txt_bad, h->d_key_off, h->d_key_len};
    txt_parse_kernel<<<(unsigned)((lines + 255) / 256), 256, 0, h->stream>>>(a); LAUNCHED(h, "txt_parse_kernel");
    if (h->dt_hash && lines > 0) {                                         // strings -> dense ids, in order of first sight
        DictArgs da{h->dt_hash, h->dt_id, h->dt_mask, h->dn_off, h->dn_len, h->dn_arena, h->dn_arena_cap, h->cfg.text_dict_capacity, h->dt_counters,
                    d, h->d_key_off, h->d_key_len, lines, h->d_in_key, h->d_txt_status, h->d_txt_bad};
        dict_insert_kernel<<<(unsigned)((lines + 255) / 256), 256, 0, h->stream>>>(da); LAUNCHED(h, "dict_insert_kernel");
        dict_lookup_kernel<<<(unsigned)((lines + 255) / 256), 256, 0, h->stream>>>(da); LAUNCHED(h, "dict_lookup_kernel");
    }
    unsigned long long bad = 0;
    CU(h, cudaMemcpyAsync(&bad, h->d_txt_bad, 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *n_lines = lines; *n_bad = (size_t)bad;
    return FMQ_OK;
}
int32_t fmq_parse_text(fmq_handle* h, const char* text, size_t bytes, int32_t text_mem, int32_t format, int32_t utc, int64_t* ts_out, uint64_t* key_out,
                       void* val_out, uint32_t* status_out, size_t cap, int32_t out_mem, size_t* n_lines, size_t* n_bad) {
    if (!h || !n_lines || !n_bad) return FMQ_ERR_BAD_ARG;
    int s = parse_text_to_staging(h, text, bytes, text_mem, format, utc, n_lines, n_bad); if (s) return s;
    if (*n_lines > cap) return set_err(h, FMQ_ERR_BAD_ARG, "output buffers smaller than the number of lines");
    const cudaMemcpyKind kind = out_mem == FMQ_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    const size_t n = *n_lines;
    if (n) {
        if (ts_out) CU(h, cudaMemcpyAsync(ts_out, h->d_in_ts, n * 8, kind, h->stream));
        if (key_out) CU(h, cudaMemcpyAsync(key_out, h->d_in_key, n * 8, kind, h->stream));
        if (val_out) CU(h, cudaMemcpyAsync(val_out, h->d_in_val, n * 8, kind, h->stream));
        if (status_out) CU(h, cudaMemcpyAsync(status_out, h->d_txt_status, n * 4, kind, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
    }
    return FMQ_OK;
}
int32_t fmq_push_text(fmq_handle* h, const char* text, size_t bytes, int32_t text_mem, int32_t format, int32_t utc, int64_t wm_after, size_t* n_lines) {
    if (!h || !n_lines) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    size_t bad = 0;
    int s = parse_text_to_staging(h, text, bytes, text_mem, format, utc, n_lines, &bad); if (s) return s;
    if (bad) { h->err = std::to_string(bad) + " line(s) are malformed or need a host-side value parse: use fmq_parse_text and fix them"; return FMQ_ERR_BAD_ARG; }
    return fmq_push_batch(h, (const int64_t*)h->d_in_ts, (const uint64_t*)h->d_in_key, h->d_in_val, *n_lines, wm_after, FMQ_MEM_DEVICE);
}

int32_t fmq_dict_size(fmq_handle* h, size_t* n_ids) {
    if (!h || !n_ids) return FMQ_ERR_BAD_ARG;
    if (!h->dt_hash) return set_err(h, FMQ_ERR_BAD_ARG, "the device dictionary is not enabled (fmq_config.text_dict_capacity)");
    CU(h, cudaSetDevice(h->device));
    unsigned long long c[4];
    CU(h, cudaMemcpyAsync(c, h->dt_counters, 32, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    *n_ids = (size_t)(c[0] < h->cfg.text_dict_capacity ? c[0] : h->cfg.text_dict_capacity);
    return FMQ_OK;
}
int32_t fmq_dict_lookup(fmq_handle* h, const uint64_t* ids, size_t n, char* names_out, size_t stride, uint32_t* lens_out) {
    if (!h || (n && (!ids || !names_out || !lens_out)) || stride == 0) return FMQ_ERR_BAD_ARG;
    if (!h->dt_hash) return set_err(h, FMQ_ERR_BAD_ARG, "the device dictionary is not enabled (fmq_config.text_dict_capacity)");
    if (n == 0) return FMQ_OK;
    size_t n_ids = 0;
    int s = fmq_dict_size(h, &n_ids); if (s) return s;
    struct Tmp { void* p = nullptr; ~Tmp() { if (p) cudaFree(p); } } tmp;
    const size_t bytes = n * 8 + n * 4 + n * stride;
    CU(h, cudaMalloc(&tmp.p, bytes));
    u64* d_ids = static_cast<u64*>(tmp.p); u32* d_lens = reinterpret_cast<u32*>(d_ids + n); char* d_rows = reinterpret_cast<char*>(d_lens + n);
    CU(h, cudaMemcpyAsync(d_ids, ids, n * 8, cudaMemcpyHostToDevice, h->stream));
    dict_names_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(d_ids, n, h->dn_off, h->dn_len, h->dn_arena, (u64)n_ids, d_rows, stride, d_lens); LAUNCHED(h, "dict_names_kernel");
    CU(h, cudaMemcpyAsync(lens_out, d_lens, n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(names_out, d_rows, n * stride, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return FMQ_OK;
}

// ---- snapshot / restore (Flink snapshotState / initializeState) --------------------
namespace {
struct SnapHeader {
    uint64_t magic; uint32_t version; uint32_t header_bytes;
    fmq_config cfg; uint64_t capacity; uint32_t ring; uint32_t has_window, has_keyed, anchored_known;
    uint64_t total_bytes;
};
constexpr uint64_t SNAP_MAGIC = 0x31504E53514D46ull;   // "FMQSNP1"
size_t snap_bytes(const fmq_handle* h) {
    size_t b = sizeof(SnapHeader) + sizeof(DevState) + h->capacity * 8;
    if (h->cfg.window_enabled) b += (size_t)h->ring * h->capacity * sizeof(Acc32);
    if (keyed_enabled(h->cfg)) b += h->capacity * 4 + h->capacity * 8;
    if (h->ml_cap) b += (size_t)h->ml_cap * 20;                  // the window median's value log (pane, slot, value)
    return b;
}
bool same_job(const fmq_config& a, const fmq_config& b) {
    return a.value_type == b.value_type && a.key_mode == b.key_mode && a.window_enabled == b.window_enabled &&
           a.window_size_ms == b.window_size_ms && a.window_slide_ms == b.window_slide_ms && a.window_offset_ms == b.window_offset_ms &&
           (a.agg_mask | FMQ_AGG_COUNT) == (b.agg_mask | FMQ_AGG_COUNT) && a.rolling_max_enabled == b.rolling_max_enabled &&
           (a.consec_n > 0) == (b.consec_n > 0);
}
}  // namespace

int32_t fmq_snapshot_size(fmq_handle* h, size_t* bytes) {
    if (!h || !bytes) return FMQ_ERR_BAD_ARG;
    *bytes = snap_bytes(h);
    return FMQ_OK;
}
int32_t fmq_snapshot(fmq_handle* h, void* buf, size_t cap, size_t* written) {
    if (!h || !buf || !written) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (h->world > 1) return set_err(h, FMQ_ERR_BAD_ARG, "snapshot of a combiner-mode handle is not supported");
    const size_t need = snap_bytes(h);
    if (cap < need) return set_err(h, FMQ_ERR_BAD_ARG, "snapshot buffer too small (see fmq_snapshot_size)");
    int st = sync_state(h); if (st) return st;
    const DevState& d = *h->h_st;
    if (d.n_fired || d.n_walert || d.n_ealert || d.n_calert || d.n_median || d.n_late) return set_err(h, FMQ_ERR_BAD_ARG, "poll every output before taking a snapshot: emitted rows are not operator state");
    char* p = static_cast<char*>(buf);
    SnapHeader hd{}; hd.magic = SNAP_MAGIC; hd.version = 1; hd.header_bytes = sizeof(SnapHeader); hd.cfg = h->cfg; hd.capacity = h->capacity; hd.ring = h->ring;
    hd.has_window = h->cfg.window_enabled; hd.has_keyed = keyed_enabled(h->cfg); hd.anchored_known = h->anchored_known; hd.total_bytes = need;
    memcpy(p, &hd, sizeof(hd)); p += sizeof(hd);
    memcpy(p, &d, sizeof(DevState)); p += sizeof(DevState);
    CU(h, cudaMemcpyAsync(p, h->d_keys, h->capacity * 8, cudaMemcpyDeviceToHost, h->stream)); p += h->capacity * 8;
    if (hd.has_window) { const size_t b = (size_t)h->ring * h->capacity * sizeof(Acc32); CU(h, cudaMemcpyAsync(p, h->d_acc, b, cudaMemcpyDeviceToHost, h->stream)); p += b; }
    if (hd.has_keyed) {
        CU(h, cudaMemcpyAsync(p, h->d_kflags, h->capacity * 4, cudaMemcpyDeviceToHost, h->stream)); p += h->capacity * 4;
        CU(h, cudaMemcpyAsync(p, h->d_kmax, h->capacity * 8, cudaMemcpyDeviceToHost, h->stream)); p += h->capacity * 8;
    }
    if (h->ml_cap) {
        const int c = h->ml_cur;
        CU(h, cudaMemcpyAsync(p, h->ml_pane[c], h->ml_cap * 8, cudaMemcpyDeviceToHost, h->stream)); p += h->ml_cap * 8;
        CU(h, cudaMemcpyAsync(p, h->ml_val[c], h->ml_cap * 8, cudaMemcpyDeviceToHost, h->stream)); p += h->ml_cap * 8;
        CU(h, cudaMemcpyAsync(p, h->ml_slot[c], h->ml_cap * 4, cudaMemcpyDeviceToHost, h->stream)); p += h->ml_cap * 4;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    *written = need;
    return FMQ_OK;
}
int32_t fmq_restore(fmq_handle* h, const void* buf, size_t bytes) {
    if (!h || !buf) return FMQ_ERR_BAD_ARG;
    if (h->world > 1) return set_err(h, FMQ_ERR_BAD_ARG, "restore into a combiner-mode handle is not supported");
    if (bytes < sizeof(SnapHeader)) return set_err(h, FMQ_ERR_BAD_ARG, "not a snapshot");
    SnapHeader hd; memcpy(&hd, buf, sizeof(hd));
    if (hd.magic != SNAP_MAGIC || hd.version != 1 || hd.header_bytes != sizeof(SnapHeader)) return set_err(h, FMQ_ERR_BAD_ARG, "not a snapshot of this library version");
    if (hd.capacity != h->capacity || hd.ring != h->ring || !same_job(hd.cfg, h->cfg) || hd.total_bytes != snap_bytes(h) || bytes < hd.total_bytes)
        return set_err(h, FMQ_ERR_BAD_ARG, "snapshot was taken with a different job geometry (capacity, ring, window, value type, key mode)");
    if (h->n_staged > 0 || h->out_pending) return set_err(h, FMQ_ERR_BAD_ARG, "restore while host batches are staged or an overlapped poll is pending: consume them first");
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->fire_stream) CU(h, cudaStreamSynchronize(h->fire_stream));
    if (h->copy_stream) CU(h, cudaStreamSynchronize(h->copy_stream));
    const char* p = static_cast<const char*>(buf) + sizeof(SnapHeader);
    DevState d; memcpy(&d, p, sizeof(d)); p += sizeof(d);
    d.n_fired = d.n_walert = d.n_ealert = d.n_calert = 0; d.n_late = 0; d.n_median = 0; d.n_mwork = 0; d.n_dirty = 0; d.err_flags = 0;
    if (!h->ml_cap) d.n_mlog = 0;
    CU(h, cudaMemcpyAsync(h->d_st, &d, sizeof(d), cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(h->d_keys, p, h->capacity * 8, cudaMemcpyHostToDevice, h->stream)); p += h->capacity * 8;
    if (hd.has_window) { const size_t b = (size_t)h->ring * h->capacity * sizeof(Acc32); CU(h, cudaMemcpyAsync(h->d_acc, p, b, cudaMemcpyHostToDevice, h->stream)); p += b; }
    if (hd.has_keyed) {
        CU(h, cudaMemcpyAsync(h->d_kflags, p, h->capacity * 4, cudaMemcpyHostToDevice, h->stream)); p += h->capacity * 4;
        CU(h, cudaMemcpyAsync(h->d_kmax, p, h->capacity * 8, cudaMemcpyHostToDevice, h->stream)); p += h->capacity * 8;
    }
    if (h->ml_cap) {
        h->ml_cur = 0;
        CU(h, cudaMemcpyAsync(h->ml_pane[0], p, h->ml_cap * 8, cudaMemcpyHostToDevice, h->stream)); p += h->ml_cap * 8;
        CU(h, cudaMemcpyAsync(h->ml_val[0], p, h->ml_cap * 8, cudaMemcpyHostToDevice, h->stream)); p += h->ml_cap * 8;
        CU(h, cudaMemcpyAsync(h->ml_slot[0], p, h->ml_cap * 4, cudaMemcpyHostToDevice, h->stream)); p += h->ml_cap * 4;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    h->cur_fired = h->cur_walert = h->cur_ealert = h->cur_calert = 0; h->cur_late = 0; h->cur_median = 0;
    h->rolling_n = 0; h->pending_n = 0; h->sticky = 0;
    h->anchored_known = hd.anchored_known != 0;
    return FMQ_OK;
}

int32_t fmq_synchronize(fmq_handle* h) {
    if (!h) return FMQ_ERR_BAD_ARG;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    if (h->fire_stream) CU(h, cudaStreamSynchronize(h->fire_stream));
    return FMQ_OK;
}
int32_t fmq_last_device_ms(fmq_handle* h, float* ms) {
    float in = 0, ke = 0, fi = 0;
    return fmq_get_push_timing(h, 0, ms, &in, &ke, &fi);
}
int32_t fmq_get_push_timing(fmq_handle* h, uint32_t back, float* total_ms, float* ingest_ms, float* keyed_ms, float* fire_ms) {
    if (!h || !total_ms || !ingest_ms || !keyed_ms || !fire_ms) return FMQ_ERR_BAD_ARG;
    if (back >= fmq_handle::TIMING_RING || back >= h->pushes) return set_err(h, FMQ_ERR_BAD_ARG, "no such push in the timing ring");
    CU(h, cudaSetDevice(h->device));
    cudaEvent_t* e = h->tev[(h->pushes - 1 - back) % fmq_handle::TIMING_RING];
    CU(h, cudaEventSynchronize(e[3]));
    CU(h, cudaEventElapsedTime(total_ms, e[0], e[3]));
    CU(h, cudaEventElapsedTime(ingest_ms, e[0], e[1]));
    CU(h, cudaEventElapsedTime(keyed_ms, e[1], e[2]));
    CU(h, cudaEventElapsedTime(fire_ms, e[2], e[3]));
    return FMQ_OK;
}
void* fmq_stream(fmq_handle* h) { return h ? (void*)h->stream : nullptr; }

// ---- keyBy routing ---------------------------------------------------------------------------------
uint32_t fmq_owner_of(uint64_t key, uint32_t world) { return world ? (uint32_t)(mix64(key) % world) : 0; }
int32_t fmq_java_string_hash(const char* s, size_t len) {
    uint32_t hsh = 0; for (size_t i = 0; i < len; i++) hsh = 31u * hsh + (uint8_t)s[i]; return (int32_t)hsh;   // ASCII keys
}
static int32_t flink_murmur(int32_t in) {                      // org.apache.flink.util.MathUtils#murmurHash
    uint32_t c = (uint32_t)in;
    c *= 0xcc9e2d51u; c = (c << 15) | (c >> 17); c *= 0x1b873593u;
    c = (c << 13) | (c >> 19); c = c * 5u + 0xe6546b64u;
    c ^= 4u; c ^= c >> 16; c *= 0x85ebca6bu; c ^= c >> 13; c *= 0xc2b2ae35u; c ^= c >> 16;
    int32_t s = (int32_t)c;
    return s >= 0 ? s : (s != INT32_MIN ? -s : 0);
}
int32_t fmq_flink_key_group(int32_t key_hash, int32_t max_parallelism) { return max_parallelism > 0 ? flink_murmur(key_hash) % max_parallelism : 0; }
int32_t fmq_flink_subtask(int32_t key_group, int32_t parallelism, int32_t max_parallelism) { return max_parallelism > 0 ? key_group * parallelism / max_parallelism : 0; }

int32_t fmq_partition_by_owner(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, uint32_t world,
                               int64_t* ts_out, uint64_t* key_out, void* val_out, uint64_t* counts) {
    if (!h || !counts || world == 0 || world > RS_DIGITS) return h ? set_err(h, FMQ_ERR_BAD_ARG, "need 1 <= world <= 256") : FMQ_ERR_BAD_ARG;
    if (n > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "batch larger than fmq_config.max_batch");
    for (uint32_t w = 0; w < world; ++w) counts[w] = 0;
    if (n == 0) return FMQ_OK;
    if (!ts || !key || !val || !ts_out || !key_out || !val_out) return set_err(h, FMQ_ERR_BAD_ARG, "null column pointer");
    CU(h, cudaSetDevice(h->device));
    const u32 dense = h->cfg.key_mode == FMQ_KEYS_DENSE;
    SplitDst dst{}; dst.out_ts = (u64*)ts_out; dst.out_key = (u64*)key_out; dst.out_val = (u64*)val_out;
    cudaError_t e = owner_split(h->stream, SplitScratch{h->d_counts, h->d_tiles}, (const u64*)ts, (const u64*)key, (const u64*)val, n, world, dense, dst, false, nullptr);
    if (e != cudaSuccess) { h->err = std::string("owner split: ") + cudaGetErrorString(e); return FMQ_ERR_CUDA; }
    h->launches += 5;
    // bucket starts = scanned offsets of chunk 0 for each digit
    const u32 nchunks = xs_num_chunks(n);
    u32 starts[RS_DIGITS];
    CU(h, cudaMemcpy2DAsync(starts, 4, h->d_counts, (size_t)nchunks * 4, 4, world, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    for (uint32_t w = 0; w < world; ++w) counts[w] = (w + 1 < world ? starts[w + 1] : (u32)n) - starts[w];
    return FMQ_OK;
}

}  // extern "C"

// ---- keyBy exchange over NVLink peer memory ("peer pull") ----------------------------------------
// The all-to-all of the reference's keyBy (BandwidthMonitorWithEventTime.java:45) without a send side:
// every rank splits its slice into owner buckets inside a buffer all peers have mapped (CUDA IPC), the
// ranks exchange the world x world bucket-size matrix, and each rank then PULLS its buckets straight
// out of the peers' buffers into its own staging columns (device-to-device copies over NVLink, in
// source-rank order so that per-key arrival order survives) and runs the operator chain on them.
// Two parities of send buffers: a peer may already split batch b+1 while this rank still reads batch b.
extern "C" {
static size_t xsend_col_bytes(const fmq_handle* h) { return h->max_batch * 8; }
int32_t fmq_exchange_alloc(fmq_handle* h, void* handle64) {
    if (!h || !handle64) return FMQ_ERR_BAD_ARG;
    CU(h, cudaSetDevice(h->device));
    if (!h->d_xsend) CU(h, cudaMalloc((void**)&h->d_xsend, 6 * xsend_col_bytes(h)));
    cudaIpcMemHandle_t ih;
    CU(h, cudaIpcGetMemHandle(&ih, h->d_xsend));
    memcpy(handle64, &ih, 64);
    return FMQ_OK;
}
void* fmq_exchange_base(fmq_handle* h) { return h ? (void*)h->d_xsend : nullptr; }
int32_t fmq_exchange_attach(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_bases) {
    if (!h || !peer_bases) return FMQ_ERR_BAD_ARG;
    if (world < 1 || world > FMQ_MAX_PEERS || rank < 0 || rank >= world) return set_err(h, FMQ_ERR_BAD_ARG, "need 1 <= world <= 8 and 0 <= rank < world");
    if (!h->d_xsend) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_exchange_alloc first");
    for (int r = 0; r < world; ++r) {
        if (r != rank && !peer_bases[r]) return set_err(h, FMQ_ERR_BAD_ARG, "null peer buffer");
        h->peer_xsend[r] = (r == rank) ? h->d_xsend : static_cast<const char*>(peer_bases[r]);
    }
    h->x_world = world; h->x_rank = rank;
    return FMQ_OK;
}
int32_t fmq_exchange_split(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int32_t parity, uint64_t* counts) {
    if (!h || !counts) return FMQ_ERR_BAD_ARG;
    if (h->x_world < 1) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_exchange_attach first");
    char* base = h->d_xsend + (size_t)(parity & 1) * 3 * xsend_col_bytes(h);
    return fmq_partition_by_owner(h, ts, key, val, n, (uint32_t)h->x_world, (int64_t*)base, (uint64_t*)(base + xsend_col_bytes(h)),
                                  (void*)(base + 2 * xsend_col_bytes(h)), counts);
}
int32_t fmq_exchange_pull_push(fmq_handle* h, int32_t parity, const uint64_t* matrix, int64_t watermark_after, size_t* n_received) {
    if (!h || !matrix) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (h->x_world < 1) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_exchange_attach first");
    const int W = h->x_world, me = h->x_rank;
    size_t n_recv = 0;
    for (int s_ = 0; s_ < W; ++s_) n_recv += matrix[(size_t)s_ * W + me];
    if (n_recv > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "received more events than fmq_config.max_batch");
    CU(h, cudaSetDevice(h->device));
    const size_t col = xsend_col_bytes(h), poff = (size_t)(parity & 1) * 3 * col;
    if (n_received) *n_received = n_recv;
    // (Letting the ingest kernel read the buckets directly out of peer memory instead of copying them was
    // measured slower at N=2: 1.72 vs 1.21 ms per step — 8-byte peer loads are latency-bound.)
    void* dst_cols[3] = {h->d_in_ts, h->d_in_key, h->d_in_val};
    size_t dst_off = 0;
    for (int src = 0; src < W; ++src) {                                   // source-rank order = arrival order per key
        const size_t cnt = matrix[(size_t)src * W + me];
        size_t start = 0;
        for (int d = 0; d < me; ++d) start += matrix[(size_t)src * W + d];    // buckets are laid out by destination
        if (cnt) for (int c = 0; c < 3; ++c)
            CU(h, cudaMemcpyAsync((char*)dst_cols[c] + dst_off * 8, h->peer_xsend[src] + poff + c * col + start * 8, cnt * 8, cudaMemcpyDefault, h->stream));
        dst_off += cnt;
    }
    return fmq_push_batch(h, (const int64_t*)h->d_in_ts, (const uint64_t*)h->d_in_key, h->d_in_val, n_recv, watermark_after, FMQ_MEM_DEVICE);
}
}  // extern "C"

// ---- keyBy exchange on the device (fmq_shard_*; protocol in fmq_shard.cuh) ---------------------------------------
static u64 peer_timeout_ns() {
    static const u64 v = [] { const char* e = getenv("FMQ_PEER_TIMEOUT_MS"); const long ms = e ? atol(e) : 20000; return (u64)(ms < 1 ? 1 : ms) * 1000000ull; }();
    return v;
}
// How the SMs are shared between the scatter of batch b+1 and the ingest of batch b (measurement knobs):
// scatter blocks per SM (0 = one block per chunk, i.e. whatever fits) and ingest blocks per SM while a rank of the exchange.
static u32 xs_blocks_per_sm() {
    static const u32 v = [] { const char* e = getenv("FMQ_XS_BLOCKS_PER_SM"); const int b = e ? atoi(e) : 2; return (u32)(b < 0 ? 0 : (b > 4 ? 4 : b)); }();
    return v;
}
static u32 cmb_ingest_blocks() {
    static const u32 v = [] { const char* e = getenv("FMQ_CMB_INGEST_BLOCKS"); const int b = e ? atoi(e) : FMQ_IW_BLOCKS; return (u32)(b < 1 ? 1 : (b > FMQ_IW_BLOCKS ? FMQ_IW_BLOCKS : b)); }();
    return v;
}
static u32 xs_ingest_blocks() {
    static const u32 v = [] { const char* e = getenv("FMQ_XS_INGEST_BLOCKS"); const int b = e ? atoi(e) : 2; return (u32)(b < 1 ? 1 : (b > FMQ_IW_BLOCKS ? FMQ_IW_BLOCKS : b)); }();
    return v;
}
static u64* xs_segment(char* base, int world, size_t seg_cap, u32 parity, int src) {
    return reinterpret_cast<u64*>(base + XCTL_BYTES) + ((size_t)parity * world + src) * 3 * seg_cap;
}
static XPeers xs_peers(const fmq_handle* h) { XPeers p{}; for (int r = 0; r < h->xs_world; ++r) p.ctl[r] = h->xs_ctl[r]; return p; }

extern "C" {
int32_t fmq_shard_alloc(fmq_handle* h, int32_t world, size_t max_send, void* handle64) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (world < 1 || world > FMQ_MAX_PEERS) return set_err(h, FMQ_ERR_BAD_ARG, "need 1 <= world <= 8");
    if (max_send == 0 || max_send >= 0xFFFFFFF0ull) return set_err(h, FMQ_ERR_BAD_ARG, "max_send must be in [1, 2^32)");
    if (h->d_xbuf) return set_err(h, FMQ_ERR_BAD_ARG, "the exchange buffer of this handle exists already");
    CU(h, cudaSetDevice(h->device));
    h->xs_seg_cap = (max_send + 3) & ~(size_t)3;                              // whole 32-byte groups: segments stay aligned for the bulk copies
    h->xs_world = world;
    const size_t bytes = XCTL_BYTES + (size_t)2 * world * 3 * h->xs_seg_cap * 8;
    CU(h, cudaMalloc((void**)&h->d_xbuf, bytes));
    CU(h, cudaMemset(h->d_xbuf, 0, XCTL_BYTES));
    const size_t nc = (size_t)8 * xs_num_chunks(h->xs_seg_cap);
    CU(h, dalloc(&h->d_xcounts, nc)); CU(h, dalloc(&h->d_xtiles, (nc + SCAN_TILE - 1) / SCAN_TILE + 1));
    CU(h, dalloc(&h->d_send_max, 1)); CU(h, dalloc(&h->d_seg_n, FMQ_MAX_PEERS + 1));
    CU(h, dalloc(&h->d_xfill, FMQ_MAX_PEERS)); CU(h, cudaMemset(h->d_xfill, 0, FMQ_MAX_PEERS * 8));
    const i64 imin = I64_MIN_;
    CU(h, cudaMemcpy(h->d_send_max, &imin, 8, cudaMemcpyHostToDevice));
    CU(h, cudaMemset(h->d_seg_n, 0, (FMQ_MAX_PEERS + 1) * 8));
    CU(h, cudaMallocHost((void**)&h->h_ntotal, 8));
    CU(h, cudaStreamCreateWithFlags(&h->send_stream, cudaStreamNonBlocking));
    if (handle64) { cudaIpcMemHandle_t ih; CU(h, cudaIpcGetMemHandle(&ih, h->d_xbuf)); memcpy(handle64, &ih, 64); }
    CU(h, cudaDeviceSynchronize());
    return FMQ_OK;
}
void* fmq_shard_base(fmq_handle* h) { return h ? (void*)h->d_xbuf : nullptr; }

static int shard_attach_common(fmq_handle* h, int32_t rank, int32_t world) {
    if (!h->d_xbuf) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_shard_alloc first");
    if (world != h->xs_world || rank < 0 || rank >= world) return set_err(h, FMQ_ERR_BAD_ARG, "world must equal fmq_shard_alloc's and 0 <= rank < world");
    if (h->xs_attached) return set_err(h, FMQ_ERR_BAD_ARG, "peers are attached already");
    return FMQ_OK;
}
int32_t fmq_shard_attach(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_bases) {
    if (!h || !peer_bases) return FMQ_ERR_BAD_ARG;
    int s = shard_attach_common(h, rank, world); if (s) return s;
    CU(h, cudaSetDevice(h->device));
    for (int r = 0; r < world; ++r) {
        char* b = (r == rank) ? h->d_xbuf : (char*)peer_bases[r];
        if (!b) return set_err(h, FMQ_ERR_BAD_ARG, "null peer buffer");
        if (r != rank) {                                                      // a buffer on another GPU of this process: make it reachable
            cudaPointerAttributes at{};
            if (cudaPointerGetAttributes(&at, b) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != h->device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { h->err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return FMQ_ERR_CUDA; }
            }
            cudaGetLastError();
        }
        h->xs_base[r] = b; h->xs_ctl[r] = reinterpret_cast<XCtl*>(b);
    }
    h->xs_rank = rank; h->xs_attached = true;
    return FMQ_OK;
}
int32_t fmq_shard_attach_ipc(fmq_handle* h, int32_t rank, int32_t world, const void* blobs) {
    if (!h || !blobs) return FMQ_ERR_BAD_ARG;
    int s = shard_attach_common(h, rank, world); if (s) return s;
    CU(h, cudaSetDevice(h->device));
    const void* bases[FMQ_MAX_PEERS] = {};
    for (int r = 0; r < world; ++r) {
        if (r == rank) { bases[r] = h->d_xbuf; continue; }
        cudaIpcMemHandle_t ih; memcpy(&ih, (const char*)blobs + (size_t)r * 64, 64);
        void* p = nullptr;
        CU(h, cudaIpcOpenMemHandle(&p, ih, cudaIpcMemLazyEnablePeerAccess));
        h->xs_opened[r] = p; bases[r] = p;
    }
    return fmq_shard_attach(h, rank, world, bases);
}

int32_t fmq_shard_send(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int32_t mem) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (!h->xs_attached) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_shard_attach first");
    if (n > h->xs_seg_cap) return set_err(h, FMQ_ERR_BAD_ARG, "batch larger than fmq_shard_alloc's max_send");
    if (n > 0 && (!ts || !key || !val)) return set_err(h, FMQ_ERR_BAD_ARG, "null column pointer");
    if (mem != FMQ_MEM_HOST && mem != FMQ_MEM_DEVICE) return set_err(h, FMQ_ERR_BAD_ARG, "bad mem");
    CU(h, cudaSetDevice(h->device));
    cudaStream_t ss = h->send_stream;
    const u64 seq = ++h->xs_send_seq; const u32 parity = (u32)(seq & 1);
    const int W = h->xs_world, me = h->xs_rank;
    const u64 *dts = (const u64*)ts, *dkey = (const u64*)key, *dval = (const u64*)val;
    if (mem == FMQ_MEM_HOST && n > 0) {
        if (!h->xs_stage) CU(h, dalloc(&h->xs_stage, 3 * h->xs_seg_cap));
        // the previous send's kernels have read the staging columns before these copies overwrite them (same stream)
        CU(h, cudaMemcpyAsync(h->xs_stage, ts, n * 8, cudaMemcpyHostToDevice, ss));
        CU(h, cudaMemcpyAsync(h->xs_stage + h->xs_seg_cap, key, n * 8, cudaMemcpyHostToDevice, ss));
        CU(h, cudaMemcpyAsync(h->xs_stage + 2 * h->xs_seg_cap, val, n * 8, cudaMemcpyHostToDevice, ss));
        dts = h->xs_stage; dkey = h->xs_stage + h->xs_seg_cap; dval = h->xs_stage + 2 * h->xs_seg_cap;
    }
    // the peers must have consumed batch seq-2, the last one scattered into the segments of this parity
    xs_wait_acks_kernel<<<1, 32, 0, ss>>>(h->xs_ctl[me], (u32)W, parity, seq >= 3 ? seq - 2 : 0, h->d_st, peer_timeout_ns()); LAUNCHED(h, "xs_wait_acks_kernel");
    const u32 nchunks = xs_num_chunks(n);
    if (n > 0) {
        SplitDst dst{}; dst.col_stride = h->xs_seg_cap;
        for (int d = 0; d < W; ++d) dst.ts[d] = xs_segment(h->xs_base[d], W, h->xs_seg_cap, parity, me);
        // Order-dependent keyed state needs every key's events in arrival order: stable two-pass split.  Window aggregates
        // and the filter are commutative: one pass, runs placed in completion order.  Either way the scatter takes at
        // most xs_blocks_per_sm blocks per SM, so that the ingest of the previous batch keeps running beside it.
        unsigned long long* fill = (keyed_enabled(h->cfg) || h->ml_cap) ? nullptr : h->d_xfill;
        cudaError_t e = owner_split(ss, SplitScratch{h->d_xcounts, h->d_xtiles}, dts, dkey, dval, n, (u32)W, h->cfg.key_mode == FMQ_KEYS_DENSE ? 1u : 0u, dst, true, h->d_send_max,
                                    fill, (u32)h->num_sms * xs_blocks_per_sm());
        if (e != cudaSuccess) { h->err = std::string("owner split: ") + cudaGetErrorString(e); return FMQ_ERR_CUDA; }
        h->launches += fill ? 1 : 5;
    }
    xs_publish_kernel<<<1, 32, 0, ss>>>(h->d_xcounts, nchunks, (u64)n, (n > 0 && !(keyed_enabled(h->cfg) || h->ml_cap)) ? h->d_xfill : nullptr, (u32)W, (u32)me, parity, seq,
                                        h->d_send_max, xs_peers(h)); LAUNCHED(h, "xs_publish_kernel");
    if (mem == FMQ_MEM_HOST && n > 0) CU(h, cudaStreamSynchronize(ss));       // the caller owns host columns only until this call returns
    return FMQ_OK;
}

}  // extern "C"

template <typename V>
static int shard_recv_t(fmq_handle* h, int64_t wm_after, size_t* n_received) {
    const fmq_config& c = h->cfg;
    const u64 seq = ++h->xs_recv_seq; const u32 parity = (u32)(seq & 1);
    const int W = h->xs_world, me = h->xs_rank;
    cudaEvent_t* tev = h->tev[h->pushes % fmq_handle::TIMING_RING];
    CU(h, cudaEventRecord(tev[0], h->stream));
    xs_gather_kernel<<<1, 32, 0, h->stream>>>(h->xs_ctl[me], (u32)W, parity, seq, h->d_seg_n, h->d_seg_n + FMQ_MAX_PEERS, h->d_st, peer_timeout_ns()); LAUNCHED(h, "xs_gather_kernel");
    u64* seg0 = xs_segment(h->d_xbuf, W, h->xs_seg_cap, parity, 0);
    SegIn seg; seg.nseg = (u32)W; seg.stride = 3 * h->xs_seg_cap; seg.seg_n = h->d_seg_n;
    int s;
    if (keyed_enabled(c) || n_received) {
        // The keyed chain sizes its launches on the host: the one number it needs, the batch's event count, is read back.
        CU(h, cudaMemcpyAsync(h->h_ntotal, h->d_seg_n + FMQ_MAX_PEERS, 8, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        if (n_received) *n_received = (size_t)*h->h_ntotal;
    }
    if (keyed_enabled(c)) {
        const size_t n = (size_t)*h->h_ntotal;
        if (n > h->max_batch) return set_err(h, FMQ_ERR_BAD_ARG, "received more events than fmq_config.max_batch");
        if (n > 0) {
            xs_compact_kernel<<<(unsigned)h->num_sms * 4, 256, 0, h->stream>>>(seg0, seg.stride, h->xs_seg_cap, h->d_seg_n, (u32)W, (u64*)h->d_in_ts, h->d_in_key, h->d_in_val, h->max_batch);
            LAUNCHED(h, "xs_compact_kernel");
        }
        s = stage_ingest<V>(h, h->d_in_ts, h->d_in_key, h->d_in_val, n); if (s) return s;
        s = stage_fire<V>(h, n, wm_after, true); if (s) return s;
    } else {
        s = stage_ingest<V>(h, (const i64*)seg0, seg0 + h->xs_seg_cap, seg0 + 2 * h->xs_seg_cap, h->xs_seg_cap, seg); if (s) return s;
        s = stage_fire<V>(h, 0, wm_after, true, seg, h->xs_seg_cap); if (s) return s;
    }
    xs_ack_kernel<<<1, 32, 0, h->stream>>>(xs_peers(h), (u32)W, (u32)me, parity, seq); LAUNCHED(h, "xs_ack_kernel");
    CU(h, cudaEventRecord(tev[3], h->stream));
    h->pushes++;
    return FMQ_OK;
}
extern "C" {
int32_t fmq_shard_recv(fmq_handle* h, int64_t watermark_after, size_t* n_received) {
    if (!h) return FMQ_ERR_BAD_ARG;
    if (h->sticky) return FMQ_ERR_POISONED;
    if (!h->xs_attached) return set_err(h, FMQ_ERR_BAD_ARG, "call fmq_shard_attach first");
    if (h->world > 1) return set_err(h, FMQ_ERR_BAD_ARG, "combiner peers attached: this handle does not take the exchange route");
    if (h->xs_recv_seq >= h->xs_send_seq) return set_err(h, FMQ_ERR_BAD_ARG, "fmq_shard_recv without a matching fmq_shard_send on this rank");
    CU(h, cudaSetDevice(h->device));
    return h->cfg.value_type == FMQ_VAL_F64 ? shard_recv_t<double>(h, watermark_after, n_received) : shard_recv_t<i64>(h, watermark_after, n_received);
}
}  // extern "C"

// ---- owner split for callers that must not allocate ------------------------------------------------
// Same stable multi-split as fmq_partition_by_owner, but its scratch (the per-chunk digit counters
// and the scan's tile sums) lives in a fixed file-scope device array instead of a handle's
// cudaMalloc'ed buffers, so the call performs no allocation of any kind.  See include/fmq_b200.h.
namespace {
constexpr size_t NOALLOC_MAX_CHUNKS = FMQ_NOALLOC_MAX_EVENTS / XS_CHUNK;                 // 2048
constexpr size_t NOALLOC_COUNTS = (size_t)RS_DIGITS * NOALLOC_MAX_CHUNKS;                 // 524288 counters
constexpr size_t NOALLOC_TILES = (NOALLOC_COUNTS + SCAN_TILE - 1) / SCAN_TILE;            // 128 tile sums
// ~2 MB per GPU that loads this library, held until the process exits.
static __device__ __align__(256) unsigned int g_noalloc_scratch[NOALLOC_COUNTS + NOALLOC_TILES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
u32* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};   // device address of g_noalloc_scratch, per device

// In host code `g_noalloc_scratch` is only a handle for the symbol, never a device address: the
// address each device gave its copy is asked for once, with that device current.
int noalloc_scratch_for(int device, u32** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            g_create_err = std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e);
            return FMQ_ERR_CUDA;
        }
        g_noalloc_ptr[device] = static_cast<u32*>(p);
    }
    *out = g_noalloc_ptr[device];
    return FMQ_OK;
}
#define NA_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { g_create_err = std::string(#call) + ": " + cudaGetErrorString(e_); return FMQ_ERR_CUDA; } } while (0)
}  // namespace

extern "C" int32_t fmq_partition_by_owner_noalloc(int32_t device, void* stream_, const int64_t* ts, const uint64_t* key, const void* val,
                                                  size_t n, uint32_t world, int64_t* ts_out, uint64_t* key_out, void* val_out,
                                                  uint64_t* counts) {
    if (!counts || world == 0 || world > RS_DIGITS) { g_create_err = "need counts != NULL and 1 <= world <= 256"; return FMQ_ERR_BAD_ARG; }
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) { g_create_err = "device ordinal out of range"; return FMQ_ERR_BAD_ARG; }
    if (n > FMQ_NOALLOC_MAX_EVENTS) {
        g_create_err = "fmq_partition_by_owner_noalloc: n exceeds FMQ_NOALLOC_MAX_EVENTS (" FMQ_STR(FMQ_NOALLOC_MAX_EVENTS)
                       " events, the size of the built-in scratch); split the batch or use fmq_partition_by_owner";
        return FMQ_ERR_BAD_ARG;
    }
    for (uint32_t w = 0; w < world; ++w) counts[w] = 0;
    if (n == 0) return FMQ_OK;
    if (!ts || !key || !val || !ts_out || !key_out || !val_out) { g_create_err = "null column pointer"; return FMQ_ERR_BAD_ARG; }
    // the caller's current device is put back before returning (success or failure)
    struct DeviceGuard { int prev = -1; ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); } } guard;
    int cur = -1;
    NA_CU(cudaGetDevice(&cur));
    if (cur != device) { NA_CU(cudaSetDevice(device)); guard.prev = cur; }
    u32* scratch = nullptr;
    int st = noalloc_scratch_for(device, &scratch); if (st) return st;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    u32* d_counts = scratch; u32* d_tiles = scratch + NOALLOC_COUNTS;
    // Nothing in the scratch has to start at zero: the counting kernel stores (not adds) every one of the
    // digits x nchunks counters it owns (8 digits up to world = 8, 256 beyond) and the scan stores every tile
    // sum before reading it, so whatever the previous call left behind is overwritten before it can be read.
    const u32 nchunks = xs_num_chunks(n);
    SplitDst dst{}; dst.out_ts = (u64*)ts_out; dst.out_key = (u64*)key_out; dst.out_val = (u64*)val_out;
    NA_CU(owner_split(stream, SplitScratch{d_counts, d_tiles}, (const u64*)ts, (const u64*)key, (const u64*)val, n, world, 0u, dst, false, nullptr));
    u32 starts[RS_DIGITS];
    NA_CU(cudaMemcpy2DAsync(starts, 4, d_counts, (size_t)nchunks * 4, 4, world, cudaMemcpyDeviceToHost, stream));
    NA_CU(cudaStreamSynchronize(stream));
    for (uint32_t w = 0; w < world; ++w) counts[w] = (w + 1 < world ? starts[w + 1] : (u32)n) - starts[w];
    return FMQ_OK;
}