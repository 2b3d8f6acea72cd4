/*
 * fmq_b200.h — C ABI of the B200-native keyed window-aggregate + alert engine.
 *
 * This is the drop-in boundary for the ONE hot path of the Flink monitoring quickstart
 * (reference = /root/reference, Jax-Rene/monitor-systam-flink-quickstart): what a Flink
 * `OneInputStreamOperator` / `KeyedProcessFunction` shim (JNI / Panama / ctypes / C++) binds.
 * Plain C: pointers, sizes, POD structs.  No C++ types, no torch types, no exceptions cross it.
 *
 * Data model: one micro-batch = n events in three columns
 *     ts[n]  : int64   event time, epoch milliseconds, ts >= 0
 *     key[n] : uint64  dictionary-encoded key (host, channel, or host<<32|metric);
 *                      0xFFFFFFFFFFFFFFFF is reserved
 *     val[n] : float64 (FMQ_VAL_F64) or int64 (FMQ_VAL_I64)
 * Watermarks arrive only at micro-batch boundaries: every record of batch b is judged against
 * the watermark in force after batch b-1 (deterministic replacement of Flink's 200 ms
 * wall-clock watermark timer, chapter3/README.md:379-386).
 *
 * Threading contract (inherited from Flink's operator model): one handle per operator
 * subtask, calls on one handle are serialised by the caller; different handles may be used
 * concurrently from different threads.  Each handle owns one CUDA stream and all of its
 * device state.
 *
 * All functions return an fmq_status (0 = OK).  A human-readable message for the last
 * failure on a handle is available from fmq_last_error().
 */
#ifndef FMQ_B200_H
#define FMQ_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FMQ_ABI_VERSION 1

typedef enum fmq_status {
    FMQ_OK = 0,
    FMQ_ERR_BAD_ARG = 1,        /* null pointer, bad enum, size not a multiple of slide ...   */
    FMQ_ERR_TABLE_FULL = 2,     /* key directory is full (raise table_capacity)               */
    FMQ_ERR_RING_OVERFLOW = 3,  /* open event-time span exceeds pane_ring panes               */
    FMQ_ERR_OUTPUT_OVERFLOW = 4,/* an output buffer filled up before it was polled            */
    FMQ_ERR_CUDA = 5,           /* a CUDA call failed; see fmq_last_error                     */
    FMQ_ERR_BAD_TS = 6,         /* negative timestamp or reserved key seen                    */
    FMQ_ERR_NO_DEVICE = 7,      /* no usable CUDA device: there is NO CPU fallback            */
    FMQ_ERR_POISONED = 8,       /* a previous sticky error invalidated this handle's state    */
    FMQ_ERR_PEER_TIMEOUT = 9    /* a peer rank did not publish / acknowledge a batch in time  */
} fmq_status;

typedef enum fmq_value_type { FMQ_VAL_F64 = 0, FMQ_VAL_I64 = 1 } fmq_value_type;
typedef enum fmq_mem        { FMQ_MEM_HOST = 0, FMQ_MEM_DEVICE = 1 } fmq_mem;
/* How the uint64 key column indexes the device state.
 * HASHED: arbitrary 64-bit keys, found through a bucketized hash directory (one extra memory access
 *         per event).  DENSE: the caller's dictionary hands out ids 0..table_capacity-1 (what a shim
 *         that dictionary-encodes String keys, SURVEY.md §8b, naturally produces) and the id IS the
 *         table slot: no directory, no hashing.  An id >= table_capacity is an input error. */
typedef enum fmq_key_mode   { FMQ_KEYS_HASHED = 0, FMQ_KEYS_DENSE = 1 } fmq_key_mode;
typedef enum fmq_cmp        { FMQ_CMP_GT = 0, FMQ_CMP_LT = 1, FMQ_CMP_GE = 2, FMQ_CMP_LE = 3 } fmq_cmp;

/* aggregate mask bits (count is always maintained: a window fires only if count > 0) */
#define FMQ_AGG_COUNT 1u
#define FMQ_AGG_SUM   2u
#define FMQ_AGG_MIN   4u
#define FMQ_AGG_MAX   8u

/* which fired-window quantity the window alert expression is evaluated on */
typedef enum fmq_alert_source {
    FMQ_ALERT_NONE = 0,
    FMQ_ALERT_SUM = 1,   /* (double)sum * mul / div0 / div1 / div2   BandwidthMonitor.java:39    */
    FMQ_ALERT_AVG = 2,   /* count == 0 ? 0.0 : sum / count           ComputeCpuAvg.java:49       */
    FMQ_ALERT_MIN = 3,
    FMQ_ALERT_MAX = 4,
    FMQ_ALERT_COUNT = 5
} fmq_alert_source;

/* watermark_after value meaning "derive it from the data": wm = max(wm, maxTsSeen - bound)
 * (BoundedOutOfOrdernessTimestampExtractor, chapter3/README.md:379-396). */
#define FMQ_WM_AUTO INT64_MIN
/* watermark_after value meaning "no watermark arrives with this batch": the records are ingested (and
 * judged against the watermark already in force), the max-timestamp tracking of the extractor moves on,
 * but the watermark itself stays where it is and nothing fires.  This is what a shim behind Flink's own
 * watermark assigner passes when its buffer fills up BETWEEN two upstream watermarks: WindowOperator only
 * moves event time in processWatermark ([Flink 1.8.0] WindowOperator#processElement never advances it). */
#define FMQ_WM_KEEP (INT64_MIN + 1)
/* watermark that fires every remaining window (Flink's end-of-input Watermark(Long.MAX_VALUE)) */
#define FMQ_WM_FINAL INT64_MAX

/*
 * Job description.  One POD replaces the reference's hard-coded DSL chain
 *   keyBy(k).timeWindow(size[,slide]).reduce/aggregate(...).map(...).filter(...)
 * (chapter3/.../BandwidthMonitorWithEventTime.java:45-55, chapter2/.../ComputeCpuAvg.java:27-31),
 * the un-keyed filter (chapter1/.../Main.java:27-32) and the keyed ValueState operators
 * (chapter2/.../ComputeCpuMax.java:26 and the consecutive-breach rule of BASELINE.json).
 * Zero-initialise, set struct_size = sizeof(fmq_config), then fill in what the job uses.
 */
typedef struct fmq_config {
    uint32_t struct_size;
    int32_t  device;                 /* CUDA device ordinal                                      */
    int32_t  value_type;             /* fmq_value_type                                           */
    int32_t  key_mode;               /* fmq_key_mode                                             */

    /* --- event-time window aggregation (A2, A3, A5, A5f) --- */
    int32_t  window_enabled;
    uint32_t agg_mask;               /* FMQ_AGG_*                                                */
    int64_t  window_size_ms;         /* Time.minutes(5) -> 300000                                */
    int64_t  window_slide_ms;        /* == size for tumbling; size % slide must be 0             */
    int64_t  window_offset_ms;       /* 0 <= offset < slide; timeWindow(...) uses 0              */
    int64_t  max_out_of_orderness_ms;/* bound of BoundedOutOfOrdernessTimestampExtractor         */
    int32_t  emit_fired;             /* 1: every fired (key, window) row is emitted              */

    /* --- alert on fired windows (A6 ii/iii) --- */
    int32_t  walert_source;          /* fmq_alert_source                                         */
    int32_t  walert_op;              /* fmq_cmp                                                  */
    int32_t  reserved1;
    double   walert_mul;             /* 8.0                                                      */
    double   walert_div0;            /* 60     (use 1.0 for "no division")                       */
    double   walert_div1;            /* 1024                                                     */
    double   walert_div2;            /* 1024                                                     */
    double   walert_threshold;       /* 100.0                                                    */

    /* --- per-event threshold filter (A6 i, chapter1 Main.java:31) --- */
    int32_t  efilter_enabled;
    int32_t  efilter_op;             /* fmq_cmp; the reference uses strict >                     */
    double   efilter_threshold;      /* 90                                                       */

    /* --- keyed ValueState operators (A7), per-key arrival order --- */
    int32_t  rolling_max_enabled;    /* keyBy(0).max(2): one output per input                    */
    int32_t  consec_n;               /* 0 = off; N consecutive over-threshold readings           */
    int32_t  consec_op;              /* fmq_cmp, strict > by default                             */
    int32_t  consec_reset_on_fire;   /* 1: counter returns to 0 after an alert                   */
    double   consec_threshold;

    /* --- capacities (device memory is allocated once, in fmq_create; only the two staging sets
     *     for host-resident input columns wait for the first host push) --- */
    uint64_t table_capacity;         /* key slots (any size; ~1.5x the key count when HASHED,
                                        max id + 1 when DENSE)                                   */
    uint32_t pane_ring;              /* open panes kept per key; rounded up to a power of two    */
    uint32_t reserved2;
    uint64_t max_batch;              /* largest n accepted by fmq_push_batch                     */
    uint64_t fired_capacity;         /* fired rows buffered between polls                        */
    uint64_t walert_capacity;        /* window-alert rows buffered between polls                 */
    uint64_t ealert_capacity;        /* per-event filter rows buffered between polls             */
    uint64_t calert_capacity;        /* consecutive-breach alert rows buffered between polls     */
    uint64_t text_max_bytes;         /* > 0: device buffers for fmq_parse_text / fmq_push_text    */
    uint64_t median_log_capacity;    /* > 0: also compute the per-(key, window) median
                                        (ComputeCpuMiddle.java:34-47); this many records can be
                                        buffered while their windows are open (f64 only).  Size it
                                        for every record pushed in that time, not only the accepted
                                        ones: log space is reserved ahead, and what a dropped (late)
                                        record would have used stays empty until the next expiry   */
    uint64_t median_capacity;        /* median rows buffered between polls                        */
    uint64_t late_capacity;          /* > 0: dropped late records are kept as a side output
                                        (sideOutputLateData), this many rows between polls        */
    int32_t  overlapped_polls;       /* 1: second set of output columns for fmq_poll_all_begin/end */
    int32_t  reserved3;
    uint64_t text_dict_capacity;     /* > 0: fmq_parse_text / fmq_push_text dictionary-encode the host / channel strings on the
                                        device: the key column carries dense ids 0, 1, 2, ... in order of first sight (what
                                        FMQ_KEYS_DENSE wants), this many distinct strings at most; names come back through
                                        fmq_dict_lookup.  0: the key column carries fmq_text_key_hash(string)           */
    uint64_t text_dict_arena_bytes;  /* bytes kept for the strings themselves (default 32 per entry)                   */
    int64_t  allowed_lateness_ms;    /* .allowedLateness(Time) — option 2 of the lateness handling the tutorial lists
                                        (chapter3/README.md:211-212): a window's state is kept until
                                        window_end - 1 + T <= watermark; a record that arrives for a window that has
                                        fired but is still kept is added to it and the window fires AGAIN.  Flink
                                        re-fires once per such record; here watermarks and outputs exist at micro-batch
                                        boundaries only, so a (key, window) that took late records in a batch re-fires
                                        ONCE at the end of that batch, with everything it holds by then.  Records
                                        older than that are dropped / side-output as before.  0 = Flink's default.   */
} fmq_config;

/* one fired (key, window): 48 B per row */
typedef struct fmq_fired_cols {
    uint64_t* key;
    int64_t*  window_end;            /* exclusive end; Flink's record timestamp is window_end-1  */
    int64_t*  count;
    void*     sum;                   /* double* or int64_t* according to value_type              */
    void*     min;
    void*     max;
} fmq_fired_cols;

/* one window alert: the fired rows that pass the alert expression, 24 B per row */
typedef struct fmq_walert_cols {
    uint64_t* key;
    int64_t*  window_end;
    double*   value;                 /* the evaluated expression, e.g. Mbps                      */
} fmq_walert_cols;

/* one event row (per-event filter output, consecutive-breach alert): 24 B per row */
typedef struct fmq_event_cols {
    int64_t*  ts;
    uint64_t* key;
    void*     val;                   /* double* or int64_t*                                      */
} fmq_event_cols;

typedef struct fmq_counters {
    uint64_t events_in;
    uint64_t late_dropped;           /* WindowOperator.numLateRecordsDropped                     */
    uint64_t windows_fired;          /* fired (key, window) pairs                                */
    uint64_t window_alerts;
    uint64_t event_alerts;
    uint64_t consec_alerts;
    uint64_t keys;                   /* distinct keys in the directory (FMQ_KEYS_HASHED and keyed
                                        state only; a FMQ_KEYS_DENSE handle has no directory: 0) */
    int64_t  watermark;              /* current watermark                                        */
    uint64_t gpu_launches;           /* kernels launched by this handle so far                   */
} fmq_counters;

typedef struct fmq_handle fmq_handle;

/* Library / build identification; usable without a GPU. */
int32_t     fmq_abi_version(void);
const char* fmq_build_info(void);

/* Replaces: operator construction + open() (OneInputStreamOperator#open) for the DSL chain at
 * BandwidthMonitorWithEventTime.java:45-55 / ComputeCpuAvg.java:27-31 / ComputeCpuMax.java:26 /
 * Main.java:27-32.  Allocates every device buffer the handle will ever use. */
int32_t fmq_create(const fmq_config* cfg, fmq_handle** out);

/* Replaces: a run of WindowOperator#processElement / StreamFilter#processElement /
 * StreamGroupedReduce#processElement calls followed by one processWatermark(watermark_after)
 * (call stack SURVEY.md §3.1).  `mem` says where ts/key/val live (host or device pointers).
 * Asynchronous on the handle's stream; results become visible through the poll functions. */
int32_t fmq_push_batch(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val,
                       size_t n, int64_t watermark_after, int32_t mem);

/* Same call as fmq_push_batch with a narrower wire format for the columns, for shims whose batches
 * come from host memory: the host path is bound by PCIe (24 B per event at ~50 GB/s), so a shim that
 * already maps host names to small ids and whose micro-batch spans less than 49 days can send
 * 4-byte offsets and ids.  The columns are widened on the device (exactly: zero/sign extension and
 * one integer add) before the same kernels run; results are identical to fmq_push_batch on the
 * widened columns.  The record fields it carries are the reference's (timestamp, key, value) tuple
 * (chapter3/.../BandwidthMonitorWithEventTime.java:39-44); the encoding itself has no reference
 * counterpart. */
typedef struct fmq_packed_columns {
    uint32_t    struct_size;         /* = sizeof(fmq_packed_columns)                             */
    uint32_t    ts_width;            /* 8: int64 epoch ms;  4: uint32 ms offset from ts_base     */
    uint32_t    key_width;           /* 8: uint64 key id;   4: uint32 key id, zero-extended      */
    uint32_t    val_width;           /* 8: double / int64 (fmq_config.value_type);
                                        4: int32, sign-extended — FMQ_VAL_I64 handles only       */
    int64_t     ts_base;             /* added to every 4-byte timestamp offset                   */
    const void* ts;
    const void* key;
    const void* val;
} fmq_packed_columns;
int32_t fmq_push_packed(fmq_handle* h, const fmq_packed_columns* cols, size_t n,
                        int64_t watermark_after, int32_t mem);

/* Starts the host->device copy of the NEXT micro-batch and returns at once, so that the copy runs
 * while the kernels of the batch just pushed finish and its results are polled (a shim that fills
 * buffer k+1 while buffer k is in flight calls this as soon as k+1 is full).  `cols` describes host
 * columns (pinned memory for a truly asynchronous copy; all widths 8 = the fmq_push_batch format).
 * The columns must stay untouched until the consuming call returns: the next fmq_push_batch /
 * fmq_push_packed with FMQ_MEM_HOST must pass the same pointers, widths, ts_base and n, and finds
 * them staged.  Up to two batches can be staged ahead (first in, first out), which keeps the copy engine
 * busy across the host's own latency between calls.  Network-buffer hand-over in Flink
 * (RecordWriter -> InputGate) is the nearest reference notion; it has no direct counterpart. */
int32_t fmq_stage_host(fmq_handle* h, const fmq_packed_columns* cols, size_t n);

/* Replaces: processWatermark(wm) with no records (and Watermark(Long.MAX_VALUE) at end of a
 * finite source when wm == FMQ_WM_FINAL). */
int32_t fmq_flush(fmq_handle* h, int64_t watermark);

/* Replaces: Output#collect of fired window results (WindowOperator#emitWindowContents).
 * Copies up to `cap` rows into caller buffers (`mem` says host or device); *n_out = rows copied,
 * *n_left = rows still buffered.  Row order is unspecified (compare as sets). */
int32_t fmq_poll_fired(fmq_handle* h, const fmq_fired_cols* out, size_t cap, int32_t mem,
                       size_t* n_out, size_t* n_left);
/* Replaces: the chained map + filter after the window (BandwidthMonitorWithEventTime.java:48-55). */
int32_t fmq_poll_window_alerts(fmq_handle* h, const fmq_walert_cols* out, size_t cap, int32_t mem,
                               size_t* n_out, size_t* n_left);
/* Replaces: Output#collect of everything a micro-batch produced (fired rows, window alerts, per-event
 * filter rows), for shims that overlap PCIe directions: fmq_poll_all_begin waits for the kernels,
 * returns the three row counts, queues the device->host copies on a second stream and returns;
 * the caller may push the next micro-batch (host->device) while they run, and calls
 * fmq_poll_all_end before reading the buffers.  Host buffers (pinned) only; every buffer must hold
 * all waiting rows.  Needs fmq_config.overlapped_polls = 1. */
int32_t fmq_poll_all_begin(fmq_handle* h, const fmq_fired_cols* fired, size_t fired_cap,
                           const fmq_walert_cols* walerts, size_t walert_cap,
                           const fmq_event_cols* ealerts, size_t ealert_cap, size_t* counts3);
int32_t fmq_poll_all_end(fmq_handle* h);
/* Replaces: DataStream#getSideOutput(lateOutputTag) after .sideOutputLateData(tag) — option 3 of the
 * lateness handling the tutorial lists (chapter3/README.md:209-228): the records the window operator
 * drops as late (every window they belong to has fired), as (ts, key, val) rows.  Needs
 * fmq_config.late_capacity > 0.  late_dropped counts them either way. */
int32_t fmq_poll_late(fmq_handle* h, const fmq_event_cols* out, size_t cap, int32_t mem,
                      size_t* n_out, size_t* n_left);
/* Replaces: the ProcessWindowFunction output of ComputeCpuMiddle.java:36-47: one row
 * (key, window_end, median) per fired (key, window); `value` is the middle element of the window's
 * sorted values, or the mean of the two middle ones.  Needs fmq_config.median_log_capacity > 0. */
int32_t fmq_poll_medians(fmq_handle* h, const fmq_walert_cols* out, size_t cap, int32_t mem,
                         size_t* n_out, size_t* n_left);
/* Replaces: StreamFilter output of Main.java:27-32 (row order unspecified, compare as sets). */
int32_t fmq_poll_event_alerts(fmq_handle* h, const fmq_event_cols* out, size_t cap, int32_t mem,
                              size_t* n_out, size_t* n_left);
/* Replaces: KeyedProcessFunction output of the consecutive-breach rule. */
int32_t fmq_poll_consec_alerts(fmq_handle* h, const fmq_event_cols* out, size_t cap, int32_t mem,
                               size_t* n_out, size_t* n_left);
/* Replaces: StreamGroupedReduce output of keyBy(0).max(2) (ComputeCpuMax.java:26): the rolling
 * per-key maximum after each record of the LAST pushed batch, in that batch's input order. */
int32_t fmq_poll_rolling_max(fmq_handle* h, void* out, size_t cap, int32_t mem, size_t* n_out);

/* Replaces: Flink metric group of the operator (numLateRecordsDropped etc.).  Synchronises. */
int32_t fmq_get_counters(fmq_handle* h, fmq_counters* out);

/* Keyed-state inspection for parity tests ("state transitions" in BASELINE.json): looks up
 * n keys; found[i] = 0/1, consec[i] = breach counter, rolling[i] = rolling max bits. */
int32_t fmq_get_key_state(fmq_handle* h, const uint64_t* keys, size_t n, int32_t* found,
                          int32_t* consec, void* rolling);

/* Text ingest on the device.  Replaces: the `map` functions in front of every job —
 * FMQ_TEXT_CPU  "<epoch-s> <host> <cpu> <usage>"     Main.java:21-25, ComputeCpuAvg.java:22-25
 *               -> ts = epoch-s * 1000, key = fmq_text_key_hash(host), val = (double) usage
 * FMQ_TEXT_FLOW "<ISO local time> <channel> <flow>"  BandwidthMonitorWithEventTime.java:33-34,39-44
 *               -> ts = (int) epochSecond(UTC+utc_offset_hours) * 1000 (the reference's 32-bit cast),
 *                  key = fmq_text_key_hash(channel), val = (int64) flow
 * Lines are '\n'-separated (a final unterminated line counts); fields are split on single spaces like
 * String.split(" ").  Doubles are parsed on the device only where that is exact (<= 15 significant
 * digits, |decimal exponent| <= 22: one correctly rounded IEEE operation, identical to
 * Double.parseDouble); any other value is flagged FMQ_TEXT_VALUE_ON_HOST for the caller to parse.
 * fmq_parse_text writes the columns (and a per-line status) to caller buffers of `cap` lines.
 * fmq_push_text parses and feeds the result straight into the operator chain without leaving the
 * device; it refuses the batch (FMQ_ERR_BAD_ARG) if any line is not FMQ_TEXT_OK.
 * Needs fmq_config.text_max_bytes >= bytes and max_batch >= number of lines. */
#define FMQ_TEXT_CPU 0
#define FMQ_TEXT_FLOW 1
#define FMQ_TEXT_OK 0u
#define FMQ_TEXT_MALFORMED 1u
#define FMQ_TEXT_VALUE_ON_HOST 2u
#define FMQ_TEXT_KEY_ON_HOST 3u      /* dictionary mode: the string shares its 64-bit hash with another one, or the dictionary is full */
uint64_t fmq_text_key_hash(const char* s, size_t len);
int32_t fmq_parse_text(fmq_handle* h, const char* text, size_t bytes, int32_t text_mem, int32_t format,
                       int32_t utc_offset_hours, int64_t* ts_out, uint64_t* key_out, void* val_out,
                       uint32_t* status_out, size_t cap, int32_t out_mem, size_t* n_lines, size_t* n_bad);
int32_t fmq_push_text(fmq_handle* h, const char* text, size_t bytes, int32_t text_mem, int32_t format,
                      int32_t utc_offset_hours, int64_t watermark_after, size_t* n_lines);
/* Replaces: the String the reference prints with every result (`value.f0` / `value.f1` in the sinks, e.g.
 * BandwidthMonitorWithEventTime.java:55) — the names behind dictionary ids (fmq_config.text_dict_capacity > 0).
 * fmq_dict_size: number of ids handed out so far.  fmq_dict_lookup: names of n ids into host rows of `stride` bytes
 * (zero padded, truncated if longer); lens[i] = the name's full length, UINT32_MAX for an id never handed out. */
int32_t fmq_dict_size(fmq_handle* h, size_t* n_ids);
int32_t fmq_dict_lookup(fmq_handle* h, const uint64_t* ids, size_t n, char* names_out, size_t stride, uint32_t* lens_out);

/* Replaces: AbstractStreamOperator#snapshotState / #initializeState (the checkpoint hooks the
 * tutorial's promised-but-unwritten next chapter is about, chapter3/README.md:454-456).  A snapshot is
 * the operator's keyed state only — key directory, open-pane accumulators, per-key ValueState,
 * watermark and counters — copied to a host buffer; rows already emitted must be polled first.
 * fmq_restore loads it into a handle created with the same job geometry. */
int32_t fmq_snapshot_size(fmq_handle* h, size_t* bytes);
int32_t fmq_snapshot(fmq_handle* h, void* host_buf, size_t cap, size_t* written);
int32_t fmq_restore(fmq_handle* h, const void* host_buf, size_t bytes);

/* Blocks until all work queued on the handle has finished. */
int32_t fmq_synchronize(fmq_handle* h);
/* Device time in ms of the last fmq_push_batch / fmq_flush kernel chain (CUDA events on the
 * handle's stream); synchronises. */
int32_t fmq_last_device_ms(fmq_handle* h, float* ms);
/* Measurement hook (no counterpart in the reference): device time of the push `back` calls ago
 * (0 = latest; a ring of 64 is kept), split by stage with CUDA events on the handle's stream:
 * ingest = H2D staging copies (host inputs only) + the window/filter kernel; keyed = ValueState
 * chain; fire = plan + fire + commit.  Synchronises on that push. */
int32_t fmq_get_push_timing(fmq_handle* h, uint32_t back, float* total_ms, float* ingest_ms,
                            float* keyed_ms, float* fire_ms);
/* Raw cudaStream_t of the handle, for callers that enqueue their own device work. */
void*   fmq_stream(fmq_handle* h);

const char* fmq_last_error(fmq_handle* h);   /* h may be NULL: last create() failure */
/* Replaces: OneInputStreamOperator#close / dispose. */
int32_t fmq_destroy(fmq_handle* h);

/* keyBy routing (A4).  fmq_owner_of: the GPU-side routing function owner = mix64(key) % world.
 * fmq_flink_key_group / fmq_flink_subtask: Flink 1.8 KeyGroupRangeAssignment restated
 * (murmurHash(key.hashCode()) % maxParallelism; kg * p / maxParallelism), pinned by golden G8. */
uint32_t fmq_owner_of(uint64_t key, uint32_t world);
int32_t  fmq_java_string_hash(const char* utf8, size_t len);
int32_t  fmq_flink_key_group(int32_t key_hash, int32_t max_parallelism);
int32_t  fmq_flink_subtask(int32_t key_group, int32_t parallelism, int32_t max_parallelism);

/* Multi-GPU keyBy exchange, device side (A4): stable partition of a device-resident batch into
 * `world` contiguous buckets by owner.  HASHED handles route by fmq_owner_of; DENSE handles route
 * by id % world and rewrite the key column to the owner's local id (id / world), so that every
 * rank's table stays dense — the global id of an output row of rank r is local * world + r.  counts[world] is written to HOST memory.
 * All pointers except counts are device pointers; runs on the handle's stream and
 * synchronises before returning (the counts are needed to size the all-to-all). */
int32_t fmq_partition_by_owner(fmq_handle* h, const int64_t* ts, const uint64_t* key,
                               const void* val, size_t n, uint32_t world,
                               int64_t* ts_out, uint64_t* key_out, void* val_out,
                               uint64_t* counts);

/* ---- combiner mode: multi-GPU window aggregation without an event shuffle (A4/A5/A5f) -------------
 * Replaces the keyBy network exchange in front of the WindowOperator
 * (BandwidthMonitorWithEventTime.java:45) for the commutative window aggregates: every rank folds its
 * own share of the events into a full-size local table (DENSE ids, so a slot means the same key on
 * every rank) and a window is fired by the rank that owns the slot range, which reads that range of
 * every rank's table over NVLink peer memory inside the fire kernel.  Per batch the ranks exchange
 * 32 bytes (max timestamp and live pane range, one all-reduce(MAX) over fmq_exchange_ptr) instead of
 * 24 bytes per event.  Order-dependent keyed state (A7) cannot be combined this way and stays on
 * the all-to-all route (fmq_partition_by_owner).  Call sequence per micro-batch, on every rank:
 *     fmq_push_local(..) / fmq_push_local_packed(..)
 *                                              ingest own events (host or device columns; host
 *                                              columns may have been staged with fmq_stage_host)
 *     all-reduce(MAX) of the 4 int64 at fmq_exchange_ptr(h), on fmq_stream(h)   [also the barrier
 *                                               "every rank has ingested"]
 *     fmq_fire_global(h, watermark_after)      same watermark on every rank; fires own slot range
 *     barrier on fmq_stream(h)                 [every rank has finished reading its peers]
 *     fmq_retire(h)                            clear expired panes, advance the pane origin
 * Peer tables are plain device pointers: within one process pass fmq_table_ptr of the other
 * handles; across processes exchange fmq_ipc_export blobs (64 bytes) and map them with fmq_ipc_open. */
void*   fmq_table_ptr(fmq_handle* h);
void*   fmq_exchange_ptr(fmq_handle* h);
int32_t fmq_ipc_export(fmq_handle* h, void* handle64);
int32_t fmq_ipc_open(fmq_handle* h, const void* handle64, void** ptr);
int32_t fmq_attach_peers(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_tables);
int32_t fmq_push_local(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val,
                       size_t n, int32_t mem);
int32_t fmq_push_local_packed(fmq_handle* h, const fmq_packed_columns* cols, size_t n, int32_t mem);
int32_t fmq_fire_global(fmq_handle* h, int64_t watermark_after);
int32_t fmq_retire(fmq_handle* h);
/* Tumbling windows only: run everything from the fire kernel on (fire, the barrier, fmq_retire) on a second
 * stream, fmq_fire_stream(h), so that it overlaps the NEXT micro-batch's fmq_push_local and first
 * all-reduce.  A tumbling fire only reads panes no later record can reach, and the ingest kernel keeps
 * away from ring cells whose clear is still pending, so the results are the same; what changes is the
 * schedule.  Per micro-batch:
 *     fmq_push_local(..); all-reduce(MAX) of fmq_exchange_ptr(h) on fmq_stream(h);
 *     fmq_fire_global(h, wm)     [waits for the previous fmq_retire, plans on fmq_stream(h), fires on fmq_fire_stream(h)]
 *     barrier on fmq_fire_stream(h) — through a SECOND communicator, the two streams' collectives interleave;
 *     fmq_retire(h)              [on fmq_fire_stream(h)]
 * A fire with no preceding local push (end-of-input flush) calls fmq_push_local with n = 0 first.
 * pane_ring must cover the live panes plus the panes that expire in one step (else FMQ_ERR_RING_OVERFLOW).
 * Also accepted on a handle without peers (world = 1, fmq_push_local / fmq_fire_global / fmq_retire with no all-reduce):
 * same rows, but no faster than fmq_push_batch — on one GPU the fire kernel and the ingest kernel load the same L2
 * (tools/exp_overlap_1gpu.py, profiles/r02_notes.md).
 * No reference counterpart (Flink fires windows on the task thread between records). */
int32_t fmq_combiner_overlap(fmq_handle* h, int32_t enable);
void*   fmq_fire_stream(fmq_handle* h);

/* ---- keyBy exchange over NVLink peer memory ("peer pull", A4) --------------------------------------
 * Replaces: the keyBy network exchange (KeyGroupStreamPartitioner + Netty/local channels behind
 * BandwidthMonitorWithEventTime.java:45) for ranks on one NVSwitch node.  Per micro-batch, on every rank:
 *     fmq_exchange_split(h, ts, key, val, n, parity, counts)    owner buckets into a buffer all peers map
 *     all-gather the `world` counts of every rank into matrix[src * world + dst]   (caller, e.g. NCCL)
 *     fmq_exchange_pull_push(h, parity, matrix, watermark_after, &n)   pull own buckets out of every
 *                       peer's buffer (source-rank order) and run the operator chain on them
 * parity alternates 0/1 per micro-batch.  Works for every operator (per-key arrival order survives) and
 * both key modes.  Set-up: fmq_exchange_alloc (returns a 64-byte CUDA IPC blob to hand to the peers),
 * fmq_ipc_open on the peers' blobs, fmq_exchange_attach (within one process: fmq_exchange_base). */
int32_t fmq_exchange_alloc(fmq_handle* h, void* handle64);
void*   fmq_exchange_base(fmq_handle* h);
int32_t fmq_exchange_attach(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_bases);
int32_t fmq_exchange_split(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val,
                           size_t n, int32_t parity, uint64_t* counts);
int32_t fmq_exchange_pull_push(fmq_handle* h, int32_t parity, const uint64_t* matrix,
                               int64_t watermark_after, size_t* n_received);

/* ---- keyBy exchange inside the library (A4; SURVEY.md §8b: rank / world and the peers' memory belong to the boundary) ----
 * Replaces: the keyBy network exchange in front of the keyed operator — KeyGroupStreamPartitioner + the Netty / local
 * channels behind BandwidthMonitorWithEventTime.java:45 (`.keyBy(1)`), in front of :46-47 (window + reduce on the owning
 * subtask) — for ranks on one NVSwitch node, with no collective library and no host round trip per micro-batch.
 * Every rank owns a receive buffer the peers map (2 parities x `world` segments of three 8-byte columns).  Set-up, once:
 *     fmq_shard_alloc(h, world, max_send, blob64)   allocate; blob64 = 64-byte CUDA IPC handle to give to the peers
 *                                                   (any transport: Flink's own RPC, a file, a collective)
 *     fmq_shard_attach_ipc(h, rank, world, blobs)   blobs = world x 64 bytes in rank order (own entry ignored); or, for
 *     fmq_shard_attach(h, rank, world, bases)       handles living in one process, the fmq_shard_base() pointers
 *     -- a barrier of the caller's: every rank has attached before the first send --
 * Per micro-batch, on every rank (same order of batches everywhere):
 *     fmq_shard_send(h, ts, key, val, n, mem)       this rank's slice (n <= max_send; host or device columns): the owner
 *                                                   split's scatter kernel writes every owner's run straight into that
 *                                                   owner's segment — local stores or NVLink peer stores — then tells
 *                                                   each owner {count, largest timestamp of the slice} behind a flag.
 *                                                   Runs on its own stream, beside the previous batch's kernels.
 *     fmq_shard_recv(h, watermark_after, n_recv)    waits ON THE DEVICE for all `world` flags, then runs the operator
 *                                                   chain on the segments in source-rank order (per-key arrival order
 *                                                   as fmq_push_batch on the concatenated slices would see it) and
 *                                                   acknowledges.  FMQ_WM_AUTO derives the watermark from the largest
 *                                                   timestamp of the WHOLE batch (one assigner for the job,
 *                                                   BandwidthMonitorWithEventTime.java:30-45), so all ranks agree.
 *                                                   n_recv may be NULL; non-NULL reads the count back (synchronises).
 * Routing and key rewriting are those of fmq_partition_by_owner.  Works for every operator and both key modes; jobs with
 * order-dependent keyed state (A7) use the stable two-pass split and read one 8-byte count back per batch (their sort
 * sizes its launches on the host), all others a single-pass split and nothing at all.  fmq_config.max_batch must cover
 * what a rank can RECEIVE (keyed jobs).  A peer that does not show up within FMQ_PEER_TIMEOUT_MS (environment, default
 * 20000) poisons the handle with FMQ_ERR_PEER_TIMEOUT instead of hanging the GPU.  Results are polled as usual; a DENSE
 * handle's rows carry the rank-local id (global id = local * world + rank). */
int32_t fmq_shard_alloc(fmq_handle* h, int32_t world, size_t max_send, void* handle64);
void*   fmq_shard_base(fmq_handle* h);
int32_t fmq_shard_attach(fmq_handle* h, int32_t rank, int32_t world, const void* const* peer_bases);
int32_t fmq_shard_attach_ipc(fmq_handle* h, int32_t rank, int32_t world, const void* blobs64);
int32_t fmq_shard_send(fmq_handle* h, const int64_t* ts, const uint64_t* key, const void* val, size_t n, int32_t mem);
int32_t fmq_shard_recv(fmq_handle* h, int64_t watermark_after, size_t* n_received);

/* Combiner mode without a collective library: after fmq_attach_peers AND fmq_shard_alloc(h, world, 1, ..) +
 * fmq_shard_attach[_ipc] (the control block at the head of that buffer is all that is used), the 32 bytes of batch
 * scalars and both barriers of the call sequence above travel as peer stores behind device flags inside
 * fmq_push_local / fmq_fire_global / fmq_retire: the caller issues no all-reduce and no barrier.  Every
 * fmq_fire_global then follows an fmq_push_local (n = 0 for an end-of-input flush).  No reference counterpart. */
int32_t fmq_combiner_use_flags(fmq_handle* h, int32_t enable);

/* Largest batch fmq_partition_by_owner_noalloc accepts (events). */
#define FMQ_NOALLOC_MAX_EVENTS 4194304

/* No counterpart in the reference.  The stable owner split of fmq_partition_by_owner in its
 * FMQ_KEYS_HASHED form — owner = fmq_owner_of(key, world) = mix64(key) % world, key column copied
 * unchanged — for callers that may not allocate device memory while they run.  (It has no handle and
 * therefore no key mode: a FMQ_KEYS_DENSE job, whose handle-based split routes by id % world and
 * rewrites the ids to id / world, must not switch to this call.)  Such callers are: an operator thread that shares its
 * GPU with other subtasks (cudaMalloc / cudaFree synchronise the whole device and would stall
 * their kernels), or a process that has handed all device memory to its own arena, where a
 * driver allocation can simply fail.  The call needs no handle and calls no cudaMalloc, cudaFree
 * or any other allocating routine: the only scratch it uses is a fixed ~2 MB array built into the
 * library's device image (every GPU that loads the library carries it until the process exits).
 *   - n may not exceed FMQ_NOALLOC_MAX_EVENTS; larger batches are refused with FMQ_ERR_BAD_ARG
 *     (message via fmq_last_error(NULL)) — split them, or use fmq_partition_by_owner.
 *   - ts/key/val and the *_out columns are device pointers owned by the caller; counts[world] is
 *     host memory.  The work is queued on `stream` (a cudaStream_t, NULL = default stream) of
 *     `device`, and the call returns after that stream has drained (the counts are read back).
 *   - The first call on each device looks up the device address of the built-in scratch once, and
 *     like any first use of a kernel it makes the CUDA runtime load this library's device code
 *     (which includes the scratch array) onto that GPU — work the driver may allocate for.  Make
 *     that first call per device while the job is being set up: before any stream capture, before
 *     an arena takes the rest of the device memory, and before other subtasks depend on the GPU
 *     being undisturbed.  Every later call allocates nothing.
 *   - The call makes `device` current while it runs and puts the caller's current device back
 *     before it returns.
 *   - All calls on one device share the one scratch array, which keeps whatever the previous call
 *     left in it.  Nothing in it needs clearing (every counter is stored before it is read), but
 *     two calls whose kernels can overlap on the same GPU — different streams, different
 *     threads — race on it: run them one after the other. */
int32_t fmq_partition_by_owner_noalloc(int32_t device, void* stream, const int64_t* ts,
                                       const uint64_t* key, const void* val, size_t n,
                                       uint32_t world, int64_t* ts_out, uint64_t* key_out,
                                       void* val_out, uint64_t* counts);

#ifdef __cplusplus
}
#endif
#endif /* FMQ_B200_H */
