// Text ingest on the device: the `map` step in front of every reference job —
//   "<epoch-s> <host> <cpu> <usage>"      chapter1/.../Main.java:21-25, chapter2/.../ComputeCpuAvg.java:22-25
//   "<ISO local time> <channel> <flow>"   chapter3/.../BandwidthMonitorWithEventTime.java:33-34,39-44
// turned from newline-separated bytes into the three 8-byte columns of the C ABI.
// Three steps: count newlines per 4 KB chunk -> exclusive scan -> record every line start;
// then one thread parses one line.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

constexpr u32 TXT_CHUNK = 4096;          // bytes per block in the newline passes
constexpr int TXT_THREADS = 128;         // 32 bytes per thread

// per-line status written to `status_out`
constexpr u32 TXT_OK = 0, TXT_MALFORMED = 1, TXT_VALUE_ON_HOST = 2, TXT_KEY_ON_HOST = 3;

__global__ void __launch_bounds__(TXT_THREADS) txt_count_kernel(const char* __restrict__ text, size_t bytes, u32* __restrict__ counts) {
    __shared__ u32 wsum[TXT_THREADS / 32];
    const size_t base = (size_t)blockIdx.x * TXT_CHUNK + (size_t)threadIdx.x * 32;
    u32 c = 0;
    for (int i = 0; i < 32; ++i) { const size_t p = base + i; if (p < bytes && text[p] == '\n') ++c; }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { u32 t = 0; for (int w = 0; w < TXT_THREADS / 32; ++w) t += wsum[w]; counts[blockIdx.x] = t; }
}

// line_start[0] = 0; line_start[k+1] = position after the k-th newline.  `offsets` = scanned counts.
__global__ void __launch_bounds__(TXT_THREADS) txt_mark_kernel(const char* __restrict__ text, size_t bytes, const u32* __restrict__ offsets,
                                                              u64* __restrict__ line_start, size_t cap_lines) {
    __shared__ u32 wsum[TXT_THREADS / 32];
    const size_t base = (size_t)blockIdx.x * TXT_CHUNK + (size_t)threadIdx.x * 32;
    u32 c = 0;
    for (int i = 0; i < 32; ++i) { const size_t p = base + i; if (p < bytes && text[p] == '\n') ++c; }
    // exclusive prefix of c inside the block
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = c;
    for (int o = 1; o < 32; o <<= 1) { const u32 t = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += t; }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    u32 wbase = 0;
    for (u32 i = 0; i < w; ++i) wbase += wsum[i];
    u64 k = (u64)offsets[blockIdx.x] + wbase + inc - c;             // index of this thread's first newline
    if (blockIdx.x == 0 && threadIdx.x == 0) line_start[0] = 0;
    for (int i = 0; i < 32; ++i) {
        const size_t p = base + i;
        if (p < bytes && text[p] == '\n') { if (k + 1 < cap_lines) line_start[k + 1] = p + 1; ++k; }
    }
}

// FNV-1a over the key bytes, finished with mix64; never the reserved KEY_EMPTY.  text.py restates it.
__device__ __forceinline__ u64 txt_key_hash(const char* s, int len) {
    u64 h = 0xcbf29ce484222325ull;
    for (int i = 0; i < len; ++i) { h ^= (unsigned char)s[i]; h *= 0x100000001b3ull; }
    h = mix64(h);
    return h == KEY_EMPTY ? 0 : h;
}

__device__ __forceinline__ bool txt_digit(char c) { return c >= '0' && c <= '9'; }

// decimal -> int64 (Long.parseLong: optional sign, digits only); false if malformed / overflow-prone
__device__ __forceinline__ bool txt_parse_i64(const char* s, int len, i64* out) {
    if (len <= 0 || len > 19) return false;
    int i = 0; bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; if (len == 1) return false; }
    u64 v = 0;
    for (; i < len; ++i) { if (!txt_digit(s[i])) return false; v = v * 10 + (u64)(s[i] - '0'); }
    if (len - (neg ? 1 : 0) > 18) return false;
    *out = neg ? -(i64)v : (i64)v;
    return true;
}

// decimal -> double, exactly like Double.parseDouble WHEN the exact (Clinger) fast path applies:
// at most 15 significant digits and a decimal exponent within [-22, 22]: mantissa and power of ten
// are both exact doubles, so one IEEE multiply or divide is the correctly rounded result.
// Returns TXT_OK, TXT_MALFORMED, or TXT_VALUE_ON_HOST (well-formed but outside the exact path: the
// host parses that one value).
__device__ __forceinline__ u32 txt_parse_f64(const char* s, int len, double* out) {
    const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    // anything this routine does not recognise (NaN, Infinity, hex floats, a trailing d/f ...) is left to
    // the host's parser rather than declared malformed
    if (len <= 0) return TXT_MALFORMED;
    int i = 0; bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; }
    u64 m = 0; int ndig = 0, nsig = 0, dec_exp = 0; bool seen_point = false, any = false;
    for (; i < len; ++i) {
        const char c = s[i];
        if (txt_digit(c)) {
            any = true;
            if (nsig > 0 || c != '0') { if (nsig < 19) { m = m * 10 + (u64)(c - '0'); ++nsig; if (seen_point) --dec_exp; } else if (!seen_point) ++dec_exp; }
            else if (seen_point) --dec_exp;
            ++ndig;
        } else if (c == '.' && !seen_point) seen_point = true;
        else break;
    }
    if (!any) return TXT_VALUE_ON_HOST;
    if (i < len) {
        if (s[i] != 'e' && s[i] != 'E') return TXT_VALUE_ON_HOST;
        ++i; bool eneg = false; if (i < len && (s[i] == '-' || s[i] == '+')) { eneg = s[i] == '-'; ++i; }
        if (i >= len) return TXT_VALUE_ON_HOST;
        int e = 0;
        for (; i < len; ++i) { if (!txt_digit(s[i])) return TXT_VALUE_ON_HOST; if (e < 10000) e = e * 10 + (s[i] - '0'); }
        dec_exp += eneg ? -e : e;
    }
    if (m == 0) { *out = neg ? -0.0 : 0.0; return TXT_OK; }
    if (nsig > 15 || dec_exp < -22 || dec_exp > 22) return TXT_VALUE_ON_HOST;
    double v = (double)m;
    v = dec_exp >= 0 ? v * P10[dec_exp] : v / P10[-dec_exp];
    *out = neg ? -v : v;
    return TXT_OK;
}

// days since 1970-01-01 of a proleptic Gregorian date
__device__ __forceinline__ i64 txt_days_from_civil(i64 y, u32 m, u32 d) {
    y -= m <= 2;
    const i64 era = (y >= 0 ? y : y - 399) / 400;
    const u32 yoe = (u32)(y - era * 400);
    const u32 doy = (153 * (m > 2 ? m - 3 : m + 9) + 2) / 5 + d - 1;
    const u32 doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + (i64)doe - 719468;
}

struct TextArgs {
    const char* text; size_t bytes; const u64* line_start; size_t n_lines;
    int format;                 // 0: "<epoch-s> <host> <cpu> <usage>" -> f64;  1: "<ISO> <channel> <flow>" -> i64
    int utc_offset_hours;       // the reference uses ZoneOffset.ofHours(8)
    i64* ts; u64* key; u64* val; u32* status; unsigned long long* n_bad;
    u64* key_off; u32* key_len;     // optional: where the key field of every line lies in the text (device dictionary)
};

__global__ void __launch_bounds__(256) txt_parse_kernel(const TextArgs a) {
    const size_t li = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= a.n_lines) return;
    const size_t lo = a.line_start[li];
    size_t hi = (li + 1 < a.n_lines || true) ? a.line_start[li + 1] : a.bytes;    // line_start has n_lines + 1 entries
    if (hi > lo && a.text[hi - 1] == '\n') --hi;
    if (hi > lo && a.text[hi - 1] == '\r') --hi;
    const char* s = a.text + lo; const int len = (int)(hi - lo);
    // String.split(" "): fields separated by single spaces
    int f0 = 0, fl[4] = {0, 0, 0, 0}, fs[4] = {0, 0, 0, 0}, nf = 0;
    for (int i = 0; i <= len && nf < 4; ++i) if (i == len || s[i] == ' ') { fs[nf] = f0; fl[nf] = i - f0; ++nf; f0 = i + 1; }
    u32 st = TXT_OK; i64 ts = 0; u64 key = 0, vb = 0;
    if (a.format == 0) {
        i64 sec;
        if (nf < 4 || !txt_parse_i64(s + fs[0], fl[0], &sec) || fl[1] <= 0) st = TXT_MALFORMED;
        else {
            ts = sec * 1000; key = txt_key_hash(s + fs[1], fl[1]);
            double v = 0.0; st = txt_parse_f64(s + fs[3], fl[3], &v); vb = (u64)__double_as_longlong(v);
        }
    } else {
        const char* t = s + fs[0];
        bool ok = nf >= 3 && fl[0] == 19 && t[4] == '-' && t[7] == '-' && t[10] == 'T' && t[13] == ':' && t[16] == ':' && fl[1] > 0;
        int num[6] = {0, 0, 0, 0, 0, 0};
        if (ok) {
            const int pos[6] = {0, 5, 8, 11, 14, 17}, nd[6] = {4, 2, 2, 2, 2, 2};
            for (int q = 0; q < 6 && ok; ++q) for (int d = 0; d < nd[q]; ++d) { const char c = t[pos[q] + d]; if (!txt_digit(c)) { ok = false; break; } num[q] = num[q] * 10 + (c - '0'); }
        }
        i64 flow = 0;
        if (!ok || !txt_parse_i64(s + fs[2], fl[2], &flow)) st = TXT_MALFORMED;
        else {
            const i64 sec = txt_days_from_civil(num[0], (u32)num[1], (u32)num[2]) * 86400 + num[3] * 3600 + num[4] * 60 + num[5] - (i64)a.utc_offset_hours * 3600;
            ts = (i64)(int)sec * 1000;                               // the reference's (int) cast, …EventTime.java:33-34
            key = txt_key_hash(s + fs[1], fl[1]); vb = (u64)flow;
        }
    }
    a.ts[li] = ts; a.key[li] = key; a.val[li] = vb; a.status[li] = st;
    if (a.key_off) { a.key_off[li] = lo + (size_t)fs[1]; a.key_len[li] = (nf > 1 && fl[1] > 0) ? (u32)fl[1] : 0u; }
    if (st != TXT_OK) atomicAdd(a.n_bad, 1ull);
}

// ---- device dictionary: host / channel strings -> dense ids 0, 1, 2, ... -------------------------
// What the reference's `map` + keyBy(String) amounts to for the state behind it (Main.java:21-25,
// BandwidthMonitorWithEventTime.java:39-45): every distinct string gets the next free id the first time it is
// seen, the id indexes the FMQ_KEYS_DENSE tables directly, and the string is kept in an arena so that rows can be
// printed with the name again (fmq_dict_lookup).  Open addressing on the 64-bit FNV/mix hash of the string; the
// strings themselves are compared on lookup, so two strings that share a hash are told apart (the second one is
// handed back to the host, TXT_KEY_ON_HOST, rather than merged).
// Two kernels, so that nobody ever waits for another thread: `insert` claims hash slots (the CAS winner allocates
// the id and copies the string), `lookup` runs after it and reads only published entries.
struct DictArgs {
    u64* hash; u32* id; u64 mask;                       // table of mask + 1 slots (a power of two), KEY_EMPTY = free
    u64* name_off; u32* name_len; char* arena; u64 arena_cap; u64 id_cap;
    unsigned long long* counters;                       // [0] next id, [1] arena bytes used, [2] error flags (1 = ids, 2 = arena)
    const char* text; const u64* key_off; const u32* key_len; size_t n_lines;
    u64* key; u32* status; unsigned long long* n_bad;
};
__global__ void __launch_bounds__(256) dict_insert_kernel(const DictArgs a) {
    const size_t li = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= a.n_lines || a.status[li] == TXT_MALFORMED) return;
    const u64 h = a.key[li];
    for (u64 s = h & a.mask, probe = 0; probe <= a.mask; ++probe, s = (s + 1) & a.mask) {
        u64 cur = a.hash[s];
        if (cur == KEY_EMPTY) cur = atomicCAS(&a.hash[s], KEY_EMPTY, h);
        if (cur == KEY_EMPTY) {                                            // first sight of this hash: this thread enters the string
            const u64 id = atomicAdd(&a.counters[0], 1ull);
            const u32 len = a.key_len[li];
            const u64 off = atomicAdd(&a.counters[1], (unsigned long long)len);
            if (id >= a.id_cap) { atomicOr(&a.counters[2], 1ull); a.id[s] = 0xFFFFFFFFu; return; }
            if (off + len > a.arena_cap) { atomicOr(&a.counters[2], 2ull); a.id[s] = 0xFFFFFFFFu; return; }
            const char* src = a.text + a.key_off[li];
            for (u32 i = 0; i < len; ++i) a.arena[off + i] = src[i];
            a.name_off[id] = off; a.name_len[id] = len; a.id[s] = (u32)id;
            return;
        }
        if (cur == h) return;                                              // known (or being entered by the winner right now)
    }
    atomicOr(&a.counters[2], 1ull);
}
__global__ void __launch_bounds__(256) dict_lookup_kernel(const DictArgs a) {
    const size_t li = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= a.n_lines || a.status[li] == TXT_MALFORMED) return;
    const u64 h = a.key[li];
    u64 out = KEY_EMPTY;
    for (u64 s = h & a.mask, probe = 0; probe <= a.mask; ++probe, s = (s + 1) & a.mask) {
        const u64 cur = a.hash[s];
        if (cur == KEY_EMPTY) break;
        if (cur != h) continue;
        const u32 id = a.id[s];
        if (id == 0xFFFFFFFFu) break;                                      // dictionary full
        const u32 len = a.key_len[li];
        bool same = a.name_len[id] == len;
        const char* x = a.arena + a.name_off[id]; const char* y = a.text + a.key_off[li];
        for (u32 i = 0; same && i < len; ++i) same = x[i] == y[i];
        if (same) out = id;
        break;                                                             // same hash, another string: not merged, the host takes it
    }
    if (out == KEY_EMPTY) { if (a.status[li] == TXT_OK) atomicAdd(a.n_bad, 1ull); a.status[li] = TXT_KEY_ON_HOST; out = 0; }
    a.key[li] = out;
}
// names of `n` ids into rows of `stride` bytes (zero padded); lens[i] = full length, 0xFFFFFFFF for an unknown id
__global__ void __launch_bounds__(256) dict_names_kernel(const u64* ids, size_t n, const u64* name_off, const u32* name_len, const char* arena, u64 n_ids,
                                                         char* out, size_t stride, u32* lens) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 id = ids[i];
    char* row = out + i * stride;
    if (id >= n_ids) { lens[i] = 0xFFFFFFFFu; for (size_t j = 0; j < stride; ++j) row[j] = 0; return; }
    const u32 len = name_len[id]; lens[i] = len;
    const char* src = arena + name_off[id];
    for (size_t j = 0; j < stride; ++j) row[j] = j < len ? src[j] : 0;
}

}  // namespace fmq
