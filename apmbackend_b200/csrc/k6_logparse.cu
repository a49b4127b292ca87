// K6: the CommonTiming lines of a WildFly log, parsed on the device.
//
// Replaces, for a buffer of raw text of one log file, the per-line work of
// stream_parse_transactions.js: readLine's dispatch on the four CommonTiming patterns (:733-777),
// parseEjbCommonTimingEntry / Exit (:378-446), parseCommonTimingEntry / Exit (:451-565),
// convertStringDateToMs (:242-256) and outputRecord's rewrite / fallback (:264-293), and pairs the
// Start and Stop lines of a (logId, service) inside the buffer (recordCache, :393-400, :471-480).
// The exits that are complete go, as transaction columns that never leave the device, into the
// hot path (apm_feed_parsed); every recognised line also comes back as an apm_log_record for the
// host-side bookkeeping the drop-in keeps (account numbers, the need-number cache, database rows).
//
// Pipeline: newline count per 4 KB tile -> scan -> line table -> one thread per line (classify,
// tokenise, convert, hash, dictionary lookup) -> ordered compaction of the recognised lines ->
// radix sort of (pair hash, record index) -> pairing against the neighbour in sorted order ->
// ordered compaction of the fed exits into service / jvm / endTs / elapsed columns.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ctx.h"

namespace {

constexpr int LP_THREADS = 256;
constexpr int LP_TILE_BYTES = 4096;                       // 16 bytes per thread
constexpr int LP_TILE_ITEMS = 1024;                       // flags per tile in the ordered compactions (4 per thread)
constexpr uint64_t FNV_OFFSET = 0xcbf29ce484222325ull, FNV_PRIME = 0x100000001b3ull;
constexpr uint32_t LP_MAX_RECORDS = 1u << 24;             // record index field of the pairing sort word
constexpr uint64_t LP_SUB_BYTES = 32ull << 20;            // sub-buffer of a parse call: copy / parse overlap granularity

struct LogParseState {
    unsigned char* d_text = nullptr;
    uint64_t text_cap = 0;
    uint32_t* tile_counts = nullptr;   // newline / flag counts per tile, then their exclusive scan in place
    uint64_t tiles_cap = 0;
    unsigned long long* d_total = nullptr;  // [4] totals of the scans, result counters ride behind: [4..9]
    unsigned long long* h_total = nullptr;  // pinned mirror
    uint64_t* nlpos = nullptr;         // position of every newline
    uint64_t lines_cap = 0;
    apm_log_record* rec_sparse = nullptr;   // one per line
    uint64_t* rawh_sparse = nullptr;        // identity hash of (logId, raw service) per line
    apm_log_record* rec = nullptr;          // recognised lines, line order
    uint64_t* rawh = nullptr;
    uint64_t rec_cap = 0;
    uint64_t* words[2] = {nullptr, nullptr};  // pairing sort ping-pong
    uint32_t* sort_hist = nullptr;
    uint32_t* sort_base = nullptr;
    uint64_t sort_tiles_cap = 0;
    apm_tx_columns cols = {nullptr, nullptr, nullptr, nullptr};  // fed exits
    uint64_t cols_cap = 0, n_fed = 0;
    // dictionary: open addressing, key 0 = empty (a hash of 0 is stored as 1)
    uint64_t* dict_keys = nullptr;
    uint32_t* dict_vals = nullptr;
    uint32_t dict_mask = 0;
    std::vector<cudaEvent_t> ev_sub;   // one per sub-buffer: its copy has landed
};

__host__ __device__ __forceinline__ uint64_t fnv_step(uint64_t h, unsigned char c) { return (h ^ c) * FNV_PRIME; }

__device__ __forceinline__ bool lp_is_ws(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }

// ---- ordered scans ------------------------------------------------------------------------------
// exclusive scan of per-tile counts by one block (the tile count is at most a few hundred thousand)
__global__ void __launch_bounds__(1024) lp_scan_kernel(uint32_t* __restrict__ counts, uint64_t ntiles, unsigned long long* __restrict__ total) {
    __shared__ unsigned long long s_warp[32];
    const uint64_t per = (ntiles + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per, hi = lo + per < ntiles ? lo + per : ntiles;
    unsigned long long sum = 0;
    for (uint64_t i = lo; i < hi; ++i) sum += counts[i];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    if (w == 0) {
        unsigned long long x = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        s_warp[lane] = x;
    }
    __syncthreads();
    unsigned long long run = incl - sum + (w ? s_warp[w - 1] : 0ull);
    for (uint64_t i = lo; i < hi; ++i) {
        const uint32_t c = counts[i];
        counts[i] = (uint32_t)run;  // tile bases fit 32 bits: at most 2^32 bytes / lines per call
        run += c;
    }
    if (threadIdx.x == 1023) *total = s_warp[31];
}

// block-wide exclusive scan of one value per thread
__device__ __forceinline__ uint32_t lp_block_scan(uint32_t v, uint32_t* s_warp) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int j = 0; j < LP_THREADS / 32; ++j)
        if (j < w) woff += s_warp[j];
    __syncthreads();
    return woff + incl - v;
}

// ---- line table ---------------------------------------------------------------------------------
// Both kernels work on the byte range [lo, hi) of the text buffer; tiles are anchored at lo rounded down to the
// tile size, so the 16-byte loads stay aligned whatever the range.
__global__ void __launch_bounds__(LP_THREADS) lp_count_newlines_kernel(const unsigned char* __restrict__ text, uint64_t lo, uint64_t hi,
                                                                       uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (lo & ~(uint64_t)(LP_TILE_BYTES - 1)) + (uint64_t)blockIdx.x * LP_TILE_BYTES + (uint64_t)threadIdx.x * 16;
    uint32_t cnt = 0;
    if (base >= lo && base + 16 <= hi) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + base));  // the buffer is 256-byte aligned
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t x = w[k] ^ 0x0A0A0A0Au;  // a zero byte where the text has a newline
#pragma unroll
            for (int b = 0; b < 4; ++b) cnt += ((x >> (8 * b)) & 0xFFu) == 0u;
        }
    } else {
        for (uint64_t i = base > lo ? base : lo; i < hi && i < base + 16; ++i) cnt += text[i] == 10;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int j = 0; j < LP_THREADS / 32; ++j) t += s_warp[j];
        tile_counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(LP_THREADS) lp_write_newlines_kernel(const unsigned char* __restrict__ text, uint64_t lo, uint64_t hi,
                                                                       const uint32_t* __restrict__ tile_base,
                                                                       uint64_t* __restrict__ nlpos) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (lo & ~(uint64_t)(LP_TILE_BYTES - 1)) + (uint64_t)blockIdx.x * LP_TILE_BYTES + (uint64_t)threadIdx.x * 16;
    unsigned char b[16];
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        b[k] = (base + k >= lo && base + k < hi) ? __ldg(text + base + k) : 0;
        cnt += b[k] == 10;
    }
    uint32_t pos = tile_base[blockIdx.x] + lp_block_scan(cnt, s_warp);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        if (b[k] == 10) nlpos[pos++] = base + k;  // absolute position in the buffer
}

// ---- one line -------------------------------------------------------------------------------------
struct ParseParams {
    const unsigned char* text;
    uint64_t lo, hi;           // byte range of this sub-buffer; its lines are numbered from 0
    const uint64_t* nlpos;
    uint64_t nl, nlines;
    uint32_t file_type, jvm_id, num_services;
    apm_tz_rule tz[APM_MAX_TZ_RULES];
    uint32_t num_tz;
    const uint64_t* dict_keys;
    const uint32_t* dict_vals;
    uint32_t dict_mask;
    apm_log_record* rec_sparse;
    uint64_t* rawh_sparse;
};

struct Tok { uint32_t s, e; };  // [s, e) inside the line; s == e: empty string; s == 0xFFFFFFFF: undefined

__device__ __forceinline__ bool lp_match(const unsigned char* t, uint32_t len, uint32_t at, const char* lit, int n) {
    if (at + n > len) return false;
    for (int i = 0; i < n; ++i)
        if (t[at + i] != (unsigned char)lit[i]) return false;
    return true;
}

// floor division helpers for the Date constructor's field normalisation
__device__ __forceinline__ long long lp_floordiv(long long a, long long b) { long long q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }

__device__ __forceinline__ long long lp_days_from_civil(long long y, int m, int d) {
    y -= m <= 2;
    const long long era = lp_floordiv(y, 400);
    const long long yoe = y - era * 400;
    const long long doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const long long doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}

// `${arr[1]} ${arr[2]}`.trim().split(/-|[\s]+|:|,/) -> new Date(f0, f1 - 1, f2, f3, f4, f5, f6).getTime()
__device__ bool lp_parse_time(const unsigned char* t, Tok a, Tok b, const ParseParams& p, int64_t* out) {
    if (a.s == 0xFFFFFFFFu || b.s == 0xFFFFFFFFu) return false;  // "undefined" is not a number
    long long f[7];
    int nf = 0;
    long long cur = 0;
    bool any_bad = false;
    for (int part = 0; part < 2; ++part) {
        const Tok k = part ? b : a;
        if (k.s == k.e) {
            // an empty token can only be arr[0]-style leading / trailing artefacts; the joined string is trimmed,
            // so it contributes no field
            continue;
        }
        cur = 0;
        for (uint32_t i = k.s; i < k.e; ++i) {
            const unsigned char c = t[i];
            if (c == '-' || c == ':' || c == ',') {
                if (nf < 7) f[nf] = cur;
                ++nf;
                cur = 0;
            } else if (c >= '0' && c <= '9') {
                cur = cur * 10 + (c - '0');
                if (cur > 100000000000ll) any_bad = true;
            } else {
                any_bad = true;
            }
        }
        if (nf < 7) f[nf] = cur;
        ++nf;
    }
    if (any_bad || nf < 7) return false;
    long long y = f[0];
    if (y >= 0 && y <= 99) y += 1900;  // the Date constructor's two-digit-year rule
    const long long mo = f[1] - 1;      // 0-based, may be out of range: carried into the year
    const long long yy = y + lp_floordiv(mo, 12);
    const int mm = (int)(mo - lp_floordiv(mo, 12) * 12) + 1;
    long long local = lp_days_from_civil(yy, mm, 1) * 86400000ll;
    local += (f[2] - 1) * 86400000ll + f[3] * 3600000ll + f[4] * 60000ll + f[5] * 1000ll + f[6];
    long long off = p.num_tz ? p.tz[0].offset_ms : 0;
    for (uint32_t r = 1; r < p.num_tz; ++r)
        if (local >= p.tz[r].local_from_ms) off = p.tz[r].offset_ms;
    *out = local - off;
    return true;
}

// parseInt(token): optional sign, decimal digits, stops at the first other character
__device__ bool lp_parse_int(const unsigned char* t, Tok k, int32_t* out) {
    if (k.s == 0xFFFFFFFFu || k.s == k.e) return false;
    uint32_t i = k.s;
    bool neg = false;
    if (t[i] == '+' || t[i] == '-') { neg = t[i] == '-'; ++i; }
    long long v = 0;
    bool any = false;
    for (; i < k.e && t[i] >= '0' && t[i] <= '9'; ++i) {
        v = v * 10 + (t[i] - '0');
        any = true;
        if (v > 0x7fffffffll) v = 0x7fffffffll;
    }
    if (!any) return false;
    *out = (int32_t)(neg ? -v : v);
    return true;
}

// Hashes of a service: `raw` over the cache key (the token, "S:"-prefixed for the EJB forms), `name` over
// what outputRecord makes of it: first /Provider\[/i -> "Provider:", then the first ']' removed (:274).
__device__ void lp_service_hashes(const unsigned char* t, Tok k, bool ejb, uint64_t* raw, uint64_t* name, bool* top) {
    const char* undef = "undefined";
    const uint32_t body = k.s == 0xFFFFFFFFu ? 9u : k.e - k.s;
    const uint32_t pre = ejb ? 2u : 0u, len = pre + body;
    auto at = [&](uint32_t i) -> unsigned char {
        if (i < pre) return (unsigned char)(i == 0 ? 'S' : ':');
        return k.s == 0xFFFFFFFFu ? (unsigned char)undef[i - pre] : t[k.s + (i - pre)];
    };
    int pm = -1;
    const char* pat = "provider[";
    for (uint32_t i = 0; i + 9 <= len && pm < 0; ++i) {
        bool ok = true;
        for (int j = 0; j < 9 && ok; ++j) {
            unsigned char c = at(i + j);
            if (c >= 'A' && c <= 'Z') c += 32;
            ok = c == (unsigned char)pat[j];
        }
        if (ok) pm = (int)i;
    }
    uint64_t hr = FNV_OFFSET, hn = FNV_OFFSET;
    bool dropped = false;
    unsigned char first2[2] = {0, 0};
    uint32_t emitted = 0;
    for (uint32_t i = 0; i < len; ++i) {
        unsigned char c = at(i);
        hr = fnv_step(hr, c);
        if (pm >= 0 && (int)i >= pm && (int)i < pm + 9) c = (unsigned char)"Provider:"[(int)i - pm];  // the literal replaces the match, case included
        if (c == ']' && !dropped) { dropped = true; continue; }
        if (emitted < 2) first2[emitted] = c;
        ++emitted;
        hn = fnv_step(hn, c);
    }
    *raw = hr;
    *name = hn;
    *top = emitted >= 2 && first2[0] == 'S' && first2[1] == ':';
}

__global__ void __launch_bounds__(LP_THREADS) lp_parse_kernel(const ParseParams p) {
    const uint64_t line = (uint64_t)blockIdx.x * LP_THREADS + threadIdx.x;
    if (line >= p.nlines) return;
    const uint64_t ls = line == 0 ? p.lo : p.nlpos[line - 1] + 1;
    const uint64_t le = line < p.nl ? p.nlpos[line] : p.hi;
    apm_log_record r;
    memset(&r, 0, sizeof r);
    r.line_offset = ls;
    const uint32_t len = (uint32_t)(le - ls > 0xFFFFu ? 0xFFFFu : le - ls);  // lines are examined up to 64 KB
    r.line_length = (uint32_t)(le - ls);
    r.service_id = APM_NO_SERVICE;
    p.rawh_sparse[line] = 0;
    if (len == 0) { p.rec_sparse[line] = r; return; }  // `if (!line) return` (:742)
    const unsigned char* t = p.text + ls;

    // -- one pass: the first 14 white-space tokens (arr[] of line.split(/[\s]+/)), the INFO occurrences
    Tok arr[14];
#pragma unroll
    for (int i = 0; i < 14; ++i) arr[i].s = arr[i].e = 0xFFFFFFFFu;
    int ntok = 0;
    bool in_tok = false;
    if (lp_is_ws(t[0])) { arr[0].s = arr[0].e = 0; ntok = 1; }  // a leading separator yields an empty arr[0]
    uint32_t info1 = 0xFFFFFFFFu, info2 = 0xFFFFFFFFu;      // first two non-overlapping "INFO"
    uint32_t next_info_ok = 0;
    bool m_ejb_entry = false, m_ejb_exit = false, m_ct_start = false, m_ct_stop = false, m_baf = false;
    for (uint32_t i = 0; i < len; ++i) {
        const unsigned char c = t[i];
        const bool ws = lp_is_ws(c);
        if (!ws && !in_tok) { if (ntok < 14) arr[ntok].s = i; in_tok = true; }
        if (ws && in_tok) { if (ntok < 14) arr[ntok].e = i; ++ntok; in_tok = false; }
        if (c == 'I' && lp_match(t, len, i, "INFO", 4)) {
            if (i >= next_info_ok) {  // String.split scans non-overlapping matches from the left
                if (info1 == 0xFFFFFFFFu) info1 = i; else if (info2 == 0xFFFFFFFFu) info2 = i;
                next_info_ok = i + 4;
            }
            uint32_t j = i + 4;
            while (j < len && t[j] == ' ') ++j;  // ` *`
            m_ejb_entry |= lp_match(t, len, j, "[CommonTiming] The EJB", 22);
            m_ejb_exit |= lp_match(t, len, j, "[CommonTiming] Total time", 25);
            m_ct_start |= lp_match(t, len, j, "CommonTiming::Start", 19);
            m_ct_stop |= lp_match(t, len, j, "CommonTiming::Stop", 18);
            // /\[[^ ]+] +INFO /: "] " (one or more spaces) before, a space after, and a '[' earlier in the same
            // space-free run with at least one character between the brackets
            if (i + 4 < len && t[i + 4] == ' ' && i >= 1 && t[i - 1] == ' ') {
                int q = (int)i - 1;
                while (q >= 0 && t[q] == ' ') --q;
                if (q >= 0 && t[q] == ']') {
                    for (int b = q - 1; b >= 0 && t[b] != ' '; --b)
                        if (t[b] == '[' && b <= q - 2) { m_baf = true; break; }
                }
            }
        }
    }
    if (in_tok) { if (ntok < 14) arr[ntok].e = len; ++ntok; }
    else if (ntok < 14) { arr[ntok].s = arr[ntok].e = len; ++ntok; }  // a trailing separator yields an empty last element

    int kind = 0;
    if (p.file_type == 1) {  // server.log (:764-776)
        if (m_ejb_entry) kind = APM_LK_EJB_ENTRY;
        else if (m_ejb_exit) kind = APM_LK_EJB_EXIT;
        else if (m_ct_start) kind = APM_LK_CT_START;
        else if (m_ct_stop) kind = APM_LK_CT_STOP;
    } else {                 // application logs (:777-787); parseAppLine stays with the host
        if (m_ct_start) kind = APM_LK_CT_START;
        else if (m_ct_stop) kind = APM_LK_CT_STOP;
    }
    if (kind == 0) { p.rec_sparse[line] = r; return; }
    const bool ejb = kind == APM_LK_EJB_ENTRY || kind == APM_LK_EJB_EXIT;
    const bool is_exit = kind == APM_LK_CT_STOP || kind == APM_LK_EJB_EXIT;

    // -- service and elapsed tokens
    Tok svc, el;
    svc.s = svc.e = el.s = el.e = 0xFFFFFFFFu;
    if (ejb) {
        svc = arr[kind == APM_LK_EJB_ENTRY ? 13 : 9];
        if (is_exit) el = arr[11];
    } else {
        // line.split(/INFO/)[1].trim().split(/[\s]+/): tokens [1] and [5] of the text between the first two INFOs
        uint32_t a = info1 + 4, b = info2 == 0xFFFFFFFFu ? len : info2;
        while (a < b && lp_is_ws(t[a])) ++a;
        while (b > a && lp_is_ws(t[b - 1])) --b;
        int k = 0;
        bool in = false;
        uint32_t s0 = a;
        if (a == b) { k = 1; }  // "" -> [""]: a single empty element
        for (uint32_t i = a; i <= b && a < b; ++i) {
            const bool ws = i == b || lp_is_ws(t[i]);
            if (!ws && !in) { s0 = i; in = true; }
            if (ws && in) {
                if (k == 1) { svc.s = s0; svc.e = i; }
                if (k == 5) { el.s = s0; el.e = i; }
                ++k;
                in = false;
            }
        }
    }

    // -- logId: arr[0] without its brackets
    uint64_t hl = FNV_OFFSET;
    uint32_t idlen = 0;
    for (uint32_t i = arr[0].s; i < arr[0].e; ++i)
        if (t[i] != '[' && t[i] != ']') { hl = fnv_step(hl, t[i]); ++idlen; }
    uint8_t flags = 0;
    if (idlen == 0) flags |= APM_LR_LOGID_EMPTY;
    if (m_baf) flags |= APM_LR_BAF_INFO;

    int64_t ts = 0;
    if (!lp_parse_time(t, arr[1], arr[2], p, &ts)) flags |= APM_LR_BAD_TIME;
    int32_t elapsed = 0;
    if (is_exit && !lp_parse_int(t, el, &elapsed)) flags |= APM_LR_BAD_ELAPSED;

    uint64_t hraw, hname;
    bool top;
    lp_service_hashes(t, svc, ejb, &hraw, &hname, &top);
    if (top) flags |= APM_LR_TOP_LEVEL;

    uint32_t id = APM_NO_SERVICE;
    if (p.dict_keys) {
        const uint64_t key = hname ? hname : 1;
        uint32_t h = (uint32_t)(key ^ (key >> 32)) & p.dict_mask;
        for (uint32_t probe = 0; probe <= p.dict_mask; ++probe) {
            const uint64_t kx = p.dict_keys[h];
            if (kx == key) { id = p.dict_vals[h]; break; }
            if (kx == 0) break;
            h = (h + 1) & p.dict_mask;
        }
    }
    r.kind = (uint8_t)kind;
    r.flags = flags;
    r.service_offset = svc.s == 0xFFFFFFFFu ? 0 : svc.s;
    r.service_length = svc.s == 0xFFFFFFFFu ? 0 : (uint16_t)(svc.e - svc.s);
    r.service_id = id;
    r.elapsed_ms = elapsed;
    r.logid_length = arr[0].e - arr[0].s;
    r.ts_ms = ts;
    r.start_ms = 0;
    r.logid_hash = hl;
    r.name_hash = hname;
    p.rec_sparse[line] = r;
    // identity of the recordCache slot: (logId, service key); never 0 for a recognised line
    uint64_t ident = (hl * 0x9E3779B97F4A7C15ull) ^ hraw;
    p.rawh_sparse[line] = ident ? ident : 1;
}

// ---- ordered compaction of the recognised lines ---------------------------------------------------
__global__ void __launch_bounds__(LP_THREADS) lp_count_records_kernel(const apm_log_record* __restrict__ sparse, uint64_t n,
                                                                      uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * LP_TILE_ITEMS + (uint64_t)threadIdx.x * 4;
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (base + k < n) cnt += sparse[base + k].kind != 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int j = 0; j < LP_THREADS / 32; ++j) t += s_warp[j];
        tile_counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(LP_THREADS) lp_gather_records_kernel(const apm_log_record* __restrict__ sparse,
                                                                       const uint64_t* __restrict__ rawh_sparse, uint64_t n,
                                                                       const uint32_t* __restrict__ tile_base, uint32_t rec_offset,
                                                                       apm_log_record* __restrict__ rec, uint64_t* __restrict__ rawh,
                                                                       uint64_t* __restrict__ words) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * LP_TILE_ITEMS + (uint64_t)threadIdx.x * 4;
    uint32_t cnt = 0;
    bool take[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        take[k] = base + k < n && sparse[base + k].kind != 0;
        cnt += take[k];
    }
    uint32_t pos = rec_offset + tile_base[blockIdx.x] + lp_block_scan(cnt, s_warp);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (take[k]) {
            rec[pos] = sparse[base + k];
            const uint64_t id = rawh_sparse[base + k];
            rawh[pos] = id;
            // pairing word: the top 40 bits of the identity, then the record index (line order)
            words[pos] = (id & ~0xFFFFFFull) | pos;
            ++pos;
        }
}

// ---- pairing (recordCache semantics inside the buffer) ---------------------------------------------
// In the order (identity, line) the record before an exit with the same identity is the cache slot's last
// writer: a Start means the exit finds it (map.get, then map.delete), another exit means the slot is empty.
// A Start is still open at the end of the buffer when nothing with its identity follows it.
__global__ void __launch_bounds__(LP_THREADS) lp_pair_kernel(const uint64_t* __restrict__ words, uint64_t n,
                                                             apm_log_record* __restrict__ rec, const uint64_t* __restrict__ rawh) {
    const uint64_t j = (uint64_t)blockIdx.x * LP_THREADS + threadIdx.x;
    if (j >= n) return;
    const uint64_t w = words[j];
    const uint32_t r = (uint32_t)(w & 0xFFFFFFu);
    apm_log_record& me = rec[r];
    if (me.flags & APM_LR_LOGID_EMPTY) return;  // never enters the cache (:386, :466) / never looks it up (:411, :526)
    const uint64_t id = rawh[r];
    const bool is_exit = me.kind == APM_LK_CT_STOP || me.kind == APM_LK_EJB_EXIT;
    if (is_exit) {
        for (uint64_t i = j; i-- > 0;) {
            if ((words[i] >> 24) != (w >> 24)) break;
            const uint32_t q = (uint32_t)(words[i] & 0xFFFFFFu);
            if (rawh[q] != id || (rec[q].flags & APM_LR_LOGID_EMPTY)) continue;  // a 40-bit neighbour of another identity
            const bool q_exit = rec[q].kind == APM_LK_CT_STOP || rec[q].kind == APM_LK_EJB_EXIT;
            if (!q_exit && !(rec[q].flags & APM_LR_BAD_TIME)) {
                me.start_ms = rec[q].ts_ms;
                me.flags |= APM_LR_PAIRED;
            } else if (!q_exit) {
                me.flags |= APM_LR_PAIRED;  // found, but its startTs does not convert: start falls back to end - elapsed
                me.start_ms = 0;
            }
            break;
        }
    } else {
        bool open = true;
        for (uint64_t i = j + 1; i < n; ++i) {
            if ((words[i] >> 24) != (w >> 24)) break;
            const uint32_t q = (uint32_t)(words[i] & 0xFFFFFFu);
            if (rawh[q] != id || (rec[q].flags & APM_LR_LOGID_EMPTY)) continue;
            open = false;  // consumed by an exit or overwritten by a later Start (map.set)
            break;
        }
        if (open) me.flags |= APM_LR_OPEN_START;
    }
}

// ---- which exits feed the hot path, their start time, the result counters ---------------------------
__global__ void __launch_bounds__(LP_THREADS) lp_finish_kernel(apm_log_record* __restrict__ rec, uint64_t n, uint32_t num_services,
                                                               unsigned long long* __restrict__ counters) {
    const uint64_t i = (uint64_t)blockIdx.x * LP_THREADS + threadIdx.x;
    if (i >= n) return;
    apm_log_record r = rec[i];
    const bool is_exit = r.kind == APM_LK_CT_STOP || r.kind == APM_LK_EJB_EXIT;
    if (!is_exit) return;
    atomicAdd(&counters[0], 1ull);
    const bool bad = r.flags & (APM_LR_BAD_TIME | APM_LR_BAD_ELAPSED);
    // `if (!startMs) startMs = endMs - parseInt(elapsed)` (:279-281)
    if (!bad && (!(r.flags & APM_LR_PAIRED) || r.start_ms == 0)) r.start_ms = r.ts_ms - r.elapsed_ms;
    // an EJB exit with a logId is only output when its entry is in the cache (:416-425); one whose entry may be in
    // an earlier buffer is the host's to resolve.  Every other exit is output (complete, salvaged or need-number).
    const bool emitted = r.kind == APM_LK_CT_STOP || (r.flags & (APM_LR_LOGID_EMPTY | APM_LR_PAIRED));
    if (emitted) {
        if (bad) atomicAdd(&counters[2], 1ull);
        else if (r.service_id == APM_NO_SERVICE || r.service_id >= num_services) atomicAdd(&counters[1], 1ull);
        else r.flags |= APM_LR_FED;
    }
    rec[i] = r;
}

__global__ void __launch_bounds__(LP_THREADS) lp_count_fed_kernel(const apm_log_record* __restrict__ rec, uint64_t n,
                                                                  uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * LP_TILE_ITEMS + (uint64_t)threadIdx.x * 4;
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (base + k < n) cnt += (rec[base + k].flags & APM_LR_FED) != 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int j = 0; j < LP_THREADS / 32; ++j) t += s_warp[j];
        tile_counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(LP_THREADS) lp_gather_fed_kernel(const apm_log_record* __restrict__ rec, uint64_t n,
                                                                   const uint32_t* __restrict__ tile_base, uint32_t jvm_id,
                                                                   apm_tx_columns cols) {
    __shared__ uint32_t s_warp[LP_THREADS / 32];
    const uint64_t base = (uint64_t)blockIdx.x * LP_TILE_ITEMS + (uint64_t)threadIdx.x * 4;
    uint32_t cnt = 0;
    bool take[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        take[k] = base + k < n && (rec[base + k].flags & APM_LR_FED);
        cnt += take[k];
    }
    uint32_t pos = tile_base[blockIdx.x] + lp_block_scan(cnt, s_warp);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (take[k]) {
            const apm_log_record& r = rec[base + k];
            cols.service_id[pos] = r.service_id;
            cols.jvm_id[pos] = jvm_id;
            cols.end_ts_ms[pos] = r.ts_ms;
            cols.elapsed_ms[pos] = r.elapsed_ms;
            ++pos;
        }
}

LogParseState* lp_state(apm_ctx* c) {
    if (!c->lp) c->lp = new LogParseState();
    return static_cast<LogParseState*>(c->lp);
}

}  // namespace

void apm_lp_destroy(apm_ctx* c) {
    if (!c->lp) return;
    LogParseState* s = static_cast<LogParseState*>(c->lp);
    cudaFree(s->d_text); cudaFree(s->tile_counts); cudaFree(s->d_total); cudaFreeHost(s->h_total); cudaFree(s->nlpos);
    cudaFree(s->rec_sparse); cudaFree(s->rawh_sparse); cudaFree(s->rec); cudaFree(s->rawh);
    cudaFree(s->words[0]); cudaFree(s->words[1]); cudaFree(s->sort_hist); cudaFree(s->sort_base);
    cudaFree(s->cols.service_id); cudaFree(s->cols.jvm_id); cudaFree(s->cols.end_ts_ms); cudaFree(s->cols.elapsed_ms);
    cudaFree(s->dict_keys); cudaFree(s->dict_vals);
    for (cudaEvent_t ev : s->ev_sub) cudaEventDestroy(ev);
    delete s;
    c->lp = nullptr;
}

#define LP_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess) return apm_fail(c, APM_E_DEVICE, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

extern "C" {

uint64_t apm_name_hash(const char* s, uint64_t len) {
    uint64_t h = FNV_OFFSET;
    for (uint64_t i = 0; i < len; ++i) h = fnv_step(h, (unsigned char)s[i]);
    return h;
}

apm_status apm_dict_set(apm_ctx* c, const uint64_t* name_hashes, const uint32_t* service_ids, uint32_t n) {
    if (!c || (n && (!name_hashes || !service_ids))) return APM_E_INVALID;
    APM_ENTER(c);
    LogParseState* s = lp_state(c);
    uint32_t size = 16;
    while (size < 2u * n + 2u) size <<= 1;
    std::vector<uint64_t> keys(size, 0);
    std::vector<uint32_t> vals(size, APM_NO_SERVICE);
    for (uint32_t i = 0; i < n; ++i) {
        if (service_ids[i] >= c->cfg.num_services) return apm_fail(c, APM_E_INVALID, "dict_set: service id outside num_services");
        const uint64_t key = name_hashes[i] ? name_hashes[i] : 1;
        uint32_t h = (uint32_t)(key ^ (key >> 32)) & (size - 1);
        while (keys[h] != 0 && keys[h] != key) h = (h + 1) & (size - 1);
        if (keys[h] == key && vals[h] != service_ids[i])
            return apm_fail(c, APM_E_INVALID, "dict_set: two service names share a 64-bit hash");
        keys[h] = key;
        vals[h] = service_ids[i];
    }
    LP_CUDA(cudaStreamSynchronize(c->s_main));
    cudaFree(s->dict_keys); cudaFree(s->dict_vals);
    s->dict_keys = nullptr; s->dict_vals = nullptr;
    LP_CUDA(cudaMalloc((void**)&s->dict_keys, size * sizeof(uint64_t)));
    LP_CUDA(cudaMalloc((void**)&s->dict_vals, size * sizeof(uint32_t)));
    LP_CUDA(cudaMemcpy(s->dict_keys, keys.data(), size * sizeof(uint64_t), cudaMemcpyHostToDevice));
    LP_CUDA(cudaMemcpy(s->dict_vals, vals.data(), size * sizeof(uint32_t), cudaMemcpyHostToDevice));
    s->dict_mask = size - 1;
    return APM_OK;
}

apm_status apm_parse_log_host(apm_ctx* c, const char* text, uint64_t nbytes, uint32_t file_type, uint32_t jvm_id,
                              const apm_tz_rule* tz, uint32_t num_tz, apm_log_record* out, uint64_t cap, uint64_t* n_out,
                              apm_parse_result* result) {
    if (!c || (nbytes && !text) || (num_tz && !tz)) return APM_E_INVALID;
    APM_ENTER(c);
    if (n_out) *n_out = 0;
    if (result) memset(result, 0, sizeof *result);
    if (num_tz > APM_MAX_TZ_RULES) return apm_fail(c, APM_E_INVALID, "parse_log: more than APM_MAX_TZ_RULES zone rules");
    if (file_type > 1) return apm_fail(c, APM_E_INVALID, "parse_log: file_type is 0 (application log) or 1 (server.log)");
    if (jvm_id >= c->cfg.num_jvms) return apm_fail(c, APM_E_INVALID, "parse_log: jvm_id outside num_jvms");
    if (nbytes >= (1ull << 32)) return apm_fail(c, APM_E_CAPACITY, "parse_log: at most 4 GiB of text per call");
    LogParseState* s = lp_state(c);
    s->n_fed = 0;
    if (nbytes == 0) return APM_OK;
    cudaStream_t st = c->s_main;
    if (!s->d_total) {
        LP_CUDA(cudaMalloc((void**)&s->d_total, 16 * sizeof(unsigned long long)));
        LP_CUDA(cudaMallocHost((void**)&s->h_total, 16 * sizeof(unsigned long long)));
    }
    // -- text to the device, in sub-buffers that end after a newline: the copy of the later ones (copy stream)
    //    hides behind the line table / parse / compaction of the earlier ones
    if (nbytes + 16 > s->text_cap) {
        cudaFree(s->d_text);
        s->d_text = nullptr;
        s->text_cap = 0;
        const uint64_t want = nbytes + nbytes / 4 + 4096;
        LP_CUDA(cudaMalloc((void**)&s->d_text, want));
        s->text_cap = want;
    }
    static uint64_t sub_bytes = 0;  // APM_LP_SUB_BYTES overrides the 32 MB default (tests cut small logs into many pieces)
    if (!sub_bytes) { const char* e = getenv("APM_LP_SUB_BYTES"); sub_bytes = e && atoll(e) > 0 ? (uint64_t)atoll(e) : LP_SUB_BYTES; }
    std::vector<uint64_t> cut(1, 0);
    while (cut.back() < nbytes) {
        const uint64_t lo = cut.back();
        uint64_t e = std::min<uint64_t>(lo + sub_bytes, nbytes);
        if (e < nbytes) {
            const void* q = memrchr(text + lo, '\n', e - lo);
            if (q) e = (uint64_t)((const char*)q - text) + 1;
            else {  // one line longer than a sub-buffer: extend to its end
                const void* f = memchr(text + e, '\n', nbytes - e);
                e = f ? (uint64_t)((const char*)f - text) + 1 : nbytes;
            }
        }
        cut.push_back(e);
    }
    const size_t nsub = cut.size() - 1;
    while (s->ev_sub.size() < nsub) {
        cudaEvent_t ev;
        LP_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        s->ev_sub.push_back(ev);
    }
    LP_CUDA(cudaStreamSynchronize(c->s_copy));
    for (size_t i = 0; i < nsub; ++i) {
        LP_CUDA(cudaMemcpyAsync(s->d_text + cut[i], text + cut[i], cut[i + 1] - cut[i], cudaMemcpyHostToDevice, c->s_copy));
        LP_CUDA(cudaEventRecord(s->ev_sub[i], c->s_copy));
    }
    c->counters.h2d_bytes += nbytes;
    auto ensure_tiles = [&](uint64_t tiles) -> cudaError_t {
        if (tiles <= s->tiles_cap) return cudaSuccess;
        cudaFree(s->tile_counts);
        s->tile_counts = nullptr;
        s->tiles_cap = 0;
        const uint64_t want = tiles + tiles / 4 + 64;
        const cudaError_t e = cudaMalloc((void**)&s->tile_counts, want * sizeof(uint32_t));
        if (e == cudaSuccess) s->tiles_cap = want;
        return e;
    };
    // record arrays grow with what has been gathered so far (the first guess is one record per 96 bytes of text)
    auto ensure_records = [&](uint64_t need, uint64_t have) -> apm_status {
        if (need <= s->rec_cap) return APM_OK;
        const uint64_t want = need + need / 2 + 1024;
        apm_log_record* rec = nullptr;
        uint64_t *rawh = nullptr, *w0 = nullptr;
        LP_CUDA(cudaMalloc((void**)&rec, want * sizeof(apm_log_record)));
        LP_CUDA(cudaMalloc((void**)&rawh, want * sizeof(uint64_t)));
        LP_CUDA(cudaMalloc((void**)&w0, want * sizeof(uint64_t)));
        if (have) {
            LP_CUDA(cudaMemcpyAsync(rec, s->rec, have * sizeof(apm_log_record), cudaMemcpyDeviceToDevice, st));
            LP_CUDA(cudaMemcpyAsync(rawh, s->rawh, have * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
            LP_CUDA(cudaMemcpyAsync(w0, s->words[0], have * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
            LP_CUDA(cudaStreamSynchronize(st));
        }
        cudaFree(s->rec); cudaFree(s->rawh); cudaFree(s->words[0]); cudaFree(s->words[1]); cudaFree(s->sort_hist); cudaFree(s->sort_base);
        cudaFree(s->cols.service_id); cudaFree(s->cols.jvm_id); cudaFree(s->cols.end_ts_ms); cudaFree(s->cols.elapsed_ms);
        s->rec = rec; s->rawh = rawh; s->words[0] = w0;
        s->words[1] = nullptr; s->sort_hist = s->sort_base = nullptr;
        s->cols = {nullptr, nullptr, nullptr, nullptr};
        s->rec_cap = 0;  // stays 0 if an allocation below fails: the next call starts over
        s->sort_tiles_cap = apm_radix_sort_tiles(want);
        LP_CUDA(cudaMalloc((void**)&s->words[1], want * sizeof(uint64_t)));
        LP_CUDA(cudaMalloc((void**)&s->sort_hist, 256 * s->sort_tiles_cap * sizeof(uint32_t)));
        LP_CUDA(cudaMalloc((void**)&s->sort_base, 256 * sizeof(uint32_t)));
        LP_CUDA(cudaMalloc((void**)&s->cols.service_id, want * sizeof(uint32_t)));
        LP_CUDA(cudaMalloc((void**)&s->cols.jvm_id, want * sizeof(uint32_t)));
        LP_CUDA(cudaMalloc((void**)&s->cols.end_ts_ms, want * sizeof(int64_t)));
        LP_CUDA(cudaMalloc((void**)&s->cols.elapsed_ms, want * sizeof(int32_t)));
        s->rec_cap = want;
        return APM_OK;
    };
    LP_CUDA(cudaMemsetAsync(s->d_total, 0, 16 * sizeof(unsigned long long), st));
    { apm_status est = ensure_records(nbytes / 96 + 1024, 0); if (est) return est; }
    ParseParams p;
    memset(&p, 0, sizeof p);
    p.text = s->d_text;
    p.file_type = file_type; p.jvm_id = jvm_id; p.num_services = c->cfg.num_services;
    for (uint32_t i = 0; i < num_tz; ++i) p.tz[i] = tz[i];
    p.num_tz = num_tz;
    p.dict_keys = s->dict_keys; p.dict_vals = s->dict_vals; p.dict_mask = s->dict_mask;
    uint64_t nrec = 0, total_lines = 0;
    for (size_t i = 0; i < nsub; ++i) {
        const uint64_t lo = cut[i], hi = cut[i + 1];
        if (hi == lo) continue;
        LP_CUDA(cudaStreamWaitEvent(st, s->ev_sub[i], 0));
        // -- line table of this sub-buffer
        const uint64_t t0 = lo & ~(uint64_t)(LP_TILE_BYTES - 1);
        const uint64_t btiles = (hi - t0 + LP_TILE_BYTES - 1) / LP_TILE_BYTES;
        LP_CUDA(ensure_tiles(btiles));
        lp_count_newlines_kernel<<<(unsigned)btiles, LP_THREADS, 0, st>>>(s->d_text, lo, hi, s->tile_counts);
        lp_scan_kernel<<<1, 1024, 0, st>>>(s->tile_counts, btiles, s->d_total + 0);
        c->launches += 2;
        LP_CUDA(cudaMemcpyAsync(s->h_total, s->d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        LP_CUDA(cudaStreamSynchronize(st));
        const uint64_t nl = s->h_total[0];
        const uint64_t nlines = nl + (text[hi - 1] != '\n' ? 1 : 0);  // only the last sub-buffer can end without a newline
        total_lines += nlines;
        if (nlines > s->lines_cap) {
            cudaFree(s->nlpos); cudaFree(s->rec_sparse); cudaFree(s->rawh_sparse);
            s->nlpos = nullptr; s->rec_sparse = nullptr; s->rawh_sparse = nullptr;
            s->lines_cap = 0;  // stays 0 if an allocation below fails: the next call starts over
            const uint64_t want = nlines + nlines / 4 + 1024;
            LP_CUDA(cudaMalloc((void**)&s->nlpos, want * sizeof(uint64_t)));
            LP_CUDA(cudaMalloc((void**)&s->rec_sparse, want * sizeof(apm_log_record)));
            LP_CUDA(cudaMalloc((void**)&s->rawh_sparse, want * sizeof(uint64_t)));
            s->lines_cap = want;
        }
        if (nl) {
            lp_write_newlines_kernel<<<(unsigned)btiles, LP_THREADS, 0, st>>>(s->d_text, lo, hi, s->tile_counts, s->nlpos);
            c->launches += 1;
        }
        // -- one thread per line
        p.lo = lo; p.hi = hi; p.nlpos = s->nlpos; p.nl = nl; p.nlines = nlines;
        p.rec_sparse = s->rec_sparse; p.rawh_sparse = s->rawh_sparse;
        lp_parse_kernel<<<(unsigned)((nlines + LP_THREADS - 1) / LP_THREADS), LP_THREADS, 0, st>>>(p);
        c->launches += 1;
        // -- its recognised lines, appended in line order
        const uint64_t ltiles = (nlines + LP_TILE_ITEMS - 1) / LP_TILE_ITEMS;
        LP_CUDA(ensure_tiles(ltiles));
        lp_count_records_kernel<<<(unsigned)ltiles, LP_THREADS, 0, st>>>(s->rec_sparse, nlines, s->tile_counts);
        lp_scan_kernel<<<1, 1024, 0, st>>>(s->tile_counts, ltiles, s->d_total + 1);
        c->launches += 2;
        LP_CUDA(cudaMemcpyAsync(s->h_total + 1, s->d_total + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        LP_CUDA(cudaStreamSynchronize(st));
        const uint64_t got = s->h_total[1];
        if (got == 0) continue;
        if (nrec + got > LP_MAX_RECORDS) return apm_fail(c, APM_E_CAPACITY, "parse_log: more than 2^24 CommonTiming lines in one buffer: split it");
        { apm_status est = ensure_records(nrec + got, nrec); if (est) return est; }
        lp_gather_records_kernel<<<(unsigned)ltiles, LP_THREADS, 0, st>>>(s->rec_sparse, s->rawh_sparse, nlines, s->tile_counts,
                                                                          (uint32_t)nrec, s->rec, s->rawh, s->words[0]);
        c->launches += 1;
        nrec += got;
    }
    if (result) { result->lines = total_lines; result->records = nrec; }
    if (nrec == 0) { LP_CUDA(cudaStreamSynchronize(st)); return APM_OK; }
    if (out && cap < nrec) return apm_fail(c, APM_E_CAPACITY, "parse_log: record buffer too small");
    const uint64_t ltiles_all = (nrec + LP_TILE_ITEMS - 1) / LP_TILE_ITEMS;
    LP_CUDA(ensure_tiles(ltiles_all));
    // -- Start / Stop pairing: sort by (identity, line), look at the neighbour
    uint64_t* sorted = s->words[0];
    if (nrec > 1) sorted = apm_radix_sort_words(st, s->words[0], s->words[1], nrec, s->sort_hist, s->sort_base, 32, 32, &c->launches);
    const unsigned rblocks = (unsigned)((nrec + LP_THREADS - 1) / LP_THREADS);
    lp_pair_kernel<<<rblocks, LP_THREADS, 0, st>>>(sorted, nrec, s->rec, s->rawh);
    lp_finish_kernel<<<rblocks, LP_THREADS, 0, st>>>(s->rec, nrec, c->cfg.num_services, s->d_total + 4);
    c->launches += 2;
    // -- the fed exits as transaction columns
    const uint64_t rtiles = (nrec + LP_TILE_ITEMS - 1) / LP_TILE_ITEMS;
    lp_count_fed_kernel<<<(unsigned)rtiles, LP_THREADS, 0, st>>>(s->rec, nrec, s->tile_counts);
    lp_scan_kernel<<<1, 1024, 0, st>>>(s->tile_counts, rtiles, s->d_total + 2);
    lp_gather_fed_kernel<<<(unsigned)rtiles, LP_THREADS, 0, st>>>(s->rec, nrec, s->tile_counts, jvm_id, s->cols);
    c->launches += 3;
    LP_CUDA(cudaGetLastError());
    LP_CUDA(cudaMemcpyAsync(s->h_total, s->d_total, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    if (out) {
        LP_CUDA(cudaMemcpyAsync(out, s->rec, nrec * sizeof(apm_log_record), cudaMemcpyDeviceToHost, st));
        c->counters.d2h_bytes += nrec * sizeof(apm_log_record);
    }
    LP_CUDA(cudaStreamSynchronize(st));
    s->n_fed = s->h_total[2];
    if (n_out) *n_out = out ? nrec : 0;
    if (result) {
        result->exits = s->h_total[4];
        result->unknown_service = s->h_total[5];
        result->malformed = s->h_total[6];
        result->fed = s->n_fed;
    }
    return APM_OK;
}

apm_status apm_feed_parsed(apm_ctx* c, apm_tick_result* result) {
    if (!c) return APM_E_INVALID;
    APM_ENTER(c);
    if (result) memset(result, 0, sizeof *result);
    if (!c->lp) return apm_fail(c, APM_E_STATE, "feed_parsed: nothing was parsed");
    LogParseState* s = static_cast<LogParseState*>(c->lp);
    const uint64_t n = s->n_fed;
    s->n_fed = 0;
    if (n == 0) return APM_OK;
    return apm_process_device(c, &s->cols, n, result);
}

}  // extern "C"

// Loads this file's kernels on the current device (see apm_preload_kernels in capi.cu).
cudaError_t apm_preload_k6_logparse() {
    const void* fns[] = {(const void*)lp_count_fed_kernel, (const void*)lp_count_newlines_kernel, (const void*)lp_count_records_kernel, (const void*)lp_finish_kernel, (const void*)lp_gather_fed_kernel, (const void*)lp_gather_records_kernel, (const void*)lp_pair_kernel, (const void*)lp_parse_kernel, (const void*)lp_scan_kernel, (const void*)lp_write_newlines_kernel};
    cudaError_t first = cudaSuccess;
    for (const void* f : fns) {
        cudaFuncAttributes a;
        const cudaError_t e = apm_preload_one(&a, f);
        if (e != cudaSuccess && first == cudaSuccess) first = e;
    }
    return first;
}
