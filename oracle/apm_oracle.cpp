/*
 * ORACLE - test infrastructure.  Never linked into, called from, or shipped
 * with the product library.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load liboracle.so.
 *
 * CPU restatement (C++17, build with -O2 -ffp-contract=off) of the reference's
 * JavaScript for the hot path.  Each function cites the reference lines it
 * follows (paths relative to /root/reference/).
 *
 * PARITY UNPINNED: the reference has no tests or golden vectors for this path
 * (package.json:6-8) and this image has no JavaScript engine, so the
 * restatement is checked against (i) an independent statement-by-statement
 * Python twin (oracle/pyoracle.py), (ii) the hand-derived known answers of
 * SURVEY.md section 8(c), (iii) the one literal `fs` line the reference holds
 * (stream_process_alerts.js:638) - not against the reference running.
 *
 * Two algorithmic structures give byte-identical results:
 *   fast == 0 : the reference's own structure (sorted insert per element into
 *               the window array, four O(lag) passes per list statistic,
 *               shift() as a memmove) - this is the timed CPU baseline;
 *   fast == 1 : sort once per window, one pass per list statistic, ring
 *               indices instead of shift() - used to check large inputs.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../include/apm_b200.h" /* record layouts only */

namespace {

const double kNaN = std::numeric_limits<double>::quiet_NaN();

/* ------------------------------------------------------------------ */
/* ECMAScript built-ins                                                */
/* ------------------------------------------------------------------ */

/* Number.prototype.toFixed (ECMA-262 21.1.3.3) for 0 <= digits <= 2 and
 * |x| < 1e21: the integer n minimising |n / 10^f - x| on the exact binary
 * value, the larger n on a tie.  Returns n and the sign separately. */
static bool to_fixed_int(double x, int digits, bool* neg, unsigned __int128* n_out) {
    if (std::isnan(x) || std::isinf(x) || std::fabs(x) >= 1e21) return false;
    *neg = x < 0; /* -0 is not < 0: prints without sign, as in JS */
    if (*neg) x = -x;
    int e;
    double fr = std::frexp(x, &e); /* x = fr * 2^e, fr in [0.5,1) */
    uint64_t m = (uint64_t)std::ldexp(fr, 53);
    e -= 53; /* x = m * 2^e exactly */
    unsigned __int128 p = 1;
    for (int i = 0; i < digits; ++i) p *= 10;
    unsigned __int128 v = (unsigned __int128)m * p;
    unsigned __int128 n;
    if (e >= 0) {
        n = v << e; /* < 1e21 * 100 < 2^77 */
    } else if (-e >= 120) {
        n = 0; /* x * 10^f < 2^-60: rounds to 0 */
    } else {
        int s = -e;
        unsigned __int128 q = v >> s;
        unsigned __int128 r = v - (q << s);
        unsigned __int128 half = (unsigned __int128)1 << (s - 1);
        n = q + (r >= half ? 1 : 0); /* tie -> larger n */
    }
    *n_out = n;
    return true;
}

static std::string u128_to_string(unsigned __int128 n) {
    if (n == 0) return "0";
    std::string s;
    while (n) {
        s.push_back(char('0' + (int)(n % 10)));
        n /= 10;
    }
    std::reverse(s.begin(), s.end());
    return s;
}

static std::string js_to_fixed(double x, int digits) {
    if (std::isnan(x)) return "NaN";
    bool neg;
    unsigned __int128 n;
    if (!to_fixed_int(x, digits, &neg, &n)) {
        char buf[64];
        snprintf(buf, sizeof buf, "%.17g", x);
        return buf;
    }
    std::string s = u128_to_string(n);
    if (digits > 0) {
        if ((int)s.size() < digits + 1) s.insert(0, digits + 1 - s.size(), '0');
        s.insert(s.size() - digits, ".");
    }
    return neg ? "-" + s : s;
}

/* nf() (entries.js:65-69): `undefined` for NaN, else parseFloat(num).toFixed(f). */
static std::string nf(double x, int digits) {
    if (std::isnan(x)) return "undefined";
    return js_to_fixed(x, digits);
}

/* parseFloat(nf(x)) - what the next stage sees after the CSV hop
 * (entries.js:58-61,179-180).  strtod is correctly rounded in glibc. */
static double quantize(double x, int digits) {
    if (std::isnan(x)) return kNaN;
    std::string s = js_to_fixed(x, digits);
    return strtod(s.c_str(), nullptr);
}

/* ------------------------------------------------------------------ */
/* util_methods.js:10-142                                              */
/* ------------------------------------------------------------------ */

/* average(), util_methods.js:10-24.  NaN result == `undefined`. */
static double js_average(const double* a, size_t n) {
    size_t cnt = 0;
    double sum = 0;
    for (size_t i = 0; i < n; ++i) {
        double v = a[i];
        if (!std::isnan(v)) { /* value != null && !isNaN(parseFloat(value)) */
            cnt += 1;
            sum = sum + v;
        }
    }
    if (cnt > 0) return sum / (double)cnt;
    return kNaN;
}

/* standardDeviation(), util_methods.js:26-50.  `this.average(squareDiffs)`
 * ignores its argument, so the value is sqrt(average()).  *undefined_out is
 * set when JS would return `undefined` (empty list or mean == 0); a negative
 * mean yields a defined NaN, which the caller treats the same way
 * (stream_calc_z_score.js:80,85).  passes: the reference walks the list four
 * times (:27, :31, :41 plus the caller's own average()); the faithful mode
 * keeps the three of this function. */
static double js_standard_deviation(const double* a, size_t n, bool faithful, bool* undefined_out) {
    double avg = js_average(a, n);
    *undefined_out = false;
    if (std::isnan(avg)) { /* !avg && avg !== 0 */
        *undefined_out = true;
        return kNaN;
    }
    double avg_sq = avg;
    if (faithful) {
        /* the map() pass over the list (:31-39) whose result is discarded */
        volatile double sink = 0;
        for (size_t i = 0; i < n; ++i)
            if (!std::isnan(a[i])) {
                double d = a[i] - avg;
                sink = d * d;
            }
        (void)sink;
        avg_sq = js_average(a, n); /* :41, argument ignored */
    }
    if (avg_sq != 0 && !std::isnan(avg_sq)) return std::sqrt(avg_sq); /* sqrt(<0) = NaN */
    *undefined_out = true;
    return kNaN;
}

/* binarySearch / binaryInsert(duplicate = true), util_methods.js:57-95. */
static void js_binary_insert_dup(std::vector<int32_t>& a, int32_t target) {
    long l = 0, h = (long)a.size() - 1;
    long pos = -1;
    while (l <= h) {
        long m = (l + h) >> 1;
        if (a[m] < target) l = m + 1;
        else if (a[m] > target) h = m - 1;
        else { pos = m; break; }
    }
    if (pos < 0) pos = l;
    a.insert(a.begin() + pos, target); /* splice(i, 0, target) */
}

/* calcPercentile(), util_methods.js:112-142, double index arithmetic as written. */
static double js_calc_percentile(const int32_t* a, size_t n, double percentile) {
    if (n == 0) return kNaN;
    if (percentile == 0) return a[0];
    if (percentile == 100) return a[n - 1];
    double index = ((percentile / 100.0) * (double)n) - 1.0;
    if (n == 1 || std::fmod(index, 1.0) == 0) {
        long i = (long)index; /* parseInt truncates toward zero; n==1 gives -0 -> 0 */
        if (i < 0) i = 0;
        return a[i];
    }
    long i = (long)std::ceil(index);
    if ((size_t)i == n - 1) return a[i];
    return ((double)a[i] + (double)a[i + 1]) / 2;
}

/* processZScoreStats(), stream_calc_z_score.js:66-104.
 * out = {inflNewValue, avg, lb, ub}; NaN == undefined. */
/* EXTENSION (APM_SD_TRUE, not in the reference): what standardDeviation() was evidently meant to
 * compute - the population standard deviation of the valid entries, summed left to right. */
static double true_standard_deviation(const double* a, size_t n, bool* undefined_out) {
    const double avg = js_average(a, n);
    *undefined_out = std::isnan(avg);
    if (*undefined_out) return kNaN;
    size_t cnt = 0;
    double ss = 0;
    for (size_t i = 0; i < n; ++i)
        if (!std::isnan(a[i])) {
            const double d = a[i] - avg;
            ss = ss + d * d;
            cnt++;
        }
    return std::sqrt(ss / (double)cnt);
}

static int js_process_zscore(long lag, double threshold, double influence, double new_value,
                             const double* list, size_t len, bool faithful, double out[4], int sd_mode = APM_SD_REFERENCE) {
    double infl = new_value, avg = kNaN, sd = kNaN, lb = kNaN, ub = kNaN;
    int signal = 0;
    if ((long)len >= lag) {
        avg = js_average(list, len);
        bool sd_undef;
        sd = sd_mode == APM_SD_TRUE ? true_standard_deviation(list, len, &sd_undef)
                                    : js_standard_deviation(list, len, faithful, &sd_undef);
        bool avg_ok = !std::isnan(avg);           /* (avg || avg == 0)       */
        bool sd_ok = !sd_undef && !std::isnan(sd); /* (stdDev || stdDev == 0) */
        if (avg_ok && sd_ok) {
            lb = avg - threshold * sd;
            ub = avg + threshold * sd;
        }
        if (!avg_ok || !sd_ok) {
            signal = 0;
        } else if (std::isnan(new_value)) { /* !newValue && newValue != 0 */
            signal = 0;
        } else if (std::fabs(new_value - avg) > threshold * sd) {
            signal = new_value > avg ? 1 : -1;
            double last = list[len - 1];
            if (!std::isnan(last)) infl = influence * new_value + (1 - influence) * last;
        } else {
            signal = 0;
        }
    }
    out[0] = infl; out[1] = avg; out[2] = lb; out[3] = ub;
    return signal;
}

/* processFSEntry(), stream_process_alerts.js:348-434: the deterministic counter core (the
 * wall-clock cooldown and e-mail that follow, :436-468, are host business).  `en` holds what
 * the fs line carries after the toFixed hop; *count is recentAlertCounts[server][service][lag].
 * Returns triggerAlert, causes as APM_CAUSE_* bits. */
static bool js_process_fs_entry(const apm_alert_rules& R, bool lag_suppressed, bool service_suppressed, double hard_max_override,
                                const apm_fs_record& en, uint32_t* count, uint32_t* causes_out) {
    bool trigger = false, incremented = false;
    uint32_t causes = 0;
    const long window = R.window, required = R.required;
    auto alert = [&](uint32_t cause) {
        if (!incremented) {
            if ((long)*count <= window) (*count)++;
            incremented = true;
        }
        if (window && window > 1 && required && required > 1) {
            if ((long)*count >= required) { trigger = true; causes |= cause; }
        } else {
            trigger = true;
            causes |= cause;
        }
    };
    if (!lag_suppressed) {
        if (!service_suppressed) {
            /* (serviceOverrides.hardMaxMsAlertThreshold && != 0) ? override : default, :396 */
            const double hard_max = (hard_max_override > 0) ? hard_max_override : R.hard_max_ms;
            const double average = en.val[0], per75 = en.val[1], tpm = en.tpm;
            if (average > hard_max) alert(APM_CAUSE_AVG_HARD_MAX);
            if (per75 > hard_max) alert(APM_CAUSE_P75_HARD_MAX);
            int both = 0;
            if (en.signal[0] > 0 && average > R.hard_min_ms && tpm > R.hard_min_tpm) {
                if (!R.alert_on_both_only) alert(APM_CAUSE_AVG_UB); else both++;
            }
            if (en.signal[1] > 0 && per75 > R.hard_min_ms && tpm > R.hard_min_tpm) {
                if (!R.alert_on_both_only) alert(APM_CAUSE_P75_UB); else both++;
            }
            if (R.alert_on_both_only && both >= 2) alert(APM_CAUSE_BOTH_UB);
        }
    }
    if (!incremented && *count > 0) (*count)--;
    *causes_out = causes;
    return trigger;
}

/* ------------------------------------------------------------------ */
/* The two stages                                                      */
/* ------------------------------------------------------------------ */

struct LagState { /* one entry of lags{} in stream_calc_z_score.js:205-229 */
    double thr, infl;
    uint32_t alert_count = 0; /* recentAlertCounts[server][service][lag], stream_process_alerts.js:362 */
    std::vector<double> list[APM_NUM_SERIES]; /* faithful: front = oldest          */
    size_t head[APM_NUM_SERIES] = {0, 0, 0};  /* fast: ring start once len == lag  */
};

struct KeyState {
    uint32_t key;
    std::map<int64_t, std::vector<int32_t>> buckets; /* integer-like keys enumerate ascending */
    std::vector<LagState> lags;
    bool z_init = false;
};

struct Pipeline {
    apm_config cfg;
    bool fast;
    int64_t latest = 0; /* latestBucket, :42 */
    std::vector<uint8_t> filter; /* optional: 1 = track this key */
    /* servers in first-seen order, services in first-seen order inside (:159-160) */
    std::vector<uint32_t> server_order;
    std::unordered_map<uint32_t, std::vector<uint32_t>> services_of; /* jvm -> key ids in first-seen order */
    std::unordered_map<uint32_t, KeyState> keys;
    std::vector<double> thr_override, infl_override; /* [lag][key], NaN = default */
    std::unordered_map<uint64_t, std::vector<double>> preload; /* (key,lag,series) -> list */
    apm_alert_rules rules = {};
    std::unordered_map<uint32_t, std::pair<double, bool>> alert_key; /* key -> (hardMax override, suppressed) */
    std::vector<apm_stat_record> stats;
    std::vector<apm_fs_record> fs;
    uint64_t ticks = 0, ingested = 0, invalid = 0;
    bool keep_outputs = true;

    uint32_t K() const { return cfg.num_services * cfg.num_jvms; }

    void z_process(const apm_stat_record& st, KeyState& ks);
    void tick_emit(int64_t edge_ts);
    void consume(uint32_t svc, uint32_t jvm, int64_t ts, int32_t el);
};

/* processData(), stream_calc_z_score.js:195-311, fed with what the CSV hop
 * delivers: parseFloat of the toFixed strings (entries.js:58-73). */
void Pipeline::z_process(const apm_stat_record& st, KeyState& ks) {
    const int nl = cfg.num_lags;
    if (!ks.z_init) { /* first sight: updateServiceSettings, :201-208 */
        ks.z_init = true;
        ks.lags.resize(nl);
        for (int l = 0; l < nl; ++l) {
            double t = cfg.default_threshold[l], f = cfg.default_influence[l];
            if (!thr_override.empty()) {
                double to = thr_override[(size_t)l * K() + ks.key], fo = infl_override[(size_t)l * K() + ks.key];
                if (!std::isnan(to)) t = to;
                if (!std::isnan(fo)) f = fo;
            }
            ks.lags[l].thr = t;
            ks.lags[l].infl = f;
            for (int s = 0; s < APM_NUM_SERIES; ++s) {
                auto it = preload.find(((uint64_t)ks.key << 8) | (uint64_t)(l * 4 + s));
                if (it != preload.end()) ks.lags[l].list[s] = it->second;
            }
        }
    }
    double val[APM_NUM_SERIES];
    val[0] = quantize(st.avg, 1);
    for (int s = 1; s < APM_NUM_SERIES; ++s) {
        int pi = cfg.z_series_pct[s - 1];
        val[s] = (pi >= 0 && pi < cfg.num_percentiles) ? quantize(st.pct[pi], 1) : kNaN;
    }
    double tpm_q = quantize(st.tpm, 2);
    for (int l = 0; l < nl; ++l) { /* Object.entries(lags): ascending lag, :217 */
        LagState& L = ks.lags[l];
        const long lag = cfg.lags[l];
        apm_fs_record r;
        memset(&r, 0, sizeof r);
        r.tick_ts = st.tick_ts; r.key = st.key; r.lag = (uint32_t)lag; r.tpm = tpm_q; r.flags = 1;
        double infl_v[APM_NUM_SERIES];
        for (int s = 0; s < APM_NUM_SERIES; ++s) {
            std::vector<double>& lst = L.list[s];
            double o[4];
            int sig;
            if (!fast || lst.size() < (size_t)lag) {
                sig = js_process_zscore(lag, L.thr, L.infl, val[s], lst.data(), lst.size(), !fast, o, cfg.sd_mode);
            } else {
                /* ring: rotate view so that index 0 is the oldest; summation order is unchanged */
                size_t h = L.head[s];
                static thread_local std::vector<double> tmp;
                tmp.resize(lst.size());
                std::copy(lst.begin() + h, lst.end(), tmp.begin());
                std::copy(lst.begin(), lst.begin() + h, tmp.begin() + (lst.size() - h));
                sig = js_process_zscore(lag, L.thr, L.infl, val[s], tmp.data(), tmp.size(), false, o, cfg.sd_mode);
            }
            infl_v[s] = o[0];
            r.val[s] = val[s]; r.avg[s] = o[1]; r.lb[s] = o[2]; r.ub[s] = o[3];
            r.signal[s] = (int8_t)sig;
        }
        for (int s = 0; s < APM_NUM_SERIES; ++s) { /* shift() then push(), :285-306 */
            std::vector<double>& lst = L.list[s];
            if (lst.size() >= (size_t)lag) {
                if (!fast) {
                    lst.erase(lst.begin());
                    lst.push_back(infl_v[s]);
                } else {
                    lst[L.head[s]] = infl_v[s];
                    L.head[s] = (L.head[s] + 1) % lst.size();
                }
            } else {
                lst.push_back(infl_v[s]);
            }
        }
        if (rules.enabled) { /* the alert stage reads the fs line next (stream_process_alerts.js:625) */
            auto ak = alert_key.find(ks.key);
            uint32_t causes = 0;
            const bool trig = js_process_fs_entry(rules, (rules.suppressed_lag_mask >> l) & 1u, ak != alert_key.end() && ak->second.second,
                                                  ak != alert_key.end() ? ak->second.first : 0.0, r, &L.alert_count, &causes);
            r.flags = (uint8_t)(1u | (trig ? 2u : 0u));
            r.pad_[0] = (uint8_t)causes;
            r.pad_[1] = (uint8_t)(L.alert_count > 255u ? 255u : L.alert_count);
        }
        if (keep_outputs) fs.push_back(r);
    }
}

/* generateAllStatsToQueue(), stream_calc_stats.js:157-203. */
void Pipeline::tick_emit(int64_t edge_ts) {
    const int64_t keep = (int64_t)cfg.window_buckets + cfg.buffer_buckets; /* NUM_KEEP_INTERVALS, :233 */
    std::vector<int32_t> sorted_vals;
    for (uint32_t jvm : server_order) {
        for (uint32_t key : services_of[jvm]) {
            KeyState& ks = keys[key];
            uint64_t cnt = 0;
            double rt_total = 0; /* JS number; exact while < 2^53 */
            unsigned __int128 sum_i = 0, sumsq = 0;
            int32_t mn = 0, mx = 0;
            bool have_minmax = false;
            sorted_vals.clear();
            for (auto& kv : ks.buckets) {
                if (kv.first >= latest - keep && kv.first <= latest - cfg.buffer_buckets) {
                    cnt += kv.second.size();
                    for (int32_t el : kv.second) {
                        rt_total += el;
                        sum_i += (uint64_t)el;
                        sumsq += (unsigned __int128)((uint64_t)el * (uint64_t)el);
                    }
                    if (cfg.num_percentiles > 0) {
                        if (!fast) for (int32_t el : kv.second) js_binary_insert_dup(sorted_vals, el); /* binaryConcat, :176 */
                        else sorted_vals.insert(sorted_vals.end(), kv.second.begin(), kv.second.end());
                    }
                    for (int32_t el : kv.second) { /* extension fields */
                        if (!have_minmax) { mn = mx = el; have_minmax = true; }
                        mn = std::min(mn, el);
                        mx = std::max(mx, el);
                    }
                }
            }
            if (fast && cfg.num_percentiles > 0) std::sort(sorted_vals.begin(), sorted_vals.end());
            apm_stat_record r;
            memset(&r, 0, sizeof r);
            r.tick_ts = edge_ts; r.key = key; r.flags = 1 | (cnt ? 2 : 0);
            r.count = cnt; r.sum = (uint64_t)sum_i;
            r.avg = kNaN; r.stddev = kNaN;
            for (int p = 0; p < APM_MAX_PCT; ++p) r.pct[p] = kNaN;
            if (cnt != 0) {
                r.avg = rt_total / (double)cnt; /* :181 */
                for (int p = 0; p < cfg.num_percentiles; ++p)
                    r.pct[p] = js_calc_percentile(sorted_vals.data(), sorted_vals.size(), cfg.percentiles[p]);
                r.min = mn; r.max = mx;
                /* extension: population stddev from exact integer moments */
                unsigned __int128 num = (unsigned __int128)cnt * sumsq - sum_i * sum_i;
                double numd = (double)(uint64_t)(num >> 64) * 18446744073709551616.0 + (double)(uint64_t)num;
                r.stddev = std::sqrt(numd) / (double)cnt;
            }
            r.tpm = (double)cnt / ((double)cfg.window_buckets * ((double)cfg.bucket_ms / 1000.0) / 60.0); /* :186 */
            if (keep_outputs && (cfg.flags & APM_F_KEEP_STATS)) stats.push_back(r);
            if ((cfg.flags & APM_F_ZSCORE) && cfg.num_lags > 0) z_process(r, ks);
        }
    }
}

/* consumeMsg(), stream_calc_stats.js:331-371. */
void Pipeline::consume(uint32_t svc, uint32_t jvm, int64_t ts, int32_t el) {
    if (svc >= cfg.num_services || jvm >= cfg.num_jvms || el < 0 || ts < 0) { invalid++; return; }
    const int64_t label = ts / cfg.bucket_ms; /* generateBucketLabel, :89-91 */
    const int64_t keep = (int64_t)cfg.window_buckets + cfg.buffer_buckets;
    if (label > latest) { /* :348 */
        latest = label;
        for (auto& kv : keys) { /* removeOldBuckets, :103-113 */
            auto& b = kv.second.buckets;
            while (!b.empty() && b.begin()->first < latest - keep) b.erase(b.begin());
        }
        int64_t edge_ts = (latest - cfg.buffer_buckets - 1) * cfg.bucket_ms; /* :356 */
        tick_emit(edge_ts);
        ticks++;
    }
    uint32_t key = jvm * cfg.num_services + svc;
    if (!filter.empty() && !filter[key]) return;
    auto it = keys.find(key); /* addData, :115-134 */
    if (it == keys.end()) {
        if (services_of.find(jvm) == services_of.end()) server_order.push_back(jvm);
        services_of[jvm].push_back(key);
        it = keys.emplace(key, KeyState()).first;
        it->second.key = key;
    }
    it->second.buckets[label].push_back(el);
    ingested++;
}

} /* namespace */

extern "C" {

int orc_to_fixed(double x, int digits, char* out, int cap) {
    std::string s = js_to_fixed(x, digits);
    snprintf(out, cap, "%s", s.c_str());
    return (int)s.size();
}
double orc_quantize(double x, int digits) { return quantize(x, digits); }
double orc_average(const double* a, uint64_t n) { return js_average(a, n); }
double orc_standard_deviation(const double* a, uint64_t n, int* undefined_out) {
    bool u;
    double r = js_standard_deviation(a, n, true, &u);
    *undefined_out = u;
    return r;
}
double orc_calc_percentile(const int32_t* sorted, uint64_t n, double p) { return js_calc_percentile(sorted, n, p); }
void orc_sorted_insert(int32_t* a, uint64_t n, int32_t v) { /* a has room for n+1 */
    std::vector<int32_t> t(a, a + n);
    js_binary_insert_dup(t, v);
    memcpy(a, t.data(), (n + 1) * sizeof(int32_t));
}
int orc_process_zscore(int64_t lag, double thr, double infl, double new_value, const double* list, uint64_t len,
                       double out[4]) {
    return js_process_zscore(lag, thr, infl, new_value, list, len, true, out);
}

void* orc_new(const apm_config* cfg, int fast) {
    Pipeline* p = new Pipeline();
    p->cfg = *cfg;
    p->fast = fast != 0;
    return p;
}
void orc_free(void* h) { delete (Pipeline*)h; }
void orc_set_filter(void* h, const uint8_t* keep) {
    Pipeline* p = (Pipeline*)h;
    p->filter.assign(keep, keep + p->K());
}
void orc_keep_outputs(void* h, int keep) { ((Pipeline*)h)->keep_outputs = keep != 0; }
void orc_set_key_params(void* h, uint32_t key, int lag_idx, double thr, double infl) {
    Pipeline* p = (Pipeline*)h;
    if (p->thr_override.empty()) {
        p->thr_override.assign((size_t)APM_MAX_LAGS * p->K(), kNaN);
        p->infl_override.assign((size_t)APM_MAX_LAGS * p->K(), kNaN);
    }
    p->thr_override[(size_t)lag_idx * p->K() + key] = thr;
    p->infl_override[(size_t)lag_idx * p->K() + key] = infl;
    auto it = p->keys.find(key);
    if (it != p->keys.end() && it->second.z_init) { it->second.lags[lag_idx].thr = thr; it->second.lags[lag_idx].infl = infl; }
}
void orc_ring_load(void* h, uint32_t key, int lag_idx, int series, const double* v, uint32_t len) {
    Pipeline* p = (Pipeline*)h;
    p->preload[((uint64_t)key << 8) | (uint64_t)(lag_idx * 4 + series)] = std::vector<double>(v, v + len);
}
void orc_set_alert_rules(void* h, const apm_alert_rules* r) { ((Pipeline*)h)->rules = *r; }
void orc_set_alert_key_params(void* h, uint32_t key, double hard_max, int suppressed) {
    ((Pipeline*)h)->alert_key[key] = std::make_pair(hard_max, suppressed != 0);
}
int orc_process_fs_entry(const apm_alert_rules* r, int lag_suppressed, int service_suppressed, double hard_max_override,
                         const apm_fs_record* en, uint32_t* count, uint32_t* causes) {
    return js_process_fs_entry(*r, lag_suppressed != 0, service_suppressed != 0, hard_max_override, *en, count, causes) ? 1 : 0;
}
void orc_consume(void* h, const uint32_t* svc, const uint32_t* jvm, const int64_t* ts, const int32_t* el, uint64_t n) {
    Pipeline* p = (Pipeline*)h;
    for (uint64_t i = 0; i < n; ++i) p->consume(svc[i], jvm[i], ts[i], el[i]);
}
uint64_t orc_num_stats(void* h) { return ((Pipeline*)h)->stats.size(); }
uint64_t orc_num_fs(void* h) { return ((Pipeline*)h)->fs.size(); }
void orc_copy_stats(void* h, apm_stat_record* out) {
    Pipeline* p = (Pipeline*)h;
    if (!p->stats.empty()) memcpy(out, p->stats.data(), p->stats.size() * sizeof(apm_stat_record));
}
void orc_copy_fs(void* h, apm_fs_record* out) {
    Pipeline* p = (Pipeline*)h;
    if (!p->fs.empty()) memcpy(out, p->fs.data(), p->fs.size() * sizeof(apm_fs_record));
}
void orc_clear_outputs(void* h) {
    Pipeline* p = (Pipeline*)h;
    p->stats.clear();
    p->fs.clear();
}
uint64_t orc_ticks(void* h) { return ((Pipeline*)h)->ticks; }
int64_t orc_latest(void* h) { return ((Pipeline*)h)->latest; }
uint64_t orc_ingested(void* h) { return ((Pipeline*)h)->ingested; }
uint32_t orc_ring_read(void* h, uint32_t key, int lag_idx, int series, double* out, uint32_t cap) {
    Pipeline* p = (Pipeline*)h;
    auto it = p->keys.find(key);
    if (it == p->keys.end() || !it->second.z_init) return 0;
    LagState& L = it->second.lags[lag_idx];
    std::vector<double>& lst = L.list[series];
    size_t n = lst.size(), hd = (p->fast && n >= (size_t)p->cfg.lags[lag_idx]) ? L.head[series] : 0;
    for (size_t i = 0; i < n && i < cap; ++i) out[i] = lst[(hd + i) % n];
    return (uint32_t)n;
}

/* StatEntry.toCSVString(), entries.js:71-73. */
int orc_format_st(const apm_stat_record* r, const char* server, const char* service, int pa, int pb, char* out, int cap) {
    std::string s = "st|" + std::to_string(r->tick_ts) + "|" + server + "|" + service + "|" + nf(r->tpm, 2) + "|" +
                    nf(r->avg, 1) + "|" + nf(pa >= 0 ? r->pct[pa] : kNaN, 1) + "|" + nf(pb >= 0 ? r->pct[pb] : kNaN, 1);
    snprintf(out, cap, "%s", s.c_str());
    return (int)s.size();
}
/* FullStatEntry.toCSVString(), entries.js:116-118: averageSignal is printed raw,
 * the other two signals go through nf() -> toFixed(1). */
int orc_format_fs(const apm_fs_record* r, const char* server, const char* service, char* out, int cap) {
    std::string s = "fs|" + std::to_string(r->tick_ts) + "|" + server + "|" + service + "|" + std::to_string(r->lag) +
                    "|" + nf(r->tpm, 2);
    for (int i = 0; i < APM_NUM_SERIES; ++i) {
        s += "|" + nf(r->val[i], 1) + ":" + nf(r->avg[i], 1) + ":" + nf(r->lb[i], 1) + ":" + nf(r->ub[i], 1) + ":";
        s += (i == 0) ? std::to_string((int)r->signal[i]) : nf((double)r->signal[i], 1);
    }
    snprintf(out, cap, "%s", s.c_str());
    return (int)s.size();
}

} /* extern "C" */
