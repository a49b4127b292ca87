"""
ORACLE (test infrastructure, never shipped, never on the product path).

Pure-Python twin of the reference's JavaScript for the hot path, written to
mirror the JS *statement by statement* (ECMAScript semantics restated by
hand), so it can cross-check the C++ restatement in apm_oracle.cpp.  Small
cases only (pure-Python loops).

PARITY UNPINNED: the reference ships no tests, fixtures or golden vectors for
this path (package.json:6-8), and no JavaScript engine exists in this image,
so neither this twin nor the C++ oracle could be checked against the
reference running.  They are pinned against each other, against the
hand-derived known-answer vectors of SURVEY.md section 8(c), and against the
one literal `fs` line the reference holds (stream_process_alerts.js:638).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.

Reference citations are relative to /root/reference/.
"""
from __future__ import annotations

import math
from fractions import Fraction

NaN = float("nan")
UNDEF = None  # JS `undefined`


# --------------------------------------------------------------------------
# ECMAScript built-ins the path leans on
# --------------------------------------------------------------------------
def js_to_fixed(x: float, digits: int) -> str:
    """Number.prototype.toFixed (ECMA-262 21.1.3.3): pick the integer n that
    makes n / 10^f - x as close to zero as possible, the larger n on a tie,
    evaluated on the EXACT binary value of x; negative x formats -x and
    prefixes '-'."""
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if abs(x) >= 1e21:
        return repr(x)
    sign = ""
    if x < 0:
        sign = "-"
        x = -x
    scaled = Fraction(x) * (10 ** digits)
    n = math.floor(scaled + Fraction(1, 2))  # ties -> larger n
    s = str(n)
    if digits:
        s = s.rjust(digits + 1, "0")
        s = s[:-digits] + "." + s[-digits:]
    return sign + s


def js_parse_float(v) -> float:
    """parseFloat of a number, `undefined`, or one of the strings this path
    produces (decimal literals or 'undefined')."""
    if v is None:
        return NaN
    if isinstance(v, (int, float)):
        f = float(v)
        return 0.0 if f == 0 else f  # Number->String drops the sign of -0
    s = str(v).strip()
    try:
        return float(s)
    except ValueError:
        return NaN


def _truthy_num(v) -> bool:
    """JS truthiness of a number-or-undefined."""
    return v is not None and not math.isnan(v) and v != 0


def nf(num, fixed: int = 1):
    """StatEntry.nf / FullStatEntry.nf (entries.js:65-69, 110-114)."""
    if (not _truthy_num(num)) and not (num is not None and num == 0):
        return UNDEF
    return js_to_fixed(js_parse_float(num), fixed)


def _s(v) -> str:
    """Template-literal interpolation for the few types on the path."""
    if v is None:
        return "undefined"
    if isinstance(v, float):
        if math.isnan(v):
            return "NaN"
        if v == int(v) and abs(v) < 1e21:
            return str(int(v))
        return repr(v)
    return str(v)


# --------------------------------------------------------------------------
# util_methods.js:10-142 (Array.prototype patches)
# --------------------------------------------------------------------------
def average(lst):
    """util_methods.js:10-24.  Left-to-right sum of the valid elements from 0."""
    cnt = 0
    total = 0
    for value in lst:
        if value is not None and not math.isnan(js_parse_float(value)):
            cnt += 1
            total = total + js_parse_float(value)
    if cnt > 0:
        return total / cnt
    return UNDEF


def standard_deviation(lst):
    """util_methods.js:26-50.  `this.average(squareDiffs)` ignores its argument
    (average() takes none and reduces over `this`), so avgSquareDiff == avg and
    the result is sqrt(avg)."""
    avg = average(lst)
    if (not _truthy_num(avg)) and not (avg is not None and avg == 0):
        return UNDEF
    avg_square_diff = average(lst)  # the argument is ignored in the reference
    if _truthy_num(avg_square_diff) and avg_square_diff != 0:
        return math.sqrt(avg_square_diff) if avg_square_diff >= 0 else NaN
    return UNDEF


def binary_search(arr, target):
    """util_methods.js:57-76."""
    lo, hi = 0, len(arr) - 1
    while lo <= hi:
        m = (lo + hi) >> 1
        if arr[m] < target:
            lo = m + 1
        elif arr[m] > target:
            hi = m - 1
        else:
            return m
    return ~lo


def binary_insert(arr, target, duplicate=False):
    """util_methods.js:84-95."""
    i = binary_search(arr, target)
    if i >= 0:
        if not duplicate:
            return i
    else:
        i = ~i
    arr.insert(i, target)
    return i


def binary_concat(arr, source, duplicate=False):
    """util_methods.js:102-106."""
    for el in source:
        binary_insert(arr, el, duplicate)


def calc_percentile(arr, percentile):
    """util_methods.js:112-142 (double arithmetic kept as written)."""
    n = len(arr)
    if n == 0:
        return UNDEF
    if percentile == 0:
        return arr[0]
    if percentile == 100:
        return arr[n - 1]
    index = ((percentile / 100.0) * n) - 1.0
    if n == 1 or math.fmod(index, 1) == 0:
        return arr[int(index)]  # parseInt truncates toward zero
    index = int(math.ceil(index))
    if index == n - 1:
        return arr[index]
    return (arr[index] + arr[index + 1]) / 2


# --------------------------------------------------------------------------
# entries.js
# --------------------------------------------------------------------------
class StatEntry:
    """entries.js:52-73."""

    def __init__(self, timestamp, server, service, tpm, average_, per75, per95):
        self.timestamp = int(timestamp)
        self.server = server
        self.service = service
        self.tpm = js_parse_float(tpm)
        self.average = js_parse_float(average_)
        self.per75 = js_parse_float(per75)
        self.per95 = js_parse_float(per95)
        self.type = "st"

    def to_csv(self) -> str:
        return (f"{self.type}|{self.timestamp}|{self.server}|{self.service}|"
                f"{_s(nf(self.tpm, 2))}|{_s(nf(self.average))}|"
                f"{_s(nf(self.per75))}|{_s(nf(self.per95))}")

    @staticmethod
    def from_csv(line: str) -> "StatEntry":
        a = line.split("|")
        assert a[0] == "st"
        return StatEntry(a[1], a[2], a[3], a[4], a[5], a[6], a[7])


def _parse_int_signal(v) -> int:
    return int(v)


class FullStatEntry(StatEntry):
    """entries.js:86-118."""

    def __init__(self, timestamp, server, service, tpm, lag, average_, a_avg, a_lb, a_ub, a_sig,
                 per75, p75_avg, p75_lb, p75_ub, p75_sig, per95, p95_avg, p95_lb, p95_ub, p95_sig):
        super().__init__(timestamp, server, service, tpm, average_, per75, per95)
        self.lag = lag
        self.averageAvg, self.averageLB, self.averageUB = map(js_parse_float, (a_avg, a_lb, a_ub))
        self.averageSignal = _parse_int_signal(a_sig)
        self.per75Avg, self.per75LB, self.per75UB = map(js_parse_float, (p75_avg, p75_lb, p75_ub))
        self.per75Signal = _parse_int_signal(p75_sig)
        self.per95Avg, self.per95LB, self.per95UB = map(js_parse_float, (p95_avg, p95_lb, p95_ub))
        self.per95Signal = _parse_int_signal(p95_sig)
        self.type = "fs"

    def to_csv(self) -> str:
        f = lambda v: _s(nf(v))  # noqa: E731
        return (f"{self.type}|{self.timestamp}|{self.server}|{self.service}|{self.lag}|{_s(nf(self.tpm, 2))}|"
                f"{f(self.average)}:{f(self.averageAvg)}:{f(self.averageLB)}:{f(self.averageUB)}:{self.averageSignal}|"
                f"{f(self.per75)}:{f(self.per75Avg)}:{f(self.per75LB)}:{f(self.per75UB)}:{f(float(self.per75Signal))}|"
                f"{f(self.per95)}:{f(self.per95Avg)}:{f(self.per95LB)}:{f(self.per95UB)}:{f(float(self.per95Signal))}")


# --------------------------------------------------------------------------
# stream_calc_stats.js: class StatParser (28-204) + consumeMsg (331-371)
# --------------------------------------------------------------------------
class StatParser:
    """Generalised by (bucket_ms, window, buffer, percentiles); the reference
    is the point (10000, 30, 6, (75, 95)).  `percentiles` beyond the first two
    are an extension and are reported only through `records`."""

    def __init__(self, bucket_ms=10000, window=30, buffer=6, percentiles=(75, 95)):
        self.servers = {}
        self.latestBucket = 0
        self.bucket_ms = bucket_ms
        self.window = window
        self.buffer = buffer
        self.num_keep = window + buffer  # stream_calc_stats.js:233
        self.percentiles = tuple(percentiles)
        self.out = []       # `stats` queue lines
        self.records = []   # (ts, server, service, cnt, tpm, avg, [pcts], min, max) un-rounded

    def generateBucketLabel(self, ms):
        """:89-91.  slice(0,-4) of the decimal string == floor(ms / 10000)."""
        if self.bucket_ms == 10000:
            return str(int(ms))[:-4]
        return str(int(ms) // self.bucket_ms)

    def generateTsFromBucketLabel(self, lab):
        """:98-101."""
        if self.bucket_ms == 10000:
            return int(f"{int(lab)}0000")
        return int(lab) * self.bucket_ms

    def _label_gt_latest(self, label):
        """:348.  Number compare while latestBucket is the initial 0, string
        compare afterwards."""
        if isinstance(self.latestBucket, int):
            return (int(label) if label != "" else 0) > self.latestBucket
        return label > self.latestBucket

    def removeOldBuckets(self, num_keep):
        """:103-113."""
        edge = int(self.latestBucket) - num_keep
        for srv in self.servers.values():
            for svc in srv["services"].values():
                for lab in [l for l in svc["buckets"] if int(l) < edge]:
                    del svc["buckets"][lab]

    def addData(self, server, service, elapsed, label):
        """:115-134 (heap push left to the host)."""
        srv = self.servers.setdefault(server, {"services": {}})
        svc = srv["services"].setdefault(service, {"buckets": {}})
        svc["buckets"].setdefault(label, []).append(int(elapsed))

    def generateAllStatsToQueue(self, ts):
        """:157-203."""
        latest = int(self.latestBucket)
        for server, srv in self.servers.items():
            for service, svc in srv["services"].items():
                cnt = 0
                total = 0
                sorted_vals = []
                mn, mx, sumsq = None, None, 0
                for lab in sorted(svc["buckets"], key=int):  # integer-like keys enumerate ascending
                    if latest - self.num_keep <= int(lab) <= latest - self.buffer:
                        arr = svc["buckets"][lab]
                        cnt += len(arr)
                        for el in arr:
                            total += el
                            sumsq += el * el
                        binary_concat(sorted_vals, arr, True)
                avg = UNDEF
                pcts = [UNDEF] * len(self.percentiles)
                if cnt != 0:
                    avg = total / cnt
                    pcts = [calc_percentile(sorted_vals, p) for p in self.percentiles]
                    mn, mx = sorted_vals[0], sorted_vals[-1]
                tpm = cnt / (self.window * (self.bucket_ms / 1000) / 60.0)
                st = StatEntry(ts, server, service, tpm, avg, pcts[0], pcts[1] if len(pcts) > 1 else UNDEF)
                self.out.append(st.to_csv())
                self.records.append((ts, server, service, cnt, tpm, avg, pcts, mn, mx, total, sumsq))

    def consume(self, server, service, end_ts, elapsed):
        """consumeMsg, :331-371: tick BEFORE the triggering transaction is added."""
        label = self.generateBucketLabel(end_ts)
        if self._label_gt_latest(label):
            self.latestBucket = label
            self.removeOldBuckets(self.num_keep)
            edge_ts = self.generateTsFromBucketLabel(int(self.latestBucket) - self.buffer - 1)
            self.generateAllStatsToQueue(edge_ts)
        self.addData(server, service, elapsed, label)


# --------------------------------------------------------------------------
# stream_calc_z_score.js: class ZScoreParser (26-312)
# --------------------------------------------------------------------------
class ZScoreParser:
    def __init__(self, defaults=None, overrides=None):
        # config/apm_config.json:126-172
        self.defaults = defaults if defaults is not None else [
            {"LAG": 360, "THRESHOLD": 20.0, "INFLUENCE": 0.1},
            {"LAG": 8640, "THRESHOLD": 15.0, "INFLUENCE": 0.0},
        ]
        self.overrides = overrides if overrides is not None else {}
        self.servers = {}

    def processZScoreStats(self, lag, threshold, influence, new_value, prev):
        """:66-104."""
        infl = new_value
        avg = sd = lb = ub = UNDEF
        signal = 0
        if len(prev) >= int(lag):
            avg = average(prev)
            sd = standard_deviation(prev)
            if (_truthy_num(avg) or (avg is not None and avg == 0)) and \
               (_truthy_num(sd) or (sd is not None and sd == 0)):
                lb = avg - threshold * sd
                ub = avg + threshold * sd
            if avg is None or math.isnan(avg) or sd is None or math.isnan(sd):
                signal = 0
            elif (not _truthy_num(new_value)) and not (new_value is not None and new_value == 0):
                signal = 0
            elif abs(new_value - avg) > threshold * sd:
                signal = 1 if new_value > avg else -1
                last = prev[-1]
                if _truthy_num(last) or (last is not None and last == 0):
                    infl = influence * new_value + (1 - influence) * last
            else:
                signal = 0
        return [infl, avg, lb, ub, signal]

    def getServiceSettingsFromConfig(self, service):
        """:106-132.  `settings` ALIASES the defaults array: overrides are
        written into the defaults and stay there for later services."""
        settings = self.defaults
        if service in self.overrides:
            ov = self.overrides[service]
            for el in settings:
                for lag in ov:
                    if str(el["LAG"]) == str(lag):
                        if ov[lag].get("THRESHOLD"):
                            el["THRESHOLD"] = ov[lag]["THRESHOLD"]
                        if ov[lag].get("INFLUENCE"):
                            el["INFLUENCE"] = ov[lag]["INFLUENCE"]
        return settings

    def updateServiceSettings(self, server, service):
        """:134-150."""
        lags = self.servers[server]["services"][service]["lags"]
        for el in self.getServiceSettingsFromConfig(service):
            k = str(el["LAG"])
            if k in lags:
                lags[k]["THRESHOLD"] = el["THRESHOLD"]
                lags[k]["INFLUENCE"] = el["INFLUENCE"]
            else:
                lags[k] = {"THRESHOLD": el["THRESHOLD"], "INFLUENCE": el["INFLUENCE"]}

    def processData(self, st: StatEntry):
        """:195-311."""
        srv = self.servers.setdefault(st.server, {"services": {}})
        if st.service not in srv["services"]:
            srv["services"][st.service] = {"lags": {}}
            self.updateServiceSettings(st.server, st.service)
        out = []
        lags = srv["services"][st.service]["lags"]
        for lag in sorted(lags, key=int):  # integer-like keys enumerate ascending
            lo = lags[lag]
            thr, inf = lo["THRESHOLD"], lo["INFLUENCE"]
            for name in ("avgList", "per75List", "per95List"):
                lo.setdefault(name, [])
            a, p75, p95 = js_parse_float(st.average), js_parse_float(st.per75), js_parse_float(st.per95)
            ra = self.processZScoreStats(lag, thr, inf, a, lo["avgList"])
            r75 = self.processZScoreStats(lag, thr, inf, p75, lo["per75List"])
            r95 = self.processZScoreStats(lag, thr, inf, p95, lo["per95List"])
            for name in ("avgList", "per75List", "per95List"):
                if len(lo[name]) >= int(lag):
                    lo[name].pop(0)
            out.append(FullStatEntry(st.timestamp, st.server, st.service, st.tpm, lag,
                                     st.average, ra[1], ra[2], ra[3], ra[4],
                                     st.per75, r75[1], r75[2], r75[3], r75[4],
                                     st.per95, r95[1], r95[2], r95[3], r95[4]))
            lo["avgList"].append(ra[0])
            lo["per75List"].append(r75[0])
            lo["per95List"].append(r95[0])
        return out


def run_pipeline(txs, bucket_ms=10000, window=30, buffer=6, percentiles=(75, 95),
                 defaults=None, overrides=None):
    """txs: iterable of (server, service, end_ts, elapsed).  Returns (st lines, fs lines)
    with the stats -> z-score hop done through the CSV strings, as the queues do."""
    sp = StatParser(bucket_ms, window, buffer, percentiles)
    zp = ZScoreParser(defaults, overrides)
    for server, service, end_ts, elapsed in txs:
        sp.consume(server, service, end_ts, elapsed)
    fs = []
    for line in sp.out:
        for e in zp.processData(StatEntry.from_csv(line)):
            fs.append(e.to_csv())
    return sp.out, fs, sp, zp


# --------------------------------------------------------------------------
# stream_process_alerts.js: processFSEntry (348-434), the deterministic counter core
# --------------------------------------------------------------------------
CAUSES = ("average exceeded hard ms threshold", "per75 exceeded hard ms threshold", "average UB exceeded",
          "per75 UB exceeded", "average and per75 UB exceeded")


class AlertsCounter:
    """ALERTSCONFIG is the dict of config/apm_config.json:176-214 the function reads."""

    def __init__(self, config=None):
        self.cfg = {"hardMinMsAlertThreshold": 200, "hardMaxMsAlertThreshold": 10000, "hardMinTpmAlertThreshold": 1.0,
                    "alertOnBothOnly": True, "overrides": {"services": {}}, "suppressedLags": [],
                    "rollingAlertWindowSizeInIntervals": 60, "requiredNumberBadIntervalsInAlertWindowToTrigger": 45,
                    "suppressedServices": ["S:getSalesServiceOffers"]}
        if config:
            self.cfg.update(config)
        self.recentAlertCounts = {}

    def processFSEntry(self, en: FullStatEntry):
        """Returns (triggerAlert, causes list); the cooldown / AlertEntry part (:436-468) is left out."""
        cfg = self.cfg
        trigger, incremented, causes = False, False, []
        ov = cfg["overrides"]["services"].get(en.service)
        counts = self.recentAlertCounts.setdefault(en.server, {}).setdefault(en.service, {})
        if not counts.get(en.lag):
            counts[en.lag] = 0
        window = int(cfg["rollingAlertWindowSizeInIntervals"])
        required = int(cfg["requiredNumberBadIntervalsInAlertWindowToTrigger"])

        def alert(s):
            nonlocal trigger, incremented
            if not incremented:
                if counts[en.lag] <= window:
                    counts[en.lag] += 1
                incremented = True
            if window and window > 1 and required and required > 1:
                if counts[en.lag] >= required:
                    trigger = True
                    causes.append(s)
            else:
                trigger = True
                causes.append(s)

        if int(en.lag) not in cfg["suppressedLags"]:
            if en.service not in cfg["suppressedServices"]:
                hard_max = ov["hardMaxMsAlertThreshold"] if (ov and ov.get("hardMaxMsAlertThreshold")) \
                    else cfg["hardMaxMsAlertThreshold"]
                if en.average > hard_max:
                    alert(CAUSES[0])
                if en.per75 > hard_max:
                    alert(CAUSES[1])
                both = 0
                if en.averageSignal > 0 and en.average > cfg["hardMinMsAlertThreshold"] and en.tpm > cfg["hardMinTpmAlertThreshold"]:
                    if not cfg["alertOnBothOnly"]:
                        alert(CAUSES[2])
                    else:
                        both += 1
                if en.per75Signal > 0 and en.per75 > cfg["hardMinMsAlertThreshold"] and en.tpm > cfg["hardMinTpmAlertThreshold"]:
                    if not cfg["alertOnBothOnly"]:
                        alert(CAUSES[3])
                    else:
                        both += 1
                if cfg["alertOnBothOnly"] and both >= 2:
                    alert(CAUSES[4])
        if not incremented and counts[en.lag] > 0:
            counts[en.lag] -= 1
        if counts[en.lag] < 0:
            counts[en.lag] = 0
        return trigger, causes
