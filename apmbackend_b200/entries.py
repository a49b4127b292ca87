"""Wire / record types of the hot path, mirroring entries.js: TxEntry (1-21), StatEntry
(52-84), FullStatEntry (86-152), EntryFactory (154-198).  Same field names, same
pipe-delimited CSV, same rounding: `nf` is Number.prototype.toFixed on the exact binary
value (round half up), `undefined` for NaN."""
from __future__ import annotations

import math

NaN = float("nan")


def to_fixed(x: float, digits: int) -> str:
    """Number.prototype.toFixed for |x| < 1e21: the integer n closest to x * 10^digits on the
    exact value of the double, the larger n on a tie; a negative x formats -x behind '-'."""
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if abs(x) >= 1e21:
        return repr(x)
    sign = "-" if x < 0 else ""
    num, den = abs(x).as_integer_ratio()  # exact
    scaled = num * 10 ** digits
    n = (2 * scaled + den) // (2 * den)   # floor(x * 10^d + 1/2)
    s = str(n)
    if digits:
        s = s.rjust(digits + 1, "0")
        s = s[:-digits] + "." + s[-digits:]
    return sign + s


def parse_float(v) -> float:
    """parseFloat for what travels on these queues: numbers, decimal strings, 'undefined'."""
    if v is None:
        return NaN
    if isinstance(v, (int, float)):
        f = float(v)
        return 0.0 if f == 0 else f
    try:
        return float(str(v).strip())
    except ValueError:
        return NaN


def parse_int(v) -> float:
    """parseInt: integer prefix, NaN when there is none (kept as float NaN like JS)."""
    if isinstance(v, int):
        return v
    if isinstance(v, float):
        return NaN if math.isnan(v) or math.isinf(v) else int(v)
    s = str(v).strip()
    i, sign = 0, 1
    if s[:1] and s[:1] in "+-":
        sign = -1 if s[0] == "-" else 1
        s = s[1:]
    while i < len(s) and s[i].isdigit():
        i += 1
    return sign * int(s[:i]) if i else NaN


def _fmt(v) -> str:
    """Template-literal interpolation of a number / string / undefined."""
    if v is None:
        return "undefined"
    if isinstance(v, float):
        if math.isnan(v):
            return "NaN"
        if v == int(v) and abs(v) < 1e21:
            return str(int(v))
        return repr(v)
    return str(v)


class TxEntry:
    """entries.js:1-21."""

    def __init__(self, server, service, logId, acctNum, startTs, endTs, elapsed, topLevel):
        self.server = server
        self.service = service
        self.logId = logId
        self.acctNum = parse_int(acctNum)
        self.startTs = parse_int(startTs)
        self.endTs = parse_int(endTs)
        self.elapsed = parse_int(elapsed)
        self.topLevel = topLevel
        self.type = "tx"

    def toCSVString(self) -> str:
        return (f"{self.type}|{self.server}|{self.service}|{self.logId}|{_fmt(self.acctNum)}|{_fmt(self.startTs)}|"
                f"{_fmt(self.endTs)}|{_fmt(self.elapsed)}|{self.topLevel}")

    def __lt__(self, other):  # heap tie-break only
        return False


class BaseEntry:
    """entries.js:45-50."""

    def __init__(self, timestamp):
        self.timestamp = parse_int(timestamp)
        self.type = "bs"


class StatEntry(BaseEntry):
    """entries.js:52-84."""

    def __init__(self, timestamp, server, service, tpm, average, per75, per95):
        super().__init__(timestamp)
        self.server = server
        self.service = service
        self.tpm = parse_float(tpm)
        self.average = parse_float(average)
        self.per75 = parse_float(per75)
        self.per95 = parse_float(per95)
        self.type = "st"

    @staticmethod
    def nf(num, fixedVal=1):
        """:65-69 - `undefined` for NaN / undefined, else parseFloat(num).toFixed(fixedVal)."""
        if num is None or (isinstance(num, float) and math.isnan(num)):
            return None
        return to_fixed(parse_float(num), fixedVal)

    def toCSVString(self) -> str:
        nf = self.nf
        return (f"{self.type}|{_fmt(self.timestamp)}|{self.server}|{self.service}|{_fmt(nf(self.tpm, 2))}|"
                f"{_fmt(nf(self.average))}|{_fmt(nf(self.per75))}|{_fmt(nf(self.per95))}")


class FullStatEntry(StatEntry):
    """entries.js:86-152."""

    def __init__(self, timestamp, server, service, tpm, lag, average, averageAvg, averageLB, averageUB, averageSignal,
                 per75, per75Avg, per75LB, per75UB, per75Signal, per95, per95Avg, per95LB, per95UB, per95Signal):
        super().__init__(timestamp, server, service, tpm, average, per75, per95)
        self.lag = lag
        self.averageAvg, self.averageLB, self.averageUB = map(parse_float, (averageAvg, averageLB, averageUB))
        self.averageSignal = parse_int(averageSignal)
        self.per75Avg, self.per75LB, self.per75UB = map(parse_float, (per75Avg, per75LB, per75UB))
        self.per75Signal = parse_int(per75Signal)
        self.per95Avg, self.per95LB, self.per95UB = map(parse_float, (per95Avg, per95LB, per95UB))
        self.per95Signal = parse_int(per95Signal)
        self.type = "fs"

    def toCSVString(self) -> str:
        f = lambda v: _fmt(self.nf(v))  # noqa: E731
        return (f"{self.type}|{_fmt(self.timestamp)}|{self.server}|{self.service}|{self.lag}|{_fmt(self.nf(self.tpm, 2))}|"
                f"{f(self.average)}:{f(self.averageAvg)}:{f(self.averageLB)}:{f(self.averageUB)}:{_fmt(self.averageSignal)}|"
                f"{f(self.per75)}:{f(self.per75Avg)}:{f(self.per75LB)}:{f(self.per75UB)}:{f(self.per75Signal)}|"
                f"{f(self.per95)}:{f(self.per95Avg)}:{f(self.per95LB)}:{f(self.per95UB)}:{f(self.per95Signal)}")


class EntryFactory:
    """entries.js:154-198 (the tx / st / fs branches; al and jx never enter the hot path)."""

    def getEntryFromObject(self, obj):
        if obj and (obj.get("timestamp") or obj.get("startTs")):
            if obj.get("type") == "tx":
                return TxEntry(obj.get("server"), obj.get("service"), obj.get("logId"), obj.get("acctNum"),
                               obj.get("startTs"), obj.get("endTs"), obj.get("elapsed"), obj.get("topLevel"))
        return None

    def getEntryFromCSV(self, line, delim="|"):
        arr = str(line).split(delim)
        if arr[0] == "tx":
            return TxEntry(*arr[1:9])
        if arr[0] == "st":
            return StatEntry(*arr[1:8])
        if arr[0] == "fs":
            a, p75, p95 = arr[6].split(":"), arr[7].split(":"), arr[8].split(":")
            return FullStatEntry(arr[1], arr[2], arr[3], arr[5], arr[4], *a, *p75, *p95)
        return None
