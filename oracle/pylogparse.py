"""TEST INFRASTRUCTURE ONLY - CPU restatement of the CommonTiming part of the reference's log parser
(SURVEY 8 f-3).  Only tests/ may import this; the product path is apmbackend_b200/csrc/k6_logparse.cu.

Restates, statement by statement, from /root/reference/stream_parse_transactions.js:
  readLine dispatch (SERVER / APP file types)            :733-791
  parseEjbCommonTimingEntry / parseEjbCommonTimingExit   :378-446
  parseCommonTimingEntry / parseCommonTimingExit         :451-565
  convertStringDateToMs                                  :242-256
  outputRecord (Provider[ rewrite, startMs fallback)     :264-293
and TxEntry (entries.js:1-21) for the emitted line.

Parity status: UNPINNED (the reference ships no tests or fixtures for its parser and no JavaScript
engine exists in this environment); the known answers in tests/test_logparse_cpu.py are derived by
hand from the JavaScript above.

Out of scope here as in the product: SOAP logs, parseAppLine (audit trail / account numbers), the
TTL of the node-cache objects (120 s record cache, 30 s need-number cache): `flush()` stands for
"every need-number entry expired".  With no account-number source an exit line therefore goes
through needNumRecordCache and is emitted at flush() with acctNum '' (-> "NaN" in the tx line).
"""
import math
import re

_WS = re.compile(r"[\s]+")
_EJB_ENTRY = re.compile(r"INFO *\[CommonTiming] The EJB")
_EJB_EXIT = re.compile(r"INFO *\[CommonTiming] Total time")
_CT_ENTRY = re.compile(r"INFO *CommonTiming::Start")
_CT_EXIT = re.compile(r"INFO *CommonTiming::Stop")
_PROVIDER = re.compile(r"Provider\[", re.I)
_TOP = re.compile(r"^S:")
_BAF = re.compile(r"\[[^ ]+] +INFO ")


def js_parse_int(s):
    """parseInt(string): leading whitespace, optional sign, decimal digits; NaN when there are none."""
    m = re.match(r"\s*([+-]?\d+)", str(s))
    return int(m.group(1)) if m else math.nan


def js_str(v):
    if isinstance(v, float) and math.isnan(v):
        return "NaN"
    return str(v)


def days_from_civil(y, m, d):
    y -= m <= 2
    era = (y if y >= 0 else y - 399) // 400
    yoe = y - era * 400
    doy = (153 * (m + (-3 if m > 2 else 9)) + 2) // 5 + d - 1
    doe = yoe * 365 + yoe // 4 - yoe // 100 + doy
    return era * 146097 + doe - 719468


class TimeZone:
    """Local time -> epoch ms the way `new Date(y, m, d, H, M, S, ms).getTime()` does for a zone whose
    UTC offset changes at known local instants: offset_ms applies from local_from_ms (a local wall-clock
    value expressed as if it were UTC) on.  A single entry is a fixed offset."""

    def __init__(self, transitions=((-(1 << 62), 0),)):
        self.transitions = sorted(transitions)

    def to_epoch_ms(self, local_ms):
        off = self.transitions[0][1]
        for frm, o in self.transitions:
            if local_ms >= frm:
                off = o
        return local_ms - off


def convert_string_date_to_ms(date_str, tz):
    """:242-256, the standard-log branch (`2020-01-07 10:00:02,669`)."""
    if not date_str:
        return ""
    arr = re.split(r"-|[\s]+|:|,", date_str.strip())
    try:
        y, mo, d, H, M, S, ms = (int(x) for x in arr[:7])
    except ValueError:
        return math.nan
    # `new Date(y, mo - 1, d, H, M, S, ms)` normalises out-of-range fields by carrying
    local = days_from_civil(y + (mo - 1) // 12, (mo - 1) % 12 + 1, 1) * 86400000
    local += (d - 1) * 86400000 + H * 3600000 + M * 60000 + S * 1000 + ms
    return tz.to_epoch_ms(local)


class TxLine:
    def __init__(self, server, service, log_id, acct, start_ts, end_ts, elapsed, top):
        self.server, self.service, self.logId = server, service, log_id
        self.acctNum = js_parse_int(acct)
        self.startTs = js_parse_int(start_ts)
        self.endTs = js_parse_int(end_ts)
        self.elapsed = js_parse_int(elapsed)
        self.topLevel = top

    def to_csv(self):
        return "|".join(["tx", js_str(self.server), self.service, self.logId, js_str(self.acctNum), js_str(self.startTs),
                         js_str(self.endTs), js_str(self.elapsed), self.topLevel])


class LogParserOracle:
    """The caches and the four CommonTiming parsers; `out` collects the tx lines in emission order."""

    def __init__(self, tz=None):
        self.tz = tz or TimeZone()
        self.recordCache = {}          # logId -> {service -> partial record}
        self.needNumRecordCache = {}   # logId -> {service -> record waiting for an account number}
        self.acctCache = {}
        self.out = []

    # ---- :264-293
    def outputRecord(self, server, service, log_id, acct, start_ts, end_ts, elapsed):
        start_ms = convert_string_date_to_ms(start_ts, self.tz)
        end_ms = convert_string_date_to_ms(end_ts, self.tz)
        service = _PROVIDER.sub("Provider:", service, count=1).replace("]", "", 1)
        if not start_ms or (isinstance(start_ms, float) and math.isnan(start_ms)):
            e = js_parse_int(elapsed)
            start_ms = math.nan if (isinstance(e, float) or isinstance(end_ms, float) or end_ms == "") else end_ms - e
        self.out.append(TxLine(server, service, log_id, acct, start_ms, end_ms, elapsed,
                               "Y" if _TOP.search(service) else "N").to_csv())

    # ---- :378-401
    def parseEjbCommonTimingEntry(self, line, server):
        arr = _WS.split(line)
        log_id = re.sub(r"[\[\]]", "", arr[0])
        start_ts = f"{_at(arr, 1)} {_at(arr, 2)}"
        if log_id == "":
            return
        service = f"S:{_at(arr, 13)}"
        self.recordCache.setdefault(log_id, {})[service] = {"server": server, "startTs": start_ts}

    # ---- :403-446
    def parseEjbCommonTimingExit(self, line, server):
        arr = _WS.split(line)
        log_id = re.sub(r"[\[\]]", "", arr[0])
        end_ts = f"{_at(arr, 1)} {_at(arr, 2)}"
        service = f"S:{_at(arr, 9)}"
        elapsed = _at(arr, 11)
        if log_id == "":
            self.outputRecord(server, service, "", "", "", end_ts, elapsed)
            return
        m = self.recordCache.get(log_id)
        if m is None:
            return                                   # logged as an error, nothing emitted (:418-419)
        part = m.get(service)
        if part is None:
            return                                   # (:423-424)
        acct = self.acctCache.get(log_id)
        if acct:
            self.outputRecord(server, service, log_id, acct, part["startTs"], end_ts, elapsed)
        else:
            self.needNumRecordCache.setdefault(log_id, {})[service] = dict(part, endTs=end_ts, elapsed=elapsed)
        del m[service]

    # ---- :451-481
    def parseCommonTimingEntry(self, line, server):
        arr = _WS.split(line)
        log_id = re.sub(r"[\[\]]", "", arr[0])
        start_ts = f"{_at(arr, 1)} {_at(arr, 2)}"
        second = _WS.split(line.split("INFO")[1].strip())
        service = _at(second, 1)
        if log_id == "":
            return
        self.recordCache.setdefault(log_id, {})[service] = {"server": server, "startTs": start_ts}

    # ---- :498-503 + :483-496: without a BAF account number this is an incomplete record
    def salvageRecordAndOutput(self, line, log_id, arr, server, service, end_ts, elapsed):
        acct = self.attemptReadAccountNumberFromBAFInfo(line, log_id, arr)
        self.outputRecord(server, service, "", acct, "", end_ts, elapsed)

    def attemptReadAccountNumberFromBAFInfo(self, line, log_id, arr):
        acct = ""
        if _BAF.search(line):
            info = re.sub(r"[\[\]]", "", re.sub(r".*]\[", "", _at(arr, 3))).split(":")
            acct = info[-1]
            if acct and re.match(r"^[0-9]+$", acct.strip()) and log_id:
                self.acctCache[log_id] = acct.strip()                       # saveAcctNum('bafmetainfo') :294-311
                need = self.needNumRecordCache.get(log_id)
                if need:
                    for svc, rec in list(need.items()):                      # (:313-323)
                        self.outputRecord(rec["server"], svc, log_id, acct.strip(), rec["startTs"], rec["endTs"], rec["elapsed"])
                        del need[svc]
        return acct

    # ---- :505-565
    def parseCommonTimingExit(self, line, server):
        arr = _WS.split(line)
        second = _WS.split(line.split("INFO")[1].strip())
        log_id = re.sub(r"[\[\]]", "", arr[0])
        end_ts = f"{_at(arr, 1)} {_at(arr, 2)}"
        service = _at(second, 1)
        elapsed = _at(second, 5)
        m = self.recordCache.get(log_id)
        if log_id == "":
            self.salvageRecordAndOutput(line, log_id, arr, server, service, end_ts, elapsed)
            return
        if m is None:
            self.salvageRecordAndOutput(line, log_id, arr, server, service, end_ts, elapsed)
            return
        part = m.get(service)
        if part is None:
            self.salvageRecordAndOutput(line, log_id, arr, server, service, end_ts, elapsed)
            return
        acct = self.acctCache.get(log_id)
        if acct:
            self.outputRecord(server, service, log_id, acct, part["startTs"], end_ts, elapsed)
        else:
            alt = self.attemptReadAccountNumberFromBAFInfo(line, log_id, arr)
            self.needNumRecordCache.setdefault(log_id, {})[service] = dict(part, endTs=end_ts, elapsed=elapsed, altAcctNum=alt)
        del m[service]

    # ---- :741-777 (the SOAP branch and parseAppLine are out of scope)
    def readLine(self, curr_log_fp, line):
        if not line:
            return
        curr_log = curr_log_fp.split("/")[-1]
        parts = curr_log_fp.split("/")
        server = parts[2] if len(parts) > 2 else "undefined"
        if "soap_io" in curr_log:
            return
        if re.search(r"server\.log", curr_log):
            if _EJB_ENTRY.search(line):
                self.parseEjbCommonTimingEntry(line, server)
            elif _EJB_EXIT.search(line):
                self.parseEjbCommonTimingExit(line, server)
            elif _CT_ENTRY.search(line):
                self.parseCommonTimingEntry(line, server)
            elif _CT_EXIT.search(line):
                self.parseCommonTimingExit(line, server)
        else:
            if _CT_ENTRY.search(line):
                self.parseCommonTimingEntry(line, server)
            elif _CT_EXIT.search(line):
                self.parseCommonTimingExit(line, server)

    # every need-number entry expires (:227-241): emitted with the BAF account number or ''
    def flush(self):
        for log_id, need in self.needNumRecordCache.items():
            for svc, rec in need.items():
                self.outputRecord(rec["server"], svc, log_id, rec.get("altAcctNum") or "", rec["startTs"], rec["endTs"], rec["elapsed"])
        self.needNumRecordCache = {}


def _at(arr, i):
    """JavaScript array indexing: out of range is `undefined`, which template strings print as such."""
    return arr[i] if 0 <= i < len(arr) else "undefined"
