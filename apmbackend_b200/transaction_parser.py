"""Host mirror of the CommonTiming part of stream_parse_transactions.js on top of the
device parser (csrc/k6_logparse.cu, `apm_parse_log_host` / `apm_feed_parsed`).

The device does the per-line work for a whole buffer of log text - readLine's dispatch (:733-777), the
four CommonTiming parsers (:378-565), convertStringDateToMs (:242-256), outputRecord's rewrite and
start-time fallback (:264-293), Start/Stop pairing inside the buffer - and feeds the complete exits
straight into the stats / z-score path.  This class keeps what the reference keeps in host state: the
server and service dictionaries, Start lines whose Stop has not arrived yet (recordCache), records waiting
for an account number (needNumRecordCache) and the account numbers found in BAF meta info (acctCache),
and it prints the `tx|...` lines outputRecord would have queued.  SOAP logs and parseAppLine are not
handled (no account-number source besides the BAF info), node-cache's wall-clock TTLs become explicit
calls: `expireNeedNumRecords()` and `expireRecordCache()`.
"""
from __future__ import annotations

import re

import numpy as np

from . import _capi
from .entries import TxEntry

_PROVIDER = re.compile(r"Provider\[", re.I)
_WS = re.compile(r"[\s]+")
_EXITS = (_capi.LK_CT_STOP, _capi.LK_EJB_EXIT)


def rewrite_service(service: str) -> str:
    """outputRecord (:274): `Provider[x]` -> `Provider:x`."""
    return _PROVIDER.sub("Provider:", service, count=1).replace("]", "", 1)


class TransactionParser:
    def __init__(self, stat_parser, tz_rules=((-(1 << 62), 0),)):
        self.sp = stat_parser
        self.engine = stat_parser.engine
        self.tz_rules = tuple(tz_rules)
        self.recordCache = {}          # logId -> {service key -> {"server", "startMs"}}: Starts still open
        self.needNumRecordCache = {}   # logId -> {service key -> record}
        self.acctCache = {}
        self._dict = {}                # rewritten service name -> service id (what the device was given)
        self.lastParseResult = None

    # ---- dictionary upkeep: names the device did not know are registered the way addData would (:117-126)
    def _sync_dict(self, server, names):
        changed = False
        for name in names:
            if name not in self._dict:
                sid, _ = self.sp._ids(server, name)
                self._dict[name] = sid
                changed = True
        if changed:
            self.engine.dict_set(self._dict)

    @staticmethod
    def _service_key(text, rec):
        """The recordCache key: the raw token, `S:`-prefixed for the EJB forms (:391, :407)."""
        o = int(rec["line_offset"]) + int(rec["service_offset"])
        raw = text[o:o + int(rec["service_length"])].decode("utf-8", "replace") if rec["service_length"] else "undefined"
        return ("S:" + raw) if rec["kind"] in (_capi.LK_EJB_ENTRY, _capi.LK_EJB_EXIT) else raw

    @staticmethod
    def _log_id(text, rec):
        o = int(rec["line_offset"])
        return text[o:o + int(rec["logid_length"])].decode("utf-8", "replace").replace("[", "").replace("]", "")

    def outputRecord(self, out, server, service, logId, acctNum, startMs, endMs, elapsed):
        service = rewrite_service(service)
        out.append(TxEntry(server, service, logId, acctNum, startMs, endMs, elapsed,
                           "Y" if service.startswith("S:") else "N").toCSVString())

    def _baf_account(self, line, logId, out):
        """attemptReadAccountNumberFromBAFInfo (:483-496) + saveAcctNum('bafmetainfo') (:294-323)."""
        arr = _WS.split(line)
        tok = arr[3] if len(arr) > 3 else "undefined"
        info = re.sub(r"[\[\]]", "", re.sub(r".*]\[", "", tok)).split(":")
        acct = info[-1]
        if acct and re.match(r"^[0-9]+$", acct.strip()) and logId:
            self.acctCache[logId] = acct.strip()
            need = self.needNumRecordCache.get(logId)
            if need:
                for svc, r in list(need.items()):
                    self.outputRecord(out, r["server"], svc, logId, acct.strip(), r["startMs"], r["endMs"], r["elapsed"])
                    del need[svc]
        return acct

    # ---- readLine for a whole buffer of one file (:741-777)
    def readLines(self, currLogFp, text: bytes, feed=True):
        """Returns (tx lines in the reference's emission order, TickResult of the feed or None)."""
        currLog = currLogFp.split("/")[-1]
        parts = currLogFp.split("/")
        server = parts[2] if len(parts) > 2 else "undefined"
        if "soap_io" in currLog:
            return [], None
        file_type = 1 if re.search(r"server\.log", currLog) else 0
        jvm = self.sp._server(server)["id"]
        recs, res = self.engine.parse_log(text, file_type, jvm, self.tz_rules)
        exits = recs[np.isin(recs["kind"], _EXITS)]
        unknown = exits[exits["service_id"] == _capi.NO_SERVICE]
        if len(unknown):
            self._sync_dict(server, {rewrite_service(self._service_key(text, r)) for r in unknown})
            recs, res = self.engine.parse_log(text, file_type, jvm, self.tz_rules)
        self.lastParseResult = res
        out, late = [], []
        for r in recs:
            kind, flags = int(r["kind"]), int(r["flags"])
            key = self._service_key(text, r)
            logId = "" if flags & _capi.LR_LOGID_EMPTY else self._log_id(text, r)
            bad_time = bool(flags & _capi.LR_BAD_TIME)
            if kind not in _EXITS:
                if not flags & _capi.LR_LOGID_EMPTY:     # map.set: whatever an earlier buffer left under this key is gone
                    m = self.recordCache.setdefault(logId, {})
                    m.pop(key, None)
                    if flags & _capi.LR_OPEN_START:      # still in recordCache when the buffer ends
                        m[key] = {"server": server, "startMs": "" if bad_time else int(r["ts_ms"])}
                continue
            endMs = float("nan") if bad_time else int(r["ts_ms"])
            elapsed = float("nan") if flags & _capi.LR_BAD_ELAPSED else int(r["elapsed_ms"])
            fallback = float("nan") if (bad_time or flags & _capi.LR_BAD_ELAPSED) else endMs - elapsed
            line = None
            if flags & _capi.LR_BAF_INFO:
                o = int(r["line_offset"])
                line = text[o:o + int(r["line_length"])].decode("utf-8", "replace")
            if flags & _capi.LR_LOGID_EMPTY:
                acct = self._baf_account(line, logId, out) if (line is not None and kind == _capi.LK_CT_STOP) else ""
                self.outputRecord(out, server, key, "", acct, fallback, endMs, elapsed)
                continue
            part = None
            if flags & _capi.LR_PAIRED:
                part = {"server": server, "startMs": int(r["start_ms"]) if int(r["start_ms"]) else fallback}
            else:                                        # its Start may be in an earlier buffer
                m = self.recordCache.get(logId)
                if m is not None and key in m:
                    part = m.pop(key)
                    if not part["startMs"]:
                        part["startMs"] = fallback
                    if kind == _capi.LK_EJB_EXIT and not (flags & _capi.LR_FED):
                        late.append((self._dict.get(rewrite_service(key)), jvm, endMs, elapsed, rewrite_service(key)))
            if part is None:
                if kind == _capi.LK_CT_STOP:             # salvageRecordAndOutput (:498-503)
                    acct = self._baf_account(line, logId, out) if line is not None else ""
                    self.outputRecord(out, server, key, "", acct, fallback, endMs, elapsed)
                continue                                 # EJB: error logged, nothing emitted (:418-424)
            acct = self.acctCache.get(logId)
            if acct:
                self.outputRecord(out, server, key, logId, acct, part["startMs"], endMs, elapsed)
            else:
                alt = self._baf_account(line, logId, out) if (line is not None and kind == _capi.LK_CT_STOP) else ""
                self.needNumRecordCache.setdefault(logId, {})[key] = {
                    "server": part["server"], "startMs": part["startMs"], "endMs": endMs, "elapsed": elapsed, "altAcctNum": alt}
        tick = None
        if feed:
            tick = self.engine.feed_parsed()
            known = [x for x in late if x[0] is not None and x[2] == x[2] and x[3] == x[3]]
            if known:                                    # EJB exits whose entry was in an earlier buffer
                a = np.array([(x[0], x[1], x[2], x[3]) for x in known], np.int64)
                t2 = self.engine.process_host(a[:, 0].astype(np.uint32), a[:, 1].astype(np.uint32), a[:, 2], a[:, 3].astype(np.int32))
                tick.ticks += t2.ticks
                tick.tx_ingested += t2.tx_ingested
                tick.latest_bucket = t2.latest_bucket
        return out, tick

    # ---- node-cache expiry, made explicit
    def expireNeedNumRecords(self):
        """needNumRecordCache.on('expired') (:227-241): emitted with the BAF account number or ''."""
        out = []
        for logId, need in self.needNumRecordCache.items():
            for svc, r in need.items():
                self.outputRecord(out, r["server"], svc, logId, r.get("altAcctNum") or "", r["startMs"], r["endMs"], r["elapsed"])
        self.needNumRecordCache = {}
        return out

    def expireRecordCache(self):
        """recordCache.on('expired') (:219-225): partial records without an exit are discarded."""
        n = sum(len(m) for m in self.recordCache.values())
        self.recordCache = {}
        return n
