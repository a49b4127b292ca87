"""Interop with the reference's JSON resume files, so a running deployment can move
to the GPU engine - and back - without losing its one-day baselines.

  stats stage   stream_calc_stats.js:54-87   JSON.stringify(StatParser):
      {"servers": {srv: {"services": {svc: {"buckets": {label: [elapsed, ...]}}}}},
       "latestBucket": "157840920", "minHeap": {"content": [TxEntry, ...]}}
  z-score stage stream_calc_z_score.js:37-64 JSON.stringify(ZScoreParser):
      {"servers": {srv: {"services": {svc: {"lags": {"360": {"THRESHOLD", "INFLUENCE",
                                                               "avgList", "per75List", "per95List"}}}}}}}
NaN is written as null by JSON.stringify and average() skips null (util_methods.js:13), so
null <-> NaN is lossless.  Settings are re-resolved by the host after a load, as
resumeFromFileIfExists does (stream_calc_z_score.js:63)."""
from __future__ import annotations

import json
import math

import numpy as np

from .binary_heap import BinaryHeap
from .entries import EntryFactory

LISTS = ("avgList", "per75List", "per95List")


def _null_to_nan(lst):
    return np.array([np.nan if v is None else float(v) for v in lst], dtype=np.float64)


def _nan_to_null(arr):
    return [None if math.isnan(v) else (int(v) if float(v).is_integer() and abs(v) < 1e15 else float(v)) for v in arr]


def load_stats_resume(stat_parser, doc):
    """doc: the parsed JSON of a stats resume file.  Rebuilds names, latestBucket, the reorder
    heap and puts every retained bucket's values back on the device."""
    eng = stat_parser.engine
    bucket_ms = eng.cfg.bucket_ms
    latest = int(doc.get("latestBucket") or 0)
    if latest:
        eng.set_latest(latest)
        stat_parser.latestBucket = str(doc["latestBucket"])
    svc_col, jvm_col, ts_col, el_col = [], [], [], []
    for server, srv in doc.get("servers", {}).items():
        for service, svc in srv.get("services", {}).items():
            sid, jid = stat_parser._ids(server, service)
            eng.mark_seen(jid * eng.cfg.num_services + sid, 1)
            for label, values in svc.get("buckets", {}).items():
                for v in values:
                    svc_col.append(sid); jvm_col.append(jid); ts_col.append(int(label) * bucket_ms); el_col.append(int(v))
    if svc_col:
        eng.add_batch(np.array(svc_col, np.uint32), np.array(jvm_col, np.uint32), np.array(ts_col, np.int64),
                      np.array(el_col, np.int32))
    factory = EntryFactory()
    stat_parser.minHeap = BinaryHeap(lambda tx: tx.endTs)
    for obj in (doc.get("minHeap") or {}).get("content", []):
        e = factory.getEntryFromObject(obj)
        if e is not None:
            stat_parser.minHeap.push(e)


def dump_stats_resume(stat_parser) -> dict:
    """The stats resume document of the reference, rebuilt from the device's bucket logs."""
    eng = stat_parser.engine
    stat_parser._flush()
    S = eng.cfg.num_services
    latest = int(stat_parser.latestBucket or 0)
    servers = {srv: {"services": {svc: {"buckets": {}} for svc in s["services"]}} for srv, s in stat_parser.servers.items()}
    keep = eng.cfg.window + eng.cfg.buffer
    for label in range(max(0, latest - keep), latest + 1):
        words = eng.log_read(label)
        if words.size == 0:
            continue
        keys, vals = (words >> np.uint64(32)).astype(np.int64), (words & np.uint64(0xFFFFFFFF)).astype(np.int64)
        for key in np.unique(keys):
            srv, svc = stat_parser._jvm_names[key // S], stat_parser._svc_names[key % S]
            servers[srv]["services"][svc]["buckets"][str(label)] = vals[keys == key].tolist()
    return {"servers": servers, "latestBucket": stat_parser.latestBucket,
            "minHeap": {"content": [vars(e) for e in stat_parser.minHeap.content]}}


def load_zscore_resume(zscore_parser, doc):
    """doc: the parsed JSON of a z-score resume file.  Lists go to the device rings, then the
    settings are resolved again from the current config (updateAllServiceSettings)."""
    sp, eng = zscore_parser.sp, zscore_parser.engine
    lag_index = {str(l): i for i, l in enumerate(eng.cfg.lags)}
    servers = {}
    for server, srv in doc.get("servers", {}).items():
        for service, svc in srv.get("services", {}).items():
            sid, jid = sp._ids(server, service)
            key = jid * eng.cfg.num_services + sid
            servers.setdefault(server, {"services": {}})["services"][service] = {"lags": {}}
            for lag, lo in svc.get("lags", {}).items():
                servers[server]["services"][service]["lags"][lag] = {"THRESHOLD": lo.get("THRESHOLD"),
                                                                     "INFLUENCE": lo.get("INFLUENCE")}
                if lag not in lag_index:
                    continue  # a lag that is no longer configured: removeStaleLagData would delete it
                for s, name in enumerate(LISTS):
                    lst = lo.get(name) or []
                    if lst:
                        eng.ring_load(key, lag_index[lag], s, _null_to_nan(lst))
    zscore_parser.setProperties(servers)
    zscore_parser.updateAllServiceSettings()


def dump_zscore_resume(zscore_parser) -> dict:
    sp, eng = zscore_parser.sp, zscore_parser.engine
    out = {"servers": {}}
    for server, srv in zscore_parser.servers.items():
        for service, svc in srv["services"].items():
            key = sp.servers[server]["services"][service]
            lags = {}
            for li, lag in enumerate(eng.cfg.lags):
                st = svc["lags"].get(str(lag), {})
                lo = {"THRESHOLD": st.get("THRESHOLD"), "INFLUENCE": st.get("INFLUENCE")}
                for s, name in enumerate(LISTS):
                    lo[name] = _nan_to_null(eng.ring_read(key, li, s))
                lags[str(lag)] = lo
            out["servers"].setdefault(server, {"services": {}})["services"][service] = {"lags": lags}
    return out


def read_json(path):
    with open(path) as f:
        return json.load(f)


def write_json(path, doc):
    with open(path, "w") as f:
        json.dump(doc, f, allow_nan=False)
