"""the reference's JSON resume files in and out of the GPU engine.  The Python twin
of the JavaScript holds exactly the reference's state shape, so its state *is* a resume document:
a stream is cut in two, the state crosses over through JSON in the middle (both directions), and
the second half must produce the same `st` / `fs` lines as an uninterrupted reference run."""
import json
import math

import pytest

from apmbackend_b200 import Engine, EngineConfig, StatParser, ZScoreParser, resume_interop as ri
from apmbackend_b200.entries import EntryFactory, TxEntry
from apmbackend_b200.stat_parser import ListQueue
from oracle import pyoracle as po
from tests.streams import make_stream, names

pytestmark = pytest.mark.gpu

LAGS, THR, INFL = (8, 30), (2.0, 1.5), (0.1, 0.0)


def _txs(seed=61, n=2400, S=5, J=2, span=80):
    svc, jvm, ts, el = make_stream(seed, n, S, J, 10000, span, shift_services=(0, 1))
    jn, sn = names(S, J)
    return [TxEntry(jn[j], sn[s], f"log{i}", 1000 + i, int(t) - int(e), int(t), int(e), "true")
            for i, (s, j, t, e) in enumerate(zip(svc, jvm, ts, el))]


def _twin_defaults():
    return [{"LAG": l, "THRESHOLD": t, "INFLUENCE": i} for l, t, i in zip(LAGS, THR, INFL)]


def _twin_run(sp, zp, txs):
    st_lines, fs_lines = [], []
    for tx in txs:
        before = len(sp.out)
        sp.consume(tx.server, tx.service, tx.endTs, tx.elapsed)
        for line in sp.out[before:]:
            st_lines.append(line)
            fs_lines += [e.to_csv() for e in zp.processData(po.StatEntry.from_csv(line))]
    return st_lines, fs_lines


def _mirror():
    eng = Engine(EngineConfig(num_services=8, num_jvms=2, lags=LAGS, thresholds=THR, influences=INFL, out_ticks=8))
    sp = StatParser(eng)
    return eng, sp, ZScoreParser(sp)


def _mirror_run(sp, zp, txs):
    out, db, fs_lines = ListQueue(), ListQueue(), []
    factory = EntryFactory()
    for tx in txs:
        before = len(out.lines)
        sp.consumeMsg(tx.toCSVString(), out, db)
        for line in out.lines[before:]:
            fs_lines += [e.toCSVString() for e in zp.processData(factory.getEntryFromCSV(line))]
    return out.lines, fs_lines


def _json_roundtrip(doc):
    def clean(o):
        if isinstance(o, float) and math.isnan(o):
            return None  # JSON.stringify(NaN) === 'null'
        if isinstance(o, dict):
            return {k: clean(v) for k, v in o.items()}
        if isinstance(o, list):
            return [clean(v) for v in o]
        return o
    return json.loads(json.dumps(clean(doc), allow_nan=False))


def test_reference_resume_files_load_into_the_engine():
    txs = _txs()
    half = len(txs) // 2
    sp_t, zp_t = po.StatParser(), po.ZScoreParser(_twin_defaults())
    _twin_run(sp_t, zp_t, txs[:half])
    stats_doc = _json_roundtrip({"servers": sp_t.servers, "latestBucket": sp_t.latestBucket, "minHeap": {"content": []}})
    z_doc = _json_roundtrip({"servers": zp_t.servers})
    eng, sp, zp = _mirror()
    ri.load_stats_resume(sp, stats_doc)
    ri.load_zscore_resume(zp, z_doc)
    want_st, want_fs = _twin_run(sp_t, zp_t, txs[half:])
    got_st, got_fs = _mirror_run(sp, zp, txs[half:])
    assert got_st == want_st and len(want_st) > 100
    assert got_fs == want_fs
    assert any(":1|" in l or ":-1|" in l for l in got_fs)
    eng.close()


def test_engine_state_exports_to_reference_resume_files():
    txs = _txs(seed=62)
    half = len(txs) // 2
    eng, sp, zp = _mirror()
    _mirror_run(sp, zp, txs[:half])
    stats_doc = _json_roundtrip(ri.dump_stats_resume(sp))
    z_doc = _json_roundtrip(ri.dump_zscore_resume(zp))
    # a reference process resumes from these files ...
    sp_t, zp_t = po.StatParser(), po.ZScoreParser(_twin_defaults())
    sp_t.servers, sp_t.latestBucket = stats_doc["servers"], stats_doc["latestBucket"]
    for srv in z_doc["servers"].values():
        for svc in srv["services"].values():
            for lo in svc["lags"].values():
                for name in ri.LISTS:
                    lo[name] = [float("nan") if v is None else v for v in lo[name]]
    zp_t.servers = z_doc["servers"]
    # ... and both continue
    want_st, want_fs = _twin_run(sp_t, zp_t, txs[half:])
    got_st, got_fs = _mirror_run(sp, zp, txs[half:])
    assert got_st == want_st and len(want_st) > 100
    assert got_fs == want_fs
    # and the document has the reference's shape
    some_srv = next(iter(stats_doc["servers"].values()))
    some_svc = next(iter(some_srv["services"].values()))
    assert isinstance(stats_doc["latestBucket"], str) and all(isinstance(v, list) for v in some_svc["buckets"].values())
    eng.close()
