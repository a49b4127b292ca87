"""The host mirror of the reference's classes (StatParser / ZScoreParser, same method
names) over the CUDA engine must put the same lines on the `stats` and `z_score` queues as
the statement-by-statement Python twin of the JavaScript - byte for byte."""
import numpy as np
import pytest

from apmbackend_b200 import Engine, EngineConfig, StatParser, ZScoreParser
from apmbackend_b200.entries import EntryFactory, TxEntry
from apmbackend_b200.stat_parser import ListQueue
from oracle import pyoracle as po
from tests.streams import make_stream, names

pytestmark = pytest.mark.gpu


def _stream(seed=21, n=2500, S=5, J=2, span=80):
    svc, jvm, ts, el = make_stream(seed, n, S, J, 10000, span, shift_services=(0, 1))
    jn, sn = names(S, J)
    return [TxEntry(jn[j], sn[s], f"log{i}", 1000 + i, int(t) - int(e), int(t), int(e), "true")
            for i, (s, j, t, e) in enumerate(zip(svc, jvm, ts, el))]


def _twin(txs, lags, thr, infl, overrides=None):
    defaults = [{"LAG": l, "THRESHOLD": t, "INFLUENCE": i} for l, t, i in zip(lags, thr, infl)]
    st, fs, _, _ = po.run_pipeline([(t.server, t.service, t.endTs, t.elapsed) for t in txs], defaults=defaults,
                                   overrides=overrides or {})
    return st, fs


def test_consume_msg_protocol_emits_the_reference_lines():
    txs = _stream()
    lags, thr, infl = (8, 30), (2.0, 1.5), (0.1, 0.0)
    eng = Engine(EngineConfig(num_services=8, num_jvms=2, lags=lags, thresholds=thr, influences=infl, out_ticks=8))
    sp = StatParser(eng)
    zp = ZScoreParser(sp)
    out, db, zq = ListQueue(), ListQueue(), ListQueue()
    factory = EntryFactory()
    for tx in txs:
        before = len(out.lines)
        sp.consumeMsg(tx.toCSVString(), out, db)           # stream_calc_stats.js:331-371
        for line in out.lines[before:]:                     # stream_calc_z_score.js:462-475
            for fs in zp.processData(factory.getEntryFromCSV(line)):
                zq.writeLineToQueue(fs.toCSVString())
    st_want, fs_want = _twin(txs, lags, thr, infl)
    assert out.lines == st_want and len(st_want) > 100
    assert zq.lines == fs_want
    assert any(":1|" in l or ":-1|" in l for l in zq.lines)
    # the reorder buffer forwards every transaction up to the tick edge, oldest first
    got_ts = [int(l.split("|")[6]) for l in db.lines]
    assert got_ts == sorted(got_ts) and len(got_ts) > 0
    eng.close()


def test_device_reorder_buffer_forwards_the_same_transactions():
    """With APM_F_PASSTHROUGH the mirror keeps no heap: the db_insert lines must be the reference heap's."""
    txs = _stream(seed=25, n=2000)
    lags, thr, infl = (8,), (2.0,), (0.1,)
    host = StatParser(Engine(EngineConfig(num_services=8, num_jvms=2, lags=lags, thresholds=thr, influences=infl, out_ticks=8)))
    dev = StatParser(Engine(EngineConfig(num_services=8, num_jvms=2, lags=lags, thresholds=thr, influences=infl, out_ticks=8,
                                         flags=1 | 2 | 4 | 8 | 16 | 64, max_tx_per_bucket=1 << 14)))
    out_h, db_h, out_d, db_d = ListQueue(), ListQueue(), ListQueue(), ListQueue()
    for tx in txs:
        host.consumeMsg(tx.toCSVString(), out_h, db_h)
        dev.consumeMsg(tx.toCSVString(), out_d, db_d)
    assert out_d.lines == out_h.lines
    ts_d = [int(l.split("|")[6]) for l in db_d.lines]
    assert ts_d == sorted(ts_d) and sorted(db_d.lines) == sorted(db_h.lines) and len(db_d.lines) > 1000
    assert dev.minHeap.size() == 0 and len(dev._tx_by_seq) == host.minHeap.size()


def test_batch_fast_path_emits_the_same_lines():
    txs = _stream(seed=22)
    lags, thr, infl = (8, 30), (2.0, 1.5), (0.1, 0.0)
    eng = Engine(EngineConfig(num_services=8, num_jvms=2, lags=lags, thresholds=thr, influences=infl, out_ticks=256))
    sp = StatParser(eng)
    out = ListQueue()
    sp.consumeColumns([t.server for t in txs], [t.service for t in txs], np.array([t.endTs for t in txs]),
                      np.array([t.elapsed for t in txs]), out)
    st_want, _ = _twin(txs, lags, thr, infl)
    assert out.lines == st_want
    eng.close()


def test_service_overrides_follow_the_reference_aliasing():
    """getServiceSettingsFromConfig writes an override INTO the shared defaults
    (stream_calc_z_score.js:107-120): services first seen afterwards inherit it."""
    txs = _stream(seed=23, n=2000, S=4, J=1, span=70)
    lags, thr, infl = (6,), (3.0,), (0.2,)
    overrides = {"S:svc0002": {"6": {"THRESHOLD": 0.5}}}
    eng = Engine(EngineConfig(num_services=4, num_jvms=1, lags=lags, thresholds=thr, influences=infl, out_ticks=8))
    sp = StatParser(eng)
    zp = ZScoreParser(sp, {"defaults": [{"LAG": 6, "THRESHOLD": 3.0, "INFLUENCE": 0.2}],
                           "overrides": {"services": overrides}})
    out, db, zq = ListQueue(), ListQueue(), ListQueue()
    factory = EntryFactory()
    for tx in txs:
        before = len(out.lines)
        sp.consumeMsg(tx.toCSVString(), out, db)
        # first sight of a (server, service): its settings go to the device before the tick that first emits it
        zp.noteService(tx.server, tx.service)
        for line in out.lines[before:]:
            for fs in zp.processData(factory.getEntryFromCSV(line)):
                zq.writeLineToQueue(fs.toCSVString())
    _, fs_want = _twin(txs, lags, thr, infl, overrides)
    assert zq.lines == fs_want


def test_resume_file_round_trip(tmp_path):
    txs = _stream(seed=24, n=1500)
    lags, thr, infl = (8,), (2.0,), (0.1,)
    cfg = dict(num_services=8, num_jvms=2, lags=lags, thresholds=thr, influences=infl, out_ticks=8)
    a = StatParser(Engine(EngineConfig(**cfg)))
    out_a, db = ListQueue(), ListQueue()
    half = len(txs) // 2
    for tx in txs[:half]:
        a.consumeMsg(tx.toCSVString(), out_a, db)
    path = str(tmp_path / "stats.resume")
    a.saveToResumeFile(path)
    b = StatParser(Engine(EngineConfig(**cfg)))
    assert b.resumeFromFileIfExists(path)
    out_b = ListQueue()
    out_a.lines.clear()
    for tx in txs[half:]:
        a.consumeMsg(tx.toCSVString(), out_a, db)
        b.consumeMsg(tx.toCSVString(), out_b, ListQueue())
    assert out_a.lines == out_b.lines and len(out_b.lines) > 0
    assert b.minHeap.size() == a.minHeap.size()
