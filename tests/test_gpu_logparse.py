"""Log feeder: the device parser + host mirror against the CPU restatement of
stream_parse_transactions.js (oracle/pylogparse.py), through the C ABI."""
import math

import numpy as np
import pytest

from tests.logs import make_log

pytestmark = pytest.mark.gpu

SRV = "/logs/jvm02/server.log"
APP = "/logs/jvm02/app.log"


def _oracle(fp, text, tz=None):
    from oracle.pylogparse import LogParserOracle
    o = LogParserOracle(tz)
    for l in text.decode().split("\n"):
        o.readLine(fp, l)
    o.flush()
    return o.out


def _mirror(**kw):
    from apmbackend_b200.engine import Engine, EngineConfig
    from apmbackend_b200.stat_parser import StatParser
    from apmbackend_b200.transaction_parser import TransactionParser
    cfg = dict(num_services=64, num_jvms=2, lags=(8,), thresholds=(2.0,), influences=(0.1,), out_ticks=64)
    cfg.update(kw)
    sp = StatParser(Engine(EngineConfig(**cfg)))
    return sp, TransactionParser(sp)


@pytest.mark.parametrize("fp,seed", [(SRV, 1), (SRV, 2), (APP, 3)])
def test_tx_lines_match_the_oracle(fp, seed):
    text = make_log(seed=seed, n=500, server_log=fp == SRV)
    sp, tp = _mirror()
    out, _ = tp.readLines(fp, text, feed=False)
    out += tp.expireNeedNumRecords()
    want = _oracle(fp, text)
    assert out == want and len(want) > 300
    res = tp.lastParseResult
    assert res.lines == text.count(b"\n") + 1 and res.exits >= len(want) and res.unknown_service == 0
    sp.engine.close()


def test_split_buffers_give_the_same_lines():
    text = make_log(seed=5, n=600)
    lines = text.split(b"\n")
    sp, tp = _mirror()
    got = []
    for a, b in ((0, 211), (211, 212), (212, 700), (700, len(lines))):
        chunk = b"\n".join(lines[a:b]) + (b"\n" if b < len(lines) else b"")
        if chunk:
            got += tp.readLines(SRV, chunk, feed=False)[0]
    got += tp.expireNeedNumRecords()
    # emission order differs only by when a buffer ends; the multiset of records must be the reference's
    assert sorted(got) == sorted(_oracle(SRV, text))
    sp.engine.close()


def test_time_zone_rules_and_crlf():
    from oracle.pylogparse import TimeZone
    rules = ((-(1 << 62), -6 * 3600000), (1583632800000, -5 * 3600000))
    text = (b"[a] 2020-03-08 01:59:59,000 INFO  CommonTiming::Start Svc\r\n"
            b"[a] 2020-03-08 03:00:01,500 INFO  CommonTiming::Stop Svc elapsed time = 2500 ms\r\n")
    sp, tp = _mirror()
    tp.tz_rules = rules
    out, _ = tp.readLines(APP, text, feed=False)
    out += tp.expireNeedNumRecords()
    assert out == _oracle(APP, text, TimeZone(list(rules)))
    assert out == ["tx|jvm02|Svc|a|NaN|1583654399000|1583654401500|2500|N"]
    sp.engine.close()


def test_fed_exits_drive_the_hot_path_like_the_reference_lines():
    """Stats and z-score records after feeding the parsed exits on the device == the same transactions,
    taken from the oracle's tx lines, pushed through apm_process_host."""
    from apmbackend_b200.engine import Engine, EngineConfig
    # unique log ids: the reference's need-number cache keeps ONE record per (logId, service) until it expires
    # (needMap.set overwrites, :549-556), a wall-clock-dependent loss the feed path does not reproduce
    text = make_log(seed=9, n=3000, span_ms=900_000, unique_ids=True)
    sp, tp = _mirror(out_ticks=256)
    _, tick = tp.readLines(SRV, text, feed=True)
    assert tick is not None and tick.ticks > 50 and tick.tx_ingested > 2000
    stats_a, fs_a = sp.engine.fetch_stats().copy(), sp.engine.fetch_fs().copy()
    sp.engine.release_ticks()
    rows = []
    for l in _oracle(SRV, text):
        f = l.split("|")
        if "NaN" in (f[6], f[7]):
            continue                      # the device holds malformed exits back (APM_LR_BAD_*)
        rows.append((sp._svc_ids[f[2]], sp.servers[f[1]]["id"], int(f[6]), int(f[7])))
    rows.sort(key=lambda r: r[2])         # the log is in end-time order; ties share a bucket
    a = np.array(rows, np.int64)
    assert len(rows) == tick.tx_ingested + tick.tx_dropped + tick.tx_invalid
    ref = Engine(EngineConfig(num_services=64, num_jvms=2, lags=(8,), thresholds=(2.0,), influences=(0.1,), out_ticks=256))
    t2 = ref.process_host(a[:, 0].astype(np.uint32), a[:, 1].astype(np.uint32), a[:, 2], a[:, 3].astype(np.int32))
    assert (t2.ticks, t2.tx_ingested, t2.latest_bucket) == (tick.ticks, tick.tx_ingested, tick.latest_bucket)
    stats_b, fs_b = ref.fetch_stats(), ref.fetch_fs()
    assert stats_a.tobytes() == stats_b.tobytes() and fs_a.tobytes() == fs_b.tobytes() and len(stats_a) > 100
    ref.close()
    sp.engine.close()


def test_empty_and_unterminated_buffers():
    sp, tp = _mirror()
    assert tp.readLines(APP, b"", feed=True)[0] == []
    assert tp.readLines(APP, b"\n\n\n", feed=False)[0] == []
    out, _ = tp.readLines(APP, b"[k] 2020-01-07 10:00:02,000 INFO  CommonTiming::Stop Svc elapsed time = 5 ms", feed=False)
    assert out == ["tx|jvm02|Svc||NaN|1578391201995|1578391202000|5|N"]
    sp.engine.close()


def test_error_behaviour_of_the_feeder_entry_points():
    from apmbackend_b200 import _capi
    from apmbackend_b200._capi import ApmError
    sp, tp = _mirror()
    eng = sp.engine
    with pytest.raises(ApmError, match="file_type"):
        eng.parse_log(b"x\n", 2, 0)
    with pytest.raises(ApmError, match="jvm_id"):
        eng.parse_log(b"x\n", 0, 7)
    with pytest.raises(ApmError, match="zone rules"):
        eng.parse_log(b"x\n", 0, 0, tz_rules=[(i, 0) for i in range(9)])
    with pytest.raises(ApmError, match="outside num_services"):
        eng.dict_set({"Svc": 64})
    # a buffer without a CommonTiming line: nothing recognised, nothing to feed
    recs, res = eng.parse_log(b"plain line\nanother INFO line\n", 0, 0)
    assert len(recs) == 0 and (res.lines, res.records, res.fed) == (2, 0, 0)
    assert eng.feed_parsed().ticks == 0
    # an exit whose service is not in the dictionary is held back and reported
    recs, res = eng.parse_log(b"[k] 2020-01-07 10:00:02,000 INFO  CommonTiming::Stop Nobody elapsed time = 5 ms\n", 0, 0)
    assert res.unknown_service == 1 and res.fed == 0 and recs[0]["service_id"] == _capi.NO_SERVICE
    eng.close()


def test_packed_entry_point_rejects_a_negative_base():
    from apmbackend_b200._capi import ApmError
    sp, _ = _mirror()
    with pytest.raises(ApmError, match="base_ts_ms"):
        sp.engine.process_host_packed(np.zeros(1, np.uint32), np.zeros(1, np.uint32), np.zeros(1, np.uint32), -1)
    sp.engine.close()
