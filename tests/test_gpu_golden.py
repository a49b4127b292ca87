"""The CUDA path through the C ABI against the committed fixtures of tests/golden/."""
import numpy as np
import pytest

from oracle import capi as ocapi  # formatting of records into the reference's CSV lines only
from tests.test_golden_cpu import STREAMS, columns, load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", STREAMS)
def test_engine_reproduces_the_golden_lines(name):
    from apmbackend_b200 import Engine, EngineConfig
    g = load(name)
    c = g["config"]
    eng = Engine(EngineConfig(num_services=c["S"], num_jvms=c["J"], bucket_ms=c["bucket_ms"], window=c["window"], buffer=c["buffer"],
                              percentiles=tuple(float(x) for x in c["pcts"]), lags=tuple(c["lags"]), thresholds=tuple(c["thr"]),
                              influences=tuple(c["infl"]), out_ticks=1024, max_batch=1 << 12, max_tx_per_bucket=1 << 12))
    eng.process_host(*columns(g))
    S = c["S"]
    st = [ocapi.format_st(r, g["servers"][r["key"] // S], g["services"][r["key"] % S]) for r in eng.fetch_stats()]
    fs = [ocapi.format_fs(r, g["servers"][r["key"] // S], g["services"][r["key"] % S]) for r in eng.fetch_fs()]
    # the engine emits key-ascending, the reference in first-seen order: the lines of the run must be the same multiset
    assert sorted(st) == sorted(g["st_lines"]) and sorted(fs) == sorted(g["fs_lines"])
    eng.close()


def test_log_feeder_reproduces_the_golden_lines():
    from apmbackend_b200.engine import Engine, EngineConfig
    from apmbackend_b200.stat_parser import StatParser
    from apmbackend_b200.transaction_parser import TransactionParser
    g = load("log_server")
    sp = StatParser(Engine(EngineConfig(num_services=64, num_jvms=2, lags=(8,), thresholds=(2.0,), influences=(0.1,), out_ticks=64)))
    tp = TransactionParser(sp)
    out, _ = tp.readLines(g["path"], g["text"].encode(), feed=False)
    out += tp.expireNeedNumRecords()
    assert out == g["tx_lines"]
    sp.engine.close()
