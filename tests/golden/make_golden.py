"""Generates the fixtures of tests/golden/.  Run from the repository root:  python tests/golden/make_golden.py

The reference is JavaScript and no JavaScript engine exists in this environment, so these vectors are NOT outputs
of the reference running: they are the outputs of the statement-by-statement Python twins of its source
(oracle/pyoracle.py for the stats / z-score stages, oracle/pylogparse.py for the log parser), frozen so that the
C++ oracle, the CUDA path and any later edit of the twins themselves are held to the same bytes.  The one vector
that does come from the reference is `reference_fs_line.json`: the literal `fs` line quoted in
stream_process_alerts.js:638.  Parity status stays "unpinned" (DESIGN.md section 2).
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import pyoracle as po          # noqa: E402
from oracle.pylogparse import LogParserOracle  # noqa: E402
from tests.logs import make_log            # noqa: E402
from tests.streams import make_stream, names  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_FS_LINE = ("fs&1579895320000&undefined&reportAbilityToPay&6&0.00&undefined:undefined:undefined:undefined:0&"
                     "undefined:undefined:undefined:undefined:0.0&undefined:undefined:undefined:undefined:0.0")

STREAMS = [
    # the reference's own settings (10-s buckets, 31-bucket window, buffer 6, p75/p95), short lags so that they fill
    dict(name="stream_reference_settings", seed=101, n=900, S=4, J=2, bucket_ms=10000, window=30, buffer=6, span=80,
         lags=(8, 40), thr=(2.0, 1.5), infl=(0.1, 0.0), pcts=(75, 95), gap=None),
    # one-bucket window, 60-s buckets, a gap without traffic, late arrivals
    dict(name="stream_single_bucket_gap", seed=102, n=700, S=3, J=1, bucket_ms=60000, window=1, buffer=0, span=50,
         lags=(6,), thr=(1.2,), infl=(0.3,), pcts=(75, 95), gap=(20, 27)),
]


def stream_fixture(c):
    svc, jvm, ts, el = make_stream(c["seed"], c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], gap_at=c["gap"],
                                   shift_services=(0, 2), late_frac=0.03)
    jn, sn = names(c["S"], c["J"])
    defaults = [{"LAG": l, "THRESHOLD": t, "INFLUENCE": i} for l, t, i in zip(c["lags"], c["thr"], c["infl"])]
    st, fs, _, _ = po.run_pipeline([(jn[j], sn[s], int(t), int(e)) for s, j, t, e in zip(svc, jvm, ts, el)],
                                   c["bucket_ms"], c["window"], c["buffer"], c["pcts"], defaults, {})
    return {"config": {k: c[k] for k in ("S", "J", "bucket_ms", "window", "buffer", "lags", "thr", "infl", "pcts")},
            "servers": jn, "services": sn,
            "tx": {"service_id": svc.tolist(), "jvm_id": jvm.tolist(), "end_ts_ms": ts.tolist(), "elapsed_ms": el.tolist()},
            "st_lines": st, "fs_lines": fs}


def log_fixture():
    text = make_log(seed=77, n=120, span_ms=60_000).decode()
    o = LogParserOracle()
    for line in text.split("\n"):
        o.readLine("/logs/jvm02/server.log", line)
    o.flush()
    return {"path": "/logs/jvm02/server.log", "text": text, "tx_lines": o.out}


def main():
    for c in STREAMS:
        with open(os.path.join(HERE, c["name"] + ".json"), "w") as f:
            json.dump(stream_fixture(c), f, indent=0)
    with open(os.path.join(HERE, "log_server.json"), "w") as f:
        json.dump(log_fixture(), f, indent=0)
    with open(os.path.join(HERE, "reference_fs_line.json"), "w") as f:
        json.dump({"source": "stream_process_alerts.js:638 - a line of the fs queue quoted there ('&' for '|')",
                   "line": REFERENCE_FS_LINE}, f, indent=0)

if __name__ == "__main__":
    main()
