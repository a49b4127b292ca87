"""Throughput of the log feeder: pinned log text -> apm_parse_log_host -> apm_feed_parsed.
Timed with the library's CUDA-event timer around the calls (H2D of the text inside the timed region).
    python scripts/bench_logparse.py [--mb 256] [--reps 5]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from apmbackend_b200.engine import Engine, EngineConfig  # noqa: E402
from apmbackend_b200.transaction_parser import rewrite_service  # noqa: E402
from tests.logs import SERVICES, make_log  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--records", action="store_true", help="also copy the per-line records back")
    a = ap.parse_args()
    block = make_log(seed=11, n=20000, span_ms=600_000, unique_ids=True) + b"\n"
    reps = max(1, a.mb * (1 << 20) // len(block))
    text = block * reps
    eng = Engine(EngineConfig(num_services=64, num_jvms=2, lags=(8,), thresholds=(2.0,), influences=(0.1,), out_ticks=4096,
                              max_batch=1 << 24, max_tx_per_bucket=1 << 24, flags=1 | 2))
    names = {}
    for s in SERVICES + ["Svc", "undefined", "Login"]:
        for n in (rewrite_service(s), rewrite_service("S:" + s)):
            names.setdefault(n, len(names))
    eng.dict_set(names)
    pin = eng.pinned_bytes(len(text))
    pin[:] = np.frombuffer(text, np.uint8)
    best = None
    for r in range(a.reps + 1):
        eng.sync()
        eng.timer_start()
        t0 = time.perf_counter()
        _, res = eng.parse_log(pin, 1, 0, want_records=a.records)
        t1 = time.perf_counter()
        tick = eng.feed_parsed()
        ms = eng.timer_stop()
        wall = (time.perf_counter() - t0) * 1e3
        parse_wall = (t1 - t0) * 1e3
        eng.release_ticks()
        eng.set_latest(1)  # the same text again: start the clock over
        if r and (best is None or ms < best[0]):
            best = (ms, wall, res, tick, parse_wall)
    ms, wall, res, tick, parse_wall = best
    print(json.dumps({"text_mb": len(text) / 2 ** 20, "lines": int(res.lines), "records": int(res.records), "fed": int(res.fed),
                      "ms": ms, "wall_ms": wall, "parse_ms": parse_wall, "feed_ms": wall - parse_wall,
                      "parse_only_text_gb_per_s": len(text) / parse_wall / 1e6, "text_gb_per_s": len(text) / ms / 1e6, "lines_per_s": res.lines / ms * 1e3,
                      "fed_tx_per_s": res.fed / ms * 1e3, "ticks": int(tick.ticks)}))
    eng.close()


if __name__ == "__main__":
    main()
