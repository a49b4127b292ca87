import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from apmbackend_b200 import Engine, EngineConfig, _capi
from apmbackend_b200 import synthetic as syn
from oracle import capi as ocapi
from tests.gpu_util import sort_records
S, J, n = 200, 2, int(os.environ.get("N", 1000000))
spec = syn.StreamSpec(seed=0xA9B200, n=1000000, num_services=S, num_jvms=J, span_ms=6 * 3600 * 1000, jitter_ms=40000, zipf_s=1.1, shift_mod=40)
stream = syn.SynthStream(spec, device=False)
svc, jvm, ts, el = stream.generate_host(count=n)
lags, thr, infl = (360, 8640), (20.0, 15.0), (0.1, 0.0)
eng = Engine(EngineConfig(num_services=S, num_jvms=J, lags=lags, thresholds=thr, influences=infl, out_ticks=256, max_tx_per_bucket=1 << 16, max_batch=1 << 17, max_alerts_per_tick=4096))
orc = ocapi.Pipeline(ocapi.make_config(S, J, 10000, 30, 6, (75.0, 95.0), (0, 1), lags, thr, infl), fast=True)
rng = np.random.default_rng(5)
if os.environ.get("PRELOAD", "1") == "1":
    for key in rng.choice(S * J, 40, replace=False):
        for s in range(3):
            hist = np.round(rng.uniform(50, 400) + rng.normal(0, 3, 8640), 1)
            hist[rng.integers(0, 8640, 20)] = np.nan
            eng.ring_load(int(key), 1, s, hist); orc.ring_load(int(key), 1, s, hist)
orc.consume(svc, jvm, ts, el)
fs_parts = []
step = 50000
for a in range(0, n, step):
    eng.process_host(svc[a:a + step], jvm[a:a + step], ts[a:a + step], el[a:a + step])
    fs_parts.append(eng.fetch_fs().copy()); eng.release_ticks()
fs = sort_records(np.concatenate(fs_parts), True); ofs = sort_records(orc.fs(), True)
print(fs.size, ofs.size)
bad = np.flatnonzero((fs["signal"] != ofs["signal"]).any(axis=1))
print("mismatching records:", bad.size)
ticks = np.unique(ofs["tick_ts"])
for i in bad[:8]:
    g, w = fs[i], ofs[i]
    print("tick#", int(np.searchsorted(ticks, w["tick_ts"])), "key", w["key"], "lag", w["lag"])
    print("  got ", g["val"], g["avg"], g["lb"], g["ub"], g["signal"])
    print("  want", w["val"], w["avg"], w["lb"], w["ub"], w["signal"])
for f in ("val", "avg", "lb", "ub"):
    a, b = fs[f], ofs[f]
    neq = ~((a == b) | (np.isnan(a) & np.isnan(b)))
    print(f, "mismatch count", int(neq.sum()))
a, b = fs["avg"], ofs["avg"]
neq = ~((a == b) | (np.isnan(a) & np.isnan(b)))
for lag in (360, 8640):
    m = fs["lag"] == lag
    print("lag", lag, "avg mismatches per series", neq[m].sum(axis=0), "of", int(m.sum()))
rows = np.flatnonzero(neq.any(axis=1))
keys = np.unique(fs["key"][rows])
print("keys with mismatches:", keys.size, keys[:20])
k = keys[0]
mk = fs["key"] == k
first = np.flatnonzero(neq.any(axis=1) & mk)[0]
print("first mismatch for key", k, "at tick#", int(np.searchsorted(ticks, fs["tick_ts"][first])), "lag", fs["lag"][first])
first_seen_tick = int(np.searchsorted(ticks, fs["tick_ts"][mk].min()))
print("key first emitted at tick#", first_seen_tick)
for i in np.flatnonzero(mk & (fs["lag"] == fs["lag"][first]))[:3]:
    print("  t#", int(np.searchsorted(ticks, fs["tick_ts"][i])), fs["avg"][i], ofs["avg"][i])
