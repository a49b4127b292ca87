import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np
from apmbackend_b200 import Engine, EngineConfig
from apmbackend_b200 import synthetic as syn
from oracle import capi as ocapi
S, J = 200, 2
spec = syn.StreamSpec(seed=0xA9B200, n=1000000, num_services=S, num_jvms=J, span_ms=6 * 3600 * 1000, jitter_ms=40000, zipf_s=1.1, shift_mod=40)
stream = syn.SynthStream(spec, device=False)
n = 3000
svc, jvm, ts, el = stream.generate_host(count=n)
lags, thr, infl = (360, 8640), (20.0, 15.0), (0.1, 0.0)
eng = Engine(EngineConfig(num_services=S, num_jvms=J, lags=lags, thresholds=thr, influences=infl, out_ticks=256, max_tx_per_bucket=1 << 16, max_batch=1 << 17))
orc = ocapi.Pipeline(ocapi.make_config(S, J, 10000, 30, 6, (75.0, 95.0), (0, 1), lags, thr, infl), fast=True)
key = int(os.environ.get("KEY", 0))
for s in range(3):
    hist = np.arange(8640, dtype=np.float64) + 10000 * (s + 1)
    eng.ring_load(key, 1, s, hist); orc.ring_load(key, 1, s, hist)
orc.consume(svc, jvm, ts, el)
eng.process_host(svc, jvm, ts, el)
print("ticks", orc.ticks)
for s in range(3):
    g = eng.ring_read(key, 1, s); w = orc.ring_read(key, 1, s)
    neq = ~((g == w) | (np.isnan(g) & np.isnan(w)))
    print("series", s, "len", g.size, w.size, "mismatch", int(neq.sum()), "first", np.flatnonzero(neq)[:5], g[np.flatnonzero(neq)[:3]], w[np.flatnonzero(neq)[:3]])
    print("   head of list", g[:4], w[:4], "tail", g[-4:], w[-4:])
