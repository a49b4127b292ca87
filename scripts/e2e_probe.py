"""Dev probe: where does the end-to-end step time go (H2D, compute, fetch)?"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from apmbackend_b200 import Engine, EngineConfig, _capi, synthetic as syn
import bench

wl = bench.WORKLOADS["cfg3"]
K = wl["services"]; n = wl["n"]; span_ms = wl["span_buckets"] * wl["bucket_ms"]
per_bucket = n // wl["span_buckets"]
flags = _capi.F_KEEP_FS | _capi.F_ALERTS | _capi.F_ZSCORE | _capi.F_PERCENTILES
ecfg = EngineConfig(num_services=K, bucket_ms=wl["bucket_ms"], window=1, buffer=0, percentiles=wl["percentiles"],
                    z_series_pct=wl["z_series_pct"], lags=wl["lags"], thresholds=wl["thresholds"], influences=wl["influences"],
                    flags=flags, max_tx_per_bucket=int(per_bucket * 1.5) + 4096, max_batch=int(os.environ.get("BATCH", 1 << 25)),
                    out_ticks=72, max_alerts_per_tick=K)
eng = Engine(ecfg)
L = wl["lags"][0]
spec = syn.StreamSpec(seed=2, n=n, num_services=K, span_ms=span_ms, shift_mod=97)
stream = syn.SynthStream(spec)
buf = syn.DeviceBuffer(L * 3 * K * 8); syn.ring_fill_device(buf, L, stream, 77); eng.ring_load_all(0, length=L, device_ptr=buf.ptr); buf.free()
tmp = syn.DeviceColumns(n)
steps = []
for s in range(3):
    stream.generate_device(tmp, t0_ms=spec.t0_ms + s * span_ms, seed=2 + s)
    pins = [syn.PinnedArray(np.uint32, n), syn.PinnedArray(np.uint32, n), syn.PinnedArray(np.int64, n), syn.PinnedArray(np.int32, n)]
    for pa, b, dt in zip(pins, tmp.bufs, (np.uint32, np.uint32, np.int64, np.int32)):
        _capi.synth().synth_d2h(pa.ptr, b.ptr, n * np.dtype(dt).itemsize)
    steps.append(pins)
# raw H2D bandwidth
d = syn.DeviceBuffer(n * 8)
t = time.perf_counter(); _capi.synth().synth_h2d(d.ptr, steps[0][2].ptr, n * 8); dt = time.perf_counter() - t
print(f"raw pinned H2D 800MB: {dt*1e3:.1f} ms = {0.8/dt:.1f} GB/s")
for s in range(3):
    p = steps[s]
    t0 = time.perf_counter(); eng.process_host(p[0].array, p[1].array, p[2].array, p[3].array); t1 = time.perf_counter()
    fs = eng.fetch_fs(pinned=True); t2 = time.perf_counter()
    al = eng.fetch_alerts(pinned=True); t3 = time.perf_counter()
    eng.release_ticks(); t4 = time.perf_counter()
    print(f"step {s}: process_host {1e3*(t1-t0):.1f} ms, fetch_fs {1e3*(t2-t1):.1f} ms ({fs.nbytes/1e6:.0f} MB), fetch_alerts {1e3*(t3-t2):.1f} ms ({al.size}), release {1e3*(t4-t3):.2f} ms")

# the 12-byte form, fs records streamed to the bound pinned buffer
eng.bind_fs_output()
for s in range(3):
    p = steps[s]
    ts = p[2].array + 3 * span_ms            # keep time moving forward
    base = int(ts.min())
    key = (p[1].array << 16) | p[0].array
    dtv = (ts - base).astype(np.uint32)
    el = p[3].array.view(np.uint32)
    pk = [syn.PinnedArray(np.uint32, n) for _ in range(3)]
    pk[0].array[:] = key; pk[1].array[:] = dtv; pk[2].array[:] = el
    eng.profile_read(); eng.profile_enable(True)
    t0 = time.perf_counter(); eng.process_host_packed(pk[0].array, pk[1].array, pk[2].array, base); t1 = time.perf_counter()
    prof = eng.profile_read(); eng.profile_enable(False)
    print("   spans ms:", {g: round(v["ms"], 2) for g, v in prof.items() if v["ms"] > 0.05})
    fs = eng.fetch_fs(pinned=True); t2 = time.perf_counter()
    al = eng.fetch_alerts(pinned=True); t3 = time.perf_counter()
    eng.release_ticks(); t4 = time.perf_counter()
    print(f"packed step {s}: process {1e3*(t1-t0):.1f} ms, fetch_fs {1e3*(t2-t1):.2f} ms, fetch_alerts {1e3*(t3-t2):.2f} ms ({al.size}), release {1e3*(t4-t3):.2f} ms")
    for a in pk: a.free()
