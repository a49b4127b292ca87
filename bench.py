#!/usr/bin/env python
"""
bench.py - the hot path on BASELINE.json's metric: transactions/sec aggregated + z-scored.

A step is one pass of the hot path over one batch of synthetic input: the whole workload's
transaction stream (cfg3: 100 M transactions, 10 k service keys, 1-min buckets, p50/p95/p99
and z-score against a pre-filled 10 080-bucket baseline), i.e. 60 ticks.  Every step gets its
own pre-generated stream one hour later than the previous one, resident in HBM (`value`) or
in pinned host memory (`e2e`) before the timed region starts; inputs are 2 GB per step, far
above the 126 MB L2, so no L2 flush is needed between steps.

  python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
  python bench.py --impl reference ...                      # the reference's algorithm on the host cores

N > 1 (launched by torchrun, one rank per GPU): weak scaling - every rank owns its own hash
shard of the key space (10 k keys and 100 M transactions per rank), and the only exchange is
the NCCL all-gather of the compact alert records after each step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[2]: the configuration the metric ("aggregated+z-scored") is quoted on
    "cfg3": dict(n=100_000_000, services=10_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                 percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                 influences=(0.0,), zscore=True, pct=True, seed=2,
                 desc="100M transactions, 10k service keys, 1-min buckets, p50/p95/p99 + z-score vs 7-day "
                      "rolling baseline (10080 buckets/key, pre-filled), exact-order baseline"),
    # BASELINE.json configs[0]: the reference's own settings (10 s buckets, window 30, buffer 6, p75/p95, lags 360 + 8640)
    "cfg1": dict(n=1_000_000, services=200, jvms=2, bucket_ms=10_000, window=30, buffer=6, span_buckets=2160,
                 percentiles=(75.0, 95.0), z_series_pct=(0, 1), lags=(360, 8640), thresholds=(20.0, 15.0),
                 influences=(0.1, 0.0), zscore=True, pct=True, seed=0xA9B200, cpu_sample_buckets=180,
                 desc="1M transactions, 200 services x 2 JVMs, reference settings (10-s buckets, 31-bucket sliding window, "
                      "p75/p95, lags 360 and 8640, both pre-filled: the steady state of a deployment older than a day), "
                      "2160 ticks per step"),
    # configs[0] read literally ("1-min buckets"; the reference itself is hard-wired to 10 s, config/apm_config.json:105):
    # the same 1 M transactions over 360 one-minute buckets
    "cfg1_60s": dict(n=1_000_000, services=200, jvms=2, bucket_ms=60_000, window=30, buffer=6, span_buckets=360,
                     percentiles=(75.0, 95.0), z_series_pct=(0, 1), lags=(360, 8640), thresholds=(20.0, 15.0),
                     influences=(0.1, 0.0), zscore=True, pct=True, seed=0xA9B200, cpu_sample_buckets=60,
                     desc="1M transactions, 200 services x 2 JVMs, the reference's window / lags with 1-min buckets as BASELINE.json "
                          "words config 1 (31-bucket sliding window, p75/p95, lags 360 and 8640, both pre-filled), 360 ticks per step"),
    # one rank's share of BASELINE.json configs[3] at 8 GPUs: 1 B transactions / 8, 1 M keys / 8
    "cfg4_shard": dict(n=125_000_000, services=125_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                       percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                       influences=(0.0,), zscore=True, pct=True, seed=3,
                       desc="one of 8 key-hash shards of 1B transactions / 1M keys: 125M transactions, 125k keys, "
                            "p50/p95/p99 + z-score vs 10080-bucket baseline (30 GB of rings)"),
    # BASELINE.json configs[4]: heavy skew - the top 10 services carry 80 % of the traffic; cfg3's shape otherwise
    # (what one rank sees when all hot keys land in its shard)
    "cfg5": dict(n=100_000_000, services=10_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                 percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                 influences=(0.0,), zscore=True, pct=True, seed=5, hot=(10, 0.8), log_cap_factor=1.5,
                 desc="100M transactions, 10k service keys, top 10 services = 80% of the traffic, p50/p95/p99 + z-score vs "
                      "10080-bucket baseline (pre-filled), exact-order baseline"),
    # configs[4] as BASELINE.json states it (1 B transactions, 10 k keys, 8 GPUs), seen from one rank: splitmix64 ownership
    # gives the heaviest rank 1316 keys with 2 of the 10 hot services = 18.6 % of the transactions (ideal 12.5 %) ...
    "cfg5_rank8_heavy": dict(n=186_000_000, services=1316, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                             percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                             influences=(0.0,), zscore=True, pct=True, seed=5, hot=(2, 0.86), log_cap_factor=1.5,
                             desc="heaviest of 8 key-hash shards of 1B transactions / 10k keys with top 10 = 80%: 186M "
                                  "transactions, 1316 keys, 2 hot services = 86% of the shard"),
    # ... and an even shard of the same job (1250 keys, 125 M transactions, one hot service)
    "cfg5_rank8_even": dict(n=125_000_000, services=1250, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                            percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                            influences=(0.0,), zscore=True, pct=True, seed=5, hot=(1, 0.64), log_cap_factor=1.5,
                            desc="an even 1/8 shard of 1B transactions / 10k keys with top 10 = 80%: 125M transactions, "
                                 "1250 keys, 1 hot service = 64% of the shard"),
    # the same job with count-aware key placement (sharding.balanced_owner_table): two ranks hold two hot services each
    # and nothing else (16 % of the transactions), six hold one hot service and a sixth of the other keys (11.3 %)
    "cfg5_lpt8_hot": dict(n=160_000_000, services=2, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                          percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                          influences=(0.0,), zscore=True, pct=True, seed=5, hot=(2, 1.0), log_cap_factor=1.5,
                          desc="count-aware placement of 1B transactions / 10k keys with top 10 = 80% on 8 ranks, a rank "
                               "that holds two hot services only: 160M transactions, 2 keys"),
    "cfg5_lpt8_rest": dict(n=113_300_000, services=1667, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                           percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                           influences=(0.0,), zscore=True, pct=True, seed=5, hot=(1, 0.706), log_cap_factor=1.5,
                           desc="count-aware placement of 1B transactions / 10k keys with top 10 = 80% on 8 ranks, a rank "
                                "with one hot service and a sixth of the other keys: 113M transactions, 1667 keys"),
    # ... and with a key's own cost in the balance (balanced_owner_table(key_cost=160e3)): the two-hot ranks keep 917 keys
    "cfg5_cost8_two_hot": dict(n=178_300_000, services=917, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                               percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                               influences=(0.0,), zscore=True, pct=True, seed=5, hot=(2, 0.897), log_cap_factor=1.5,
                               desc="key-cost-aware placement of 1B transactions / 10k keys with top 10 = 80% on 8 ranks, a rank "
                                    "with two hot services: 178M transactions, 917 keys"),
    "cfg5_cost8_one_hot": dict(n=107_200_000, services=1361, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                               percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,),
                               influences=(0.0,), zscore=True, pct=True, seed=5, hot=(1, 0.746), log_cap_factor=1.5,
                               desc="key-cost-aware placement of 1B transactions / 10k keys with top 10 = 80% on 8 ranks, a rank "
                                    "with one hot service: 107M transactions, 1361 keys"),
    # BASELINE.json configs[3] as a job: ONE stream of 1 B transactions over 1 M keys per step, sharded by key over the ranks
    # (--gpus 4 or 8: 242 GB of lag rings in total)
    "cfg4": dict(n=1_000_000_000, services=1_000_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                 percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,), influences=(0.0,),
                 zscore=True, pct=True, seed=3, job=True, min_gpus=4, alert_frac=0.1, alert_rules=True,
                 desc="1B transactions, 1M keys, key-sharded over the ranks with NCCL alert all-gather, p50/p95/p99 + z-score vs "
                      "10080-bucket baseline (242 GB of rings over all ranks), alert rules of the reference's stage [4] on "
                      "(most of the 1M keys see about one transaction per bucket: every other record carries a signal, the "
                      "rules' tpm / bad-interval gates make them the rare alerts they are in a deployment), exact-order baseline"),
    # BASELINE.json configs[4] as a job: 1 B transactions, 10 k keys, top 10 services = 80 % of the traffic (load-balance stress)
    "cfg5_job": dict(n=1_000_000_000, services=10_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                     percentiles=(50.0, 95.0, 99.0), z_series_pct=(1, 2), lags=(10080,), thresholds=(15.0,), influences=(0.0,),
                     zscore=True, pct=True, seed=4, hot=(10, 0.8), log_cap_factor=1.5, job=True, min_gpus=2,
                     desc="1B transactions, 10k keys, top 10 services = 80% of the traffic, key-sharded over the ranks "
                          "(placement table), p50/p95/p99 + z-score vs 10080-bucket baseline, exact-order baseline"),
    # BASELINE.json configs[1]
    "cfg2": dict(n=100_000_000, services=10_000, jvms=1, bucket_ms=60_000, window=1, buffer=0, span_buckets=60,
                 percentiles=(), z_series_pct=(0, 1), lags=(), thresholds=(), influences=(), zscore=False, pct=False,
                 seed=1, desc="100M transactions, 10k service keys, 1-min buckets, stats-only"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=float, default=1.0, help="shrink n (debug only; the line says so)")
    ap.add_argument("--baseline", default="exact", choices=["exact", "incremental"],
                    help="exact: left-to-right re-sum of every lag list per tick (bit-identical means, the parity default); "
                         "incremental: guarded running sums (APM_F_INCREMENTAL: signals and lists bit-identical, means to 1e-9)")
    ap.add_argument("--placement", default="auto", choices=["auto", "hash", "table", "capped"],
                    help="N > 1: which rank owns a key. hash: apm_owner_of_key (splitmix64 mod ranks) for every key; capped: "
                         "apm_owner_table_build_capped, transactions balanced with even key counts; table: apm_owner_table_build, "
                         "count- and key-cost-aware; auto: table for the two-tier skew of config 5, capped otherwise")
    ap.add_argument("--no-incremental", action="store_true", help="skip the extra timed pass with APM_F_INCREMENTAL")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-profile", action="store_true",
                    help="no per-kernel CUDA-event spans inside the timed region (they cost host time when ticks are "
                         "tiny, e.g. cfg1); the roofline then comes from the isolated pass only")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons of one GPU, sampled DURING the timed region: NVML in a thread (a sample
    every few ms; nvidia-smi needs a second to start on an 8-GPU box and missed short runs), nvidia-smi -lms as
    the fallback when the NVML binding is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.nvml = None
        self.samples = []      # (sm_mhz, reasons bitmask)
        self.max_mhz = None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    def _poll(self):
        nv = self.nvml
        while not self._stop.is_set():
            try:
                self.samples.append((float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)),
                                     int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))))
            except Exception:
                pass
            self._stop.wait(0.005)

    def start(self):
        if self.nvml is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.nvml is not None:
            self._stop.set()
            self.t.join(timeout=1)
            nv = self.nvml
            bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            sm = [x[0] for x in self.samples]
            reasons = sorted(n for n, b in bits.items() if any(x[1] & b for x in self.samples))
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                    "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# ---------------------------------------------------------------------------------------------
# the reference's algorithm on the host cores (oracle port; test/baseline infrastructure)
# ---------------------------------------------------------------------------------------------
def cpu_reference_run(wl, cols, ring_seed, threads, stream, prefill=True, fast=False):
    """Times the reference-structure CPU port (oracle fast=0: sorted insert per element, four
    list passes per statistic, shift() memmove) on the given columns.  Keys are independent,
    so the host threads each own key % threads of them while all of them walk the whole
    stream for the ticks - the reference itself runs each stage as one single-threaded process."""
    from oracle import capi as ocapi
    from apmbackend_b200 import synthetic as syn

    svc, jvm, ts, el = cols
    K = wl["services"] * wl["jvms"]
    flags = ocapi.F_KEEP_FS | (ocapi.F_ZSCORE if wl["zscore"] else 0) | (ocapi.F_PERCENTILES if wl["pct"] else 0)
    cfg = ocapi.make_config(wl["services"], wl["jvms"], wl["bucket_ms"], wl["window"], wl["buffer"], wl["percentiles"],
                            wl["z_series_pct"], wl["lags"], wl["thresholds"], wl["influences"], flags=flags)
    pipes = []
    for t in range(threads):
        p = ocapi.Pipeline(cfg, fast=fast)
        keep = np.zeros(K, dtype=np.uint8)
        keep[t::threads] = 1
        p.set_filter(keep)
        p.keep_outputs(False)
        if prefill and wl["zscore"] and wl.get("prefill", True):
            mine = np.arange(t, K, threads)
            for li, L in enumerate(wl["lags"]):
                for s in range(3):
                    vals = syn.ring_fill_host(L, stream, ring_seed + li, s * K + mine)
                    for j, key in enumerate(mine):
                        p.ring_load(int(key), li, s, vals[:, j])
        pipes.append(p)
    t0 = time.perf_counter()
    th = [threading.Thread(target=p.consume, args=(svc, jvm, ts, el)) for p in pipes]
    for x in th:
        x.start()
    for x in th:
        x.join()
    dt = time.perf_counter() - t0
    ticks = pipes[0].ticks
    for p in pipes:
        p.close()
    return dt, ticks


def sample_for_cpu(wl, stream, ticks=2):
    """A bounded sample of the same workload: the first `ticks` buckets' worth of the stream
    (all 10 k keys, same rates), plus one transaction of the next bucket to close the last tick."""
    per_bucket = wl["n"] // wl["span_buckets"]
    ticks = wl.get("cpu_sample_buckets", ticks)
    if wl["n"] <= 2_000_000 and "cpu_sample_buckets" not in wl:
        return stream.generate_host(first=0, count=wl["n"]), wl["n"]  # small enough to run whole
    n = int(per_bucket * ticks + per_bucket // 50)
    return stream.generate_host(first=0, count=n), n


# ---------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def emit(line: dict):
    """The one JSON line goes to the process's real stdout; everything else libraries print while the bench
    runs (NCCL's version banner, for one) was sent to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def resolve_job(wl, world):
    """Totals of the job one step runs.  Workloads quoted per GPU (cfg3 ...) scale weakly: `world` times the keys and
    the transactions, in ONE global stream.  Workloads quoted as a job (cfg4, cfg5_job) keep their totals (strong)."""
    if wl.get("job"):
        total_n, total_services, scaling = wl["n"], wl["services"], "strong"
    else:
        total_n, total_services, scaling = wl["n"] * world, wl["services"] * world, "weak"
    return total_n, total_services, scaling


class Shard:
    """What this rank runs: its keys of the global stream (all of them at world 1)."""

    def __init__(self, args, wl, rank, world, syn, _capi):
        self.syn, self.wl, self.rank, self.world = syn, wl, rank, world
        self.total_n, self.total_services, self.scaling = resolve_job(wl, world)
        self.span_ms = wl["span_buckets"] * wl["bucket_ms"]
        self.spec = syn.StreamSpec(seed=wl["seed"], n=self.total_n, num_services=self.total_services, num_jvms=wl["jvms"],
                                   span_ms=self.span_ms, shift_mod=97, hot=wl.get("hot"))
        self.stream = syn.SynthStream(self.spec, device=True)
        self.K_total = self.total_services * wl["jvms"]
        self.plan = None
        self.placement = "none"
        if world > 1:
            exp = syn.expected_key_counts(self.stream)
            placement = args.placement
            if placement == "auto":
                # Zipf traffic is not hash-friendly: with 80 k keys the top key alone is 8 % of the stream, as much as a
                # rank's whole fair share at 8 ranks.  The table keeps the hash for the light keys and places the heavy ones.
                placement = "table" if wl.get("hot") else "capped"
            owner = None
            if placement == "capped":
                counts = np.ascontiguousarray(np.rint(exp), dtype=np.uint64)
                owner = np.empty(self.K_total, np.uint32)
                st = _capi.lib().apm_owner_table_build_capped(counts.ctypes.data, self.K_total, world, 0.01, owner.ctypes.data)
                if st != 0:
                    raise SystemExit("apm_owner_table_build_capped failed")
                self.placement = ("apm_owner_table_build_capped: transactions balanced longest-first, every rank's key count within "
                                  "1 % of the mean (Zipf traffic: the top key alone is most of a rank's fair share, the hash "
                                  "would leave one rank with 1.5 x the mean)")
            elif placement == "table":
                # count- and key-cost-aware placement (apm_owner_table_build): what a key costs per step, in transactions
                key_cost = int(wl.get("key_cost", 160_000 if args.baseline == "exact" else 4_000))
                counts = np.ascontiguousarray(np.rint(exp), dtype=np.uint64)
                owner = np.empty(self.K_total, np.uint32)
                st = _capi.lib().apm_owner_table_build(counts.ctypes.data, self.K_total, world, key_cost, owner.ctypes.data)
                if st != 0:
                    raise SystemExit("apm_owner_table_build failed")
                self.placement = (f"apm_owner_table_build: key hash (splitmix64 mod ranks) for the light keys, the heavy ones placed "
                                  f"longest-first on the least loaded rank, a key weighing its transactions + {key_cost}")
            else:
                self.placement = "key hash (apm_owner_of_key: splitmix64(key) mod ranks)"
            self.plan = syn.ShardPlan(self.K_total, world, owner)
            mine = self.plan.keys_of[rank]
            self.K = int(mine.size)
            self.n_expected = float(exp[mine].sum())
            self.rank_loads = np.bincount(self.plan.owner, weights=exp, minlength=world)
            base_local = np.ascontiguousarray(self.stream.base[mine % self.spec.num_services], dtype=np.uint32)
            self.base_local_dev = syn.DeviceBuffer(max(base_local.nbytes, 4)).upload(base_local) if self.K else None
        else:
            self.K = self.K_total
            self.n_expected = float(self.total_n)
            self.rank_loads = np.array([float(self.total_n)])
            self.base_local_dev = None
        self.cap_n = int(self.n_expected * 1.03) + (1 << 16)   # transactions one step of this shard may hold
        self.latest = 0                                        # latestBucket of the global stream generated so far

    def new_columns(self):
        return self.syn.DeviceColumns(self.cap_n if self.world > 1 else self.total_n)

    def generate(self, dc, t0_ms, seed):
        """One step of this rank's shard into dc: (transactions, tick marks or None)."""
        if self.world == 1:
            self.stream.generate_device(dc, t0_ms=t0_ms, seed=seed)
            return self.total_n, None
        n, marks, self.latest = self.syn.generate_shard_device(self.stream, self.plan, self.rank, dc, self.wl["bucket_ms"],
                                                               self.latest, t0_ms=t0_ms, seed=seed)
        return n, marks

    def prefill(self, eng, ring_seed):
        # every lag's lists start full, as in a deployment that has run for longer than its longest lag
        for li, L in enumerate(self.wl["lags"]):
            buf = self.syn.DeviceBuffer(L * 3 * self.K * 8)
            if self.world == 1:
                self.syn.ring_fill_device(buf, L, self.stream, ring_seed + li)
            else:
                self.syn.ring_fill_device(buf, L, self.stream, ring_seed + li, base_dev=self.base_local_dev.ptr, num_keys=self.K)
            eng.ring_load_all(li, length=L, device_ptr=buf.ptr)
            buf.free()


def make_engine(wl, shard, flags, local_rank, Engine, EngineConfig):
    per_bucket = shard.n_expected / wl["span_buckets"]
    ecfg = EngineConfig(num_services=shard.K if shard.world > 1 else wl["services"], num_jvms=1 if shard.world > 1 else wl["jvms"],
                        bucket_ms=wl["bucket_ms"], window=wl["window"],
                        buffer=wl["buffer"], percentiles=wl["percentiles"], z_series_pct=wl["z_series_pct"],
                        lags=wl["lags"], thresholds=wl["thresholds"], influences=wl["influences"], flags=flags,
                        max_tx_per_bucket=int(per_bucket * wl.get("log_cap_factor", 1.5)) + 4096 if per_bucket > 4096 else 1 << 16,
                        max_batch=min(shard.cap_n if shard.world > 1 else wl["n"], int(os.environ.get("APM_BENCH_MAX_BATCH", 1 << 25))),
                        out_ticks=wl["span_buckets"] + 12, max_alerts_per_tick=max(shard.K * max(len(wl["lags"]), 1), 1024),
                        device=local_rank)
    return Engine(ecfg)


def broadcast_unique_id(rank, dist, Engine):
    import torch
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.frombuffer(bytearray(Engine.comm_unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(uid, 0)
    return bytes(uid.cpu().numpy().tobytes())


def sharded_parity(rank, world, local_rank, dist, Engine, EngineConfig, _capi, syn):
    """Untimed correctness leg of the N-rank job: a small global stream is binned by owner on the host
    (apm_partition_host), every rank runs its shard on its GPU with the global ticks as marks, the alert records are
    all-gathered over NCCL - on demand for the first half of the stream, streamed per tick for the second - and every
    rank compares (a) what it gathered with the alert records and (b) its own keys' fs records with the fs records of
    ONE unsharded context it runs over the whole stream on the same GPU.  Bit for bit, or the bench exits non-zero."""
    from apmbackend_b200 import sharding
    S, bucket_ms, nb = 48 * world, 10_000, 90
    spec = syn.StreamSpec(seed=4242, n=90_000, num_services=S, num_jvms=1, span_ms=nb * bucket_ms, jitter_ms=1500, shift_mod=5)
    svc, jvm, ts, el = syn.SynthStream(spec, device=False).generate_host()
    kw = dict(bucket_ms=bucket_ms, window=3, buffer=1, percentiles=(75.0, 95.0), z_series_pct=(0, 1), lags=(6, 15),
              thresholds=(1.5, 2.5), influences=(0.3, 0.0), max_tx_per_bucket=1 << 15, max_batch=1 << 12, out_ticks=nb + 8,
              max_alerts_per_tick=2 * S, device=local_rank)
    plan = syn.ShardPlan(S, world)
    mine = plan.keys_of[rank]
    eng = Engine(EngineConfig(num_services=max(int(mine.size), 1), num_jvms=1, **kw))
    eng.comm_init(rank, world, broadcast_unique_id(rank, dist, Engine))
    eng.comm_set_capacity(1 << 13)
    full = Engine(EngineConfig(num_services=S, num_jvms=1, **kw))
    calls = np.linspace(0, svc.size, 7).astype(int)
    latest = 0
    gathered, own_fs, want_fs, want_al = [], [], [], []
    for ci, (a, b) in enumerate(zip(calls[:-1], calls[1:])):
        note(f"parity call {ci}")
        if ci == 3:
            eng.bind_gather_output(1 << 16)  # second half: the gather is streamed tick by tick
        parts, latest = sharding.partition_native(svc[a:b], jvm[a:b], ts[a:b], el[a:b], bucket_ms, S, 1, world, latest=latest,
                                                  local_id=plan.local_id)
        (s_, j_, t_, e_), marks = parts[rank]
        eng.set_tick_marks(marks)
        eng.process_host(s_, j_, t_, e_)
        own_fs.append(eng.fetch_fs().copy())
        gathered.append(eng.gather_alerts(1 << 16, pinned=True).copy())
        eng.release_ticks()
        full.process_host(svc[a:b], jvm[a:b], ts[a:b], el[a:b])
        want_fs.append(full.fetch_fs().copy())
        want_al.append(full.fetch_alerts().copy())
        full.release_ticks()
    ok = True
    g = np.concatenate(gathered)
    for r in range(world):  # gathered records carry their sender's dense ids: back to global keys
        m = g["pad"][:, 2] == r
        g["key"][m] = plan.keys_of[r][g["key"][m]]
    g["pad"][:, 2] = 0
    order = lambda a: a[np.lexsort((a["key"], a["lag"], a["tick_ts"]))]  # noqa: E731
    wa = np.concatenate(want_al)
    ok &= wa.size > 0 and order(g).tobytes() == order(wa).tobytes()
    f = np.concatenate(own_fs)
    f["key"] = mine[f["key"]] if mine.size else f["key"]
    wf = np.concatenate(want_fs)
    wf = wf[np.isin(wf["key"], mine)]
    ok &= order(f).tobytes() == order(wf).tobytes()
    ok &= eng.counters().ticks == full.counters().ticks
    eng.close(); full.close()
    import torch
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) != 1:
        raise SystemExit(f"[rank {rank}] sharded_parity FAILED: gathered / sharded records differ from the unsharded run")
    return "ok", int(wa.size)


def note(msg):
    """progress on stderr (APM_BENCH_VERBOSE=1)"""
    if os.environ.get("APM_BENCH_VERBOSE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} +{time.perf_counter() - _T0:.1f}s] {msg}", file=sys.stderr, flush=True)


_T0 = time.perf_counter()


def main():
    global _REAL_STDOUT
    args = parse_args()
    if os.environ.get("APM_BENCH_WATCHDOG"):  # a hung run dumps every thread's Python stack and exits
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["APM_BENCH_WATCHDOG"]), exit=True)
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # stdout of everything below (C libraries included) joins stderr
    wl = dict(WORKLOADS[args.workload])
    if args.scale != 1.0:
        wl["n"] = int(wl["n"] * args.scale)
        if os.environ.get("APM_BENCH_SCALE_KEYS"):  # debug: shrink the key space with the stream (the line says scaled_down)
            wl["services"] = max(int(wl["services"] * args.scale), 16)
            wl["min_gpus"] = 1
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != max(args.gpus, 1) and world > 1:
        print(f"WORLD_SIZE={world} but --gpus {args.gpus}", file=sys.stderr)
    if world < wl.get("min_gpus", 1) and args.impl != "reference":
        raise SystemExit(f"--workload {args.workload} needs at least {wl['min_gpus']} GPUs (its lag rings do not fit fewer)")
    span_ms = wl["span_buckets"] * wl["bucket_ms"]
    total_n, total_services, scaling = resolve_job(wl, world)
    config = {"workload": f"{args.workload}: {wl['desc']}", "transactions_per_step_total": total_n,
              "keys_total": total_services * wl["jvms"], "ticks_per_step": wl["span_buckets"],
              "l2_policy": "inputs (2 GB/step/GPU) larger than L2, fresh stream each step"}
    if args.scale != 1.0:
        config["scaled_down"] = args.scale

    from apmbackend_b200 import synthetic as syn

    if args.impl == "reference":
        if wl.get("job"):
            raise SystemExit("--impl reference runs the per-GPU workloads (cfg1, cfg2, cfg3, cfg4_shard, cfg5 ...), not the N-rank jobs")
        config["sharding"] = "none"
        return run_reference(args, wl, rank, world, config, span_ms)

    from apmbackend_b200 import Engine, EngineConfig, _capi
    from apmbackend_b200.engine import pack8
    if _capi.device_count() == 0:
        raise SystemExit("bench.py needs a CUDA device: apmbackend_b200 has no CPU fallback")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    _capi.synth().synth_set_device(local_rank)

    def allreduce(x, op="max"):
        if dist is None:
            return float(x)
        import torch
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN, "sum": dist.ReduceOp.SUM}[op])
        return float(t.item())

    parity = None
    if world > 1 and wl["zscore"]:
        note("sharded parity leg")
        parity = sharded_parity(rank, world, local_rank, dist, Engine, EngineConfig, _capi, syn)
        note(f"sharded parity {parity}")

    shard = Shard(args, wl, rank, world, syn, _capi)
    K = shard.K
    flags = _capi.F_KEEP_FS | _capi.F_ALERTS if wl["zscore"] else _capi.F_KEEP_STATS
    if wl["zscore"]:
        flags |= _capi.F_ZSCORE
    if wl["pct"]:
        flags |= _capi.F_PERCENTILES
    inc_flag = _capi.F_INCREMENTAL if (args.baseline == "incremental" and wl["zscore"]) else 0
    if inc_flag:
        config["workload"] = config["workload"].replace("exact-order baseline", "guarded incremental baseline")
        config["baseline_mode"] = "incremental (APM_F_INCREMENTAL): signals and lag lists bit-identical, rolling mean to 1e-9"
    if world == 1:
        config["sharding"] = "none"
    else:
        lo, hi = float(shard.rank_loads.min()), float(shard.rank_loads.max())
        config["sharding"] = (f"ONE global stream per step ({total_n} transactions, {shard.K_total} keys) sharded by key over {world} "
                              f"ranks - {shard.placement}; every rank ingests only the transactions of the keys it owns "
                              f"(expected {lo:.3g} .. {hi:.3g} per rank), ticks are global (apm_set_tick_marks), the alert records "
                              f"are all-gathered over NCCL tick by tick (apm_bind_gather_output)")
        config["keys_per_rank"] = [int(x) for x in shard.plan.counts]
        config["sharded_parity"] = parity[0]
        config["sharded_parity_alerts_compared"] = parity[1]
    ring_seed = 77 + rank
    nsteps = args.warmup + args.steps
    t0 = shard.spec.t0_ms

    def step_t0(s):
        return t0 + s * span_ms

    # records a step's gather may deliver to every rank: all ranks' keys x lags x ticks x the share of records that alert
    # (a few % for 10 k busy keys; sparse keys - cfg4's tail sees one transaction per bucket - alert far more often)
    kmax = int(shard.plan.counts.max()) if shard.plan else K
    gather_out_cap = max(int(kmax * max(len(wl["lags"]), 1) * (wl["span_buckets"] + 2) * world * wl.get("alert_frac", 0.1)), 4096)

    def setup_engine(extra_flags):
        e = make_engine(wl, shard, flags | extra_flags, local_rank, Engine, EngineConfig)
        if world > 1 and wl["zscore"]:
            e.comm_init(rank, world, broadcast_unique_id(rank, dist, Engine))
            e.comm_set_capacity(max(kmax * len(wl["lags"]), 1024))  # records per rank and tick: a tick always fits
            e.bind_gather_output(gather_out_cap)
        if wl["zscore"] and wl.get("prefill", True) and K:
            shard.prefill(e, ring_seed)
        if wl.get("alert_rules"):
            # the alert stage's counter core behind every tick, with the reference's own settings
            # (config/apm_config.json:189-204): the gathered block then holds triggered alerts, not every record with a signal
            e.set_alert_rules()
        return e

    def barrier(e):
        if dist is not None:
            dist.barrier()
        e.sync()

    # ---- value: inputs resident in HBM before the timed region
    dev_steps = []
    for s in range(nsteps + 1):  # one more for the isolated pass
        dc = shard.new_columns()
        n_s, marks = shard.generate(dc, step_t0(s), shard.spec.seed + s)
        dev_steps.append((dc, n_s, marks))

    def timed_device_pass(e, profile):
        gather_n = [0]

        def run_step(s):
            dc, n_s, marks = dev_steps[s]
            if marks is not None:
                e.set_tick_marks(marks)
            e.process_device(dc.columns(), n_s)
            if world > 1 and wl["zscore"]:
                gather_n[0] = len(e.gather_alerts(gather_out_cap, pinned=True))
            e.release_ticks()

        for s in range(args.warmup):
            run_step(s)
        e.profile_read()
        e.profile_enable(profile)
        c0 = e.counters()
        clocks = ClockSampler(local_rank)
        barrier(e)
        clocks.start()
        e.timer_start()
        w0 = time.perf_counter()
        for s in range(args.warmup, nsteps):
            run_step(s)
        ms = e.timer_stop()
        barrier(e)
        wall_ms = (time.perf_counter() - w0) * 1e3
        clk = clocks.stop()
        prof = e.profile_read()
        e.profile_enable(False)
        c1 = e.counters()
        # one more, untimed step with the two-stream overlap switched off: the dominant kernel timed alone
        prof_iso = None
        if wl["zscore"]:
            e.set_overlap(False)
            e.profile_enable(True)
            run_step(nsteps)
            prof_iso = e.profile_read()
            e.profile_enable(False)
            e.set_overlap(True)
        ms = allreduce(ms, "max")
        return dict(ms=ms, wall_ms=wall_ms, clk=clk, prof=prof, prof_iso=prof_iso, launches=c1.kernel_launches - c0.kernel_launches,
                    ticks=c1.ticks - c0.ticks, gathered=gather_n[0], inc_fallbacks=c1.inc_fallbacks - c0.inc_fallbacks)

    note(f"{len(dev_steps)} device-resident steps generated")
    eng = setup_engine(inc_flag)
    note("engine ready")
    main_pass = timed_device_pass(eng, not args.no_profile)
    note("device pass done")
    ms = main_pass["ms"]
    tx_timed = allreduce(sum(dev_steps[s][1] for s in range(args.warmup, nsteps)), "sum")
    value = tx_timed / (ms / 1e3)
    if world > 1:
        print(f"[rank {rank}] {dev_steps[0][1]} transactions, {K} keys per step; {main_pass['gathered']} alert records gathered "
              f"in the last step", file=sys.stderr)

    # ---- e2e: host buffers, H2D inside the timed region, fs + alert records read back.  Measured through the three host
    # entry points: apm_process_host_packed8 (8 B per transaction; the headline where it applies), apm_process_host_packed
    # (12 B) and apm_process_host (20 B, `e2e_columns20`).
    e2e_all = {}
    h2d_ceiling = None
    if not args.no_e2e:
        if wl["zscore"]:
            eng.bind_fs_output()  # fs records stream to the pinned result buffer tick by tick

        def measure_e2e(form, epoch):
            """form: wide | packed12 | packed8.  The packed forms carry endTs relative to a base, so a ring of R distinct
            pinned steps serves all steps (the base moves one span per step); the wide form needs every step distinct and
            is timed over at most 5 steps."""
            warm, steps = (args.warmup, args.steps) if form != "wide" else (min(args.warmup, 2), min(args.steps, 5))
            ns = warm + steps
            R = min(ns, 4) if form != "wide" else ns
            E = step_t0((epoch + 1) * (nsteps + 2))          # time keeps moving forward after everything run so far
            host_steps = []
            tmp = shard.new_columns()
            for r in range(R):
                n_r, marks = shard.generate(tmp, E + r * span_ms, shard.spec.seed + 100 * (epoch + 1) + r)
                pins = [syn.PinnedArray(np.uint32, n_r), syn.PinnedArray(np.uint32, n_r), syn.PinnedArray(np.int64, n_r),
                        syn.PinnedArray(np.int32, n_r)]
                for pa, b, dt in zip(pins, tmp.bufs, (np.uint32, np.uint32, np.int64, np.int32)):
                    _capi.synth().synth_d2h(pa.ptr, b.ptr, n_r * np.dtype(dt).itemsize)
                base = E + r * span_ms
                if form == "packed12":  # the same transactions as three pinned uint32 columns (built outside the timed region)
                    pk = [syn.PinnedArray(np.uint32, n_r) for _ in range(3)]
                    np.left_shift(pins[1].array, 16, out=pk[0].array)
                    np.bitwise_or(pk[0].array, pins[0].array, out=pk[0].array)
                    pk[1].array[:] = pins[2].array - base
                    pk[2].array[:] = pins[3].array.view(np.uint32)
                    for pa in pins:
                        pa.free()
                    host_steps.append((pk, base, marks, None))
                elif form == "packed8":
                    pw = syn.PinnedArray(np.uint64, n_r)
                    _, esc = pack8(pins[0].array, pins[1].array, pins[2].array, pins[3].array, eng.cfg.num_services,
                                   eng.cfg.num_jvms, base, out=pw.array)
                    for pa in pins:
                        pa.free()
                    host_steps.append(([pw], base, marks, esc))
                else:
                    host_steps.append((pins, base, marks, None))
            tmp.free()
            out_bytes = [0]

            def run_step_host(s):
                p, base, marks, esc = host_steps[s % R]
                k = s // R
                shift = k * R * span_ms
                if marks is not None:
                    m = marks.copy()
                    m["label"] += k * R * wl["span_buckets"]
                    # a tick the previous step's last transactions (event-time jitter) already triggered is not repeated;
                    # latestBucket is global, so every rank drops the same marks
                    eng.set_tick_marks(m[m["label"] > eng.latest])
                if form == "packed8":
                    eng.process_host_packed8(p[0].array, esc, base + shift)
                elif form == "packed12":
                    eng.process_host_packed(p[0].array, p[1].array, p[2].array, base + shift)
                else:
                    eng.process_host(p[0].array, p[1].array, p[2].array, p[3].array)
                if wl["zscore"]:
                    fs = eng.fetch_fs(pinned=True)
                    al = eng.gather_alerts(gather_out_cap, pinned=True) if world > 1 else eng.fetch_alerts(pinned=True)
                    out_bytes[0] = fs.nbytes + al.nbytes
                else:
                    st = eng.fetch_stats(pinned=True)
                    out_bytes[0] = st.nbytes
                eng.release_ticks()

            ceiling = None
            if form != "wide" and dist is not None:
                # what the host can deliver: the same pinned bytes copied by all ranks at once, no kernels running
                nbytes = sum(a.array.nbytes for a in host_steps[0][0])
                dst = syn.DeviceBuffer(nbytes)
                barrier(eng)
                tq = time.perf_counter()
                for _ in range(3):
                    off = 0
                    for a in host_steps[0][0]:
                        _capi.synth().synth_h2d(dst.ptr + off, a.ptr, a.array.nbytes)
                        off += a.array.nbytes
                dt_copy = time.perf_counter() - tq
                dst.free()
                worst = allreduce(dt_copy, "max")
                total_bytes = allreduce(3 * nbytes, "sum")
                ceiling = {"aggregate_gbs": total_bytes / worst / 1e9, "per_rank_gbs": 3 * nbytes / dt_copy / 1e9,
                           "bytes_per_rank": nbytes, "how": "all ranks copy one step's pinned input to their GPU at the same time, 3 times, "
                                                            "no kernels running; aggregate = bytes of all ranks / slowest rank's time"}
            for s in range(warm):
                run_step_host(s)
            ce0 = eng.counters()
            barrier(eng)
            eng.timer_start()
            for s in range(warm, ns):
                run_step_host(s)
            ems = eng.timer_stop()
            barrier(eng)
            ce1 = eng.counters()
            ems = allreduce(ems, "max")
            tx = allreduce(sum(host_steps[s % R][0][0].array.size for s in range(warm, ns)), "sum")
            for pins, _, _, _ in host_steps:
                for pa in pins:
                    pa.free()
            api = {"packed8": "apm_process_host_packed8 (8 B per transaction: key 16 | endTs - base 24 | elapsed 24 bits)",
                   "packed12": "apm_process_host_packed (12 B per transaction: key u32, endTs - base u32, elapsed u32)",
                   "wide": "apm_process_host (20 B per transaction: service u32, jvm u32, endTs i64, elapsed i32)"}[form]
            res = {"value": tx / (ems / 1e3), "unit": "transactions/s",
                   "h2d_bytes_per_step": int(allreduce(ce1.h2d_bytes - ce0.h2d_bytes, "sum") // steps),
                   "d2h_bytes_per_step": int(allreduce(ce1.d2h_bytes - ce0.d2h_bytes, "sum") // steps),
                   "ms_per_step": ems / steps, "steps": steps, "warmup": warm, "records_bytes_per_step_per_rank": int(out_bytes[0]),
                   "api": api, "host_buffers": f"{R} distinct pinned steps" + (", reused with the time base moved on by one span per step" if R < ns else "")}
            if ceiling:
                res["h2d_ceiling"] = ceiling
                need = res["h2d_bytes_per_step"] / (ems / steps / 1e3) / 1e9
                res["h2d_achieved_gbs"] = need
                res["h2d_frac_of_ceiling"] = need / ceiling["aggregate_gbs"]
            return res

        can12 = eng.cfg.num_services < 65536 and eng.cfg.num_jvms < 65536 and span_ms + 4000 < (1 << 32)
        can8 = eng.cfg.num_services * eng.cfg.num_jvms < 65536 and span_ms + 4000 < (1 << 24)
        e2e_all["wide"] = measure_e2e("wide", 0)
        note("e2e wide done")
        if can12:
            e2e_all["packed12"] = measure_e2e("packed12", 1)
        if can8:
            e2e_all["packed8"] = measure_e2e("packed8", 2)
    eng.close()
    note("e2e done")

    # ---- the guarded incremental baseline on the same device-resident steps (one more timed pass, fresh context)
    inc_pass = None
    if wl["zscore"] and not inc_flag and not args.no_incremental:
        eng2 = setup_engine(_capi.F_INCREMENTAL)
        inc_pass = timed_device_pass(eng2, not args.no_profile)
        eng2.close()
    for dc, _, _ in dev_steps:
        dc.free()

    # ---- roofline of the dominant kernel (live CUDA-event spans recorded inside the timed region)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    n_rank_step = sum(dev_steps[s][1] for s in range(args.warmup, nsteps)) / max(args.steps, 1)

    def roofline_of(p):
        prof, prof_iso = p["prof"], p["prof_iso"]
        if args.no_profile and prof_iso:
            prof = prof_iso  # nothing was recorded in the timed region: fall back to the isolated pass
        dom = max(prof, key=lambda g: prof[g]["ms"])
        spans = max(prof[dom]["spans"], 1)
        if dom == "rolling_baseline":
            L = wl["lags"][0]
            alg = K * L * 3 * 8 + K * 3 * 8 + K * 128  # ring read + push + fs record, per launch
            if p is inc_pass or inc_flag:
                alg = K * 3 * (8 + 8 + 32) + K * 128     # evicted read + push + running state, + fs record
        elif dom == "keyed_reduce":
            alg = 20 * n_rank_step * args.steps / spans
        elif dom == "radix_sort":
            alg = 16 * n_rank_step * args.steps / spans
        else:
            alg = 0
        avg_ms = prof[dom]["ms"] / spans
        achieved = alg / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        r = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
             "frac": achieved / peak if peak else None, "traffic": None, "peak_source": peak_src,
             "algorithmic_bytes_per_launch": alg, "avg_launch_ms": avg_ms, "launches_timed": prof[dom]["spans"],
             "share_of_step": prof[dom]["ms"] / p["ms"] if p["ms"] else None,
             "kernel_ms": {g: round(v["ms"], 3) for g, v in prof.items()}}
        if prof_iso and prof_iso[dom]["spans"]:
            iso_ms = prof_iso[dom]["ms"] / prof_iso[dom]["spans"]
            r["isolated"] = {"avg_launch_ms": iso_ms, "achieved": alg / (iso_ms * 1e-3) / 1e9,
                             "frac": alg / (iso_ms * 1e-3) / 1e9 / peak,
                             "note": "same kernel, one extra untimed step with the baseline stream's overlap off; "
                                     "`achieved`/`frac` above are live in the timed region, where the kernel shares "
                                     "HBM with the overlapped ingest/sort stream"}
        prof_json = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof_json):
            try:
                r["traffic"] = json.load(open(prof_json)).get(dom if not (p is inc_pass or inc_flag) else dom + "_incremental")
            except Exception:
                pass
        return r

    roofline = roofline_of(main_pass)
    line = {"metric": "transactions/sec aggregated+z-scored", "value": value, "unit": "transactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u32/u64 + f64",
            "data": "synthetic", "config": config, "clocks": main_pass["clk"], "gpu_launches": int(main_pass["launches"]),
            "ticks": int(main_pass["ticks"]), "wall_ms_per_step": main_pass["wall_ms"] / args.steps, "roofline": roofline}
    if world > 1:
        line["alerts_gathered_last_step"] = int(main_pass["gathered"])
    if e2e_all:
        head = "packed8" if "packed8" in e2e_all else ("packed12" if "packed12" in e2e_all else "wide")
        line["e2e"] = e2e_all[head]
        if head != "packed12" and "packed12" in e2e_all:
            line["e2e_packed12"] = e2e_all["packed12"]
        if head != "wide":
            line["e2e_columns20"] = e2e_all["wide"]
    if inc_pass:
        line["incremental"] = {"value": tx_timed / (inc_pass["ms"] / 1e3), "unit": "transactions/s",
                               "ms_per_step": inc_pass["ms"] / args.steps, "roofline": roofline_of(inc_pass),
                               "exact_fallback_columns": int(inc_pass["inc_fallbacks"]),
                               "mode": "APM_F_INCREMENTAL: guarded running sums instead of the exact-order re-sum; signals and lag "
                                       "lists bit-identical, rolling mean / bounds within 1e-9 (north_star's tolerance); same "
                                       "device-resident steps, fresh context"}

    # ---- CPU baseline: the reference-structure port on a bounded sample, rank 0 at N = 1 only
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        hs = syn.SynthStream(shard.spec, device=False)
        cols, n_s = sample_for_cpu(wl, hs, ticks=6)
        threads = 1
        dt, tks = cpu_reference_run(wl, cols, ring_seed, threads, hs)
        line["cpu_baseline"] = {"value": n_s / dt, "unit": "transactions/s", "cores": threads, "kind": "port",
                                "sample": f"first {n_s} transactions of the same stream ({wl.get('cpu_sample_buckets', 6)} of its buckets + the tick "
                                          f"that closes the last, all {K} keys, rings pre-filled): {tks} ticks in {dt:.1f} s, "
                                          "reference-structure C++ port, 1 thread (each reference stage is one "
                                          "single-threaded Node process; Node itself is absent on this box)"}
        # SURVEY 8(d)'s second CPU line: the same results from the optimised variant of the port (one sort per window instead
        # of sorted inserts, one pass per list statistic, a ring instead of shift(); bit-identical,
        # tests/test_oracle_twins.py) with the keys partitioned over all host threads - so that the GPU figure is not
        # only measured against the reference's quadratic insert
        threads = os.cpu_count() or 1
        dt, tks = cpu_reference_run(wl, cols, ring_seed, threads, hs, fast=True)
        line["cpu_baseline_optimised"] = {"value": n_s / dt, "unit": "transactions/s", "cores": threads, "kind": "port, optimised (sort once per window, one pass per list statistic, ring)",
                                          "sample": f"the same {n_s} transactions: {tks} ticks in {dt:.2f} s, keys over {threads} threads"}
    if rank == 0:
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def run_reference(args, wl, rank, world, config, span_ms):
    """--impl reference: the reference's own algorithm (oracle port, fast=0) on the host cores."""
    if rank != 0:
        return
    from apmbackend_b200 import synthetic as syn
    spec = syn.StreamSpec(seed=wl["seed"], n=wl["n"], num_services=wl["services"], num_jvms=wl["jvms"], span_ms=span_ms,
                          shift_mod=97, hot=wl.get("hot"))
    hs = syn.SynthStream(spec, device=False)
    threads = os.cpu_count() or 1
    (svc, jvm, ts, el), n_s = sample_for_cpu(wl, hs, ticks=6)
    times, ticks = [], 0
    for s in range(args.warmup + args.steps):
        # prefill only matters for the z stage's cost; it is rebuilt each step so every step does the same work
        dt, tk = cpu_reference_run(wl, (svc, jvm, ts + s * span_ms, el), 77, threads, hs)
        if s >= args.warmup:
            times.append(dt); ticks += tk
    total = float(np.sum(times))
    v = n_s * args.steps / total
    sample = (f"each step = the first {n_s} transactions of the workload's stream ({wl.get('cpu_sample_buckets', 6)} of its {wl['span_buckets']} buckets + "
              f"the tick that closes the last, all keys, rings pre-filled), reference-structure C++ port of "
              f"stream_calc_stats.js + stream_calc_z_score.js, keys partitioned over {threads} host threads")
    line = {"impl": "reference", "metric": "transactions/sec aggregated+z-scored", "value": v, "unit": "transactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 (JS numbers)",
            "data": "synthetic", "config": config,
            "cpu_baseline": {"value": v, "unit": "transactions/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "transactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "Node.js is absent on this box (node --version: not found), so the reference's own JS cannot be "
                    "timed; this is its CPU restatement"}
    emit(line)


if __name__ == "__main__":
    main()
