"""Host side of the stats stage with the reference's own class and method names
(stream_calc_stats.js:28-204) over the CUDA engine.  The per-bucket arrays the reference
keeps in `servers[srv].services[svc].buckets` live on the device as (key, bucket)
accumulators and value logs; this class keeps what must stay on the host: the
name -> id dictionaries in first-seen order (they define the emission order of
Object.entries, :159-160), latestBucket, and the min-heap passthrough."""
from __future__ import annotations

import json
import os
from typing import List, Optional

import numpy as np

from .binary_heap import BinaryHeap
from .engine import Engine, EngineConfig
from .entries import EntryFactory, StatEntry, TxEntry

NaN = float("nan")


class ListQueue:
    """Stand-in for a ProducerQueue (queue.js:206-262): collects the lines written to it."""

    def __init__(self):
        self.lines: List[str] = []

    def writeLineToQueue(self, line, verbose=False):
        self.lines.append(line)


class StatParser:
    def __init__(self, engine: Optional[Engine] = None, *, max_services=256, max_jvms=16, bucket_ms=10000,
                 window=30, buffer=6, engine_kwargs=None):
        if engine is None:
            engine = Engine(EngineConfig(num_services=max_services, num_jvms=max_jvms, bucket_ms=bucket_ms,
                                         window=window, buffer=buffer, **(engine_kwargs or {})))
        self.engine = engine
        self.bucket_ms = engine.cfg.bucket_ms
        self.entryFactory = EntryFactory()
        self.setProperties({}, 0, BinaryHeap(lambda tx: tx.endTs))
        self._svc_ids = {}        # service name -> service id
        self._svc_names = []
        self._jvm_names = []
        self._pending = ([], [], [], [])  # service_id, jvm_id, endTs, elapsed
        self.lastFsRecords = None  # fs records of the most recent tick (read by ZScoreParser)
        # With APM_F_PASSTHROUGH the device keeps the reorder buffer: the host only remembers each
        # transaction by its arrival sequence number until the device releases it.
        self._device_heap = bool(engine.cfg.flags & 64)
        self._tx_by_seq = {}
        self._seq = 0
        self._released = []

    # ---- :40-52
    def setProperties(self, servers, latestBucket, minHeap):
        self.servers = servers          # server -> {"id": jvm id, "services": {service: key}}
        self.latestBucket = latestBucket
        self.minHeap = BinaryHeap(lambda tx: tx.endTs)
        for obj in getattr(minHeap, "content", []) or []:
            e = obj if isinstance(obj, TxEntry) else self.entryFactory.getEntryFromObject(obj)
            if e is not None:
                self.minHeap.push(e)

    # ---- :89-101
    def generateBucketLabel(self, stringMsTs):
        if self.bucket_ms == 10000:
            return str(stringMsTs)[:-4]   # the reference's slice(0,-4)
        return str(int(stringMsTs) // self.bucket_ms)

    def generateTsFromBucketLabel(self, bucketLab):
        if self.bucket_ms == 10000:
            return int(f"{int(bucketLab)}0000")
        return int(bucketLab) * self.bucket_ms

    # ---- :103-113.  Eviction is fused into the device tick (slots of labels below
    # latest - numKeep are cleared there); nothing is left to do on the host.
    def removeOldBuckets(self, numKeepIntervals):
        if numKeepIntervals != self.engine.cfg.window + self.engine.cfg.buffer:
            raise ValueError("numKeepIntervals differs from the engine's window + buffer")

    # ---- :115-134
    def _server(self, server):
        srv = self.servers.get(server)
        if srv is None:
            if len(self._jvm_names) >= self.engine.cfg.num_jvms:
                raise OverflowError("more servers than EngineConfig.num_jvms")
            srv = self.servers[server] = {"id": len(self._jvm_names), "services": {}}
            self._jvm_names.append(server)
        return srv

    def _ids(self, server, service):
        srv = self._server(server)
        sid = self._svc_ids.get(service)
        if sid is None:
            if len(self._svc_names) >= self.engine.cfg.num_services:
                raise OverflowError("more services than EngineConfig.num_services")
            sid = self._svc_ids[service] = len(self._svc_names)
            self._svc_names.append(service)
        if service not in srv["services"]:
            srv["services"][service] = srv["id"] * self.engine.cfg.num_services + sid
        return sid, srv["id"]

    def addData(self, txEntry, bucketLabel):
        sid, jid = self._ids(txEntry.server, txEntry.service)
        p = self._pending
        p[0].append(sid); p[1].append(jid); p[2].append(int(txEntry.endTs)); p[3].append(int(txEntry.elapsed))
        if self._device_heap:
            self._tx_by_seq[self._seq & 0xFFFFFF] = txEntry
            self._seq += 1
        else:
            self.minHeap.push(txEntry)

    def _flush(self):
        p = self._pending
        if p[0]:
            self.engine.add_batch(np.array(p[0], np.uint32), np.array(p[1], np.uint32), np.array(p[2], np.int64),
                                  np.array(p[3], np.int32))
            self._pending = ([], [], [], [])

    # ---- :136-155
    def writeHeapToQueue(self, ts, outQueue):
        if self._device_heap:
            # the device releases them during the tick that follows (same edgeTs): remember where they go
            self._released_queue = outQueue
            return
        for el in self.minHeap.popAllLessOrEqualToScore(self.minHeap.scoreFunction(_Score(ts))):
            outQueue.writeLineToQueue(el.toCSVString(), False)

    # ---- :157-203
    def generateAllStatsToQueue(self, ts, intervalLengthSec, windowSz, numKeepIntervals, intervalBufferSz, outQueue):
        cfg = self.engine.cfg
        if (intervalLengthSec * 1000, windowSz, intervalBufferSz) != (cfg.bucket_ms, cfg.window, cfg.buffer):
            raise ValueError("interval/window/buffer differ from the engine's configuration")
        self._flush()
        self.engine.tick(int(self.latestBucket))
        if self._device_heap:
            _, seqs, _ = self.engine.fetch_passthrough()
            q = getattr(self, "_released_queue", None)
            for sq in seqs.tolist():
                tx = self._tx_by_seq.pop(sq)
                if q is not None:
                    q.writeLineToQueue(tx.toCSVString(), False)
        stats = self.engine.fetch_stats()
        self.lastFsRecords = self.engine.fetch_fs().copy() if cfg.lags else None
        self.engine.release_ticks()
        by_key = {int(r["key"]): r for r in stats}
        pa, pb = cfg.z_series_pct
        out = []
        for server, srv in self.servers.items():            # first-seen order, as Object.entries
            for service, key in srv["services"].items():
                r = by_key.get(key)
                if r is None:
                    continue  # named by addData but its batch holds no valid transaction yet
                if int(r["tick_ts"]) != int(ts):
                    raise ValueError("ts differs from the tick timestamp the engine derived from latestBucket")
                e = StatEntry(ts, server, service, float(r["tpm"]), _u(r["avg"]), _u(r["pct"][pa]), _u(r["pct"][pb]))
                out.append(e)
                outQueue.writeLineToQueue(e.toCSVString(), False)
        return out

    # ---- consumeMsg, :331-371 (a closure in the reference's main; a method here)
    def consumeMsg(self, line, outQueue, dbQueue):
        txEntry = self.entryFactory.getEntryFromCSV(line)
        cfg = self.engine.cfg
        bucketLabel = self.generateBucketLabel(txEntry.endTs)
        newer = (int(bucketLabel or 0) > self.latestBucket) if isinstance(self.latestBucket, int) \
            else bucketLabel > self.latestBucket   # number compare first, string compare afterwards (:348)
        if newer:
            self.latestBucket = bucketLabel
            self.removeOldBuckets(cfg.window + cfg.buffer)
            edgeTs = self.generateTsFromBucketLabel(int(self.latestBucket) - cfg.buffer - 1)
            self.writeHeapToQueue(edgeTs, dbQueue)
            self.generateAllStatsToQueue(edgeTs, cfg.bucket_ms // 1000, cfg.window, cfg.window + cfg.buffer, cfg.buffer,
                                         outQueue)
        self.addData(txEntry, bucketLabel)

    # ---- batch fast path (addBatch / tick of SURVEY 8b): columns in, lines out, ticks found on the device
    def consumeColumns(self, servers, services, endTs, elapsed, outQueue):
        sid = np.empty(len(servers), np.uint32)
        jid = np.empty(len(servers), np.uint32)
        for i, (srv, svc) in enumerate(zip(servers, services)):
            sid[i], jid[i] = self._ids(srv, svc)
        self._flush()
        res = self.engine.process_host(sid, jid, endTs, elapsed)
        stats = self.engine.fetch_stats()
        self.lastFsRecords = self.engine.fetch_fs().copy() if self.engine.cfg.lags else None
        self.engine.release_ticks()
        if res.ticks:
            self.latestBucket = str(res.latest_bucket)
        pa, pb = self.engine.cfg.z_series_pct
        order = {}
        for server, srv in self.servers.items():
            for service, key in srv["services"].items():
                order[key] = (len(order), server, service)
        for ts in np.unique(stats["tick_ts"]):
            tick = stats[stats["tick_ts"] == ts]
            for r in sorted(tick, key=lambda r: order[int(r["key"])][0]):
                _, server, service = order[int(r["key"])]
                e = StatEntry(int(ts), server, service, float(r["tpm"]), _u(r["avg"]), _u(r["pct"][pa]), _u(r["pct"][pb]))
                outQueue.writeLineToQueue(e.toCSVString(), False)
        return res

    def name_of_key(self, key):
        S = self.engine.cfg.num_services
        return self._jvm_names[key // S], self._svc_names[key % S]

    # ---- :54-87.  The device state goes to `<resumeFile>.bin`; names, latestBucket and the heap to JSON.
    def saveToResumeFile(self, resumeFile, quiet=True):
        self._flush()
        self.engine.state_save(resumeFile + ".bin")
        doc = {"servers": self.servers, "latestBucket": self.latestBucket, "services": self._svc_names,
               "jvms": self._jvm_names, "minHeap": {"content": [vars(e) for e in self.minHeap.content]}}
        with open(resumeFile, "w") as f:
            json.dump(doc, f)

    def resumeFromFileIfExists(self, resumeFile):
        if not (os.path.exists(resumeFile) and os.path.exists(resumeFile + ".bin")):
            return False
        with open(resumeFile) as f:
            doc = json.load(f)
        self.engine.state_load(resumeFile + ".bin")
        self._svc_names, self._jvm_names = doc["services"], doc["jvms"]
        self._svc_ids = {s: i for i, s in enumerate(self._svc_names)}

        class _H:
            content = doc["minHeap"]["content"]
        self.setProperties(doc["servers"], doc["latestBucket"], _H)
        return True


class _Score:
    def __init__(self, ts):
        self.endTs = ts


def _u(x):
    """NaN record fields are JS `undefined`."""
    x = float(x)
    return None if x != x else x
