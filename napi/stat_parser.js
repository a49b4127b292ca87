'use strict';
/*
 * StatParser over the CUDA engine: the class of stream_calc_stats.js:28-204 with the reference's own method names,
 * its bucket arithmetic replaced by calls into apm_b200.node (napi/apm_b200_addon.c -> include/apm_b200.h).
 *
 * NEVER EXECUTED HERE: there is no node binary in this image or on the GPU box.  The logic is the one of the Python
 * mirror that the test suite does run (apmbackend_b200/stat_parser.py, tests/test_gpu_host_mirror.py), written out in
 * JavaScript for the maintainer who wires the addon into stream_calc_stats.js; the addon's methods it calls are
 * exercised one by one through the fake Node-API (tests/test_napi_shim.py).
 *
 * What stays on the host, as in the reference: the name -> id dictionaries in first-seen order (they define the
 * emission order of Object.entries, :159-160), latestBucket, the min-heap passthrough (unless APM_F_PASSTHROUGH),
 * the entry classes and all queue traffic.  What moves to the device: buckets[label] arrays, removeOldBuckets, the
 * window combine, the sorted multiset and calcPercentile.
 */
const { ApmB200 } = require('./apm_b200.node');
const { StatEntry, EntryFactory } = require('../entries.js');       // the reference's own files, unchanged
const BinaryHeap = require('../binary_heap.js');

const STAT_BYTES = 96;   // apm_stat_record
const F_PASSTHROUGH = 64;

class Columns {          // a growable SoA batch: what addData appends to between two ticks
  constructor(cap = 4096) { this.n = 0; this._alloc(cap); }
  _alloc(cap) {
    const svc = new Uint32Array(cap), jvm = new Uint32Array(cap), ts = new BigInt64Array(cap), el = new Int32Array(cap);
    if (this.n) { svc.set(this.svc.subarray(0, this.n)); jvm.set(this.jvm.subarray(0, this.n)); ts.set(this.ts.subarray(0, this.n)); el.set(this.el.subarray(0, this.n)); }
    this.svc = svc; this.jvm = jvm; this.ts = ts; this.el = el; this.cap = cap;
  }
  push(svc, jvm, endTs, elapsed) {
    if (this.n === this.cap) this._alloc(this.cap * 2);
    this.svc[this.n] = svc; this.jvm[this.n] = jvm; this.ts[this.n] = BigInt(endTs); this.el[this.n] = elapsed;
    this.n++;
  }
  views() { return [this.svc.subarray(0, this.n), this.jvm.subarray(0, this.n), this.ts.subarray(0, this.n), this.el.subarray(0, this.n)]; }
  clear() { this.n = 0; }
}

// one apm_stat_record out of the ArrayBuffer the addon returns
function statRecordAt(view, i) {
  const o = i * STAT_BYTES;
  const nan2undef = (x) => (Number.isNaN(x) ? undefined : x);
  return {
    tickTs: Number(view.getBigInt64(o, true)), key: view.getUint32(o + 8, true), flags: view.getUint32(o + 12, true),
    count: Number(view.getBigUint64(o + 16, true)), sum: Number(view.getBigUint64(o + 24, true)),
    tpm: view.getFloat64(o + 32, true), avg: nan2undef(view.getFloat64(o + 40, true)),
    pct: [0, 1, 2, 3].map((q) => nan2undef(view.getFloat64(o + 48 + 8 * q, true))),
    min: view.getInt32(o + 80, true), max: view.getInt32(o + 84, true), stddev: nan2undef(view.getFloat64(o + 88, true)),
  };
}

class StatParser {
  // options: what `new ApmB200(...)` takes; defaults are config/apm_config.json:105-107, 126-172
  constructor(options = {}) {
    this.options = Object.assign({ numServices: 512, numJvms: 16, bucketMs: 10000, window: 30, buffer: 6,
                                   percentiles: [75, 95], zSeriesPct: [0, 1], lags: [360, 8640], thresholds: [20, 15],
                                   influences: [0.1, 0] }, options);
    this.gpu = new ApmB200(this.options);
    this.entryFactory = new EntryFactory();
    this.setProperties({}, 0, new BinaryHeap((e) => e.endTs));
    this._svcIds = new Map();           // service name -> service id (first-seen order)
    this._svcNames = [];
    this._jvmNames = [];
    this._pending = new Columns();
    this.lastFs = null;                 // fs records of the most recent tick: read by ZScoreParser.processData
    this._deviceHeap = Boolean((this.options.flags || 0) & F_PASSTHROUGH);
    this._txBySeq = new Map();          // APM_F_PASSTHROUGH: arrival sequence number -> TxEntry until the device releases it
    this._seq = 0;
  }

  // :40-52
  setProperties(servers, latestBucket, minHeap) {
    this.servers = servers;             // server -> { id: jvm id, services: { service: key } }
    this.latestBucket = latestBucket;
    this.minHeap = new BinaryHeap((e) => e.endTs);
    for (const obj of (minHeap && minHeap.content) || []) {
      const e = this.entryFactory.getEntryFromObject(obj);
      if (e) this.minHeap.push(e);
    }
  }

  // :89-101, unchanged
  generateBucketLabel(stringMsTs) { return stringMsTs.toString().slice(0, -4); }
  generateTsFromBucketLabel(bucketLab) { return parseInt(`${bucketLab}0000`); }

  // :103-113: eviction is fused into the device tick (slots of labels below latest - numKeep are cleared there)
  removeOldBuckets(numKeepIntervals) {
    if (numKeepIntervals !== this.options.window + this.options.buffer) throw new Error('numKeepIntervals differs from window + buffer of the engine');
  }

  // ---- :115-134
  _ids(server, service) {
    let srv = this.servers[server];
    if (!srv) {
      if (this._jvmNames.length >= this.options.numJvms) throw new Error('more servers than numJvms');
      srv = this.servers[server] = { id: this._jvmNames.length, services: {} };
      this._jvmNames.push(server);
    }
    let sid = this._svcIds.get(service);
    if (sid === undefined) {
      if (this._svcNames.length >= this.options.numServices) throw new Error('more services than numServices');
      sid = this._svcNames.length;
      this._svcIds.set(service, sid);
      this._svcNames.push(service);
    }
    if (!(service in srv.services)) srv.services[service] = srv.id * this.options.numServices + sid;
    return [sid, srv.id];
  }

  addData(txEntry, bucketLabel) {      // eslint-disable-line no-unused-vars
    const [sid, jid] = this._ids(txEntry.server, txEntry.service);
    this._pending.push(sid, jid, txEntry.endTs, parseInt(txEntry.elapsed));
    if (this._deviceHeap) { this._txBySeq.set(this._seq & 0xFFFFFF, txEntry); this._seq++; } else this.minHeap.push(txEntry);
  }

  _flush() {
    if (this._pending.n) { this.gpu.addBatch(...this._pending.views()); this._pending.clear(); }
  }

  // ---- :136-155
  writeHeapToQueue(ts, outQueue) {
    if (this._deviceHeap) { this._releasedQueue = outQueue; return; }   // the device releases them during the tick that follows
    for (const el of this.minHeap.popAllLessOrEqualToScore(ts)) outQueue.writeLineToQueue(el.toCSVString(), false);
  }

  // ---- :157-203: one tick of the device replaces the three nested loops
  generateAllStatsToQueue(ts, intervalLengthSec, windowSz, numKeepIntervals, intervalBufferSz, outQueue) {
    const o = this.options;
    if (intervalLengthSec * 1000 !== o.bucketMs || windowSz !== o.window || intervalBufferSz !== o.buffer) {
      throw new Error('interval / window / buffer differ from the engine configuration');
    }
    this._flush();
    const res = this.gpu.tick(parseInt(this.latestBucket));   // apm_tick + fetches + apm_release_ticks
    if (this._deviceHeap && res.passthrough) {
      const seqs = new Uint32Array(res.passthrough.seq);
      for (const sq of seqs) {
        const tx = this._txBySeq.get(sq);
        this._txBySeq.delete(sq);
        if (tx && this._releasedQueue) this._releasedQueue.writeLineToQueue(tx.toCSVString(), false);
      }
    }
    this.lastFs = res.fs || null;
    this.lastGathered = res.gathered || null;                  // multi-GPU: every rank's alert records of this tick
    const view = new DataView(res.stats);
    const byKey = new Map();
    for (let i = 0; i < res.stats.byteLength / STAT_BYTES; i++) { const r = statRecordAt(view, i); byKey.set(r.key, r); }
    const [pa, pb] = o.zSeriesPct;
    const out = [];
    for (const [server, srv] of Object.entries(this.servers)) {           // first-seen order, :159-160
      for (const [service, key] of Object.entries(srv.services)) {
        const r = byKey.get(key);
        if (!r) continue;                                                  // named by addData, no valid transaction yet
        const e = new StatEntry(ts, server, service, r.tpm, r.avg, r.pct[pa], r.pct[pb]);
        out.push(e);
        outQueue.writeLineToQueue(e.toCSVString(), false);
      }
    }
    return out;
  }

  nameOfKey(key) { const S = this.options.numServices; return [this._jvmNames[Math.floor(key / S)], this._svcNames[key % S]]; }

  // ---- :54-87: the device state goes to `<resumeFile>.bin`; names, latestBucket and the heap stay JSON
  saveToResumeFile(resumeFile, quiet) {   // eslint-disable-line no-unused-vars
    const fs = require('fs');
    this._flush();
    this.gpu.saveState(`${resumeFile}.bin`);
    fs.writeFileSync(resumeFile, JSON.stringify({ servers: this.servers, latestBucket: this.latestBucket, services: this._svcNames,
                                                  jvms: this._jvmNames, minHeap: { content: this.minHeap.content } }));
  }

  resumeFromFileIfExists(resumeFile) {
    const fs = require('fs');
    if (!fs.existsSync(resumeFile) || !fs.existsSync(`${resumeFile}.bin`)) return false;
    const doc = JSON.parse(fs.readFileSync(resumeFile, 'utf8'));
    this.gpu.loadState(`${resumeFile}.bin`);
    this._svcNames = doc.services; this._jvmNames = doc.jvms;
    this._svcIds = new Map(this._svcNames.map((s, i) => [s, i]));
    this.setProperties(doc.servers, doc.latestBucket, doc.minHeap);
    return true;
  }
}

module.exports = { StatParser, Columns, statRecordAt };
