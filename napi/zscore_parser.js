'use strict';
/*
 * ZScoreParser over the CUDA engine: the class of stream_calc_z_score.js:26-312 with the reference's own method names.
 * The lag lists (avgList / per75List / per95List) live on the device as time-major rings and processZScoreStats runs
 * there, fused behind every tick of the stats stage; this class keeps what the reference keeps on the host - the
 * settings resolution (getServiceSettingsFromConfig and friends, with the defaults-aliasing behaviour of :107-120) -
 * and turns the device's 128-byte fs records into FullStatEntry objects.
 *
 * NEVER EXECUTED HERE: no node binary exists in this image or on the GPU box.  Same logic as the Python mirror the
 * test suite runs (apmbackend_b200/zscore_parser.py); the addon methods it calls are covered through the fake
 * Node-API (tests/test_napi_shim.py).
 */
const { FullStatEntry } = require('../entries.js');   // the reference's own file, unchanged

const FS_BYTES = 128;   // apm_fs_record

// one apm_fs_record out of the ArrayBuffer the addon returns; NaN stands for JS `undefined`
function fsRecordAt(view, i) {
  const o = i * FS_BYTES;
  const u = (x) => (Number.isNaN(x) ? undefined : x);
  const f = (off) => u(view.getFloat64(o + off, true));
  const rec = { tickTs: Number(view.getBigInt64(o, true)), key: view.getUint32(o + 8, true), lag: view.getUint32(o + 12, true),
                tpm: view.getFloat64(o + 16, true), val: [], avg: [], lb: [], ub: [], signal: [],
                flags: view.getUint8(o + 123), causes: view.getUint8(o + 124), badIntervals: view.getUint8(o + 125),
                rank: view.getUint8(o + 126) };
  for (let s = 0; s < 3; s++) {
    rec.val.push(f(24 + 8 * s)); rec.avg.push(f(48 + 8 * s)); rec.lb.push(f(72 + 8 * s)); rec.ub.push(f(96 + 8 * s));
    rec.signal.push(view.getInt8(o + 120 + s));
  }
  return rec;
}

class ZScoreParser {
  // statParser: the StatParser whose ticks produce the fs records; zscoreConfig mirrors ZSCORECONFIG
  // ({ defaults: [{LAG, THRESHOLD, INFLUENCE}, ...], overrides: { services: { name: { lag: {...} } } } }, config/apm_config.json:126-172)
  constructor(statParser, zscoreConfig) {
    this.sp = statParser;
    this.gpu = statParser.gpu;
    const o = statParser.options;
    this.ZSCORECONFIG = zscoreConfig || { defaults: o.lags.map((l, i) => ({ LAG: l, THRESHOLD: o.thresholds[i], INFLUENCE: o.influences[i] })),
                                          overrides: { services: {} } };
    this.setProperties({});
  }

  setProperties(servers) { this.servers = servers; }

  // the arithmetic of :66-104 for one list, for callers that still hold a plain array (not used by processData):
  // average() = left-to-right sum of the valid entries, "stddev" = sqrt(mean) (util_methods.js:10-50)
  processZScoreStats(lag, threshold, influence, newValue, prevValuesList) {
    let inflNewValue = newValue, avg, stdDev, lb, ub, signal = 0;
    if (prevValuesList.length >= lag) {
      let sum = 0, cnt = 0;
      for (const v of prevValuesList) if (v != null && !Number.isNaN(parseFloat(v))) { sum += parseFloat(v); cnt++; }
      avg = cnt ? sum / cnt : undefined;
      stdDev = avg ? Math.sqrt(avg) : undefined;
      if (avg !== undefined && stdDev !== undefined) { lb = avg - threshold * stdDev; ub = avg + threshold * stdDev; }
      if (avg === undefined || stdDev === undefined || Number.isNaN(avg) || Number.isNaN(stdDev)) signal = 0;
      else if (Number.isNaN(newValue)) signal = 0;
      else if (Math.abs(newValue - avg) > threshold * stdDev) {
        signal = newValue > avg ? 1 : -1;
        const last = prevValuesList[prevValuesList.length - 1];
        if (last != null && !Number.isNaN(last)) inflNewValue = influence * newValue + (1 - influence) * last;
      }
    }
    return [inflNewValue, avg, lb, ub, signal];
  }

  // ---- :106-132 (aliases and mutates the defaults, like the reference)
  getServiceSettingsFromConfig(service) {
    const settings = this.ZSCORECONFIG.defaults;
    const ov = this.ZSCORECONFIG.overrides && this.ZSCORECONFIG.overrides.services && this.ZSCORECONFIG.overrides.services[service];
    if (ov) {
      for (const el of settings) {
        for (const lag of Object.keys(ov)) {
          if (String(el.LAG) === String(lag)) {
            if (ov[lag].THRESHOLD) el.THRESHOLD = ov[lag].THRESHOLD;
            if (ov[lag].INFLUENCE) el.INFLUENCE = ov[lag].INFLUENCE;
          }
        }
      }
    }
    return settings;
  }

  // ---- :134-150: the resolved values go to the device arrays of this key
  updateServiceSettings(server, service) {
    const lags = this.servers[server].services[service].lags;
    const key = this.sp.servers[server].services[service];
    for (const el of this.getServiceSettingsFromConfig(service)) {
      const k = String(el.LAG);
      lags[k] = lags[k] || {};
      lags[k].THRESHOLD = el.THRESHOLD; lags[k].INFLUENCE = el.INFLUENCE;
      const li = this.sp.options.lags.indexOf(parseInt(el.LAG));
      if (li >= 0) this.gpu.setKeyParams(key, li, Number(el.THRESHOLD), Number(el.INFLUENCE));
    }
  }

  // ---- :152-167
  updateAllServiceSettings() {
    for (const [server, srv] of Object.entries(this.servers)) for (const service of Object.keys(srv.services)) this.updateServiceSettings(server, service);
  }

  // ---- :170-193: a lag that left the configuration loses its lists
  removeStaleLagData() {
    const configured = new Set(this.ZSCORECONFIG.defaults.map((el) => parseInt(el.LAG)));
    this.sp.options.lags.forEach((lag, li) => { if (!configured.has(parseInt(lag))) this.gpu.dropLag(li); });
    for (const srv of Object.values(this.servers)) for (const svc of Object.values(srv.services)) {
      for (const lag of Object.keys(svc.lags)) if (!configured.has(parseInt(lag))) delete svc.lags[lag];
    }
  }

  // first sight of a (server, service), :201-208: call before the tick that first emits it, so that its overrides are
  // on the device when that tick runs
  noteService(server, service) {
    const srv = this.servers[server] || (this.servers[server] = { services: {} });
    if (!srv.services[service]) { srv.services[service] = { lags: {} }; this.updateServiceSettings(server, service); }
  }

  // ---- :195-311: the arithmetic already ran on the device behind the tick; this finds the records of
  // (timestamp, server, service) and builds one FullStatEntry per lag, lags ascending (:217)
  processData(statEntry) {
    this.noteService(statEntry.server, statEntry.service);
    const buf = this.sp.lastFs;
    if (!buf) return [];
    if (this._indexed !== buf) {      // one pass per tick: (key, lag) -> record index
      this._view = new DataView(buf);
      this._index = new Map();
      for (let i = 0; i < buf.byteLength / FS_BYTES; i++) {
        const o = i * FS_BYTES;
        this._index.set(`${this._view.getUint32(o + 8, true)}:${this._view.getUint32(o + 12, true)}:${this._view.getBigInt64(o, true)}`, i);
      }
      this._indexed = buf;
    }
    const key = this.sp.servers[statEntry.server].services[statEntry.service];
    const out = [];
    for (const lag of [...this.sp.options.lags].sort((a, b) => a - b)) {
      const i = this._index.get(`${key}:${lag}:${BigInt(statEntry.timestamp)}`);
      if (i === undefined) continue;
      const r = fsRecordAt(this._view, i);
      const a = [];
      for (let s = 0; s < 3; s++) a.push(r.val[s], r.avg[s], r.lb[s], r.ub[s], r.signal[s]);
      out.push(new FullStatEntry(r.tickTs, statEntry.server, statEntry.service, r.tpm, String(r.lag), ...a));
    }
    return out;
  }

  // the lists are part of the engine image StatParser.saveToResumeFile writes
  saveToResumeFile(resumeFile, quiet) { this.sp.saveToResumeFile(resumeFile, quiet); }
  resumeFromFileIfExists(resumeFile) {
    const ok = this.sp.resumeFromFileIfExists(resumeFile);
    this.updateAllServiceSettings();
    return ok;
  }
}

module.exports = { ZScoreParser, fsRecordAt };
