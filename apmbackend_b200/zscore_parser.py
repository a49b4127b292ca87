"""Host side of the z-score stage with the reference's class and method names
(stream_calc_z_score.js:26-312).  The lag lists live on the device as time-major rings;
the host keeps the settings resolution (getServiceSettingsFromConfig and friends, with the
reference's defaults-aliasing behaviour) and turns the device's fs records into
FullStatEntry objects."""
from __future__ import annotations

import numpy as np

from .entries import FullStatEntry
from .stat_parser import StatParser, _u


class ZScoreParser:
    def __init__(self, stat_parser: StatParser, zscore_config=None):
        """zscore_config mirrors ZSCORECONFIG: {"defaults": [{"LAG","THRESHOLD","INFLUENCE"}...],
        "overrides": {"services": {name: {lag: {...}}}}} (config/apm_config.json:126-172)."""
        self.sp = stat_parser
        self.engine = stat_parser.engine
        cfg = self.engine.cfg
        self.ZSCORECONFIG = zscore_config or {
            "defaults": [{"LAG": l, "THRESHOLD": t, "INFLUENCE": i}
                         for l, t, i in zip(cfg.lags, cfg.thresholds, cfg.influences)],
            "overrides": {"services": {}}}
        self.setProperties()

    def setProperties(self, servers=None):
        self.servers = servers if servers is not None else {}

    # ---- :106-132 (aliases and mutates the defaults, like the reference)
    def getServiceSettingsFromConfig(self, service):
        settings = self.ZSCORECONFIG["defaults"]
        ov = self.ZSCORECONFIG.get("overrides", {}).get("services", {}).get(service)
        if ov:
            for el in settings:
                for lag in ov:
                    if str(el["LAG"]) == str(lag):
                        if ov[lag].get("THRESHOLD"):
                            el["THRESHOLD"] = ov[lag]["THRESHOLD"]
                        if ov[lag].get("INFLUENCE"):
                            el["INFLUENCE"] = ov[lag]["INFLUENCE"]
        return settings

    # ---- :134-150: resolved values go to the device arrays of this key
    def updateServiceSettings(self, server, service):
        lags = self.servers[server]["services"][service]["lags"]
        key = self.sp.servers[server]["services"][service]
        for el in self.getServiceSettingsFromConfig(service):
            k = str(el["LAG"])
            lags.setdefault(k, {})
            lags[k]["THRESHOLD"], lags[k]["INFLUENCE"] = el["THRESHOLD"], el["INFLUENCE"]
            if int(el["LAG"]) in self.engine.cfg.lags:
                li = list(self.engine.cfg.lags).index(int(el["LAG"]))
                self.engine.set_key_params(key, li, [float(el["THRESHOLD"])], [float(el["INFLUENCE"])])

    # ---- :152-167
    def updateAllServiceSettings(self):
        for server, srv in self.servers.items():
            for service in srv["services"]:
                self.updateServiceSettings(server, service)

    # ---- :170-193
    def removeStaleLagData(self):
        configured = {int(el["LAG"]) for el in self.ZSCORECONFIG["defaults"]}
        for li, lag in enumerate(self.engine.cfg.lags):
            if int(lag) not in configured:
                self.engine.drop_lag(li)
        for srv in self.servers.values():
            for svc in srv["services"].values():
                for lag in [l for l in svc["lags"] if int(l) not in configured]:
                    del svc["lags"][lag]

    def noteService(self, server, service):
        """First sight of a (server, service): :201-208.  Call before the tick that first
        emits it so its overrides are on the device when the tick runs."""
        srv = self.servers.setdefault(server, {"services": {}})
        if service not in srv["services"]:
            srv["services"][service] = {"lags": {}}
            self.updateServiceSettings(server, service)

    # ---- :195-311: the arithmetic already ran on the device behind the tick; this looks the
    # records of (timestamp, server, service) up and builds the FullStatEntry per lag.
    def processData(self, statEntry):
        self.noteService(statEntry.server, statEntry.service)
        fs = self.sp.lastFsRecords
        key = self.sp.servers[statEntry.server]["services"][statEntry.service]
        rows = fs[(fs["key"] == key) & (fs["tick_ts"] == int(statEntry.timestamp))]
        out = []
        for r in rows[np.argsort(rows["lag"])]:
            a = []
            for s in range(3):
                a += [_u(r["val"][s]), _u(r["avg"][s]), _u(r["lb"][s]), _u(r["ub"][s]), int(r["signal"][s])]
            out.append(FullStatEntry(int(r["tick_ts"]), statEntry.server, statEntry.service, float(r["tpm"]),
                                     str(int(r["lag"])), *a))
        return out

    def saveToResumeFile(self, resumeFile, quiet=True):
        """The lists are part of the engine image StatParser.saveToResumeFile writes."""
        self.sp.saveToResumeFile(resumeFile, quiet)

    def resumeFromFileIfExists(self, resumeFile):
        ok = self.sp.resumeFromFileIfExists(resumeFile)
        self.updateAllServiceSettings()
        return ok
