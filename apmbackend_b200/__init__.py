"""
apmbackend_b200 - B200-native drop-in for APMBackend's two hot stream stages
(stream_calc_stats.js, stream_calc_z_score.js).

The arithmetic lives in hand-written sm_100a CUDA kernels behind the C ABI of
include/apm_b200.h (libapm_b200.so, built in-tree from csrc/).  This package is
the host side above that ABI: a ctypes binding (`_capi`) and mirrors of the
reference's own classes (`StatParser`, `ZScoreParser`, entry types) so the host
keeps its module / message API.  There is no CPU fallback: importing works
anywhere, but every compute call needs the CUDA library and a device.
"""
from .entries import EntryFactory, FullStatEntry, StatEntry, TxEntry  # noqa: F401
from .engine import Engine, EngineConfig  # noqa: F401
from .stat_parser import StatParser  # noqa: F401
from .zscore_parser import ZScoreParser  # noqa: F401

__all__ = ["Engine", "EngineConfig", "StatParser", "ZScoreParser", "TxEntry", "StatEntry", "FullStatEntry",
           "EntryFactory"]
