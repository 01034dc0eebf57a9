"""B200-native replacement for the solve hot path of KumarRobotics/multicam_calibration.

The product is the C-ABI shared library ``libmcc.so`` (declared in ``include/mcc.h``, built from
``multicam_calibration_b200/csrc``); this package is the thin Python host layer used by the tests and
``bench.py``: a flat problem container (``rig``), input generators (``synth``) and the ctypes binding (``capi``).
"""
from .rig import Camera, Rig, shard_bounds  # noqa: F401
