"""ppqsflhe_b200 -- B200-native (sm_100a) implementation of the PPQSFLHE encrypted-aggregation hot path.

The product is the C-ABI shared library `ppqsflhe_b200/lib/libppq_b200.so` (include/ppq_b200.h); this package is
the thin host-side mirror of the OpenFHE CryptoContext surface the reference calls (context.py).  There is no CPU
fallback and nothing here imports `oracle/`.
"""
from ._capi import LIB_PATH, PPQError, build, lib  # noqa: F401


def __getattr__(name):  # torch is imported lazily so that `build()` works without touching CUDA
    if name in ("CryptoContext", "CiphertextBatch", "PlaintextBatch", "EvalKey", "CtMeta", "to_device_u64",
                "to_host_u64", "Aggregator", "Communicator"):
        from . import context
        return getattr(context, name)
    raise AttributeError(name)
