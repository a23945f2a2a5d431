"""ctypes mirror of `FlybModel` (include/flyb_model.h), built from model_fields.py."""
from __future__ import annotations

import ctypes
import os
import numpy as np

from .model_fields import SCALAR_INT, SCALAR_DBL, ARRAYS

_CT = {"i": ctypes.POINTER(ctypes.c_int32), "d": ctypes.POINTER(ctypes.c_double)}
_NP = {"i": np.int32, "d": np.float64}


class FlybModel(ctypes.Structure):
    _fields_ = ([(n, ctypes.c_int32) for n in SCALAR_INT] + [(n, ctypes.c_double) for n in SCALAR_DBL]
                + [(n, _CT[t]) for n, t, _ in ARRAYS])


def pack_model(M: dict):
    """dict of arrays (compile_fly.compile_task / load_model) -> (FlybModel, keepalive list)."""
    s = FlybModel()
    keep = []
    for n in SCALAR_INT:
        setattr(s, n, int(M[n]))
    for n in SCALAR_DBL:
        setattr(s, n, float(M[n]))
    for n, t, shape in ARRAYS:
        dims = tuple(int(M[d]) if isinstance(d, str) else d for d in shape)
        a = np.ascontiguousarray(np.asarray(M[n], dtype=_NP[t]).reshape(dims))
        keep.append(a)
        setattr(s, n, a.ctypes.data_as(_CT[t]))
    return s, keep


def asset_path(task: str) -> str:
    return os.path.join(os.path.dirname(__file__), "assets", f"{task}.npz")
