"""Host-side mirror of the reference's estep.py (the py_func boundary, estep.py:18-31).

Same names, argument order, defaults and in/out types as the reference module, so
`from em_adapt_tensorflow_b200.estep import estep as _estep` replaces
`from estep import estep as _estep` (deeplab.py:9).  Everything is routed to
libweaklabel_cuda.so's `e_step` (the C-ABI symbol of estep/weaklabel.cpp:7); there is no NumPy
path: `use_c` is accepted for signature compatibility only (the reference's use_c=False branch,
estep_py, is a numerically different algorithm -- SURVEY.md 8-a -- and not the parity target).
"""
from ctypes import POINTER, c_float, c_int

import numpy as np

from . import _lib


def estep(feature_map, label, suppress_others=True, num_iter=5, margin_others=1e-5, bg_p=0.5,
          fg_p=0.25, use_c=True):
    """estep.py:18-27 -- returns a NEW float32 array [B,H,W,C]; inputs are left untouched
    (the astype copies of estep.py:21-22)."""
    feature_map = np.asarray(feature_map).astype(c_float)
    label = np.asarray(label).astype(c_int)
    f = estep_c(feature_map, label, suppress_others, num_iter, margin_others, bg_p, fg_p)
    return f.astype(np.float32)


def estep_c(feature_map, label, suppress_others, num_iter, margin_others, bg_p, fg_p):
    """estep.py:29-31 -- in place on `feature_map` (float32, strides honoured as
    weaklabel.cpp:50 does); `label` is a dense int32 [B,H,W] ground-truth map."""
    assert feature_map.dtype == np.float32 and feature_map.ndim == 4
    assert label.dtype == np.int32 and label.flags["C_CONTIGUOUS"]
    _lib.load().e_step(feature_map.ctypes.data_as(POINTER(c_float)),
                       feature_map.ctypes.shape_as(c_int), feature_map.ctypes.strides_as(c_int),
                       label.ctypes.data_as(POINTER(c_int)), bool(suppress_others), int(num_iter),
                       float(margin_others), float(bg_p), float(fg_p))
    return feature_map
