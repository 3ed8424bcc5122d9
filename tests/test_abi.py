"""The C-ABI shared library loads and exports every symbol include/weaklabel_cuda.h declares
(and nothing leaks C++ names); the drop-in aliases exist under the names the reference's
ctypes callers look for."""
import ctypes
import os
import re
import subprocess

import em_adapt_tensorflow_b200 as wl
from em_adapt_tensorflow_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    h = open(os.path.join(ROOT, "include", "weaklabel_cuda.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"WL_API\s+[\w\s\*]+?\b(\w+)\s*\(", h)))


def test_header_symbols_all_exported():
    lib = ctypes.CDLL(build.build())
    names = _declared()
    assert "e_step" in names and len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
    assert sorted(_lib.EXPORTS) == names


def test_no_cxx_symbols_exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", build.build()], text=True)
    syms = [l.split()[-1] for l in out.splitlines() if " T " in l]
    ours = [s for s in syms if not s.startswith("_")]
    assert sorted(ours) == _declared()


def test_legacy_signature_is_the_reference_prototype():
    """estep.py:15 -- argtypes the reference's callers set; binding them must work."""
    from ctypes import POINTER, c_bool, c_float, c_int
    lib = ctypes.cdll.LoadLibrary(build.build())
    lib.e_step.argtypes = [POINTER(c_float), POINTER(c_int), POINTER(c_int), POINTER(c_int), c_bool, c_int,
                           c_float, c_float, c_float]
    assert lib.weaklabel_version() >= 100


def test_dropin_aliases():
    lib = build.build()
    d = os.path.join(os.path.dirname(lib), "dropin")
    for name in build.DROPIN_NAMES:      # estep.py:9-12, estep/estep.py:11, estep_with_ctype.py:16
        p = os.path.join(d, name)
        assert os.path.realpath(p) == os.path.realpath(lib)
        assert hasattr(ctypes.CDLL(p), "e_step")


def test_built_for_sm_100a_only():
    out = subprocess.check_output(["cuobjdump", "-lelf", build.build()], text=True)
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
