#!/usr/bin/env python
"""tools/gpu_check.py -- verbose parity run of every case in tests/cases.py on cuda:0.
Compares the device API (labels, probes, finalized map) and the legacy host-pointer e_step
against the oracle, and prints WHERE they diverge.  Debug aid; the judged tests are tests/."""
import ctypes
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import torch  # noqa: E402
import em_adapt_tensorflow_b200 as wl  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

libc = ctypes.CDLL("libc.so.6")


def run_case(name, tune=None, verbose=True):
    f, l, p = cases.build_case(name)
    B, H, W, C = f.shape
    orc = Oracle(1)
    fo = np.array(f, dtype=np.float32, order="C", copy=True)
    order_o, bias_o, m0_o, m1_o = orc.e_step_inplace(fo, l, probes=True, **p)
    lab_o = orc.argmax_labels(fo)
    ok = True
    msgs = []
    # ---- device API (contiguous copy) ----
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(B, H, W, C)
    if tune:
        plan.tune(*tune)
    plan.prepare(wl.label_set(l, C), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                 rng=wl.OrderStream(1))
    t0 = time.time()
    lab = plan.labels(t)
    torch.cuda.synchronize()
    dt = time.time() - t0
    order_g, bias_g, m0_g, m1_g = plan.probes()
    if not np.array_equal(order_g, order_o):
        ok = False; msgs.append("order differs")
    if bias_g.tobytes() != bias_o.tobytes():
        ok = False
        bad = np.argwhere(bias_g.view(np.uint32) != bias_o.view(np.uint32))
        msgs.append("bias differs at %d entries, first (b,j,t)=%s gpu=%r oracle=%r" % (
            len(bad), bad[0].tolist(), bias_g[tuple(bad[0])], bias_o[tuple(bad[0])]))
    if m0_g.tobytes() != m0_o.tobytes():
        ok = False; msgs.append("mean0 differs gpu=%s oracle=%s" % (m0_g[:3], m0_o[:3]))
    if m1_g.tobytes() != m1_o.tobytes():
        ok = False; msgs.append("mean1 differs gpu=%s oracle=%s" % (m1_g[:3], m1_o[:3]))
    lab_h = lab.cpu().numpy()
    nd = int((lab_h != lab_o).sum())
    if nd:
        ok = False; msgs.append("labels differ at %d pixels" % nd)
    plan.finalize_map(t)
    fm = t.cpu().numpy()
    if fm.tobytes() != np.ascontiguousarray(fo).tobytes():
        ok = False; msgs.append("finalized map differs at %d elements" % int((fm.view(np.uint32) != fo.view(np.uint32)).sum()))
    plan.close()
    # ---- legacy host-pointer e_step with the case's own strides ----
    if tune is None:
        f2, l2, _ = cases.build_case(name)
        libc.srand(1)
        wl.estep_c(f2, l2, p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"])
        if np.ascontiguousarray(f2).tobytes() != np.ascontiguousarray(fo).tobytes():
            ok = False; msgs.append("legacy e_step map differs at %d elements" % int(
                (np.ascontiguousarray(f2).view(np.uint32) != np.ascontiguousarray(fo).view(np.uint32)).sum()))
    if verbose or not ok:
        print("%-18s %s  %.1f ms  %s" % (name, "OK  " if ok else "FAIL", dt * 1e3, "; ".join(msgs)), flush=True)
    return ok


def main():
    names = sys.argv[1:] or [c["name"] for c in cases.CASES]
    print("device:", torch.cuda.get_device_name(0), flush=True)
    bad = [n for n in names if not run_case(n)]
    # forced geometries on small cases: multi-tile, multi-round, clusters of 2/4/8
    for tune in [(1, 64, 0), (2, 64, 0), (4, 128, 0), (8, 32, 2), (2, 0, 0), (16, 16, 0)]:
        for n in ["c1_41_p2", "p5_seq3", "nosuppress_p3", "p21_41", "allneg_p3", "ties_int", "shape_7x5", "mag_big"]:
            if not run_case(n, tune, verbose=False):
                bad.append("%s@%s" % (n, tune))
        print("tune", tune, "done", flush=True)
    print("FAILED:" if bad else "ALL OK", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
