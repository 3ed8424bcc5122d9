#!/usr/bin/env python
"""Staged bring-up check of the fast kernel on a B200: every (case, tuning, flags) runs in its own
process (a device-side trap poisons the CUDA context), compares bias / mean0 / mean1 / labels / map
with the oracle and prints one line.  usage: python tools/gpu_check4.py [quick|all]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

JOBS = [
    ("c1_41_p2", "1,0,0", 0), ("c1_41_p2", "1,0,0", 127), ("c1_41_p2", "2,0,0", 0), ("p5_seq3", "1,0,0", 0),
    ("p5_seq3", "4,0,0", 0), ("p21_41", "2,0,0", 0), ("ties_int", "2,0,0", 0), ("mag_tiny", "2,0,0", 0),
    ("c2_41_b30", "1,0,3", 0), ("c2_41_b30", "9,0,2", 0), ("shape_100x37", "2,0,0", 0),
    ("c3_321_randn", "0,0,0", 0), ("c3_321_randn", "4,0,1", 16), ("c4_513_b4", "9,0,1", 0), ("c4_513_b4", "9,0,2", 0),
    ("c4_513_p21", "9,0,1", 0), ("c4_513_scale5", "9,0,1", 0),
]


def child(name, tune, flags):
    import numpy as np
    import torch
    import cases
    import em_adapt_tensorflow_b200 as wl
    from oracle.oracle import Oracle
    os.environ["WEAKLABEL_FLAGS"] = str(flags)
    f, l, p = cases.build_case(name)
    orc = Oracle(1)
    fo = np.array(f, dtype=np.float32, order="C", copy=True)
    order_o, bias_o, m0_o, m1_o = orc.e_step_inplace(fo, l, probes=True, **p)
    lab_o = orc.argmax_labels(fo)
    B, H, W, C = f.shape
    t = torch.from_numpy(np.ascontiguousarray(f)).cuda()
    plan = wl.EStepPlan(B, H, W, C)
    tn = [int(x) for x in tune.split(",")]
    if any(tn):
        plan.tune(*tn)
    plan.set_path("fast")
    plan.prepare(wl.label_set(l, C), p["suppress_others"], p["num_iter"], p["margin_others"], p["bg"], p["fg"],
                 rng=wl.OrderStream(1))
    res = []
    for rep in range(2):
        lab = plan.labels(t)
        torch.cuda.synchronize()
        order_g, bias_g, m0_g, m1_g = plan.probes()
        res.append("bias=%s m0=%s m1=%s lab=%d" % (
            "ok" if bias_g.tobytes() == bias_o.tobytes() else "BAD(%d)" % int((bias_g != bias_o).sum()),
            "ok" if m0_g.tobytes() == m0_o.tobytes() else "BAD", "ok" if m1_g.tobytes() == m1_o.tobytes() else "BAD",
            int((lab.cpu().numpy() != lab_o).sum())))
        if "BAD" in res[-1] and rep == 0:
            bad = np.argwhere(bias_g != bias_o)
            if len(bad):
                b, j, tt = bad[0]
                res.append("first bad bias at image %d iter %d pos %d: got %r want %r" % (b, j, tt, bias_g[b, j, tt], bias_o[b, j, tt]))
            if m0_g.tobytes() != m0_o.tobytes():
                i = int(np.argmax(m0_g != m0_o)); res.append("m0[%d] got %r want %r" % (i, m0_g[i], m0_o[i]))
            if m1_g.tobytes() != m1_o.tobytes():
                i = int(np.argmax(m1_g != m1_o)); res.append("m1[%d] got %r want %r" % (i, m1_g[i], m1_o[i]))
    plan.finalize_map(t)
    res.append("map=%s" % ("ok" if t.cpu().numpy().tobytes() == fo.tobytes() else "BAD"))
    print("RESULT %s tune=%s flags=%d: %s" % (name, tune, flags, " | ".join(res)), flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 5 and sys.argv[1] == "child":
        child(sys.argv[2], sys.argv[3], int(sys.argv[4]))
        sys.exit(0)
    jobs = JOBS[:6] if (len(sys.argv) > 1 and sys.argv[1] == "quick") else JOBS
    env = dict(os.environ, WEAKLABEL_TRAIL="1")
    for name, tune, flags in jobs:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "child", name, tune, str(flags)],
                           capture_output=True, text=True, timeout=180, env=env)
        out = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
        if out:
            print(out[0])
        else:
            print("CRASH %s tune=%s flags=%d rc=%d" % (name, tune, flags, r.returncode))
            tail = (r.stderr or "").strip().splitlines()
            for ln in tail[-14:]:
                print("    " + ln[:300])
        sys.stdout.flush()
