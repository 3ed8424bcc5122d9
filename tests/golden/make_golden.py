#!/usr/bin/env python
"""tests/golden/make_golden.py -- regenerate the golden vectors (needs /root/reference).

The reference ships no tests or fixtures for its E-step, so the golden
vectors are OUTPUTS OF THE REFERENCE ITSELF: estep/weaklabel.cpp compiled unmodified into
oracle/_ref/libweaklabel_ref.so (oracle/Makefile), cross-checked against the binary the
reference ships (estep/libweaklabel.so).  Each case runs in a FRESH subprocess so that the
process-global rand() the reference shuffles with (weaklabel.cpp:83) starts at glibc's
default seed; inside a case the images are fed one per call in index order because the
reference crashes for batch > 1 (weaklabel.cpp:117-118).  Labels are
np.argmax(map, axis=3) (first maximum), the pinned reading of tf.argmax (deeplab.py:123).

    python tests/golden/make_golden.py            # all cases
    python tests/golden/make_golden.py NAME ...   # some cases (merged into the files)
"""
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402


def _child(name, which, out_path):
    from oracle.oracle import Reference
    ref = Reference(which)
    f, l, p = cases.build_case(name)
    for b in range(f.shape[0]):
        ref.e_step_image_inplace(f[b:b + 1], l[b:b + 1], **p)
    labels = np.argmax(f, axis=3).astype(np.uint8)
    np.savez(out_path, map=np.ascontiguousarray(f), labels=labels)


def main(argv):
    if len(argv) >= 2 and argv[1] == "--child":
        return _child(argv[2], argv[3], argv[4])
    names = argv[1:] or [c["name"] for c in cases.CASES]
    jpath = os.path.join(HERE, "golden.json")
    zpath = os.path.join(HERE, "golden_arrays.npz")
    meta = json.load(open(jpath)) if os.path.exists(jpath) else {}
    arrays = dict(np.load(zpath)) if os.path.exists(zpath) else {}
    tmp = "/tmp/_golden_%d.npz" % os.getpid()
    for name in names:
        res = {}
        for which in ("ref", "shipped"):
            subprocess.check_call([sys.executable, os.path.abspath(__file__), "--child", name, which, tmp])
            z = np.load(tmp)
            res[which] = (z["map"], z["labels"])
        m, lab = res["ref"]
        assert m.tobytes() == res["shipped"][0].tobytes(), "ref and shipped builds disagree: " + name
        assert np.isfinite(m).all(), name
        case = cases.CASE_BY_NAME[name]
        meta[name] = dict(map_sha1=cases.sha1(m), labels_sha1=cases.sha1(lab),
                          shape=list(m.shape), store=case["store"],
                          label_hist=np.bincount(lab.ravel(), minlength=21).tolist())
        if case["store"] in ("full", "labels"):
            arrays[name + "/labels"] = lab
        if case["store"] == "full":
            arrays[name + "/map"] = m
        print("%-18s map %s labels %s" % (name, meta[name]["map_sha1"][:12], meta[name]["labels_sha1"][:12]))
    os.remove(tmp)
    json.dump(meta, open(jpath, "w"), indent=1, sort_keys=True)
    np.savez_compressed(zpath, **arrays)
    print("wrote", jpath, zpath, "%.1f KB" % (os.path.getsize(zpath) / 1e3))


if __name__ == "__main__":
    main(sys.argv)
