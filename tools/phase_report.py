#!/usr/bin/env python
"""Summarise per-image phase stamps written by bench.py --phase-profile (weaklabel_plan_profile)."""
import sys
import numpy as np
t = np.load(sys.argv[1]).astype(np.int64)
names = ["stream-in", "exchange+mean0", "iterations", "mean1", "labels"]
d = np.diff(t[:, :6], axis=1) / 1e3
tot = (t[:, 5] - t[:, 0]) / 1e3
P = t[:, 7]
print("images %d  kernel span %.1f us  sum of image times %.1f us" % (len(t), (t[:, 5].max() - t[:, 0].min()) / 1e3, tot.sum()))
print("%-4s %5s %9s " % ("P", "n", "total") + " ".join("%14s" % n for n in names) + "  selections")
for p in sorted(set(P)):
    m = P == p
    print("%-4d %5d %9.1f " % (p, m.sum(), tot[m].mean()) + " ".join("%14.1f" % d[m, k].mean() for k in range(5)) + "  %.1f" % t[m, 6].mean())
print("%-4s %5d %9.1f " % ("all", len(t), tot.mean()) + " ".join("%14.1f" % d[:, k].mean() for k in range(5)) + "  %.1f" % t[:, 6].mean())
