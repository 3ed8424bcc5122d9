#!/usr/bin/env python
"""Summarise per-image phase stamps written by bench.py --phase-profile (weaklabel_plan_profile):
[B][32] uint64 = 0 claimed, 1 E0 in shared memory, 2 mean_max_0 digest done, 3 iterations done, 4 sums known,
5 labels written, 6 selections, 7 P; fine stamps of the lead thread: 8 planes ready, 9 plane loads issued,
10 chunk prefix, 11 first exchange, 12 float digest built, 16-18 accumulated plane acquire / select / fold ns,
20 chunk prefix, 21 exchange, 22 double digest built, 23 own tile walked, 24 peers' digests landed, 25 peers walked."""
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
if t.shape[1] >= 32 and t[:, 8].any():
    fine = [("wait planes", 0, 8), ("plane loads", 8, 9), ("E0 landed", 9, 1), ("chunk prefix", 1, 10), ("exchange", 10, 11),
            ("float digest", 11, 12), ("walk/compact", 12, 2), ("chunk prefix1", 3, 20), ("exchange1", 20, 21),
            ("double digest", 21, 22), ("own walk", 22, 23), ("digests land", 23, 24), ("peer walk", 24, 25), ("result", 25, 4)]
    print("fine (us, mean over images): " + "  ".join("%s %.1f" % (n, ((t[:, b] - t[:, a]) / 1e3).mean()) for n, a, b in fine))
    sel = np.maximum(t[:, 6], 1)
    has = t[:, 6] > 0
    print("per selection (us): acquire %.1f  select %.1f  fold %.1f" % tuple((t[has, k] / sel[has] / 1e3).mean() for k in (16, 17, 18)))
    if t[:, 27].any():
        print("own double digest: segments %.1f (max %d)  raw chunks %.1f (max %d)  segments before the cap %.1f (max %d)" % (
            t[:, 27].mean(), t[:, 27].max(), t[:, 28].mean(), t[:, 28].max(), t[:, 29].mean(), t[:, 29].max()))
    if t[:, 14].any():
        print("producer warp 0 of CTA 0, per image (us): streaming %.1f, of which waiting for its stages %.1f; chunks %.0f -> %.2f us per chunk, %.2f waiting" % (
            t[:, 14].mean() / 1e3, t[:, 13].mean() / 1e3, t[:, 15].mean(), (t[:, 14] / np.maximum(t[:, 15], 1)).mean() / 1e3,
            (t[:, 13] / np.maximum(t[:, 15], 1)).mean() / 1e3))
        for pp in sorted(set(P)):
            m = P == pp
            print("   P=%d: streaming %.1f us, waiting %.1f us" % (pp, t[m, 14].mean() / 1e3, t[m, 13].mean() / 1e3))
