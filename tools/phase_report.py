#!/usr/bin/env python
"""Summarise per-image phase stamps written by bench.py --phase-profile (weaklabel_plan_profile):
[B][32] uint64 = 0 claimed, 1 E0 in shared memory, 2 first exchange done, 3 iterations done, 4 sums known (or
skipped), 5 labels written, 6 selections, 7 P; fine stamps of the lead thread: 8 planes ready, 9 plane loads
issued, 13-15 producer warp 0: ns waiting / ns streaming / chunks, 16-18 accumulated plane acquire / select /
fold ns, 19 exact sums formed (0/1), 20 iterations left, 21 pass A + flag exchange done, 22 double sum done."""
import sys
import numpy as np
t = np.load(sys.argv[1]).astype(np.int64)
names = ["stream-in", "exchange", "iterations", "labels+sums", "pass B"]
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
    fine = [("wait planes", 0, 8), ("plane loads", 8, 9), ("E0 landed", 9, 1), ("exchange", 1, 2), ("pass A + flag exchange", 20, 21),
            ("double sum", 21, 22), ("float sum", 22, 4)]
    print("fine (us, mean over images): " + "  ".join("%s %.1f" % (n, ((t[:, b] - t[:, a]) / 1e3).mean()) for n, a, b in fine))
    print("images that formed the exact sums: %d of %d" % (int(t[:, 19].sum()), len(t)))
    sel = np.maximum(t[:, 6], 1)
    has = t[:, 6] > 0
    print("per selection (us): acquire %.1f  select %.1f  fold %.1f" % tuple((t[has, k] / sel[has] / 1e3).mean() for k in (16, 17, 18)))
    if t[:, 14].any():
        print("producer warp 0 of CTA 0, per image (us): streaming %.1f, of which waiting for its stages %.1f; chunks %.0f -> %.2f us per chunk" % (
            t[:, 14].mean() / 1e3, t[:, 13].mean() / 1e3, t[:, 15].mean(), (t[:, 14] / np.maximum(t[:, 15], 1)).mean() / 1e3))
    if t[:, 23:32].any():          # library built with -DWL_SELPROF: laps of the lead thread inside select / fold
        lap = ["hist pass", "exchange 1 (super bins)", "locate + fine bins", "gather pass", "gather exchange", "local select",
               "fold pass", "fold exchange", "fold tail"]
        for p in sorted(set(P)):
            m = (P == p) & has
            if m.any():
                print("P=%d laps per selection (us): " % p + "  ".join("%s %.2f" % (n, (t[m, 23 + i] / sel[m] / 1e3).mean()) for i, n in enumerate(lap)))
    if t[:, 14].any():
        for p in sorted(set(P)):
            m = P == p
            print("P=%d producer warp 0 of CTA 0: streaming %.1f us (waiting for stages %.1f), %.0f chunks -> %.2f us per chunk; stream start to image claimed %.1f us" % (
                p, t[m, 14].mean() / 1e3, t[m, 13].mean() / 1e3, t[m, 15].mean(), (t[m, 14] / np.maximum(t[m, 15], 1)).mean() / 1e3,
                ((t[m, 0] - t[m, 12]) / 1e3).mean()))
