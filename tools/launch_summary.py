#!/usr/bin/env python
"""Summarise the ncu launch list (gpu__time_duration.sum per launch) into profiles/rNN_launches_summary.txt.
usage: python tools/launch_summary.py gpurun_out/launches.csv [round]"""
import collections, csv, os, sys
path = sys.argv[1]
rnd = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.defaultdict(float); cnt = collections.Counter(); per = collections.defaultdict(list)
for r in rows[1:]:
    ns = float(r[vi].replace(",", ""))
    tot[r[ki]] += ns; cnt[r[ki]] += 1; per[r[ki]].append(ns / 1e3)
total = sum(tot.values())
lines = ["ncu --metrics gpu__time_duration.sum --clock-control none : python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e "
         "--no-legs --parity-images 0",
         "(the timed region of bench.py launches ONE kernel per step, estep_onchip_kernel<896,9>; the torch kernels below build "
         "the synthetic inputs before it; rows are cold and serialised under ncu)", ""]
for k, t in sorted(tot.items(), key=lambda kv: -kv[1]):
    lines.append("%6d launches  total %10.1f us  mean %9.1f us  %5.1f%%  %s" % (cnt[k], t / 1e3, t / 1e3 / cnt[k], 100 * t / total, k[:120]))
for k in per:
    if "estep_onchip_kernel" in k:
        lines += ["", "%s launches over 256 images of 513x513x21 (warm-up, timed, prepare legs): %s us"
                  % (k, ", ".join("%.1f" % v for v in per[k]))]
out = os.path.join(ROOT, "profiles", "r%02d_launches_summary.txt" % rnd)
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
