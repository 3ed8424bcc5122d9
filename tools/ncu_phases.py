#!/usr/bin/env python
"""Attribute an ncu report's warp-stall samples of the E-step kernel to its phases: SASS instructions are
walked in address order and carry the phase of the last call-site line of the kernel body seen before them.
usage: python tools/ncu_phases.py report.ncu-rep body_first_line [top_n]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]; body0 = int(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
recs = []; cur = None; fname = None; hdr = None; src = {}
for row in csv.reader(out.splitlines()):
    if not row: continue
    if row[0] == "File Path": fname = row[1].split("/")[-1]; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No":
        hdr = row; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed")
        st = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if row[0] != "": cur = (fname, int(row[0])); src[cur] = row[1].strip(); continue
    if len(row) > si and row[2].startswith("0x"):
        recs.append((int(row[2], 16), cur, int(row[si]), int(row[ii]), row[3].strip(), [int(row[i]) for i, _ in st]))
recs.sort()
phase = None; agg = collections.Counter(); inst = collections.Counter(); lines = collections.defaultdict(collections.Counter)
stalls = collections.defaultdict(lambda: [0] * len(st))
tot = 0
for addr, key, n, k, sass, sv in recs:
    if key and key[0].startswith("estep_onchip") and key[1] >= body0: phase = key[1]
    agg[phase] += n; inst[phase] += k; tot += n
    lines[phase][key] += n
    for j, v in enumerate(sv): stalls[phase][j] += v
print("total samples", tot)
for ph, n in sorted(agg.items(), key=lambda kv: (kv[0] or 0)):
    if n < tot * 0.004: continue
    s = sorted(zip(stalls[ph], [h for _, h in st]), reverse=True)[:3]
    print("%5.1f%% inst=%10d body line %s: %s   [%s]" % (100.0 * n / tot, inst[ph], ph, src.get(("estep_onchip.cu", ph), "")[:60],
          ", ".join("%s %.0f%%" % (h[6:], 100.0 * v / max(n, 1)) for v, h in s)))
    for key, m in lines[ph].most_common(3):
        if m >= tot * 0.004: print("          %5.1f%%  %s:%d %s" % (100.0 * m / tot, key[0], key[1], src[key][:80]))
