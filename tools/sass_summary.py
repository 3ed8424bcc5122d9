#!/usr/bin/env python
"""SASS opcode histogram of the built library -> profiles/sass_summary.txt (evidence of tcgen05 / bulk-copy / mbarrier use).
usage: python tools/sass_summary.py"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join("em_adapt_tensorflow_b200", "lib", "libweaklabel_cuda.so")
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, so)], capture_output=True, text=True).stdout
KEY = ["LDTM", "STTM", "UBLKCP", "UBLKPF", "SYNCS", "LDGSTS", "ATOMS", "ATOM", "LDS", "STS", "LDG", "STG", "LD", "ST", "LDL", "STL",
       "BAR", "UCGABAR_ARV", "UCGABAR_WAIT", "MUFU", "DADD", "FADD", "SHFL", "MEMBAR", "FENCE", "CCTL", "ELECT", "NANOSLEEP"]
allc = collections.Counter(); per = collections.OrderedDict(); cur = None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1); per[cur] = collections.Counter(); continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1); allc[op] += 1; per[cur][op] += 1
L = ["SASS opcode histogram of %s (cuobjdump -sass, sm_100a only)" % so, "", "Blackwell / async-path opcodes over the whole library:"]
L += ["  %-16s %6d" % (k, allc[k]) for k in KEY if allc[k]]
L += ["  (LDTM / STTM = tcgen05.ld / st: class planes in tensor memory;  UBLKCP = cp.async.bulk, UBLKPF = cp.async.bulk.prefetch.L2;",
      "   SYNCS = mbarrier;  LDGSTS = cp.async;  no UTMALDG: linear bulk copies, no tensor maps;  no UTC*MMA: nothing on this path is a contraction)",
      "", "per kernel: instructions, then the opcodes above"]
for fn, c in sorted(per.items(), key=lambda kv: -sum(kv[1].values())):
    L.append("  %-80s %6d  %s" % (fn, sum(c.values()), " ".join("%s=%d" % (k, c[k]) for k in KEY if c[k])))
L += ["", "top 40 opcodes:"] + ["  %-16s %6d" % kv for kv in allc.most_common(40)]
open(os.path.join(ROOT, "profiles", "sass_summary.txt"), "w").write("\n".join(L) + "\n")
print("\n".join(L[:50]))
