#!/usr/bin/env python
"""Summarise an ncu report's warp-stall samples per CUDA source line (cuda,sass source page).
usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur_file = None; hdr = None; agg = collections.Counter(); src = {}; total = 0
inst = collections.Counter()
for row in csv.reader(out.splitlines()):
    if not row: continue
    if row[0] == "File Path": cur_file = row[1].split("/")[-1]; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No": hdr = row; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
    if hdr is None or len(row) <= si: continue
    if row[0] != "":          # a source line row (aggregated over its SASS)
        try: n = int(row[si]); k = int(row[ii])
        except ValueError: continue
        key = (cur_file, int(row[0]))
        agg[key] += n; inst[key] += k; src[key] = row[1].strip(); total += n
print("total samples", total)
for key, n in agg.most_common(top):
    print("%5.1f%% %8d inst=%10d %s:%d  %s" % (100.0 * n / total, n, inst[key], key[0], key[1], src[key][:110]))
