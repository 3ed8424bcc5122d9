#!/usr/bin/env python
"""Turn one `ncu --set full` capture of the bench kernel into the tracked files under profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [round]
writes profiles/rNN_ncu_full.txt, profiles/rNN_ncu_source_hotlines.txt and profiles/traffic.json
(bench.py reads traffic.json for roofline.traffic)."""
import csv, json, os, subprocess, sys
rep = sys.argv[1]
rnd = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
names, units, vals = rows[0], rows[1], rows[2]
col = {n: (u, v) for n, u, v in zip(names, units, vals)}
WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__cluster_size", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
lines = ["ncu --set full --clock-control none --import-source on -k regex:estep -s 3 -c 1 : python bench.py --steps 1 --warmup 3 "
         "--no-cpu --no-e2e --no-legs --parity-images 0",
         "(one launch of the bench kernel over the full 256 images of 513x513x21, labels-only output mode)", ""]
for n in WANT:
    u, v = col.get(n, ("", "n/a"))
    lines.append("%-70s %-12s %s" % (n, u, v))
for n in names:
    if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio"):
        u, v = col[n]
        lines.append("%-70s %-12s %s" % ("stall " + n[len("smsp__average_warps_issue_stalled_"):], u, v))
out = os.path.join(ROOT, "profiles", "r%02d_ncu_full.txt" % rnd)
open(out, "w").write("\n".join(lines) + "\n")
def gb(name):
    u, v = col[name]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}[u]
    return float(v) * scale
rd, wr = gb("dram__bytes_read.sum"), gb("dram__bytes_write.sum")
kname = col["Kernel Name"][1]
traffic = {"round": rnd, "kernel": kname, "kernel_name": "estep_onchip_kernel",
           "source": "profiles/r%02d_ncu_full.txt (ncu --set full, one launch over the 256 images of the bench)" % rnd,
           "captured_images": 256, "dram_bytes_read": rd, "dram_bytes_write": wr,
           "algorithmic_bytes_captured": 256 * 513 * 513 * 85, "estep_bytes_per_launch": rd + wr}
json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
hot = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), rep, "45"], capture_output=True, text=True).stdout
open(os.path.join(ROOT, "profiles", "r%02d_ncu_source_hotlines.txt" % rnd), "w").write(hot)
print(open(out).read())
print(hot[:3000])
