#!/bin/bash
# L2 prefetch distance of the producers (WEAKLABEL_PREFETCH, chunks ahead of the staging ring): one bench line each
for pf in 0 4 8 16 24 32 48; do
  WEAKLABEL_PREFETCH=$pf timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 0 > gpurun_out/pf_$pf.json 2>/dev/null
  echo "prefetch $pf: $(python -c "import json;d=json.load(open('gpurun_out/pf_$pf.json'));print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4))")"
done
