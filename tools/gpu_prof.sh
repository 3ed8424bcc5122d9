#!/bin/bash
# one ncu --set full capture of the E-step kernel (64 images) + per-line summary
mkdir -p gpurun_out
ARGS="${@:---path fast}"
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --parity-images 0 --images 64 $ARGS > gpurun_out/plain_full.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_full.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:estep -s 3 -c 1 -f -o gpurun_out/prof \
   python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --parity-images 0 --images 64 $ARGS > gpurun_out/ncu_full.log 2>&1
echo "full exit $?"; tail -2 gpurun_out/ncu_full.log
