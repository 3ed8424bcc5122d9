mkdir -p gpurun_out
WEAKLABEL_V3_TRAIL=1 timeout -k 5 200 python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --parity-images 0 > gpurun_out/e.json 2> gpurun_out/e.err; echo "exit $?"
grep "^trail" gpurun_out/e.err > gpurun_out/trail_t.txt; wc -l gpurun_out/trail_t.txt
