mkdir -p gpurun_out
for pf in 0 16; do
WEAKLABEL_V3_PREFETCH=$pf timeout -k 5 120 python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --images 40 > gpurun_out/b40_$pf.json 2> gpurun_out/b40_$pf.err; echo "bench40 pf=$pf exit $?"; cut -c1-200 gpurun_out/b40_$pf.json; tail -2 gpurun_out/b40_$pf.err
done
WEAKLABEL_V3_PREFETCH=0 timeout -k 5 600 compute-sanitizer --tool memcheck --print-limit 20 python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --images 40 --parity-images 0 > gpurun_out/san.log 2>&1; echo "san exit $?"; grep -v "^$" gpurun_out/san.log | head -60
