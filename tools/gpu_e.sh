mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fast_path_forced or device_api or sharded or repeat" > gpurun_out/pytest_v3.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_v3.log
n=0
for i in $(seq 1 8); do
  timeout -k 5 200 python bench.py --steps 5 --warmup 2 --no-cpu --no-e2e --parity-images 4 > gpurun_out/e.json 2> gpurun_out/e.err; rc=$?
  if [ $rc -ne 0 ]; then n=$((n+1)); echo "run $i crashed"; tail -3 gpurun_out/e.err; fi
done
echo "crashes $n of $i"; grep -o '"value": [0-9.]*' gpurun_out/e.json | head -1; grep -o 'mismatched_pixels": [0-9]*' gpurun_out/e.json
timeout -k 5 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --phase-profile gpurun_out/phase_v3.npy > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err; echo "bench exit $?"
python tools/phase_report.py gpurun_out/phase_v3.npy
timeout -k 5 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --tune 8,0,0 > gpurun_out/bench_v3_g8.json 2> gpurun_out/bench_v3_g8.err; echo "bench G8 exit $?";  grep -o '"value": [0-9.]*' gpurun_out/bench_v3_g8.json | head -1
