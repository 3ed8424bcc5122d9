#!/bin/bash
# tools/gpu_round.sh -- one gpurun call: GPU tests, microbench, bench, ncu launch list.
# usage: gpurun --timeout 1700 -- bash tools/gpu_round.sh [tests] [micro] [bench] [launches] [full]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/smi.txt 2>&1
for what in "$@"; do
case $what in
tests)
  timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  tail -5 gpurun_out/pytest_gpu.log ;;
testsall)
  timeout 1200 python -m pytest tests -q -m gpu -x --timeout 120 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -60 ;;
smoke)
  timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log ;;
micro)
  timeout 300 ./tools/microbench > gpurun_out/microbench.txt 2>&1; echo "micro exit $?"; cat gpurun_out/microbench.txt ;;
bench)
  timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err ;;
benchref)
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "benchref exit $?"; cat gpurun_out/bench_ref.json ;;
launches)
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --parity-images 0 > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --parity-images 0 > gpurun_out/ncu_launches.log 2>&1
  echo "launches exit $?"; tail -12 gpurun_out/launches.csv ;;
full)
  timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --parity-images 0 --images 64 --path fast > gpurun_out/plain_full.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:estep -s 3 -c 1 -f -o gpurun_out/prof \
      python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --parity-images 0 --images 64 --path fast > gpurun_out/ncu_full.log 2>&1
  echo "full exit $?"; tail -3 gpurun_out/ncu_full.log ;;
esac
done
