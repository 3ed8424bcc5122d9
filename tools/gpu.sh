#!/bin/bash
# tools/gpu.sh -- one gpurun call: selected steps, everything logged under gpurun_out/.
# usage: gpurun --timeout 1500 -- bash tools/gpu.sh step [step ...]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/smi.txt 2>&1
for what in "$@"; do
echo "=== $what"
case $what in
chk)    # staged bring-up check, one process per case
  timeout -k 10 900 python tools/gpu_check.py all > gpurun_out/check.log 2>&1; echo "check exit $?"; cat gpurun_out/check.log ;;
tfast)  # the on-chip kernel's own tests
  timeout -k 10 900 python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 120 \
     -k "device_api_bit_exact or fast_path_forced or cycles_images or kernel_switches or full_size or one_call or sharded" \
     > gpurun_out/pytest_fast.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_fast.log
  grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_fast.log | head -40 ;;
soak)
  timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 300 -k "soak" > gpurun_out/pytest_soak.log 2>&1
  echo "soak exit $?"; tail -5 gpurun_out/pytest_soak.log ;;
tests)
  timeout -k 10 1500 python -m pytest tests -q -m gpu --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
  grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -40 ;;
tapi)   # the C-ABI around the kernel
  timeout -k 10 1200 python -m pytest tests/test_gpu_api.py -q -m gpu --timeout 600 > gpurun_out/pytest_api.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_api.log
  grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_api.log | head -40 ;;
c5)     # BASELINE.json configs[4]: the end-to-end EM-Adapt step, device loss vs the py_func round trip
  timeout 900 python tools/config5.py --steps 10 --warmup 3 > gpurun_out/config5_n1.json 2> gpurun_out/config5_n1.err
  echo "config5 exit $?"; cat gpurun_out/config5_n1.json; tail -3 gpurun_out/config5_n1.err
  timeout 900 python tools/config5.py --steps 10 --warmup 3 --tf32 > gpurun_out/config5_n1_tf32.json 2>> gpurun_out/config5_n1.err
  echo "config5 tf32 exit $?"; cat gpurun_out/config5_n1_tf32.json ;;
c5n)    # the same under DDP/NCCL on NG GPUs: NG=8 bash tools/gpu.sh c5n
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port 29531 \
      tools/config5.py --steps 10 --warmup 3 > gpurun_out/config5_n${NG:-2}.json 2> gpurun_out/config5_n${NG:-2}.err
  echo "config5 x${NG:-2} exit $?"; cat gpurun_out/config5_n${NG:-2}.json; tail -3 gpurun_out/config5_n${NG:-2}.err ;;
benchn) # the bench under torchrun on NG GPUs, as the driver launches it
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus ${NG:-2} > gpurun_out/bench_n${NG:-2}.json 2> gpurun_out/bench_n${NG:-2}.err
  echo "bench x${NG:-2} exit $?"; cat gpurun_out/bench_n${NG:-2}.json; tail -3 gpurun_out/bench_n${NG:-2}.err ;;
lat)
  ./tools/lat 2>&1 | tee gpurun_out/lat.txt ;;
smoke)
  timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log ;;
bk)     # kernel-only bench + per-image phase stamps
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-legs --phase-profile gpurun_out/phase.npy \
     > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err; echo "bench exit $?"; cat gpurun_out/bench_k.json; tail -3 gpurun_out/bench_k.err
  python tools/phase_report.py gpurun_out/phase.npy 2>&1 | tee gpurun_out/phase.txt ;;
ab)     # switches of the kernel, one bench line each: AB="1 2 4" bash tools/gpu.sh ab
  for fl in ${AB:-1 2 4 8 16 32}; do
    WEAKLABEL_FLAGS=$fl timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 2 \
       --phase-profile gpurun_out/phase_f$fl.npy > gpurun_out/bench_f$fl.json 2> gpurun_out/bench_f$fl.err
    echo "flags $fl exit $?: $(python -c "import json;d=json.load(open('gpurun_out/bench_f$fl.json'));print(d['ms_per_step'], d['roofline']['frac'])" 2>&1)"
    python tools/phase_report.py gpurun_out/phase_f$fl.npy 2>&1 | tail -4
  done ;;
bench)
  timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err ;;
benchref)
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "benchref exit $?"; cat gpurun_out/bench_ref.json ;;
launches)
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 0 > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 0 > gpurun_out/ncu_launches.log 2>&1
  echo "launches exit $?"; tail -12 gpurun_out/launches.csv ;;
full)
  timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 0 > gpurun_out/plain_full.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:estep -s 3 -c 1 -f -o gpurun_out/prof \
      python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --no-legs --parity-images 0 > gpurun_out/ncu_full.log 2>&1
  echo "full exit $?"; tail -3 gpurun_out/ncu_full.log ;;
*)  echo "unknown step $what" ;;
esac
done
