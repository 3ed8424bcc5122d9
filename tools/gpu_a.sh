mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
for th in 1024 512; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --threads $th --phase-profile gpurun_out/phase_$th.npy > gpurun_out/bench_v2_$th.json 2> gpurun_out/bench_v2_$th.err; echo "bench $th exit $?"; cat gpurun_out/bench_v2_$th.json | cut -c1-300; tail -3 gpurun_out/bench_v2_$th.err
python tools/phase_report.py gpurun_out/phase_$th.npy
done
