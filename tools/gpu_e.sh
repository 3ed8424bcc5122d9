mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fast_path_forced or device_api or sharded or repeat" > gpurun_out/pytest_v3.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_v3.log
timeout -k 5 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --phase-profile gpurun_out/phase_v3.npy > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err; echo "bench exit $?"; grep -o '"value": [0-9.]*' gpurun_out/bench_v3.json | head -1; grep -o 'mismatched_pixels": [0-9]*' gpurun_out/bench_v3.json
python tools/phase_report.py gpurun_out/phase_v3.npy
