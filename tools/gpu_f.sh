mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "cycles_images" > gpurun_out/pytest_cyc.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_cyc.log
