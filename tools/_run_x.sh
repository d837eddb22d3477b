mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02x_pytest_parity.log 2>&1; echo "parity rc=$?"; tail -12 gpurun_out/r02x_pytest_parity.log
