mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02m_pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -3 gpurun_out/r02m_pytest_no_alloc_alone.log
timeout 600 python -m pytest tests/test_gpu_swap.py -m gpu -x -q > gpurun_out/r02m_pytest_swap.log 2>&1; echo "swap rc=$?"; tail -3 gpurun_out/r02m_pytest_swap.log
timeout 2400 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r02m_pytest_gpu.log 2>&1; echo "full rc=$?"; tail -25 gpurun_out/r02m_pytest_gpu.log
timeout 1200 python bench.py > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r02m_bench.err
timeout 300 python tools/time_aux.py > gpurun_out/r02m_aux_timings.json 2> gpurun_out/r02m_aux.err; echo "aux rc=$?"
