mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider --durations=8 > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02g_pytest_gpu.log); tail -25 gpurun_out/r02g_pytest_gpu.log
timeout 600 python bench.py --mode sharded --shard-mode composite --steps 20 --warmup 5 > gpurun_out/r02g_sharded1_composite.json 2> gpurun_out/r02g_sharded1_composite.err; tail -3 gpurun_out/r02g_sharded1_composite.err; head -c 1500 gpurun_out/r02g_sharded1_composite.json; echo
timeout 600 python bench.py --mode sharded --shard-mode lookup --steps 20 --warmup 5 > gpurun_out/r02g_sharded1_lookup.json 2> gpurun_out/r02g_sharded1_lookup.err; tail -3 gpurun_out/r02g_sharded1_lookup.err; head -c 600 gpurun_out/r02g_sharded1_lookup.json; echo
