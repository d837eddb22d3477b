mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_swap.py tests/test_gpu_zz_swap_full_size.py -m gpu -x -q > gpurun_out/r02t_pytest_swap.log 2>&1; echo "swap rc=$?"; tail -5 gpurun_out/r02t_pytest_swap.log
for m in composite lookup; do
timeout 600 python bench.py --mode sharded --shard-mode $m > gpurun_out/r02t_sharded1_$m.json 2> gpurun_out/r02t_sh.err; echo "sharded $m rc=$?"; tail -2 gpurun_out/r02t_sh.err
python - <<PY
import json
r=json.loads(open("gpurun_out/r02t_sharded1_$m.json").read().strip().splitlines()[-1])
print("$m", round(r["value"]), r["step_ms_rank0"], r["blocks_allocated_total"])
PY
done
