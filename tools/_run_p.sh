mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_config.py tests/test_gpu_swap.py tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/r02p_pytest_deep.log 2>&1; echo "parity rc=$?"; tail -6 gpurun_out/r02p_pytest_deep.log
for d in 0 1; do
DSB200_DEEP_LOOKAHEAD=$d timeout 600 python bench.py --no-baselines --no-sharded > gpurun_out/r02p_bench_deep$d.json 2> gpurun_out/r02p_bench_deep$d.err; echo "bench deep=$d rc=$?"; tail -2 gpurun_out/r02p_bench_deep$d.err
python - <<PY
import json
r=json.loads(open("gpurun_out/r02p_bench_deep$d.json").read().strip().splitlines()[-1])
print("deep=$d C2", round(r["value"]), round(r["e2e"]["value"]), r["parity_check"], r["gpu_launches"])
for k,v in r.get("workloads",{}).items(): print("   ",k, round(v["value"]), round(v["e2e"]["value"]), v["parity_check"])
PY
done
