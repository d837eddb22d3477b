mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r02ad_pytest_gpu.log 2>&1; echo "full rc=$?"; tail -14 gpurun_out/r02ad_pytest_gpu.log
timeout 1200 python bench.py > gpurun_out/r02ad_bench.json 2> gpurun_out/r02ad_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02ad_bench.err
python - <<'PY'
import json
r=json.loads(open("gpurun_out/r02ad_bench.json").read().strip().splitlines()[-1])
print("C2", round(r["value"]), round(r["e2e"]["value"]), r["parity_check"], r["gpu_launches"], r["roofline"]["frac"], r["gpu_comparator"]["value"], r["cpu_baseline"]["value"])
for k,v in r.get("workloads",{}).items(): print("   ",k, round(v["value"]), round(v["e2e"]["value"]), v["parity_check"])
for k,v in r.get("sharded",{}).items(): print("   ",k, round(v["value"]), v["step_ms_rank0"])
PY
B="python bench.py --steps 6 --warmup 3 --single-workload --no-baselines --no-sharded"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02ad_launches_bench_c2_steps6.csv $B > gpurun_out/r02ad_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_fuse_warp_flat -s 42 -c 1 -o gpurun_out/r02ad_fuse -f python tools/tune.py --stage integrate --workload C2 --frames 40 --reps 3 > gpurun_out/r02ad_ncu_fuse.log 2>&1; echo "ncu fuse rc=$?"; tail -2 gpurun_out/r02ad_ncu_fuse.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_raycast -s 42 -c 1 -o gpurun_out/r02ad_raycast -f python tools/tune.py --stage frame --workload C2 --frames 40 --reps 3 > gpurun_out/r02ad_ncu_ray.log 2>&1; echo "ncu raycast rc=$?"; tail -2 gpurun_out/r02ad_ncu_ray.log
ls -la gpurun_out/*.ncu-rep
