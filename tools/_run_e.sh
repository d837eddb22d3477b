mkdir -p gpurun_out
(timeout 400 python -m pytest tests/test_gpu_refcuda.py -m gpu -q -p no:cacheprovider > gpurun_out/r02e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02e_pytest.log); tail -6 gpurun_out/r02e_pytest.log
B="python bench.py --single-workload --no-baselines --steps 20 --warmup 5"
run() { name=$1; shift; env "$@" timeout 200 $B > gpurun_out/r02e_bench_$name.json 2>gpurun_out/r02e_$name.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02e_bench_$name.json"))
    print("$name", round(d["value"]), round(d["e2e"]["value"]), {k:round(v["ms"],4) for k,v in d["stages"].items()}, d["parity_check"]["resident_pass"], d["parity_check"]["e2e_pass"])
except Exception as e:
    print("$name failed", e)
PY
}
run A DSB200_RAY_STEPS=0
run B DSB200_RAY_STEPS=12 DSB200_RAY_GROUP=2
run C DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4
run D DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=8
run E DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=4
run F DSB200_RAY_STEPS=32 DSB200_RAY_GROUP=4
run G DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4 DSB200_RAY_FINISH_CTAS=8
B="python bench.py --single-workload --no-baselines --steps 20 --warmup 5 --workload C1"
run H_C1 DSB200_RAY_STEPS=0
run I_C1 DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4
run J_C1 DSB200_RAY_STEPS=32 DSB200_RAY_GROUP=4
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_fuse_warp_flat -s 42 -c 1 -o gpurun_out/r02e_fuse_flat -f python tools/tune.py --stage integrate --workload C2 --frames 40 --reps 3 > gpurun_out/r02e_ncu_fuse.log 2>&1; tail -2 gpurun_out/r02e_ncu_fuse.log
