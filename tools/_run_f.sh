mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 200 $B > gpurun_out/r02f_bench_$name.json 2>gpurun_out/r02f_$name.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02f_bench_$name.json"))
    print("$name", round(d["value"]), round(d["e2e"]["value"]), {k:round(v["ms"],4) for k,v in d["stages"].items()}, d["parity_check"]["resident_pass"], d["parity_check"]["e2e_pass"])
except Exception as e:
    print("$name failed", e)
PY
}
B="python bench.py --single-workload --no-baselines --steps 20 --warmup 5"
run A DSB200_RAY_STEPS=0
run C DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4
run D DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=8
run E DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=4
run G DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4 DSB200_RAY_FINISH_CTAS=8
run K DSB200_RAY_STEPS=32 DSB200_RAY_GROUP=8
run L DSB200_RAY_STEPS=12 DSB200_RAY_GROUP=4
run M DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=16 DSB200_RAY_FINISH_CTAS=8
B="python bench.py --single-workload --no-baselines --steps 20 --warmup 5 --workload C1"
run H_C1 DSB200_RAY_STEPS=0
run I_C1 DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4
run J_C1 DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=8
(timeout 300 python -m pytest tests/test_gpu_refcuda.py -m gpu -q -p no:cacheprovider > gpurun_out/r02f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02f_pytest.log); tail -6 gpurun_out/r02f_pytest.log
