mkdir -p gpurun_out
(DSB200_FUSE_KERNEL=3 timeout 500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_config.py tests/test_exactdiv.py -m gpu -q -p no:cacheprovider -x > gpurun_out/r02h_pytest_tma.log 2>&1; echo "pytest tma rc=$?" >> gpurun_out/r02h_pytest_tma.log); tail -5 gpurun_out/r02h_pytest_tma.log
run() { name=$1; shift; env "$@" timeout 200 $B > gpurun_out/r02h_bench_$name.json 2>gpurun_out/r02h_$name.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02h_bench_$name.json"))
    print("$name", round(d["value"]), round(d["e2e"]["value"]), {k:round(v["ms"],4) for k,v in d["stages"].items()}, d["parity_check"]["resident_pass"], d["parity_check"]["e2e_pass"])
except Exception as e:
    print("$name failed", e)
PY
}
B="python bench.py --single-workload --no-baselines --no-sharded --steps 20 --warmup 5"
run F2 DSB200_FUSE_KERNEL=2
run F3a DSB200_FUSE_KERNEL=3 DSB200_FUSE_MINB=5 DSB200_FUSE_CTAS=5
run F3b DSB200_FUSE_KERNEL=3 DSB200_FUSE_MINB=5 DSB200_FUSE_CTAS=10
run F3c DSB200_FUSE_KERNEL=3 DSB200_FUSE_MINB=4 DSB200_FUSE_CTAS=8
B="python bench.py --single-workload --no-baselines --no-sharded --steps 20 --warmup 5 --workload C1"
run F2_C1 DSB200_FUSE_KERNEL=2
run F3_C1 DSB200_FUSE_KERNEL=3 DSB200_FUSE_MINB=5 DSB200_FUSE_CTAS=10
(timeout 300 python -m pytest tests/test_gpu_refcuda.py -m gpu -q -p no:cacheprovider > gpurun_out/r02h_pytest_refcuda.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02h_pytest_refcuda.log); grep -E "AssertionError|passed|failed" gpurun_out/r02h_pytest_refcuda.log | cut -c1-700
