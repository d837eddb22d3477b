mkdir -p gpurun_out
(timeout 500 python -m pytest tests/test_gpu_refcuda.py tests/test_meshing.py tests/test_gpu_adapters.py -m gpu -q -p no:cacheprovider > gpurun_out/r02d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02d_pytest.log); tail -12 gpurun_out/r02d_pytest.log
B="python bench.py --single-workload --no-baselines --steps 40 --warmup 5"
DSB200_RAY_STEPS=0 DSB200_FUSE_KERNEL=1 timeout 200 $B > gpurun_out/r02d_bench_A.json 2>gpurun_out/r02d_A.err
DSB200_RAY_STEPS=12 DSB200_RAY_GROUP=2 DSB200_FUSE_KERNEL=1 timeout 200 $B > gpurun_out/r02d_bench_B.json 2>gpurun_out/r02d_B.err
DSB200_RAY_STEPS=16 DSB200_RAY_GROUP=4 DSB200_FUSE_KERNEL=2 DSB200_FUSE_CTAS=12 timeout 200 $B > gpurun_out/r02d_bench_C.json 2>gpurun_out/r02d_C.err
DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=8 DSB200_FUSE_KERNEL=2 DSB200_FUSE_CTAS=6 timeout 200 $B > gpurun_out/r02d_bench_D.json 2>gpurun_out/r02d_D.err
DSB200_RAY_STEPS=24 DSB200_RAY_GROUP=4 DSB200_FUSE_KERNEL=2 DSB200_FUSE_CTAS=12 timeout 200 $B --workload C1 > gpurun_out/r02d_bench_E_C1.json 2>gpurun_out/r02d_E.err
DSB200_RAY_STEPS=0 DSB200_FUSE_KERNEL=1 timeout 200 $B --workload C1 > gpurun_out/r02d_bench_F_C1.json 2>gpurun_out/r02d_F.err
for f in A B C D E_C1 F_C1; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02d_bench_$f.json"))
    print("$f", round(d["value"]), round(d["e2e"]["value"]), {k:round(v["ms"],4) for k,v in d["stages"].items()}, d["parity_check"])
except Exception as e:
    print("$f failed", e)
PY
done
