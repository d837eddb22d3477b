mkdir -p gpurun_out
nvidia-smi -L | head -3
(timeout 400 python -m pytest tests/test_gpu_sharded.py -m gpu -q -p no:cacheprovider > gpurun_out/r02i_pytest_sharded_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02i_pytest_sharded_2gpu.log); tail -5 gpurun_out/r02i_pytest_sharded_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r02i_bench_n2.json 2> gpurun_out/r02i_bench_n2.err; echo "bench rc=$?"; tail -4 gpurun_out/r02i_bench_n2.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02i_bench_n2.json"))
print("replicas", round(d["value"]), round(d["e2e"]["value"]), d["clocks"])
for m,r in d.get("sharded",{}).items():
    print(m, round(r["value"]), r["ms_per_step"], r["blocks_allocated_per_shard"], r["visible_blocks_per_shard"], r["ms_per_shard"], r["clocks"].get("per_gpu"))
PY
