mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29551 bench.py --gpus 8 > gpurun_out/r02ae_bench_n8.json 2> gpurun_out/r02ae_bench_n8.err; echo "bench n8 rc=$?"; tail -3 gpurun_out/r02ae_bench_n8.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02ae_bench_n8.json").read().strip().splitlines()[-1])
print("replicas", round(d["value"]), round(d["e2e"]["value"]), d["parity_check"], d["clocks"].get("per_gpu"))
for m,r in d.get("sharded",{}).items():
    print(m, round(r["value"]), r["ms_per_step"], r["step_ms_rank0"], r["blocks_allocated_total"], [round(x) for x in r["visible_blocks_per_shard"]], [round(x,3) for x in r["ms_per_shard"]])
PY
