mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc; free -g | head -2 | tail -1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29541 bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/r02j_bench_n8.json 2> gpurun_out/r02j_bench_n8.err; echo "bench n8 rc=$?"; tail -3 gpurun_out/r02j_bench_n8.err
timeout 600 $TR --master-port 29542 bench.py --gpus 8 --steps 30 --warmup 5 --mode sharded --shard-mode composite --brick-log2 2 > gpurun_out/r02j_sharded8_composite_b2.json 2> gpurun_out/r02j_sharded8_b2.err; echo "sharded b2 rc=$?"
timeout 600 $TR --master-port 29543 bench.py --gpus 8 --steps 20 --warmup 3 --mode batched --scenes 8 --workload C1 > gpurun_out/r02j_batched_c5_8x8.json 2> gpurun_out/r02j_batched.err; echo "batched rc=$?"; tail -3 gpurun_out/r02j_batched.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02j_bench_n8.json"))
print("replicas", round(d["value"]), round(d["e2e"]["value"]), d["clocks"].get("per_gpu"))
for m,r in d.get("sharded",{}).items():
    print(m, round(r["value"]), r["ms_per_step"], r["blocks_allocated_total"], [round(x) for x in r["visible_blocks_per_shard"]], [round(x,3) for x in r["ms_per_shard"]], r["clocks"].get("per_gpu"))
try:
    r=json.load(open("gpurun_out/r02j_sharded8_composite_b2.json")); print("composite brick4", round(r["value"]), r["blocks_allocated_total"], [round(x) for x in r["visible_blocks_per_shard"]])
except Exception as e: print("b2 failed", e)
try:
    r=json.load(open("gpurun_out/r02j_batched_c5_8x8.json")); print("C5 64 sequences", round(r["value"]), r["config"]["workload"][:80], r["clocks"])
except Exception as e: print("batched failed", e)
PY
