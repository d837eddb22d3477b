mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for b in 2 3; do
timeout 400 $TR --master-port 2955$b bench.py --gpus 2 --steps 30 --warmup 5 --mode sharded --shard-mode composite --brick-log2 $b > gpurun_out/r02k_sharded2_composite_b$b.json 2> gpurun_out/r02k_b$b.err; echo "b$b rc=$?"
python - <<PY
import json
r=json.load(open("gpurun_out/r02k_sharded2_composite_b$b.json")); print("brick_log2 $b", round(r["value"]), r["blocks_allocated_per_shard"], [round(x) for x in r["visible_blocks_per_shard"]], r["config"]["per_shard_capacity"])
PY
done
