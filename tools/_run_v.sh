mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29561 bench.py --gpus 2 --mode sharded --shard-mode composite > gpurun_out/r02v_sharded2_composite.json 2> gpurun_out/r02v.err; echo "rc=$?"
python - <<'PY'
import json
r=json.loads(open("gpurun_out/r02v_sharded2_composite.json").read().strip().splitlines()[-1])
print(round(r["value"]), {k:v for k,v in r["step_ms_rank0"].items() if k!="all"})
print(r["step_ms_rank0"]["all"])
PY
