mkdir -p gpurun_out
for rep in 1 2; do for d in 0 1; do
DSB200_DEEP_LOOKAHEAD=$d timeout 300 python bench.py --workload C3 --single-workload --no-baselines --no-sharded > gpurun_out/r02q_c3_deep${d}_$rep.json 2> gpurun_out/r02q.err; echo "rc=$?"
python - <<PY
import json
r=json.loads(open("gpurun_out/r02q_c3_deep${d}_$rep.json").read().strip().splitlines()[-1])
print("deep=$d rep=$rep", round(r["value"]), round(r["e2e"]["value"]), r["parity_check"], r["e2e"]["detail"])
PY
done; done
