mkdir -p gpurun_out
timeout 500 python bench.py --no-cpu-baseline > gpurun_out/r2_bench_3extras.json 2> gpurun_out/r2_bench_3extras.err; echo "rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_3extras.json"))
print("bench", d["value"], d["e2e"]["value"], d["ms_per_step"])
for x in d["extra"]["configs"]: print("  extra", x["config"]["segments"], x["config"]["workload"][:40], x["value"], x["e2e"]["value"], x["ms_per_step"], x["roofline"]["frac"])
PY
