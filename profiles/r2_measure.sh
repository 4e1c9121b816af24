# FP32-vs-oracle error measurement at contract size + a trial of the full bench line
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q ${R2_K:+-k "$R2_K"} > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log); tail -4 gpurun_out/r2_tests.log
timeout 150 python profiles/r2_fp32_err.py 4096 5 10 > gpurun_out/r2_err_n5.log 2>&1; tail -13 gpurun_out/r2_err_n5.log
timeout 150 python profiles/r2_fp32_err.py 16384 8 6 512 > gpurun_out/r2_err_n8.log 2>&1; tail -9 gpurun_out/r2_err_n8.log
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_try.json 2> gpurun_out/r2_bench_try.err; tail -c 600 gpurun_out/r2_bench_try.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_try.json"))
print(d["value"], d["e2e"], d["roofline"]["frac"], d["stats"], d["config"]["timestep"])
for x in d["extra"]["configs"]:
    print(x["config"]["segments"], x["value"], x["e2e"]["value"], x["roofline"]["frac"], x["roofline"]["mean_contacts_per_substep"], x["stats"]["calibration_s"], x["ms_per_step"])
PY
