# round-2 final evidence: bench lines of both arms, launch list of the bench command, full captures of k_step at the two
# bench shapes and of the wide kernel (compute-sanitizer is closed on this pool: profiles/r2_sanitizer_closed.txt)
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2_bench_full.json 2> gpurun_out/r2_bench_full.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/r2_bench_short.json 2> gpurun_out/r2_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_bench_launches.csv $B > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
N=5 ENVS=4096 TAG=r2_k_step_n5 timeout 900 bash profiles/r2_ncu.sh 2>&1 | tail -1
N=8 ENVS=16384 TAG=r2_k_step_n8 timeout 1200 bash profiles/r2_ncu.sh 2>&1 | tail -1
python profiles/profile_step.py 592 24 5 > gpurun_out/p24.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step_wide --launch-skip 8 --launch-count 1 -o gpurun_out/r2_k_step_wide_n24p -f python profiles/profile_step.py 592 24 5 > gpurun_out/ncu24.log 2>&1; tail -1 gpurun_out/ncu24.log
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_bench_full.json"))
print("bench", d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["gpu_launches"], d["clocks"])
for x in d["extra"]["configs"]: print("  extra", x["config"]["segments"], x["value"], x["e2e"]["value"], x["ms_per_step"], x["roofline"]["frac"])
r = json.load(open("gpurun_out/r2_bench_ref.json")); print("ref", r["value"], r["cpu_baseline"]["cores"], r["config"] == d["config"])
PY
