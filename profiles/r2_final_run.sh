# round-2 evidence: launch list of the bench command, full captures of k_step at the two bench shapes, compute-sanitizer logs
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $B > gpurun_out/r2_bench_short.json 2> gpurun_out/r2_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_bench_launches.csv $B > gpurun_out/ncu_launch.log 2>&1
echo "launch list rc=$?"
N=5 ENVS=4096 TAG=r2_k_step_n5 timeout 900 bash profiles/r2_ncu.sh
N=8 ENVS=16384 TAG=r2_k_step_n8 timeout 1200 bash profiles/r2_ncu.sh
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" 2>&1 | tail -2
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 900 $CS --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_sanitizer_memcheck_smoke.log 2>&1; echo "memcheck smoke rc=$?"
timeout 900 $CS --tool memcheck --error-exitcode 9 python -m pytest tests -m gpu -q -x -k "finger_geometry_sweep_n8_grasp and fp32 and 0.001" > gpurun_out/r2_sanitizer_memcheck_n8.log 2>&1; echo "memcheck n8 rc=$?"
timeout 900 $CS --tool memcheck --error-exitcode 9 python -m pytest tests -m gpu -q -x -k "wide_objects_on_ground_and_grasp and fp32" > gpurun_out/r2_sanitizer_memcheck_n24.log 2>&1; echo "memcheck n24 rc=$?"
timeout 900 $CS --tool racecheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_sanitizer_racecheck_smoke.log 2>&1; echo "racecheck smoke rc=$?"
timeout 900 $CS --tool synccheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_sanitizer_synccheck_smoke.log 2>&1; echo "synccheck smoke rc=$?"
timeout 900 $CS --tool racecheck --error-exitcode 9 python -m pytest tests -m gpu -q -x -k "wide_objects_on_ground_and_grasp and fp32" > gpurun_out/r2_sanitizer_racecheck_n24.log 2>&1; echo "racecheck n24 rc=$?"
for f in gpurun_out/r2_sanitizer_*.log; do echo "== $f"; tail -4 $f; done
