mkdir -p gpurun_out
timeout 600 python bench.py --preset termination_action_v1 --steps 10 --warmup 3 --cpu-seconds 6 --no-extras > gpurun_out/r2_bench_preset.json 2> gpurun_out/r2_bench_preset.err; echo "rc=$?"; tail -c 300 gpurun_out/r2_bench_preset.err; cut -c1-1200 gpurun_out/r2_bench_preset.json
