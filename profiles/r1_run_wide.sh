# round-1 evidence run: wide kernel tests + benches (configs[4] sweep points)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/tw.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tw.log)
tail -5 gpurun_out/tw.log
timeout 300 python profiles/profile_step.py 2048 24 5 > gpurun_out/p24.log 2>&1; cat gpurun_out/p24.log
timeout 300 python profiles/profile_step.py 4096 16 5 > gpurun_out/p16.log 2>&1; cat gpurun_out/p16.log
( time timeout 1200 python bench.py --segments 24 --envs 4096 --steps 3 --warmup 3 --no-cpu-baseline ) > gpurun_out/b24.log 2> gpurun_out/b24.err
cat gpurun_out/b24.log
