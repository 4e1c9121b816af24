# round-1 final evidence: bench lines, launch list, full captures of the two step kernels
mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/bench_n5.json 2> gpurun_out/bench_n5.err
timeout 300 python bench.py --segments 8 --envs 16384 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
timeout 300 python bench.py --segments 5 --envs 16384 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n5_16k.json 2> gpurun_out/bench_n5_16k.err
timeout 600 python bench.py --segments 24 --envs 1024 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n24_1k.json 2> gpurun_out/bench_n24_1k.err
timeout 900 python bench.py --segments 24 --envs 16384 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n24_16k.json 2> gpurun_out/bench_n24_16k.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r1_bench_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python profiles/profile_step.py 4096 5 16 > gpurun_out/p5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step --launch-skip 26 --launch-count 1 -o gpurun_out/k_step_r1_final_n5 -f python profiles/profile_step.py 4096 5 16 > gpurun_out/ncu5.log 2>&1
python profiles/profile_step.py 16384 8 12 > gpurun_out/p8.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step --launch-skip 18 --launch-count 1 -o gpurun_out/k_step_r1_final_n8 -f python profiles/profile_step.py 16384 8 12 > gpurun_out/ncu8.log 2>&1
python profiles/profile_step.py 592 24 5 > gpurun_out/p24.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step_wide --launch-skip 8 --launch-count 1 -o gpurun_out/k_step_wide_r1_final_n24 -f python profiles/profile_step.py 592 24 5 > gpurun_out/ncu24.log 2>&1
cat gpurun_out/bench_n5.json; for f in n8 n5_16k n24_1k n24_16k; do cut -c1-260 gpurun_out/bench_$f.json; done; cat gpurun_out/p5.log gpurun_out/p8.log gpurun_out/p24.log
