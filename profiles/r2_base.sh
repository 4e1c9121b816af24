# round-2 baseline: GPU tests + A/B timing of the unchanged round-1 build
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_base_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_base_tests.log)
tail -3 gpurun_out/r2_base_tests.log
R=gpurun_out/r2_base_ab.log; : > $R
for cfg in "4096 5 40 20" "16384 5 30 10" "16384 8 25 10"; do
  python profiles/ab_bench.py $cfg 2>&1 | tail -1 >> $R
done
cat $R
