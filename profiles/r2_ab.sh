# GPU tests + A/B timing of the current build; extra variants: VARIANTS="name1 name2" (build/variants/lib_<name>.so)
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then
(timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log)
tail -5 gpurun_out/r2_tests.log
fi
R=gpurun_out/r2_ab.log; : > $R
for cfg in "4096 5 40 20" "16384 5 30 10" "16384 8 25 10"; do
  python profiles/ab_bench.py $cfg 2>&1 | tail -1 >> $R
  for v in $VARIANTS; do
    GMJ_LIB=$PWD/build/variants/lib_$v.so python profiles/ab_bench.py $cfg 2>&1 | tail -1 >> $R
  done
done
cat $R
