mkdir -p gpurun_out
R=gpurun_out/r2_ab.log; : > $R
python profiles/ab_bench.py 4096 5 30 12 2>&1 | tail -1 >> $R
for v in nocol nodense neither; do
  GMJ_LIB=$PWD/build/variants/lib_$v.so python profiles/ab_bench.py 4096 5 30 12 2>&1 | tail -1 >> $R
done
cat $R
