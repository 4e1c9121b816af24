mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t.log)
tail -4 gpurun_out/t.log
R=gpurun_out/ab.log; : > $R
for cfg in "4096 5 40 20" "16384 8 25 10"; do
  GMJ_BALANCE=0 GMJ_ASIN_TAB=0 python profiles/ab_bench.py $cfg 2>&1 | tail -1 | sed 's/^/base(bal0,tab0) /' >> $R
  GMJ_BALANCE=0 python profiles/ab_bench.py $cfg 2>&1 | tail -1 | sed 's/^/bal0,tab1 /' >> $R
  GMJ_ASIN_TAB=0 python profiles/ab_bench.py $cfg 2>&1 | tail -1 | sed 's/^/bal1,tab0 /' >> $R
  python profiles/ab_bench.py $cfg 2>&1 | tail -1 | sed 's/^/bal1,tab1 /' >> $R
  for v in nb5 nb8 nb58 w14 w10; do
    GMJ_LIB=$PWD/build/variants/lib_$v.so python profiles/ab_bench.py $cfg 2>&1 | tail -1 >> $R
  done
done
cat $R
