for v in s10 s12; do GMJ_LIB=$PWD/build/variants/lib_$v.so python profiles/ab_bench.py 4096 5 40 20 2>&1 | tail -1; done
python profiles/ab_bench.py 4096 5 40 20 2>&1 | tail -1
