R=gpurun_out/r2_env_cost.log; : > $R
GMJ_LIB=$PWD/build/variants/lib_prof.so python profiles/r2_env_cost.py 1776 0 >> $R 2>&1
cat $R
