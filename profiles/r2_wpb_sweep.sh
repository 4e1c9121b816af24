R=gpurun_out/r2_wpb_sweep.log; : > $R
python profiles/r2_wpb_sweep.py 12 1 2>&1 | grep wpb >> $R
GMJ_LIB=$PWD/build/variants/lib_w8.so python profiles/r2_wpb_sweep.py 8 1 2>&1 | grep wpb >> $R
GMJ_LIB=$PWD/build/variants/lib_w4.so python profiles/r2_wpb_sweep.py 4 1 2>&1 | grep wpb >> $R
GMJ_LIB=$PWD/build/variants/lib_w6x2.so python profiles/r2_wpb_sweep.py 6 2 2>&1 | grep wpb >> $R
GMJ_LIB=$PWD/build/variants/lib_w4x3.so python profiles/r2_wpb_sweep.py 4 3 2>&1 | grep wpb >> $R
cat $R
