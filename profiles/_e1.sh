mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_par.log 2>&1; echo "rc=$?" >> gpurun_out/r2_par.log)
tail -4 gpurun_out/r2_par.log
SKIP_TESTS=1 VARIANTS="base" bash profiles/r2_ab.sh
