mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_self_collision.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r2_sc.log 2>&1; echo "rc=$?" >> gpurun_out/r2_sc.log)
tail -4 gpurun_out/r2_sc.log
