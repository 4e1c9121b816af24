mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_contract.py -m gpu -x -q -s > gpurun_out/r2_contract.log 2>&1; echo "rc=$?" >> gpurun_out/r2_contract.log)
grep -n "self-overlap\|passed\|failed\|rc=" gpurun_out/r2_contract.log
