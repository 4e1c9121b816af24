mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log); grep "MEASURED\|passed\|failed\|rc=\|^E  " gpurun_out/r2_tests.log | cut -c1-420
