mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log); tail -4 gpurun_out/r2_tests.log | cut -c1-250
timeout 300 python profiles/ab_bench.py 1024 24 6 3 2>&1 | tail -1
timeout 300 python profiles/ab_bench.py 4096 24 4 2 2>&1 | tail -1
timeout 300 python profiles/ab_bench.py 4096 16 4 2 2>&1 | tail -1
