mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log); grep "MEASURED\|passed\|failed\|rc=\|^E  " gpurun_out/r2_tests.log | cut -c1-700
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_try.json 2> gpurun_out/r2_bench_try.err; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_try.json')); print(d['value'], d['e2e']['value'], d['stats'])"
