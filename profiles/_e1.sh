mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_tests.log); tail -5 gpurun_out/r2_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_try.json 2> gpurun_out/r2_bench_try.err; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_try.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['gpu_launches'])
for x in d['extra']['configs']: print(x['config']['segments'], x['value'], x['e2e']['value'], x['ms_per_step'], x['roofline']['frac'])"
