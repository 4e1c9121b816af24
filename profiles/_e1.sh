mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_presets.py tests/test_gpu_contract.py -m gpu -q -s > gpurun_out/r2_presets.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_presets.log); grep "MEASURED\|passed\|failed\|^E  " gpurun_out/r2_presets.log | cut -c1-420
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_try.json 2> gpurun_out/r2_bench_try.err; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_try.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])
for x in d['extra']['configs']: print(x['config']['segments'], x['value'], x['e2e']['value'], x['ms_per_step'], x['roofline']['frac'])"
