mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_mjenv_sequence.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2_seq.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_seq.log); tail -40 gpurun_out/r2_seq.log | cut -c1-300
