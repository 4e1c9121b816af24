# 8-GPU weak-scaling lines of round 2 (launched the way the driver does)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_scale8_default.json 2> gpurun_out/r2_scale8_default.err; echo "rc=$?"
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --envs 16384 --segments 5 --no-cpu-baseline --no-extras > gpurun_out/r2_scale8_n5_16k.json 2> gpurun_out/r2_scale8_n5_16k.err; echo "rc=$?"
python - <<'PY'
import json
for f in ("r2_scale8_default", "r2_scale8_n5_16k"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, d["n_gpus"], round(d["value"]), round(d["e2e"]["value"]), d["ms_per_step"], d["clocks"])
        for x in d.get("extra", {}).get("configs", []):
            print("   extra", x["config"]["segments"], x["config"]["envs_per_gpu"], round(x["value"]), round(x["e2e"]["value"]), x["ms_per_step"])
    except Exception as ex:
        print(f, "failed", ex)
PY
tail -3 gpurun_out/r2_scale8_default.err
