#!/bin/bash
# Round 2, GPU call A: the no-alloc test alone in a fresh process, the whole GPU suite, the bench line.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/a_zz_alone.log 2>&1; echo "zz_alone rc=$?" | tee -a gpurun_out/a_status.txt
timeout 2400 python -m pytest tests -m gpu -x -q -s > gpurun_out/a_gpu_tests.log 2>&1; echo "gpu_tests rc=$?" | tee -a gpurun_out/a_status.txt
tail -5 gpurun_out/a_gpu_tests.log
grep -h "\[auto\]\|\[convergence\]\|\[auto/train\]" gpurun_out/a_gpu_tests.log | head -40
timeout 1500 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?" | tee -a gpurun_out/a_status.txt
tail -3 gpurun_out/a_bench.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/a_bench.json").read().strip().splitlines()[-1])
    r = d["roofline"]
    print("slices/s", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "attn6 ms", r["avg_launch_ms"], "frac", r["frac"])
    print("plan", d["config"].get("precision_plan"))
    for k in ("train", "train_gan"):
        t = d[k]
        print(k, round(t["ms_per_step"], 2), "ms; vs_gpu_eager", t.get("vs_gpu_eager"), "roofline", {a: round(b, 4) for a, b in t["roofline"].items() if isinstance(b, float)})
    print("cpu", d["cpu_baseline"]["value"], "gpu_eager", d["cpu_baseline"]["gpu_eager"]["value"], "bf16", d["cpu_baseline"]["gpu_eager_autocast_bf16"]["value"])
except Exception as e:
    print("bench parse failed:", e)
PY
