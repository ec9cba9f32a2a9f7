#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_bf16_stream.py -m gpu -x -q -s > gpurun_out/f_bf16_tests.log 2>&1; echo "bf16 tests rc=$?"
grep -h "bf16-stream\]" gpurun_out/f_bf16_tests.log; tail -15 gpurun_out/f_bf16_tests.log | cut -c1-200
timeout 900 python bench.py --no-train --no-cpu --no-highres --steps 10 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/f_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/f_bench.json").read().strip().splitlines()[-1])
print("default slices/s", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "| bf16_stream", {k: (round(v, 5) if isinstance(v, float) else v) for k, v in d["bf16_stream"].items() if k != "note"})
PY
