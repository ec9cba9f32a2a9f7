#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 --no-highres > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "rc=$?"
tail -5 gpurun_out/n2_bench.err | cut -c1-300
python - <<'PY'
import json
d = json.loads(open("gpurun_out/n2_bench.json").read().strip().splitlines()[-1])
print("N=2 slices/s", round(d["value"], 1), "train", round(d["train"]["ms_per_step"], 2), "ms allreduce", d["train"]["allreduce_ms"], "gan", round(d["train_gan"]["ms_per_step"], 2), d["train_gan"]["allreduce_ms"])
print("dp_check", d["train"]["dp_check"])
PY
