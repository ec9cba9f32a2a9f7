#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err; echo "rc=$?"
grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/n${N}_bench.err | tail -5 | cut -c1-300
python - <<PY
import json
d = json.loads(open("gpurun_out/n${N}_bench.json").read().strip().splitlines()[-1])
print("N", d["n_gpus"], "slices/s", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "bf16", round(d["bf16_stream"]["slices_per_s"], 1), "highres", round(d["highres"]["slices_per_s"], 1))
print("train", round(d["train"]["ms_per_step"], 2), "ms allreduce", d["train"]["allreduce_ms"], "| gan", round(d["train_gan"]["ms_per_step"], 2), d["train_gan"]["allreduce_ms"])
print("dp_check", d["train"]["dp_check"])
PY
