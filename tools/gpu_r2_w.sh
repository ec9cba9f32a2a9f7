#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/w_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -3 gpurun_out/w_tests.log | cut -c1-200
python - <<'PY'
import torch, transmar_gan_b200 as T
from oracle import ngswin_oracle as O
ops = T.ops
sd = {k: v.cuda() for k, v in O.synthetic_state_dict("W1", seed=1).items()}
for nh, pre, (B, ph, pw) in ((6, "encoder_layer1.blocks.1.", (32, 416, 416)), (4, "encoder_layer2.blocks.0.", (32, 208, 208))):
    P = [sd[pre + n].contiguous() for n in ops.NSTB_PARAM_NAMES]
    z = torch.randn(B, ph * pw, 64, device="cuda")
    for _ in range(3): c = ops.ngram_ctx_infer(z, *P[0:10], ph, pw, nh)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): c = ops.ngram_ctx_infer(z, *P[0:10], ph, pw, nh)
    e1.record(); torch.cuda.synchronize()
    print("ngram_ctx_infer (unigram_qkv + ctx) nh", nh, round(e0.elapsed_time(e1) / 10, 4), "ms")
PY
python bench.py --no-train --no-cpu --no-highres --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('W1', round(d['value'],1), round(d['ms_per_step'],2))"
