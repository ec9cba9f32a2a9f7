"""ADVICE (round 1): launches must follow the tensors' device, not the caller's current device, and per-function
attributes / SM counts are per device.  Needs two GPUs (skipped otherwise; run with `gpurun --gpus 2`)."""
import pytest
import torch

from oracle import ngswin_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_model_on_cuda1_while_current_device_is_cuda0():
    import torch.nn.functional as F
    import transmar_gan_b200 as T
    sd = O.synthetic_state_dict("W1", seed=1)
    x = O.synthetic_input(1, 64, 96, seed=999)
    t = O.synthetic_input(1, 64, 96, seed=1000)
    outs, grads = [], []
    for dev in ("cuda:0", "cuda:1"):
        torch.cuda.set_device(0)                         # the caller's current device stays cuda:0
        m = T.NGswin()
        m.load_state_dict(sd, strict=True)
        m = m.to(dev).eval()
        with torch.no_grad():
            outs.append(m(x.to(dev)).cpu())              # inference kernels (> 48 KB smem attributes per device)
        m.train()
        loss = F.l1_loss(m(x.to(dev)), t.to(dev))
        loss.backward()
        grads.append(torch.cat([p.grad.flatten().cpu() for p in m.parameters()]))
        assert torch.cuda.current_device() == 0
    assert torch.equal(outs[0], outs[1])                 # forward is bitwise repeatable, whatever the device
    assert (grads[0] - grads[1]).norm() <= 2e-2 * grads[0].norm()
    with pytest.raises(RuntimeError):                    # tensors on two devices in one op: refused
        T.ops.unigram_fwd(torch.zeros(1, 64 * 64, 64, device="cuda:0"), m.encoder_layer1.blocks[0].kernel_params()[0],
                          m.encoder_layer1.blocks[0].kernel_params()[1], 64, 64)
    assert T.ops.tc_error() == 0
