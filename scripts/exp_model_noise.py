"""Experiment harness (GPU box): forward differences between the ATen network, the NCHW network with
the glue kernels and the channels_last network (same parameters, TF32 off)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import dvs_of_training_framework_b200.evflownet as m  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(3)
a = m.Model('cuda')
with torch.no_grad():
    for p in a.parameters():
        if p.dim() == 1:
            p.add_(torch.randn_like(p) * 0.05)
b = m.Model('cuda')
b.load_state_dict(a.state_dict())
b = b.to(memory_format=torch.channels_last)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 3
x = torch.randn(N, 4, 64, 96, device='cuda')
ts = torch.tensor([0., 1.] * N, device='cuda')
sidx = torch.arange(N, device='cuda').repeat_interleave(2)


def run(model, aten):
    ok, up = m._fused_bn_ok, m.upsample2x_reflect_pad
    cl = m._is_cl
    if aten:
        m._fused_bn_ok = lambda z, bn: False
        m._is_cl = lambda t: False
        m.upsample2x_reflect_pad = lambda t: torch.nn.functional.pad(
            torch.nn.functional.interpolate(t, scale_factor=2), (1, 1, 1, 1), mode='reflect')
    try:
        sd = {k: v.clone() for k, v in model.state_dict().items()}
        out = model(x, ts, sidx, (64, 96), raw=False, intermediate=False)[0]
        model.load_state_dict(sd)          # undo the running-statistics update
        return [o.detach().double() for o in out]
    finally:
        m._fused_bn_ok, m.upsample2x_reflect_pad, m._is_cl = ok, up, cl


A, B, C = run(a, True), run(a, False), run(b, False)
A64 = None
for name, u, v in (('aten vs nchw-glue', A, B), ('aten vs channels_last', A, C), ('nchw-glue vs channels_last', B, C)):
    print(name, ['%.2e' % float((p - q).abs().max()) for p, q in zip(u, v)],
          'rel L2', ['%.2e' % float((p - q).norm() / p.norm()) for p, q in zip(u, v)])
