"""Train-step glue kernels against the ATen ops they replace (EV_FlowNet/net.py:72-85)."""
import pytest
import torch
import torch.nn.functional as F


def _ref(x):
    return F.pad(F.interpolate(x, scale_factor=2), (1, 1, 1, 1), mode='reflect')


@pytest.mark.gpu
@pytest.mark.parametrize('shape', [(2, 3, 1, 1), (1, 2, 1, 5), (2, 5, 4, 1), (3, 7, 16, 20), (2, 130, 33, 47),
                                   (1, 1, 128, 128)])
@pytest.mark.parametrize('channels_last', [False, True])
def test_upsample_reflect_pad_matches_aten(shape, channels_last):
    from dvs_of_training_framework_b200.evflownet import upsample2x_reflect_pad
    g = torch.Generator(device='cuda').manual_seed(sum(shape))
    x = torch.randn(shape, generator=g, device='cuda')
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    x.requires_grad_()
    y = upsample2x_reflect_pad(x)
    if channels_last and not x.is_contiguous():          # 1x1 planes are both layouts at once
        assert y.is_contiguous(memory_format=torch.channels_last)
    xr = x.detach().clone().requires_grad_()
    yr = _ref(xr)
    assert y.shape == yr.shape and torch.equal(y, yr)                 # a gather: exact
    gy = torch.randn(y.shape, generator=g, device='cuda')
    y.backward(gy)
    yr.backward(gy)
    # the adjoint sums up to 16 values per element; the order differs from ATen's two kernels
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-5, atol=1e-5)
    # and it is the exact adjoint: <y(x), gy> == <x, grad>
    lhs = (y.detach().double() * gy.double()).sum()
    rhs = (x.detach().double() * x.grad.double()).sum()
    torch.testing.assert_close(lhs, rhs, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
def test_upsample_reflect_pad_non_contiguous_input_and_determinism():
    from dvs_of_training_framework_b200.evflownet import upsample2x_reflect_pad
    x = torch.randn(2, 6, 9, 11, device='cuda')[:, ::2]
    assert torch.equal(upsample2x_reflect_pad(x), _ref(x))
    xs = torch.randn(4, 8, 32, 32, device='cuda', requires_grad=True)
    gy = torch.randn(4, 8, 66, 66, device='cuda')
    grads = []
    for _ in range(2):
        xs.grad = None
        upsample2x_reflect_pad(xs).backward(gy)
        grads.append(xs.grad.clone())
    assert torch.equal(grads[0], grads[1])


@pytest.mark.gpu
@pytest.mark.parametrize('shape', [(4, 3, 5, 7), (32, 32, 64, 64), (2, 512, 16, 16), (1, 8, 1, 2), (3, 5, 8, 8), (5, 96, 9, 7),
                                   (8, 130, 20, 12)])
@pytest.mark.parametrize('channels_last', [False, True])
def test_relu_batchnorm_matches_aten(shape, channels_last):
    from torch import nn
    from dvs_of_training_framework_b200.evflownet import relu_batchnorm
    g = torch.Generator(device='cuda').manual_seed(sum(shape))
    C = shape[1]
    bn = nn.BatchNorm2d(C).cuda()
    ref = nn.BatchNorm2d(C).cuda()
    with torch.no_grad():
        bn.weight.copy_(torch.rand(C, generator=g, device='cuda') + 0.5)
        bn.bias.copy_(torch.randn(C, generator=g, device='cuda'))
        bn.running_mean.copy_(torch.randn(C, generator=g, device='cuda'))
        bn.running_var.copy_(torch.rand(C, generator=g, device='cuda') + 0.5)
    ref.load_state_dict(bn.state_dict())
    z = torch.randn(shape, generator=g, device='cuda')
    if channels_last:
        z = z.contiguous(memory_format=torch.channels_last)
    z.requires_grad_()
    zr = z.detach().clone().requires_grad_()
    cb = cbr = None
    if sum(shape) % 2 == 0:                   # half of the cases: a convolution bias folded in
        cb = torch.randn(C, generator=g, device='cuda').requires_grad_()
        cbr = cb.detach().clone().requires_grad_()
    y = relu_batchnorm(z, bn, cb)
    yr = ref(torch.relu(zr if cbr is None else zr + cbr.view(1, -1, 1, 1)))
    torch.testing.assert_close(y, yr, rtol=2e-5, atol=2e-5)
    torch.testing.assert_close(bn.running_mean, ref.running_mean, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-5, atol=1e-6)
    assert int(bn.num_batches_tracked) == int(ref.num_batches_tracked) == 1
    gy = torch.randn(shape, generator=g, device='cuda')
    y.backward(gy)
    yr.backward(gy)
    torch.testing.assert_close(z.grad, zr.grad, rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(bn.weight.grad, ref.weight.grad, rtol=2e-4, atol=2e-4)
    torch.testing.assert_close(bn.bias.grad, ref.bias.grad, rtol=2e-5, atol=2e-4)
    if cb is not None:
        torch.testing.assert_close(cb.grad, cbr.grad, rtol=2e-4, atol=2e-4)
    # evaluation mode goes through ATen (running statistics)
    bn.eval()
    assert torch.equal(relu_batchnorm(z.detach(), bn), bn(torch.relu(z.detach())))
    assert int(bn.num_batches_tracked) == 1


@pytest.mark.gpu
def test_relu_batchnorm_single_value_per_channel_raises_like_torch():
    from torch import nn
    from dvs_of_training_framework_b200.evflownet import relu_batchnorm
    with pytest.raises(ValueError):
        relu_batchnorm(torch.randn(1, 8, 1, 1, device='cuda'), nn.BatchNorm2d(8).cuda())


@pytest.mark.gpu
@pytest.mark.parametrize('chans,pad_in', [((8, 8), 0), ((256, 2, 256), 6), ((6, 3), 7), ((4,), 4), ((130, 64, 2, 4), 0)])
def test_upsample_reflect_pad_cat_matches_aten(chans, pad_in):
    from dvs_of_training_framework_b200.evflownet import upsample2x_reflect_pad_cat
    g = torch.Generator(device='cuda').manual_seed(sum(chans) + pad_in)
    N, H, W = 3, 9, 13
    parts = [torch.randn(N, c, H, W, generator=g, device='cuda').contiguous(memory_format=torch.channels_last)
             for c in chans]
    for i, p in enumerate(parts):
        p.requires_grad_(i != 1 or len(parts) == 1)          # one part without gradient
    refs = [p.detach().clone().requires_grad_(p.requires_grad) for p in parts]
    y = upsample2x_reflect_pad_cat(parts, pad_in)
    zeros = [torch.zeros(N, pad_in, H, W, device='cuda')] if pad_in else []
    yr = _ref(torch.cat(refs + zeros, 1))
    assert y.shape == yr.shape and torch.equal(y, yr)
    gy = torch.randn(y.shape, generator=g, device='cuda')
    y.backward(gy)
    yr.backward(gy)
    for p, r in zip(parts, refs):
        if p.requires_grad:
            assert p.grad.is_contiguous(memory_format=torch.channels_last)
            torch.testing.assert_close(p.grad, r.grad, rtol=1e-5, atol=1e-5)
        else:
            assert p.grad is None


def _model_run(model, x, aten):
    """Flows and parameter gradients of one training-mode pass; aten=True disables every glue kernel."""
    import dvs_of_training_framework_b200.evflownet as m
    saved = (m._fused_bn_ok, m.upsample2x_reflect_pad, m._is_cl)
    if aten:
        m._fused_bn_ok = lambda z, bn: False
        m._is_cl = lambda t: False
        m.upsample2x_reflect_pad = _ref
    try:
        N = x.shape[0]
        ts = torch.tensor([0., 1.] * N, device='cuda')
        sidx = torch.arange(N, device='cuda').repeat_interleave(2)
        state = {k: v.clone() for k, v in model.state_dict().items()}
        model.zero_grad(set_to_none=True)
        flows, _, _, feats = model(x, ts, sidx, tuple(x.shape[-2:]), raw=False, intermediate=True)
        sum((f * torch.linspace(-1, 1, f.numel(), device='cuda').view_as(f)).sum() for f in flows).backward()
        grads = {n: p.grad.detach().double().clone() for n, p in model.named_parameters()}
        model.load_state_dict(state)                         # undo the running-statistics update
        return [f.detach().double() for f in flows], grads, feats
    finally:
        m._fused_bn_ok, m.upsample2x_reflect_pad, m._is_cl = saved


@pytest.mark.gpu
def test_channels_last_model_equals_nchw_model():
    """The channels_last network (parts fed to the gather in the order (h, skip, flow + zeros), weight
    rearranged on the fly, bias folded into ReLU+BatchNorm) computes what the network built from
    ATen ops computes from the same parameters: flows and every parameter gradient, to within the
    rounding noise that the NCHW network with the glue kernels shows against ATen as well (the
    13-layer chain with batch statistics amplifies fp32 rounding; scripts/exp_model_noise.py)."""
    from dvs_of_training_framework_b200.evflownet import Model
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        torch.manual_seed(3)
        a = Model('cuda')
        with torch.no_grad():                       # non-trivial biases (initialised to zero)
            for p in a.parameters():
                if p.dim() == 1:
                    p.add_(torch.randn_like(p) * 0.05)
        b = Model('cuda')
        b.load_state_dict(a.state_dict())
        b = b.to(memory_format=torch.channels_last)
        x = torch.randn(4, 4, 64, 96, device='cuda')
        f0, g0, feats0 = _model_run(a, x, aten=True)
        f1, g1, _ = _model_run(a, x, aten=False)
        f2, g2, feats2 = _model_run(b, x, aten=False)
        assert feats0.keys() == feats2.keys() and all(feats0[k].shape == feats2[k].shape for k in feats0)
        rel = lambda u, v: float((u - v).norm() / (u.norm() + 1e-30))       # noqa: E731
        for u, v, w in zip(f0, f1, f2):
            assert rel(u, v) < 1e-4 and rel(u, w) < 1e-4
        for n in g0:
            noise = rel(g0[n], g1[n])
            assert rel(g0[n], g2[n]) <= 4 * noise + 1e-5, (n, rel(g0[n], g2[n]), noise)
            assert noise < 5e-2, (n, noise)
    finally:
        torch.backends.cudnn.allow_tf32 = old
