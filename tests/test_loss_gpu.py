"""GPU parity: csrc/loss.cu (fused multiscale warp loss, forward + analytic backward) through
the C ABI vs (1) the reference's own outputs committed under tests/golden (values AND gradients
of utils/loss.py run unmodified, see make_golden.py), (2) the reference's known-answer tests
tests/loss/test_loss.py:20,41,63, (3) the torch-fp32 CPU oracle on seeded inputs.
Bar (north_star): rtol 1e-5 / atol 1e-6 in fp32."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import dvs_of_training_framework_b200 as dv
from oracle import loss_oracle as lo
from tests.helpers import synth_loss_inputs

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-6
WEIGHTS = (0.5, 1.0, 1.0)          # utils/options.py loss weights (smooth, photo, outborder)


def _combined(terms):
    return sum(w * sum(t) / len(t) for w, t in zip(WEIGHTS, terms))    # utils/training.py:12-24


def _run_gpu(inp, flows_cpu, batch_size=None):
    shapes = [tuple(f.shape[-2:]) for f in flows_cpu]
    P = flows_cpu[0].shape[0]
    flows = [f.detach().clone().cuda().requires_grad_() for f in flows_cpu]
    ev = dv.Losses(shapes, batch_size or P, 'cuda')
    terms = ev(flows, inp['flow_ts'].cuda(), inp['flow_sample_idx'].cuda(), inp['images'].cuda(),
               inp['timestamps'].cuda(), inp['sample_idx'].cuda())
    return flows, terms


def _oracle_grads(inp, flows_cpu, scale=1.0, cot=None):
    fl = [f.detach().clone().requires_grad_() for f in flows_cpu]
    terms = lo.multiscale_loss(fl, inp['flow_ts'], inp['flow_sample_idx'], inp['images'] * scale,
                               inp['timestamps'], inp['sample_idx'])
    total = lo.combined(terms) if cot is None else sum(cot[k, s] * terms[k][s] for k in range(3) for s in range(4))
    total.backward()
    return terms, [f.grad.numpy() for f in fl]


WIDENED = {}        # test id -> the largest count of gradient elements that needed the allowance


def assert_grads_close(got, want, inp, flows_cpu, cot=None):
    """rtol 1e-5 / atol 1e-6 (north_star) on every gradient element, widened ONLY where the
    reference itself is ill-conditioned: where the warped frame nearly equals the previous one the
    Charbonnier derivative c'(d) = 0.9 d (d^2+1e-6)^-0.55 has slope ~1e3, so a 1-ulp change of an
    interpolated grey value moves the reference's own gradient by more than the tolerance.  The
    allowance is that measured sensitivity: |grad_ref(images) - grad_ref(images * (1 +- 2^-22))|
    (a 1-2 ulp relative perturbation of the frames), times 4.  It is exactly zero for
    well-conditioned pixels, and it must stay rare.  Returns (elements that needed the allowance,
    elements compared); the per-test maximum is printed in the test log (`-rA` / the summary
    line written by the session hook below) so the relaxation is reported next to the claim."""
    import os
    _, up = _oracle_grads(inp, flows_cpu, 1.0 + 2.0 ** -22, cot)
    _, dn = _oracle_grads(inp, flows_cpu, 1.0 - 2.0 ** -22, cot)
    total_w, total_n = 0, 0
    for g, w, a, b in zip(got, want, up, dn):
        g = g.cpu().numpy() if hasattr(g, 'cpu') else g
        sens = 4.0 * np.maximum(np.abs(a - w), np.abs(b - w))
        err = np.abs(g - w)
        base = ATOL + RTOL * np.abs(w)
        widened = int((err > base).sum())
        assert widened <= max(2, g.size // 20000), f'{widened} of {g.size} elements need the allowance'
        bad = err > base + sens
        assert not bad.any(), (f'{int(bad.sum())} / {g.size} gradient elements off; worst abs '
                               f'{float(err[bad].max()):.3e}')
        total_w += widened
        total_n += g.size
    test = os.environ.get('PYTEST_CURRENT_TEST', '?').split(' ')[0]
    prev = WIDENED.get(test, (0, 0))
    WIDENED[test] = (max(prev[0], total_w), max(prev[1], total_n))
    print(f'[grad tolerance] {test}: {total_w} of {total_n} gradient elements outside rtol 1e-5 / atol 1e-6 '
          f'(inside 4x the reference\'s own 1-ulp sensitivity)')
    return total_w, total_n


def _as_array(terms):
    return np.array([[float(v.detach()) for v in t] for t in terms])


@pytest.mark.parametrize('tag', ['a', 'b', 'c'])
def test_matches_reference_golden_values_and_grads(golden, tag):
    g = np.load(golden / f'loss_{tag}.npz')
    inp = {k: torch.tensor(g[k]) for k in ('flow_ts', 'flow_sample_idx', 'timestamps', 'sample_idx')}
    inp['images'] = torch.tensor(g['images']).float()
    flows, terms = _run_gpu(inp, [torch.tensor(g[f'flow{k}']) for k in range(4)])
    np.testing.assert_allclose(_as_array(terms), g['terms'], rtol=RTOL, atol=ATOL)
    total = _combined(terms)
    np.testing.assert_allclose(float(total), float(g['total']), rtol=RTOL, atol=ATOL)
    total.backward()
    flows_cpu = [torch.tensor(g[f'flow{k}']) for k in range(4)]
    assert_grads_close([f.grad for f in flows], [g[f'grad{k}'] for k in range(4)], inp, flows_cpu)


def test_reference_kat_no_changes():
    # tests/loss/test_loss.py:8-22: zero images, zero flow, 5x6 -> (0.002, 0.002, 0)
    H, W = 5, 6
    ts = torch.tensor([0, 0.4]).cuda()
    ev = dv.Losses([(H, W)], 2, 'cuda')
    out = ev([torch.zeros(1, 2, H, W).cuda()], ts.view(1, 2), torch.LongTensor([0]).cuda(),
             torch.zeros(2, 1, H, W).cuda(), ts, torch.LongTensor([0, 0]).cuda())
    for term, gt in zip(out, [0.002, 0.002, 0]):
        assert len(term) == 1 and abs(float(term[0]) - gt) < 5e-6


def test_reference_kat_fixture(golden):
    # tests/loss/test_loss.py:25-65: goldens 0.622660; 0.002120 / 0.652659 / 0.007802 (+-5e-6)
    H, W = 246, 340
    f = np.load(golden / 'fixture_000001.npz')
    images = torch.tensor(np.stack([f['image1'][:H, :W], f['image2'][:H, :W]])[:, None]).float().cuda()
    ts = torch.tensor([0, float(f['stop']) - float(f['start'])], dtype=torch.float32).cuda()
    sidx, fs = torch.LongTensor([0, 0]).cuda(), torch.LongTensor([0]).cuda()
    ev = dv.Losses([(H, W)], 2, 'cuda')
    out = ev([torch.zeros(1, 2, H, W).cuda()], ts.view(1, 2), fs, images, ts, sidx)
    for term, gt in zip(out, [0.002, 0.622660, 0]):
        assert abs(float(term[0]) - gt) < 5e-6
    flow = torch.tensor(f['flow'][:H, :W]).permute(2, 0, 1)[None].contiguous().cuda()
    out = ev([flow], ts.view(1, 2), fs, images, ts, sidx)
    for term, gt in zip(out, [0.002120, 0.652659, 0.007802]):
        assert abs(float(term[0]) - gt) < 5e-6
    kat = np.load(golden / 'loss_kat.npz')
    np.testing.assert_allclose([float(t[0]) for t in out], kat['pred'], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize('B,H,W,sigma', [
    (1, 260, 346, 3.0),       # cfg1 DAVIS
    (8, 256, 256, 3.0),       # cfg2 MVSEC
    (8, 256, 256, 64.0),      # stress: most pixels warp out of the image
    (3, 72, 104, 3.0),        # ragged tiles (not multiples of 32), odd sizes at coarse scales
    (2, 480, 640, 3.0),       # cfg4 DSEC shape, small batch so the CPU oracle stays quick
])
def test_matches_oracle_values_and_grads(B, H, W, sigma):
    inp = synth_loss_inputs(100 + B + H, B, H, W, sigma)
    flows, terms = _run_gpu(inp, inp['flows'])
    oflows = [f.clone().requires_grad_() for f in inp['flows']]
    oterms = lo.multiscale_loss(oflows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                inp['timestamps'], inp['sample_idx'])
    np.testing.assert_allclose(_as_array(terms), _as_array(oterms), rtol=RTOL, atol=ATOL)
    _combined(terms).backward()
    lo.combined(oterms).backward()
    assert_grads_close([f.grad for f in flows], [f.grad.numpy() for f in oflows], inp, inp['flows'])


def test_per_term_gradients_with_unequal_weights():
    """backward contracts with an arbitrary grad_terms vector (12 independent cotangents)."""
    inp = synth_loss_inputs(7, 2, 64, 96, 6.0)
    flows, terms = _run_gpu(inp, inp['flows'])
    oflows = [f.clone().requires_grad_() for f in inp['flows']]
    oterms = lo.multiscale_loss(oflows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                inp['timestamps'], inp['sample_idx'])
    g = torch.Generator().manual_seed(3)
    cot = torch.rand(3, 4, generator=g) + 0.1
    sum(cot[k, s] * terms[k][s] for k in range(3) for s in range(4)).backward()
    sum(cot[k, s] * oterms[k][s] for k in range(3) for s in range(4)).backward()
    assert_grads_close([f.grad for f in flows], [f.grad.numpy() for f in oflows], inp, inp['flows'], cot)


def test_image_flow_matching_with_shuffled_images():
    """utils/loss.py:182-206: predictions find their frames by (timestamp, sample) equality, in
    any image order; distinct timestamps per sample."""
    B, H, W = 4, 48, 64
    inp = synth_loss_inputs(17, B, H, W, 2.0)
    ts = inp['timestamps'].clone()
    for b in range(B):
        ts[2 * b:2 * b + 2] += 0.1 * b
    inp['timestamps'], inp['flow_ts'] = ts, ts.view(-1, 2).clone()
    perm = torch.tensor([5, 0, 3, 6, 1, 7, 2, 4])
    shuf = dict(inp, images=inp['images'][perm], timestamps=ts[perm], sample_idx=inp['sample_idx'][perm])
    _, t0 = _run_gpu(inp, inp['flows'])
    _, t1 = _run_gpu(shuf, inp['flows'])
    assert np.array_equal(_as_array(t0), _as_array(t1))
    oterms = lo.multiscale_loss(inp['flows'], shuf['flow_ts'], shuf['flow_sample_idx'], shuf['images'],
                                shuf['timestamps'], shuf['sample_idx'])
    np.testing.assert_allclose(_as_array(t1), _as_array(oterms), rtol=RTOL, atol=ATOL)


def test_missing_or_duplicate_frame_raises_assertion():
    inp = synth_loss_inputs(1, 2, 32, 32, 1.0)
    bad = dict(inp, flow_ts=inp['flow_ts'] + 1.0)                        # no frame has that time
    with pytest.raises(AssertionError):
        _run_gpu(bad, inp['flows'])
    dup = dict(inp, sample_idx=torch.zeros_like(inp['sample_idx']))       # two frames match
    with pytest.raises(AssertionError):
        _run_gpu(dup, inp['flows'])


def test_single_scale_loss_object_and_partial_batch():
    # Loss(pred_shape, batch_size, device)(prev, next, flow), N <= batch_size (loss.py:122-124)
    H, W, N = 40, 56, 3
    g = torch.Generator().manual_seed(5)
    prev = torch.randint(0, 256, (N, 1, H, W), generator=g).float()
    nxt = torch.randint(0, 256, (N, 1, H, W), generator=g).float()
    flow = (torch.randn(N, 2, H, W, generator=g) * 4).requires_grad_()
    fl = flow.detach().clone().cuda().requires_grad_()
    sm, ph, ob = dv.Loss((H, W), 8, 'cuda')(prev.cuda(), nxt.cuda(), fl)
    osm, oph, oob = lo.scale_loss(prev, nxt, flow)
    np.testing.assert_allclose([float(sm), float(ph), float(ob)],
                               [float(osm), float(oph), float(oob)], rtol=RTOL, atol=ATOL)
    (0.5 * sm + ph + ob).backward()
    (0.5 * osm + oph + oob).backward()
    np.testing.assert_allclose(fl.grad.cpu().numpy(), flow.grad.numpy(), rtol=RTOL, atol=ATOL)
    with pytest.raises(AssertionError):
        dv.Loss((H, W), 2, 'cuda')(prev.cuda(), nxt.cuda(), fl)           # N > batch_size
    with pytest.raises(AssertionError):
        dv.Loss((H + 1, W), 8, 'cuda')(prev.cuda(), nxt.cuda(), fl)       # wrong height


@pytest.mark.parametrize('hin,win,hout,wout', [(260, 346, 32, 43), (32, 43, 65, 86), (7, 9, 7, 9),
                                               (480, 640, 60, 80), (5, 6, 1, 1)])
def test_interpolate_matches_aten(hin, win, hout, wout):
    g = torch.Generator().manual_seed(hin)
    img = torch.randint(0, 256, (3, 1, hin, win), generator=g).float()
    want = F.interpolate(img, size=(hout, wout), mode='bilinear', align_corners=True)
    got = dv.interpolate(img.cuda(), (hout, wout)).cpu()
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=RTOL, atol=1e-4)


def test_deterministic_bitwise():
    inp = synth_loss_inputs(9, 4, 128, 160, 5.0)
    f0, t0 = _run_gpu(inp, inp['flows'])
    f1, t1 = _run_gpu(inp, inp['flows'])
    assert np.array_equal(_as_array(t0), _as_array(t1))
    _combined(t0).backward()
    _combined(t1).backward()
    for a, b in zip(f0, f1):
        assert torch.equal(a.grad, b.grad)


def test_full_size_properties_cfg4():
    """BASELINE cfg4 size (480x640, batch 16) -- beyond a quick CPU oracle run: size-independent
    properties.  (1) identical frames + constant flow: photometric == smoothness == eps^(2 alpha);
    (2) permuting the batch permutes gradients and keeps every term (means over the batch);
    (3) backward is linear in the cotangent."""
    B, H, W = 16, 480, 640
    inp = synth_loss_inputs(44, B, H, W, 3.0, device='cuda')
    shapes = inp['shapes']
    ev = dv.Losses(shapes, B, 'cuda')
    args = (inp['flow_ts'], inp['flow_sample_idx'], inp['images'], inp['timestamps'], inp['sample_idx'])
    c0 = (1e-3 ** 2) ** 0.45
    same = inp['images'].clone()
    same[1::2] = same[0::2]
    zero = [torch.zeros(B, 2, h, w, device='cuda') for h, w in shapes]
    sm, ph, ob = ev(zero, inp['flow_ts'], inp['flow_sample_idx'], same, inp['timestamps'], inp['sample_idx'])
    np.testing.assert_allclose([float(v) for v in sm], [c0] * 4, rtol=RTOL, atol=ATOL)
    # photometric: the zero-flow grid lands within 1 ulp of the pixel centres (x / h * h != x in
    # fp32, exactly as in the reference), so warped - prev is ~1e-5 grey levels, not 0
    for v in ph:
        assert c0 * (1 - 1e-5) <= float(v) <= c0 * 1.01
    assert all(float(v) == 0 for v in ob)
    flows = [f.requires_grad_() for f in inp['flows']]
    terms = ev(flows, *args)
    _combined(terms).backward()
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(1)).cuda()
    pflows = [f.detach()[perm].contiguous().requires_grad_() for f in flows]
    pterms = ev(pflows, inp['flow_ts'][perm], inp['flow_sample_idx'][perm], inp['images'],
                inp['timestamps'], inp['sample_idx'])
    np.testing.assert_allclose(_as_array(pterms), _as_array(terms), rtol=RTOL, atol=ATOL)
    (3.0 * _combined(pterms)).backward()
    for f, pf in zip(flows, pflows):
        torch.testing.assert_close(pf.grad, 3.0 * f.grad[perm], rtol=RTOL, atol=ATOL)
        assert torch.isfinite(f.grad).all()


# ------------------------------------------------------------------------------------------
# fused training pass (dvsof_multiscale_loss_fused): loss, terms and gradients from ONE sweep
# ------------------------------------------------------------------------------------------
def _run_fused(inp, flows_cpu, weights=WEIGHTS, scale=None):
    shapes = [tuple(f.shape[-2:]) for f in flows_cpu]
    flows = [f.detach().clone().cuda().requires_grad_() for f in flows_cpu]
    ev = dv.Losses(shapes, flows_cpu[0].shape[0], 'cuda')
    loss, flat = ev.combined(flows, inp['flow_ts'].cuda(), inp['flow_sample_idx'].cuda(),
                             inp['images'].cuda(), inp['timestamps'].cuda(), inp['sample_idx'].cuda(),
                             weights)
    (loss if scale is None else loss * scale).backward()
    return flows, loss, flat


@pytest.mark.parametrize('tag', ['a', 'b', 'c'])
def test_fused_matches_reference_golden(golden, tag):
    g = np.load(golden / f'loss_{tag}.npz')
    inp = {k: torch.tensor(g[k]) for k in ('flow_ts', 'flow_sample_idx', 'timestamps', 'sample_idx')}
    inp['images'] = torch.tensor(g['images']).float()
    flows_cpu = [torch.tensor(g[f'flow{k}']) for k in range(4)]
    flows, loss, flat = _run_fused(inp, flows_cpu)
    np.testing.assert_allclose(flat.cpu().numpy().reshape(3, 4), g['terms'], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(float(loss), float(g['total']), rtol=RTOL, atol=ATOL)
    assert not flat.requires_grad
    assert_grads_close([f.grad for f in flows], [g[f'grad{k}'] for k in range(4)], inp, flows_cpu)


@pytest.mark.parametrize('B,H,W,sigma,scale', [
    (1, 260, 346, 3.0, None),     # cfg1 DAVIS: odd widths at every scale
    (8, 256, 256, 3.0, 0.25),     # cfg2 MVSEC, loss / accumulation_steps
    (8, 256, 256, 64.0, None),    # stress: most pixels warp out of the image
    (3, 72, 104, 3.0, 3.0),       # ragged tiles
    (2, 480, 640, 3.0, None),     # cfg4 DSEC shape
    (2, 33, 65, 2.0, None),       # one pixel past a tile edge in both directions
])
def test_fused_matches_oracle_and_generic_path(B, H, W, sigma, scale):
    inp = synth_loss_inputs(300 + B + H, B, H, W, sigma)
    flows, loss, flat = _run_fused(inp, inp['flows'], scale=scale)
    gflows, gterms = _run_gpu(inp, inp['flows'])
    np.testing.assert_allclose(flat.cpu().numpy().reshape(3, 4), _as_array(gterms), rtol=1e-6, atol=1e-7)
    oflows = [f.clone().requires_grad_() for f in inp['flows']]
    oterms = lo.multiscale_loss(oflows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                inp['timestamps'], inp['sample_idx'])
    np.testing.assert_allclose(flat.cpu().numpy().reshape(3, 4), _as_array(oterms), rtol=RTOL, atol=ATOL)
    ototal = lo.combined(oterms)
    np.testing.assert_allclose(float(loss), float(ototal), rtol=RTOL, atol=ATOL)
    k = 1.0 if scale is None else scale
    (ototal * k).backward()
    cot = torch.tensor(WEIGHTS).view(3, 1).expand(3, 4) / 4 * k
    assert_grads_close([f.grad for f in flows], [f.grad.numpy() for f in oflows], inp, inp['flows'], cot)


def test_fused_unequal_weights_and_no_grad_path():
    inp = synth_loss_inputs(11, 2, 64, 96, 6.0)
    w = (0.3, 1.7, 0.9)
    flows, loss, flat = _run_fused(inp, inp['flows'], weights=w)
    oflows = [f.clone().requires_grad_() for f in inp['flows']]
    oterms = lo.multiscale_loss(oflows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                inp['timestamps'], inp['sample_idx'])
    ototal = sum(wk * sum(t) / len(t) for wk, t in zip(w, oterms))
    ototal.backward()
    np.testing.assert_allclose(float(loss), float(ototal), rtol=RTOL, atol=ATOL)
    cot = torch.tensor(w).view(3, 1).expand(3, 4) / 4
    assert_grads_close([f.grad for f in flows], [f.grad.numpy() for f in oflows], inp, inp['flows'], cot)
    # validation: no flow requires grad -> forward-only kernels, same numbers
    ev = dv.Losses(inp['shapes'], 2, 'cuda')
    with torch.no_grad():
        l2, f2 = ev.combined([f.cuda() for f in inp['flows']], inp['flow_ts'].cuda(),
                             inp['flow_sample_idx'].cuda(), inp['images'].cuda(),
                             inp['timestamps'].cuda(), inp['sample_idx'].cuda(), w)
    np.testing.assert_allclose(float(l2), float(loss), rtol=1e-6)
    np.testing.assert_allclose(f2.cpu().numpy(), flat.cpu().numpy(), rtol=1e-6, atol=1e-7)


def test_fused_deterministic_and_single_backward():
    inp = synth_loss_inputs(9, 4, 128, 160, 5.0)
    f0, l0, t0 = _run_fused(inp, inp['flows'])
    f1, l1, t1 = _run_fused(inp, inp['flows'])
    assert torch.equal(l0, l1) and torch.equal(t0, t1)
    for a, b in zip(f0, f1):
        assert torch.equal(a.grad, b.grad)
    flows = [f.detach().clone().cuda().requires_grad_() for f in inp['flows']]
    ev = dv.Losses(inp['shapes'], 4, 'cuda')
    loss, _ = ev.combined(flows, inp['flow_ts'].cuda(), inp['flow_sample_idx'].cuda(),
                          inp['images'].cuda(), inp['timestamps'].cuda(), inp['sample_idx'].cuda())
    loss.backward(retain_graph=True)
    with pytest.raises(RuntimeError):
        loss.backward()


def test_fused_full_size_cfg4_agrees_with_generic_path():
    """BASELINE cfg4 size (480x640, batch 16): the fused sweep against the two-kernel path
    (itself oracle-checked at smaller sizes): same terms, same gradients."""
    B, H, W = 16, 480, 640
    inp = synth_loss_inputs(44, B, H, W, 3.0, device='cuda')
    ev = dv.Losses(inp['shapes'], B, 'cuda')
    args = (inp['flow_ts'], inp['flow_sample_idx'], inp['images'], inp['timestamps'], inp['sample_idx'])
    fa = [f.clone().requires_grad_() for f in inp['flows']]
    fb = [f.clone().requires_grad_() for f in inp['flows']]
    loss, flat = ev.combined(fa, *args)
    loss.backward()
    terms = ev(fb, *args)
    total = _combined(terms)
    total.backward()
    np.testing.assert_allclose(flat.cpu().numpy().reshape(3, 4), _as_array(terms), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(float(loss), float(total), rtol=1e-6)
    for a, b in zip(fa, fb):
        torch.testing.assert_close(a.grad, b.grad, rtol=RTOL, atol=ATOL)


def test_cfg3_shape_batch32_matches_oracle():
    """BASELINE configs[2] loss shape (256x256, batch 32 -- 2048 + 512 + 128 + 32 tiles, 64 frames in
    the one-launch prelude): terms, loss and every gradient against the CPU oracle, through both
    the two-kernel path and the fused training pass."""
    B, H, W = 32, 256, 256
    inp = synth_loss_inputs(1234 + 3, B, H, W, 3.0)
    oterms, ograds = _oracle_grads(inp, inp['flows'])
    want = _as_array(oterms)
    flows, terms = _run_gpu(inp, inp['flows'])
    np.testing.assert_allclose(_as_array(terms), want, rtol=RTOL, atol=ATOL)
    _combined(terms).backward()
    assert_grads_close([f.grad for f in flows], ograds, inp, inp['flows'])
    f2, loss, flat = _run_fused(inp, inp['flows'])
    np.testing.assert_allclose(flat.cpu().numpy().reshape(3, 4), want, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(float(loss), float(lo.combined(oterms)), rtol=RTOL, atol=ATOL)
    assert_grads_close([f.grad for f in f2], ograds, inp, inp['flows'])


def _run_fused_abi(inp):
    import ctypes
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200.loss import _fill_frames
    L = dv._lib
    dev = torch.device('cuda', 0)
    fl = [f.cuda().contiguous() for f in inp['flows']]
    images = inp['images'].cuda()
    B, S = fl[0].shape[0], len(fl)
    D, _, H, W = images.shape
    ev = dv.Losses(inp['shapes'], B, 'cuda')
    plan = ev._plan(dev, B, D, H, W, fl)
    o = plan.offs
    buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
    gbuf = torch.zeros(plan.gtotal, dtype=torch.float32, device=dev)
    dev_in = {k: inp[k].cuda() for k in ('flow_ts', 'flow_sample_idx', 'timestamps', 'sample_idx')}
    fr = _fill_frames(plan, images, dev_in['flow_ts'], dev_in['flow_sample_idx'], dev_in['timestamps'],
                      dev_in['sample_idx'])
    plan.fill(plan.bwd, buf.data_ptr(), fl, gbuf.data_ptr())
    coef = (ctypes.c_float * (3 * S))(*([0.5 / S] * S + [1.0 / S] * S + [1.0 / S] * S))
    st = L.status_for(dev)
    base = buf.data_ptr()
    L.call('dvsof_multiscale_loss_fused', plan.bwd, fr, coef, base + 4 * o[S], base + 4 * o[S + 2],
           base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), L.stream_ptr(dev))
    st.check()
    levels = [buf[o[s]:o[s] + D * h * w].view(D, 1, h, w).clone() for s, (h, w) in enumerate(inp['shapes'])]
    n_bad = buf[o[S + 1]:o[S + 1] + S * B].view(torch.int32).view(S, B).clone()
    return levels, n_bad, images, fl


@pytest.mark.gpu
@pytest.mark.parametrize('B,H,W,n_scales', [(1, 260, 346, 4), (3, 97, 131, 4), (2, 64, 48, 3), (1, 33, 70, 1),
                                            (40, 64, 64, 4), (1, 1024, 1536, 4), (1, 2048, 2304, 4)])
def test_pyramid_chain_and_tile_counts_through_the_abi(B, H, W, n_scales):
    """The one-launch prelude (pyramid_chain_kernel: row slabs, levels in shared memory, per-tile
    border counts) against the level-by-level resize kernel and a direct count, for odd widths
    (scalar stores), a single scale, many frames (one slab per frame) and a frame whose later
    levels do not fit in shared memory (falls back to the resize kernel for them)."""
    import dvs_of_training_framework_b200 as dv
    inp = synth_loss_inputs(31 + B + H, B, H, W, sigma=6.0, n_scales=n_scales)
    levels, n_bad, images, fl = _run_fused_abi(inp)
    img = images
    for s, (h, w) in enumerate(inp['shapes']):
        img = dv.interpolate(img, (h, w))                       # utils/loss.py:207-210, chained
        assert torch.equal(levels[s], img), f'level {s}'
        # out-of-border count of utils/loss.py:92-113 per (scale, prediction)
        f = fl[s]
        xs = torch.arange(w, device='cuda', dtype=torch.float32).view(1, 1, w)
        ys = torch.arange(h, device='cuda', dtype=torch.float32).view(1, h, 1)
        gx = (xs + f[:, 0]) / ((w - 1) / 2) - 1 if w > 1 else torch.full_like(f[:, 0], float('nan'))
        gy = (ys + f[:, 1]) / ((h - 1) / 2) - 1 if h > 1 else torch.full_like(f[:, 1], float('nan'))
        bad = ((gx < -1) | (gx > 1) | (gy < -1) | (gy > 1)).flatten(1).sum(1).int()
        assert torch.equal(n_bad[s], bad), f'border count, scale {s}'


@pytest.mark.gpu
@pytest.mark.parametrize('fused', [False, True])
def test_sweep_reads_match_idx_only_after_the_prelude(fused):
    """The sweep starts under the prelude (programmatic dependent launch) and may touch what the
    prelude writes only after griddepcontrol.wait.  The compiler once hoisted the `match_idx`
    loads of the forward sweep above the wait (invariant loads through a const __restrict__
    pointer): at cfg4 size, where the prelude leaves SMs free for early sweep CTAs, the sweep then
    warped whatever frames the buffer's previous contents named (or faulted on garbage).  Here the
    buffer is poisoned before every call with an index that is valid (no fault) but wrong, and the
    frames are shuffled so that a wrong index changes the photometric term.  Expected values: the
    single-scale kernels (no dependent launch) on explicitly gathered frames."""
    import ctypes
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200.loss import _fill_frames
    L = dv._lib
    dev = torch.device('cuda', 0)
    B, H, W = 16, 480, 640
    inp = synth_loss_inputs(77, B, H, W, 3.0, device='cuda')
    ts = inp['timestamps'].clone()
    for b in range(B):
        ts[2 * b:2 * b + 2] += 0.1 * b
    perm = torch.randperm(2 * B, generator=torch.Generator().manual_seed(5)).cuda()
    inv = torch.argsort(perm)
    images = inp['images'][perm].contiguous()
    fl = [f.contiguous() for f in inp['flows']]
    S, D = len(fl), 2 * B
    want = np.zeros((3, S), dtype=np.float32)
    level = inp['images']
    for s, (h, w) in enumerate(inp['shapes']):
        level = dv.interpolate(level, (h, w))                   # utils/loss.py:207-210, chained
        terms_s = dv.Loss((h, w), B, 'cuda')(level[0::2].contiguous(), level[1::2].contiguous(), fl[s])
        want[:, s] = [float(v) for v in terms_s]
    ev = dv.Losses(inp['shapes'], B, 'cuda')
    plan = ev._plan(dev, B, D, H, W, fl)
    o = plan.offs
    buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
    gbuf = torch.zeros(plan.gtotal, dtype=torch.float32, device=dev)
    flow_ts, tsp, sidx = ts.view(-1, 2).contiguous(), ts[perm].contiguous(), inp['sample_idx'][perm].contiguous()
    fr = _fill_frames(plan, images, flow_ts, inp['flow_sample_idx'], tsp, sidx)
    plan.fill(plan.bwd, buf.data_ptr(), fl, gbuf.data_ptr())
    coef = (ctypes.c_float * (3 * S))(*([0.5 / S] * S + [1.0 / S] * S + [1.0 / S] * S))
    st = L.status_for(dev)
    base = buf.data_ptr()
    match = buf[o[S]:o[S] + 2 * B].view(torch.int32)
    terms = buf[o[S + 2]:o[S + 2] + 3 * S]
    for it in range(40):
        match.fill_(it % D)                                   # stale contents: valid, wrong
        if fused:
            L.call('dvsof_multiscale_loss_fused', plan.bwd, fr, coef, base + 4 * o[S], base + 4 * o[S + 2],
                   base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), L.stream_ptr(dev))
        else:
            L.call('dvsof_multiscale_loss_forward', plan.bwd, fr, base + 4 * o[S], base + 4 * o[S + 2],
                   base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), L.stream_ptr(dev))
        st.check()
        assert torch.equal(match.view(2, B), torch.stack((inv[0::2], inv[1::2])).int()), f'call {it}'
        np.testing.assert_allclose(terms.cpu().numpy().reshape(3, S), want, rtol=RTOL, atol=ATOL,
                                   err_msg=f'call {it}')
