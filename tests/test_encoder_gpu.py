"""GPU parity: csrc/encode.cu through the C ABI vs the oracle / the reference's golden outputs.
Bar: bit-exact (integer counts; fp32 timestamps are one IEEE sub + one IEEE div + max)."""
import numpy as np
import pytest
import torch

import dvs_of_training_framework_b200 as dv
from oracle import encoder_oracle as eo
from tests.helpers import np_events, synth_events

pytestmark = pytest.mark.gpu


def _cuda(ev):
    return {k: torch.as_tensor(v).cuda() for k, v in ev.items()}


@pytest.fixture(params=['direct', 'binned'])
def path(request, monkeypatch):
    """Both device paths of the encoder (csrc/encode.cu direct REDs, csrc/encode_binned.cuh two-pass
    shared-memory slabs); the library reads DVSOF_ENCODE_PATH at every call."""
    monkeypatch.setenv('DVSOF_ENCODE_PATH', request.param)
    return request.param


@pytest.mark.parametrize('name', ['fixture', 'dense', 'ragged'])
def test_matches_reference_golden(golden, name, path):
    g = np.load(golden / f'encoder_{name}.npz')
    ev = _cuda({k[3:]: g[k] for k in g.files if k.startswith('ev_')})
    out = dv.compute_event_image(ev, torch.tensor(g['timestamps']).cuda(),
                                 torch.tensor(g['sample_idx']).cuda(), (int(g['H']), int(g['W'])))
    assert out.dtype == torch.float32 and out.is_cuda
    assert np.array_equal(out.cpu().numpy(), g['out'])


@pytest.mark.parametrize('B,H,W,n,contended', [
    (1, 272, 352, 32768, False),      # cfg1 DAVIS 260x346 -> 272x352
    (8, 256, 256, 131072, False),     # cfg2 MVSEC
    (8, 256, 256, 131072, True),      # contended pixels
    (3, 48, 80, 1001, False),         # n % 4 != 0 tail
    (2, 480, 640, 200000, False),     # DSEC shape
])
def test_matches_oracle_synthetic(B, H, W, n, contended, path):
    ev, ts, sidx = synth_events(7 + B, B, H, W, n, contended=contended)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    got = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert np.array_equal(got.cpu().numpy(), want)


def test_unaligned_views_take_scalar_path(path):
    ev, ts, sidx = synth_events(3, 2, 64, 64, 5000)
    want = eo.compute_event_image(np_events({k: v[1:] for k, v in ev.items()}), ts.numpy(),
                                  sidx.numpy(), (64, 64))
    cu = {k: v.cuda()[1:] for k, v in ev.items()}     # 8-byte / 4-byte offset views
    got = dv.compute_event_image(cu, ts.cuda(), sidx.cuda(), (64, 64))
    assert np.array_equal(got.cpu().numpy(), want)


def test_ragged_and_empty_samples(path):
    ev, ts, sidx = synth_events(11, 6, 32, 48, 4000, ragged=True)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (32, 48))
    got = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (32, 48))
    assert np.array_equal(got.cpu().numpy(), want)


def test_empty_events_give_zeros():
    # EV_FlowNet/net.py:115-116 (init_losses relies on it)
    ev = {k: torch.zeros(0, dtype=torch.long).cuda() for k in ('x', 'y', 'polarity', 'sample_index')}
    ev['timestamp'] = torch.zeros(0).cuda()
    out = dv.compute_event_image(ev, torch.tensor([0., .04]).cuda(), torch.tensor([0, 0]).cuda(), (32, 32))
    assert out.shape == (1, 4, 32, 32) and not out.any()


def test_events_before_start_and_late_events(path):
    # t < start -> negative t^ loses against 0; t > last -> t^ > 1 kept (dt uses the LAST event)
    ev = {'x': [1, 1, 2, 2], 'y': [0, 0, 0, 0], 'polarity': [1, 1, -1, -1],
          'sample_index': [0, 0, 0, 0], 'timestamp': np.array([0.5, 3.0, 0.2, 1.0], np.float32)}
    ev = {k: np.asarray(v) for k, v in ev.items()}
    ts = np.array([1.0, 2.0], np.float32)
    want = eo.compute_event_image(ev, ts, np.array([0, 0]), (16, 16))
    got = dv.compute_event_image(_cuda(ev), torch.tensor(ts).cuda(), torch.tensor([0, 0]).cuda(), (16, 16))
    assert np.array_equal(got.cpu().numpy(), want)
    assert want[0, 3, 0, 2] == 0 and want[0, 2, 0, 1] > 1


def test_quantize_extends_to_multiple_of_16():
    ev, ts, sidx = synth_events(5, 1, 260, 346, 2000)
    out = dv.quantize(_cuda(ev), ts.cuda(), sidx.cuda(), (260, 346))
    assert out.shape == (1, 4, 272, 352)
    assert np.array_equal(out.cpu().numpy(),
                          eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (272, 352)))


@pytest.mark.parametrize('field,value', [('polarity', 0), ('sample_index', -1), ('x', 9999)])
def test_invalid_input_raises_assertion(field, value, path):
    ev, ts, sidx = synth_events(1, 2, 32, 32, 100)
    ev[field][150] = value
    with pytest.raises(AssertionError):
        dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (32, 32))


def test_ungrouped_samples_raise(path):
    ev, ts, sidx = synth_events(1, 2, 32, 32, 100)
    ev['sample_index'] = ev['sample_index'].flip(0)
    with pytest.raises(AssertionError):
        dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (32, 32))


def test_compact_storage_form_equals_decoded_form():
    B, H, W = 4, 64, 96
    ev, ts, sidx = synth_events(21, B, H, W, 7001, ragged=True)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    eps = torch.bincount(ev['sample_index'], minlength=B)
    got = dv.compute_event_image_compact(ev['x'].to(torch.int16).cuda(), ev['y'].to(torch.int16).cuda(),
                                         ev['timestamp'].cuda(), (ev['polarity'] > 0).cuda(),
                                         eps.cuda(), ts.cuda(), (H, W))
    assert np.array_equal(got.cpu().numpy(), want)


@pytest.mark.parametrize('H,W', [(64, 96), (272, 640)])
def test_compact_storage_form_binned_equals_direct_and_oracle(monkeypatch, H, W):
    """The storage form on both device paths (a dense, ragged stream with an empty sample: the
    library picks the binned path by itself once the stream is large; here it is forced both ways)
    against the oracle on the decoded events, and against the int64 entry point."""
    B, n = 3, 400_000
    ev, ts, sidx = synth_events(23 + W, B, H, W, n, ragged=True, contended=True)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    eps = torch.bincount(ev['sample_index'], minlength=B)
    args = (ev['x'].to(torch.int16).cuda(), ev['y'].to(torch.int16).cuda(), ev['timestamp'].cuda(),
            (ev['polarity'] > 0).cuda(), eps.cuda(), ts.cuda(), (H, W))
    outs = {}
    for p in ('direct', 'binned'):
        monkeypatch.setenv('DVSOF_ENCODE_PATH', p)
        outs[p] = dv.compute_event_image_compact(*args)
    assert torch.equal(outs['direct'], outs['binned'])
    assert np.array_equal(outs['binned'].cpu().numpy(), want)
    monkeypatch.delenv('DVSOF_ENCODE_PATH')
    big, tsb, sidxb = synth_events(5, 2, 64, 64, 2_200_000)                # >= 4 M events: default = binned
    got = dv.compute_event_image_compact(big['x'].to(torch.int16).cuda(), big['y'].to(torch.int16).cuda(),
                                         big['timestamp'].cuda(), (big['polarity'] > 0).cuda(),
                                         torch.bincount(big['sample_index'], minlength=2).cuda(), tsb.cuda(), (64, 64))
    assert torch.equal(got, dv.compute_event_image(_cuda(big), tsb.cuda(), sidxb.cuda(), (64, 64)))


@pytest.mark.parametrize('case', ['offset', 'dense', 'shuffled', 'mixed', 'inf'])
def test_binned_record_formats_bitwise(monkeypatch, case):
    """Timestamp layouts on the binned path, whose pass 2 keeps the maximum of the RAW t under an
    order-preserving integer key: a large time offset (coarse fp32 steps, many equal keys), a dense
    burst, timestamps in random order, a jump in step size mid-stream, infinite timestamps.  Every
    one bit-identical to the direct path and the oracle.  (Written with the packed 4-byte record
    experiment of profiles/r02zf_packed_records_negative.*: it passed these and measured slower.)"""
    B, H, W, n = 3, 64, 96, 300_000
    ev, ts, sidx = synth_events(91, B, H, W, n, ragged=(case == 'mixed'))
    if case == 'offset':
        ev['timestamp'] = (ev['timestamp'] + 100.0).float()
        ts = (ts + 100.0).float()
    elif case == 'dense':
        ev['timestamp'] = (ev['timestamp'] * 0.01 + 0.02).float()       # 300 k events in 0.32 ms
    elif case == 'shuffled':
        perm = torch.cat([torch.randperm(n) + b * n for b in range(B)])
        ev['timestamp'] = ev['timestamp'][perm].contiguous()
    elif case == 'mixed':                                                 # the first half coarse, the rest fine steps
        m = ev['timestamp'].numel() // 2
        ev['timestamp'][:m] = (ev['timestamp'][:m] + 8.0).float() - 8.0
    elif case == 'inf':
        ev['timestamp'][1000] = float('inf')
        ev['timestamp'][5000] = float('-inf')
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    outs = {}
    for p in ('direct', 'binned'):
        monkeypatch.setenv('DVSOF_ENCODE_PATH', p)
        outs[p] = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert torch.equal(outs['direct'], outs['binned'])
    assert np.array_equal(outs['binned'].cpu().numpy(), want, equal_nan=True)


def test_full_size_properties_cfg3_like():
    """At sizes the oracle cannot finish quickly: size-independent checks (count conservation,
    linearity over a split of the stream, max of maxima)."""
    B, H, W, n = 4, 256, 256, 2_000_000
    ev, ts, sidx = synth_events(33, B, H, W, n)
    cu = _cuda(ev)
    full = dv.compute_event_image(cu, ts.cuda(), sidx.cuda(), (H, W))
    assert int(full[:, :2].sum()) == B * n
    pos = int((ev['polarity'] > 0).sum())
    assert int(full[:, 0].sum()) == pos
    assert float(full[:, 2:].max()) <= 1.0 and float(full[:, 2:].min()) >= 0.0
    # per-sample last event has t^ == 1 exactly (dt = t_last - start)
    last = torch.cumsum(torch.bincount(ev['sample_index']), 0) - 1
    for b in range(B):
        i = int(last[b])
        c = 2 + (0 if ev['polarity'][i] > 0 else 1)
        assert float(full[b, c, int(ev['y'][i]), int(ev['x'][i])]) == 1.0
    # counts are invariant under any permutation inside a sample
    perm = torch.cat([torch.randperm(n) + b * n for b in range(B)])
    sh = {k: v[perm].contiguous() for k, v in ev.items()}
    cnt2 = dv.compute_event_image(_cuda(sh), ts.cuda(), sidx.cuda(), (H, W))[:, :2]
    assert torch.equal(cnt2, full[:, :2])


def test_binned_equals_direct_bitwise_dense(monkeypatch):
    """The two device paths on a stream dense enough for real slab contention (40 events per
    cell) and with negative / late timestamps: bit-identical images."""
    B, H, W, n = 3, 80, 112, 700_000
    ev, ts, sidx = synth_events(55, B, H, W, n, contended=True)
    ev['timestamp'] = ev['timestamp'] - 0.004            # some events before `start` -> clamp at 0
    ts = ts.clone()
    ts[0::2] = 0.001
    outs = {}
    for p in ('direct', 'binned'):
        monkeypatch.setenv('DVSOF_ENCODE_PATH', p)
        outs[p] = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert torch.equal(outs['direct'], outs['binned'])
    assert int(outs['binned'][:, :2].sum()) == B * n and float(outs['binned'][:, 2:].min()) == 0.0
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    assert np.array_equal(outs['binned'].cpu().numpy(), want)


@pytest.mark.parametrize('H,W', [(272, 640), (256, 1024), (480, 640)])
def test_binned_wide_images_bitwise(monkeypatch, H, W):
    """Wide images: more than 127 two-row bands -> the plan takes 4 rows per slab (40-64 KB of
    shared memory, 13-bit cell codes) and pass 2 its short-run variant (one record per lane and
    batch); (256, 1024) fills the 64 KB / 8192-cell limit exactly."""
    B, n = 2, 150_000
    ev, ts, sidx = synth_events(91 + W, B, H, W, n, contended=True)
    outs = {}
    for p in ('direct', 'binned'):
        monkeypatch.setenv('DVSOF_ENCODE_PATH', p)
        outs[p] = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert torch.equal(outs['direct'], outs['binned'])
    assert int(outs['binned'][:, :2].sum()) == B * n
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    assert np.array_equal(outs['binned'].cpu().numpy(), want)


def test_default_path_selection_large_stream_is_exact():
    """>= 4 M events and >= 8 events per cell -> the library picks the binned path by itself."""
    B, H, W, n = 2, 64, 64, 2_200_000
    ev, ts, sidx = synth_events(77, B, H, W, n)
    got = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    assert np.array_equal(got.cpu().numpy(), want)


def test_cfg3_full_shape_default_path_binned_equals_direct_and_oracle(monkeypatch):
    """BASELINE configs[2] at full size: batch 32, 256x256, 5 M events per sample (160 M events,
    ~39 k pass-1 chunks, 32 x 32 pass-2 slabs).  The library picks the binned path by itself here;
    it must be bit-identical to the direct path, conserve the event count per sample and polarity,
    and equal the oracle on a 1-in-32 subset of the samples (each sample's image depends on its
    own events only, so one sample can be checked alone)."""
    B, H, W, n = 32, 256, 256, 5_000_000
    dev = torch.device('cuda')
    g = torch.Generator(device=dev).manual_seed(1234 + 3)
    N = B * n
    ev = {'x': torch.randint(0, W, (N,), generator=g, device=dev),
          'y': torch.randint(0, H, (N,), generator=g, device=dev),
          'polarity': (torch.rand(N, generator=g, device=dev) < 0.7).long() * 2 - 1,
          'sample_index': torch.arange(B, device=dev).repeat_interleave(n)}
    ev['timestamp'] = torch.sort(torch.rand(B, n, generator=g, device=dev) * 0.032, dim=1)[0].reshape(-1).float()
    ts = torch.tensor([0, 0.032] * B, dtype=torch.float32, device=dev)
    sidx = torch.arange(B, device=dev).repeat_interleave(2)
    monkeypatch.delenv('DVSOF_ENCODE_PATH', raising=False)
    default = dv.compute_event_image(ev, ts, sidx, (H, W))
    monkeypatch.setenv('DVSOF_ENCODE_PATH', 'binned')
    binned = dv.compute_event_image(ev, ts, sidx, (H, W))
    monkeypatch.setenv('DVSOF_ENCODE_PATH', 'direct')
    direct = dv.compute_event_image(ev, ts, sidx, (H, W))
    assert torch.equal(direct, binned) and torch.equal(default, binned)
    per_sample = default[:, :2].sum(dim=(2, 3), dtype=torch.float64)
    pos = (ev['polarity'] > 0).view(B, n).sum(1).double()
    assert torch.equal(per_sample[:, 0], pos) and torch.equal(per_sample[:, 1], n - pos)
    assert float(default[:, 2:].max()) == 1.0 and float(default[:, 2:].min()) >= 0.0
    for b in (7,):                                    # 1 in 32: the oracle on that sample alone
        one = {k: v[b * n:(b + 1) * n].cpu().numpy() for k, v in ev.items()}
        one['sample_index'] = one['sample_index'] - b
        want = eo.compute_event_image(one, ts[2 * b:2 * b + 2].cpu().numpy(), np.zeros(2, np.int64), (H, W))
        assert np.array_equal(default[b].cpu().numpy(), want[0])


def test_deferred_validation_reports_without_blocking():
    """Status.mode 'deferred': the call returns without a host sync; the error surfaces at
    `synchronize()` (or at a later call once the async copy of the status word has landed)."""
    from dvs_of_training_framework_b200 import _lib
    ev, ts, sidx = synth_events(1, 2, 32, 32, 100)
    ev['polarity'][7] = 0
    st = _lib.status_for(torch.device('cuda', torch.cuda.current_device()))
    old = _lib.Status.mode
    _lib.Status.mode = 'deferred'
    try:
        dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (32, 32))     # no raise here
        with pytest.raises(AssertionError, match='polarity'):
            st.synchronize()
        good, ts2, sidx2 = synth_events(2, 2, 32, 32, 100)
        dv.compute_event_image(_cuda(good), ts2.cuda(), sidx2.cuda(), (32, 32))
        st.synchronize()                                                         # clean again
    finally:
        _lib.Status.mode = old


@pytest.mark.parametrize('B', [512, 513])
def test_many_samples_at_the_binned_path_limit(monkeypatch, B):
    """The binned path looks a chunk's sample up in a per-CTA table of at most 512 samples
    (csrc/encode_binned.cuh kBinMaxSamples); one more and the call must take the direct path even
    when the binned one is asked for -- same image either way."""
    H, W, n = 16, 32, 300
    ev, ts, sidx = synth_events(77, B, H, W, n, ragged=True)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    lib = dv._lib.load()
    N = ev['x'].numel()
    monkeypatch.setenv('DVSOF_ENCODE_PATH', 'binned')
    assert lib.dvsof_encode_path(N, B, H, W, lib.dvsof_encode_workspace_bytes(N, B, H, W)) == (1 if B <= 512 else 0)
    got = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert np.array_equal(got.cpu().numpy(), want)


def test_narrow_image_binned_path(monkeypatch):
    """Slabs with fewer than 32 cells per polarity row (W = 8): lanes without a record park their
    no-op atomics on cell `lane` of the slab (csrc/encode_binned.cuh slab_apply)."""
    B, H, W, n = 2, 40, 8, 9000
    ev, ts, sidx = synth_events(5, B, H, W, n)
    want = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (H, W))
    monkeypatch.setenv('DVSOF_ENCODE_PATH', 'binned')
    got = dv.compute_event_image(_cuda(ev), ts.cuda(), sidx.cuda(), (H, W))
    assert np.array_equal(got.cpu().numpy(), want)
