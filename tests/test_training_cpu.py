"""CPU (no GPU): host logic of the data-parallel step -- batch sharding
by sample, flat-bucket gradient averaging over `gloo` at world size 2, and the `train` loop
contract of utils/training.py:89-235.  The CUDA loss has no CPU path, so the loss used here is
the torch oracle (test infrastructure) behind the same evaluator interface."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

from dvs_of_training_framework_b200 import training as tr
from oracle import loss_oracle as lo
from tests.helpers import synth_events, synth_loss_inputs


def _batch(B=4, H=32, W=32, n=500, seed=3):
    ev, ts, sidx = synth_events(seed, B, H, W, n, ragged=True)
    inp = synth_loss_inputs(seed, B, H, W, 2.0)
    return {'events': ev, 'timestamps': ts, 'sample_idx': sidx, 'images': inp['images'],
            'augmentation_params': {'idx': torch.arange(B)}, 'size': B}


def test_shard_bounds_cover_and_balance():
    for B in (1, 5, 8, 256):
        for world in (1, 2, 3, 8):
            edges = [tr.shard_bounds(B, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize('world', [2, 3, 4])
def test_shard_batch_partitions_by_sample(world):
    batch = _batch()
    shards = [tr.shard_batch(batch, r, world) for r in range(world)]
    assert sum(s['size'] for s in shards) == batch['size']
    for k in ('x', 'y', 'timestamp', 'polarity'):
        assert torch.equal(torch.cat([s['events'][k] for s in shards]), batch['events'][k])
    assert torch.equal(torch.cat([s['images'] for s in shards]), batch['images'])
    lo_ = 0
    for s in shards:
        b = s['events']['sample_index']
        if b.numel():
            assert int(b.min()) >= 0 and int(b.max()) < s['size']       # rebased to the shard
        assert torch.equal(s['sample_idx'], torch.arange(s['size']).repeat_interleave(2))
        assert s['timestamps'].numel() == 2 * s['size']
        assert torch.equal(s['augmentation_params']['idx'], torch.arange(lo_, lo_ + s['size']))
        # the encoder's contract (net.py:101): (sample_idx[-1] + 1) * 2 == len(timestamps)
        assert (int(s['sample_idx'][-1]) + 1) * 2 == s['timestamps'].numel()
        lo_ += s['size']
    assert tr.shard_batch(batch, 0, 1) is batch


class _ViewLoss(torch.autograd.Function):
    """Like loss._FusedTrainingLoss: the loss is a VIEW of a buffer the node owns (an in-place
    `loss /= k` on it is what torch refuses), the terms are non-differentiable."""
    @staticmethod
    def forward(ctx, weights, args, *flows):
        with torch.enable_grad():
            fl = [f.detach().requires_grad_() for f in flows]
            terms = lo.multiscale_loss(fl, *args)
            loss = lo.combined(terms, weights)
            ctx.grads = torch.autograd.grad(loss, fl)
        buf = torch.cat([loss.detach().reshape(1), torch.stack([t.detach() for g in terms for t in g])])
        flat = buf[1:]
        ctx.mark_non_differentiable(flat)
        return buf[0:1].view(()), flat

    @staticmethod
    def backward(ctx, g, _):
        return (None, None) + tuple(g * x for x in ctx.grads)

class Fused:                     # the (loss, flat_terms) interface of Losses.combined
    def __call__(self, flows, *a):
        return lo.multiscale_loss(flows, *a)

    def combined(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx, weights):
        return _ViewLoss.apply(tuple(weights), (flow_ts, flow_sample_idx, images, timestamps, sample_idx),
                               *flows)


def test_combined_loss_generic_and_fused_paths_agree():
    inp = synth_loss_inputs(5, 2, 32, 48, 2.0)

    class Generic:
        def __call__(self, flows, *a):
            return lo.multiscale_loss(flows, *a)

    args = (inp['flows'], inp['flow_ts'], inp['flow_sample_idx'], inp['images'], inp['timestamps'],
            inp['sample_idx'], None)
    l0, t0 = tr.combined_loss(Generic(), *args, weights=[0.5, 1, 1])
    l1, t1 = tr.combined_loss(Fused(), *args, weights=[0.5, 1, 1])
    assert abs(float(l0) - float(l1)) < 1e-6 * max(1.0, abs(float(l0)))
    assert len(t1) == 3 and all(len(g) == 4 for g in t1)
    for a, b in zip(t0, t1):
        for u, v in zip(a, b):
            assert float(u) == float(v)
    assert float(l0) == pytest.approx(float(lo.combined(t0)), rel=1e-6)
    # the fused branch returns (loss, 3*S terms) and its loss survives the accumulation division
    # and backward the way `train` applies them (utils/training.py:155-158)
    fl = [f.clone().requires_grad_() for f in inp['flows']]
    l2, t2 = tr.combined_loss(Fused(), fl, *args[1:], weights=[0.5, 1, 1])
    assert l2.dim() == 0 and l2._is_view()
    with pytest.raises(RuntimeError):
        l2 /= 2
    (l2 / 2).backward()
    fr = [f.clone().requires_grad_() for f in inp['flows']]
    l3, _ = tr.combined_loss(Generic(), fr, *args[1:], weights=[0.5, 1, 1])
    (l3 / 2).backward()
    for a, b in zip(fl, fr):
        torch.testing.assert_close(a.grad, b.grad, rtol=1e-6, atol=1e-8)


class TinyFlowNet(nn.Module):
    """4-scale flow predictor with the Model.forward contract (no BatchNorm: replicas == one
    big batch exactly)."""

    def __init__(self):
        super().__init__()
        self.heads = nn.ModuleList(nn.Conv2d(1, 2, 3, padding=1) for _ in range(4))
        self.unused = nn.Parameter(torch.zeros(3))        # never gets a gradient

    def forward(self, events, timestamps, sample_idx, imsize, raw=True, intermediate=False):
        B = timestamps.numel() // 2
        H, W = imsize
        x = torch.zeros(B, 1, H, W)
        if events['x'].numel():
            x.view(B, -1).index_put_((events['sample_index'], events['y'] * W + events['x']),
                                     torch.ones(events['x'].numel()), accumulate=True)
        flows = tuple(h(nn.functional.interpolate(x, size=(H // 2 ** k, W // 2 ** k)))
                      for h, k in zip(self.heads, (3, 2, 1, 0)))
        return flows, timestamps.reshape(-1, 2), sample_idx[0::2], {}


class OracleEvaluator:
    def __call__(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
        return lo.multiscale_loss(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx)


class Log:
    def __init__(self):
        self.rows = []

    def add_scalar(self, tag, value, step):
        self.rows.append((tag, float(value), step))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _reference_grads(batch, steps=1):
    torch.manual_seed(0)
    model = TinyFlowNet()
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    tr.train(model, 'cpu', [batch] * steps, opt, steps, torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0),
             Log(), OracleEvaluator(), max_events_per_batch=10 ** 9)
    return [p.detach().clone() for p in model.parameters()]


def _worker(rank, world, port, batch, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        torch.manual_seed(rank)                              # different init: broadcast must fix it
        model = TinyFlowNet()
        red = tr.GradientAllreducer(model.parameters(), bucket_mb=1e-4)   # many small buckets
        red.broadcast_parameters(model)
        opt = torch.optim.SGD(model.parameters(), lr=0.1)
        log = Log()
        steps = tr.train(model, 'cpu', [batch, batch], opt, 2,
                         torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0), log, OracleEvaluator(),
                         max_events_per_batch=10 ** 9, reducer=red)
        q.put((rank, steps, len(red.buckets), [p.detach().numpy().copy() for p in model.parameters()],
               len(log.rows)))
    finally:
        dist.destroy_process_group()


def test_data_parallel_train_equals_single_process_gloo():
    """world 2 over gloo: after 2 optimizer steps both replicas hold the SAME weights, equal to
    the single-process run on the whole batch (equal shards, no BatchNorm => mean of shard
    gradients == full-batch gradient)."""
    torch.manual_seed(0)
    init = TinyFlowNet()
    batch = _batch(B=4, n=300)
    # equal-size shards are required for mean-of-means == global mean
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1] == 2 and res[0][2] > 1
    for a, b in zip(res[0][3], res[1][3]):
        assert (a == b).all()                                 # replicas in lock-step
    want = _reference_grads(batch, steps=2)
    for a, w in zip(res[0][3], want):
        torch.testing.assert_close(torch.from_numpy(a), w, rtol=2e-4, atol=1e-6)
    assert res[0][4] > 0 and res[1][4] == 0                   # only rank 0 logs
    del init


def test_train_loop_contract_single_process():
    """utils/training.py:89-235: skip of over-large batches, accumulation, hooks, logging tags."""
    torch.manual_seed(0)
    model = TinyFlowNet()
    opt = torch.optim.SGD(model.parameters(), lr=0.01)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0)
    log, calls = Log(), []
    batch = _batch(B=2, n=200)
    big = _batch(B=2, n=5000, seed=9)
    steps = tr.train(model, 'cpu', [batch, big, batch, batch, batch], opt, 2, sched, log,
                     OracleEvaluator(), accumulation_steps=2, hooks={'h': lambda s, n: calls.append((s, n))},
                     max_events_per_batch=3000)
    assert steps == 2 and calls == [(1, 4), (2, 8)]
    tags = {t for t, _, _ in log.rows}
    assert 'General/Train loss' in tags and 'Train/photometric loss/32x32' in tags
    assert 'Train/smoothness loss/4x4' in tags and 'General/learning rate/0' in tags
    v = tr.validate(model, 'cpu', [batch], 8, log, OracleEvaluator())
    assert v > 0 and any(t == 'General/Validation loss' for t, _, _ in log.rows)


def test_prefetching_loader_is_transparent():
    """train(..., prefetch=True) (DevicePrefetcher: shard + upload one batch ahead) takes the same
    steps, hooks and weights as the plain loop, including the over-large-batch skip and the
    sample counter, which must keep seeing the GLOBAL batch."""
    def run(prefetch):
        torch.manual_seed(0)
        model = TinyFlowNet()
        opt = torch.optim.SGD(model.parameters(), lr=0.01)
        sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0)
        calls = []
        batch, big = _batch(B=2, n=200), _batch(B=2, n=5000, seed=9)
        steps = tr.train(model, 'cpu', [batch, big, batch, batch, batch], opt, 2, sched, Log(),
                         OracleEvaluator(), accumulation_steps=2,
                         hooks={'h': lambda s, n: calls.append((s, n))}, max_events_per_batch=3000,
                         prefetch=prefetch)
        return steps, calls, [p.detach().clone() for p in model.parameters()]
    s0, c0, w0 = run(False)
    s1, c1, w1 = run(True)
    assert s0 == s1 == 2 and c0 == c1 == [(1, 4), (2, 8)]
    for a, b in zip(w0, w1):
        assert torch.equal(a, b)
    # the wrapper alone: shards by sample, keeps order, reports the global size
    batch = _batch(B=4, n=100)
    got = list(tr.DevicePrefetcher([batch, batch], 'cpu', True, rank=1, world=2))
    assert len(got) == 2 and got[0]['size'] == 2 and got[0]['global_size'] == 4
    assert got[0]['global_events'] == batch['events']['x'].numel()
    assert int(got[0]['events']['sample_index'].min()) == 0 and got[0]['images'].shape[0] == 4


@pytest.mark.parametrize('accum', [1, 2])
def test_train_with_fused_evaluator_and_wall_clock_timers(accum, capsys):
    """`train` end to end with an evaluator of the `Losses.combined` kind (loss = view of a buffer
    its autograd node owns) and the `--timers` group: same weights as with the generic evaluator,
    one timer line per micro-batch with the phases of utils/training.py:222-233."""
    from dvs_of_training_framework_b200.timer import SynchronizedWallClockTimer
    batch = _batch(B=2, n=200)
    out = []
    for ev in (OracleEvaluator(), Fused()):
        torch.manual_seed(0)
        model = TinyFlowNet()
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        steps = tr.train(model, 'cpu', [batch] * (2 * accum), opt, 2,
                         torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0), Log(), ev,
                         accumulation_steps=accum, timers=SynchronizedWallClockTimer(),
                         max_events_per_batch=10 ** 9)
        assert steps == 2
        out.append([p.detach().clone() for p in model.parameters()])
    for a, b in zip(*out):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-7)
    lines = [ln for ln in capsys.readouterr().out.splitlines() if ln.startswith('rank=0 time (ms)')]
    assert len(lines) == 2 * 2 * accum
    for name in ('batch_construction', 'batch2gpu', 'forward', 'loss', 'backprop', 'optimizer_step', 'logging'):
        assert f'{name}: ' in lines[0]


def test_evflownet_checkpoint_keys_follow_the_reference():
    """A `step_N.pt`['model'] written by the reference's EV_FlowNet/net.py Model loads into the
    stand-in network and vice versa (names: enc.N.{conv,bn}, tr.N.conv.M.{conv,bn},
    dec.N.conv.{conv,bn}, flow.N -- net.py:20-86,153-190)."""
    from dvs_of_training_framework_b200.evflownet import Model
    torch.manual_seed(1)
    a, b = Model('cpu'), Model('cpu')
    ref = a.reference_state_dict()
    assert len(ref) == 92
    for k in ('enc.0.conv.weight', 'enc.3.bn.running_mean', 'tr.1.conv.1.bn.running_var',
              'tr.0.conv.0.conv.bias', 'dec.3.conv.conv.bias', 'dec.0.conv.bn.weight', 'flow.2.weight'):
        assert k in ref, k
    assert not any('.block.' in k or '.body.' in k for k in ref)
    res = b.load_state_dict({k: v.clone() for k, v in ref.items()})      # reference names -> strict load
    assert not res.missing_keys and not res.unexpected_keys
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)
