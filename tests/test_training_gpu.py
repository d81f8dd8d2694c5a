"""GPU: the training step the way the reference's callers drive it --
`init_losses` + `process_minibatch` with a DummyNet-contract model on BASELINE configs[0]
(260x346, batch 1) against the golden produced by the reference's own DummyNet / init_losses /
process_minibatch (tests/golden/make_golden_dummynet.py), and `training.train()` end to end with
the package's own fused `Losses` evaluator (accumulation 1 and 2) against the same loop run on
the CPU with the oracle encoder + oracle loss."""
import numpy as np
import pytest
import torch
from torch import nn

import dvs_of_training_framework_b200 as dv
from dvs_of_training_framework_b200 import training as tr
from oracle import encoder_oracle as eo
from oracle import loss_oracle as lo
from tests.helpers import np_events, synth_events, synth_loss_inputs

pytestmark = pytest.mark.gpu


class DummyNet(nn.Module):
    """The contract of DummyNet/net.py:43-80 on a device: zero flows at the four scales
    imsize // 8, 4, 2, 1 (coarse to fine), the (start, stop) timestamp pair and the sample of each
    prediction, and an empty feature tuple when `intermediate`."""

    def __init__(self, device):
        super().__init__()
        self.device = device

    def forward(self, events, timestamps, sample_idx, imsize, raw=True, intermediate=False):
        outsize = [tuple(x // 2 ** i for x in imsize) for i in range(4)][::-1]
        B = int(sample_idx[-1]) + 1
        result = tuple(torch.zeros([B, 2, h, w], dtype=torch.float32, device=self.device) for h, w in outsize)
        assert timestamps.numel() == 2 * B            # sequence length 1: two frames per sample
        return (result, timestamps.view(-1, 2), sample_idx[0::2]) + (((),) if intermediate else ())


def test_dummynet_cfg1_process_minibatch_matches_reference(golden):
    g = np.load(golden / 'dummynet_cfg1.npz')
    dev = torch.device('cuda')
    model = DummyNet(dev)
    H, W = g['images'].shape[-2:]
    evaluator = dv.init_losses((H, W), 1, model, dev, 1)
    assert [(l.H, l.W) for l in evaluator.losses] == [tuple(s) for s in g['shapes']]
    batch = {'events': {k[3:]: torch.tensor(g[k]) for k in g.files if k.startswith('ev_')},
             'timestamps': torch.tensor(g['timestamps']), 'sample_idx': torch.tensor(g['sample_idx']),
             'images': torch.tensor(g['images']).float(), 'augmentation_params': None, 'size': 1}
    loss, terms, tags, info = tr.process_minibatch(model, batch, dv.timer.FakeTimer(), dev, True, evaluator,
                                                   [0.5, 1, 1], return_prediction=True)
    assert list(tags) == list(g['tags'])
    got = np.array([[v for v in group] for group in terms])
    np.testing.assert_allclose(got, g['terms'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(float(loss), float(g['loss']), rtol=1e-5, atol=1e-6)
    assert torch.equal(info['flow_ts'].cpu(), torch.tensor(g['flow_ts']))
    assert torch.equal(info['flow_sample_idx'].cpu(), torch.tensor(g['flow_sample_idx']))
    # the events of the batch through the encoder at this shape (272 x 352), against the oracle
    img = dv.quantize({k: v.to(dev) for k, v in batch['events'].items()}, batch['timestamps'].to(dev),
                      batch['sample_idx'].to(dev), (H, W))
    want = eo.compute_event_image(np_events(batch['events']), g['timestamps'], g['sample_idx'],
                                  dv.extend_size((H, W)))
    assert np.array_equal(img.cpu().numpy(), want)


class TinyFlowNet(nn.Module):
    """Event image -> 4 flows (one 3x3 head per scale on the average-pooled image); no BatchNorm,
    smooth activations: the CPU twin (oracle encoder) and the GPU model (dv.quantize) see the
    same numbers."""

    def __init__(self, device, encode):
        super().__init__()
        self.heads = nn.ModuleList(nn.Conv2d(4, 2, 3, padding=1) for _ in range(4))
        self.encode = encode
        self.to(device)

    def forward(self, events, timestamps, sample_idx, imsize, raw=True, intermediate=False):
        x = self.encode(events, timestamps, sample_idx, imsize)
        flows = []
        for head, k in zip(self.heads, (3, 2, 1, 0)):
            h, w = imsize[0] // 2 ** k, imsize[1] // 2 ** k
            z = nn.functional.avg_pool2d(x, 2 ** k) if k else x
            flows.append(4.0 * torch.tanh(head(z))[..., :h, :w])
        return (tuple(flows), timestamps.reshape(-1, 2), sample_idx[0::2]) + (({},) if intermediate else ())


def _cpu_encode(events, timestamps, sample_idx, imsize):
    if events['x'].numel() == 0:
        return torch.zeros((timestamps.numel() // 2, 4) + dv.extend_size(imsize))
    return torch.from_numpy(eo.compute_event_image(np_events(events), timestamps.numpy(), sample_idx.numpy(),
                                                   dv.extend_size(imsize)))


class OracleEvaluator:
    def __call__(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
        return lo.multiscale_loss(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx)


class Log:
    def __init__(self):
        self.rows = []

    def add_scalar(self, tag, value, step):
        self.rows.append((tag, float(value), step))


def _batches(n, B=2, H=64, W=96):
    out = []
    for i in range(n):
        ev, ts, sidx = synth_events(40 + i, B, H, W, 3000)
        inp = synth_loss_inputs(40 + i, B, H, W, 2.0)
        out.append({'events': ev, 'timestamps': ts, 'sample_idx': sidx, 'images': inp['images'],
                    'augmentation_params': None, 'size': B})
    return out


@pytest.mark.parametrize('accum', [1, 2])
def test_train_end_to_end_with_fused_losses(accum):
    """`training.train` with `dv.init_losses(...)` (the fused one-sweep evaluator whose loss is a
    view of a buffer its autograd node owns) for 2 optimizer steps: runs, and ends with the weights
    the same loop reaches on the CPU with the oracle encoder and the oracle loss."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    H, W = 64, 96
    batches = _batches(2 * accum, H=H, W=W)
    results = []
    for device in ('cpu', 'cuda'):
        torch.manual_seed(0)
        model = TinyFlowNet('cpu', _cpu_encode if device == 'cpu' else dv.quantize)
        model.to(device)
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        evaluator = OracleEvaluator() if device == 'cpu' else dv.init_losses((H, W), 2, model, device, 1)
        log = Log()
        steps = tr.train(model, device, batches, opt, 2, torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0),
                         log, evaluator, accumulation_steps=accum, max_events_per_batch=10 ** 9)
        assert steps == 2
        results.append(([p.detach().cpu() for p in model.parameters()], log.rows))
    (w_cpu, rows_cpu), (w_gpu, rows_gpu) = results
    moved = 0.0
    torch.manual_seed(0)
    init = [p.detach() for p in TinyFlowNet('cpu', _cpu_encode).parameters()]
    for a, b, i in zip(w_cpu, w_gpu, init):
        torch.testing.assert_close(b, a, rtol=1e-4, atol=1e-6)
        moved = max(moved, float((a - i).abs().max()))
    assert moved > 1e-4                                        # the steps did change the weights
    assert [r[0] for r in rows_cpu] == [r[0] for r in rows_gpu]
    np.testing.assert_allclose([r[1] for r in rows_gpu], [r[1] for r in rows_cpu], rtol=2e-5, atol=1e-6)
