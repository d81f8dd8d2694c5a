"""EV-FlowNet stand-in for the data-parallel train-step measurement (SURVEY.md section 8 row
A16 / (e), BASELINE.json configs[1] and configs[4]).

The network itself is OUT of the hot-path scope: its convolutions stay on cuDNN, as the
north_star says.  What matters here is the call site the reference's `Model.forward` has
(EV_FlowNet/net.py:207-264): `forward(events, timestamps, sample_idx, imsize, raw, intermediate)`
starts with `quantize` -> the CUDA event-image encoder of this package, and returns
`(flows coarse->fine, flow_ts (P,2), flow_sample_idx (P), [features])`, which is what
`training.process_minibatch` and `loss.Losses` consume.  Same layer plan as the reference
(4 stride-2 encoders 64..512, 2 residual blocks, 4 nearest-upsample decoders with a 1x1 flow
head `tanh * 256` each; 14,136,680 parameters) so the step cost is representative; weights are
random (there are no checkpoints offline).  A user of the reference keeps the reference's own
`EV_FlowNet/net.py` and only swaps `compute_event_image`."""
import torch
import torch.nn.functional as F
from torch import nn

from . import _lib
from .encoder import quantize

_WIDTHS = (64, 128, 256, 512)


class _ReluBN(torch.autograd.Function):
    """`bn(relu(z + conv_bias))` in training mode (net.py:36,54) through csrc/glue.cu: batch
    statistics and the running-statistics update like torch.nn.BatchNorm2d, the bias add and the
    ReLU folded into every pass (the convolution itself then runs without a bias: no separate add,
    no separate reduction for its gradient)."""

    @staticmethod
    def forward(ctx, z, conv_bias, weight, bias, running_mean, running_var, momentum, eps):
        z, cl = _layout(z)
        N, C = z.shape[0], z.shape[1]
        if cl and (C % 4 or C > 1024):
            z, cl = z.contiguous(), 0
        HW = z.numel() // max(N * C, 1)
        dev = z.device
        y = torch.empty_like(z)
        stats = torch.empty((2, C), dtype=torch.float32, device=dev)
        ws = torch.empty(_lib.load().dvsof_relu_bn_workspace_bytes(N, C, HW, cl), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.call('dvsof_relu_bn_forward', _lib.ptr(z), _lib.ptr(conv_bias), N, C, HW, _lib.ptr(weight),
                      _lib.ptr(bias), _lib.ptr(running_mean), _lib.ptr(running_var), float(momentum), float(eps),
                      _lib.ptr(y), _lib.ptr(stats[0]), _lib.ptr(stats[1]), _lib.ptr(ws), cl, _lib.stream_ptr(dev))
        ctx.save_for_backward(z, conv_bias, weight, stats)
        ctx.ws, ctx.cl = ws, cl
        return y

    @staticmethod
    def backward(ctx, gy):
        z, conv_bias, weight, stats = ctx.saved_tensors
        cl = ctx.cl
        gy = gy.contiguous(memory_format=torch.channels_last) if cl else gy.contiguous()
        N, C = z.shape[0], z.shape[1]
        HW = z.numel() // max(N * C, 1)
        dev = z.device
        gz = torch.empty_like(z)
        g3 = torch.empty((3, C), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.call('dvsof_relu_bn_backward', _lib.ptr(z), _lib.ptr(conv_bias), _lib.ptr(gy), N, C, HW,
                      _lib.ptr(weight), _lib.ptr(stats[0]), _lib.ptr(stats[1]), _lib.ptr(gz), _lib.ptr(g3[0]),
                      _lib.ptr(g3[1]), _lib.ptr(g3[2]) if conv_bias is not None else None, _lib.ptr(ctx.ws), cl,
                      _lib.stream_ptr(dev))
        return gz, (g3[2] if conv_bias is not None else None), g3[0], g3[1], None, None, None, None


def _fused_bn_ok(z, bn):
    return (bn.training and z.is_cuda and z.dtype == torch.float32 and bn.track_running_stats
            and bn.affine and bn.momentum is not None and z.numel() > 0)


def relu_batchnorm(z, bn, conv_bias=None):
    """bn(relu(z + conv_bias)).  Training on CUDA fp32 takes the fused kernels; evaluation mode and
    CPU tensors (the world-size-2 tests of the host logic) the ATen ops."""
    if _fused_bn_ok(z, bn):
        if z.numel() // z.shape[1] <= 1:       # torch.nn.functional.batch_norm's own check
            raise ValueError(f'Expected more than 1 value per channel when training, got input size {z.size()}')
        bn.num_batches_tracked.add_(1)
        return _ReluBN.apply(z, conv_bias, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.momentum, bn.eps)
    if conv_bias is not None:
        z = z + conv_bias.view(1, -1, 1, 1)
    return bn(F.relu(z))


class _ConvReluBN(nn.Module):
    """conv -> relu -> batchnorm (the reference normalises AFTER the activation, net.py:36,54)."""

    def __init__(self, cin, cout, stride=1, pad=0):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 3, stride=stride, padding=pad)
        self.bn = nn.BatchNorm2d(cout)
        nn.init.kaiming_normal_(self.conv.weight, a=4.4, nonlinearity='relu')
        nn.init.zeros_(self.conv.bias)
        nn.init.constant_(self.bn.weight, 0.1)

    def forward(self, x, w=None):
        """w: the weight rearranged by the caller to the channel order of x (see _Up.forward)."""
        if w is None:
            w = self.conv.weight
        if _fused_bn_ok(x, self.bn):     # the bias is added inside the ReLU+BatchNorm kernels
            z = F.conv2d(x, w, None, self.conv.stride, self.conv.padding)
            return relu_batchnorm(z, self.bn, self.conv.bias)
        return relu_batchnorm(F.conv2d(x, w, self.conv.bias, self.conv.stride, self.conv.padding), self.bn)


class _Down(nn.Module):
    """stride-2 3x3 conv, input padded on the right/bottom only (net.py:39-54)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.block = _ConvReluBN(cin, cout, stride=2)

    def forward(self, x):
        return self.block(F.pad(x, (0, 2, 0, 2)))


class _Residual(nn.Module):
    def __init__(self, ch, depth=2):
        super().__init__()
        self.body = nn.Sequential(*[_ConvReluBN(ch, ch, pad=1) for _ in range(depth)])

    def forward(self, x):
        return x + self.body(x)


def _layout(x):
    """(tensor in a layout the kernels take, channels_last flag): a channels_last tensor stays as it
    is (the kernels have (N, H, W, C) variants), anything else is made NCHW-contiguous."""
    if x.dim() == 4 and x.shape[1] > 1 and x.is_contiguous(memory_format=torch.channels_last) \
            and not x.is_contiguous():
        return x, 1
    return x.contiguous(), 0


def _empty(shape, cl, dev):
    return torch.empty(shape, dtype=torch.float32, device=dev,
                       memory_format=torch.channels_last if cl else torch.contiguous_format)


class _UpsamplePad(torch.autograd.Function):
    """`F.pad(F.interpolate(x, scale_factor=2), (1, 1, 1, 1), mode='reflect')` as one gather per
    direction (csrc/glue.cu): the 4x upsampled intermediate is never written."""

    @staticmethod
    def forward(ctx, x):
        x, cl = _layout(x)
        N, C, H, W = x.shape
        y = _empty((N, C, 2 * H + 2, 2 * W + 2), cl, x.device)
        with torch.cuda.device(x.device):
            _lib.call('dvsof_upsample2x_reflect_pad_forward', _lib.ptr(x), N, C, H, W, _lib.ptr(y), cl,
                      _lib.stream_ptr(x.device))
        ctx.shape, ctx.cl = (N, C, H, W), cl
        return y

    @staticmethod
    def backward(ctx, gy):
        cl = ctx.cl
        gy = gy.contiguous(memory_format=torch.channels_last) if cl else gy.contiguous()
        N, C, H, W = ctx.shape
        gx = _empty(ctx.shape, cl, gy.device)
        with torch.cuda.device(gy.device):
            _lib.call('dvsof_upsample2x_reflect_pad_backward', _lib.ptr(gy), N, C, H, W, _lib.ptr(gx), cl,
                      _lib.stream_ptr(gy.device))
        return gx


class _UpsamplePadCat(torch.autograd.Function):
    """upsample2x_reflect_pad(cat(parts + zero channels, 1)) for channels_last parts in one gather;
    backward writes one contiguous gradient per part (csrc/glue.cu)."""

    @staticmethod
    def forward(ctx, pad_in, *parts):
        import ctypes
        N, _, H, W = parts[0].shape
        chans = [p.shape[1] for p in parts]
        dev = parts[0].device
        y = _empty((N, sum(chans) + pad_in, 2 * H + 2, 2 * W + 2), 1, dev)
        ptrs = (ctypes.c_void_p * len(parts))(*[p.data_ptr() for p in parts])
        cc = (ctypes.c_int32 * len(parts))(*chans)
        with torch.cuda.device(dev):
            _lib.call('dvsof_upsample2x_reflect_pad_cat_forward', ptrs, cc, len(parts), pad_in, N, H, W,
                      _lib.ptr(y), _lib.stream_ptr(dev))
        ctx.meta = (N, H, W, chans, pad_in)
        return y

    @staticmethod
    def backward(ctx, gy):
        import ctypes
        N, H, W, chans, pad_in = ctx.meta
        gy = gy.contiguous(memory_format=torch.channels_last)
        need = ctx.needs_input_grad[1:]
        grads = [_empty((N, c, H, W), 1, gy.device) if nd else None for c, nd in zip(chans, need)]
        ptrs = (ctypes.c_void_p * len(chans))(*[g.data_ptr() if g is not None else None for g in grads])
        cc = (ctypes.c_int32 * len(chans))(*chans)
        with torch.cuda.device(gy.device):
            _lib.call('dvsof_upsample2x_reflect_pad_cat_backward', _lib.ptr(gy), ptrs, cc, len(chans), pad_in,
                      N, H, W, _lib.stream_ptr(gy.device))
        return (None,) + tuple(grads)


def _is_cl(t):
    return t.is_cuda and t.dtype == torch.float32 and t.dim() == 4 and \
        t.is_contiguous(memory_format=torch.channels_last) and (not t.is_contiguous() or t.shape[1] == 1)


def upsample2x_reflect_pad_cat(parts, pad_in=0):
    """upsample2x_reflect_pad of the channel concatenation of `parts` (+ pad_in zero channels)."""
    if len(parts) <= 4 and all(_is_cl(p) for p in parts) and parts[0].shape[2] * parts[0].shape[3] > 1:
        return _UpsamplePadCat.apply(pad_in, *parts)
    parts = tuple(parts)
    if pad_in:
        N, _, H, W = parts[0].shape
        parts += (torch.empty((N, pad_in, H, W), dtype=parts[0].dtype, device=parts[0].device,
                              memory_format=torch.channels_last).zero_(),)
    return upsample2x_reflect_pad(torch.cat(parts, 1))


def upsample2x_reflect_pad(x):
    """Head of a decoder block (net.py:72-85).  CUDA fp32 tensors take the fused kernel; anything
    else (the CPU world-size-2 tests of the host logic) the two ATen ops."""
    if x.is_cuda and x.dtype == torch.float32:
        return _UpsamplePad.apply(x)
    return F.pad(F.interpolate(x, scale_factor=2), (1, 1, 1, 1), mode='reflect')


class _Up(nn.Module):
    """nearest x2, reflect pad 1, 3x3 conv (net.py:72-85)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.block = _ConvReluBN(cin, cout)

    def forward(self, h, f, skip):
        """The block's input is cat((h, f, skip), 1) (net.py:246-262; f = the previous flow, None
        for the first block).  In channels_last the parts go straight into the gather, in the order
        (h, skip, f + zero channels) with the total filled up to a multiple of 8 channels -- 130 / 258
        / 514 input channels otherwise make cuDNN copy every operand into a padded buffer around each
        of its three kernels (3.8 ms per step), and a 2-channel part in the middle would break the
        16-byte vectors of the gather.  The weight is rearranged to that order on the fly (input
        channels are permuted and zero slices appended: the parameters stay as they are)."""
        if _is_cl(h) and _is_cl(skip) and (f is None or f.is_cuda):
            w = self.block.conv.weight
            ch, cs = h.shape[1], skip.shape[1]
            if f is None:
                pad_in = (-(ch + cs)) % 8
                parts = (h, skip)
                if pad_in:
                    w = F.pad(w, (0, 0, 0, 0, 0, pad_in))
            else:
                cf = f.shape[1]
                pad_in = (-(ch + cs + cf)) % 8
                N, _, H, W = f.shape
                tail = torch.zeros((N, cf + pad_in, H, W), dtype=f.dtype, device=f.device).contiguous(
                    memory_format=torch.channels_last)
                tail[:, :cf] = f
                parts = (h, skip, tail)
                w = torch.cat((w[:, :ch], w[:, ch + cf:], w[:, ch:ch + cf],
                               w.new_zeros((w.shape[0], pad_in) + tuple(w.shape[2:]))), 1)
            w = w.contiguous(memory_format=torch.channels_last)
            return self.block(upsample2x_reflect_pad_cat(parts, 0), w)
        x = torch.cat((h, skip) if f is None else (h, f, skip), 1)
        return self.block(upsample2x_reflect_pad(x))


class Model(nn.Module):
    def __init__(self, device='cuda'):
        super().__init__()
        self.device = device
        chans = (4,) + _WIDTHS
        self.enc = nn.ModuleList(_Down(chans[i], chans[i + 1]) for i in range(4))
        self.tr = nn.ModuleList(_Residual(_WIDTHS[-1]) for _ in range(2))
        outs = (256, 128, 64, 32)
        self.dec = nn.ModuleList()
        self.flow = nn.ModuleList()
        for i, cout in enumerate(outs):
            cin = 2 * _WIDTHS[3 - i] + (2 if i else 0)     # features + skip (+ previous flow)
            self.dec.append(_Up(cin, cout))
            head = nn.Conv2d(cout, 2, 1)
            nn.init.kaiming_normal_(head.weight, a=4.4, nonlinearity='relu')
            nn.init.zeros_(head.bias)
            self.flow.append(head)
        self._register_load_state_dict_pre_hook(self._accept_reference_keys)
        self.to(device)

    # parameter names of the reference's EV_FlowNet/net.py (GConv2d `conv`/`bn` directly under
    # enc.N, ResNetBlock.conv.M, UpsampleBlock.conv) <-> the names of this module tree
    _KEY_MAP = ((r'^enc\.(\d+)\.(conv|bn)\.', r'enc.\1.block.\2.'),
                (r'^tr\.(\d+)\.conv\.(\d+)\.', r'tr.\1.body.\2.'),
                (r'^dec\.(\d+)\.conv\.(conv|bn)\.', r'dec.\1.block.\2.'))

    @classmethod
    def _accept_reference_keys(cls, state_dict, prefix, *_):
        """`load_state_dict` also takes a checkpoint written by the reference's Model
        (utils/serializer.py `step_N.pt`['model']): its keys are renamed in place."""
        import re
        for key in list(state_dict.keys()):
            local = key[len(prefix):]
            for pat, rep in cls._KEY_MAP:
                new = re.sub(pat, rep, local)
                if new != local:
                    state_dict[prefix + new] = state_dict.pop(key)
                    break

    def reference_state_dict(self):
        """state_dict() under the reference's parameter names (loadable by EV_FlowNet/net.py Model)."""
        import re
        inv = ((r'^enc\.(\d+)\.block\.', r'enc.\1.'), (r'^tr\.(\d+)\.body\.(\d+)\.', r'tr.\1.conv.\2.'),
               (r'^dec\.(\d+)\.block\.', r'dec.\1.conv.'))
        out = {}
        for k, v in self.state_dict().items():
            for pat, rep in inv:
                k2 = re.sub(pat, rep, k)
                if k2 != k:
                    k = k2
                    break
            out[k] = v
        return out

    def quantize(self, events, timestamps, sample_idx, imsize):
        return quantize(events, timestamps, sample_idx, imsize)           # net.py:197-205

    def forward(self, events, timestamps, sample_idx, imsize, raw=True, intermediate=False):
        outsize = [(imsize[0] // 2 ** k, imsize[1] // 2 ** k) for k in (3, 2, 1, 0)]
        x = self.quantize(events, timestamps, sample_idx, imsize) if raw else events
        feats = {'input': x} if intermediate else None
        skips = []
        h = x
        for i, enc in enumerate(self.enc):
            h = enc(h)
            skips.append(h)
            if intermediate:
                feats[f'enc_{i}'] = h
        for i, res in enumerate(self.tr):
            h = res(h)
            if intermediate:
                feats[f'tr_{i}'] = h
        flows = []
        for i, (dec, head) in enumerate(zip(self.dec, self.flow)):
            h = dec(h, flows[-1] if flows else None, skips[3 - i])
            arth = head(h)
            f = torch.tanh(arth) * 256.
            flows.append(f)
            if intermediate:             # net.py:246-258 (`dec_cat_i` is never materialised here: the
                feats[f'dec_op_{i}'] = h  # concatenation happens inside the decoder's gather kernel)
                feats[f'dec_flow_arth_{i}'] = arth
        result = tuple(f[..., :s[0], :s[1]] for f, s in zip(flows, outsize))
        extra = (feats,) if intermediate else ()
        return (result, timestamps.reshape(-1, 2), sample_idx[0::2]) + extra
