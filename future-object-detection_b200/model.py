"""Drop-in modules for the reference's hot path, computing on the fod_b200 CUDA kernels.

Module tree, constructor arguments, `state_dict` keys and forward signatures mirror
  future_od/models/st_detr.py   (SpatioTemporalDETR, SpatioTemporalDETRArgs)
  future_od/models/paper.py     (PositionalEncoder, CDetrBackbone, SeparateEncoder, CDetrDetectorSpatioTemporal,
                                 FuturePredCore)
  future_od/models/transformer.py (MLP, Encoder/Decoder layers and attention blocks)
so a reference checkpoint loads with `load_state_dict` and `runs/_model.py:build_model` can construct this model
unchanged (see fod_b200.build_model).  Internally tokens are batch-first (rows = batch*tokens, 256 channels, bf16)
and every Linear / LayerNorm / attention / loss op is a C-ABI kernel call (autograd.py); torch is used for the
ResNet body (cuDNN), for weight casts and for a few 128-row element-wise ops in the decoder.
Dropout (p = 0.1 in training, as the reference hard-codes) is drawn inside the kernels from a counter-based generator
(csrc/dropout.cuh): every dropout site of a forward pass takes a call id from ops.drop_site(), the masks are a function
of (seed, step, call id, element) and backward regenerates them.  `set_dropout(model, 0.0)` gives the reference's
dropout-free numerics, which is what the parity tests compare.
"""
import math

import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import autograd as A
from . import ops

BF16 = torch.bfloat16


def _q(w):
    """bf16 copy of an fp32 weight for the kernels (no autograd edge; the Function routes the gradient).  Under a
    FlatTrainState (optim.py) parameters carry `_fod_shadow`, a view of the bf16 arena the fused optimizer keeps
    current, and no cast is launched."""
    s = getattr(w, "_fod_shadow", None)
    if s is not None:
        return s
    return w.detach().to(BF16).contiguous()


def _rows(t):
    return t.reshape(-1, t.shape[-1])


def _xavier_(module):
    """transformer.py:12-15 — xavier_uniform_ on every parameter with dim > 1 in the subtree."""
    for p in module.parameters():
        if p.dim() > 1:
            nn.init.xavier_uniform_(p)


def _lin(x, layer, relu=False, out_f32=False, residual=None, res_row_mod=0):
    return A.linear(x, layer.weight, layer.bias, _q(layer.weight), residual, res_row_mod, relu, out_f32)


def _lin_padded(x, layer, n_pad=64):
    """Linear with fewer than 8 outputs (class / box / reference-point heads): weight rows zero-padded to n_pad so the
    tensor-core tile and the MN-major backward apply; fp32 output, sliced back."""
    n = layer.weight.shape[0]
    w = F.pad(layer.weight, (0, 0, 0, n_pad - n))
    b = F.pad(layer.bias, (0, n_pad - n))
    return A.linear(x, w, b, F.pad(_q(layer.weight), (0, 0, 0, n_pad - n)), None, 0, False, True)[:, :n]


class MLP(nn.Module):
    """transformer.py:18-32 — Linear/ReLU stack (attribute name `layers`)."""

    def __init__(self, input_dim, hidden_dim, output_dim, num_layers):
        super().__init__()
        h = [hidden_dim] * (num_layers - 1)
        self.num_layers = num_layers
        self.layers = nn.ModuleList(nn.Linear(n, k) for n, k in zip([input_dim] + h, h + [output_dim]))

    def forward(self, x, out_f32_last=False):
        for i, layer in enumerate(self.layers):
            last = i == self.num_layers - 1
            if last and layer.weight.shape[0] < 8:
                x = _lin_padded(x, layer)
            else:
                x = _lin(x, layer, relu=not last, out_f32=last and out_f32_last)
        return x


# ------------------------------------------------------------------------------------------------ positional encodings
def _sincos(angles):
    return torch.stack((angles[..., 0::2].sin(), angles[..., 1::2].cos()), dim=-1).flatten(-2)


class PositionalEncoder(nn.Module):
    """paper.py:32-80.  The 2-D sine encoding is input-independent, so it is built once per (h, w) and cached."""

    def __init__(self, no_temporal=False, temperature=10000.0, extra_temporal_offset=0.0):
        super().__init__()
        self.no_temporal = no_temporal
        self.temperature = temperature
        self.extra_temporal_offset = extra_temporal_offset
        self._cache = {}

    def spatial_tokens(self, h, w, c, device):
        """(h*w, c) fp32: tokens in (h w) order, first c/2 channels encode y (paper.py:57-64, 75-80)."""
        key = (h, w, c, str(device))
        if key not in self._cache:
            half = c // 2
            ones = torch.ones((h, w), dtype=torch.float32, device=device)
            ys, xs = ones.cumsum(0), ones.cumsum(1)
            ys = ys / (ys[-1:, :] + 1e-6) * (2 * math.pi)
            xs = xs / (xs[:, -1:] + 1e-6) * (2 * math.pi)
            idx = torch.arange(half, dtype=torch.float32, device=device)
            dim_t = self.temperature ** (2 * torch.div(idx, 2, rounding_mode="floor") / half)
            enc = torch.cat((_sincos(ys[..., None] / dim_t), _sincos(xs[..., None] / dim_t)), dim=-1)
            self._cache[key] = enc.reshape(h * w, c).contiguous()
        return self._cache[key]

    def temporal_channels(self, offsets, b, t, c, device):
        """paper.py:66-73 — (B, T, c) fp32 sine code of the frame offsets (None: the frame counter 1..T), normalised
        by the LAST frame's value; it is constant over a frame's tokens, so it is kept as one vector per frame."""
        if offsets is not None:
            te = offsets.to(device=device, dtype=torch.float32) + self.extra_temporal_offset
        else:
            te = torch.arange(1, t + 1, dtype=torch.float32, device=device).expand(b, t)
        te = te / (te[:, -1:] + 1e-6) * (2 * math.pi)
        idx = torch.arange(c, dtype=torch.float32, device=device)
        dim_t = self.temperature ** (2 * torch.div(idx, 2, rounding_mode="floor") / c)
        return _sincos(te[..., None] / dim_t)


def gen_sineembed_for_position(ref_xy, d=256):
    """transformer.py:35-48 — ref_xy (..., 2) in [0,1] -> (..., d) = [sine(y) ‖ sine(x)]."""
    half = d // 2
    idx = torch.arange(half, dtype=torch.float32, device=ref_xy.device)
    dim_t = 10000 ** (2 * torch.div(idx, 2, rounding_mode="floor") / half)
    ax = ref_xy[..., 0, None] * (2 * math.pi) / dim_t
    ay = ref_xy[..., 1, None] * (2 * math.pi) / dim_t
    return torch.cat((_sincos(ay), _sincos(ax)), dim=-1)


def inverse_sigmoid(x, eps=1e-5):
    """ConditionalDETR/util/misc.py:477-481."""
    x = x.clamp(0, 1)
    return torch.log(x.clamp(min=eps) / (1 - x).clamp(min=eps))


# ------------------------------------------------------------------------------------------------ backbone (cuDNN)
def _is_main_process():
    """ConditionalDETR/util/misc.py:405-406."""
    import torch.distributed as dist
    return not (dist.is_available() and dist.is_initialized()) or dist.get_rank() == 0


def _load_pretrained_backbone(backbone, name):
    """torchvision's ImageNet weights for `name` (what `pretrained=True` means upstream, paper.py:94-98): a local file
    named by FOD_B200_BACKBONE_WEIGHTS, else torchvision's own cache / download.  Fails with the remedy spelled out."""
    import os

    import torchvision
    path = os.environ.get("FOD_B200_BACKBONE_WEIGHTS")
    try:
        if path:
            sd = torch.load(path, map_location="cpu", weights_only=True)
        else:
            sd = torchvision.models.get_model_weights(name).DEFAULT.get_state_dict(progress=False)
    except Exception as e:  # no network and no cached file
        raise RuntimeError(
            "pretrained_backbone=True needs torchvision's ImageNet weights for %s and they could not be loaded (%s: %s). "
            "Either point FOD_B200_BACKBONE_WEIGHTS at a local torchvision state_dict file, or build with "
            "SpatioTemporalDETRArgs(pretrained_backbone=False) and load_state_dict() a checkpoint (INTEGRATION.md §2)."
            % (name, type(e).__name__, str(e)[:200])) from e
    missing, unexpected = backbone.load_state_dict(sd, strict=False)
    bad = [k for k in missing if "num_batches_tracked" not in k]
    if bad:
        raise RuntimeError("backbone weights for %s lack %s" % (name, bad[:5]))


class FrozenBatchNorm2d(nn.Module):
    """ConditionalDETR/models/backbone.py:27-63 — statistics and affine are buffers; eps 1e-5."""

    def __init__(self, n):
        super().__init__()
        self.register_buffer("weight", torch.ones(n))
        self.register_buffer("bias", torch.zeros(n))
        self.register_buffer("running_mean", torch.zeros(n))
        self.register_buffer("running_var", torch.ones(n))

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        state_dict.pop(prefix + "num_batches_tracked", None)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def forward(self, x):
        scale = self.weight * (self.running_var + 1e-5).rsqrt()
        shift = self.bias - self.running_mean * scale
        return x * scale.view(1, -1, 1, 1).to(x.dtype) + shift.view(1, -1, 1, 1).to(x.dtype)

    def fold(self):
        """(scale, shift) of the frozen affine map, cached: the four buffers only change through load_state_dict / .to,
        which bump their version counters.  Refreshed IN PLACE so a captured CUDA graph keeps reading valid memory."""
        bufs = (self.weight, self.bias, self.running_mean, self.running_var)
        key = tuple(b._version for b in bufs) + (self.weight.device, self.weight.data_ptr())
        if getattr(self, "_fold_key", None) != key:
            with torch.no_grad():
                s = self.weight * (self.running_var + 1e-5).rsqrt()
                t = self.bias - self.running_mean * s
                old = getattr(self, "_fold_st", None)
                if old is not None and old[0].device == s.device and old[0].shape == s.shape:
                    old[0].copy_(s)
                    old[1].copy_(t)
                    self._fold_tq.copy_(t)
                else:
                    self._fold_st = (s, t)
                    self._fold_tq = t.to(BF16)  # the fused conv's bias operand, cast once instead of per step
            self._fold_key = key
            if not getattr(self, "_fold_hooked", False):
                self.register_load_state_dict_post_hook(lambda m, keys: (m.fold(), None)[1])
                self._fold_hooked = True
        return self._fold_st


class _ConvAffineAct(torch.autograd.Function):
    """y = [relu](conv(x, W·s) + t [+ z]) on cuDNN with the frozen BatchNorm (scale s, shift t) folded into the conv —
    library code for the backbone (not a hand-written kernel).  Forward is ONE fused cuDNN launch where
    the runtime offers it (cudnn_convolution_relu / cudnn_convolution_add_relu); backward is one mask pass + cuDNN
    dgrad / wgrad.  Replaces conv -> x*scale -> +shift -> (+identity) -> relu, i.e. 3-4 full passes per layer.
    The fold is done here (dW = dW_folded · s), so an arena-backed weight gets its gradient written in place on the
    weight-gradient stream (autograd._OffCriticalPath) instead of through MulBackward + AccumulateGrad."""

    fused_ok = True

    @staticmethod
    def forward(ctx, x, w, scale, shift, stride, padding, relu, z, wq_pre=None):
        """shift None: bias-free (the downsample branch, whose shift rides on the consuming conv3).  wq_pre: W·s already
        folded for this step by ops.ConvFoldTable (one launch for the whole backbone)."""
        s4 = scale.view(-1, 1, 1, 1)
        if wq_pre is not None:
            wq = wq_pre
            bq = shift.to(x.dtype) if shift is not None else None
        elif w.requires_grad:
            wq = (w.detach() * s4).to(x.dtype).contiguous(memory_format=torch.channels_last)
            bq = shift.to(x.dtype) if shift is not None else None
        else:  # frozen stem / layer1 (paper.py:102-109): folded operands cached on the version counters
            key = (w._version, scale._version, x.dtype, w.device, w.data_ptr(), scale.data_ptr())
            c = getattr(w, "_fod_folded", None)
            if c is None or c[0] != key:
                c = (key, (w.detach() * s4).to(x.dtype).contiguous(memory_format=torch.channels_last))
                w._fod_folded = c
            wq = c[1]
            bq = shift.to(x.dtype) if shift is not None else None
        y = None
        if relu and _ConvAffineAct.fused_ok:
            try:
                if z is None:
                    y = torch.cudnn_convolution_relu(x, wq, bq, stride, padding, (1, 1), 1)
                else:
                    y = torch.cudnn_convolution_add_relu(x, wq, z, 1.0, bq, stride, padding, (1, 1), 1)
            except RuntimeError as e:
                _ConvAffineAct.fused_ok = False
                import warnings
                warnings.warn("fod_b200: the fused cuDNN conv+bias(+add)+ReLU call failed (%s); the backbone uses the "
                              "unfused conv -> add -> relu path for the rest of this process" % str(e)[:200])
        if y is None:
            y = F.conv2d(x, wq, bq, stride, padding)
            if z is not None:
                y = y + z
            if relu:
                y = y.relu_()
        ctx.save_for_backward(x, wq, y if relu else None, s4)
        ctx.cfg = (stride, padding, relu, z is not None)
        ctx.direct = A._use(w)[0] if ctx.needs_input_grad[1] else False
        ctx.w = w
        return y

    @staticmethod
    def backward(ctx, dy):
        x, wq, y, s4 = ctx.saved_tensors
        stride, padding, relu, has_z = ctx.cfg
        g = torch.ops.aten.threshold_backward(dy, y, 0) if relu else dy
        g = g.contiguous(memory_format=torch.channels_last)
        cb = torch.ops.aten.convolution_backward
        dw = None
        if ctx.direct:
            w = ctx.w
            dx = cb(g, x, wq, None, stride, padding, (1, 1), False, (0, 0), 1, (True, False, False))[0] \
                if ctx.needs_input_grad[0] else None
            with A._OffCriticalPath(w._fod_reducer, g, x):
                dwq = cb(g, x, wq, None, stride, padding, (1, 1), False, (0, 0), 1, (False, True, False))[1]
                if dwq.is_cuda and dwq.dtype == BF16 and dwq.permute(0, 2, 3, 1).is_contiguous() and w.grad.is_contiguous():
                    ops.conv_wgrad_unfold(dwq, s4.view(-1), w.grad)   # dW += dW_folded · s, one launch
                else:
                    w.grad.addcmul_(dwq.float(), s4)
            A._done(w)
        else:
            dx, dwq, _ = cb(g, x, wq, None, stride, padding, (1, 1), False, (0, 0), 1,
                            (ctx.needs_input_grad[0], ctx.needs_input_grad[1], False))
            if dwq is not None:
                dw = dwq.float() * s4
        return dx, dw, None, None, None, None, None, (g if has_z else None), None


class _BottleneckStackFn(torch.autograd.Function):
    """All bottleneck blocks of the trainable ResNet stages (layer2-4) as ONE autograd node — library convolutions (cuDNN
    fused conv+bias(+add)+ReLU forward, dgrad / wgrad backward; SURVEY.md §8 f2: not hand-written kernels) with the
    element-wise glue of the backward pass done by fod_relu_bwd_add.

    Per block autograd would run: threshold_backward on the block output (a full-size pass), two more on the mid
    activations, and an add of the two input-gradient contributions (conv1's dgrad and the identity / downsample path) —
    at batch 32 and 448x800 these passes move ~1.1 GB each on layer2.  Here the add is deferred into the NEXT mask pass
    (out = (a + b) * (y > 0): 4 tensors of traffic instead of 6) and the block-to-block gradients never become autograd
    tensors.  Weight gradients are produced exactly as in _ConvAffineAct (dW = dW_folded * s, written into the arena on
    the weight-gradient stream when the parameter owns a slot)."""

    @staticmethod
    def forward(ctx, x, blocks, *weights):
        """blocks: list of dicts {convs: [(widx, scale, shift_bf16 or None, stride, padding, wq)] for conv1, conv2, conv3,
        ds: the same tuple for the downsample conv or None}; `weights` are the fp32 parameters (widx indexes them)."""
        acts = []
        for blk in blocks:
            c1, c2, c3 = blk["convs"]
            if blk["ds"] is not None:
                d = blk["ds"]
                idt = F.conv2d(x, d[5], None, d[3], d[4])
            else:
                idt = x
            o1 = torch.cudnn_convolution_relu(x, c1[5], c1[2], c1[3], c1[4], (1, 1), 1)
            o2 = torch.cudnn_convolution_relu(o1, c2[5], c2[2], c2[3], c2[4], (1, 1), 1)
            out = torch.cudnn_convolution_add_relu(o2, c3[5], idt, 1.0, c3[2], c3[3], c3[4], (1, 1), 1)
            acts.append((x, o1, o2, out))
            x = out
        ctx.blocks, ctx.acts, ctx.weights = blocks, acts, weights
        ctx.direct = [A._use(w)[0] if ctx.needs_input_grad[2 + i] else False for i, w in enumerate(weights)]
        return x

    @staticmethod
    def _wgrad(ctx, grads, conv, g, x):
        widx, scale, _, stride, padding, wq = conv
        w = ctx.weights[widx]
        if not ctx.needs_input_grad[2 + widx]:
            return
        cb = torch.ops.aten.convolution_backward
        s4 = scale.view(-1, 1, 1, 1)
        if ctx.direct[widx]:
            with A._OffCriticalPath(w._fod_reducer, g, x):
                dwq = cb(g, x, wq, None, stride, padding, (1, 1), False, (0, 0), 1, (False, True, False))[1]
                if dwq.dtype == BF16 and dwq.permute(0, 2, 3, 1).is_contiguous() and w.grad.is_contiguous():
                    ops.conv_wgrad_unfold(dwq, s4.view(-1), w.grad)
                else:
                    w.grad.addcmul_(dwq.float(), s4)
            A._done(w)
        else:
            dwq = cb(g, x, wq, None, stride, padding, (1, 1), False, (0, 0), 1, (False, True, False))[1]
            grads[widx] = dwq.float() * s4

    @staticmethod
    def backward(ctx, dy):
        cb = torch.ops.aten.convolution_backward
        grads = [None] * len(ctx.weights)

        def dgrad(conv, g, x):
            return cb(g, x, conv[5], None, conv[3], conv[4], (1, 1), False, (0, 0), 1, (True, False, False))[0]

        def like(t, ref):  # the fused mask kernel wants identical dense layouts
            return t if t.stride() == ref.stride() else t.contiguous(memory_format=torch.channels_last)

        ga, gb = dy, None
        for bi in range(len(ctx.blocks) - 1, -1, -1):
            blk = ctx.blocks[bi]
            x, o1, o2, out = ctx.acts[bi]
            c1, c2, c3 = blk["convs"]
            g3 = ops.relu_bwd_add(like(ga, out), None if gb is None else like(gb, out), out)
            need_dx = bi > 0 or ctx.needs_input_grad[0]
            g2 = ops.relu_bwd_add(like(dgrad(c3, g3, o2), o2), None, o2)
            _BottleneckStackFn._wgrad(ctx, grads, c3, g3, o2)
            g1 = ops.relu_bwd_add(like(dgrad(c2, g2, o1), o1), None, o1)
            _BottleneckStackFn._wgrad(ctx, grads, c2, g2, o1)
            ga = dgrad(c1, g1, x) if need_dx else None
            _BottleneckStackFn._wgrad(ctx, grads, c1, g1, x)
            if blk["ds"] is not None:
                gb = dgrad(blk["ds"], g3, x) if need_dx else None
                _BottleneckStackFn._wgrad(ctx, grads, blk["ds"], g3, x)
            else:
                gb = g3
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ga + gb
        ctx.acts = None
        return (dx, None) + tuple(grads)


def _conv_bn(x, conv, bn, relu, z=None, wq=None, shift=True, extra_shift=None):
    """Frozen-BN fold: conv(x, W)·s + t == conv(x, W·s) + t (backbone.py:53-63).  shift=False leaves t out (the caller
    adds it elsewhere); extra_shift is added to t."""
    s, t = bn.fold()
    if x.dtype == BF16:
        t = bn._fold_tq
    if not shift:
        t = None
    elif extra_shift is not None:
        t = t + extra_shift.to(t.dtype)
    return _ConvAffineAct.apply(x, conv.weight, s, t, conv.stride, conv.padding, relu, z, wq)


class CDetrBackbone(nn.Module):
    """paper.py:83-116.  The ResNet body stays on torch/cuDNN (channels_last, bf16 autocast); the 1x1 `input_proj`
    is a GEMM over NHWC rows and runs on the tcgen05 kernel, emitting encoder tokens directly."""

    def __init__(self, name, train_backbone, dilation, hidden_dim, pretrained=True):
        super().__init__()
        import torchvision
        from torchvision.models._utils import IntermediateLayerGetter

        # paper.py:94-98: rank 0 alone loads torchvision's ImageNet weights (`pretrained=is_main_process() and pretrained`);
        # the other ranks get them through the construction-time broadcast (DDP there, FlatTrainState here).
        # FOD_B200_BACKBONE_WEIGHTS=<file> names a local torchvision state_dict for boxes without network access.
        backbone = getattr(torchvision.models, name)(
            replace_stride_with_dilation=[False, False, dilation], weights=None, norm_layer=FrozenBatchNorm2d)
        if pretrained and _is_main_process():
            _load_pretrained_backbone(backbone, name)
        for pname, p in backbone.named_parameters():
            if not train_backbone or ("layer2" not in pname and "layer3" not in pname and "layer4" not in pname):
                p.requires_grad_(False)
        self.body = IntermediateLayerGetter(backbone, return_layers={"layer4": "0"})
        self.num_channels = 512 if name in ("resnet18", "resnet34") else 2048
        self.input_proj = nn.Conv2d(self.num_channels, hidden_dim, kernel_size=(1, 1))
        self.compute_dtype = BF16
        self.fold_frozen_bn = name not in ("resnet18", "resnet34")  # Bottleneck layout; BasicBlock uses the stock path
        self.fuse_backbone_backward = True  # layer2-4 as one autograd node with fused mask(+add) passes (training)

    def _body_folded(self, x):
        """torchvision ResNet (Bottleneck) forward with every conv+FrozenBN(+add)+ReLU group as one fused op."""
        b = self.body
        pre = self._prefold(x)
        c1 = b["conv1"]
        pooled = False
        padded = isinstance(x, ops.PaddedFrames)
        if padded and (c1.weight.requires_grad or c1.bias is not None or tuple(c1.weight.shape) != (64, 3, 7, 7)
                       or (c1.stride, c1.padding) != ((2, 2), (3, 3)) or x.h % 2):
            x, padded = x.nchw().contiguous(memory_format=torch.channels_last), False  # not the frozen ResNet stem
        if padded:
            # frozen stem, operand tiles fetched by TMA from the padded 4-channel frames (csrc/stemconv_sm100.cu)
            s1, t1 = b["bn1"].fold()
            w1 = c1.weight
            key = (w1._version, s1._version, t1._version, w1.data_ptr(), s1.data_ptr())
            c = getattr(w1, "_fod_stem4", None)
            if c is None or c[0] != key:
                c = (key, ops.stem_weights4(w1, s1), t1.detach().float().contiguous())
                w1._fod_stem4 = c
            mp = b["maxpool"]
            if (mp.kernel_size, mp.stride, mp.padding) == (3, 2, 1) and _STEM_POOL:
                x, pooled = ops.stem_conv_pool_p4(x, c[1], c[2]), True  # conv + bn + relu + maxpool, one kernel
            else:
                x = ops.stem_conv7x7s2_p4(x, c[1], c[2])
        elif (x.is_cuda and x.dtype == BF16 and not x.requires_grad and not c1.weight.requires_grad and c1.bias is None
                and tuple(c1.weight.shape) == (64, 3, 7, 7) and (c1.stride, c1.padding) == ((2, 2), (3, 3))
                and x.shape[3] % 8 == 0 and x.is_contiguous(memory_format=torch.channels_last)):
            # frozen stem (paper.py:102-109): hand-written implicit-GEMM kernel, operands cached on the version counters
            s1, t1 = b["bn1"].fold()
            w1 = c1.weight
            key = (w1._version, s1._version, t1._version, w1.data_ptr(), s1.data_ptr())
            c = getattr(w1, "_fod_stem", None)
            if c is None or c[0] != key:
                c = (key, ops.stem_weights(w1, s1), t1.detach().float().contiguous())
                w1._fod_stem = c
            x = ops.stem_conv7x7s2(x, c[1], c[2])
        else:
            x = _conv_bn(x, c1, b["bn1"], True)
        mp = b["maxpool"]
        if pooled:
            pass
        elif (not x.requires_grad and x.is_cuda and x.dtype == BF16 and (mp.kernel_size, mp.stride, mp.padding) == (3, 2, 1)
                and x.is_contiguous(memory_format=torch.channels_last)):
            x = ops.maxpool3x3s2_nhwc(x)  # frozen stem (paper.py:102-109): forward only, one HBM-bound kernel
        else:
            x = mp(x)
        stack, weights = [], []
        for name in ("layer1", "layer2", "layer3", "layer4"):
            for blk in b[name]:
                trainable = blk.conv1.weight.requires_grad
                if (trainable and torch.is_grad_enabled() and x.is_cuda and x.dtype == BF16 and _ConvAffineAct.fused_ok
                        and self.fuse_backbone_backward and all(c.weight in pre for c in self._block_convs(blk))):
                    # trainable stages (layer2-4) in training: collected and run as ONE autograd node below
                    def spec(conv, bn, shift=True, extra=None):
                        s_, _ = bn.fold()
                        t = bn._fold_tq if shift else None
                        if t is not None and extra is not None:
                            t = t + extra
                        weights.append(conv.weight)
                        return (len(weights) - 1, s_, t, conv.stride, conv.padding, pre[conv.weight])
                    ds = blk.downsample
                    stack.append({"convs": [spec(blk.conv1, blk.bn1), spec(blk.conv2, blk.bn2),
                                            spec(blk.conv3, blk.bn3, extra=ds[1]._fold_tq if ds is not None else None)],
                                  "ds": spec(ds[0], ds[1], shift=False) if ds is not None else None})
                    continue
                if stack:  # (not reached with torchvision's layer order: trainable stages come last)
                    x = _BottleneckStackFn.apply(x, stack, *weights)
                    stack, weights = [], []
                extra = None
                if blk.downsample is None:
                    idt = x
                else:
                    # relu(conv3(o)·s3 + t3 + convd(x)·sd + td): the downsample shift td rides on conv3's fused bias, so
                    # the (bias-free) downsample conv needs no separate strided bias-add pass
                    ds = blk.downsample
                    idt = _conv_bn(x, ds[0], ds[1], False, wq=pre.get(ds[0].weight), shift=False)
                    extra = ds[1]._fold_tq if x.dtype == BF16 else ds[1].fold()[1]
                o = _conv_bn(x, blk.conv1, blk.bn1, True, wq=pre.get(blk.conv1.weight))
                o = _conv_bn(o, blk.conv2, blk.bn2, True, wq=pre.get(blk.conv2.weight))
                x = _conv_bn(o, blk.conv3, blk.bn3, True, idt, wq=pre.get(blk.conv3.weight), extra_shift=extra)
        if stack:
            x = _BottleneckStackFn.apply(x, stack, *weights)
        return x

    @staticmethod
    def _block_convs(blk):
        convs = [blk.conv1, blk.conv2, blk.conv3]
        if blk.downsample is not None:
            convs.append(blk.downsample[0])
        return convs

    def _prefold(self, x):
        """W·s (bf16, channels_last) of every TRAINABLE body conv for this step in one launch (ops.ConvFoldTable);
        {} when not on the CUDA bf16 path.  Frozen convs keep their cached fold (_ConvAffineAct.forward)."""
        if not (x.is_cuda and self.compute_dtype == BF16):
            return {}
        b = self.body
        pairs = []
        for name in ("layer1", "layer2", "layer3", "layer4"):
            for blk in b[name]:
                group = [(blk.conv1, blk.bn1), (blk.conv2, blk.bn2), (blk.conv3, blk.bn3)]
                if blk.downsample is not None:
                    group.append((blk.downsample[0], blk.downsample[1]))
                pairs += [(c.weight, bn.fold()[0]) for c, bn in group if c.weight.requires_grad]
        if not pairs:
            return {}
        ws, ss = [w for w, _ in pairs], [s_ for _, s_ in pairs]
        tab = getattr(self, "_fold_table", None)
        if tab is None or not tab.valid_for(ws, ss):
            tab = ops.ConvFoldTable([w.detach() for w in ws], ss)
            self._fold_table = tab
        outs = tab.run()
        return dict(zip(ws, outs))

    def forward_tokens(self, images):
        """(N,3,H,W) -> tokens (N*h*w, hidden) bf16 in (n, h, w) row order, plus (h, w)."""
        if isinstance(images, ops.PaddedFrames):
            x = images if (self.fold_frozen_bn and self.compute_dtype == BF16) else images.nchw()
        else:
            x = images
        if not isinstance(x, ops.PaddedFrames):
            x = x.contiguous(memory_format=torch.channels_last)
        if self.fold_frozen_bn:
            f = self._body_folded(x if isinstance(x, ops.PaddedFrames) else x.to(self.compute_dtype))
        else:
            with torch.autocast("cuda", dtype=self.compute_dtype, enabled=self.compute_dtype != torch.float32):
                f = self.body(x)["0"]
        n, c, h, w = f.shape
        rows = f.permute(0, 2, 3, 1).to(BF16).contiguous().view(n * h * w, c)
        wgt = self.input_proj.weight.view(self.input_proj.weight.shape[0], c)
        wq = _q(self.input_proj.weight).view(wgt.shape)
        wsrc = self.input_proj.weight if A._direct(self.input_proj.weight) else wgt
        return A.linear(rows, wsrc, self.input_proj.bias, wq), h, w

    def forward(self, images):
        t, h, w = self.forward_tokens(images)
        return t.view(images.shape[0], h, w, -1).permute(0, 3, 1, 2).float()


# ------------------------------------------------------------------------------------------------ encoder
_EGO_BIAS_ROW = os.environ.get("FOD_EGO_BIAS_ROW", "1") != "0"  # 0: out_proj bias of the egodeep table product as baddbmm addend
_EGO_K = 16
_STEM_POOL = os.environ.get("FOD_STEM_POOL", "1") != "0"  # 0: stem convolution and max-pool as two kernels (A/B switch)


class _InProjMHA(nn.Module):
    """Parameter holder with torch.nn.MultiheadAttention's names (transformer.py:404): packed in-proj + out_proj."""

    def __init__(self, d, heads):
        super().__init__()
        self.embed_dim, self.num_heads = d, heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * d, d))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * d))
        self.out_proj = nn.Linear(d, d)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.zeros_(self.out_proj.bias)


def _pos_mod(pos_bf, seq_len):
    """Row modulus of a positional residual: seq_len rows are shared by all sequences, otherwise one row per token."""
    return seq_len if pos_bf.shape[0] == seq_len else 0


class EncoderAttention(nn.Module):
    """transformer.py:401-419: x1 = LN(x + MHA(q=k=x+pos, v=x)); x2 = LN(x1 + W2 relu(W1 x1))."""

    def __init__(self, D, Nhead, Dff, dropout=0.1):
        super().__init__()
        self.attn = _InProjMHA(D, Nhead)
        self.norm1 = nn.LayerNorm(D)
        self.mlp = nn.Sequential(nn.Linear(D, Dff), nn.ReLU(inplace=True), nn.Dropout(dropout), nn.Linear(Dff, D))
        self.norm2 = nn.LayerNorm(D)
        self.heads = Nhead
        self.p = dropout

    def forward(self, x, pos_bf, n_seq, seq_len):
        """x (n_seq*seq_len, D) bf16; pos_bf (seq_len, D) bf16 shared by every sequence, or (n_seq*seq_len, D) when the
        encoding differs per sequence (temporal code in joint mode)."""
        w, b = self.attn.in_proj_weight, self.attn.in_proj_bias
        wq = _q(w)
        d = w.shape[1]
        # (x + pos)·Wᵀ = x·Wᵀ + pos·Wᵀ: the positional half is batch-invariant -> a (seq_len, 3D) row-modular residual
        if A._direct(w):  # gradient of the positional half accumulates into rows [0, 2d) of in_proj_weight.grad
            pqk = A.linear(pos_bf, w, None, wq[:2 * d], w_rows=(0, 2 * d))
        else:
            pqk = A.linear(pos_bf, w[:2 * d], None, wq[:2 * d])
        res = torch.cat((pqk, pqk.new_zeros(pqk.shape[0], d)), dim=1)
        qkv = A.linear(x, w, b, wq, residual=res, res_row_mod=_pos_mod(pos_bf, seq_len))
        site = lambda: ops.drop_site(self.p, self.training)  # noqa: E731 — transformer.py:403-412 dropout sites
        a = A.self_attention(qkv.view(n_seq, seq_len, 3 * d), self.heads, (d // self.heads) ** -0.5, drop=site())
        op = self.attn.out_proj
        x = A.linear_ln(a.view(-1, d), x, op.weight, op.bias, _q(op.weight), self.norm1.weight, self.norm1.bias,
                        self.norm1.eps, drop=site())
        l1, l2 = self.mlp[0], self.mlp[3]
        return A.ffn_ln(x, l1.weight, l1.bias, _q(l1.weight), l2.weight, l2.bias, _q(l2.weight), self.norm2.weight,
                        self.norm2.bias, self.norm2.eps, drop_h=site(), drop_o=site())


class _OutProjOnly(nn.Module):
    """ConditionalDETR MultiheadAttention parameters (attention.py:52-172): no in-projection, `out_proj` only."""

    def __init__(self, vdim):
        super().__init__()
        self.out_proj = nn.Linear(vdim, vdim)


class _ZeroGrad(torch.autograd.Function):
    """Gives numerically dead parameters an exact zero gradient (what the reference produces with dropout off), so
    DDP(find_unused_parameters=False) and the 'got no gradient' check of trainer.py:181-185 behave as upstream."""

    @staticmethod
    def forward(ctx, x, *dead):
        ctx.shapes = [(p.shape, p.dtype, p.device) for p in dead]
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return (g,) + tuple(torch.zeros(s, dtype=dt, device=dv) for s, dt, dv in ctx.shapes)


class EgodeepAttention(nn.Module):
    """transformer.py:85-119.  With ONE IMU key per sequence the softmax weight is exactly 1, so the block reduces to
    o = LN2(o1 + mlp(o1)), o1 = LN1(2·out_proj(value(e))) per sequence; query_content / query_pos /
    key cannot influence the output and receive zero gradients.  With several keys (JointEncoder) the general
    attention path is used."""

    def __init__(self, D, Nhead, Dff=None, dropout=0.1):
        super().__init__()
        self.query_content = nn.Linear(D, D)
        self.query_pos = nn.Linear(D, D)
        self.key = nn.Linear(D, D)
        self.value = nn.Linear(D, D)
        self.fun = _OutProjOnly(D)
        self.heads = Nhead
        self.p = dropout
        self.use_mlp = Dff is not None
        if self.use_mlp:
            self.norm1 = nn.LayerNorm(D)
            self.mlp = nn.Sequential(nn.Linear(D, Dff), nn.ReLU(inplace=True), nn.Dropout(dropout), nn.Linear(Dff, D))
            self.norm2 = nn.LayerNorm(D)

    def _site(self):
        return ops.drop_site(self.p, self.training)

    def _post(self, o):
        if not self.use_mlp:
            return o
        # transformer.py:116-118: out = norm1(out + attn_dropout(out)); out = norm2(out + mlp(out))
        site = self._site()
        if site is not None:  # out + attn_dropout(out): residual, dropout and norm in one kernel
            o = A.add_ln_drop(o, o, self.norm1.weight, self.norm1.bias, eps=self.norm1.eps, drop=site)
        else:
            o = A.add_ln(o * 2, self.norm1.weight, self.norm1.bias, eps=self.norm1.eps)
        l1, l2 = self.mlp[0], self.mlp[3]
        return A.ffn_ln(o, l1.weight, l1.bias, _q(l1.weight), l2.weight, l2.bias, _q(l2.weight), self.norm2.weight,
                        self.norm2.bias, self.norm2.eps, drop_h=self._site(), drop_o=self._site())

    def forward_single_key(self, ego):
        """ego (n_seq, D) bf16 -> per-sequence update vector (n_seq, D) bf16."""
        o = _lin(_lin(ego, self.value), self.fun.out_proj)
        o = self._post(o)
        dead = [p for m in (self.query_content, self.query_pos, self.key) for p in m.parameters() if p.requires_grad]
        return _ZeroGrad.apply(o, *dead) if dead and torch.is_grad_enabled() else o

    def _out_proj_of_masked_heads(self, v, keep):
        """v (n_seq, D) value rows, keep (n_seq, seq_len, H) per-token head multipliers -> out_proj(v with heads scaled)
        for every token, (n_seq*seq_len, D) bf16, through the (n_seq, H, D) table of per-head out_proj contributions."""
        n_seq, seq_len, kw = keep.shape
        h, d = self.heads, v.shape[1]
        op = self.fun.out_proj
        table = torch.einsum("shk,ohk->sho", v.float().view(n_seq, h, d // h), op.weight.view(d, h, d // h)).to(BF16)
        if not _EGO_BIAS_ROW:
            return torch.baddbmm(op.bias.to(BF16).view(1, 1, d), keep[..., :h], table).view(n_seq * seq_len, d)
        # The bias rides as one more table row against a column of ones (K padded to 16 for the library's aligned
        # kernels): as baddbmm's broadcast addend its gradient was a separate reduction of the whole (n_seq, seq_len, D)
        # gradient — 35 us per layer at 64 x 350 tokens — which the table-gradient product now does on the way.
        if kw == h:
            keep = F.pad(keep, (0, _EGO_K - h))
            keep[..., h] = 1
        rows = (table, op.bias.to(BF16).view(1, 1, d).expand(n_seq, 1, d), table.new_zeros((n_seq, _EGO_K - h - 1, d)))
        return torch.bmm(keep, torch.cat(rows, dim=1)).view(n_seq * seq_len, d)

    def forward_single_key_dropout(self, ego, n_seq, seq_len):
        """ego (n_seq, D) -> per-TOKEN update (n_seq*seq_len, D) with dropout on.  The softmax weight over one key is
        still exactly 1, so dropout(P)·V is the sequence's value row with every HEAD kept with probability 1-p and
        scaled by 1/(1-p), independently per token (attention.py:370 drops entries of the (B·heads, Lq, Lk) weights);
        the sub-block dropouts after it (transformer.py:95-103,117) are per element, hence per-token rows from here on."""
        d, h = ego.shape[1], self.heads
        v = _lin(ego, self.value)
        # out_proj is linear and a token's attention output is its sequence's value row with heads switched off, so
        #   out_proj(a_token) = sum_h keep[token, h] * (W_o[:, head h] · v_seq[head h]) + b_o = keep_token (1 x H) · M_seq (H x D) + b_o:
        # the per-sequence table M is 8 rows (tiny torch ops), and the per-token out_proj GEMM (K = 256), its input, its
        # dX / dW GEMMs and the mask multiply become one K = 8 batched library product and its transpose in backward.
        if _EGO_BIAS_ROW:  # columns [0, h): head multipliers; column h: the ones that carry the out_proj bias; rest: zero
            keep = (torch.rand((n_seq, seq_len, _EGO_K), device=ego.device) >= self.p).to(BF16) * (1.0 / (1.0 - self.p))
            keep[..., h:] = 0
            keep[..., h] = 1
        else:
            keep = (torch.rand((n_seq, seq_len, h), device=ego.device) >= self.p).to(BF16) * (1.0 / (1.0 - self.p))
        o = self._post(self._out_proj_of_masked_heads(v, keep))
        dead = [p for m in (self.query_content, self.query_pos, self.key) for p in m.parameters() if p.requires_grad]
        return _ZeroGrad.apply(o, *dead) if dead and torch.is_grad_enabled() else o

    def forward_multi_key(self, x, pos_bf, ego, n_seq, seq_len):
        """x (n_seq*seq_len, D), ego (n_seq, K, D): attention of every token over K IMU keys (JointEncoder)."""
        d = x.shape[1]
        pq = _lin(pos_bf, self.query_pos)
        q = _lin(x, self.query_content, residual=pq, res_row_mod=_pos_mod(pos_bf, seq_len))
        k = _lin(_rows(ego), self.key).view(n_seq, -1, d)
        v = _lin(_rows(ego), self.value).view(n_seq, -1, d)
        qkv_free = A.PlainAttnFn.apply(q.view(n_seq, seq_len, d), k, v, self.heads, (d // self.heads) ** -0.5,
                                       self._site())
        o = _lin(qkv_free.view(-1, d), self.fun.out_proj)
        return self._post(o)


class TransformerEncoderLayer(nn.Module):
    """transformer.py:422-487."""

    def __init__(self, D, Nhead, Dff, dropout=0.1, use_egodeep=False):
        super().__init__()
        self.self_attn = EncoderAttention(D, Nhead, Dff, dropout)
        self.use_egodeep = use_egodeep
        self.p = dropout
        if use_egodeep:
            self.egodeep_attend = EgodeepAttention(D, Nhead, Dff, dropout)
            self.norm_eda = nn.LayerNorm(D)

    def forward(self, x, pos_bf, ego, n_seq, seq_len):
        x = self.self_attn(x, pos_bf, n_seq, seq_len)
        if self.use_egodeep and ego is not None:
            drop = self.training and self.p > 0
            if ego.dim() == 2 and not drop:
                o = self.egodeep_attend.forward_single_key(ego)
                x = A.add_ln(x, self.norm_eda.weight, self.norm_eda.bias, bcast=o, rows_per_group=seq_len,
                             eps=self.norm_eda.eps)
            else:
                # With dropout on, the masks differ per token (attention weights, transformer.py:92; attn_dropout and
                # the MLP, :95-103; dropout_eda, :485), so the one-key collapse no longer holds: every token runs the
                # general path, as in the reference.
                if ego.dim() == 2:
                    o = self.egodeep_attend.forward_single_key_dropout(ego, n_seq, seq_len)
                else:
                    o = self.egodeep_attend.forward_multi_key(x, pos_bf, ego, n_seq, seq_len)
                site = ops.drop_site(self.p, self.training)
                if site is not None:  # features + dropout_eda(new) -> norm_eda in one kernel
                    x = A.add_ln_drop(x, o, self.norm_eda.weight, self.norm_eda.bias, eps=self.norm_eda.eps, drop=site)
                else:
                    x = A.add_ln(x + o, self.norm_eda.weight, self.norm_eda.bias, eps=self.norm_eda.eps)
        return x


class TransformerEncoder(nn.Module):
    """transformer.py:490-512."""

    def __init__(self, layers):
        super().__init__()
        self.layers = layers
        _xavier_(self)

    def forward(self, x, pos_bf, ego, n_seq, seq_len):
        for layer in self.layers:
            x = layer(x, pos_bf, ego, n_seq, seq_len)
        return x
