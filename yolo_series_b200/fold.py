"""Load-time algebraic folding: every conv-like block becomes one (weight, bias, k, s, act) tuple.

Works on duck-typed modules, so it accepts both this package's containers (modules.py) and
instances of the reference's own classes (an unpickled checkpoint), fused or not.

Reference formulas followed:
  Conv+BN          utils/torch_utils.py:181-201   W' = diag(g/sqrt(var+eps)) W ; b' = beta - g*mu/sqrt(var+eps)
  RepConv          models/common.py:561-643       fold BN into 3x3 and 1x1, zero-pad 1x1 -> 3x3, BN-identity
                                                  branch as an identity 3x3 kernel, sum
  ImplicitA/M      models/yolo.py:178-190         b += W ia ; b *= im ; W *= im^T
All arithmetic is fp32 on the CPU and runs under torch.no_grad().
"""
import torch
import torch.nn as nn

from ._lib import ACT_LEAKY, ACT_NONE, ACT_SILU


class FoldedConv:
    """One conv primitive after folding: y = act(conv2d(x, weight, bias, stride, pad=k//2))."""
    __slots__ = ('weight', 'bias', 'k', 's', 'act', 'slope')

    def __init__(self, weight, bias, k, s, act=ACT_NONE, slope=0.0):
        self.weight, self.bias, self.k, self.s, self.act, self.slope = weight, bias, k, s, act, slope

    @property
    def c_in(self):
        return self.weight.shape[1]

    @property
    def c_out(self):
        return self.weight.shape[0]


def act_code(act):
    """Map an activation module to (code, slope)."""
    if act is None or isinstance(act, nn.Identity):
        return ACT_NONE, 0.0
    if isinstance(act, nn.SiLU) or type(act).__name__ == 'SiLU':
        return ACT_SILU, 0.0
    if isinstance(act, nn.LeakyReLU):
        return ACT_LEAKY, float(act.negative_slope)
    raise NotImplementedError(f'activation {type(act).__name__} is not on the yolov7 hot path')


def _f(t):
    """fp32 CPU copy (folding is host-side, whatever device the module lives on)."""
    return t.detach().float().cpu()


def _bn_scale_shift(bn):
    std = torch.sqrt(_f(bn.running_var) + bn.eps)
    gamma = _f(bn.weight) if bn.weight is not None else torch.ones_like(std)
    beta = _f(bn.bias) if bn.bias is not None else torch.zeros_like(std)
    scale = gamma / std
    return scale, beta - _f(bn.running_mean) * scale


def _conv_geometry(conv):
    k = conv.kernel_size[0] if isinstance(conv.kernel_size, tuple) else conv.kernel_size
    s = conv.stride[0] if isinstance(conv.stride, tuple) else conv.stride
    p = conv.padding[0] if isinstance(conv.padding, tuple) else conv.padding
    if conv.groups != 1 or k not in (1, 3) or s not in (1, 2) or p != k // 2:
        raise NotImplementedError(f'conv k={k} s={s} p={p} g={conv.groups} is outside the yolov7 hot path')
    return k, s


@torch.no_grad()
def fold_conv_bn(conv, bn):
    """(W', b') of conv followed by eval-mode BatchNorm; bn may be None (already fused module)."""
    w = _f(conv.weight)
    b = _f(conv.bias) if conv.bias is not None else torch.zeros(w.shape[0])
    if bn is None:
        return w.clone(), b.clone()
    scale, shift = _bn_scale_shift(bn)
    return w * scale.view(-1, 1, 1, 1), b * scale + shift


@torch.no_grad()
def fold_conv_block(m):
    """``Conv`` (common.py:99-111): fused (no .bn / forward swapped) or unfused."""
    k, s = _conv_geometry(m.conv)
    bn = getattr(m, 'bn', None)
    if bn is not None and m.conv.bias is not None and not hasattr(bn, 'running_var'):
        bn = None
    w, b = fold_conv_bn(m.conv, bn)
    a, sl = act_code(m.act)
    return FoldedConv(w, b, k, s, a, sl)


@torch.no_grad()
def fold_repconv(m):
    """``RepConv`` (common.py:463-643) -> a single 3x3 conv."""
    a, sl = act_code(m.act)
    if hasattr(m, 'rbr_reparam'):
        k, s = _conv_geometry(m.rbr_reparam)
        return FoldedConv(_f(m.rbr_reparam.weight).clone(), _f(m.rbr_reparam.bias).clone(), k, s, a, sl)
    dense_conv, dense_bn = m.rbr_dense[0], m.rbr_dense[1]
    one_conv, one_bn = m.rbr_1x1[0], m.rbr_1x1[1]
    s = dense_conv.stride[0]
    w3, b3 = fold_conv_bn(dense_conv, dense_bn)
    w1, b1 = fold_conv_bn(one_conv, one_bn)
    w = w3 + torch.nn.functional.pad(w1, [1, 1, 1, 1])
    b = b3 + b1
    ident = getattr(m, 'rbr_identity', None)
    if ident is not None:
        if isinstance(ident, nn.BatchNorm2d) or hasattr(ident, 'running_var'):
            scale, shift = _bn_scale_shift(ident)
            c = scale.numel()
            wi = torch.zeros(c, c, 3, 3)
            wi[torch.arange(c), torch.arange(c), 1, 1] = scale
            w, b = w + wi, b + shift
        else:  # reference fuse_repvgg_block turns it into a conv (common.py:608-626)
            wi, bi = fold_conv_bn(ident, None)
            w, b = w + torch.nn.functional.pad(wi, [1, 1, 1, 1]) if wi.shape[-1] == 1 else w + wi, b + bi
    return FoldedConv(w, b, 3, s, a, sl)


@torch.no_grad()
def fold_head(head, fused):
    """1x1 output convs of Detect / IDetect / IAuxDetect with ImplicitA/ImplicitM folded in.

    ``fused`` says the module already went through the reference's ``fuse()`` (implicit layers
    already folded into ``m``); otherwise ``ia``/``im`` are folded here.  Only the first ``nl``
    convs are live (IAuxDetect.fuseforward, yolo.py:364-399).
    """
    out = []
    for i in range(head.nl):
        conv = head.m[i]
        w = _f(conv.weight).clone()
        b = _f(conv.bias).clone()
        if not fused and hasattr(head, 'ia'):
            ia = _f(head.ia[i].implicit).view(-1)
            im = _f(head.im[i].implicit).view(-1)
            b = b + w.view(w.shape[0], -1) @ ia
            b = b * im
            w = w * im.view(-1, 1, 1, 1)
        out.append(FoldedConv(w, b, 1, 1, ACT_NONE, 0.0))
    return out
