"""Host-side parameter containers mirroring the reference's YOLOv7 building blocks.

These classes exist so that (a) ``state_dict`` keys are interchangeable with the
reference (``model.N.conv.weight``, ``model.N.rbr_dense.0.weight``,
``model.N.m.0.bias``, ``model.N.ia.0.implicit`` ...), so reference checkpoints
load unchanged, and (b) callers that poke at ``model.model[-1].nl/.na/.stride``
keep working.  They carry *no* compute: the forward pass is the
CUDA engine (engine.py).  Calling ``forward`` on a block raises.

Reference shapes followed: models/common.py:30-111 (MP/SP/ReOrg/Concat/Shortcut/
Conv), 181-192 (DownC), 262-280 (SPPCSPC), 433-456 (ImplicitA/M), 463-507
(RepConv); models/yolo.py:23-41 (Detect), 97-117 (IDetect), 311-332 (IAuxDetect).
"""
import torch
import torch.nn as nn


class _NoCompute(nn.Module):
    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError(f'{type(self).__name__} is a parameter container; run the model through '
                           'yolo_series_b200.Model.forward (CUDA engine)')


def _same_pad(k, p=None):
    return k // 2 if p is None else p


def _activation(act):
    if act is True:
        return nn.SiLU()
    return act if isinstance(act, nn.Module) else nn.Identity()


class Conv(_NoCompute):
    """conv(no bias) + BN + act; k in {1,3}, s in {1,2} in the yolov7 cfgs."""

    def __init__(self, c1, c2, k=1, s=1, p=None, g=1, act=True):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, _same_pad(k, p), groups=g, bias=False)
        self.bn = nn.BatchNorm2d(c2)
        self.act = _activation(act)


class RepConv(_NoCompute):
    """3x3 + 1x1 (+BN identity when c1==c2 and s==1) re-parameterisable conv."""

    def __init__(self, c1, c2, k=3, s=1, p=None, g=1, act=True, deploy=False):
        super().__init__()
        if k != 3 or _same_pad(k, p) != 1:
            raise ValueError('RepConv is 3x3/pad 1 only')
        self.deploy, self.groups, self.in_channels, self.out_channels = deploy, g, c1, c2
        self.act = _activation(act)
        if deploy:
            self.rbr_reparam = nn.Conv2d(c1, c2, k, s, 1, groups=g, bias=True)
        else:
            self.rbr_identity = nn.BatchNorm2d(c1) if (c1 == c2 and s == 1) else None
            self.rbr_dense = nn.Sequential(nn.Conv2d(c1, c2, 3, s, 1, groups=g, bias=False), nn.BatchNorm2d(c2))
            self.rbr_1x1 = nn.Sequential(nn.Conv2d(c1, c2, 1, s, 0, groups=g, bias=False), nn.BatchNorm2d(c2))


class MP(_NoCompute):
    def __init__(self, k=2):
        super().__init__()
        self.k = k
        self.m = nn.MaxPool2d(kernel_size=k, stride=k)


class SP(_NoCompute):
    def __init__(self, k=3, s=1):
        super().__init__()
        self.k, self.s = k, s
        self.m = nn.MaxPool2d(kernel_size=k, stride=s, padding=k // 2)


class ReOrg(_NoCompute):
    pass


class Concat(_NoCompute):
    def __init__(self, dimension=1):
        super().__init__()
        self.d = dimension


class Shortcut(_NoCompute):
    def __init__(self, dimension=0):
        super().__init__()
        self.d = dimension


class Upsample(_NoCompute):
    """nn.Upsample(None, 2, 'nearest') stand-in (no parameters)."""

    def __init__(self, size=None, scale_factor=2, mode='nearest'):
        super().__init__()
        if size is not None or int(scale_factor) != 2 or mode != 'nearest':
            raise ValueError('only 2x nearest upsampling is on the yolov7 path')
        self.scale_factor, self.mode = scale_factor, mode


class DownC(_NoCompute):
    def __init__(self, c1, c2, n=1, k=2):
        super().__init__()
        self.cv1 = Conv(c1, c1, 1, 1)
        self.cv2 = Conv(c1, c2 // 2, 3, k)
        self.cv3 = Conv(c1, c2 // 2, 1, 1)
        self.mp = nn.MaxPool2d(kernel_size=k, stride=k)


class SPPCSPC(_NoCompute):
    def __init__(self, c1, c2, n=1, shortcut=False, g=1, e=0.5, k=(5, 9, 13)):
        super().__init__()
        c_ = int(2 * c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c1, c_, 1, 1)
        self.cv3 = Conv(c_, c_, 3, 1)
        self.cv4 = Conv(c_, c_, 1, 1)
        self.m = nn.ModuleList([nn.MaxPool2d(kernel_size=x, stride=1, padding=x // 2) for x in k])
        self.cv5 = Conv(4 * c_, c_, 1, 1)
        self.cv6 = Conv(c_, c_, 3, 1)
        self.cv7 = Conv(2 * c_, c2, 1, 1)


class ImplicitA(_NoCompute):
    def __init__(self, channel, mean=0., std=.02):
        super().__init__()
        self.channel = channel
        self.implicit = nn.Parameter(torch.zeros(1, channel, 1, 1).normal_(mean, std))


class ImplicitM(_NoCompute):
    def __init__(self, channel, mean=1., std=.02):
        super().__init__()
        self.channel = channel
        self.implicit = nn.Parameter(torch.ones(1, channel, 1, 1).normal_(mean, std))


class _Head(_NoCompute):
    stride = None
    export = False
    end2end = False
    include_nms = False
    concat = False

    def _common(self, nc, anchors):
        self.nc, self.no = nc, nc + 5
        self.nl, self.na = len(anchors), len(anchors[0]) // 2
        a = torch.tensor(anchors).float().view(self.nl, -1, 2)
        self.register_buffer('anchors', a)
        self.register_buffer('anchor_grid', a.clone().view(self.nl, 1, -1, 1, 1, 2))


class Detect(_Head):
    def __init__(self, nc=80, anchors=(), ch=()):
        super().__init__()
        self._common(nc, anchors)
        self.m = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch)


class IDetect(_Head):
    def __init__(self, nc=80, anchors=(), ch=()):
        super().__init__()
        self._common(nc, anchors)
        self.m = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch)
        self.ia = nn.ModuleList(ImplicitA(x) for x in ch)
        self.im = nn.ModuleList(ImplicitM(self.no * self.na) for _ in ch)


class IAuxDetect(_Head):
    def __init__(self, nc=80, anchors=(), ch=()):
        super().__init__()
        self._common(nc, anchors)
        nl = self.nl
        self.m = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch[:nl])
        self.m2 = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch[nl:])
        self.ia = nn.ModuleList(ImplicitA(x) for x in ch[:nl])
        self.im = nn.ModuleList(ImplicitM(self.no * self.na) for _ in ch[:nl])


HEADS = (Detect, IDetect, IAuxDetect)
CONV_LIKE = {'Conv': Conv, 'RepConv': RepConv, 'DownC': DownC, 'SPPCSPC': SPPCSPC}
REGISTRY = {
    **CONV_LIKE, 'MP': MP, 'SP': SP, 'ReOrg': ReOrg, 'Concat': Concat, 'Shortcut': Shortcut,
    'nn.Upsample': Upsample, 'Upsample': Upsample, 'Detect': Detect, 'IDetect': IDetect, 'IAuxDetect': IAuxDetect,
}
