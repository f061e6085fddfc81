"""ORACLE (test infrastructure, NOT product code): CPU fp32 restatement of the reference's YOLOv7
forward pass.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product path (yolo_series_b200/) never does.

Parity status: PINNED -- tests/test_oracle_golden.py checks this restatement against outputs of the
real reference (models.yolo.Model imported from /root/reference, fused, fp32 CPU) stored in
tests/golden/ by oracle/make_golden.py.

What is restated (file:line in /root/reference):
  forward_once interpreter            models/yolo.py:601-631
  Conv / fused Conv                   models/common.py:99-111, utils/torch_utils.py:181-201
  RepConv (3 branches | reparam)      models/common.py:498-507, 584-643
  MP / SP / ReOrg / Concat / Shortcut models/common.py:30-86
  DownC / SPPCSPC                     models/common.py:181-192, 262-280
  IDetect / IAuxDetect / Detect       models/yolo.py:42-77, 119-176, 334-399  (+ fuse 178-190)
The function takes the cfg dict and an *unfused* state_dict (reference key names) and applies the
same fold the reference's Model.fuse() applies when ``fused=True`` (the detect.py / test.py path).
"""
import math

import torch
import torch.nn.functional as F

BN_EPS = 1e-3  # utils/torch_utils.py:150


def _act(x, spec):
    if spec is None or spec == 'silu':
        return F.silu(x)
    if spec == 'identity':
        return x
    if isinstance(spec, tuple) and spec[0] == 'leaky':
        return F.leaky_relu(x, spec[1])
    raise ValueError(spec)


def _act_spec(args):
    """Conv args after [c2, k, s]: p, g, act (common.py:101)."""
    if len(args) >= 6:
        a = args[5]
        if isinstance(a, str) and a.startswith('nn.LeakyReLU'):
            return ('leaky', float(a[a.index('(') + 1:a.rindex(')')]))
        if a is False or a == 'False':
            return 'identity'
    return 'silu'


class _Weights:
    def __init__(self, sd, fused):
        self.sd, self.fused = sd, fused

    def bn(self, p):
        g, b = self.sd[p + '.weight'].float(), self.sd[p + '.bias'].float()
        m, v = self.sd[p + '.running_mean'].float(), self.sd[p + '.running_var'].float()
        return g, b, m, v

    def conv_bn(self, x, p_conv, p_bn, stride, pad):
        w = self.sd[p_conv + '.weight'].float()
        g, b, m, v = self.bn(p_bn)
        if self.fused:  # fuse_conv_and_bn: W' = diag(g/sqrt(v+eps)) W ; b' = b - g m / sqrt(v+eps)
            scale = g / torch.sqrt(v + BN_EPS)
            return F.conv2d(x, w * scale.view(-1, 1, 1, 1), b - m * scale, stride, pad)
        return F.batch_norm(F.conv2d(x, w, None, stride, pad), m, v, g, b, False, 0.0, BN_EPS)

    def conv_block(self, x, p, k, s, act):
        return _act(self.conv_bn(x, p + '.conv', p + '.bn', s, k // 2), act)


def _repconv(W, x, p, s):
    if p + '.rbr_reparam.weight' in W.sd:
        return F.silu(F.conv2d(x, W.sd[p + '.rbr_reparam.weight'].float(), W.sd[p + '.rbr_reparam.bias'].float(), s, 1))
    y = W.conv_bn(x, p + '.rbr_dense.0', p + '.rbr_dense.1', s, 1) + W.conv_bn(x, p + '.rbr_1x1.0', p + '.rbr_1x1.1', s, 0)
    if p + '.rbr_identity.weight' in W.sd:
        g, b, m, v = W.bn(p + '.rbr_identity')
        y = y + F.batch_norm(x, m, v, g, b, False, 0.0, BN_EPS)
    return F.silu(y)


def _head(W, xs, p, nc, anchors, strides, kind):
    """Returns (pred, raw list).  anchors: per-level pixel anchors (anchor_grid)."""
    nl, na, no = len(anchors), len(anchors[0]) // 2, nc + 5
    z, raws = [], []
    for i in range(nl):
        w = W.sd[f'{p}.m.{i}.weight'].float()
        b = W.sd[f'{p}.m.{i}.bias'].float()
        x = xs[i]
        if kind in ('IDetect', 'IAuxDetect'):
            ia = W.sd[f'{p}.ia.{i}.implicit'].float()
            im = W.sd[f'{p}.im.{i}.implicit'].float()
            if W.fused:  # IDetect.fuse yolo.py:178-190
                b = (b + (w.view(w.shape[0], -1) @ ia.view(-1))) * im.view(-1)
                w = w * im.view(-1, 1, 1, 1)
                x = F.conv2d(x, w, b)
            else:
                x = F.conv2d(x + ia, w, b) * im
        else:
            x = F.conv2d(x, w, b)
        bs, _, ny, nx = x.shape
        x = x.view(bs, na, no, ny, nx).permute(0, 1, 3, 4, 2).contiguous()
        raws.append(x)
        yv, xv = torch.meshgrid(torch.arange(ny), torch.arange(nx), indexing='ij')
        grid = torch.stack((xv, yv), 2).view(1, 1, ny, nx, 2).float()
        ag = torch.tensor(anchors[i]).float().view(1, na, 1, 1, 2)
        y = x.sigmoid()
        y[..., 0:2] = (y[..., 0:2] * 2. - 0.5 + grid) * strides[i]
        y[..., 2:4] = (y[..., 2:4] * 2) ** 2 * ag
        z.append(y.view(bs, -1, no))
    return torch.cat(z, 1), raws


@torch.no_grad()
def ref_forward(cfg, sd, x, fused=True, return_layers=False):
    """cfg: dict (zoo / yaml); sd: unfused state_dict with reference key names; x: (B,3,H,W) fp32 in [0,1]."""
    W = _Weights(sd, fused)
    rows = cfg['backbone'] + cfg['head']
    nc, anchors = cfg['nc'], cfg['anchors']
    ys, red = [], []
    x = x.float()
    H_in = x.shape[2]
    for i, (f, n, mname, args) in enumerate(rows):
        srcs = [i + j if j < 0 else j for j in ([f] if isinstance(f, int) else f)]
        xin = [x if i == 0 else ys[s] for s in srcs]
        p = f'model.{i}'
        if mname == 'Conv':
            y = W.conv_block(xin[0], p, args[1], args[2], _act_spec(args))
        elif mname == 'RepConv':
            y = _repconv(W, xin[0], p, args[2])
        elif mname == 'Concat':
            y = torch.cat(xin, 1)
        elif mname == 'MP':
            y = F.max_pool2d(xin[0], 2, 2)
        elif mname == 'SP':
            y = F.max_pool2d(xin[0], args[0], 1, args[0] // 2)
        elif mname in ('nn.Upsample', 'Upsample'):
            y = F.interpolate(xin[0], scale_factor=2, mode='nearest')
        elif mname == 'ReOrg':
            a = xin[0]
            y = torch.cat([a[..., ::2, ::2], a[..., 1::2, ::2], a[..., ::2, 1::2], a[..., 1::2, 1::2]], 1)
        elif mname == 'Shortcut':
            y = xin[0] + xin[1]
        elif mname == 'DownC':
            a = W.conv_block(W.conv_block(xin[0], p + '.cv1', 1, 1, 'silu'), p + '.cv2', 3, 2, 'silu')
            b = W.conv_block(F.max_pool2d(xin[0], 2, 2), p + '.cv3', 1, 1, 'silu')
            y = torch.cat((a, b), 1)
        elif mname == 'SPPCSPC':
            x1 = W.conv_block(W.conv_block(W.conv_block(xin[0], p + '.cv1', 1, 1, 'silu'), p + '.cv3', 3, 1, 'silu'),
                              p + '.cv4', 1, 1, 'silu')
            cat = torch.cat([x1] + [F.max_pool2d(x1, k, 1, k // 2) for k in (5, 9, 13)], 1)
            y1 = W.conv_block(W.conv_block(cat, p + '.cv5', 1, 1, 'silu'), p + '.cv6', 3, 1, 'silu')
            y2 = W.conv_block(xin[0], p + '.cv2', 1, 1, 'silu')
            y = W.conv_block(torch.cat((y1, y2), 1), p + '.cv7', 1, 1, 'silu')
        elif mname in ('Detect', 'IDetect', 'IAuxDetect'):
            nl = len(anchors)
            strides = [H_in / xin[l].shape[2] for l in range(nl)]
            order = list(range(nl))
            a0, a1 = anchors[0], anchors[-1]
            if math.copysign(1, a1[-2] * a1[-1] - a0[0] * a0[1]) != math.copysign(1, strides[-1] - strides[0]):
                order = order[::-1]  # check_anchor_order autoanchor.py:12-20
            pred, raws = _head(W, xin, p, nc, [anchors[o] for o in order], strides, mname)
            if return_layers:
                return pred, raws, ys
            return pred, raws
        else:
            raise NotImplementedError(mname)
        ys.append(y)
    raise ValueError('no head')
