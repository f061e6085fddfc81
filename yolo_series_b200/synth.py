"""Deterministic synthetic weights and inputs (there are no checkpoints or datasets offline).
Shared by bench.py, the tests and the golden-vector generator.  Everything is a pure function of
integer seeds on the CPU torch generator, so every machine and every implementation sees the same
bits.  No compute of the hot path lives here.
"""
import math

import torch


def seeded_state_dict(model, seed=0, head_obj_shift=4.0, head_cls_shift=3.0):
    """A non-trivial state_dict for ``model`` (this package's Model or the reference's): conv weights
    ~ N(0, 1/fan_in)-scaled so activations stay O(1) through ~100 layers, BN statistics randomised so
    the fold is non-trivial, and Detect biases shifted so a controlled fraction of anchors clears
    conf_thres (random init gives zero detections -- SURVEY.md section 0 item 3)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    full = model.state_dict()
    for k, v in full.items():
        if not torch.is_floating_point(v):
            sd[k] = v.clone()
            continue
        if k.endswith('anchors') or k.endswith('anchor_grid'):
            sd[k] = v.clone()
        elif k.endswith('running_mean'):
            sd[k] = torch.randn(v.shape, generator=g) * 0.1
        elif k.endswith('running_var'):
            sd[k] = torch.rand(v.shape, generator=g) * 0.5 + 0.75
        elif k.rsplit('.', 1)[0] + '.running_mean' in full:
            # BN affine: gamma ~ U(0.8,1.2), beta ~ N(0,0.1)
            if k.endswith('weight'):
                sd[k] = torch.rand(v.shape, generator=g) * 0.4 + 0.8
            else:
                sd[k] = torch.randn(v.shape, generator=g) * 0.1
        elif k.endswith('implicit'):
            base = 1.0 if '.im.' in k else 0.0
            sd[k] = base + torch.randn(v.shape, generator=g) * 0.02
        elif v.dim() == 4:
            fan_in = v.shape[1] * v.shape[2] * v.shape[3]
            sd[k] = torch.randn(v.shape, generator=g) * math.sqrt(2.0 / fan_in)
        elif k.endswith('bias'):   # head conv biases (viewed as (na, no))
            b = torch.randn(v.shape, generator=g) * 0.5
            na = 3
            bb = b.view(na, -1)
            bb[:, 4] += -head_obj_shift
            bb[:, 5:] += -head_cls_shift
            sd[k] = bb.view(-1)
        else:
            sd[k] = torch.randn(v.shape, generator=g) * 0.1
    return sd


def synth_image(B, H, W, seed=2, ch=3, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    if dtype == torch.uint8:
        return torch.randint(0, 256, (B, ch, H, W), generator=g, dtype=torch.uint8)
    return torch.rand((B, ch, H, W), generator=g)


def synth_pred(B, A, nc, seed=0, p_obj=0.3, clusters=(40, 80), extent=640.0, box_wh=(40.0, 90.0), jitter=8.0,
               unique_scores=True):
    """Clustered pedestrian-like predictions (B, A, nc+5) fp32."""
    g = torch.Generator().manual_seed(seed)
    out = torch.empty((B, A, nc + 5))
    for b in range(B):
        K = int(torch.randint(clusters[0], clusters[1] + 1, (1,), generator=g))
        centres = torch.rand((K, 2), generator=g) * extent
        idx = torch.randint(0, K, (A,), generator=g)
        xy = centres[idx] + torch.randn((A, 2), generator=g) * jitter
        wh = (torch.rand((A, 2), generator=g) * 0.5 + 0.75) * torch.tensor(box_wh)
        is_obj = torch.rand((A,), generator=g) < p_obj
        obj = torch.where(is_obj, torch.rand((A,), generator=g) * 0.9 + 0.1, torch.rand((A,), generator=g) * 9e-4)
        cls = torch.rand((A, nc), generator=g) * 0.05
        hot = torch.randint(0, nc, (A,), generator=g)
        cls[torch.arange(A), hot] = torch.rand((A,), generator=g) * 0.8 + 0.2
        if unique_scores:  # distinct obj per row => distinct best-class confidences with overwhelming probability
            obj = obj + torch.arange(A).float() * 1e-9
        out[b, :, :2], out[b, :, 2:4], out[b, :, 4], out[b, :, 5:] = xy, wh, obj, cls
    return out


def synth_pred_wide(B, A, nc, seed=0, mixed=False, extent=1280.0):
    """Predictions whose boxes are 6 000-9 000 px wide AND tall (centres inside ``extent``, so x1 / y1 go negative):
    the reference adds ``cls * 4096`` to all four coordinates (utils/general.py:702-704), so such boxes of
    NEIGHBOURING classes overlap after the offset, i.e. suppression crosses classes (IoU ~0.05-0.2).
    ``mixed``: half of the rows are narrow 40x90 boxes of classes 0-2 in 60 tight clusters with the highest scores
    (several 256-candidate chunks of the sorted order, few survivors) and only the lower-scoring rows are wide -- the
    x extent of what has been seen crosses 4 000 px in the middle of an image."""
    g = torch.Generator().manual_seed(seed)
    out = torch.empty((B, A, nc + 5))
    for b in range(B):
        x = torch.rand((A,), generator=g) * extent
        y = torch.rand((A,), generator=g) * extent
        w = torch.rand((A,), generator=g) * 3000.0 + 6000.0
        h = torch.rand((A,), generator=g) * 3000.0 + 6000.0
        obj = torch.rand((A,), generator=g) * 0.7 + 0.3
        hot = torch.randint(0, nc, (A,), generator=g)
        if mixed:
            narrow = torch.rand((A,), generator=g) < 0.5
            cxy = torch.rand((60, 2), generator=g) * extent
            pick = torch.randint(0, 60, (A,), generator=g)
            jit = torch.randn((A, 2), generator=g) * 2.0
            x = torch.where(narrow, cxy[pick, 0] + jit[:, 0], x)
            y = torch.where(narrow, cxy[pick, 1] + jit[:, 1], y)
            w = torch.where(narrow, torch.rand((A,), generator=g) * 20.0 + 30.0, w)
            h = torch.where(narrow, torch.rand((A,), generator=g) * 45.0 + 68.0, h)
            obj = torch.where(narrow, obj * 0.4 + 0.6, obj * 0.5)       # narrow rows score higher than every wide row
            hot = torch.where(narrow, torch.randint(0, min(3, nc), (A,), generator=g), hot)
        cls = torch.rand((A, nc), generator=g) * 0.05
        cls[torch.arange(A), hot] = torch.rand((A,), generator=g) * 0.1 + 0.9
        obj = obj + torch.arange(A).float() * 1e-9
        out[b, :, 0], out[b, :, 1], out[b, :, 2], out[b, :, 3], out[b, :, 4], out[b, :, 5:] = x, y, w, h, obj, cls
    return out
