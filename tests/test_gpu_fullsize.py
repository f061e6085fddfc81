"""Parity at BASELINE.json's FULL sizes through size-independent properties (the oracle takes minutes there):
batch independence, fused == unfused, bf16 vs the on-device fp32 reference mode, NMS invariants (descending scores,
no surviving pair above the IoU threshold within a class, survivors are candidates), permutation invariance, and an
oracle spot check on a bounded sample."""
import numpy as np
import pytest
import torch

from oracle.nms_ref import non_max_suppression_ref
from oracle.synth import seeded_state_dict, synth_image

pytestmark = pytest.mark.gpu
CONF, IOU = 0.001, 0.65


def _build(name, mode='bf16'):
    from yolo_series_b200 import Model
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m.set_precision(mode)
    return m.cuda().eval()


def _check_nms_invariants(out, counts, iou_thres, max_det, agnostic=False):
    """out (B, max_det, 6), counts (B,) on the device."""
    from yolo_series_b200.evaluate import box_iou
    cnt = counts.tolist()
    for b, n in enumerate(cnt):
        assert 0 <= n <= max_det
        d = out[b, :n]
        if n < 2:
            continue
        assert bool((d[1:, 4] <= d[:-1, 4]).all()), 'scores must be descending'
        off = d[:, :4] + (0 if agnostic else d[:, 5:6] * 4096.0)
        iou = box_iou(off.contiguous(), off.contiguous())
        iou = iou - torch.eye(n, device=iou.device) * 2
        assert float(iou.max()) <= iou_thres + 1e-6, f'image {b}: kept pair with IoU {float(iou.max())}'


def test_config2_yolov7_640_b32_properties():
    """BASELINE configs[1]: yolov7 640x640 batch 32 bf16, conf 0.001 / iou 0.65."""
    from yolo_series_b200 import non_max_suppression
    m = _build('yolov7')
    x = synth_image(32, 640, 640, seed=100, dtype=torch.uint8).cuda()
    out, meta = m.detect(x, CONF, IOU, packed=True)
    out, meta = out.clone(), meta.clone()
    assert int(meta[-1]) == 0
    _check_nms_invariants(out, meta[:-1], IOU, 300)
    # fused decode+NMS == forward -> stand-alone NMS (bit for bit)
    pred, raws = m(x)
    assert pred.shape == (32, 25200, 85) and len(raws) == 3
    dets = non_max_suppression(pred, CONF, IOU)
    for i, d in enumerate(dets):
        assert torch.equal(d, out[i, :int(meta[i])]), i
    # images are independent units: image 7 / 31 alone == inside the batch
    for i in (7, 31):
        o1, m1 = m.detect(x[i:i + 1], CONF, IOU, packed=True)
        assert int(m1[0]) == int(meta[i]) and torch.equal(o1[0, :int(m1[0])], out[i, :int(meta[i])]), i
    # oracle spot check of the NMS stage on two images of the real head output
    ref = non_max_suppression_ref(pred[:2].cpu(), CONF, IOU)
    for i in range(2):
        assert np.array_equal(dets[i].cpu().numpy(), ref[i].numpy())
    # multi_label=True variant (test.py:126): fused == unfused on the whole batch, oracle on one image
    om, mm = m.detect(x, CONF, IOU, multi_label=True, packed=True)
    assert int(mm[-1]) == 0
    _check_nms_invariants(om, mm[:-1], IOU, 300)
    dm = non_max_suppression(pred, CONF, IOU, multi_label=True)
    for i, d in enumerate(dm):
        assert torch.equal(d, om[i, :int(mm[i])]), i
    refm = non_max_suppression_ref(pred[5:6].cpu(), CONF, IOU, multi_label=True)[0]
    assert np.array_equal(dm[5].cpu().numpy(), refm.numpy())
    # bf16 trunk vs the fp32 reference mode on the device (stated bf16 tolerance, tests/test_gpu_e2e.py), 8 images
    m32 = _build('yolov7', 'fp32')
    p32 = m32(x[:8])[0]
    size = p32[..., 2:4].clamp(min=1.0).max(-1, keepdim=True)[0]
    rel = ((pred[:8, :, :4] - p32[..., :4]).abs() / (0.02 * size + 1.0)).flatten()
    sc = (pred[:8, :, 4:] - p32[..., 4:]).abs().flatten()
    idx = torch.randperm(rel.numel(), device=rel.device)[:2_000_000]
    assert float(rel[idx].quantile(0.99)) <= 1.0
    idx = torch.randperm(sc.numel(), device=sc.device)[:2_000_000]
    assert float(sc[idx].quantile(0.99)) <= 0.03


@pytest.mark.parametrize('name,B', [('yolov7-w6-custom', 8), ('yolov7-e6e', 8)])
def test_config34_p6_1280_properties(name, B):
    """BASELINE configs[2]/[3] per-GPU shards (w6-custom 1280 B=16 -> 8/GPU at 2 GPUs; e6e B=64 -> 8/GPU at 8)."""
    from yolo_series_b200 import non_max_suppression
    m = _build(name)
    x = synth_image(B, 1280, 1280, seed=3, dtype=torch.uint8).cuda()
    out, meta = m.detect(x, CONF, IOU, packed=True)
    out, meta = out.clone(), meta.clone()
    assert int(meta[-1]) == 0
    _check_nms_invariants(out, meta[:-1], IOU, 300)
    pred = m(x)[0]
    head = m.model[-1]
    assert pred.shape == (B, 102000, head.no)
    dets = non_max_suppression(pred, CONF, IOU)
    for i, d in enumerate(dets):
        assert torch.equal(d, out[i, :int(meta[i])]), i
    o1, m1 = m.detect(x[B - 1:B], CONF, IOU, packed=True)
    assert torch.equal(o1[0, :int(m1[0])], out[B - 1, :int(meta[B - 1])])
    ref = non_max_suppression_ref(pred[:1].cpu(), CONF, IOU)[0]
    assert np.array_equal(dets[0].cpu().numpy(), ref.numpy())


def test_config5_nms_stress_full_size_properties():
    """BASELINE configs[4]: (256, 100 000, 85) fp32 = 8.7 GB, max_det 1000.  Invariants on every image, permutation
    invariance of the row order (scores are distinct), oracle on a bounded sample of 2 images."""
    import bench
    from yolo_series_b200.nms import _run_nms
    dev = torch.device('cuda', torch.cuda.current_device())
    B, A, no = 256, 100000, 85
    pred = _stress_pred(dev, B, A)
    # make every candidate score distinct (the generator's obj*cls products collide ~50 times per image, and equal
    # scores are ordered by row index, which a permutation legitimately changes): hot class -> 1, obj -> distinct values
    g = torch.Generator(device=dev).manual_seed(1)
    for b0 in range(0, B, 32):
        blk = pred[b0:b0 + 32]
        cls = blk[..., 5:]
        cls[cls > 0.15] = 1.0
        rank = torch.rand((blk.shape[0], A), device=dev, generator=g).argsort(1).float()
        obj = blk[..., 4]
        blk[..., 4] = torch.where(obj > 0.05, (rank + 1.0) / (A + 1.0) * 0.9 + 0.1, obj)
    run = lambda p: _run_nms(p, B, A, no, CONF, IOU, None, False, False, 30000, 1000, dev, return_packed=True)
    out, meta = run(pred)
    out, meta = out.clone(), meta.clone()
    assert int(meta[-1]) == 0
    _check_nms_invariants(out[:32], meta[:32], IOU, 1000)
    cnt = meta[:B]
    assert int(cnt.min()) >= 1 and int(cnt.max()) <= 1000
    valid = torch.arange(1000, device=dev)[None, :] < cnt[:, None]
    conf = out[..., 4]
    assert bool(((conf[:, 1:] <= conf[:, :-1]) | ~valid[:, 1:]).all())
    assert bool((conf[valid] > CONF).all())
    # survivors are rows of the input: xyxy -> xywh centre must be one of the image's rows (checked via the score)
    obj_cls = (pred[:4, :, 4:5] * pred[:4, :, 5:]).max(2)[0]
    for b in range(4):
        n = int(cnt[b])
        assert bool(torch.isin(out[b, :n, 4], obj_cls[b]).all())
    del obj_cls
    # permutation invariance: a different row order gives the same detections
    perm = torch.randperm(A, device=dev)
    for b0 in range(0, B, 32):
        pred[b0:b0 + 32] = pred[b0:b0 + 32][:, perm]
    out2, meta2 = run(pred)
    assert torch.equal(meta2[:B], cnt)
    assert torch.equal(torch.where(valid[..., None], out2, torch.zeros_like(out2)),
                       torch.where(valid[..., None], out, torch.zeros_like(out)))
    # oracle on a bounded sample (the permuted rows of images 0 and 255)
    sample = torch.stack((pred[0], pred[255])).cpu()
    ref = non_max_suppression_ref(sample, CONF, IOU, max_det=1000)
    for r, b in zip(ref, (0, 255)):
        assert np.array_equal(out2[b, :int(cnt[b])].cpu().numpy(), r.numpy())


def _stress_pred(dev, B, A, nc=80, K=600):
    """Same generator as bench.nms_stress, scores distinct per image."""
    g = torch.Generator(device=dev).manual_seed(0)
    pred = torch.empty((B, A, nc + 5), device=dev)
    for b0 in range(0, B, 32):
        nb = min(32, B - b0)
        centres = torch.rand((nb, K, 2), device=dev, generator=g) * 1280
        idx = torch.randint(0, K, (nb, A), device=dev, generator=g)
        xy = torch.gather(centres, 1, idx.unsqueeze(-1).expand(-1, -1, 2)) + torch.randn((nb, A, 2), device=dev, generator=g) * 8
        wh = (torch.rand((nb, A, 2), device=dev, generator=g) * 0.5 + 0.75) * torch.tensor([40., 90.], device=dev)
        is_obj = torch.rand((nb, A), device=dev, generator=g) < 0.3
        obj = torch.where(is_obj, torch.rand((nb, A), device=dev, generator=g) * 0.9 + 0.1,
                          torch.rand((nb, A), device=dev, generator=g) * 9e-4)
        cls = torch.rand((nb, A, nc), device=dev, generator=g) * 0.05
        hot = torch.randint(0, nc, (nb, A, 1), device=dev, generator=g)
        cls.scatter_(2, hot, torch.rand((nb, A, 1), device=dev, generator=g) * 0.8 + 0.2)
        pred[b0:b0 + nb, :, :2], pred[b0:b0 + nb, :, 2:4], pred[b0:b0 + nb, :, 4], pred[b0:b0 + nb, :, 5:] = xy, wh, obj, cls
        del cls
    return pred
