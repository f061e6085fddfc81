"""GPU end-to-end parity: Model.forward / detect (CUDA engine) vs the reference's golden outputs
(tests/golden/fwd_*.npz) and the oracle.  Tolerances (BASELINE.json north_star):
  fp32 mode: boxes <= 1e-2 px, scores <= 1e-3 (absolute)
  bf16 mode: stated below from measurement (bf16 activations through ~60-250 conv layers)."""
import numpy as np
import pytest
import torch

from oracle.make_golden import E2E_CASE, FWD_CASES
from oracle.nms_ref import non_max_suppression_ref
from oracle.ref_forward import ref_forward
from oracle.synth import seeded_state_dict, synth_image

pytestmark = pytest.mark.gpu


def _build(name, mode):
    from yolo_series_b200 import Model
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m.set_precision(mode)
    return m.cuda().eval()


def _match_dets(got, ref, box_tol, conf_tol):
    """Row-by-row agreement after ordering by conf (both are already in descending conf)."""
    n = min(got.shape[0], ref.shape[0])
    ok_box = (got[:n, :4] - ref[:n, :4]).abs().max(1)[0] <= box_tol
    ok_conf = (got[:n, 4] - ref[:n, 4]).abs() <= conf_tol
    ok_cls = got[:n, 5] == ref[:n, 5]
    return float((ok_box & ok_conf & ok_cls).float().mean()) if n else 1.0


@pytest.mark.parametrize('case', FWD_CASES, ids=[c[0] for c in FWD_CASES])
def test_forward_fp32_mode_matches_reference(case, gold):
    from yolo_series_b200 import non_max_suppression
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    m = _build(name, 'fp32')
    pred, raws = m(synth_image(B, H, W, seed=2).cuda())
    ref = torch.from_numpy(g['pred'])
    p = pred.cpu()
    assert p.shape == ref.shape
    assert float((p[..., :4] - ref[..., :4]).abs().max()) <= 1e-2
    assert float((p[..., 4:] - ref[..., 4:]).abs().max()) <= 1e-3
    assert torch.allclose(raws[0].cpu(), torch.from_numpy(g['raw0']), rtol=1e-3, atol=1e-3)
    dets = non_max_suppression(pred, conf, iou)
    refd = torch.from_numpy(g['dets'])
    got = dets[0].cpu()
    # index-level equality modulo rows whose conf / IoU sits within rounding of a threshold
    assert abs(got.shape[0] - refd.shape[0]) <= max(2, refd.shape[0] // 50)
    assert _match_dets(got, refd, 1e-2, 1e-3) >= 0.95
    # and the NMS stage itself is bit-exact on identical input
    r = non_max_suppression_ref(pred.cpu(), conf, iou)[0]
    assert np.array_equal(got.numpy(), r.numpy())


@pytest.mark.parametrize('case', FWD_CASES, ids=[c[0] for c in FWD_CASES])
def test_forward_bf16_mode_tolerance(case, gold):
    """bf16 trunk (tcgen05), fp32 accumulate, fp32 head logits.  Stated tolerance: 99th-percentile box error
    <= 2% of the box size + 1 px, score error <= 0.03 at the 99th percentile."""
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    m = _build(name, 'bf16')
    pred, _ = m(synth_image(B, H, W, seed=2).cuda())
    ref = torch.from_numpy(g['pred'])
    p = pred.cpu()
    size = ref[..., 2:4].clamp(min=1.0).max(-1, keepdim=True)[0]
    rel = ((p[..., :4] - ref[..., :4]).abs() / (0.02 * size + 1.0)).flatten()
    sc = (p[..., 4:] - ref[..., 4:]).abs().flatten()
    print(f'bf16 {name}: box err/tol p99 {float(rel.quantile(0.99)):.3f} max {float(rel.max()):.3f}; '
          f'score err p99 {float(sc.quantile(0.99)):.4f} max {float(sc.max()):.4f}')
    assert float(rel.quantile(0.99)) <= 1.0
    assert float(sc.quantile(0.99)) <= 0.03


def test_config1_tiny_640_detect_path(gold):
    """BASELINE.json configs[0]: yolov7-tiny 1x3x640x640, conf 0.25 / iou 0.45, fp32 forward+decode+NMS."""
    from yolo_series_b200 import non_max_suppression
    name, B, H, W, conf, iou = E2E_CASE
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    m = _build(name, 'fp32')
    x = synth_image(B, H, W, seed=2).cuda()
    pred, _ = m(x)
    sub = pred[:, ::37].cpu()
    ref = torch.from_numpy(g['pred_sub'])
    assert float((sub[..., :4] - ref[..., :4]).abs().max()) <= 1e-2
    assert float((sub[..., 4:] - ref[..., 4:]).abs().max()) <= 1e-3
    dets = non_max_suppression(pred, conf, iou)[0].cpu()
    refd = torch.from_numpy(g['dets'])
    assert abs(dets.shape[0] - refd.shape[0]) <= 3
    assert _match_dets(dets, refd, 1e-2, 1e-3) >= 0.95
    fused = m.detect(x, conf, iou)[0].cpu()          # fused decode+NMS fast path == two-step path
    assert np.array_equal(fused.numpy(), dets.numpy())
    # uint8 input path (detect.py:100-104) gives the same result as float(img)/255
    u8 = synth_image(B, H, W, seed=2, dtype=torch.uint8)
    a = m(u8.cuda())[0]
    b = m((u8.float() / 255.0).cuda())[0]
    assert torch.equal(a, b)


def test_batch_consistency_and_oracle_b2():
    """B=2 forward equals two B=1 forwards (images are independent units) and the oracle."""
    from yolo_series_b200 import zoo
    m = _build('yolov7-tiny', 'fp32')
    x = synth_image(2, 96, 128, seed=5)
    p2 = m(x.cuda())[0].cpu()
    p1 = torch.cat([m(x[i:i + 1].cuda())[0].cpu() for i in range(2)], 0)
    assert torch.equal(p1, p2)
    sd = seeded_state_dict(m, seed=0)
    o, _ = ref_forward(zoo.get_cfg('yolov7-tiny'), sd, x, fused=True)
    assert float((p2[..., :4] - o[..., :4]).abs().max()) <= 1e-2
    assert float((p2[..., 4:] - o[..., 4:]).abs().max()) <= 1e-3


def test_detect_batches_stream_equals_detect():
    """The three-stream pipeline (Model.detect_batches) returns, batch by batch and in order, exactly what the
    blocking Model.detect returns; ragged last batch and a dtype change rebuild the slot buffers."""
    m = _build('yolov7-tiny', 'bf16')
    batches = [synth_image(b, 96, 128, seed=20 + i, dtype=torch.uint8).pin_memory()
               for i, b in enumerate([4, 4, 4, 4, 4, 2])]
    want = []
    for xb in batches:
        out, meta = m.detect(xb, 0.05, 0.5, max_det=50, packed=True)
        want.append((out.cpu(), meta[:-1].cpu()))
    got = []
    for out, cnt in m.detect_batches(iter(batches), 0.05, 0.5, max_det=50):
        assert out.is_pinned() and out.device.type == 'cpu'
        got.append((out.clone(), cnt.clone()))
    assert len(got) == len(want)
    for (o, c), (wo, wc) in zip(got, want):
        assert torch.equal(c, wc)
        for i in range(o.shape[0]):
            assert torch.equal(o[i, :c[i]], wo[i, :c[i]])
    assert sum(int(c.sum()) for _, c in got) > 0
    assert list(m.detect_batches(iter([]), 0.05, 0.5)) == []


def test_two_batches_in_flight_equal_serial_results():
    """Engines and NMS scratch are per stream: two streams running Model.detect at the same time (what
    detect_batches(compute_streams=2) and bench.py --streams 2 do) give, batch by batch, exactly the serial results --
    a shared activation arena or workspace would corrupt them."""
    m = _build('yolov7-tiny', 'bf16')
    xs = [synth_image(8, 256, 320, seed=40 + i, dtype=torch.uint8).cuda() for i in range(10)]
    want = []
    for x in xs:
        out, meta = m.detect(x, 0.05, 0.5, max_det=100, packed=True)
        want.append((out.clone(), meta.clone()))
    torch.cuda.synchronize()
    side = m.compute_streams(2)
    assert side[0] is m.compute_streams(2)[0] and side[0].cuda_stream != side[1].cuda_stream
    got = []
    for rep in range(3):                     # back to back, no host synchronisation between the calls
        got = []
        for i, x in enumerate(xs):
            with torch.cuda.stream(side[i & 1]):
                got.append(m.detect(x, 0.05, 0.5, max_det=100, packed=True))
    torch.cuda.synchronize()
    keys = {k[-1] for k in m._engines}
    assert {side[0].cuda_stream, side[1].cuda_stream} <= keys          # one engine per stream
    for (o, me), (wo, wm) in zip(got, want):
        assert torch.equal(me, wm) and int(me[-1]) == 0
        for i in range(o.shape[0]):
            assert torch.equal(o[i, :int(me[i])], wo[i, :int(me[i])])
    # the streaming API on one and on two compute streams
    host = [x.cpu().pin_memory() for x in xs]
    one = [(o.clone(), c.clone()) for o, c in m.detect_batches(iter(host), 0.05, 0.5, max_det=100, compute_streams=1)]
    two = [(o.clone(), c.clone()) for o, c in m.detect_batches(iter(host), 0.05, 0.5, max_det=100, compute_streams=2)]
    assert len(one) == len(two) == len(xs)
    for (o1, c1), (o2, c2), (wo, wm) in zip(one, two, want):
        assert torch.equal(c1, c2) and torch.equal(c1, wm[:-1].cpu())
        for i in range(o1.shape[0]):
            assert torch.equal(o1[i, :c1[i]], o2[i, :c1[i]]) and torch.equal(o1[i, :c1[i]], wo[i, :c1[i]].cpu())


def test_peer_gather_single_rank_equals_detect():
    """The gather fused into the NMS kernel (yb_decode_nms_gather + yb_peer_wait, dist.PeerGather) on a one-rank group:
    rows go through the peer-store path into the (self-mapped) receive buffer, both buffer sets are exercised and the
    result equals the plain fused path bit for bit.  (2 and 8 GPUs: scripts/check_dist_gpu.py, profiles/round1_dist_check.txt.)"""
    import os
    import socket

    import torch.distributed as dist

    from yolo_series_b200 import dist as ydist
    if dist.is_initialized():
        pytest.skip('a process group already exists in this process')
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    dist.init_process_group('gloo', rank=0, world_size=1)
    try:
        m = _build('yolov7-tiny', 'bf16')
        x = synth_image(4, 96, 128, seed=8, dtype=torch.uint8).cuda()
        want_out, want_meta = m.detect(x, 0.05, 0.5, max_det=50, packed=True)
        pg = ydist.PeerGather(4, 50)
        for _ in range(3):
            out, meta = m.detect(x, 0.05, 0.5, max_det=50, packed=True, gather=pg)
            assert torch.equal(meta, want_meta)
            for i in range(4):
                n = int(meta[i])
                assert torch.equal(out[i, :n], want_out[i, :n])
        lst = m.detect(x, 0.05, 0.5, max_det=50, gather=pg)
        assert [d.shape[0] for d in lst] == want_meta[:4].tolist()
        pg.check()
        with pytest.raises(ValueError):
            pg.begin(8, 50)
        # two batches in flight (bench.py --streams 2 on N > 1 GPUs): one PeerGather per compute stream, because the slot
        # protocol of a gather is ordered by one stream; a single gather with two compute streams is refused
        pg2 = ydist.PeerGather(4, 50)
        xs = [synth_image(4, 96, 128, seed=60 + i, dtype=torch.uint8).pin_memory() for i in range(7)]
        ref = [m.detect(xb.cuda(), 0.05, 0.5, max_det=50, packed=True) for xb in xs]
        got = [(o.clone(), c.clone()) for o, c in m.detect_batches(iter(xs), 0.05, 0.5, max_det=50, gather=[pg, pg2],
                                                                   compute_streams=2)]
        assert len(got) == len(xs)
        for (o, c), (wo, wm) in zip(got, ref):
            assert torch.equal(c, wm[:4].cpu())
            for i in range(4):
                assert torch.equal(o[i, :c[i]], wo[i, :c[i]].cpu())
        with pytest.raises(ValueError):
            list(m.detect_batches(iter(xs[:1]), 0.05, 0.5, max_det=50, gather=pg, compute_streams=2))
        pg.check(), pg2.check()
        pg.close(), pg2.close()
        assert int(want_meta[:4].sum()) > 0
        # a sender that never arrives: the bounded wait must poison its block (counts = -1) and set the status word, and
        # the host-side read-back paths must raise instead of returning stale rows
        late = ydist.PeerGather(4, 50, timeout_s=0.05)
        _, slot, step = late.begin(4, 50)          # no NMS launch follows: the flag for `step` is never raised
        _, cnt = late.finish(slot, step)
        assert cnt.tolist() == [-1, -1, -1, -1]
        with pytest.raises(RuntimeError, match='timed out'):
            late.check()
        with pytest.raises(RuntimeError, match='timed out'):
            late.raise_if_poisoned(cnt.cpu())
        late.close()
    finally:
        dist.destroy_process_group()


def test_accelerate_wrapper_runs_the_engine():
    """accelerate(model) (the INTEGRATION.md route for checkpoints unpickled by attempt_load) drives the same engine:
    wrapping a module tree gives bit-identical predictions and detections (the duck-typing against the reference's own
    classes is pinned on the CPU by tests/test_zoo.py::test_accelerate_lowers_a_real_reference_model)."""
    from yolo_series_b200 import accelerate, non_max_suppression
    m = _build('yolov7-tiny', 'bf16')
    acc = accelerate(m).half()
    x = synth_image(2, 96, 128, seed=5, dtype=torch.uint8).cuda()
    p0, r0 = m(x)
    p1, r1 = acc(x)
    assert torch.equal(p0, p1) and all(torch.equal(a, b) for a, b in zip(r0, r1))
    d0, d1 = non_max_suppression(p0, 0.05, 0.5), acc.detect(x, 0.05, 0.5)
    assert all(torch.equal(a, b) for a, b in zip(d0, d1))
    assert acc.stride.tolist() == [8., 16., 32.] and next(acc.parameters()).is_cuda


def test_engine_cache_is_a_bounded_lru_sharing_packed_weights():
    """Model._engines: shape-specialised engines are evicted least-recently-used and destroyed; the packed
    weights are shared between engines of one model; results do not depend on cache state."""
    m = _build('yolov7-tiny', 'bf16')
    m.engine_cache_size = 2
    xs = [synth_image(1, 64 + 32 * i, 96, seed=30 + i, dtype=torch.uint8).cuda() for i in range(3)]
    first = m.detect(xs[0], 0.05, 0.5, max_det=50, packed=True)
    first = (first[0].clone(), first[1].clone())
    n_packed = len(m._packed)
    e0 = next(iter(m._engines.values()))
    m.detect(xs[1], 0.05, 0.5, max_det=50)
    m.detect(xs[2], 0.05, 0.5, max_det=50)
    assert len(m._engines) == 2 and len(m._packed) == n_packed
    assert e0.handle is None, 'the evicted engine must have been destroyed'
    engines = list(m._engines.values())
    assert engines[0].weights[0].data_ptr() == engines[1].weights[0].data_ptr()
    again = m.detect(xs[0], 0.05, 0.5, max_det=50, packed=True)      # rebuilt after eviction: same bits
    assert torch.equal(again[1], first[1])
    n = int(first[1][0])
    assert torch.equal(again[0][0, :n], first[0][0, :n])


def test_forward_profile_prints_per_op_table(capsys):
    """model(x, profile=True) (yolo.py:614-623): per-op device times are printed, results are unchanged."""
    m = _build('yolov7-tiny', 'bf16')
    x = synth_image(1, 64, 96, seed=3, dtype=torch.uint8).cuda()
    a = m(x)[0]
    b = m(x, profile=True)[0]
    assert torch.equal(a, b)
    txt = capsys.readouterr().out
    assert 'k_conv' in txt and 'ms total' in txt


def test_bench_line_contract_no_extras():
    """`bench.py` (b200 arm) on a small workload: ONE JSON line with the contract keys, the metric string shared with the
    reference arm, e2e with its copy sizes, the clocks sample and the in-flight configuration."""
    import json
    import os
    import subprocess
    import sys

    import bench
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, 'bench.py'), '--model', 'yolov7-tiny', '--img', '320', '--batch', '4',
                        '--steps', '5', '--warmup', '3', '--no-extras'], capture_output=True, text=True, timeout=600, cwd=root)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.strip().splitlines() if ln.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['metric'] == bench.metric_name('yolov7-tiny', 320) and 'impl' not in d
    for k in ('value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling', 'vs_baseline',
              'dtype', 'data', 'config', 'e2e', 'gpu_launches', 'clocks', 'value_one_stream'):
        assert k in d, k
    assert d['steps'] == 5 and d['n_gpus'] == 1 and d['dtype'] == 'bf16' and d['scaling'] == 'weak'
    assert d['value'] > 0 and d['gpu_launches'] > 0 and d['config']['streams'] == 2
    e = d['e2e']
    assert e['value'] > 0 and e['h2d_bytes_per_step'] == 4 * 3 * 320 * 320 and e['d2h_bytes_per_step'] > 0
    assert abs(d['value'] - 4 * 5 / (d['ms_per_step'] * 5 / 1e3)) / d['value'] < 1e-6
