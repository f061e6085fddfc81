"""Multi-process host logic of the image-sharded path on CPU: world_size-2 gloo.
Checks shard bounds, the pack/unpack of the fixed-size detection message and that the gathered
result equals the single-process result in original image order."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from yolo_series_b200 import dist as ydist


def test_shard_bounds_cover_batch():
    for B in (1, 7, 16, 32, 64):
        for world in (1, 2, 4, 8):
            spans = [ydist.shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1


def test_pack_unpack_roundtrip():
    g = torch.Generator().manual_seed(0)
    out = torch.randn((5, 300, 6), generator=g)
    counts = torch.tensor([0, 300, 17, 1, 299], dtype=torch.int32)
    o2, c2 = ydist.unpack_detections(ydist.pack_detections(out, counts), 300)
    assert torch.equal(o2, out) and torch.equal(c2, counts)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, B, max_det, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    g = torch.Generator().manual_seed(123)
    full_out = torch.randn((B, max_det, 6), generator=g)
    full_cnt = torch.randint(0, max_det + 1, (B,), generator=g, dtype=torch.int32)
    lo, hi = ydist.shard_bounds(B, rank, world)
    out, cnt = ydist.gather_detections(full_out[lo:hi].clone(), full_cnt[lo:hi].clone())
    ok = torch.equal(out, full_out) and torch.equal(cnt, full_cnt)
    lst = ydist.to_list(out, cnt)
    ok = ok and all(lst[i].shape[0] == int(full_cnt[i]) for i in range(B))
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_gather_world2_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 8, 50, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
