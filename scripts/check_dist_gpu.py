"""SURVEY.md 8(e) correctness check on real GPUs (run under torchrun, one rank per GPU): the image-sharded result
gathered over NCCL must equal the single-replica result bit for bit, in original image order.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        scripts/check_dist_gpu.py [model] [img] [global_batch]
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from yolo_series_b200 import Model, dist as ydist                    # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image    # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else 'yolov7-tiny'
    img = int(sys.argv[2]) if len(sys.argv) > 2 else 320
    G = int(sys.argv[3]) if len(sys.argv) > 3 else 16
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m = m.to(dev).eval().half()
    x = synth_image(G, img, img, seed=5, dtype=torch.uint8)
    lo, hi = ydist.shard_bounds(G, rank, world)
    out, meta = m.detect(x[lo:hi].to(dev), 0.001, 0.65, packed=True)
    out_all, cnt_all = ydist.gather_detections(out, meta[:hi - lo])
    # the same exchange fused into the NMS kernel (peer-to-peer stores), three steps to exercise both buffer sets
    pg = ydist.PeerGather(hi - lo, 300, device=dev)
    ok_peer = True
    for _ in range(3):
        o2, m2 = m.detect(x[lo:hi].to(dev), 0.001, 0.65, packed=True, gather=pg)
        ok_peer = ok_peer and torch.equal(m2[:-1], cnt_all)
        valid = torch.arange(300, device=dev)[None, :] < cnt_all[:, None]
        ok_peer = ok_peer and torch.equal(o2[valid], out_all[valid])
    pg.check()
    pg.close()
    flag = torch.tensor([int(ok_peer)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok_peer = bool(flag.item())
    torch.cuda.synchronize()
    ok = ok_peer
    if rank == 0:
        ref_out, ref_meta = m.detect(x.to(dev), 0.001, 0.65, packed=True)
        cnt = ref_meta[:G]
        ok = ok and torch.equal(cnt_all, cnt)
        for i in range(G):
            n = int(cnt[i])
            ok = ok and torch.equal(out_all[i, :n], ref_out[i, :n])
        print(f'dist check {name} {img} G={G} world={world}: {"bit-identical" if ok else "MISMATCH"} '
              f'(NCCL gather == in-kernel peer gather on every rank: {ok_peer}); '
              f'detections/image mean {float(cnt.float().mean()):.1f}', flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == '__main__':
    main()
