"""Throughput of N independent replicas of the step (trunk graph + decode + NMS) on N streams of ONE GPU vs one stream:
the tails of one stream's persistent kernels are filled by the other stream's CTAs.  Usage: python scripts/two_stream_probe.py [B] [streams]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import Model  # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
NS = int(sys.argv[2]) if len(sys.argv) > 2 else 2
name = sys.argv[3] if len(sys.argv) > 3 else 'yolov7'
HW = int(sys.argv[4]) if len(sys.argv) > 4 else 640
models, xs, streams = [], [], []
for s in range(NS):
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, 0))
    models.append(m.cuda().eval().half())
    xs.append(synth_image(B, HW, HW, seed=1 + s, dtype=torch.uint8).cuda())
    streams.append(torch.cuda.Stream())
torch.cuda.synchronize()


def run(n_streams, steps):
    for s in range(n_streams):
        with torch.cuda.stream(streams[s]):
            for _ in range(3):
                models[s].detect(xs[s], 0.001, 0.65, packed=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(steps):
        s = k % n_streams
        with torch.cuda.stream(streams[s]):
            models[s].detect(xs[s], 0.001, 0.65, packed=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return steps * B / dt, dt / steps * 1e3


STEPS = int(os.environ.get('STEPS', 60))
for rep in range(int(os.environ.get('ROUNDS', 2))):
    for ns in range(1, NS + 1):
        ips, ms = run(ns, STEPS)
        print(f'{name} B={B} streams={ns}: {ips:.1f} img/s, {ms:.4f} ms/step', flush=True)
