"""Where does the e2e pipeline lose time vs the device-resident loop?  (diagnostic, not a bench)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolo_series_b200 import Model
from yolo_series_b200.synth import seeded_state_dict, synth_image

dev = torch.device('cuda', 0)
m = Model('yolov7'); m.load_state_dict(seeded_state_dict(m, seed=0)); m = m.to(dev).eval().half()
B = 32
xh = synth_image(B, 640, 640, seed=100, dtype=torch.uint8).pin_memory()
xd = xh.to(dev)
K = 40
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

def timed(name, fn):
    fn(3); torch.cuda.synchronize()
    t0 = time.perf_counter(); e0.record(); fn(K); e1.record(); torch.cuda.synchronize()
    print(f'{name:40s} {e0.elapsed_time(e1)/K:.3f} ms/step  (wall {(time.perf_counter()-t0)*1e3/K:.3f})', flush=True)

def dev_loop(n):
    for _ in range(n): m.detect(xd, 0.001, 0.65, max_det=300, packed=True)
timed('device loop', dev_loop)

def dev_loop_sync_prev(n):
    prev = None
    for _ in range(n):
        ev = torch.cuda.Event()
        m.detect(xd, 0.001, 0.65, max_det=300, packed=True); ev.record()
        if prev is not None: prev.synchronize()
        prev = ev
timed('device loop, host waits on step i-1', dev_loop_sync_prev)

def pipe(src):
    def f(n):
        for _ in m.detect_batches((src for _ in range(n)), 0.001, 0.65, max_det=300): pass
    return f
timed('detect_batches, device source', pipe(xd))
timed('detect_batches, pinned host source', pipe(xh))

def h2d_only(n):
    st = torch.empty_like(xd)
    for _ in range(n): st.copy_(xh, non_blocking=True)
timed('H2D only', h2d_only)

def trunk_only(n):
    for _ in range(n): m.run_trunk(xd)
timed('trunk only (stage+graph)', trunk_only)

eng = next(iter(m._engines.values()))
def graph_only(n):
    for _ in range(n): eng.run()
timed('graph only', graph_only)

# host cost of issuing one step
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): m.detect(xd, 0.001, 0.65, max_det=300, packed=True)
t1 = time.perf_counter(); torch.cuda.synchronize()
print(f'host issue time per detect() call: {(t1-t0)/20*1e3:.3f} ms')
