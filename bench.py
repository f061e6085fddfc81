#!/usr/bin/env python
"""Headline benchmark: yolov7 640x640 bf16 end-to-end images/s (forward + decode + NMS).

    python bench.py --gpus N --steps K --warmup W            # this framework (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...             # the reference algorithm on the host cores

Workload (BASELINE.json configs[1]): yolov7 (cfg/training/yolov7.yaml), 640x640, batch 32 per GPU,
bf16 trunk (RepConv/BN/Implicit folded), conf_thres 0.001, iou_thres 0.65, max_det 300, seeded
synthetic weights and images (no checkpoints / datasets offline).  A "step" is one pass of the hot
path over one batch.  N > 1 (torchrun): image-sharded replicas, batch 32 per rank (weak scaling),
one NCCL gather of the fixed-size detection buffers per step.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODEL, IMG, BATCH = 'yolov7', 640, 32
CONF, IOU, MAX_DET = 0.001, 0.65, 300
FALLBACK_PEAKS = {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0}


def load_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return {k: float(d[k]) for k in FALLBACK_PEAKS if k in d}, 'measured'
        except Exception:
            pass
    return dict(FALLBACK_PEAKS), 'fallback'


class ClockSampler:
    """nvidia-smi sampling DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-lms', '50', '-i', str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([time.time()] + [c.strip() for c in line.split(',')])

    def window(self, t0, t1):
        """Keep the samples taken inside [t0, t1] (the sampler is started before the warm-up so that nvidia-smi's own
        start-up -- NVML initialisation stalls launches for milliseconds -- does not fall into the timed region)."""
        self.t0, self.t1 = t0, t1

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        t0, t1 = getattr(self, 't0', 0.0), getattr(self, 't1', float('inf'))
        inside = [r[1:] for r in self.rows if t0 <= r[0] <= t1 + 0.12]
        for r in (inside if inside else [r[1:] for r in self.rows[-3:]]):
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def cpu_reference_step(cfg, sd, imgs, threads):
    """One bounded sample of the reference algorithm on the host cores: the oracle's restatement of
    Model.fuse()+forward (fp32, torch CPU kernels = what the reference itself dispatches to) and
    non_max_suppression called per image.  Returns seconds."""
    import torch
    from oracle.nms_ref import non_max_suppression_ref
    from oracle.ref_forward import ref_forward
    torch.set_num_threads(threads)
    t0 = time.perf_counter()
    pred, _ = ref_forward(cfg, sd, imgs, fused=True)
    for i in range(pred.shape[0]):
        non_max_suppression_ref(pred[i:i + 1], CONF, IOU, max_det=MAX_DET)
    return time.perf_counter() - t0


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores (rank 0 only)."""
    if rank != 0:
        return
    import torch
    from yolo_series_b200 import Model, zoo
    from yolo_series_b200.synth import seeded_state_dict, synth_image
    threads = os.cpu_count() or 1
    sample = 8  # images per step: bounded so that K steps finish within minutes (~1 s/step on 16 cores)
    cfg = zoo.get_cfg(MODEL)
    sd = seeded_state_dict(Model(cfg), seed=0)
    imgs = synth_image(sample, IMG, IMG, seed=2)
    for _ in range(args.warmup):
        cpu_reference_step(cfg, sd, imgs, threads)
    t = sum(cpu_reference_step(cfg, sd, imgs, threads) for _ in range(args.steps))
    v = sample * args.steps / t
    line = {
        'impl': 'reference', 'metric': 'yolov7 640 e2e images/sec (forward+decode+NMS)', 'value': v,
        'unit': 'images/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * t / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': f'{MODEL} {IMG}x{IMG} conf {CONF} iou {IOU} max_det {MAX_DET}, {sample} images/step on host CPU',
                   'batch_per_step': sample},
        'cpu_baseline': {'value': v, 'unit': 'images/s', 'cores': threads, 'kind': 'port',
                         'sample': f'{sample} images/step x {args.steps} steps, fp32 torch CPU kernels + oracle NMS'},
        'e2e': {'value': v, 'unit': 'images/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def nms_stress(dev, peaks):
    """BASELINE.json configs[4]: 100k candidates/image, 80 classes, max_det 1000, batch 256 (8.7 GB)."""
    import torch
    from yolo_series_b200.nms import _run_nms
    B, A, nc = 256, 100000, 80
    g = torch.Generator(device=dev).manual_seed(0)
    pred = torch.empty((B, A, nc + 5), device=dev)
    K = 600
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
    run = lambda: _run_nms(pred, B, A, nc + 5, 0.001, 0.65, None, False, False, 30000, 1000, dev, return_packed=True)
    for _ in range(3):
        out, meta = run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        out, meta = run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    algo_bytes = B * A * (nc + 5) * 4 + B * 1000 * 6 * 4 + B * 4
    return {'us_per_image': 1e3 * ms / B, 'batch': B, 'rows_per_image': A, 'max_det': 1000,
            'kept_mean': float(meta[:B].float().mean()), 'algorithmic_GBps': algo_bytes / ms / 1e6,
            'frac_of_hbm_peak': algo_bytes / ms / 1e6 / peaks['hbm_gbs']}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=BATCH, help='images per GPU per step')
    ap.add_argument('--no-extras', action='store_true', help='skip cpu_baseline / NMS stress / per-op profile')
    # secondary BASELINE.json configs (configs[2], configs[3]); the defaults are the headline workload
    ap.add_argument('--model', default=MODEL, help='cfg name (yolov7 | yolov7-tiny | yolov7-w6-custom | yolov7-e6e)')
    ap.add_argument('--img', type=int, default=IMG)
    ap.add_argument('--global-batch', type=int, default=0,
                    help='fixed global batch sharded image-wise over the ranks (strong scaling); 0 = --batch per rank (weak)')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from yolo_series_b200 import Model, dist as ydist
    from yolo_series_b200.synth import seeded_state_dict, synth_image
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    peaks, peak_src = load_peaks()
    B = args.batch
    model_name, img = args.model, args.img
    scaling = 'weak'
    if args.global_batch:
        lo, hi = ydist.shard_bounds(args.global_batch, rank, world)
        if (hi - lo) * world != args.global_batch:
            raise SystemExit('--global-batch must divide evenly over the ranks (equal shards for the all-gather)')
        B, scaling = hi - lo, 'strong'
    headline = (model_name, img) == (MODEL, IMG)
    model = Model(model_name)
    model.load_state_dict(seeded_state_dict(model, seed=0))
    model = model.to(dev).eval().half()          # bf16 trunk (reference: model.half(), detect.py:49)
    x_host = synth_image(B, img, img, seed=100 + rank, dtype=torch.uint8).pin_memory()
    x_dev = x_host.to(dev)

    # N > 1: the one exchange of the path (every rank's detections to every rank) happens inside the NMS kernel through
    # peer-to-peer stores over NVLink (dist.PeerGather); YB_GATHER=nccl selects the plain NCCL all-gather instead.
    use_peer = world > 1 and os.environ.get('YB_GATHER', 'peer') != 'nccl'
    pg = None
    if use_peer:
        try:
            pg = ydist.PeerGather(B, MAX_DET, device=dev)
        except RuntimeError as ex:          # raised on every rank together: CUDA IPC unavailable -> NCCL all-gather
            use_peer = False
            if rank == 0:
                print(f'[bench] {ex}', file=sys.stderr)

    def step_device():
        out, meta = model.detect(x_dev, CONF, IOU, max_det=MAX_DET, packed=True, gather=pg)
        if world > 1 and pg is None:
            out, cnt = ydist.gather_detections(out, meta[:B])
        return out, meta

    def step_e2e():
        # the pinned host batch is copied H2D straight into the engine's input buffer, every step
        out, meta = model.detect(x_host, CONF, IOU, max_det=MAX_DET, packed=True, gather=pg)
        cnt = meta[:-1]
        if world > 1 and pg is None:
            out, cnt = ydist.gather_detections(out, cnt)
        return out.cpu(), cnt.cpu()                               # D2H read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record()
    for _ in range(args.steps):
        out, meta = step_device()
    e1.record()
    barrier()
    sampler.window(t_wall0, time.time())
    ms_dev = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    # ---- end-to-end through the public API with host buffers
    for _ in range(2):
        step_e2e()
    barrier()
    e0.record()
    for _ in range(args.steps):
        res, cnt = step_e2e()
    e1.record()
    barrier()
    ms_e2e_serial = e0.elapsed_time(e1)
    # the same through the streaming API (Model.detect_batches): every step still copies its own pinned host
    # batch H2D and reads its own detections back D2H, but on copy streams that overlap the neighbouring steps
    post = (lambda o, c: ydist.gather_detections(o, c)) if (world > 1 and pg is None) else None

    def run_stream(n):
        last = None
        for last in model.detect_batches((x_host for _ in range(n)), CONF, IOU, max_det=MAX_DET, post=post, gather=pg):
            pass
        return last
    run_stream(3)
    barrier()
    e0.record()
    res, cnt = run_stream(args.steps)
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_e2e_serial], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_e2e_serial = float(t[0]), float(t[1]), float(t[2])
    if pg is not None:
        pg.check()
        pg.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    eng = next(iter(model._engines.values()))
    n_imgs = B * world * args.steps
    launches_per_step = 1 + eng.num_launches() + 2     # stage + trunk ops + k_filter + k_sort_nms
    line = {
        'metric': f'{model_name} {img} bf16 e2e images/sec (forward+decode+NMS)', 'value': n_imgs / (ms_dev / 1e3),
        'unit': 'images/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': ms_dev / args.steps, 'higher_is_better': True, 'scaling': scaling, 'vs_baseline': None,
        'dtype': 'bf16', 'data': 'synthetic',
        'config': {'workload': f'{model_name} {img}x{img} batch {B}/GPU bf16, RepConv/BN folded, conf {CONF} iou {IOU} max_det {MAX_DET}',
                   'batch_per_gpu': B, 'global_batch': B * world, 'parallelism': f'image-sharded replicas x{world}',
                   'gather': ('none' if world == 1 else 'in-kernel P2P stores over NVLink (yb_decode_nms_gather)' if use_peer
                              else 'NCCL all_gather'),
                   'l2': 'no flush: per-step working set (activations, GBs) >> 126 MB L2'},
        'e2e': {'value': n_imgs / (ms_e2e / 1e3), 'unit': 'images/s', 'h2d_bytes_per_step': x_host.numel(),
                'd2h_bytes_per_step': int(res.numel() * 4 + cnt.numel() * 4 + 4), 'ms_per_step': ms_e2e / args.steps,
                'api': 'Model.detect_batches (H2D / compute / D2H on three streams, two slots)',
                'serial_value': n_imgs / (ms_e2e_serial / 1e3),
                'serial_api': 'Model.detect(x_host) + .cpu() per step, one stream'},
        'gpu_launches': launches_per_step * args.steps,
        'clocks': clocks,
        'detections_per_image': float(cnt.float().mean()),
    }
    if not args.no_extras:
        # ---- roofline of the dominant kernel (k_conv_tc): CUDA events around every op of an eager pass
        prof = [eng.profile() for _ in range(3)][-1]
        conv_ms = sum(r[2] for r in prof if r[1] == 1)
        all_ms = sum(r[2] for r in prof)
        flops = sum(r[3] for r in prof)
        ach = flops / (conv_ms / 1e3) / 1e12
        n_conv = sum(1 for r in prof if r[1] == 1)
        traffic, traffic_src = None, None
        try:   # (headline workload only) DRAM bytes of the conv launches of one step, from the committed ncu launch list (cold-cache per launch)
            if not headline:
                raise KeyError('launch list exists for the headline workload only')
            summ = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles',
                                               'round1_launch_summary.json')))
            traffic = summ['conv_dram_bytes_per_step']
            traffic_src = 'profiles/round1_launch_summary.json (ncu dram__bytes_read+write, sum over the conv launches of one step)'
        except Exception:  # noqa: BLE001
            pass
        line['roofline'] = {'bound': 'tensor', 'kernel': f'tcgen05 conv kernels k_conv_tc / k_conv3_tc / k_conv3s2p_tc / k_stem_tc ({n_conv} launches/step)',
                            'achieved': ach,
                            'peak': peaks['bf16_tflops_sustained'], 'peak_source': peak_src + ' (sustained bf16 cuBLAS)',
                            'unit': 'TFLOP/s', 'frac': ach / peaks['bf16_tflops_sustained'], 'traffic': traffic,
                            'traffic_source': traffic_src,
                            'flops_per_step': flops, 'conv_ms_per_step': conv_ms, 'trunk_ms_per_step_eager': all_ms,
                            'conv_share_of_trunk': conv_ms / all_ms}
        top = sorted(prof, key=lambda r: -r[2])[:8]
        line['top_ops'] = [{'op': r[0], 'ms': round(r[2], 4), 'tflops': round(r[3] / r[2] / 1e9, 1) if r[3] else None} for r in top]
        if world == 1 and headline:
            try:
                line['nms_stress'] = nms_stress(dev, peaks)
            except Exception as ex:  # noqa: BLE001
                line['nms_stress'] = {'error': str(ex)[:200]}
            from yolo_series_b200 import zoo
            threads = os.cpu_count() or 1
            cfg = zoo.get_cfg(MODEL)
            sd = seeded_state_dict(Model(cfg), seed=0)
            nimg = 8
            imgs = synth_image(nimg, IMG, IMG, seed=2)
            cpu_reference_step(cfg, sd, imgs, threads)
            reps = 10      # ~10-15 s of host work
            t = sum(cpu_reference_step(cfg, sd, imgs, threads) for _ in range(reps))
            line['cpu_baseline'] = {'value': nimg * reps / t, 'unit': 'images/s', 'cores': threads, 'kind': 'port',
                                    'sample': f'{nimg * reps} images ({reps} steps of {nimg}) of the same workload, fp32 torch CPU + oracle NMS'}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
