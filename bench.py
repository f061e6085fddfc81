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


def metric_name(model, img):
    """ONE string for both arms (the driver divides the two lines only when metric / unit / direction agree); the
    arithmetic type lives in `dtype`, not in the name."""
    return f'{model} {img} e2e images/sec (forward+decode+NMS)'


def workload_name(model, img, batch):
    return f'{model} {img}x{img} batch {batch}/GPU, RepConv/BN folded, conf {CONF} iou {IOU} max_det {MAX_DET}'


def cpu_model():
    try:
        for l in open('/proc/cpuinfo'):
            if l.startswith('model name'):
                return l.split(':', 1)[1].strip()
    except OSError:
        pass
    return 'unknown'
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
        sm, mx, pw, reasons = [], [], [], set()
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
            try:
                pw.append(float(r[2]))
            except ValueError:
                pass
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'power_w': statistics.median(pw) if pw else None, 'samples': len(sm), 'reasons': sorted(reasons)}


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
        non_max_suppression_ref(pred[i:i + 1], CONF, IOU, max_det=MAX_DET, early_exit=False)   # torchvision visits every box
    return time.perf_counter() - t0


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores (rank 0 only), same metric / config strings as the
    repo arm; each step is a bounded sample (8 images) of the configured batch."""
    if rank != 0:
        return
    import torch
    from yolo_series_b200 import Model, zoo
    from yolo_series_b200.synth import seeded_state_dict, synth_image
    threads = os.cpu_count() or 1
    model_name, img = args.model, args.img
    # images per step: bounded so that K steps finish within minutes (yolov7 640: ~1 s/step on 16 cores)
    sample = 8 if img <= 640 else 2
    cfg = zoo.get_cfg(model_name)
    sd = seeded_state_dict(Model(cfg), seed=0)
    imgs = synth_image(sample, img, img, seed=2)
    for _ in range(args.warmup):
        cpu_reference_step(cfg, sd, imgs, threads)
    t = sum(cpu_reference_step(cfg, sd, imgs, threads) for _ in range(args.steps))
    v = sample * args.steps / t
    line = {
        'impl': 'reference', 'metric': metric_name(model_name, img), 'value': v,
        'unit': 'images/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * t / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': workload_name(model_name, img, args.batch), 'batch_per_gpu': args.batch,
                   'sample': f'{sample} images/step of that workload on the host CPU ({threads} threads, {cpu_model()})',
                   'images_per_step': sample},
        'cpu_baseline': {'value': v, 'unit': 'images/s', 'cores': threads, 'kind': 'port', 'cpu': cpu_model(),
                         'sample': f'{sample} images/step x {args.steps} steps, fp32 torch CPU kernels + oracle NMS'},
        'e2e': {'value': v, 'unit': 'images/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def nms_stress(dev, peaks, cpu_images=8):
    """BASELINE.json configs[4]: 100k candidates/image, 80 classes, max_det 1000, batch 256 (8.7 GB) -- the GPU path in
    us/image next to the reference algorithm on the host cores (second half of BASELINE.json's metric)."""
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
    res = {'us_per_image': 1e3 * ms / B, 'batch': B, 'rows_per_image': A, 'max_det': 1000,
           'kept_mean': float(meta[:B].float().mean()), 'algorithmic_GBps': algo_bytes / ms / 1e6,
           'frac_of_hbm_peak': algo_bytes / ms / 1e6 / peaks['hbm_gbs']}
    # ---- CPU leg of "NMS us/image vs CPU ref" (SURVEY.md 8d): the reference algorithm (oracle restatement of
    # general.py:628-720 with max_det as a parameter, one image per call so no time_limit applies) on a fixed subset of
    # the SAME batch, with all host threads and with one; the GPU rows of those images are checked against it.
    import numpy as np
    from oracle.nms_ref import non_max_suppression_ref
    sub = list(range(0, B, B // cpu_images))[:cpu_images]
    host = pred[sub].cpu()
    counts = meta[:B].cpu()
    gpu_rows = [out[i, :int(counts[i])].cpu().numpy() for i in sub]
    del pred
    threads = os.cpu_count() or 1

    def cpu_leg(nthreads, idx):
        torch.set_num_threads(nthreads)
        t0, outs = time.perf_counter(), []
        for j in idx:
            outs.append(non_max_suppression_ref(host[j:j + 1], 0.001, 0.65, max_det=1000, max_nms=30000,
                                                early_exit=False)[0].numpy())
        return (time.perf_counter() - t0) / len(idx), outs
    t_all, ref_rows = cpu_leg(threads, range(len(sub)))
    t_one, _ = cpu_leg(1, range(min(2, len(sub))))
    torch.set_num_threads(threads)
    res.update({'cpu_us_per_image': 1e6 * t_all, 'cpu_threads': threads, 'cpu_us_per_image_1thread': 1e6 * t_one,
                'cpu_model': cpu_model(), 'cpu_kind': 'port (oracle/nms_ref.py + oracle/nms_greedy.c)',
                'cpu_sample': f'images {sub} of the {B}-image batch ({len(sub)} images, all threads; first '
                              f'{min(2, len(sub))} with 1 thread)',
                'gpu_vs_cpu': t_all * 1e6 / (1e3 * ms / B),
                'parity_on_sample': 'bit-identical' if all(np.array_equal(a, b) for a, b in zip(gpu_rows, ref_rows))
                                    else 'MISMATCH'})
    return res


def trunk_and_nms_ms(model, eng, reps=10):
    """Device time of the captured trunk graph alone and of the fused decode+NMS alone (CUDA events on the
    launching stream, after warm-up).  Both run back to back inside every timed step, so each is <= ms_per_step."""
    import torch
    from yolo_series_b200.nms import _run_nms
    lv, head = model.head_levels(eng)
    A = sum(head.na * v.H * v.W for v in eng.head_views)
    nms = lambda: _run_nms(None, eng.B, A, head.no, CONF, IOU, None, False, False, 30000, MAX_DET, eng.device,
                           levels=(lv, head.nl, head.na), return_packed=True)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for _ in range(2):
        eng.run(); nms()
    torch.cuda.synchronize()
    t_trunk = t_nms = 0.0
    for _ in range(reps):
        e[0].record(); eng.run(); e[1].record(); nms(); e[2].record()
        torch.cuda.synchronize()
        t_trunk += e[0].elapsed_time(e[1]); t_nms += e[1].elapsed_time(e[2])
    return t_trunk / reps, t_nms / reps, A, head.no


def in_graph_spans(model, x_dev, reps=3):
    """Per-launch time of the tcgen05 conv kernels INSIDE one replay of the captured trunk graph, from per-CTA
    %globaltimer stamps the kernels write when a plan was created with YB_CONV_TS set (first griddepcontrol.wait
    returned -> last CTA exit; csrc/tc_common.cuh YB_CTS, scripts/step_timeline.py).  The eager per-op pass puts an
    event -- and a full dependency -- between consecutive ops, which inflates the many short launches; these spans
    are what the launches cost in the graph.  Returns {op index: span ms} (median of `reps` replays) and the engine."""
    import torch
    ctas, slots = 320, 16
    stride = ctas * slots
    ts = torch.zeros(512 * stride, dtype=torch.int64, device=x_dev.device)
    os.environ['YB_CONV_TS'], os.environ['YB_CONV_TS_STRIDE'] = hex(ts.data_ptr()), str(stride)
    try:
        st = torch.cuda.Stream(x_dev.device)          # engines are cached per stream: this one is built with the stamps on
        with torch.cuda.stream(st):
            eng = model.run_trunk(x_dev)
            for _ in range(2):
                model.run_trunk(x_dev)
            stamped = [i for i in range(len(eng.ops)) if eng.op_kernel(i) in
                       ('k_conv_tc', 'k_conv3_tc', 'k_conv3w_tc', 'k_conv3s2p_tc')]
            runs = []
            for _ in range(reps):
                st.synchronize()
                ts.zero_()
                st.synchronize()
                model.run_trunk(x_dev)
                st.synchronize()
                t = ts.view(-1, ctas, slots)[:len(stamped)].cpu()
                spans = {}
                for slot, i in enumerate(stamped):
                    a = t[slot]
                    a = a[a[:, 0] > 0]
                    w, ex = a[:, 2][a[:, 2] > 0], a[:, 9][a[:, 9] > 0]
                    if w.numel() and ex.numel():
                        spans[i] = (int(ex.max()) - int(w.min())) / 1e6
                runs.append(spans)
    finally:
        os.environ.pop('YB_CONV_TS', None)
        os.environ.pop('YB_CONV_TS_STRIDE', None)
    out = {i: sorted(r[i] for r in runs if i in r)[len(runs) // 2] for i in runs[0]}
    return out, eng, ts


def kernel_rooflines(model, eng, peaks, peak_src, traffic_by_kernel, step_ms=None, x_dev=None):
    """One roofline entry per kernel of the step.  Per-op times come from an eager pass with a CUDA event between
    consecutive ops (yb_engine_profile); that pass runs without the graph's programmatic-dependent-launch overlap, so
    its per-kernel SHARES are applied to the measured time of the captured trunk graph -- every `ms` below is a slice
    of the timed step.  Bound of a kernel = the slower of FLOPs / sustained bf16 peak and algorithmic bytes / HBM copy
    peak, summed over its launches (per-layer bound); frac = bound time / measured time."""
    trunk_ms, nms_ms, A, no = trunk_and_nms_ms(model, eng)
    measured = (trunk_ms, nms_ms)
    if step_ms:     # slices of the TIMED step: the separate trunk / NMS timings run later, on a warmer GPU
        sc = step_ms / (trunk_ms + nms_ms)
        trunk_ms, nms_ms = trunk_ms * sc, nms_ms * sc
    table = eng.profile_table(3)
    eager_ms = sum(r['ms'] for r in table)
    tf_peak, hbm_peak = peaks['bf16_tflops_sustained'], peaks['hbm_gbs']
    groups = {}
    for r in table:
        g = groups.setdefault(r['kernel'], {'launches': 0, 'ms_eager': 0.0, 'flops': 0, 'bytes': 0, 'bound_ms': 0.0,
                                            't_tensor': 0.0, 't_hbm': 0.0})
        tt, th = r['flops'] / (tf_peak * 1e9), r['bytes'] / (hbm_peak * 1e6)
        g['launches'] += 1; g['ms_eager'] += r['ms']; g['flops'] += r['flops']; g['bytes'] += r['bytes']
        g['bound_ms'] += max(tt, th); g['t_tensor'] += tt; g['t_hbm'] += th
    spans = {}
    if x_dev is not None:
        try:
            spans, _eng2, _ts = in_graph_spans(model, x_dev)
        except Exception as ex:  # noqa: BLE001
            print(f'[bench] in-graph spans unavailable: {ex}', file=sys.stderr)
    for i, r in enumerate(table):
        if i in spans:
            g = groups[r['kernel']]
            g['span_ms'] = g.get('span_ms', 0.0) + spans[i]
            g['span_bound_ms'] = g.get('span_bound_ms', 0.0) + max(r['flops'] / (tf_peak * 1e9), r['bytes'] / (hbm_peak * 1e6))
    out = []
    for k, g in groups.items():
        ms = g['ms_eager'] * trunk_ms / eager_ms
        tensor = g['t_tensor'] >= g['t_hbm']
        ach = g['flops'] / ms / 1e9 if tensor else g['bytes'] / ms / 1e6
        out.append({'kernel': k, 'launches': g['launches'], 'ms': ms, 'bound': 'tensor' if tensor else 'hbm',
                    'achieved': ach, 'peak': tf_peak if tensor else hbm_peak, 'unit': 'TFLOP/s' if tensor else 'GB/s',
                    'frac': ach / (tf_peak if tensor else hbm_peak), 'frac_of_layer_bound': g['bound_ms'] / ms,
                    'tflops': g['flops'] / ms / 1e9, 'algorithmic_GBps': g['bytes'] / ms / 1e6,
                    'flops': g['flops'], 'bytes': g['bytes'], 'traffic': traffic_by_kernel.get(k)})
        if g.get('span_ms'):      # the same launches inside the graph (timer stamps): see in_graph_spans
            out[-1]['in_graph_ms'] = round(g['span_ms'], 4)
            out[-1]['in_graph_frac_of_layer_bound'] = round(g['span_bound_ms'] / g['span_ms'], 4)
    nms_bytes = eng.B * A * no * 4 + eng.B * MAX_DET * 24 + eng.B * 4
    out.append({'kernel': 'k_filter_cells+k_sort_nms', 'launches': 2, 'ms': nms_ms, 'bound': 'hbm',
                'achieved': nms_bytes / nms_ms / 1e6, 'peak': hbm_peak, 'unit': 'GB/s',
                'frac': nms_bytes / nms_ms / 1e6 / hbm_peak, 'frac_of_layer_bound': nms_bytes / nms_ms / 1e6 / hbm_peak,
                'flops': 0, 'bytes': nms_bytes,
                'traffic': (traffic_by_kernel.get('k_filter_cells', 0) + traffic_by_kernel.get('k_sort_nms', 0)) or None})
    out.sort(key=lambda r: -r['ms'])
    for r in out:
        for k in ('ms', 'achieved', 'frac', 'frac_of_layer_bound', 'tflops', 'algorithmic_GBps'):
            if k in r:
                r[k] = round(r[k], 4)
    conv = [r for r in out if r['kernel'].startswith(('k_conv', 'k_stem'))]
    summary = {'trunk_graph_ms': round(trunk_ms, 4), 'decode_nms_ms': round(nms_ms, 4),
               'eager_per_op_sum_ms': round(eager_ms, 4),
               'conv_ms': round(sum(r['ms'] for r in conv), 4),
               'conv_tflops': round(sum(r['flops'] for r in conv) / sum(r['ms'] for r in conv) / 1e9, 1),
               'conv_frac_of_sustained_bf16': round(sum(r['flops'] for r in conv) / sum(r['ms'] for r in conv) / 1e9 / tf_peak, 4),
               'step_frac_of_layer_bounds': round(sum(r['frac_of_layer_bound'] * r['ms'] for r in out) / sum(r['ms'] for r in out), 4),
               'in_graph': ('in_graph_ms / in_graph_frac_of_layer_bound: the conv launches timed INSIDE a graph replay by per-CTA '
                            '%globaltimer stamps (first griddepcontrol.wait passed -> last CTA exit, summed per kernel), '
                            'measured in a separate replay at that moment\'s clocks -- the eager per-op pass serialises '
                            'consecutive launches and overstates the many short ones'),
               'trunk_graph_ms_alone': round(measured[0], 4), 'decode_nms_ms_alone': round(measured[1], 4),
               'method': ('per-kernel shares of an eager per-op event pass applied to the captured trunk graph; trunk graph '
                          'and decode+NMS are timed alone with CUDA events (.._alone) and scaled by one factor so that '
                          'their sum is the timed step (ms_per_step) -- every ms is a slice of the timed step; with two '
                          'batches in flight that slice is the kernel\'s share of the wall time, its rate an effective one'),
               'peaks': f'{peak_src}: bf16 sustained {tf_peak} TFLOP/s, HBM copy {hbm_peak} GB/s'}
    return out, summary, table


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=BATCH, help='images per GPU per step')
    ap.add_argument('--no-extras', action='store_true', help='skip cpu_baseline / NMS stress / per-op profile')
    ap.add_argument('--streams', type=int, default=2, choices=[1, 2, 3, 4],
                    help='batches in flight on one GPU: n = consecutive steps rotate over n streams / engines')
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
    pg, pgs = None, [None] * args.streams
    if use_peer:
        try:                                # one gather per compute stream (its slot protocol is ordered by one stream)
            pgs = [ydist.PeerGather(B, MAX_DET, device=dev) for _ in range(args.streams)]
            pg = pgs[0]
        except RuntimeError as ex:          # raised on every rank together: CUDA IPC unavailable -> NCCL all-gather
            use_peer, pg, pgs = False, None, [None] * args.streams
            if rank == 0:
                print(f'[bench] {ex}', file=sys.stderr)
    # Two batches in flight (args.streams = 2): consecutive steps are independent, so step k+1 runs on a second stream
    # with its own engine and its CTAs take the SMs that step k's persistent kernels leave idle (last rounds, dependent
    # launches, the 32-CTA NMS).  Every step is still one full pass over one batch; --streams 1 = strictly serial.
    n_streams = args.streams if (world == 1 or use_peer) else 1
    side = model.compute_streams(n_streams, dev) if n_streams > 1 else None

    def step_device(k=0, serial=False):
        if side is None or serial:
            out, meta = model.detect(x_dev, CONF, IOU, max_det=MAX_DET, packed=True, gather=pg)
            if world > 1 and pg is None:
                out, cnt = ydist.gather_detections(out, meta[:B])
            return out, meta
        with torch.cuda.stream(side[k % n_streams]):
            return model.detect(x_dev, CONF, IOU, max_det=MAX_DET, packed=True, gather=pgs[k % n_streams])

    def join_streams():
        if side is not None:
            for st in side:
                torch.cuda.current_stream().wait_stream(st)

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
    for k in range(max(args.warmup, 3) * n_streams):
        step_device(k)
    for _ in range(3):
        step_device(serial=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record()
    for k in range(args.steps):
        out, meta = step_device(k)
    join_streams()
    e1.record()
    barrier()
    sampler.window(t_wall0, time.time())
    ms_dev = e0.elapsed_time(e1)
    # the same number of strictly serial steps on one stream (reported as value_one_stream)
    e0.record()
    for _ in range(args.steps):
        step_device(serial=True)
    e1.record()
    barrier()
    ms_serial = e0.elapsed_time(e1)
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
        for last in model.detect_batches((x_host for _ in range(n)), CONF, IOU, max_det=MAX_DET, post=post,
                                         gather=(pgs if (pg is not None and n_streams > 1) else pg),
                                         compute_streams=n_streams):
            pass
        return last
    run_stream(3 * n_streams + 2)   # every slot of the pipeline (3 per compute stream) has been through one batch
    barrier()
    e0.record()
    res, cnt = run_stream(args.steps)
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_e2e_serial, ms_serial], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_e2e_serial, ms_serial = float(t[0]), float(t[1]), float(t[2]), float(t[3])
    # ---- N > 1: correctness of the exchange, outside the timed region.  Rank 0 regenerates every rank's seeded shard,
    # runs it through its own replica (no gather) and compares with the block that arrived for that rank.
    gather_check = None
    if world > 1:
        if pg is not None:
            out_all, meta_all = step_device()
            cnt_all = meta_all[:B * world]
        else:
            o, m = model.detect(x_dev, CONF, IOU, max_det=MAX_DET, packed=True)
            out_all, cnt_all = ydist.gather_detections(o, m[:B])
        torch.cuda.synchronize()
        if rank == 0:
            ok = True
            for r in range(world):
                xr = synth_image(B, img, img, seed=100 + r, dtype=torch.uint8).to(dev)
                o_r, m_r = model.detect(xr, CONF, IOU, max_det=MAX_DET, packed=True)
                c_r = m_r[:B]
                blk_o, blk_c = out_all[r * B:(r + 1) * B], cnt_all[r * B:(r + 1) * B]
                ok &= bool(torch.equal(blk_c.to(torch.int32), c_r.to(torch.int32)))
                for i in range(B):
                    n = int(c_r[i])
                    ok &= bool(torch.equal(blk_o[i, :n], o_r[i, :n]))
            gather_check = 'bit-identical' if ok else 'MISMATCH'
        barrier()
    if pg is not None:
        for g in pgs:
            g.check()
            g.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    eng = model.run_trunk(x_dev)         # the engine of the main stream (per-op profile, serial trunk / NMS times)
    torch.cuda.synchronize()
    n_imgs = B * world * args.steps
    launches_per_step = 1 + eng.num_launches() + 2     # stage + trunk ops + k_filter + k_sort_nms
    line = {
        'metric': metric_name(model_name, img), 'value': n_imgs / (ms_dev / 1e3),
        'unit': 'images/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': ms_dev / args.steps, 'higher_is_better': True, 'scaling': scaling, 'vs_baseline': None,
        'dtype': 'bf16', 'data': 'synthetic',
        'value_note': 'inputs resident in HBM (uint8 batch on the device); the host-to-host figure is e2e.value',
        'value_one_stream': n_imgs / (ms_serial / 1e3), 'ms_per_step_one_stream': ms_serial / args.steps,
        'config': {'workload': workload_name(model_name, img, B),
                   'batch_per_gpu': B, 'global_batch': B * world, 'parallelism': f'image-sharded replicas x{world}',
                   'gather': ('none' if world == 1 else 'in-kernel P2P stores over NVLink (yb_decode_nms_gather)' if use_peer
                              else 'NCCL all_gather'),
                   'streams': n_streams,
                   'in_flight': (f'{n_streams} batches in flight: steps alternate between {n_streams} streams, one engine '
                                 'each; every step is one full pass over one batch (value_one_stream = strictly serial)'
                                 if n_streams > 1 else 'strictly serial steps on one stream'),
                   'l2': 'no flush: per-step working set (activations, GBs) >> 126 MB L2',
                   'input': 'one seeded pinned uint8 batch per rank, re-sent H2D every e2e step'},
        'e2e': {'value': n_imgs / (ms_e2e / 1e3), 'unit': 'images/s', 'h2d_bytes_per_step': x_host.numel(),
                'd2h_bytes_per_step': int(res.numel() * 4 + cnt.numel() * 4 + 4), 'ms_per_step': ms_e2e / args.steps,
                'api': f'Model.detect_batches (H2D / D2H copy streams + {n_streams} compute stream(s), {2 if n_streams == 1 else 3 * n_streams} slots)',
                'serial_value': n_imgs / (ms_e2e_serial / 1e3),
                'serial_api': 'Model.detect(x_host) + .cpu() per step, one stream'},
        'gpu_launches': launches_per_step * args.steps,
        'clocks': clocks,
        'detections_per_image': float(cnt.float().mean()),
    }
    if gather_check is not None:
        line['gather_check'] = gather_check
    if not args.no_extras:
        # ---- per-kernel rooflines; `roofline` = the kernel with the largest share of the step
        traffic, traffic_src = {}, None
        try:   # DRAM bytes per kernel of one step from the committed ncu launch list (cold-cache per launch)
            if not headline:
                raise KeyError('launch list exists for the headline workload only')
            tp = os.path.join(ROOT, 'profiles', 'round2_launch_summary.json')
            if not os.path.exists(tp):
                tp = os.path.join(ROOT, 'profiles', 'round1_launch_summary.json')
            for k in json.load(open(tp))['kernels']:        # template instances (k_conv_tc<0> / <1>) add up
                base = k['kernel'].split('<')[0]
                traffic[base] = traffic.get(base, 0) + k['dram_bytes']
            traffic_src = os.path.relpath(tp, ROOT) + ' (ncu dram__bytes_read+write summed over the kernel\'s launches of one step)'
        except Exception:  # noqa: BLE001
            pass
        kernels, summary, table = kernel_rooflines(model, eng, peaks, peak_src, traffic, step_ms=ms_dev / args.steps, x_dev=x_dev)
        top = dict(kernels[0])
        line['roofline'] = {'bound': top['bound'], 'achieved': top['achieved'], 'peak': top['peak'], 'unit': top['unit'],
                            'frac': top['frac'], 'traffic': top['traffic'], 'kernel': top['kernel'],
                            'launches_per_step': top['launches'], 'ms_per_step': top['ms'],
                            'frac_of_layer_bound': top['frac_of_layer_bound'],
                            'in_graph_ms': top.get('in_graph_ms'),
                            'in_graph_frac_of_layer_bound': top.get('in_graph_frac_of_layer_bound'),
                            'peak_source': peak_src + (' (sustained bf16 cuBLAS)' if top['bound'] == 'tensor' else ' (HBM copy)'),
                            'traffic_source': traffic_src, **summary}
        line['roofline_kernels'] = kernels
        line['top_ops'] = [{'op': r['op'], 'kernel': r['kernel'], 'ms_eager': round(r['ms'], 4),
                            'tflops': round(r['flops'] / r['ms'] / 1e9, 1) if r['flops'] else None,
                            'GBps': round(r['bytes'] / r['ms'] / 1e6)} for r in sorted(table, key=lambda r: -r['ms'])[:8]]
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
                                    'cpu': cpu_model(),
                                    'sample': f'{nimg * reps} images ({reps} steps of {nimg}) of the same workload, fp32 torch CPU + oracle NMS'}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
