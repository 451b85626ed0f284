#!/usr/bin/env python
"""Benchmark of the DeltaConv hot path on B200 (contract: see DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W]           # this repo's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]     # the CPU oracle port on the host cores

A step = one training step (forward + backward + gradient all-reduce + SGD) of
DeltaNetClassification on BASELINE.json configs[1]: 32 clouds x 1024 points per GPU, k=20, 40
classes, synthetic data, random-init weights.  One JSON line is printed by rank 0.
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

import torch                                                   # noqa: E402

CLOUDS, POINTS, K_NN, CLASSES = 32, 1024, 20, 40
WORKLOAD = ("configs[1]: DeltaNetClassification training (fwd+bwd+SGD), 32 clouds x 1024 pts per GPU, k=20, "
            "40 classes, convs [64,64,128,256], normals given")
METRIC = "DeltaNetClassification training throughput (fwd+bwd+SGD), clouds/s"


# ------------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------------

def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p['hbm_gbs'], tensor_burst=p['bf16_tflops'], tensor=p.get('bf16_tflops_sustained', p['bf16_tflops']),
                    source='measured (MEASURED_PEAKS.json)')
    return dict(hbm=6650.0, tensor_burst=1590.0, tensor=1400.0, source='fallback (B200_PROFILING.md)')


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed region runs (a default run's timed region is ~0.25 s)."""
    QUERY = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                                          '-lms', '50', '-i', str(gpu_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(',')]))

    def summary(self, t0, t1):
        if self.proc is not None:
            self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 - 0.25 <= ts <= t1 + 0.25 and len(r) >= 6] or [r for _, r in self.rows if len(r) >= 6]
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['unavailable'])
        sm = [float(r[0]) for r in rows if r[0].replace('.', '').isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith('active') for r in rows)]
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=float(rows[0][1]) if rows[0][1].replace('.', '').isdigit() else None,
                    reasons=reasons, samples=len(rows))


# algorithmic work per entry point: name -> (args -> (bytes, flops)), family
def _w_knn(a): N, k = a[3], a[4]; return N * (12 + 4 * k + (16 * k if a[6] else 0)), 0
def _w_build(a): N, k = a[7], a[8]; return N * (48 + 20 * k), 0
def _w_grad(a): N, k, C = a[2], a[3], a[4]; return N * (12 * k + 12 * C), 0
def _w_grad_t(a): N, k, C = a[3], a[4], a[5]; return N * (12 * k + 12 * C), 0
def _w_dcn(a): N, k, C = a[2], a[3], a[4]; return N * (12 * k + 20 * C + (8 * C if a[13] else 0)), 0     # + x copy-through
def _w_dcn_t(a): N, k, C = a[3], a[4], a[5]; return N * (12 * k + 20 * C + (8 * C if a[12] else 0)), 0
def _w_hodge(a): N, k, C = a[2], a[3], a[4]; return N * (12 * k + 16 * C + (16 * C if a[11] else 0)), 0  # + v copy-through
def _w_hodge_t(a): N, k, C = a[3], a[4], a[5]; return N * (12 * k + 16 * C), 0
def _w_gmax(a): N, k, C = a[1], a[2], a[3]; return N * (4 * k + 9 * C), 0
def _w_gmax_t(a): N, k, C = a[2], a[3], a[4]; return N * (4 * k + 9 * C), 0
def _w_stats(a): M, C, mode = a[2], a[3], a[4]; return M * C * 4 * (2 if mode else 1), 0
def _w_bnf(a): M, C = a[2], a[3]; return M * C * (8 + (4 if a[9] else 0)), 0
def _w_bnb(a): M, C = a[4], a[5]; return M * C * 12, 0
def _w_vnf(a): M, C = a[2], a[3]; return M * C * 16, 0
def _w_vnb(a): M, C = a[4], a[5]; return M * C * 24, 0
def _w_lin(a): M, K, N = a[5], a[6], a[7]; return 4 * (M * K + K * N + M * N), 2 * M * K * N
def _w_linb(a): M, K, N = a[4], a[5], a[6]; return 4 * (M * K + K * N + M * N), 2 * M * K * N
def _w_ediff(a): N, k, C = a[1], a[2], a[3]; return N * (4 * k + 4 * C + 4 * k * C), 0
def _w_segmax(a): N, k, C = a[1], a[2], a[3]; return N * C * (4 * k + 5), 0
def _w_segmax_t(a): N, k, C = a[2], a[3], a[4]; return N * C * (4 * k + 5), 0
def _w_pool(a): return 0, 0
def _w_sgd(a): return a[3] * 20, 0                               # p, buf read and written, g read
def _w_bnpool(a): return CLOUDS * POINTS * a[4] * 4, 0          # y read once, only [B,2C] written
def _w_bnpool_t(a): M, C = a[2], a[3]; return M * C * 8, 0    # y read, dy written


WORK = {
    'dc_knn': (_w_knn, 'knn'), 'dc_build_grad_div': (_w_build, 'build_grad_div'),
    'dc_grad_fwd': (_w_grad, 'spmm'), 'dc_grad_bwd': (_w_grad_t, 'spmm'),
    'dc_divcurlnorm_fwd': (_w_dcn, 'spmm'), 'dc_divcurlnorm_bwd': (_w_dcn_t, 'spmm'),
    'dc_hodge_fwd': (_w_hodge, 'spmm'), 'dc_hodge_bwd': (_w_hodge_t, 'spmm'),
    'dc_gathermax_fwd': (_w_gmax, 'gathermax'), 'dc_gathermax_bwd': (_w_gmax_t, 'gathermax'),
    'dc_bn_act_pool_fwd': (_w_bnpool, 'batchnorm'), 'dc_bn_act_pool_bwd': (_w_bnpool_t, 'batchnorm'),
    'dc_sgd_step': (_w_sgd, 'optimizer'),
    'dc_bn_stats': (_w_stats, 'batchnorm'), 'dc_bn_act_fwd': (_w_bnf, 'batchnorm'), 'dc_bn_act_bwd': (_w_bnb, 'batchnorm'),
    'dc_vecnonlin_fwd': (_w_vnf, 'vecnonlin'), 'dc_vecnonlin_bwd': (_w_vnb, 'vecnonlin'),
    'dc_linear_fwd': (_w_lin, 'gemm'), 'dc_linear_bwd_input': (_w_linb, 'gemm'), 'dc_linear_bwd_weight': (_w_linb, 'gemm'),
    'dc_edge_diff_fwd': (_w_ediff, 'centralized'), 'dc_segmax_fwd': (_w_segmax, 'centralized'),
    'dc_segmax_bwd': (_w_segmax_t, 'centralized'),
}
KERNEL_OF = {'gemm': 'tc_pair_gemm_kernel / tc_gemm_kernel / sgemm_kernel (dc_linear_fwd / _bwd_input / _bwd_weight)', 'spmm': 'spmm_kernel (grad/div/curl/hodge and transposes)',
             'batchnorm': 'col_reduce/bn_act kernels', 'knn': 'knn_kernel', 'build_grad_div': 'fit_kernel (+ edge/cloud means)',
             'gathermax': 'gathermax kernels', 'vecnonlin': 'vecnonlin kernels', 'centralized': 'edge_diff/segmax kernels'}


class EntryPointProfiler:
    """CUDA-event timing of every library entry point on the launching stream."""

    def __init__(self):
        self.records = []

    def begin(self, name, args):
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        work = WORK.get(name)
        wb, wf = work[0](args) if work else (0, 0)
        return (name, work[1] if work else 'other', wb, wf, s, e, tuple(a for a in args if isinstance(a, int) and 0 < a < (1 << 31)))

    def end(self, token):
        token[5].record()
        self.records.append(token)

    def families(self):
        torch.cuda.synchronize()
        fam = {}
        for name, family, wb, wf, s, e, _dims in self.records:
            f = fam.setdefault(family, dict(ms=0.0, bytes=0, flops=0, launches=0))
            f['ms'] += s.elapsed_time(e)
            f['bytes'] += wb
            f['flops'] += wf
            f['launches'] += 1
        return fam


    def detail(self, steps, top=40):
        torch.cuda.synchronize()
        calls = {}
        for name, family, wb, wf, s, e, dims in self.records:
            c = calls.setdefault((name, dims), [0, 0.0, wb, wf])
            c[0] += 1
            c[1] += s.elapsed_time(e)
        lines = []
        for (name, dims), (n, ms, wb, wf) in sorted(calls.items(), key=lambda kv: -kv[1][1])[:top]:
            sec = ms / n / 1000
            lines.append(f"{ms / steps:8.4f} ms/step  x{n // steps:<3d} {name:24s} {str(dims):44s} "
                         f"{wb / sec / 1e9:8.1f} GB/s {wf / sec / 1e12:7.2f} TF/s")
        return '\n'.join(lines)


def smoothed_loss(logits, target):
    from deltaconv_b200.utils.losses import calc_loss
    return calc_loss(logits, target)


# ------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle port of the reference's algorithm on the host cores
# ------------------------------------------------------------------------------------------------

def cpu_port_steps(steps, warmup, clouds):
    """Training steps of the oracle port (oracle/network.py, fast cdist+topk kNN) on the CPU."""
    from oracle import network as on
    from deltaconv_b200.utils import synthetic_batch, fill_deterministic
    torch.manual_seed(1)
    model = fill_deterministic(on.DeltaNetClassification(3, CLASSES, num_neighbors=K_NN), 1)
    model.deltanet_base.fast_knn = True
    model.train()
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    data = synthetic_batch(clouds, POINTS, config=2, num_classes=CLASSES)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()
        loss = on.smoothed_cross_entropy(model(data), data.y)
        loss.backward()
        opt.step()
        loss.item()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def run_reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)          # torchrun pins OMP_NUM_THREADS=1; use all host cores
    cores = torch.get_num_threads()
    clouds = 4
    probe = cpu_port_steps(1, 0, 2)[0] / 2                                   # size the sample: seconds per cloud
    budget = 150.0
    while clouds > 2 and probe * clouds * (args.steps + args.warmup) > budget:   # >= 2 clouds: BatchNorm in the head
        clouds //= 2
    steps = args.steps
    while steps > 1 and probe * clouds * (steps + args.warmup) > budget:
        steps -= 1
    times = cpu_port_steps(steps, args.warmup, clouds)
    ms = 1000 * sum(times) / len(times)
    value = clouds / (ms / 1000)
    sample = f"{steps} training steps of {clouds} clouds x {POINTS} pts (k={K_NN}) through the oracle port (torch CPU, cdist+topk kNN)"
    line = dict(metric=METRIC, value=round(value, 3), unit='clouds/s', impl='reference', n_gpus=args.gpus, steps=steps,
                warmup=args.warmup, ms_per_step=round(ms, 2), higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='f32', data='synthetic', config=dict(workload=WORKLOAD, sample_clouds_per_step=clouds),
                cpu_baseline=dict(value=round(value, 3), unit='clouds/s', cores=cores, kind='port', sample=sample),
                e2e=dict(value=round(value, 3), unit='clouds/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------

def run_cuda_arm(args):
    import torch.distributed as dist
    import deltaconv_b200 as D
    from deltaconv_b200 import _lib
    from deltaconv_b200.parallel import FlatGradients, FlatSGD

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device — the CUDA arm has no CPU fallback (use --impl reference for the CPU port)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # stdout must carry exactly one JSON line: NCCL writes its version banner (at WARN and above) to stdout, so its log is
        # off unless asked for through DC_NCCL_DEBUG, and then it goes to stderr
        os.environ.pop('NCCL_DEBUG', None)
        if os.environ.get('DC_NCCL_DEBUG'):
            os.environ['NCCL_DEBUG'] = os.environ['DC_NCCL_DEBUG']
            os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=dev)
    lib = _lib.load()

    torch.manual_seed(1)
    model = D.utils.fill_deterministic(D.models.DeltaNetClassification(3, CLASSES, num_neighbors=K_NN), 1).to(dev).train()
    flat = FlatGradients(model.parameters())
    opt = FlatSGD(flat, lr=0.1, momentum=0.9, weight_decay=1e-4)          # torch.optim.SGD's update as one kernel on flat buffers
    host = D.utils.synthetic_batch(CLOUDS, POINTS, config=2, rank=rank, num_classes=CLASSES).pin_memory()
    host._num_clouds, host._max_points = CLOUDS, POINTS
    resident = host.to(dev)
    resident._num_clouds, resident._max_points = CLOUDS, POINTS
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)             # > 126 MB L2

    def step(data):
        flat.zero()
        loss = smoothed_loss(model(data), data.y)
        loss.backward()
        flat.all_reduce()
        opt.step()
        return loss

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        """K steps bracketed by barrier + synchronize; L2 flushed between steps; device time, max over ranks."""
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            flush.zero_()
            fn()
        t1.record()
        barrier()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    for _ in range(max(args.warmup, 3)):
        step(resident)

    graphed = None
    if not args.no_graph:
        static = host.to(dev)
        static._num_clouds, static._max_points = CLOUDS, POINTS
        graphed = D.utils.GraphedTrainStep(model, opt, smoothed_loss, static, flat=flat)
        for _ in range(3):
            graphed()

    # (1) device-resident throughput
    sampler = ClockSampler(local) if rank == 0 else None
    wall0 = time.time()
    lib.dc_reset_launch_count()
    eager_ms = timed(lambda: step(resident), args.steps)
    launches = int(lib.dc_launch_count())
    total_ms = timed(graphed, args.steps) if graphed else eager_ms
    wall1 = time.time()
    clocks = sampler.summary(wall0, wall1) if sampler else None

    # (2) end to end: pinned host buffers -> device inside the timed region, loss read back every step
    def e2e_step():
        data = host.to(dev, non_blocking=True)
        data._num_clouds, data._max_points = CLOUDS, POINTS
        return step(data).item()
    # Graph replay: the next batch's host->device copy runs on a side stream while the current step computes (prefetch /
    # commit), and each step's loss is copied to pinned host memory asynchronously and read one step later — the usual
    # input pipeline of a training loop.  Every step still moves its own inputs H2D and its own loss D2H inside the timed region.
    loss_host = torch.empty(2, dtype=torch.float32).pin_memory()
    loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
    pipe = dict(i=0, last=None)
    def e2e_graph_step():
        i = pipe['i']
        graphed.commit()
        loss = graphed()
        loss_host[i & 1].copy_(loss, non_blocking=True)
        loss_ready[i & 1].record()
        graphed.prefetch(host)
        if i > 0:
            loss_ready[(i - 1) & 1].synchronize()
            pipe['last'] = float(loss_host[(i - 1) & 1])
        pipe['i'] = i + 1
        return pipe['last']
    e2e_fn = e2e_graph_step if graphed else e2e_step
    if graphed:
        graphed.prefetch(host)
    for _ in range(2):
        e2e_fn()
    e2e_ms = timed(e2e_fn, args.steps)
    e2e_eager_ms = timed(e2e_step, args.steps) if graphed else e2e_ms

    # (3) per-entry-point device time (CUDA events on the launching stream), same steps
    prof = EntryPointProfiler()
    _lib.PROFILER = prof
    barrier()
    # The device is parked on a spin kernel while the host enqueues the whole step, so that each event pair brackets
    # device time only (an eager step is host-bound: without this the gaps between launches would be counted).
    park_cycles = int(3.0 * eager_ms / args.steps * 1e-3 * 2.0e9)
    spans = []
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda._sleep(park_cycles)
        w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0.record()
        step(resident)
        w1.record()
        spans.append((w0, w1))
        torch.cuda.synchronize()
    _lib.PROFILER = None
    fam = prof.families()
    prof_total_ms = sum(a.elapsed_time(b) for a, b in spans)
    if args.detail and rank == 0:
        open(args.detail, "w").write(prof.detail(args.steps, top=120) + "\n")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    ms_per_step = total_ms / args.steps
    value = world * CLOUDS / (ms_per_step / 1000)
    e2e_value = world * CLOUDS / (e2e_ms / args.steps / 1000)
    top = max(fam, key=lambda f: fam[f]['ms'])
    kernels = {}
    for f, r in sorted(fam.items(), key=lambda kv: -kv[1]['ms']):
        sec = r['ms'] / 1000
        kernels[f] = dict(ms_per_step=round(r['ms'] / args.steps, 4), share=round(r['ms'] / prof_total_ms, 4),
                          calls_per_step=r['launches'] // args.steps,
                          gbps=round(r['bytes'] / sec / 1e9, 1) if sec > 0 else None,
                          tflops=round(r['flops'] / sec / 1e12, 2) if r['flops'] else None)
    tr = fam[top]
    traffic, ncu = None, {}
    tpath = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    if os.path.exists(tpath):
        ncu = json.load(open(tpath)).get(top) or {}
        traffic = ncu.get('dram_bytes_per_launch_avg')
    if tr['flops']:
        ach = tr['flops'] / (tr['ms'] / 1000) / 1e12
        mode = int(lib.dc_get_gemm_mode())
        roof = dict(bound='tensor', kernel=KERNEL_OF.get(top, top), achieved=round(ach, 3), peak=pk['tensor'], unit='TFLOP/s',
                    frac=round(ach / pk['tensor'], 5), traffic=traffic, peak_source=pk['source'] + ', sustained bf16 (kernel timed inside a long step)',
                    share_of_step=kernels[top]['share'],
                    note=('achieved counts the 2*M*K*N algorithmic flops; in 3xTF32 mode the tensor cores execute 3x that in TF32 '
                          '(whose dense peak is half the bf16 peak)') if mode == 3 else None,
                    executed_tf32_tflops=round(ach * (3 if mode == 3 else 1), 2),
                    algorithmic_bytes_per_launch=round(tr['bytes'] / max(tr['launches'], 1)),
                    tensor_busy_share_in_kernel=ncu.get('in_kernel_mma_share_large_layers'))
    else:
        ach = tr['bytes'] / (tr['ms'] / 1000) / 1e9
        roof = dict(bound='hbm', kernel=KERNEL_OF.get(top, top), achieved=round(ach, 1), peak=pk['hbm'], unit='GB/s',
                    frac=round(ach / pk['hbm'], 5), traffic=traffic, peak_source=pk['source'], share_of_step=kernels[top]['share'])
    hbm_fams = {f: r for f, r in fam.items() if not r['flops'] and r['bytes']}
    hbm_bytes = sum(r['bytes'] for r in hbm_fams.values())
    hbm_ms = sum(r['ms'] for r in hbm_fams.values())

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        cores = torch.get_num_threads()
        t = cpu_port_steps(2, 1, 4)
        v = 4 / (sum(t) / len(t))
        cpu = dict(value=round(v, 3), unit='clouds/s', cores=cores, kind='port',
                   sample=f"2 training steps (after 1 warm-up) of 4 clouds x {POINTS} pts, k={K_NN}, oracle port (torch CPU, cdist+topk kNN)")

    line = dict(metric=METRIC, value=round(value, 2), unit='clouds/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=round(ms_per_step, 4), higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                data='synthetic',
                config=dict(workload=WORKLOAD, clouds_per_gpu=CLOUDS, points=POINTS, k=K_NN, parallelism=f'dp{world} (shard by cloud, flat-gradient all-reduce)',
                            l2='256 MiB buffer rewritten between steps (inside the timed region)',
                            gemm='tcgen05 kind::tf32, 3xTF32 split (fp32-grade accuracy), fp32 accumulate; FFMA kernel for ineligible shapes',
                            launch='CUDA-graph replay of the whole step' if graphed else 'eager (one Python launch per kernel)',
                            eager_ms_per_step=round(eager_ms / args.steps, 4), e2e_eager_ms_per_step=round(e2e_eager_ms / args.steps, 4)),
                e2e=dict(value=round(e2e_value, 2), unit='clouds/s', h2d_bytes_per_step=host.nbytes(), d2h_bytes_per_step=4,
                         ms_per_step=round(e2e_ms / args.steps, 4),
                         pipeline='next batch copied on a side stream during the step, loss read one step later' if graphed else 'serial'),
                gpu_launches=launches, roofline=roof, cpu_baseline=cpu, clocks=clocks,
                kernels=kernels,
                hbm_bound_kernels=dict(gbps=round(hbm_bytes / (hbm_ms / 1000) / 1e9, 1) if hbm_ms else None,
                                       frac_of_peak=round(hbm_bytes / (hbm_ms / 1000) / 1e9 / pk['hbm'], 4) if hbm_ms else None))
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# second workload (not the contract line): the dataset pre-transform, geodesic farthest-point sampling
# ------------------------------------------------------------------------------------------------

FPS_CLOUDS, FPS_POINTS, FPS_SAMPLES = 296, 8192, 1024       # train_modelnet.py:32-33: 8 x 1024 sampled points -> 1024


def _fps_clouds(count, seed=5):
    import numpy as np
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(count):
        u = rng.standard_normal((FPS_POINTS, 3))
        u /= np.linalg.norm(u, axis=1, keepdims=True)
        r = 1 + 0.15 * np.sin(3 * np.arctan2(u[:, 1], u[:, 0])) * u[:, 2]
        out.append((u * r[:, None] * rng.uniform(0.8, 1.25, 3)).astype(np.float32))
    return out


def run_fps_workload(args):
    """One step = sampling FPS_CLOUDS clouds of 8192 points down to 1024 in one call of ops.geodesic_fps.  CPU beside it: the
    reference's own sampler (oracle/_ref/libref_fps.so) where it was built, else the C restatement, one cloud per host thread."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    import deltaconv_b200 as D
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    base = _fps_clouds(8)
    host = torch.from_numpy(np.concatenate([base[b % 8] for b in range(FPS_CLOUDS)])).pin_memory()
    batch = torch.arange(FPS_CLOUDS, device=dev).repeat_interleave(FPS_POINTS)
    start = torch.arange(FPS_CLOUDS, dtype=torch.int32, device=dev) * 7 % FPS_POINTS
    pos = host.to(dev)
    kw = dict(batch=batch, num_clouds=FPS_CLOUDS, max_points=FPS_POINTS, start=start)
    out_host = torch.empty(FPS_CLOUDS, FPS_SAMPLES, dtype=torch.int64).pin_memory()
    for _ in range(max(args.warmup, 3)):
        D.ops.geodesic_fps(pos, FPS_SAMPLES, **kw)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    D._lib.load().dc_reset_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    t0 = time.time()
    ev[0].record()
    for _ in range(args.steps):
        D.ops.geodesic_fps(pos, FPS_SAMPLES, **kw)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / args.steps
    launches = int(D._lib.load().dc_launch_count())
    t_e2e = time.perf_counter()
    for _ in range(args.steps):                       # host buffers in, host buffer out
        out_host.copy_(D.ops.geodesic_fps(host.to(dev, non_blocking=True), FPS_SAMPLES, **kw), non_blocking=True)
        torch.cuda.synchronize()
    e2e_ms = 1000 * (time.perf_counter() - t_e2e) / args.steps
    clocks = sampler.summary(t0, time.time())
    line = dict(metric='geodesic farthest-point sampling (pre-transform), clouds/s', value=round(FPS_CLOUDS / (ms / 1000), 1), unit='clouds/s',
                n_gpus=1, steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=round(ms, 3), higher_is_better=True, scaling='weak',
                vs_baseline=None, dtype='f64', data='synthetic',
                config=dict(workload=f'{FPS_CLOUDS} clouds x {FPS_POINTS} pts -> {FPS_SAMPLES} samples per call (train_modelnet.py:32-33), 10-NN graph in fp64',
                            l2='no flush: one call streams 58 MB of positions and builds a 310 MB graph (more than L2); the sampler is bound by DRAM latency per relaxation round, not by bandwidth'),
                e2e=dict(value=round(FPS_CLOUDS / (e2e_ms / 1000), 1), unit='clouds/s', h2d_bytes_per_step=host.numel() * 4,
                         d2h_bytes_per_step=out_host.numel() * 8),
                gpu_launches=launches, clocks=clocks)
    pk = peaks()
    algorithmic = host.numel() * 8 + out_host.numel() * 4          # fp64 positions read once, int32 samples written once
    line['roofline'] = dict(bound='hbm', kernel='dc_geodesic_fps (fps_knn_kernel + fps_sample_kernel)', achieved=round(algorithmic / (ms / 1000) / 1e9, 2),
                            peak=pk['hbm'], unit='GB/s', frac=round(algorithmic / (ms / 1000) / 1e9 / pk['hbm'], 5), traffic=None,
                            peak_source=pk['source'],
                            note='not a bandwidth-bound path: the graph kernel is bound by instruction issue, the sampler by DRAM latency per '
                                 'relaxation round (profiles/r01_fps_ncu.md); the fraction is reported for completeness')
    if not args.no_cpu_baseline:
        from oracle import fps as ofps
        cores = os.cpu_count() or 1
        kind = 'reference' if ofps.have_reference() else 'port'
        one = (lambda x: ofps.reference_fps(x, FPS_SAMPLES)) if kind == 'reference' else (lambda x: ofps.restated_fps(x, FPS_SAMPLES, 0))
        sample = [base[i % 8] for i in range(4 * cores)]
        t = time.perf_counter()
        with ThreadPoolExecutor(cores) as pool:
            list(pool.map(one, sample))
        cpu = len(sample) / (time.perf_counter() - t)
        line['cpu_baseline'] = dict(value=round(cpu, 1), unit='clouds/s', cores=cores, kind=kind,
                                    sample=f'{len(sample)} clouds of the same shape, one cloud per host thread ({cores} threads)')
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='cuda', choices=['cuda', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-graph', action='store_true', help='time eager launches instead of CUDA-graph replays')
    ap.add_argument('--detail', default=None, help='write the per-call device-time table to this file')
    ap.add_argument('--workload', default='train', choices=['train', 'fps'],
                    help="'fps': time the geodesic farthest-point sampling pre-transform instead of the training step (1 GPU)")
    args = ap.parse_args()
    if args.workload == 'fps':
        return run_fps_workload(args)
    if args.impl == 'reference':
        run_reference_arm(args)
    else:
        run_cuda_arm(args)


if __name__ == '__main__':
    main()
