#!/usr/bin/env python
"""Benchmark of the DeltaConv hot path on B200 (contract: see DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload train|seg|infer]   # this repo's CUDA path
    python bench.py --impl reference [--steps K] [--warmup W] [--workload ...]          # the CPU arm on the host cores

A step = one pass of the workload over one synthetic batch with random-init weights:
  train (default, the contract line)  configs[1]: DeltaNetClassification training step (forward + backward + gradient
                                      all-reduce + SGD), 32 clouds x 1024 points per GPU, k=20, 40 classes
  seg                                 configs[2]: DeltaNetSegmentation training step, 16 clouds x 2048 points per GPU, k=30
  infer                               configs[4]: DeltaNetClassification evaluation forward, 32 clouds x 2048 points per GPU,
                                      k=20, no normals (estimated frames), no collective
One JSON line is printed by rank 0.
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

# The three model workloads BASELINE.json names (c2, c3, c5).  `train` is the contract line the driver runs.
WORKLOADS = {
    'train': dict(
        tag='configs[1]', kind='cls', training=True, clouds=32, points=1024, k=20, classes=40, convs=[64, 64, 128, 256], normals=True,
        synth=2, smoothing=True,
        workload=("configs[1]: DeltaNetClassification training (fwd+bwd+SGD), 32 clouds x 1024 pts per GPU, k=20, "
                  "40 classes, convs [64,64,128,256], normals given"),
        metric="DeltaNetClassification training throughput (fwd+bwd+SGD), clouds/s"),
    'seg': dict(
        tag='configs[2]', kind='seg', training=True, clouds=16, points=2048, k=30, classes=50, convs=[64, 128, 256], normals=True,
        synth=3, smoothing=False,
        workload=("configs[2]: DeltaNetSegmentation training (fwd+bwd+SGD), ShapeNet-part shape: 16 clouds x 2048 pts per GPU, "
                  "k=30, convs [64,128,256] with depth-2 MLPs, embedding 1024, categorical vector, 50 part classes, normals given "
                  "(experiments/train_shapenet.py:79-89,169)"),
        metric="DeltaNetSegmentation training throughput (fwd+bwd+SGD), clouds/s"),
    'infer': dict(
        tag='configs[4]', kind='cls', training=False, clouds=32, points=2048, k=20, classes=15, convs=[64, 64, 64, 128], normals=False,
        synth=5, smoothing=True,
        workload=("configs[4]: DeltaNetClassification inference (eval forward), ScanObjectNN shape: 32 clouds x 2048 pts per GPU, "
                  "k=20, 15 classes, convs [64,64,64,128], no normals (second kNN k=10 + estimated frames), sharded by cloud, "
                  "no collective (experiments/train_scanobjectnn.py:69-76)"),
        metric="DeltaNetClassification inference throughput (eval forward), clouds/s"),
    'infer_depth2': dict(
        tag='configs[4] (depth-2 reading)', kind='cls', training=False, clouds=32, points=2048, k=20, classes=15, convs=[64, 64, 64, 128],
        normals=False, synth=5, smoothing=True, mlp_depth=2,
        workload=("configs[4] read literally ('depth-2 vector MLPs'): the same ScanObjectNN-shape inference with the backbone built "
                  "as DeltaNetBase(mlp_depth=2) under the classification head (DeltaNetClassification itself fixes depth 1, "
                  "deltanet_classification.py:31); 32 clouds x 2048 pts per GPU, k=20, no normals"),
        metric="DeltaNetClassification (depth-2 backbone) inference throughput (eval forward), clouds/s"),
}
WL = WORKLOADS['train']                 # set by main()


def make_model(pkg, wl):
    """The workload's architecture from `pkg` (deltaconv_b200.models, oracle.network or the reference's deltaconv.models)."""
    if wl['kind'] == 'seg':
        return pkg.DeltaNetSegmentation(3, wl['classes'], conv_channels=list(wl['convs']), mlp_depth=2, embedding_size=1024,
                                        categorical_vector=True, num_neighbors=wl['k'])
    model = pkg.DeltaNetClassification(3, wl['classes'], conv_channels=list(wl['convs']), num_neighbors=wl['k'])
    if wl.get('mlp_depth', 1) != 1:                  # same head over a deeper backbone (deltanet_base.py:9,39)
        base_cls = type(model.deltanet_base)
        model.deltanet_base = base_cls(3, list(wl['convs']), wl['mlp_depth'], wl['k'], 1e-3, 1)
    return model


def make_batch(wl, rank=0, clouds=None):
    from deltaconv_b200.utils import synthetic_batch
    seg = wl['kind'] == 'seg'
    return synthetic_batch(clouds or wl['clouds'], wl['points'], config=wl['synth'], rank=rank, normals=wl['normals'],
                           num_classes=wl['classes'], per_point_labels=seg, categories=16 if seg else 0)


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
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw')

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
        power = [float(r[6]) for r in rows if len(r) > 6 and r[6].replace('.', '').isdigit()]
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=float(rows[0][1]) if rows[0][1].replace('.', '').isdigit() else None,
                    reasons=reasons, samples=len(rows), power_w_median=statistics.median(power) if power else None,
                    power_w_max=max(power) if power else None)


# algorithmic work per entry point: name -> (args -> (bytes, flops)), family
def _w_knn(a): N, k = a[3], a[4]; return N * (12 + 4 * k + (16 * k if a[6] else 0) + (4 * k if a[7] else 0)), 0      # + edge_index, + d2
def _w_knn_grid(a): N, k = a[3], a[4]; return N * (12 + 4 * k + (16 * k if a[6] else 0) + (4 * k if a[7] else 0)), 0
def _w_basis(a): N, k = a[2], a[3]; return N * (48 + 4 * k), 0
def _w_build(a): N, k = a[8], a[9]; return N * (48 + 20 * k), 0
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
def _w_segmax_bn(a): N, k, C = a[1], a[2], a[3]; return N * C * (4 * k + 5), 0                  # y [N*k,C] read, out + arg written
def _w_segmax_bn_t(a): N, k, C = a[1], a[2], a[3]; return N * C * (8 * k + 5), 0               # y read, dy written
def _w_cedge(a): N, k, Cout = a[1], a[2], a[4]; return N * (12 + 4 * k + 4 * k * Cout), 0      # coordinates + idx read, h [N*k,Cout] written
def _w_cedge_t(a): N, k, Cout = a[1], a[2], a[4]; return N * (12 + 4 * k + 4 * k * Cout), 0    # dh read once
def _w_pool(a): return 0, 0
def _w_sgd(a): return a[3] * 20, 0                               # p, buf read and written, g read
def _w_bnpool(a): return WL['clouds'] * WL['points'] * a[4] * 4, 0          # y read once, only [B,2C] written
def _w_bnpool_t(a): M, C = a[2], a[3]; return M * C * 8, 0    # y read, dy written


WORK = {
    'dc_knn': (_w_knn, 'knn'), 'dc_knn_grid': (_w_knn_grid, 'knn'), 'dc_build_grad_div': (_w_build, 'build_grad_div'),
    'dc_estimate_basis': (_w_basis, 'frames'),
    'dc_grad_fwd': (_w_grad, 'spmm'), 'dc_grad_bwd': (_w_grad_t, 'spmm'),
    'dc_divcurlnorm_fwd': (_w_dcn, 'spmm'), 'dc_divcurlnorm_bwd': (_w_dcn_t, 'spmm'),
    'dc_hodge_fwd': (_w_hodge, 'spmm'), 'dc_hodge_bwd': (_w_hodge_t, 'spmm'),
    'dc_gathermax_fwd': (_w_gmax, 'gathermax'), 'dc_gathermax_bwd': (_w_gmax_t, 'gathermax'), 'dc_gathermax_bn_fwd': (_w_gmax, 'gathermax'),
    'dc_bn_act_pool_fwd': (_w_bnpool, 'batchnorm'), 'dc_bn_act_pool_bwd': (_w_bnpool_t, 'batchnorm'),
    'dc_sgd_step': (_w_sgd, 'optimizer'), 'dc_pool_fwd': (_w_pool, 'pool'), 'dc_pool_bwd': (_w_pool, 'pool'), 'dc_bcast_rows': (_w_pool, 'pool'),
    'dc_bn_stats': (_w_stats, 'batchnorm'), 'dc_bn_act_fwd': (_w_bnf, 'batchnorm'), 'dc_bn_act_bwd': (_w_bnb, 'batchnorm'),
    'dc_vecnonlin_fwd': (_w_vnf, 'vecnonlin'), 'dc_vecnonlin_bwd': (_w_vnb, 'vecnonlin'),
    'dc_linear_fwd': (_w_lin, 'gemm'), 'dc_linear_bwd_input': (_w_linb, 'gemm'), 'dc_linear_bwd_weight': (_w_linb, 'gemm'),
    'dc_edge_diff_fwd': (_w_ediff, 'centralized'), 'dc_segmax_fwd': (_w_segmax, 'centralized'),
    'dc_segmax_bwd': (_w_segmax_t, 'centralized'),
    'dc_segmax_bn_fwd': (_w_segmax_bn, 'centralized'), 'dc_segmax_bn_bwd': (_w_segmax_bn_t, 'centralized'),
    'dc_central_edge_fwd': (_w_cedge, 'centralized'), 'dc_central_edge_bwd': (_w_cedge_t, 'centralized'),
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


def workload_loss(logits, target):
    """calc_loss as the workload's script calls it: label smoothing for the classifiers (train_modelnet.py:104), plain cross
    entropy for the part segmentation (train_shapenet.py:107)."""
    from deltaconv_b200.utils.losses import calc_loss
    return calc_loss(logits, target, smoothing=WL['smoothing'])


# ------------------------------------------------------------------------------------------------
# CPU baseline / reference arm.  Where a checkout of the reference is importable (DELTACONV_REFERENCE, default
# /root/reference — the build container), its own unmodified deltaconv/**/*.py is timed through the fast stand-ins of
# oracle/ref_standins_fast (kind "reference"); on the GPU box the checkout does not exist and the oracle port of the same
# algorithm is timed (kind "port").  Both run the SAME workload as the CUDA arm: same model, same batch size.
# ------------------------------------------------------------------------------------------------

def reference_checkout():
    path = os.environ.get('DELTACONV_REFERENCE', '/root/reference')
    return path if os.path.isdir(os.path.join(path, 'deltaconv', 'models')) else None


def _cpu_model_and_loss(wl):
    """(model, loss_fn, kind, how) on the CPU with the deterministic weights every arm uses."""
    from deltaconv_b200.utils import fill_deterministic
    from deltaconv_b200.utils.losses import calc_loss
    ref = reference_checkout()
    if ref is not None and not os.environ.get('DC_BENCH_FORCE_PORT'):
        here = os.path.join(ROOT, 'oracle')
        for extra in (ref, os.path.join(here, 'ref_standins'), os.path.join(here, 'ref_standins_fast')):
            if extra not in sys.path:
                sys.path.insert(0, extra)
        import deltaconv.models as ref_models                      # the reference itself
        model = fill_deterministic(make_model(ref_models, wl), 1)
        kind, how = 'reference', ("the unmodified reference's deltaconv/**/*.py (torch CPU) with cdist+topk kNN, CSR torch.sparse.mm "
                                  "and index_add/scatter_reduce in place of the absent torch_cluster/torch_sparse/torch_scatter")
    else:
        from oracle import network as on
        model = fill_deterministic(make_model(on, wl), 1)
        model.deltanet_base.fast_knn = True
        kind, how = 'port', 'the oracle port of the reference algorithm (torch CPU, ELL operators, cdist+topk kNN)'
    return model, (lambda out, y: calc_loss(out, y, smoothing=wl['smoothing'])), kind, how


def cpu_steps(wl, steps, warmup, budget_s=None):
    """Steps of the workload on the host cores; returns (seconds of every timed step, kind, how).  With `budget_s` the number
    of timed steps is trimmed (never the batch) so that the run ends in time; the warm-up steps size it."""
    model, loss_fn, kind, how = _cpu_model_and_loss(wl)
    data = make_batch(wl)
    if wl['training']:
        model.train()
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    else:
        model.eval()

    def one():
        t0 = time.perf_counter()
        if wl['training']:
            opt.zero_grad()
            loss = loss_fn(model(data), data.y)
            loss.backward()
            opt.step()
            loss.item()
        else:
            with torch.no_grad():
                model(data).sum().item()
        return time.perf_counter() - t0

    warm = [one() for _ in range(max(warmup, 1))]          # discarded (the first one pays allocator, thread pools, lazy imports)
    if budget_s is not None:
        steps = max(1, min(steps, int((budget_s - sum(warm)) / max(warm[-1], 1e-9))))
    return [one() for _ in range(steps)], kind, how


def run_reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    wl = WL
    torch.set_num_threads(os.cpu_count() or 1)          # torchrun pins OMP_NUM_THREADS=1; use all host cores
    cores = torch.get_num_threads()
    warm = max(args.warmup, 1)            # as many untimed steps as asked for; only the TIMED steps are trimmed to the budget
    times, kind, how = cpu_steps(wl, args.steps, warm, budget_s=args.cpu_budget)
    ms = 1000 * sum(times) / len(times)
    value = wl['clouds'] / (ms / 1000)
    what = 'training steps' if wl['training'] else 'evaluation forwards'
    sample = (f"{len(times)} {what} (after {warm} warm-up) of the full {wl['clouds']} clouds x {wl['points']} pts "
              f"(k={wl['k']}): {how}")
    line = dict(metric=wl['metric'], value=round(value, 3), unit='clouds/s', impl='reference', n_gpus=args.gpus, steps=len(times),
                warmup=warm, ms_per_step=round(ms, 2), higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='f32', data='synthetic',
                config=dict(workload=wl['workload'], clouds_per_gpu=wl['clouds'], points=wl['points'], k=wl['k'],
                            steps_requested=args.steps, steps_timed=len(times), cpu_budget_s=args.cpu_budget),
                cpu_baseline=dict(value=round(value, 3), unit='clouds/s', cores=cores, kind=kind, sample=sample),
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

    wl = WL
    CLOUDS, POINTS, training = wl['clouds'], wl['points'], wl['training']
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device — the CUDA arm has no CPU fallback (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    nccl_log = None
    if world > 1:
        # stdout carries exactly one JSON line (printed last).  The caller's NCCL_DEBUG setting is kept; unless the caller
        # chose a file, every rank logs to a file of its own and copies it to stderr when it is done (NCCL opens its log file
        # itself: N ranks opening /dev/stderr would each write from offset 0 of a redirected stderr, over one another).
        if os.environ.get('NCCL_DEBUG') and not os.environ.get('NCCL_DEBUG_FILE'):
            import tempfile
            nccl_log = os.path.join(tempfile.gettempdir(), f'dc_bench_nccl.{os.getpid()}.rank{rank}.log')
            os.environ['NCCL_DEBUG_FILE'] = nccl_log
        dist.init_process_group('nccl', device_id=dev)
    lib = _lib.load()
    if args.gemm_mode is not None:
        lib.dc_set_gemm_mode(int(args.gemm_mode))

    torch.manual_seed(1)
    model = D.utils.fill_deterministic(make_model(D.models, wl), 1).to(dev)
    model = model.train() if training else model.eval()
    flat = opt = None
    if training:
        flat = FlatGradients(model.parameters())
        opt = FlatSGD(flat, lr=0.1, momentum=0.9, weight_decay=1e-4)      # torch.optim.SGD's update as one kernel on flat buffers

    def hinted(data):
        data._num_clouds, data._max_points = CLOUDS, POINTS
        return data
    batch0 = make_batch(wl, rank)
    if args.morton:                                  # a pre-transform: applied once, outside every timed region
        batch0._num_clouds = CLOUDS
        batch0 = D.transforms.MortonOrder()(batch0)
        batch0 = D.utils.Data(**{k_: v_ for k_, v_ in batch0.__dict__.items() if k_ != 'morton_perm'})
    host = hinted(batch0.pin_memory())
    resident = hinted(host.to(dev))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)             # > 126 MB L2

    if training:
        def step(data):
            flat.zero()
            loss = workload_loss(model(data), data.y)
            loss.backward()
            flat.all_reduce(defer_average=True)          # SUM over NCCL; FlatSGD folds 1 / world into its update kernel
            opt.step()
            return loss
    else:
        def step(data):
            with torch.no_grad():
                return model(data)

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
        static = hinted(host.to(dev))
        graphed = (D.utils.GraphedTrainStep(model, opt, workload_loss, static, flat=flat) if training
                   else D.utils.GraphedInference(model, static))
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

    # (2) end to end: pinned host buffers -> device inside the timed region, result (loss / logits) read back every step
    result_numel = 1 if training else CLOUDS * wl['classes']

    def e2e_step():
        out = step(hinted(host.to(dev, non_blocking=True)))
        return out.item() if training else out.cpu()
    # Graph replay: the next batch's host->device copy runs on a side stream while the current step computes (prefetch /
    # commit), and each step's result is copied to pinned host memory asynchronously and read one step later — the usual
    # input pipeline of a training / serving loop.  Every step still moves its own inputs H2D and its own result D2H inside
    # the timed region.
    out_host = torch.empty(2, result_numel, dtype=torch.float32).pin_memory()
    out_ready = [torch.cuda.Event(), torch.cuda.Event()]
    pipe = dict(i=0, last=None)

    def e2e_graph_step():
        i = pipe['i']
        graphed.commit()
        out = graphed()
        out_host[i & 1].copy_(out.reshape(-1), non_blocking=True)
        out_ready[i & 1].record()
        graphed.prefetch(host)
        if i > 0:
            out_ready[(i - 1) & 1].synchronize()
            pipe['last'] = float(out_host[(i - 1) & 1][0])
        pipe['i'] = i + 1
        return pipe['last']
    e2e_fn = e2e_graph_step if graphed else e2e_step
    if graphed:
        graphed.prefetch(host)
    for _ in range(2):
        e2e_fn()
    e2e_ms = timed(e2e_fn, args.steps)
    e2e_eager_ms = timed(e2e_step, args.steps) if graphed else e2e_ms

    # (3) sustained: the same replay for >= 2 s with the clock / power record (the default timed region is a burst of ~0.1 s)
    sustained = None
    if graphed and args.sustain > 0:
        n_sus = max(args.steps, int(args.sustain * 1000 / max(total_ms / args.steps, 1e-3)) + 1)
        sus_sampler = ClockSampler(local) if rank == 0 else None
        w0 = time.time()
        sus_ms = timed(graphed, n_sus)
        w1 = time.time()
        sus_clocks = sus_sampler.summary(w0, w1) if sus_sampler else None
        sus_value = world * CLOUDS / (sus_ms / n_sus / 1000)
        sustained = dict(value=round(sus_value, 2), unit='clouds/s', steps=n_sus, seconds=round(sus_ms / 1000, 3),
                         ms_per_step=round(sus_ms / n_sus, 4), clocks=sus_clocks)

    # (4) per-entry-point device time (CUDA events on the launching stream), same steps
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

    def flush_nccl_log():
        if nccl_log and os.path.exists(nccl_log):
            with open(nccl_log, errors='replace') as f:
                sys.stderr.write(f.read())
            sys.stderr.flush()
            os.remove(nccl_log)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
            flush_nccl_log()
        return

    pk = peaks()
    ms_per_step = total_ms / args.steps
    value = world * CLOUDS / (ms_per_step / 1000)
    e2e_value = world * CLOUDS / (e2e_ms / args.steps / 1000)
    if sustained:
        sustained['vs_burst'] = round(sustained['value'] / value, 4)
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
        ncu = json.load(open(tpath)).get(top if args.workload == 'train' else f'{args.workload}:{top}') or {}
        traffic = ncu.get('dram_bytes_per_launch_avg')
    mode = int(lib.dc_get_gemm_mode())
    if tr['flops']:
        ach = tr['flops'] / (tr['ms'] / 1000) / 1e12
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
        t, kind, how = cpu_steps(wl, 2, 1)
        v = CLOUDS / (sum(t) / len(t))
        what = 'training steps' if training else 'evaluation forwards'
        cpu = dict(value=round(v, 3), unit='clouds/s', cores=cores, kind=kind,
                   sample=f"{len(t)} {what} (after 1 warm-up) of the full {CLOUDS} clouds x {POINTS} pts, k={wl['k']}: {how}")

    gemm_modes = {0: 'FFMA fp32 only', 1: 'tcgen05 kind::tf32, single TF32 product (~1e-3 relative per product), fp32 accumulate',
                  3: 'tcgen05 kind::tf32, 3xTF32 split (fp32-grade accuracy), fp32 accumulate; FFMA kernel for ineligible shapes'}
    par = (f'dp{world} (shard by cloud, flat-gradient all-reduce)' if training else
           f'{world} independent shards by cloud, no collective')
    line = dict(metric=wl['metric'], value=round(value, 2), unit='clouds/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=round(ms_per_step, 4), higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                data='synthetic',
                config=dict(workload=wl['workload'] + ('; points of every cloud pre-ordered along a Morton curve (transforms.MortonOrder)' if args.morton else ''),
                            point_order='morton' if args.morton else 'as generated (random)', clouds_per_gpu=CLOUDS, points=POINTS, k=wl['k'], parallelism=par,
                            l2='256 MiB buffer rewritten between steps (inside the timed region)',
                            gemm=gemm_modes.get(mode, str(mode)),
                            launch='CUDA-graph replay of the whole step' if graphed else 'eager (one Python launch per kernel)',
                            eager_ms_per_step=round(eager_ms / args.steps, 4), e2e_eager_ms_per_step=round(e2e_eager_ms / args.steps, 4)),
                e2e=dict(value=round(e2e_value, 2), unit='clouds/s', h2d_bytes_per_step=host.nbytes(), d2h_bytes_per_step=4 * result_numel,
                         ms_per_step=round(e2e_ms / args.steps, 4),
                         pipeline='next batch copied on a side stream during the step, result read one step later' if graphed else 'serial'),
                gpu_launches=launches, roofline=roof, cpu_baseline=cpu, clocks=clocks, sustained=sustained,
                kernels=kernels,
                hbm_bound_kernels=dict(gbps=round(hbm_bytes / (hbm_ms / 1000) / 1e9, 1) if hbm_ms else None,
                                       frac_of_peak=round(hbm_bytes / (hbm_ms / 1000) / 1e9 / pk['hbm'], 4) if hbm_ms else None))
    if world > 1:
        dist.destroy_process_group()
        flush_nccl_log()
    print(json.dumps(line), flush=True)


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
    ap.add_argument('--workload', default='train', choices=['train', 'seg', 'infer', 'infer_depth2', 'fps'],
                    help="'train' (default, the contract line): configs[1] classifier training; 'seg': configs[2] part-segmentation "
                         "training; 'infer': configs[4] ScanObjectNN-shape inference; 'fps': the geodesic farthest-point sampling "
                         "pre-transform (1 GPU)")
    ap.add_argument('--morton', action='store_true',
                    help='apply the MortonOrder pre-transform to the batch (untimed, like any dataset pre-transform): points of every '
                         'cloud along a Z-order curve; named in config.workload')
    ap.add_argument('--sustain', type=float, default=2.0, help='seconds of back-to-back replays for the `sustained` record (0 = skip)')
    ap.add_argument('--cpu-budget', type=float, default=240.0, help='--impl reference: seconds after which timed steps are trimmed')
    ap.add_argument('--gemm-mode', type=int, default=None, choices=[0, 1, 3],
                    help='opt-in GEMM arithmetic: 3 = 3xTF32 split (default, fp32-grade), 1 = single TF32 product, 0 = FFMA')
    args = ap.parse_args()
    if args.workload == 'fps':
        return run_fps_workload(args)
    global WL
    WL = WORKLOADS[args.workload]
    if args.impl == 'reference':
        run_reference_arm(args)
    else:
        run_cuda_arm(args)


if __name__ == '__main__':
    main()
