#!/usr/bin/env python
"""Operator microbenchmark (BASELINE.json configs[3]): kNN, build_grad_div and the grad/div/curl SpMMs on
2^20 points per launch, achieved HBM GB/s (algorithmic bytes of SURVEY.md §8d / device time) against the
measured HBM peak.  Writes a markdown table and prints one JSON line per row.

    python bench_ops.py [--points-log2 10 12 14] [--k 10 20 30 40] [--channels 16 64 128] [--out profiles/xyz.md]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch                                                    # noqa: E402

TOTAL = 1 << 20          # points per launch (B * P), SURVEY.md §8d


def time_ms(fn, flush, reps=5):
    """median device time of fn(), L2 flushed before every repetition"""
    out = []
    for _ in range(reps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        out.append(s.elapsed_time(e))
    return sorted(out)[len(out) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--points-log2', type=int, nargs='+', default=[10, 12])
    ap.add_argument('--k', type=int, nargs='+', default=[10, 20, 30, 40])
    ap.add_argument('--channels', type=int, nargs='+', default=[16, 64, 128])
    ap.add_argument('--out', default=None)
    ap.add_argument('--morton', action='store_true', help='points of every cloud in Morton order (transforms.MortonOrder) before the graph is built')
    args = ap.parse_args()
    import deltaconv_b200 as D
    from deltaconv_b200 import ops
    peak_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    peak = json.load(open(peak_path))['hbm_gbs'] if os.path.exists(peak_path) else 6650.0
    dev = torch.device('cuda', 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for lg in args.points_log2:
        P = 1 << lg
        B = TOTAL // P
        data = D.utils.synthetic_batch(B, P, config=4).to(dev)
        if args.morton:
            data._num_clouds = B
            data = D.transforms.MortonOrder()(data)
        N = B * P
        for k in args.k:
            graph = ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False)      # writes idx and d2
            t_knn = time_ms(lambda: ops.knn(data.pos, k, data.batch, num_clouds=B, want_edge_index=False), flush, 3)
            xb, yb = ops.tangent_basis(data.norm)
            cp = graph.cloud_ptr
            # the build as the models call it (geometry.build_grad_div): mean edge lengths from the kNN distances that ride on
            # the graph, per-cloud normalisation deferred to the operator applications
            t_build = time_ms(lambda: ops.build_operators(data.pos, data.norm, xb, yb, graph, cp, B, deferred=True), flush)
            G, Dv, scale = ops.build_operators(data.pos, data.norm, xb, yb, graph, cp, B, deferred=True)
            grad, div = ops.EllOperator('grad', graph, G, Dv, scale), ops.EllOperator('div', graph, G, Dv, scale)
            graph.transpose()
            rec = dict(P=P, B=B, k=k,
                       knn_ms=round(t_knn, 4), knn_gbps=round(N * (12 + 4 * k) / t_knn / 1e6, 1),
                       knn_gpairs_per_s=round(N * P / t_knn / 1e6, 1),
                       build_ms=round(t_build, 4), build_gbps=round(N * (48 + 20 * k) / t_build / 1e6, 1))
            rec['build_frac'] = round(rec['build_gbps'] / peak, 4)
            for C in args.channels:
                x = torch.randn(N, C, device=dev)
                v = torch.randn(2 * N, C, device=dev)
                with torch.no_grad():
                    t_g = time_ms(lambda: grad @ x, flush)
                    t_d = time_ms(lambda: ops.scalar_cat(x, v, div), flush)
                    xr = x.clone().requires_grad_(True)
                yv = grad @ xr
                gy = torch.randn_like(yv)
                t_gt = time_ms(lambda: torch.autograd.grad(yv, xr, gy, retain_graph=True), flush)
                rec[f'grad_C{C}_gbps'] = round(N * (12 * k + 12 * C) / t_g / 1e6, 1)
                rec[f'gradT_C{C}_gbps'] = round(N * (12 * k + 12 * C) / t_gt / 1e6, 1)
                # scalar_cat = fused div/curl/norm (12k + 20C) with x copied through into the concatenation (8C)
                rec[f'divcurlnorm_C{C}_gbps'] = round(N * (12 * k + 28 * C) / t_d / 1e6, 1)
                del x, v, xr, yv, gy
            rows.append(rec)
            print(json.dumps(rec), flush=True)
            del graph, G, Dv, grad, div
        del data
    if args.out:
        keys = list(rows[0].keys())
        with open(args.out, 'w') as f:
            f.write(f'Operator microbenchmark, {TOTAL} points per launch, L2 flushed, median of 5; HBM peak {peak} GB/s (measured); '
                    f'point order: {"Morton curve inside every cloud (transforms.MortonOrder)" if args.morton else "as generated (random)"}.\n\n')
            f.write('| ' + ' | '.join(keys) + ' |\n|' + '---|' * len(keys) + '\n')
            for r in rows:
                f.write('| ' + ' | '.join(str(r[k_]) for k_ in keys) + ' |\n')


if __name__ == '__main__':
    main()
