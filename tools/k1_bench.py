#!/usr/bin/env python
"""Micro-benchmark of the aggregation kernels on the products-shaped graph
(BASELINE configs[4]): K1 v1 (one warp per row) vs K1 v2 (streaming) at several
feature widths, CUDA-event timing, algorithmic GB/s and the fraction
of the measured HBM peak.  Also cross-checks v2 against v1.

  python tools/k1_bench.py [--workload products] [--widths 256,100,48] [--dtype f32]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='products')
    ap.add_argument('--widths', default='256,100,48,64,16')
    ap.add_argument('--dtype', default='f32')
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--T', default='')
    ap.add_argument('--skip-v1', action='store_true')
    ap.add_argument('--ld', type=int, default=0, help='row stride of x in elements (0: F padded to 16 bytes)')
    args = ap.parse_args()
    from bench import WORKLOADS, k1_bytes, peaks
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    from torch_geometric.graph import Adj
    from torch_geometric.nn import GCNConv

    dev = torch.device('cuda:0')
    n, pairs, fin, hid, ncls = WORKLOADS[args.workload]
    data = synthetic.products_like(seed=4, device=dev, num_nodes=n, num_pairs=pairs,
                                   num_features=4)
    ei, norm = GCNConv.norm(data.edge_index, n, None, False, torch.float32)
    adj = Adj(ei, 1, (n, n))
    E = ei.size(1)
    w = adj.permuted(norm, 'dst')
    dtype = torch.float32 if args.dtype == 'f32' else torch.bfloat16
    peak, _ = peaks()
    res = []
    Ts = [None] + [int(t) for t in args.T.split(',') if t]
    for F in [int(f) for f in args.widths.split(',')]:
        vec = 16 // (4 if dtype == torch.float32 else 2)
        ld = max(args.ld, (F + vec - 1) // vec * vec)
        x = torch.randn(n, ld, device=dev, dtype=torch.float32).to(dtype)[:, :F]
        nbytes = k1_bytes(E, n, F, x.element_size())
        variants = [] if args.skip_v1 else [('v1', False, None)]
        variants += [('v2' + ('' if t is None else '_T%d' % t), True, t) for t in Ts]
        outs = {}
        for name, stream, T in variants:
            ops.STREAM, ops.STREAM_T = stream, T
            for _ in range(3):
                out = ops.seg_gather_reduce_raw(x, adj.dst, adj.dst.col, 'sum', w, None, None)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                out = ops.seg_gather_reduce_raw(x, adj.dst, adj.dst.col, 'sum', w, None, None)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            outs[name] = out
            gbs = nbytes / ms / 1e6
            res.append({'F': F, 'dtype': args.dtype, 'kernel': name, 'ms': round(ms, 4),
                        'GBps': round(gbs, 1), 'frac_measured': round(gbs / peak, 4),
                        'Gedges_s': round(E / ms / 1e6, 3)})
            print(json.dumps(res[-1]), flush=True)
        if 'v1' in outs:
            for k, o in outs.items():
                if k == 'v1':
                    continue
                a, b = o.float(), outs['v1'].float()
                err = (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)
                print('  check %s vs v1: rel err %.3e' % (k, err), flush=True)
                assert err < (1e-5 if dtype == torch.float32 else 2e-2)
        del x, outs
    return res


if __name__ == '__main__':
    main()
