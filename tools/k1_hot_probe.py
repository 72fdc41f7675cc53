#!/usr/bin/env python
"""K1 with L2 residency hints for the most frequently gathered source rows
(pygb_seg_mark_hot / pygb_seg_gather_reduce_stream_hinted) on the products-shaped graph:
time per launch for several hint budgets, output bit-compared with the unhinted launch.

  python tools/k1_hot_probe.py [--mb 0,32,64,96] [--widths 256,100,48]
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
    ap.add_argument('--mb', default='0,32,64,96')
    ap.add_argument('--widths', default='256,100,48')
    ap.add_argument('--iters', type=int, default=10)
    args = ap.parse_args()
    from bench import WORKLOADS
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    from torch_geometric.graph import Adj
    from torch_geometric.nn import GCNConv
    dev = torch.device('cuda:0')
    n, pairs = WORKLOADS['products'][:2]
    data = synthetic.products_like(seed=4, device=dev, num_nodes=n, num_pairs=pairs, num_features=4)
    ei, norm = GCNConv.norm(data.edge_index, n, None, False, torch.float32)
    adj = Adj(ei, 1, (n, n))
    w = adj.permuted(norm, 'dst')
    for F in [int(f) for f in args.widths.split(',')]:
        ld = (F + 31) // 32 * 32                      # 128-byte row stride, as the layers see it
        x = torch.randn(n, ld, device=dev)[:, :F]
        ref = None
        for mb in [int(m) for m in args.mb.split(',')]:
            ops.K1_HOT_MB = mb
            for _ in range(3):
                out = ops.seg_gather_reduce_raw(x, adj.dst, adj.dst.col, 'sum', w, None, None)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                out = ops.seg_gather_reduce_raw(x, adj.dst, adj.dst.col, 'sum', w, None, None)
            e1.record()
            torch.cuda.synchronize()
            if ref is None:
                ref = out
            hot = adj.dst._items.get(('hot', (mb << 20) // (ld * 4))) if mb else None
            print(json.dumps({'F': F, 'hot_mb': mb, 'ms': round(e0.elapsed_time(e1) / args.iters, 4),
                              'same_bits': bool(torch.equal(out, ref)),
                              'hot_slots_frac': None if hot is None else round(
                                  float(((hot >> 30) & 1).float().mean().item()), 4)}), flush=True)
        del x, ref


if __name__ == '__main__':
    main()
