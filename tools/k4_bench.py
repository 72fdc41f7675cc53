#!/usr/bin/env python
"""First-call cost of a graph (K4, SURVEY 8d "separately timed as first-call"): every native
call GCNConv / GATConv issue the first time they see an edge_index — self-loop handling,
degrees and D^-1/2 A D^-1/2 weights, the stable radix sort into the destination-sorted CSR,
the weight permutation and the work tables of the streaming kernels — with CUDA events, against
the algorithmic bytes of SURVEY 8d (K4 row).

  python tools/k4_bench.py [--cfg 5,3] [--iters 5]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch  # noqa: E402


def first_call(fn, iters):
    """Median over `iters` cold builds (caches cleared) of total ms and per-call ms."""
    from torch_geometric import _native as nat
    from torch_geometric import graph
    totals, per = [], {}
    for it in range(iters + 1):
        graph.clear_caches()
        torch.cuda.synchronize()
        nat.event_log = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        log, nat.event_log = nat.event_log, None
        if it == 0:
            continue                      # warm-up: allocator, module loading
        totals.append(e0.elapsed_time(e1))
        agg = {}
        for name, a, s, e in log:
            agg[name] = agg.get(name, 0.0) + s.elapsed_time(e)
        for k, v in agg.items():
            per.setdefault(k, []).append(v)
    med = lambda v: sorted(v)[len(v) // 2]
    return med(totals), {k: round(med(v), 4) for k, v in sorted(per.items(), key=lambda kv: -med(kv[1]))}


def k4_bytes(E, Ep, N):
    # read edge_index, write edge_index', col + perm, rowptr, deg, norm
    return 2 * E * 8 + 2 * Ep * 8 + Ep * (4 + 4) + (N + 1) * 4 + N * 4 + Ep * 4


def cfg5(dev, iters):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    data = synthetic.products_like(seed=4, device=dev)
    ei, N = data.edge_index, data.x.size(0)
    x = torch.randn(N, 48, device=dev)

    def build():
        conv = GCNConv(48, 48, cached=True).to(dev)
        with torch.no_grad():
            conv(x, ei)                      # norm + CSR + tables + one aggregation
        return conv

    conv = build()
    Ep = conv.cached_result[0].size(1)
    t, per = first_call(build, iters)
    agg = per.pop('pygb_seg_gather_reduce_stream', 0.0) + per.pop('pygb_gemm_xw', 0.0)
    b = k4_bytes(ei.size(1), Ep, N)
    k4 = sum(per.values())
    return {'cfg': 5, 'what': 'GCNConv first call on the products-shaped graph (N=%d, E=%d, E\'=%d)' % (
        N, ei.size(1), Ep), 'first_call_ms': round(t, 3), 'k4_native_ms': round(k4, 3),
        'aggregation_and_gemm_ms': round(agg, 3), 'per_call_ms': per,
        'k4_algorithmic_bytes': b, 'k4_GBps': b / k4 / 1e6}


def cfg2(dev, iters):
    """PubMed-shaped graph: the first call is launch-latency bound (what a DataLoader that
    produces a new edge_index per mini-batch pays every step)."""
    import time
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    data = synthetic.pubmed_like(seed=1).to(dev)
    ei, N = data.edge_index, data.x.size(0)
    x = torch.randn(N, 16, device=dev)

    def build():
        conv = GCNConv(16, 16, cached=True).to(dev)
        with torch.no_grad():
            conv(x, ei)
        return conv

    conv = build()
    Ep = conv.cached_result[0].size(1)
    t, per = first_call(build, iters)
    from torch_geometric import graph
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    for _ in range(20):
        graph.clear_caches()
        build()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - w0) / 20 * 1e3
    return {'cfg': 2, 'what': 'GCNConv first call on the PubMed-shaped graph (N=%d, E=%d, E\'=%d)' % (
        N, ei.size(1), Ep), 'first_call_ms': round(t, 3), 'first_call_wall_ms': round(wall, 3),
        'native_calls': len(per), 'per_call_ms': per}


def cfg3(dev, iters):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GATConv
    data = synthetic.reddit_like(seed=2, device=dev)
    ei, N = data.edge_index, data.x.size(0)
    x = torch.randn(N, 64, device=dev)

    def build():
        gat = GATConv(64, 8, heads=8).to(dev)
        with torch.no_grad():
            gat(x, ei)
        return gat

    gat = build()
    Ep = gat.__dict__['_loop_cache'][2].size(1)
    t, per = first_call(build, iters)
    agg = per.pop('pygb_gat_forward_stream', 0.0) + per.pop('pygb_gemm_xw', 0.0) + \
        per.pop('pygb_gat_scores', 0.0)
    k4 = sum(per.values())
    b = 2 * ei.size(1) * 8 + 2 * Ep * 8 + Ep * 8 + (N + 1) * 4
    return {'cfg': 3, 'what': 'GATConv first call on the Reddit-shaped graph (N=%d, E=%d, E\'=%d)' % (
        N, ei.size(1), Ep), 'first_call_ms': round(t, 3), 'k4_native_ms': round(k4, 3),
        'aggregation_and_gemm_ms': round(agg, 3), 'per_call_ms': per,
        'k4_algorithmic_bytes': b, 'k4_GBps': b / k4 / 1e6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--cfg', default='5,3,2')
    ap.add_argument('--iters', type=int, default=5)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    for c in args.cfg.split(','):
        print(json.dumps({'5': cfg5, '3': cfg3, '2': cfg2}[c](dev, args.iters)), flush=True)


if __name__ == '__main__':
    main()
