#!/usr/bin/env python
"""Timings of the parity configs of BASELINE.json on one B200 (the driver's bench line
is cfg5, bench.py): cfg2 PruningGCN hidden 512 on a PubMed-shaped graph (forward and
forward+backward), cfg3 GATConv 8x8 on a Reddit-shaped graph, cfg4 EdgeConv max on
kNN point clouds.  CUDA events; per-kernel breakdown from the native event log.

  python tools/bench_configs.py [--cfg 2,3,4] [--iters 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def breakdown(fn):
    from torch_geometric import _native as nat
    nat.event_log = []
    fn()
    torch.cuda.synchronize()
    log, nat.event_log = nat.event_log, None
    agg = {}
    for name, a, e0, e1 in log:
        if name == 'pygb_gat_backward_stream':
            name += ':mode%d' % a[0]          # 0 = destination-sorted walk, 1 = source-sorted
        agg[name] = agg.get(name, 0.0) + e0.elapsed_time(e1)
    return {k: round(v, 4) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}


def cfg2(dev, iters):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    data = synthetic.pubmed_like(seed=1).to(dev)
    torch.manual_seed(0)
    c1, c2 = GCNConv(500, 512, cached=True).to(dev), GCNConv(512, 3, cached=True).to(dev)

    def net(x):
        h = c1(x, data.edge_index)
        h = h * torch.sigmoid(h)          # examples/ConvexPruning.py:121 (x*sigmoid(x))
        return F.log_softmax(c2(h, data.edge_index), dim=1)

    def fwd():
        with torch.no_grad():
            return net(data.x)

    params = list(c1.parameters()) + list(c2.parameters())

    def fwdbwd():
        for p in params:
            p.grad = None
        loss = F.nll_loss(net(data.x)[data.train_mask], data.y[data.train_mask])
        loss.backward()

    fwd()
    e_prime = c1.cached_result[0].size(1)
    t_f, t_fb = timeit(fwd, iters), timeit(fwdbwd, iters)
    from torch_geometric.cuda_graph import CapturedForward
    cap = CapturedForward(net, data.x)
    assert torch.equal(cap(data.x), fwd())
    t_g = timeit(lambda: cap.graph.replay(), iters)
    return {'cfg': 2, 'workload': 'PruningGCN hidden 512, PubMed-shaped (19,717 nodes, 88,648 edges, '
            '%d after self-loops, 500 feats)' % e_prime,
            'fwd_ms': round(t_f, 4), 'fwd_cuda_graph_ms': round(t_g, 4), 'fwd_bwd_ms': round(t_fb, 4),
            'fwd_edges_per_s': 2 * e_prime / t_f * 1e3, 'fwd_bwd_edges_per_s': 2 * e_prime / t_fb * 1e3,
            'fwd_breakdown_ms': breakdown(fwd), 'fwd_bwd_breakdown_ms': breakdown(fwdbwd)}


def cfg3(dev, iters):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GATConv
    data = synthetic.reddit_like(seed=2, device=dev)
    torch.manual_seed(0)
    gat = GATConv(602, 8, heads=8).to(dev)

    def fwd():
        with torch.no_grad():
            return gat(data.x, data.edge_index)

    fwd()
    ei = gat.__dict__['_loop_cache'][2]
    e_prime = ei.size(1)
    n, H, C = data.x.size(0), 8, 8
    t = timeit(fwd, iters)
    bd = breakdown(fwd)
    xg = data.x.clone().requires_grad_(True)

    def fwdbwd():
        xg.grad = None
        for p_ in gat.parameters():
            p_.grad = None
        gat(xg, data.edge_index).square().sum().backward()

    t_fb = timeit(fwdbwd, max(iters // 2, 2))
    bd_fb = breakdown(fwdbwd)
    k2 = bd.get('pygb_gat_forward_stream', bd.get('pygb_gat_forward', None))
    k2_bytes = e_prime * (4 + 4 * H + 4 * H * C) + n * H * C * 4 + 2 * n * H * 4 + (n + 1) * 4
    return {'cfg': 3, 'workload': 'GATConv 8x8, Reddit-shaped (%d nodes, %d edges, %d after '
            'remove+add self-loops, 602 feats)' % (n, data.edge_index.size(1), e_prime),
            'fwd_ms': round(t, 4), 'fwd_edges_per_s': e_prime / t * 1e3, 'breakdown_ms': bd,
            'fwd_bwd_ms': round(t_fb, 4), 'fwd_bwd_breakdown_ms': bd_fb,
            'k2_algorithmic_bytes': k2_bytes,
            'k2_GBps': (k2_bytes / k2 / 1e6) if k2 else None}


def cfg4(dev, iters):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import EdgeConv
    data = synthetic.knn_clouds(device=dev)
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(6, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64))
    conv = EdgeConv(mlp, aggr='max').to(dev)

    def fwd():
        with torch.no_grad():
            return conv(data.x, data.edge_index)

    fwd()
    E = data.edge_index.size(1)
    t = timeit(fwd, iters)
    bd = breakdown(fwd)
    from torch_geometric.cuda_graph import CapturedForward
    cap = CapturedForward(lambda x: conv(x, data.edge_index), data.x)
    assert torch.equal(cap(data.x), fwd())
    t_g = timeit(lambda: cap.graph.replay(), iters)
    k3 = bd.get('pygb_seg_max')
    k3_bytes = E * 64 * 4 + E * 4 + data.x.size(0) * 64 * (4 + 8)
    return {'cfg': 4, 'workload': 'EdgeConv max, 32 x 1,024-point clouds, k=20 (%d edges, 3->64)' % E,
            'fwd_ms': round(t, 4), 'fwd_cuda_graph_ms': round(t_g, 4), 'fwd_edges_per_s': E / t * 1e3,
            'breakdown_ms': bd, 'k3_algorithmic_bytes': k3_bytes, 'k3_GBps': (k3_bytes / k3 / 1e6) if k3 else None}


def cfg5(dev, iters):
    """The bench.py workload (3-layer GCN on the products-shaped graph) as a training step."""
    from bench import WORKLOADS
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    n, pairs, fin, hid, ncls = WORKLOADS['products']
    data = synthetic.products_like(seed=4, device=dev, num_nodes=n, num_pairs=pairs, num_features=fin)
    torch.manual_seed(0)
    convs = [GCNConv(fin, hid, cached=True).to(dev), GCNConv(hid, hid, cached=True).to(dev),
             GCNConv(hid, ncls, cached=True).to(dev)]
    params = [p_ for c in convs for p_ in c.parameters()]

    def net(x):
        h = x
        for k, c in enumerate(convs):
            h = c(h, data.edge_index)
            if k < 2:
                h = F.relu(h)
        return F.log_softmax(h, dim=1)

    def fwd():
        with torch.no_grad():
            return net(data.x)

    def fwdbwd():
        for p_ in params:
            p_.grad = None
        F.nll_loss(net(data.x), data.y).backward()

    fwd()
    e_prime = convs[0].cached_result[0].size(1)
    for c in convs[1:]:
        c.cached_result = convs[0].cached_result
    iters = max(iters // 2, 3)
    t_f, t_fb = timeit(fwd, iters), timeit(fwdbwd, iters)
    return {'cfg': 5, 'workload': '3-layer GCN 100->256->256->47, products-shaped (%d nodes, %d edges '
            'after self-loops)' % (n, e_prime), 'fwd_ms': round(t_f, 3), 'fwd_bwd_ms': round(t_fb, 3),
            'fwd_edges_per_s': 3 * e_prime / t_f * 1e3, 'fwd_bwd_edges_per_s': 3 * e_prime / t_fb * 1e3,
            'fwd_bwd_breakdown_ms': breakdown(fwdbwd)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--cfg', default='2,3,4')
    ap.add_argument('--iters', type=int, default=10)
    args = ap.parse_args()
    dev = torch.device('cuda:0')
    fns = {'2': cfg2, '3': cfg3, '4': cfg4, '5': cfg5}
    for c in args.cfg.split(','):
        print(json.dumps(fns[c](dev, args.iters)), flush=True)
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
