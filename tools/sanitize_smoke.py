#!/usr/bin/env python
"""Small exerciser for compute-sanitizer: the streaming kernels whose
cross-lane shared-memory traffic is ordered only by __syncwarp + cp.async.wait_group
(seg_stream_kernel, gat_stream_kernel, gat_bwd_stream_kernel), the radix scatter of the CSR
build, the SDDMM edge-dot kernel, the k-NN / fps / radius kernels and the tcgen05 GEMMs, on
graphs with hub rows, empty rows and cut rows.

    compute-sanitizer --tool memcheck  python tools/sanitize_smoke.py
    compute-sanitizer --tool racecheck python tools/sanitize_smoke.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import torch  # noqa: E402


def main():
    from util import skewed_graph
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    from torch_geometric.nn import GCNConv, GATConv, knn_graph, fps, radius_graph
    dev = torch.device('cuda:0')
    torch.manual_seed(0)
    n, E = 600, 9000
    ei = skewed_graph(n, E, 3).to(dev)
    for F in (4, 47, 100, 256):
        x = torch.randn(n, F, device=dev)
        adj = Adj(ei, 1, (n, n))
        w = torch.rand(E, device=dev)
        for T in (32, None):
            ops.STREAM_T = T
            ops.gather_reduce(x, adj, w, 'sum', torch.randn(F, device=dev))
            ops.gather_reduce(x, adj, None, 'mean')
        ops.STREAM_T = None
        ops.edge_dot_raw(torch.randn(n, F, device=dev), x, adj.dst)
    for H, C in ((8, 8), (1, 16), (2, 64)):
        conv = GATConv(40, C, heads=H).to(dev)
        xin = torch.randn(n, 40, device=dev, requires_grad=True)
        conv(xin, ei).square().sum().backward()
    gcn = GCNConv(100, 48).to(dev)
    xin = torch.randn(n, 100, device=dev, requires_grad=True)
    gcn(xin, ei).sum().backward()
    big = torch.randn(3000, 256, device=dev, requires_grad=True)
    wt = torch.randn(256, 64, device=dev, requires_grad=True)
    ops.linear(big, wt).sum().backward()
    ops.linear(big.detach().bfloat16(), wt.detach().bfloat16())
    pts = torch.rand(512, 3, device=dev)
    batch = torch.arange(4, device=dev).repeat_interleave(128)
    knn_graph(pts, 20, batch)
    fps(pts, batch, 0.5, random_start=False)
    radius_graph(pts, 0.3, batch)
    from torch_geometric.utils import scatter_, softmax
    scatter_('max', torch.randn(E, 33, device=dev), ei[1], 0, n)
    softmax(torch.randn(E, 4, device=dev), ei[1], n)
    torch.cuda.synchronize()
    print('sanitize_smoke ok')


if __name__ == '__main__':
    main()
