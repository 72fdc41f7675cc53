#!/usr/bin/env python
"""The launches whose DRAM traffic / pipe utilisation the bench lines quote, in a fixed order,
for ONE ncu capture (B200_PROFILING.md recipe):

    python tools/ncu_targets.py > gpurun_out/ncu_targets_plain.log 2>&1 && \
    ncu --set full --clock-control none --import-source on \
        -k regex:'seg_stream_kernel|gat_stream_kernel|gat_bwd_stream_kernel|seg_max_kernel|gemm_tf32_kernel|edge_dot_kernel' \
        -o gpurun_out/r02_targets python tools/ncu_targets.py

Order of the profiled launches (also written to gpurun_out/ncu_targets_order.json):
  K1 single GPU F=256, F=100 (line-aligned copy), F=100 (400-byte rows), F=47->48;
  K1 routed, half graph of the 2 x 4 plan (what one rank of the 8-GPU run executes), F=32, 64, 16;
  K5 256x256 (3xTF32), 100x256, 256x47, bf16 256x256;
  cfg2 K1 F=500; cfg4 K3 seg_max; cfg3 K2 forward, K2b mode 0 and mode 1; SDDMM F=256.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch  # noqa: E402


def main():
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    from torch_geometric.dist.sliced import SlicePlan
    from torch_geometric.graph import Adj, clear_caches
    from torch_geometric.nn import GCNConv, GATConv
    dev = torch.device('cuda:0')
    order = []
    n, pairs = 2449029, 31000000
    data = synthetic.products_like(seed=4, device=dev, num_nodes=n, num_pairs=pairs, num_features=4)
    ei2, norm = GCNConv.norm(data.edge_index, n, None, False, torch.float32)
    adj = Adj(ei2, 1, (n, n))
    g = torch.Generator(device=dev).manual_seed(1)
    # ---- K1 single GPU -------------------------------------------------------------------
    x256 = torch.randn(n, 256, device=dev, generator=g)
    ops.gather_reduce(x256, adj, norm, 'sum')
    order.append('K1 F=256 fp32, cfg5 whole graph')
    x100p = torch.zeros(n, 128, device=dev)[:, :100]
    x100p.copy_(x256[:, :100])
    ops.LINE_ALIGN = False
    ops.gather_reduce(x100p, adj, norm, 'sum')
    order.append('K1 F=100 fp32, rows at a 512-byte stride')
    x100 = x256[:, :100].contiguous()
    ops.gather_reduce(x100, adj, norm, 'sum')
    order.append('K1 F=100 fp32, 400-byte rows')
    x48 = x256[:, :48].contiguous()
    ops.gather_reduce(x48[:, :47], adj, norm, 'sum')
    order.append('K1 F=47 (padded 48) fp32')
    # ---- K5 ---------------------------------------------------------------------------------
    for K, N in ((256, 256), (100, 256), (256, 47)):
        w = torch.randn(K, N, device=dev, generator=g) / K ** 0.5
        ops.gemm_xw_raw(x256[:, :K] if K == 256 else x100, w, torch.zeros(N, device=dev))
        order.append('K5 3xTF32 M=2449029 K=%d N=%d' % (K, N))
    xb = x256.bfloat16()
    ops.gemm_xw_raw(xb, (torch.randn(256, 256, device=dev, generator=g) / 16).bfloat16())
    order.append('K5 bf16 M=2449029 K=256 N=256')
    # ---- SDDMM -----------------------------------------------------------------------------
    ops.edge_dot_raw(x256, x256, adj.dst)
    order.append('SDDMM edge_dot F=256 fp32, cfg5 whole graph')
    del xb, x100, x100p, x48
    # ---- K1 routed: one rank's share of the 2 x 4 plan -----------------------------------
    plan = SlicePlan(n, 8, 0, 2, blk_align=8)
    t = ei2[1]
    mask = (t >= plan.g_lo) & (t < plan.g_hi)
    eig = ei2[:, mask].clone()
    eig[1] -= plan.g_lo
    adj_g = Adj(eig, 1, (n, plan.n_group))
    w_g = adj_g.permuted(norm[mask].contiguous(), 'dst')
    del adj, mask, t
    clear_caches()
    for Fc in (32, 64, 16):
        xs = x256[:, :Fc].contiguous()
        bufs = [torch.empty(plan.blk, Fc, device=dev) for _ in range(plan.Gc)]
        tab = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64).to(dev)
        ops.seg_gather_reduce_routed_raw(xs, adj_g.dst, w_g, None, tab, plan.blk, Fc)
        order.append('K1 routed F=%d fp32, row group 0 of the 2x4 plan (%d edges)' % (Fc, eig.size(1)))
    del adj_g, w_g, eig, x256, ei2, norm, data
    clear_caches()
    torch.cuda.empty_cache()
    # ---- cfg2 K1 F=500 -----------------------------------------------------------------------
    d2 = synthetic.pubmed_like(seed=1).to(dev)
    c1 = GCNConv(500, 512, cached=True).to(dev)
    with torch.no_grad():
        c1(d2.x, d2.edge_index)
    order.append('cfg2: K1 F=500 fp32 (then K5 500x512, not listed)')
    order.append('cfg2: K5 3xTF32 M=19717 K=500 N=512')
    # ---- cfg4 K3 ---------------------------------------------------------------------------
    d4 = synthetic.knn_clouds(32, 1024, 20, seed=3, device=dev)
    msg = torch.randn(d4.edge_index.size(1), 64, device=dev, generator=g)
    ops.scatter_max(msg, d4.edge_index[1], 32768)
    order.append('cfg4: K3 seg_max F=64, 655360 messages')
    # ---- cfg3 K2 / K2b ---------------------------------------------------------------------
    d3 = synthetic.reddit_like(seed=2, device=dev, with_features=False)
    n3 = 232965
    gat = GATConv(64, 8, heads=8).to(dev)
    xw = torch.randn(n3, 64, device=dev, generator=g, requires_grad=True)
    ei3 = gat._with_self_loops(d3.edge_index, n3)
    adj3 = Adj(ei3, 1, (n3, n3))
    out = ops.gat_propagate(xw, xw, gat.att, gat.bias, adj3, 8, 8, 0.2)
    order.append('cfg3: K2 gat_stream forward H=8 C=8 (%d edges)' % ei3.size(1))
    out.square().sum().backward()
    order.append('cfg3: K2b gat_bwd_stream mode 0 (destination-sorted walk)')
    order.append('cfg3: K2b gat_bwd_stream mode 1 (source-sorted walk)')
    torch.cuda.synchronize()
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    json.dump(order, open(os.path.join(ROOT, 'gpurun_out', 'ncu_targets_order.json'), 'w'), indent=1)
    print('ncu_targets ok:', len(order), 'labelled launches')


if __name__ == '__main__':
    main()
