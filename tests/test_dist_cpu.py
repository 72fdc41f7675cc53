"""CPU, world_size 2, gloo: the host logic of the destination-partitioned GCN
(partition plan, renumbering into the padded gather layout, in-place halo
all-gather, exchange at min(in, out) width) against the oracle on the whole graph.
The aggregation itself is injected (the oracle's scatter) — the CUDA kernel has no
CPU path; it is covered on the GPU by tests/test_dist_gpu.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_cpu as O
from util import random_graph


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_aggregate(graph, full, bias):
    ei = graph.edge_index_local
    msg = full[ei[1 - graph.i]]
    if graph.norm is not None:
        msg = graph.norm.view(-1, 1) * msg
    out = O.scatter_('add', msg, ei[graph.i], 0, graph.plan.n_local)
    return out + bias if bias is not None else out


def _cpu_transform(x, weight, bias, out=None):
    y = x @ weight
    if bias is not None:
        y = y + bias
    if out is not None:
        out[:, :y.size(1)].copy_(y)
        return out[:, :y.size(1)]
    return y


def _worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torch_geometric.dist import DistGraph, dist_gcn_forward
        torch.manual_seed(0)
        n = 301
        ei = random_graph(n, n, 4000, 5)
        # skew: a hub target so the edge-balanced bounds differ from node-balanced ones
        ei[1, :600] = 7
        x = torch.randn(n, 8)
        dims = [(8, 16), (16, 16), (16, 5)]          # aggregate-first, transform-first, narrow
        layers = [(torch.randn(a, b) / a ** 0.5, torch.randn(b)) for a, b in dims]
        ei2, norm = O.gcn_norm(ei, n, dtype=torch.float32)
        g = DistGraph(ei2, norm, n, rank, world, build_adj=False)
        assert g.plan.bounds[0] == 0 and g.plan.bounds[-1] == n
        out_local = dist_gcn_forward(g, g.local_rows(x).contiguous(), layers,
                                     aggregate=_oracle_aggregate, transform=_cpu_transform)
        parts = [torch.zeros(g.plan.block, 5) for _ in range(world)]
        pad = torch.zeros(g.plan.block, 5)
        pad[:g.plan.n_local] = out_local
        dist.all_gather(parts, pad)
        if rank == 0:
            got = torch.cat([parts[p][:g.plan.bounds[p + 1] - g.plan.bounds[p]]
                             for p in range(world)])
            h = x
            for k, (w, b) in enumerate(layers):
                h = O.scatter_('add', norm.view(-1, 1) * (h @ w)[ei2[0]], ei2[1], 0, n) + b
                if k + 1 < len(layers):
                    h = torch.relu(h)
            err = (got - h).abs().max().item() / h.abs().max().item()
            edges = [int(((ei2[1] >= g.plan.bounds[p]) & (ei2[1] < g.plan.bounds[p + 1])).sum())
                     for p in range(world)]
            q.put(('ok', err, edges, g.halo.bytes_received_total))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e), None, None))
        raise
    finally:
        dist.destroy_process_group()


def test_dist_gcn_two_ranks_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    status, err, edges, nbytes = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 1e-5, err
    # edge-balanced partition: the two ranks hold about the same number of in-edges
    assert abs(edges[0] - edges[1]) <= 0.2 * sum(edges), edges
    assert nbytes > 0


def test_partition_plan_properties():
    from torch_geometric.dist import PartitionPlan, balanced_bounds
    deg = torch.tensor([0, 5, 1, 1, 1, 100, 1, 1, 0, 0, 3])
    for world in (1, 2, 3, 4, 8):
        b = balanced_bounds(deg, world)
        assert b[0] == 0 and b[-1] == deg.numel() and all(b[i] <= b[i + 1] for i in range(world))
        plans = [PartitionPlan(b, r) for r in range(world)]
        assert sum(p.n_local for p in plans) == deg.numel()
        node = torch.arange(deg.numel())
        padded = plans[0].to_padded(node)
        assert padded.unique().numel() == deg.numel()          # injective
        own = plans[0].owner(node)
        for r, p in enumerate(plans):
            assert torch.equal(own == r, (node >= p.lo) & (node < p.hi))
