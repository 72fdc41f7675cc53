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


def _train_worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torch_geometric.dist import DistGraph, all_reduce_grads, dist_gcn_forward
        torch.manual_seed(0)
        n = 257
        ei = random_graph(n, n, 3000, 6)
        ei[1, :500] = 11                                  # a hub target
        x = torch.randn(n, 8, dtype=torch.float64)
        dims = [(8, 16), (16, 16), (16, 5)]               # aggregate-first, transform-first, narrow
        layers = [((torch.randn(a, b, dtype=torch.float64) / a ** 0.5).requires_grad_(True),
                   torch.randn(b, dtype=torch.float64).requires_grad_(True)) for a, b in dims]
        coef = torch.randn(n, 5, dtype=torch.float64)     # loss = <out, coef>: a distinct weight per element
        ei2, norm = O.gcn_norm(ei, n, dtype=torch.float64)
        g = DistGraph(ei2, norm, n, rank, world, build_adj=False)
        x_loc = g.local_rows(x).clone().requires_grad_(True)
        # x * sigmoid(x): smooth, so the comparison has no activation kinks
        out_local = dist_gcn_forward(g, x_loc, layers, activation=lambda t: t * torch.sigmoid(t),
                                     aggregate=_oracle_aggregate, transform=_cpu_transform)
        loss = (out_local * g.local_rows(coef)).sum()
        loss.backward()
        all_reduce_grads([p for wb in layers for p in wb])
        total = loss.detach().clone()
        dist.all_reduce(total)
        gx = torch.zeros(g.plan.block, 8, dtype=torch.float64)
        gx[:g.plan.n_local] = x_loc.grad
        parts = [torch.zeros_like(gx) for _ in range(world)]
        dist.all_gather(parts, gx)
        if rank == 0:
            got_gx = torch.cat([parts[p][:g.plan.bounds[p + 1] - g.plan.bounds[p]] for p in range(world)])
            xr = x.clone().requires_grad_(True)
            ref_layers = [(w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True))
                          for w, b in layers]
            h = xr
            for k, (w, b) in enumerate(ref_layers):
                h = O.scatter_('add', norm.view(-1, 1) * (h @ w)[ei2[0]], ei2[1], 0, n) + b
                if k + 1 < len(ref_layers):
                    h = h * torch.sigmoid(h)
            ref_loss = (h * coef).sum()
            ref_loss.backward()
            errs = [abs(total.item() - ref_loss.item()) / abs(ref_loss.item()),
                    ((got_gx - xr.grad).abs().max() / xr.grad.abs().max()).item()]
            for (w, b), (rw, rb) in zip(layers, ref_layers):
                errs.append(((w.grad - rw.grad).abs().max() / rw.grad.abs().max()).item())
                errs.append(((b.grad - rb.grad).abs().max() / rb.grad.abs().max()).item())
            q.put(('ok', max(errs)))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e)))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_dist_gcn_training_step_gloo(world):
    """Forward + backward on the destination-partitioned graph: the exchange is an autograd node
    (all-gather forward, reduce-scatter backward), weight gradients are all-reduced; loss, d x and
    every d W / d b equal the single-process autograd of the oracle (fp64)."""
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_train_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    status, err = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 1e-10, err


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


# ---- feature-sliced layout (dist/sliced.py) ------------------------------------------
def _sliced_oracle_aggregate(graph, cols, bias):
    ei, n = graph.edge_index, graph.plan.n_group        # in-edges of the rank's row group
    msg = graph.norm.view(-1, 1) * cols[ei[1 - graph.i]]
    out = O.scatter_('add', msg, ei[graph.i], 0, n)
    return out + bias if bias is not None else out


def _sliced_transform(x, weight, bias, out, relu=False):
    y = x @ weight
    y = y + bias if bias is not None else y
    out[:, :y.size(1)].copy_(torch.relu(y) if relu else y)
    return out


def _sliced_worker(rank, world, port, q, x_layout, row_groups):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from torch_geometric.dist import SlicedGraph, sliced_gcn_forward
        torch.manual_seed(0)
        n = 301                                        # not a multiple of the world size
        ei = random_graph(n, n, 4000, 5)
        ei[1, :600] = 7
        x = torch.randn(n, 10)                         # 10 columns: padded to 4 * world
        dims = [(10, 16), (16, 16), (16, 5), (5, 7)]   # (AX)W, (AX)W, A(XW), (AX)W
        layers = [(torch.randn(a, b) / a ** 0.5, torch.randn(b)) for a, b in dims]
        ei2, norm = O.gcn_norm(ei, n, dtype=torch.float32)
        g = SlicedGraph(ei2, norm, n, rank, world, transport='collective', build_adj=False,
                        row_groups=row_groups)
        pl = g.plan
        assert pl.n_padded >= n and pl.blk * (world - 1) < n
        assert g.num_edges_local == int(((ei2[1] >= pl.g_lo) & (ei2[1] < pl.g_hi)).sum())
        xin = {'cols': lambda: g.slice_cols(x), 'full': lambda: x,
               'rows': lambda: g.local_rows(x).contiguous()}[x_layout]()
        out_local = sliced_gcn_forward(g, xin, layers, relu=True, x_layout=x_layout,
                                       aggregate=_sliced_oracle_aggregate,
                                       transform=_sliced_transform)
        assert out_local.shape == (pl.n_local, 7)
        parts = [torch.zeros(pl.blk, 7) for _ in range(world)]
        pad = torch.zeros(pl.blk, 7)
        pad[:pl.n_local] = out_local
        dist.all_gather(parts, pad)
        if rank == 0:
            got = torch.cat(parts)[:n]
            h = x
            for k, (w, b) in enumerate(layers):
                h = O.scatter_('add', norm.view(-1, 1) * (h @ w)[ei2[0]], ei2[1], 0, n) + b
                if k + 1 < len(layers):
                    h = torch.relu(h)
            err = (got - h).abs().max().item() / h.abs().max().item()
            q.put(('ok', err, g.transport.bytes_sent_total, g.transport.collectives))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e), None, None))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world,x_layout,row_groups', [(2, 'cols', 1), (2, 'rows', 1),
                                                       (3, 'cols', 1), (4, 'cols', 2),
                                                       (2, 'rows', 2), (2, 'full', 1),
                                                       (4, 'full', 2)])
def test_sliced_gcn_gloo(world, x_layout, row_groups):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sliced_worker, args=(r, world, port, q, x_layout, row_groups))
             for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    status, err, nbytes, ncoll = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 1e-5, err
    # layout changes: one per aggregation (cols -> rows) + one per rows -> cols
    # ('full' with an injected aggregate: the replicated input is column-sliced locally)
    assert ncoll == (8 if x_layout == 'rows' else 7) and nbytes > 0


def test_slice_plan_properties():
    from torch_geometric.dist import SlicePlan
    for n in (2, 7, 301, 2449029):
        for world, gr in ((1, 1), (2, 1), (3, 1), (4, 1), (8, 1), (2, 2), (4, 2), (8, 2), (8, 4)):
            plans = [SlicePlan(n, world, r, gr) for r in range(world)]
            gc = world // gr
            assert sum(p.n_local for p in plans) == n
            assert all(p.lo == min(r * p.blk, n) for r, p in enumerate(plans))
            # a row group is the union of the row blocks of its ranks
            for p in plans:
                mine = [q for q in plans if q.rg == p.rg]
                assert p.g_lo == min(q.lo for q in mine) and p.g_hi == max(q.hi for q in mine)
                assert p.rank == p.rg * gc + p.cg
            assert sum(p.n_group for p in plans if p.cg == 0) == n
            for F in (1, 5, 47, 100, 256):
                Fp, Fc = plans[0].widths(F, 4)
                assert Fp >= F and Fp == Fc * gc and Fc % 8 == 0 and Fp - F < 32 * gc
                cover = [plans[0].col_slice(F, 4, c) for c in range(gc)]
                assert cover[0].start == 0 and cover[-1].stop == Fp
    # blk_align (peer transport): row blocks that split into equal row chunks
    for n in (7, 301, 2449029):
        pa = [SlicePlan(n, 8, r, 2, blk_align=8) for r in range(8)]
        assert pa[0].blk % 8 == 0 and sum(p.n_local for p in pa) == n
        assert pa[0].blk - SlicePlan(n, 8, 0, 2).blk < 8
    # exchange volume: P times less than the all-gather of the row partition
    p8 = SlicePlan(2449029, 8, 0)
    assert p8.exchange_bytes(256) * 8 <= 7 * p8.blk * 256 * 4 + 8
