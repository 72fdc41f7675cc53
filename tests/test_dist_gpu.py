"""GPU, 2 ranks, NCCL: the destination-partitioned GCN forward (native kernels on
each rank + halo all-gather over NVLink) against the oracle on the whole graph.
Skipped when fewer than two GPUs are visible."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, 'pytorch_geometric-1_b200'), os.path.join(root, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        from oracle import ref_cpu as O
        from util import skewed_graph
        from torch_geometric.dist import DistGraph, dist_gcn_forward
        from torch_geometric.nn import GCNConv
        torch.manual_seed(0)
        n = 3001
        ei = skewed_graph(n, 60000, 11)
        x = torch.randn(n, 100)
        dims = [(100, 256), (256, 256), (256, 47)]
        layers = [(torch.randn(a, b) / a ** 0.5, torch.randn(b)) for a, b in dims]
        ei2, norm = GCNConv.norm(ei.to(dev), n, None, False, torch.float32)
        g = DistGraph(ei2, norm, n, rank, world)
        out_local = dist_gcn_forward(g, g.local_rows(x.to(dev)).contiguous(),
                                     [(w.to(dev), b.to(dev)) for w, b in layers])
        parts = [torch.zeros(g.plan.block, 47, device=dev) for _ in range(world)]
        pad = torch.zeros(g.plan.block, 47, device=dev)
        pad[:g.plan.n_local] = out_local
        dist.all_gather(parts, pad)
        if rank == 0:
            got = torch.cat([parts[p][:g.plan.bounds[p + 1] - g.plan.bounds[p]]
                             for p in range(world)]).cpu()
            e2, nrm = O.gcn_norm(ei, n, dtype=torch.float32)
            assert torch.equal(e2, ei2.cpu())
            h = x
            for k, (w, b) in enumerate(layers):
                h = O.scatter_('add', nrm.view(-1, 1) * (h @ w)[e2[0]], e2[1], 0, n) + b
                if k + 1 < len(layers):
                    h = torch.relu(h)
            q.put(('ok', (got - h).abs().max().item() / h.abs().max().item()))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def test_dist_gcn_two_gpus_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    status, err = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 2e-5, err
