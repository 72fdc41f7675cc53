#!/usr/bin/env python
"""Where does the routed aggregation lose time when its rows go to peer HBM?

Two ranks (torchrun), whole products-shaped graph on each, K1 at width F with the
output table pointing at (local) this rank's own arena for every row block,
(remote) the peer's arena for every row block, (mixed) the production mapping —
first with only rank 0 running (the peer idle: no incoming traffic), then with both
ranks running at once.

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29533 tools/peer_k1_bench.py --widths 64,32
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--widths', default='64,32')
    ap.add_argument('--iters', type=int, default=10)
    args = ap.parse_args()
    from bench import WORKLOADS
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    from torch_geometric.dist import SlicedGraph
    from torch_geometric.nn import GCNConv

    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    dev = torch.device('cuda', int(os.environ['LOCAL_RANK']))
    dist.init_process_group('nccl', device_id=dev)
    n, pairs, _, _, _ = WORKLOADS['products']
    data = synthetic.products_like(seed=4, device=dev, num_nodes=n, num_pairs=pairs, num_features=4)
    ei, norm = GCNConv.norm(data.edge_index, n, None, False, torch.float32)
    g = SlicedGraph(ei, norm, n, rank, world, transport='peer')
    tr, p = g.transport, g.plan
    tr.prepare(256 * world)
    off = tr._r_off[0]
    for F, xsrc in [(int(f), s_) for f in args.widths.split(',')
                    for s_ in ('torch', 'arena', 'arena-peer-written')]:
        x = torch.randn(p.n_padded, F, device=dev)
        if xsrc != 'torch':
            # the column buffer of the arena, filled locally or by the peer's scatter kernel
            xa = tr._view(tr._c_off, p.n_padded, F)
            if xsrc == 'arena':
                xa.copy_(x)
            else:
                rows = torch.randn(p.n_local, F * world, device=dev)
                ops.cols_scatter_raw(rows, F, world, tr._cols_table(), p.lo, F, False, 1)
                torch.cuda.synchronize()
                dist.barrier()
            x = xa
        ldo = F * world
        maps = {'local': [tr._peer[rank]] * world,
                'remote': [tr._peer[(rank + 1) % world]] * world,
                'mixed': list(tr._peer), 'mixed-compact': list(tr._peer)}
        for who in ('all',):
            for name, bases in maps.items():
                # compact: this rank's slices of a block are contiguous (slice-major row buffer)
                compact = name == 'mixed-compact'
                ld = F if compact else ldo
                tab = torch.tensor([b + off + (rank * p.blk * F * 4 if compact else rank * F * 4)
                                    for b in bases], dtype=torch.int64).to(dev)
                active = who == 'all' or rank == 0
                start = 0
                torch.cuda.synchronize()
                dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if active:
                    for _ in range(2):
                        ops.seg_gather_reduce_routed_raw(x, g.adj.dst, g.w_csr, None, tab, p.blk, ld,
                                                         start_row=start)
                    e0.record()
                    for _ in range(args.iters):
                        ops.seg_gather_reduce_routed_raw(x, g.adj.dst, g.w_csr, None, tab, p.blk, ld,
                                                         start_row=start)
                    e1.record()
                torch.cuda.synchronize()
                dist.barrier()
                if rank == 0:
                    print(json.dumps({'F': F, 'x': xsrc, 'ranks_running': who, 'rows_go_to': name,
                                      'ms': round(e0.elapsed_time(e1) / args.iters, 4)}), flush=True)
    tr.close()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
