"""Halo exchange for the destination-partitioned graph: an all-gather of the
source-node feature blocks.  ``torch.distributed`` is the plumbing
(NCCL over NVLink/NVSwitch on GPUs; gloo in the CPU unit tests).

The gather buffer is ``[P*block, ld]`` (rank-major, every rank's block padded to
``plan.block`` rows; ``ld`` = the width padded to 16 bytes so the aggregation
kernel keeps its 128-bit path).  Producers write their rows straight into
``slot()``, so the collective runs in place."""
import torch
import torch.distributed as dist


class HaloExchange(object):
    def __init__(self, plan, group=None):
        self.plan = plan
        self.group = group
        self._buf = {}
        self.bytes_received_total = 0
        self.event_log = None  # list -> (start, end, bytes) CUDA events per collective (bench.py)

    def _ld(self, F, itemsize):
        vec = max(16 // itemsize, 1)
        return (F + vec - 1) // vec * vec if F > 1 else F

    def _buffer(self, F, dtype, device):
        key = (F, dtype, str(device))
        buf = self._buf.get(key)
        if buf is None:
            ld = self._ld(F, torch.empty((), dtype=dtype).element_size())
            buf = torch.zeros((self.plan.n_padded, ld), dtype=dtype, device=device)
            self._buf[key] = buf
        return buf

    def slot(self, F, dtype, device):
        """View ``[n_local, ld]`` of this rank's rows inside the gather buffer."""
        plan = self.plan
        full = self._buffer(F, dtype, device)
        r0 = plan.rank * plan.block
        return full[r0:r0 + plan.n_local]

    def all_gather_slot(self, F):
        """All-gather in place after the rank's rows were written through ``slot()``;
        returns the gathered ``[P*block, F]`` view."""
        plan = self.plan
        full = [b for (f, _, _), b in self._buf.items() if f == F][-1]
        if plan.world_size > 1:
            mine = full[plan.rank * plan.block:(plan.rank + 1) * plan.block]
            nbytes = self.bytes_received(full.size(1), full.element_size())
            if self.event_log is not None and full.is_cuda:
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.all_gather_into_tensor(full, mine, group=self.group)
                e1.record()
                self.event_log.append((e0, e1, nbytes))
            else:
                dist.all_gather_into_tensor(full, mine, group=self.group)
            self.bytes_received_total += nbytes
        return full[:, :F] if full.size(1) != F else full

    def all_gather(self, x_local):
        """x_local [n_local, F] -> gathered [P*block, F]."""
        F = x_local.size(1)
        self.slot(F, x_local.dtype, x_local.device)[:, :F].copy_(x_local)
        return self.all_gather_slot(F)

    def bytes_received(self, ld, itemsize):
        p = self.plan
        return (p.world_size - 1) * p.block * ld * itemsize


    def reduce_scatter(self, g_full):
        """Transpose of ``all_gather``: ``g_full [P*block, F]`` (every rank's gradient for ALL source
        rows) -> the sum over ranks of the rows this rank owns, ``[block, F]``.  NCCL reduce-scatter;
        gloo (CPU unit tests) has none, so there it is an all-reduce and a slice."""
        plan = self.plan
        r0 = plan.rank * plan.block
        if plan.world_size == 1:
            return g_full[r0:r0 + plan.block]
        g_full = g_full.contiguous()
        if g_full.is_cuda:
            out = torch.empty((plan.block, g_full.size(1)), dtype=g_full.dtype, device=g_full.device)
            dist.reduce_scatter_tensor(out, g_full, op=dist.ReduceOp.SUM, group=self.group)
            return out
        g_full = g_full.clone()
        dist.all_reduce(g_full, op=dist.ReduceOp.SUM, group=self.group)
        return g_full[r0:r0 + plan.block]


class AllGatherRows(torch.autograd.Function):
    """``halo.all_gather`` with its transpose as the backward (training on the partitioned graph):
    forward gathers the rank blocks ``[n_local, F] -> [P*block, F]``, backward reduce-scatters the
    gradient of the gathered matrix back to the owners of the rows.  The output is a copy — the
    gather buffer is reused by the next layer of the same width, autograd must own what it saved."""

    @staticmethod
    def forward(ctx, x_local, halo):
        ctx.halo = halo
        ctx.n_local = x_local.size(0)
        return halo.all_gather(x_local).clone()

    @staticmethod
    def backward(ctx, g_full):
        return ctx.halo.reduce_scatter(g_full)[:ctx.n_local], None
