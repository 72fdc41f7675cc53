"""Halo exchange for the destination-partitioned graph: an all-gather of the
source-node feature blocks.  ``torch.distributed`` is the plumbing
(NCCL over NVLink/NVSwitch on GPUs; gloo in the CPU unit tests)."""
import torch
import torch.distributed as dist


class HaloExchange(object):
    def __init__(self, plan, group=None):
        self.plan = plan
        self.group = group
        self._buf = {}

    def _buffers(self, F, dtype, device):
        key = (F, dtype, device)
        if key not in self._buf:
            full = torch.zeros((self.plan.n_padded, F), dtype=dtype, device=device)
            self._buf = {key: full}  # keep one shape alive
        return self._buf[key]

    def all_gather(self, x_local):
        """x_local [n_local, F] -> gathered [P*block, F] (rank-major, each block
        padded to ``plan.block`` rows).  The rank's own block is written in place
        into the gather buffer, so no extra copy of it is made."""
        plan = self.plan
        F = x_local.size(1)
        full = self._buffers(F, x_local.dtype, x_local.device)
        mine = full[plan.rank * plan.block:(plan.rank + 1) * plan.block]
        mine[:plan.n_local].copy_(x_local)
        if plan.world_size > 1:
            dist.all_gather_into_tensor(full, mine, group=self.group)
        return full

    def bytes_received(self, F, itemsize):
        p = self.plan
        return (p.world_size - 1) * p.block * F * itemsize
