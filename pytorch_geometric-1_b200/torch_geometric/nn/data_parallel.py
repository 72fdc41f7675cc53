import warnings

import torch

from ..data import Batch


class DataParallel(torch.nn.DataParallel):
    r"""List-of-``Data`` front end used by the fork (ConvexPruning.py:530,559).

    The reference splits the list across the visible GPUs inside one process
    (SURVEY A.12).  That strategy is out of scope here: multi-GPU runs use one
    process per GPU with a destination-partitioned graph (``torch_geometric.dist``).
    This class keeps the single-device behaviour: batch the list, move it to the
    module's device, call the module."""

    def __init__(self, module, device_ids=None, output_device=None):
        torch.nn.Module.__init__(self)
        self.module = module
        self.device_ids = list(device_ids) if device_ids is not None else \
            list(range(torch.cuda.device_count()))[:1]
        self.output_device = output_device
        self.dim = 0
        if len(self.device_ids) > 1:
            warnings.warn('torch_geometric.nn.DataParallel (B200 build) runs on one device; '
                          'use torch_geometric.dist for multi-GPU graphs.')
            self.device_ids = self.device_ids[:1]
        if not self.device_ids:
            raise RuntimeError('DataParallel needs a CUDA device; there is no CPU path')
        self.src_device = torch.device('cuda:{}'.format(self.device_ids[0]))

    def forward(self, data_list):
        if len(data_list) == 0:
            warnings.warn('DataParallel received an empty data list, which may result in '
                          'unexpected behaviour.')
            return None
        data = Batch.from_data_list(data_list).to(self.src_device)
        return self.module(data)
