#!/usr/bin/env python
"""The three X·W launches of the products step (cfg5) once each, for an ncu capture of
`gemm_ts_kernel` (profiles/r02_k5_ts_ncu.md):

    python tools/ncu_k5.py && ncu --set full --clock-control none --import-source on \
        -k regex:gemm_ts_kernel -o gpurun_out/r02_k5_ts python tools/ncu_k5.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch  # noqa: E402


def main():
    from torch_geometric import ops
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(1)
    M = 2449029
    x = torch.randn(M, 256, device=dev, generator=g)
    x100 = torch.zeros(M, 128, device=dev)[:, :100]
    x100.copy_(x[:, :100])
    for a, K, N in ((x100, 100, 256), (x, 256, 256), (x, 256, 47)):
        w = torch.randn(K, N, device=dev, generator=g) / K ** 0.5
        out = ops.gemm_xw_raw(a, w, torch.zeros(N, device=dev))
        torch.cuda.synchronize()
        print('K5 3xTF32 M=%d K=%d N=%d' % (M, K, N), tuple(out.shape), flush=True)


if __name__ == '__main__':
    main()
