#!/usr/bin/env python
"""Time K5 (tcgen05 3xTF32 / TF32) against the library fp32 GEMM on the X·W shapes of
the BASELINE configs."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch  # noqa: E402


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument('--only', default='')
    ap.add_argument('--kinds', default='lib,3xtf32,tf32')
    args = ap.parse_args()
    from torch_geometric import ops
    dev = torch.device('cuda:0')
    shapes = [('cfg5 L1', 2449029, 100, 256), ('cfg5 L2', 2449029, 256, 256),
              ('cfg5 L3', 2449029, 256, 47), ('cfg3 (K padded to 604)', 232965, 604, 64), ('cfg2 L1', 19717, 500, 512)]
    for name, M, K, N in shapes:
        if args.only and args.only not in name:
            continue
        if 'xtg' in args.kinds.split(',') and K <= 256 and N <= 256:
            # dW = x^T g (K5t) against the library's fp32 GEMM
            x = torch.randn(M, K, device=dev)
            # gradients reach the GEMM with rows padded to 16 bytes (outputs of K1 / K5)
            g = torch.randn(M, (N + 3) // 4 * 4, device=dev)[:, :N]
            ref = torch.matmul(x.t(), g)
            for kind, fn in (('lib x^T g', lambda: torch.matmul(x.t(), g)),
                             ('xtg 3xtf32', lambda: ops.gemm_xtg_raw(x, g, passes=3)),
                             ('xtg tf32', lambda: ops.gemm_xtg_raw(x, g, passes=1))):
                ms = timeit(fn)
                err = (fn() - ref).abs().max().item() / ref.abs().max().item()
                print(json.dumps({'shape': name + ' dW', 'R': M, 'Kx': K, 'Ng': N, 'kernel': kind,
                                  'ms': round(ms, 4), 'TFLOPs': round(2.0 * M * K * N / ms / 1e9, 1),
                                  'GBps': round(4.0 * M * (K + N) / ms / 1e6, 1),
                                  'rel_vs_lib': float('%.3e' % err)}), flush=True)
            del x, g, ref
        x = torch.randn(M, K, device=dev)
        w = torch.randn(K, N, device=dev) / K ** 0.5
        ref = None
        for kind in args.kinds.split(','):
            if kind == 'xtg':
                continue
            if kind == 'lib':
                fn = lambda: torch.matmul(x, w)
            else:
                p = 3 if kind == '3xtf32' else 1
                fn = lambda: ops.gemm_xw_raw(x, w, None, passes=p)
            ms = timeit(fn)
            out = fn()
            if ref is None:
                ref = out
            err = (out - ref).abs().max().item() / max(ref.abs().max().item(), 1e-30)
            flops = 2.0 * M * K * N
            bytes_ = 4.0 * (M * K + M * N)
            print(json.dumps({'shape': name, 'M': M, 'K': K, 'N': N, 'kernel': kind,
                              'ms': round(ms, 4), 'TFLOPs': round(flops / ms / 1e9, 1),
                              'GBps': round(bytes_ / ms / 1e6, 1), 'rel_vs_lib': float('%.3e' % err)}),
                  flush=True)


if __name__ == '__main__':
    main()
