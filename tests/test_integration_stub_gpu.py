"""GPU: the reference-side binding printed in INTEGRATION.md section 2 (a ``torch_scatter``
module over ``libpygb.so`` through ctypes) is executed verbatim and checked against the
oracle's ``torch_scatter`` restatement — the stub a maintainer would add really works."""
import os
import re
import types

import pytest
import torch

from oracle import ref_cpu as O
from util import assert_close, skewed_graph

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _stub_module():
    from torch_geometric import _native
    text = open(os.path.join(ROOT, 'INTEGRATION.md')).read()
    blocks = re.findall(r'```python\n(.*?)```', text, flags=re.S)
    code = [b for b in blocks if 'torch_scatter/__init__.py' in b]
    assert len(code) == 1, 'stub code block not found in INTEGRATION.md'
    src = code[0].replace("ctypes.CDLL('libpygb.so')", 'ctypes.CDLL(%r)' % _native.LIB_PATH)
    mod = types.ModuleType('torch_scatter_stub')
    exec(compile(src, 'INTEGRATION.md:torch_scatter', 'exec'), mod.__dict__)
    return mod


def test_integration_stub_matches_oracle(dev):
    ts = _stub_module()
    n, E, F = 700, 9000, 24
    idx = skewed_graph(n, E, 3)[1]
    src = torch.randn(E, F, generator=torch.Generator().manual_seed(1))
    got = ts.scatter_add(src.to(dev), idx.to(dev), 0, None, n)
    assert_close(got, O.scatter_('add', src, idx, 0, n), 1e-5, 'stub scatter_add')
    got = ts.scatter_mean(src.to(dev), idx.to(dev), 0, None, n)
    assert_close(got, O.scatter_('mean', src, idx, 0, n), 1e-5, 'stub scatter_mean')
    val, arg = ts.scatter_max(src.to(dev), idx.to(dev), 0, None, n)
    rv, ra = O.scatter_max(src, idx, 0, None, n)
    bad = (val.cpu() != rv).nonzero()
    assert bad.numel() == 0, ('stub scatter_max values', bad[:5].tolist(),
                              val.cpu()[bad[0, 0], bad[0, 1]].item(), rv[bad[0, 0], bad[0, 1]].item(),
                              arg.cpu()[bad[0, 0], bad[0, 1]].item(), ra[bad[0, 0], bad[0, 1]].item())
    assert torch.equal(arg.cpu(), ra), 'stub scatter_max argmax (bit-exact; empty rows E)'
    # dim_size inferred from the index, 3-D sources
    s3 = torch.randn(50, 2, 3, generator=torch.Generator().manual_seed(2))
    i3 = torch.randint(0, 9, (50,), generator=torch.Generator().manual_seed(3))
    got = ts.scatter_add(s3.to(dev), i3.to(dev))
    assert_close(got, O.scatter_('add', s3, i3, 0, int(i3.max()) + 1), 1e-5, 'stub 3-D')
