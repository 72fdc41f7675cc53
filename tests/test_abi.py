"""CPU: libpygb.so loads and exports exactly the symbols include/pygb.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, 'include', 'pygb.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(pygb_[a-z0-9_]+)\s*\(', src)))


def test_header_symbols_exported():
    from torch_geometric import _native
    if not os.path.exists(_native.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(_native.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), 'missing export {}'.format(n)
    # and the ctypes table covers the header one to one
    assert sorted(_native.SIGNATURES) == names
    assert lib.pygb_abi_version() == 4


def test_no_cpu_fallback():
    import torch
    from torch_geometric import _native
    from torch_geometric.nn import GCNConv
    conv = GCNConv(4, 3)
    with pytest.raises(_native.NativeError):
        conv(torch.randn(5, 4), torch.tensor([[0, 1], [1, 0]]))
    from torch_geometric.utils import scatter_, softmax, degree
    with pytest.raises(_native.NativeError):
        scatter_('add', torch.randn(3, 2), torch.tensor([0, 1, 1]), 0, 2)
    with pytest.raises(_native.NativeError):
        softmax(torch.randn(3), torch.tensor([0, 1, 1]), 2)
    with pytest.raises(_native.NativeError):
        degree(torch.tensor([0, 1, 1]), 2)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'pytorch_geometric-1_b200')
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                txt = open(os.path.join(dp, f)).read()
                assert 'oracle' not in txt.replace('the oracle', '').replace("oracle's", ''), \
                    '{} mentions oracle'.format(f)


def test_integration_doc_lists_every_entry_point():
    """INTEGRATION.md tells a maintainer of the reference what each entry point replaces: a new
    declaration in the header without a line there is documentation drift."""
    doc = open(os.path.join(ROOT, 'INTEGRATION.md')).read()
    listed = set(re.findall(r'pygb_[a-z0-9_]+', doc))
    for m in re.finditer(r'(pygb_[a-z0-9_]*)\{([^}]*)\}', doc):           # pygb_peer_{alloc,free,...}
        listed.update(m.group(1) + part.strip() for part in m.group(2).split(','))
    for m in re.finditer(r'(pygb_[a-z0-9_]+)\(_([a-z0-9_]+)\)', doc):     # pygb_edge_softmax(_backward)
        listed.add(m.group(1) + '_' + m.group(2))
    missing = [n for n in _declared() if n not in listed]
    assert not missing, missing
