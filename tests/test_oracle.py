"""CPU: the oracle against the Appendix A.8 known-answer vectors, its own
literal-loop restatement and the C restatement."""
import ctypes
import os

import pytest
import torch

from oracle import ref_cpu as O
from util import kat, random_graph

K = kat()
T = torch.tensor


def test_softmax_kat():
    k = K['softmax']
    assert torch.allclose(O.softmax(T(k['src']), T(k['index'])), T(k['out']))


def test_degree_kat():
    k = K['degree']
    out = O.degree(T(k['index']), dtype=torch.long)
    assert out.dtype == torch.long and out.tolist() == k['out']


def test_loops_kat():
    k = K['remove_self_loops']
    assert O.remove_self_loops(T(k['edge_index']))[0].tolist() == k['out']
    k = K['add_self_loops']
    ei, w = O.add_self_loops(T(k['edge_index']), T(k['w']))
    assert ei.tolist() == k['out'] and w.tolist() == k['w_out']
    k = K['add_remaining_self_loops']
    ei, w = O.add_remaining_self_loops(T(k['edge_index']), T(k['w']))
    assert ei.tolist() == k['out'] and w.tolist() == k['w_out']


def test_gcn_norm_kat():
    k = K['gcn_norm']
    ei, norm = O.gcn_norm(T(k['edge_index']), k['num_nodes'], dtype=torch.float32)
    assert ei.tolist() == k['out']
    assert torch.allclose(norm, T(k['norm']), rtol=1e-6, atol=0)


def test_scatter_max_kat():
    k = K['scatter_max']
    out, arg = O.scatter_max(T(k['src']), T(k['index']), 0, None, k['dim_size'])
    assert out.tolist() == k['out'] and arg.tolist() == k['arg']
    assert O.scatter_('max', T(k['src']), T(k['index']), 0, k['dim_size']).tolist() == k['scatter_']
    k = K['scatter_mean']
    assert O.scatter_('mean', T(k['src']), T(k['index']), 0, k['dim_size']).tolist() == k['out']


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_scatter_max_vectorised_equals_loop(seed):
    g = torch.Generator().manual_seed(seed)
    src = (torch.randn(300, 4, generator=g) * 3).round() / 2  # many ties
    src[5, 1] = float('nan')
    src[9, 0] = float('-inf')
    src[11, 2] = -30000.0
    idx = torch.randint(0, 40, (300,), generator=g)
    a = O.scatter_max(src, idx, 0, None, 45)
    b = O.scatter_max_loop(src, idx, 45)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_remaining_loops_last_write_wins():
    ei = T([[0, 1, 1, 1, 2], [1, 1, 0, 1, 2]])
    w = T([1., 2., 3., 4., 5.])
    out_ei, out_w = O.add_remaining_self_loops(ei, w, 9.0, 4)
    assert out_ei.tolist() == [[0, 1, 0, 1, 2, 3], [1, 0, 0, 1, 2, 3]]
    assert out_w.tolist() == [1., 3., 9., 4., 5., 9.]


def _clib():
    path = os.path.join(os.path.dirname(O.__file__), '_ref', 'libscatter_ref.so')
    if not os.path.exists(path):
        import subprocess
        subprocess.check_call(['make', '-C', os.path.dirname(O.__file__)])
    return ctypes.CDLL(path)


def test_c_restatement_matches_python_oracle():
    lib = _clib()
    E, K_, N = 500, 7, 60
    ei = random_graph(N, N, E, 3)
    g = torch.Generator().manual_seed(3)
    src = (torch.randn(E, K_, generator=g) * 2).round() / 2
    P = ctypes.c_void_p
    L = ctypes.c_int64
    out = torch.zeros(N, K_)
    lib.ref_scatter_add(P(src.data_ptr()), P(ei[1].contiguous().data_ptr()), L(E), L(K_),
                        P(out.data_ptr()))
    assert torch.equal(out, O.scatter_add(src, ei[1], 0, None, N))  # same sequential order
    outm = torch.empty(N, K_)
    arg = torch.empty(N, K_, dtype=torch.long)
    lib.ref_scatter_max(P(src.data_ptr()), P(ei[1].contiguous().data_ptr()), L(E), L(K_), L(N),
                        P(outm.data_ptr()), P(arg.data_ptr()))
    ov, oa = O.scatter_max(src, ei[1], 0, None, N)
    assert torch.equal(outm, ov) and torch.equal(arg, oa)
    mean = torch.empty(N, K_)
    cnt = torch.empty(N)
    lib.ref_scatter_mean(P(src.data_ptr()), P(ei[1].contiguous().data_ptr()), L(E), L(K_), L(N),
                         P(mean.data_ptr()), P(cnt.data_ptr()))
    assert torch.allclose(mean, O.scatter_mean(src, ei[1], 0, None, N), rtol=1e-6, atol=1e-7)


def test_message_passing_oracle_shapes_and_flow():
    class Conv(O.MessagePassing):
        def message(self, x_i, x_j, edge_index_i, size_i):
            assert size_i == 6 and edge_index_i.numel() == x_i.size(0)
            return x_j - x_i

    ei = random_graph(6, 6, 20, 0)
    x = torch.randn(6, 3)
    out = Conv(aggr='mean')(ei, x=x) if False else Conv(aggr='mean').propagate(ei, x=x)
    assert out.shape == (6, 3)
    out_t = Conv(aggr='mean', flow='target_to_source').propagate(ei.flip(0), x=x)
    assert torch.allclose(out, out_t)
    with pytest.raises(ValueError):
        Conv().propagate(ei, x=(torch.randn(6, 3), torch.randn(7, 3)), size=(6, 6))
