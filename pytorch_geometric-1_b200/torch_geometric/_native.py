"""ctypes binding of the C ABI in ``include/pygb.h`` (``lib/libpygb.so``).

This is the only place the Python host touches native code.  There is no CPU
fallback: if the shared library is missing, or a tensor is not on a CUDA
device, the call raises.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'lib', 'libpygb.so')

F32, BF16 = 0, 1
SUM, MEAN = 0, 1
LOOPS_REMOVE, LOOPS_ADD, LOOPS_REMOVE_ADD, LOOPS_ADD_REMAINING = 0, 1, 2, 3

_c = ctypes
_P, _I, _L, _F, _Z = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_float, _c.c_size_t

# name -> (restype, argtypes) ; mirrors include/pygb.h one to one
SIGNATURES = {
    'pygb_abi_version': (_I, []),
    'pygb_last_error': (_c.c_char_p, []),
    'pygb_check_device': (_I, [_P, _P, _P]),
    'pygb_count_self_loops': (_I, [_P, _L, _P, _P]),
    'pygb_scan_edge_index': (_I, [_P, _L, _L, _P, _P]),
    'pygb_apply_loops_workspace': (_Z, [_L, _L]),
    'pygb_apply_loops': (_I, [_P, _L, _L, _I, _L, _P, _F, _P, _P, _P, _Z, _P]),
    'pygb_csr_build_workspace': (_Z, [_L, _L]),
    'pygb_csr_build': (_I, [_P, _P, _L, _L, _L, _P, _P, _P, _P, _P, _Z, _P]),
    'pygb_permute_f32': (_I, [_P, _P, _L, _P, _P]),
    'pygb_degree': (_I, [_P, _L, _L, _I, _P, _P]),
    'pygb_weighted_degree': (_I, [_P, _P, _P, _L, _P, _P]),
    'pygb_gcn_norm': (_I, [_P, _P, _L, _L, _P, _P, _P]),
    'pygb_gcn_norm_csr': (_I, [_P, _P, _P, _L, _P, _P]),
    'pygb_seg_gather_reduce': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _P, _L, _P, _L, _P]),
    'pygb_seg_plan': (_I, [_L, _P, _P]),
    'pygb_seg_tables_workspace': (_Z, [_L]),
    'pygb_seg_tables': (_I, [_P, _L, _L, _I, _L, _P, _P, _P, _Z, _P]),
    'pygb_seg_flag_idx': (_I, [_P, _L, _P, _L, _P, _P]),
    'pygb_seg_stream_workspace': (_Z, [_I, _L, _L]),
    'pygb_seg_gather_reduce_stream': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _L, _L, _P, _P, _I,
                                           _L, _P, _L, _P, _Z, _P]),
    'pygb_seg_mark_hot': (_I, [_P, _L, _P, _L, _I, _P]),
    'pygb_seg_gather_reduce_stream_hinted': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _L, _L, _P, _P,
                                                  _I, _L, _P, _L, _P, _Z, _P]),
    'pygb_seg_gather_reduce_stream_routed_hinted': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _L, _L,
                                                         _P, _P, _I, _L, _P, _L, _L, _L, _L, _P, _Z,
                                                         _P]),
    'pygb_seg_gather_reduce_stream_static': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _L, _L, _P,
                                                  _P, _I, _L, _P, _L, _P]),
    'pygb_seg_gather_reduce_stream_routed': (_I, [_I, _I, _P, _L, _L, _P, _P, _P, _P, _L, _L, _P, _P,
                                                  _I, _L, _P, _L, _L, _L, _L, _P, _Z, _P]),
    'pygb_peer_alloc': (_I, [_Z, _P]),
    'pygb_peer_free': (_I, [_P]),
    'pygb_peer_export': (_I, [_P, _P, _Z]),
    'pygb_peer_open': (_I, [_P, _Z, _P]),
    'pygb_peer_close': (_I, [_P]),
    'pygb_peer_barrier': (_I, [_P, _I, _I, _L, _L, _P]),
    'pygb_cols_scatter': (_I, [_I, _P, _L, _L, _L, _I, _I, _P, _L, _L, _I, _P]),
    'pygb_gemm_xw_routed': (_I, [_P, _L, _L, _L, _L, _L, _P, _L, _L, _P, _I, _P, _L, _I, _L, _L, _I,
                                 _L, _P, _Z, _P]),
    'pygb_seg_gather_reduce_heads': (_I, [_I, _P, _L, _I, _I, _P, _P, _P, _P, _P, _L, _P, _L, _P]),
    'pygb_gather_rows': (_I, [_I, _P, _L, _L, _P, _L, _P, _L, _P]),
    'pygb_edge_dot': (_I, [_I, _P, _L, _P, _L, _L, _P, _P, _P, _L, _L, _P, _P]),
    'pygb_seg_max': (_I, [_I, _P, _L, _L, _P, _P, _P, _L, _L, _F, _I, _P, _L, _P, _P]),
    'pygb_seg_max_backward': (_I, [_P, _L, _P, _L, _L, _L, _P, _L, _P]),
    'pygb_edge_softmax': (_I, [_P, _L, _P, _P, _L, _P, _P]),
    'pygb_edge_softmax_backward': (_I, [_P, _P, _L, _P, _P, _L, _P, _P]),
    'pygb_gat_scores': (_I, [_I, _P, _L, _P, _L, _I, _I, _P, _P, _P]),
    'pygb_gat_forward': (_I, [_I, _P, _L, _P, _P, _I, _I, _F, _P, _P, _L, _P, _P, _L, _P, _P, _P]),
    'pygb_gat_stream_workspace': (_Z, [_L, _I, _I]),
    'pygb_gat_forward_stream': (_I, [_I, _P, _L, _P, _P, _I, _I, _F, _P, _P, _L, _L, _P, _P, _I, _L,
                                     _P, _P, _L, _P, _P, _P, _Z, _P]),
    'pygb_gat_backward_prep': (_I, [_P, _P, _P, _P, _L, _P, _P, _L, _L, _I, _I, _P, _P]),
    'pygb_gat_backward_stream_workspace': (_Z, [_I, _L, _I, _I]),
    'pygb_gat_backward_stream': (_I, [_I, _P, _L, _P, _L, _P, _P, _I, _I, _F, _P, _P, _L, _L, _P, _P,
                                      _I, _L, _P, _L, _P, _P, _Z, _P]),
    'pygb_gat_alpha': (_I, [_P, _P, _I, _F, _P, _P, _P, _L, _P, _P, _P, _P]),
    'pygb_gemm_weight_workspace': (_Z, [_L, _L]),
    'pygb_gemm_xw': (_I, [_P, _L, _L, _L, _L, _L, _P, _L, _L, _P, _I, _P, _L, _I, _P, _Z, _P]),
    'pygb_gemm_xw_bf16': (_I, [_P, _L, _L, _L, _P, _L, _L, _P, _I, _P, _L, _P, _Z, _P]),
    'pygb_knn_graph': (_I, [_P, _L, _L, _I, _P, _P, _I, _I, _P, _P]),
    'pygb_fps': (_I, [_P, _L, _L, _I, _P, _P, _P, _L, _P, _P, _P]),
    'pygb_radius': (_I, [_P, _L, _P, _L, _I, _P, _P, _L, _F, _I, _I, _P, _P, _P]),
    'pygb_gemm_xtg_workspace': (_Z, [_L, _L, _L]),
    'pygb_gemm_xtg': (_I, [_P, _L, _P, _L, _L, _L, _L, _P, _L, _I, _P, _Z, _P]),
    'pygb_gat_edge_grad': (_I, [_P, _L, _P, _P, _I, _I, _F, _P, _P, _P, _L, _P, _L, _P, _P, _L,
                                _P, _P, _P, _P, _P]),
}

_lib = None
_lock = threading.Lock()
_checked_devices = set()
launches = 0  # number of native calls issued (bench.py reports it)
# Optional profiler: a list that receives (name, info, start_event, end_event) for
# every native call — CUDA events recorded on the launching stream (bench.py).
event_log = None


class NativeError(RuntimeError):
    pass


if os.environ.get('PYGB_LOG_NATIVE_CALLS', '0') not in ('', '0'):
    import atexit

    @atexit.register
    def _report_calls():          # evidence that a user script ran on libpygb.so (tests)
        print('[pygb] native calls: {}'.format(launches), flush=True)


def load():
    """Load libpygb.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                'libpygb.so not found at {} — build it with `python -c "import '
                '__graft_entry__ as g; g.build()"` or `make -C '
                'pytorch_geometric-1_b200/csrc`. There is no CPU fallback.'.format(LIB_PATH))
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        if lib.pygb_abi_version() != 4:
            raise NativeError('libpygb.so ABI version mismatch')
        _lib = lib
    return _lib


def _check_device(device):
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx in _checked_devices:
        return
    lib = load()
    with torch.cuda.device(idx):
        rc = lib.pygb_check_device(None, None, None)
    if rc != 0:
        raise NativeError(lib.pygb_last_error().decode())
    _checked_devices.add(idx)


def require_cuda(*tensors):
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise NativeError(
                'torch_geometric (B200 build) runs on CUDA tensors only; got a {} '
                'tensor. There is no CPU fallback.'.format(t.device))
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise NativeError('tensors on different devices: {} vs {}'.format(dev, t.device))
    _check_device(dev)
    return dev


def ptr(t):
    return None if t is None else t.data_ptr()


def _raw_stream(idx):
    """cudaStream_t of the current stream of device ``idx`` without building a Stream object."""
    try:
        return torch._C._cuda_getCurrentRawStream(idx)
    except AttributeError:  # pragma: no cover - older/newer torch builds
        return torch.cuda.current_stream(idx).cuda_stream


def call(name, device, *args):
    """Invoke ``name(*args, stream)`` on ``device``'s current stream."""
    global launches
    lib = load()
    fn = getattr(lib, name)
    idx = device.index
    if event_log is None and (idx is None or idx == torch.cuda.current_device()):
        # fast path (small graphs are launch-latency bound): no device guard, no Stream object
        rc = fn(*args, _raw_stream(torch.cuda.current_device() if idx is None else idx))
    else:
        with torch.cuda.device(device):
            cur = torch.cuda.current_stream()
            if event_log is not None:
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record(cur)
                rc = fn(*args, cur.cuda_stream)
                e1.record(cur)
                event_log.append((name, args, e0, e1))
            else:
                rc = fn(*args, cur.cuda_stream)
    launches += 1
    if rc != 0:
        raise NativeError('{} failed ({}): {}'.format(name, rc, lib.pygb_last_error().decode()))


def call_host(name, *args):
    """Invoke a stream-less entry point (allocation / IPC plumbing); raises on error."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise NativeError('{} failed ({}): {}'.format(name, rc, lib.pygb_last_error().decode()))


_ws_cache = {}


def workspace(name, *args):
    """Workspace size query (pure functions of their arguments: memoised)."""
    key = (name,) + args
    hit = _ws_cache.get(key)
    if hit is None:
        hit = int(getattr(load(), name)(*args))
        if len(_ws_cache) > 4096:
            _ws_cache.clear()
        _ws_cache[key] = hit
    return hit


def dtype_code(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise NativeError('unsupported feature dtype {} (float32 or bfloat16)'.format(t.dtype))
