"""GPU: K5 (tcgen05 X·W) against an fp64 CPU matmul, through the C ABI."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(1000, 100, 256), (777, 256, 256), (300, 256, 47), (130, 500, 512), (5000, 604, 64),
          (129, 32, 16), (1, 8, 3), (128, 36, 300), (4097, 64, 8), (256, 1000, 7)]


def _rel(a, b):
    return (a.double().cpu() - b).abs().max().item() / max(b.abs().max().item(), 1e-30)


@pytest.mark.parametrize('M,K,N', SHAPES)
def test_gemm_3xtf32_matches_fp64(dev, M, K, N):
    from torch_geometric import ops
    g = torch.Generator().manual_seed(M + K + N)
    x = torch.randn(M, K, generator=g)
    w = torch.randn(K, N, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    ref = x.double() @ w.double() + b.double()
    assert ops._gemm_eligible(x.to(dev), w.to(dev))
    got = ops.gemm_xw_raw(x.to(dev), w.to(dev), b.to(dev), passes=3)
    assert got.shape == (M, N)
    # 3xTF32 drops the lo*lo term (2^-22) and the tensor core's fp32 accumulator truncates:
    # 1e-6 .. 5e-6 of the largest entry for K = 100 .. 600, against ~5e-7 for the library's
    # fp32 FMA chain.  The bar is north_star's fp32 tolerance, 1e-5 relative.
    assert _rel(got, ref) <= 1e-5, _rel(got, ref)
    got_nb = ops.gemm_xw_raw(x.to(dev), w.to(dev), None, passes=3)
    assert _rel(got_nb, x.double() @ w.double()) <= 1e-5


@pytest.mark.parametrize('M,K,N', [(1000, 104, 256), (300, 256, 47)])
def test_gemm_fused_relu_and_padded_output(dev, M, K, N):
    """relu flag == torch.relu of the unfused result, bit for bit; columns of a wider
    output matrix beyond N are left untouched."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(3)
    x = torch.randn(M, K, generator=g).to(dev)
    w = (torch.randn(K, N, generator=g) / K ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    plain = ops.gemm_xw_raw(x, w, b)
    out = torch.full((M, N + 9), -7.0, device=dev)
    fused = ops.gemm_xw_raw(x, w, b, out=out, relu=True)
    assert torch.equal(fused, torch.relu(plain))
    assert float(out[:, N:].max()) == -7.0 and float(out[:, N:].min()) == -7.0


@pytest.mark.parametrize('M,S,Kc,N', [(1000, 4, 32, 256), (777, 2, 64, 48), (129, 8, 32, 16)])
def test_gemm_slice_major_operand(dev, M, S, Kc, N):
    """A stored as S planes of [rows, Kc] columns (3-D TMA map) == the row-major product,
    bit for bit; rows of the planes beyond M are never read into the result."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(S * Kc)
    x = torch.randn(M, S * Kc, generator=g).to(dev)
    w = (torch.randn(S * Kc, N, generator=g) / (S * Kc) ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    planes = torch.full((S, M + 37, Kc), float('nan'), device=dev)
    planes[:, :M] = x.view(M, S, Kc).permute(1, 0, 2)
    assert ops._gemm_eligible(planes, w)
    ref = ops.gemm_xw_raw(x, w, b, relu=True)
    got = ops.gemm_xw_raw(planes, w, b, relu=True, rows=M)
    assert got.shape == (M, N) and torch.equal(got, ref)


@pytest.mark.parametrize('M,K,N', [(1000, 100, 256), (777, 256, 256), (300, 256, 47),
                                   (130, 500, 512), (4097, 64, 128), (1, 8, 3), (257, 16, 16)])
def test_gemm_tile_shapes_agree(dev, monkeypatch, M, K, N):
    """The 256-row tile (two halves sharing each W stage, 64-byte swizzle, k-step 16) gives
    the same bits as the 128-row tile: same products, same accumulation order per element."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(M + K + N)
    x = torch.randn(M, K, generator=g).to(dev)
    w = (torch.randn(K, N, generator=g) / K ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    # the split accumulator (corrections in their own TMEM columns) is chosen from the TMEM
    # budget, which depends on the tile height: pin it off so that the shapes are comparable
    monkeypatch.setenv('PYGB_GEMM_SPLIT_ACC', '0')
    monkeypatch.setenv('PYGB_GEMM_TILE', '128x32')
    ref = ops.gemm_xw_raw(x, w, b, relu=True)
    ref1 = ops.gemm_xw_raw(x, w, None, passes=1)
    for tile in ('256x16', '128x16'):
        monkeypatch.setenv('PYGB_GEMM_TILE', tile)
        got = ops.gemm_xw_raw(x, w, b, relu=True)
        got1 = ops.gemm_xw_raw(x, w, None, passes=1)
        assert torch.equal(got, ref) and torch.equal(got1, ref1), tile
    # the default 3xTF32 kernel (Ah / Al in tensor memory, BN <= 128, several n-tiles per row
    # block) runs the same products in the same order per output element
    monkeypatch.delenv('PYGB_GEMM_TILE')
    assert torch.equal(ops.gemm_xw_raw(x, w, b, relu=True), ref), 'A-in-TMEM kernel'
    monkeypatch.setenv('PYGB_GEMM_TS', '0')
    assert torch.equal(ops.gemm_xw_raw(x, w, b, relu=True), ref), 'PYGB_GEMM_TS=0'
    monkeypatch.delenv('PYGB_GEMM_TS')
    assert _rel(ops.gemm_xw_raw(x, w, b), x.double().cpu() @ w.double().cpu() + b.double().cpu()) <= 1e-5


@pytest.mark.parametrize('M,K,N', [(3000, 602, 64), (2708, 1433, 16), (2000, 1000, 128)])
def test_gemm_split_accumulator_is_more_accurate(dev, monkeypatch, M, K, N):
    """Long-K transforms (cfg3: K = 602, cfg1: K = 1433): keeping the 3xTF32 correction terms
    in their own TMEM accumulator cuts the truncation bias of the tensor core's fp32
    accumulation chain; error against fp64 must not get worse and must stay below 5e-6."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(K)
    x = torch.randn(M, K, generator=g)
    w = torch.randn(K, N, generator=g) / K ** 0.5
    ref = x.double() @ w.double()
    xd = ops._aligned_rows(x.to(dev))
    monkeypatch.setenv('PYGB_GEMM_SPLIT_ACC', '0')
    e_off = _rel(ops.gemm_xw_raw(xd, w.to(dev)), ref)
    monkeypatch.setenv('PYGB_GEMM_SPLIT_ACC', '1')
    e_on = _rel(ops.gemm_xw_raw(xd, w.to(dev)), ref)
    print('K={} split off {:.3e} on {:.3e}'.format(K, e_off, e_on))
    assert e_on <= e_off * 1.05 and e_on <= 5e-6, (e_off, e_on)


def test_gemm_tile_256_sliced_and_routed(dev, monkeypatch):
    from torch_geometric import ops
    monkeypatch.setenv('PYGB_GEMM_TILE', '256x16')
    g = torch.Generator().manual_seed(5)
    M, S, Kc, N, Fc = 900, 4, 32, 256, 64
    x = torch.randn(M, S * Kc, generator=g).to(dev)
    w = (torch.randn(S * Kc, N, generator=g) / (S * Kc) ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    ref = ops.gemm_xw_raw(x, w, b, relu=True)
    planes = torch.zeros((S, M + 11, Kc), device=dev)
    planes[:, :M] = x.view(M, S, Kc).permute(1, 0, 2)
    assert torch.equal(ops.gemm_xw_raw(planes, w, b, relu=True, rows=M), ref)
    outs = [torch.full((M + 2, Fc), -7.0, device=dev) for _ in range(N // Fc)]
    tab = torch.tensor([o.data_ptr() for o in outs], dtype=torch.int64).to(dev)
    ops.gemm_xw_routed_raw(x, w, b, Fc, tab, 2, Fc, relu=True)
    for q, o in enumerate(outs):
        assert torch.equal(o[2:], ref[:, q * Fc:(q + 1) * Fc]) and float(o[:2].max()) == -7.0


@pytest.mark.parametrize('M', [900, 130])
def test_gemm_default_kernel_sliced_and_routed(dev, monkeypatch, M):
    """The default 3-pass kernel (A in tensor memory; CTA pairs when M > 256) with a slice-major
    A and a routed epilogue (the multi-GPU transform's local case) == the plain call, bit for bit;
    and the TMA-store epilogue == the transposing one."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(6)
    S, Kc, N, Fc = 4, 32, 256, 64
    x = torch.randn(M, S * Kc, generator=g).to(dev)
    w = (torch.randn(S * Kc, N, generator=g) / (S * Kc) ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    ref = ops.gemm_xw_raw(x, w, b, relu=True)
    monkeypatch.setenv('PYGB_GEMM_TMA_STORE', '0')
    assert torch.equal(ops.gemm_xw_raw(x, w, b, relu=True), ref)
    monkeypatch.delenv('PYGB_GEMM_TMA_STORE')
    monkeypatch.setenv('PYGB_GEMM_CG', '1')
    assert torch.equal(ops.gemm_xw_raw(x, w, b, relu=True), ref)
    monkeypatch.delenv('PYGB_GEMM_CG')
    planes = torch.zeros((S, M + 11, Kc), device=dev)
    planes[:, :M] = x.view(M, S, Kc).permute(1, 0, 2)
    assert torch.equal(ops.gemm_xw_raw(planes, w, b, relu=True, rows=M), ref)
    outs = [torch.full((M + 2, Fc), -7.0, device=dev) for _ in range(N // Fc)]
    tab = torch.tensor([o.data_ptr() for o in outs], dtype=torch.int64).to(dev)
    ops.gemm_xw_routed_raw(x, w, b, Fc, tab, 2, Fc, relu=True)
    for q, o in enumerate(outs):
        assert torch.equal(o[2:], ref[:, q * Fc:(q + 1) * Fc]) and float(o[:2].max()) == -7.0


@pytest.mark.parametrize('M,K,N', SHAPES[:5])
def test_gemm_tf32_single_pass(dev, M, K, N):
    from torch_geometric import ops
    g = torch.Generator().manual_seed(7)
    x = torch.randn(M, K, generator=g)
    w = torch.randn(K, N, generator=g) / K ** 0.5
    got = ops.gemm_xw_raw(x.to(dev), w.to(dev), None, passes=1)
    assert _rel(got, x.double() @ w.double()) <= 3e-3   # 10-bit mantissa inputs


def test_linear_autograd_and_fallbacks(dev):
    from torch_geometric import ops
    g = torch.Generator().manual_seed(3)
    x = torch.randn(500, 64, generator=g).to(dev).requires_grad_(True)
    w = torch.randn(64, 47, generator=g).to(dev).requires_grad_(True)
    b = torch.randn(47, generator=g).to(dev).requires_grad_(True)
    out = ops.linear(x, w, b)
    ref = x.detach().double() @ w.detach().double() + b.detach().double()
    assert _rel(out, ref.cpu()) <= 1e-5
    out.square().sum().backward()
    xr, wr, br = (t.detach().clone().requires_grad_(True) for t in (x, w, b))
    (xr @ wr + br).square().sum().backward()
    for a, r in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert _rel(a, r.double().cpu()) <= 1e-4
    # rows that are not 16-byte multiples (K = 7) and bf16 go to the library GEMM
    x7 = torch.randn(50, 7, generator=g).to(dev)
    w7 = torch.randn(7, 5, generator=g).to(dev)
    assert not ops._gemm_eligible(x7, w7)
    assert _rel(ops.linear(x7, w7), (x7.double() @ w7.double()).cpu()) <= 1e-5
    xb = x.detach().bfloat16()
    assert ops.linear(xb, w.detach().bfloat16()).dtype == torch.bfloat16


XTG_SHAPES = [(5000, 100, 256), (70001, 256, 256), (33000, 256, 48), (1024, 16, 16), (9999, 128, 4),
              (40000, 132, 200), (2500, 4, 256)]


@pytest.mark.parametrize('R,Kx,Ng', XTG_SHAPES)
def test_gemm_xtg_matches_fp64(dev, R, Kx, Ng):
    """K5t: dW = X^T G (split-K over the rows, MN-major tcgen05 operands) against fp64,
    run-to-run bit equality, and padded row strides."""
    from torch_geometric import ops
    gen = torch.Generator().manual_seed(R + Kx + Ng)
    x = torch.randn(R, Kx, generator=gen)
    g = torch.randn(R, Ng, generator=gen)
    ref = x.double().t() @ g.double()
    xd, gd = x.to(dev), g.to(dev)
    assert ops._xtg_eligible(xd, gd)
    got = ops.gemm_xtg_raw(xd, gd, passes=3)
    assert got.shape == (Kx, Ng)
    assert _rel(got, ref) <= 1e-5, _rel(got, ref)
    assert torch.equal(got, ops.gemm_xtg_raw(xd, gd, passes=3)), 'not deterministic'
    # plain TF32: 10-bit mantissas
    assert _rel(ops.gemm_xtg_raw(xd, gd, passes=1), ref) <= 3e-3
    # views with padded rows (what the padded GEMM / aggregation outputs look like)
    xp = torch.zeros(R, Kx + 4, device=dev)
    xp[:, :Kx] = xd
    gp = torch.zeros(R, Ng + 8, device=dev)
    gp[:, :Ng] = gd
    assert torch.equal(ops.gemm_xtg_raw(xp[:, :Kx], gp[:, :Ng], passes=3), got)


def test_linear_weight_grad_on_tensor_cores(dev):
    """_Linear.backward routes dW through K5t when eligible and through the library otherwise."""
    from torch_geometric import ops, _native as nat
    gen = torch.Generator().manual_seed(5)
    x = torch.randn(3000, 100, generator=gen).to(dev).requires_grad_(True)
    w = (torch.randn(100, 256, generator=gen) / 10).to(dev).requires_grad_(True)
    nat.event_log = []
    ops.linear(x, w).square().sum().backward()
    names = [e[0] for e in nat.event_log]
    nat.event_log = None
    assert 'pygb_gemm_xtg' in names
    xr, wr = (t.detach().double().cpu().requires_grad_(True) for t in (x, w))
    (xr @ wr).square().sum().backward()
    assert _rel(w.grad, wr.grad) <= 1e-5
    assert _rel(x.grad, xr.grad) <= 1e-5
    # too few rows: library GEMM; wider than the 256-column TMEM partial: tiled (round 2)
    assert not ops._xtg_eligible(x[:100], torch.zeros(100, 256, device=dev))
    assert ops._xtg_eligible(torch.zeros(2000, 300, device=dev), torch.zeros(2000, 8, device=dev))


def test_gemm_xtg_full_size_accuracy(dev):
    """cfg5's layer-2 weight gradient (2,449,029 rows, 256 x 256): the chunked accumulation keeps
    the truncating tensor-core accumulator inside the fp32 bar against an fp64 product."""
    if torch.cuda.get_device_properties(dev).total_memory < 60 * 2 ** 30:
        pytest.skip('needs a large-memory GPU')
    from torch_geometric import ops
    gen = torch.Generator(device=dev).manual_seed(17)
    R = 2449029
    x = torch.randn(R, 256, generator=gen, device=dev)
    g = torch.randn(R, 256, generator=gen, device=dev)
    got = ops.gemm_xtg_raw(x, g, passes=3)
    ref = torch.zeros(256, 256, dtype=torch.float64, device=dev)
    for r0 in range(0, R, 1 << 19):                      # fp64 in slabs: bounded memory
        ref += x[r0:r0 + (1 << 19)].double().t() @ g[r0:r0 + (1 << 19)].double()
    err = (got.double() - ref).abs().max().item() / ref.abs().max().item()
    assert err <= 1e-5, err
    lib = torch.matmul(x.t(), g)
    err_lib = (lib.double() - ref).abs().max().item() / ref.abs().max().item()
    print('xtg rel err %.2e, library fp32 GEMM %.2e' % (err, err_lib))


BF16_SHAPES = [(1000, 100, 256), (777, 256, 256), (300, 256, 47), (5000, 602, 64), (129, 32, 16),
               (1, 8, 3), (4097, 64, 8), (256, 1433, 7), (2000, 96, 300), (333, 72, 40)]


@pytest.mark.parametrize('M,K,N', BF16_SHAPES)
def test_gemm_bf16_matches_fp64_of_the_bf16_inputs(dev, M, K, N):
    """kind::f16 kernel (bf16 had no tensor-core transform).  Reference: fp64
    product of the SAME bf16-rounded operands — what is left is fp32 accumulation order and
    the bf16 rounding of the result (2^-9 relative), far inside north_star's 1e-2 bf16 bar."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(M + K + N)
    x = torch.randn(M, K, generator=g).bfloat16()
    w = (torch.randn(K, N, generator=g) / K ** 0.5).bfloat16()
    b = torch.randn(N, generator=g)
    xd = ops._aligned_rows(x.to(dev))
    assert ops._gemm_eligible(xd, w.to(dev))
    got = ops.gemm_xw_raw(xd, w.to(dev), b.to(dev))
    assert got.dtype == torch.bfloat16 and got.shape == (M, N)
    ref = x.double() @ w.double() + b.double()
    err = (got.double().cpu() - ref).abs()
    bound = 2.0 ** -8 * ref.abs() + 1e-3 * ref.abs().max()
    assert bool((err <= bound).all()), float((err / bound).max())
    fused = ops.gemm_xw_raw(xd, w.to(dev), b.to(dev), relu=True)
    assert torch.equal(fused, torch.relu(got))


@pytest.mark.parametrize('K', [602, 1433, 7, 33, 101])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_linear_unaligned_k_runs_on_tcgen05(dev, K, dtype):
    """BASELINE cfg3 (K = 602) and cfg1 (K = 1433): rows that are not 16-byte multiples reach
    the tensor-core kernel through one memoised padded copy — asserted on the native event
    log, under PYGB_STRICT_GEMM semantics (no library GEMM), forward and backward."""
    from torch_geometric import _native as nat
    from torch_geometric import ops
    M, N = 1500, 64
    g = torch.Generator().manual_seed(K)
    x = torch.randn(M, K, generator=g).to(dtype)
    w = (torch.randn(K, N, generator=g) / K ** 0.5).to(dtype)
    b = torch.randn(N, generator=g).to(dtype)
    xd = x.to(dev)
    wd = w.to(dev).requires_grad_(True)
    bd = b.to(dev).requires_grad_(True)
    old, ops.STRICT_GEMM = ops.STRICT_GEMM, True
    nat.event_log = []
    try:
        out = ops.linear(xd, wd, bd)
        names = [e[0] for e in nat.event_log]
        again = ops.linear(xd, wd, bd)
    finally:
        ops.STRICT_GEMM = old
        log, nat.event_log = nat.event_log, None
    want = 'pygb_gemm_xw_bf16' if dtype == torch.bfloat16 else 'pygb_gemm_xw'
    assert names == [want], names
    assert torch.equal(out, again)
    assert out.shape == (M, N)
    ref = x.double() @ w.double() + b.double()
    tol = 1e-5 if dtype == torch.float32 else 2.0 ** -7
    err = (out.double().cpu() - ref).abs()
    assert bool((err <= tol * ref.abs() + tol * ref.abs().max()).all())
    # the padded copy of a constant input is made once
    if K % (16 // x.element_size()):
        assert ops._aligned_rows(xd).data_ptr() == ops._aligned_rows(xd).data_ptr()
    gout = torch.randn(M, N, generator=g).to(dtype)
    out.backward(gout.to(dev))
    ref_gw = x.double().t() @ gout.double()
    err = (wd.grad.double().cpu() - ref_gw).abs()
    gtol = 1e-5 if dtype == torch.float32 else 2e-2
    assert bool((err <= gtol * ref_gw.abs() + gtol * ref_gw.abs().max()).all())
    assert wd.grad.shape == (K, N) and bd.grad.shape == (N,)


def test_linear_unaligned_input_that_requires_grad(dev):
    from torch_geometric import ops
    g = torch.Generator().manual_seed(5)
    x = torch.randn(900, 602, generator=g)
    w = torch.randn(602, 40, generator=g) / 25
    xd = x.to(dev).requires_grad_(True)
    wd = w.to(dev).requires_grad_(True)
    out = ops.linear(xd, wd)
    gout = torch.randn(900, 40, generator=g)
    out.backward(gout.to(dev))
    xr, wr = x.double().requires_grad_(True), w.double().requires_grad_(True)
    (xr @ wr).backward(gout.double())
    for got, ref, what in ((xd.grad, xr.grad, 'dx'), (wd.grad, wr.grad, 'dw')):
        assert got.shape == ref.shape
        err = (got.double().cpu() - ref).abs()
        assert bool((err <= 1e-5 * ref.abs() + 1e-5 * ref.abs().max()).all()), what


def test_gemm_xtg_tiles_wide_outputs(dev):
    """dW for K > 256 (cfg3: 602 x 64) stays on the split-K tcgen05 kernel, tiled over column
    blocks of x (views, nothing copied)."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(9)
    x = torch.randn(3000, 604, generator=g)
    gr = torch.randn(3000, 300, generator=g)
    got = ops.gemm_xtg_raw(x.to(dev), gr.to(dev))
    ref = x.double().t() @ gr.double()
    assert got.shape == (604, 300)
    assert (got.double().cpu() - ref).abs().max().item() / ref.abs().max().item() <= 1e-5
