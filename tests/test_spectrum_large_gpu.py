"""GPU: the image-sized spectrum stage -- row-side Gram (K < D) and the cooperative whole-GPU eigensolver (n > 1024) --
against torch.linalg.svdvals in fp64 of the same centred matrix (the reference's dim_reduction.py:190-199 in exact
arithmetic).  Tolerance 1e-3 relative per value as in BASELINE.json; measured errors are ~1e-6.  When K <= D the
centred matrix has rank K-1: its last singular value is structurally zero (the reference returns rounding noise
there), so it is compared absolutely."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _matrix(K, D, rank, seed, decay=1e-3):
    g = torch.Generator().manual_seed(seed)
    r = min(rank, K, D)
    U = torch.randn(K, r, generator=g, dtype=torch.float64)
    V = torch.randn(r, D, generator=g, dtype=torch.float64)
    s = torch.logspace(0, np.log10(decay), r, dtype=torch.float64)
    X = (U * s) @ V / np.sqrt(K) + 3.0 + 1e-4 * torch.randn(K, D, generator=g, dtype=torch.float64)
    return X.float()


@pytest.mark.parametrize("K,D,rank", [(96, 300, 40), (257, 640, 300), (700, 701, 50),      # row side, one-CTA solver
                                      (2000, 1500, 200), (1300, 3000, 2000), (1100, 1100, 64), (1400, 1201, 90)])  # whole-GPU solver
def test_spectrum_from_rows_all_shapes(K, D, rank):
    from iddiff_b200 import engine as E
    from oracle import id_oracle as O
    A = 2
    X = torch.stack([_matrix(K, D, rank, 10 + a) for a in range(A)])
    res = E.spectrum_from_rows(X.cuda(), K, D, -1.0 / 0.01)
    torch.cuda.synchronize()
    n = min(K, D)
    assert tuple(res.singular_values.shape) == (A, n)
    for a in range(A):
        Xc = X[a].double()
        Xc = (Xc - Xc.mean(0, keepdim=True)) * (-1.0 / 0.01)
        want = torch.linalg.svdvals(Xc).numpy()
        got = res.singular_values[a].cpu().numpy().astype(np.float64)
        m = n - 1 if K <= D else n
        # Gram-side floor: eigenvalues of G carry eps64 * s_max^2, i.e. sqrt(eps64) * s_max ~ 1e-8 s_max on a singular value
        np.testing.assert_allclose(got[:m], want[:m], rtol=1e-3, atol=1e-7 * want[0])
        if m < n:
            assert abs(got[-1]) < 1e-4 * want[0]
        assert np.all(np.diff(got) <= 0)
        if K > D:
            assert int(res.dims[a]) == O.id_from_spectrum(got.tolist())


def test_spectrum_large_matches_small_solver_on_same_gram():
    """n = 900 through both solvers (the one-CTA kernel and the cooperative one) -- same Gram matrix, same spectrum."""
    from iddiff_b200 import engine as E
    X = _matrix(1200, 900, 300, 3)[None].cuda()
    small = E.spectrum_from_rows(X, 1200, 900, 1.0).singular_values.cpu().numpy()
    old = E.LARGE_N
    try:
        E.LARGE_N = 256
        large = E.spectrum_from_rows(X, 1200, 900, 1.0).singular_values.cpu().numpy()
    finally:
        E.LARGE_N = old
    np.testing.assert_allclose(large, small, rtol=2e-5, atol=1e-6 * small[0, 0])


@pytest.mark.parametrize("n", [1201, 1536, 2050])
def test_tridiagonalisation_variants_agree(n):
    """The three whole-GPU tridiagonalisation kernels (unfused two-pass, fused, fused with two column classes -- the one
    n = 12288 runs) on the same fp64 Gram matrix, through the C-ABI, against fp64 eigvalsh of that matrix."""
    from iddiff_b200 import _lib
    g = torch.Generator().manual_seed(n)
    B = torch.randn(n, n + 40, generator=g, dtype=torch.float64) * torch.logspace(0, -3, n, dtype=torch.float64)[:, None]
    G = (B @ B.T).cuda()
    want = torch.linalg.eigvalsh(G.cpu()).clamp_min(0).sqrt().flip(0).numpy()
    lib = _lib.load()
    ws = torch.empty(lib.idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    out = {}
    for variant in (-1, 1, 2) + ((3,) if n % 2 == 0 else ()):     # 3: the panel kernel with deferred updates (even n)
        A = G.clone()
        sv = torch.empty(n, dtype=torch.float32, device="cuda")
        _lib.call("idb_spectrum_large_variant", A.data_ptr(), n, n, sv.data_ptr(), ws.data_ptr(), variant, st)
        torch.cuda.synchronize()
        out[variant] = sv.cpu().numpy().astype(np.float64)
        np.testing.assert_allclose(out[variant], want, rtol=1e-5, atol=1e-7 * want[0])
    np.testing.assert_allclose(out[2], out[1], rtol=1e-6, atol=1e-8 * want[0])
    np.testing.assert_allclose(out[-1], out[1], rtol=1e-6, atol=1e-8 * want[0])
    if 3 in out:
        np.testing.assert_allclose(out[3], out[1], rtol=1e-6, atol=1e-8 * want[0])


@pytest.mark.parametrize("n", [4096, 12288])
def test_spectrum_large_at_baseline_sizes(n):
    """The sizes BASELINE configs 4 / 5 produce: n = 4096 (K = 4096 rows of a 64x64x3 image, one column class) and
    n = 12288 (the reference's own K = 16768: two column classes, the path that only runs above n = 9000), through the
    production entry point idb_spectrum_large, against fp64 eigenvalues of the same Gram matrix.  The checker is
    cuSOLVER's fp64 syevd through torch (LAPACK on the host would take minutes at n = 12288)."""
    from iddiff_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(n)
    B = torch.randn(n, n + 64, generator=g, device="cuda", dtype=torch.float64)
    B *= torch.logspace(0, -3, n, dtype=torch.float64, device="cuda")[:, None]
    G = B @ B.T
    del B
    want = torch.linalg.eigvalsh(G).clamp_min(0).sqrt().flip(0).cpu().numpy()
    lib = _lib.load()
    ws = torch.empty(lib.idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for variant in (0, 3, 2 if n > 9000 else 1):       # production choice, panel kernel, fused kernel
        A = G.clone()
        sv = torch.empty(n, dtype=torch.float32, device="cuda")
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.call("idb_spectrum_large_variant", A.data_ptr(), n, n, sv.data_ptr(), ws.data_ptr(), variant, st)
        b.record()
        torch.cuda.synchronize()
        got = sv.cpu().numpy().astype(np.float64)
        print(f"n = {n}, variant {variant}: {a.elapsed_time(b):.1f} ms, max rel err {np.abs(got / np.maximum(want, 1e-7 * want[0]) - 1).max():.2e}")
        assert np.all(np.diff(got) <= 0)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-7 * want[0])
        del A
    del G, ws
    torch.cuda.empty_cache()


@pytest.mark.parametrize("n", [3586, 5002, 9002])
def test_panel_kernel_at_ragged_sizes(n):
    """The production entry point takes the panel kernel for every even n >= 3584: sizes that are not multiples of the
    64-column chunk, of the 128 / 256-column tiles or of the CTA count (ragged last chunk, ragged row ownership), and one
    above the fused kernel's one-class limit, against cuSOLVER fp64 eigenvalues; n_out < n returns the leading values.
    Rank-deficient on purpose (K = n - 40 rows): the trailing 40 singular values are structural zeros."""
    from iddiff_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(n)
    B = torch.randn(n, n - 40, generator=g, device="cuda", dtype=torch.float64)
    B *= torch.logspace(0, -2, n, dtype=torch.float64, device="cuda")[:, None]
    G = B @ B.T
    del B
    want = torch.linalg.eigvalsh(G).clamp_min(0).sqrt().flip(0).cpu().numpy()
    lib = _lib.load()
    ws = torch.empty(lib.idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    n_out = n - 41
    out = {}
    for variant in (0, 3):
        A = G.clone()
        sv = torch.full((n_out,), -1.0, dtype=torch.float32, device="cuda")
        _lib.call("idb_spectrum_large_variant", A.data_ptr(), n, n_out, sv.data_ptr(), ws.data_ptr(), variant, st)
        torch.cuda.synchronize()
        out[variant] = sv.cpu().numpy().astype(np.float64)
        assert np.all(np.diff(out[variant]) <= 0)
        np.testing.assert_allclose(out[variant], want[:n_out], rtol=1e-5, atol=1e-7 * want[0])
    assert np.array_equal(out[0], out[3])               # the automatic choice IS the panel kernel at these sizes
    del G, ws
    torch.cuda.empty_cache()
