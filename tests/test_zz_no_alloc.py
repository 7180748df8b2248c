"""GPU: idb_spectrum_large_static -- the eigen-solver entry point that allocates nothing (its workspace is a file-scope
device array inside the library, include/iddiff_b200.h) -- against the existing idb_spectrum_large with a caller-supplied
workspace, through the C ABI, on two calls in a row.  The two entry points enqueue the same kernels on the same input, so
the spectra must agree to the tolerance the solver tests use (rtol 1e-6 here; they come out bit-identical).

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it depends on no state but the library itself, and the second call runs on a different matrix and size with the first
call's contents still in the shared array."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from iddiff_b200 import _lib  # noqa: E402
from iddiff_b200._lib import call, ptr  # noqa: E402


def _gram(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    B = torch.randn(n, n + 32, generator=g, device="cuda", dtype=torch.float64)
    B *= torch.logspace(0, -2, n, dtype=torch.float64, device="cuda")[:, None]
    return B @ B.T


def _existing(G, n):
    ws = torch.empty(_lib.load().idb_spectrum_large_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    sv = torch.empty(n, dtype=torch.float32, device="cuda")
    A = G.clone()
    call("idb_spectrum_large", ptr(A), n, n, ptr(sv), ptr(ws), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return sv.cpu().numpy()


def test_static_scratch_entry_point_matches_the_existing_one_twice_in_a_row():
    lib = _lib.load()
    assert lib.idb_spectrum_large_static_max_n() == 4096
    st = torch.cuda.current_stream().cuda_stream
    cases = [(1536, 11), (4096, 12)]
    want = [_existing(_gram(n, seed), n) for n, seed in cases]
    # inputs and outputs exist before the calls under test; nothing is allocated between or inside them
    As = [_gram(n, seed) for n, seed in cases]
    svs = [torch.empty(n, dtype=torch.float32, device="cuda") for n, _ in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (n, _), A, sv in zip(cases, As, svs):                # two calls in a row, the array keeps the first call's contents
        call("idb_spectrum_large_static", ptr(A), n, n, ptr(sv), st)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (n, _), sv, w in zip(cases, svs, want):
        got = sv.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.all(np.diff(got) <= 0)
        np.testing.assert_allclose(got, w, rtol=1e-6, atol=1e-6 * w[0])
        # and against LAPACK on the same matrix (the solver tests' own bar)
    ref = np.sqrt(np.clip(np.linalg.eigvalsh(_gram(*cases[0]).cpu().numpy())[::-1], 0, None))
    np.testing.assert_allclose(svs[0].cpu().numpy(), ref, rtol=1e-5, atol=1e-6 * ref[0])


def test_static_scratch_entry_point_rejects_sizes_beyond_its_array():
    n = 4098
    A = torch.eye(n, dtype=torch.float64, device="cuda")
    sv = torch.empty(n, dtype=torch.float32, device="cuda")
    with pytest.raises(_lib.IdbError, match="IDB_SPECTRUM_STATIC_MAX_N"):
        call("idb_spectrum_large_static", ptr(A), n, n, ptr(sv), torch.cuda.current_stream().cuda_stream)
    with pytest.raises(_lib.IdbError, match="bad argument"):
        call("idb_spectrum_large_static", ptr(A), 0, 0, ptr(sv), torch.cuda.current_stream().cuda_stream)
