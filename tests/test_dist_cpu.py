"""CPU, gloo, world_size 2: the anchor sharding and the single all-gather of the path."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from iddiff_b200.dim_reduction import all_gather_spectra, shard_range
    lo, hi = shard_range(n_total, rank, world)
    # a rank's "spectra": row a holds the global anchor id, so order and ownership are checkable
    sv = torch.arange(lo, hi, dtype=torch.float32)[:, None] * torch.ones(1, 5)
    dims = torch.arange(lo, hi, dtype=torch.int32) + 100
    g_sv, g_dims = all_gather_spectra(sv, dims, n_total)
    q.put((rank, lo, hi, g_sv[:, 0].tolist(), g_dims.tolist()))
    dist.destroy_process_group()


def test_shard_and_all_gather_world2():
    for n_total in (7, 8):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = sorted(q.get(timeout=120) for _ in procs)
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        covered = []
        for rank, lo, hi, sv0, dims in res:
            covered += list(range(lo, hi))
            assert sv0 == [float(i) for i in range(n_total)]          # global order, nothing lost or duplicated
            assert dims == [100 + i for i in range(n_total)]
        assert covered == list(range(n_total))                          # shards tile the anchors exactly


def test_shard_range_edges():
    from iddiff_b200.dim_reduction import shard_range
    assert [shard_range(4096, r, 8) for r in (0, 7)] == [(0, 512), (3584, 4096)]
    assert [shard_range(3, r, 4) for r in range(4)] == [(0, 1), (1, 2), (2, 3), (3, 3)]
    assert shard_range(0, 0, 2) == (0, 0)


def _rows_worker(rank, world, port, K, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from iddiff_b200.dim_reduction import all_gather_rows, shard_range
    lo, hi = shard_range(K, rank, world)
    A, D = 2, 3
    # S[a, k, d] = 1000 a + 10 k + d for the rows this rank owns
    k = torch.arange(lo, hi, dtype=torch.float32)
    S_loc = (1000 * torch.arange(A, dtype=torch.float32)[:, None, None] + 10 * k[None, :, None]
             + torch.arange(D, dtype=torch.float32)[None, None, :])
    S = all_gather_rows(S_loc, K)
    q.put((rank, S.tolist()))
    dist.destroy_process_group()


def test_row_sharding_of_one_anchor_world2():
    """SURVEY 8e second row: K perturbation rows of one anchor sharded over ranks, one all-gather, global row order."""
    for K in (7, 8):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_rows_worker, args=(r, 2, port, K, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = sorted(q.get(timeout=120) for _ in procs)
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        want = (1000 * torch.arange(2.)[:, None, None] + 10 * torch.arange(float(K))[None, :, None]
                + torch.arange(3.)[None, None, :]).tolist()
        for _, S in res:
            assert S == want


class _HostGramOps:
    """fp64 torch stand-in for the three device Gram pieces (collective bookkeeping test only)."""

    @staticmethod
    def column_sums(X):
        return X.double().sum(0)

    @staticmethod
    def gram_about(X, center, scale):
        Y = (X.double() - center[None]) * scale
        return Y.T @ Y

    @staticmethod
    def row_gram(X, scale):
        Y = (X.double() - X.double().mean(0, keepdim=True)) * scale
        return Y @ Y.T


def _gram_worker(rank, world, port, K, D, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from iddiff_b200.dim_reduction import shard_range, sharded_centred_gram, rows_to_column_panels
    S = torch.randn(K, D, generator=torch.Generator().manual_seed(K * 1000 + D)) + 3.0
    lo, hi = shard_range(K, rank, world)
    panel = rows_to_column_panels(S[lo:hi].contiguous(), K)
    clo, chi = shard_range(D, rank, world)
    ok_panel = torch.equal(panel, S[:, clo:chi])
    G = sharded_centred_gram(S[lo:hi].contiguous(), K, -2.0, ops=_HostGramOps)
    q.put((rank, ok_panel, G))
    dist.destroy_process_group()


def test_sharded_partial_gram_world2():
    """SURVEY 8e as specified: partial Grams + ONE all-reduce; both sides (K >= D: D x D about the all-reduced mean;
    K < D: all-to-all to column panels, K x K), ragged shards included, against the centred Gram of the whole matrix."""
    for K, D in ((9, 4), (8, 8), (5, 11), (6, 7)):
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_gram_worker, args=(r, 2, port, K, D, q)) for r in range(2)]
        for p in procs:
            p.start()
        res = sorted((q.get(timeout=120) for _ in procs), key=lambda t: t[0])
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        S = (torch.randn(K, D, generator=torch.Generator().manual_seed(K * 1000 + D)) + 3.0).double()
        Y = (S - S.mean(0, keepdim=True)) * -2.0
        want = Y.T @ Y if K >= D else Y @ Y.T
        for rank, ok_panel, G in res:
            assert ok_panel
            assert G.shape == want.shape
            torch.testing.assert_close(G, want, rtol=1e-12, atol=1e-12)
