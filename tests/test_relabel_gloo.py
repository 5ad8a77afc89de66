"""World-size-2 gloo test (CPU) of the sharded relabel's partition + exchange logic.  The three
compute stages are replaced by oracle-backed stand-ins (test infrastructure); what is under test
is exorl_b200.relabel's shard bounds, all_gather / all_to_all layouts, sums all-reduce and the
final reward gather."""
import os
import socket
import types

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _OracleStages:
    """CPU stand-ins with the same contracts as the CUDA stages."""

    @staticmethod
    def local_knn(rep, lo, hi, k, mode):
        from oracle import knn_smallest
        d, i = knn_smallest(rep, rep[lo:hi], k)
        return d, (i + lo).to(torch.int32)

    @staticmethod
    def block_knn(rep, lo, hi, k, mode):
        from oracle import knn_smallest
        d, i = knn_smallest(rep[lo:hi], rep, k)
        return d, i.to(torch.int32)

    # "lists" exchange: stand-ins with the CUDA contracts — per shard, every query keeps its k + 2
    # best bank rows as (score = -dist^2, LOCAL index) in two lists of 24 slots (as the 8-warp
    # screen does: even / odd positions), counts in cnt; the rerank merges all shards' lists.
    @staticmethod
    def lists_available(n, nb, dim, k, world):
        return 2 if 2 * world <= 16 else 0

    @staticmethod
    def screen_lists(rep, lo, hi, k):
        from oracle import knn_smallest
        n = rep.size(0)
        kk = min(k + 2, hi - lo)
        d, i = knn_smallest(rep, rep[lo:hi], kk)
        val = torch.full((n, 2, 24), -3.0e38)
        idx = torch.zeros((n, 2, 24), dtype=torch.int32)
        cnt = torch.zeros((n, 2), dtype=torch.int32)
        for h in range(2):
            sel = torch.arange(h, kk, 2)
            val[:, h, :len(sel)] = -(d[:, sel] ** 2)
            idx[:, h, :len(sel)] = i[:, sel].to(torch.int32)
            cnt[:, h] = len(sel)
        sc = torch.tensor([1, 2 + lo, 0, 0], dtype=torch.int32)     # rank-dependent: MAX-reduced
        return val, idx, cnt, torch.zeros(n), torch.zeros(n), sc

    @staticmethod
    def rerank_lists(rep, lo, hi, k, val, idx, cnt, shard_lo, ha, ea, sc):
        g, my_n, lists, cap = val.shape
        assert my_n == hi - lo and len(shard_lo) == g and ha.numel() == my_n
        assert sc.tolist()[0] == 1 and sc.tolist()[1] == 2 + max(shard_lo)   # MAX over ranks
        q = rep[lo:hi]
        out_d = torch.empty(my_n, k)
        out_i = torch.empty(my_n, k, dtype=torch.int32)
        for r in range(my_n):
            cand = []
            for s in range(g):
                for h in range(lists):
                    c = int(cnt[s, r, h])
                    cand += [int(x) + shard_lo[s] for x in idx[s, r, h, :c]]
            cand = torch.tensor(sorted(set(cand)))
            dd = torch.norm(q[r][None, :] - rep[cand], dim=1, p=2)
            key = dd.double() * 1e7 + cand.double() * 1e-3
            pos = key.argsort()[:k]
            out_d[r] = dd[pos]
            out_i[r] = cand[pos].to(torch.int32)
        return out_d, out_i

    @staticmethod
    def merge(d_lists, i_lists, k):
        g, n, _ = d_lists.shape
        d = d_lists.permute(1, 0, 2).reshape(n, g * k)
        i = i_lists.permute(1, 0, 2).reshape(n, g * k)
        key = d.double() * 1e7 + i.double() * 1e-3          # smaller distance, then smaller index
        pos = key.argsort(dim=1)[:, :k]
        return torch.gather(d, 1, pos), torch.gather(i, 1, pos)

    @staticmethod
    def sums(knn_dist, k, avg):
        v = (knn_dist if avg else knn_dist[:, -1]).double().reshape(-1)
        return torch.tensor([v.numel(), v.sum(), (v * v).sum()], dtype=torch.float64)

    @staticmethod
    def finish(knn_dist, k, avg, use_rms, clip, sums, rms):
        r = (knn_dist if avg else knn_dist[:, -1:]).clone()
        if use_rms:
            bs, s, ss = (float(x) for x in sums)
            mean = s / bs
            var = (ss - s * mean) / (bs - 1)
            m, sv, n = float(rms.M), float(rms.S), rms.n
            delta = mean - m
            rms.M = torch.tensor([m + delta * bs / (n + bs)], dtype=torch.float32)
            rms.S = torch.tensor([(sv * n + var * bs + delta * delta * n * bs / (n + bs))
                                  / (n + bs)], dtype=torch.float32)
            rms.n = n + bs          # the stage owns the whole state update (M, S and n)
            r = r / rms.M
        if clip >= 0:
            r = torch.clamp(r - clip, min=0)
        if avg:
            r = r.mean(dim=1, keepdim=True)
        return torch.log(r + 1.0)

    @staticmethod
    def reward(knn_dist, k, avg, use_rms, clip, rms):
        sums = _OracleStages.sums(knn_dist, k, avg)
        return _OracleStages.finish(knn_dist, k, avg, use_rms, clip, sums, rms)


def _worker(rank, world, port, exchange, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exorl_b200.relabel import relabel_apt_reward
        g = torch.Generator().manual_seed(5)
        rep = torch.tanh(torch.randn(n, 16, generator=g))
        args = types.SimpleNamespace(knn_k=4, knn_avg=True, knn_rms=True, knn_clip=0.0)
        rms = types.SimpleNamespace(M=torch.zeros(1), S=torch.ones(1), n=1e-4)
        r = relabel_apt_reward(rep, args, rms=rms, exchange=exchange, stages=_OracleStages)
        q.put((rank, r.numpy(), float(rms.M), float(rms.S), rms.n))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("exchange", ["all_gather", "all_to_all", "lists", "queries", "auto"])
@pytest.mark.parametrize("n", [64, 51])
def test_two_rank_relabel_equals_unsharded(exchange, n):
    from oracle import RmsState, pbe_reward
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, exchange, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(5)
    rep = torch.tanh(torch.randn(n, 16, generator=g))
    st = RmsState()
    want = pbe_reward(rep, rep, st, 4, True, True, 0.0).numpy()
    for rank, r, m, s, cnt in res:
        assert r.shape == (n, 1)
        np.testing.assert_allclose(r, want, rtol=1e-5, atol=1e-7)
        assert m == pytest.approx(float(st.M), rel=1e-5) and s == pytest.approx(float(st.S), rel=1e-4)
        assert cnt == pytest.approx(st.n)


def _host_worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exorl_b200.relabel import reps_from_host
        g = torch.Generator().manual_seed(9)
        host = torch.randn(n, 8, generator=g)
        out = reps_from_host(host, out=torch.full((n, 8), float("nan")))
        q.put((rank, bool(torch.equal(out, host))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [64, 51])
def test_reps_from_host_shards_then_gathers(n):
    """Each rank copies only its rows and the collective fills in the rest (even and ragged)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_host_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res)


class _FakeWatch:
    """Stopwatch whose 'device time' is rows x a per-rank cost: rank 1 is twice as slow."""
    cost = {0: 1.0, 1: 2.0}

    def __init__(self):
        self.rows = 0
        self.t = 0.0

    def start(self):
        pass

    def stop(self):
        from exorl_b200 import relabel
        b = relabel.last_query_block
        self.t = (b["hi"] - b["lo"]) * self.cost[dist.get_rank()]

    def elapsed_ms(self):
        t, self.t = self.t, 0.0
        return t


def _balance_worker(rank, world, port, n, steps, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exorl_b200 import relabel

        class Stages(_OracleStages):
            stopwatch = _FakeWatch

        g = torch.Generator().manual_seed(5)
        rep = torch.tanh(torch.randn(n, 16, generator=g))
        args = types.SimpleNamespace(knn_k=4, knn_avg=True, knn_rms=True, knn_clip=0.0)
        blocks, rewards = [], []
        for _ in range(steps):
            rms = types.SimpleNamespace(M=torch.zeros(1), S=torch.ones(1), n=1e-4)
            r = relabel.relabel_apt_reward(rep, args, rms=rms, exchange="queries", stages=Stages,
                                           balance=True)
            rewards.append(r.numpy())
            blocks.append((relabel.last_query_block["lo"], relabel.last_query_block["hi"]))
        q.put((rank, blocks, rewards))
    finally:
        dist.destroy_process_group()


def test_query_blocks_follow_the_measured_rates_and_results_do_not_change():
    """QueryBalance: rank 1 takes twice as long per row, so its query block shrinks towards one
    third of the rows (cut at multiples of the 256-row quantum); rewards stay those of the
    unsharded computation at every step and identical on both ranks."""
    from oracle import RmsState, pbe_reward
    n, steps = 4096, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_balance_worker, args=(r, 2, port, n, steps, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(5)
    rep = torch.tanh(torch.randn(n, 16, generator=g))
    want = pbe_reward(rep, rep, RmsState(), 4, True, True, 0.0).numpy()
    (_, b0, r0), (_, b1, r1) = res
    assert b0[0] == (0, n // 2) and b1[0] == (n // 2, n)               # first step: equal split
    for s in range(steps):
        assert b0[s][1] == b1[s][0] and b0[s][0] == 0 and b1[s][1] == n    # a partition
        assert b0[s][1] % 256 == 0
        np.testing.assert_allclose(r0[s], want, rtol=1e-5, atol=1e-7)
        np.testing.assert_array_equal(r0[s], r1[s])
    share1 = (b1[-1][1] - b1[-1][0]) / n
    assert abs(share1 - 1 / 3) < 0.07, (b0, b1)


def test_balanced_bounds_quantum_and_fallback():
    from exorl_b200.relabel import balanced_bounds, shard_bounds
    b = balanced_bounds(1_000_000, [1.0, 1.1, 0.9, 1.0], quantum=256, min_rows=12)
    assert b[0][0] == 0 and b[-1][1] == 1_000_000
    assert all(x[1] == y[0] for x, y in zip(b, b[1:]))
    assert all(x[1] % 256 == 0 for x in b[:-1])
    sizes = [hi - lo for lo, hi in b]
    assert sizes[1] > sizes[0] > sizes[2] and abs(sizes[0] - 250_000) < 512
    # a block that would fall below min_rows -> equal split
    assert balanced_bounds(100, [1.0, 1e-6], quantum=256, min_rows=12) == \
        [shard_bounds(100, 2, r) for r in range(2)]


# ---------------------------------------------------------------- symmetric screen over the ranks
class _OracleSymBackend:
    """CPU stand-in of relabel._CudaSymBackend with tiles of 4 rows: direct candidates are kept
    per row, transposed hits are returned as (query row, -dist^2, bank row) for ANY row, exactly
    like the CUDA screen; what is under test is `_sym_queries_knn`'s span assignment, routing of
    the hits to the owner of their row, and the all-to-all layouts."""
    tile = 4

    def __init__(self, rep, k):
        self.rep, self.k, self.n = rep, k, rep.size(0)
        self.direct = {}

    def sample_rows(self):
        return max(2 * self.k, self.n // 4)

    def sample_distances(self, lo, hi, sample):
        d = torch.cdist(self.rep[lo:hi].double(), sample.double())
        return d.topk(self.k, dim=1, largest=False).values[:, -1].float() * 1.0001

    def room(self, expected):
        return True

    def screen(self, d_all, span):
        pair_lo, pair_hi, span_end, mid, extra = span
        n, t = self.n, self.tile
        tiles = (n + t - 1) // t
        full = torch.cdist(self.rep.double(), self.rep.double())
        tg, va, sr = [], [], []
        for u in range(pair_lo, pair_hi):
            length = span_end - u + (extra if u >= mid else 0)
            for j in range(length):
                tt = (u + j) % tiles
                for i in range(u * t, min(n, (u + 1) * t)):
                    for c in range(tt * t, min(n, (tt + 1) * t)):
                        dd = float(full[i, c])
                        if dd <= float(d_all[i]):
                            self.direct.setdefault(i, []).append((dd, c))
                        if tt != u and dd <= float(d_all[c]):
                            tg.append(c); va.append(-dd * dd); sr.append(i)
        return (torch.tensor(tg, dtype=torch.int32), torch.tensor(va, dtype=torch.float32),
                torch.tensor(sr, dtype=torch.int32)), False

    def scatter_rerank(self, r_t, r_v, r_s, lo, hi):
        cand = {i: list(self.direct.get(i, [])) for i in range(lo, hi)}
        for tgt, v, s in zip(r_t.tolist(), r_v.tolist(), r_s.tolist()):
            assert lo <= tgt < hi, "a hit was routed to a rank that does not own its row"
            cand[tgt].append(((-v) ** 0.5, s))
        d = torch.empty(hi - lo, self.k)
        ix = torch.empty(hi - lo, self.k, dtype=torch.int32)
        for i in range(lo, hi):
            rows = sorted(set(c for _, c in cand[i]))
            assert len(rows) == len(cand[i]), "a (row, bank row) pair was covered twice"
            dd = torch.norm(self.rep[i][None, :] - self.rep[rows], dim=1, p=2)
            o = torch.argsort(dd, stable=True)[:self.k]
            d[i - lo], ix[i - lo] = dd[o], torch.tensor(rows, dtype=torch.int32)[o]
        return d, ix, 0, 0

    def repair(self, lo, hi, d_mine, i_mine):
        raise AssertionError("nothing to repair in the stand-in")


class _OracleStagesSym(_OracleStages):
    sym_backend = _OracleSymBackend


def _sym_worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from exorl_b200 import knn as xknn
        from exorl_b200.relabel import relabel_apt_reward, last_query_block
        g = torch.Generator().manual_seed(5)
        rep = torch.tanh(torch.randn(n, 16, generator=g))
        args = types.SimpleNamespace(knn_k=4, knn_avg=True, knn_rms=True, knn_clip=0.0)
        rms = types.SimpleNamespace(M=torch.zeros(1), S=torch.ones(1), n=1e-4)
        xknn.sym_last.clear()
        r, d, i = relabel_apt_reward(rep, args, rms=rms, exchange="queries",
                                     stages=_OracleStagesSym, return_knn=True)
        assert xknn.sym_last.get("world") == world
        q.put((rank, r.numpy(), float(rms.M), float(rms.S), rms.n, d.numpy(), i.numpy(),
               last_query_block["lo"], last_query_block["hi"]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 64), (2, 51), (3, 70), (4, 96)])
def test_symmetric_screen_over_ranks_equals_unsharded(world, n):
    """Every unordered tile pair on exactly one rank, transposed hits routed to the owner of
    their row with an all-to-all: rewards, RMS state and every rank's own (dist, idx) block equal
    the unsharded oracle."""
    from oracle import RmsState, pbe_reward, knn_smallest
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sym_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(5)
    rep = torch.tanh(torch.randn(n, 16, generator=g))
    st = RmsState()
    want = pbe_reward(rep, rep, st, 4, True, True, 0.0).numpy()
    od, oi = knn_smallest(rep, rep, 4)
    covered = 0
    for rank, r, m, s, cnt, d, i, lo, hi in res:
        np.testing.assert_allclose(r, want, rtol=1e-5, atol=1e-7)
        assert m == pytest.approx(float(st.M), rel=1e-5) and s == pytest.approx(float(st.S), rel=1e-4)
        assert cnt == pytest.approx(st.n)
        np.testing.assert_allclose(d, od[lo:hi].numpy(), rtol=1e-5, atol=1e-6)
        assert (np.sort(i, axis=1) == np.sort(oi[lo:hi].numpy(), axis=1)).all()
        covered += hi - lo
    assert covered == n
