"""Sharded relabel on one device: virtual shards + merge kernel == unsharded."""
import types

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from exorl_b200.relabel import shard_bounds, _cuda_merge
from oracle import RmsState, pbe_reward, knn_smallest
from helpers import reps, assert_knn_matches

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("mode", ["exact", "screened"])
@pytest.mark.parametrize("world", [2, 8])
def test_virtual_shards_merge_equals_unsharded(mode, world):
    n, k = 3000, 12
    x = reps(5, n, 64)
    xd = x.to(DEV)
    ds, is_ = [], []
    for r in range(world):
        lo, hi = shard_bounds(n, world, r)
        d, i = xb.knn_topk(xd, xd[lo:hi], k, idx_base=lo, mode=mode)
        ds.append(d)
        is_.append(i)
    d, i = _cuda_merge(torch.stack(ds), torch.stack(is_), k)
    assert_knn_matches(x, x, k, d, i)


def test_merge_tie_rule_and_padding():
    d = torch.tensor([[[1., 2., 2.]], [[0.5, 2., 9.]]], device=DEV)          # (G=2, n=1, k=3)
    i = torch.tensor([[[7, 3, 9]], [[100, 1, 5]]], dtype=torch.int32, device=DEV)
    od, oi = _cuda_merge(d, i, 3)
    assert od.flatten().tolist() == [0.5, 1.0, 2.0] and oi.flatten().tolist() == [100, 7, 1]


def test_relabel_single_process_matches_oracle():
    n = 5000
    x = reps(6, n, 128)
    args = types.SimpleNamespace(knn_k=12, knn_avg=True, knn_rms=True, knn_clip=0.0)
    rms = xb.RMS(DEV)
    got, d, i = xb.relabel_apt_reward(x.to(DEV), args, rms=rms, return_knn=True)
    st = RmsState()
    knn = knn_smallest(x, x, 12)
    want = pbe_reward(x, x, st, 12, True, True, 0.0, knn=knn)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
    assert rms.n == st.n
    assert_knn_matches(x, x, 12, d, i, oracle=knn)


def _scale_bank(g):
    """The seeded bank a scale fixture was computed on (oracle/gen_scale_golden.py:bank); the
    stored fp64 sum tells a different host random stream from a kNN defect."""
    n, dim = int(g["n"]), int(g["dim"])
    gen = torch.Generator().manual_seed(int(g["seed"]))
    x = torch.tanh(torch.randn(n, dim, generator=gen))
    assert abs(float(x.double().sum()) - float(g["bank_sum"])) <= 1e-6 * n, \
        "host RNG stream differs from the authoring container's"
    return x


def test_medium_bank_all_queries_vs_oracle_fixture(golden):
    """32 768 x 32 768 x 512 (SURVEY §8c full-N check): EVERY query's k nearest from the tcgen05
    path against the tiled reference formula (tests/golden/knn_scale_32k.npz, computed by
    oracle/gen_scale_golden.py from utils/utils.py:293-300), and the fp32 CUDA-core path too."""
    g = golden("knn_scale_32k.npz")
    x = _scale_bank(g)
    n, k = x.size(0), int(g["k"])
    oracle = (torch.from_numpy(g["dist"]), torch.from_numpy(g["idx"]).long())
    xd = x.to(DEV)
    d_s, i_s = xb.knn_topk(xd, xd, k, mode="screened")
    from exorl_b200 import knn as xknn
    und, invalid, pairs, _ = xknn.screened_status(torch.device(DEV))
    assert invalid == 0 and und < n // 100
    ties = assert_knn_matches(x, x, k, d_s, i_s, oracle=oracle)
    assert ties <= n // 1000
    d_e, i_e = xb.knn_topk(xd, xd, k, mode="exact")
    assert_knn_matches(x, x, k, d_e, i_e, oracle=oracle)
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)


@pytest.mark.parametrize("world", [2, 8])
def test_lists_exchange_form_equals_exact_kernel(world):
    """'lists' exchange on one device with virtual shards: screen-only per shard
    (exb_knn_screen_lists_f32), lists regrouped by query block as the all-to-all would, exact
    rerank per block against the whole bank (exb_knn_rerank_lists_f32) == the fp32 exact kernel;
    a 200-query subset is checked against the oracle."""
    from exorl_b200 import knn as xknn
    n, dim, k = 40000, 64, 12
    x = reps(7, n, dim)
    xd = x.to(DEV)
    bounds = [shard_bounds(n, world, r) for r in range(world)]
    per_shard = []
    sc = None
    for lo, hi in bounds:
        assert xknn.screen_lists_count(n, hi - lo, dim, k) * world <= 16
        val, idx, cnt, ha, ea, s = xknn.knn_screen_lists(xd, xd[lo:hi], k)
        per_shard.append((val.clone(), idx.clone(), cnt.clone()))
        sc = s.clone() if sc is None else torch.maximum(sc, s)
    d_e, i_e = xb.knn_topk(xd, xd, k, mode="exact")
    d_all = torch.empty(n, k, device=DEV)
    i_all = torch.empty(n, k, dtype=torch.int32, device=DEV)
    for lo, hi in bounds:
        val = torch.stack([p[0][lo:hi] for p in per_shard])
        idx = torch.stack([p[1][lo:hi] for p in per_shard])
        cnt = torch.stack([p[2][lo:hi] for p in per_shard])
        d, i = xknn.knn_rerank_lists(xd[lo:hi], xd, k, val, idx, cnt, [a for a, _ in bounds],
                                     ha[lo:hi], ea[lo:hi], sc)
        d_all[lo:hi], i_all[lo:hi] = d, i
    np.testing.assert_allclose(d_all.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)
    diff = (i_all.sort(dim=1).values != i_e.sort(dim=1).values).any(dim=1).sum().item()
    assert diff <= n // 1000
    sub = torch.arange(0, n, n // 200)[:200]
    assert_knn_matches(x[sub], x, k, d_all[sub.to(DEV)], i_all[sub.to(DEV)])


def test_lists_form_rejects_per_update_sizes():
    from exorl_b200 import knn as xknn
    assert xknn.screen_lists_count(1024, 1024, 512, 12) == 0
    x = reps(1, 1024, 64).to(DEV)
    with pytest.raises(ValueError):
        xknn.knn_screen_lists(x, x, 12)


def test_query_block_against_long_bank_wave_splits():
    """A query block with more tiles than SMs against a long bank (the 'queries' exchange of the
    sharded relabel): the plan cuts the bank into <= 4 ranges to fill the last wave; results must
    equal the exact fp32 kernel and, on a subset, the oracle."""
    nq, nb, dim, k = 25000, 600000, 64, 12
    bank = reps(13, nb, dim).to(DEV)
    q = bank[300000:300000 + nq]
    d_s, i_s = xb.knn_topk(q, bank, k, mode="screened")
    und, invalid, _, _ = xb.knn.screened_status(DEV)
    assert invalid == 0
    d_e, i_e = xb.knn_topk(q, bank, k, mode="exact")
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)
    diff = (i_s.sort(dim=1).values != i_e.sort(dim=1).values).any(dim=1).sum().item()
    assert diff <= nq // 1000
    assert torch.equal(i_s[:, 0].cpu(), torch.arange(300000, 300000 + nq, dtype=torch.int32))
    sub = torch.arange(0, nq, 250)
    qc, bc = q.cpu(), bank.cpu()
    assert_knn_matches(qc[sub], bc, k, d_s[sub.to(DEV)], i_s[sub.to(DEV)])


@pytest.mark.parametrize("nq", [20224, 26624, 19000 + 77])
def test_hybrid_tail_grids_equal_exact_kernel(nq):
    """Query blocks whose last wave of CTA pairs is partly empty (5, 30 and ~1 pairs on 74 slots):
    the screen cuts those tiles' scans into 4 / 2 / 4 bank ranges (hybrid tail).  Distances must
    equal the exact fp32 kernel bit for bit up to fp32 ulps, index sets up to near-ties, and rows of
    full units and of tail units must both be right (first, middle and last rows checked against
    the oracle)."""
    nb, dim, k = 300000, 32, 12
    bank = reps(17, nb, dim).to(DEV)
    q = reps(18, nq, dim).to(DEV)
    d_s, i_s = xb.knn_topk(q, bank, k, mode="screened")
    assert xb.knn.screened_status(DEV)[1] == 0
    d_e, i_e = xb.knn_topk(q, bank, k, mode="exact")
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)
    diff = (i_s.sort(dim=1).values != i_e.sort(dim=1).values).any(dim=1).sum().item()
    assert diff <= max(1, nq // 1000)
    sub = torch.cat([torch.arange(0, 64), torch.arange(nq // 2, nq // 2 + 64),
                     torch.arange(nq - 64, nq)])
    assert_knn_matches(q.cpu()[sub], bank.cpu(), k, d_s[sub.to(DEV)], i_s[sub.to(DEV)])


def test_randomized_shapes_screened_equals_exact():
    """40 random (nq, nb, dim, k) shapes across the kernel variants (latency chain, FILT 1/2/3,
    single CTA / CTA pair, ragged last tiles): screened == exact kernel."""
    rng = np.random.default_rng(2026)
    for it in range(40):
        nq = int(rng.integers(1, 5000))
        nb = int(rng.integers(1, 60000)) if it % 3 else int(rng.integers(1, 3000))
        dim = int(rng.choice([1, 3, 10, 24, 50, 64, 65, 128, 200, 256, 500, 512]))
        k = int(rng.integers(1, 17))
        if k > nb:
            k = nb
        g = torch.Generator().manual_seed(it)
        q = torch.tanh(torch.randn(nq, dim, generator=g)).to(DEV)
        b = torch.tanh(torch.randn(nb, dim, generator=g)).to(DEV)
        d_s, i_s = xb.knn_topk(q, b, k, mode="screened")
        assert xb.knn.screened_status(DEV)[1] == 0, (nq, nb, dim, k)
        d_e, i_e = xb.knn_topk(q, b, k, mode="exact")
        np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=2e-6, atol=1e-7,
                                   err_msg=str((nq, nb, dim, k)))
        diff = (i_s.sort(dim=1).values != i_e.sort(dim=1).values).any(dim=1).sum().item()
        assert diff <= max(1, nq // 500), (nq, nb, dim, k, diff)


def test_c5_full_size_properties(golden):
    """BASELINE config 5 at full size (10^6 x 10^6, rep 512, k = 12, avg + rms) on one GPU, checked
    through size-independent properties + a 4 096-query slice against the oracle fixture (the full
    oracle would take CPU-days): self-match first with distance exactly 0, ascending distances,
    distinct in-range indices, returned distances equal to a direct recompute of the returned
    pairs, rewards equal to the reference formula applied to the returned distances (fp64 on the
    GPU), RMS state = moments of all N*k distances, no undecided rows."""
    g = golden("knn_scale_1m.npz")
    rep_cpu = _scale_bank(g)                 # tanh(randn(10^6, 512)), seed 5
    n, dim, k = rep_cpu.size(0), rep_cpu.size(1), int(g["k"])
    assert (n, dim, k) == (1_000_000, 512, 12)
    rep = rep_cpu.to(DEV)
    args = types.SimpleNamespace(knn_k=k, knn_avg=True, knn_rms=True, knn_clip=0.0)
    rms = xb.RMS(DEV)
    reward, dist, idx = xb.relabel_apt_reward(rep, args, rms=rms, mode="screened", return_knn=True)
    # a set against itself at this size takes the symmetric screen (knn_sym.cuh): its status
    sym = xb.knn.sym_last
    assert sym.get("n") == n and "fell_back_to_full_screen" not in sym
    assert sym["status"] == 0 and sym["undecided"] == 0
    assert reward.shape == (n, 1) and dist.shape == (n, k) and idx.shape == (n, k)
    assert torch.equal(idx[:, 0], torch.arange(n, dtype=torch.int32, device=DEV))
    assert not dist[:, 0].any()
    assert bool((dist[:, 1:] >= dist[:, :-1]).all())
    assert bool((idx >= 0).all()) and bool((idx < n).all())
    srt = idx.sort(dim=1).values
    assert bool((srt[:, 1:] != srt[:, :-1]).all())
    # returned distances are the direct-difference distances of the returned pairs
    rows = torch.randint(0, n, (4096,), generator=torch.Generator(device=DEV).manual_seed(5),
                         device=DEV)
    nb_rows = idx[rows].long()
    direct = torch.norm(rep[rows][:, None, :] - rep[nb_rows.reshape(-1)].reshape(4096, k, dim),
                        dim=2, p=2)
    np.testing.assert_allclose(dist[rows].cpu().numpy(), direct.cpu().numpy(), rtol=2e-6, atol=1e-6)
    # epilogue + RMS (utils/utils.py:264-276, 309-318) recomputed in fp64 from the returned dists
    d64 = dist.double().reshape(-1)
    bs, n0 = float(d64.numel()), 1e-4
    mean, var = d64.mean(), d64.var(unbiased=True)
    m_new = 0.0 + (mean - 0.0) * bs / (n0 + bs)
    s_new = (1.0 * n0 + var * bs + mean * mean * n0 * bs / (n0 + bs)) / (n0 + bs)
    assert abs(float(rms.M) - float(m_new)) <= 1e-6 * float(m_new)
    assert abs(float(rms.S) - float(s_new)) <= 1e-5 * float(s_new)
    assert rms.n == n0 + bs
    want = torch.log((dist.double() / float(rms.M)).clamp_min(0).mean(dim=1, keepdim=True) + 1.0)
    np.testing.assert_allclose(reward.cpu().numpy(), want.float().cpu().numpy(), rtol=1e-5)
    # oracle slice: the reference formula (tiled) for a fixed 4 096-query slice against the whole
    # bank (tests/golden/knn_scale_1m.npz; ~40 CPU-minutes in oracle/gen_scale_golden.py)
    sub = torch.from_numpy(g["rows"])
    oracle = (torch.from_numpy(g["dist"]), torch.from_numpy(g["idx"]).long())
    assert_knn_matches(rep_cpu[sub], rep_cpu, k, dist[sub.to(DEV)], idx[sub.to(DEV)],
                       oracle=oracle)


@pytest.mark.parametrize("kind", ["duplicates", "outlier"])
def test_more_undecided_rows_than_the_device_fallback_holds(kind, monkeypatch):
    """Dataset-scale input on which the screen cannot certify > 4096 rows: 6000 identical rows
    (each has 6000 neighbours at distance 0, far more than a candidate list holds) or one
    element 1e15 times the others (the call's power-of-two scale then pushes every other value
    below the fp16 range: every row undecided).  The device-side fallback takes 4096 rows; the
    rest is flagged and recomputed from the host (strided sub-bank re-screen, exact kernel at the
    bottom) — results must equal the exact path, never silently wrong."""
    n, dim, k = 20000, 64, 12
    x = reps(23, n, dim)
    if kind == "duplicates":
        x[5000:11000] = x[4999]
    else:
        x[123, 7] = 1.0e15
    xd = x.to(DEV)
    from exorl_b200 import knn as xknn
    flagged = []
    real = xknn._recompute_rows

    def spy(src, tgt, kk, base, dist, idx, rows):
        flagged.append(int(rows.numel()))
        return real(src, tgt, kk, base, dist, idx, rows)
    monkeypatch.setattr(xknn, "_recompute_rows", spy)
    d_s, i_s = xb.knn_topk(xd, xd, k, mode="screened")
    assert flagged and flagged[0] > 1000          # rows beyond the 4096-row device fallback
    assert bool((i_s >= 0).all()) and bool((i_s < n).all())
    d_e, i_e = xb.knn_topk(xd, xd, k, mode="exact")
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6, atol=1e-6)
    if kind == "duplicates":
        assert not d_s[5000:11000].any()          # k duplicates at distance exactly 0
    # the reward path takes the same route (compute_apt_reward at dataset scale)
    args = types.SimpleNamespace(knn_k=k, knn_avg=True, knn_rms=True, knn_clip=0.0)
    r_s = xb.compute_apt_reward(xd, xd, args, rms=xb.RMS(DEV))
    r_e = xb.compute_apt_reward(xd, xd, args, rms=xb.RMS(DEV), mode="exact")
    np.testing.assert_allclose(r_s.cpu().numpy(), r_e.cpu().numpy(), rtol=1e-5, atol=1e-7)
    # policy "flag": the call never synchronises; the rows come back marked and are repaired later
    prev = xknn.set_overflow_policy("flag")
    try:
        before = len(flagged)
        d_f, i_f = xb.knn_topk(xd, xd, k, mode="screened")
        assert len(flagged) == before
        n_marked = int((i_f[:, 0] < 0).sum())
        assert n_marked == flagged[0]
        assert xknn.repair_flagged(xd, xd, k, d_f, i_f) == n_marked
        np.testing.assert_allclose(d_f.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6, atol=1e-6)
    finally:
        xknn.set_overflow_policy(prev)


def test_dense_trajectory_data_is_rescreened_not_scanned():
    """Trajectory-like reps whose consecutive states are near-duplicates (random walk, step 0.005):
    most rows have more bank rows inside the screening margin than a candidate list holds.  The
    flagged rows are re-screened against 8 interleaved sub-banks and merged; the result must equal
    the exact kernel, and only a small remainder may reach the exact fp32 scan."""
    from exorl_b200 import knn as xknn
    g = torch.Generator(device=DEV).manual_seed(0)
    E, L = 60, 1000
    start = torch.randn(E, 1, 24, generator=g, device=DEV)
    obs = (start + torch.cumsum(0.005 * torch.randn(E, L, 24, generator=g, device=DEV), dim=1))
    w = torch.randn(512, 24, generator=g, device=DEV) / 24 ** 0.5
    z = torch.zeros(512, device=DEV)
    p = xb.icm_get_rep((w, z, torch.ones(512, device=DEV), z), obs.reshape(-1, 24))
    calls = []
    real = xknn.knn_topk

    def spy(source, target, k, idx_base=0, mode="auto"):
        calls.append((mode, source.size(0), target.size(0)))
        return real(source, target, k, idx_base=idx_base, mode=mode)
    xknn.knn_topk, keep = spy, xknn.knn_topk
    try:
        d, i = real(p, p, 12)
    finally:
        xknn.knn_topk = keep
    d_e, i_e = xb.knn_topk(p.rep, p.rep, 12, mode="exact")
    np.testing.assert_allclose(d.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6, atol=1e-6)
    diff = (i.sort(dim=1).values != i_e.sort(dim=1).values).any(dim=1).sum().item()
    assert diff <= E * L // 500
    assert sum(1 for c in calls if c[0] == "screened") >= 8       # the strided re-screen ran
    assert sum(c[1] for c in calls if c[0] == "exact") <= E * L // 10


@pytest.mark.parametrize("n,dim", [(40000, 64), (20000, 512)])
def test_relabel_from_host_pipeline_equals_device_resident_call(n, dim):
    """relabel_apt_reward_from_host (pinned host rows -> H2D -> prep -> screen-only -> exact
    rerank; on N ranks the collectives slot in between) returns what relabel_apt_reward returns
    for device-resident reps, RMS state included."""
    x = reps(17, n, dim)
    args = types.SimpleNamespace(knn_k=12, knn_avg=True, knn_rms=True, knn_clip=0.0)
    rms_a, rms_b = xb.RMS(DEV), xb.RMS(DEV)
    want, wd, wi = xb.relabel_apt_reward(x.to(DEV), args, rms=rms_a, mode="screened",
                                         return_knn=True)
    host = x.pin_memory()
    with torch.cuda.device(DEV):
        got, gd, gi = xb.relabel_apt_reward_from_host(host, args, rms=rms_b, return_knn=True)
    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(gd.cpu().numpy(), wd.cpu().numpy(), rtol=1e-6)
    diff = (gi.sort(dim=1).values != wi.sort(dim=1).values).any(dim=1).sum().item()
    assert diff <= n // 1000
    assert rms_a.n == rms_b.n
    np.testing.assert_allclose(float(rms_a.M), float(rms_b.M), rtol=1e-6)


# ------------------------------------------------------------------ symmetric screen (knn_sym.cuh)
@pytest.fixture
def sym_on():
    from exorl_b200 import knn as xknn
    prev = (xknn.SYM_ENABLED, xknn.SYM_MIN_ROWS)
    xknn.SYM_ENABLED, xknn.SYM_MIN_ROWS = True, 1024
    xknn.sym_last.clear()
    yield xknn
    xknn.SYM_ENABLED, xknn.SYM_MIN_ROWS = prev


@pytest.mark.parametrize("n,dim,k,kind", [(20000, 64, 12, "tanh"), (9001, 512, 12, "tanh"),
                                          (12345, 50, 5, "raw"), (16384, 128, 3, "unit"),
                                          (6000, 130, 16, "tanh")])
def test_symmetric_screen_equals_full_screen_and_oracle(sym_on, n, dim, k, kind):
    """A set against itself through the symmetric screen (half the tile pairs, thresholds fixed
    from a sample pre-pass, direct + transposed hit logs): the result must be the full screen's
    — and the oracle's on a query slice — for ragged sizes, vector and scalar row layouts."""
    x = reps(41, n, dim, kind)
    xd = x.to(DEV)
    d_s, i_s = xb.knn_topk(xd, xd, k, mode="screened")
    assert sym_on.sym_last.get("n") == n and "fell_back_to_full_screen" not in sym_on.sym_last
    assert sym_on.sym_last["status"] == 0
    sym_on.SYM_ENABLED = False
    d_f, i_f = xb.knn_topk(xd, xd, k, mode="screened")
    # bit-equal unless one of the two paths sent a row to the exact kernel (other summation order)
    np.testing.assert_allclose(d_s.cpu().numpy(), d_f.cpu().numpy(), rtol=1e-6, atol=1e-7)
    assert (i_s != i_f).any(dim=1).sum().item() <= max(2, n // 2000)   # fp32 near-ties only
    sl = slice(n - 700, n)                                    # late rows: candidates arrive transposed
    assert_knn_matches(x[sl], x, k, d_s[sl], i_s[sl])
    assert_knn_matches(x[:300], x, k, d_s[:300], i_s[:300])


def test_symmetric_screen_on_shifted_scaled_and_duplicated_rows(sym_on):
    """Range-proof operands carry over (thresholds are formed in the operand space), and rows whose
    logs overflow (6000 duplicates) are recomputed — or the pass reports itself invalid and the
    full screen runs; either way the exact path's result."""
    n, dim, k = 24000, 64, 12
    x = reps(43, n, dim, "raw") * 1e5 + 3e6
    xd = x.to(DEV)
    d_s, i_s = xb.knn_topk(xd, xd, k, mode="screened")
    assert sym_on.sym_last["status"] == 0 and sym_on.sym_last["undecided"] == 0
    d_e, i_e = xb.knn_topk(xd, xd, k, mode="exact")
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)
    assert (i_s != i_e).any(dim=1).sum().item() <= 8
    y = reps(44, n, dim)
    y[5000:11000] = y[4999]
    yd = y.to(DEV)
    sym_on.sym_last.clear()
    d_s, i_s = xb.knn_topk(yd, yd, k, mode="screened")
    d_e, i_e = xb.knn_topk(yd, yd, k, mode="exact")
    np.testing.assert_allclose(d_s.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6, atol=1e-6)
    assert not d_s[5000:11000].any()


def test_symmetric_screen_relabel_reward_matches_oracle(sym_on):
    n, dim, k = 16000, 512, 12
    x = reps(45, n, dim)
    args = types.SimpleNamespace(knn_k=k, knn_avg=True, knn_rms=True, knn_clip=0.0)
    rms = xb.RMS(DEV)
    r = xb.relabel_apt_reward(x.to(DEV), args, rms=rms)
    assert sym_on.sym_last.get("n") == n and sym_on.sym_last["status"] == 0
    sl = slice(0, 400)
    st = RmsState()
    want = pbe_reward(x, x, st, k, True, True, 0.0)
    np.testing.assert_allclose(r.cpu().numpy()[sl], want.numpy()[sl], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(float(rms.M), float(st.M), rtol=1e-6)


@pytest.mark.parametrize("world", [2, 3, 4])
def test_symmetric_peer_routing_all_ranks_on_one_gpu(world):
    """The multi-GPU form with the exchange inside the kernel (exb_knn_sym_screen_peer_f32 ->
    exb_knn_sym_scatter_hits -> exb_knn_sym_rerank_f32), every "rank" played by this one GPU:
    rank r screens its span with the table of all ranks' hit buffers, each hit must land in the
    buffer of the owner of its query row, and the reranked blocks together must be the full
    screen's result.  (Across processes the buffers are CUDA-IPC mappings:
    benchmarks/relabel_nccl_check.py.)"""
    import ctypes
    from exorl_b200 import _lib, knn as xknn
    from exorl_b200.relabel import sym_tile_blocks, sym_span, _CudaSymBackend
    n, dim, k = 20000, 64, 12
    x = reps(47, n, dim)
    xd = x.to(DEV)
    lib = _lib.load()
    prev = xknn.SYM_ENABLED
    xknn.SYM_ENABLED = False
    try:
        d_f, i_f = xb.knn_topk(xd, xd, k, mode="screened")
        sample = xd[::16].contiguous()
        d_all = xb.knn_topk(xd, sample, k, mode="screened")[0][:, k - 1].contiguous()
    finally:
        xknn.SYM_ENABLED = prev
    expected = 1.3 * k * n / sample.size(0)
    log_cap = int(-(-(3 * expected / 2 + 96) // 8) * 8)
    recv_cap = int(-(-(3 * expected + 96) // 8) * 8)
    need = lib.exb_knn_sym_workspace_bytes(n, dim, k, log_cap, recv_cap)
    assert need > 0
    tiles, tb, rb = sym_tile_blocks(n, world)
    rows_max = max(rb[r + 1] - rb[r] for r in range(world))
    cap = -(-int(rows_max * (recv_cap // 3 + 16)) // 64) * 64
    with torch.cuda.device(DEV):
        ws = [torch.zeros(need, dtype=torch.uint8, device=DEV) for _ in range(world)]
        bufs = [torch.zeros(256 + 12 * cap, dtype=torch.uint8, device=DEV) for _ in range(world)]
        hits = (_lib.SymHits * world)()
        for r in range(world):
            hits[r] = _CudaSymBackend._hits_at(bufs[r].data_ptr(), cap)
        tl = (ctypes.c_int32 * (world + 1))(*tb)
        st = _lib.stream_ptr(DEV)
        for r in range(world):
            lo, hi, span_end, mid, extra = sym_span(tb, world, r)
            ready = 0
            if r % 2 == 1:
                # odd "ranks": operand prepared by the caller in shards straight into the
                # workspace's slots (what prepare_shard does across processes)
                slots, aff_p = _lib.Operand(), ctypes.c_void_p()
                _lib.check(lib.exb_knn_sym_operand_slots(n, dim, k, log_cap, recv_cap,
                                                         _lib.ptr(ws[r]), need,
                                                         ctypes.byref(slots), ctypes.byref(aff_p)))
                base = ws[r].data_ptr()
                dpad = -(-dim // 64) * 64
                aff = ws[r][aff_p.value - base:aff_p.value - base + 4 * _lib.AFFINE_FLOATS].view(
                    torch.float32)
                from exorl_b200.rep import column_affine
                aff.copy_(column_affine(xd))
                for a, b in ((0, n // 3), (n // 3, n)):
                    _lib.check(lib.exb_knn_prep_affine_f16(
                        _lib.ptr(xd[a:b]), b - a, dim, _lib.ptr(aff),
                        ctypes.c_void_p(slots.f16 + a * dpad * 2),
                        ctypes.c_void_p(slots.half_norm + 4 * a),
                        ctypes.c_void_p(slots.round_err + 4 * a),
                        ctypes.c_void_p(slots.scalars), st))
                ready = 1
            _lib.check(lib.exb_knn_sym_screen_peer_f32(
                _lib.ptr(xd), n, dim, k, _lib.ptr(d_all), 1, log_cap, recv_cap, lo, hi, span_end,
                mid, extra, world, hits, tl, ready, _lib.ptr(ws[r]), need, st))
        torch.cuda.synchronize()
        total = 0
        for r in range(world):                      # every hit sits with the owner of its row
            cnt = int(bufs[r][:8].view(torch.int64).item())
            assert 0 < cnt <= cap
            tgt = bufs[r][256:256 + 4 * cnt].view(torch.int32)
            assert int(tgt.min()) >= rb[r] and int(tgt.max()) < rb[r + 1]
            total += cnt
        assert total > n * k // 4                   # about half of all candidates travel
        d_parts, i_parts = [], []
        for r in range(world):
            _lib.check(lib.exb_knn_sym_scatter_hits(n, dim, k, log_cap, recv_cap,
                                                    ctypes.byref(hits[r]), _lib.ptr(ws[r]), st))
            d_r = torch.empty(rb[r + 1] - rb[r], k, dtype=torch.float32, device=DEV)
            i_r = torch.empty(rb[r + 1] - rb[r], k, dtype=torch.int32, device=DEV)
            _lib.check(lib.exb_knn_sym_rerank_f32(_lib.ptr(xd), n, dim, k, 0, log_cap, recv_cap,
                                                  rb[r], rb[r + 1], _lib.ptr(d_r), _lib.ptr(i_r),
                                                  _lib.ptr(ws[r]), st))
            status = (ctypes.c_int32 * 4)()
            _lib.check(lib.exb_knn_screened_status_sync(_lib.ptr(ws[r]), status, st))
            assert status[1] == 0, f"rank {r}: status flags {status[1]}"
            assert status[0] == 0 or (i_r[:, 0] < 0).sum().item() == 0   # fallbacks were recomputed
            d_parts.append(d_r)
            i_parts.append(i_r)
        d_s, i_s = torch.cat(d_parts), torch.cat(i_parts)
    np.testing.assert_allclose(d_s.cpu().numpy(), d_f.cpu().numpy(), rtol=1e-6, atol=1e-7)
    assert (i_s.long() != i_f.long()).any(dim=1).sum().item() <= max(2, n // 2000)
    assert_knn_matches(x[n - 500:], x, k, d_s[n - 500:], i_s[n - 500:])


def test_from_host_symmetric_pipeline_matches_device_path(sym_on):
    """relabel_apt_reward_from_host on the symmetric path: the rows arrive in chunks (two here)
    behind the sample pre-pass, the sample itself as one strided 2-D DMA; same rewards, same
    neighbours and the same device copy of the rows as upload-then-compute."""
    from exorl_b200 import relabel as xrel
    n, dim, k = 130000, 64, 12
    x = reps(48, n, dim)
    xh = x.pin_memory()
    args = types.SimpleNamespace(knn_k=k, knn_avg=True, knn_rms=True, knn_clip=0.0)
    rms_a, rms_b = xb.RMS(DEV), xb.RMS(DEV)
    out_a = torch.zeros(n, dim, device=DEV)
    r_a, d_a, i_a = xb.relabel_apt_reward_from_host(xh, args, rms=rms_a, rep_out=out_a,
                                                    return_knn=True)
    assert sym_on.sym_last.get("n") == n and "fell_back_to_full_screen" not in sym_on.sym_last
    assert torch.equal(out_a.cpu(), x)
    prev = xrel.SYM_HOST_PIPELINE
    xrel.SYM_HOST_PIPELINE = False
    try:
        r_b, d_b, i_b = xb.relabel_apt_reward_from_host(xh, args, rms=rms_b, return_knn=True)
    finally:
        xrel.SYM_HOST_PIPELINE = prev
    np.testing.assert_allclose(r_a.cpu().numpy(), r_b.cpu().numpy(), rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(d_a.cpu().numpy(), d_b.cpu().numpy(), rtol=1e-6, atol=1e-7)
    assert (i_a != i_b).any(dim=1).sum().item() <= n // 2000
    np.testing.assert_allclose(float(rms_a.M), float(rms_b.M), rtol=1e-6)
    assert_knn_matches(x[n - 300:], x, k, d_a[n - 300:], i_a[n - 300:])
