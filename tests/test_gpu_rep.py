"""GPU parity of the rep producers through the C ABI.

Bars: rep vs the reference's ICM.get_rep / F.normalize fixtures rtol 1e-5 + atol 2e-6 (fp32,
different summation order of the Linear / LayerNorm reductions); prepared operand BIT-EXACT equal
to exb_knn_prep_f16 of the produced rep; kNN on the prepared operand == kNN of the same rep
through the ordinary entry (bit-exact distances and indices); rewards vs the reference rel 1e-4
(the rep's 1e-5 propagates through distance differences)."""
import types

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from oracle import RmsState, icm_trunk_rep, l2_normalize, pbe_reward, proto_reward
from helpers import assert_knn_matches

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REP_TOL = dict(rtol=1e-5, atol=2e-6)


def _params(g, tag):
    return [torch.from_numpy(g[f"{tag}_{n}"]).to(DEV) for n in ("w", "b", "lnw", "lnb")] + [
        float(g[f"{tag}_eps"])]


def _assert_operand_is_prep_of_rep(p):
    q = xb.prepare_rep(p.rep.clone())
    assert torch.equal(p.f16.view(torch.int16), q.f16.view(torch.int16))
    assert torch.equal(p.half_norm, q.half_norm)
    assert torch.equal(p.round_err, q.round_err)
    assert torch.equal(p.scalars, q.scalars)


@pytest.mark.parametrize("tag", ["walker", "small", "wide", "odd"])
def test_icm_trunk_matches_reference_fixture(golden, tag):
    g = golden("rep_reference.npz")
    obs = torch.from_numpy(g[f"{tag}_obs"]).to(DEV)
    p = xb.icm_get_rep(tuple(_params(g, tag)), obs)
    np.testing.assert_allclose(p.rep.cpu().numpy(), g[f"{tag}_rep"], **REP_TOL)
    _assert_operand_is_prep_of_rep(p)


@pytest.mark.parametrize("tag", ["walker", "small"])
def test_icm_trunk_accepts_modules(golden, tag):
    g = golden("rep_reference.npz")
    w, b, lnw, lnb, eps = _params(g, tag)
    trunk = torch.nn.Sequential(torch.nn.Linear(w.size(1), w.size(0)),
                                torch.nn.LayerNorm(w.size(0), eps=eps), torch.nn.Tanh()).to(DEV)
    with torch.no_grad():
        trunk[0].weight.copy_(w); trunk[0].bias.copy_(b)
        trunk[1].weight.copy_(lnw); trunk[1].bias.copy_(lnb)

    class Icm(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.trunk = trunk
    obs = torch.from_numpy(g[f"{tag}_obs"]).to(DEV)
    a = xb.icm_get_rep(Icm(), obs).rep
    b2 = xb.icm_get_rep(trunk, obs).rep
    assert torch.equal(a, b2)
    np.testing.assert_allclose(a.cpu().numpy(), g[f"{tag}_rep"], **REP_TOL)


@pytest.mark.parametrize("tag", ["walker", "small", "wide"])
def test_apt_reward_from_obs_matches_reference(golden, tag):
    """obs -> trunk -> PBE, the whole compute_intr_reward of icm_apt.py:96-100."""
    g = golden("rep_reference.npz")
    obs = torch.from_numpy(g[f"{tag}_obs"]).to(DEV)
    p = xb.icm_get_rep(tuple(_params(g, tag)), obs)
    rms = xb.RMS(DEV)
    got = xb.PBE(rms, 0.0, 12, True, True, DEV)(p)
    assert got.shape == (obs.size(0), 1)
    np.testing.assert_allclose(got.cpu().numpy(), g[f"{tag}_reward"], rtol=1e-4, atol=1e-6)
    # same rep through the ordinary entry point: identical neighbours, rewards equal to fp32 ulps
    rms2 = xb.RMS(DEV)
    pbe2 = xb.PBE(rms2, 0.0, 12, True, True, DEV)
    pbe2.mode = "screened"
    want = pbe2(p.rep)
    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-6)
    assert rms.n == rms2.n
    # and the oracle evaluated on OUR rep (isolates the kNN/epilogue from the trunk's rounding)
    rep_cpu = p.rep.cpu()
    ora = pbe_reward(rep_cpu, rep_cpu, RmsState(), 12, True, True, 0.0)
    np.testing.assert_allclose(got.cpu().numpy(), ora.numpy(), rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("n,in_dim,dim", [(1, 24, 512), (7, 24, 512), (8, 24, 512), (1027, 24, 512),
                                          (300, 3, 4), (513, 96, 256), (4099, 24, 384),
                                          (100, 110, 512), (64, 39200, 512),
                                          (1024, 39200, 512), (300, 5000, 132), (129, 1028, 512),
                                          (77, 39201, 64),
                                          (9001, 24, 512), (8195, 17, 128), (8194, 6, 36)])
def test_icm_trunk_shapes_vs_oracle(n, in_dim, dim):
    g = torch.Generator().manual_seed(n + in_dim)
    obs = torch.randn(n, in_dim, generator=g)
    w = torch.randn(dim, in_dim, generator=g) / in_dim ** 0.5
    b = 0.1 * torch.randn(dim, generator=g)
    lnw = 1 + 0.1 * torch.randn(dim, generator=g)
    lnb = 0.1 * torch.randn(dim, generator=g)
    want = icm_trunk_rep(obs, w, b, lnw, lnb, 1e-5)
    p = xb.icm_get_rep((w.to(DEV), b.to(DEV), lnw.to(DEV), lnb.to(DEV)), obs.to(DEV))
    # wide inputs run the Linear as a 3xTF32 tensor-core GEMM (fp32-class); in_dim % 4 != 0 falls
    # back to the library GEMM, whose summation order is further from the CPU reference's
    tol = dict(rtol=1e-5, atol=2e-6) if in_dim < 1000 else (
        dict(rtol=1e-5, atol=4e-6) if in_dim % 4 == 0 else dict(rtol=1e-4, atol=2e-5))
    np.testing.assert_allclose(p.rep.cpu().numpy(), want.numpy(), **tol)
    _assert_operand_is_prep_of_rep(p)


def test_trunk_large_knn_equals_ordinary_path():
    """20k walker observations -> rep -> screened kNN on the producer's operand (streaming
    kernel): bit-identical to the ordinary screened entry on the same rep, and equal to the
    oracle's neighbours."""
    g = torch.Generator().manual_seed(11)
    n = 20000
    obs = torch.randn(n, 24, generator=g).to(DEV)
    w = (torch.randn(512, 24, generator=g) / 24 ** 0.5).to(DEV)
    z = torch.zeros(512, device=DEV)
    p = xb.icm_get_rep((w, z, torch.ones(512, device=DEV), z), obs)
    d1, i1 = xb.knn_topk(p, p, 12)
    d2, i2 = xb.knn_topk(p.rep, p.rep, 12, mode="screened")
    assert torch.equal(d1, d2) and torch.equal(i1, i2)
    assert xb.knn.screened_status(DEV)[1] == 0
    sub = torch.arange(0, n, 97)
    rep = p.rep.cpu()
    from oracle import knn_smallest
    od, oi = knn_smallest(rep[sub], rep, 12)
    assert_knn_matches(rep[sub], rep, 12, d1[sub.to(DEV)], i1[sub.to(DEV)], oracle=(od, oi))


def test_normalize_matches_reference_fixture(golden):
    g = golden("rep_reference.npz")
    x = torch.from_numpy(g["norm_x"]).to(DEV)
    p = xb.normalize_rep(x)
    np.testing.assert_allclose(p.rep.cpu().numpy(), g["norm_out"], rtol=2e-6, atol=1e-30)
    assert not p.rep[5].any()
    _assert_operand_is_prep_of_rep(p)


def test_proto_reward_from_prepared_z():
    """proto.py:102-118 with the producer: z = normalize(predictor output) prepared in one pass,
    queue prepared by the ordinary prep; same reward as the plain-tensor call and the oracle."""
    g = torch.Generator().manual_seed(3)
    raw = torch.randn(1024, 128, generator=g)
    q = torch.nn.functional.normalize(torch.randn(2048, 128, generator=g), dim=1, p=2)
    q[1500:] = 0.0
    p = xb.normalize_rep(raw.to(DEV))
    got = xb.proto_knn_reward(p, q.to(DEV), 3)
    want = xb.proto_knn_reward(p.rep, q.to(DEV), 3, mode="screened")
    assert torch.equal(got, want)
    ora = proto_reward(p.rep.cpu(), q, 3)
    np.testing.assert_allclose(got.cpu().numpy(), ora.numpy(), rtol=1e-5, atol=1e-7)
    ref = proto_reward(l2_normalize(raw), q, 3)
    np.testing.assert_allclose(got.cpu().numpy(), ref.numpy(), rtol=1e-4, atol=1e-6)


def test_ln_tanh_in_place_and_errors():
    g = torch.Generator().manual_seed(5)
    pre = torch.randn(333, 256, generator=g)
    lnw, lnb = torch.ones(256), torch.zeros(256)
    want = torch.tanh(torch.nn.functional.layer_norm(pre, (256,), lnw, lnb, 1e-5))
    p = xb.ln_tanh_rep(pre.to(DEV), lnw.to(DEV), lnb.to(DEV))
    np.testing.assert_allclose(p.rep.cpu().numpy(), want.numpy(), **REP_TOL)
    with pytest.raises(ValueError, match="rep_dim"):
        xb.ln_tanh_rep(torch.randn(8, 6, device=DEV), torch.ones(6, device=DEV),
                       torch.zeros(6, device=DEV))
    with pytest.raises(RuntimeError, match="no CPU path"):
        xb.normalize_rep(torch.randn(4, 8))
    with pytest.raises(ValueError, match="screened path"):
        xb.knn_topk(p, p, 3, mode="exact")


def _walker_trunk(seed, dim=512):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(dim, 24, generator=g) / 24 ** 0.5
    b = 0.1 * torch.randn(dim, generator=g)
    lnw = 1 + 0.1 * torch.randn(dim, generator=g)
    lnb = 0.1 * torch.randn(dim, generator=g)
    return tuple(t.to(DEV) for t in (w, b, lnw, lnb))


@pytest.mark.parametrize("world", [2, 8])
def test_prepared_shards_lists_form_equals_ordinary(world):
    """Sharded relabel from observations: the bank shard's operand is a row slice of the
    producer's operand (scalars of the whole set).  Screen-only lists per virtual shard + exact
    rerank == the ordinary (tensor) lists form == the unsharded prepared kNN, bit for bit."""
    from exorl_b200 import knn as xknn
    from exorl_b200.relabel import shard_bounds
    n, k = 40000, 12
    obs = torch.randn(n, 24, generator=torch.Generator().manual_seed(21)).to(DEV)
    p = xb.icm_get_rep(_walker_trunk(22, 128), obs)
    bounds = [shard_bounds(n, world, r) for r in range(world)]
    shards, sc = [], None
    for lo, hi in bounds:
        val, idx, cnt, ha, ea, s = xknn.knn_screen_lists(p, p.rows(lo, hi), k)
        shards.append((val.clone(), idx.clone(), cnt.clone()))
        sc = s.clone() if sc is None else torch.maximum(sc, s)
        assert torch.equal(ha, p.half_norm) and torch.equal(ea, p.round_err)
    d_all = torch.empty(n, k, device=DEV)
    i_all = torch.empty(n, k, dtype=torch.int32, device=DEV)
    for lo, hi in bounds:
        val = torch.stack([t[0][lo:hi] for t in shards])
        idx = torch.stack([t[1][lo:hi] for t in shards])
        cnt = torch.stack([t[2][lo:hi] for t in shards])
        d, i = xknn.knn_rerank_lists(p.rep[lo:hi], p.rep, k, val, idx, cnt,
                                     [a for a, _ in bounds], p.half_norm[lo:hi],
                                     p.round_err[lo:hi], sc)
        d_all[lo:hi], i_all[lo:hi] = d, i
    d_u, i_u = xb.knn_topk(p, p, k)
    assert torch.equal(d_all, d_u) and torch.equal(i_all, i_u)
    d_e, i_e = xb.knn_topk(p.rep, p.rep, k, mode="exact")
    np.testing.assert_allclose(d_all.cpu().numpy(), d_e.cpu().numpy(), rtol=1e-6)


def test_relabel_from_observations_single_process():
    """relabel_apt_reward(PreparedRep) == relabel_apt_reward(rep tensor) == oracle on a subset."""
    n = 20000
    args = types.SimpleNamespace(knn_k=12, knn_avg=True, knn_rms=True, knn_clip=0.0)
    obs = torch.randn(n, 24, generator=torch.Generator().manual_seed(31)).to(DEV)
    p = xb.icm_get_rep(_walker_trunk(32), obs)
    rms_a, rms_b = xb.RMS(DEV), xb.RMS(DEV)
    ra, da, ia = xb.relabel_apt_reward(p, args, rms=rms_a, return_knn=True)
    rb, db, ib = xb.relabel_apt_reward(p.rep, args, rms=rms_b, mode="screened", return_knn=True)
    assert torch.equal(da, db) and torch.equal(ia, ib) and torch.equal(ra, rb)
    assert rms_a.n == rms_b.n and torch.equal(rms_a.M, rms_b.M)
    rep = p.rep.cpu()
    sub = torch.arange(0, n, 101)
    from oracle import knn_smallest
    od, oi = knn_smallest(rep[sub], rep, 12)
    assert_knn_matches(rep[sub], rep, 12, da[sub.to(DEV)], ia[sub.to(DEV)], oracle=(od, oi))


def test_prepared_rep_edge_cases():
    """Empty batch, k larger than the bank (reference: RuntimeError from topk, utils/utils.py:297),
    a prepared source against a plain-tensor target (compute_apt_reward with source != target)."""
    tr = _walker_trunk(41, 128)
    empty = xb.icm_get_rep(tr, torch.empty(0, 24, device=DEV))
    assert empty.rep.shape == (0, 128)
    small = xb.icm_get_rep(tr, torch.randn(8, 24, device=DEV))
    with pytest.raises(RuntimeError, match="selected index k out of range"):
        xb.knn_topk(small, small, 12)
    with pytest.raises(RuntimeError, match="selected index k out of range"):
        xb.PBE(xb.RMS(DEV), 0.0, 12, True, True, DEV)(small)
    g = torch.Generator().manual_seed(42)
    src = xb.icm_get_rep(tr, torch.randn(300, 24, generator=g).to(DEV))
    bank = torch.tanh(torch.randn(700, 128, generator=g)).to(DEV)
    args = types.SimpleNamespace(knn_k=5, knn_avg=False, knn_rms=True, knn_clip=0.1)
    got = xb.compute_apt_reward(src, bank, args, rms=xb.RMS(DEV))
    want = pbe_reward(src.rep.cpu(), bank.cpu(), RmsState(), 5, False, True, 0.1)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
    # non-finite observation: the row is flagged (round_err = inf), nothing crashes, the other
    # rows keep their exact neighbours
    obs = torch.randn(400, 24, generator=g)
    obs[7, 3] = float("nan")
    p = xb.icm_get_rep(tr, obs.to(DEV))
    assert torch.isinf(p.round_err[7]) and int(p.scalars[2]) == 1
