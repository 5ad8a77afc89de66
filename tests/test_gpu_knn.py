"""GPU parity: exorl_b200 (CUDA, through the C ABI) vs the oracle / reference golden vectors.

Tolerances (north_star): kNN index sets bit-exact outside exact/near ties; rewards rel 1e-5."""
import pickle
import types

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from exorl_b200 import knn as xknn
from oracle import RmsState, pbe_reward, knn_smallest, proto_reward, rms_step
from helpers import reps, assert_knn_matches

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5


def _pbe(clip, k, avg, rms_on, mode="auto"):
    p = xb.PBE(xb.RMS(DEV), clip, k, avg, rms_on, DEV)
    p.mode = mode
    return p


def _close(got, want, rtol=RTOL, atol=1e-7):
    np.testing.assert_allclose(np.asarray(got, np.float64), np.asarray(want, np.float64),
                               rtol=rtol, atol=atol, equal_nan=True)


# ---------------------------------------------------------------- known answers
REP4 = torch.tensor([[0., 0.], [3., 4.], [6., 8.], [0., 1.]])


@pytest.mark.parametrize("avg,rms_on,clip,want", [
    (True, True, 0.0, [0.30438635, 0.92008394, 1.02207816, 0.30438635]),
    (True, True, 0.5, [0.10056251, 0.81514400, 0.92780966, 0.10056251]),
    (True, False, 0.0, [0.40546510, 1.13825607, 1.25276291, 0.40546510]),
    (True, False, 0.5, [0.22314355, 1.05477190, 1.17865503, 0.22314355]),
    (False, True, 0.0, [0.30438960, 0.92009133, 1.02208614, 0.30438960]),
    (False, True, 0.5, [0.0, 0.69789571, 0.82373059, 0.0]),
    (False, False, 0.0, [0.69314718, 1.65682530, 1.79175949, 0.69314718]),
])
def test_kat_rewards(avg, rms_on, clip, want):
    r = _pbe(clip, 2, avg, rms_on)(REP4.to(DEV))
    assert r.shape == (4, 1) and r.dtype == torch.float32 and r.is_cuda
    _close(r.flatten().cpu(), want)


def test_kat_state_trajectory():
    p = _pbe(0.0, 2, True, True)
    x = REP4.to(DEV)
    p(x)
    assert float(p.rms.M) == pytest.approx(1.40531242, rel=1e-6)
    assert float(p.rms.S) == pytest.approx(4.17146730, rel=1e-6)
    assert p.rms.n == pytest.approx(8.0001)
    r2 = p(x)
    _close(r2.flatten().cpu(), [0.30438468, 0.92008013, 1.02207410, 0.30438468])
    assert float(p.rms.M) == pytest.approx(1.40532124, rel=1e-6)
    assert float(p.rms.S) == pytest.approx(4.17147541, rel=1e-6)
    assert p.rms.n == pytest.approx(16.0001)


def test_kat_rms_alone():
    r = xb.RMS(DEV)
    r(torch.tensor([[1.], [2.], [3.], [6.]], device=DEV))
    assert float(r.M) == pytest.approx(2.99992490, rel=1e-6)
    assert float(r.S) == pytest.approx(4.66679955, rel=1e-6)
    m, s = r(torch.tensor([[10.], [20.]], device=DEV))
    assert float(m) == pytest.approx(6.99988317, rel=1e-6)
    assert float(s) == pytest.approx(51.77775192, rel=1e-6)
    assert r.n == pytest.approx(6.0001) and m.shape == (1,) and m.is_cuda


def test_k_out_of_range_raises_like_reference():
    with pytest.raises(RuntimeError, match="selected index k out of range"):
        _pbe(0.0, 5, True, True)(REP4.to(DEV))


# ---------------------------------------------------------------- golden fixtures from the reference
def test_golden_all_modes_three_calls(golden):
    g = golden("pbe_reference.npz")
    for c in range(int(g["n_cases"])):
        tag = f"pbe{c:03d}"
        b, d, k, avg, rms_on = (int(v) for v in g[f"{tag}_cfg"])
        clip = float(g[f"{tag}_clip"])
        p = _pbe(clip, k, bool(avg), bool(rms_on))
        for call in range(3):
            x = torch.from_numpy(g[f"rep_{b}_{d}_{call}"]).to(DEV)
            r = p(x)
            _close(r.cpu().numpy(), g[f"{tag}_reward{call}"])
            if rms_on:
                m, s, n = g[f"{tag}_state{call}"]
                _close([float(p.rms.M), float(p.rms.S)], [m, s], rtol=2e-5, atol=1e-7)
                assert p.rms.n == n


@pytest.mark.parametrize("mode", ["exact", "screened"])
def test_golden_knn_index_sets(golden, mode):
    g = golden("knn_reference.npz")
    for tag in ("sq512", "rect", "proto", "tiny"):
        s, t = torch.from_numpy(g[f"{tag}_src"]), torch.from_numpy(g[f"{tag}_tgt"])
        k = g[f"{tag}_idx"].shape[1]
        sd = s.to(DEV)
        td = sd if tag in ("sq512", "tiny") else t.to(DEV)
        d, i = xb.knn_topk(sd, td, k, mode=mode)
        torch.cuda.synchronize()
        oracle = (torch.from_numpy(g[f"{tag}_dist"]), torch.from_numpy(g[f"{tag}_idx"]))
        assert_knn_matches(s, t, k, d, i, oracle=oracle)


def test_golden_proto(golden):
    g = golden("proto_reference.npz")
    z = torch.from_numpy(g["z"]).to(DEV)
    full = torch.from_numpy(g["queue_full"]).to(DEV)
    for tag, q in (("full", full.clone()),
                   ("warm", torch.cat([full[:64], torch.zeros(192, 128, device=DEV)]))):
        ptr = xb.proto_enqueue(q, 0, z, torch.arange(0, 32, device=DEV))
        assert ptr == int(g[f"{tag}_ptr"])
        assert np.array_equal(q.cpu().numpy(), g[f"{tag}_queue_after"])
        for mode in ("exact", "screened"):
            r = xb.proto_knn_reward(z, q, 3, mode=mode)
            assert r.shape == (96, 1)
            _close(r.cpu().numpy(), g[f"{tag}_reward"])


# ---------------------------------------------------------------- oracle-diff on random inputs
@pytest.mark.parametrize("mode", ["exact", "screened"])
@pytest.mark.parametrize("b1,b2,d,k", [(300, 300, 512, 12), (257, 1000, 50, 12), (128, 2048, 128, 3),
                                       (1024, 1024, 10, 12), (130, 513, 24, 1), (64, 700, 512, 16),
                                       (5, 17, 3, 4)])
def test_knn_vs_oracle(mode, b1, b2, d, k):
    s = reps(b1 * 7 + d, b1, d)
    t = s if b1 == b2 else reps(b2 * 3 + d, b2, d)
    sd = s.to(DEV)
    td = sd if b1 == b2 else t.to(DEV)
    dist, idx = xb.knn_topk(sd, td, k, mode=mode)
    assert_knn_matches(s, t, k, dist, idx)
    if b1 == b2:   # self-match first, distance exactly 0
        assert torch.equal(idx[:, 0].cpu().long(), torch.arange(b1))
        assert torch.all(dist[:, 0] == 0)


@pytest.mark.parametrize("mode", ["exact", "screened"])
def test_c1_full_size_reward_and_indices(mode):
    """BASELINE config 1: 1024 x 512, k=12, avg + rms, clip 0."""
    x = reps(0, 1024, 512)
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, 12, mode=mode)
    assert_knn_matches(x, x, 12, dist, idx)
    p = _pbe(0.0, 12, True, True, mode)
    st = RmsState()
    for _ in range(2):
        want = pbe_reward(x, x, st, 12, True, True, 0.0)
        got = p(xd)
        _close(got.cpu().numpy(), want.numpy())
    _close([float(p.rms.M), float(p.rms.S)], [float(st.M), float(st.S)], rtol=2e-5)
    assert p.rms.n == st.n
    if mode == "screened":
        und, invalid, pairs, _ = xknn.screened_status(torch.device(DEV))
        assert invalid == 0 and und <= 8 and pairs >= (1024 - und) * 12


@pytest.mark.parametrize("mode", ["exact", "screened"])
def test_c2_proto_full_size(mode):
    z = reps(1, 1024, 128, "unit")
    q = reps(2, 2048, 128, "unit")
    for queue in (q, torch.cat([q[:512], torch.zeros(1536, 128)])):   # full / warm-up state
        want = proto_reward(z, queue, 3)
        got = xb.proto_knn_reward(z.to(DEV), queue.to(DEV), 3, mode=mode)
        _close(got.cpu().numpy(), want.numpy())


@pytest.mark.parametrize("mode", ["exact", "screened"])
def test_duplicates_and_exact_ties(mode):
    """Duplicate rows (exact ties) are tolerated: distances match, indices valid and distinct."""
    base = reps(9, 200, 64)
    x = torch.cat([base, base[:100], base[:50]])          # up to 3 copies
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, 5, mode=mode)
    od, _ = knn_smallest(x, x, 5)
    _close(dist.cpu().numpy(), od.numpy())
    assert all(len(set(r)) == 5 for r in idx.cpu().tolist())


def test_screened_handles_many_ties_through_fallback():
    """More tied candidates than list slots: rows must come back exact via the fallback."""
    z = reps(4, 256, 128, "unit")
    q = torch.zeros(1024, 128)                            # every queue row at distance 1
    got = xb.proto_knn_reward(z.to(DEV), q.to(DEV), 3, mode="screened")
    _close(got.cpu().numpy(), proto_reward(z, q, 3).numpy())
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV), key="apt_reward")
    assert und == 256 and invalid == 0


@pytest.mark.parametrize("where", [(100, 140), (40, 90), (0, 300)])
def test_per_update_rows_recomputed_over_the_overflowed_splits_only(where):
    """A tight cluster of batch rows overflows the screen's list for the bank ranges (splits) that
    hold it.  Those rows are recomputed inline from the overflowed ranges + the other splits'
    complete lists; results stay equal to the oracle, PBE included."""
    lo, hi = where
    x = reps(17, 1024, 512)
    g = torch.Generator().manual_seed(5)
    x[lo:hi] = x[lo] + 1e-3 * torch.randn(hi - lo, 512, generator=g)
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, 12, mode="screened")
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV))
    # (a cluster of 300 sits inside the screening margin of every row of concentrated random reps)
    assert invalid == 0 and 0 < und and (und < 1024 or hi - lo > 100)
    assert_knn_matches(x, x, 12, dist, idx)
    p = _pbe(0.0, 12, True, True, "screened")
    st = RmsState()
    _close(p(xd).cpu().numpy(), pbe_reward(x, x, st, 12, True, True, 0.0).numpy())


def test_per_update_batch_of_consecutive_trajectory_states():
    """One whole episode as the batch: a row's nearest neighbours are the adjacent bank columns."""
    g = torch.Generator().manual_seed(3)
    walk = torch.cumsum(0.05 * torch.randn(1024, 24, generator=g), 0)
    w = torch.randn(24, 512, generator=g) / 24 ** 0.5
    x = torch.tanh(torch.nn.functional.layer_norm(walk @ w, (512,)))
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, 12, mode="screened")
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV))
    assert invalid == 0 and und <= 128
    assert_knn_matches(x, x, 12, dist, idx)


def test_screened_range_fallback_for_unbounded_reps():
    """Values outside fp16 range cannot be screened; the call must still return exact results."""
    x = reps(12, 300, 32, "raw") * 1e5
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, 4, mode="screened")
    assert_knn_matches(x, x, 4, dist, idx)


@pytest.mark.parametrize("mode", ["exact", "screened"])
def test_unscaled_and_tiny_reps(mode):
    for scale in (30.0, 1e-3):
        x = reps(21, 400, 96, "raw") * scale
        xd = x.to(DEV)
        dist, idx = xb.knn_topk(xd, xd, 12, mode=mode)
        assert_knn_matches(x, x, 12, dist, idx)


def test_properties_translation_and_permutation():
    x = reps(33, 512, 64)
    xd = x.to(DEV)
    _, i0 = xb.knn_topk(xd, xd, 8, mode="screened")
    perm = torch.randperm(512, generator=torch.Generator().manual_seed(1))
    xp = x[perm].to(DEV)
    dp, ip = xb.knn_topk(xp, xp, 8, mode="screened")
    assert_knn_matches(x[perm], x[perm], 8, dp, ip)
    # permutation equivariance of the neighbour sets
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(512)
    a = torch.sort(i0.cpu().long(), dim=1).values
    b = torch.sort(perm[ip.cpu().long()][inv], dim=1).values
    assert (a != b).any(dim=1).sum() <= 2     # only fp32 near-ties may differ


def test_3d_rep_is_flattened_like_reference():
    x = reps(5, 64, 24).reshape(64, 4, 6)
    want = pbe_reward(x, x, RmsState(), 3, True, True, 0.0)
    got = _pbe(0.0, 3, True, True)(x.to(DEV))
    _close(got.cpu().numpy(), want.numpy())


def test_compute_apt_reward_source_ne_target_and_grad_mode():
    s, t = reps(1, 200, 64), reps(2, 900, 64)
    args = types.SimpleNamespace(knn_k=12, knn_avg=True, knn_rms=True, knn_clip=0.0001)
    st = RmsState()
    want = pbe_reward(s, t, st, 12, True, True, 0.0001)
    rms = xb.RMS(DEV)
    sd = s.to(DEV).requires_grad_(True)            # must not fail under grad mode
    got = xb.compute_apt_reward(sd, t.to(DEV), args, rms=rms)
    assert not got.requires_grad
    _close(got.cpu().numpy(), want.numpy())
    assert rms.n == st.n


def test_rms_matches_oracle_trajectory_large_and_strided():
    r, st = xb.RMS(DEV), RmsState()
    for i, n in enumerate((50000, 3, 1, 70000)):
        x = reps(100 + i, n, 1, "raw") * 3 + 20
        rms_step(st, x)
        r(x.to(DEV))
        _close([float(r.M), float(r.S)], [float(st.M), float(st.S)], rtol=2e-5)
        assert r.n == st.n


def test_pbe_and_rms_are_picklable():
    p = _pbe(0.0, 3, True, True)
    x = reps(8, 64, 16).to(DEV)
    p(x)
    q = pickle.loads(pickle.dumps(p))
    assert q.rms.n == p.rms.n and torch.equal(q.rms.M, p.rms.M)
    assert torch.equal(q(x), p(x))


def test_large_moments_path():
    """nq*k above the single-block limit exercises the multi-block moments kernel."""
    x = reps(77, 4096, 32)
    st = RmsState()
    want = pbe_reward(x, x, st, 12, True, True, 0.0)
    p = _pbe(0.0, 12, True, True)
    got = p(x.to(DEV))
    _close(got.cpu().numpy(), want.numpy())
    _close([float(p.rms.M), float(p.rms.S)], [float(st.M), float(st.S)], rtol=2e-5)
    st2 = RmsState()
    want2 = pbe_reward(x, x, st2, 12, False, True, 0.5)
    _close(_pbe(0.5, 12, False, True)(x.to(DEV)).cpu().numpy(), want2.numpy())


@pytest.mark.parametrize("avg,rms_on,clip", [(True, True, 0.0), (False, True, 0.5), (True, False, 0.0),
                                            (False, False, -1.0)])
@pytest.mark.parametrize("b,d,k", [(1024, 512, 12), (1000, 50, 12), (640, 128, 3), (2048, 64, 16)])
def test_fused_per_update_chain_matches_oracle(avg, rms_on, clip, b, d, k):
    """exb_apt_reward_f32 on per-update sizes: latency screen + rerank with the RMS/epilogue fused
    into the last rerank block; three sequential calls carry the RMS state."""
    args = types.SimpleNamespace(knn_k=k, knn_avg=avg, knn_rms=rms_on, knn_clip=clip)
    rms, st = xb.RMS(DEV), RmsState()
    for call in range(3):
        x = reps(500 + call + b, b, d)
        want = pbe_reward(x, x, st, k, avg, rms_on, clip)
        xd = x.to(DEV)
        got, dist, idx = xb.compute_apt_reward(xd, xd, args, rms=rms, mode="screened",
                                               return_knn=True)
        _close(got.cpu().numpy(), want.numpy())
        assert_knn_matches(x, x, k, dist, idx)
        if rms_on:
            _close([float(rms.M), float(rms.S)], [float(st.M), float(st.S)], rtol=2e-5)
            assert rms.n == st.n
    und, invalid, pairs, _ = xknn.screened_status(torch.device(DEV), key="apt_reward")
    assert invalid == 0 and und <= 8


def test_fused_chain_source_ne_target_and_ragged():
    s, t = reps(1, 333, 96), reps(2, 1500, 96)
    args = types.SimpleNamespace(knn_k=12, knn_avg=True, knn_rms=True, knn_clip=0.0)
    st = RmsState()
    want = pbe_reward(s, t, st, 12, True, True, 0.0)
    rms = xb.RMS(DEV)
    got = xb.compute_apt_reward(s.to(DEV), t.to(DEV), args, rms=rms, mode="screened")
    _close(got.cpu().numpy(), want.numpy())
    _close([float(rms.M), float(rms.S)], [float(st.M), float(st.S)], rtol=2e-5)


@pytest.mark.parametrize("n,dim", [(1000, 512), (777, 50), (64, 130)])
def test_prep_operand_norms_and_rigorous_error_bound(n, dim):
    """exb_knn_prep_f16: the fp16 operand is the round-to-nearest of the input, half_norm matches
    the fp64 value to the budgeted 2^-19, and round_err is an UPPER bound of ||x - fp16(x)||
    (the screening certificate needs the bound, not an estimate) that is tight to 1e-5."""
    import ctypes
    from exorl_b200 import _lib
    lib = _lib.load()
    x = reps(11, n, dim) * 3.0
    xd = x.to(DEV).contiguous()
    dpad = (dim + 63) // 64 * 64
    out = torch.full((n, dpad), 7.0, dtype=torch.float16, device=DEV)
    hn = torch.empty(n, device=DEV)
    er = torch.empty(n, device=DEV)
    sc = torch.zeros(4, dtype=torch.int32, device=DEV)
    _lib.check(lib.exb_knn_prep_f16(_lib.ptr(xd), n, dim, _lib.ptr(out), _lib.ptr(hn),
                                    _lib.ptr(er), _lib.ptr(sc), _lib.stream_ptr(torch.device(DEV))))
    torch.cuda.synchronize()
    xh = x.half()
    assert torch.equal(out[:, :dim].cpu(), xh) and torch.all(out[:, dim:] == 0)
    xh64 = xh.double()
    want_hn = 0.5 * (xh64 ** 2).sum(dim=1)
    want_er = ((x.double() - xh64) ** 2).sum(dim=1).sqrt()
    got_hn, got_er = hn.cpu().double(), er.cpu().double()
    assert torch.all((got_hn - want_hn).abs() <= want_hn * 2.0 ** -19)
    assert torch.all(got_er >= want_er) and torch.all(got_er <= want_er * (1 + 1e-5) + 1e-12)
    s = sc.cpu()
    assert float(s[:1].view(torch.float32)) == float(er.max()) and int(s[2]) == 0
    assert float(s[1:2].view(torch.float32)) == float(hn.max())


@pytest.mark.parametrize("tag", ["walker", "odd"])
def test_rnd_reward_matches_reference_fixture(golden, tag):
    """exb_rnd_reward_f32 against the reference's RND.forward error + RMS-normalised reward
    (rnd.py:52-54,95-98; fixtures from oracle/gen_golden.py:gen_rnd), three sequential calls
    carrying the RMS state.  fp32 bar: errors / rewards rel 1e-5, state rel 1e-6."""
    g = golden("rnd_reference.npz")
    rms = xb.RMS(DEV)
    for call in range(3):
        pred = torch.from_numpy(g[f"{tag}_pred{call}"]).to(DEV)
        tgt = torch.from_numpy(g[f"{tag}_target{call}"]).to(DEV)
        r, e = xb.rnd_intr_reward(pred, tgt, rms, float(g[f"{tag}_scale"]))
        assert r.shape == (pred.size(0), 1) and e.shape == (pred.size(0), 1)
        np.testing.assert_allclose(e.cpu().numpy(), g[f"{tag}_error{call}"], rtol=1e-5)
        np.testing.assert_allclose(r.cpu().numpy(), g[f"{tag}_reward{call}"], rtol=1e-5)
        m, s, n = g[f"{tag}_state{call}"]
        assert abs(float(rms.M) - m) <= 1e-6 * abs(m) and abs(float(rms.S) - s) <= 2e-6 * abs(s)
        assert rms.n == n


def test_rnd_reward_large_and_unaligned():
    from oracle import rnd_reward
    gen = torch.Generator().manual_seed(8)
    for n, dim in ((100000, 512), (3001, 50), (1, 7)):
        p = torch.randn(n, dim, generator=gen)
        t = torch.randn(n, dim, generator=gen)
        st = RmsState()
        want_r, want_e = rnd_reward(p, t, st, 2.0)
        rms = xb.RMS(DEV)
        r, e = xb.rnd_intr_reward(p.to(DEV), t.to(DEV), rms, 2.0)
        np.testing.assert_allclose(e.cpu().numpy(), want_e.numpy(), rtol=1e-5)
        if n > 1:     # a single value has NaN variance in the reference too
            np.testing.assert_allclose(r.cpu().numpy(), want_r.numpy(), rtol=2e-5)
            assert abs(float(rms.S) - float(st.S)) <= 1e-5 * abs(float(st.S))
        else:
            assert torch.isnan(r).all() and torch.isnan(want_r).all()


@pytest.mark.parametrize("b,d,k", [(300, 600, 12), (260, 1024, 5), (400, 64, 20), (300, 32, 32),
                                   (130, 3 * 8 * 8, 12)])
def test_auto_mode_outside_the_screen_limits(b, d, k):
    """rep_dim > 512 or k > 16 are outside the tcgen05 screen: mode='auto' must route them to the
    exact fp32 kernel (same results), mode='screened' must refuse loudly; a 3-D rep is flattened
    like the reference's .view(b, 1, -1) (utils/utils.py:293-294)."""
    x = reps(77, b, d)
    if d == 3 * 8 * 8:
        x = x.reshape(b, 3, 8, 8)
    xd = x.to(DEV)
    dist, idx = xb.knn_topk(xd, xd, k)
    assert_knn_matches(x.reshape(b, -1), x.reshape(b, -1), k, dist, idx)
    pbe = xb.PBE(xb.RMS(DEV), 0.0, k, True, True, DEV)
    got = pbe(xd)
    want = pbe_reward(x, x, RmsState(), k, True, True, 0.0)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-5, atol=1e-7)
    if d > 512 or k > 16:
        with pytest.raises(ValueError):
            xb.knn_topk(xd, xd, k, mode="screened")
    with pytest.raises(ValueError):
        xb.knn_topk(xd, xd, 33)


def test_library_graph_cache_hits_misses_and_bypass():
    """The per-update chain is replayed from a graph cached inside the library, keyed by the
    call's pointers and shapes.  Same tensor again -> cache hit (the RMS state still advances,
    it lives on the device); a stream of never-repeating tensors -> the cache gives up and the
    plain launch chain serves the calls; results equal the oracle's either way."""
    p = _pbe(0.0, 12, True, True)
    st = RmsState()
    x = reps(321, 512, 64)
    xd = x.to(DEV)
    out = torch.empty(0)
    for call in range(6):                       # first call captures, the rest replay
        got = p(xd)
        want = pbe_reward(x, x, st, 12, True, True, 0.0)
        _close(got.cpu().numpy(), want.numpy())
        assert p.rms.n == st.n
    keep = []                                   # hold every tensor: no address is ever reused
    for call in range(40):
        y = reps(400 + call, 256, 32)
        keep.append(y.to(DEV))
        got = p(keep[-1])
        keep.append(got)
        want = pbe_reward(y, y, st, 12, True, True, 0.0)
        _close(got.cpu().numpy(), want.numpy())
    got = p(xd)                                 # and the first tensor still works afterwards
    _close(got.cpu().numpy(), pbe_reward(x, x, st, 12, True, True, 0.0).numpy())
    assert p.rms.n == st.n
    # inside a caller's CUDA graph the kernels are captured into THAT graph
    pg = _pbe(0.0, 12, True, True)
    stg = RmsState()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        pg(xd)
    torch.cuda.current_stream().wait_stream(s)
    rms_step_calls = 1
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        r = pg(xd)
    for _ in range(3):
        g.replay()
        rms_step_calls += 1
    torch.cuda.synchronize()
    for _ in range(rms_step_calls):
        want = pbe_reward(x, x, stg, 12, True, True, 0.0)
    _close(r.cpu().numpy(), want.numpy())
    assert pg.rms.n == stg.n


# ------------------------------------------------------------- range-proof operands (affine)
@pytest.mark.parametrize("n,dim", [(5000, 512), (3001, 50), (700, 130), (300, 4)])
def test_column_stats_and_affine_record(n, dim):
    """exb_knn_colstats_f32 / exb_knn_affine_from_stats: per-column sum / min / max of the rows,
    mu = sum / n, and a power-of-two scale that puts max |x - mu| into [64, 128)."""
    from exorl_b200 import _lib
    x = reps(3, n, dim, "raw") * 37.0 + 1000.0
    xd = x.to(DEV)
    st = xb.column_stats(xd).cpu()
    np.testing.assert_allclose(st[0].numpy(), x.double().sum(0).numpy(), rtol=1e-6)
    assert torch.equal(st[1], x.min(0).values) and torch.equal(st[2], x.max(0).values)
    assert torch.equal(xb.column_stats(xd).cpu(), st)            # deterministic
    aff = xb.column_affine(xd).cpu()
    mu, s, on = aff[:dim], float(aff[_lib.SCREEN_MAX_DIM]), float(aff[_lib.SCREEN_MAX_DIM + 1])
    assert on == 1.0 and torch.all(aff[dim:_lib.SCREEN_MAX_DIM] == 0)
    np.testing.assert_allclose(mu.numpy(), x.double().mean(0).numpy(), rtol=1e-6)
    spread = float((x.double() - mu.double()).abs().max())
    assert np.log2(s) == round(np.log2(s)) and 64.0 <= s * spread * (1 - 1e-6) and s * spread < 128.0 * (1 + 1e-6)
    # non-finite rows switch the transform off (the prep's range flag then routes the call)
    xd2 = xd.clone()
    xd2[n // 2, dim // 2] = float("inf")
    a2 = xb.column_affine(xd2).cpu()
    assert float(a2[_lib.SCREEN_MAX_DIM + 1]) == 0.0 and float(a2[_lib.SCREEN_MAX_DIM]) == 1.0
    xd2[n // 2, dim // 2] = float("nan")
    assert float(xb.column_affine(xd2).cpu()[_lib.SCREEN_MAX_DIM + 1]) == 0.0


@pytest.mark.parametrize("n,dim,offset,scale", [(1000, 512, 1000.0, 1.0), (777, 50, -3e4, 1e5),
                                                (640, 130, 0.0, 1e-7)])
def test_affine_operand_is_the_rounded_transform_with_a_rigorous_error_bound(n, dim, offset, scale):
    """exb_knn_prep_affine_f16: operand = fp16(s * fl(x - mu)); round_err bounds the distance of
    the fp16 row from the EXACT s * (x - mu) (fp64), including the fp32 subtraction rounding."""
    from exorl_b200 import _lib, rep as xrep
    x = reps(17, n, dim, "raw") * scale + offset
    xd = x.to(DEV)
    pre = xrep.prepare_rep(xd, affine="auto")
    aff = pre.affine.cpu()
    mu, s = aff[:dim], float(aff[_lib.SCREEN_MAX_DIM])
    assert float(aff[_lib.SCREEN_MAX_DIM + 1]) == 1.0
    y32 = (x - mu) * s                                   # fp32, same two roundings as the kernel
    assert torch.equal(pre.f16[:, :dim].cpu(), y32.half())
    exact = (x.double() - mu.double()) * s
    want_er = ((exact - y32.half().double()) ** 2).sum(1).sqrt()
    got_er = pre.round_err.cpu().double()
    assert torch.all(got_er >= want_er)
    assert torch.all(got_er <= want_er * 1.01 + 1e-3 * float(want_er.max()))
    assert int(pre.scalars.cpu()[2]) == 0
    want_hn = 0.5 * (y32.half().double() ** 2).sum(1)
    assert torch.all((pre.half_norm.cpu().double() - want_hn).abs() <= want_hn * 2.0 ** -19)


@pytest.mark.parametrize("offset,scale", [(0.0, 1e5), (1000.0, 1.0), (0.0, 1e-6), (-5e6, 300.0)])
def test_dataset_scale_screen_is_range_proof(offset, scale):
    """Reps with a large common offset, beyond the fp16 range or far below it: the dataset-scale
    screened call centres and scales its operands, so NO row leaves the tensor-core
    path (0 undecided rows, no host repair) and the result is the oracle's."""
    n, dim, k = 12000, 64, 12
    x = reps(23, n, dim, "raw") * scale + offset
    xd = x.to(DEV)
    assert xknn._streaming(n, n, dim, k)
    dist, idx = xb.knn_topk(xd, xd, k, mode="screened")
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV))
    assert und == 0 and invalid == 0
    sl = slice(0, 1500)
    assert_knn_matches(x[sl], x, k, dist[sl], idx[sl])
    de, ie = xb.knn_topk(xd, xd, k, mode="exact")
    np.testing.assert_allclose(dist.cpu().numpy(), de.cpu().numpy(), rtol=1e-6)
    assert (idx != ie).any(dim=1).sum().item() <= 4          # fp32 near-ties only
    # source != target takes the statistics of both row sets
    q = x[:3000] * 1.0 + 0.25 * scale
    d2, i2 = xb.knn_topk(q.to(DEV), xd, k, mode="screened")
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV))
    assert und <= 15 and invalid == 0      # a few rows with > 24 near-equidistant candidates
    assert_knn_matches(q[:800], x, k, d2[:800], i2[:800])


def test_prepared_operands_must_share_one_affine_record():
    from exorl_b200 import rep as xrep
    x = (reps(5, 9000, 64, "raw") + 500.0).to(DEV)
    y = (reps(6, 9000, 64, "raw") + 500.0).to(DEV)
    px = xrep.prepare_rep(x, affine="auto")
    # a raw second operand is prepared with px's record
    d1, i1 = xb.knn_topk(y, px, 5, mode="screened")
    d0, i0 = xb.knn_topk(y, x, 5, mode="exact")
    np.testing.assert_allclose(d1.cpu().numpy(), d0.cpu().numpy(), rtol=1e-6)
    py = xrep.prepare_rep(y, affine="auto")
    with pytest.raises(ValueError, match="affine"):
        xb.knn_topk(py, px, 5)
    with pytest.raises(ValueError, match="affine"):
        xb.knn_topk(xrep.prepare_rep(y), px, 5)


@pytest.mark.parametrize("block_scan", [True, False])
@pytest.mark.parametrize("n,nb,dim,k,scale", [(1024, 1024, 512, 12, 1.0), (300, 700, 50, 5, 1e5),
                                              (257, 1024, 130, 12, 1e6), (640, 640, 64, 3, 1.0)])
def test_per_update_block_cooperative_scan(n, nb, dim, k, scale, block_scan):
    """Every row of a per-update batch undecided (a dense random walk: all pairs inside the fp16
    margin; or values outside the fp16 range).  set_per_update_scan(True): the rerank blocks scan
    the bank ranges their rows lost TOGETHER through shared memory (coop_scan); False (default):
    every row's own warp does.  Result = the exact path's either way, vector and scalar row
    layouts, source != bank, ragged sizes."""
    prev = xknn.set_per_update_scan(block_scan)
    try:
        _check_all_rows_undecided(n, nb, dim, k, scale)
    finally:
        xknn.set_per_update_scan(prev)


def _check_all_rows_undecided(n, nb, dim, k, scale):
    g = torch.Generator().manual_seed(n + dim)
    if scale == 1.0:
        walk = torch.randn(1, 24, generator=g) + torch.cumsum(
            0.005 * torch.randn(max(n, nb), 24, generator=g), 0)
        w = torch.randn(24, dim, generator=g) / 24 ** 0.5
        x = torch.tanh(torch.nn.functional.layer_norm(walk @ w, (dim,)))
    else:
        x = torch.randn(max(n, nb), dim, generator=g) * scale
    src, bank = x[:n].contiguous(), x[:nb].contiguous()
    sd, bd = src.to(DEV), bank.to(DEV)
    if n == nb:
        bd = sd
    assert not xknn._streaming(n, nb, dim, k)
    dist, idx = xb.knn_topk(sd, bd, k, mode="screened")
    und, invalid, _, _ = xknn.screened_status(torch.device(DEV))
    assert invalid == 0 and und >= n // 2
    de, ie = xb.knn_topk(sd, bd, k, mode="exact")
    np.testing.assert_allclose(dist.cpu().numpy(), de.cpu().numpy(), rtol=1e-6)
    assert_knn_matches(src, bank, k, dist, idx)
