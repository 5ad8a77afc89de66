"""Pin the CPU oracle (oracle/) to the unmodified reference.

The fixtures in tests/golden/ were produced by oracle/gen_golden.py running
/root/reference/utils/utils.py and /root/reference/utils/replay_buffer.py in the authoring
container; the hard-coded numbers are the known-answer vectors of SURVEY.md §8c.
"""
import numpy as np
import pytest
import torch

from oracle import (RmsState, rms_step, pbe_reward, knn_smallest, proto_reward, proto_enqueue,
                    proto_candidates, nstep_transition, onestep_transition)

REP4 = torch.tensor([[0., 0.], [3., 4.], [6., 8.], [0., 1.]])


def _run(avg, rms_on, clip, calls=1, rep=REP4, k=2):
    st = RmsState()
    outs = [pbe_reward(rep, rep, st, k, avg, rms_on, clip).flatten().tolist()
            for _ in range(calls)]
    return outs, st


def test_kat_distances():
    d, i = knn_smallest(REP4, REP4, 4)
    want = torch.tensor([[0, 1, 5, 10], [0, 4.2426405, 5, 5], [0, 5, 9.2195444, 10],
                         [0, 1, 4.2426405, 9.2195444]])
    assert torch.allclose(d, want, rtol=1e-7)
    assert i[:, 0].tolist() == [0, 1, 2, 3]


@pytest.mark.parametrize("avg,rms_on,clip,want", [
    (True, True, 0.0, [0.30438635, 0.92008394, 1.02207816, 0.30438635]),
    (True, True, 0.5, [0.10056251, 0.81514400, 0.92780966, 0.10056251]),
    (True, False, 0.0, [0.40546510, 1.13825607, 1.25276291, 0.40546510]),
    (True, False, 0.5, [0.22314355, 1.05477190, 1.17865503, 0.22314355]),
    (False, True, 0.0, [0.30438960, 0.92009133, 1.02208614, 0.30438960]),
    (False, True, 0.5, [0.0, 0.69789571, 0.82373059, 0.0]),
    (False, False, 0.0, [0.69314718, 1.65682530, 1.79175949, 0.69314718]),
])
def test_kat_rewards_survey(avg, rms_on, clip, want):
    outs, _ = _run(avg, rms_on, clip)
    np.testing.assert_allclose(outs[0], want, rtol=2e-7, atol=1e-8)


def test_kat_state_trajectory_survey():
    outs, st = _run(True, True, 0.0, calls=2)
    np.testing.assert_allclose(outs[1], [0.30438468, 0.92008013, 1.02207410, 0.30438468],
                               rtol=2e-7)
    assert abs(float(st.M) - 1.40532124) < 2e-7 and abs(float(st.S) - 4.17147541) < 1e-6
    assert st.n == pytest.approx(16.0001)
    outs, st = _run(False, True, 0.0, calls=2)
    assert abs(float(st.M) - 2.81062484) < 5e-7 and abs(float(st.S) - 4.46697426) < 1e-6
    assert st.n == pytest.approx(8.0001)


def test_kat_rms_alone_survey():
    st = RmsState()
    rms_step(st, torch.tensor([[1.], [2.], [3.], [6.]]))
    assert float(st.M) == pytest.approx(2.99992490, abs=3e-7)
    assert float(st.S) == pytest.approx(4.66679955, abs=1e-6)
    rms_step(st, torch.tensor([[10.], [20.]]))
    assert float(st.M) == pytest.approx(6.99988317, abs=1e-6)
    assert float(st.S) == pytest.approx(51.77775192, abs=1e-5)
    assert st.n == pytest.approx(6.0001)


def test_clip_negative_equals_clip_zero_on_nonnegative_data():
    a, _ = _run(True, True, -1.0)
    b, _ = _run(True, True, 0.0)
    assert a == b


def test_k_larger_than_bank_raises():
    with pytest.raises(RuntimeError):
        knn_smallest(REP4, REP4, 5)


def test_golden_kat_file_bitwise(golden):
    g = golden("pbe_kat.npz")
    rep = torch.from_numpy(g["rep"])
    modes = {"avg_rms_c0": (True, True, 0.0), "avg_rms_c05": (True, True, 0.5),
             "avg_norms_c0": (True, False, 0.0), "avg_norms_c05": (True, False, 0.5),
             "kth_rms_c0": (False, True, 0.0), "kth_rms_c05": (False, True, 0.5),
             "kth_norms_c0": (False, False, 0.0), "avg_rms_cneg": (True, True, -1.0)}
    for name, (avg, rms_on, clip) in modes.items():
        st = RmsState()
        for call in range(2):
            r = pbe_reward(rep, rep, st, 2, avg, rms_on, clip)
            assert np.array_equal(r.numpy(), g[f"{name}_r{call}"]), (name, call)
            s = g[f"{name}_s{call}"]
            assert float(st.M) == s[0] and float(st.S) == s[1] and st.n == s[2]


def test_golden_pbe_all_modes_bitwise(golden):
    """144 (shape, k, avg, rms, clip) cases x 3 sequential calls: oracle == reference bitwise."""
    g = golden("pbe_reference.npz")
    for c in range(int(g["n_cases"])):
        tag = f"pbe{c:03d}"
        b, d, k, avg, rms_on = (int(v) for v in g[f"{tag}_cfg"])
        clip = float(g[f"{tag}_clip"])
        st = RmsState()
        for call in range(3):
            x = torch.from_numpy(g[f"rep_{b}_{d}_{call}"])
            r = pbe_reward(x, x, st, k, bool(avg), bool(rms_on), clip)
            assert np.array_equal(r.numpy(), g[f"{tag}_reward{call}"], equal_nan=True), (tag, call)
            s = g[f"{tag}_state{call}"]
            assert (float(st.M), float(st.S), st.n) == (s[0], s[1], s[2]), (tag, call)


def test_golden_knn_indices_and_distances(golden):
    g = golden("knn_reference.npz")
    for tag in ("sq512", "rect", "proto", "tiny"):
        s, t = torch.from_numpy(g[f"{tag}_src"]), torch.from_numpy(g[f"{tag}_tgt"])
        k = g[f"{tag}_idx"].shape[1]
        d, i = knn_smallest(s, t, k, q_tile=32, t_tile=48)   # ragged tiles on purpose
        assert np.array_equal(d.numpy(), g[f"{tag}_dist"]), tag
        assert np.array_equal(i.numpy(), g[f"{tag}_idx"]), tag


def test_golden_proto(golden):
    g = golden("proto_reference.npz")
    z = torch.from_numpy(g["z"])
    full = torch.from_numpy(g["queue_full"])
    for tag, q in (("full", full.clone()), ("warm", torch.cat([full[:64], torch.zeros(192, 128)]))):
        ptr = proto_enqueue(q, 0, z[torch.arange(0, 32)])
        assert ptr == int(g[f"{tag}_ptr"])
        assert np.array_equal(q.numpy(), g[f"{tag}_queue_after"])
        assert np.array_equal(proto_reward(z, q, 3).numpy(), g[f"{tag}_reward"])


def _episodes(g, tag):
    lens = g[f"{tag}_lens"]
    names = sorted({k.split("_", 3)[3] for k in g.files if k.startswith(f"{tag}_ep0_")})
    return [{n: g[f"{tag}_ep{e}_{n}"] for n in names} for e in range(len(lens))]


@pytest.mark.parametrize("tag", ["state_n3", "state_n1", "meta_n3", "pix_n3", "state_n5"])
def test_golden_replay_nstep_bitwise(golden, tag):
    g = golden("replay_reference.npz")
    eps = _episodes(g, tag)
    nstep = int(g[f"{tag}_nstep"])
    meta = ("skill", "done") if tag.startswith("meta") else ()
    for s, (e, idx) in enumerate(g[f"{tag}_draws"]):
        got = nstep_transition(eps[e], int(idx), nstep, 0.99, meta)
        for j, v in enumerate(got):
            want = g[f"{tag}_out{j}"][s]
            assert np.array_equal(np.asarray(v), want) and np.asarray(v).dtype == want.dtype


def test_golden_replay_offline_bitwise(golden):
    g = golden("replay_reference.npz")
    eps = _episodes(g, "state_n3")
    for s, (e, idx) in enumerate(g["offline_draws"]):
        got = onestep_transition(eps[e], int(idx), 0.99)
        for j, v in enumerate(got):
            assert np.array_equal(np.asarray(v), g[f"offline_out{j}"][s])


def test_nstep_discount_kat():
    ep = {"observation": np.zeros((5, 2), np.float32), "action": np.zeros((5, 1), np.float32),
          "reward": np.ones((5, 1), np.float32), "discount": np.ones((5, 1), np.float32)}
    out = nstep_transition(ep, 1, 3, 0.99)
    assert out[3].dtype == np.float32 and float(out[3][0]) == pytest.approx(0.97029907, abs=1e-8)


@pytest.mark.parametrize("tag", ["walker", "small", "wide", "odd"])
def test_golden_icm_trunk_rep_bitwise(golden, tag):
    """oracle.icm_trunk_rep == the reference's own ICM.get_rep (icm_apt.py:17-18,50-52) and the
    reference PBE on that rep == oracle.pbe_reward (fixtures: gen_golden.py:gen_rep)."""
    from oracle import icm_trunk_rep
    g = golden("rep_reference.npz")
    t = {n: torch.from_numpy(g[f"{tag}_{n}"]) for n in ("obs", "w", "b", "lnw", "lnb")}
    rep = icm_trunk_rep(t["obs"], t["w"], t["b"], t["lnw"], t["lnb"], float(g[f"{tag}_eps"]))
    assert np.array_equal(rep.numpy(), g[f"{tag}_rep"])
    r = pbe_reward(rep, rep, RmsState(), 12, True, True, 0.0)
    assert np.array_equal(r.numpy(), g[f"{tag}_reward"])


def test_golden_l2_normalize_bitwise(golden):
    from oracle import l2_normalize
    g = golden("rep_reference.npz")
    out = l2_normalize(torch.from_numpy(g["norm_x"]))
    assert np.array_equal(out.numpy(), g["norm_out"])
    assert not np.any(out[5].numpy())   # zero row stays zero (eps clamp, no NaN)


@pytest.mark.parametrize("tag", ["walker", "odd"])
def test_golden_rnd_reward_bitwise(golden, tag):
    """oracle.rnd_reward == the reference's RND.forward error (rnd.py:52-54) + the
    compute_intr_reward lines (:95-98) with utils.RMS, three sequential calls."""
    from oracle import rnd_reward
    g = golden("rnd_reference.npz")
    st = RmsState()
    for call in range(3):
        r, e = rnd_reward(torch.from_numpy(g[f"{tag}_pred{call}"]),
                          torch.from_numpy(g[f"{tag}_target{call}"]), st,
                          float(g[f"{tag}_scale"]))
        assert np.array_equal(e.numpy(), g[f"{tag}_error{call}"])
        assert np.array_equal(r.numpy(), g[f"{tag}_reward{call}"])
        m, s, n = g[f"{tag}_state{call}"]
        assert float(st.M) == m and float(st.S) == s and st.n == n


@pytest.mark.parametrize("tag", ["small", "peaked", "odd"])
def test_proto_candidate_sampling_pinned_to_reference(golden, tag):
    """oracle.proto_candidates (softmax -> Categorical renormalisation -> exponential race on the
    recorded noise) returns the candidates the reference's proto.py:105-107 sampled, bitwise equal
    probabilities included, and the queue write of :110-112 gives the recorded queue."""
    g = golden("proto_sampling_reference.npz")
    z, w = torch.from_numpy(g[f"{tag}_z"]), torch.from_numpy(g[f"{tag}_w"])
    cand, prob = proto_candidates(z, w, torch.from_numpy(g[f"{tag}_noise"]))
    assert np.array_equal(cand.numpy(), g[f"{tag}_candidates"])
    assert np.array_equal(prob.numpy(), g[f"{tag}_prob"])
    queue = torch.from_numpy(g[f"{tag}_queue_before"]).clone()
    ptr, new_ptr = (int(v) for v in g[f"{tag}_ptr"])
    assert proto_enqueue(queue, ptr, z[cand]) == new_ptr
    assert np.array_equal(queue.numpy(), g[f"{tag}_queue_after"])
