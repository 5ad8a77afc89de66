"""GPU parity of the replay gather: bit-exact tuples vs the reference's ReplayBuffer._sample /
OfflineReplayBuffer._sample (golden fixtures) and vs the oracle on boundary positions."""
import pathlib
import random
import types

import numpy as np
import pytest
import torch

import exorl_b200 as xb
from oracle import (nstep_transition, onestep_transition, write_episode_npz,
                    draw_indices_philox, flatten_episodes)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _episodes(g, tag):
    lens = g[f"{tag}_lens"]
    names = sorted({k.split("_", 3)[3] for k in g.files if k.startswith(f"{tag}_ep0_")})
    return [{n: g[f"{tag}_ep{e}_{n}"] for n in names} for e in range(len(lens))]


def _buffer(eps, nstep, meta=(), max_size=10 ** 6):
    buf = xb.DeviceReplayBuffer(DEV, max_size, nstep, 0.99, meta_names=meta)
    for e, ep in enumerate(eps):
        buf.add_episode(ep, key=f"episode_{e}_{len(ep['reward']) - 1}.npz")
    return buf


@pytest.mark.parametrize("tag", ["state_n3", "state_n1", "meta_n3", "pix_n3", "state_n5"])
def test_golden_nstep_bit_exact(golden, tag):
    g = golden("replay_reference.npz")
    eps = _episodes(g, tag)
    nstep = int(g[f"{tag}_nstep"])
    meta = ("skill", "done") if tag.startswith("meta") else ()
    buf = _buffer(eps, nstep, meta)
    out = buf.gather(buf.rows_from_draws(g[f"{tag}_draws"]))
    torch.cuda.synchronize()
    assert len(out) == 5 + len(meta)
    for j, t in enumerate(out):
        want = g[f"{tag}_out{j}"]
        got = t.cpu().numpy()
        assert got.dtype == want.dtype and got.shape == want.shape, (tag, j)
        assert np.array_equal(got, want), (tag, j)


def test_golden_offline_bit_exact(golden):
    g = golden("replay_reference.npz")
    buf = _buffer(_episodes(g, "state_n3"), 1)
    out = buf.gather(buf.rows_from_draws(g["offline_draws"]), offline=True)
    for j, t in enumerate(out):
        assert np.array_equal(t.cpu().numpy(), g[f"offline_out{j}"]), j


def _synth(rng, L, obs_shape=(24,), dtype=np.float32, terminal=False):
    n = L + 1
    obs = (rng.integers(0, 256, (n, *obs_shape), dtype=np.uint8) if dtype == np.uint8
           else rng.standard_normal((n, *obs_shape)).astype(np.float32))
    d = np.ones((n, 1), np.float32)
    if terminal:
        d[-1] = 0
        d[L // 2] = 0.5
    return {"observation": obs, "action": rng.uniform(-1, 1, (n, 6)).astype(np.float32),
            "reward": rng.uniform(0, 1, (n, 1)).astype(np.float32), "discount": d}


@pytest.mark.parametrize("nstep", [1, 3, 7])
def test_boundaries_and_terminal_discounts_vs_oracle(nstep):
    rng = np.random.default_rng(0)
    eps = [_synth(rng, L, terminal=(i % 2 == 0)) for i, L in enumerate((nstep, 9, 30, nstep + 1))]
    buf = _buffer(eps, nstep)
    draws = []
    for e, ep in enumerate(eps):
        L = len(ep["reward"]) - 1
        draws += [(e, 1), (e, L - nstep + 1)] + [(e, i) for i in range(1, L - nstep + 2)]
    out = buf.gather(buf.rows_from_draws(draws))
    for s, (e, idx) in enumerate(draws):
        want = nstep_transition(eps[e], idx, nstep, 0.99)
        for j in range(5):
            assert np.array_equal(out[j][s].cpu().numpy(), np.asarray(want[j])), (e, idx, j)


def test_pixel_rows_large_vectorised_copy():
    rng = np.random.default_rng(4)
    eps = [_synth(rng, 12, (9, 84, 84), np.uint8) for _ in range(3)]
    buf = _buffer(eps, 3)
    draws = [(e, i) for e in range(3) for i in (1, 5, 10)]
    out = buf.gather(buf.rows_from_draws(draws))
    assert out[0].dtype == torch.uint8 and out[0].shape == (9, 9, 84, 84)
    for s, (e, idx) in enumerate(draws):
        want = nstep_transition(eps[e], idx, 3, 0.99)
        assert np.array_equal(out[0][s].cpu().numpy(), want[0])
        assert np.array_equal(out[4][s].cpu().numpy(), want[4])


def test_unaligned_rows_take_narrow_vector_paths():
    rng = np.random.default_rng(5)
    eps = []
    for _ in range(2):
        ep = _synth(rng, 10, (7,))                       # 28-byte obs rows
        ep["action"] = rng.uniform(-1, 1, (11, 3)).astype(np.float32)   # 12-byte rows
        ep["flag"] = rng.integers(0, 2, (11, 5)).astype(np.uint8)       # 5-byte rows
        eps.append(ep)
    buf = _buffer(eps, 2, meta=("flag",))
    draws = [(0, 1), (1, 9), (0, 4)]
    out = buf.gather(buf.rows_from_draws(draws))
    for s, (e, idx) in enumerate(draws):
        want = nstep_transition(eps[e], idx, 2, 0.99, ("flag",))
        for j in range(6):
            assert np.array_equal(out[j][s].cpu().numpy(), np.asarray(want[j]))


def test_philox_draw_matches_oracle_and_ranges():
    rng = np.random.default_rng(7)
    lens = [5, 40, 13, 1000, 8]
    eps = [_synth(rng, L) for L in lens]
    buf = _buffer(eps, 3)
    sorted_eps, starts, elen = buf._tables()
    rows = buf.draw_rows_philox(4096, seed=1234567890123, call=42).cpu().numpy()
    _, _, want = draw_indices_philox(1234567890123, 42, 4096, starts.cpu().numpy(),
                                     elen.cpu().numpy(), 3)
    assert np.array_equal(rows, want)
    # every draw is a valid start: idx in [1, L-nstep+1]
    st, ln = starts.cpu().numpy(), elen.cpu().numpy()
    ep = np.searchsorted(np.sort(st), rows, side="right") - 1
    order = np.argsort(st)
    idx = rows - st[order][ep]
    assert idx.min() >= 1 and np.all(idx <= ln[order][ep] - 3 + 1)
    # uniform over episodes (not over transitions)
    counts = np.bincount(ep, minlength=5) / 4096
    assert np.all(np.abs(counts - 0.2) < 0.03)


def test_loader_dropin_signature_and_reference_rng_parity(tmp_path):
    """make_replay_loader with the reference's positional signature; with sampler
    'reference_rng' the stream of batches equals the reference algorithm's single-process stream
    for the same seeds (oracle arithmetic on the reference's own RNG call order)."""
    rng = np.random.default_rng(3)
    lens = [20, 7, 33, 12]
    eps = []
    for e, L in enumerate(lens):
        ep = _synth(rng, L, terminal=(e == 1))
        ep["skill"] = rng.standard_normal((L + 1, 4)).astype(np.float32)
        write_episode_npz(ep, tmp_path / f"episode_{e}_{L}.npz")
        eps.append(ep)
    storage = types.SimpleNamespace(_replay_dir=pathlib.Path(tmp_path),
                                    _meta_specs=(types.SimpleNamespace(name="skill"),))
    loader = xb.make_replay_loader(storage, 1000000, 16, 1, True, 3, 0.99, device=DEV,
                                   sampler="reference_rng")
    random.seed(11)
    np.random.seed(11)
    it = iter(loader)
    batches = [next(it) for _ in range(3)]
    random.seed(11)
    np.random.seed(11)
    order = sorted(range(4), key=lambda e: f"episode_{e}_{lens[e]}.npz")
    for batch in batches:
        assert len(batch) == 6 and all(t.is_cuda for t in batch)
        for s in range(16):
            e = random.choice(order)
            idx = np.random.randint(0, lens[e] - 3 + 1) + 1
            want = nstep_transition(eps[e], idx, 3, 0.99, ("skill",))
            for j in range(6):
                assert np.array_equal(batch[j][s].cpu().numpy(), np.asarray(want[j]))
    # philox loader with the reference's default of 4 workers (pretrain.yaml:33): yields forever,
    # right shapes/dtypes, every batch from one worker partition
    loader2 = xb.make_replay_loader(storage, 1000000, 1024, 4, True, 3, 0.99, device=DEV, seed=5)
    assert len(loader2.buffers) == 4
    obs, act, rew, disc, nobs, skill = next(iter(loader2))
    assert obs.shape == (1024, 24) and act.shape == (1024, 6) and rew.shape == (1024, 1)
    assert disc.shape == (1024, 1) and nobs.shape == (1024, 24) and skill.shape == (1024, 4)


def test_offline_loader_and_sorted_name_eviction(tmp_path):
    rng = np.random.default_rng(8)
    eps = []
    for e in range(6):
        ep = _synth(rng, 10)
        write_episode_npz(ep, tmp_path / f"episode_{e}_10.npz")
        eps.append(ep)
    loader = xb.make_offline_replay_loader(None, pathlib.Path(tmp_path), 1000, 32, 1, 0.99,
                                           device=DEV, seed=1)
    obs, act, rew, disc, nobs = next(iter(loader))
    assert obs.shape == (32, 24) and torch.all(disc == np.float32(0.99))
    # eviction by episode at max_size, first of the sorted names (utils/replay_buffer.py:178-185)
    buf = xb.DeviceReplayBuffer(DEV, 35, 3, 0.99)
    for e, ep in enumerate(eps):
        buf.add_episode(ep, key=f"episode_{e}_10.npz")
        assert len(buf) <= 35
    assert [x["key"] for x in buf.episodes] == [f"episode_{e}_10.npz" for e in (3, 4, 5)]
    out = buf.gather(buf.rows_from_draws([(0, 1), (2, 8)]))
    assert np.array_equal(out[0][0].cpu().numpy(), eps[3]["observation"][0])
    assert np.array_equal(out[4][1].cpu().numpy(), eps[5]["observation"][10])


@pytest.mark.parametrize("offline", [False, True])
def test_fused_sample_equals_draw_then_gather(offline):
    """exb_replay_sample (draw + gather in one launch) == exb_replay_draw + exb_replay_gather."""
    rng = np.random.default_rng(21)
    eps = []
    for L in (5, 40, 13, 300, 8):
        ep = _synth(rng, L, terminal=True)
        ep["skill"] = rng.standard_normal((L + 1, 4)).astype(np.float32)
        eps.append(ep)
    buf = _buffer(eps, 3, meta=("skill",))
    for batch, call in ((1, 0), (257, 3), (2048, 2 ** 33 + 5)):
        rows = buf.draw_rows_philox(batch, 987654321, call, offline=offline)
        want = buf.gather(rows, offline=offline)
        got, got_rows = buf.sample_philox(batch, 987654321, call, offline=offline,
                                          return_rows=True)
        assert torch.equal(got_rows, rows)
        assert len(got) == len(want)
        for a, b in zip(got, want):
            assert a.dtype == b.dtype and torch.equal(a, b)


@pytest.mark.parametrize("obs_shape,dtype", [((129 * 4,), np.float32), ((1024 * 4 + 4,), np.float32),
                                             ((3, 84, 84), np.uint8), ((4099,), np.uint8),
                                             ((2, 1031), np.float32)])
def test_long_rows_are_copied_in_pieces(obs_shape, dtype):
    """Rows longer than 128 vectors are split into 1024-vector pieces (one block each); sizes
    just above the short-row limit, just above one piece, ragged last pieces and odd byte counts
    (1-byte vector path) must all arrive intact."""
    rng = np.random.default_rng(22)
    eps = [_synth(rng, 9, obs_shape, dtype) for _ in range(3)]
    if dtype != np.uint8:
        for ep in eps:
            ep["observation"] = rng.standard_normal((10, *obs_shape)).astype(np.float32)
    buf = _buffer(eps, 2)
    draws = [(e, i) for e in range(3) for i in (1, 4, 8)]
    rows = buf.rows_from_draws(draws)
    out = buf.gather(rows)
    for s, (e, idx) in enumerate(draws):
        want = nstep_transition(eps[e], idx, 2, 0.99)
        for j in range(5):
            assert np.array_equal(out[j][s].cpu().numpy(), np.asarray(want[j])), (s, j)


def test_full_size_gather_properties():
    """BASELINE C4 shape at reduced depth: 1024-batch of (9,84,84) uint8 frames; every gathered
    frame equals a direct index of the device column (size-independent identity)."""
    n_rows = 4004
    g = torch.Generator(device=DEV).manual_seed(4)
    buf = xb.DeviceReplayBuffer(DEV, 4000, 3, 0.99, capacity_rows=n_rows)
    buf.cols["observation"] = torch.randint(0, 256, (n_rows, 9, 84, 84), generator=g,
                                            dtype=torch.uint8, device=DEV)
    buf.cols["action"] = torch.rand((n_rows, 6), generator=g, device=DEV)
    buf.cols["reward"] = torch.rand((n_rows, 1), generator=g, device=DEV)
    buf.cols["discount"] = torch.ones((n_rows, 1), device=DEV)
    buf.episodes = [{"key": f"episode_{e:06d}_1000.npz", "start": e * 1001, "len": 1000}
                    for e in range(4)]
    buf.size = 4000
    (obs, act, rew, disc, nobs), rows = buf.sample_philox(1024, 3, 9, return_rows=True)
    assert torch.equal(obs, buf.cols["observation"][rows - 1])
    assert torch.equal(nobs, buf.cols["observation"][rows + 2])
    assert torch.equal(act, buf.cols["action"][rows])
    r = buf.cols["reward"]
    g1 = torch.tensor(0.99, device=DEV)
    want = (r[rows] + g1 * r[rows + 1]) + (g1 * g1) * r[rows + 2]
    assert torch.equal(rew, want)
    assert torch.all(disc == np.float32(0.99) * np.float32(0.99) * np.float32(0.99))


def test_relabel_rewards_and_export_roundtrip(tmp_path):
    """§8f.4: rewards computed for every stored state are written back into the reward column
    (relabel_episode's in-memory effect, utils/replay_buffer.py:31-42), sampled n-step returns use
    them, and the exported files are episodes in the reference's format."""
    rng = np.random.default_rng(31)
    eps = [_synth(rng, L) for L in (6, 11, 9)]
    buf = _buffer(eps, 3)
    rows = buf.live_rows()
    assert rows.numel() == sum(L + 1 for L in (6, 11, 9))
    new_r = torch.arange(rows.numel(), dtype=torch.float32, device=DEV).reshape(-1, 1) * 0.5
    buf.relabel_rewards(new_r)
    # the same rewards laid over the episodes on the host
    off = 0
    relabelled = []
    for e in buf._tables()[0]:
        k = int(str(e["key"]).split("_")[1])
        ep = dict(eps[k])
        ep["reward"] = new_r[off:off + e["len"] + 1].cpu().numpy()
        off += e["len"] + 1
        relabelled.append((k, ep))
    draws = [(pos, 1) for pos in range(3)] + [(1, 5)]
    out = buf.gather(buf.rows_from_draws(draws))
    for s, (pos, idx) in enumerate(draws):
        want = nstep_transition(relabelled[pos][1], idx, 3, 0.99)
        for j in range(5):
            assert np.array_equal(out[j][s].cpu().numpy(), np.asarray(want[j])), (s, j)
    files = buf.export_episodes(tmp_path / "relabelled")
    assert sorted(f.name for f in files) == ["episode_0_6.npz", "episode_1_11.npz",
                                             "episode_2_9.npz"]
    from oracle import read_episode_npz
    for k, ep in relabelled:
        got = read_episode_npz(tmp_path / "relabelled" / f"episode_{k}_{len(ep['reward']) - 1}.npz")
        assert set(got) == set(ep)
        for name in ep:
            assert np.array_equal(got[name], ep[name]), (k, name)
    with pytest.raises(ValueError):
        buf.relabel_rewards(new_r[:-1])


def _residency_scenarios(kind):
    import json
    path = pathlib.Path(__file__).resolve().parent / "golden" / "residency_reference.json"
    return sorted(json.loads(path.read_text())[kind].items())


def _check_device_rows(buf, written):
    for e in buf.episodes:
        want = written[pathlib.Path(str(e["key"])).name]
        lo, hi = e["start"], e["start"] + e["len"] + 1
        for name, col in buf.cols.items():
            assert np.array_equal(col[lo:hi].cpu().numpy(), want[name]), (e["key"], name)


@pytest.mark.parametrize("name,sc", _residency_scenarios("online"))
def test_online_residency_matches_reference_fixture(tmp_path, name, sc):
    """make_replay_loader over a growing directory: resident files per worker partition, sizes
    and on-disk leftovers after every fetch round equal what the reference's ReplayBuffer did
    (fixture recorded by oracle/gen_golden.py:gen_residency; utils/replay_buffer.py:172-212)."""
    from oracle.gen_golden import residency_episode
    storage = types.SimpleNamespace(_replay_dir=pathlib.Path(tmp_path), _meta_specs=())
    loader = xb.make_replay_loader(storage, sc["max_size"], 1000, sc["num_workers"],
                                   sc["save_snapshot"], 3, 0.99, device=DEV, seed=3)
    written = {}
    for r, rnd in enumerate(sc["rounds"]):
        for eps_idx, eps_len in rnd["written"]:
            fn = f"episode_{eps_idx}_{eps_len}.npz"
            written[fn] = residency_episode(eps_idx, eps_len)
            write_episode_npz(written[fn], tmp_path / fn)
        # batch 1000 = fetch_every: every batch of a partition is preceded by a fetch
        batches = [next(iter(loader)) for _ in range(sc["num_workers"])]
        assert all(b[0].shape == (1000, 3) for b in batches)
        assert loader.resident_files() == rnd["resident"], (name, r)
        assert [b.size for b in loader.buffers] == rnd["size"], (name, r)
        assert sorted(f.name for f in tmp_path.glob("*.npz")) == rnd["on_disk"], (name, r)
        for b in loader.buffers:
            _check_device_rows(b, written)


@pytest.mark.parametrize("name,sc", _residency_scenarios("offline"))
def test_offline_residency_matches_reference_fixture(tmp_path, name, sc):
    """make_offline_replay_loader over a directory larger than the cap keeps the FIRST episodes
    (sorted names), overshooting the cap by at most one episode, like
    OfflineReplayBuffer._load (utils/replay_buffer.py:57-75)."""
    from oracle.gen_golden import residency_episode
    written = {}
    for eps_idx, eps_len in sc["files"]:
        fn = f"episode_{eps_idx}_{eps_len}.npz"
        written[fn] = residency_episode(eps_idx, eps_len)
        write_episode_npz(written[fn], tmp_path / fn)
    loader = xb.make_offline_replay_loader(None, pathlib.Path(tmp_path), sc["max_size"], 64,
                                           sc["num_workers"], 0.99, device=DEV, seed=1)
    it = iter(loader)
    for _ in range(sc["num_workers"]):
        obs, act, rew, disc, nobs = next(it)
        assert obs.shape == (64, 3)
    assert loader.resident_files() == sc["resident"]
    assert [b.size for b in loader.buffers] == sc["size"]
    assert len(list(tmp_path.glob("*.npz"))) == len(sc["files"])
    for b in loader.buffers:
        _check_device_rows(b, written)


def test_batches_of_a_partitioned_loader_come_from_one_partition_each(tmp_path):
    """num_workers = 2: batch i is drawn from partition i % 2, whose episodes are the files with
    eps_idx % 2 == partition (utils/replay_buffer.py:204-205 + DataLoader's round-robin)."""
    for e in range(6):
        ep = _synth(np.random.default_rng(e), 10)
        ep["observation"][:] = e            # the observation value names its episode
        write_episode_npz(ep, tmp_path / f"episode_{e}_10.npz")
    storage = types.SimpleNamespace(_replay_dir=pathlib.Path(tmp_path), _meta_specs=())
    loader = xb.make_replay_loader(storage, 1000, 256, 2, True, 3, 0.99, device=DEV, seed=9)
    it = iter(loader)
    for i in range(4):
        obs = next(it)[0]
        seen = set(obs[:, 0].cpu().numpy().astype(int).tolist())
        assert seen == {e for e in range(6) if e % 2 == i % 2}, (i, seen)


def test_episode_shorter_than_nstep_raises_like_the_reference():
    """np.random.randint(0, eps_len - nstep + 1) raises ValueError when an episode has no valid
    start (utils/replay_buffer.py:222); every sampler does the same before launching."""
    rng = np.random.default_rng(12)
    buf = _buffer([_synth(rng, 10), _synth(rng, 2)], 3)
    for call in (lambda: buf.sample_philox(8, 1, 0), lambda: buf.draw_rows_philox(8, 1, 0)):
        with pytest.raises(ValueError, match="low >= high"):
            call()
    random.seed(0)
    np.random.seed(0)
    with pytest.raises(ValueError):
        buf.draw_rows_reference_rng(64)
    # the same episodes are fine for a one-step (offline) draw and for nstep 2
    out = buf.sample_philox(8, 1, 0, offline=True)
    assert out[0].shape == (8, 24)
    ok = _buffer([_synth(rng, 10), _synth(rng, 2)], 2)
    (obs, act, rew, disc, nobs), rows = ok.sample_philox(512, 1, 0, return_rows=True)
    st = {e["start"]: e["len"] for e in ok.episodes}
    short_start = next(s for s, L in st.items() if L == 2)
    in_short = (rows >= short_start) & (rows <= short_start + 2)
    assert in_short.any() and torch.all(rows[in_short] == short_start + 1)


def test_store_growth_and_compaction_on_device():
    """The device columns grow from a small first allocation and compact when fragmented;
    sampled transitions still equal the oracle's afterwards."""
    rng = np.random.default_rng(13)
    buf = xb.DeviceReplayBuffer(DEV, 2000, 3, 0.99)
    eps = {}
    cap0 = None
    for i in rng.permutation(200):
        L = int(rng.choice((5, 12, 40)))
        key = f"episode_{i:03d}_{L}.npz"
        eps[key] = _synth(rng, L)
        buf.add_episode(eps[key], key=key)
        cap0 = cap0 or buf.capacity_rows
    assert buf.capacity_rows > cap0 and 1900 < buf.size <= 2000
    sorted_eps, _, _ = buf._tables()
    draws = [(pos, 1) for pos in range(len(sorted_eps))]
    out = buf.gather(buf.rows_from_draws(draws))
    for s, (pos, idx) in enumerate(draws):
        want = nstep_transition(eps[sorted_eps[pos]["key"]], idx, 3, 0.99)
        for j in range(5):
            assert np.array_equal(out[j][s].cpu().numpy(), np.asarray(want[j]))
