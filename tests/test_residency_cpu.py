"""Replay residency on the host: which episode files are resident, in which
order, what is evicted/unlinked and what stays on disk — against fixtures recorded from the
UNMODIFIED reference (`oracle/gen_golden.py:gen_residency`: ReplayBuffer._try_fetch /
_store_episode, utils/replay_buffer.py:172-212, and OfflineReplayBuffer._load, :57-75).

The loader logic runs here without a GPU over a numpy-backed `EpisodeStore` (the same class
`DeviceReplayBuffer` derives from), which also checks that rows survive growth and compaction.
tests/test_gpu_replay.py repeats the scenarios through make_replay_loader on the device."""
import json
import pathlib
import random

import numpy as np
import pytest

from exorl_b200.replay import DeviceReplayLoader
from exorl_b200.residency import EpisodeIndex, EpisodeStore, RowAllocator
from oracle import write_episode_npz
from oracle.gen_golden import residency_episode

GOLDEN = pathlib.Path(__file__).resolve().parent / "golden" / "residency_reference.json"


class NumpyStore(EpisodeStore):
    """EpisodeStore over host arrays: the CPU stand-in for the device columns."""

    def __init__(self, max_size, capacity_rows=None):
        super().__init__(max_size, capacity_rows)
        self.cols = {}
        self.grow_calls = 0
        self.moves = 0

    @property
    def _has_columns(self):
        return bool(self.cols)

    def _alloc_columns(self, episode, capacity):
        for name, a in episode.items():
            a = np.asarray(a)
            self.cols[name] = np.full((capacity, *a.shape[1:]), 0, a.dtype)

    def _grow_columns(self, capacity):
        self.grow_calls += 1
        for name, old in self.cols.items():
            new = np.zeros((capacity, *old.shape[1:]), old.dtype)
            new[:old.shape[0]] = old
            self.cols[name] = new

    def _move_rows(self, old, new, rows):
        self.moves += 1
        for col in self.cols.values():
            col[new:new + rows] = col[old:old + rows].copy()

    def _write_rows(self, start, episode, rows):
        for name, col in self.cols.items():
            col[start:start + rows] = np.asarray(episode[name])

    def check_rows(self, episodes_by_name):
        """Every resident episode's rows still hold that episode's arrays; ranges are disjoint."""
        taken = np.zeros(self._alloc.capacity, bool)
        for e in self.episodes:
            lo, hi = e["start"], e["start"] + e["len"] + 1
            assert not taken[lo:hi].any(), "row ranges overlap"
            taken[lo:hi] = True
            want = episodes_by_name[pathlib.Path(str(e["key"])).name]
            for name, col in self.cols.items():
                assert np.array_equal(col[lo:hi], want[name]), (e["key"], name)
        assert taken.sum() + self._alloc.free_rows == self._alloc.capacity


def _scenarios(kind):
    return sorted(json.loads(GOLDEN.read_text())[kind].items())


@pytest.mark.parametrize("name,sc", _scenarios("online"))
def test_online_fetch_eviction_unlink_match_reference(tmp_path, name, sc):
    workers = sc["num_workers"]
    stores = [NumpyStore(sc["max_size"] // workers) for _ in range(workers)]
    loader = DeviceReplayLoader(stores, 1000, replay_dir=tmp_path, offline=False,
                                save_snapshot=sc["save_snapshot"], fetch_every=1000)
    written = {}
    for r, rnd in enumerate(sc["rounds"]):
        for eps_idx, eps_len in rnd["written"]:
            fn = f"episode_{eps_idx}_{eps_len}.npz"
            written[fn] = residency_episode(eps_idx, eps_len)
            write_episode_npz(written[fn], tmp_path / fn)
        for w in loader.workers:
            w.since_fetch = loader.fetch_every
            loader._try_fetch(w)
        assert loader.resident_files() == rnd["resident"], (name, r)
        assert [s.size for s in stores] == rnd["size"], (name, r)
        assert sorted(f.name for f in tmp_path.glob("*.npz")) == rnd["on_disk"], (name, r)
        for s in stores:
            s.check_rows(written)


@pytest.mark.parametrize("name,sc", _scenarios("offline"))
def test_offline_load_keeps_first_episodes_like_reference(tmp_path, name, sc):
    workers = sc["num_workers"]
    written = {}
    for eps_idx, eps_len in sc["files"]:
        fn = f"episode_{eps_idx}_{eps_len}.npz"
        written[fn] = residency_episode(eps_idx, eps_len)
        write_episode_npz(written[fn], tmp_path / fn)
    stores = [NumpyStore(sc["max_size"] // workers) for _ in range(workers)]
    loader = DeviceReplayLoader(stores, 32, replay_dir=tmp_path, offline=True)
    for w in loader.workers:
        loader._load_offline(w)
    assert loader.resident_files() == sc["resident"]
    assert [s.size for s in stores] == sc["size"]
    assert len(list(tmp_path.glob("*.npz"))) == len(sc["files"])      # offline never unlinks
    for s in stores:
        s.check_rows(written)
        assert s.grow_calls == 0, "the offline load sizes its columns once, from the file names"


def test_damaged_file_ends_the_fetch_like_a_failed_load_episode(tmp_path):
    """utils/replay_buffer.py:173-176,211-212: a file that cannot be loaded stops the fetch; the
    files scanned before it (newer names) are stored, older ones wait for the next fetch."""
    eps = {}
    for e in range(4):
        fn = f"episode_{e}_10.npz"
        eps[fn] = residency_episode(e, 10)
        write_episode_npz(eps[fn], tmp_path / fn)
    (tmp_path / "episode_1_10.npz").write_bytes(b"PK\x03\x04 not a zip archive")
    store = NumpyStore(1000)
    loader = DeviceReplayLoader([store], 16, replay_dir=tmp_path, fetch_every=1000)
    loader._try_fetch(loader.workers[0])
    assert loader.resident_files() == [["episode_2_10.npz", "episode_3_10.npz"]]
    write_episode_npz(eps["episode_1_10.npz"], tmp_path / "episode_1_10.npz")
    loader.workers[0].since_fetch = 1000
    loader._try_fetch(loader.workers[0])      # 3 is resident -> the scan stops at once (:206)
    assert loader.resident_files() == [["episode_2_10.npz", "episode_3_10.npz"]]


def test_row_allocator_against_a_bitmap_model():
    rng = random.Random(0)
    cap = 257
    alloc, used, live = RowAllocator(cap), np.zeros(cap, bool), []
    for step in range(4000):
        if live and (rng.random() < 0.45 or alloc.free_rows < 8):
            s, n = live.pop(rng.randrange(len(live)))
            alloc.release(s, n)
            used[s:s + n] = False
        else:
            n = rng.randint(1, 40)
            s = alloc.alloc(n)
            if s is None:
                runs = "".join("1" if u else "0" for u in used).split("1")
                assert max(len(r) for r in runs) < n, "first fit missed a free run"
                if alloc.free_rows >= n and rng.random() < 0.5:
                    moves = alloc.compact(live)
                    placed = {old: new for old, new, _ in moves}
                    live = [(placed.get(s0, s0), n0) for s0, n0 in live]
                    used[:] = False
                    for s0, n0 in live:
                        assert not used[s0:s0 + n0].any()
                        used[s0:s0 + n0] = True
                    assert all(new < old for old, new, _ in moves)
                    assert moves == sorted(moves)
                continue
            assert not used[s:s + n].any()
            used[s:s + n] = True
            live.append((s, n))
        assert alloc.free_rows == cap - used.sum()
        free = alloc._free
        assert free == sorted(free) and all(a + n < b for (a, n), (b, _) in zip(free, free[1:]))


def test_store_grows_compacts_and_keeps_rows_intact():
    """Variable-length episodes, sorted-name eviction from the middle of the address space,
    fragmentation -> compaction, growth beyond the first estimate: rows stay intact and the
    transition count obeys the reference's rule `eps_len + size <= max_size` after each store."""
    rng = random.Random(1)
    store = NumpyStore(400)
    eps = {}
    order = list(range(120))
    rng.shuffle(order)                    # names arrive out of order: evictions hit random rows
    for i in order:
        L = rng.choice((3, 10, 25, 60))
        fn = f"episode_{i:04d}_{L}.npz"
        eps[fn] = residency_episode(i, L)
        model = EpisodeIndex(400)         # what the reference's list would do
        for e in store.episodes:
            model.insert(e["key"], e["start"], e["len"])
        gone = [e["key"] for e in model.evictions_for(L)]
        evicted = store.add_episode(eps[fn], key=fn)
        assert [e["key"] for e in evicted] == gone
        assert store.size <= 400
        assert [e["key"] for e in store.episodes] == sorted(e["key"] for e in store.episodes)
        store.check_rows(eps)
    assert store.grow_calls >= 1 and store.moves >= 1
    with pytest.raises(ValueError):
        store.add_episode(eps[store.episodes[0]["key"]], key=store.episodes[0]["key"])


def test_fixed_capacity_store_refuses_to_grow():
    store = NumpyStore(1000, capacity_rows=30)
    store.add_episode(residency_episode(0, 10), key="episode_0_10.npz")
    store.add_episode(residency_episode(1, 10), key="episode_1_10.npz")
    with pytest.raises(RuntimeError):
        store.add_episode(residency_episode(2, 10), key="episode_2_10.npz")
