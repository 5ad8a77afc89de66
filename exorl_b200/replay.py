"""Device-resident replay buffer + loaders with the reference's loader signatures.

Replaces /root/reference/utils/replay_buffer.py:153-277: the per-sample python of
ReplayBuffer._sample / OfflineReplayBuffer._sample, DataLoader worker processes, collate,
pin_memory and the H2D copy of utils.to_torch.  Episodes are read from the same on-disk format
(`episode_{idx}_{len}.npz`, :13-29) into flat per-key device arrays; a batch is one index draw
plus one gather kernel (exorl_b200/csrc/replay_gather.cu) and is yielded as CUDA tensors, which
`utils.to_torch` passes through unchanged (utils/utils.py:55-56).
"""
import ctypes
import pathlib
import random

import numpy as np
import torch

from . import _lib
from . import ingest

_CORE = ("observation", "action", "reward", "discount")


def _episode_len(episode):
    # rows - 1: row 0 is the dummy first transition (utils/replay_buffer.py:13-15)
    return next(iter(episode.values())).shape[0] - 1


def _load_npz(path):
    """One episode file -> dict of (pinned) CPU tensors, via the native reader."""
    return ingest.read_episode(path)


def _as_tensor(a):
    return a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))


class DeviceReplayBuffer:
    """Flat struct-of-arrays episode store in HBM with FIFO eviction by episode.

    Row layout: an episode of length L occupies L+1 consecutive rows of every column (row 0 of
    the episode is the dummy first step).  Rows are allocated from a ring so that evicting the
    oldest episode (utils/replay_buffer.py:178-182) frees the space that is reused next.
    """

    def __init__(self, device, max_size, nstep, discount, meta_names=(), capacity_rows=None):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("exorl_b200: DeviceReplayBuffer needs a CUDA device (no CPU path)")
        self.max_size = int(max_size)
        self.nstep = int(nstep)
        self.discount = float(discount)
        self.meta_names = tuple(meta_names)
        self.capacity_rows = capacity_rows
        self.cols = {}            # name -> (capacity_rows, *shape) device tensor
        self.size = 0             # transitions currently held
        self.episodes = []        # FIFO of dicts(key, start, len) in insertion order
        self._head = 0            # next free row in the ring
        self._tables_dirty = True
        self._eps_start = None
        self._eps_len = None
        self._sorted_keys = []

    # -- storage ----------------------------------------------------------------------------
    def _allocate(self, episode):
        eps_len = _episode_len(episode)
        if self.capacity_rows is None:
            per = eps_len + 1
            n_eps = self.max_size // max(eps_len, 1) + 2
            self.capacity_rows = n_eps * per + per
        for name in (*_CORE, *self.meta_names):
            a = _as_tensor(episode[name])
            self.cols[name] = torch.empty((self.capacity_rows, *a.shape[1:]), dtype=a.dtype,
                                          device=self.device)

    def _evict_oldest(self):
        old = self.episodes.pop(0)
        self.size -= old["len"]
        self._tables_dirty = True
        return old

    def _reserve(self, rows):
        """Start row of a free contiguous range of `rows`, evicting oldest episodes if needed.

        Ring invariant: live rows run from tail = episodes[0].start forward (wrapping) to
        _head.  head > tail: live is [tail, head), free is [head, cap) and [0, tail).
        head <= tail: the allocator has wrapped, free is [head, tail)."""
        if rows > self.capacity_rows:
            raise RuntimeError(f"episode of {rows} rows exceeds buffer capacity {self.capacity_rows}")
        while True:
            if not self.episodes:
                self._head = 0
                return 0
            tail = self.episodes[0]["start"]
            if self._head > tail:
                if self._head + rows <= self.capacity_rows:
                    return self._head
                if rows <= tail:
                    return 0
            elif self._head + rows <= tail:
                return self._head
            self._evict_oldest()

    def add_episode(self, episode, key=None):
        """Append one episode (dict of numpy arrays with L+1 rows), evicting from the front while
        `eps_len + size > max_size` like utils/replay_buffer.py:178-182."""
        eps_len = _episode_len(episode)
        if not self.cols:
            self._allocate(episode)
        while self.episodes and eps_len + self.size > self.max_size:
            self._evict_oldest()
        start = self._reserve(eps_len + 1)
        for name, col in self.cols.items():
            a = _as_tensor(episode[name])
            if a.shape[0] != eps_len + 1:
                raise ValueError(f"column {name}: {a.shape[0]} rows, expected {eps_len + 1}")
            # pinned staging (ingest.read_episode) -> asynchronous DMA on the current stream; the
            # staging tensor stays referenced by `episode` until the caller drops it, and the
            # caching host allocator does not reuse a pinned block with a copy in flight
            col[start:start + eps_len + 1].copy_(a, non_blocking=a.is_pinned())
        self._head = start + eps_len + 1
        self.episodes.append({"key": key if key is not None else len(self.episodes),
                              "start": start, "len": eps_len})
        self.size += eps_len
        self._tables_dirty = True

    def _tables(self):
        """Episode tables ordered like the reference's sorted `_episode_fns` (:184-185)."""
        if self._tables_dirty:
            order = sorted(range(len(self.episodes)), key=lambda i: str(self.episodes[i]["key"]))
            self._sorted = [self.episodes[i] for i in order]
            starts = np.array([e["start"] for e in self._sorted], np.int64)
            lens = np.array([e["len"] for e in self._sorted], np.int32)
            self._eps_start = torch.from_numpy(starts).to(self.device)
            self._eps_len = torch.from_numpy(lens).to(self.device)
            self._tables_dirty = False
        return self._sorted, self._eps_start, self._eps_len

    def __len__(self):
        return self.size

    # -- relabelling (the step after the kNN reward path) ------------------------
    def live_rows(self):
        """Global row ids of every stored step (all L+1 rows of every episode, in the sorted
        episode order of `_tables`), as a device int64 vector."""
        eps, _, _ = self._tables()
        parts = [torch.arange(e["start"], e["start"] + e["len"] + 1, device=self.device)
                 for e in eps]
        return torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64, device=self.device)

    def relabel_rewards(self, rewards, rows=None):
        """Overwrite the stored reward of `rows` (default: every live row, `live_rows()` order)
        with `rewards` (n, 1) — the in-memory effect of relabel_episode
        (utils/replay_buffer.py:31-42: episode['reward'] = one reward per stored state), with the
        rewards coming from the kNN reward path instead of the simulator."""
        rows = self.live_rows() if rows is None else rows
        r = rewards.reshape(-1, 1).to(device=self.device, dtype=torch.float32)
        if r.size(0) != rows.numel():
            raise ValueError(f"{r.size(0)} rewards for {rows.numel()} rows")
        self.cols["reward"].index_copy_(0, rows, r)

    def export_episodes(self, out_dir):
        """Write every stored episode back in the reference's format (save_episode,
        utils/replay_buffer.py:18-23), one `.npz` per episode under its original file name."""
        out_dir = pathlib.Path(out_dir)
        out_dir.mkdir(parents=True, exist_ok=True)
        written = []
        for e in self.episodes:
            lo, hi = e["start"], e["start"] + e["len"] + 1
            ep = {name: col[lo:hi].cpu().numpy() for name, col in self.cols.items()}
            name = pathlib.Path(str(e["key"])).name
            if not name.endswith(".npz"):
                name = f"episode_{name}_{e['len']}.npz"
            with open(out_dir / name, "wb") as f:
                np.savez_compressed(f, **ep)
            written.append(out_dir / name)
        return written

    # -- sampling ---------------------------------------------------------------------------
    def draw_rows_philox(self, batch, seed, call, offline=False):
        """Device index draw: uniform episode, then uniform start (range of :222 / :90)."""
        lib = _lib.load()
        eps, starts, lens = self._tables()
        if not eps:
            raise RuntimeError("replay buffer is empty")
        rows = torch.empty(batch, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.exb_replay_draw(
                ctypes.c_uint64(seed), ctypes.c_uint64(call), _lib.ptr(starts), _lib.ptr(lens),
                len(eps), 1 if offline else self.nstep, batch, _lib.ptr(rows), None,
                _lib.stream_ptr(self.device)))
        return rows

    def draw_rows_reference_rng(self, batch, offline=False):
        """Host index draw with the reference's own RNG calls, in its order: python
        `random.choice` over the sorted episodes then `np.random.randint`
        (utils/replay_buffer.py:168-170,222 and :78-80,90) — the parity mode."""
        eps, _, _ = self._tables()
        rows = np.empty(batch, np.int64)
        for b in range(batch):
            e = random.choice(eps)
            span = e["len"] if offline else e["len"] - self.nstep + 1
            rows[b] = e["start"] + np.random.randint(0, span) + 1
        return torch.from_numpy(rows).pin_memory().to(self.device, non_blocking=True)

    def rows_from_draws(self, draws):
        """(episode position in sorted order, 1-based idx) pairs -> global rows (tests)."""
        eps, _, _ = self._tables()
        rows = np.array([eps[int(e)]["start"] + int(i) for e, i in draws], np.int64)
        return torch.from_numpy(rows).to(self.device)

    def _outputs(self, batch, offline):
        """Output tensors + column specs of one batch (the tuple of utils/replay_buffer.py:235)."""
        nstep = 1 if offline else self.nstep
        obs_src, act_src = self.cols["observation"], self.cols["action"]
        dev = self.device
        if self.cols["reward"][0].numel() != 1 or self.cols["reward"].dtype != torch.float32:
            raise ValueError("reward/discount columns must be float32 of shape (1,) per step")
        obs = torch.empty((batch, *obs_src.shape[1:]), dtype=obs_src.dtype, device=dev)
        nobs = torch.empty_like(obs)
        act = torch.empty((batch, *act_src.shape[1:]), dtype=act_src.dtype, device=dev)
        rew = torch.empty((batch, *self.cols["reward"].shape[1:]), dtype=torch.float32, device=dev)
        disc = torch.empty_like(rew)
        specs = [(obs_src, obs, -1), (act_src, act, 0), (obs_src, nobs, nstep - 1)]
        metas = []
        if not offline:
            for name in self.meta_names:
                src = self.cols[name]
                dst = torch.empty((batch, *src.shape[1:]), dtype=src.dtype, device=dev)
                metas.append(dst)
                specs.append((src, dst, -1))
        cols = (_lib.GatherCol * len(specs))()
        for c, (src, dst, off) in zip(cols, specs):
            c.src, c.dst = src.data_ptr(), dst.data_ptr()
            c.row_bytes = src[0].numel() * src.element_size()
            c.row_offset = off
        return (obs, act, rew, disc, nobs, *metas), cols, len(specs), nstep

    def gather(self, rows, offline=False):
        """(obs, action, reward, discount, next_obs, *meta) for global `rows`, as CUDA tensors —
        the tuple of utils/replay_buffer.py:235 (:96 offline)."""
        lib = _lib.load()
        batch = rows.numel()
        out, cols, n_cols, nstep = self._outputs(batch, offline)
        with torch.cuda.device(self.device):
            _lib.check(lib.exb_replay_gather(
                _lib.ptr(rows), batch, cols, n_cols, _lib.ptr(self.cols["reward"]),
                _lib.ptr(self.cols["discount"]), nstep, 0 if offline else 1, self.discount,
                _lib.ptr(out[2]), _lib.ptr(out[3]), _lib.stream_ptr(self.device)))
        return out

    def sample_philox(self, batch, seed, call, offline=False, return_rows=False):
        """Index draw + gather as ONE kernel launch (exb_replay_sample): same result as
        `gather(draw_rows_philox(batch, seed, call))`."""
        lib = _lib.load()
        eps, starts, lens = self._tables()
        if not eps:
            raise RuntimeError("replay buffer is empty")
        out, cols, n_cols, nstep = self._outputs(batch, offline)
        rows = torch.empty(batch, dtype=torch.int64, device=self.device) if return_rows else None
        with torch.cuda.device(self.device):
            _lib.check(lib.exb_replay_sample(
                ctypes.c_uint64(seed), ctypes.c_uint64(call), _lib.ptr(starts), _lib.ptr(lens),
                len(eps), batch, cols, n_cols, _lib.ptr(self.cols["reward"]),
                _lib.ptr(self.cols["discount"]), nstep, 0 if offline else 1, self.discount,
                _lib.ptr(out[2]), _lib.ptr(out[3]), _lib.ptr(rows), _lib.stream_ptr(self.device)))
        return (out, rows) if return_rows else out


class DeviceReplayLoader:
    """Iterable that yields batches forever, like the DataLoader the reference builds
    (utils/replay_buffer.py:246-277).  `num_workers` only scales `max_size` the way the
    reference's per-worker split does; there are no worker processes."""

    def __init__(self, buffer, batch_size, replay_dir=None, offline=False, save_snapshot=True,
                 fetch_every=1000, sampler="philox", seed=None):
        if sampler not in ("philox", "reference_rng"):
            raise ValueError("sampler must be 'philox' or 'reference_rng'")
        self.buffer = buffer
        self.batch_size = int(batch_size)
        self.replay_dir = pathlib.Path(replay_dir) if replay_dir is not None else None
        self.offline = offline
        self.save_snapshot = save_snapshot
        self.fetch_every = fetch_every
        self.sampler = sampler
        self.seed = int(seed if seed is not None else np.random.randint(0, 2 ** 31 - 1))
        self._since_fetch = fetch_every
        self._seen = set()
        self._calls = 0
        self._loaded = False

    # -- ingestion ---------------------------------------------------------------------------
    def _try_fetch(self):
        """Online: newest-first scan of the replay dir for unseen episodes
        (utils/replay_buffer.py:192-212)."""
        if self._since_fetch < self.fetch_every or self.replay_dir is None:
            return
        self._since_fetch = 0
        new = []
        fetched = 0
        for fn in sorted(self.replay_dir.glob("*.npz"), reverse=True):
            eps_len = int(fn.stem.split("_")[2])
            if fn in self._seen:
                break
            if fetched + eps_len > self.buffer.max_size:
                break
            fetched += eps_len
            new.append(fn)
        try:              # the reference stores them in this (newest-first) order too
            for fn, episode in ingest.read_episodes(new):
                self.buffer.add_episode(episode, key=str(fn))
                self._seen.add(fn)
                if not self.save_snapshot:
                    fn.unlink(missing_ok=True)
        except (ValueError, RuntimeError, OSError):
            pass          # a half-written file: stop here, like the `break` of :206-209

    def _load_offline(self):
        """Offline: sorted scan until the size cap (utils/replay_buffer.py:57-75)."""
        for fn, episode in ingest.read_episodes(sorted(self.replay_dir.glob("*.npz"))):
            if self.buffer.size > self.buffer.max_size:
                break
            self.buffer.add_episode(episode, key=str(fn))
        self._loaded = True

    # -- iteration ---------------------------------------------------------------------------
    def sample(self):
        if self.offline:
            if not self._loaded:
                self._load_offline()
        else:
            self._try_fetch()
            self._since_fetch += self.batch_size
        self._calls += 1
        if self.sampler == "philox":
            return self.buffer.sample_philox(self.batch_size, self.seed, self._calls - 1,
                                             offline=self.offline)
        rows = self.buffer.draw_rows_reference_rng(self.batch_size, offline=self.offline)
        return self.buffer.gather(rows, offline=self.offline)

    def __iter__(self):
        while True:
            yield self.sample()


def _device(device):
    return torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")


def make_replay_loader(storage, max_size, batch_size, num_workers, save_snapshot, nstep, discount,
                       device=None, sampler="philox", seed=None):
    """Same positional signature as utils/replay_buffer.py:260-277.  `storage` exposes
    `_replay_dir` and `_meta_specs` (objects with `.name`)."""
    workers = max(1, num_workers)
    cap = (max_size // workers) * workers      # the reference caps each worker at max_size//W
    meta = tuple(spec.name for spec in storage._meta_specs)
    buf = DeviceReplayBuffer(_device(device), cap, nstep, discount, meta_names=meta)
    return DeviceReplayLoader(buf, batch_size, replay_dir=storage._replay_dir, offline=False,
                              save_snapshot=save_snapshot, fetch_every=1000, sampler=sampler,
                              seed=seed)


def make_offline_replay_loader(env, replay_dir, max_size, batch_size, num_workers, discount,
                               device=None, sampler="philox", seed=None):
    """Same positional signature as utils/replay_buffer.py:246-258.  `env` is accepted and
    unused: extrinsic relabelling through MuJoCo physics (:31-42) is outside this path."""
    workers = max(1, num_workers)
    cap = (max_size // workers) * workers
    buf = DeviceReplayBuffer(_device(device), cap, 1, discount)
    return DeviceReplayLoader(buf, batch_size, replay_dir=replay_dir, offline=True,
                              sampler=sampler, seed=seed)
