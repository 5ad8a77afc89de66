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
from . import ops  # noqa: F401  (registers torch.ops.exorl_b200.*)
from .residency import EpisodeStore

_CORE = ("observation", "action", "reward", "discount")


def _as_tensor(a):
    return a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))


class DeviceReplayBuffer(EpisodeStore):
    """Flat struct-of-arrays episode store in HBM.

    Row layout: an episode of length L occupies L+1 consecutive rows of every column (row 0 of
    the episode is the dummy first step).  Which episodes are resident, where their rows are and
    when the columns grow or compact is `residency.EpisodeStore` (the reference's sorted-name
    eviction, utils/replay_buffer.py:172-190); this class holds the device columns and the
    sampling kernels."""

    def __init__(self, device, max_size, nstep, discount, meta_names=(), capacity_rows=None):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("exorl_b200: DeviceReplayBuffer needs a CUDA device (no CPU path)")
        super().__init__(max_size, capacity_rows)
        self.nstep = int(nstep)
        self.discount = float(discount)
        self.meta_names = tuple(meta_names)
        self.cols = {}            # name -> (capacity_rows, *shape) device tensor
        self._eps_start = None
        self._eps_len = None
        self._min_len = 0

    # -- column memory (EpisodeStore hooks) ---------------------------------------------------
    @property
    def _has_columns(self):
        return bool(self.cols)

    def _alloc_columns(self, episode, capacity):
        for name in (*_CORE, *self.meta_names):
            a = _as_tensor(episode[name])
            self.cols[name] = torch.empty((capacity, *a.shape[1:]), dtype=a.dtype,
                                          device=self.device)

    def _grow_columns(self, capacity):
        for name, old in self.cols.items():
            new = torch.empty((capacity, *old.shape[1:]), dtype=old.dtype, device=self.device)
            new[:old.shape[0]].copy_(old)
            self.cols[name] = new

    def _move_rows(self, old, new, rows):
        for col in self.cols.values():
            src = col[old:old + rows]
            col[new:new + rows].copy_(src.clone() if new + rows > old else src)

    def _write_rows(self, start, episode, rows):
        for name, col in self.cols.items():
            a = _as_tensor(episode[name])
            if a.shape[0] != rows:
                raise ValueError(f"column {name}: {a.shape[0]} rows, expected {rows}")
            # pinned staging (ingest.read_episode) -> asynchronous DMA on the current stream; the
            # staging tensor stays referenced by `episode` until the caller drops it, and the
            # caching host allocator does not reuse a pinned block with a copy in flight
            col[start:start + rows].copy_(a, non_blocking=a.is_pinned())

    def _tables(self):
        """Device episode tables in the order of the reference's sorted `_episode_fns` (:184)."""
        if self._tables_dirty:
            eps = self.episodes
            starts = np.array([e["start"] for e in eps], np.int64)
            lens = np.array([e["len"] for e in eps], np.int32)
            self._eps_start = torch.from_numpy(starts).to(self.device)
            self._eps_len = torch.from_numpy(lens).to(self.device)
            self._min_len = int(lens.min()) if len(eps) else 0
            self._tables_dirty = False
        return self.episodes, self._eps_start, self._eps_len

    def _check_spans(self, offline):
        """The reference draws `np.random.randint(0, eps_len - nstep + 1)` (:222; `eps_len`
        offline, :90), which raises ValueError for an episode with no valid start.  There the
        error depends on which episode is drawn; here any resident episode that short raises."""
        eps, _, _ = self._tables()
        if not eps:
            raise RuntimeError("replay buffer is empty")
        need = 1 if offline else self.nstep
        if self._min_len < need:
            short = next(e for e in eps if e["len"] < need)
            raise ValueError(f"low >= high: episode {short['key']} has {short['len']} steps, "
                             f"fewer than nstep={need}")

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
        self._check_spans(offline)
        eps, starts, lens = self._tables()
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
            rows[b] = e["start"] + np.random.randint(0, span) + 1   # ValueError if span <= 0
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
        `gather(draw_rows_philox(batch, seed, call))`.  Thin caller of
        torch.ops.exorl_b200.replay_sample."""
        self._check_spans(offline)
        _, starts, lens = self._tables()
        nstep = 1 if offline else self.nstep
        obs_src, act_src = self.cols["observation"], self.cols["action"]
        if self.cols["reward"][0].numel() != 1 or self.cols["reward"].dtype != torch.float32:
            raise ValueError("reward/discount columns must be float32 of shape (1,) per step")
        metas = [] if offline else [self.cols[name] for name in self.meta_names]
        cols = [obs_src, act_src, obs_src, *metas]
        offsets = [-1, 0, nstep - 1] + [-1] * len(metas)
        out = ops.call(
            torch.ops.exorl_b200.replay_sample, _replay_sample_impl, cols, offsets, self.cols["reward"], self.cols["discount"], starts, lens, int(batch),
            _as_i64(seed), _as_i64(call), nstep, 0 if offline else 1, self.discount)
        obs, act, nobs = out[0], out[1], out[2]
        rew, disc, rows = out[-3], out[-2], out[-1]
        batch_out = (obs, act, rew, disc, nobs, *out[3:-3])
        return (batch_out, rows) if return_rows else batch_out


def _as_i64(v):
    """64-bit seed / call counter as the signed int an op schema carries (two's complement)."""
    v = int(v) & (2 ** 64 - 1)
    return v - 2 ** 64 if v >= 2 ** 63 else v


def _replay_sample_impl(columns, row_offsets, reward, discount, eps_start, eps_len, batch, seed,
                        call, nstep, nstep_mode, gamma):
    """CUDA implementation of torch.ops.exorl_b200.replay_sample: one exb_replay_sample launch."""
    lib = _lib.load()
    dev = reward.device
    outs = [torch.empty((batch, *c.shape[1:]), dtype=c.dtype, device=dev) for c in columns]
    rew = torch.empty((batch, *reward.shape[1:]), dtype=torch.float32, device=dev)
    disc = torch.empty_like(rew)
    rows = torch.empty(batch, dtype=torch.int64, device=dev)
    cols = (_lib.GatherCol * len(columns))()
    for c, src, dst, off in zip(cols, columns, outs, row_offsets):
        c.src, c.dst = src.data_ptr(), dst.data_ptr()
        c.row_bytes = src[0].numel() * src.element_size()
        c.row_offset = int(off)
    with torch.cuda.device(dev):
        _lib.check(lib.exb_replay_sample(
            ctypes.c_uint64(seed & (2 ** 64 - 1)), ctypes.c_uint64(call & (2 ** 64 - 1)),
            _lib.ptr(eps_start), _lib.ptr(eps_len), eps_len.numel(), batch, cols, len(columns),
            _lib.ptr(reward), _lib.ptr(discount), nstep, nstep_mode, gamma, _lib.ptr(rew),
            _lib.ptr(disc), _lib.ptr(rows), _lib.stream_ptr(dev)))
    return [*outs, rew, disc, rows]


def _eps_idx_len(fn):
    # `episode_{idx}_{len}.npz` (utils/replay_buffer.py:146-149); ValueError on other names, as
    # the reference's `int(x) for x in eps_fn.stem.split('_')[1:]` (:67, :203)
    idx, eps_len = [int(x) for x in fn.stem.split("_")[1:]]
    return idx, eps_len


class _Worker:
    """One DataLoader worker of the reference: its own episode set (files with
    `eps_idx % num_workers == worker_id`, :68-69, :204-205), size cap `max_size // num_workers`
    (:262) and fetch counter (:164, :195-197)."""

    def __init__(self, wid, buffer, fetch_every):
        self.wid = wid
        self.buffer = buffer
        self.since_fetch = fetch_every
        self.loaded = False


class DeviceReplayLoader:
    """Iterable that yields batches forever, like the DataLoader the reference builds
    (utils/replay_buffer.py:246-277).  The reference's `num_workers` worker processes become
    `num_workers` partitions of the device store: batch i is drawn from partition i % W (a
    DataLoader takes whole batches from its workers in turn), and each partition fetches, caps
    and evicts on its own like a worker does.  There are no processes; `buffer` is partition 0."""

    def __init__(self, buffers, batch_size, replay_dir=None, offline=False, save_snapshot=True,
                 fetch_every=1000, sampler="philox", seed=None):
        if sampler not in ("philox", "reference_rng"):
            raise ValueError("sampler must be 'philox' or 'reference_rng'")
        if isinstance(buffers, DeviceReplayBuffer):
            buffers = [buffers]
        self.workers = [_Worker(w, b, fetch_every) for w, b in enumerate(buffers)]
        self.batch_size = int(batch_size)
        self.replay_dir = pathlib.Path(replay_dir) if replay_dir is not None else None
        self.offline = offline
        self.save_snapshot = save_snapshot
        self.fetch_every = fetch_every
        self.sampler = sampler
        self.seed = int(seed if seed is not None else np.random.randint(0, 2 ** 31 - 1))
        self._calls = 0

    @property
    def buffer(self):
        return self.workers[0].buffer

    @property
    def buffers(self):
        return [w.buffer for w in self.workers]

    def resident_files(self):
        """Names of the resident episode files per partition, in the reference's sorted order
        (`ReplayBuffer._episode_fns` of worker w)."""
        return [[pathlib.Path(str(e["key"])).name for e in w.buffer.episodes]
                for w in self.workers]

    # -- ingestion ---------------------------------------------------------------------------
    def _try_fetch(self, worker):
        """`ReplayBuffer._try_fetch` + `_store_episode` (utils/replay_buffer.py:172-212) for one
        partition: newest-first scan of the directory, stop at the first file that is already
        resident (:206) or that would take this fetch over the cap (:208); every stored episode
        evicts the first of the sorted resident names while over the cap and UNLINKS the evicted
        file (:178-182); a file that cannot be read ends the fetch (:211-212)."""
        if worker.since_fetch < self.fetch_every or self.replay_dir is None:
            return
        worker.since_fetch = 0
        buf, n_workers = worker.buffer, len(self.workers)
        todo, fetched = [], 0
        for fn in sorted(self.replay_dir.glob("*.npz"), reverse=True):
            eps_idx, eps_len = _eps_idx_len(fn)
            if eps_idx % n_workers != worker.wid:
                continue
            if str(fn) in buf._index:
                break
            if fetched + eps_len > buf.max_size:
                break
            fetched += eps_len
            todo.append(fn)
        if todo and not buf._has_columns:
            buf.reserve_rows(sum(_eps_idx_len(fn)[1] + 1 for fn in todo))
        it = ingest.read_episodes(todo)
        while True:
            try:
                fn, episode = next(it)
            except StopIteration:
                break
            except (ValueError, OSError, EOFError):
                break     # a half-written or damaged file: `_store_episode` returned False
            for old in buf.add_episode(episode, key=str(fn)):
                pathlib.Path(old["key"]).unlink(missing_ok=True)
            if not self.save_snapshot:
                fn.unlink(missing_ok=True)

    def _load_offline(self, worker):
        """`OfflineReplayBuffer._load` (utils/replay_buffer.py:57-75): sorted scan; the size test
        comes BEFORE each file (:65-66), so loading stops once the partition holds more than its
        cap — the first episodes are kept, the cap is overshot by at most one episode, nothing
        is evicted and nothing further is read."""
        buf, n_workers = worker.buffer, len(self.workers)
        todo, size = [], 0
        for fn in sorted(self.replay_dir.glob("*.npz")):
            if size > buf.max_size:
                break
            eps_idx, eps_len = _eps_idx_len(fn)
            if eps_idx % n_workers != worker.wid:
                continue
            todo.append(fn)
            size += eps_len     # the name's length field; the loop below re-checks on real sizes
        if todo:
            buf.reserve_rows(sum(_eps_idx_len(fn)[1] + 1 for fn in todo))
        for fn, episode in ingest.read_episodes(todo):
            if buf.size > buf.max_size:
                break
            buf.add_episode(episode, key=str(fn), evict=False)
        worker.loaded = True

    # -- iteration ---------------------------------------------------------------------------
    def sample(self):
        worker = self.workers[self._calls % len(self.workers)]
        if self.offline:
            if not worker.loaded:
                self._load_offline(worker)
        else:
            self._try_fetch(worker)
            worker.since_fetch += self.batch_size
        self._calls += 1
        buf = worker.buffer
        if self.sampler == "philox":
            return buf.sample_philox(self.batch_size, self.seed, self._calls - 1,
                                     offline=self.offline)
        rows = buf.draw_rows_reference_rng(self.batch_size, offline=self.offline)
        return buf.gather(rows, offline=self.offline)

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
    meta = tuple(spec.name for spec in storage._meta_specs)
    bufs = [DeviceReplayBuffer(_device(device), max_size // workers, nstep, discount,
                               meta_names=meta) for _ in range(workers)]
    return DeviceReplayLoader(bufs, batch_size, replay_dir=storage._replay_dir, offline=False,
                              save_snapshot=save_snapshot, fetch_every=1000, sampler=sampler,
                              seed=seed)


def make_offline_replay_loader(env, replay_dir, max_size, batch_size, num_workers, discount,
                               device=None, sampler="philox", seed=None):
    """Same positional signature as utils/replay_buffer.py:246-258.  `env` is accepted and
    unused: extrinsic relabelling through MuJoCo physics (:31-42) is outside this path."""
    workers = max(1, num_workers)
    bufs = [DeviceReplayBuffer(_device(device), max_size // workers, 1, discount)
            for _ in range(workers)]
    return DeviceReplayLoader(bufs, batch_size, replay_dir=replay_dir, offline=True,
                              sampler=sampler, seed=seed)
