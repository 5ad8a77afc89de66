"""Host-side bookkeeping of which episodes are resident and where their rows live.

Pure Python (no device work), so the residency rules of the reference are testable on CPU against
fixtures generated from the reference itself (tests/golden/residency_reference.npz):

* `EpisodeIndex` mirrors `ReplayBuffer._episode_fns` / `_size` / `_store_episode`
  (/root/reference/utils/replay_buffer.py:172-190): the list is kept SORTED by file name after
  every insert (:184), eviction pops index 0 = the lexicographically smallest name (:179), while
  `eps_len + size > max_size` (:178).
* `RowAllocator` hands out contiguous row ranges of the device columns.  Because the reference
  evicts by sorted name rather than by age, freed ranges are not a ring tail: a first-fit free
  list with coalescing, compaction (when the free rows are fragmented) and growth.
"""
import bisect


class RowAllocator:
    def __init__(self, capacity=0):
        self.capacity = int(capacity)
        self._free = [(0, self.capacity)] if capacity else []   # sorted, coalesced (start, rows)

    @property
    def free_rows(self):
        return sum(n for _, n in self._free)

    def alloc(self, rows):
        """First fit; start row or None."""
        for i, (s, n) in enumerate(self._free):
            if n >= rows:
                if n == rows:
                    del self._free[i]
                else:
                    self._free[i] = (s + rows, n - rows)
                return s
        return None

    def alloc_at(self, start, rows):
        """Claim exactly [start, start+rows) (episodes laid out by the caller); grows the
        capacity if the range ends beyond it."""
        if start + rows > self.capacity:
            self.grow(start + rows)
        for i, (s, n) in enumerate(self._free):
            if s <= start and start + rows <= s + n:
                parts = [(s, start - s), (start + rows, s + n - start - rows)]
                self._free[i:i + 1] = [p for p in parts if p[1] > 0]
                return start
        raise ValueError(f"rows [{start}, {start + rows}) are not free")

    def release(self, start, rows):
        i = bisect.bisect_left(self._free, (start, 0))
        self._free.insert(i, (start, rows))
        if i + 1 < len(self._free) and start + rows == self._free[i + 1][0]:       # merge right
            self._free[i] = (start, rows + self._free[i + 1][1])
            del self._free[i + 1]
        if i > 0 and sum(self._free[i - 1]) == self._free[i][0]:                   # merge left
            self._free[i - 1] = (self._free[i - 1][0], self._free[i - 1][1] + self._free[i][1])
            del self._free[i]

    def grow(self, capacity):
        if capacity <= self.capacity:
            return
        self.release(self.capacity, capacity - self.capacity)
        self.capacity = capacity

    def compact(self, live):
        """Pack `live` = [(start, rows)] to the front in address order.  Returns the moves
        [(old_start, new_start, rows)] (only ranges that change place), ascending, each moving
        DOWN, so applying them in order never overwrites a range that has not moved yet."""
        moves, at = [], 0
        for s, n in sorted(live):
            if s != at:
                moves.append((s, at, n))
            at += n
        self._free = [(at, self.capacity - at)] if at < self.capacity else []
        return moves


class EpisodeIndex:
    """Sorted resident-episode list with the reference's size accounting and eviction rule."""

    def __init__(self, max_size):
        self.max_size = int(max_size)
        self.size = 0
        self.entries = []      # dicts(key, start, len), sorted by str(key)
        self._names = []       # str(key) of entries, same order (bisect key)

    def __len__(self):
        return len(self.entries)

    def __contains__(self, key):
        i = bisect.bisect_left(self._names, str(key))
        return i < len(self._names) and self._names[i] == str(key)

    def evictions_for(self, eps_len):
        """Pop and return the entries `_store_episode` would evict before storing an episode of
        `eps_len` transitions (:178-182).  Stops when the list is empty (the reference's fetch
        loop never stores an episode longer than max_size, :206-207)."""
        out = []
        while self.entries and eps_len + self.size > self.max_size:
            self._names.pop(0)
            e = self.entries.pop(0)
            self.size -= e["len"]
            out.append(e)
        return out

    def insert(self, key, start, eps_len):
        name = str(key)
        i = bisect.bisect_left(self._names, name)
        if i < len(self._names) and self._names[i] == name:
            raise ValueError(f"episode {name} is already resident")
        self._names.insert(i, name)
        self.entries.insert(i, {"key": key, "start": start, "len": eps_len})
        self.size += eps_len


def episode_len(episode):
    # rows - 1: row 0 is the dummy first transition (utils/replay_buffer.py:13-15)
    return next(iter(episode.values())).shape[0] - 1


class EpisodeStore:
    """Residency + row placement of a flat episode store; the column memory itself is behind five
    hooks (`_has_columns`, `_alloc_columns`, `_grow_columns`, `_move_rows`, `_write_rows`) that
    `replay.DeviceReplayBuffer` implements on device tensors (and the CPU tests on numpy arrays).

    The resident set follows `ReplayBuffer._store_episode` (utils/replay_buffer.py:172-190):
    episodes are kept sorted by file name, and while `eps_len + size > max_size` the FIRST of the
    sorted list is evicted — the lexicographically smallest name, not the oldest insertion.  The
    columns grow geometrically towards the `max_size`-derived estimate instead of being allocated
    at full size up front, and are compacted when the free rows are fragmented."""

    def __init__(self, max_size, capacity_rows=None):
        self.max_size = int(max_size)
        self.capacity_rows = capacity_rows      # fixed by the caller, or grown on demand
        self._fixed_capacity = capacity_rows is not None
        self._index = EpisodeIndex(self.max_size)
        self._alloc = RowAllocator(capacity_rows or 0)
        self._tables_dirty = True

    # hooks ---------------------------------------------------------------------------------
    @property
    def _has_columns(self):
        raise NotImplementedError

    def _alloc_columns(self, episode, capacity):
        raise NotImplementedError

    def _grow_columns(self, capacity):
        raise NotImplementedError

    def _move_rows(self, old, new, rows):
        raise NotImplementedError

    def _write_rows(self, start, episode, rows):
        raise NotImplementedError

    # `episodes` / `size`: the sorted resident list and the transition count of the reference's
    # `_episode_fns` / `_size`.  Assignable, so that a benchmark can lay synthetic columns under
    # a store without going through add_episode.
    @property
    def episodes(self):
        return self._index.entries

    @episodes.setter
    def episodes(self, entries):
        self._index = EpisodeIndex(self.max_size)
        self._alloc = RowAllocator(self.capacity_rows or 0)
        for e in sorted(entries, key=lambda e: e["start"]):
            start = self._alloc.alloc_at(e["start"], e["len"] + 1)
            self._index.insert(e["key"], start, e["len"])
        self._tables_dirty = True

    @property
    def size(self):
        return self._index.size

    @size.setter
    def size(self, value):
        self._index.size = int(value)

    def __len__(self):
        return self.size

    def _estimate_rows(self, eps_len):
        """Rows a full store of `eps_len`-step episodes needs (+1 dummy row per episode, two
        episodes of slack: the offline load overshoots the cap by one episode, :65)."""
        return (self.max_size // max(eps_len, 1) + 2) * (eps_len + 1)

    def reserve_rows(self, rows):
        """Make room for `rows` more rows without evicting (grows the columns).  Loaders call it
        with the exact need of a fetch — file names carry the episode lengths — so that a load is
        one allocation."""
        if self._alloc.free_rows < rows:
            self._grow(self._alloc.capacity + rows - self._alloc.free_rows)

    def _grow(self, capacity):
        if self._fixed_capacity:
            raise RuntimeError(f"episode store of {self.capacity_rows} rows (fixed by the caller) "
                               f"cannot hold {capacity} rows")
        if self._has_columns:
            self._grow_columns(capacity)
        self._alloc.grow(capacity)
        self.capacity_rows = capacity

    def _compact(self):
        moves = self._alloc.compact([(e["start"], e["len"] + 1) for e in self.episodes])
        where = {e["start"]: e for e in self.episodes}
        for old, new, rows in moves:
            self._move_rows(old, new, rows)
            where[old]["start"] = new
        self._tables_dirty = True

    def _take_rows(self, rows, eps_len):
        start = self._alloc.alloc(rows)
        if start is not None:
            return start
        if self._alloc.free_rows >= rows:
            self._compact()
        else:
            need = self._alloc.capacity + rows - self._alloc.free_rows
            target = max(2 * self._alloc.capacity, need)
            est = self._estimate_rows(eps_len)
            if need <= est < 2 * target:      # close to the steady-state size: go there at once
                target = est
            self._grow(target)
        return self._alloc.alloc(rows)

    def add_episode(self, episode, key=None, evict=True):
        """Store one episode (dict of arrays with L+1 rows).  `evict=True` is the online rule of
        utils/replay_buffer.py:178-182 (evict the first of the sorted names while
        `eps_len + size > max_size`); `evict=False` is the offline load (:57-75), which never
        evicts.  Returns the evicted entries (the loader unlinks their files, :182)."""
        eps_len = episode_len(episode)
        if key is None:
            key = f"episode_{len(self.episodes):08d}_{eps_len}.npz"
        if key in self._index:
            raise ValueError(f"episode {key} is already resident")
        if not self._has_columns:
            if not self._alloc.capacity:
                self._alloc.grow(min(self._estimate_rows(eps_len), 16 * (eps_len + 1)))
                self.capacity_rows = self._alloc.capacity
            self._alloc_columns(episode, self._alloc.capacity)
        evicted = self._index.evictions_for(eps_len) if evict else []
        for e in evicted:
            self._alloc.release(e["start"], e["len"] + 1)
        start = self._take_rows(eps_len + 1, eps_len)
        self._write_rows(start, episode, eps_len + 1)
        self._index.insert(key, start, eps_len)
        self._tables_dirty = True
        return evicted
