"""Dataset-scale APT relabel: kNN reward of a whole buffer against itself, bank sharded over ranks.

New composition of the reference's PBE semantics (utils/utils.py:289-319) at N ~ 10^6, where the
reference formula cannot be evaluated (it materialises an (N, N, c) tensor); SURVEY.md §3.5, §8e.

  every rank holds the full query set `rep` (N, c) and owns bank rows [lo, hi) = its shard;
  1. local:    k nearest bank rows of the shard for every query (exact distances, global ids)
  2. exchange: "all_gather"  — every rank receives all G per-shard (N, k) lists and merges all
                               N rows (north_star's formulation), or
               "all_to_all"  — rank r receives the G lists of query block r only, merges N/G
                               rows, and the RMS sums are all-reduced (8x less traffic/merge work)
               "lists"       — step 1 stops after the tensor-core screen; the CANDIDATE lists of
                               query block r go to rank r (all-to-all), which recomputes them
                               exactly against the whole fp32 bank it holds anyway: the exact
                               recompute runs once per query instead of once per (query, shard)
                               and there is no merge
               "queries"     — no bank sharding at all: every rank already holds the whole
                               set, so rank r computes the kNN of QUERY block r against the whole
                               bank; nothing but the RMS sums and the rewards is exchanged, and
                               every row's k-th-best chain sees the whole bank (fewest hits)
               "auto"        — "queries"
  3. one global RMS update with bs = N*k (or N), then the epilogue
  4. all_to_all only: rewards all-gathered so every rank returns the full (N, 1).

The three compute stages are injectable so the partition/exchange logic can be exercised with
the gloo backend on CPU (tests/test_relabel_gloo.py); the defaults are the CUDA kernels.
"""
import ctypes

import torch
import torch.distributed as dist

from . import _lib
from .rep import PreparedRep, column_affine
from .knn import (knn_topk, _reward_from_knn, _pbe_ws, knn_screen_lists, knn_rerank_lists,
                  screen_lists_count)


def shard_bounds(n, world, rank):
    """Contiguous near-equal row split: rows [lo, hi) of rank `rank`."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _shard(rep, lo, hi):
    """Bank shard [lo, hi) of a rep tensor or PreparedRep; the object itself when the shard is the
    whole set (`is` marks source = target)."""
    if lo == 0 and hi == rep.size(0):
        return rep
    return rep.rows(lo, hi) if isinstance(rep, PreparedRep) else rep[lo:hi]


def _fp32(rep):
    return rep.rep if isinstance(rep, PreparedRep) else rep


def _cuda_local_knn(rep, lo, hi, k, mode):
    if isinstance(rep, PreparedRep) and mode == "auto":
        mode = "screened"
    return knn_topk(rep, _shard(rep, lo, hi), k, idx_base=lo, mode=mode)


def _cuda_merge(dist_lists, idx_lists, k):
    """dist_lists/idx_lists: (G, n, k) -> (n, k) global k smallest."""
    lib = _lib.load()
    g, n, _ = dist_lists.shape
    dev = dist_lists.device
    out_d = torch.empty(n, k, dtype=torch.float32, device=dev)
    out_i = torch.empty(n, k, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.exb_topk_merge(_lib.ptr(dist_lists), _lib.ptr(idx_lists), g, n, k,
                                      _lib.ptr(out_d), _lib.ptr(out_i), _lib.stream_ptr(dev)))
    return out_d, out_i


def _cuda_sums(knn_dist, k, avg):
    lib = _lib.load()
    dev = knn_dist.device
    out = torch.empty(3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        ws = _pbe_ws(dev)
        _lib.check(lib.exb_pbe_moments(_lib.ptr(knn_dist), knn_dist.size(0), k, int(avg),
                                       _lib.ptr(out), _lib.ptr(ws), ws.numel(),
                                       _lib.stream_ptr(dev)))
    return out


def _cuda_finish(knn_dist, k, avg, use_rms, clip, sums, rms):
    lib = _lib.load()
    dev = knn_dist.device
    n = knn_dist.size(0)
    out = torch.empty(n, 1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        # RMS state (M, S, n) is updated in place on the device from the combined sums
        st = (_lib.ptr(rms.M), _lib.ptr(rms.S), _lib.ptr(rms._n)) if use_rms else (None,) * 3
        _lib.check(lib.exb_pbe_finish(_lib.ptr(knn_dist), n, k, int(avg), int(use_rms),
                                      float(clip), _lib.ptr(sums) if use_rms else None, *st,
                                      _lib.ptr(out), _lib.stream_ptr(dev)))
    return out


def _cuda_screen_lists(rep, lo, hi, k):
    return knn_screen_lists(rep, _shard(rep, lo, hi), k)


def _cuda_rerank_lists(rep, lo, hi, k, val, idx, cnt, shard_lo, ha, ea, sc):
    rep = _fp32(rep)     # the exact recompute reads the fp32 rows only
    return knn_rerank_lists(rep[lo:hi], rep, k, val, idx, cnt, shard_lo, ha, ea, sc)


def _cuda_lists_available(n, nb, dim, k, world):
    """Candidate lists per (row, shard) of the screen-only form, 0 if it cannot be used."""
    lists = screen_lists_count(n, nb, dim, k)
    return lists if 0 < lists * world <= 16 else 0


def _cuda_block_knn(rep, lo, hi, k, mode):
    """Exact k nearest rows of the WHOLE set for the query block [lo, hi) (global ids)."""
    if isinstance(rep, PreparedRep) and mode == "auto":
        mode = "screened"
    return knn_topk(_shard(rep, lo, hi), rep, k, mode=mode)


class _CudaStopwatch:
    """Device time of the block kNN of one step, read without a sync at the start of the next."""

    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.armed = False

    def start(self):
        self.e0.record()

    def stop(self):
        self.e1.record()
        self.armed = True

    def elapsed_ms(self):
        """Milliseconds of the last start/stop pair, or 0.0 if there is none or it has not
        finished on the device yet."""
        if not self.armed or not self.e1.query():
            return 0.0
        self.armed = False
        return float(self.e0.elapsed_time(self.e1))


class _CudaStages:
    local_knn = staticmethod(_cuda_local_knn)
    block_knn = staticmethod(_cuda_block_knn)
    screen_lists = staticmethod(_cuda_screen_lists)
    rerank_lists = staticmethod(_cuda_rerank_lists)
    lists_available = staticmethod(_cuda_lists_available)
    merge = staticmethod(_cuda_merge)
    sums = staticmethod(_cuda_sums)
    finish = staticmethod(_cuda_finish)
    stopwatch = _CudaStopwatch

    @staticmethod
    def reward(knn_dist, k, avg, use_rms, clip, rms):
        st = (rms.M, rms.S, rms._n) if use_rms else (None,) * 3
        return _reward_from_knn(knn_dist, k, avg, use_rms, clip, *st)


def balanced_bounds(n, shares, quantum=256, min_rows=1):
    """Contiguous row blocks [cut[r], cut[r+1]) with sizes proportional to `shares`, cut at
    multiples of `quantum` rows (one CTA pair of the screening kernel owns 256 query rows).
    Falls back to the equal split when a block would get fewer than `min_rows` rows."""
    world = len(shares)
    tot = float(sum(shares))
    cuts, acc = [0], 0.0
    for r in range(world - 1):
        acc += shares[r] / tot
        cuts.append(min(n, max(cuts[-1], int(round(acc * n / quantum)) * quantum)))
    cuts.append(n)
    if any(b - a < max(min_rows, 1) for a, b in zip(cuts[:-1], cuts[1:])):
        return [shard_bounds(n, world, r) for r in range(world)]
    return list(zip(cuts[:-1], cuts[1:]))


class QueryBalance:
    """Per-rank query-block sizes of the 'queries' partition, proportional to each rank's
    measured kNN rate (rows per millisecond) in the previous step.

    Meant for ranks that differ PERSISTENTLY (mixed parts, a throttled GPU).  On one box of
    8 x B200 under the 1 kW power cap they do not: measured per-rank block times at equal blocks
    are 89-98 ms with a different slowest rank every step (12 steps x 8 ranks in
    profiles/r02b_balance_observe_n8.json; per-rank means within 1.5 %), so the step pays for the
    maximum of eight noisy times whatever the split, and chasing last step's rates only adds
    noise (97.3 vs 96.3 ms per step).  Hence off by default.  Each rank times its block kNN with
    CUDA events;
    the time of step s rides along in the RMS-sums all-reduce of step s+1 (no extra collective,
    no host sync), so every rank derives the SAME bounds for step s+2.  Damped (`gain`) because
    the kernel's time is not exactly linear in its rows (waves of 148 CTA pairs)."""

    def __init__(self, n, world, quantum=256, min_rows=1, gain=0.7):
        self.n, self.world, self.quantum, self.min_rows, self.gain = n, world, quantum, min_rows, gain
        self.observe_only = False     # collect the times, keep the equal split (measurements)
        self.log = []                 # (rows per rank, ms per rank) of every settled step
        self.shares = [1.0 / world] * world
        self.bounds = [shard_bounds(n, world, r) for r in range(world)]
        self.prev_bounds = None       # bounds of the previous call (what the stopwatch timed)
        self.pending = []             # (step, host copy of the ranks' times, bounds they timed)
        self.step = 0
        self.watch = None

    def push(self, times, timed_bounds):
        """Queue the all-reduced per-rank times of this call (a device vector) for a later call:
        an asynchronous copy into pinned memory, never a sync."""
        if times.is_cuda:
            host = torch.empty(times.shape, dtype=times.dtype, pin_memory=True)
            host.copy_(times, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        else:
            host, ev = times.clone(), None
        self.pending.append((self.step, host, ev, timed_bounds))

    def settle(self):
        """Fold in the times all-reduced TWO calls ago.  A fixed lag, not "whatever has arrived",
        because every rank must apply the same updates at the same step to derive the same
        bounds; two calls back the copy has always landed (the host runs at most one step
        ahead of the device through the rewards it returns), so the wait below is a formality."""
        self.step += 1
        while self.pending and self.pending[0][0] <= self.step - 2:
            _, host, ev, timed_bounds = self.pending.pop(0)
            if ev is not None:
                ev.synchronize()
            self.update(host.tolist(), timed_bounds)

    def update(self, times_ms, timed_bounds):
        """`times_ms[r]` = rank r's block time under `timed_bounds` (0 = not measured)."""
        if timed_bounds is None or any(t <= 0.0 for t in times_ms):
            return
        self.log.append(([b - a for a, b in timed_bounds], [round(t, 3) for t in times_ms]))
        if self.observe_only:
            return
        rates = [(b - a) / t for (a, b), t in zip(timed_bounds, times_ms)]
        tot = sum(rates)
        target = [r / tot for r in rates]
        self.shares = [(1 - self.gain) * s + self.gain * t for s, t in zip(self.shares, target)]
        self.bounds = balanced_bounds(self.n, self.shares, self.quantum, self.min_rows)


_balances = {}


def query_balance(group, n, world, k):
    key = (id(group) if group is not None else 0, n, world, k)
    if key not in _balances:
        _balances[key] = QueryBalance(n, world, min_rows=k)
    return _balances[key]


last_query_block = {}     # {"lo", "hi"} of this rank's query block in the last balanced call


def reps_from_host(rep_host, out=None, group=None):
    """Full (N, c) device copy of host-resident reps for a sharded relabel: every rank copies only
    ITS rows over PCIe (pinned host memory -> asynchronous DMA) and the shards are all-gathered over
    NVLink, instead of G ranks each pulling all N rows through the host links."""
    distributed = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if distributed else 1
    rank = dist.get_rank(group) if distributed else 0
    n = rep_host.size(0)
    if out is None:
        out = torch.empty(rep_host.shape, dtype=rep_host.dtype,
                          device=torch.device("cuda", torch.cuda.current_device()))
    lo, hi = shard_bounds(n, world, rank)
    out[lo:hi].copy_(rep_host[lo:hi], non_blocking=True)
    if world > 1:
        if n % world == 0:
            dist.all_gather_into_tensor(out, out[lo:hi], group=group)     # in place
        else:
            for r in range(world):            # ragged shards: one broadcast per owner
                a, b = shard_bounds(n, world, r)
                dist.broadcast(out[a:b], src=dist.get_global_rank(group, r) if group else r,
                               group=group)
    return out


def sample_parts(n, world, stride):
    """The rows 0, stride, 2 stride, ... of an n-row set split by the ranks' shards:
    (first sample row, number of sample rows) inside shard_bounds(n, world, r) for every r."""
    first = [-(-shard_bounds(n, world, r)[0] // stride) * stride for r in range(world)]
    counts = [max(0, -(-(shard_bounds(n, world, r)[1] - first[r]) // stride))
              for r in range(world)]
    return first, counts


def feed_chunk_cuts(rows):
    """Chunk boundaries [0, ..., rows] of a host upload that the pre-pass follows: a small first
    chunk (the pre-pass starts after ~1 ms of PCIe instead of after 1/8 of the rows), then sizes
    that grow by 1.4x: one GPU uploads 27 rows/us (55 GB/s, 2 KB rows) and pre-passes 19 rows/us
    at 1M rows, so chunk i + 1 has arrived when the pre-pass of chunk i ends as long as the sizes
    grow by less than 27 / 19 — capped at 256k rows."""
    if rows < 120000:
        return [0, rows]
    cuts, size = [0], 32768
    while rows - cuts[-1] > 3 * size // 2:
        cuts.append(cuts[-1] + size)
        size = min(size * 7 // 5, 262144)
    cuts.append(rows)
    return cuts


class _HostFeed:
    """Rows of a pinned host matrix on their way to `rep_out` while the symmetric path already
    works (relabel_apt_reward_from_host): the sample (every stride-th row) goes first as one 2-D
    DMA, then this rank's rows in chunks on a copy stream with an event each — the sample
    pre-pass of a chunk waits for that chunk only — and, for several ranks, the all-gather of the
    fp32 rows over NVLink behind the last chunk; only the symmetric pass waits for everything."""

    def __init__(self, rep_host, rep_out, group, world, rank):
        self.h, self.d = rep_host, rep_out
        self.n, self.dim = rep_host.shape
        self.dev = rep_out.device
        self.group, self.world, self.rank = group, world, rank
        self.lo, self.hi = shard_bounds(self.n, world, rank)
        self.copy = _side_stream(self.dev)
        self.copy.wait_stream(torch.cuda.current_stream(self.dev))   # earlier readers of rep_out
        self.events, self.gather = [], None

    @staticmethod
    def usable(rep_host, n, world):
        return (rep_host.is_pinned() and rep_host.dtype == torch.float32
                and rep_host.is_contiguous() and rep_host.dim() == 2 and n % world == 0)

    def sample(self, stride):
        """rep[::stride] on the device: every rank uploads the sample rows of ITS shard and the
        parts are all-gathered (padded to the largest part)."""
        lib, dev, world, dim = _lib.load(), self.dev, self.world, self.dim
        first, counts = sample_parts(self.n, world, stride)
        s_max = max(counts)
        buf = torch.empty((world, s_max, dim), dtype=torch.float32, device=dev)
        mine = buf[self.rank]
        with torch.cuda.stream(self.copy):
            _lib.check(lib.exb_h2d_rows_2d(
                _lib.ptr(mine), ctypes.c_void_p(self.h.data_ptr() + first[self.rank] * dim * 4),
                counts[self.rank], dim * 4, stride * dim * 4,
                ctypes.c_void_p(self.copy.cuda_stream)))
            ev = torch.cuda.Event()
            ev.record(self.copy)
        torch.cuda.current_stream(dev).wait_event(ev)
        if world == 1:
            return buf[0]
        dist.all_gather_into_tensor(buf.view(world * s_max, dim), mine, group=self.group)
        if all(c == s_max for c in counts):
            return buf.view(world * s_max, dim)
        return torch.cat([buf[r, :counts[r]] for r in range(world)], dim=0)

    def start(self, qlo, qhi):
        """Enqueue the upload of this rank's shard plus the rows of its query block [qlo, qhi)
        outside it (tile alignment: < 256 rows; the all-gather rewrites them with the same
        bytes) and the all-gather behind it.  -> the chunks of [qlo, qhi) in arrival order."""
        a, b = min(self.lo, qlo), max(self.hi, qhi)
        cuts = [a + c for c in feed_chunk_cuts(b - a)]
        with torch.cuda.stream(self.copy):
            for ca, cb in zip(cuts[:-1], cuts[1:]):
                self.d[ca:cb].copy_(self.h[ca:cb], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy)
                self.events.append((ca, cb, ev))
            if self.world > 1:
                self.gather = dist.all_gather_into_tensor(self.d, self.d[self.lo:self.hi],
                                                          group=self.group, async_op=True)
        return [(max(ca, qlo), min(cb, qhi)) for ca, cb, _ in self.events
                if max(ca, qlo) < min(cb, qhi)]

    def wait(self, a, b):
        cur = torch.cuda.current_stream(self.dev)
        for ca, cb, ev in self.events:
            if ca < b and cb > a:
                cur.wait_event(ev)

    def wait_all(self):
        self.wait(0, self.n)
        if self.gather is not None:
            self.gather.wait()
            self.gather = None


class StageTimer:
    """Optional CUDA-event stopwatch for the relabel stages (bench.py --stage-times)."""

    def __init__(self):
        self.marks = []

    def mark(self, name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.marks.append((name, e))

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for (_, a), (name, b) in zip(self.marks[:-1], self.marks[1:]):
            out[name] = out.get(name, 0.0) + a.elapsed_time(b)
        return out


def _finish_block(d_mine, i_mine, n, lo, hi, in_split, world, k, avg, use_rms, clip, rms, group,
                  stages, tm, return_knn, balance=None, prev_ms=0.0):
    """Global RMS update from all-reduced sums, epilogue of this rank's query block, rewards
    all-gathered so that every rank returns the full (N, 1).  With `balance`, this rank's block
    time of the previous step rides along in the same all-reduce and re-sizes the next blocks."""
    my_n = hi - lo
    sums = None
    if balance is not None:
        rank = dist.get_rank(group)
        vec = torch.zeros(4 + world, dtype=torch.float64, device=d_mine.device)
        if use_rms:
            vec[:3] = stages.sums(d_mine, k, avg)               # {count, sum, sumsq} of my block
        vec[4 + rank] = prev_ms
        dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
        sums = vec[:3] if use_rms else None
        times = vec[4:]
    elif use_rms:
        sums = stages.sums(d_mine, k, avg)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    tm("rms_allreduce")
    r_mine = stages.finish(d_mine, k, avg, use_rms, clip, sums, rms)
    tm("finish")       # `finish` also advances rms.n by the global count (sums[0])
    # variable block sizes: gather padded to the largest block
    mx = max(in_split)
    pad = torch.zeros((mx, 1), dtype=r_mine.dtype, device=r_mine.device)
    pad[:my_n] = r_mine
    gathered = torch.empty((world * mx, 1), dtype=r_mine.dtype, device=r_mine.device)
    dist.all_gather_into_tensor(gathered, pad, group=group)
    reward = torch.cat([gathered[r * mx:r * mx + in_split[r]] for r in range(world)], dim=0)
    tm("reward_allgather")
    if balance is not None:
        # read at the start of the next call (no sync in this one); the times are those of the
        # PREVIOUS call's blocks
        balance.push(times, balance.prev_bounds)
    if return_knn:
        return reward, d_mine, i_mine
    return reward


# ------------------------------------------------------------------------------------------
# Symmetric screen over several GPUs (csrc/knn_sym.cuh, include/exorl_b200.h "three calls")
# ------------------------------------------------------------------------------------------
# Exchange of the symmetric screen's transposed hits between the ranks: True = inside the screen
# kernel over NVLink peer memory (exb_knn_sym_screen_peer_f32), False = counting partition +
# all_to_all_single after it.  Falls back to the latter when peer mapping is refused.
SYM_PEER_MEMORY = True
# relabel_apt_reward_from_host on the symmetric path: True = upload in chunks behind the sample
# pre-pass (_HostFeed), False = upload (+ all-gather), then the device path.
SYM_HOST_PIPELINE = True
# Symmetric screen over several GPUs: size the ranks' tile blocks by the rate each GPU's screen
# showed in the previous step (sym_balanced_blocks) instead of equally.
SYM_RATE_BALANCE = True
# ... and prepare the screen's fp16 operand N / G rows per rank + all-gather (behind the sample
# pre-pass) instead of all N rows on every rank (device-resident reps, equal shards).
SYM_SHARDED_PREP = True
SYM_BALANCE_MAX_SKEW = 0.12


def sym_tile_blocks(n, world, tile=256):
    """Tile-aligned row blocks: rank r owns the query tile pairs [tb[r], tb[r + 1]) (tiles of
    `tile` rows) = the rows [min(n, tile * tb[r]), min(n, tile * tb[r + 1]))."""
    tiles = (n + tile - 1) // tile
    tb = [(r * tiles) // world for r in range(world + 1)]
    return tiles, tb, [min(n, tile * t) for t in tb]


def sym_span(tb, world, rank):
    """(pair_lo, pair_hi, span_end, mid_pair, extra_tiles) of `rank`: its pairs scan their own
    block from the diagonal on, then the next blocks cyclically — (world - 1) // 2 whole blocks
    and, for an even world, half of the antipodal block pair: ranks below world / 2 take the
    first half of the antipodal block's tiles with ALL their pairs, the others take the WHOLE
    antipodal block with the second half of their pairs only.  Every unordered tile pair is
    multiplied exactly once (tests/test_relabel_gloo.py checks the coverage)."""
    size = [tb[r + 1] - tb[r] for r in range(world)]
    lo, hi = tb[rank], tb[rank + 1]
    whole = (world - 1) // 2
    span_end = hi + sum(size[(rank + s) % world] for s in range(1, whole + 1))
    mid, extra = hi, 0
    if world % 2 == 0 and world > 1:
        anti = (rank + world // 2) % world
        if rank < world // 2:
            span_end += size[anti] // 2                    # first half of its tiles, every pair
        else:
            half_anti_of_me = size[rank] // 2              # what rank `anti` takes of MY tiles
            mid = lo + half_anti_of_me                     # ... so my pairs from there on take
            extra = size[anti]                             # the whole antipodal block
    return lo, hi, span_end, mid, extra


def _sym_work(size, world, r):
    """Tile products of rank r under sym_span for (real-valued) block sizes `size`."""
    whole = (world - 1) // 2
    w = size[r] * size[r] / 2 + size[r] * sum(size[(r + s) % world] for s in range(1, whole + 1))
    if world % 2 == 0 and world > 1:
        anti = (r + world // 2) % world
        w += size[r] * size[anti] / 2       # either all pairs x half block or half the pairs x block
    return w


def sym_balanced_blocks(tiles, world, ms, max_skew=SYM_BALANCE_MAX_SKEW):
    """Tile blocks tb[0..world] sized so that every rank's symmetric screen takes the same time,
    from the time per tile product `ms[r]` each rank's screen needed in the previous step: B200s
    under their power cap differ by several per cent in sustained clock, persistently, and the
    slowest rank sets the step time.  Sizes
    stay within (1 +- max_skew) of the equal split (the peer hit buffers are sized for that).
    Pure and deterministic: every rank computes the same blocks from the all-gathered times."""
    eq = tiles / world
    if world == 1 or min(ms) <= 0 or eq < 8:
        return [(r * tiles) // world for r in range(world + 1)]
    rate = [1.0 / m for m in ms]
    mean_rate = sum(rate) / world
    rate = [x / mean_rate for x in rate]
    lo_s, hi_s = eq * (1 - max_skew), eq * (1 + max_skew)
    size = [eq] * world
    for _ in range(60):
        t = [_sym_work(size, world, r) / rate[r] for r in range(world)]
        mean_t = sum(t) / world
        size = [sz * (mean_t / tt) for sz, tt in zip(size, t)]
        for _ in range(world + 1):        # back onto {sum = tiles, lo_s <= size <= hi_s}
            size = [min(hi_s, max(lo_s, sz)) for sz in size]
            excess = tiles - sum(size)
            free = [i for i, sz in enumerate(size) if (sz < hi_s if excess > 0 else sz > lo_s)]
            if abs(excess) < 1e-9 or not free:
                break
            for i in free:
                size[i] += excess / len(free)
    cuts = [0.0]
    for sz in size:
        cuts.append(cuts[-1] + sz)
    tb = [int(round(c)) for c in cuts]
    tb[0], tb[-1] = 0, tiles
    return tb


def _sym_span_work(span):
    """Tile products of a span (pair_lo, pair_hi, span_end, mid_pair, extra)."""
    lo, hi, end, mid, extra = span
    return sum(end - u + (extra if u >= mid else 0) for u in range(lo, hi))


_sym_pending = {}      # key -> (start event, end event, tile products) of the last screen
_sym_unit = {}         # key -> smoothed ms per tile product of this rank's screen


def _sym_unit_time(key):
    """This rank's screen time per tile product, from the stopwatch of the previous step with
    the same sizes (0.0 = no history yet -> equal blocks)."""
    pend = _sym_pending.pop(key, None)
    if pend is not None:
        e0, e1, work = pend
        e1.synchronize()                  # long done: the previous step's results were read
        u = e0.elapsed_time(e1) / max(work, 1)
        old = _sym_unit.get(key)
        # first measurement as it is, then a slow average: the rates differ persistently from GPU
        # to GPU (what the blocks should follow) AND fluctuate by a few per cent from step to step
        # (what they should not)
        _sym_unit[key] = u if old is None else 0.7 * old + 0.3 * u
    return _sym_unit.get(key, 0.0)


class _CudaSymBackend:
    """The device side of the symmetric screen over several ranks (C entries exb_knn_sym_*).
    tests/test_relabel_gloo.py swaps in a CPU stand-in with the same contract to exercise the
    partition / exchange logic of `_sym_queries_knn` under gloo."""
    tile = 256

    def __init__(self, rep, k):
        from . import knn as xknn
        self.xknn, self.lib = xknn, _lib.load()
        self.rep, self.k = rep, k
        self.n, self.dim = rep.shape
        self.dev = rep.device

    def sample_rows(self):
        return self.xknn._sym_sample_rows(self.n)

    def sample_distances(self, lo, hi, sample):
        d, _ = self.xknn._knn_topk_impl(self.rep[lo:hi], sample, self.k, 0, "screened", False,
                                        allow_sym=False)
        return d[:, self.k - 1]

    def room(self, expected):
        self.log_cap = int(-(-(3 * expected / 2 + 96) // 8) * 8)
        self.recv_cap = int(-(-(3 * expected + 96) // 8) * 8)
        self.need = self.lib.exb_knn_sym_workspace_bytes(self.n, self.dim, self.k, self.log_cap,
                                                         self.recv_cap)
        return bool(self.need) and self.xknn._sym_fits(self.dev, self.need)

    def screen(self, d_all, span):
        """-> (target, val, src) int32 / f32 / int32 device vectors of the transposed hits this
        rank found (for ANY row), and whether the append buffer overflowed.  Synchronises."""
        lib, dev = self.lib, self.dev
        pair_lo, pair_hi, span_end, mid, extra = span
        with torch.cuda.device(dev):
            self.ws, _ = _lib.Workspace.get(dev, "knn_sym", self.need)
            ws = self.ws
            hits = _lib.SymHits()
            _lib.check(lib.exb_knn_sym_screen_f32(
                _lib.ptr(self.rep), self.n, self.dim, self.k, _lib.ptr(d_all), 1, self.log_cap,
                self.recv_cap, pair_lo, pair_hi, span_end, mid, extra, _lib.ptr(ws), ws.numel(),
                ctypes.byref(hits), _lib.stream_ptr(dev)))
            base = ws.data_ptr()
            cnt = ws[hits.count - base:hits.count - base + 8].view(torch.int64)
            n_hits = int(cnt.item())                               # sync: the screen is done
            over = n_hits > hits.capacity
            n_hits = min(n_hits, hits.capacity)

            def view(ptr, dtype):
                off = ptr - base
                return ws[off:off + 4 * n_hits].view(dtype)
            return (view(hits.target, torch.int32), view(hits.val, torch.float32),
                    view(hits.src, torch.int32)), over

    # ---- peer-memory exchange: the screen kernel appends every hit straight into the buffer of
    # the rank that owns its query row (NVLink stores), so no all-to-all follows the screen
    _peer = {}     # device index -> {"cap", "own" (ptr), "handle", "mapped" {rank: ptr}, "hits"}

    @staticmethod
    def _hits_at(base, cap):
        """exb_sym_hits of a peer buffer: count word, then target / val / src arrays of `cap`."""
        return _lib.SymHits(base + 256, base + 256 + 4 * cap, base + 256 + 8 * cap, base, cap)

    def open_peers(self, group, world, rank, rows_max):
        """Make sure every rank owns a peer-visible hit buffer for `rows_max` rows and that all
        are mapped here; zero the own count on the current stream.  The capacity is the same
        function of (n, world, recv_cap) on every rank, so all ranks (re)allocate together and
        the handle exchange (all_gather_object) is collective.  False when the hardware or the
        driver refuses peer mapping (the caller falls back to the all-to-all path)."""
        lib, dev = self.lib, self.dev
        cap = -(-int(rows_max * (self.recv_cap // 3 + 16)) // 64) * 64
        st = self._peer.get(dev.index)
        with torch.cuda.device(dev):
            if st is None or st["cap"] < cap or st["world"] != world:
                if st is not None:
                    torch.cuda.synchronize(dev)
                    for q in st["mapped"].values():
                        lib.exb_peer_buffer_close(q)
                    dist.barrier(group=group)                    # every peer unmapped
                    lib.exb_peer_buffer_free(st["own"])
                    self._peer.pop(dev.index)
                own = ctypes.c_void_p()
                handle = ctypes.create_string_buffer(64)
                rc = lib.exb_peer_buffer_alloc(256 + 12 * cap, ctypes.byref(own), handle)
                mine = (rc, handle.raw, cap)
                everyone = [None] * world
                dist.all_gather_object(everyone, mine, group=group)
                ok = all(e[0] == 0 and e[2] == cap for e in everyone)
                mapped = {}
                if ok:
                    for r, (_, h, _) in enumerate(everyone):
                        if r == rank:
                            continue
                        q = ctypes.c_void_p()
                        if lib.exb_peer_buffer_open(h, ctypes.byref(q)) != 0:
                            ok = False
                            break
                        mapped[r] = q.value
                flag = [None] * world
                dist.all_gather_object(flag, ok, group=group)
                if not all(flag):
                    for q in mapped.values():
                        lib.exb_peer_buffer_close(q)
                    dist.barrier(group=group)
                    if rc == 0:
                        lib.exb_peer_buffer_free(own)
                    return False
                hits = (_lib.SymHits * world)()
                for r in range(world):
                    hits[r] = self._hits_at(own.value if r == rank else mapped[r], cap)
                st = {"cap": cap, "world": world, "own": own.value, "mapped": mapped,
                      "hits": hits, "rank": rank}
                self._peer[dev.index] = st
            _lib.check(lib.exb_peer_buffer_clear(st["own"], 256, _lib.stream_ptr(dev)))
        self.peer = st
        return True

    operand_work = None    # pending all-gathers of a sharded operand preparation

    def prepare_shard(self, group, world, rank):
        """Operand of the symmetric screen, every rank its N / world rows: one affine record from
        all-reduced column statistics, exb_knn_prep_affine_f16 straight into the workspace's
        operand arrays, the shards all-gathered over NVLink (asynchronously: the sample pre-pass
        runs meanwhile; `screen_peer` waits) and the scalars MAX-reduced — instead of every rank
        preparing all N rows.  Needs equal shards."""
        lib, dev, n, dim = self.lib, self.dev, self.n, self.dim
        if n % world != 0:
            return False
        lo, hi = shard_bounds(n, world, rank)
        with torch.cuda.device(dev):
            self.ws, _ = _lib.Workspace.get(dev, "knn_sym", self.need)
            ws = self.ws
            slots, aff_p = _lib.Operand(), ctypes.c_void_p()
            _lib.check(lib.exb_knn_sym_operand_slots(n, dim, self.k, self.log_cap, self.recv_cap,
                                                     _lib.ptr(ws), ws.numel(),
                                                     ctypes.byref(slots), ctypes.byref(aff_p)))
            base = ws.data_ptr()
            dpad = -(-dim // 64) * 64

            def view(ptr, nbytes, dtype):
                return ws[ptr - base:ptr - base + nbytes].view(dtype)
            f16 = view(slots.f16, n * dpad * 2, torch.float16).view(n, dpad)
            hn = view(slots.half_norm, n * 4, torch.float32)
            er = view(slots.round_err, n * 4, torch.float32)
            sc = view(slots.scalars, 16, torch.int32)
            aff = view(aff_p.value, _lib.AFFINE_FLOATS * 4, torch.float32)
            sc.zero_()
            aff.copy_(column_affine(self.rep[lo:hi], distributed=True, group=group, n_total=n))
            _lib.check(lib.exb_knn_prep_affine_f16(
                _lib.ptr(self.rep[lo:hi]), hi - lo, dim, _lib.ptr(aff), _lib.ptr(f16[lo:hi]),
                _lib.ptr(hn[lo:hi]), _lib.ptr(er[lo:hi]), _lib.ptr(sc), _lib.stream_ptr(dev)))
            self.operand_work = [
                dist.all_gather_into_tensor(f16, f16[lo:hi], group=group, async_op=True),
                dist.all_gather_into_tensor(hn, hn[lo:hi], group=group, async_op=True),
                dist.all_gather_into_tensor(er, er[lo:hi], group=group, async_op=True),
                dist.all_reduce(sc, op=dist.ReduceOp.MAX, group=group, async_op=True)]
        return True

    def screen_peer(self, d_all, span, tile_lo):
        """The screen with the exchange inside the kernel; asynchronous."""
        lib, dev, st = self.lib, self.dev, self.peer
        pair_lo, pair_hi, span_end, mid, extra = span
        with torch.cuda.device(dev):
            self.ws, _ = _lib.Workspace.get(dev, "knn_sym", self.need)
            ws = self.ws
            ready = self.operand_work is not None
            if ready:
                for w in self.operand_work:
                    w.wait()
                self.operand_work = None
            tl = (ctypes.c_int32 * len(tile_lo))(*tile_lo)
            _lib.check(lib.exb_knn_sym_screen_peer_f32(
                _lib.ptr(self.rep), self.n, self.dim, self.k, _lib.ptr(d_all), 1, self.log_cap,
                self.recv_cap, pair_lo, pair_hi, span_end, mid, extra, st["world"], st["hits"],
                tl, int(ready), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))

    def scatter_rerank(self, r_t, r_v, r_s, lo, hi):
        """Received hits into the rows' lists, then the exact top-k of rows [lo, hi).
        r_t None: the hits are this rank's peer buffer (count on the device).
        -> (dist, idx, undecided_rows, status_flags)"""
        lib, dev, k = self.lib, self.dev, self.k
        with torch.cuda.device(dev):
            if r_t is None:
                st = self.peer
                _lib.check(lib.exb_knn_sym_scatter_hits(
                    self.n, self.dim, k, self.log_cap, self.recv_cap,
                    ctypes.byref(st["hits"][st["rank"]]), _lib.ptr(self.ws), _lib.stream_ptr(dev)))
            else:
                _lib.check(lib.exb_knn_sym_scatter(self.n, self.dim, k, self.log_cap,
                                                   self.recv_cap, _lib.ptr(r_t), _lib.ptr(r_v),
                                                   _lib.ptr(r_s), r_t.numel(), _lib.ptr(self.ws),
                                                   _lib.stream_ptr(dev)))
            d_mine = torch.empty(hi - lo, k, dtype=torch.float32, device=dev)
            i_mine = torch.empty(hi - lo, k, dtype=torch.int32, device=dev)
            _lib.check(lib.exb_knn_sym_rerank_f32(
                _lib.ptr(self.rep), self.n, self.dim, k, 0, self.log_cap, self.recv_cap, lo, hi,
                _lib.ptr(d_mine), _lib.ptr(i_mine), _lib.ptr(self.ws), _lib.stream_ptr(dev)))
        st = self.xknn.screened_status(dev, key="knn_sym")
        return d_mine, i_mine, st[0], st[1]

    def repair(self, lo, hi, d_mine, i_mine):
        rows = (i_mine[:, 0] < 0).nonzero().flatten()
        self.xknn._recompute_rows(self.rep[lo:hi], self.rep, self.k, 0, d_mine, i_mine, rows)


def _sym_queries_knn(rep, k, world, rank, group, tm, backend=None, feed=None):
    """kNN of this rank's (tile-aligned) query block through the symmetric screen spread over
    the ranks.  Returns (dist, idx, row_bounds) or None when any rank reports the pass invalid
    (the caller then runs the full screen).  Collectives: d_sample all-reduce (4N bytes); the
    hits (12 bytes each, ~ k N^2 / (2 S world) per rank) travel INSIDE the screen kernel — it
    appends each one to the owner rank's peer-mapped buffer over NVLink — with two one-word
    all-reduces as barriers around it; without peer memory (SYM_PEER_MEMORY off, mapping
    refused, CPU stand-in): hit counts all-to-all + the hits all-to-all after a partition."""
    from . import knn as xknn
    be = backend if backend is not None else _CudaSymBackend(rep, k)
    n = rep.size(0)
    dev = rep.device
    tiles, tb, rb = sym_tile_blocks(n, world, be.tile)
    stride = max(1, n // be.sample_rows())
    n_sample = -(-n // stride)
    # one collective + one host sync up front: "does the workspace fit everywhere" and, for the
    # rate balance, the time per tile product every rank's screen needed in the PREVIOUS step
    rebalance = SYM_RATE_BALANCE and rep.is_cuda and world > 1
    key = (n, rep.size(1), k, world, rank)
    unit = _sym_unit_time(key) if rebalance else 0.0
    mine = torch.tensor([1.0 if be.room(1.3 * k * n / n_sample) else 0.0, unit],
                        dtype=torch.float64, device=dev)
    everyone = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(everyone, mine, group=group)
    everyone = torch.stack(everyone).tolist()
    if not all(e[0] for e in everyone):                               # all ranks or none
        xknn.sym_last.update(n=n, no_room=True)
        return None
    units = [e[1] for e in everyone]
    if rebalance and min(units) > 0:
        # B200s under their power cap sustain clocks that differ by several per cent from GPU to
        # GPU, persistently: slow ranks get fewer tiles so that the screens end together
        tb = sym_balanced_blocks(tiles, world, units)
        rb = [min(n, be.tile * t) for t in tb]
    lo, hi = rb[rank], rb[rank + 1]
    xknn.sym_last.update(rows_per_rank=[rb[r + 1] - rb[r] for r in range(world)],
                         screen_ns_per_tile_pair=[round(1e6 * u, 3) for u in units])
    rows_bound = -(-n // world)
    if rebalance:                         # blocks may grow by SYM_BALANCE_MAX_SKEW (+ rounding)
        rows_bound = int(rows_bound * (1 + SYM_BALANCE_MAX_SKEW))
    rows_bound += 2 * be.tile
    # the own hit buffer is cleared on this stream BEFORE the all-reduce below, which therefore
    # doubles as the barrier "nobody appends to a buffer that is not cleared yet"
    peer = (SYM_PEER_MEMORY and hasattr(be, "open_peers") and world <= _lib.SYM_MAX_RANKS
            and be.open_peers(group, world, rank, rows_bound))
    if peer and feed is None and SYM_SHARDED_PREP:
        be.prepare_shard(group, world, rank)      # all-gathers run behind the pre-pass
    d_all = torch.zeros(n, dtype=torch.float32, device=dev)
    if feed is None:
        sample = rep[::stride].contiguous()
        d_all[lo:hi] = be.sample_distances(lo, hi, sample)
    else:
        # rows still arriving from the host (_HostFeed): pre-pass chunk by chunk behind the upload
        sample = feed.sample(stride)
        for a, b in feed.start(lo, hi):
            feed.wait(a, b)
            d_all[a:b] = be.sample_distances(a, b, sample)
    dist.all_reduce(d_all, op=dist.ReduceOp.SUM, group=group)
    if feed is not None:
        feed.wait_all()                   # the symmetric pass reads every row
    tm("sym_sample_prepass")
    span = sym_span(tb, world, rank)
    if rebalance:                         # stopwatch around this rank's screen, read next step
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    if peer:
        be.screen_peer(d_all, span, tb)
        if rebalance:
            e1.record()
            _sym_pending[key] = (e0, e1, _sym_span_work(span))
        bad = torch.zeros(1, dtype=torch.int64, device=dev)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=group)   # barrier: every screen is done
        tm("sym_screen")
        d_mine, i_mine, und, status = be.scatter_rerank(None, None, None, lo, hi)
        bad[0] = 1 if (status & 2) else 0
        dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=group)
        xknn.sym_last.update(n=n, sample_rows=sample.size(0), undecided=und, status=status,
                             world=world, exchange="peer_memory")
        if int(bad.item()):
            xknn.sym_last["fell_back_to_full_screen"] = True
            return None
        if status & 1:
            be.repair(lo, hi, d_mine, i_mine)
        tm("sym_rerank")
        return d_mine, i_mine, rb
    (tgt, val, src), over = be.screen(d_all, span)
    if rebalance:
        e1.record()
        _sym_pending[key] = (e0, e1, _sym_span_work(span))
    tm("sym_screen")
    bad = torch.tensor([1 if over else 0], device=dev)
    # route every hit to the rank that owns its query row (a counting partition: one mask +
    # compaction per destination is cheaper than a sort of 10^7..10^8 keys)
    edges = torch.tensor(rb[1:-1], dtype=torch.int32, device=dev)
    dest = torch.bucketize(tgt, edges, right=True)
    packed = torch.stack([tgt, val.view(torch.int32), src], dim=1)
    parts = [packed[(dest == r).nonzero().flatten()] for r in range(world)]
    sc = [p.size(0) for p in parts]
    counts = torch.tensor(sc, dtype=torch.int64, device=dev)
    recv_counts = torch.empty_like(counts)
    dist.all_to_all_single(recv_counts, counts, group=group)
    rc_ = recv_counts.tolist()
    recv = torch.empty((sum(rc_), 3), dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv, torch.cat(parts, dim=0), rc_, sc, group=group)
    tm("sym_hit_exchange")
    d_mine, i_mine, und, status = be.scatter_rerank(
        recv[:, 0].contiguous(), recv[:, 1].contiguous().view(torch.float32),
        recv[:, 2].contiguous(), lo, hi)
    bad[0] |= 1 if (status & 2) else 0
    dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=group)
    xknn.sym_last.update(n=n, sample_rows=sample.size(0), undecided=und, status=status,
                         world=world, hits_sent=int(tgt.numel()), exchange="all_to_all")
    if int(bad.item()):
        xknn.sym_last["fell_back_to_full_screen"] = True
        return None
    if status & 1:
        be.repair(lo, hi, d_mine, i_mine)
    tm("sym_rerank")
    return d_mine, i_mine, rb


def relabel_apt_reward(rep, args, rms=None, group=None, exchange="auto", mode="auto",
                       return_knn=False, stages=_CudaStages, timer=None, balance=False):
    """APT reward (N, 1) of every row of `rep` against all N rows, identical on every rank.

    `rep` is the (N, c) fp32 rep tensor or a PreparedRep (exorl_b200.rep: e.g. icm_get_rep of the
    dataset's observations — the trunk then emits the kNN operand itself and the relabel never
    re-reads the rep to prepare it).

    `args` exposes knn_k, knn_avg, knn_rms, knn_clip like PBE; `rms` is the running-moment state
    (updated once with bs = N*k in avg mode, N otherwise).  Works without an initialised process
    group (single shard).  `balance`: in the 'queries' partition, size the per-rank query blocks
    by the ranks' measured rates (QueryBalance) instead of equally — off by default: on 8 x B200
    the per-rank times fluctuate from step to step (power cap) rather than differing persistently,
    so following them loses (profiles/r02b_balance_observe_n8.json); "observe" only records the
    times.  With `return_knn` the rows of the returned block are `relabel.last_query_block`."""
    if exchange not in ("auto", "lists", "all_gather", "all_to_all", "queries"):
        raise ValueError("exchange must be 'auto', 'lists', 'all_gather', 'all_to_all' or "
                         "'queries'")
    if rms is None:
        rms = getattr(args, "rms", None)
    k, avg, use_rms, clip = int(args.knn_k), bool(args.knn_avg), bool(args.knn_rms), args.knn_clip
    distributed = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if distributed else 1
    rank = dist.get_rank(group) if distributed else 0
    n = rep.size(0)
    lo, hi = shard_bounds(n, world, rank)
    if hi - lo < k:
        raise RuntimeError("selected index k out of range")   # a shard smaller than k rows

    tm = timer.mark if timer is not None else (lambda name: None)
    tm("start")
    if world > 1 and exchange == "auto":
        # every rank holds the whole set (it is the query set), so the cheapest partition is by
        # query block: no candidate exchange, no merge, and each row's k-th-best chain sees the
        # whole bank.  Measured on B200: 95.5 ms per rank at G = 8 against 106.6 ms for its bank
        # shard (+ exchange); the bank-sharded forms stay for banks that do not fit one GPU.
        exchange = "queries"
    if world > 1 and exchange == "lists":
        cnt = 0
        if mode != "exact" and hasattr(stages, "lists_available"):
            cnt = int(stages.lists_available(n, hi - lo, _fp32(rep)[0].numel(), k, world))
        # every rank must take the same branch with the same list count (shard sizes differ by
        # up to one row, which can change the kernel variant): agree on min and max
        flag = torch.tensor([cnt, -cnt], device=_fp32(rep).device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        lo_cnt, hi_cnt = (int(v) for v in flag.tolist())
        ok = lo_cnt > 0 and lo_cnt == -hi_cnt
        if exchange == "lists" and not ok:
            raise ValueError("exchange='lists' is not available for these sizes / this mode")
        exchange = "lists" if ok else "all_to_all"
    elif exchange == "auto":
        exchange = "all_to_all"

    sym_backend = getattr(stages, "sym_backend", None)     # tests: CPU stand-in (gloo)
    if (world > 1 and exchange == "queries" and (stages is _CudaStages or sym_backend)
            and mode != "exact" and not balance and not isinstance(rep, PreparedRep)):
        from . import knn as xknn
        if sym_backend or (
                xknn.SYM_ENABLED and n >= xknn.SYM_MIN_ROWS and xknn.OVERFLOW_POLICY == "sync"
                and rep.size(1) <= _lib.SCREEN_MAX_DIM and k <= _lib.SCREEN_MAX_K):
            # a set against itself: every unordered tile pair on exactly one rank (K1s)
            rows = _as_rows_f32(rep)
            out = _sym_queries_knn(rows, k, world, rank, group, tm,
                                   backend=sym_backend(rows, k) if sym_backend else None)
            if out is not None:
                d_mine, i_mine, rb = out
                lo, hi = rb[rank], rb[rank + 1]
                last_query_block.update(lo=lo, hi=hi)
                in_split = [rb[r + 1] - rb[r] for r in range(world)]
                return _finish_block(d_mine, i_mine, n, lo, hi, in_split, world, k, avg, use_rms,
                                     clip, rms, group, stages, tm, return_knn)
    if world > 1 and exchange == "queries":
        bal = query_balance(group, n, world, k) if balance else None
        if bal is not None:
            bal.observe_only = balance == "observe"
            bal.settle()
        bounds = bal.bounds if bal else [shard_bounds(n, world, r) for r in range(world)]
        in_split = [b - a for a, b in bounds]
        lo, hi = bounds[rank]
        last_query_block.update(lo=lo, hi=hi)
        prev_ms = 0.0
        if bal is not None:
            if bal.watch is None and hasattr(stages, "stopwatch"):
                bal.watch = stages.stopwatch()
            if bal.watch is not None:
                prev_ms = bal.watch.elapsed_ms()     # last step's block, under bal.timed_bounds
                bal.watch.start()
        d_mine, i_mine = stages.block_knn(rep, lo, hi, k, mode)
        if bal is not None and bal.watch is not None:
            bal.watch.stop()
        tm("block_knn")
        out = _finish_block(d_mine, i_mine, n, lo, hi, in_split, world, k, avg, use_rms, clip,
                            rms, group, stages, tm, return_knn, balance=bal, prev_ms=prev_ms)
        if bal is not None:
            bal.prev_bounds = bounds      # the time read at the next call was measured on these
        return out

    if world > 1 and exchange == "lists":
        bounds = [shard_bounds(n, world, r) for r in range(world)]
        in_split = [(b - a) for a, b in bounds]
        my_n = hi - lo
        val, idx, cnt, ha, ea, sc = stages.screen_lists(rep, lo, hi, k)   # (N, L, 24) / (N, L)
        tm("screen")
        lists = val.size(1)
        r_val = torch.empty((world * my_n, *val.shape[1:]), dtype=val.dtype, device=val.device)
        r_idx = torch.empty((world * my_n, *idx.shape[1:]), dtype=idx.dtype, device=idx.device)
        r_cnt = torch.empty((world * my_n, lists), dtype=cnt.dtype, device=cnt.device)
        dist.all_to_all_single(r_val, val, [my_n] * world, in_split, group=group)
        dist.all_to_all_single(r_idx, idx, [my_n] * world, in_split, group=group)
        dist.all_to_all_single(r_cnt, cnt, [my_n] * world, in_split, group=group)
        dist.all_reduce(sc, op=dist.ReduceOp.MAX, group=group)   # int32 view of >= 0 floats
        tm("all_to_all")
        d_mine, i_mine = stages.rerank_lists(
            rep, lo, hi, k, r_val.view(world, my_n, *val.shape[1:]),
            r_idx.view(world, my_n, *idx.shape[1:]), r_cnt.view(world, my_n, lists),
            [a for a, _ in bounds], ha[lo:hi], ea[lo:hi], sc)
        tm("rerank")
        return _finish_block(d_mine, i_mine, n, lo, hi, in_split, world, k, avg, use_rms, clip, rms,
                             group, stages, tm, return_knn)

    d_loc, i_loc = stages.local_knn(rep, lo, hi, k, mode)      # (N, k) exact, global ids
    tm("local_knn")
    if world == 1:
        reward = stages.reward(d_loc, k, avg, use_rms, clip, rms)
        return (reward, d_loc, i_loc) if return_knn else reward

    if exchange == "all_gather":
        d_all = torch.empty((world * n, k), dtype=d_loc.dtype, device=d_loc.device)
        i_all = torch.empty((world * n, k), dtype=i_loc.dtype, device=i_loc.device)
        dist.all_gather_into_tensor(d_all, d_loc.contiguous(), group=group)
        dist.all_gather_into_tensor(i_all, i_loc.contiguous(), group=group)
        d_glob, i_glob = stages.merge(d_all.view(world, n, k), i_all.view(world, n, k), k)
        reward = stages.reward(d_glob, k, avg, use_rms, clip, rms)   # same update on every rank
        return (reward, d_glob, i_glob) if return_knn else reward

    # all_to_all: query block b of every rank's list goes to rank b
    bounds = [shard_bounds(n, world, r) for r in range(world)]
    in_split = [(b - a) for a, b in bounds]
    my_n = hi - lo
    d_blk = torch.empty((world * my_n, k), dtype=d_loc.dtype, device=d_loc.device)
    i_blk = torch.empty((world * my_n, k), dtype=i_loc.dtype, device=i_loc.device)
    dist.all_to_all_single(d_blk, d_loc.contiguous(), [my_n] * world, in_split, group=group)
    dist.all_to_all_single(i_blk, i_loc.contiguous(), [my_n] * world, in_split, group=group)
    tm("all_to_all")
    d_mine, i_mine = stages.merge(d_blk.view(world, my_n, k), i_blk.view(world, my_n, k), k)
    tm("merge")
    return _finish_block(d_mine, i_mine, n, lo, hi, in_split, world, k, avg, use_rms, clip, rms,
                         group, stages, tm, return_knn)


_side_streams = {}


def _as_rows_f32(rep):
    rep = rep.detach()
    if rep.dtype != torch.float32:
        rep = rep.float()
    return rep.reshape(rep.size(0), -1).contiguous()


def _side_stream(dev):
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(dev)
    return _side_streams[key]


def relabel_apt_reward_from_host(rep_host, args, rms=None, group=None, rep_out=None,
                                 return_knn=False, balance=False, timer=None):
    """`relabel_apt_reward` for reps that live in (pinned) HOST memory — the end-to-end form of
    config 5: (N, c) fp32 rows in, (N, 1) rewards out, identical on every rank.

    Pipeline per rank r of G (query-block partition, whole bank on every rank):
      1. H2D of ITS N/G rows only (PCIe), 2. exb_knn_prep_affine_f16 of those rows with the
         set's affine record (column statistics of its rows, all-reduced),
      3. all-gather of the fp16 operand + row norms over NVLink (1 GB instead of the 2 GB of fp32
         rows; the scalars are MAX-reduced), 4. the tcgen05 screen of its query block against the
         whole bank starts, while 5. the fp32 rows — needed only by the exact rerank — are
         all-gathered behind it on the NCCL stream, 6. rerank, global RMS, epilogue, rewards.
    So the screen waits for N/G rows of PCIe + 1 GB of NVLink instead of for all 2 GB twice.
    Uneven shards (N % G != 0) take the plain path (reps_from_host + relabel_apt_reward)."""
    from .knn import knn_screen_lists, knn_rerank_lists
    if rms is None:
        rms = getattr(args, "rms", None)
    k, avg, use_rms, clip = int(args.knn_k), bool(args.knn_avg), bool(args.knn_rms), args.knn_clip
    distributed = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if distributed else 1
    rank = dist.get_rank(group) if distributed else 0
    n, dim = rep_host.shape
    dev = torch.device("cuda", torch.cuda.current_device()) if rep_out is None else rep_out.device
    if rep_out is None:
        rep_out = torch.empty((n, dim), dtype=torch.float32, device=dev)
    lists = screen_lists_count(max(n // world, 1), n, dim, k)
    from . import knn as xknn
    sym = (xknn.SYM_ENABLED and n >= xknn.SYM_MIN_ROWS and xknn.OVERFLOW_POLICY == "sync"
           and dim <= _lib.SCREEN_MAX_DIM and k <= _lib.SCREEN_MAX_K and not balance)
    tm = timer.mark if timer is not None else (lambda name: None)
    if sym and SYM_HOST_PIPELINE and _HostFeed.usable(rep_host, n, world):
        # symmetric screen with the upload (and the NVLink all-gather of the fp32 rows) hidden
        # behind its sample pre-pass
        with torch.cuda.device(dev):
            feed = _HostFeed(rep_host, rep_out, group, world, rank)
            tm("start")
            if world == 1:
                out = xknn._knn_topk_sym(rep_out, k, 0, feed=feed)
                tm("sym_knn")
                if out is not None:
                    reward = _CudaStages.reward(out[0], k, avg, use_rms, clip, rms)
                    return (reward, out[0], out[1]) if return_knn else reward
            else:
                out = _sym_queries_knn(rep_out, k, world, rank, group, tm, feed=feed)
                if out is not None:
                    d_mine, i_mine, rb = out
                    last_query_block.update(lo=rb[rank], hi=rb[rank + 1])
                    return _finish_block(d_mine, i_mine, n, rb[rank], rb[rank + 1],
                                         [rb[r + 1] - rb[r] for r in range(world)], world, k, avg,
                                         use_rms, clip, rms, group, _CudaStages, tm, return_knn)
            # the symmetric pass could not certify its result (or had no room): full screen
            xknn.sym_last["fell_back_to_full_screen"] = True
            if feed.events:
                feed.wait_all()
            else:
                reps_from_host(rep_host, out=rep_out, group=group)
            prev = xknn.SYM_ENABLED
            xknn.SYM_ENABLED = False
            try:
                return relabel_apt_reward(rep_out, args, rms=rms, group=group, mode="screened",
                                          return_knn=return_knn, balance=balance)
            finally:
                xknn.SYM_ENABLED = prev
    if world == 1 or sym or n % world != 0 or lists == 0 or lists > 16:
        # symmetric screen (half the tensor-core work) beats the overlap of the pipeline below
        # one GPU: nothing to pipeline against — copy, then the device path (which takes the
        # symmetric screen for a set against itself)
        rep = reps_from_host(rep_host, out=rep_out, group=group)
        return relabel_apt_reward(rep, args, rms=rms, group=group, mode="screened",
                                  return_knn=return_knn, balance=balance)
    lib = _lib.load()
    lo, hi = shard_bounds(n, world, rank)
    with torch.cuda.device(dev):
        pre = PreparedRep(n, dim, dev, rep=rep_out)
        tm("start")
        rep_out[lo:hi].copy_(rep_host[lo:hi], non_blocking=True)                        # 1
        tm("h2d_own_rows")
        # range-proof operand: one affine record (column means + power-of-two scale) for the
        # whole set — per-rank column statistics, three dim-sized all-reduces
        aff = column_affine(rep_out[lo:hi], distributed=world > 1, group=group, n_total=n)
        pre.affine = aff
        _lib.check(lib.exb_knn_prep_affine_f16(                                          # 2
            _lib.ptr(rep_out[lo:hi]), hi - lo, dim, _lib.ptr(aff), _lib.ptr(pre.f16[lo:hi]),
            _lib.ptr(pre.half_norm[lo:hi]), _lib.ptr(pre.round_err[lo:hi]),
            _lib.ptr(pre.scalars), _lib.stream_ptr(dev)))
        tm("prep_own_rows")
        fp32_rows = None
        if world > 1:
            dist.all_gather_into_tensor(pre.f16, pre.f16[lo:hi], group=group)            # 3
            dist.all_gather_into_tensor(pre.half_norm, pre.half_norm[lo:hi], group=group)
            dist.all_gather_into_tensor(pre.round_err, pre.round_err[lo:hi], group=group)
            dist.all_reduce(pre.scalars, op=dist.ReduceOp.MAX, group=group)   # >= 0 floats as i32
            side = _side_stream(dev)                                                     # 5
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                fp32_rows = dist.all_gather_into_tensor(rep_out, rep_out[lo:hi], group=group,
                                                        async_op=True)
            tm("allgather_f16_operand")
        bal = query_balance(group, n, world, k) if (balance and world > 1) else None
        if bal is not None:
            bal.observe_only = balance == "observe"
            bal.settle()
        bounds = bal.bounds if bal else [shard_bounds(n, world, r) for r in range(world)]
        qlo, qhi = bounds[rank]
        last_query_block.update(lo=qlo, hi=qhi)
        prev_ms = 0.0
        if bal is not None:
            if bal.watch is None:
                bal.watch = _CudaStopwatch()
            prev_ms = bal.watch.elapsed_ms()
            bal.watch.start()
        block = pre if (qlo, qhi) == (0, n) else pre.rows(qlo, qhi)
        val, idx, cnt, ha, ea, sc = knn_screen_lists(block, pre, k)                      # 4
        if bal is not None:
            bal.watch.stop()
        tm("screen")
        if fp32_rows is not None:
            fp32_rows.wait()              # the current stream waits for the fp32 bank rows
        tm("wait_f32_rows")
        d_mine, i_mine = knn_rerank_lists(rep_out if block is pre else rep_out[qlo:qhi], rep_out,
                                          k, val[None], idx[None], cnt[None], [0], ha, ea, sc)
        tm("rerank")
        if world == 1:
            reward = _CudaStages.reward(d_mine, k, avg, use_rms, clip, rms)
            return (reward, d_mine, i_mine) if return_knn else reward
        out = _finish_block(d_mine, i_mine, n, qlo, qhi, [b - a for a, b in bounds], world, k,
                            avg, use_rms, clip, rms, group, _CudaStages, tm, return_knn,
                            balance=bal, prev_ms=prev_ms)
        if bal is not None:
            bal.prev_bounds = bounds
        return out
