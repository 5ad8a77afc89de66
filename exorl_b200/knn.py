"""Host-side mirror of the reference's particle-based reward interface over the C ABI.

Same names, argument meaning and error behaviour as the reference:
  RMS  <- /root/reference/utils/utils.py:257-276
  PBE  <- /root/reference/utils/utils.py:279-319
  compute_apt_reward(source, target, args) <- the same computation with source != target
  proto_knn_reward / proto_enqueue <- agents/unsupervised_learning/proto.py:110-118
All arithmetic runs in hand-written sm_100a kernels (exorl_b200/csrc); there is no CPU path.
"""
import ctypes

import torch

from . import _lib
from . import ops  # noqa: F401  (registers torch.ops.exorl_b200.*)

_MODES = ("auto", "exact", "screened")


def _as_rows(x, name):
    """(b, ...) -> contiguous fp32 (b, c); the reference flattens trailing dims with .view
    (utils/utils.py:293-294)."""
    _lib.require_cuda(x, name)
    x = x.detach()
    if x.dtype != torch.float32:
        x = x.float()
    return x.reshape(x.size(0), -1).contiguous()


def _prepared_pair(source, target):
    """(PreparedRep, PreparedRep) if either side is one (the other side is prepared here with
    exb_knn_prep_f16), else (None, None)."""
    from .rep import PreparedRep, prepare_rep
    ps, pt = isinstance(source, PreparedRep), isinstance(target, PreparedRep)
    if not (ps or pt):
        return None, None
    if target is source:
        return source, source
    if ps and pt:
        a, b = source.affine, target.affine
        if (a is None) != (b is None) or (a is not None and a is not b and not torch.equal(a, b)):
            raise ValueError("exorl_b200: the two PreparedReps were prepared with different "
                             "affine records (rep.prepare_rep(..., affine=other.affine))")
        return source, target
    # the raw side is prepared with the other side's transform: scores are only comparable
    # between operands of one (mu, s)
    if ps:
        return source, prepare_rep(target, affine=source.affine)
    return prepare_rep(source, affine=target.affine), target


def knn_topk(source, target, k, idx_base=0, mode="auto"):
    """k smallest L2 distances (ascending) of every source row to the target rows.

    Returns (dist (b1, k) float32, idx (b1, k) int32), both on source's device.  Replaces the
    (b1, b2, c) broadcast + torch.norm + topk of utils/utils.py:293-300; raises RuntimeError
    "selected index k out of range" when k > b2 like the reference's topk.
    mode: "exact" = fp32 CUDA-core kernel; "screened" = tcgen05 fp16 screening + exact fp32
    recompute of the survivors (same results); "auto" picks by shape.
    Thin caller of torch.ops.exorl_b200.knn_topk (ops.py)."""
    if mode not in _MODES:
        raise ValueError(f"mode must be one of {_MODES}")
    same = target is source
    pre_s, pre_t = _prepared_pair(source, target)
    if pre_s is not None:
        return _knn_topk_prepared(pre_s, pre_t, int(k), idx_base, mode)
    src = _as_rows(source, "source")
    tgt = src if same else _as_rows(target, "target")
    _check_pair(src, tgt, int(k))
    return ops.call(torch.ops.exorl_b200.knn_topk, _knn_topk_impl, src, tgt, int(k),
                    int(idx_base), mode, same)


def _check_pair(src, tgt, k):
    if src.size(1) != tgt.size(1):
        raise RuntimeError(f"feature sizes differ: {src.size(1)} vs {tgt.size(1)}")
    if src.device != tgt.device:
        raise RuntimeError("source and target must be on the same device")
    if k > tgt.size(0):
        raise RuntimeError("selected index k out of range")
    if k < 1 or k > _lib.MAX_K:
        raise ValueError(f"exorl_b200: knn_k={k} outside [1, {_lib.MAX_K}]")


def _knn_topk_prepared(pre_s, pre_t, k, idx_base, mode):
    """knn_topk on operands a rep producer already prepared (rep.PreparedRep)."""
    lib = _lib.load()
    if mode == "exact":
        raise ValueError("exorl_b200: a PreparedRep is consumed by the screened path")
    src, tgt = _as_rows(pre_s.rep, "source"), _as_rows(pre_t.rep, "target")
    _check_pair(src, tgt, k)
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    dev = src.device
    with torch.cuda.device(dev):
        dist = torch.empty(nq, k, dtype=torch.float32, device=dev)
        idx = torch.empty(nq, k, dtype=torch.int32, device=dev)
        if nq == 0:
            return dist, idx
        need = lib.exb_knn_screened_workspace_bytes(nq, nb, dim, k)
        if need == 0:
            raise ValueError(f"exorl_b200: screened kNN does not support dim={dim}, k={k}")
        ws, _ = _lib.Workspace.get(dev, "knn_screened", need)
        _lib.check(lib.exb_knn_screened_pre_f32(
            _lib.ptr(src), nq, pre_s.operand, _lib.ptr(tgt), nb, pre_t.operand, dim, k,
            int(idx_base), _lib.ptr(dist), _lib.ptr(idx), _lib.ptr(ws), ws.numel(),
            _lib.stream_ptr(dev)))
        _repair_overflow_rows(src, tgt, k, idx_base, dist, idx, "knn_screened")
    return dist, idx


# Symmetric screen (csrc/knn_sym.cuh): a set against itself at dataset scale.
SYM_ENABLED = True
SYM_MIN_ROWS = 200_000        # measured on B200 (rep 512, k 12): 1.06x at 131k rows, 1.33x at
                              # 262k, 1.50x at 1M; below, the sample pre-pass and the hit logs
                              # cost what the halved MMA work saves
SYM_SAMPLE_MAX = 1 << 22      # (measurement scripts lower it to study the hit rate)
SYM_MEM_FRACTION = 0.6        # of the free device memory the hit logs may take; else full screen
sym_last = {}                 # diagnostics of the last symmetric call (sizes, fallback reason)


def _sym_sample_rows(n):
    """Rows of the sample pre-pass: 1/16 of the set.  The rate of hits per accumulator element is
    k / sample_rows whatever n is, and the symmetric kernel's epilogue pays per hit (measured at
    1M rows: 407 / 475 / 517 ms for 62 500 / 33 334 / 16 394 sample rows), while the pre-pass
    costs sample_rows / n of a full screen — 1/16 is the measured optimum."""
    return int(min(SYM_SAMPLE_MAX, max(4096, n // 16)))


def _sym_fits(dev, need):
    """The hit logs are ~18 GB per million rows (rep 512, k 12): take the symmetric path only
    when they fit beside what is already resident (a workspace of that size that already exists is
    reused)."""
    have = _lib.Workspace.size(dev, "knn_sym")
    if have >= need:
        return True
    free, _ = torch.cuda.mem_get_info(dev)
    return need - have <= SYM_MEM_FRACTION * free


def _knn_topk_sym(src, k, idx_base, feed=None):
    """kNN of `src` against itself through the symmetric screen, or None when the pass reports
    that it cannot certify the result (a transposed log overflowed: hub rows, heavy duplication)
    — the caller then runs the full screen.  `feed` (relabel._HostFeed): the rows of `src` are
    still on their way from the host; the sample pre-pass runs chunk by chunk behind the
    upload and only the symmetric pass itself waits for all of them."""
    lib = _lib.load()
    n, dim = src.shape
    dev = src.device
    stride = max(1, n // _sym_sample_rows(n))
    # exact distance of every row to its k-th nearest SAMPLE row: an upper bound of its true
    # k-th-neighbour distance, which fixes the row's keep threshold for the symmetric pass
    if feed is None:
        sample = src[::stride].contiguous()
        d_s, _ = _knn_topk_impl(src, sample, k, 0, "screened", False, allow_sym=False)
    else:
        sample = feed.sample(stride)
        d_s = torch.empty(n, k, dtype=torch.float32, device=dev)
        for a, b in feed.start(0, n):
            feed.wait(a, b)
            d_s[a:b] = _knn_topk_impl(src[a:b], sample, k, 0, "screened", False,
                                      allow_sym=False)[0]
        feed.wait_all()
    expected = 1.3 * k * n / sample.size(0)          # candidates per row the thresholds admit
    log_cap = int(-(-(3 * expected / 2 + 96) // 8) * 8)      # direct log per (row, column half)
    recv_cap = int(-(-(3 * expected + 96) // 8) * 8)         # received list per row
    with torch.cuda.device(dev):
        dist = torch.empty(n, k, dtype=torch.float32, device=dev)
        idx = torch.empty(n, k, dtype=torch.int32, device=dev)
        need = lib.exb_knn_sym_workspace_bytes(n, dim, k, log_cap, recv_cap)
        if need == 0 or not _sym_fits(dev, need):
            sym_last.update(n=n, workspace_bytes=need, no_room=True)
            return None
        ws, _ = _lib.Workspace.get(dev, "knn_sym", need)
        _lib.check(lib.exb_knn_sym_f32(
            _lib.ptr(src), n, dim, k, int(idx_base), _lib.ptr(d_s[:, k - 1:]), k, log_cap,
            recv_cap, _lib.ptr(dist), _lib.ptr(idx), _lib.ptr(ws), ws.numel(),
            _lib.stream_ptr(dev)))
    st = screened_status(dev, key="knn_sym")
    sym_last.update(n=n, sample_rows=sample.size(0), log_cap=log_cap, recv_cap=recv_cap,
                    undecided=st[0], status=st[1])
    if st[1] & 2:
        return None
    if st[1] & 1:
        rows = (idx[:, 0] < 0).nonzero().flatten()
        _recompute_rows(src, src, k, idx_base, dist, idx, rows)
    return dist, idx


def _knn_topk_impl(src, tgt, k, idx_base, mode, same, allow_sym=True):
    """CUDA implementation of torch.ops.exorl_b200.knn_topk: contiguous fp32 (b, c) rows in."""
    lib = _lib.load()
    if same:
        tgt = src
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    dev = src.device
    if (same and allow_sym and SYM_ENABLED and mode != "exact" and nq >= SYM_MIN_ROWS
            and OVERFLOW_POLICY == "sync" and _repair_depth == 0
            and dim <= _lib.SCREEN_MAX_DIM and k <= _lib.SCREEN_MAX_K and nq > 4 * k):
        out = _knn_topk_sym(src, k, idx_base)
        if out is not None:
            return out
        sym_last["fell_back_to_full_screen"] = True
    with torch.cuda.device(dev):
        dist = torch.empty(nq, k, dtype=torch.float32, device=dev)
        idx = torch.empty(nq, k, dtype=torch.int32, device=dev)
        if nq == 0:
            return dist, idx
        use_screen = mode == "screened" or (mode == "auto" and _screen_ok(nq, nb, dim, k))
        if use_screen:
            need = lib.exb_knn_screened_workspace_bytes(nq, nb, dim, k)
            if need == 0:
                raise ValueError(f"exorl_b200: screened kNN does not support dim={dim}, k={k}")
            ws, _ = _lib.Workspace.get(dev, "knn_screened", need)
            _lib.check(lib.exb_knn_screened_f32(
                _lib.ptr(src), nq, _lib.ptr(tgt), nb, dim, k, int(idx_base), int(same),
                _lib.ptr(dist), _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
            _repair_overflow_rows(src, tgt, k, idx_base, dist, idx, "knn_screened")
        else:
            need = lib.exb_knn_exact_workspace_bytes(nq, nb, dim, k)
            ws, _ = _lib.Workspace.get(dev, "knn_exact", need)
            _lib.check(lib.exb_knn_exact_f32(
                _lib.ptr(src), nq, _lib.ptr(tgt), nb, dim, k, int(idx_base), _lib.ptr(dist),
                _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return dist, idx


def _streaming(nq, nb, dim, k):
    """True when the screened call runs the dataset-scale (streaming) kernels; the per-update
    chain recomputes undecided rows inline and needs no host check."""
    if dim > _lib.SCREEN_MAX_DIM or k > _lib.SCREEN_MAX_K:
        return False
    return int(_lib.load().exb_knn_screen_lists_count(nq, nb, dim, k)) > 0


_repair_depth = 0        # nesting of the strided re-screen below (host-side, single thread)
STRIDED_BANKS = 8


def _topk_merge(dist_lists, idx_lists, k):
    """(G, n, k) per-part lists -> (n, k) global k smallest (exb_topk_merge)."""
    lib = _lib.load()
    g, n, _ = dist_lists.shape
    dev = dist_lists.device
    out_d = torch.empty(n, k, dtype=torch.float32, device=dev)
    out_i = torch.empty(n, k, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.exb_topk_merge(_lib.ptr(dist_lists.contiguous()),
                                      _lib.ptr(idx_lists.contiguous()), g, n, k, _lib.ptr(out_d),
                                      _lib.ptr(out_i), _lib.stream_ptr(dev)))
    return out_d, out_i


def _recompute_rows(src, tgt, k, idx_base, dist, idx, rows):
    """Exact result for the flagged `rows`.  Few rows: the exact fp32 kernel.  Many rows (dense
    data: more bank rows inside a row's screening margin than its candidate lists hold — e.g.
    trajectories whose consecutive states are near-duplicates): re-screen them against
    STRIDED_BANKS interleaved sub-banks bank[r::P].  Each sub-bank is P times less dense, so the
    lists fit again; the union of the per-sub-bank k nearest contains the global k nearest, and
    exb_topk_merge selects them — still exact, at about the cost of one more screened pass instead
    of a CUDA-core scan of the whole bank per row."""
    global _repair_depth
    nb = tgt.size(0)
    p = STRIDED_BANKS
    if rows.numel() <= 4096 or _repair_depth >= 2 or nb < p * max(256, k):
        for c in range(0, rows.numel(), 8192):
            sel = rows[c:c + 8192]
            d, i = knn_topk(src[sel], tgt, k, idx_base=idx_base, mode="exact")
            dist[sel] = d
            idx[sel] = i
        return
    q = src[rows].contiguous()
    dl, il = [], []
    _repair_depth += 1
    try:
        for r in range(p):
            d_r, i_r = knn_topk(q, tgt[r::p].contiguous(), k, mode="screened")
            dl.append(d_r)
            il.append(i_r * p + (r + int(idx_base)))
    finally:
        _repair_depth -= 1
    d, i = _topk_merge(torch.stack(dl), torch.stack(il), k)
    dist[rows] = d
    idx[rows] = i


OVERFLOW_POLICY = "sync"


def set_overflow_policy(policy):
    """What a dataset-scale screened call does about rows beyond the device-side fallback (more
    than 4096 undecided rows: heavily duplicated or very dense data, non-finite values).
      "sync" (default): read the call's status word — ONE stream synchronisation per
               dataset-scale call, after all of its kernels are enqueued — and recompute the
               flagged rows before returning, so the result is always the exact path's;
      "flag":  never synchronise (the call only enqueues work and can sit inside a CUDA graph or
               a multi-stream pipeline); such rows come back with idx[:, 0] == -1 and are the
               caller's to repair with `repair_flagged(...)` once it synchronises anyway.
    Per-update calls recompute undecided rows inline and never synchronise under either policy.
    Returns the previous policy."""
    global OVERFLOW_POLICY
    if policy not in ("sync", "flag"):
        raise ValueError("policy must be 'sync' or 'flag'")
    prev, OVERFLOW_POLICY = OVERFLOW_POLICY, policy
    return prev


def set_per_update_scan(block_cooperative):
    """How the per-update chain recomputes rows the latency screen could not certify: False
    (default) each row's own warp scans the bank ranges it lost; True the rows of a rerank block
    scan them together (exb_knn_set_per_update_scan: 273 vs 369 us when every row of a 1024-row
    batch is undecided, +1.3 us on ordinary batches).  Returns the previous setting."""
    return bool(_lib.load().exb_knn_set_per_update_scan(1 if block_cooperative else 0))


def repair_flagged(source, target, k, dist, idx, idx_base=0):
    """Recompute in place the rows a "flag"-policy call marked with idx[:, 0] == -1 (synchronises).
    Returns the number of rows repaired."""
    src = _as_rows(source, "source")
    tgt = src if target is source else _as_rows(target, "target")
    rows = (idx[:, 0] < 0).nonzero().flatten()
    if rows.numel():
        prev = set_overflow_policy("sync")     # the re-screens inside must repair their own rows
        try:
            _recompute_rows(src, tgt, int(k), idx_base, dist, idx, rows)
        finally:
            set_overflow_policy(prev)
    return int(rows.numel())


def _repair_overflow_rows(src, tgt, k, idx_base, dist, idx, ws_key):
    """Dataset-scale screened calls only: rows the screen could not certify are recomputed on
    the device by the exact kernel, up to 4096 of them; if there were more (heavily duplicated
    or very dense rows, values outside the fp16 range) the rerank marks the rest with
    idx[:, 0] = -1 and raises the status flag.  Policy "sync": read the flag (one stream sync per
    dataset-scale call) and recompute those rows (_recompute_rows), so the result is always the
    exact path's.  Policy "flag": leave them marked (set_overflow_policy)."""
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    if OVERFLOW_POLICY == "flag" or not _streaming(nq, nb, dim, k):
        return
    st = screened_status(src.device, key=ws_key)
    if st[1] == 0:
        return
    rows = (idx[:, 0] < 0).nonzero().flatten()
    _recompute_rows(src, tgt, k, idx_base, dist, idx, rows)


SCREEN_CAP = 24   # EXB_SCREEN_CAP: candidate slots per list


def screen_lists_count(nq, nb, dim, k):
    """Candidate lists per (query, shard) the screen-only form produces; 0 = not available for
    these sizes (per-update inputs use the fused chain instead)."""
    if dim > _lib.SCREEN_MAX_DIM or k > _lib.SCREEN_MAX_K:
        return 0
    return int(_lib.load().exb_knn_screen_lists_count(nq, nb, dim, k))


def knn_screen_lists(source, bank, k):
    """Screen every `source` row against `bank` (one shard) and stop: returns
    (val (nq, L, 24) f32, idx (nq, L, 24) i32 local to bank, cnt (nq, L) i32, ha (nq,), ea (nq,),
    scalars (4,) i32 view of {err_max, hn_max, range_flag, 0}).  exb_knn_screen_lists_f32."""
    lib = _lib.load()
    same = bank is source
    pre_s, pre_t = _prepared_pair(source, bank)
    if pre_s is not None:
        source, bank = pre_s.rep, pre_t.rep
    src = _as_rows(source, "source")
    tgt = src if same else _as_rows(bank, "bank")
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    k = int(k)
    if k > nb:
        raise RuntimeError("selected index k out of range")
    lists = screen_lists_count(nq, nb, dim, k)
    if lists == 0:
        raise ValueError(f"exorl_b200: no screen-only form for nq={nq}, nb={nb}, dim={dim}, k={k}")
    dev = src.device
    with torch.cuda.device(dev):
        val = torch.empty(nq, lists, SCREEN_CAP, dtype=torch.float32, device=dev)
        idx = torch.empty(nq, lists, SCREEN_CAP, dtype=torch.int32, device=dev)
        cnt = torch.empty(nq, lists, dtype=torch.int32, device=dev)
        ha = torch.empty(nq, dtype=torch.float32, device=dev)
        ea = torch.empty(nq, dtype=torch.float32, device=dev)
        sc = torch.empty(4, dtype=torch.int32, device=dev)
        need = lib.exb_knn_screened_workspace_bytes(nq, nb, dim, k)
        ws, _ = _lib.Workspace.get(dev, "knn_screened", need)
        if pre_s is not None:
            _lib.check(lib.exb_knn_screen_lists_pre_f32(
                _lib.ptr(src), nq, pre_s.operand, _lib.ptr(tgt), nb, pre_t.operand, dim, k,
                _lib.ptr(val), _lib.ptr(idx), _lib.ptr(cnt), _lib.ptr(ha), _lib.ptr(ea),
                _lib.ptr(sc), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
        else:
            _lib.check(lib.exb_knn_screen_lists_f32(
                _lib.ptr(src), nq, _lib.ptr(tgt), nb, dim, k, int(same), _lib.ptr(val),
                _lib.ptr(idx), _lib.ptr(cnt), _lib.ptr(ha), _lib.ptr(ea), _lib.ptr(sc),
                _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return val, idx, cnt, ha, ea, sc


def knn_rerank_lists(source, bank, k, val, idx, cnt, shard_lo, ha, ea, scalars):
    """Exact top-k of `source` rows (a query block) from the candidate lists of all shards:
    val/idx (G, nq, L, 24), cnt (G, nq, L), shard_lo = first bank row of every shard, `bank` the
    whole fp32 bank, scalars the MAX-combined (4,) i32 word.  exb_knn_rerank_lists_f32."""
    lib = _lib.load()
    src = _as_rows(source, "source")
    tgt = src if bank is source else _as_rows(bank, "bank")
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    g, lists = val.size(0), val.size(2)
    if val.size(1) != nq or cnt.shape != (g, nq, lists) or len(shard_lo) != g:
        raise ValueError("candidate lists do not match the query block / shard count")
    dev = src.device
    lo = (ctypes.c_int32 * g)(*[int(x) for x in shard_lo])
    with torch.cuda.device(dev):
        dist = torch.empty(nq, k, dtype=torch.float32, device=dev)
        out_i = torch.empty(nq, k, dtype=torch.int32, device=dev)
        need = lib.exb_knn_rerank_lists_workspace_bytes(nq, nb, int(k))
        ws, _ = _lib.Workspace.get(dev, "knn_rerank_lists", need)
        _lib.check(lib.exb_knn_rerank_lists_f32(
            _lib.ptr(src), nq, _lib.ptr(tgt), nb, dim, int(k), g, lists,
            _lib.ptr(val.contiguous()), _lib.ptr(idx.contiguous()), _lib.ptr(cnt.contiguous()),
            lo, _lib.ptr(ha.contiguous()), _lib.ptr(ea.contiguous()), _lib.ptr(scalars),
            _lib.ptr(dist), _lib.ptr(out_i), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
        if screened_status(dev, key="knn_rerank_lists")[1] != 0:   # see _repair_overflow_rows
            rows = (out_i[:, 0] < 0).nonzero().flatten()
            _recompute_rows(src, tgt, k, 0, dist, out_i, rows)
    return dist, out_i


def _screen_ok(nq, nb, dim, k):
    return (dim <= _lib.SCREEN_MAX_DIM and k <= _lib.SCREEN_MAX_K and nq >= 128 and nb >= 256)


def screened_status(device=None, key="knn_screened"):
    """(undecided_rows, invalid_flag, reranked_pairs, _) of the last screened call that used the
    workspace `key` ("knn_screened": knn_topk / relabel, "apt_reward": PBE / compute_apt_reward);
    synchronises."""
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    ws, fresh = _lib.Workspace.get(dev, key, 1024)
    if fresh:
        return (0, 0, 0, 0)
    out = (ctypes.c_int32 * 4)()
    _lib.check(lib.exb_knn_screened_status_sync(_lib.ptr(ws), out, _lib.stream_ptr(dev)))
    return tuple(out)


def _pbe_ws(dev):
    lib = _lib.load()
    need = lib.exb_pbe_workspace_bytes(0, 0)
    ws, fresh = _lib.Workspace.get(dev, "pbe", need, zero_init=True)
    return ws


class RMS(object):
    """running mean and std — drop-in for utils/utils.py:257-276.

    M and S are tensors on `device`; `n` reads as the reference's python float.  Its value is
    held in one fp64 word on the device (`_n`) next to M and S, and all three are updated IN
    PLACE by the kernels: an update has no host-side value that changes from call to call, so it
    can sit inside a `torch.compile`d graph or a CUDA graph.  Reading `.n` copies the word back
    (a sync; logging and pickling only — the hot path never reads it).  The object pickles as
    {M, S, n} (snapshots pickle the whole agent)."""

    def __init__(self, device, epsilon=1e-4, shape=(1,)):
        self.M = torch.zeros(shape).to(device)
        self.S = torch.ones(shape).to(device)
        self._n = torch.full((1,), epsilon, dtype=torch.float64).to(device)

    @property
    def n(self):
        return float(self._n.item())

    @n.setter
    def n(self, value):
        self._n.fill_(float(value))

    def __getstate__(self):
        return {"M": self.M, "S": self.S, "n": self.n}

    def __setstate__(self, state):
        self.M, self.S = state["M"], state["S"]
        self._n = torch.full((1,), state["n"], dtype=torch.float64, device=self.M.device)

    def __call__(self, x):
        bs = x.size(0)
        if (self.M.numel() == 1 and x.is_cuda and x.dtype == torch.float32
                and x.numel() == bs and self.M.is_cuda):
            ops.call(torch.ops.exorl_b200.rms_update, _rms_update_impl,
                     x.detach().reshape(-1).contiguous(), self.M, self.S, self._n)
        else:
            # general-shape states (not on the kNN path): same formula with torch ops on
            # whatever device the state lives on
            n = self.n
            delta = torch.mean(x, dim=0) - self.M
            new_m = self.M + delta * bs / (n + bs)
            new_s = (self.S * n + torch.var(x, dim=0) * bs
                     + torch.square(delta) * n * bs / (n + bs)) / (n + bs)
            self.M = new_m
            self.S = new_s
            self._n += bs
        return self.M, self.S


def _rms_update_impl(x, rms_m, rms_s, rms_n):
    lib = _lib.load()
    dev = x.device
    with torch.cuda.device(dev):
        ws = _pbe_ws(dev)
        _lib.check(lib.exb_rms_update(_lib.ptr(x), x.numel(), _lib.ptr(rms_m), _lib.ptr(rms_s),
                                      _lib.ptr(rms_n), _lib.ptr(ws), ws.numel(),
                                      _lib.stream_ptr(dev)))


def _check_rms(rms):
    if rms is None:
        raise ValueError("knn_rms=True needs an RMS state object")
    if rms.M.numel() != 1 or not rms.M.is_cuda:
        raise ValueError("the RMS state used by PBE must have shape (1,) on the GPU")


def _rms_ptrs(knn_rms, rms_m, rms_s, rms_n):
    if not knn_rms:
        return (None, None, None)
    return (_lib.ptr(rms_m), _lib.ptr(rms_s), _lib.ptr(rms_n))


def _reward_from_knn(dist, k, knn_avg, knn_rms, knn_clip, rms_m, rms_s, rms_n):
    """RMS update + PBE epilogue on (nq, k) distances; the state tensors are updated in place."""
    lib = _lib.load()
    dev = dist.device
    nq = dist.size(0)
    out = torch.empty(nq, 1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        ws = _pbe_ws(dev)
        _lib.check(lib.exb_pbe_reward(
            _lib.ptr(dist), nq, k, int(bool(knn_avg)), int(bool(knn_rms)), float(knn_clip),
            *_rms_ptrs(knn_rms, rms_m, rms_s, rms_n), _lib.ptr(out), _lib.ptr(ws), ws.numel(),
            _lib.stream_ptr(dev)))
    return out


_MODE_ID = {"auto": 0, "exact": 1, "screened": 2}


def compute_apt_reward(source, target, args, rms=None, mode="auto", return_knn=False):
    """APT particle-based reward of every `source` row against the `target` bank, (b1, 1).

    `args` exposes knn_k, knn_avg, knn_rms, knn_clip (and optionally .rms); this is the
    signature the computation had before the reference folded it into PBE.__call__
    (utils/utils.py:289-319, where source = target = rep).  Thin caller of
    torch.ops.exorl_b200.apt_reward: one C call (exb_apt_reward_f32), for per-update sizes a
    short chain prep -> tcgen05 screen -> rerank + RMS/epilogue."""
    if mode not in _MODES:
        raise ValueError(f"mode must be one of {_MODES}")
    if rms is None:
        rms = getattr(args, "rms", None)
    k = int(args.knn_k)
    knn_avg, knn_rms, knn_clip = bool(args.knn_avg), bool(args.knn_rms), float(args.knn_clip)
    same = target is source
    pre_s, pre_t = _prepared_pair(source, target)
    if pre_s is not None:
        if mode == "exact":
            raise ValueError("exorl_b200: a PreparedRep is consumed by the screened path")
        source, target = pre_s.rep, pre_t.rep
    src = _as_rows(source, "source")
    tgt = src if same else _as_rows(target, "target")
    _check_pair(src, tgt, k)
    if mode == "screened" and (src.size(1) > _lib.SCREEN_MAX_DIM or k > _lib.SCREEN_MAX_K):
        raise ValueError(f"exorl_b200: screened kNN does not support dim={src.size(1)}, k={k}")
    if knn_rms:
        _check_rms(rms)
        m, s_, n = rms.M, rms.S, rms._n
    else:
        m, s_, n = src.new_empty(0), src.new_empty(0), src.new_empty(0)   # unused placeholders
    if pre_s is not None:
        out, dist, idx = _apt_reward_impl(src, tgt, m, s_, n, k, knn_avg, knn_rms, knn_clip,
                                          mode, same, pre=(pre_s, pre_t))
    else:
        out, dist, idx = ops.call(torch.ops.exorl_b200.apt_reward, _apt_reward_impl, src, tgt,
                                  m, s_, n, k, knn_avg, knn_rms, knn_clip, mode, same)
    return (out, dist, idx) if return_knn else out


def _apt_reward_impl(src, tgt, rms_m, rms_s, rms_n, k, knn_avg, knn_rms, knn_clip, mode, same,
                     pre=None):
    """CUDA implementation of torch.ops.exorl_b200.apt_reward (and, with `pre`, of the same call
    on operands a rep producer prepared).  RMS state tensors are updated in place."""
    lib = _lib.load()
    if same:
        tgt = src
    pre_s, pre_t = pre if pre is not None else (None, None)
    nq, nb, dim = src.size(0), tgt.size(0), src.size(1)
    dev = src.device
    if nq > 0 and mode != "exact" and _streaming(nq, nb, dim, k):
        # dataset-scale input: kNN (with the host-side overflow check) then the epilogue; the
        # one-call fused chain below is for per-update sizes
        if pre_s is not None:
            d_, i_ = _knn_topk_prepared(pre_s, pre_t, k, 0, "screened")
        else:
            d_, i_ = _knn_topk_impl(src, tgt, k, 0, "screened", same)
        out = _reward_from_knn(d_, k, knn_avg, knn_rms, knn_clip, rms_m, rms_s, rms_n)
        return out, d_, i_
    with torch.cuda.device(dev):
        out = torch.empty(nq, 1, dtype=torch.float32, device=dev)
        dist = torch.empty(nq, k, dtype=torch.float32, device=dev)
        idx = torch.empty(nq, k, dtype=torch.int32, device=dev)
        if nq == 0:
            return out, dist, idx
        need = lib.exb_apt_reward_workspace_bytes(nq, nb, dim, k)
        ws, _ = _lib.Workspace.get(dev, "apt_reward", need, zero_init=True)
        st = _rms_ptrs(knn_rms, rms_m, rms_s, rms_n)
        if pre_s is not None:
            if dim > _lib.SCREEN_MAX_DIM or k > _lib.SCREEN_MAX_K:
                raise ValueError(f"exorl_b200: screened kNN does not support dim={dim}, k={k}")
            _lib.check(lib.exb_apt_reward_pre_f32(
                _lib.ptr(src), nq, pre_s.operand, _lib.ptr(tgt), nb, pre_t.operand, dim, k,
                int(knn_avg), int(knn_rms), knn_clip, *st, _lib.ptr(out), _lib.ptr(dist),
                _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
        else:
            _lib.check(lib.exb_apt_reward_f32(
                _lib.ptr(src), nq, _lib.ptr(tgt), nb, dim, k, int(same), _MODE_ID[mode],
                int(knn_avg), int(knn_rms), knn_clip, *st, _lib.ptr(out), _lib.ptr(dist),
                _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return out, dist, idx


class PBE(object):
    """particle-based entropy based on knn normalized by running mean — drop-in for
    utils/utils.py:279-319 (same constructor, same call, same RMS side effects)."""

    def __init__(self, rms, knn_clip, knn_k, knn_avg, knn_rms, device):
        self.rms = rms
        self.knn_rms = knn_rms
        self.knn_k = knn_k
        self.knn_avg = knn_avg
        self.knn_clip = knn_clip
        self.device = device
        self.mode = "auto"

    def __call__(self, rep):
        return compute_apt_reward(rep, rep, self, rms=self.rms,
                                  mode=getattr(self, "mode", "auto"))


def rnd_intr_reward(prediction, target, rms, rnd_scale=1.0):
    """RND reward from the predictor / target network outputs — replaces the arithmetic of
    RND.forward's last line (agents/unsupervised_learning/rnd.py:52-54) and of
    RNDAgent.compute_intr_reward (:94-99): prediction_error = mean((target - prediction)^2, -1),
    `rms(prediction_error)` (state M, S, n updated like utils/utils.py:272-274),
    reward = rnd_scale * prediction_error / (sqrt(var) + 1e-8).
    Returns (reward (n, 1), prediction_error (n, 1)).  torch.ops.exorl_b200.rnd_reward."""
    pred = _as_rows(prediction, "prediction")
    tgt = _as_rows(target, "target")
    if pred.shape != tgt.shape:
        raise RuntimeError(f"The size of tensor a {tuple(tgt.shape)} must match the size of "
                           f"tensor b {tuple(pred.shape)}")
    if rms.M.numel() != 1 or not rms.M.is_cuda:
        raise ValueError("the RMS state used by RND must have shape (1,) on the GPU")
    return ops.call(torch.ops.exorl_b200.rnd_reward, _rnd_reward_impl, pred, tgt, rms.M, rms.S,
                    rms._n, float(rnd_scale))


def _rnd_reward_impl(pred, tgt, rms_m, rms_s, rms_n, rnd_scale):
    lib = _lib.load()
    n, dim = pred.shape
    dev = pred.device
    with torch.cuda.device(dev):
        err = torch.empty(n, 1, dtype=torch.float32, device=dev)
        reward = torch.empty(n, 1, dtype=torch.float32, device=dev)
        if n == 0:
            return reward, err
        ws = _pbe_ws(dev)
        _lib.check(lib.exb_rnd_reward_f32(
            _lib.ptr(pred), _lib.ptr(tgt), n, dim, float(rnd_scale), _lib.ptr(rms_m),
            _lib.ptr(rms_s), _lib.ptr(rms_n), _lib.ptr(err), _lib.ptr(reward), _lib.ptr(ws),
            ws.numel(), _lib.stream_ptr(dev)))
    return reward, err


def proto_enqueue(queue, queue_ptr, z, candidates):
    """FIFO write of z[candidates] into the queue; returns the new pointer (proto.py:110-112)."""
    n = candidates.numel()
    queue[queue_ptr:queue_ptr + n] = z[candidates]
    return (queue_ptr + n) % queue.shape[0]


_proto_calls = 0


def proto_sample_enqueue(z, protos, queue, queue_ptr, noise=None, seed=0, call=None):
    """Candidate sampling + queue write of ProtoAgent.compute_intr_reward (proto.py:105-112) in
    one kernel: `scores = protos(z).T; prob = softmax(scores, 1); candidates =
    Categorical(prob).sample(); queue[ptr:ptr+P] = z[candidates]`.  `protos` is the prototype
    nn.Linear (bias-free) or its (P, dim) weight.  torch draws the sample as the exponential race
    argmax(probs / q), q ~ Exp(1); `noise` (P, B) supplies q (parity with a recorded reference
    stream), otherwise q comes from the device Philox stream keyed by (seed, call) — `call`
    defaults to a process-wide counter so that successive calls draw fresh noise.
    Returns (new queue_ptr, candidates (P,) int64).  torch.ops.exorl_b200.proto_sample."""
    global _proto_calls
    w = protos.weight if hasattr(protos, "weight") else protos
    zr = _as_rows(z, "z")
    wr = _as_rows(w, "protos")
    if wr.size(1) != zr.size(1):
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({zr.size(0)}x{zr.size(1)} "
                           f"and {wr.size(1)}x{wr.size(0)})")
    _lib.require_cuda(queue, "queue")
    if (queue.dtype != torch.float32 or not queue.is_contiguous() or queue.dim() != 2
            or queue.size(1) != zr.size(1)):
        raise ValueError("queue must be a contiguous float32 (queue_size, dim) tensor")
    n_protos = wr.size(0)
    ptr = int(queue_ptr)
    if ptr + n_protos > queue.size(0):
        # the reference's slice assignment fails the same way when the slice is clipped
        raise RuntimeError(f"The expanded size of the tensor ({queue.size(0) - ptr}) must match "
                           f"the existing size ({n_protos}) at non-singleton dimension 0")
    if noise is None:
        noise_t = zr.new_empty(0)
        if call is None:
            call = _proto_calls
            _proto_calls += 1
    else:
        noise_t = _as_rows(noise, "noise")
        if noise_t.shape != (n_protos, zr.size(0)):
            raise ValueError(f"noise must have shape ({n_protos}, {zr.size(0)})")
        call = 0 if call is None else call
    cand = ops.call(torch.ops.exorl_b200.proto_sample, _proto_sample_impl, zr, wr, queue, ptr,
                    noise_t, int(seed), int(call))
    return (ptr + n_protos) % queue.size(0), cand


def _proto_sample_impl(z, protos, queue, queue_ptr, noise, seed, call):
    lib = _lib.load()
    dev = z.device
    n_protos, dim = protos.shape
    cand = torch.empty(n_protos, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.exb_proto_sample_f32(
            _lib.ptr(z), z.size(0), _lib.ptr(protos), n_protos, dim,
            _lib.ptr(noise) if noise.numel() else None, ctypes.c_uint64(seed & (2 ** 64 - 1)),
            ctypes.c_uint64(call & (2 ** 64 - 1)), _lib.ptr(queue), queue.size(0),
            int(queue_ptr), _lib.ptr(cand), None, _lib.stream_ptr(dev)))
    return cand


def proto_knn_reward(z, queue, topk, mode="auto"):
    """Distance of every z row to its topk-th nearest queue row, (B, 1) — replaces
    proto.py:115-118 (no RMS, no clip, no log).  torch.ops.exorl_b200.proto_reward."""
    if mode not in _MODES:
        raise ValueError(f"mode must be one of {_MODES}")
    pre_s, pre_t = _prepared_pair(z, queue)
    if pre_s is not None:
        dist, _ = _knn_topk_prepared(pre_s, pre_t, int(topk), 0, mode)
        return _take_kth(dist, int(topk))
    src = _as_rows(z, "z")
    tgt = _as_rows(queue, "queue")
    _check_pair(src, tgt, int(topk))
    return ops.call(torch.ops.exorl_b200.proto_reward, _proto_reward_impl, src, tgt, int(topk),
                    mode)


def _take_kth(dist, k):
    lib = _lib.load()
    out = torch.empty(dist.size(0), 1, dtype=torch.float32, device=dist.device)
    if dist.size(0) == 0:
        return out
    with torch.cuda.device(dist.device):
        _lib.check(lib.exb_take_kth(_lib.ptr(dist), dist.size(0), k, _lib.ptr(out),
                                    _lib.stream_ptr(dist.device)))
    return out


def _proto_reward_impl(z, queue, topk, mode):
    """CUDA implementation of torch.ops.exorl_b200.proto_reward (exb_proto_reward_f32)."""
    lib = _lib.load()
    nq, nb, dim = z.size(0), queue.size(0), z.size(1)
    if nq == 0 or (mode != "exact" and _streaming(nq, nb, dim, topk)):
        # dataset-scale input: kNN with the host-side overflow check, then the k-th column
        dist, _ = _knn_topk_impl(z, queue, topk, 0, mode, False)
        return _take_kth(dist, topk)
    dev = z.device
    with torch.cuda.device(dev):
        out = torch.empty(nq, 1, dtype=torch.float32, device=dev)
        dist = torch.empty(nq, topk, dtype=torch.float32, device=dev)
        idx = torch.empty(nq, topk, dtype=torch.int32, device=dev)
        need = lib.exb_apt_reward_workspace_bytes(nq, nb, dim, topk)
        ws, _ = _lib.Workspace.get(dev, "apt_reward", need, zero_init=True)
        _lib.check(lib.exb_proto_reward_f32(
            _lib.ptr(z), nq, _lib.ptr(queue), nb, dim, topk, _MODE_ID[mode], _lib.ptr(out),
            _lib.ptr(dist), _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return out
