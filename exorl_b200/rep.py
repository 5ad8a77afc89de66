"""Rep producers: the step in front of the kNN reward.

  icm_get_rep(icm, obs)     <- ICM.get_rep / ICM.trunk = Linear -> LayerNorm -> Tanh,
                               agents/unsupervised_learning/icm_apt.py:17-18,50-52
  normalize_rep(z)          <- F.normalize(z, dim=1, p=2), agents/unsupervised_learning/proto.py:104

Each returns a PreparedRep: the fp32 rep the reference would hand to PBE.__call__ / the Proto kNN
lines, plus the prepared fp16 operand + row norms the screening kernel consumes, written by ONE
kernel (csrc/rep_producer.cu) — the kNN then skips its own prep pass over the rep.
PBE.__call__, compute_apt_reward, knn_topk and proto_knn_reward accept a PreparedRep wherever they
accept a rep tensor.  No CPU path.
"""
import torch

from . import _lib


def _dim_pad(dim):
    return (dim + 63) // 64 * 64


class PreparedRep:
    """`.rep` (n, dim) fp32 + the prepared operand (`f16` (n, dim_pad), `half_norm`, `round_err`
    (n,), `scalars` 16 bytes).  Behaves like the rep for shape queries."""

    _zero_pool = {}   # device -> [int32 tensor of pre-zeroed 16-byte slots, next slot]

    @classmethod
    def _zeroed_scalars(cls, dev):
        """A fresh all-zero 16-byte scalars word without a fill launch per call (the per-update
        chain is launch-bound): slots of a zeroed pool, each handed out once."""
        key = dev.index if dev.index is not None else torch.cuda.current_device()
        ent = cls._zero_pool.get(key)
        if ent is None or ent[1] >= ent[0].size(0):
            ent = [torch.zeros(4096, 4, dtype=torch.int32, device=dev), 0]
            cls._zero_pool[key] = ent
        ent[1] += 1
        return ent[0][ent[1] - 1]

    def __init__(self, n, dim, device, rep=None):
        dev = torch.device(device)
        self.rep = rep if rep is not None else torch.empty(n, dim, dtype=torch.float32,
                                                           device=dev)
        # one allocation; the fp16 block comes first (TMA needs its rows 128-byte aligned)
        dpad = _dim_pad(dim)
        buf = torch.empty(n * (2 * dpad + 8), dtype=torch.uint8, device=dev)
        self.f16 = buf[:n * 2 * dpad].view(torch.float16).view(n, dpad)
        self.half_norm = buf[n * 2 * dpad:n * (2 * dpad + 4)].view(torch.float32)
        self.round_err = buf[n * (2 * dpad + 4):].view(torch.float32)
        self.scalars = self._zeroed_scalars(dev)
        self.affine = None    # affine record the operand was prepared with (prepare_rep)
        self._c = _lib.Operand(self.f16.data_ptr(), self.half_norm.data_ptr(),
                               self.round_err.data_ptr(), self.scalars.data_ptr())

    def rows(self, lo, hi):
        """Row-slice view [lo, hi) (a bank shard): shares storage and the scalars word of the
        whole set — maxima over a superset keep the screening certificate rigorous."""
        v = object.__new__(PreparedRep)
        v.rep, v.f16 = self.rep[lo:hi], self.f16[lo:hi]
        v.half_norm, v.round_err = self.half_norm[lo:hi], self.round_err[lo:hi]
        v.scalars = self.scalars
        v.affine = self.affine
        v._c = _lib.Operand(v.f16.data_ptr(), v.half_norm.data_ptr(), v.round_err.data_ptr(),
                            v.scalars.data_ptr())
        return v

    @property
    def operand(self):
        return self._c

    @property
    def device(self):
        return self.rep.device

    @property
    def shape(self):
        return self.rep.shape

    def size(self, *a):
        return self.rep.size(*a)


def _f32c(t, name):
    _lib.require_cuda(t, name)
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _trunk_params(trunk):
    """(weight, bias, ln_weight, ln_bias, eps) of nn.Sequential(Linear, LayerNorm, Tanh)
    (icm_apt.py:17-18) or of a module exposing it as `.trunk`."""
    if hasattr(trunk, "trunk"):
        trunk = trunk.trunk
    if isinstance(trunk, (tuple, list)):
        w, b, g, be, *rest = trunk
        return w, b, g, be, (float(rest[0]) if rest else 1e-5)
    lin, ln, act = trunk[0], trunk[1], trunk[2]
    if not (isinstance(lin, torch.nn.Linear) and isinstance(ln, torch.nn.LayerNorm)
            and isinstance(act, torch.nn.Tanh)) or lin.bias is None or ln.weight is None:
        raise ValueError("exorl_b200: expected nn.Sequential(Linear, LayerNorm, Tanh)")
    return lin.weight, lin.bias, ln.weight, ln.bias, float(ln.eps)


def ln_tanh_rep(pre, ln_weight, ln_bias, eps=1e-5):
    """tanh(LayerNorm(pre)) for given pre-activations (n, dim) -> PreparedRep."""
    lib = _lib.load()
    pre = _f32c(pre, "pre")
    n, dim = pre.shape
    dev = pre.device
    g, be = _f32c(ln_weight, "ln_weight"), _f32c(ln_bias, "ln_bias")
    with torch.cuda.device(dev):
        out = PreparedRep(n, dim, dev)
        if n:
            _lib.check(lib.exb_rep_ln_tanh_f32(_lib.ptr(pre), n, dim, _lib.ptr(g), _lib.ptr(be),
                                               float(eps), _lib.ptr(out.rep), out.operand,
                                               _lib.stream_ptr(dev)))
    return out


def icm_get_rep(icm, obs):
    """ICM.get_rep(obs, action) = trunk(obs) (icm_apt.py:50-52) -> PreparedRep.

    `icm`: the ICM module, its trunk, or (weight, bias, ln_weight, ln_bias[, eps]).  Narrow inputs
    (state observations) run the Linear inside the kernel with W^T in shared memory; wide inputs
    (the 39200-wide pixel encoder output) run it as a 3xTF32 tcgen05 GEMM (csrc/rep_gemm.cuh)
    followed by one bias + LayerNorm + Tanh + operand pass."""
    lib = _lib.load()
    w, b, g, be, eps = _trunk_params(icm)
    obs = _f32c(obs, "obs")
    w, b, g, be = _f32c(w, "weight"), _f32c(b, "bias"), _f32c(g, "ln_weight"), _f32c(be, "ln_bias")
    obs = obs.reshape(obs.size(0), w.size(1) if obs.numel() == 0 else -1)
    n, in_dim = obs.shape
    dim = w.size(0)
    if w.size(1) != in_dim:
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({n}x{in_dim} and "
                           f"{w.size(1)}x{dim})")
    dev = obs.device
    if lib.exb_rep_trunk_smem_bytes(in_dim, dim) > 232448:
        need = lib.exb_rep_trunk_wide_workspace_bytes(n, in_dim, dim) if n else 0
        if n and need == 0:
            # in_dim % 4 != 0 (no TMA row pitch) or rep_dim outside the producers' range: the
            # Linear as a library GEMM, the rest fused
            return ln_tanh_rep(torch.addmm(b, obs, w.t()), g, be, eps)
        with torch.cuda.device(dev):
            out = PreparedRep(n, dim, dev)
            if n:
                ws, _ = _lib.Workspace.get(dev, "trunk_wide", need)
                _lib.check(lib.exb_rep_trunk_wide_f32(
                    _lib.ptr(obs), n, in_dim, _lib.ptr(w), _lib.ptr(b), _lib.ptr(g), _lib.ptr(be),
                    eps, dim, _lib.ptr(out.rep), out.operand, _lib.ptr(ws), ws.numel(),
                    _lib.stream_ptr(dev)))
        return out
    with torch.cuda.device(dev):
        out = PreparedRep(n, dim, dev)
        if n:
            _lib.check(lib.exb_rep_trunk_f32(_lib.ptr(obs), n, in_dim, _lib.ptr(w), _lib.ptr(b),
                                             _lib.ptr(g), _lib.ptr(be), eps, dim,
                                             _lib.ptr(out.rep), out.operand,
                                             _lib.stream_ptr(dev)))
    return out


def normalize_rep(x, eps=1e-12):
    """F.normalize(x, dim=1, p=2, eps) (proto.py:104) -> PreparedRep."""
    lib = _lib.load()
    x = _f32c(x, "x")
    n, dim = x.shape
    dev = x.device
    with torch.cuda.device(dev):
        out = PreparedRep(n, dim, dev)
        if n:
            _lib.check(lib.exb_rep_normalize_f32(_lib.ptr(x), n, dim, float(eps),
                                                 _lib.ptr(out.rep), out.operand,
                                                 _lib.stream_ptr(dev)))
    return out


def column_stats(rows):
    """(3, dim) fp32 per-column {sum, min, max} of `rows` (exb_knn_colstats_f32; deterministic).
    Shards combine with SUM / MIN / MAX."""
    lib = _lib.load()
    rows = _f32c(rows, "rows")
    rows = rows.reshape(rows.size(0), -1)
    n, dim = rows.shape
    dev = rows.device
    with torch.cuda.device(dev):
        stats = torch.empty(3, dim, dtype=torch.float32, device=dev)
        need = lib.exb_knn_colstats_workspace_bytes(n, dim)
        if need == 0:
            raise ValueError(f"exorl_b200: column statistics need 1 <= dim <= "
                             f"{_lib.SCREEN_MAX_DIM} and n >= 1 (got n={n}, dim={dim})")
        ws, _ = _lib.Workspace.get(dev, "colstats", need)
        _lib.check(lib.exb_knn_colstats_f32(_lib.ptr(rows), n, dim, _lib.ptr(stats),
                                            _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)))
    return stats


def column_affine(rows, distributed=False, group=None, n_total=None):
    """Affine record (mu, s) of the range-proof screen for a row set (include/exorl_b200.h,
    "Range-proof operands"): the screen then runs on s * (x - mu).  With `group` the statistics
    (a process group; None = this process's rows only) the statistics of every rank's rows are
    combined (three small all-reduces), so all ranks derive the SAME
    record for the union of their rows; `n_total` = rows in the union (default: summed)."""
    import torch.distributed as dist
    lib = _lib.load()
    stats = column_stats(rows)
    n = float(rows.size(0))
    dev = stats.device
    if distributed and dist.get_world_size(group) > 1:
        dist.all_reduce(stats[0], op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(stats[1], op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(stats[2], op=dist.ReduceOp.MAX, group=group)
        if n_total is None:
            cnt = torch.tensor([n], dtype=torch.float64, device=dev)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
            n_total = float(cnt.item())
    n = float(n_total) if n_total is not None else n
    with torch.cuda.device(dev):
        aff = torch.empty(_lib.AFFINE_FLOATS, dtype=torch.float32, device=dev)
        _lib.check(lib.exb_knn_affine_from_stats(_lib.ptr(stats), n, stats.size(1), _lib.ptr(aff),
                                                 _lib.stream_ptr(dev)))
    return aff


def prepare_rep(rep, affine=None, out=None):
    """Prepared operand of an existing fp32 rep (exb_knn_prep_f16), e.g. the Proto queue.
    `affine`: None = the rows as they are; "auto" = range-proof operand from the rep's own column
    statistics (column_affine); a record tensor = that transform (both operands of a kNN call
    must share one — knn_topk prepares the other side with the PreparedRep's record)."""
    lib = _lib.load()
    rep = _f32c(rep, "rep")
    rep = rep.reshape(rep.size(0), -1)
    n, dim = rep.shape
    dev = rep.device
    if isinstance(affine, str):
        if affine != "auto":
            raise ValueError("affine must be None, 'auto' or an affine record tensor")
        affine = column_affine(rep) if (n and dim <= _lib.SCREEN_MAX_DIM) else None
    with torch.cuda.device(dev):
        if out is None:
            out = PreparedRep(n, dim, dev, rep=rep)
        out.affine = affine
        if n and affine is not None:
            _lib.check(lib.exb_knn_prep_affine_f16(
                _lib.ptr(rep), n, dim, _lib.ptr(affine), _lib.ptr(out.f16),
                _lib.ptr(out.half_norm), _lib.ptr(out.round_err), _lib.ptr(out.scalars),
                _lib.stream_ptr(dev)))
        elif n:
            _lib.check(lib.exb_knn_prep_f16(_lib.ptr(rep), n, dim, _lib.ptr(out.f16),
                                            _lib.ptr(out.half_norm), _lib.ptr(out.round_err),
                                            _lib.ptr(out.scalars), _lib.stream_ptr(dev)))
    return out
