"""ctypes binding of libexorl_b200.so (include/exorl_b200.h).  No CPU fallback: if the library is
missing and cannot be built, every op raises."""
import ctypes
import os
import pathlib
import threading

import torch

from . import _build

_c = ctypes
_i64, _i32, _f32, _f64, _vp, _sz, _u64 = (_c.c_int64, _c.c_int32, _c.c_float, _c.c_double,
                                          _c.c_void_p, _c.c_size_t, _c.c_uint64)


class GatherCol(_c.Structure):
    _fields_ = [("src", _vp), ("dst", _vp), ("row_bytes", _i64), ("row_offset", _i32)]


class NpzArray(_c.Structure):
    _fields_ = [("name", _c.c_char * 64), ("dtype", _c.c_char * 8), ("ndim", _i32),
                ("fortran_order", _i32), ("shape", _i64 * 8), ("nbytes", _i64)]


class Operand(_c.Structure):
    """exb_operand: a prepared fp16 kNN operand (device pointers)."""
    _fields_ = [("f16", _vp), ("half_norm", _vp), ("round_err", _vp), ("scalars", _vp)]


_op_p = _c.POINTER(Operand)

# name -> (restype, argtypes); mirrors include/exorl_b200.h one to one
class SymHits(_c.Structure):
    """exb_sym_hits: the transposed hits a symmetric screen call left in its workspace."""
    _fields_ = [("target", _c.c_void_p), ("val", _c.c_void_p), ("src", _c.c_void_p),
                ("count", _c.c_void_p), ("capacity", _c.c_longlong)]


SYM_MAX_RANKS = 8      # EXB_SYM_MAX_RANKS

SIGNATURES = {
    "exb_abi_version": (_c.c_int, []),
    "exb_last_error": (_c.c_char_p, []),
    "exb_device_info": (_c.c_int, [_c.POINTER(_c.c_int), _c.POINTER(_c.c_int)]),
    "exb_knn_exact_workspace_bytes": (_sz, [_i64, _i64, _c.c_int, _c.c_int]),
    "exb_knn_exact_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _i32, _vp, _vp,
                                     _vp, _sz, _vp]),
    "exb_knn_prep_f16": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _vp, _vp, _vp]),
    "exb_knn_screened_workspace_bytes": (_sz, [_i64, _i64, _c.c_int, _c.c_int]),
    "exb_knn_screened_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _i32,
                                        _c.c_int, _vp, _vp, _vp, _sz, _vp]),
    "exb_knn_screen_lists_count": (_c.c_int, [_i64, _i64, _c.c_int, _c.c_int]),
    "exb_knn_screen_lists_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _c.c_int,
                                            _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "exb_knn_screen_lists_pre_f32": (_c.c_int, [_vp, _i64, _op_p, _vp, _i64, _op_p, _c.c_int,
                                                _c.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz,
                                                _vp]),
    "exb_knn_rerank_lists_workspace_bytes": (_sz, [_i64, _i64, _c.c_int]),
    "exb_knn_rerank_lists_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _c.c_int,
                                            _c.c_int, _vp, _vp, _vp, _c.POINTER(_i32), _vp, _vp,
                                            _vp, _vp, _vp, _vp, _sz, _vp]),
    "exb_knn_screened_status_sync": (_c.c_int, [_vp, _c.POINTER(_i32), _vp]),
    "exb_debug_trace_sync": (_c.c_int, [_vp, _c.POINTER(_i64), _vp]),
    "exb_knn_set_per_update_scan": (_c.c_int, [_c.c_int]),
    "exb_debug_knob_set": (_c.c_int, [_c.c_char_p, _c.c_int]),
    "exb_profile_begin": (_c.c_int, [_c.c_int]),
    "exb_profile_collect_sync": (_c.c_int, [_c.POINTER(_f32), _c.c_int, _c.POINTER(_c.c_int)]),
    "exb_topk_merge": (_c.c_int, [_vp, _vp, _c.c_int, _i64, _c.c_int, _vp, _vp, _vp]),
    "exb_pbe_workspace_bytes": (_sz, [_i64, _c.c_int]),
    "exb_pbe_workspace_init": (_c.c_int, [_vp, _sz, _vp]),
    "exb_pbe_reward": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _c.c_int, _f32, _vp, _vp, _vp,
                                  _vp, _vp, _sz, _vp]),
    "exb_pbe_reward_noalloc": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _c.c_int, _f32, _vp,
                                          _vp, _vp, _vp, _vp]),
    "exb_pbe_moments": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _vp, _vp, _sz, _vp]),
    "exb_pbe_finish": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _c.c_int, _f32, _vp, _vp, _vp,
                                  _vp, _vp, _vp]),
    "exb_apt_reward_workspace_bytes": (_sz, [_i64, _i64, _c.c_int, _c.c_int]),
    "exb_apt_reward_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                      _c.c_int, _c.c_int, _f32, _vp, _vp, _vp, _vp, _vp, _vp,
                                      _vp, _sz, _vp]),
    "exb_rep_trunk_smem_bytes": (_sz, [_c.c_int, _c.c_int]),
    "exb_rep_trunk_f32": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _vp, _vp, _f32, _c.c_int,
                                     _vp, _op_p, _vp]),
    "exb_rep_ln_tanh_f32": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _f32, _vp, _op_p, _vp]),
    "exb_rep_trunk_wide_workspace_bytes": (_sz, [_i64, _c.c_int, _c.c_int]),
    "exb_rep_trunk_wide_f32": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _vp, _vp, _f32, _c.c_int,
                                          _vp, _op_p, _vp, _sz, _vp]),
    "exb_rep_normalize_f32": (_c.c_int, [_vp, _i64, _c.c_int, _f32, _vp, _op_p, _vp]),
    "exb_knn_screened_pre_f32": (_c.c_int, [_vp, _i64, _op_p, _vp, _i64, _op_p, _c.c_int,
                                            _c.c_int, _i32, _vp, _vp, _vp, _sz, _vp]),
    "exb_apt_reward_pre_f32": (_c.c_int, [_vp, _i64, _op_p, _vp, _i64, _op_p, _c.c_int, _c.c_int,
                                          _c.c_int, _c.c_int, _f32, _vp, _vp, _vp, _vp, _vp,
                                          _vp, _vp, _sz, _vp]),
    "exb_rms_update": (_c.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "exb_rnd_reward_f32": (_c.c_int, [_vp, _vp, _i64, _c.c_int, _f32, _vp, _vp, _vp, _vp, _vp,
                                      _vp, _sz, _vp]),
    "exb_take_kth": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp]),
    "exb_proto_reward_f32": (_c.c_int, [_vp, _i64, _vp, _i64, _c.c_int, _c.c_int, _c.c_int, _vp, _vp,
                                        _vp, _vp, _sz, _vp]),
    "exb_proto_sample_f32": (_c.c_int, [_vp, _i64, _vp, _c.c_int, _c.c_int, _vp, _u64, _u64, _vp,
                                        _i64, _i64, _vp, _vp, _vp]),
    "exb_replay_gather": (_c.c_int, [_vp, _i64, _c.POINTER(GatherCol), _c.c_int, _vp, _vp,
                                     _c.c_int, _c.c_int, _f32, _vp, _vp, _vp]),
    "exb_replay_draw": (_c.c_int, [_u64, _u64, _vp, _vp, _i64, _c.c_int, _i64, _vp, _vp, _vp]),
    "exb_replay_sample": (_c.c_int, [_u64, _u64, _vp, _vp, _i64, _i64, _c.POINTER(GatherCol),
                                     _c.c_int, _vp, _vp, _c.c_int, _c.c_int, _f32, _vp, _vp, _vp,
                                     _vp]),
    "exb_knn_sym_workspace_bytes": (_sz, [_i64, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "exb_knn_sym_f32": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _i32, _vp, _i64, _c.c_int,
                                   _c.c_int, _vp, _vp, _vp, _sz, _vp]),
    "exb_knn_sym_screen_f32": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _vp, _i64, _c.c_int,
                                          _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                          _c.c_int, _vp, _sz, _c.POINTER(SymHits), _vp]),
    "exb_knn_sym_scatter": (_c.c_int, [_i64, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _vp, _vp,
                                       _vp, _i64, _vp, _vp]),
    "exb_knn_sym_screen_peer_f32": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _vp, _i64, _c.c_int,
                                               _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                               _c.c_int, _c.c_int, _c.POINTER(SymHits),
                                               _c.POINTER(_c.c_int32), _c.c_int, _vp, _sz, _vp]),
    "exb_knn_sym_operand_slots": (_c.c_int, [_i64, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _vp,
                                             _sz, _op_p, _c.POINTER(_vp)]),
    "exb_knn_sym_scatter_hits": (_c.c_int, [_i64, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                            _c.POINTER(SymHits), _vp, _vp]),
    "exb_peer_buffer_alloc": (_c.c_int, [_sz, _c.POINTER(_vp), _c.c_char_p]),
    "exb_peer_buffer_open": (_c.c_int, [_c.c_char_p, _c.POINTER(_vp)]),
    "exb_peer_buffer_close": (_c.c_int, [_vp]),
    "exb_peer_buffer_free": (_c.c_int, [_vp]),
    "exb_peer_buffer_clear": (_c.c_int, [_vp, _sz, _vp]),
    "exb_h2d_rows_2d": (_c.c_int, [_vp, _vp, _i64, _i64, _i64, _vp]),
    "exb_knn_sym_rerank_f32": (_c.c_int, [_vp, _i64, _c.c_int, _c.c_int, _i32, _c.c_int, _c.c_int,
                                          _i64, _i64, _vp, _vp, _vp, _vp]),
    "exb_knn_colstats_workspace_bytes": (_sz, [_i64, _c.c_int]),
    "exb_knn_colstats_f32": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _sz, _vp]),
    "exb_knn_affine_from_stats": (_c.c_int, [_vp, _c.c_double, _c.c_int, _vp, _vp]),
    "exb_knn_prep_affine_f16": (_c.c_int, [_vp, _i64, _c.c_int, _vp, _vp, _vp, _vp, _vp, _vp]),
    "exb_npz_open": (_c.c_int, [_c.c_char_p, _c.POINTER(_vp), _c.POINTER(_c.c_int)]),
    "exb_npz_info": (_c.c_int, [_vp, _c.c_int, _c.POINTER(NpzArray)]),
    "exb_npz_read": (_c.c_int, [_vp, _c.c_int, _vp, _i64]),
    "exb_npz_close": (None, [_vp]),
    "exb_npz_set_inflate_mode": (_c.c_int, [_c.c_int]),
}

EXB_OK, EXB_EINVAL, EXB_EKRANGE, EXB_EUNSUPPORTED, EXB_ECUDA, EXB_EFALLBACK = range(6)
MAX_K = 32
SCREEN_MAX_K = 16
SCREEN_MAX_DIM = 512
AFFINE_FLOATS = SCREEN_MAX_DIM + 8

_lock = threading.Lock()
_lib = None


def load():
    """Return the loaded library, building it in-tree first if the .so is absent."""
    global _lib
    with _lock:
        if _lib is None:
            path = _build.LIB
            if os.environ.get("EXORL_B200_LIB"):        # A/B measurements of two builds
                path = pathlib.Path(os.environ["EXORL_B200_LIB"])
            if not path.exists():
                try:
                    _build.build_library()
                except Exception as e:  # no nvcc on this machine
                    raise RuntimeError(
                        f"exorl_b200: {path} is missing and could not be built ({e}); "
                        "run `python -c 'import __graft_entry__ as g; g.build()'`") from e
            lib = _c.CDLL(str(path))
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)   # AttributeError if the ABI drifted: fail loudly
                fn.restype, fn.argtypes = res, args
            _lib = lib
    return _lib


def check(rc):
    if rc == EXB_OK:
        return
    msg = load().exb_last_error().decode("utf-8", "replace")
    if rc == EXB_EKRANGE:
        # same exception type/text as torch.topk in the reference (utils/utils.py:297)
        raise RuntimeError("selected index k out of range")
    if rc in (EXB_EINVAL, EXB_EUNSUPPORTED):
        raise ValueError(f"exorl_b200: {msg}")
    raise RuntimeError(f"exorl_b200 (status {rc}): {msg}")


def stream_ptr(device=None):
    return _c.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    return _c.c_void_p(t.data_ptr()) if t is not None else _c.c_void_p(0)


class Workspace:
    """Grow-only device scratch, one per (device, purpose); reuse is stream-ordered."""
    _cache = {}

    @classmethod
    def get(cls, device, key, nbytes, zero_init=False):
        dev = torch.device(device)
        k = (dev.index if dev.index is not None else torch.cuda.current_device(), key)
        buf = cls._cache.get(k)
        fresh = buf is None or buf.numel() < nbytes
        if fresh:
            buf = torch.empty(max(int(nbytes), 1024), dtype=torch.uint8, device=dev)
            if zero_init:
                buf.zero_()
            cls._cache[k] = buf
        return buf, fresh

    @classmethod
    def size(cls, device, key):
        """Bytes of the cached workspace (0 if none)."""
        dev = torch.device(device)
        k = (dev.index if dev.index is not None else torch.cuda.current_device(), key)
        buf = cls._cache.get(k)
        return 0 if buf is None else int(buf.numel())


def require_cuda(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise RuntimeError(f"exorl_b200: `{name}` must be a CUDA tensor — this implementation has "
                           "no CPU path (use the reference for CPU tensors)")
