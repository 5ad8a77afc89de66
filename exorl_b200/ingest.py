"""Episode ingestion: the reference's `episode_{idx}_{len}.npz` files -> host staging tensors.

Replaces np.load in load_episode (/root/reference/utils/replay_buffer.py:26-29) and the
DataLoader worker processes that call it (:192-212, :57-75): the archive is parsed and inflated by
the native reader (csrc/npz_ingest.cu, exb_npz_*), many files at a time on a thread pool (the
calls release the GIL), into pinned tensors when a CUDA device is present so that the copy into
the device-resident columns is an asynchronous DMA.
"""
import concurrent.futures
import ctypes
import os

import numpy as np
import torch

from . import _lib

_DTYPES = {"<f4": torch.float32, "<f8": torch.float64, "<f2": torch.float16, "|u1": torch.uint8,
           "|i1": torch.int8, "|b1": torch.bool, "<i2": torch.int16, "<i4": torch.int32,
           "<i8": torch.int64}


def read_episode(path, pin=None, keys=None):
    """dict name -> CPU tensor with the arrays of one episode file, like load_episode (:26-29)
    (which returns numpy arrays of the same dtypes and shapes; `.numpy()` gives those)."""
    lib = _lib.load()
    if pin is None:
        pin = torch.cuda.is_available()
    handle, n = ctypes.c_void_p(), ctypes.c_int(0)
    _lib.check(lib.exb_npz_open(os.fsencode(str(path)), ctypes.byref(handle), ctypes.byref(n)))
    try:
        out = {}
        info = _lib.NpzArray()
        for i in range(n.value):
            _lib.check(lib.exb_npz_info(handle, i, ctypes.byref(info)))
            name = info.name.decode()
            if keys is not None and name not in keys:
                continue
            descr = info.dtype.decode()
            if info.fortran_order or descr not in _DTYPES:
                raise ValueError(f"{path}: array '{name}' has unsupported layout/dtype {descr}")
            shape = tuple(info.shape[d] for d in range(info.ndim))
            t = torch.empty(shape, dtype=_DTYPES[descr], pin_memory=bool(pin))
            if info.nbytes:
                _lib.check(lib.exb_npz_read(handle, i, ctypes.c_void_p(t.data_ptr()), info.nbytes))
            out[name] = t
        return out
    finally:
        lib.exb_npz_close(handle)


def read_episodes(paths, workers=None, pin=None, keys=None):
    """Yield (path, episode dict) in the order of `paths`, decoding up to `workers` files
    concurrently (default: all host cores, at most 16)."""
    paths = list(paths)
    if workers is None:
        workers = min(16, os.cpu_count() or 1)
    if workers <= 1 or len(paths) <= 1:
        for p in paths:
            yield p, read_episode(p, pin=pin, keys=keys)
        return
    with concurrent.futures.ThreadPoolExecutor(max_workers=workers) as ex:
        window = 2 * workers                     # bounded look-ahead: staging memory stays small
        futs = {}
        nxt = 0
        for i, p in enumerate(paths):
            while nxt < len(paths) and nxt < i + window:
                futs[nxt] = ex.submit(read_episode, paths[nxt], pin, keys)
                nxt += 1
            yield p, futs.pop(i).result()


def as_numpy(episode):
    """The reference's in-memory episode: dict of numpy arrays."""
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
            for k, v in episode.items()}
