"""`torch.ops.exorl_b200.*` — the hot path as PyTorch custom ops over the C ABI.

north_star: "the host side is a thin PyTorch custom op over a C-ABI".  Each op below is the CUDA
implementation of one reference call site, registered with `torch.library` together with a fake
(meta) kernel, so that shapes/dtypes are known without running and an agent update compiled with
`torch.compile(fullgraph=True)` keeps the reward inside the graph instead of breaking at it:

  exorl_b200::knn_topk      utils/utils.py:293-300 (distance + topk, indices kept)
  exorl_b200::apt_reward    utils/utils.py:289-319 = PBE.__call__ / compute_apt_reward;
                            call sites icm_apt.py:96-100, aps.py:167-172
  exorl_b200::proto_reward  proto.py:115-118
  exorl_b200::proto_sample  proto.py:105-111 (candidate per prototype + queue write)
  exorl_b200::rms_update    utils/utils.py:264-276 (RMS.__call__ on an (n, 1) column; rnd.py:96)
  exorl_b200::rnd_reward    rnd.py:52-54, 94-99
  exorl_b200::replay_sample utils/replay_buffer.py:214-235 / :87-96 (index draw + gather of a batch)

The RMS state (M, S fp32 (1,), n fp64 (1,)) lives on the device and is declared MUTATED by the
ops that update it — there is no host-side value that changes from call to call, which is also
what lets a CUDA graph replay an update.  The implementations are the functions of knn.py /
replay.py that bind the C ABI through ctypes (`_lib`); the Python classes (RMS, PBE,
DeviceReplayBuffer) are thin callers of these ops.  CUDA only: there is no CPU kernel.
"""
from typing import List, Sequence, Tuple

import torch

_NS = "exorl_b200"


def _knn():
    from . import knn
    return knn


def call(op, impl, *args):
    """How the Python classes reach an op: through the dispatcher while a graph is being traced
    (torch.compile sees one opaque node with a fake kernel), straight into the implementation in
    eager mode — the dispatcher's Python-side schema handling costs more host time than the
    per-update kernels run (tens of microseconds per call)."""
    if torch.compiler.is_compiling():
        return op(*args)
    return impl(*args)


@torch.library.custom_op(f"{_NS}::knn_topk", mutates_args=(), device_types="cuda")
def knn_topk(source: torch.Tensor, target: torch.Tensor, k: int, idx_base: int, knn_mode: str,
             same: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    return _knn()._knn_topk_impl(source, target, k, idx_base, knn_mode, same)


@knn_topk.register_fake
def _(source, target, k, idx_base, knn_mode, same):
    nq = source.shape[0]
    return (source.new_empty((nq, k), dtype=torch.float32),
            source.new_empty((nq, k), dtype=torch.int32))


@torch.library.custom_op(f"{_NS}::apt_reward", mutates_args=("rms_m", "rms_s", "rms_n"),
                         device_types="cuda")
def apt_reward(source: torch.Tensor, target: torch.Tensor, rms_m: torch.Tensor,
               rms_s: torch.Tensor, rms_n: torch.Tensor, k: int, knn_avg: bool, knn_rms: bool,
               knn_clip: float, knn_mode: str,
               same: bool) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return _knn()._apt_reward_impl(source, target, rms_m, rms_s, rms_n, k, knn_avg, knn_rms,
                                   knn_clip, knn_mode, same)


@apt_reward.register_fake
def _(source, target, rms_m, rms_s, rms_n, k, knn_avg, knn_rms, knn_clip, knn_mode, same):
    nq = source.shape[0]
    return (source.new_empty((nq, 1), dtype=torch.float32),
            source.new_empty((nq, k), dtype=torch.float32),
            source.new_empty((nq, k), dtype=torch.int32))


@torch.library.custom_op(f"{_NS}::proto_reward", mutates_args=(), device_types="cuda")
def proto_reward(z: torch.Tensor, queue: torch.Tensor, topk: int, knn_mode: str) -> torch.Tensor:
    return _knn()._proto_reward_impl(z, queue, topk, knn_mode)


@proto_reward.register_fake
def _(z, queue, topk, knn_mode):
    return z.new_empty((z.shape[0], 1), dtype=torch.float32)


@torch.library.custom_op(f"{_NS}::proto_sample", mutates_args=("queue",), device_types="cuda")
def proto_sample(z: torch.Tensor, protos: torch.Tensor, queue: torch.Tensor, queue_ptr: int,
                 noise: torch.Tensor, seed: int, call: int) -> torch.Tensor:
    """proto.py:105-111: one batch row per prototype (exponential race on softmax(protos(z).T)),
    written into queue[queue_ptr : queue_ptr + P].  `noise` of numel 0 = Philox (seed, call)."""
    return _knn()._proto_sample_impl(z, protos, queue, queue_ptr, noise, seed, call)


@proto_sample.register_fake
def _(z, protos, queue, queue_ptr, noise, seed, call):
    return z.new_empty((protos.shape[0],), dtype=torch.int64)


@torch.library.custom_op(f"{_NS}::rms_update", mutates_args=("rms_m", "rms_s", "rms_n"),
                         device_types="cuda")
def rms_update(x: torch.Tensor, rms_m: torch.Tensor, rms_s: torch.Tensor,
               rms_n: torch.Tensor) -> None:
    _knn()._rms_update_impl(x, rms_m, rms_s, rms_n)


@rms_update.register_fake
def _(x, rms_m, rms_s, rms_n):
    return None


@torch.library.custom_op(f"{_NS}::rnd_reward", mutates_args=("rms_m", "rms_s", "rms_n"),
                         device_types="cuda")
def rnd_reward(prediction: torch.Tensor, target: torch.Tensor, rms_m: torch.Tensor,
               rms_s: torch.Tensor, rms_n: torch.Tensor,
               rnd_scale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    return _knn()._rnd_reward_impl(prediction, target, rms_m, rms_s, rms_n, rnd_scale)


@rnd_reward.register_fake
def _(prediction, target, rms_m, rms_s, rms_n, rnd_scale):
    n = prediction.shape[0]
    return (prediction.new_empty((n, 1), dtype=torch.float32),
            prediction.new_empty((n, 1), dtype=torch.float32))


@torch.library.custom_op(f"{_NS}::replay_sample", mutates_args=(), device_types="cuda")
def replay_sample(columns: Sequence[torch.Tensor], row_offsets: Sequence[int],
                  reward: torch.Tensor, discount: torch.Tensor, eps_start: torch.Tensor,
                  eps_len: torch.Tensor, batch: int, seed: int, call: int, nstep: int,
                  nstep_mode: int, gamma: float) -> List[torch.Tensor]:
    """One batch: Philox index draw + gather in ONE launch.  `columns[i]` is gathered at row
    `draw + row_offsets[i]`; returns [*gathered columns, reward (B,1), discount (B,1), rows (B,)]."""
    from . import replay
    return replay._replay_sample_impl(columns, row_offsets, reward, discount, eps_start, eps_len,
                                      batch, seed, call, nstep, nstep_mode, gamma)


@replay_sample.register_fake
def _(columns, row_offsets, reward, discount, eps_start, eps_len, batch, seed, call, nstep,
      nstep_mode, gamma):
    out = [c.new_empty((batch, *c.shape[1:])) for c in columns]
    out.append(reward.new_empty((batch, *reward.shape[1:])))
    out.append(reward.new_empty((batch, *reward.shape[1:])))
    out.append(reward.new_empty((batch,), dtype=torch.int64))
    return out
