"""exorl_b200 — B200-native particle-based kNN intrinsic reward + n-step replay gather.

Drop-in surface of the reference's hot path:
  RMS, PBE                      <- utils/utils.py:257-319
  compute_apt_reward            <- the pre-refactor name of the same computation (source != target)
  proto_knn_reward, proto_enqueue <- agents/unsupervised_learning/proto.py:110-118
  proto_sample_enqueue          <- proto.py:105-112 (candidate sampling fused with the queue write)
  rnd_intr_reward               <- RND error + RMS normalisation, agents/unsupervised_learning/rnd.py:52-54,94-99
  make_replay_loader, make_offline_replay_loader <- utils/replay_buffer.py:246-277
  icm_get_rep, normalize_rep    <- ICM.trunk (icm_apt.py:17-18,50-52), F.normalize (proto.py:104):
                                   rep producers that also emit the prepared kNN operand
  read_episode(s)               <- load_episode, utils/replay_buffer.py:26-29 (native .npz reader)
"""
from .knn import (RMS, PBE, compute_apt_reward, knn_topk, proto_knn_reward, proto_enqueue,
                  proto_sample_enqueue, rnd_intr_reward)
from .replay import (DeviceReplayBuffer, DeviceReplayLoader, make_replay_loader,
                     make_offline_replay_loader)
from .relabel import relabel_apt_reward, relabel_apt_reward_from_host, reps_from_host
from .ingest import read_episode, read_episodes
from .rep import (PreparedRep, icm_get_rep, ln_tanh_rep, normalize_rep, prepare_rep,
                  column_affine, column_stats)

__all__ = ["RMS", "PBE", "compute_apt_reward", "knn_topk", "proto_knn_reward", "proto_enqueue",
           "proto_sample_enqueue", "rnd_intr_reward",
           "DeviceReplayBuffer", "DeviceReplayLoader", "make_replay_loader",
           "make_offline_replay_loader", "relabel_apt_reward", "relabel_apt_reward_from_host", "reps_from_host", "read_episode", "read_episodes",
           "PreparedRep", "icm_get_rep", "ln_tanh_rep", "normalize_rep", "prepare_rep"]
