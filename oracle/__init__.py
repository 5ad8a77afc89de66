"""CPU oracle for the particle-based kNN reward + n-step replay gather hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``exorl_b200/`` may import this package; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs use it, and only as the checker / CPU baseline.

Parity status: PINNED against the unmodified reference run in the authoring container
(``oracle/gen_golden.py`` imports ``/root/reference/utils/utils.py`` and
``/root/reference/utils/replay_buffer.py`` by file path and writes ``tests/golden/*.npz``), and
against the known-answer vectors recorded in SURVEY.md §8c.  The reference ships no tests of its
own for this path, so there are no upstream golden vectors to pin against.
"""
from .pbe_oracle import (rms_step, pbe_reward, knn_smallest, proto_reward, proto_enqueue,
                         proto_candidates, rnd_reward, RmsState)
from .rep_oracle import icm_trunk_rep, l2_normalize
from .replay_oracle import (nstep_transition, onestep_transition, episode_length,
                            flatten_episodes, write_episode_npz, read_episode_npz,
                            philox4x32_10, draw_indices_philox)
