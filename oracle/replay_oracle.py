"""NumPy restatement of the reference's replay sampling arithmetic and episode layout.

Test infrastructure only (see oracle/__init__.py).  Each function cites the reference lines it
follows in /root/reference/utils/replay_buffer.py.
"""
import io

import numpy as np


def episode_length(episode) -> int:
    """Transitions in an episode = rows - 1 (row 0 is the dummy first step).

    Follows utils/replay_buffer.py:13-15.
    """
    return next(iter(episode.values())).shape[0] - 1


def write_episode_npz(episode, path):
    """Compressed .npz with one array per spec name (utils/replay_buffer.py:18-23)."""
    with io.BytesIO() as buf:
        np.savez_compressed(buf, **episode)
        buf.seek(0)
        with open(path, "wb") as f:
            f.write(buf.read())


def read_episode_npz(path):
    """Inverse of write_episode_npz (utils/replay_buffer.py:25-29)."""
    with open(path, "rb") as f:
        z = np.load(f)
        return {k: z[k] for k in z.keys()}


def nstep_transition(episode, idx, nstep, gamma, meta_names=()):
    """One n-step sample at 1-based position ``idx`` in [1, eps_len - nstep + 1].

    Follows utils/replay_buffer.py:223-235: meta rows at idx-1, obs at idx-1, action at idx,
    next_obs at idx+nstep-1, and the fp32 accumulation
    ``reward += discount * reward[idx+i]; discount *= discount[idx+i] * gamma`` in that order,
    starting from zeros_like / ones_like of the stored (1,)-shaped rows.
    Returns (obs, action, reward, discount, next_obs, *meta).
    """
    meta = [episode[name][idx - 1] for name in meta_names]
    obs = episode["observation"][idx - 1]
    action = episode["action"][idx]
    next_obs = episode["observation"][idx + nstep - 1]
    reward = np.zeros_like(episode["reward"][idx])
    discount = np.ones_like(episode["discount"][idx])
    for i in range(nstep):
        reward += discount * episode["reward"][idx + i]
        discount *= episode["discount"][idx + i] * gamma
    return (obs, action, reward, discount, next_obs, *meta)


def onestep_transition(episode, idx, gamma):
    """Offline 1-step sample at 1-based ``idx`` in [1, eps_len].

    Follows utils/replay_buffer.py:87-96.
    """
    obs = episode["observation"][idx - 1]
    action = episode["action"][idx]
    next_obs = episode["observation"][idx]
    reward = episode["reward"][idx]
    discount = episode["discount"][idx] * gamma
    return (obs, action, reward, discount, next_obs)


def flatten_episodes(episodes, names):
    """Concatenate episodes row-wise into one array per name plus (start_row, eps_len) tables.

    This is the device-resident layout of the new implementation; the reference
    keeps a dict of per-episode arrays (utils/replay_buffer.py:172-190).  Global row of the
    1-based in-episode position ``idx`` of episode e is ``start_row[e] + idx``.
    """
    starts, lens, cols = [], [], {n: [] for n in names}
    row = 0
    for ep in episodes:
        starts.append(row)
        lens.append(episode_length(ep))
        for n in names:
            cols[n].append(ep[n])
        row += lens[-1] + 1
    flat = {n: np.concatenate(cols[n], axis=0) for n in names}
    return flat, np.asarray(starts, np.int64), np.asarray(lens, np.int32)


# ---------------------------------------------------------------------------------------------
# Counter-based index draw used by the device sampler's "fast mode" (no reference counterpart:
# the reference draws with python `random.choice` + legacy `np.random.randint`,
# utils/replay_buffer.py:168-170,222; SURVEY.md §7.2 "Sampling RNG parity").  Uniform over
# episodes, then uniform over valid starts inside the chosen episode — the same two-stage law.
# ---------------------------------------------------------------------------------------------
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(counter, key):
    """Philox-4x32 with 10 rounds (Salmon et al., SC'11).  counter (..., 4) u32, key (2,) u32."""
    c = np.array(counter, dtype=np.uint32, copy=True)
    k0, k1 = np.uint32(key[0]), np.uint32(key[1])
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = c[..., 0].astype(np.uint64) * _M0
            p1 = c[..., 2].astype(np.uint64) * _M1
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            n0 = hi1 ^ c[..., 1] ^ k0
            n2 = hi0 ^ c[..., 3] ^ k1
            c = np.stack([n0, lo1, n2, lo0], axis=-1)
            k0 = np.uint32(k0 + _W0)
            k1 = np.uint32(k1 + _W1)
    return c


def draw_indices_philox(seed, call, batch, eps_start, eps_len, nstep):
    """(episode, idx, global_row) for ``batch`` samples of draw number ``call``.

    counter = (sample, call_lo, call_hi, 0), key = (seed_lo, seed_hi);
    episode = mulhi(r0, n_episodes); idx = mulhi(r1, eps_len - nstep + 1) + 1  (range of
    utils/replay_buffer.py:222); global_row = eps_start[episode] + idx.
    """
    s = np.arange(batch, dtype=np.uint32)
    ctr = np.stack([s, np.full(batch, call & 0xFFFFFFFF, np.uint32),
                    np.full(batch, (call >> 32) & 0xFFFFFFFF, np.uint32),
                    np.zeros(batch, np.uint32)], axis=-1)
    r = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    n_eps = np.uint64(len(eps_len))
    ep = ((r[:, 0].astype(np.uint64) * n_eps) >> np.uint64(32)).astype(np.int64)
    span = (np.asarray(eps_len, np.int64)[ep] - nstep + 1).astype(np.uint64)
    idx = ((r[:, 1].astype(np.uint64) * span) >> np.uint64(32)).astype(np.int64) + 1
    return ep, idx, np.asarray(eps_start, np.int64)[ep] + idx
