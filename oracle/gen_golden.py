"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the authoring container.

Run from the repo root:  python oracle/gen_golden.py
Needs /root/reference (absent on the GPU box, which only reads the committed fixtures).
The reference modules are loaded by file path; `omegaconf` (imported but unused by
utils/utils.py:10, absent from this image) is stubbed.  Nothing is copied from the reference.
"""
import importlib.util
import pathlib
import random
import sys
import tempfile
import types

import numpy as np
import torch

REF = pathlib.Path("/root/reference")
OUT = pathlib.Path(__file__).resolve().parent.parent / "tests" / "golden"


def load_reference():
    sys.modules.setdefault("omegaconf", types.SimpleNamespace(OmegaConf=object))
    mods = {}
    for name, rel in (("ref_utils", "utils/utils.py"), ("ref_replay", "utils/replay_buffer.py")):
        spec = importlib.util.spec_from_file_location(name, REF / rel)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        mods[name] = m
    return mods["ref_utils"], mods["ref_replay"]


def reps(seed, b, d):
    g = torch.Generator().manual_seed(seed)
    return torch.tanh(torch.randn(b, d, generator=g))


def gen_pbe(ref_utils):
    """PBE / RMS trajectories: every (avg, rms, clip) mode, several k and D, 3 sequential calls.
    The three input reps of a (b, d) shape are stored once and shared by all of its mode cases."""
    out = {}
    case = 0
    for (b, d) in ((96, 512), (64, 50), (80, 10), (72, 24), (64, 128)):
        xs = [reps(1000 + 31 * d + call, b, d) for call in range(3)]
        for call, x in enumerate(xs):
            out[f"rep_{b}_{d}_{call}"] = x.numpy()
        for k in (1, 3, 12):
            for avg in (True, False):
                for rms_on in (True, False):
                    for clip in (0.0, 0.0001, 0.5, -1.0):
                        if d != 512 and clip in (0.0001, -1.0):
                            continue
                        rms = ref_utils.RMS("cpu")
                        pbe = ref_utils.PBE(rms, clip, k, avg, rms_on, "cpu")
                        tag = f"pbe{case:03d}"
                        out[f"{tag}_cfg"] = np.array([b, d, k, int(avg), int(rms_on)], np.int64)
                        out[f"{tag}_clip"] = np.array(clip, np.float64)
                        for call, x in enumerate(xs):
                            r = pbe(x.clone())
                            out[f"{tag}_reward{call}"] = r.numpy()
                            out[f"{tag}_state{call}"] = np.array(
                                [float(rms.M), float(rms.S), rms.n], np.float64)
                        case += 1
    out["n_cases"] = np.array(case)
    np.savez_compressed(OUT / "pbe_reference.npz", **out)
    print("pbe cases:", case)


def gen_kat(ref_utils):
    """The 4-point known-answer vectors of SURVEY.md §8c, re-derived from the reference."""
    rep = torch.tensor([[0., 0.], [3., 4.], [6., 8.], [0., 1.]])
    out = {"rep": rep.numpy()}
    for name, (avg, rms_on, clip) in {
            "avg_rms_c0": (True, True, 0.0), "avg_rms_c05": (True, True, 0.5),
            "avg_norms_c0": (True, False, 0.0), "avg_norms_c05": (True, False, 0.5),
            "kth_rms_c0": (False, True, 0.0), "kth_rms_c05": (False, True, 0.5),
            "kth_norms_c0": (False, False, 0.0), "avg_rms_cneg": (True, True, -1.0)}.items():
        rms = ref_utils.RMS("cpu")
        pbe = ref_utils.PBE(rms, clip, 2, avg, rms_on, "cpu")
        for call in range(2):
            out[f"{name}_r{call}"] = pbe(rep.clone()).numpy()
            out[f"{name}_s{call}"] = np.array([float(rms.M), float(rms.S), rms.n], np.float64)
    rms = ref_utils.RMS("cpu")
    rms(torch.tensor([[1.], [2.], [3.], [6.]]))
    out["rms_a"] = np.array([float(rms.M), float(rms.S), rms.n])
    rms(torch.tensor([[10.], [20.]]))
    out["rms_b"] = np.array([float(rms.M), float(rms.S), rms.n])
    np.savez_compressed(OUT / "pbe_kat.npz", **out)


def gen_knn_dists(ref_utils):
    """Distance/top-k of the reference expression itself (utils.py:293-300) with indices kept,
    including B1 != B2 (compute_apt_reward form) and a 3-D rep."""
    out = {}
    for tag, (b1, b2, d, k) in {"sq512": (128, 128, 512, 12), "rect": (40, 200, 64, 12),
                                "proto": (64, 256, 128, 3), "tiny": (16, 16, 10, 5)}.items():
        s = reps(50 + b1, b1, d)
        t = s if b1 == b2 else reps(51 + b2, b2, d)
        sim = torch.norm(s[:, None, :].view(b1, 1, -1) - t[None, :, :].view(1, b2, -1),
                         dim=-1, p=2)
        dist, idx = sim.topk(k, dim=1, largest=False, sorted=True)
        kp1 = sim.topk(min(k + 1, b2), dim=1, largest=False, sorted=True)[0]
        out[f"{tag}_src"], out[f"{tag}_tgt"] = s.numpy(), t.numpy()
        out[f"{tag}_dist"], out[f"{tag}_idx"] = dist.numpy(), idx.numpy()
        out[f"{tag}_kp1"] = kp1.numpy()
    np.savez_compressed(OUT / "knn_reference.npz", **out)


def gen_proto():
    """proto.py:110-118 evaluated verbatim on seeded z / queue (the module itself imports the
    `utils` package, which needs dm_control, so its five lines are executed here directly)."""
    out = {}
    g = torch.Generator().manual_seed(1)
    z = torch.nn.functional.normalize(torch.randn(96, 128, generator=g), dim=1, p=2)
    g = torch.Generator().manual_seed(2)
    full = torch.nn.functional.normalize(torch.randn(256, 128, generator=g), dim=1, p=2)
    for tag, queue in (("full", full.clone()),
                       ("warm", torch.cat([full[:64], torch.zeros(192, 128)]))):
        ptr = 0
        cand = torch.arange(0, 32)
        queue[ptr:ptr + 32] = z[cand]
        ptr = (ptr + 32) % queue.shape[0]
        z_to_q = torch.norm(z[:, None, :] - queue[None, :, :], dim=2, p=2)
        all_dists, _ = torch.topk(z_to_q, 3, dim=1, largest=False)
        out[f"{tag}_queue_after"] = queue.numpy()
        out[f"{tag}_reward"] = all_dists[:, -1:].numpy()
        out[f"{tag}_ptr"] = np.array(ptr)
    out["z"] = z.numpy()
    out["queue_full"] = full.numpy()
    np.savez_compressed(OUT / "proto_reference.npz", **out)


def gen_proto_sampling():
    """proto.py:102-112 evaluated verbatim (normalize, protos(z).T, softmax, Categorical.sample,
    queue slice write) under a recorded seed; the Exp(1) tensor torch.multinomial consumed is
    re-drawn from the same seed and stored, and checked to reproduce the sample."""
    import torch.distributions as pyd
    out = {}
    for tag, (b, n_protos, dim, qsize, ptr, temp) in {
            "small": (192, 64, 128, 256, 64, 1.0), "peaked": (160, 32, 64, 128, 96, 20.0),
            "odd": (77, 24, 50, 96, 0, 3.0)}.items():
        g = torch.Generator().manual_seed(500 + b)
        z = torch.nn.functional.normalize(torch.randn(b, dim, generator=g), dim=1, p=2)
        protos = torch.nn.Linear(dim, n_protos, bias=False)
        with torch.no_grad():
            w = torch.nn.functional.normalize(torch.randn(n_protos, dim, generator=g), dim=1, p=2)
            protos.weight.copy_(w * temp)       # temp > 1: sharper softmax (trained prototypes)
            queue = torch.nn.functional.normalize(torch.randn(qsize, dim, generator=g), dim=1)
            queue_before = queue.clone()
            torch.manual_seed(900 + b)
            scores = protos(z).T
            prob = torch.nn.functional.softmax(scores, dim=1)
            candidates = pyd.Categorical(prob).sample()
            queue[ptr:ptr + n_protos] = z[candidates]
            new_ptr = (ptr + n_protos) % queue.shape[0]
            torch.manual_seed(900 + b)
            noise = torch.empty_like(prob).exponential_(1)
            probs = pyd.Categorical(prob).probs
            assert torch.equal(torch.argmax(probs / noise, dim=1), candidates)
        out[f"{tag}_z"], out[f"{tag}_w"] = z.numpy(), protos.weight.detach().numpy()
        out[f"{tag}_noise"], out[f"{tag}_prob"] = noise.numpy(), probs.numpy()
        out[f"{tag}_candidates"] = candidates.numpy()
        out[f"{tag}_queue_before"], out[f"{tag}_queue_after"] = queue_before.numpy(), queue.numpy()
        out[f"{tag}_ptr"] = np.array([ptr, new_ptr])
    np.savez_compressed(OUT / "proto_sampling_reference.npz", **out)


def synth_episode(rng, eps_len, obs_shape, obs_dtype, terminal, with_meta):
    n = eps_len + 1
    if obs_dtype == np.uint8:
        obs = rng.integers(0, 256, size=(n, *obs_shape), dtype=np.uint8)
    else:
        obs = rng.standard_normal((n, *obs_shape)).astype(np.float32)
    disc = np.ones((n, 1), np.float32)
    if terminal:
        disc[-1] = 0.0
    ep = {"observation": obs,
          "action": rng.uniform(-1, 1, (n, 6)).astype(np.float32),
          "reward": rng.uniform(0, 1, (n, 1)).astype(np.float32),
          "discount": disc}
    if with_meta:
        ep["skill"] = rng.standard_normal((n, 16)).astype(np.float32)
        ep["done"] = (rng.uniform(size=(n,)) < 0.1)
    return ep


def gen_replay(ref_replay):
    """ReplayBuffer._sample / OfflineReplayBuffer._sample of the reference over synthetic episode
    directories written with the reference's own save_episode; the (episode, idx) draws are
    recorded by replaying the same RNG streams."""
    out = {}
    for tag, (obs_shape, obs_dtype, with_meta, nstep) in {
            "state_n3": ((24,), np.float32, False, 3), "state_n1": ((24,), np.float32, False, 1),
            "meta_n3": ((24,), np.float32, True, 3), "pix_n3": ((9, 12, 12), np.uint8, False, 3),
            "state_n5": ((24,), np.float32, False, 5)}.items():
        rng = np.random.default_rng(3)
        with tempfile.TemporaryDirectory() as td:
            td = pathlib.Path(td)
            lens = [20, 7, 33, 12, 9, 50]
            eps = []
            for e, L in enumerate(lens):
                ep = synth_episode(rng, L, obs_shape, obs_dtype, terminal=(e % 2 == 0),
                                   with_meta=with_meta)
                ref_replay.save_episode(ep, td / f"episode_{e}_{L}.npz")
                eps.append(ep)
            meta_specs = ((types.SimpleNamespace(name="skill"), types.SimpleNamespace(name="done"))
                          if with_meta else ())
            storage = types.SimpleNamespace(_replay_dir=td, _meta_specs=meta_specs)
            buf = ref_replay.ReplayBuffer(storage, 10 ** 6, 1, nstep, 0.99, fetch_every=1000,
                                          save_snapshot=True)
            random.seed(11)
            np.random.seed(11)
            samples = [buf._sample() for _ in range(64)]
            # replay the identical draw sequence to recover (episode, idx)
            random.seed(11)
            np.random.seed(11)
            fns = list(buf._episode_fns)
            draws = []
            for _ in range(64):
                fn = random.choice(fns)
                e = int(fn.stem.split("_")[1])
                idx = np.random.randint(0, lens[e] - nstep + 1) + 1
                draws.append((e, idx))
            out[f"{tag}_draws"] = np.array(draws, np.int64)
            out[f"{tag}_lens"] = np.array(lens, np.int64)
            out[f"{tag}_nstep"] = np.array(nstep)
            for name in eps[0]:
                for e, ep in enumerate(eps):
                    out[f"{tag}_ep{e}_{name}"] = ep[name]
            for j in range(len(samples[0])):
                out[f"{tag}_out{j}"] = np.stack([np.asarray(s[j]) for s in samples])
            # boundary positions, evaluated by the reference's own arithmetic through _sample
            # cannot be forced; they are covered through the oracle (pinned on the draws above).
            if tag == "state_n3":
                off = ref_replay.OfflineReplayBuffer(None, td, 10 ** 6, 1, 0.99)
                off._load(relable=False)
                off._loaded = True
                random.seed(5)
                np.random.seed(5)
                osamples = [off._sample() for _ in range(64)]
                random.seed(5)
                np.random.seed(5)
                ofns = list(off._episode_fns)
                odraws = []
                for _ in range(64):
                    fn = random.choice(ofns)
                    e = int(fn.stem.split("_")[1])
                    odraws.append((e, np.random.randint(0, lens[e]) + 1))
                out["offline_draws"] = np.array(odraws, np.int64)
                for j in range(5):
                    out[f"offline_out{j}"] = np.stack([np.asarray(s[j]) for s in osamples])
    np.savez_compressed(OUT / "replay_reference.npz", **out)


RESIDENCY_ONLINE = {
    # name: (max_size, num_workers, save_snapshot, rounds of [(eps_idx, eps_len), ...])
    # Every round first writes its files, then each worker fetches once (:192-212).
    "w1_snapshot": (50, 1, True, [[(0, 10), (1, 12), (2, 9), (3, 11), (4, 10)],
                                  [(5, 10), (6, 14)],
                                  [(7, 8), (8, 10), (9, 10), (10, 10), (11, 12)],   # 9 -> 10: names sort as text
                                  [(12, 30)],
                                  [(13, 10), (14, 10), (15, 10), (16, 10), (17, 10), (18, 10), (19, 10)]]),
    "w1_no_snapshot": (40, 1, False, [[(0, 10), (1, 10), (2, 10)], [(3, 15), (4, 15)],
                                      [(5, 10), (6, 7), (7, 13)], [(8, 10), (9, 10), (10, 10)]]),
    "w2_snapshot": (60, 2, True, [[(0, 10), (1, 10), (2, 10), (3, 10)],
                                  [(4, 12), (5, 9), (6, 11), (7, 10), (8, 10)],
                                  [(9, 10), (10, 10), (11, 10), (12, 10), (13, 10)],
                                  [(14, 25), (15, 25), (16, 5)]]),
    "first_fetch_over_cap": (35, 1, True, [[(e, 10) for e in range(8)], [(8, 10)], [(9, 10), (10, 10)]]),
}
RESIDENCY_OFFLINE = {
    # name: (max_size as passed to the factory, num_workers, [(eps_idx, eps_len), ...])
    "w1_cap_inside": (45, 1, [(e, 10) for e in range(12)]),
    "w1_ragged": (30, 1, [(0, 7), (1, 12), (2, 9), (3, 3), (4, 20), (5, 8), (6, 10)]),
    "w2": (50, 2, [(e, 10 + e % 3) for e in range(14)]),
    "under_cap": (1000, 1, [(e, 10) for e in range(5)]),
}


def residency_episode(eps_idx, eps_len):
    """The (tiny) episode written for a residency scenario; the tests rebuild the same files."""
    rng = np.random.default_rng(1000 + eps_idx)
    return synth_episode(rng, eps_len, (3,), np.float32, terminal=False, with_meta=False)


def gen_residency(ref_replay):
    """Which episode files the reference keeps resident and which it leaves on disk.

    Online: ReplayBuffer._try_fetch/_store_episode (utils/replay_buffer.py:172-212) over a
    directory that grows between fetches — sorted-name eviction with unlink, save_snapshot on and
    off, one and two DataLoader workers (the worker id the reference reads from
    torch.utils.data.get_worker_info() is supplied by patching that function, nothing in the
    reference is modified).  Offline: OfflineReplayBuffer._load (:57-75) over directories larger
    and smaller than the cap."""
    import json
    import torch.utils.data as tud
    real_info = tud.get_worker_info
    out = {"online": {}, "offline": {}}
    try:
        for name, (max_size, workers, snapshot, rounds) in RESIDENCY_ONLINE.items():
            with tempfile.TemporaryDirectory() as td:
                td = pathlib.Path(td)
                storage = types.SimpleNamespace(_replay_dir=td, _meta_specs=())
                bufs = [ref_replay.ReplayBuffer(storage, max_size // workers, workers, 3, 0.99,
                                                fetch_every=1000, save_snapshot=snapshot)
                        for _ in range(workers)]
                log = []
                for files in rounds:
                    for eps_idx, eps_len in files:
                        ref_replay.save_episode(residency_episode(eps_idx, eps_len),
                                                td / f"episode_{eps_idx}_{eps_len}.npz")
                    for w, buf in enumerate(bufs):
                        tud.get_worker_info = lambda w=w: types.SimpleNamespace(id=w)
                        buf._samples_since_last_fetch = buf._fetch_every
                        buf._try_fetch()
                    log.append({"written": files,
                                "resident": [[fn.name for fn in b._episode_fns] for b in bufs],
                                "size": [b._size for b in bufs],
                                "on_disk": sorted(fn.name for fn in td.glob("*.npz"))})
                out["online"][name] = {"max_size": max_size, "num_workers": workers,
                                       "save_snapshot": snapshot, "rounds": log}
        for name, (max_size, workers, files) in RESIDENCY_OFFLINE.items():
            with tempfile.TemporaryDirectory() as td:
                td = pathlib.Path(td)
                for eps_idx, eps_len in files:
                    ref_replay.save_episode(residency_episode(eps_idx, eps_len),
                                            td / f"episode_{eps_idx}_{eps_len}.npz")
                resident, size = [], []
                for w in range(workers):
                    tud.get_worker_info = lambda w=w: types.SimpleNamespace(id=w)
                    off = ref_replay.OfflineReplayBuffer(None, td, max_size // workers, workers,
                                                         0.99)
                    off._load(relable=False)
                    resident.append([fn.name for fn in off._episode_fns])
                    size.append(off._size)
                out["offline"][name] = {"max_size": max_size, "num_workers": workers,
                                        "files": files, "resident": resident, "size": size}
    finally:
        tud.get_worker_info = real_info
    (OUT / "residency_reference.json").write_text(json.dumps(out, indent=1) + "\n")
    print("residency scenarios:", len(out["online"]), "online,", len(out["offline"]), "offline")


def load_reference_icm(ref_utils):
    """class ICM of agents/unsupervised_learning/icm_apt.py:10-52, executed from the reference's
    own source text (the module imports hydra / the `utils` package / .ddpg, absent here, so only
    the class statement is compiled, with `utils` bound to the reference's utils/utils.py)."""
    import ast
    path = REF / "agents/unsupervised_learning/icm_apt.py"
    tree = ast.parse(path.read_text())
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ICM"]
    ns = {"torch": torch, "nn": torch.nn, "F": torch.nn.functional, "utils": ref_utils}
    exec(compile(ast.Module(body=cls, type_ignores=[]), str(path), "exec"), ns)
    return ns["ICM"]


def gen_rep(ref_utils):
    """Rep producers: ICM.get_rep (trunk Linear -> LayerNorm -> Tanh) on seeded
    observations, then the reference PBE on that rep; F.normalize as used by proto.py:104."""
    ICM = load_reference_icm(ref_utils)
    out = {}
    for tag, (n, obs_dim, rep_dim, scale) in {"walker": (160, 24, 512, 1.0),
                                              "small": (96, 17, 128, 3.0),
                                              "wide": (64, 300, 512, 0.5),
                                              "odd": (40, 5, 36, 1.0)}.items():
        torch.manual_seed(100 + obs_dim)
        icm = ICM(obs_dim, 6, 64, rep_dim)
        g = torch.Generator().manual_seed(200 + obs_dim)
        with torch.no_grad():   # trained-looking parameters (weight_init leaves bias 0, LN 1/0)
            icm.trunk[0].bias.copy_(0.1 * torch.randn(rep_dim, generator=g))
            icm.trunk[1].weight.copy_(1.0 + 0.2 * torch.randn(rep_dim, generator=g))
            icm.trunk[1].bias.copy_(0.1 * torch.randn(rep_dim, generator=g))
            obs = scale * torch.randn(n, obs_dim, generator=g)
            rep = icm.get_rep(obs, None)
            rms = ref_utils.RMS("cpu")
            reward = ref_utils.PBE(rms, 0.0, 12, True, True, "cpu")(rep.clone())
        out[f"{tag}_obs"] = obs.numpy()
        out[f"{tag}_w"] = icm.trunk[0].weight.detach().numpy()
        out[f"{tag}_b"] = icm.trunk[0].bias.detach().numpy()
        out[f"{tag}_lnw"] = icm.trunk[1].weight.detach().numpy()
        out[f"{tag}_lnb"] = icm.trunk[1].bias.detach().numpy()
        out[f"{tag}_eps"] = np.array(icm.trunk[1].eps, np.float64)
        out[f"{tag}_rep"] = rep.numpy()
        out[f"{tag}_reward"] = reward.numpy()
    g = torch.Generator().manual_seed(7)
    x = torch.randn(200, 128, generator=g) * torch.logspace(-3, 3, 200)[:, None]
    x[5] = 0.0        # all-zero row: F.normalize's eps clamp (the warm-up queue rows)
    out["norm_x"] = x.numpy()
    out["norm_out"] = torch.nn.functional.normalize(x, dim=1, p=2).numpy()
    np.savez_compressed(OUT / "rep_reference.npz", **out)


def gen_rnd(ref_utils):
    """RND (rnd.py:13-55) run unmodified (class statement compiled from the reference source, like
    ICM above; predictor / target outputs captured with forward hooks), then the three lines of
    RNDAgent.compute_intr_reward (:95-98) with the reference RMS, over three sequential calls."""
    import ast
    import copy
    path = REF / "agents/unsupervised_learning/rnd.py"
    tree = ast.parse(path.read_text())
    cls = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "RND"]
    ns = {"torch": torch, "nn": torch.nn, "F": torch.nn.functional, "utils": ref_utils,
          "copy": copy}
    exec(compile(ast.Module(body=cls, type_ignores=[]), str(path), "exec"), ns)
    out = {}
    for tag, (n, rep_dim, scale) in {"walker": (96, 512, 1.0), "odd": (77, 50, 0.5)}.items():
        torch.manual_seed(300 + rep_dim)
        rnd = ns["RND"](24, 64, rep_dim, torch.nn.Identity(), torch.nn.Identity(), (24,), "states")
        cap = {}
        rnd.predictor.register_forward_hook(lambda m, i, o: cap.__setitem__("pred", o.detach()))
        rnd.target.register_forward_hook(lambda m, i, o: cap.__setitem__("target", o.detach()))
        rms = ref_utils.RMS("cpu")
        g = torch.Generator().manual_seed(400 + rep_dim)
        for call in range(3):
            obs = torch.randn(n, 24, generator=g) * (1.0 + call)
            with torch.no_grad():
                prediction_error = rnd(obs)
                _, intr_reward_var = rms(prediction_error)
                reward = scale * prediction_error / (torch.sqrt(intr_reward_var) + 1e-8)
            out[f"{tag}_pred{call}"] = cap["pred"].numpy()
            out[f"{tag}_target{call}"] = cap["target"].numpy()
            out[f"{tag}_error{call}"] = prediction_error.numpy()
            out[f"{tag}_reward{call}"] = reward.numpy()
            out[f"{tag}_state{call}"] = np.array([float(rms.M), float(rms.S), rms.n], np.float64)
        out[f"{tag}_scale"] = np.array(scale, np.float64)
    np.savez_compressed(OUT / "rnd_reference.npz", **out)


if __name__ == "__main__":
    torch.set_num_threads(8)
    OUT.mkdir(parents=True, exist_ok=True)
    ref_utils, ref_replay = load_reference()
    if "--rep-only" in sys.argv:
        gen_rep(ref_utils)
        sys.exit(0)
    if "--proto-sampling-only" in sys.argv:
        gen_proto_sampling()
        sys.exit(0)
    if "--residency-only" in sys.argv:
        gen_residency(ref_replay)
        sys.exit(0)
    if "--rnd-only" in sys.argv:
        gen_rnd(ref_utils)
        sys.exit(0)
    gen_rep(ref_utils)
    gen_rnd(ref_utils)
    gen_kat(ref_utils)
    gen_pbe(ref_utils)
    gen_knn_dists(ref_utils)
    gen_proto()
    gen_proto_sampling()
    gen_replay(ref_replay)
    gen_residency(ref_replay)
    for f in sorted(OUT.glob("*.npz")):
        print(f.name, f.stat().st_size)
