"""Measurements for BASELINE.json configs 1-4 (the headline config 5 is bench.py's main line).

  C1  PBE 1024 x 512, k=12, avg+rms, clip 0         per-call latency (north_star target < 50 us)
  C2  Proto 1024 z (128) vs 2048 queue, topk 3      per-call latency
  C3  walker-state replay sample B=1024, nstep 3    per-batch latency, + sample -> trunk -> PBE
  C4  pixel replay sample B=1024 / 256 from a 1M-transition uint8 (9,84,84) buffer: GB/s vs HBM peak
  C4b the pixel path's rep producer: ICM trunk 39200 -> 512 as a 3xTF32 tcgen05 GEMM, vs torch fp32
  C5r range-proof screen: 131072 x 131072 kNN on bounded reps vs the same reps * 1e5 + offset

`run_all()` is what `bench.py --gpus 1` embeds as the `"secondary"` object of its JSON line; run as
a script it prints one JSON line per case.  Every case: the public call a user makes, CUDA events
on the launching stream, >= 10 warm-ups, the L2 flushed before EVERY timed iteration (a 256 MB
write followed by a 256 MB read sweep, so the lines left in L2 are clean), a timed region of
>= 2 s so that the nvidia-smi clock sampler (200 ms period) sees it, the clocks record of that
region, and the reference's CPU arithmetic (oracle port) timed on the host cores beside it.
"""
import argparse
import json
import os
import statistics
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

L2_POLICY = ("flushed before every timed iteration (256 MB write + 256 MB read sweep); "
             "median over the timed iterations")


def hbm_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], \
            "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6400.0, "fallback (B200_PROFILING.md)"


class Timer:
    """Per-iteration CUDA-event timing with an L2 flush in front of every iteration."""

    def __init__(self, dev, sampler_factory=None, min_seconds=2.0, min_iters=50, max_iters=20000):
        self.dev = dev
        self.sampler_factory = sampler_factory
        self.min_seconds, self.min_iters, self.max_iters = min_seconds, min_iters, max_iters
        self.scratch = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        self.scratch2 = torch.zeros(64 << 20, dtype=torch.int32, device=dev)

    def flush(self):
        self.scratch.zero_()
        self.scratch2.max()

    def run(self, fn, warm=10):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(self.dev)
        sampler = self.sampler_factory() if self.sampler_factory else None
        ts = []
        t_end = time.perf_counter() + self.min_seconds
        while (len(ts) < self.min_iters or time.perf_counter() < t_end) and len(ts) < self.max_iters:
            self.flush()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        clocks = sampler.stop() if sampler else None
        return {"us_median": statistics.median(ts), "us_min": min(ts), "iters": len(ts),
                "clocks": clocks}

    def run_graph(self, fn):
        """The same call replayed from a CUDA graph captured by the CALLER (torch.cuda.graph)."""
        s = torch.cuda.Stream(self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            for _ in range(3):
                fn()
        torch.cuda.current_stream(self.dev).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return self.run(g.replay, warm=5)


def time_cpu(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e6)
    return statistics.median(ts)


def _graph_or_none(tm, fn):
    try:
        return tm.run_graph(fn)
    except Exception as e:  # caller-side capture is optional evidence
        print("graph capture failed:", e, file=sys.stderr)
        return None


def c1(tm, cpu=True):
    import exorl_b200 as xb
    from oracle import RmsState, pbe_reward
    dev = tm.dev
    g = torch.Generator().manual_seed(0)
    x = torch.tanh(torch.randn(1024, 512, generator=g))
    xd = x.to(dev)
    out = {"workload": "C1 ind_apt PBE: rep 1024 x 512 = tanh(randn) seed 0, knn_k=12, knn_avg, "
                       "knn_rms, clip 0; one PBE.__call__ (utils/utils.py:289-319)",
           "target_us": 50, "l2": L2_POLICY, "algorithmic_flops": 2 * 1024 * 1024 * 512}
    p = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
    out["stream"] = tm.run(lambda: p(xd))
    pg = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
    out["caller_graph"] = _graph_or_none(tm, lambda: pg(xd))
    pe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
    pe.mode = "exact"
    out["exact_kernel_stream"] = tm.run(lambda: pe(xd))
    out["us"] = out["stream"]["us_median"]
    out["meets_target"] = out["us"] < 50
    if cpu:
        out["cpu_reference_us"] = time_cpu(
            lambda: pbe_reward(x, x, RmsState(), 12, True, True, 0.0))
        out["cpu_cores"] = torch.get_num_threads()
    return out


def c2(tm, cpu=True):
    import exorl_b200 as xb
    from oracle import proto_reward
    dev = tm.dev
    g = torch.Generator().manual_seed(1)
    z = torch.nn.functional.normalize(torch.randn(1024, 128, generator=g), dim=1)
    g = torch.Generator().manual_seed(2)
    q = torch.nn.functional.normalize(torch.randn(2048, 128, generator=g), dim=1)
    zd, qd = z.to(dev), q.to(dev)
    out = {"workload": "C2 ProtoRL reward: 1024 z (dim 128, unit rows, seed 1) vs 2048-entry "
                       "queue (seed 2), topk=3 (proto.py:115-118)",
           "l2": L2_POLICY, "algorithmic_flops": 2 * 1024 * 2048 * 128}
    out["stream"] = tm.run(lambda: xb.proto_knn_reward(zd, qd, 3))
    out["exact_kernel_stream"] = tm.run(lambda: xb.proto_knn_reward(zd, qd, 3, mode="exact"))
    out["us"] = out["stream"]["us_median"]
    if cpu:
        out["cpu_reference_us"] = time_cpu(lambda: proto_reward(z, q, 3))
        out["cpu_cores"] = torch.get_num_threads()
    return out


def _synthetic_buffer(dev, n_eps, L, obs_shape, obs_dtype, seed):
    """A full buffer laid straight into the device columns (headless)."""
    import exorl_b200 as xb
    rows = n_eps * (L + 1)
    buf = xb.DeviceReplayBuffer(dev, n_eps * L, 3, 0.99, capacity_rows=rows)
    g = torch.Generator(device=dev).manual_seed(seed)
    if obs_dtype == torch.uint8:
        obs = torch.empty((rows, *obs_shape), dtype=torch.uint8, device=dev)
        chunk = 20000
        for r0 in range(0, rows, chunk):
            r1 = min(r0 + chunk, rows)
            obs[r0:r1] = torch.randint(0, 256, (r1 - r0, *obs_shape), generator=g,
                                       dtype=torch.uint8, device=dev)
    else:
        obs = torch.randn((rows, *obs_shape), generator=g, device=dev)
    buf.cols["observation"] = obs
    buf.cols["action"] = torch.rand((rows, 6), generator=g, device=dev) * 2 - 1
    buf.cols["reward"] = torch.rand((rows, 1), generator=g, device=dev)
    buf.cols["discount"] = torch.ones((rows, 1), device=dev)
    buf.episodes = [{"key": f"episode_{e:06d}_{L}.npz", "start": e * (L + 1), "len": L}
                    for e in range(n_eps)]
    return buf


def _cpu_sample_us(obs_shape, obs_dtype, batch, L=200):
    """The reference's per-sample arithmetic (ReplayBuffer._sample, utils/replay_buffer.py:214-235,
    oracle port) for one batch on one host core — its DataLoader workers each run this loop."""
    from oracle import nstep_transition
    rng = np.random.default_rng(3)
    eps = []
    for _ in range(4):
        obs = (rng.integers(0, 256, (L + 1, *obs_shape), dtype=np.uint8) if obs_dtype == np.uint8
               else rng.standard_normal((L + 1, *obs_shape)).astype(np.float32))
        eps.append({"observation": obs,
                    "action": rng.uniform(-1, 1, (L + 1, 6)).astype(np.float32),
                    "reward": rng.uniform(0, 1, (L + 1, 1)).astype(np.float32),
                    "discount": np.ones((L + 1, 1), np.float32)})

    def one_batch():
        out = [nstep_transition(eps[s % 4], 1 + (s * 37) % (L - 3), 3, 0.99) for s in range(batch)]
        return [np.stack([o[j] for o in out]) for j in range(5)]     # the collate step
    return time_cpu(one_batch, iters=3, warm=1)


def c3(tm, cpu=True):
    import exorl_b200 as xb
    dev = tm.dev
    buf = _synthetic_buffer(dev, 1000, 1000, (24,), torch.float32, 3)
    call = [0]

    def sample():
        call[0] += 1
        return buf.sample_philox(1024, 7, call[0])
    out = {"workload": "C3 APT walker-state update: ReplayBuffer sample B=1024, nstep=3, "
                       "gamma 0.99 from a 1M-transition buffer (1000 episodes x 1001 rows, obs 24 "
                       "f32, action 6 f32), then ICM trunk 24->512 and PBE k=12",
           "l2": L2_POLICY, "algorithmic_bytes_per_batch": 464 * 1024,
           "note": "475 KB per batch: latency-bound, reported as us"}
    out["sample"] = tm.run(sample)
    trunk = torch.nn.Sequential(torch.nn.Linear(24, 512), torch.nn.LayerNorm(512),
                                torch.nn.Tanh()).to(dev)
    pbe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)

    def update_path():
        call[0] += 1
        batch = buf.sample_philox(1024, 7, call[0])
        return pbe(xb.icm_get_rep(trunk, batch[0]))
    out["sample_trunk_pbe"] = tm.run(update_path)
    out["sample_trunk_pbe_caller_graph"] = _graph_or_none(tm, update_path)
    out["us"] = out["sample"]["us_median"]
    if cpu:
        out["cpu_reference_us"] = _cpu_sample_us((24,), np.float32, 1024)
        out["cpu_cores"] = 1
    return out


def c4(tm, n_trans=1_000_000, cpu=True):
    dev = tm.dev
    obs_bytes = 9 * 84 * 84
    L = 1000
    buf = _synthetic_buffer(dev, n_trans // L, L, (9, 84, 84), torch.uint8, 4)
    peak, peak_src = hbm_peak()
    out = {"workload": f"C4 APT pixel update: ReplayBuffer sample nstep=3 from {n_trans} "
                       f"transitions of uint8 (9,84,84) = {buf.cols['observation'].numel() / 1e9:.1f}"
                       f" GB resident, random bytes seed 4; fresh Philox rows every iteration",
           "l2": L2_POLICY}
    for batch in (1024, 256):
        call = [0]

        def sample():
            call[0] += 1
            return buf.sample_philox(batch, 11, call[0])
        r = tm.run(sample)
        # per transition read 2*obs + act + nstep*8, write 2*obs + act + 8
        alg = (2 * obs_bytes + 24 + 3 * 8 + 2 * obs_bytes + 24 + 8) * batch
        gbs = alg / (r["us_median"] * 1e-6) / 1e9
        r["roofline"] = {"bound": "hbm", "kernel": "replay_gather_kernel", "achieved": gbs,
                         "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": alg,
                         "traffic": None}
        out[f"B{batch}"] = r
    out["us"] = out["B1024"]["us_median"]
    out["roofline"] = out["B1024"]["roofline"]
    # correctness spot check on a fresh batch: gathered frames == direct index of the column
    rows = buf.draw_rows_philox(64, 11, 999)
    got = buf.gather(rows)
    col = buf.cols["observation"]
    assert torch.equal(got[0], col[rows - 1]) and torch.equal(got[4], col[rows + 2])
    if cpu:
        out["cpu_reference_us"] = _cpu_sample_us((9, 84, 84), np.uint8, 1024)
        out["cpu_cores"] = 1
    del buf
    torch.cuda.empty_cache()
    return out


def tensor_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"], \
            "measured (MEASURED_PEAKS.json bf16_tflops burst) / 2: tf32 runs at half the bf16 rate"
    except Exception:
        return 2250.0, "fallback (B200_PROFILING.md nominal bf16) / 2"


def c4b(tm, cpu=True):
    """C4's rep producer: the 39 200-wide pixel-encoder output through the ICM trunk
    (icm_apt.py:17-18) into PBE — the wide Linear as a 3xTF32 tcgen05 GEMM (csrc/rep_gemm.cuh)."""
    import exorl_b200 as xb
    from exorl_b200 import rep as xrep
    dev = tm.dev
    g = torch.Generator().manual_seed(6)
    in_dim, dim, n = 39200, 512, 1024
    obs = torch.relu(torch.randn(n, in_dim, generator=g)).to(dev)
    w = (torch.randn(dim, in_dim, generator=g) / in_dim ** 0.5).to(dev)
    b = (0.1 * torch.randn(dim, generator=g)).to(dev)
    lnw, lnb = torch.ones(dim, device=dev), torch.zeros(dim, device=dev)
    peak, peak_src = tensor_peak()
    out = {"workload": f"C4b pixel APT rep producer: ICM trunk Linear({in_dim} -> {dim}) + "
                       f"LayerNorm + Tanh + kNN operand on a {n}-row batch of encoder features "
                       "(relu(randn) seed 6), fp32 parity (rep rtol 1e-5)",
           "l2": L2_POLICY, "algorithmic_flops": 2.0 * n * in_dim * dim}
    r = tm.run(lambda: xb.icm_get_rep((w, b, lnw, lnb), obs))
    alg_tf = out["algorithmic_flops"] / (r["us_median"] * 1e-6) / 1e12
    r["roofline"] = {"bound": "tensor", "kernel": "trunk_gemm_kernel (3xTF32 tcgen05)",
                     "achieved": 3 * alg_tf, "peak": peak / 2, "unit": "TFLOP/s",
                     "frac": 3 * alg_tf / (peak / 2), "peak_source": peak_src,
                     "algorithmic_tflops": alg_tf,
                     "note": "achieved = EXECUTED tf32 flops (3 MMAs per product term: hi*hi, "
                             "hi*lo, lo*hi) over the whole call (weight split + GEMM + LayerNorm "
                             "pass); algorithmic_tflops counts 2*n*K*D once", "traffic": None}
    out["tcgen05_3xtf32"] = r
    out["torch_fp32_addmm_then_ln"] = tm.run(
        lambda: xrep.ln_tanh_rep(torch.addmm(b, obs, w.t()), lnw, lnb))
    pbe = xb.PBE(xb.RMS(dev), 0.0, 12, True, True, dev)
    out["trunk_then_pbe"] = tm.run(lambda: pbe(xb.icm_get_rep((w, b, lnw, lnb), obs)))
    ref = torch.tanh(torch.nn.functional.layer_norm(obs.double() @ w.double().t() + b.double(),
                                                    (dim,), lnw.double(), lnb.double(), 1e-5))
    got = xb.icm_get_rep((w, b, lnw, lnb), obs).rep.double()
    out["max_abs_err_vs_fp64"] = float((got - ref).abs().max())
    out["us"] = r["us_median"]
    out["roofline"] = r["roofline"]
    if cpu:
        oc, wc, bc = obs.cpu(), w.cpu(), b.cpu()
        t0 = time.perf_counter()
        torch.tanh(torch.nn.functional.layer_norm(torch.addmm(bc, oc, wc.t()), (dim,)))
        out["cpu_reference_us"] = (time.perf_counter() - t0) * 1e6
        out["cpu_cores"] = torch.get_num_threads()
    return out


def c5r(tm, cpu=True):
    """Range-proof screen A/B at dataset scale: the same reps bounded (tanh) and times 1e5 +
    offset (outside the fp16 range: before the affine operands every row fell to the CUDA-core
    exact kernel)."""
    import exorl_b200 as xb
    from exorl_b200 import knn as xknn
    dev = tm.dev
    g = torch.Generator().manual_seed(8)
    n, dim, k = 131072, 512, 12
    x = torch.tanh(torch.randn(n, dim, generator=g)).to(dev)
    y = x * 1e5 + 3e6
    out = {"workload": f"C5r range-proof screen: {n} x {n} kNN (rep {dim}, k={k}) on tanh(randn) "
                       "reps (seed 8) and on the same reps * 1e5 + 3e6",
           "l2": "inputs (268 MB fp32 + 134 MB fp16) larger than the 126 MB L2"}
    small = Timer(dev, tm.sampler_factory, min_seconds=1.0, min_iters=5, max_iters=40)
    small.flush = lambda: None
    out["bounded"] = small.run(lambda: xb.knn_topk(x, x, k, mode="screened"), warm=2)
    out["times_1e5_plus_offset"] = small.run(lambda: xb.knn_topk(y, y, k, mode="screened"), warm=2)
    st = xknn.screened_status(dev)
    out["undecided_rows_scaled"] = int(st[0])
    out["ratio"] = out["times_1e5_plus_offset"]["us_median"] / out["bounded"]["us_median"]
    d0, i0 = xb.knn_topk(x[:4096], x, k, mode="screened")
    d1, i1 = xb.knn_topk(y[:4096], y, k, mode="screened")
    out["same_neighbours"] = bool((i0 == i1).float().mean() > 0.999)
    out["us"] = out["times_1e5_plus_offset"]["us_median"]
    return out


def run_all(dev, sampler_factory=None, cases=("C1", "C2", "C3", "C4", "C4b", "C5r"), min_seconds=2.0,
            c4_transitions=1_000_000, cpu=True):
    tm = Timer(dev, sampler_factory, min_seconds=min_seconds)
    fns = {"C1": c1, "C2": c2, "C3": c3,
           "C4": lambda t, cpu=True: c4(t, c4_transitions, cpu=cpu), "C4b": c4b, "C5r": c5r}
    out = {}
    for name in cases:
        out[name] = fns[name](tm, cpu=cpu)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="C1,C2,C3,C4,C4b,C5r")
    ap.add_argument("--c4-transitions", type=int, default=1_000_000)
    ap.add_argument("--min-seconds", type=float, default=2.0)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    sys.path.insert(0, ROOT)
    from bench import ClockSampler
    dev = torch.device("cuda:0")
    res = run_all(dev, lambda: ClockSampler(0), tuple(a.cases.split(",")), a.min_seconds,
                  a.c4_transitions, cpu=not a.no_cpu)
    for name, r in res.items():
        print(json.dumps({"case": name, **r}), flush=True)
