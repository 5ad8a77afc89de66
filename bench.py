"""Benchmark of the north-star hot path: APT kNN reward relabel of an N-transition buffer
against itself (BASELINE.json metric: kNN reward queries/sec, 1M x 1M, rep 512, k=12).

  python bench.py --gpus 1 --steps K --warmup W            (our arm, one process per GPU)
  torchrun ... bench.py --gpus N ...                       (bank sharded over N ranks, strong scaling)
  python bench.py --impl reference ...                     (reference algorithm on the host cores)

One "step" = one full relabel pass: fp32 reps resident in HBM -> (N,1) rewards (prep fp16 operand,
tcgen05 screen, exact fp32 rerank, [exchange + merge], RMS update, epilogue).  Prints ONE JSON line.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import time
import types

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "kNN reward queries/sec (1M x 1M, rep 512, k=12)"
UNIT = "queries/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--k", type=int, default=12)
    ap.add_argument("--exchange", default="auto",
                    choices=["auto", "lists", "all_to_all", "all_gather", "queries"])
    ap.add_argument("--cpu-queries", type=int, default=128,
                    help="queries of the bounded CPU-baseline sample (vs the full bank)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the C1-C4 block (BASELINE configs 1-4) of the 1-GPU line")
    ap.add_argument("--no-full-ab", action="store_true",
                    help="skip the full-screen A/B steps (symmetric path switched off)")
    ap.add_argument("--no-exchange-ab", action="store_true",
                    help="skip the one-step timings of the bank-sharded exchanges (N > 1)")
    ap.add_argument("--balance", action="store_true",
                    help="rate-proportional query blocks per rank instead of equal ones (A/B)")
    ap.add_argument("--observe-balance", action="store_true",
                    help="equal query blocks, but collect and report the per-rank block times")
    ap.add_argument("--stage-times", action="store_true",
                    help="after the timed runs, one extra step with per-stage CUDA events (stderr)")
    return ap.parse_args()


def workload(a):
    return (f"C5 APT relabel: {a.n} x {a.n} kNN reward, rep {a.dim}, k={a.k}, knn_avg+knn_rms, "
            f"clip 0, reps = tanh(randn) seed 5")


def screen_traffic(nq, nb, dim, name="screen_traffic.json"):
    """dram__bytes_read.sum + dram__bytes_write.sum of one screening launch from the committed
    `ncu --set full` capture (profiles/screen_traffic.json; sym_screen_traffic.json for the
    symmetric screen) when it is of this launch shape."""
    try:
        t = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles",
                                        name)))
        if (t["n_queries"], t["n_bank"], t["dim"]) == (nq, nb, dim):
            return {"traffic": t["traffic_bytes_per_launch"], "traffic_unit": "bytes",
                    "traffic_source": t["source"]}
    except Exception:
        pass
    return {"traffic": None}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["bf16_tflops_sustained"], p["bf16_tflops"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 1400.0, 1590.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------
# reference algorithm on the host cores (oracle port; /root/reference is absent on the GPU box)
# ------------------------------------------------------------------------------------------
def cpu_reference_pass(rep_cpu, q, k):
    """q queries against the full bank with the reference formula (tiled), + RMS + epilogue."""
    from oracle import RmsState, knn_smallest, pbe_reward
    src = rep_cpu[:q]
    t0 = time.perf_counter()
    knn = knn_smallest(src, rep_cpu, k, q_tile=max(1, min(q, 16)), t_tile=8192)
    pbe_reward(src, rep_cpu, RmsState(), k, True, True, 0.0, knn=knn)
    return time.perf_counter() - t0


def make_rep_cpu(a):
    g = torch.Generator().manual_seed(5)
    return torch.tanh(torch.randn(a.n, a.dim, generator=g))


def run_reference(a, rank):
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    rep = make_rep_cpu(a)
    q = max(1, a.cpu_queries // 4)
    for _ in range(a.warmup):
        cpu_reference_pass(rep, 1, a.k)
    times = [cpu_reference_pass(rep, q, a.k) for _ in range(a.steps)]
    dt = sum(times)
    value = q * a.steps / dt
    sample = (f"{q} queries per step against the full {a.n}-row bank (reference formula "
              f"utils/utils.py:293-319, tiled), scaled per query; the per-query cost is flat in "
              f"the slice size (every query scans the whole bank), so SURVEY 8d's 4096-query "
              f"slice (minutes per step) would measure the same rate")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": workload(a), "n": a.n, "dim": a.dim, "k": a.k},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(),
                         "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        # median over the samples taken under load (above half of the top power seen)
        if sm:
            hi = max(pw) if pw else 0
            load = [s for s, p in zip(sm, pw) if p >= 0.5 * hi] or sm
            return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx),
                    "reasons": sorted(reasons), "samples": len(sm),
                    "power_w_max": max(pw) if pw else None}
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def run_ours(a, rank, world, local_rank):
    import torch.distributed as dist
    import exorl_b200 as xb
    from exorl_b200 import _lib

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    # stdout carries exactly ONE line (the contract): libraries that write to fd 1 (NCCL prints
    # its version banner there at some NCCL_DEBUG levels) are sent to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        # NCCL prints its version banner to STDOUT at NCCL_DEBUG=VERSION; keep stdout to the one
        # JSON line of the contract
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    args = types.SimpleNamespace(knn_k=a.k, knn_avg=True, knn_rms=True, knn_clip=0.0)

    # synthetic reps: generated on the device (seed 5), mirrored into pinned host memory for the
    # end-to-end leg and the CPU baseline
    g = torch.Generator(device=dev).manual_seed(5)
    rep = torch.tanh(torch.randn(a.n, a.dim, generator=g, device=dev))
    rep_host = torch.empty((a.n, a.dim), dtype=torch.float32, pin_memory=True)
    rep_host.copy_(rep)
    reward_host = torch.empty((a.n, 1), dtype=torch.float32, pin_memory=True)
    torch.cuda.synchronize()

    cpu_baseline = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        cpu_reference_pass(rep_host, 1, a.k)
        dt = cpu_reference_pass(rep_host, a.cpu_queries, a.k)
        cpu_baseline = {"value": a.cpu_queries / dt, "unit": UNIT,
                        "cores": torch.get_num_threads(), "kind": "port",
                        "sample": (f"{a.cpu_queries} queries against the full {a.n}-row bank with "
                                   f"the reference formula (tiled torch CPU), {dt:.1f} s, scaled "
                                   f"per query; SURVEY 8d's 4096-query slice would take "
                                   f"{4096 / a.cpu_queries * dt / 60:.0f} min on these cores, the "
                                   f"per-query cost is flat in the slice size (every query "
                                   f"scans the whole bank), so the bounded sample measures the "
                                   f"same rate")}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)
        return ms

    rms = xb.RMS(dev)
    balance = "observe" if a.observe_balance else a.balance

    def step():
        return xb.relabel_apt_reward(rep, args, rms=rms, exchange=a.exchange, mode="screened",
                                     balance=balance)

    def step_e2e():
        # host buffers in, host buffer out: every rank copies ITS rows from pinned host memory,
        # prepares them, the fp16 operand is all-gathered over NVLink and the screen starts while
        # the fp32 rows follow behind it (relabel_apt_reward_from_host)
        r = xb.relabel_apt_reward_from_host(rep_host, args, rms=rms, rep_out=rep,
                                            balance=balance)
        reward_host.copy_(r, non_blocking=True)

    for _ in range(a.warmup):
        step()
    barrier()
    sampler = ClockSampler(dev.index) if rank == 0 else None
    _lib.check(lib.exb_profile_begin(1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    kms = (ctypes.c_float * 64)()
    kn = ctypes.c_int(0)
    _lib.check(lib.exb_profile_collect_sync(kms, 64, ctypes.byref(kn)))
    from exorl_b200 import knn as xknn
    if a.stage_times:          # every rank's screen launches of the timed steps, in launch order
        print(f"rank {rank} timed-loop screen kernels (ms): "
              + json.dumps([round(kms[i], 3) for i in range(kn.value)])
              + f" rows_per_rank={xknn.sym_last.get('rows_per_rank')}", file=sys.stderr, flush=True)
    _lib.check(lib.exb_profile_begin(0))
    clocks = sampler.stop() if sampler else None
    from exorl_b200 import knn as xknn
    und, invalid, pairs_lo, pairs_hi = xknn.screened_status(dev)

    # end to end through the public API with host buffers
    step_e2e()
    barrier()
    e0.record()
    for _ in range(a.steps):
        step_e2e()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))

    # the FULL screen on the same inputs (symmetric path switched off): one warm + two timed steps,
    # so that the line itself shows what halving the tile pairs buys and that the full kernel still
    # sits at the tensor roofline
    full_ab = None
    if not a.no_full_ab and xknn.sym_last.get("n") == a.n:
        prev_sym = xknn.SYM_ENABLED
        xknn.SYM_ENABLED = False
        try:
            step()
            barrier()
            _lib.check(lib.exb_profile_begin(1))
            e0.record()
            step()
            step()
            e1.record()
            barrier()
            ms_full = max_over_ranks(e0.elapsed_time(e1)) / 2
            fk = (ctypes.c_float * 64)()
            fn_ = ctypes.c_int(0)
            _lib.check(lib.exb_profile_collect_sync(fk, 64, ctypes.byref(fn_)))
            _lib.check(lib.exb_profile_begin(0))
            k_ms = sum(fk[i] for i in range(fn_.value)) / max(fn_.value, 1)
            full_ab = {"ms_per_step": ms_full, "queries_per_s": a.n / (ms_full * 1e-3),
                       "kernel": "knn_screen_kernel (all tile pairs)", "kernel_ms": k_ms,
                       "kernel_tflops": 2.0 * a.n * (a.n // world) * a.dim / (k_ms * 1e-3) / 1e12}
        finally:
            xknn.SYM_ENABLED = prev_sym

    # N > 1: the bank-sharded partitions of north_star (per-shard top-k -> NCCL exchange -> merge),
    # one warm + one timed step each, beside the partition the headline uses
    exchange_ab = None
    if world > 1 and not a.no_exchange_ab:
        exchange_ab = {}
        for ex in ("lists", "all_to_all", "all_gather"):
            try:
                xb.relabel_apt_reward(rep, args, rms=rms, exchange=ex, mode="screened")
                barrier()
                e0.record()
                xb.relabel_apt_reward(rep, args, rms=rms, exchange=ex, mode="screened")
                e1.record()
                barrier()
                exchange_ab[ex] = max_over_ranks(e0.elapsed_time(e1))
            except Exception as e:          # keep the headline line even if one form fails
                exchange_ab[ex] = f"failed: {type(e).__name__}: {e}"[:200]
                barrier()

    if a.stage_times:
        from exorl_b200.relabel import StageTimer
        tmr = StageTimer()
        _lib.check(lib.exb_profile_begin(1))       # this rank's screen-kernel times beside them
        xb.relabel_apt_reward(rep, args, rms=rms, exchange=a.exchange, mode="screened", timer=tmr)
        st = tmr.summary()
        sk = (ctypes.c_float * 64)()
        sn = ctypes.c_int(0)
        _lib.check(lib.exb_profile_collect_sync(sk, 64, ctypes.byref(sn)))
        _lib.check(lib.exb_profile_begin(0))
        st["screen_kernels_ms"] = [round(sk[i], 3) for i in range(sn.value)]
        tmr2 = StageTimer()
        xb.relabel_apt_reward_from_host(rep_host, args, rms=rms, rep_out=rep, balance=balance,
                                        timer=tmr2)
        st2 = tmr2.summary()
        print(f"rank {rank} stage times (ms): " + json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()})
              + " | from_host: " + json.dumps({k: round(v, 3) for k, v in st2.items()}),
              file=sys.stderr, flush=True)

    secondary = None
    if world == 1 and not a.no_secondary:
        # BASELINE configs 1-4 (per-update latencies, replay gather GB/s), each with its own
        # clock record; the C5 tensors are released first (C4 holds a 63.6 GB pixel buffer)
        del rep
        torch.cuda.empty_cache()
        from benchmarks import secondary as sec
        torch.set_num_threads(os.cpu_count() or 1)
        secondary = sec.run_all(dev, lambda: ClockSampler(dev.index),
                                cpu=not a.no_cpu_baseline)

    query_blocks = None
    if world > 1 and a.exchange in ("auto", "queries"):
        from exorl_b200 import relabel as xrel
        bal = xrel.query_balance(None, a.n, world, a.k)
        query_blocks = {"balance": balance,
                        "rows_per_rank": [hi - lo for lo, hi in bal.bounds]}
        if a.observe_balance or a.stage_times:
            query_blocks["log_rows_ms"] = bal.log[-12:]
    if rank == 0:
        sustained, burst, src = peaks()
        nb_local = a.n // world
        sym = (xknn.sym_last.get("n") == a.n and xknn.sym_last.get("world", 1) == world
               and a.exchange in ("auto", "queries")
               and "fell_back_to_full_screen" not in xknn.sym_last and kn.value == 2 * a.steps)
        if sym and query_blocks is not None and xknn.sym_last.get("rows_per_rank"):
            # symmetric path: tile blocks re-cut every step from the ranks' screen rates of the step before
            query_blocks.update(balance="screen rates of the previous step (sym_balanced_blocks)",
                                rows_per_rank=xknn.sym_last["rows_per_rank"],
                                screen_ns_per_tile_pair=xknn.sym_last["screen_ns_per_tile_pair"])
        flops = 2.0 * a.n * nb_local * a.dim          # algorithmic: every (query, bank row) pair
        sym_info = None
        if sym:
            # symmetric screen: per step one sample pre-pass launch (n x S, full screen kernel) and
            # one launch of knn_screen_kernel<2,4> over the tile pairs I <= J.  The roofline object
            # is the latter's, on the flops it EXECUTES; the algorithmic figure stands beside it.
            pre_ms = sum(kms[i] for i in range(0, kn.value, 2)) / a.steps
            kernel_ms = sum(kms[i] for i in range(1, kn.value, 2)) / a.steps
            t = (a.n + 255) // 256
            executed = 2.0 * 256 * 256 * a.dim * (t * (t + 1) / 2) / world   # this rank's tile pairs
            if world > 1 and xknn.sym_last.get("rows_per_rank"):
                # rate-balanced blocks: rank 0's own tile products under the LAST step's blocks
                # (they move by ~1 % from step to step with the measured rates)
                from exorl_b200.relabel import sym_span
                tb_ = [0]
                for r_ in xknn.sym_last["rows_per_rank"]:
                    tb_.append(tb_[-1] + (r_ + 255) // 256)
                lo_, hi_, end_, mid_, extra_ = sym_span(tb_, world, 0)
                executed = 2.0 * 256 * 256 * a.dim * sum(
                    end_ - u + (extra_ if u >= mid_ else 0) for u in range(lo_, hi_))
            flops = flops                                  # (algorithmic share of a rank: N * N/G)
            s_rows = xknn.sym_last["sample_rows"]
            achieved = executed / (kernel_ms * 1e-3) / 1e12
            und, invalid = xknn.sym_last["undecided"], xknn.sym_last["status"]
            sym_info = {
                "kernel": "knn_screen_kernel<2,4> (symmetric: tile pairs I <= J)",
                "executed_flops_per_launch": executed,
                "executed_over_algorithmic": executed / flops,
                "sample_prepass": {"rows": s_rows, "kernel_ms": pre_ms,
                                   "flops": 2.0 * (a.n / world) * s_rows * a.dim},
                "algorithmic_tflops_equivalent_of_both_launches":
                    flops / ((kernel_ms + pre_ms) * 1e-3) / 1e12,
                "note": "frac = EXECUTED flops of the symmetric kernel / its time / peak; the "
                        "screen does all the work of the full one with half the MMAs because "
                        "a^_i . a^_j serves the pairs (i, j) and (j, i) (DESIGN.md K1s)"}
        else:
            kernel_ms = sum(kms[i] for i in range(kn.value)) / max(kn.value, 1)
            achieved = flops / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else 0.0
        # kernels per step: colstats (one per row set), colstats-reduce, affine, prep (one per row
        # set), screen, rerank, exact-fallback, fallback-merge, [merge, moments, rms-apply] or
        # [moments], finish
        # symmetric path (1 GPU): sample gather, [colstats x2, reduce, affine, prep x2, screen,
        # rerank, exact, merge] of the pre-pass, [colstats, reduce, affine, prep, threshold, screen,
        # scatter, rerank, exact, merge], moments, finish
        peer = sym and xknn.sym_last.get("exchange") == "peer_memory"   # + the routing-table kernel
        launches_per_step = ((23 if sym else 10) if world == 1 else
                             ((26 if peer else 25) if sym else 14))
        out = {
            "metric": METRIC, "value": a.n * a.steps / (ms * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f16 tensor-core screen (f32 accumulate) + f32 exact recompute",
            "launches_timed_note": "roofline.launches_timed counts the bracketed screen launches",
            "data": "synthetic",
            "config": {"workload": workload(a), "n": a.n, "dim": a.dim, "k": a.k,
                       "parallelism": (f"{world} GPU(s), exchange={a.exchange}" + (
                           "" if world == 1 else
                           " (symmetric screen: every unordered tile pair on one rank, transposed "
                           "hits appended to the owner rank's buffer over NVLink peer memory "
                           "inside the screen kernel)" if peer else
                           " (symmetric screen: every unordered tile pair on one rank, transposed "
                           "hits exchanged all-to-all)" if sym else
                           " (query blocks over the ranks, whole bank on every rank)"
                           if a.exchange in ("auto", "queries") else " (bank sharded)")),
                       "l2": "inputs (2 GB fp32 + 1 GB fp16 bank) larger than the 126 MB L2; "
                             "no flush needed",
                       "undecided_rows_last_step": und, "invalid": invalid,
                       **({"query_blocks": query_blocks} if query_blocks else {})},
            "clocks": clocks,
            "e2e": {"value": a.n * a.steps / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": rep_host.numel() * 4,   # whole job: each rank its shard
                    "d2h_bytes_per_step": reward_host.numel() * 4,
                    "ms_per_step": ms_e2e / a.steps},
            "gpu_launches": launches_per_step * a.steps,
            "roofline": {"bound": "tensor",
                         "kernel": sym_info["kernel"] if sym else "knn_screen_kernel",
                         "achieved": achieved,
                         "peak": sustained, "unit": "TFLOP/s",
                         "frac": achieved / sustained if sustained else None,
                         "frac_of_burst_peak": achieved / burst if burst else None,
                         "peak_source": src + ", sustained bf16 (kernel timed inside a long step)",
                         "kernel_ms": kernel_ms, "launches_timed": kn.value,
                         "algorithmic_flops_per_launch": flops,
                         **({k: v for k, v in sym_info.items() if k != "kernel"} if sym else {}),
                         **(screen_traffic(a.n, nb_local, a.dim, "sym_screen_traffic.json") if sym
                            else screen_traffic(a.n, nb_local, a.dim))},
            "cpu_baseline": cpu_baseline,
        }
        if full_ab is not None:
            sustained_, _, _ = peaks()
            full_ab["frac_of_sustained_peak"] = full_ab["kernel_tflops"] / sustained_
            out["full_screen_ab"] = full_ab
        if exchange_ab is not None:
            out["exchange_ab"] = {"unit": "ms per step (one timed step each, max over ranks)",
                                  "headline_exchange": a.exchange, **exchange_ab}
        if secondary is not None:
            out["secondary"] = secondary
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank)
        return
    if world != a.gpus and world == 1 and a.gpus > 1:
        print(f"bench.py: --gpus {a.gpus} needs torchrun with {a.gpus} ranks", file=sys.stderr)
        sys.exit(2)
    run_ours(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
