#!/usr/bin/env python
"""bench.py -- GP log-marginal-likelihood evaluations/s at n=4096 (BASELINE.json configs[1]) on
N B200s, through libgpb200.so's C ABI.  Prints ONE JSON line (rank 0).

  step      = one gpb_loglik_batch call: --batch fresh proposals (new hyper-parameters every step,
              K is rebuilt and refactored every time; nothing is cached) on the resident dataset.
  value     = evals/s from the device time of the step's kernels (CUDA events on the library's own
              launching stream: build + factor/solve + epilogue), max over ranks.
  e2e       = evals/s through the reference-facing call with HOST buffers: hyper-parameters go H2D
              and the (Q, U1, S, info) results come D2H inside the timed region, every step.
  roofline  = the factorisation kernel (k_chol_tasks): algorithmic FP64 flops / its average launch
              duration vs the measured DMMA issue peak of this pool's B200 (37.1 TFLOP/s,
              profiles/r01_dmma_shapes_microbench.log; MEASURED_PEAKS.json has no FP64 entry).
  N > 1     : proposals are sharded across ranks (independent, no data-path collective); each rank
              then all-gathers its log-likelihoods through the library's own NCCL communicator
              (gpb_comm_* / gpb_gather behind the C ABI; north_star: gather only). scaling=weak.
              torch.distributed only carries the barrier, the max-over-ranks of the clocks and the
              128-byte communicator id.
  extra     : the other BASELINE.json configs (C1 shipped example at n=300 and thinned to n=20, C3
              256 chains x n=1024, C4 n=32768 + 10^4 predictive points, C5 SE+SE+white n=8192 x 64
              chains per GPU), each with its own fraction of the FP64 DMMA peak; MH runs go through
              gpb_mh_run_ex (device Philox streams, ring flushed every 32 iterations, NCCL gather
              of the ring segments when N > 1).
  --impl reference : the CPU oracle (port of the Scala/Breeze path; the JVM is not in this image),
              same config, metric and step as the GPU arm: a step = the same batch of proposals,
              one single-threaded evaluation per host thread.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))

N_TRAIN = 4096
DIM = 2
SEED = 2
TRUE_TERMS = [("se", 3.0, 5.5), ("white", 0.5)]  # SimulatedGp.scala:30-33
FP64_PEAK_TFLOPS = 37.1                          # measured DMMA.8x8x4 issue peak, B200 @1965 MHz
METRIC = "gp_loglik_evals_per_sec_n4096"


def chol_flops(n):
    return n ** 3 / 3.0 + n ** 2 / 2.0 + n / 6.0 + float(n) ** 2  # dpotrf + one forward solve


def make_dataset():
    """n=4096, 2-D U(-10,10)^2, y = L z with the true kernel, seed 2."""
    import scipy.linalg as sl
    rng = np.random.default_rng(SEED)
    x = rng.uniform(-10, 10, size=(N_TRAIN, DIM))
    d2 = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)
    K = 3.0 * np.exp(-d2 / (5.5 * 5.5)) + (0.5 + 1e-3) * np.eye(N_TRAIN)
    y = sl.cholesky(K, lower=True) @ rng.standard_normal(N_TRAIN)
    return x, y


def proposals(rng, batch):
    """Fresh hyper-parameters around the truth (log-normal jitter 0.05, as the example proposal)."""
    out = []
    for _ in range(batch):
        z = rng.standard_normal(3) * 0.05
        out.append([("se", 3.0 * math.exp(z[0]), 5.5 * math.exp(z[1])), ("white", 0.5 * math.exp(z[2]))])
    return out


def gterms(terms):
    return [(0, t[1], t[2]) if t[0] == "se" else (1, t[1], 0.0) for t in terms]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def count_since(self, t):
        return sum(1 for ts, _ in self.rows if ts >= t)

    def stop(self, t0=None, t1=None, note="timed region"):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ts, r in self.rows:
            if (t0 is not None and ts < t0) or (t1 is not None and ts > t1):
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons), "window": note}


def cpu_evals(x, y, term_lists, threads):
    """Oracle (test infrastructure; here only as the timed CPU baseline): one independent n=4096
    evaluation per host thread, each single-threaded like the reference's F2J path."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as o
    o.lib()
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        res = list(ex.map(lambda t: o.loglik(x, y, t, nugget=1e-3), term_lists))
    dt = time.perf_counter() - t0
    assert all(r["info"] == 0 for r in res)
    return len(term_lists) / dt, dt, res


def workload_config(B, world):
    """The `config` object both arms print (the driver compares them)."""
    return {"workload": f"C2: n={N_TRAIN} 2-D SE+white log-marginal-likelihood, batch of {B} independent proposals per step per GPU",
            "batch": B, "n": N_TRAIN, "dim": DIM, "parallelism": f"proposal-sharded x{world}",
            "l2": "inputs larger than L2 (each proposal rebuilds and refactors its own 134 MB matrix)"}


def run_reference(args):
    """--impl reference: the reference's CPU path restated (oracle port; JVM unavailable on this
    box), same config / metric / unit / step as the GPU arm: a step = one batch of --batch fresh
    proposals, each evaluation single-threaded like the reference's F2J path, one per host thread.
    Rank 0 only.  A step takes seconds: the run is bounded to a few minutes and the line reports the
    steps it actually executed as `steps` (the metric is a rate)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    B = args.batch
    # all the host threads the box has: `lanes` steps run side by side, one single-threaded evaluation
    # per thread (a step is still one batch of B proposals; the rate does not depend on the grouping)
    lanes = max(1, cores // B)
    threads = min(cores, lanes * B)
    x, y = make_dataset()
    rng = np.random.default_rng(1234)
    t_start = time.perf_counter()
    warm_done = 0
    while warm_done < args.warmup:
        if warm_done >= 1 and time.perf_counter() - t_start > 25.0:
            break
        k = min(lanes, args.warmup - warm_done)
        cpu_evals(x, y, proposals(rng, k * B), threads); warm_done += k
    total_t, steps_done = 0.0, 0
    while steps_done < max(args.steps, 1):
        if steps_done >= 1 and time.perf_counter() - t_start > 150.0:
            break
        k = min(lanes, max(args.steps, 1) - steps_done)
        _, dt, _ = cpu_evals(x, y, proposals(rng, k * B), threads)
        total_t += dt; steps_done += k
    v = steps_done * B / total_t
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": steps_done, "warmup": warm_done, "steps_requested": args.steps, "warmup_requested": args.warmup,
        "ms_per_step": 1e3 * total_t / steps_done,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(B, args.gpus),
        "cpu_baseline": {"value": v, "unit": "evals/s", "cores": threads, "kind": "port",
                         "sample": f"{steps_done} steps x {B} full n={N_TRAIN} evaluations, 1 thread each, {lanes} step(s) side by side on {threads} of {cores} host cores (time-bounded; reference JVM path: unavailable on this box)"},
        "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def run_ours(args):
    import torch
    from gpb200 import _lib as G
    from gpb200 import chains as CH

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- gpb200 has no CPU fallback")
    torch.cuda.set_device(local)
    sampler = ClockSampler(local)
    sampler.start()                 # nvidia-smi needs seconds to come up: start it before the set-up
    ctx = G.Context(local)
    if world > 1:
        CH.init_comm(ctx, dist)     # the library's own NCCL communicator (gpb_comm_init_rank)
    x, y = make_dataset()
    ds = ctx.dataset(x, y)
    B = args.batch
    rng = np.random.default_rng(1000 + rank)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        props = [gterms(p) for p in proposals(rng, B)]
        rs = ds.loglik_batch(props, nugget=1e-3)       # H2D params, kernels, D2H results
        t = ctx.last_timing()
        if dist is not None:                            # gather the log-likelihoods (north_star): gpb_gather
            ctx.gather([G.ll_value(r, 0) for r in rs])
        return rs, t

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    l0 = ctx.launch_count()
    dev_ms, chol_ms = 0.0, 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        rs, t = step()
        dev_ms += t[0] + t[1] + t[2]
        chol_ms += t[1]
        assert all(r.info == 0 for r in rs)
    barrier()
    t1 = time.perf_counter()
    wall = t1 - t0
    launches = ctx.launch_count() - l0
    note = "timed region"
    # a short timed region can fall between two 100 ms samples: then every rank keeps the same load
    # running (untimed, same number of extra steps on all ranks: step() holds a collective) for
    # ~0.6 s so that the sampler sees it, and the record says so
    flag = torch.tensor([1.0 if sampler.count_since(t0) < 3 else 0.0, wall], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    if float(flag[0]) > 0:
        note = "same load continued after the timed region (timed region shorter than the sampling period)"
        for _ in range(min(2000, int(0.6 / max(float(flag[1]) / args.steps, 1e-4)) + 1)):
            step()
        barrier()
        t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1, note)

    # max over ranks of both clocks
    tt = torch.tensor([dev_ms, wall, chol_ms], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dev_ms, wall, chol_ms = (float(v) for v in tt.cpu())
    evals = args.steps * B * world
    value = evals / (dev_ms * 1e-3)
    e2e = evals / wall
    achieved = args.steps * B * chol_flops(N_TRAIN) / (chol_ms * 1e-3) * 1e-12

    extra = {}
    if not args.no_extras:
        if world == 1:
            extra = extras_single_gpu(ctx, ds, G, CH)
        else:
            extra = extras_multi_gpu(ctx, G, CH, dist, local)   # collective: every rank takes part
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            v, dt, res = cpu_evals(x, y, proposals(np.random.default_rng(7), cores), cores)
            cpu = {"value": v, "unit": "evals/s", "cores": cores, "kind": "port",
                   "sample": f"{cores} full n={N_TRAIN} evaluations, one per thread, {dt:.1f} s (reference JVM path: unavailable on this box)"}
            lap = extra.get("cpu_lapack_b2_n4096", {}) if isinstance(extra, dict) else {}
            if "evals_per_sec_factor_only_path" in lap:
                # the strong CPU baseline beside the port: OpenBLAS dpotrf + dtrsv on all cores
                cpu["lapack_b2_evals_per_sec"] = lap["evals_per_sec_factor_only_path"]
                extra["speedup_vs_cpu"] = {"e2e_over_port": e2e / v, "e2e_over_lapack_b2": e2e / lap["evals_per_sec_factor_only_path"],
                                           "note": "port = line-by-line restatement of the reference's single-threaded F2J path, one evaluation per core; B2 = scipy/OpenBLAS LAPACK on all cores, covariance build excluded"}
        traffic = None
        tp = os.path.join(ROOT, "profiles", "chol_traffic_bytes_per_launch.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(str(B))
        line = {
            "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(B, world),
            "e2e": {"value": e2e, "unit": "evals/s", "h2d_bytes_per_step": B * 192,
                    "d2h_bytes_per_step": B * 40 + 8},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "k_chol_tasks", "achieved": achieved,
                         "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": achieved / FP64_PEAK_TFLOPS,
                         "traffic": traffic,
                         "peak_source": "measured FP64 DMMA issue peak on this pool (profiles/r01_dmma_shapes_microbench.log); MEASURED_PEAKS.json has no FP64 entry; cuBLAS DGEMM reaches 35.4"},
            "cpu_baseline": cpu,
            "extra": extra,
        }
        emit(line)
    ds.close()
    if dist is not None:
        dist.barrier()
        ctx.comm_destroy()
        dist.destroy_process_group()


IG3 = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]                       # SimulatedGp.scala:132-134
IG5 = [[3.0, 4.0], [3.0, 6.0], [3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
C5_TERMS = [("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)]


def synth_dataset(ctx, n, dim, lo, hi, seed, terms):
    """x ~ U(lo, hi)^dim, y = L z with K = k(true terms) + 1e-3 I.  L z is formed on the
    GPU (gpb_draw: build + factor + L z), so that n = 8192 / 32768 need no host Cholesky."""
    r = np.random.default_rng(seed)
    x = r.uniform(lo, hi, size=(n, dim)) if dim > 1 else np.sort(r.uniform(lo, hi, n))
    y, info = ctx.draw(x, gterms(terms), r.standard_normal(n), None, nugget=1e-3)
    assert info == 0
    return x, y


def prior_draws(terms, ig, count, seed):
    """Initial states = draws from the inverse-gamma priors: IG(a, b) = b / Gamma(a, 1)."""
    r = np.random.default_rng(seed)
    out = []
    for _ in range(count):
        tl, p = [], 0
        for t in terms:
            if t[0] == "se":
                tl.append((0, ig[p][1] / r.gamma(ig[p][0]), ig[p + 1][1] / r.gamma(ig[p + 1][0]))); p += 2
            else:
                tl.append((1, ig[p][1] / r.gamma(ig[p][0]), 0.0)); p += 1
        out.append(tl)
    return out


def chol_frac(ctx, ds, term_lists):
    """Factor+solve kernel of one batch at this shape: ms, algorithmic TFLOP/s, fraction of the DMMA peak."""
    for _ in range(2):
        ds.loglik_batch(term_lists)
    ts = []
    for _ in range(5):
        ds.loglik_batch(term_lists); ts.append(ctx.last_timing())
    t = np.median(np.array(ts), axis=0)
    tf = len(term_lists) * chol_flops(ds.n) / (t[1] * 1e-3) * 1e-12
    return {"build_ms": float(t[0]), "chol_ms": float(t[1]), "chol_tflops": tf, "frac_of_dmma_peak": tf / FP64_PEAK_TFLOPS}


def mh_bench(ctx, G, CH, n, chains_total, iters, warm, terms, ig, seed, lo, hi, dist=None, local=0):
    """MCMC iterations/s for `chains_total` chains block-partitioned over the ranks, through
    gpb_mh_run_ex: device Philox streams (keyed by the global chain index), ring flushed every 32
    iterations and -- for N > 1 -- all-gathered device to device by the library's NCCL communicator
    so that every rank ends up with every chain.  Timed end to end through the ABI call with host
    buffers (initial states H2D, all chains' outputs D2H), barrier on both sides, max over ranks; the
    device time of the loop (CUDA events inside the library) is reported too."""
    import torch
    world = dist.get_world_size() if dist is not None else 1
    x, y = synth_dataset(ctx, n, 2, lo, hi, seed, terms)
    ds = ctx.dataset(x, y)
    inits_all = prior_draws(terms, ig, chains_total, seed + 17)
    first, count = CH.shard(chains_total, world, dist.get_rank() if dist is not None else 0)
    dev = f"cuda:{local}"

    def run(nit, seed2):
        if world > 1:
            return CH.run_sharded(ctx, ds, inits_all, ig, 0.05, nit, seed2, on_notpd=1, formula=0)
        return ds.mh_run_ex(inits_all, ig, 0.05, nit, rng=G.RNG_PHILOX, seed=seed2, on_notpd=1, formula=0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run(warm, 100)
    barrier()
    t0 = time.perf_counter()
    r = run(iters, 101)
    barrier()
    wall = time.perf_counter() - t0
    tt = torch.tensor([wall, r["device_ms"] * 1e-3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    wall, tdev = (float(v) for v in tt.cpu())
    shape = chol_frac(ctx, ds, inits_all[first:first + count])
    ds.close()
    P = len(ig)
    # the timed call also evaluates the initial state once: iters + 1 factorisations per chain
    evals = (iters + 1) * chains_total
    return {"iters_per_sec_e2e": iters / wall, "iters_per_sec_device": iters / tdev,
            "evals_per_sec_device": evals / tdev, "evaluations_timed": evals,
            "chains_total": chains_total, "chains_per_gpu": count, "n": n, "iters": iters,
            "accept_rate": float(np.mean(r["accept"])), "factorisations_per_iter_per_chain": 1,
            "flushes": r["flushes"], "gathered_on_every_rank": bool(world > 1),
            "d2h_bytes": int(iters * chains_total * (P + 2) * 8),
            "loop_tflops": evals * chol_flops(n) / world / tdev * 1e-12,
            "loop_frac_of_dmma_peak": evals * chol_flops(n) / world / tdev * 1e-12 / FP64_PEAK_TFLOPS,
            "factor_kernel_at_this_shape": shape,
            "inputs": f"SURVEY 8d: x ~ U({lo},{hi})^2, y = L z (seed {seed}), inits = prior draws"}


def extras_single_gpu(ctx, ds, G, CH):
    """The other numbers BASELINE.json's metric string / configs name (not the headline)."""
    out = {}
    rng = np.random.default_rng(5)

    def guarded(name, fn):
        try:
            out[name] = fn()
        except Exception as e:  # never lose the headline line to a secondary measurement
            out[name] = {"error": repr(e)[:200]}

    def latency():
        for _ in range(3):
            ds.loglik_batch([gterms(proposals(rng, 1)[0])])
        t0 = time.perf_counter(); dev = 0.0; chol = 0.0
        for _ in range(20):
            ds.loglik_batch([gterms(proposals(rng, 1)[0])]); t = ctx.last_timing(); dev += sum(t); chol += t[1]
        tf = 20 * chol_flops(N_TRAIN) / (chol * 1e-3) * 1e-12
        return {"evals_per_sec_e2e": 20 / (time.perf_counter() - t0), "evals_per_sec_device": 20 / (dev * 1e-3),
                "chol_ms": chol / 20, "chol_tflops": tf, "frac_of_dmma_peak": tf / FP64_PEAK_TFLOPS,
                "note": "the reference's own call shape: one proposal per call (chain-bound, one CTA per SM)"}

    def c2_predict():
        m = 4096
        xs = rng.uniform(-10, 10, size=(m, 2))
        ds.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        t0 = time.perf_counter()
        ds.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        wall = time.perf_counter() - t0
        t = ctx.last_timing()
        fl = chol_flops(N_TRAIN) + float(N_TRAIN) ** 2 * m
        tf = fl / (t[1] * 1e-3) * 1e-12
        return {"m": m, "points_per_sec_e2e": m / wall, "device_ms": sum(t), "factor_solve_tflops": tf,
                "frac_of_dmma_peak": tf / FP64_PEAK_TFLOPS}

    def c4():
        n, m = 32768, 10000
        x4, y4 = synth_dataset(ctx, n, 2, -30, 30, 4, TRUE_TERMS)
        xs = np.random.default_rng(44).uniform(-30, 30, size=(m, 2))
        ds4 = ctx.dataset(x4, y4)
        res = {}
        ds4.loglik_batch([gterms(TRUE_TERMS)])
        t0 = time.perf_counter(); r = ds4.loglik_batch([gterms(TRUE_TERMS)])[0]; w = time.perf_counter() - t0
        t = ctx.last_timing()
        tf = chol_flops(n) / (t[1] * 1e-3) * 1e-12
        res["loglik_n32768"] = {"info": r.info, "wall_s": w, "build_ms": t[0], "factor_solve_ms": t[1],
                                "factor_solve_tflops": tf, "frac_of_dmma_peak": tf / FP64_PEAK_TFLOPS}
        ds4.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        t0 = time.perf_counter(); mean, var, info = ds4.predict(gterms(TRUE_TERMS), xs, nugget=1e-6); w = time.perf_counter() - t0
        t = ctx.last_timing()
        fl = chol_flops(n) + float(n) ** 2 * m
        tf = fl / (t[1] * 1e-3) * 1e-12
        res["predict_n32768_m10000"] = {"info": info, "wall_s": w, "points_per_sec_e2e": m / w, "build_ms": t[0],
                                        "factor_solve_ms": t[1], "epilogue_ms": t[2], "factor_solve_tflops": tf,
                                        "frac_of_dmma_peak": tf / FP64_PEAK_TFLOPS}
        res["inputs"] = "SURVEY 8d C4: x ~ U(-30,30)^2, y = L z generated on the GPU (gpb_draw), seed 4"
        ds4.close()
        return res

    def c1():
        """configs[0]: the shipped example (SimulatedGp.scala): 1-D, n=300 sorted U(-10,10) AND the thinning
        to n=20 it actually runs (:112), SE+white, 2 chains (:164), log-normal random walk delta=0.05.  GPU =
        gpb_mh_run_ex (n=20: the one-launch kernel), 1 factorisation / iteration; CPU = the oracle running
        Breeze's 2 evaluations per iteration on one core per chain (cpu_baseline leg: the oracle is only timed)."""
        from oracle import oracle as o
        x1, y1 = synth_dataset(ctx, 300, 1, -10, 10, 1, TRUE_TERMS)
        keep = [(i + 1) % 15 == 0 for i in range(300)]
        res = {}
        for label, xx, yy, it_gpu, it_cpu in (("n300", x1, y1, 2000, 40), ("n20_thinned", x1[keep], y1[keep], 20000, 4000)):
            d1 = ctx.dataset(xx, yy)
            inits = [gterms(TRUE_TERMS)] * 2
            d1.mh_run_ex(inits, IG3, 0.05, 50, rng=G.RNG_PHILOX, seed=11, on_notpd=1)
            t0 = time.perf_counter(); r = d1.mh_run_ex(inits, IG3, 0.05, it_gpu, rng=G.RNG_PHILOX, seed=12, on_notpd=1); w = time.perf_counter() - t0
            z, u = G.philox_streams(11, 0, 1, it_cpu, 3)
            t0 = time.perf_counter()
            o.mh_run(xx, yy, TRUE_TERMS, IG3, 0.05, z[:, 0], u[:, 0], evals_per_iter=2)
            wc = time.perf_counter() - t0
            entry = {"n": len(xx), "chains": 2, "gpu_path": "one-launch small-n kernel" if r["path"] == 1 else "general",
                     "gpu_iters_per_sec_e2e_2_chains": it_gpu / w, "gpu_us_per_iter_device": 1e3 * r["device_ms"] / it_gpu,
                     "cpu_oracle_iters_per_sec_per_chain_one_core": it_cpu / wc,
                     "gpu_over_cpu_at_2_chains": (it_gpu / w) / (it_cpu / wc)}
            if len(xx) <= 64:
                # where the GPU starts to pay at the shipped size: iterations/s stay flat until the chains fill the SMs
                sweep = {}
                for ch in (2, 148, 1184, 4736):
                    d1.mh_run_ex([gterms(TRUE_TERMS)] * ch, IG3, 0.05, 200, rng=G.RNG_PHILOX, seed=5, on_notpd=1)
                    t0 = time.perf_counter(); d1.mh_run_ex([gterms(TRUE_TERMS)] * ch, IG3, 0.05, 4000, rng=G.RNG_PHILOX, seed=6, on_notpd=1)
                    sweep[str(ch)] = 4000 / (time.perf_counter() - t0)
                entry["gpu_iters_per_sec_e2e_by_chains"] = sweep
                entry["chain_iterations_per_sec_at_4736_chains"] = sweep["4736"] * 4736
            res[label] = entry
            d1.close()
        res["note"] = "CPU = oracle port, 2 likelihood evaluations per iteration like Breeze metropolis, one core per chain (the reference runs its 2 chains on 2 threads); GPU = 1 evaluation per iteration"
        return res

    def cpu_lapack():
        """scipy/OpenBLAS dpotrf + dtrsv on all host cores, same n=4096 system."""
        import scipy.linalg as sl
        xx, yy = make_dataset()
        d2 = ((xx[:, None, :] - xx[None, :, :]) ** 2).sum(-1)
        K = 3.0 * np.exp(-d2 / 30.25) + 0.501 * np.eye(N_TRAIN)
        sl.cholesky(K, lower=True)
        t0 = time.perf_counter()
        for _ in range(3):
            L = sl.cholesky(K, lower=True); sl.solve_triangular(L, yy, lower=True)
        dt = (time.perf_counter() - t0) / 3
        return {"factor_plus_solve_ms": dt * 1e3, "evals_per_sec_factor_only_path": 1.0 / dt, "cores": os.cpu_count(),
                "note": "LAPACK time excludes the covariance build"}

    def c2_gradient():
        """mllGradient at C2's size, 8 parameter vectors per call (log parameters)."""
        BATCH = 8
        th = [[(0, math.log(3.0 + 0.1 * b), math.log(5.5)), (1, math.log(0.5), 0.0)] for b in range(BATCH)]
        for _ in range(2):
            ds.mll_gradient(th, mode=G.GRAD_ANALYTIC)
        t0 = time.perf_counter(); g, info = ds.mll_gradient(th, mode=G.GRAD_ANALYTIC); w = time.perf_counter() - t0
        t = ctx.last_timing()
        n = float(N_TRAIN)
        return {"batch": BATCH, "info": int(info.max()), "gradients_per_sec_e2e": BATCH / w, "build_ms": t[0],
                "factor_and_inverse_rows_ms": t[1], "contraction_ms": t[2],
                "factor_and_inverse_rows_tflops": BATCH * (2 * n ** 3 / 3) / (t[1] * 1e-3) * 1e-12,
                "contraction_tflops": BATCH * (n ** 3 / 3) / (t[2] * 1e-3) * 1e-12}

    def c3_mean_step():
        """Mcmc.sample with a Plane mean (mean Gibbs step + kernel step, 3 factorisations per chain and
        iteration like the reference): 64 chains x n=1024, device resident."""
        n, chains, iters = 1024, 64, 40
        x3, y3 = synth_dataset(ctx, n, 2, -10, 10, 3, TRUE_TERMS)
        y3 = y3 + 0.7 + 0.2 * x3[:, 0] - 0.1 * x3[:, 1]
        X = np.hstack([np.ones((n, 1)), x3.reshape(-1).reshape((2, n)).T])   # makeDesignMatrix layout
        d3 = ctx.dataset(x3, y3)
        inits = prior_draws(TRUE_TERMS, IG3, chains, 31)
        kw = dict(rng=G.RNG_PHILOX, beta0=np.zeros((chains, 3)), design=X, mean_step=True, mean_prior=(0.0, 4.0), on_notpd=1)
        d3.mh_run_ex(inits, IG3, 0.05, 4, seed=1, **kw)
        t0 = time.perf_counter(); r = d3.mh_run_ex(inits, IG3, 0.05, iters, seed=2, **kw); w = time.perf_counter() - t0
        d3.close()
        return {"n": n, "chains": chains, "iters": iters, "iters_per_sec_e2e": iters / w, "iters_per_sec_device": iters / (r["device_ms"] * 1e-3),
                "factorisations_per_iter_per_chain": 3, "accept_rate": float(np.mean(r["accept"]))}

    guarded("c1_shipped_example_mcmc", c1)
    guarded("cpu_lapack_b2_n4096", cpu_lapack)
    guarded("c2_single_proposal_latency", latency)
    guarded("c2_predict_m4096", c2_predict)
    guarded("c2_mll_gradient_batch8", c2_gradient)
    guarded("c3_mcmc_256chains_n1024", lambda: mh_bench(ctx, G, CH, 1024, 256, 200, 20, TRUE_TERMS, IG3, 3, -10, 10))
    guarded("c3_mcmc_32chains_n1024_per_gpu_shape_of_8_gpus", lambda: mh_bench(ctx, G, CH, 1024, 32, 400, 20, TRUE_TERMS, IG3, 3, -10, 10))
    guarded("c3_plane_mean_gibbs_64chains_n1024", c3_mean_step)
    guarded("c5_mcmc_64chains_n8192_se_se_white", lambda: mh_bench(ctx, G, CH, 8192, 64, 32, 2, C5_TERMS, IG5, 5, -15, 15))
    guarded("c4", c4)
    return out


def extras_multi_gpu(ctx, G, CH, dist, local):
    """N > 1: the chain-parallel configs (C3 strong: 256 chains over N GPUs; C5 weak: 64 per GPU)."""
    out = {}
    world = dist.get_world_size()
    for name, fn in (
            ("c3_mcmc_256chains_n1024", lambda: mh_bench(ctx, G, CH, 1024, 256, 400, 20, TRUE_TERMS, IG3, 3, -10, 10, dist, local)),
            ("c5_mcmc_64chains_per_gpu_n8192_se_se_white", lambda: mh_bench(ctx, G, CH, 8192, 64 * world, 32, 2, C5_TERMS, IG5, 5, -15, 15, dist, local))):
        try:
            out[name] = fn()
        except Exception as e:
            out[name] = {"error": repr(e)[:200]}
    return out


_JSON_OUT = None


def claim_stdout():
    """stdout carries ONE JSON line: native libraries that write to fd 1 on their own (NCCL's version
    banner under NCCL_DEBUG=VERSION/WARN/INFO without NCCL_DEBUG_FILE) are pointed at stderr, and the
    line goes out through a private duplicate of the original stdout."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = 3 if args.steps is None else args.steps
        args.warmup = 1 if args.warmup is None else args.warmup
        run_reference(args)
    else:
        args.steps = 100 if args.steps is None else args.steps
        args.warmup = 3 if args.warmup is None else args.warmup
        run_ours(args)


if __name__ == "__main__":
    main()
