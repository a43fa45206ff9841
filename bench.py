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
              then all-gathers its log-likelihoods over NCCL (north_star: gather only). scaling=weak.
  --impl reference : the CPU oracle (port of the Scala/Breeze path; the JVM is not in this image)
              on all host cores, one independent evaluation per thread, same config and metric.
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


def run_reference(args):
    """--impl reference: the reference's CPU path restated (oracle port; JVM unavailable on this
    box), all host cores, same config / metric / unit.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    x, y = make_dataset()
    rng = np.random.default_rng(1234)
    # A step (one full n=4096 evaluation per host thread) takes ~10 s of wall clock: whatever K and W
    # the caller passes, the run is bounded to a few minutes -- warm-up steps while under 30 s, timed
    # steps while under 170 s (at least one), and the line says how many were executed.  The metric
    # is a rate, so it does not depend on the count.
    t_start = time.perf_counter()
    warm_done = 0
    for _ in range(args.warmup):
        if warm_done >= 1 and time.perf_counter() - t_start > 30.0:
            break
        cpu_evals(x, y, proposals(rng, cores), cores); warm_done += 1
    total_t, total_n, steps_done = 0.0, 0, 0
    for _ in range(max(args.steps, 1)):
        if steps_done >= 1 and time.perf_counter() - t_start > 170.0:
            break
        _, dt, _ = cpu_evals(x, y, proposals(rng, cores), cores)
        total_t += dt; total_n += cores; steps_done += 1
    v = total_n / total_t
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / steps_done,
        "steps_executed": steps_done, "warmup_executed": warm_done,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"C2: n={N_TRAIN} 2-D SE+white log-marginal-likelihood, {cores} independent proposals per step (one per host thread)"},
        "cpu_baseline": {"value": v, "unit": "evals/s", "cores": cores, "kind": "port",
                         "sample": f"{steps_done} of {args.steps} steps executed (time-bounded) x {cores} full n={N_TRAIN} evaluations, 1 thread each (reference JVM path: unavailable on this box)"},
        "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    from gpb200 import _lib as G

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
    x, y = make_dataset()
    ds = ctx.dataset(x, y)
    B = args.batch
    rng = np.random.default_rng(1000 + rank)
    gather_in = torch.zeros(B, dtype=torch.float64, device=f"cuda:{local}")
    gather_out = torch.zeros(B * world, dtype=torch.float64, device=f"cuda:{local}") if world > 1 else None

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        props = [gterms(p) for p in proposals(rng, B)]
        rs = ds.loglik_batch(props, nugget=1e-3)       # H2D params, kernels, D2H results
        t = ctx.last_timing()
        if dist is not None:                            # gather the log-likelihoods (north_star)
            gather_in.copy_(torch.tensor([G.ll_value(r, 0) for r in rs], dtype=torch.float64))
            dist.all_gather_into_tensor(gather_out, gather_in)
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
            extra = extras_single_gpu(ctx, ds, G)
        else:
            extra = extras_multi_gpu(ctx, G, dist, local)   # collective: every rank takes part
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            v, dt, res = cpu_evals(x, y, proposals(np.random.default_rng(7), cores), cores)
            cpu = {"value": v, "unit": "evals/s", "cores": cores, "kind": "port",
                   "sample": f"{cores} full n={N_TRAIN} evaluations, one per thread, {dt:.1f} s (reference JVM path: unavailable on this box)"}
        traffic = None
        tp = os.path.join(ROOT, "profiles", "chol_traffic_bytes_per_launch.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(str(B))
        line = {
            "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C2: n={N_TRAIN} 2-D SE+white log-marginal-likelihood, batch of {B} independent proposals per step per GPU",
                       "batch": B, "n": N_TRAIN, "dim": DIM, "parallelism": f"proposal-sharded x{world}",
                       "l2": "inputs larger than L2 (each proposal rebuilds and refactors its own 134 MB matrix)"},
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
        print(json.dumps(line), flush=True)
    ds.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def _mh_setup(n, lo, hi, seed, terms):
    """Synthetic chain workload: shared dataset (as writeParallelChain shares it), SURVEY.md 8d."""
    r = np.random.default_rng(seed)
    x = r.uniform(lo, hi, size=(n, 2))
    y = r.standard_normal(n) * 1.8  # timing only: parity of the chain itself is in tests/
    return x, y, r


def mh_bench(ctx, G, n, chains_total, iters, warm, terms, ig, seed, lo, hi, dist=None, local=0):
    """MCMC iterations/s for `chains_total` chains sharded over the ranks (gpb_mh_run per rank,
    then ONE NCCL all_gather of accepted states + ll + accept flags).  Timed end to end through
    the ABI call with host buffers (streams H2D, chain outputs D2H), barrier on both sides, max
    over ranks; the device time of the loop (CUDA events inside the library) is reported too."""
    import torch
    from gpb200 import chains as CH
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    x, y, r = _mh_setup(n, lo, hi, seed, terms)
    ds = ctx.dataset(x, y)
    first, count = CH.shard(chains_total, world, rank)
    P = sum(2 if t[0] == "se" else 1 for t in terms)
    rr = np.random.default_rng(seed + 17)
    inits_all = []
    for _ in range(chains_total):
        z = rr.standard_normal(P) * 0.05
        k, tl = 0, []
        for t in terms:
            if t[0] == "se":
                tl.append((0, t[1] * math.exp(z[k]), t[2] * math.exp(z[k + 1]))); k += 2
            else:
                tl.append((1, t[1] * math.exp(z[k]), 0.0)); k += 1
        inits_all.append(tl)
    inits = inits_all[first:first + count]
    dev = f"cuda:{local}"

    def run(nit, seed2):
        z, u = G.philox_streams(seed2, first, count, nit, P)
        st, ll, acc, info = ds.mh_run(inits, ig, 0.05, z, u, on_notpd=1)
        tdev = ctx.last_timing()[1]
        if dist is not None and world > 1:
            width = -(-chains_total // world)
            block = np.zeros((nit, width, P + 2))
            block[:, :count, :P] = st; block[:, :count, P] = ll; block[:, :count, P + 1] = acc
            t = torch.from_numpy(block).to(dev)
            out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=dev)
            dist.all_gather_into_tensor(out, t)
            out.cpu()
        return acc, tdev

    def barrier():
        torch.cuda.synchronize()
        if dist is not None and world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run(warm, 100)
    barrier()
    t0 = time.perf_counter()
    acc, tdev = run(iters, 101)
    barrier()
    wall = time.perf_counter() - t0
    tt = torch.tensor([wall, tdev * 1e-3], dtype=torch.float64, device=dev)
    if dist is not None and world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    wall, tdev = (float(v) for v in tt.cpu())
    ds.close()
    # the timed call also evaluates the initial state once (iters + 1 factorisations per chain)
    return {"iters_per_sec_e2e": iters / wall, "iters_per_sec_device": iters / tdev,
            "evals_per_sec_device": (iters + 1) * chains_total / tdev,
            "chains_total": chains_total, "chains_per_gpu": count, "n": n, "iters": iters,
            "accept_rate": float(np.mean(acc)), "factorisations_per_iter_per_chain": 1,
            "gflop_per_iter": chains_total * chol_flops(n) * 1e-9}


IG3 = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]                       # SimulatedGp.scala:132-134
IG5 = [[3.0, 4.0], [3.0, 6.0], [3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
C5_TERMS = [("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)]


def extras_single_gpu(ctx, ds, G):
    """Secondary numbers named by BASELINE.json's metric string / configs (not the headline)."""
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
        t0 = time.perf_counter(); dev = 0.0
        for _ in range(20):
            ds.loglik_batch([gterms(proposals(rng, 1)[0])]); dev += sum(ctx.last_timing())
        return {"evals_per_sec_e2e": 20 / (time.perf_counter() - t0), "evals_per_sec_device": 20 / (dev * 1e-3)}

    def c2_predict():
        m = 4096
        xs = rng.uniform(-10, 10, size=(m, 2))
        ds.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        t0 = time.perf_counter()
        ds.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        wall = time.perf_counter() - t0
        t = ctx.last_timing()
        fl = chol_flops(N_TRAIN) + float(N_TRAIN) ** 2 * m
        return {"m": m, "points_per_sec_e2e": m / wall, "device_ms": sum(t), "factor_solve_tflops": fl / (t[1] * 1e-3) * 1e-12}

    def c4():
        import torch
        n, m = 32768, 10000
        r4 = np.random.default_rng(4)
        x4 = r4.uniform(-30, 30, size=(n, 2)); y4 = r4.standard_normal(n) * 1.8
        xs = r4.uniform(-30, 30, size=(m, 2))
        ds4 = ctx.dataset(x4, y4)
        res = {}
        ds4.loglik_batch([gterms(TRUE_TERMS)])
        t0 = time.perf_counter(); r = ds4.loglik_batch([gterms(TRUE_TERMS)])[0]; w = time.perf_counter() - t0
        t = ctx.last_timing()
        res["loglik_n32768"] = {"info": r.info, "wall_s": w, "build_ms": t[0], "factor_solve_ms": t[1],
                                "factor_solve_tflops": chol_flops(n) / (t[1] * 1e-3) * 1e-12}
        ds4.predict(gterms(TRUE_TERMS), xs, nugget=1e-6)
        t0 = time.perf_counter(); mean, var, info = ds4.predict(gterms(TRUE_TERMS), xs, nugget=1e-6); w = time.perf_counter() - t0
        t = ctx.last_timing()
        fl = chol_flops(n) + float(n) ** 2 * m
        res["predict_n32768_m10000"] = {"info": info, "wall_s": w, "points_per_sec_e2e": m / w, "build_ms": t[0],
                                        "factor_solve_ms": t[1], "epilogue_ms": t[2],
                                        "factor_solve_tflops": fl / (t[1] * 1e-3) * 1e-12}
        ds4.close()
        return res

    def c1():
        """configs[0]: the shipped example (SimulatedGp.scala): 1-D, n=300 sorted U(-10,10), SE+white,
        2 chains, log-normal random walk delta=0.05; GPU device-resident MH vs the CPU oracle running
        the reference's 2 evaluations per iteration (cpu_baseline leg: the oracle is only timed)."""
        from oracle import oracle as o
        r1 = np.random.default_rng(1)
        n, chains, iters = 300, 2, 300
        x1 = np.sort(r1.uniform(-10, 10, n)); y1 = r1.standard_normal(n) * 1.8
        ds1 = ctx.dataset(x1, y1)
        inits = [gterms(TRUE_TERMS)] * chains
        z, u = G.philox_streams(11, 0, chains, iters, 3)
        ds1.mh_run(inits, IG3, 0.05, z[:5], u[:5], on_notpd=1)
        t0 = time.perf_counter(); ds1.mh_run(inits, IG3, 0.05, z, u, on_notpd=1); w = time.perf_counter() - t0
        ds1.close()
        it_cpu = 40
        t0 = time.perf_counter()
        o.mh_run(x1, y1, TRUE_TERMS, IG3, 0.05, z[:it_cpu, 0], u[:it_cpu, 0], evals_per_iter=2)
        wc = time.perf_counter() - t0
        return {"n": n, "chains": chains, "gpu_iters_per_sec_e2e": iters / w,
                "cpu_oracle_iters_per_sec_one_chain_one_core": it_cpu / wc,
                "note": "CPU = oracle port, 2 likelihood evaluations per iteration like Breeze metropolis; GPU = 1"}

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
        import math
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

    guarded("c1_example_mcmc_n300", c1)
    guarded("cpu_lapack_b2_n4096", cpu_lapack)
    guarded("c2_single_proposal_latency", latency)
    guarded("c2_predict_m4096", c2_predict)
    guarded("c2_mll_gradient_batch8", c2_gradient)
    guarded("c3_mcmc_256chains_n1024", lambda: mh_bench(ctx, G, 1024, 256, 100, 5, TRUE_TERMS, IG3, 3, -10, 10))
    guarded("c5_mcmc_64chains_n8192_se_se_white", lambda: mh_bench(ctx, G, 8192, 64, 8, 1, C5_TERMS, IG5, 5, -15, 15))
    guarded("c4", c4)
    return out


def extras_multi_gpu(ctx, G, dist, local):
    """N > 1: the chain-parallel configs (C3 strong: 256 chains over N GPUs; C5 weak: 64 per GPU)."""
    out = {}
    world = dist.get_world_size()
    for name, fn in (
            ("c3_mcmc_256chains_n1024", lambda: mh_bench(ctx, G, 1024, 256, 100, 5, TRUE_TERMS, IG3, 3, -10, 10, dist, local)),
            ("c5_mcmc_64chains_per_gpu_n8192_se_se_white", lambda: mh_bench(ctx, G, 8192, 64 * world, 8, 1, C5_TERMS, IG5, 5, -15, 15, dist, local))):
        try:
            out[name] = fn()
        except Exception as e:
            out[name] = {"error": repr(e)[:200]}
    return out


def main():
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
