"""Chain / proposal sharding across the GPUs of one box.

The hot path shards only by independent MH chains (or batches of proposals): one process per GPU,
the dataset replicated, no communication inside an iteration.  The single collective is the gather
of accepted states, log-likelihoods and accept flags (north_star), done here with
torch.distributed (NCCL on GPUs; gloo in the CPU tests).  Streams are counter-based (Philox keyed by
the global chain index), so a rank regenerates exactly the streams of the chains it owns.
"""
import numpy as np


def shard(total: int, world: int, rank: int):
    """Contiguous block partition of `total` chains: returns (first_chain, count) for `rank`."""
    if total < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError("bad shard arguments")
    base, rem = divmod(total, world)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def run_sharded(total_chains, iters, P, seed, runner, dist=None, device=None):
    """Run `runner(first_chain, z, u) -> (states[iters,c,P], ll[iters,c], accept[iters,c])` on this
    rank's chains and gather everything on every rank (all_gather of padded blocks).

    `runner` is the device path in production (Dataset.mh_run); tests inject a CPU stand-in to
    exercise the sharding / gather logic under gloo.  Returns global (states, ll, accept)."""
    from . import _lib as G
    import torch
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    first, count = shard(total_chains, world, rank)
    z, u = G.philox_streams(seed, first, max(count, 1), iters, P)
    if count > 0:
        st, ll, acc = runner(first, z[:, :count], u[:, :count])
    else:
        st, ll, acc = np.zeros((iters, 0, P)), np.zeros((iters, 0)), np.zeros((iters, 0), dtype=np.int32)
    if dist is None or world == 1:
        return st, ll, np.asarray(acc)
    width = -(-total_chains // world)  # padded per-rank block so all_gather shapes match
    block = np.zeros((iters, width, P + 2))
    block[:, :count, :P] = st
    block[:, :count, P] = ll
    block[:, :count, P + 1] = acc
    t = torch.from_numpy(block)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    parts = []
    for r, o in enumerate(out):
        _, c = shard(total_chains, world, r)
        parts.append(o.cpu().numpy()[:, :c])
    full = np.concatenate(parts, axis=1)
    return full[..., :P], full[..., P], full[..., P + 1].astype(np.int32)


# ---- chain files -----------------------------------------------------------
def param_names(terms, nbeta=0):
    """Column names in the order the examples' `format` writes them (SimulatedGp.scala:159-160):
    kernelParameters.flatMap(_.toList) ::: meanParameters.toList."""
    names, se, wh = [], 0, 0
    for t in terms:
        kind = t[0]
        if kind in (0, "se"):
            sfx = "" if se == 0 else f"_{se}"
            names += [f"h{sfx}", f"sigma{sfx}"]; se += 1
        else:
            names.append("sigma_y" if wh == 0 else f"sigma_y_{wh}"); wh += 1
    return names + [f"beta_{i}" for i in range(nbeta)]


def write_chains(prefix, states, names, chain0=0):
    """One CSV per chain, `${prefix}_${chain}.csv` (Streaming.writeParallelChain's naming, read back
    at SimulatedGp.scala:209-219 with a header row): header = parameter names, one row per
    iteration.  states: (iters, chains, P) as gathered from the devices.  Returns the paths."""
    states = np.asarray(states, dtype=np.float64)
    iters, chains, P = states.shape
    if len(names) != P:
        raise ValueError("one column name per parameter")
    paths = []
    for c in range(chains):
        path = f"{prefix}_{chain0 + c}.csv"
        with open(path, "w") as f:
            f.write(",".join(names) + "\n")
            for it in range(iters):
                f.write(",".join(repr(float(v)) for v in states[it, c]) + "\n")
        paths.append(path)
    return paths


def read_chain(path):
    """Read a chain file back: (names, (iters, P) array); values round-trip exactly (repr)."""
    with open(path) as f:
        names = f.readline().strip().split(",")
        rows = [[float(v) for v in line.split(",")] for line in f if line.strip()]
    return names, np.array(rows, dtype=np.float64).reshape(-1, len(names))
