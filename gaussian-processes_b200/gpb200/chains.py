"""Chain / proposal sharding across the GPUs of one box.

The hot path shards only by independent MH chains (or batches of proposals): one process (or host
thread) per GPU, the dataset replicated, no communication inside an iteration.  The single
collective is the gather of accepted states, log-likelihoods and accept flags (north_star).  It lives
behind the C ABI -- gpb_comm_* + gpb_mh_run_ex(gather = 1): NCCL all-gather of every ring segment,
device to device, inside libgpb200.so -- so a JVM caller reaches it the same way this module does.
torch.distributed is used here for ONE thing: handing rank 0's communicator id to the other ranks.
Streams are counter-based (Philox keyed by the global chain index), so a rank generates exactly the
streams of the chains it owns.
"""
import numpy as np

from . import _lib as G


def shard(total: int, world: int, rank: int):
    """Contiguous block partition of `total` chains: (first_chain, count) for `rank` -- the partition
    gpb_mh_run_ex assumes (gpb_comm_shard; host only)."""
    if total < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError("bad shard arguments")
    return G.comm_shard(total, world, rank)


def exchange_id(dist=None, device=None):
    """Rank 0 makes the communicator id (gpb_comm_unique_id) and broadcasts it through the process
    group (any backend: nccl under torchrun, gloo in the CPU tests).  Returns the 128 bytes."""
    import torch
    if dist is None or dist.get_world_size() == 1:
        return G.comm_unique_id()
    dev = torch.device("cuda", device) if (dist.get_backend() == "nccl" and device is not None) else torch.device("cpu")
    buf = torch.zeros(G.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
    if dist.get_rank() == 0:
        buf.copy_(torch.frombuffer(bytearray(G.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(buf, src=0)
    return buf.cpu().numpy().tobytes()


def init_comm(ctx, dist=None):
    """Create the context's NCCL communicator (gpb_comm_init_rank) for this rank of the process group,
    or a 1-rank communicator in a single process.  Returns (rank, world)."""
    rank = 0 if dist is None else dist.get_rank()
    world = 1 if dist is None else dist.get_world_size()
    ctx.comm_init_rank(world, rank, exchange_id(dist, ctx.device))
    return rank, world


def run_sharded(ctx, ds, inits_all, ig, delta, iters, seed, **kw):
    """Run this rank's block of len(inits_all) chains (device Philox streams keyed by the global chain
    index) and gather every chain's outputs on every rank inside the library (gpb_mh_run_ex,
    gather = 1).  The context must carry a communicator (init_comm).  Returns mh_run_ex's dict with
    states (iters, total_chains, W), ll, accept (iters, total_chains)."""
    rank, world, ok = ctx.comm_info()
    if not ok:
        raise G.GpbError(G.E_NOCOMM, "init_comm(ctx, dist) first")
    total = len(inits_all)
    first, count = shard(total, world, rank)
    if count == 0:
        raise ValueError("fewer chains than ranks")
    return ds.mh_run_ex(inits_all[first:first + count], ig, delta, iters, rng=G.RNG_PHILOX, seed=seed,
                        chain0=first, gather=True, total_chains=total, **kw)


def run_sharded_host(total_chains, iters, P, seed, runner, dist=None):
    """Transport-agnostic restatement of the same partition + padded-block gather on the HOST, for the
    CPU tests of the N > 1 logic (gloo, world_size 2: tests/test_chain_sharding_gloo.py): `runner(first,
    z, u) -> (states[iters, c, P], ll[iters, c], accept[iters, c])` stands in for the device."""
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
    width = shard(total_chains, world, 0)[1]  # rank 0 owns the widest block (as in gpb_mh_run_ex)
    block = np.zeros((iters, width, P + 2))
    block[:, :count, :P] = st
    block[:, :count, P] = ll
    block[:, :count, P + 1] = acc
    t = torch.from_numpy(block)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    parts = [o.numpy()[:, :shard(total_chains, world, r)[1]] for r, o in enumerate(out)]
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
