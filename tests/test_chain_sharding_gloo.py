"""world_size-2 gloo test (CPU) of the only multi-GPU logic the path has: chain sharding, per-rank
Philox streams and the gather of states / log-likelihoods / accept flags.  The per-chain compute is
injected (the CPU oracle stands in for the device runner: test infrastructure only)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TERMS = [("se", 3.0, 5.5), ("white", 0.5)]
IG = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
N, CHAINS, ITERS, SEED = 24, 5, 12, 77


def _data():
    rng = np.random.default_rng(0)
    x = np.sort(rng.uniform(-10, 10, N))
    y = rng.standard_normal(N)
    return x, y


def _runner(first, z, u):
    from oracle import oracle as o
    x, y = _data()
    sts, lls, accs = [], [], []
    for c in range(z.shape[1]):
        st, ll, acc, rc = o.mh_run(x, y, TERMS, IG, 0.05, z[:, c], u[:, c], formula=1)
        assert rc == 0
        sts.append(st); lls.append(ll); accs.append(acc)
    return np.stack(sts, 1), np.stack(lls, 1), np.stack(accs, 1)


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "gaussian-processes_b200")):
        sys.path.insert(0, p)
    import torch.distributed as dist
    from gpb200 import chains
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    st, ll, acc = chains.run_sharded(CHAINS, ITERS, 3, SEED, _runner, dist=dist)
    q.put((rank, st, ll, acc))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_partition():
    from gpb200.chains import shard
    for total in (0, 1, 5, 256):
        for world in (1, 2, 4, 8):
            blocks = [shard(total, world, r) for r in range(world)]
            assert sum(c for _, c in blocks) == total
            assert all(blocks[i][0] + blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
    assert [shard(256, 8, r)[1] for r in range(8)] == [32] * 8  # C3: 256 chains on 8 GPUs


def test_two_rank_gather_equals_single_process():
    from gpb200 import chains
    want = chains.run_sharded(CHAINS, ITERS, 3, SEED, _runner)  # world 1
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, st, ll, acc in got:
        assert st.shape == (ITERS, CHAINS, 3)
        assert np.array_equal(st, want[0]) and np.array_equal(ll, want[1]) and np.array_equal(acc, want[2])
