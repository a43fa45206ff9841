"""world_size-2 gloo tests (CPU) of the host-side multi-GPU logic: the block partition (gpb_comm_shard),
per-rank Philox streams keyed by the global chain index, the padded-block gather of states /
log-likelihoods / accept flags (chains.run_sharded_host, the host restatement of what gpb_mh_run_ex
does with NCCL on the device) and the hand-over of the communicator id through the process group.
The per-chain compute is injected (the CPU oracle stands in for the device: test infrastructure only).
The device-to-device gather itself is tested on 2 GPUs in tests/test_gpu_mh.py."""
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
    st, ll, acc = chains.run_sharded_host(CHAINS, ITERS, 3, SEED, _runner, dist=dist)
    uid = chains.exchange_id(dist)
    q.put((rank, st, ll, acc, uid))
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
    assert [shard(9, 2, r) for r in range(2)] == [(0, 5), (5, 4)]  # the first total % world ranks own one more


def test_two_rank_gather_equals_single_process():
    from gpb200 import chains
    want = chains.run_sharded_host(CHAINS, ITERS, 3, SEED, _runner)  # world 1
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
    assert got[0][4] == got[1][4] and len(got[0][4]) == 128 and any(got[0][4])   # both ranks hold rank 0's id
    for rank, st, ll, acc, _uid in got:
        assert st.shape == (ITERS, CHAINS, 3)
        assert np.array_equal(st, want[0]) and np.array_equal(ll, want[1]) and np.array_equal(acc, want[2])
