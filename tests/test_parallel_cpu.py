"""N>1 host path on CPU: world_size-2 gloo processes shard the chains and all-reduce the
statistics exactly as the NCCL path does.  The per-rank "kernel" is replaced
here by a deterministic function of the GLOBAL chain index, so the test pins the
partition + reduction logic (sharding invariance), not the CUDA arithmetic."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lmc_atomi_b200 import parallel
from lmc_atomi_b200.mixtures import ChainStats


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_stats(offset, count):
    """Statistics of chains [offset, offset+count): theta_c = (c, -2c) 'sampled' 3 times."""
    c = torch.arange(offset, offset + count, dtype=torch.float64)
    t0, t1 = c, -2.0 * c
    mom = torch.stack([torch.tensor(3.0 * count, dtype=torch.float64), 3 * t0.sum(), 3 * t1.sum(),
                       3 * (t0 * t0).sum(), 3 * (t0 * t1).sum(), 3 * (t1 * t1).sum()])
    hist = torch.zeros((100, 100), dtype=torch.int64)
    for v in c.long().tolist():
        hist[v % 100, (7 * v) % 100] += 1
    nacc = (c.to(torch.int32) % 3)
    return ChainStats("mala", count, 3, "float32", mom, hist, None, naccept=nacc)


def _worker(rank, world_size, port, total, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        off, cnt = parallel.shard(total, rank, world_size)
        st = parallel.allreduce_chain_stats_(_fake_stats(off, cnt))
        img = torch.full((4,), float(rank + 1), dtype=torch.float64)
        n, mean, m2 = parallel.allreduce_image_sums(2 * (rank + 1), img * 2 * (rank + 1), img * img * 2 * (rank + 1))
        if rank == 0:
            torch.save({"mom": st.moments_dev, "hist": st.hist_dev, "chains": st.chains, "acc": st.global_accepted,
                        "n": n, "mean": mean, "m2": m2}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [10, 7])
def test_two_rank_gloo_allreduce_equals_single_rank(tmp_path, total):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), total, out), nprocs=2, join=True)
    got = torch.load(out)
    ref = _fake_stats(0, total)
    assert torch.equal(got["mom"], ref.moments_dev)
    assert torch.equal(got["hist"], ref.hist_dev)
    assert got["chains"] == total and got["acc"] == float(ref.naccept.sum())
    # image sums: ranks hold 2 and 4 samples of constant images 1 and 2
    assert got["n"] == 6.0
    assert torch.allclose(got["mean"], torch.full((4,), (2 * 1 + 4 * 2) / 6.0, dtype=torch.float64))
    assert torch.allclose(got["m2"], torch.full((4,), (2 * 1 + 4 * 4) / 6.0, dtype=torch.float64))


def test_shard_partition_is_exact():
    for total in (0, 1, 5, 8, 1 << 20, 4096):
        for w in (1, 2, 3, 4, 8):
            parts = [parallel.shard(total, r, w) for r in range(w)]
            assert sum(c for _, c in parts) == total
            assert parts[0][0] == 0
            for (o1, c1), (o2, _) in zip(parts, parts[1:]):
                assert o1 + c1 == o2
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1


def test_no_process_group_is_a_noop():
    st = _fake_stats(0, 5)
    before = st.moments_dev.clone()
    assert parallel.allreduce_chain_stats_(st) is st and torch.equal(st.moments_dev, before)
    n, mean, _ = parallel.allreduce_image_sums(4, torch.ones(3, dtype=torch.float64) * 8)
    assert n == 4.0 and torch.allclose(mean, torch.full((3,), 2.0, dtype=torch.float64))
