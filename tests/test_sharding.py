"""world_size-2 gloo test of the multi-rank plumbing (no GPU): shard ranges, per-replica deterministic inputs and
the observable all-gather that bench.py performs over NCCL."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, total, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bench import load_ala10, synth_inputs
    from molecular_dynamics_b200.sharding import gather_observables, shard_range

    lo, hi = shard_range(total, rank, world)
    g = load_ala10()
    pos, vel = synth_inputs(g, hi - lo, lo)
    # stand-in observables computed from the shard's own inputs: global replica id and a position checksum
    obs = torch.tensor(np.stack([np.arange(lo, hi, dtype=np.float64), pos.astype(np.float64).sum((1, 2))], 1))
    full = gather_observables(obs, total)
    if rank == 0:
        q.put(full.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions_exactly():
    from molecular_dynamics_b200.sharding import shard_range

    for total in (1, 7, 8192, 32768, 100):
        for world in (1, 2, 3, 8):
            edges = [shard_range(total, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(180)
def test_two_rank_gloo_gather_matches_single_rank():
    total, world = 13, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    full = q.get(timeout=150)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT)
    from bench import load_ala10, synth_inputs

    pos, _ = synth_inputs(load_ala10(), total, 0)
    np.testing.assert_array_equal(full[:, 0], np.arange(total))
    # inputs are keyed by global replica id: sharding does not change them
    np.testing.assert_array_equal(full[:, 1], pos.astype(np.float64).sum((1, 2)))


def _jarzynski_worker(rank, world, port, total, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from molecular_dynamics_b200.sharding import jarzynski_free_energy, shard_range

    lo, hi = shard_range(total, rank, world)
    work = _work_table(total)[:, lo:hi]
    r = jarzynski_free_energy(work, 300.0)
    if rank == 1:  # any rank holds the ensemble value
        q.put({k: v.numpy() for k, v in r.items()})
    dist.barrier()
    dist.destroy_process_group()


def _work_table(total):
    g = torch.Generator().manual_seed(3)
    return 40.0 * torch.arange(1, 6, dtype=torch.float64)[:, None] + torch.randn((5, total), generator=g, dtype=torch.float64)


@pytest.mark.timeout(180)
def test_two_rank_gloo_jarzynski_matches_single_rank():
    from molecular_dynamics_b200.sharding import jarzynski_free_energy

    total, world = 13, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_jarzynski_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=150)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    one = jarzynski_free_energy(_work_table(total), 300.0)
    for k, v in one.items():
        np.testing.assert_allclose(got[k], v.numpy(), rtol=1e-12, atol=1e-12)
    assert got["n"].tolist() == [13.0] * 5
