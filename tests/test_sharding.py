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
