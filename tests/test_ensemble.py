"""Replica ensembles sharded over ranks: the N>1 path has no data-path collective,
only a gather of per-replica statistics.  CPU: two gloo ranks on the oracle backend against one
process; GPU: shards of one ensemble on the engine reproduce the unsharded run bit for bit."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import bdb200 as bd
import ensemble as en
import scenarios as sc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shards_partition_the_ensemble():
    for n, world in ((1024, 8), (5, 2), (3, 4), (0, 2)):
        ids = [en.shard_replicas(n, world, r) for r in range(world)]
        assert sorted(np.concatenate(ids).tolist()) == list(range(n))
        sizes = [len(i) for i in ids]
        assert max(sizes) - min(sizes) <= 1
        assert all((i % world == r).all() for r, i in enumerate(ids))
    assert len({en.replica_seed(9, r) for r in range(1024)}) == 1024
    with pytest.raises(ValueError):
        en.shard_replicas(4, 2, 2)


def test_radius_of_gyration():
    x = np.array([[0.0, 0, 0], [2, 0, 0], [0, 0, 0], [0, 0, 4]])
    assert np.allclose(en.radius_of_gyration(x, [(0, 2), (2, 4)]), [1.0, 2.0])
    assert en.radius_of_gyration(x).shape == (1,)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_gloo_ranks_match_one_process(tmp_path, oracle):
    n_global, steps = 5, 20
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "ensemble_worker.py"), str(tmp_path), str(n_global), str(steps)]
    subprocess.run(cmd, check=True, timeout=600, cwd=ROOT)

    scen = sc.genome_10mb(total=320, bend_energy=0.7)
    single = en.ReplicaEnsemble(oracle, scen, n_global, seed=9)
    single.step(steps)
    want_stats, want_pos = single.statistics(), single.positions()
    assert want_stats.shape == (n_global, len(scen.chains))
    for rank in range(2):
        stats = np.load(tmp_path / f"stats_rank{rank}.npy")
        ids = np.load(tmp_path / f"ids_rank{rank}.npy")
        pos = np.load(tmp_path / f"pos_rank{rank}.npy")
        assert (ids == en.shard_replicas(n_global, 2, rank)).all()
        assert np.array_equal(stats, want_stats)      # every rank holds the whole table, bit for bit
        assert np.array_equal(pos, want_pos[ids])     # a replica's trajectory does not depend on its rank
    # different replicas did take different noise
    assert not np.array_equal(want_pos[0], want_pos[1])


@pytest.mark.gpu
def test_engine_shards_reproduce_the_unsharded_ensemble(engine):
    scen = sc.genome_10mb(total=320, bend_energy=0.7)
    n_global, steps = 6, 60
    whole = en.ReplicaEnsemble(engine, scen, n_global, seed=4)
    whole.step(steps)
    want = whole.positions()
    assert not np.array_equal(want[0], want[1])
    for world in (2, 4):
        for rank in range(world):
            shard = en.ReplicaEnsemble(engine, scen, n_global, rank=rank, world=world, seed=4)
            shard.step(steps // 2)
            shard.step(steps - steps // 2)
            assert np.array_equal(shard.positions(), want[shard.ids]), (world, rank)
            shard.close()
    stats = whole.statistics()
    assert stats.shape == (n_global, len(scen.chains)) and np.isfinite(stats).all() and (stats > 0).all()
    whole.close()


@pytest.mark.gpu
def test_host_pipeline_equals_the_single_engine_ensemble(engine):
    """ensemble.HostPipeline (replica chunks on their own streams, upload -> step -> download per chunk) advances
    every replica as one engine holding all of them does: same noise keys (global replica ids), the same
    positions after steps driven through the host buffer (to rounding: the chunks have their own cell grids,
    so forces are summed in a different order)."""
    import torch

    import ensemble as en

    R, steps = 6, 8
    scen = sc.genome_10mb(total=320, replicas=R, bend_energy=0.7)
    host = torch.empty((R, scen.n, 3), dtype=torch.float64).pin_memory()
    host.numpy()[:] = scen.positions
    pipe = en.HostPipeline(engine, scen, R, chunks=4, seed=5)
    for _ in range(steps):
        pipe.step_through_host(host, host, scen.timestep, scen.temperature)
    got = host.numpy().copy()
    pipe.close()

    one = bd.System(engine, scen.n, R)
    scen.apply(one)
    for k in range(steps):  # the same per-step host round trip, one engine
        one.step(1, scen.timestep, scen.temperature, seed=5, first_step=k + 1)
    want = np.stack([one.get_positions(r) for r in range(R)])
    one.close()
    assert np.abs(got - want).max() <= 1e-12
    assert np.abs(got - scen.positions).max() > 1e-4
