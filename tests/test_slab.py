"""Slab decomposition of one system over several ranks: ownership, migration and the
per-step halo exchange.  CPU: the protocol on a toy bead store, in process (3 slabs) and over a
world_size-2 gloo group.  GPU: slabs of the 1kb chromatin model on the engine against the undecomposed
engine run."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

import bdb200 as bd
import scenarios as sc
import slab

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_layout_geometry():
    x = np.random.default_rng(0).uniform(0, 10, 3000)
    lay = slab.SlabLayout.balanced(x, 4, period=10.0)
    owner = lay.owner_of(torch.from_numpy(x)).numpy()
    assert np.bincount(owner, minlength=4).tolist() == [750, 750, 750, 750]
    # wrapped coordinates land in the same slab; distances are minimum-image distances
    assert (lay.owner_of(torch.from_numpy(x + 30.0)).numpy() == owner).all()
    t = torch.tensor([0.05, 9.95, 5.0], dtype=torch.float64)
    d0, d3 = lay.distance_to(t, 0).numpy(), lay.distance_to(t, 3).numpy()
    assert d0[0] == 0 and abs(d0[1] - 0.05) < 1e-12 and d3[1] == 0 and abs(d3[0] - 0.05) < 1e-12
    for g in range(4):
        inside = lay.distance_to(torch.from_numpy(x), g).numpy() == 0
        assert (inside == (owner == g)).all()
    open_lay = slab.SlabLayout.balanced(x, 2)
    assert open_lay.distance_to(torch.tensor([-100.0]), 0).item() == 0
    with pytest.raises(ValueError):
        slab.SlabLayout([0, 2, 1])


def replay(x0, steps, amplitude):
    x = torch.from_numpy(x0.copy())
    for k in range(1, steps + 1):
        x += slab.ToyStore.displacement(len(x0), k, amplitude)
    return x.numpy()


def test_three_slabs_in_process_follow_the_global_motion():
    x0 = np.random.default_rng(3).uniform(0.0, 8.0, size=(600, 3))
    lay = slab.SlabLayout.balanced(x0[:, 0], 3, period=8.0)
    slabs = [slab.SlabSystem(slab.ToyStore(x0, 0.02), lay, r, halo=0.9, margin=0.5, migrate_every=5) for r in range(3)]
    slab.run_local(slabs, 120, 1e-4, 0.0, 1)
    want = replay(x0, 120, 0.02)
    owned = np.stack([s.owned.numpy() for s in slabs])
    assert (owned.sum(axis=0) == 1).all()                      # a partition
    assert (owned.sum(axis=1) != np.array([200, 200, 200])).any()  # beads did migrate (net drift along x)
    for r, s in enumerate(slabs):
        idx, xyz = s.owned_positions()
        assert np.array_equal(xyz, want[idx])
        # every bead within reach of the slab is exact on this rank, owned or not
        near = lay.distance_to(torch.from_numpy(want[:, 0]), r).numpy() < 0.9 - 0.5
        assert np.array_equal(s.store.xyz.numpy()[near], want[near])
        assert s.halo_bytes_sent > 0
    # a margin that the motion outruns is reported, not silently wrong
    fast = [slab.SlabSystem(slab.ToyStore(x0, 0.2), lay, r, halo=0.9, margin=0.5, migrate_every=5) for r in range(3)]
    with pytest.raises(RuntimeError, match="drifted"):
        slab.run_local(fast, 30, 1e-4, 0.0, 1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_gloo_ranks_exchange_halos(tmp_path):
    steps = 60
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "slab_worker.py"), str(tmp_path), str(steps)]
    subprocess.run(cmd, check=True, timeout=600, cwd=ROOT)
    x0 = np.random.default_rng(3).uniform(0.0, 8.0, size=(600, 3))
    want = replay(x0, steps, 0.02)
    owned = np.stack([np.load(tmp_path / f"slab_owned_rank{r}.npy") for r in range(2)])
    assert (owned.sum(axis=0) == 1).all()
    for r in range(2):
        assert np.array_equal(np.load(tmp_path / f"slab_all_rank{r}.npy"), want)
        local = np.load(tmp_path / f"slab_local_rank{r}.npy")
        assert np.array_equal(local[owned[r]], want[owned[r]])


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_engine_slabs_match_the_undecomposed_run(engine, world):
    scen = sc.chromatin_1kb(6, 400)
    dcut = max(ff["dcut"] for ff in scen.pair_ffs)
    steps, seed = 40, 5
    whole = bd.System(engine, scen.n)
    scen.apply(whole)
    whole.step(steps, scen.timestep, scen.temperature, seed=seed)
    want = whole.get_positions()
    assert np.abs(want - scen.positions[0]).max() > 1e-3

    lay = slab.SlabLayout.balanced(scen.positions[0][:, 0], world, period=scen.period[0])
    slabs = []
    for r in range(world):
        s = bd.System(engine, scen.n)
        scen.apply(s)
        slabs.append(slab.SlabSystem(slab.EngineStore(s, "cuda:0"), lay, r, halo=3.0 * dcut, margin=dcut, migrate_every=8,
                                     partners=slab.topology_partners(scen)))
    slab.run_local(slabs, steps, scen.timestep, scen.temperature, seed)
    got = np.full_like(want, np.nan)
    for s in slabs:
        idx, xyz = s.owned_positions()
        assert 0 < len(idx) < scen.n
        got[idx] = xyz
    # same noise per bead whatever the owner; forces are summed in a different order: not bit-equal
    assert np.abs(got - want).max() < 1e-9
    # each rank evaluated only its own beads' rows
    inter = [s.store.system.get_counters().pair_interactions for s in slabs]
    total = whole.get_counters().pair_interactions
    assert abs(sum(inter) - total) <= 0.02 * total and max(inter) < 0.8 * total
