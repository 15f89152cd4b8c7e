"""Replica ensembles sharded over GPUs (BASELINE.json configs 3 and 4).

The paper runs replicas of one model as separate processes with different seeds
(5-sim-genome/scripts/run_simulation:8-11); nothing is exchanged between them.  Here the replicas of
an ensemble are dealt round-robin to the ranks -- global replica r lives on rank r mod G -- and each
rank advances its shard as ONE engine (one launch per step for all local replicas).  There is no
data-path collective.  The only communication is the gather of per-replica statistics (radius of
gyration, contact counts) when the caller asks for them.

Results do not depend on G: the engine keys its Philox noise by the GLOBAL replica id
(bd_set_option replica_id_base / replica_id_stride), so replica r follows the same trajectory
whether the ensemble runs on 1, 2, 4 or 8 GPUs.

With a CPU backend (tests only: the oracle has single-replica systems and a sequential RNG) every
local replica is its own system seeded by (seed, global id); the sharding and gathering logic is the
same and is what the world_size-2 gloo tests exercise.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

import bdb200 as bd


def shard_replicas(n_global: int, world: int, rank: int) -> np.ndarray:
    """Global ids of the replicas rank `rank` of `world` owns (round-robin)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return np.arange(rank, n_global, world, dtype=np.int64)


def replica_seed(seed: int, replica: int) -> int:
    """Seed of one replica for backends with a sequential generator (splitmix64 of the pair)."""
    z = (seed * 0x9E3779B97F4A7C15 + replica + 1) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return z ^ (z >> 31)


def radius_of_gyration(xyz: np.ndarray, chains: Optional[np.ndarray] = None) -> np.ndarray:
    """Rg of every chain (or of the whole system) of one replica; xyz is [n, 3]."""
    spans = [(0, len(xyz))] if chains is None else [tuple(c) for c in np.asarray(chains)]
    out = np.empty(len(spans))
    for k, (a, b) in enumerate(spans):
        x = xyz[a:b]
        out[k] = np.sqrt(((x - x.mean(axis=0)) ** 2).sum(axis=1).mean())
    return out


class ReplicaEnsemble:
    """`n_global` replicas of `scenario`, of which this rank advances its shard."""

    def __init__(self, backend: bd.Backend, scenario, n_global: int, rank: int = 0, world: int = 1,
                 device: int = 0, seed: int = 0):
        self.backend, self.scenario = backend, scenario
        self.n_global, self.rank, self.world, self.seed = int(n_global), int(rank), int(world), int(seed)
        self.ids = shard_replicas(self.n_global, self.world, self.rank)
        self.steps_done = 0
        self.batched = backend.prefix == "bd_"
        self.systems: List[bd.System] = []
        start = [scenario.positions[int(r) % scenario.replicas] for r in self.ids]
        if self.batched:
            if len(self.ids):
                s = bd.System(backend, scenario.n, len(self.ids), device=device)
                scenario.apply(s, replicas=[])
                for k, x in enumerate(start):
                    s.set_positions(x, replica=k)
                s.set_option("replica_id_base", self.rank)
                s.set_option("replica_id_stride", self.world)
                self.systems.append(s)
        else:
            for x in start:
                s = bd.System(backend, scenario.n, 1, device=device)
                scenario.apply(s, replicas=[])
                s.set_positions(x)
                self.systems.append(s)

    @property
    def n_local(self) -> int:
        return len(self.ids)

    def step(self, steps: int, timestep: Optional[float] = None, temperature: Optional[float] = None):
        dt = self.scenario.timestep if timestep is None else timestep
        kT = self.scenario.temperature if temperature is None else temperature
        first = self.steps_done + 1
        if self.batched:
            for s in self.systems:
                s.step(steps, dt, kT, seed=self.seed, first_step=first)
        else:
            for s, r in zip(self.systems, self.ids):
                s.step(steps, dt, kT, seed=replica_seed(self.seed, int(r)), first_step=first)
        self.steps_done += steps

    def positions(self) -> np.ndarray:
        """[n_local, n, 3] in original bead order."""
        out = np.empty((self.n_local, self.scenario.n, 3))
        if self.batched:
            for k in range(self.n_local):
                out[k] = self.systems[0].get_positions(replica=k)
        else:
            for k, s in enumerate(self.systems):
                out[k] = s.get_positions()
        return out

    def local_statistics(self) -> np.ndarray:
        """[n_local, n_chains] radius of gyration per chain of every local replica."""
        x = self.positions()
        chains = self.scenario.chains
        width = 1 if chains is None else len(chains)
        out = np.empty((self.n_local, width))
        for k in range(self.n_local):
            out[k] = radius_of_gyration(x[k], chains)
        return out

    def gather(self, local: np.ndarray, device: str = "cpu") -> np.ndarray:
        """Per-replica rows of every rank, ordered by global replica id, on every rank.

        One all_gather of equal-sized (padded) shards over torch.distributed -- NCCL on the GPU box
        (`device="cuda"`), gloo in the CPU tests.  With world == 1 nothing is communicated."""
        local = np.asarray(local, dtype=np.float64).reshape(self.n_local, -1)
        width = local.shape[1]
        if self.world == 1:
            return local.copy()
        import torch
        import torch.distributed as dist

        per = -(-self.n_global // self.world)  # ceil: shard sizes differ by at most one
        pad = torch.zeros((per, width), dtype=torch.float64, device=device)
        pad[: self.n_local] = torch.from_numpy(local).to(device)
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        dist.all_gather(parts, pad)
        out = np.empty((self.n_global, width))
        for g, part in enumerate(parts):
            ids = shard_replicas(self.n_global, self.world, g)
            out[ids] = part[: len(ids)].cpu().numpy()
        return out

    def statistics(self, device: str = "cpu") -> np.ndarray:
        """[n_global, n_chains] on every rank."""
        return self.gather(self.local_statistics(), device)

    def close(self):
        for s in self.systems:
            s.close()
        self.systems = []


class HostPipeline:
    """A replica ensemble driven through HOST buffers every step -- the path of a driver whose per-step
    callback reads and may rewrite all positions (MD/simulation/brownian_dynamics.hpp:66-72) -- with the
    PCIe copies hidden behind the compute.

    The local replicas are split into `chunks` engines, each on its own CUDA stream.  One step is, per
    chunk, upload -> bd_step(1) -> download, all enqueued without waiting; the copy engines move chunk
    k + 1 in and chunk k - 1 out while chunk k computes (PCIe is full duplex, replicas are independent),
    and one synchronisation at the end of the step hands the host buffer back to the caller.  Noise keys
    are the global replica ids, so the trajectories equal those of the single-engine ensemble.

    host_in / host_out are pinned [n_local, n, 3] float64 buffers (torch tensors or anything with
    data_ptr()); they may be the same buffer."""

    def __init__(self, backend: bd.Backend, scenario, n_local: int, chunks: int = 8, rank: int = 0, world: int = 1,
                 device: int = 0, seed: int = 0, setup=None):
        if backend.prefix != "bd_":
            raise ValueError("HostPipeline drives the B200 engine")
        import torch

        self.scenario, self.n_local, self.seed = scenario, int(n_local), int(seed)
        chunks = max(1, min(int(chunks), self.n_local))
        per = -(-self.n_local // chunks)
        self.spans = [(a, min(a + per, self.n_local)) for a in range(0, self.n_local, per)]
        self.systems, self.streams, self.ids = [], [], []
        for a, b in self.spans:
            s = bd.System(backend, scenario.n, b - a, device=device)
            ids = scenario.apply(s, replicas=[])
            st = torch.cuda.Stream(device=device)
            s.set_stream(st.cuda_stream)
            s.set_option("replica_id_base", rank + world * a)
            s.set_option("replica_id_stride", world)
            if setup is not None:
                setup(s, ids)
            self.systems.append(s)
            self.streams.append(st)
            self.ids.append(ids)
        self.bytes_per_replica = scenario.n * 24
        self.steps_done = 0

    def upload(self, host):
        for s, (a, _) in zip(self.systems, self.spans):
            s.set_all_positions(host.data_ptr() + a * self.bytes_per_replica)

    def step_resident(self, steps: int, timestep: float, temperature: float):
        """steps without touching the host (relaxation, warm-up)"""
        for s in self.systems:
            s.step(steps, timestep, temperature, seed=self.seed, first_step=self.steps_done + 1)
        self.steps_done += steps

    def step_through_host(self, host_in, host_out, timestep: float, temperature: float, before_step=None):
        """one BD step of every replica: positions come from host_in and arrive in host_out"""
        first = self.steps_done + 1
        for s, ids, (a, _) in zip(self.systems, self.ids, self.spans):
            s.set_all_positions(host_in.data_ptr() + a * self.bytes_per_replica)
            if before_step is not None:
                before_step(s, ids)
            s.step(1, timestep, temperature, seed=self.seed, first_step=first)
            s.get_all_positions_async(host_out.data_ptr() + a * self.bytes_per_replica)
        for s in self.systems:
            s.sync()
        self.steps_done += 1

    def sync(self):
        for s in self.systems:
            s.sync()

    def launch_count(self) -> int:
        return sum(s.get_launch_count() for s in self.systems)

    def close(self):
        for s in self.systems:
            s.close()
        self.systems = []
