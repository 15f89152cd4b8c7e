"""Spatial slab decomposition of ONE large system over several GPUs (BASELINE.json
configs[4]: the 1kb chromatin model scaled to millions of beads).

The reference has no counterpart: micromd advances one system on one CPU thread.  Here the box is cut
into slabs along x, one per rank.  Every rank's engine holds all beads (6 M beads are 0.2 GB: memory
is not the constraint, the per-step force work is), but advances only the beads it OWNS -- those
inside its slab.  All other beads are ghosts on that rank (bd_set_owned): neighbours and bonded
partners whose positions arrive from their owner.  Per step each rank

    1. runs one fused force + integrate step over its owned beads (bd_step),
    2. packs the new positions of its owned beads that lie within `halo` of another rank's slab
       (bd_gather_positions_indexed), exchanges them point to point -- NCCL send/recv over NVLink on
       the GPU box -- and writes what it receives into its ghosts (bd_scatter_positions_indexed).

Everything is enqueued on one CUDA stream; the host never waits inside the step loop.  Ghosts
farther than `halo` from a slab are stale and harmless: they cannot get within interaction range of
an owned bead without first entering the halo, where they are refreshed.

Every `migrate_every` steps ownership follows the beads: a bead that crossed a slab boundary is
handed to the rank whose slab now contains it (both ranks hold its exact position, so they agree
without negotiation), and the send lists are rebuilt.  Between migrations a bead may drift by at most
`margin / 2`, which is checked.

Noise is Philox keyed by (seed, step, original bead index), so a bead draws the same noise whichever
rank owns it: the decomposed trajectory equals the single-GPU trajectory up to the summation order of
the forces.

`SlabLayout` and `SlabSystem` are backend-agnostic: the in-process `run_local` driver (several slabs
on one GPU) and the torch.distributed `DistExchange` share all of the logic; the CPU tests drive it
with a toy bead store over gloo.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch


class SlabLayout:
    """Slab boundaries along x.  Slab g covers [bounds[g], bounds[g+1]); with a periodic box x is wrapped
    into [0, period) first and distances along x are minimum-image distances."""

    def __init__(self, bounds: Sequence[float], period: Optional[float] = None):
        self.bounds = np.asarray(bounds, dtype=np.float64)
        self.world = len(self.bounds) - 1
        self.period = None if period is None else float(period)
        if self.world < 1 or not (np.diff(self.bounds) >= 0).all():
            raise ValueError("slab bounds must be ascending")
        if self.period is not None and not (self.bounds[0] == 0.0 and self.bounds[-1] == self.period):
            raise ValueError("periodic slabs must tile [0, period)")

    @staticmethod
    def balanced(x: np.ndarray, world: int, period: Optional[float] = None) -> "SlabLayout":
        """Equal bead counts per slab (quantiles of x)."""
        x = np.asarray(x, dtype=np.float64)
        if period is not None:
            x = x - np.floor(x / period) * period
        inner = np.quantile(x, np.arange(1, world) / world) if world > 1 else np.zeros(0)
        if period is not None:
            return SlabLayout(np.concatenate([[0.0], inner, [period]]), period)
        return SlabLayout(np.concatenate([[-np.inf], inner, [np.inf]]), None)

    def wrap(self, x: torch.Tensor) -> torch.Tensor:
        if self.period is None:
            return x
        return x - torch.floor(x / self.period) * self.period

    def owner_of(self, x: torch.Tensor) -> torch.Tensor:
        """Slab index of every coordinate."""
        inner = torch.as_tensor(self.bounds[1:-1], dtype=x.dtype, device=x.device)
        return torch.bucketize(self.wrap(x), inner, right=True)

    def distance_to(self, x: torch.Tensor, slab: int) -> torch.Tensor:
        """Distance along x from every coordinate to slab `slab` (0 inside it)."""
        lo, hi = float(self.bounds[slab]), float(self.bounds[slab + 1])
        xw = self.wrap(x)
        below = torch.clamp(lo - xw, min=0.0) if np.isfinite(lo) else torch.zeros_like(xw)
        above = torch.clamp(xw - hi, min=0.0) if np.isfinite(hi) else torch.zeros_like(xw)
        d = below + above
        if self.period is not None:
            # the slab's periodic images one period to the left and to the right
            d = torch.minimum(d, torch.clamp(lo + self.period - xw, min=0.0) + torch.clamp(xw - (hi + self.period), min=0.0))
            d = torch.minimum(d, torch.clamp(lo - self.period - xw, min=0.0) + torch.clamp(xw - (hi - self.period), min=0.0))
        return d


_streams: Dict[str, "torch.cuda.Stream"] = {}


def work_stream(device) -> "torch.cuda.Stream":
    """The one CUDA stream per device on which engines, torch operations and NCCL transfers of this module
    are ordered.  It becomes torch's current stream (the legacy default stream cannot be handed to the
    engine: bd_set_stream(NULL) means "the engine's own stream")."""
    key = str(torch.device(device))
    if key not in _streams:
        _streams[key] = torch.cuda.Stream(device=torch.device(device))
    torch.cuda.set_stream(_streams[key])
    return _streams[key]


class EngineStore:
    """The beads of one rank in a B200 engine (bdb200.System with one replica)."""

    def __init__(self, system, device: str):
        self.system, self.device, self.n = system, torch.device(device), system.n
        torch.cuda.set_device(self.device)
        self.stream = work_stream(self.device)
        system.set_stream(self.stream.cuda_stream)
        self._xyz = torch.empty((self.n, 3), dtype=torch.float64, device=self.device)

    def positions(self) -> torch.Tensor:
        self.system.get_positions_device(self._xyz.data_ptr())
        return self._xyz

    def gather(self, index: torch.Tensor) -> torch.Tensor:
        out = torch.empty((len(index), 3), dtype=torch.float64, device=self.device)
        self.system.gather_positions_indexed(index.data_ptr(), len(index), out.data_ptr())
        return out

    def scatter(self, index: torch.Tensor, xyz: torch.Tensor):
        assert xyz.is_contiguous() and xyz.shape == (len(index), 3)
        self.system.scatter_positions_indexed(index.data_ptr(), len(index), xyz.data_ptr())

    def set_owned(self, owned: torch.Tensor):
        self._owned = owned.to(torch.uint8).contiguous()  # kept alive until the kernel has read it
        self.system.set_owned_device(self._owned.data_ptr())

    def step(self, first_step: int, timestep: float, temperature: float, seed: int):
        self.system.step(1, timestep, temperature, seed=seed, first_step=first_step)


class SlabSystem:
    """One rank's share of a slab-decomposed system.  The methods come in send / receive halves so that
    the same logic runs under any transport (`run_local`, `DistExchange`)."""

    def __init__(self, store, layout: SlabLayout, rank: int, halo: float, margin: float, migrate_every: int = 10,
                 partners: Optional[np.ndarray] = None):
        """`partners` [m, 2]: bead pairs tied by topology (bonds, loops, glue, the three pairs of a bending
        triple).  They may be far apart in space, so the halo alone does not cover them: the owner of either
        bead always receives the other."""
        if not (0 < margin < halo):
            raise ValueError("need 0 < margin < halo")
        self.store, self.layout, self.rank = store, layout, int(rank)
        self.halo, self.margin, self.migrate_every = float(halo), float(margin), int(migrate_every)
        self.n, self.device = store.n, store.device
        self.owned = torch.zeros(self.n, dtype=torch.bool, device=self.device)
        self.accurate = torch.ones(self.n, dtype=torch.bool, device=self.device)  # before the first migration
        self.send_index: Dict[int, torch.Tensor] = {}
        self.recv_index: Dict[int, torch.Tensor] = {}
        self.anchor: Optional[torch.Tensor] = None  # positions at the last migration (drift check)
        self.steps_done = 0
        self.halo_bytes_sent = 0
        if partners is None or len(partners) == 0:
            self.pair_a = self.pair_b = torch.zeros(0, dtype=torch.long, device=self.device)
        else:
            p = np.asarray(partners, dtype=np.int64).reshape(-1, 2)
            both = np.concatenate([p, p[:, ::-1]])
            self.pair_a = torch.from_numpy(np.ascontiguousarray(both[:, 0])).to(self.device)
            self.pair_b = torch.from_numpy(np.ascontiguousarray(both[:, 1])).to(self.device)

    # -- ownership ---------------------------------------------------------------------------------
    def migrate_plan(self) -> Dict[int, torch.Tensor]:
        """New ownership from the positions this rank knows exactly; returns, per peer, the indices of the
        owned beads whose positions that peer needs from now on."""
        xyz = self.store.positions()
        x = xyz[:, 0]
        if self.anchor is not None:
            d = xyz - self.anchor
            if self.layout.period is not None:
                d[:, 0] = d[:, 0] - torch.round(d[:, 0] / self.layout.period) * self.layout.period
            # over the beads this rank advanced (every bead is advanced by exactly one rank)
            drift = torch.sqrt((d * d).sum(dim=1))[self.owned].max().item() if bool(self.owned.any()) else 0.0
            if drift > 0.5 * self.margin:
                raise RuntimeError(f"slab: a bead drifted {drift:.3g} between migrations (margin/2 = {0.5 * self.margin:.3g}); "
                                   "migrate more often or widen the margin")
        owner = self.layout.owner_of(x)
        self.owned = self.accurate & (owner == self.rank)
        self.send_index = {}
        # topological partners: my bead a goes to whoever owns its partner b (b is exact here by the same
        # rule applied on b's owner at the previous migration; everything is exact before the first one)
        mine = self.owned[self.pair_a]
        for peer in range(self.layout.world):
            if peer == self.rank:
                continue
            sel = self.owned & (self.layout.distance_to(x, peer) < self.halo)
            if len(self.pair_a):
                sel[self.pair_a[mine & (owner[self.pair_b] == peer)]] = True
            idx = torch.nonzero(sel, as_tuple=False).flatten().to(torch.int32)
            if len(idx):
                self.send_index[peer] = idx
        self.anchor = xyz.clone()
        return self.send_index

    def migrate_commit(self, received: Dict[int, torch.Tensor]):
        self.recv_index = {p: t for p, t in received.items() if len(t)}
        self.accurate = self.owned.clone()
        for t in self.recv_index.values():
            self.accurate[t.long()] = True
        self.store.set_owned(self.owned)
        # the per-step exchange as ONE gather, ONE collective and ONE scatter (DistExchange.halo): everything
        # this rank sends, concatenated by peer, and everything it receives, with the row counts per peer
        world = self.layout.world
        empty = torch.zeros(0, dtype=torch.int32, device=self.device)
        self.send_cat = torch.cat([self.send_index.get(p, empty) for p in range(world)]).contiguous()
        self.recv_cat = torch.cat([self.recv_index.get(p, empty) for p in range(world)]).contiguous()
        self.send_rows = [len(self.send_index.get(p, empty)) for p in range(world)]
        self.recv_rows = [len(self.recv_index.get(p, empty)) for p in range(world)]
        self.send_buf = torch.empty((len(self.send_cat), 3), dtype=torch.float64, device=self.device)
        self.recv_buf = torch.empty((len(self.recv_cat), 3), dtype=torch.float64, device=self.device)

    # -- halo --------------------------------------------------------------------------------------
    def halo_pack(self) -> Dict[int, torch.Tensor]:
        out = {p: self.store.gather(idx) for p, idx in self.send_index.items()}
        self.halo_bytes_sent += sum(t.numel() * 8 for t in out.values())
        return out

    def halo_unpack(self, received: Dict[int, torch.Tensor]):
        for p, idx in self.recv_index.items():
            self.store.scatter(idx, received[p])

    def advance(self, timestep: float, temperature: float, seed: int):
        self.steps_done += 1
        self.store.step(self.steps_done, timestep, temperature, seed)

    def owned_positions(self):
        """(indices, xyz) of the beads this rank owns, as numpy arrays."""
        idx = torch.nonzero(self.owned, as_tuple=False).flatten()
        return idx.cpu().numpy(), self.store.positions()[idx].cpu().numpy()


def topology_partners(scenario) -> np.ndarray:
    """[m, 2] bead pairs of a scenario's bonded and bending force fields (scenarios.Scenario)."""
    out = [np.zeros((0, 2), dtype=np.int64)]
    for ff in scenario.bond_ffs:
        out.append(np.asarray(ff["pairs"], dtype=np.int64).reshape(-1, 2))
    for ff in scenario.bend_ffs:
        t = np.asarray(ff["triples"], dtype=np.int64).reshape(-1, 3)
        out += [t[:, [0, 1]], t[:, [1, 2]], t[:, [0, 2]]]
    return np.unique(np.concatenate(out), axis=0)


def run_local(slabs: List[SlabSystem], steps: int, timestep: float, temperature: float, seed: int):
    """Several slabs in one process (one GPU, or the CPU toy store): the transport is a dictionary."""

    def deliver(outboxes):
        inboxes = [dict() for _ in slabs]
        for src, box in enumerate(outboxes):
            for dst, payload in box.items():
                inboxes[dst][src] = payload
        return inboxes

    for _ in range(steps):
        if slabs[0].steps_done % slabs[0].migrate_every == 0:
            for s, inbox in zip(slabs, deliver([s.migrate_plan() for s in slabs])):
                s.migrate_commit(inbox)
            total = sum(int(s.owned.sum()) for s in slabs)
            if total != slabs[0].n:
                raise RuntimeError(f"slab: {total} of {slabs[0].n} beads are owned after a migration")
            for s, inbox in zip(slabs, deliver([s.halo_pack() for s in slabs])):
                s.halo_unpack(inbox)
        for s in slabs:
            s.advance(timestep, temperature, seed)
        for s, inbox in zip(slabs, deliver([s.halo_pack() for s in slabs])):
            s.halo_unpack(inbox)


class DistExchange:
    """torch.distributed transport: NCCL send/recv between GPUs (gloo with CPU tensors in the tests)."""

    def __init__(self, device):
        import torch.distributed as dist

        self.dist, self.device = dist, torch.device(device)
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def _p2p(self, send: Dict[int, torch.Tensor], recv: Dict[int, torch.Tensor]):
        ops = []
        for peer in sorted(set(send) | set(recv)):
            # both ends order their operations by peer and lower rank sends first: no deadlock
            first_send = self.rank < peer
            pair = []
            if peer in send:
                pair.append(self.dist.P2POp(self.dist.isend, send[peer], peer))
            if peer in recv:
                pair.append(self.dist.P2POp(self.dist.irecv, recv[peer], peer))
            ops.extend(pair if first_send else pair[::-1])
        if ops:
            for req in self.dist.batch_isend_irecv(ops):
                req.wait()

    def indices(self, send: Dict[int, torch.Tensor]) -> Dict[int, torch.Tensor]:
        counts = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        for p, t in send.items():
            counts[p] = len(t)
        flat = torch.empty(self.world * self.world, dtype=torch.int64, device=self.device)
        self.dist.all_gather_into_tensor(flat, counts)
        table = flat.view(self.world, self.world)
        incoming = table[:, self.rank].cpu().tolist()  # one host read per migration
        recv = {p: torch.empty(c, dtype=torch.int32, device=self.device) for p, c in enumerate(incoming) if c and p != self.rank}
        self._p2p(send, recv)
        return recv

    def positions(self, send: Dict[int, torch.Tensor], recv_index: Dict[int, torch.Tensor]) -> Dict[int, torch.Tensor]:
        recv = {p: torch.empty((len(t), 3), dtype=torch.float64, device=self.device) for p, t in recv_index.items()}
        self._p2p(send, recv)
        return recv

    def halo(self, slab: "SlabSystem"):
        """The per-step ghost refresh.  On NCCL: one indexed gather of everything this rank sends (engine kernel),
        one all-to-all with per-peer row counts, one indexed scatter into the ghosts (engine kernel) -- three
        enqueues on the work stream whatever the number of peers, no host wait.  Other backends (gloo in the CPU
        tests has no all-to-all) keep the point-to-point form."""
        if self.dist.get_backend() != "nccl" or not hasattr(slab.store, "system"):
            slab.halo_unpack(self.positions(slab.halo_pack(), slab.recv_index))
            return
        system = slab.store.system
        if len(slab.send_cat):
            system.gather_positions_indexed(slab.send_cat.data_ptr(), len(slab.send_cat), slab.send_buf.data_ptr())
        self.dist.all_to_all_single(slab.recv_buf, slab.send_buf, output_split_sizes=slab.recv_rows,
                                    input_split_sizes=slab.send_rows)
        if len(slab.recv_cat):
            system.scatter_positions_indexed(slab.recv_cat.data_ptr(), len(slab.recv_cat), slab.recv_buf.data_ptr())
        slab.halo_bytes_sent += slab.send_buf.numel() * 8
        # one list-rebuild decision for the whole system, as in the undecomposed run: the largest displacement since
        # the last build over ALL ranks.  Without it every rank rebuilds on its own schedule and each rebuild (O(N))
        # stalls everybody's next exchange.
        if not hasattr(slab, "_moved"):
            slab._moved = _device_int64(system.displacement_pointer(), self.device)
            # the all-reduce must act on the engine's own word; a silent copy would leave the ranks rebuilding on
            # their own schedules with no error
            if slab._moved.data_ptr() != system.displacement_pointer():
                raise RuntimeError("slab: the displacement tensor does not alias the engine's memory")
        self.dist.all_reduce(slab._moved, op=self.dist.ReduceOp.MAX)


class _DevicePointer:
    def __init__(self, ptr: int):
        self.__cuda_array_interface__ = {"shape": (1,), "typestr": "<i8", "data": (int(ptr), False), "version": 2}


def _device_int64(ptr: int, device) -> torch.Tensor:
    """a one-element int64 tensor aliasing engine memory at `ptr`"""
    return torch.as_tensor(_DevicePointer(ptr), device=device)


def run_distributed(slab: SlabSystem, exchange: DistExchange, steps: int, timestep: float, temperature: float, seed: int):
    """One rank of the decomposed run; every rank calls this with the same arguments."""
    import os
    import time
    prof = os.environ.get("SLAB_PROFILE") and slab.device.type == "cuda"

    def lap(t0, tag):
        if not prof:
            return t0
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if slab.rank == 0:
            print(f"slab migration {tag}: {1e3 * (t1 - t0):.2f} ms", flush=True)
        return t1

    for _ in range(steps):
        if slab.steps_done % slab.migrate_every == 0:
            t = lap(time.perf_counter(), "start") if prof else 0.0
            plan = slab.migrate_plan()
            t = lap(t, "plan")
            recv = exchange.indices(plan)
            t = lap(t, "index exchange")
            slab.migrate_commit(recv)
            t = lap(t, "commit")
            owned = slab.owned.sum().to(torch.int64).reshape(1)
            exchange.dist.all_reduce(owned)
            if int(owned.item()) != slab.n:
                raise RuntimeError(f"slab: {int(owned.item())} of {slab.n} beads are owned after a migration")
            exchange.halo(slab)
            t = lap(t, "check + halo")
        slab.advance(timestep, temperature, seed)
        exchange.halo(slab)


def gather_owned(slab: SlabSystem, exchange: Optional[DistExchange] = None) -> np.ndarray:
    """All positions [n, 3] assembled from every rank's owned beads (validation, snapshots)."""
    idx, xyz = slab.owned_positions()
    out = np.full((slab.n, 3), np.nan)
    out[idx] = xyz
    if exchange is None or exchange.world == 1:
        return out
    t = torch.from_numpy(np.nan_to_num(out, nan=0.0)).to(exchange.device)
    m = torch.from_numpy(~np.isnan(out[:, 0])).to(exchange.device).to(torch.float64)
    exchange.dist.all_reduce(t)
    exchange.dist.all_reduce(m)
    if not bool((m == 1).all()):
        raise RuntimeError("slab: ownership is not a partition")
    return t.cpu().numpy()


class ToyStore:
    """TEST INFRASTRUCTURE: a bead store on the CPU with a prescribed motion -- every owned bead moves by a
    displacement that depends only on (bead index, step) -- so that the ownership and halo logic can be
    checked against a global replay without a GPU."""

    def __init__(self, xyz: np.ndarray, amplitude: float):
        self.xyz = torch.from_numpy(np.array(xyz, dtype=np.float64))
        self.n, self.device, self.amplitude = len(xyz), torch.device("cpu"), float(amplitude)
        self.owned = torch.ones(self.n, dtype=torch.bool)

    @staticmethod
    def displacement(n: int, step: int, amplitude: float) -> torch.Tensor:
        i = torch.arange(n, dtype=torch.float64)
        return amplitude * torch.stack([torch.sin(0.37 * i + 0.11 * step) + 0.6, torch.cos(0.23 * i - 0.05 * step),
                                        torch.sin(0.71 * i + 0.3 * step)], dim=1)

    def positions(self):
        return self.xyz

    def gather(self, index):
        return self.xyz[index.long()].clone()

    def scatter(self, index, xyz):
        self.xyz[index.long()] = xyz

    def set_owned(self, owned):
        self.owned = owned.clone()

    def step(self, first_step, timestep, temperature, seed):
        d = self.displacement(self.n, first_step, self.amplitude)
        self.xyz[self.owned] += d[self.owned]
