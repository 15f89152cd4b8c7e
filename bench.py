#!/usr/bin/env python
"""Benchmark of the Brownian-dynamics hot path (force evaluation + BAOAB-limit update).

    python bench.py [--gpus N] [--steps K] [--warmup W]            the B200 engine
    python bench.py --impl reference [...]                          the reference CPU code (oracle/_ref)
    torchrun --nproc-per-node N ... bench.py --gpus N ...           one rank per GPU (weak scaling)

Workload (BASELINE.json configs[3], the one north_star's roofline target is quoted on): the
5-sim-genome 100kb interphase model -- 62 000 chromatin beads in 46 chains + 600 nucleolar
side-beads, A/B/u classes, mixed softcore<2,3>+softcore<8,3> non-bonded pairs, three bonded
forcefields, inward + outward ellipsoid walls -- as an ensemble of R independent replicas per GPU in
one launch (a single 2.5 MB system sits in L2 and is launch-bound).  A "step"
is one BD step of every replica.  Inputs are synthetic (scenarios.py, fixed seeds).

One JSON line on stdout (rank 0).  `value` is device-resident throughput; `e2e` goes through the
C ABI with pinned HOST buffers every step (upload all positions, one step, download all positions).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "2022a-genome-dynamics_b200"))

import numpy as np  # noqa: E402

ALGO_BYTES_PER_BEAD_STEP = 64  # fp64: 32 B record read + 32 B record written


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--replicas", type=int, default=None, help="replicas per GPU (default: per workload)")
    ap.add_argument("--beads", type=int, default=62000, help="genome100kb: chromatin beads per replica")
    ap.add_argument("--relax", type=int, default=400, help="untimed relaxation steps before warm-up")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-mixed", action="store_true", help="skip the second, separately labelled mixed-precision line")
    ap.add_argument("--no-parity", action="store_true", help="skip the force / pair-set check against the reference")
    ap.add_argument("--chunks", type=int, default=8, help="e2e: engines (replica chunks) of the host pipeline")
    ap.add_argument("--frozen-walls", action="store_true",
                    help="genome100kb: keep the wall semiaxes fixed (round-1 workload) instead of the driver's per-step feedback")
    ap.add_argument("--verlet-factor", type=float, default=None, help="list radius / dcut (engine option; reference 1.5)")
    ap.add_argument("--step-graph", type=int, default=None, help="engine option step_graph (0: launch kernel by kernel, "
                    "e.g. under ncu)")
    ap.add_argument("--staged", type=int, default=None, help="engine option staged (tile staging of the pair walk) after the relaxation")
    ap.add_argument("--stage-capacity", type=int, default=None, help="engine option stage_capacity")
    ap.add_argument("--pair-precision", type=int, default=None, choices=[32, 64],
                    help="engine option pair_precision for the MAIN line (profiling the mixed lane; the default line is 64)")
    ap.add_argument("--workload", default="genome100kb", choices=sorted(WORKLOADS) + ["slab1kb"],
                    help="genome100kb: the headline replica ensemble (weak scaling); ab_box / ab_sphere / genome10mb / "
                         "chromatin1kb: the other BASELINE.json configurations on one engine per GPU; slab1kb: ONE 1kb "
                         "chromatin system split into x-slabs with a per-step NCCL halo exchange (strong scaling)")
    ap.add_argument("--chains", type=int, default=600, help="chromatin1kb / slab1kb: chains")
    ap.add_argument("--chain-len", type=int, default=10000, help="chromatin1kb / slab1kb: beads per chain")
    ap.add_argument("--migrate-every", type=int, default=80, help="slab1kb: steps between ownership migrations (the drift "
                    "bound that makes the interval safe is checked at every migration)")
    ap.add_argument("--no-slab", action="store_true", help="N > 1: skip the attached slab1kb pass")
    ap.add_argument("--slab-steps", type=int, default=160, help="N > 1: steps of the attached slab1kb pass")
    return ap.parse_args()


# BASELINE.json configs[0..4] as bench workloads: builder(args, distinct, seed), replicas per GPU, label
def _w_genome100kb(args, distinct, seed):
    import scenarios as sc
    return sc.genome_100kb(total=args.beads, n_nor=300, seed=seed, replicas=distinct)


def _w_ab_box(args, distinct, seed):
    import scenarios as sc
    return sc.ab_box(200, 20, seed=seed, replicas=distinct)


def _w_ab_sphere(args, distinct, seed):
    import scenarios as sc
    return sc.ab_sphere(200, 20, seed=seed, replicas=distinct)


def _w_genome10mb(args, distinct, seed):
    import scenarios as sc
    return sc.genome_10mb(640, seed=seed, replicas=distinct, bend_energy=0.7)


def _w_chromatin1kb(args, distinct, seed):
    import scenarios as sc
    return sc.chromatin_1kb(args.chains, args.chain_len, seed=seed, replicas=1)


WORKLOADS = {
    "genome100kb": dict(make=_w_genome100kb, replicas=64, label="5-sim-genome 100kb interphase"),
    "ab_box": dict(make=_w_ab_box, replicas=1024, label="4-sim-ab/box A/B blend, periodic"),
    "ab_sphere": dict(make=_w_ab_sphere, replicas=1024, label="4-sim-ab/sphere A/B blend in a spherical wall"),
    "genome10mb": dict(make=_w_genome10mb, replicas=1024, label="5-sim-genome 10Mb ana/telophase"),
    "chromatin1kb": dict(make=_w_chromatin1kb, replicas=1, label="3-sim-1kb chromatin, one periodic system"),
}


def scenario(args, distinct, seed):
    return WORKLOADS[args.workload]["make"](args, distinct, seed)


def replicas_of(args):
    return args.replicas if args.replicas else WORKLOADS[args.workload]["replicas"]


def wall_feedback(args, scen):
    """the interphase driver moves the wall semiaxes after every step from the wall reaction of that system
    (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:70-80); spring and mobility are the
    reference's defaults (config_entries.inc:33,37)"""
    if args.workload != "genome100kb" or args.frozen_walls:
        return None
    p = scen.params
    return dict(spring=np.full(3, p["wall_semiaxes_spring"]), mobility=p["wall_mobility"])


def workload_name(args, scen):
    tail = ", per-step wall feedback" if wall_feedback(args, scen) else ""
    return f"{WORKLOADS[args.workload]['label']}, {scen.n} beads x {replicas_of(args)} replicas per GPU{tail}"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        out, _ = self.proc.communicate(timeout=10)
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- reference arm -------------------------------------------------------------------------------

def cpu_backend():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import checkers  # test infrastructure: only this CPU leg of the bench executes anything under oracle/

    if checkers.have_reference():
        return checkers.load_reference(), "reference"
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
    return checkers.load_oracle(), "port"


class _GenericReplica:
    """one replica on the generic C entry points (oracle/_ref/libmicromd_ref.so or the oracle port); with
    `feedback` it runs the interphase driver's per-step callback: read both walls' reaction, move the
    semiaxes, hand them back to both walls"""

    def __init__(self, backend, scen, x, feedback):
        import bdb200 as bd

        self.s, self.scen, self.feedback = bd.System(backend, scen.n), scen, feedback
        ids = scen.apply(self.s, replicas=[0])
        self.s.set_positions(x)
        self.walls = ids["wall"]
        self.semi = np.array(scen.params["semiaxes"], dtype=np.float64) if feedback else None
        self.done = 0

    def run(self, k):
        scen, s = self.scen, self.s
        if k <= 0:
            return
        if not self.feedback:
            s.step(k, scen.timestep, scen.temperature, seed=1, first_step=self.done + 1)
        else:
            for q in range(k):
                s.step(1, scen.timestep, scen.temperature, seed=1, first_step=self.done + q + 1)
                net = (s.get_wall_stats(self.walls[0])[1:] + s.get_wall_stats(self.walls[1])[1:]
                       - self.feedback["spring"] * self.semi)
                self.semi = self.semi + scen.timestep * self.feedback["mobility"] * net
                for w in self.walls:
                    s.update_wall_ff(w, (0, 0, 0), self.semi)
        self.done += k

    def positions(self):
        return self.s.get_positions()

    def close(self):
        self.s.close()


class _DriverReplica:
    """one replica of the 100kb model registered the way the reference driver registers it -- compile-time
    expression-template potentials, the wall feedback as the simulate_brownian_dynamics callback
    (oracle/ref_drivers.cc, oracle/checkers.DriverC4)"""

    def __init__(self, scen, x, feedback, flavour):
        import checkers

        self.d, self.feedback, self.calls = checkers.DriverC4(scen, x, flavour), feedback, 0

    def run(self, k):
        if k > 0:
            self.calls += 1
            self.d.run(k, seed=self.calls, feedback=bool(self.feedback))

    def positions(self):
        return self.d.get()[0]

    def close(self):
        self.d.close()


def cpu_flavours(args, backend_kind):
    """the CPU builds to time: the drivers' own code path where oracle/_ref has it (genome100kb), else the
    generic entry points"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import checkers

    if args.workload == "genome100kb" and backend_kind == "reference" and checkers.have_driver_baseline("O3"):
        return ["driver-O3", "driver-O2", "generic"]
    return ["generic"]


def cpu_time_steps(backend, scen, start_positions, steps, threads, warm, feedback=None, flavour="generic"):
    """`threads` independent replicas, each advanced `steps` BD steps by the single-threaded reference code
    after `warm` untimed ones; returns (wall seconds, final positions of replica 0) (ctypes releases the
    GIL, so the replicas run in parallel)."""
    reps = []
    for t in range(threads):
        x = start_positions[t % len(start_positions)]
        if flavour.startswith("driver-"):
            reps.append(_DriverReplica(scen, x, feedback, flavour.split("-")[1]))
        else:
            reps.append(_GenericReplica(backend, scen, x, feedback))

    def fan(k):
        th = [threading.Thread(target=r.run, args=(k,)) for r in reps]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        return time.perf_counter() - t0

    fan(warm)
    dt = fan(steps)
    final = reps[0].positions()
    for r in reps:
        r.close()
    return dt, final


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    backend, kind = cpu_backend()
    threads = os.cpu_count() or 1
    scen = scenario(args, min(threads, 4), 4)
    feedback = wall_feedback(args, scen)
    K, W = max(args.steps, 1), max(args.warmup, 0)
    # size the per-step sample from a short probe so that the whole run stays within a few minutes; the
    # relaxation (the same number of steps the engine arm relaxes for) is done by the reference itself
    flavour = cpu_flavours(args, kind)[0]
    probe = cpu_time_steps(backend, scen, scen.positions, 3, threads, 1, feedback, flavour)[0] / 3
    relax = min(args.relax, max(10, int(60.0 / max(probe, 1e-6))))
    inner = max(1, int(100.0 / max(probe, 1e-6) / (K + W)))
    inner = min(inner, 50)
    wall, _ = cpu_time_steps(backend, scen, scen.positions, K * inner, threads, relax + W * inner, feedback, flavour)
    units = threads * scen.n * inner * K
    value = units / wall
    line = {
        "impl": "reference", "metric": "bead_steps_per_s", "value": value, "unit": "bead-steps/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": 1e3 * wall / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, scen),
                   "sample": f"{threads} replicas (one per host thread), {inner} BD steps of each per bench step, "
                             f"after {relax} relaxation steps by the reference itself"},
        "cpu_baseline": {"value": value, "unit": "bead-steps/s", "cores": threads, "kind": kind, "build": flavour,
                         "sample": f"{threads} independent replicas x {inner * K} BD steps, micromd single-threaded per replica"
                                   + ((", per-step wall feedback " + ("as the simulate_brownian_dynamics callback" if flavour.startswith("driver")
                                                                   else "through the C entry points")) if feedback else "")},
        "e2e": {"value": value, "unit": "bead-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- B200 arm ------------------------------------------------------------------------------------

def bind_near_gpu(torch, local):
    """Pin this rank to the CPUs of the NUMA node its GPU hangs off (sysfs local_cpulist of the PCI device), so
    that the rank's pinned host buffers are first-touched there and its copies do not cross the socket
    interconnect.  Best effort: returns what was found."""
    out = {"bound": False}
    try:
        p = torch.cuda.get_device_properties(local)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bus}"
        out["pci"] = bus
        out["numa_node"] = int(open(f"{base}/numa_node").read())
        cpus = open(f"{base}/local_cpulist").read().strip()
        out["local_cpulist"] = cpus
        ids = set()
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            ids.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = ids & allowed
        if use and use != allowed:
            os.sched_setaffinity(0, use)
            out["bound"] = True
    except Exception as e:  # sysfs not exposed (containers): keep the inherited affinity
        out["note"] = f"{type(e).__name__}: {e}"
    return out


def parity_check(eng, sysm, scen, R, ids, tol=1e-6):
    """Replica 0 and replica R-1 of the state about to be timed against the reference's own code (the
    unmodified micromd where oracle/_ref was built, else the CPU restatement): per-bead forces, and the
    Verlet pair set of a fresh list build on the same positions.  The checker is used as a checker only."""
    import bdb200 as bd

    backend, kind = cpu_backend()
    out = {"checker": kind, "replicas": sorted({0, R - 1}), "force_rel_err": 0.0, "pairs_equal": True}
    for r in out["replicas"]:
        x = sysm.get_positions(r)
        ref = bd.System(backend, scen.n)
        rids = scen.apply(ref, replicas=[0])
        ref.set_positions(x)
        for w_ref, w_eng in zip(rids["wall"], ids["wall"]):  # this replica's walls where the feedback has moved them
            ref.update_wall_ff(w_ref, *sysm.get_wall_geometry(w_eng, r))
        fr = ref.compute_forces()
        fe = sysm.compute_forces(r)
        scale = np.linalg.norm(fr, axis=1).max()
        out["force_rel_err"] = max(out["force_rel_err"], float(np.linalg.norm(fe - fr, axis=1).max() / scale))
        fresh = bd.System(eng, scen.n, 1, device=sysm.device)
        scen.apply(fresh, replicas=[0])
        fresh.set_positions(x)
        a, b = fresh.get_neighbor_list(), ref.get_neighbor_list()
        key = lambda p: np.sort((np.minimum(p[:, 0], p[:, 1]).astype(np.uint64) << np.uint64(32))
                                | np.maximum(p[:, 0], p[:, 1]).astype(np.uint64))
        out["pairs_equal"] = bool(out["pairs_equal"] and np.array_equal(key(a), key(b)))
        out["listed_pairs"] = int(len(b))
        fresh.close()
        ref.close()
    if not (out["force_rel_err"] <= tol and out["pairs_equal"]):
        raise SystemExit(f"bench.py: parity check failed: {out}")
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist

    import bdb200 as bd
    import ensemble as en

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(torch, local)  # before any pinned allocation: first touch then lands on the GPU's node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    eng = bd.load_engine()
    R = replicas_of(args)
    distinct = min(R, 4)
    scen = scenario(args, distinct, 4 + rank)  # every rank simulates its own replicas (weak scaling)
    n, ntot = scen.n, scen.n * R
    feedback = wall_feedback(args, scen)
    sysm = bd.System(eng, n, R, device=local)
    ids = scen.apply(sysm, replicas=[])
    sysm.device = local
    if feedback:
        sysm.set_wall_relaxation(ids["wall"], feedback["spring"], feedback["mobility"])
    if args.verlet_factor:
        sysm.set_option("verlet_factor", args.verlet_factor)
    if args.step_graph is not None:
        sysm.set_option("step_graph", args.step_graph)
    stream = torch.cuda.Stream()
    sysm.set_stream(stream.cuda_stream)
    host = torch.empty((R, n, 3), dtype=torch.float64).pin_memory()
    host.numpy()[:] = scen.positions[np.arange(R) % distinct]
    sysm.set_all_positions(host.data_ptr())
    dt, kT = scen.timestep, scen.temperature

    step_no = [1]

    def advance(k):
        sysm.step(k, dt, kT, seed=1234 + rank, first_step=step_no[0])
        step_no[0] += k

    if args.staged is not None:
        sysm.set_option("staged", 0)
    advance(args.relax)
    if args.staged is not None:
        sysm.sync()
        sysm.set_option("staged", args.staged)
    if args.stage_capacity is not None:
        sysm.set_option("stage_capacity", args.stage_capacity)
    if args.pair_precision is not None:
        sysm.set_option("pair_precision", args.pair_precision)
    advance(max(args.warmup, 3))
    sysm.sync()
    if args.staged:
        print("max_stage", sysm.get_option("max_stage"), "stage_capacity", sysm.get_option("stage_capacity"),
              "staged", sysm.get_option("staged"), file=sys.stderr)
    parity = None
    if rank == 0 and not args.no_parity:
        parity = parity_check(eng, sysm, scen, R, ids, tol=5e-5 if args.pair_precision == 32 else 1e-6)
    c0 = sysm.get_counters()
    l0 = sysm.get_launch_count()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    with torch.cuda.stream(stream):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        advance(args.steps)
        e1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    launches = sysm.get_launch_count() - l0
    c1 = sysm.get_counters()
    inter = c1.pair_interactions - c0.pair_interactions
    rebuilds = c1.list_rebuilds - c0.list_rebuilds
    semi_now = sysm.get_wall_geometry(ids["wall"][0], 0)[1].tolist() if feedback else None

    # the dominant kernel alone, CUDA events around every launch
    sysm.set_option("time_kernel", 1)
    advance(min(args.steps, 50))
    kernel_ms = sysm.get_option("kernel_ms_avg")
    sysm.set_option("time_kernel", 0)
    # a second, separately labelled line: the mixed-precision lane (pair displacements from the single-precision
    # shadows, everything else fp64), same state, same steps; its force error against the reference is measured here
    mixed = None
    if not args.no_mixed and sysm.get_option("pair_program") != 0:
        sysm.set_option("pair_precision", 32)
        advance(5)
        sysm.sync()
        with torch.cuda.stream(stream):
            m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            m0.record(stream)
            advance(args.steps)
            m1.record(stream)
        barrier()
        mixed_ms = m0.elapsed_time(m1)
        sysm.set_option("time_kernel", 1)
        advance(min(args.steps, 50))
        mixed_kernel = sysm.get_option("kernel_ms_avg")
        sysm.set_option("time_kernel", 0)
        mixed = {"dtype": "f64 state and accumulation, f32 pair displacements (bd_set_option pair_precision=32)",
                 "value": ntot * args.steps / (mixed_ms * 1e-3), "unit": "bead-steps/s (this rank)",
                 "ms_per_step": mixed_ms / args.steps, "kernel_ms": mixed_kernel}
        if rank == 0 and not args.no_parity:
            mixed["force_rel_err_vs_reference"] = parity_check(eng, sysm, scen, R, ids, tol=1e-3)["force_rel_err"]
        sysm.set_option("pair_precision", 64)
    sysm.get_all_positions(host.data_ptr())
    sysm.close()

    # end to end through the C ABI with pinned HOST buffers every step: the replicas as `chunks` engines on
    # their own streams, upload -> bd_step(1) -> download per chunk, one synchronisation per step
    e2e = None
    if not args.no_e2e:
        ke = max(3, min(args.steps, 20))

        def setup(s, fids):
            if feedback:
                s.set_wall_relaxation(fids["wall"], feedback["spring"], feedback["mobility"])
            if args.verlet_factor:
                s.set_option("verlet_factor", args.verlet_factor)

        pipe = en.HostPipeline(eng, scen, R, chunks=args.chunks, rank=rank, world=world, device=local, seed=1234 + rank,
                               setup=setup)
        pipe.steps_done = step_no[0] - 1
        pipe.upload(host)
        pipe.step_resident(3, dt, kT)  # builds the lists; the timed steps below start from host memory
        for s in pipe.systems:
            s.sync()
        le0 = pipe.launch_count()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(ke):
            pipe.step_through_host(host, host, dt, kT)
        f1.record()
        barrier()
        e2e = (ke, f0.elapsed_time(f1), (pipe.launch_count() - le0) / ke, len(pipe.systems))
        pipe.close()

    t = torch.tensor([ms, e2e[1] if e2e else 0.0, float(inter), kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, e2e_ms_max, kernel_ms = tmax[0].item(), tmax[1].item(), tmax[3].item()
        inter = tsum[2].item()
    else:
        e2e_ms_max = t[1].item()

    slab_line = None
    if world > 1 and args.workload == "genome100kb" and not args.no_slab:
        # the one workload with a data-path collective, so that the scaling record carries it too: a short pass of
        # the slab-decomposed 1kb chromatin system on the same ranks (strong scaling; its own N = 1 figure is in
        # profiles/)
        import copy
        sargs = copy.copy(args)
        # warm-up runs past the second migration: first-use costs (NCCL connections, allocator growth) stay out of
        # the timed region, which then holds steps / migrate_every ordinary migrations
        sargs.steps, sargs.warmup, sargs.relax = args.slab_steps, args.migrate_every + 5, 200
        slab_line = run_slab(sargs, embedded=True)

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        traffic, traffic_src, binding_unit = None, None, None
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tp) and args.workload == "genome100kb" and R == 64:
            # a number from a committed ncu capture, NOT measured in this run: shown only while the step kernel's
            # sources are byte-for-byte the ones that were profiled, so that it cannot go stale unnoticed
            tj = json.load(open(tp))
            import hashlib
            digest = hashlib.sha256()
            for src in tj.get("kernel_sources", []):
                digest.update(open(os.path.join(ROOT, src), "rb").read())
            if tj.get("kernel_sources") and digest.hexdigest() == tj.get("kernel_sources_sha256"):
                traffic, traffic_src = tj.get("k_step_fused_dram_bytes_per_launch"), tj.get("source")
                binding_unit = tj.get("binding_unit")
            else:
                traffic_src = "dropped: the step kernel's sources changed since " + str(tj.get("captured_at_commit"))
        achieved = ALGO_BYTES_PER_BEAD_STEP * ntot / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
        value = world * ntot * args.steps / (ms * 1e-3)
        line = {
            "metric": "bead_steps_per_s", "value": value, "unit": "bead-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if args.pair_precision != 32 else "f64 state and accumulation, f32 pair displacements",
            "data": "synthetic",
            "config": {"workload": workload_name(args, scen), "beads_per_replica": n, "replicas_per_gpu": R,
                       "timestep": dt, "temperature": kT, "noise": "philox4x32-10",
                       "wall_feedback": ("every step, per replica, on the device (bd_set_wall_relaxation)" if feedback
                                         else "none"),
                       "l2": "working set %.2f GB per GPU > 126 MB L2 (no flush needed)" % (ntot * 32 * 3 / 1e9 + ntot * 4 * 8 / 1e9),
                       "relax_steps": args.relax, "list_rebuilds_in_timed_region": rebuilds,
                       "parallelism": f"replica shards, {world} GPU(s), no data-path collective",
                       "host_binding_rank0": numa},
            "pair_interactions_per_s": inter / (ms * 1e-3),
            "pair_interactions_per_step": inter / args.steps,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_from_profile_at": traffic_src,
                         "binding_unit": binding_unit,
                         "kernel": "k_step_fused",
                         "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": ALGO_BYTES_PER_BEAD_STEP * ntot,
                         "peak_source": peak_src},
        }
        if parity:
            line["parity_check"] = parity
        if mixed:
            line["mixed_precision"] = mixed
        if slab_line:
            line["slab1kb"] = {k: slab_line[k] for k in ("value", "unit", "ms_per_step", "scaling", "steps", "halo_bytes_per_step",
                                                          "pair_interactions_per_s", "gpu_launches")}
            line["slab1kb"]["workload"] = slab_line["config"]["workload"]
            line["slab1kb"]["parity_check"] = slab_line.get("parity_check")
            line["slab1kb"]["parallelism"] = slab_line["config"]["parallelism"]
        if semi_now:
            line["config"]["wall_semiaxes_replica0_after_timed_region"] = semi_now
        if e2e:
            line["e2e"] = {"value": world * ntot * e2e[0] / (e2e_ms_max * 1e-3), "unit": "bead-steps/s",
                           "h2d_bytes_per_step": ntot * 24, "d2h_bytes_per_step": ntot * 24,
                           "steps": e2e[0], "ms_per_step": e2e_ms_max / e2e[0], "gpu_launches_per_step": e2e[2],
                           "path": f"ensemble.HostPipeline: {e2e[3]} engines (replica chunks) on their own streams, per chunk "
                                   "bd_set_all_positions -> bd_step(1) -> bd_get_all_positions_async from / to one pinned "
                                   "host buffer, one synchronisation per step"}
        if world == 1 and not args.no_cpu_baseline:
            backend, kind = cpu_backend()
            start = host.numpy()[:1].copy()  # GPU-relaxed configuration: the same work per step
            builds = {}
            for flavour in cpu_flavours(args, kind):
                probe = cpu_time_steps(backend, scen, start, 3, 1, 1, feedback, flavour)[0] / 3
                s = max(5, int(args.cpu_seconds / len(cpu_flavours(args, kind)) / max(probe, 1e-6)))
                wall, _ = cpu_time_steps(backend, scen, start, s, 1, 2, feedback, flavour)
                builds[flavour] = {"value": n * s / wall, "steps": s}
            best = max(builds, key=lambda k: builds[k]["value"])
            line["cpu_baseline"] = {
                "value": builds[best]["value"], "unit": "bead-steps/s", "cores": 1, "kind": kind, "build": best,
                "builds": {k: v["value"] for k, v in builds.items()},
                "build_notes": "driver-O3 / driver-O2: micromd with the drivers' compile-time potentials and callback "
                               "(oracle/ref_drivers.cc; g++ -O3 -march=x86-64-v3 / -O2); generic: micromd behind the C entry "
                               "points (oracle/ref_harness.cc, -O2 -ffp-contract=off, the parity build)",
                "sample": f"1 replica ({n} beads) x {builds[best]['steps']} BD steps from the GPU-relaxed state, single thread"
                          + (", per-step wall feedback" if feedback else "")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_slab(args, embedded=False):
    """One large periodic 1kb chromatin system, slab-decomposed over the ranks (slab.py).  A step is one
    BD step of the whole system: fused force+integrate on every rank's owned beads, then the halo
    exchange (NCCL send/recv).  Migrations (every --migrate-every steps) are inside the timed region.
    embedded: called from the replica-ensemble run at N > 1 (the process group exists); returns the line
    (rank 0) instead of printing it, so that the scaling record carries the decomposed workload too."""
    import torch
    import torch.distributed as dist

    import bdb200 as bd
    import scenarios as sc
    import slab

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1 and not embedded:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = bd.load_engine()
    scen = sc.chromatin_1kb(args.chains, args.chain_len)
    dcut = max(ff["dcut"] for ff in scen.pair_ffs)
    sysm = bd.System(eng, scen.n, 1, device=local)
    scen.apply(sysm)
    if args.verlet_factor:
        sysm.set_option("verlet_factor", args.verlet_factor)
    dt, kT, seed = scen.timestep, scen.temperature, 77
    sampler = ClockSampler(local)
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stream = slab.work_stream(dev)  # torch's current stream from here on: engine, events and NCCL are ordered on it
    if world == 1:
        sysm.set_stream(stream.cuda_stream)
        sysm.step(args.relax, dt, kT, seed=seed + 1)
        sysm.step(W, dt, kT, seed=seed)
        done = W
        c0, l0 = sysm.get_counters(), sysm.get_launch_count()
        barrier()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sysm.step(K, dt, kT, seed=seed, first_step=done + 1)
        e1.record()
        barrier()
        halo_bytes = 0
    else:
        lay = slab.SlabLayout.balanced(scen.positions[0][:, 0], world, period=scen.period[0])
        sl = slab.SlabSystem(slab.EngineStore(sysm, dev), lay, rank, halo=4.0 * dcut, margin=2.0 * dcut,
                             migrate_every=args.migrate_every, partners=slab.topology_partners(scen))
        ex = slab.DistExchange(dev)
        # The synthetic start (overlapping random walks) relaxes violently: beads would outrun the halo
        # margin.  Every rank relaxes the whole system first, rank 0's result is broadcast so that all
        # ranks start the decomposed run from the same bits.
        sysm.step(args.relax, dt, kT, seed=seed + 1)
        start = sl.store.positions().clone()
        dist.broadcast(start, src=0)
        sysm.set_positions_device(start.data_ptr())
        slab.run_distributed(sl, ex, W, dt, kT, seed)
        c0, l0 = sysm.get_counters(), sysm.get_launch_count()
        b0 = sl.halo_bytes_sent
        barrier()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        slab.run_distributed(sl, ex, K, dt, kT, seed)
        e1.record()
        barrier()
        halo_bytes = sl.halo_bytes_sent - b0
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    c1 = sysm.get_counters()
    t = torch.tensor([ms, float(c1.pair_interactions - c0.pair_interactions), float(halo_bytes),
                      float(sysm.get_launch_count() - l0)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax, tsum = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, inter, halo_bytes, launches = tmax[0].item(), tsum[1].item(), tsum[2].item(), tsum[3].item()
    else:
        inter, launches = t[1].item(), t[3].item()
    slab_parity = None
    if world > 1:
        # In-run check of the decomposed trajectory (outside the timed region): the ranks carry on for a few steps
        # from the state the timed region left; rank 0 replays the same steps on ONE engine that holds the whole
        # system.  Philox is keyed by (seed, step, bead), so both see the same noise; forces differ in summation
        # order only.  A short stretch, because rounding differences grow along a chaotic trajectory.
        V, tol = 32, 1e-9
        before = slab.gather_owned(sl, ex)
        first = sl.steps_done + 1
        slab.run_distributed(sl, ex, V, dt, kT, seed)
        after = slab.gather_owned(sl, ex)
        rebuilds = torch.tensor([float(sysm.get_counters().list_rebuilds)], dtype=torch.float64, device="cuda")
        rmin, rmax = rebuilds.clone(), rebuilds.clone()
        dist.all_reduce(rmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(rmax, op=dist.ReduceOp.MAX)
        bad = torch.zeros(1, dtype=torch.float64, device="cuda")
        if rank == 0:
            whole = bd.System(eng, scen.n, 1, device=local)
            scen.apply(whole)
            if args.verlet_factor:
                whole.set_option("verlet_factor", args.verlet_factor)
            whole.step(1, dt, kT, seed=seed)  # one throw-away step: the engine then knows the previous timestep
            whole.set_positions(before)
            whole.step(V, dt, kT, seed=seed, first_step=first)
            want = whole.get_positions()
            whole.close()
            slab_parity = {"max_dx": float(np.abs(after - want).max()), "moved": float(np.abs(want - before).max()),
                           "steps": V, "from_step": first, "ranks": world, "tolerance": tol,
                           "list_rebuilds_equal_on_all_ranks": bool(rmin.item() == rmax.item())}
            if not (slab_parity["max_dx"] <= tol and slab_parity["moved"] > 100 * tol
                    and slab_parity["list_rebuilds_equal_on_all_ranks"]):
                bad += 1
        dist.all_reduce(bad)
        if bad.item():
            raise SystemExit(f"bench.py: slab parity check failed: {slab_parity}")
    if rank == 0:
        line = {
            "metric": "bead_steps_per_s", "value": scen.n * K / (ms * 1e-3), "unit": "bead-steps/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"3-sim-1kb chromatin, one periodic system of {scen.n} beads ({args.chains} chains x {args.chain_len})",
                       "timestep": dt, "temperature": kT, "noise": "philox4x32-10", "relax_steps": args.relax,
                       "parallelism": (f"{world} x-slabs, per-step halo exchange (one indexed gather, one NCCL all-to-all, one indexed scatter), migration every "
                                       f"{args.migrate_every} steps" if world > 1 else "single GPU, no decomposition"),
                       "halo_width": 4.0 * dcut, "l2": "working set > 126 MB L2 for >= 2 M beads"},
            "pair_interactions_per_s": inter / (ms * 1e-3), "gpu_launches": int(launches),
            "halo_bytes_per_step": halo_bytes / K, "clocks": clocks,
        }
        if slab_parity:
            line["parity_check"] = slab_parity
        if not embedded:
            print(json.dumps(line), flush=True)
    sysm.close()
    if world > 1 and not embedded:
        dist.destroy_process_group()
    return line if rank == 0 else None


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "slab1kb":
        run_slab(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
