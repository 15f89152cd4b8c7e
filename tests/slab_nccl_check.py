"""The NCCL path of the slab decomposition against the undecomposed run, on real GPUs (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/slab_nccl_check.py

Not a pytest case: the round-end GPU test box has one GPU.  Every rank runs the decomposed system through
slab.DistExchange.halo (one indexed gather, one NCCL all-to-all, one indexed scatter per step); rank 0 also runs
the whole system on its own engine and compares (<= 1e-9: same noise per bead, different summation order).
Prints one line and exits non-zero on a mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "2022a-genome-dynamics_b200"))

import bdb200 as bd  # noqa: E402
import scenarios as sc  # noqa: E402
import slab  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    eng = bd.load_engine()
    scen = sc.chromatin_1kb(8, 500)
    dcut = max(ff["dcut"] for ff in scen.pair_ffs)
    steps, seed = 48, 5
    dev = f"cuda:{local}"
    s = bd.System(eng, scen.n, 1, device=local)
    scen.apply(s)
    lay = slab.SlabLayout.balanced(scen.positions[0][:, 0], world, period=scen.period[0])
    sl = slab.SlabSystem(slab.EngineStore(s, dev), lay, rank, halo=3.0 * dcut, margin=dcut, migrate_every=8,
                         partners=slab.topology_partners(scen))
    ex = slab.DistExchange(dev)
    slab.run_distributed(sl, ex, steps, scen.timestep, scen.temperature, seed)
    got = slab.gather_owned(sl, ex)
    ok = True
    if rank == 0:
        whole = bd.System(eng, scen.n, 1, device=local)
        scen.apply(whole)
        whole.step(steps, scen.timestep, scen.temperature, seed=seed)
        want = whole.get_positions()
        err = float(np.abs(got - want).max())
        moved = float(np.abs(want - scen.positions[0]).max())
        ok = err < 1e-9 and moved > 1e-3
        print(f"slab_nccl_check world={world} beads={scen.n} steps={steps} max|dx|={err:.3e} moved={moved:.3e} "
              f"halo_bytes_per_step={sl.halo_bytes_sent / steps:.0f} {'OK' if ok else 'MISMATCH'}", flush=True)
    # one rebuild decision for the whole system (slab.DistExchange.halo all-reduces the engines' displacement word in
    # place): every rank must have rebuilt its list the same number of times
    rebuilds = torch.tensor([float(s.get_counters().list_rebuilds)], dtype=torch.float64, device=dev)
    lo, hi = rebuilds.clone(), rebuilds.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if lo.item() != hi.item():
        ok = False
        if rank == 0:
            print(f"slab_nccl_check: list rebuild counts differ between ranks ({lo.item():.0f} .. {hi.item():.0f})", flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(int(flag.item() != 0))


if __name__ == "__main__":
    main()
