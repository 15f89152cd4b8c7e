"""One rank of the world_size-2 ensemble test (launched by torch.distributed.run, gloo, CPU).

Test infrastructure: drives ensemble.ReplicaEnsemble on the CPU oracle so that the sharding and
gathering logic of the N>1 path runs without a GPU."""
import os
import sys

import numpy as np
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "2022a-genome-dynamics_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import checkers  # noqa: E402
import ensemble as en  # noqa: E402
import scenarios as sc  # noqa: E402


def main():
    out_dir, n_global, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    scen = sc.genome_10mb(total=320, bend_energy=0.7)
    ens = en.ReplicaEnsemble(checkers.load_oracle(), scen, n_global, rank=rank, world=world, seed=9)
    ens.step(steps)
    stats = ens.statistics(device="cpu")
    np.save(os.path.join(out_dir, f"stats_rank{rank}.npy"), stats)
    np.save(os.path.join(out_dir, f"ids_rank{rank}.npy"), ens.ids)
    np.save(os.path.join(out_dir, f"pos_rank{rank}.npy"), ens.positions())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
