"""One rank of the world_size-2 slab test (torch.distributed.run, gloo, CPU toy store): the halo and
ownership protocol of slab.py over a real process group.  Test infrastructure."""
import os
import sys

import numpy as np
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "2022a-genome-dynamics_b200"))

import slab  # noqa: E402


def main():
    out_dir, steps = sys.argv[1], int(sys.argv[2])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    x0 = np.random.default_rng(3).uniform(0.0, 8.0, size=(600, 3))
    layout = slab.SlabLayout.balanced(x0[:, 0], world, period=8.0)
    store = slab.ToyStore(x0, amplitude=0.02)
    s = slab.SlabSystem(store, layout, rank, halo=0.9, margin=0.5, migrate_every=5)
    ex = slab.DistExchange("cpu")
    slab.run_distributed(s, ex, steps, 1e-4, 0.0, 1)
    np.save(os.path.join(out_dir, f"slab_all_rank{rank}.npy"), slab.gather_owned(s, ex))
    np.save(os.path.join(out_dir, f"slab_owned_rank{rank}.npy"), s.owned.numpy())
    np.save(os.path.join(out_dir, f"slab_local_rank{rank}.npy"), store.xyz.numpy())
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
