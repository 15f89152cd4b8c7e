import sys, os, numpy as np
sys.path.insert(0, "2022a-genome-dynamics_b200")
import bdb200 as bd, scenarios as sc
eng = bd.load_engine()
for R, total in ((1, 62000), (4, 62000), (64, 62000)):
    scen = sc.genome_100kb(total=total, n_nor=300, seed=4, replicas=min(R, 4))
    s = bd.System(eng, scen.n, R)
    scen.apply(s, replicas=[])
    for r in range(R):
        s.set_positions(scen.positions[r % scen.replicas], replica=r)
    e = s.compute_energy(0)
    c = s.get_counters()
    print("R", R, "energy", e, "listed", c.listed_pairs, "rebuilds", c.list_rebuilds, "cap", s.get_option("list_capacity"), flush=True)
    for k in range(5):
        s.step(100, scen.timestep, scen.temperature, seed=1, first_step=1 + 100 * k)
        try:
            s.sync()
        except Exception as ex:
            print("  sync error", ex)
        x = s.get_positions(R - 1) if True else None
        c = s.get_counters() if False else None
        print("  after", 100 * (k + 1), "finite", np.isfinite(x).all(), "max|x|", np.abs(x).max(), flush=True)
    s.close()
