#!/usr/bin/env python
"""Every BASELINE.json configuration on one B200 + the single-system latency of a relinked driver.

    python profiles/measure_configs.py [--steps 100] [--out profiles/r2_configs.json]

Runs on the GPU box (needs libbd_b200.so and, for the CPU figures, oracle/_ref built where /root/reference
exists).  Per configuration: bench.py --workload <name> (device-resident throughput, pair interactions/s,
roofline fraction of the fused step kernel, the reference CPU figure timed in the same run).  Then the path a
relinked driver takes: ONE 62 600-bead system through md::simulate_brownian_dynamics with the per-step wall
feedback callback -- oracle/ref_drivers.cc (the reference driver's set-up calls) compiled UNCHANGED against
2022a-genome-dynamics_b200/md/md.hpp and linked with libbd_b200.so, next to the same source on the unmodified
micromd (-O3).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "2022a-genome-dynamics_b200")
for p in (PKG, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)


def bench_line(extra, timeout=900):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py")] + extra
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if out.returncode != 0:
        return {"error": out.stderr[-600:]}
    return json.loads(out.stdout.strip().splitlines()[-1])


def single_system_latency(steps):
    import checkers
    import scenarios as sc

    build = os.path.join(ROOT, "profiles", "_build")
    os.makedirs(build, exist_ok=True)
    lib = os.path.join(build, "libdrv_b200.so")
    subprocess.run(["g++", "-std=c++20", "-O2", "-fPIC", "-shared", "-I" + os.path.join(PKG, "md"),
                    "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "oracle", "ref_drivers.cc"), "-o", lib,
                    "-L" + PKG, "-lbd_b200", "-Wl,-rpath," + PKG], check=True)
    scen = sc.genome_100kb(62000, 300, seed=4)
    out = {"beads": scen.n, "steps": steps,
           "path": "md::simulate_brownian_dynamics(system, {.callback = wall feedback}) on md/md.hpp + libbd_b200.so: "
                   "per step one bd_step(1), the potentials re-tabulated, both walls' statistics read back, the walls updated"}
    d = checkers.DriverC4(scen, scen.positions[0], lib_path=lib)
    d.run(50)  # relax + build lists
    t0 = time.perf_counter()
    d.run(steps)
    dt = time.perf_counter() - t0
    d.close()
    out["b200_steps_per_s"] = steps / dt
    out["b200_ms_per_step"] = 1e3 * dt / steps
    if checkers.have_driver_baseline("O3"):
        c = checkers.DriverC4(scen, scen.positions[0], "O3")
        c.run(10)
        k = max(5, steps // 20)
        t0 = time.perf_counter()
        c.run(k)
        dc = time.perf_counter() - t0
        c.close()
        out["micromd_steps_per_s"] = k / dc
        out["micromd_ms_per_step"] = 1e3 * dc / k
        out["micromd_build"] = "oracle/ref_drivers.cc on the unmodified micromd, g++ -O3 -march=x86-64-v3, one thread"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_configs.json"))
    ap.add_argument("--chains", type=int, default=600)
    args = ap.parse_args()
    table = {}
    for name, extra in (("genome100kb", []), ("ab_box", ["--no-e2e"]), ("ab_sphere", ["--no-e2e"]),
                        ("genome10mb", ["--no-e2e"]), ("chromatin1kb", ["--no-e2e", "--chains", str(args.chains)])):
        line = bench_line(["--workload", name, "--steps", str(args.steps), "--warmup", "5", "--cpu-seconds", "6"] + extra)
        table[name] = line
        if "error" in line:
            print(name, "FAILED", line["error"], flush=True)
            continue
        print(name, "%.3e bead-steps/s, %.3e pair interactions/s, %.4f ms/step, roofline %.4f, cpu %.3e (%s)" % (
            line["value"], line["pair_interactions_per_s"], line["ms_per_step"], line["roofline"]["frac"],
            line.get("cpu_baseline", {}).get("value", float("nan")), line.get("cpu_baseline", {}).get("build", "-")), flush=True)
    table["single_system_latency_c4"] = single_system_latency(200)
    print("single system:", table["single_system_latency_c4"], flush=True)
    json.dump(table, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
