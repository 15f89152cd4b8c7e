"""TEST INFRASTRUCTURE -- loaders of the two CPU checkers, behind the same Python binding as the engine.

* ``load_oracle()``     oracle/liboracle.so          the plain-C restatement (oracle/bd_oracle.c), prefix ``orc_``
* ``load_reference()``  oracle/_ref/libmicromd_ref.so  the UNMODIFIED micromd behind the same entry points
                                                      (oracle/ref_harness.cc, built by oracle/Makefile), prefix ``ref_``

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline / ``--impl reference`` legs import this
module.  Nothing under 2022a-genome-dynamics_b200/ refers to it: the product loads libbd_b200.so or fails.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "2022a-genome-dynamics_b200"))

import bdb200  # noqa: E402

ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")
REFERENCE_LIB = os.path.join(ROOT, "oracle", "_ref", "libmicromd_ref.so")

_cache = {}


def load_oracle() -> "bdb200.Backend":
    if "orc_" not in _cache:
        _cache["orc_"] = bdb200.Backend(ORACLE_LIB, "orc_")
    return _cache["orc_"]


def load_reference() -> "bdb200.Backend":
    if "ref_" not in _cache:
        _cache["ref_"] = bdb200.Backend(REFERENCE_LIB, "ref_")
    return _cache["ref_"]


def have_reference() -> bool:
    return os.path.exists(REFERENCE_LIB)


# ---- the 100kb interphase model set up as the reference driver sets it up (oracle/ref_drivers.cc) --------

import ctypes as _C

import numpy as _np

DRIVER_LIBS = {"O2": os.path.join(ROOT, "oracle", "_ref", "libmicromd_drv_O2.so"),
               "O3": os.path.join(ROOT, "oracle", "_ref", "libmicromd_drv_O3.so")}


def have_driver_baseline(flavour="O2") -> bool:
    return os.path.exists(DRIVER_LIBS[flavour])


class DriverC4:
    """One system of scenarios.genome_100kb on the unmodified micromd with compile-time potentials and the
    driver's per-step wall feedback as the simulate_brownian_dynamics callback.  TIMING ONLY."""

    def __init__(self, scen, positions, flavour="O2", lib_path=None):
        """lib_path: the same source (oracle/ref_drivers.cc) compiled against another <md.hpp> -- e.g. the
        B200 header 2022a-genome-dynamics_b200/md/md.hpp + libbd_b200.so, which is how the single-system
        latency of a relinked driver is measured (profiles/measure_configs.py)"""
        import scenarios as sc

        lib = _C.CDLL(lib_path or DRIVER_LIBS[flavour])
        lib.drv_c4_create.restype = _C.c_void_p
        lib.drv_c4_create.argtypes = [_C.c_int64, _C.c_void_p, _C.c_void_p, _C.c_int64, _C.c_void_p, _C.c_int64,
                                      _C.c_void_p, _C.c_void_p, _C.c_int64]
        lib.drv_c4_run.argtypes = [_C.c_void_p, _C.c_int64, _C.c_double, _C.c_double, _C.c_uint64, _C.c_int32]
        lib.drv_c4_get.argtypes = [_C.c_void_p, _C.c_void_p, _C.c_void_p]
        lib.drv_c4_destroy.argtypes = [_C.c_void_p]
        self.lib, self.scen = lib, scen
        p = scen.params
        ab = _np.ascontiguousarray(sc.C4_AB[scen.types], dtype=_np.float64)
        chains = _np.ascontiguousarray(scen.chains, dtype=_np.int64)
        nuc = _np.ascontiguousarray(scen.bond_ffs[2]["pairs"], dtype=_np.int64)
        params = _np.array([p["a_core_diameter"], p["b_core_diameter"], p["a_core_repulsion"], p["b_core_repulsion"],
                            p["a_core_bond_spring"], p["a_core_bond_length"], p["b_core_bond_spring"], p["b_core_bond_length"],
                            p["a_core_2nd_bond_spring"], p["b_core_2nd_bond_spring"],
                            p["nucleolus_bond_spring"], p["nucleolus_bond_length"],
                            p["wall_packing_spring"], p["wall_a_factor"], p["wall_b_factor"],
                            p["wall_mobility"], p["wall_semiaxes_spring"], *p["semiaxes"],
                            p.get("bead_scale", 1.0), p.get("bond_scale", 1.0)], dtype=_np.float64)
        x = _np.ascontiguousarray(positions, dtype=_np.float64)
        self.h = lib.drv_c4_create(scen.n, ab.ctypes.data, chains.ctypes.data, len(chains), nuc.ctypes.data, len(nuc),
                                   x.ctypes.data, params.ctypes.data, len(params))
        if not self.h:
            raise RuntimeError("drv_c4_create failed (parameter block mismatch)")

    def run(self, steps, seed=1, feedback=True):
        self.lib.drv_c4_run(self.h, int(steps), float(self.scen.temperature), float(self.scen.timestep), int(seed),
                            1 if feedback else 0)

    def get(self):
        x, s = _np.zeros((self.scen.n, 3)), _np.zeros(3)
        self.lib.drv_c4_get(self.h, x.ctypes.data, s.ctypes.data)
        return x, s

    def close(self):
        if self.h:
            self.lib.drv_c4_destroy(self.h)
            self.h = None


# ---- 3-sim-1kb's kinetic glue model, the unmodified class (oracle/ref_glue.cc + the reference's glue_simulator.cpp) ----

GLUE_LIB = os.path.join(ROOT, "oracle", "_ref", "libglue_ref.so")


def have_reference_glue() -> bool:
    return os.path.exists(GLUE_LIB)


class ReferenceGlue:
    """glue_simulator of the 1kb chromatin driver with its own std::mt19937_64; CHECKER of bd_glue_update only."""

    def __init__(self, max_glues, max_distance, binding_rate, unbinding_rate, period, seed):
        lib = _C.CDLL(GLUE_LIB)
        lib.ref_glue_create.restype = _C.c_void_p
        lib.ref_glue_create.argtypes = [_C.c_uint64, _C.c_double, _C.c_double, _C.c_double, _C.c_void_p, _C.c_uint64]
        lib.ref_glue_destroy.argtypes = [_C.c_void_p]
        lib.ref_glue_update.restype = _C.c_int64
        lib.ref_glue_update.argtypes = [_C.c_void_p, _C.c_double, _C.c_void_p, _C.c_int64]
        lib.ref_glue_pairs.restype = _C.c_int64
        lib.ref_glue_pairs.argtypes = [_C.c_void_p, _C.c_void_p, _C.c_int64]
        self.lib = lib
        box = _np.ascontiguousarray(period, dtype=_np.float64)
        self.h = lib.ref_glue_create(int(max_glues), float(max_distance), float(binding_rate), float(unbinding_rate),
                                     box.ctypes.data, int(seed))

    def update(self, timestep, xyz) -> int:
        x = _np.ascontiguousarray(xyz, dtype=_np.float64).reshape(-1, 3)
        return self.lib.ref_glue_update(self.h, float(timestep), x.ctypes.data, len(x))

    def pairs(self) -> "_np.ndarray":
        n = self.lib.ref_glue_pairs(self.h, None, 0)
        out = _np.zeros((max(n, 1), 2), dtype=_np.uint32)
        self.lib.ref_glue_pairs(self.h, out.ctypes.data, n)
        return out[:n]

    def close(self):
        if self.h:
            self.lib.ref_glue_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()
