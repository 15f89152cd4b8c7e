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
REFERENCE_FAST_LIB = os.path.join(ROOT, "oracle", "_ref", "libmicromd_fast.so")

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
