"""The micromd-compatible C++ header (2022a-genome-dynamics_b200/md/md.hpp) against the reference.

tests/cpp/shim_driver.cc is a miniature of the paper's drivers written against <md.hpp> only.  The
expected output, tests/golden/shim_driver_expected.txt, was printed by that program compiled against
the UNMODIFIED micromd headers (regenerated and compared below wherever /root/reference exists).

  * CPU: the same source compiled against md.hpp with the calls bound to the CPU oracle
    (-DMD_B200_PREFIX=orc_) checks the header's host logic -- class tabulation of the lambdas,
    parameter refresh between steps, wall statistics feedback, phase changes -- without a GPU.
  * GPU: compiled against md.hpp + libbd_b200.so, i.e. exactly what a driver links.

Temperature is zero in the driver, so trajectories do not depend on the noise generator.
"""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "2022a-genome-dynamics_b200")
SRC = os.path.join(ROOT, "tests", "cpp", "shim_driver.cc")
EXPECTED = os.path.join(ROOT, "tests", "golden", "shim_driver_expected.txt")
MICROMD = "/root/reference/submodules/github.com/snsinfu/micromd/include"

# relative to max(1, largest component of the line); the wall semiaxes integrate a feedback loop
# over 120 steps and amplify last-bit differences of the reaction statistics
TOL = {"default": 1e-9, "interphase.semiaxes": 1e-8}


def parse(text):
    rows = []
    for line in text.strip().splitlines():
        f = line.split()
        rows.append((f[0], np.array([float(x) for x in f[1:]])))
    return rows


def compare(got, want):
    # the B200 header runs the 3-sim-1kb case a second time with its pair-list force fields on the device
    # (chromatin_dev.*): same model, so the same numbers as the reference's chromatin.* lines
    dev = [(t.replace("chromatin_dev", "chromatin"), a) for t, a in got if t.startswith("chromatin_dev")]
    got = [(t, a) for t, a in got if not t.startswith("chromatin_dev")]
    ref_chromatin = [(t, a) for t, a in want if t.startswith("chromatin.")]
    assert dev and [t for t, _ in dev] == [t for t, _ in ref_chromatin]
    got, want = got + dev, want + ref_chromatin
    assert [t for t, _ in got] == [t for t, _ in want]
    for (tag, a), (_, b) in zip(got, want):
        err = np.abs(a - b).max() / max(1.0, np.abs(b).max())
        assert err <= TOL.get(tag, TOL["default"]), (tag, a, b, err)


def build_and_run(tmp_path, name, flags):
    exe = str(tmp_path / name)
    cmd = ["g++", "-std=c++20", "-O2", "-ffp-contract=off", SRC, "-o", exe] + flags
    subprocess.run(cmd, check=True)
    return subprocess.run([exe], check=True, capture_output=True, text=True, timeout=600).stdout


def test_expected_output_is_the_reference(tmp_path):
    """The committed expectation is what the unmodified micromd prints for this driver."""
    if not os.path.isdir(MICROMD):
        pytest.skip("no /root/reference here")
    out = build_and_run(tmp_path, "ref_driver", ["-I" + MICROMD])
    assert parse(out) and out == open(EXPECTED).read()


def test_header_logic_on_oracle_backend(tmp_path):
    oracle_dir = os.path.join(ROOT, "oracle")
    out = build_and_run(tmp_path, "orc_driver", [
        "-DMD_B200_PREFIX=orc_", "-I" + os.path.join(PKG, "md"), "-I" + os.path.join(ROOT, "include"),
        "-L" + oracle_dir, "-loracle", "-Wl,-rpath," + oracle_dir])
    compare(parse(out), parse(open(EXPECTED).read()))


@pytest.mark.gpu
def test_driver_links_the_engine(tmp_path, engine):
    assert shutil.which("g++"), "the driver is C++: a host compiler is needed"
    out = build_and_run(tmp_path, "b200_driver", [
        "-I" + os.path.join(PKG, "md"), "-I" + os.path.join(ROOT, "include"),
        "-L" + PKG, "-lbd_b200", "-Wl,-rpath," + PKG])
    compare(parse(out), parse(open(EXPECTED).read()))
