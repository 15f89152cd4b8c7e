"""bd_refine_spline_device_noalloc: the spline hand-off for callers that cannot allocate on the device.

Some users of the library run the 10Mb -> 100kb refinement at a moment when cudaMalloc is not an option: the card
has been filled by the engines of the next phase (or by a framework's caching allocator), or other engines are
stepping on the same GPU and would be stalled by the device-wide synchronisation every cudaMalloc / cudaFree
implies.  For them the library offers a variant of bd_refine_spline_device whose workspace is a fixed array inside
libbd_b200.so (include/bd_cuda.h).  This test calls it through the C ABI and checks it against the allocating entry
point on the same device arrays, twice in a row -- the second call is a different problem of another shape, so anything
the first call left in the shared workspace would show -- and checks that oversized calls are refused.

The file is named so that pytest collects it last: `pytest -x` stops at the first failure and a failure here must
not hide the parity tests.  It passes alone in a fresh process (the first call then does the per-device set-up) and
after the rest of the suite.
"""
import ctypes as C
import os

import numpy as np
import pytest

import bdb200 as bd

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-10  # tests/test_refine.py: relative to the extent of the conformation

I64P = C.POINTER(C.c_int64)
SIGNATURE = [C.c_void_p, I64P, C.c_int64, C.c_int32, I64P, I64P, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]


def bind(engine, name):
    f = getattr(engine.lib, name)
    f.restype = C.c_int
    f.argtypes = SIGNATURE
    return f


def last_error(engine):
    msg = engine._last_error()
    return msg.decode() if msg else ""


def refine_on_device(engine, fn, coarse, coarse_ranges, factor, fine_ranges, n_fine, attach, fill=0.0):
    """Device arrays in, device array out (torch only provides the memory); ranges and attachments stay on the host."""
    import torch

    cr = np.ascontiguousarray(coarse_ranges, dtype=np.int64).reshape(-1, 2)
    fr = np.ascontiguousarray(fine_ranges, dtype=np.int64).reshape(-1, 2)
    at = np.ascontiguousarray(attach, dtype=np.int64).reshape(-1, 2)
    d_coarse = torch.as_tensor(np.ascontiguousarray(coarse, dtype=np.float64), device="cuda:0")
    d_fine = torch.full((int(n_fine), 3), float(fill), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    rc = fn(d_coarse.data_ptr(), cr.ctypes.data_as(I64P), len(cr), int(factor), fr.ctypes.data_as(I64P),
            at.ctypes.data_as(I64P), len(at), len(coarse), int(n_fine), d_fine.data_ptr(), None)
    if rc != 0:
        raise bd.BackendError(last_error(engine))
    torch.cuda.synchronize()
    refine_on_device.device_bytes_taken = free_before - torch.cuda.mem_get_info(0)[0]  # across the C call alone
    return d_fine.cpu().numpy()


def err(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


@pytest.mark.gpu
def test_noalloc_refinement_matches_the_allocating_entry_point(engine):
    import torch

    torch.cuda.set_device(0)  # the entry point works on the current device
    noalloc = bind(engine, "bd_refine_spline_device_noalloc")
    alloc = bind(engine, "bd_refine_spline_device")

    g = np.load(os.path.join(HERE, "golden", "refine_spline.npz"))
    rng = np.random.default_rng(11)
    # call 1: ragged random chains; more coarse points than call 2, fewer chains and attachments
    lengths = [300, 4, 57, 128, 5, 211]
    starts = np.concatenate([[0], np.cumsum(lengths)])
    coarse1 = np.cumsum(rng.normal(size=(starts[-1], 3)), axis=0)
    cr1 = [[starts[i], starts[i + 1]] for i in range(len(lengths))]
    factor1 = 7
    fine_len = [n * factor1 - (3 if i % 2 else 0) for i, n in enumerate(lengths)]  # some chains are cut short (run.py:34)
    fstarts = np.concatenate([[0], np.cumsum(fine_len)])
    fr1 = [[fstarts[i], fstarts[i + 1]] for i in range(len(lengths))]
    n_fine1 = int(fstarts[-1]) + 9  # 9 rows no chain covers: they keep the caller's value
    attach1 = [[10, n_fine1 - 1], [500, n_fine1 - 2], [0, n_fine1 - 3]]
    problems = [
        (coarse1, cr1, factor1, fr1, n_fine1, attach1),
        # call 2: the reference-made golden case (also compared with the reference's own output)
        (g["coarse"], g["coarse_ranges"], int(g["factor"]), g["fine_ranges"], len(g["fine"]), g["attach"]),
    ]
    outputs = []
    for k, p in enumerate(problems):
        got = refine_on_device(engine, noalloc, *p, fill=7.0)
        if k == 1:
            # set-up is done and the library's device image is loaded: the call must not have taken device memory
            assert refine_on_device.device_bytes_taken == 0
        want = refine_on_device(engine, alloc, *p, fill=7.0)
        assert np.isfinite(got).all()
        assert err(got, want) <= TOL, (k, err(got, want))
        outputs.append(got)
    assert np.array_equal(outputs[0][-9:-3], np.full((6, 3), 7.0))
    for nor, nuc in attach1:
        assert np.array_equal(outputs[0][nuc], outputs[0][nor])
    assert err(outputs[1], g["fine"]) <= TOL
    # the first problem again: the workspace now holds the golden case's leftovers
    again = refine_on_device(engine, noalloc, *problems[0], fill=7.0)
    assert np.array_equal(again, outputs[0])


@pytest.mark.gpu
def test_noalloc_refinement_refuses_what_does_not_fit(engine):
    import torch

    torch.cuda.set_device(0)
    noalloc = bind(engine, "bd_refine_spline_device_noalloc")
    too_many = 65536 + 1  # BD_REFINE_NOALLOC_MAX_COARSE + 1
    coarse = np.zeros((too_many, 3))
    with pytest.raises(bd.BackendError, match="static workspace holds at most 65536 coarse points"):
        refine_on_device(engine, noalloc, coarse, [[0, too_many]], 1, [[0, too_many]], too_many, np.zeros((0, 2)))
    # the same argument checks as the allocating entry point
    small = np.cumsum(np.ones((10, 3)), axis=0)
    with pytest.raises(bd.BackendError):
        refine_on_device(engine, noalloc, small, [[0, 3]], 2, [[0, 6]], 6, np.zeros((0, 2)))
    with pytest.raises(bd.BackendError):
        refine_on_device(engine, noalloc, small, [[0, 10]], 2, [[0, 21]], 21, np.zeros((0, 2)))
    # exactly at the limit is accepted
    full = np.cumsum(np.random.default_rng(5).normal(size=(65536, 3)), axis=0)
    out = refine_on_device(engine, noalloc, full, [[0, 65536]], 2, [[0, 131072]], 131072, np.zeros((0, 2)))
    want = refine_on_device(engine, bind(engine, "bd_refine_spline_device"), full, [[0, 65536]], 2, [[0, 131072]], 131072,
                            np.zeros((0, 2)))
    assert err(out, want) <= TOL
