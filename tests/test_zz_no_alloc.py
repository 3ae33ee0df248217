"""lmc_img_sample_metrics_static -- the entry point that allocates nothing (scratch in a static device array inside
the library; include/lmc_b200.h) -- against lmc_img_sample_metrics on the same inputs, through the C-ABI, on two
calls in a row (the second call finds the first call's partial sums in the shared array).  Same kernels, same
summation order: the results must be equal bit for bit.  Collected last on purpose."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


def _setup(torch, dims, dtype, kernel=5):
    from lmc_atomi_b200 import imaging as I
    rng = np.random.default_rng(11)
    P = dims[0] * dims[1]
    h = np.ones((kernel, kernel)) / kernel ** 2
    eng = I.DeconvEngine(dims, h, 255 * rng.random(P), 1 / 0.75 ** 2, dtype=dtype)
    xtrue = torch.from_numpy(255 * rng.random(P)).to("cuda", eng.tdt)
    return eng, rng, xtrue


def _both(torch, eng, x, xtrue, isotropic, per_chain):
    """(result of the workspace entry point, result of the static one), each (M,4) float64 on the host."""
    from lmc_atomi_b200 import _lib, imaging as I
    lib = eng.lib
    M = x.shape[0]
    pb = eng._problem()
    dt, st = I._dt(eng.dtype), I._stream(torch, eng.dev)
    nb = lib.lmc_img_metrics_workspace_bytes(M)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    want = torch.full((M, 4), float("nan"), dtype=torch.float64, device="cuda")
    got = torch.full((M, 4), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    with torch.cuda.device(eng.dev):
        rc = lib.lmc_img_sample_metrics(C.byref(pb), dt, isotropic, M, per_chain, x.data_ptr(), xtrue.data_ptr(),
                                        want.data_ptr(), ws.data_ptr(), nb, st)
        _lib.check(rc, "lmc_img_sample_metrics")
        before = torch.cuda.memory_allocated()
        rc = lib.lmc_img_sample_metrics_static(C.byref(pb), dt, isotropic, M, per_chain, x.data_ptr(),
                                               xtrue.data_ptr(), got.data_ptr(), st)
        _lib.check(rc, "lmc_img_sample_metrics_static")
        assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_static_scratch_matches_workspace_entry_point_twice(torch, dtype):
    dims = (70, 93)
    eng, rng, xtrue = _setup(torch, dims, dtype)
    P = dims[0] * dims[1]
    # two calls in a row with different M, inputs and prior: the second one runs over the first one's leftovers
    for M, iso, per_chain in ((37, 1, 1), (12, 0, 1)):
        x = torch.from_numpy(255 * rng.random((M, P))).to("cuda", eng.tdt)
        want, got = _both(torch, eng, x, xtrue, iso, per_chain)
        assert np.isfinite(want).all() and (want[:, :3] > 0).all()
        assert np.array_equal(want, got), np.abs(want - got).max()


def test_static_scratch_at_its_maximum_and_beyond(torch):
    from lmc_atomi_b200 import _lib, imaging as I
    dims = (24, 40)
    eng, rng, xtrue = _setup(torch, dims, "float32")
    P = dims[0] * dims[1]
    MAX = 1024      # LMC_METRICS_STATIC_MAX_IMAGES (include/lmc_b200.h)
    hdr = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "include", "lmc_b200.h")
    assert f"#define LMC_METRICS_STATIC_MAX_IMAGES {MAX}" in open(hdr).read()
    x = torch.from_numpy(255 * rng.random((MAX + 1, P))).to("cuda", eng.tdt)
    want, got = _both(torch, eng, x[:MAX], xtrue, 1, 1)
    assert np.array_equal(want, got)
    out = torch.zeros((MAX + 1, 4), dtype=torch.float64, device="cuda")
    pb = eng._problem()
    rc = eng.lib.lmc_img_sample_metrics_static(C.byref(pb), I._dt(eng.dtype), 1, MAX + 1, 1, x.data_ptr(),
                                               xtrue.data_ptr(), out.data_ptr(), I._stream(torch, eng.dev))
    assert rc == -1
    assert b"exceeds the static scratch" in eng.lib.lmc_last_error()
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0      # rejected before any launch
