"""Chunked runs with a burn-in: the running mean / M2 every image kernel keeps (lmc_img_io::d_mean, d_m2) must equal
numpy's mean / variance of the samples the same launches return -- per engine, not engine against engine (all five
kernels share the count formula).  Contract (include/lmc_b200.h): burnin and iter_offset are global iteration
indices, stat_count0 = samples folded in before the launch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


def _problem(dims):
    from oracle import deconv_ref as D
    return D.build_problem(D.synthetic_image(*dims), kernel=5)


CHUNKS = (3, 4, 6)       # burn-in 5 ends inside the second launch
BURNIN = 5


def _check(torch, samples, mean, m2, tol):
    s = np.concatenate([t.double().cpu().numpy() for t in samples], axis=1)[:, BURNIN:]      # (C, n_kept, P)
    n = s.shape[1]
    assert n == sum(CHUNKS) - BURNIN
    want_mean, want_m2 = s.mean(axis=1), ((s - s.mean(axis=1, keepdims=True)) ** 2).sum(axis=1)
    got_mean, got_m2 = mean.double().cpu().numpy(), m2.double().cpu().numpy()
    assert np.abs(got_mean - want_mean).max() <= tol * max(1.0, np.abs(want_mean).max())
    assert np.abs(got_m2 - want_m2).max() <= 10 * tol * max(1.0, np.abs(want_m2).max())


@pytest.mark.parametrize("engine,dtype", [("stream", "float64"), ("band", "float64"), ("band", "float32"),
                                          ("onchip", "float32")])
def test_myula_chunked_mean_and_m2(torch, engine, dtype):
    from lmc_atomi_b200 import imaging as I
    dims, Cn = (48, 64), 3
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    x = torch.zeros((Cn, P), dtype=eng.tdt, device="cuda")
    mean, m2 = torch.zeros_like(x), torch.zeros_like(x)
    done, samples = 0, []
    for n in CHUNKS:
        s = torch.empty((Cn, n, P), dtype=eng.tdt, device="cuda")
        eng.myula(x, n, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, seed=5, samples=s, mean=mean, m2=m2,
                  burnin=BURNIN, stat_count0=max(done - BURNIN, 0), iter_offset=done, engine=engine)
        torch.cuda.synchronize()
        samples.append(s)
        done += n
    _check(torch, samples, mean, m2, 1e-12 if dtype == "float64" else 2e-6)


@pytest.mark.parametrize("engine,dtype", [("tile", "float64"), ("band", "float64"), ("band", "float32")])
def test_ulpda_chunked_mean_and_m2(torch, engine, dtype):
    from lmc_atomi_b200 import imaging as I
    dims, Cn = (48, 64), 2
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    x = torch.zeros((Cn, P), dtype=eng.tdt, device="cuda")
    xhat, warm = torch.zeros_like(x), torch.zeros_like(x)
    y = torch.zeros((Cn, 2, P), dtype=eng.tdt, device="cuda")
    mean, m2 = torch.zeros_like(x), torch.zeros_like(x)
    done, samples = 0, []
    for n in CHUNKS:
        s = torch.empty((Cn, n, P), dtype=eng.tdt, device="cuda")
        eng.ulpda(x, xhat, y, warm, n, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, seed=5, samples=s, mean=mean,
                  m2=m2, burnin=BURNIN, stat_count0=max(done - BURNIN, 0), iter_offset=done, engine=engine)
        torch.cuda.synchronize()
        samples.append(s)
        done += n
    _check(torch, samples, mean, m2, 1e-12 if dtype == "float64" else 2e-6)


def test_ncvx_one_iteration_launches_count_samples_once(torch):
    """The non-convex models advance by one-iteration launches (imaging.DeconvEngine.myula(ncvx=...)); the count
    handed to each launch must make the running mean the mean of the samples."""
    from lmc_atomi_b200 import imaging as I
    dims, Cn, N = (40, 48), 2, 7
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    x = torch.zeros((Cn, P), dtype=eng.tdt, device="cuda")
    mean, m2 = torch.zeros_like(x), torch.zeros_like(x)
    s = torch.empty((Cn, N, P), dtype=eng.tdt, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, seed=2, samples=s, mean=mean, m2=m2, burnin=2,
              ncvx=I.MoreauTerm("mc", 0.3, 15.0, True))
    torch.cuda.synchronize()
    sn = s.cpu().numpy()[:, 2:]
    assert np.abs(mean.cpu().numpy() - sn.mean(axis=1)).max() < 1e-11 * np.abs(sn).max()
    assert np.abs(m2.cpu().numpy() - ((sn - sn.mean(axis=1, keepdims=True)) ** 2).sum(axis=1)).max() < 1e-9 * np.abs(sn).max() ** 2
