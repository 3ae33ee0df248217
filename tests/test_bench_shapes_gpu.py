"""Parity of the kernel instantiations bench.py TIMES, on their own shapes:
512 x 512 band kernels -- img_ulpda_band_kernel<float,5,16,2>, img_myula_band_kernel<float,5,2> and their fp64 /
H6 / H7 siblings -- against the CPU oracle for a few iterations with the reference's host noise stream.
(32 bands per chain and exactly one column pair per thread: a regime the small-image tests never reach.)"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DIMS = (512, 512)


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(1.0, np.abs(b).max()))


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


@pytest.fixture(scope="module")
def problems():
    from oracle import deconv_ref as D
    img = D.synthetic_image(*DIMS)
    return {k: D.build_problem(img, kernel=k) for k in (5, 7)}


@pytest.fixture(scope="module")
def oracle_runs(problems):
    """Oracle trajectories, computed once per (sampler, kernel) and shared by the fp32 / fp64 cases."""
    from oracle import deconv_ref as D
    P = DIMS[0] * DIMS[1]
    out = {}
    for k, pb in problems.items():
        N = 3
        noise = D.image_noise(0, N, P)
        tv, l2 = pb["tv"](), pb["l2"]()
        my = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                              niter=N, noise=noise)
        ul = D.UnadjustedLangevinPrimalDual(l2, pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1.,
                                            niter=N, gfirst=False, noise=noise)
        out[k] = dict(noise=noise, myula=my, ulpda=ul, tv_updates=list(tv.updates), lsqr_itn=list(l2.lsqr_itn))
    return out


@pytest.mark.parametrize("kernel", [5, 7])
@pytest.mark.parametrize("dtype,tol", [("float64", 1e-12), ("float32", 1e-5)])
def test_myula_band_512(torch, problems, oracle_runs, kernel, dtype, tol):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    pb, run = problems[kernel], oracle_runs[kernel]
    P, N = DIMS[0] * DIMS[1], len(run["myula"])
    eng = I.DeconvEngine(DIMS, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    # fp32 = the timed band instantiation; a 512-wide fp64 band does not fit in shared memory twice per SM, so the
    # fp64 line of the bench runs (and is checked on) the tile kernel at this shape
    engine = "band" if eng.myula_band_supported() else "stream"
    assert dtype == "float64" or engine == "band"
    x = torch.zeros((1, P), dtype=eng.tdt, device="cuda")
    samples = torch.empty((1, N, P), dtype=eng.tdt, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=run["noise"][None], samples=samples,
              engine=engine)
    got = samples.double().cpu().numpy()[0]
    errs = [rel(got[n], run["myula"][n]) for n in range(N)]
    print(f"MYULA band 512x512 H{kernel} {dtype}: rel err per iteration {errs}")
    assert max(errs) < tol, errs
    if dtype == "float64":
        assert eng.read_counters()["tv_sweeps"] == sum(min(u + 1, 10 + D.TV_EXTRA_PASS) for u in run["tv_updates"])


@pytest.mark.parametrize("kernel", [5, 7])
@pytest.mark.parametrize("dtype,tol", [("float64", 1e-11), ("float32", 2e-5)])
def test_ulpda_band_512(torch, problems, oracle_runs, kernel, dtype, tol):
    from lmc_atomi_b200 import imaging as I
    pb, run = problems[kernel], oracle_runs[kernel]
    P, N = DIMS[0] * DIMS[1], len(run["ulpda"])
    eng = I.DeconvEngine(DIMS, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    engine = "band" if eng.band_supported() else "tile"
    assert dtype == "float64" or engine == "band"
    z = lambda *s: torch.zeros(s, dtype=eng.tdt, device="cuda")
    x, xhat, y, warm, samples = z(1, P), z(1, P), z(1, 2, P), z(1, P), z(1, N, P)
    eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50, noise=run["noise"][None],
              samples=samples, engine=engine)
    got = samples.double().cpu().numpy()[0]
    errs = [rel(got[n], run["ulpda"][n]) for n in range(N)]
    print(f"ULPDA band 512x512 H{kernel} {dtype}: rel err per iteration {errs}")
    assert max(errs) < tol, errs
    if dtype == "float64":
        assert eng.read_counters()["lsqr_iterations"] == sum(run["lsqr_itn"])


def test_myula_onchip_256_benchmark_shape(torch):
    """img_myula_onchip at 256 x 256 (FULL footprint instantiation, configs[3]) with several chains and Philox noise
    against the band kernel on the same draws: the many-chain regime of the bench (the single-chain oracle
    comparison of this shape is tests/test_img_onchip_gpu.py)."""
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims, Cn, N = (256, 256), 20, 4
    pb = D.build_problem(D.synthetic_image(*dims), kernel=5)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    out = {}
    for engine in ("band", "onchip"):
        x = torch.zeros((Cn, P), dtype=torch.float32, device="cuda")
        mean = torch.zeros_like(x)
        eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, seed=3, chain_offset=7, mean=mean,
                  engine=engine)
        torch.cuda.synchronize()
        out[engine] = (x.cpu().numpy(), mean.cpu().numpy())
    for a, b in zip(out["onchip"], out["band"]):
        assert rel(a, b) < 1e-5


@pytest.mark.parametrize("sampler", ["myula", "ulpda"])
def test_band_wide_rows_two_column_groups(torch, sampler):
    """A 24 x 776 fp32 image: rows wider than 2 x 256 threads (a second, partial column group per thread in every pass),
    one CTA per SM (139 KB of shared memory), rows staged by 3 KB bulk copies -- against the oracle for a few iterations."""
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims, N = (24, 776), 3
    pb = D.build_problem(D.synthetic_image(*dims), kernel=5)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    z = lambda *s: torch.zeros(s, dtype=eng.tdt, device="cuda")
    samples = z(1, N, P)
    if sampler == "myula":
        assert eng.myula_band_supported()
        want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                                gamma=pb["gamma_myula"], niter=N, noise=noise)
        eng.myula(z(1, P), N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples,
                  engine="band")
    else:
        assert eng.band_supported()
        want = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                              theta=1., niter=N, gfirst=False, noise=noise)
        eng.ulpda(z(1, P), z(1, P), z(1, 2, P), z(1, P), N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50,
                  noise=noise[None], samples=samples, engine="band")
    got = samples.double().cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 2e-5, errs
