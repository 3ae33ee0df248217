"""GPU parity tests of the image (deconvolution) path against the CPU oracle
(oracle/deconv_ref.py): unit operators, TV prox, LSQR-solved L2 prox, and the MYULA /
ULPDA trajectories on the same inputs and the same noise.

Tolerances (north_star): fp64 mode 1e-12 relative, fp32 mode 1e-5 relative, both
normalised by the dynamic range of the compared array (images live in [0, 255])."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = [(64, 64), (40, 52), (33, 95)]     # tile-aligned, ragged, ragged with nx % 4 != 0


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(1.0, np.abs(b).max()))


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


def dev_t(torch, a, dtype="float64"):
    return torch.as_tensor(np.ascontiguousarray(a)).to("cuda", torch.float64 if dtype == "float64" else torch.float32)


# --------------------------------------------------------------------------- operators
@pytest.mark.parametrize("k", [5, 6, 7])
@pytest.mark.parametrize("dims", SIZES)
def test_blur_matches_oracle(torch, k, dims):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    rng = np.random.default_rng(k)
    h = np.ones((k, k)) / k ** 2
    H = D.Convolve2D(dims, h, (k // 2, k // 2))
    spec = I.BlurSpec(dims, I.pad_filter(h, (k // 2, k // 2)))
    x = rng.standard_normal((3, dims[0] * dims[1])) * 50
    xd = dev_t(torch, x)
    got_f = I.blur(spec, xd).cpu().numpy()
    got_a = I.blur(spec, xd, adjoint=True).cpu().numpy()
    for c in range(3):
        assert rel(got_f[c], H.matvec(x[c])) < 1e-13
        assert rel(got_a[c], H.rmatvec(x[c])) < 1e-13
    # adjoint identity on the device operators themselves
    y = rng.standard_normal(dims[0] * dims[1])
    lhs = float(got_f[0] @ y)
    rhs = float(x[0] @ I.blur(spec, dev_t(torch, y), adjoint=True).cpu().numpy())
    assert abs(lhs - rhs) < 1e-9 * max(1.0, abs(lhs))


def test_blur_nonuniform_separable_filter(torch):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims = (37, 61)
    h = np.outer([1., 2., 3.], [0.5, 1., 0.25, 2., 1.])
    h /= h.sum()
    H = D.Convolve2D(dims, h, (1, 2))
    spec = I.BlurSpec(dims, I.pad_filter(h, (1, 2)))
    x = np.random.default_rng(0).standard_normal(dims[0] * dims[1])
    assert rel(I.blur(spec, dev_t(torch, x)).cpu().numpy(), H.matvec(x)) < 1e-13
    assert rel(I.blur(spec, dev_t(torch, x), adjoint=True).cpu().numpy(), H.rmatvec(x)) < 1e-13
    with pytest.raises(NotImplementedError):
        I.BlurSpec(dims, np.array([[1., 2.], [3., 5.]]))     # rank 2: not separable


@pytest.mark.parametrize("dims", SIZES)
def test_gradient_matches_oracle(torch, dims):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    rng = np.random.default_rng(1)
    G = D.Gradient(dims)
    x = rng.standard_normal((2, dims[0] * dims[1]))
    v = rng.standard_normal((2, 2 * dims[0] * dims[1]))
    gf = I.grad(dims, dev_t(torch, x)).cpu().numpy().reshape(2, -1)
    ga = I.grad(dims, dev_t(torch, v), adjoint=True).cpu().numpy()
    for c in range(2):
        assert np.array_equal(gf[c], G.matvec(x[c]))
        assert rel(ga[c], G.rmatvec(v[c])) < 1e-15


@pytest.mark.parametrize("iso", [True, False])
def test_proxdual_matches_oracle(torch, iso):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    rng = np.random.default_rng(2)
    v = rng.standard_normal((2, 5000)) * 0.6
    op = D.L21(2, 0.3) if iso else D.L1(0.3)
    want = op.proxdual(v.ravel().copy(), 1.0).reshape(2, -1)
    got = I.proxdual(dev_t(torch, v), 0.3, isotropic=iso).cpu().numpy()
    assert rel(got, want) < 1e-14


# --------------------------------------------------------------------------- TV prox
@pytest.mark.parametrize("dims", SIZES)
@pytest.mark.parametrize("gamma,niter", [(0.16875, 10), (15.0, 50), (0.16875, 1)])
def test_tv_prox_matches_oracle(torch, dims, gamma, niter):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    img = D.synthetic_image(*dims).astype(np.float64)
    rng = np.random.default_rng(3)
    xs = np.stack([img + rng.normal(0, s, img.shape) for s in (0.5, 5.0, 20.0)])
    counters = torch.zeros(8, dtype=torch.int64, device="cuda")
    got = I.tv_prox(dims, dev_t(torch, xs.reshape(3, -1)), gamma, niter, 1e-4, counters=counters).cpu().numpy()
    sweeps = 0
    for c in range(3):
        want, nupd = D.tv_prox_2d(xs[c], gamma, niter, 1e-4)
        sweeps += min(nupd + 1, niter + D.TV_EXTRA_PASS)
        assert rel(got[c], want.ravel()) < 1e-12, (c, nupd)
    assert int(counters[0].item()) == sweeps     # per-chain early exit reproduced


def test_tv_prox_zero_image(torch):
    """x = 0 (first MYULA iteration from x0 = 0): obj <= 1e-10 never triggers the break."""
    from lmc_atomi_b200 import imaging as I
    got = I.tv_prox((48, 48), torch.zeros((1, 48 * 48), dtype=torch.float64, device="cuda"), 0.2, 10, 1e-4)
    assert float(got.abs().max()) == 0.0


# --------------------------------------------------------------------------- L2 prox (LSQR)
@pytest.mark.parametrize("k", [5, 6])
def test_l2_prox_matches_scipy_lsqr(torch, k):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims = (40, 52)
    img = D.synthetic_image(*dims)
    pb = D.build_problem(img, kernel=k)
    l2 = pb["l2"]()
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    warm = torch.zeros((1, dims[0] * dims[1]), dtype=torch.float64, device="cuda")
    rng = np.random.default_rng(4)
    x = img.ravel().astype(np.float64)
    for it in range(3):    # warm start carried across calls, as pyproximal.L2(warm=True) does
        arg = x + rng.standard_normal(x.shape) * 3
        want = l2.prox(arg.copy(), pb["tau0"])
        got = eng.l2_prox(dev_t(torch, arg[None]), pb["tau0"], warm).cpu().numpy()[0]
        assert rel(got, want) < 1e-11, it
        x = want
    c = eng.read_counters()
    assert c["lsqr_iterations"] == sum(l2.lsqr_itn) and c["lsqr_solves"] == 3


# --------------------------------------------------------------------------- samplers
def _problem(dims, kernel=5):
    from oracle import deconv_ref as D
    img = D.synthetic_image(*dims)
    return D.build_problem(img, kernel=kernel)


@pytest.mark.parametrize("dims,kernel", [((64, 64), 5), ((40, 52), 5), ((33, 95), 6), ((64, 64), 7)])
def test_myula_trajectory_fp64(torch, dims, kernel):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    N = 25
    pb = _problem(dims, kernel)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    tv = pb["tv"]()
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                            niter=N, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    x = torch.zeros((1, P), dtype=torch.float64, device="cuda")
    samples = torch.empty((1, N, P), dtype=torch.float64, device="cuda")
    mean = torch.zeros((1, P), dtype=torch.float64, device="cuda")
    m2 = torch.zeros((1, P), dtype=torch.float64, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples,
              mean=mean, m2=m2)
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 1e-12, errs
    assert rel(mean.cpu().numpy()[0], want.mean(axis=0)) < 1e-12
    assert rel(m2.cpu().numpy()[0] / N, want.var(axis=0)) < 1e-10
    assert eng.read_counters()["tv_sweeps"] == sum(min(u + 1, 10 + D.TV_EXTRA_PASS) for u in tv.updates)


@pytest.mark.parametrize("engine", ["tile", "band"])
@pytest.mark.parametrize("dims,kernel,iso,gfirst", [((64, 64), 5, True, False), ((40, 52), 5, False, False),
                                                    ((33, 95), 6, True, True), ((64, 64), 7, True, False),
                                                    ((40, 52), 6, True, False), ((23, 30), 7, False, True)])
def test_ulpda_trajectory_fp64(torch, dims, kernel, iso, gfirst, engine):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    N = 12
    pb = _problem(dims, kernel)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    l2 = pb["l2"]()
    g = pb["l1iso"]() if iso else pb["l1"]()
    want, wanty = D.UnadjustedLangevinPrimalDual(l2, g, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1., niter=N,
                                                 gfirst=gfirst, returny=True, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    z = lambda *s: torch.zeros(s, dtype=torch.float64, device="cuda")
    x, xhat, y, warm = z(1, P), z(1, P), z(1, 2, P), z(1, P)
    samples, ys = z(1, N, P), z(1, N, 2, P)
    eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, iso, gfirst, 50, noise=noise[None],
              samples=samples, ysamples=ys, engine=engine)
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 1e-11, errs
    assert rel(ys.cpu().numpy()[0].reshape(N, -1), wanty) < 1e-11
    c = eng.read_counters()
    assert c["lsqr_iterations"] == sum(l2.lsqr_itn), (c, l2.lsqr_itn)


@pytest.mark.parametrize("engine", ["stream", "onchip"])
def test_myula_fp32_within_tolerance(torch, engine):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims, N = (64, 64), 25
    pb = _problem(dims)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                            gamma=pb["gamma_myula"], niter=N, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    x = torch.zeros((1, P), dtype=torch.float32, device="cuda")
    samples = torch.empty((1, N, P), dtype=torch.float32, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples,
              engine=engine)
    got = samples.cpu().numpy()[0]
    err = max(rel(got[n], want[n]) for n in range(N))
    print(f"MYULA fp32 ({engine}) free-run rel err over {N} iterations: {err:.3e}")
    assert err < 1e-5


@pytest.mark.parametrize("engine", ["tile", "band"])
def test_ulpda_fp32_within_tolerance(torch, engine):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims, N = (64, 64), 12
    pb = _problem(dims)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    want = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                          theta=1., niter=N, gfirst=False, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    z = lambda *s: torch.zeros(s, dtype=torch.float32, device="cuda")
    x, xhat, y, warm, samples = z(1, P), z(1, P), z(1, 2, P), z(1, P), z(1, N, P)
    eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50, noise=noise[None],
              samples=samples, engine=engine)
    got = samples.cpu().numpy()[0]
    err = max(rel(got[n], want[n]) for n in range(N))
    print(f"ULPDA fp32 ({engine}) free-run rel err over {N} iterations: {err:.3e}")
    # LSQR stops on atol=btol=1e-6: the solve itself is only 1e-6-accurate, so fp32 rounding can move
    # the stopping iteration; the tolerance reflects the solver's own accuracy.
    assert err < 2e-5


def test_batched_chains_are_independent(torch):
    """C chains with different noise == C separate single-chain runs (per-chain early exit)."""
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    dims, N, Cn = (40, 52), 8, 3
    pb = _problem(dims)
    P = dims[0] * dims[1]
    noise = np.stack([D.image_noise(s, N, P) for s in range(Cn)])
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    x = torch.zeros((Cn, P), dtype=torch.float64, device="cuda")
    samples = torch.empty((Cn, N, P), dtype=torch.float64, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise, samples=samples)
    got = samples.cpu().numpy()
    for c in range(Cn):
        want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                                gamma=pb["gamma_myula"], niter=N, noise=noise[c])
        assert rel(got[c], want) < 1e-12
    z = lambda *s: torch.zeros(s, dtype=torch.float64, device="cuda")
    x, xhat, y, warm, samples = z(Cn, P), z(Cn, P), z(Cn, 2, P), z(Cn, P), z(Cn, N, P)
    eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50, noise=noise, samples=samples)
    got = samples.cpu().numpy()
    for c in range(Cn):
        want = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                              theta=1., niter=N, gfirst=False, noise=noise[c])
        assert rel(got[c], want) < 1e-11


def test_ulpda_band_groups_match_tile_kernel(torch):
    """Band kernel with chain groups smaller than the batch (each group runs all N iterations before the
    next) == tile kernel advancing all chains in lockstep; in-kernel Philox noise, running statistics."""
    from lmc_atomi_b200 import imaging as I
    dims, N, Cn = (70, 96), 6, 5
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    assert eng.band_supported()
    out = {}
    for engine, group in (("tile", 0), ("band", 2), ("band", 0)):
        z = lambda *s: torch.zeros(s, dtype=torch.float64, device="cuda")
        x, xhat, y, warm, mean, m2 = z(Cn, P), z(Cn, P), z(Cn, 2, P), z(Cn, P), z(Cn, P), z(Cn, P)
        eng.counters.zero_()
        eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50, seed=5, chain_offset=77,
                  mean=mean, m2=m2, burnin=1, engine=engine, group=group)
        torch.cuda.synchronize()
        out[(engine, group)] = [t.cpu().numpy() for t in (x, xhat, y, warm, mean, m2)] + [eng.read_counters()]
    ref = out[("tile", 0)]
    for key in (("band", 2), ("band", 0)):
        for a, b in zip(out[key][:6], ref[:6]):
            assert rel(a, b) < 1e-11, key
        assert out[key][6] == ref[6], key           # same LSQR iteration counts
    assert np.abs(ref[0]).max() > 10


@pytest.mark.parametrize("dims,Cn,dtype,tol", [((48, 40), 150, "float64", 1e-11), ((31, 33), 40, "float64", 1e-11),
                                               ((48, 64), 130, "float32", 3e-5)])
def test_ulpda_band_work_queue_many_chains(torch, dims, Cn, dtype, tol):
    """More work items than resident CTAs (the band kernel draws them from its global queue over several
    rounds), ragged last band, odd width (one column per thread): band kernel == tile kernel."""
    from lmc_atomi_b200 import imaging as I
    N = 3
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    assert eng.band_supported()
    out = {}
    for engine in ("tile", "band"):
        z = lambda *s: torch.zeros(s, dtype=eng.tdt, device="cuda")
        x, xhat, y, warm, mean = z(Cn, P), z(Cn, P), z(Cn, 2, P), z(Cn, P), z(Cn, P)
        eng.counters.zero_()
        eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, True, False, 50, seed=9, mean=mean,
                  engine=engine)
        torch.cuda.synchronize()
        out[engine] = [t.double().cpu().numpy() for t in (x, y, warm, mean)] + [eng.read_counters()]
    for i, (a, b) in enumerate(zip(out["band"][:4], out["tile"][:4])):
        # fp32: the two kernels sum the LSQR norms in different orders, so a solve may stop one iteration apart;
        # x (values up to 255) stays within the solver's 1e-6 accuracy, the dual y (|y| <= 0.3) sees that as ~1e-4
        assert rel(a, b) < (1e-3 if (dtype == "float32" and i == 1) else tol)
    if dtype == "float64":
        assert out["band"][4] == out["tile"][4]
    assert len({tuple(np.round(out["band"][0][c][:4], 6)) for c in range(Cn)}) == Cn     # chains differ (own noise)


@pytest.mark.parametrize("dims,kernel", [((64, 64), 5), ((40, 52), 5), ((23, 30), 5), ((70, 96), 5), ((9, 4), 5),
                                         ((33, 95), 5), ((41, 77), 6), ((20, 301), 7),
                                         ((40, 52), 6), ((33, 96), 7), ((64, 64), 7)])
def test_myula_band_trajectory_fp64(torch, dims, kernel):
    """Row-band MYULA kernel (img_myula_band.inl) against the oracle: ragged last band / sub-band, narrow images,
    the H6 / H7 blurs (one column per thread in the gradient phase)."""
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    N = 14
    pb = _problem(dims, kernel)
    P = dims[0] * dims[1]
    noise = D.image_noise(2, N, P)
    tv = pb["tv"]()
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                            niter=N, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    assert eng.myula_band_supported()
    z = lambda *s: torch.zeros(s, dtype=torch.float64, device="cuda")
    x, samples, mean, m2 = z(1, P), z(1, N, P), z(1, P), z(1, P)
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples,
              mean=mean, m2=m2, engine="band")
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 1e-12, errs
    assert rel(mean.cpu().numpy()[0], want.mean(axis=0)) < 1e-12
    assert eng.read_counters()["tv_sweeps"] == sum(min(u + 1, 10 + D.TV_EXTRA_PASS) for u in tv.updates)


@pytest.mark.parametrize("dims,Cn,dtype,tol", [((48, 40), 150, "float64", 1e-11), ((64, 128), 90, "float32", 1e-5)])
def test_myula_band_many_chains_match_tile_kernel(torch, dims, Cn, dtype, tol):
    """More work items than resident CTAs, in-kernel Philox noise, running statistics: band == tile kernel."""
    from lmc_atomi_b200 import imaging as I
    N = 4
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype=dtype)
    out = {}
    for engine in ("stream", "band"):
        z = lambda *s: torch.zeros(s, dtype=eng.tdt, device="cuda")
        x, mean, m2 = z(Cn, P), z(Cn, P), z(Cn, P)
        eng.counters.zero_()
        eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, seed=11, chain_offset=5, mean=mean, m2=m2,
                  burnin=1, engine=engine)
        torch.cuda.synchronize()
        out[engine] = [t.double().cpu().numpy() for t in (x, mean, m2)] + [eng.read_counters()["tv_sweeps"]]
    for a, b in zip(out["band"][:2], out["stream"][:2]):
        assert rel(a, b) < tol
    assert rel(out["band"][2], out["stream"][2]) < 100 * tol
    if dtype == "float64":
        assert out["band"][3] == out["stream"][3]
