"""Non-convex data terms of the reference (algs.L2_ncvx_tv, algs.py:22-291; models M2/M3/M5/M6/M8/M9 of
prox_lmc_deconv.py:106-113): MC-TV and ME-TV through the CUDA unit kernels (csrc/img_ncvx.cu) and through both
samplers behind the drop-in ``algs`` surface, against the oracle on the same seeded inputs (fp64, 1e-10)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300)


def _setup(dims, k=5, gamma=15.0):
    from oracle import deconv_ref as D
    from lmc_atomi_b200 import algs, compat
    pb = D.build_problem(D.synthetic_image(*dims), kernel=k, gamma_mc=gamma, gamma_me=gamma)
    hk = np.ones((k, k)) / k ** 2
    H = compat.Convolve2D(dims, h=hk, offset=(k // 2, k // 2))
    kw = dict(dims=dims, Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, lamda=0.3, gamma=gamma, niter=50, warm=True)
    mc = algs.L2_ncvx_tv(Op2=compat.Gradient(dims), isotropic=True, **kw)
    me = algs.L2_ncvx_tv(isotropic=True, **kw)
    return D, compat, algs, pb, mc, me


@pytest.mark.parametrize("dims,k,gamma", [((37, 45), 5, 15.0), ((40, 33), 6, 0.8), ((1, 9), 5, 2.0), ((7, 1), 7, 2.0)])
def test_operator_methods_match_oracle(dims, k, gamma):
    D, compat, algs, pb, mc, me = _setup(dims, k, gamma)
    rng = np.random.default_rng(1)
    x = pb["img"].ravel().astype(np.float64) + rng.normal(0, 20, dims[0] * dims[1])
    x[3 % x.size] = x[0]                      # exact ties -> zero differences (the 1e-9 branch)
    for mine, ref in ((mc, pb["l2_mc"]()), (me, pb["l2_me"]())):
        assert rel(mine.grad_moreau(x), ref.grad_moreau(x)) < 1e-11
        assert rel(mine.grad(x), ref.grad(x)) < 1e-11
        assert abs(mine(x) - ref(x)) <= 1e-10 * abs(ref(x))
        xa, xb = x.copy(), x.copy()
        pa, pbb = mine.prox(xa, 0.53), ref.prox(xb, 0.53)
        assert rel(xa, xb) < 1e-12            # the in-place pre-step of algs.py:208-223
        assert rel(pa, pbb) < 1e-9
        assert mine.count == 1 and rel(mine.x0, ref.x0) < 1e-9


def test_anisotropic_mc_tv_matches_oracle():
    D, compat, algs, pb, mc, me = _setup((29, 31), 5, 1.5)
    dims = (29, 31)
    mine = algs.L2_ncvx_tv(dims=dims, Op=mc.Op, Op2=compat.Gradient(dims), b=mc.b, sigma=mc.sigma, lamda=0.3, gamma=1.5,
                           isotropic=False, niter=50)
    ref = D.L2_ncvx_tv(dims=dims, Op=pb["H"], Op2=pb["Gop"], b=pb["y"].ravel(), sigma=mc.sigma, lamda=0.3, gamma=1.5,
                       isotropic=False, niter=50)
    x = np.random.default_rng(2).normal(0, 3, dims[0] * dims[1])
    assert rel(mine.grad_moreau(x), ref.grad_moreau(x)) < 1e-12
    xa, xb = x.copy(), x.copy()
    assert rel(mine.prox(xa, 0.4), ref.prox(xb, 0.4)) < 1e-9 and rel(xa, xb) < 1e-12


@pytest.mark.parametrize("which,k", [("mc", 5), ("me", 5), ("mc", 6), ("me", 7)])
def test_myula_with_nonconvex_data_term(which, k):
    dims, N = (40, 36), 6
    D, compat, algs, pb, mc, me = _setup(dims, k)
    ref = pb["l2_" + which]()
    want = D.MoreauYosidaUnadjustedLangevin(ref, pb["tv"](), pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                            niter=N, seed=4)
    mine = mc if which == "mc" else me
    tv = compat.TV(dims, sigma=0.3, niter=10)
    seen = []
    got = algs.MoreauYosidaUnadjustedLangevin(mine, tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                              niter=N, seed=4, callback=lambda x: seen.append(x.copy()))
    assert got.shape == want.shape and rel(got, want) < 1e-10
    assert rel(np.array(seen), want) < 1e-10
    got2 = algs.MoreauYosidaUnadjustedLangevin(mine, tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                               niter=N, seed=4)
    assert rel(got2, got) < 1e-13


@pytest.mark.parametrize("which,k,iso", [("mc", 5, True), ("me", 5, True), ("mc", 6, False), ("me", 7, True)])
def test_ulpda_with_nonconvex_data_term(which, k, iso):
    dims, N = (40, 36), 5
    D, compat, algs, pb, mc, me = _setup(dims, k)
    ref = pb["l2_" + which]()
    g_ref = pb["l1iso"]() if iso else pb["l1"]()
    want, wy = D.UnadjustedLangevinPrimalDual(ref, g_ref, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1.,
                                              niter=N, seed=1, gfirst=False, returny=True)
    mine = mc if which == "mc" else me
    g = compat.L21(ndim=2, sigma=0.3) if iso else compat.L1(sigma=0.3)
    got, gy = algs.UnadjustedLangevinPrimalDual(mine, g, compat.Gradient(dims), pb["x0"], tau=pb["tau0"],
                                                mu=pb["mu0"], theta=1., niter=N, seed=1, gfirst=False, returny=True)
    assert rel(got, want) < 1e-9 and rel(gy, wy) < 1e-9
    assert mine.count == N and rel(mine.x0, ref.x0) < 1e-9
    # gfirst=True (algs.py:435-441) with the non-convex term: in-kernel since r02
    ref2 = pb["l2_" + which]()
    want2 = D.UnadjustedLangevinPrimalDual(ref2, g_ref, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1., niter=3,
                                           seed=2, gfirst=True)
    mine.x0, mine.count = None, 0
    got2 = algs.UnadjustedLangevinPrimalDual(mine, g, compat.Gradient(dims), pb["x0"], tau=pb["tau0"], mu=pb["mu0"],
                                             theta=1., niter=3, seed=2, gfirst=True)
    assert rel(got2, want2) < 1e-9


def test_fp32_engine_with_moreau_term_tracks_fp64():
    """fp32 production arithmetic with the MC-TV term: many chains, Philox noise, against the fp64 engine."""
    import torch
    from lmc_atomi_b200 import deconv_bench, imaging
    dev = torch.device("cuda", 0)
    dims, N, Cn = (64, 48), 3, 3
    term = imaging.MoreauTerm("mc", 0.3, 15.0, True)
    outs = []
    for dt in ("float32", "float64"):
        _, eng = deconv_bench.make_engine(dims, dev, dt)
        noise = imaging.philox_image_noise(dims, Cn, N, seed=7, dtype="float64", device=dev).to(eng.tdt)
        x = torch.zeros((Cn, dims[0] * dims[1]), dtype=eng.tdt, device=dev)
        xh, y, w = x.clone(), torch.zeros((Cn, 2, x.shape[1]), dtype=eng.tdt, device=dev), x.clone()
        eng.ulpda(x, xh, y, w, N, 0.534375, 1.0, 1.0, 0.3, True, False, 50, noise=noise, ncvx=term)
        torch.cuda.synchronize()
        outs.append(x.double().cpu().numpy())
    assert rel(outs[0], outs[1]) < 2e-4


@pytest.mark.parametrize("engine", ["stream", "band"])
@pytest.mark.parametrize("which,k,dims", [("mc", 5, (40, 36)), ("me", 5, (40, 36)), ("mc", 6, (33, 45)), ("me", 7, (48, 64))])
def test_myula_nonconvex_in_kernel_multi_iteration_launch(which, k, dims, engine):
    """The MC-TV stencil / ME-TV inner prox evaluated INSIDE the MYULA kernels (lmc_img_io::ncvx): one launch of N
    iterations, band and tile kernels, against the oracle's L2_ncvx_tv on the reference noise stream (fp64)."""
    import torch
    from lmc_atomi_b200 import imaging as I
    N = 7
    D, compat, algs, pb, mc, me = _setup(dims, k)
    P = dims[0] * dims[1]
    noise = D.image_noise(5, N, P)
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2_" + which](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                            gamma=pb["gamma_myula"], niter=N, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    term = I.MoreauTerm(which, 0.3, 15.0, True, 50, 1e-4)
    x = torch.zeros((1, P), dtype=torch.float64, device="cuda")
    samples = torch.empty((1, N, P), dtype=torch.float64, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples, ncvx=term,
              engine=engine)
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 1e-10, errs


@pytest.mark.parametrize("which", ["mc", "me"])
def test_myula_nonconvex_in_kernel_many_chains_fp32(which):
    """fp32, more chains than a wave, Philox noise, running mean: band kernel == tile kernel, and both track fp64."""
    import torch
    from lmc_atomi_b200 import deconv_bench, imaging as I
    dev = torch.device("cuda", 0)
    dims, N, Cn = (64, 96), 4, 40
    term = I.MoreauTerm(which, 0.3, 15.0, True, 50, 1e-4)
    out = {}
    for dt, engine in (("float32", "band"), ("float32", "stream"), ("float64", "band")):
        _, eng = deconv_bench.make_engine(dims, dev, dt)
        noise = I.philox_image_noise(dims, Cn, N, seed=7, dtype="float64", device=dev).to(eng.tdt)
        x = torch.zeros((Cn, dims[0] * dims[1]), dtype=eng.tdt, device=dev)
        mean = torch.zeros_like(x)
        eng.myula(x, N, 0.1125, 0.5625, 0.3, 10, 1e-4, noise=noise, mean=mean, ncvx=term, engine=engine)
        torch.cuda.synchronize()
        out[(dt, engine)] = (x.double().cpu().numpy(), mean.double().cpu().numpy())
    for a, b in zip(out[("float32", "band")], out[("float32", "stream")]):
        assert rel(a, b) < 1e-5
    for a, b in zip(out[("float32", "band")], out[("float64", "band")]):
        assert rel(a, b) < 2e-5


@pytest.mark.parametrize("engine", ["tile", "band"])
@pytest.mark.parametrize("which,k,dims,iso", [("mc", 5, (40, 36), True), ("me", 5, (40, 36), True),
                                               ("mc", 6, (33, 45), False), ("me", 7, (48, 64), True)])
def test_ulpda_nonconvex_in_kernel_multi_iteration_launch(which, k, dims, iso, engine):
    """L2_ncvx_tv.prox's pre-step evaluated INSIDE the ULPDA kernels (lmc_img_io::ncvx): one launch of N iterations
    on both kernels against the oracle (fp64, reference noise stream), with the prior the driver pairs with the term."""
    import torch
    from lmc_atomi_b200 import imaging as I
    N = 6
    D, compat, algs, pb, mc, me = _setup(dims, k)
    P = dims[0] * dims[1]
    noise = D.image_noise(3, N, P)
    ref = pb["l2_" + which]() if iso else D.L2_ncvx_tv(dims=dims, Op=pb["H"], Op2=pb["Gop"], b=pb["y"].ravel(),
                                                       sigma=1 / 0.75 ** 2, lamda=0.3, gamma=15.0, isotropic=False,
                                                       niter=50)
    g_ref = pb["l1iso"]() if which == "me" else pb["l1"]()
    want = D.UnadjustedLangevinPrimalDual(ref, g_ref, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1., niter=N,
                                          gfirst=False, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float64")
    term = I.MoreauTerm(which, 0.3, 15.0, iso, 50, 1e-4)
    z = lambda *s: torch.zeros(s, dtype=torch.float64, device="cuda")
    x, xhat, y, warm, samples = z(1, P), z(1, P), z(1, 2, P), z(1, P), z(1, N, P)
    eng.ulpda(x, xhat, y, warm, N, pb["tau0"], pb["mu0"], 1.0, 0.3, which == "me", False, 50, noise=noise[None],
              samples=samples, ncvx=term, engine=engine)
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    assert max(errs) < 1e-9, errs


@pytest.mark.parametrize("which", ["mc", "me"])
def test_ulpda_nonconvex_in_kernel_many_chains(which):
    """Several chains with their own Philox noise and running means: band == tile kernel (fp64), fp32 tracks fp64."""
    import torch
    from lmc_atomi_b200 import deconv_bench, imaging as I
    dev = torch.device("cuda", 0)
    dims, N, Cn = (64, 96), 3, 9
    term = I.MoreauTerm(which, 0.3, 15.0, True, 50, 1e-4)
    out = {}
    for dt, engine in (("float64", "band"), ("float64", "tile"), ("float32", "band")):
        _, eng = deconv_bench.make_engine(dims, dev, dt)
        noise = I.philox_image_noise(dims, Cn, N, seed=7, dtype="float64", device=dev).to(eng.tdt)
        z = lambda *s: torch.zeros(s, dtype=eng.tdt, device=dev)
        x, xhat, y, warm, mean = z(Cn, dims[0] * dims[1]), z(Cn, dims[0] * dims[1]), z(Cn, 2, dims[0] * dims[1]), \
            z(Cn, dims[0] * dims[1]), z(Cn, dims[0] * dims[1])
        eng.ulpda(x, xhat, y, warm, N, 0.534375, 1.0, 1.0, 0.3, True, False, 50, noise=noise, mean=mean, ncvx=term,
                  engine=engine)
        torch.cuda.synchronize()
        out[(dt, engine)] = (x.double().cpu().numpy(), mean.double().cpu().numpy())
    for a, b in zip(out[("float64", "band")], out[("float64", "tile")]):
        assert rel(a, b) < 1e-10
    for a, b in zip(out[("float32", "band")], out[("float64", "band")]):
        assert rel(a, b) < 2e-4
