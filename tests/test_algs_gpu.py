"""Drop-in ``algs`` surface (algs.py:295-587) against the oracle, called the way
prox_lmc_deconv.py:455-473 calls it: operator objects in, (niter, P) float64 samples out,
seed-driven noise, per-iteration callback with a host array."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300)


def _setup(dims, k=5):
    from oracle import deconv_ref as D
    from lmc_atomi_b200 import compat
    pb = D.build_problem(D.synthetic_image(*dims), kernel=k)
    hk = np.ones((k, k)) / k ** 2
    H = compat.Convolve2D(dims, h=hk, offset=(k // 2, k // 2))
    l2 = compat.L2(Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, niter=50, warm=True)
    return D, compat, pb, H, l2


@pytest.mark.parametrize("dims,k", [((48, 40), 5), ((37, 61), 6)])
def test_myula_dropin_matches_oracle_with_callback(dims, k):
    from lmc_atomi_b200 import algs
    D, compat, pb, H, l2 = _setup(dims, k)
    N = 12
    seen_ref, seen = [], []
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                            gamma=pb["gamma_myula"], niter=N, seed=3,
                                            callback=lambda x: seen_ref.append(x.copy()))
    tv = compat.TV(dims, sigma=0.3, niter=10)
    got = algs.MoreauYosidaUnadjustedLangevin(l2, tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                              niter=N, seed=3, callback=lambda x: seen.append(x.copy()))
    assert got.shape == want.shape == (N, dims[0] * dims[1]) and got.dtype == np.float64
    assert rel(got, want) < 1e-11
    assert len(seen) == N and rel(np.array(seen), np.array(seen_ref)) < 1e-11
    # no callback: one launch for all iterations, same samples
    got2 = algs.MoreauYosidaUnadjustedLangevin(l2, tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                               niter=N, seed=3)
    assert rel(got2, got) < 1e-13


@pytest.mark.parametrize("iso,gfirst", [(True, False), (False, False), (True, True)])
def test_ulpda_dropin_matches_oracle(iso, gfirst):
    from lmc_atomi_b200 import algs
    dims, N = (48, 40), 8
    D, compat, pb, H, l2 = _setup(dims)
    g_ref = pb["l1iso"]() if iso else pb["l1"]()
    want, wy = D.UnadjustedLangevinPrimalDual(pb["l2"](), g_ref, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                              theta=1., niter=N, seed=1, gfirst=gfirst, returny=True)
    g = compat.L21(ndim=2, sigma=0.3) if iso else compat.L1(sigma=0.3)
    seen = []
    got, gy = algs.UnadjustedLangevinPrimalDual(l2, g, compat.Gradient(dims), pb["x0"], tau=pb["tau0"],
                                                mu=pb["mu0"], theta=1., niter=N, seed=1, gfirst=gfirst,
                                                returny=True, callback=lambda x: seen.append(x.copy()))
    assert got.shape == (N, dims[0] * dims[1]) and gy.shape == (N, 2 * dims[0] * dims[1])
    assert rel(got, want) < 1e-10 and rel(gy, wy) < 1e-10
    assert len(seen) == N and rel(np.array(seen), want) < 1e-10
    assert l2.x0 is not None and l2.x0.shape == (dims[0] * dims[1],)   # warm start kept, like pyproximal.L2


def test_ulpda_variable_steps_and_z():
    from lmc_atomi_b200 import algs
    dims, N = (40, 44), 6
    D, compat, pb, H, l2 = _setup(dims)
    taus = pb["tau0"] * np.linspace(1.0, 0.5, N)
    mus = np.linspace(1.0, 0.8, N)
    z = np.random.default_rng(5).normal(0, 0.1, dims[0] * dims[1])
    want = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], taus, mus, z=z,
                                          theta=0.7, niter=N, seed=2, gfirst=False)
    got = algs.UnadjustedLangevinPrimalDual(l2, compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"],
                                            tau=taus, mu=mus, z=z, theta=0.7, niter=N, seed=2, gfirst=False)
    assert rel(got, want) < 1e-10


def test_compat_operators_match_oracle():
    dims = (37, 61)
    D, compat, pb, H, l2 = _setup(dims, 6)
    rng = np.random.default_rng(0)
    x = rng.normal(size=dims[0] * dims[1]) * 50
    l2r, tvr, gr = pb["l2"](), pb["tv"](), pb["Gop"]
    assert rel(H.matvec(x), pb["H"].matvec(x)) < 1e-13 and rel(H.rmatvec(x), pb["H"].rmatvec(x)) < 1e-13
    assert rel(H.H * x, pb["H"].rmatvec(x)) < 1e-13
    assert abs(l2(x) - l2r(x)) / l2r(x) < 1e-12
    assert rel(l2.grad(x), l2r.grad(x)) < 1e-12
    G = compat.Gradient(dims)
    assert rel(G.matvec(x), gr.matvec(x)) < 1e-14
    v = rng.normal(size=2 * dims[0] * dims[1])
    assert rel(G.rmatvec(v), gr.rmatvec(v)) < 1e-14
    tv = compat.TV(dims, sigma=0.3, niter=10)
    assert abs(tv(x) - tvr(x)) / tvr(x) < 1e-12
    assert rel(tv.prox(x, 0.5625), tvr.prox(x, 0.5625)) < 1e-12
    for mine, ref in ((compat.L21(2, 0.3), pb["l1iso"]()), (compat.L1(0.3), pb["l1"]())):
        assert abs(mine(v) - ref(v)) / ref(v) < 1e-12
        assert rel(mine.proxdual(v, 1.0), ref.proxdual(v, 1.0)) < 1e-14
        assert rel(mine.prox(v, 0.7), ref.prox(v, 0.7)) < 1e-13
    assert rel(l2.prox(x, 0.53), l2r.prox(x.copy(), 0.53)) < 1e-10


def test_unsupported_combination_raises():
    from lmc_atomi_b200 import algs
    dims = (32, 32)
    D, compat, pb, H, l2 = _setup(dims)
    with pytest.raises(NotImplementedError):
        algs.MoreauYosidaUnadjustedLangevin(l2, compat.L1(0.3), pb["x0"], tau=0.1, gamma=0.5, niter=2)
    with pytest.raises(NotImplementedError):
        algs.UnadjustedLangevinPrimalDual(l2, compat.TV(dims), compat.Gradient(dims), pb["x0"], 0.5, 1.0, niter=2)


def test_ulpda_callable_niter_follows_the_prox_call_count():
    """``niter`` as a function of the prox call count (algs.py:203-207, pyproximal.L2): the LSQR iteration limit of
    sampler iteration n is niter(count0 + n + 1); the wrapper advances one launch per iteration."""
    from lmc_atomi_b200 import algs
    dims, N = (40, 36), 7
    D, compat, pb, H, _ = _setup(dims)
    sched = lambda count: 2 + count                # 3, 4, 5, ...: below the 10 iterations an unrestricted solve takes here
    ref = pb["l2"]()
    ref.niter = sched
    want = D.UnadjustedLangevinPrimalDual(ref, pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1.,
                                          niter=N, seed=4, gfirst=False)
    assert ref.lsqr_itn == [3, 4, 5, 6, 7, 8, 9]   # the limits bind: the schedule is what is being tested
    l2 = compat.L2(Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, niter=sched, warm=True)
    got = algs.UnadjustedLangevinPrimalDual(l2, compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"],
                                            tau=pb["tau0"], mu=pb["mu0"], theta=1., niter=N, seed=4, gfirst=False)
    assert rel(got, want) < 1e-10
    assert l2.count == N
    # the operator's own prox honours the schedule too (count keeps running)
    x = np.random.default_rng(0).normal(100, 20, dims[0] * dims[1])
    a, b = ref.prox(x.copy(), 0.3), l2.prox(x.copy(), 0.3)
    assert ref.lsqr_itn[-1] <= sched(N + 1) and ref.count == l2.count == N + 1 and rel(b, a) < 1e-10


def test_linear_term_q_alpha_qgrad():
    """alpha q^T x of pyproximal.L2 / algs.L2_ncvx_tv (algs.py:186-187, 226-227, 288-289): unit operators, both samplers
    and the MAP loop against the oracle."""
    from lmc_atomi_b200 import algs, primaldual
    dims, N = (40, 36), 5
    D, compat, pb, H, _ = _setup(dims)
    P = dims[0] * dims[1]
    rng = np.random.default_rng(3)
    q, alpha = rng.normal(0, 0.05, P), 0.7
    y, sig = pb["y"].ravel(), 1 / 0.75 ** 2
    ref = lambda **kw: D.L2(Op=pb["H"], b=y, sigma=sig, niter=50, warm=True, q=q, alpha=alpha, **kw)
    mine = lambda **kw: compat.L2(Op=H, b=y, sigma=sig, niter=50, warm=True, q=q, alpha=alpha, **kw)
    x = rng.normal(100, 20, P)
    a, b = ref(), mine()
    assert abs(b(x) - a(x)) < 1e-9 * abs(a(x)) and rel(b.grad(x), a.grad(x)) < 1e-12
    assert rel(mine(qgrad=False).grad(x), ref(qgrad=False).grad(x)) < 1e-12
    assert rel(b.prox(x.copy(), 0.3), a.prox(x.copy(), 0.3)) < 1e-10
    # ULPDA: alpha q folds into the linear term z
    z = rng.normal(0, 0.1, P)
    want = D.UnadjustedLangevinPrimalDual(ref(), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], z=z,
                                          theta=1., niter=N, seed=6, gfirst=False)
    got = algs.UnadjustedLangevinPrimalDual(mine(), compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"],
                                            tau=pb["tau0"], mu=pb["mu0"], z=z, theta=1., niter=N, seed=6, gfirst=False)
    assert rel(got, want) < 1e-10
    # MYULA: alpha q is a constant extra gradient term (dropped when qgrad is False)
    for qgrad in (True, False):
        want = D.MoreauYosidaUnadjustedLangevin(ref(qgrad=qgrad), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                                gamma=pb["gamma_myula"], niter=N, seed=6)
        got = algs.MoreauYosidaUnadjustedLangevin(mine(qgrad=qgrad), compat.TV(dims, sigma=0.3, niter=10, rtol=1e-4),
                                                  pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"], niter=N, seed=6)
        assert rel(got, want) < 1e-10
    # MAP loop with the non-convex MC-TV term: q enters after the Moreau pre-step
    mk = dict(dims=dims, b=y, sigma=sig, lamda=0.3, gamma=15.0, isotropic=True, niter=50, warm=True, q=q, alpha=alpha)
    want, _ = D.AdaptivePrimalDual(D.L2_ncvx_tv(Op=pb["H"], Op2=pb["Gop"], **mk), pb["l1iso"](), pb["Gop"], pb["x0"],
                                   pb["tau0"], pb["mu0"], niter=N)
    got, _ = primaldual.AdaptivePrimalDual(algs.L2_ncvx_tv(Op=H, Op2=compat.Gradient(dims), **mk),
                                           compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"],
                                           pb["tau0"], pb["mu0"], niter=N)
    assert rel(got, want) < 1e-9
