"""CPU checks of the image-path oracle (oracle/deconv_ref.py).  The third-party arithmetic
(pylops / pyproximal) is PARITY UNPINNED (libraries absent); what can be pinned
is: scipy's own convolve / lsqr are the code executed, the noise idiom is the reference's
bit for bit, and every restated rule satisfies the identities that define it."""
import numpy as np
import pytest
import scipy.stats
from numpy.random import default_rng

from oracle import deconv_ref as D


def test_noise_idiom_is_standard_normal():
    # algs.py:433,565: scipy.stats.multivariate_normal.rvs(size=x.shape, random_state=rng)
    a = default_rng(7)
    b = default_rng(7)
    for _ in range(3):
        xi = scipy.stats.multivariate_normal.rvs(size=(50,), random_state=a)
        assert np.array_equal(xi, b.standard_normal(50))
    assert np.array_equal(D.image_noise(7, 3, 50)[2], xi)


@pytest.mark.parametrize("k,fwd,adj", [(5, (-2, 2), (-2, 2)), (6, (-2, 3), (-3, 2)), (7, (-3, 3), (-3, 3))])
def test_blur_windows_and_adjoint(k, fwd, adj):
    """window of the forward / adjoint box blur per kernel size (even kernel is not self-adjoint)."""
    dims = (23, 31)
    H = D.Convolve2D(dims, h=np.ones((k, k)) / k ** 2, offset=(k // 2, k // 2))
    e = np.zeros(dims)
    e[11, 15] = 1.0
    # forward: out[i] = sum_{o in window} x[i+o]  ->  impulse at c appears at rows c-o
    rows = np.nonzero(np.abs(H.matvec(e.ravel()).reshape(dims)[:, 15]) > 1e-9)[0]
    assert (11 - rows.max(), 11 - rows.min()) == fwd
    rows = np.nonzero(np.abs(H.rmatvec(e.ravel()).reshape(dims)[:, 15]) > 1e-9)[0]
    assert (11 - rows.max(), 11 - rows.min()) == adj
    rng = default_rng(0)
    x, y = rng.normal(size=dims[0] * dims[1]), rng.normal(size=dims[0] * dims[1])
    assert abs(np.dot(H.matvec(x), y) - np.dot(x, H.rmatvec(y))) < 1e-10
    # direct (non-FFT) evaluation agrees with the FFT method the reference uses
    Hd = D.Convolve2D(dims, h=np.ones((k, k)) / k ** 2, offset=(k // 2, k // 2), method="direct")
    assert np.abs(Hd.matvec(x) - H.matvec(x)).max() < 1e-13


def test_gradient_adjoint_and_boundary():
    dims = (9, 13)
    G = D.Gradient(dims)
    rng = default_rng(1)
    x, v = rng.normal(size=dims[0] * dims[1]), rng.normal(size=2 * dims[0] * dims[1])
    assert abs(np.dot(G.matvec(x), v) - np.dot(x, G.rmatvec(v))) < 1e-12
    g = G.matvec(x).reshape(2, *dims)
    assert np.all(g[0, -1, :] == 0) and np.all(g[1, :, -1] == 0)
    dx, dy = v.reshape(2, *dims)
    assert np.allclose(D.div2d(dx, dy), -D.grad2d_adj(dx, dy))


def test_moreau_identity_is_projection():
    rng = default_rng(2)
    v = rng.normal(size=2 * 40) * 2
    for op, proj in ((D.L21(2, 0.3), lambda w: (w.reshape(2, -1) / np.maximum(1, np.linalg.norm(w.reshape(2, -1), axis=0) / 0.3)).ravel()),
                     (D.L1(0.3), lambda w: np.clip(w, -0.3, 0.3))):
        for mu in (1.0, 0.4):
            assert np.abs(op.proxdual(v, mu) - proj(v)).max() < 1e-14


def test_tv_prox_converges_to_fixed_point_and_counts():
    dims = (24, 20)
    x = D.synthetic_image(*dims).astype(float)
    gam = 2.0
    primal = lambda s: 0.5 * np.sum((s - x) ** 2) + gam * np.sum(np.sqrt(sum(g ** 2 for g in D.grad2d(s))))
    objs = [primal(D.tv_prox_2d(x, gam, n, 0.0)[0]) for n in (5, 50, 500, 3000)]
    assert objs[0] > objs[1] > objs[2] >= objs[3] - 1e-9
    assert (objs[2] - objs[3]) / objs[3] < 1e-4
    tv = D.TV(dims, sigma=0.3, niter=10)
    out = tv.prox(x.ravel(), 0.5625)
    assert out.shape == (dims[0] * dims[1],) and 0 <= tv.updates[-1] <= 11
    assert np.array_equal(D.tv_prox_2d(np.zeros(dims), 1.0, 10, 1e-4)[0], np.zeros(dims))


def test_l2_prox_solves_normal_equations_and_warm_starts():
    dims = (32, 28)
    pb = D.build_problem(D.synthetic_image(*dims))
    l2 = pb["l2"]()
    x = default_rng(3).normal(size=dims[0] * dims[1]) * 30
    tau = pb["tau0"]
    z = l2.prox(x.copy(), tau)
    H = pb["H"]
    lhs = z + tau * l2.sigma * H.rmatvec(H.matvec(z))
    rhs = x + tau * l2.sigma * H.rmatvec(pb["y"].ravel())
    assert np.linalg.norm(lhs - rhs) / np.linalg.norm(rhs) < 1e-5      # lsqr atol=btol=1e-6
    k1 = l2.lsqr_itn[-1]
    l2.prox(x.copy(), tau)
    assert l2.lsqr_itn[-1] < k1                                        # warm start from the previous solution
    # grad is sigma H^T (H x - b)
    assert np.allclose(l2.grad(x), l2.sigma * H.rmatvec(H.matvec(x) - pb["y"].ravel()))


def test_samplers_are_deterministic_and_seeded():
    dims = (16, 20)
    pb = D.build_problem(D.synthetic_image(*dims))
    a = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                         gamma=pb["gamma_myula"], niter=4, seed=0)
    b = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                         gamma=pb["gamma_myula"], niter=4, noise=D.image_noise(0, 4, 320))
    assert np.array_equal(a, b) and a.shape == (4, 320)
    # first MYULA step from x0=0 in closed form: tau*sigma*H^T b + sqrt(2 tau) xi  (prox_TV(0)=0)
    l2 = pb["l2"]()
    want = -pb["tau_myula"] * l2.grad(np.zeros(320)) + np.sqrt(2 * pb["tau_myula"]) * D.image_noise(0, 1, 320)[0]
    assert np.allclose(a[0], want, rtol=0, atol=1e-12)
    u, uy = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                           niter=3, gfirst=False, returny=True)
    assert u.shape == (3, 320) and uy.shape == (3, 640)
    assert np.all(np.linalg.norm(uy.reshape(3, 2, -1), axis=1) <= 0.3 + 1e-12)   # dual iterates are projected
