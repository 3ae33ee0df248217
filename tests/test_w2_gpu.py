"""GPU 2-Wasserstein evaluation (csrc/w2_sinkhorn.cu, lmc_atomi_b200/w2.py) against the exact transport linear
programme the reference solves with POT (oracle/w2_ref.py; lmc.py:386-416) and against closed forms.
Tolerance: |W2_gpu - W2_exact| < 1e-3 at the default eps = 1e-3 (the north star's Wasserstein tolerance)."""
import time

import numpy as np
import pytest

from helpers import alg_params, make_samplers

pytestmark = pytest.mark.gpu

TOL = 1e-3


@pytest.mark.parametrize("n,m", [(300, 300), (300, 102), (150, 2), (64, 257)])
def test_sinkhorn_w2_matches_exact_emd(n, m):
    from oracle import w2_ref as W
    from lmc_atomi_b200 import w2
    rng = np.random.default_rng(n * 1000 + m)
    x = rng.normal(size=(n, 2)) * 1.5
    y = rng.normal(size=(m, 2)) * 1.3 + 0.3
    ev = w2.SinkhornW2(x)
    got = ev.distance(y)
    assert ev.last["converged"] and abs(ev.last["mass"] - 1.0) < 1e-3
    assert abs(got - W.wasserstein2(x, y)) < TOL, (got, ev.last)


def test_translation_and_sorted_matching_closed_forms():
    from lmc_atomi_b200 import w2
    rng = np.random.default_rng(5)
    x = rng.normal(size=(2000, 2)) * 2.0
    t = np.array([0.6, -0.8])                                   # translation is optimal for the quadratic cost: W2 = |t| = 1
    assert abs(w2.wasserstein2(x, x + t) - 1.0) < TOL
    assert w2.wasserstein2(x, x) < 0.05                          # entropic blur only (sqrt of a cost of order eps)
    a, b = rng.normal(size=(700, 1)), rng.normal(size=(700, 1)) * 0.5 + 1.0
    exact = np.sqrt(np.mean((np.sort(a[:, 0]) - np.sort(b[:, 0])) ** 2))      # 1-D: monotone matching
    assert abs(w2.wasserstein2(a, b) - exact) < TOL


def test_curve_of_a_cuda_chain_matches_the_exact_curve():
    """The loop of lmc.py:396-413 end to end on the device: true samples vs prefixes of a CUDA ULA chain."""
    from oracle import w2_ref as W
    from lmc_atomi_b200 import w2
    K = 320
    z, s, _ = make_samplers(K=K)["gauss"]
    chain = s.ula(float(z["gamma"]))
    true = s.true_samples()
    want = W.wasserstein2_curve(true, chain, interval=100)
    ev = w2.SinkhornW2(true)
    got = ev.curve(chain, interval=100)
    assert len(got) == len(want) == 4
    assert np.abs(np.array(got) - np.array(want)).max() < TOL, (got, want)
    warm = w2.SinkhornW2(true).curve(chain, interval=100, warm_start=True)
    assert np.abs(np.array(warm) - np.array(want)).max() < TOL           # a warm start does not change the answer


def test_smaller_eps_is_closer_and_bad_arguments_raise():
    from oracle import w2_ref as W
    from lmc_atomi_b200 import _lib, w2
    rng = np.random.default_rng(9)
    x, y = rng.normal(size=(200, 2)), rng.normal(size=(120, 2)) + 0.5
    exact = W.wasserstein2(x, y)
    errs = [abs(w2.wasserstein2(x, y, eps=e) - exact) for e in (1e-1, 1e-2, 1e-3)]
    assert errs[2] < TOL and errs[2] <= errs[0]
    with pytest.raises(ValueError):
        w2.SinkhornW2(np.zeros((4, 5)))
    with pytest.raises(ValueError):
        w2.SinkhornW2(x).distance(np.zeros((4, 3)))
    with pytest.raises(_lib.LmcError):
        w2.SinkhornW2(x, eps=-1.0).distance(y)


def test_k5000_evaluation_runs_in_seconds():
    """The reference's default driver size (K = 5000 true samples, lmc.py:194-196): one full-size distance."""
    import torch
    from lmc_atomi_b200 import w2
    rng = np.random.default_rng(1)
    x = rng.normal(size=(5000, 2)) * 1.5
    ev = w2.SinkhornW2(x)
    torch.cuda.synchronize()
    t0 = time.time()
    d = ev.distance(x[::-1] * 0.9 + 0.2)
    dt = time.time() - t0
    print("W2 5000x5000: %.5f in %.2f s, %d iterations" % (d, dt, ev.last["iters"]))
    assert ev.last["converged"] and np.isfinite(d) and dt < 120


def test_headless_main_returns_the_wasserstein_curves():
    from oracle import w2_ref as W
    from lmc_atomi_b200.lmc import lmc_gaussian_mixture
    out = lmc_gaussian_mixture(0.1, 0.1, 0.1, 0.1, 0.1, n=5, K=202, seed=0, wasserstein=True)
    w = out["wasserstein"]
    assert set(w) == {"ula", "mala", "pula", "ihpula", "mla"} and len(w["ula"]) == 3
    assert len(w["mala"]) == len([k for k in (1, 101, 201) if k < out["mala_effective"]])
    for a in ("ula", "mla"):
        want = W.wasserstein2_curve(out["true_samples"], out[a], interval=100)
        assert np.abs(np.array(w[a]) - np.array(want)).max() < TOL
    from lmc_atomi_b200.lmc_laplace import lmc_laplacian_mixture
    np.random.seed(7)                                   # the Laplace true samples use numpy's global state
    out = lmc_laplacian_mixture(0.05, 0.05, 0.05, 0.05, lamda=0.1, alpha=0.5, n=5, K=202, seed=0, wasserstein=True)
    w = out["wasserstein"]
    assert set(w) == {"ula", "mala", "pula", "mla"} and len(w["pula"]) == 3
    want = W.wasserstein2_curve(out["true_samples"], out["pula"], interval=100)
    assert np.abs(np.array(w["pula"]) - np.array(want)).max() < TOL


def test_sinkhorn_w2_at_driver_scale_matches_exact_assignment():
    """K = 5000, the drivers' default number of true samples (lmc.py:194-196), where the transport LP is out of the
    oracle's reach: for two uniformly weighted sets of EQUAL size the optimal plan is a permutation (Birkhoff), so
    scipy's linear_sum_assignment gives the exact EMD that POT's network simplex returns.  Reference-shaped sets:
    true samples of the 5-component Gaussian mixture vs a CUDA ULA chain of the same length."""
    from scipy.optimize import linear_sum_assignment
    from lmc_atomi_b200 import w2
    K = 5000
    z, s, _ = make_samplers(K=K)["gauss"]
    chain = s.ula(float(z["gamma"]))
    true = s.true_samples()
    assert true.shape == chain.shape == (K, 2)
    C = ((true[:, None, :] - chain[None, :, :]) ** 2).sum(-1)
    r, c = linear_sum_assignment(C)
    exact = float(np.sqrt(C[r, c].mean()))
    ev = w2.SinkhornW2(true)
    got = ev.distance(chain)
    print(f"W2 at K={K}: sinkhorn {got:.6f} exact {exact:.6f} diff {got - exact:+.2e} iters {ev.last['iters']}")
    assert ev.last["converged"]
    assert abs(got - exact) < TOL
