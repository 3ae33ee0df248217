"""The exact W2 evaluator of the oracle (oracle/w2_ref.py) against closed forms."""
import numpy as np

from oracle import w2_ref as W


def test_translation_and_permutation():
    rng = np.random.default_rng(0)
    x = rng.normal(size=(60, 2))
    t = np.array([0.3, -1.1])
    assert abs(W.wasserstein2(x, x[rng.permutation(60)])) < 1e-7
    assert abs(W.wasserstein2(x, x + t) - np.linalg.norm(t)) < 1e-7


def test_one_dimensional_sorted_matching():
    rng = np.random.default_rng(1)
    x, y = rng.normal(size=40), rng.normal(1.0, 2.0, size=40)
    want = np.sqrt(np.mean((np.sort(x) - np.sort(y)) ** 2))
    got = W.wasserstein2(np.c_[x, np.zeros(40)], np.c_[y, np.zeros(40)])
    assert abs(got - want) < 1e-7


def test_unequal_sizes_two_points():
    # mass 1/2 at 0 and 1 against a single point at 0.25: cost = (0.25^2 + 0.75^2) / 2
    x = np.array([[0.0, 0.0], [1.0, 0.0]])
    y = np.array([[0.25, 0.0]])
    assert abs(W.wasserstein2(x, y) ** 2 - (0.0625 + 0.5625) / 2) < 1e-9
    curve = W.wasserstein2_curve(np.r_[x, x], np.r_[y, y, y, y], interval=1)
    assert len(curve) == 3 and all(abs(c ** 2 - 0.3125) < 1e-9 for c in curve)
