"""TEST INFRASTRUCTURE ONLY (never imported by the product).

Exact 2-Wasserstein evaluation of the reference's mixture experiments (lmc.py:386-416,
lmc_laplace.py:426-451): ``ot.emd2(a, b, ot.dist(Z1, Z[:k+1]))**.5`` between the K true samples
(uniform weights a) and a chain prefix (uniform weights b).  POT (``ot``) is absent from this image
and from the reference tree (README.md:5 `pip install -U ... pot`, no pinned version), so the linear
programme POT's network simplex solves is restated here and handed to scipy's HiGHS:

    min <M, G>  s.t.  G 1 = a,  G^T 1 = b,  G >= 0,      M_ij = ||x_i - y_j||_2^2   (ot.dist default)

Parity: unpinned against POT itself (absent); pinned against closed forms in tests/test_oracle_w2.py
(translations, 1-D sorted matching, permutations).  Sizes of a few hundred points solve in well under a second.
"""
from __future__ import annotations

import numpy as np
from scipy import sparse
from scipy.optimize import linprog


def dist_sqeuclidean(x, y):
    """ot.dist(x, y) with its default metric 'sqeuclidean'."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    return ((x[:, None, :] - y[None, :, :]) ** 2).sum(axis=2)


def emd2(a, b, M):
    """Optimal transport cost  min_G <M, G>  with marginals a, b (ot.emd2)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    n, m = M.shape
    rows = sparse.kron(sparse.eye(n), np.ones((1, m)), format="csr")     # sum_j G_ij = a_i
    cols = sparse.kron(np.ones((1, n)), sparse.eye(m), format="csr")     # sum_i G_ij = b_j
    A = sparse.vstack([rows, cols[:-1]], format="csr")                   # one marginal is redundant
    rhs = np.concatenate([a, b[:-1]])
    res = linprog(np.asarray(M, dtype=np.float64).ravel(), A_eq=A, b_eq=rhs, bounds=(0, None), method="highs")
    if res.status != 0:
        raise RuntimeError("emd2: " + res.message)
    return float(res.fun)


def wasserstein2(true_samples, chain):
    """The quantity the reference appends to its wass_*_list: emd2(...)**.5 with uniform weights."""
    x = np.asarray(true_samples, dtype=np.float64)
    y = np.asarray(chain, dtype=np.float64)
    a = np.ones(len(x)) / len(x)
    b = np.ones(len(y)) / len(y)
    return emd2(a, b, dist_sqeuclidean(x, y)) ** .5


def wasserstein2_curve(true_samples, chain, interval=100):
    """lmc.py:396-413: W2 between the true samples and the chain prefixes Z[:k+1], k = 1, 1+interval, ..."""
    K = len(true_samples)
    return [wasserstein2(true_samples, chain[:k + 1]) for k in range(1, min(K, len(chain))) if (k - 1) % interval == 0]
