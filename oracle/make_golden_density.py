"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/mix_density.npz.

Evaluates the density helpers and ``true_samples`` of the *unmodified* reference classes (imported from
/root/reference through ``oracle/refshim.py``) on a small grid, for the drop-in classes' evaluation helpers
(lmc.py:39-48,82-90; lmc_laplace.py:43-65; prox_lmc.py:42-67).

    python oracle/make_golden_density.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import refshim  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden", "mix_density.npz")


def main():
    mods = refshim.load_reference()
    g = np.linspace(-5, 5, 23)
    X, Y = np.meshgrid(g, g)
    pos = np.stack([X, Y], axis=-1)
    out = {"pos": pos}
    with refshim.quiet():
        for n in (1, 4, 5):
            mus, Sg, om = refshim.gaussian_mixture_params(n)
            s = mods["lmc"].LangevinMonteCarlo(mus, Sg, om, 48, 3)
            out[f"gauss_n{n}"] = s.density_gaussian_mixture(pos)
            if n == 5:
                out["gauss_true_samples"] = s.true_samples()
                out["gauss_point"] = np.array(s.density_gaussian_mixture(np.array([0.3, -0.7])))
        mus, al, om = refshim.laplacian_mixture_params(5, 0.5)
        s = mods["lmc_laplace"].LangevinMonteCarloLaplacian(mus, al, om, 0.1, 10, 0)
        out["laplace"] = s.density_laplacian_mixture(pos)
        out["laplace_smooth"] = s.smooth_density_laplacian_mixture(pos)
        s = mods["lmc_laplace"].LangevinMonteCarloLaplacian(mus, al, om, 0.1, 40, 2)
        np.random.seed(1234)            # multivariate_laplace.rvs draws from numpy's global state
        out["laplace_true_samples"] = s.true_samples()
        mus, Sg, om = refshim.gaussian_mixture_params(5)
        s = mods["prox_lmc"].ProximalLangevinMonteCarlo(mus, Sg, om, 0.25, 0.15, np.zeros(2), 10, 0)
        out["prox"] = s.density_gaussian_mixture(pos) * s.multivariate_laplacian(pos)
        out["prox_smooth"] = s.density_gaussian_mixture(pos) * s.smooth_multivariate_laplacian(pos)
    np.savez_compressed(OUT, **out)
    print(OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
