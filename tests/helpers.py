"""Shared helpers of the GPU parity tests (no reference access: golden files only)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def make_samplers(K=1000, seed=0, dtype="float64"):
    """(golden, product sampler, oracle target) for the three mixture experiments."""
    from lmc_atomi_b200.lmc import LangevinMonteCarlo
    from lmc_atomi_b200.lmc_laplace import LangevinMonteCarloLaplacian
    from lmc_atomi_b200.prox_lmc import ProximalLangevinMonteCarlo
    from oracle import mix_ref as O
    g = load_golden("mix_gauss.npz")
    l = load_golden("mix_laplace.npz")
    p = load_golden("mix_prox.npz")
    out = {}
    out["gauss"] = (g, LangevinMonteCarlo(list(g["mus"]), list(g["Sigmas"]), g["omegas"], K, seed, dtype=dtype),
                    O.GaussMix(list(g["mus"]), list(g["Sigmas"]), g["omegas"]))
    out["laplace"] = (l, LangevinMonteCarloLaplacian(list(l["mus"]), l["alphas"], l["omegas"], float(l["lamda"]),
                                                     K, seed, dtype=dtype),
                      O.LaplaceMix(list(l["mus"]), l["alphas"], l["omegas"], float(l["lamda"])))
    out["prox"] = (p, ProximalLangevinMonteCarlo(list(p["mus"]), list(p["Sigmas"]), p["omegas"], float(p["lamda"]),
                                                 float(p["alpha"]), p["mu"], K, seed, dtype=dtype),
                   O.GaussMixL1(list(p["mus"]), list(p["Sigmas"]), p["omegas"], float(p["lamda"]),
                                float(p["alpha"]), p["mu"]))
    return out


def alg_params(z, alg):
    """keyword arguments of the product API / positional args of the oracle for `alg`."""
    if alg in ("pula",):
        return {"M": z["M"]}, (z["M"],)
    if alg == "ihpula":
        return {}, (len(z["mus"]),)
    if alg == "mla":
        return {"beta": z["beta"]}, (z["beta"],)
    if alg == "ppula":
        return {"M": z["M"], "Q": z["Q"], "t": int(z["t"])}, (z["M"], z["Q"], int(z["t"]))
    if alg == "lbmumla":
        return {"beta": z["beta"], "sigma": z["sigma"]}, (z["beta"], z["sigma"])
    return {}, ()


def rel_err(a, b):
    """max_k ||a_k - b_k||_inf / max(1, ||b_k||_inf)"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    num = np.abs(a - b).max(axis=-1)
    den = np.maximum(1.0, np.abs(b).max(axis=-1))
    return float((num / den).max())


CASES = [("gauss", a) for a in ("ula", "pula", "ihpula", "mla")] + \
        [("laplace", a) for a in ("ula", "pula", "ihpula", "mla")] + \
        [("prox", a) for a in ("pgld", "myula", "ppula", "fbula", "lbmumla")]
ADJ_CASES = [("gauss", "mala"), ("laplace", "mala"), ("prox", "mymala")]
