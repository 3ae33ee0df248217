"""Pins oracle/mix_ref.py (numpy restatement) against the reference's own outputs.

* golden vectors in tests/golden/mix_*.npz were produced by the unmodified
  reference classes (oracle/make_golden.py) -- bit-exact match required;
* in the build container the reference is additionally executed live.
"""
import os

import numpy as np
import pytest

from oracle import mix_ref as O


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def _targets(golden_dir):
    g = _load(golden_dir, "mix_gauss.npz")
    l = _load(golden_dir, "mix_laplace.npz")
    p = _load(golden_dir, "mix_prox.npz")
    tg = O.GaussMix(list(g["mus"]), list(g["Sigmas"]), g["omegas"])
    tl = O.LaplaceMix(list(l["mus"]), l["alphas"], l["omegas"], float(l["lamda"]))
    tp = O.GaussMixL1(list(p["mus"]), list(p["Sigmas"]), p["omegas"], float(p["lamda"]),
                      float(p["alpha"]), p["mu"])
    return (g, tg), (l, tl), (p, tp)


def test_noise_idioms_are_standard_normal():
    """both Gaussian draw idioms of the reference reduce to standard_normal."""
    a = np.random.default_rng(3)
    b = np.random.default_rng(3)
    for _ in range(200):
        x = a.multivariate_normal(np.zeros(2), np.identity(2))
        y = b.standard_normal(2)
        assert np.array_equal(x, y)
    a = np.random.default_rng(5)
    b = np.random.default_rng(5)
    import scipy.stats
    x = scipy.stats.multivariate_normal.rvs(size=(37,), random_state=a)
    assert np.array_equal(x, b.standard_normal(37))
    a = np.random.default_rng(7)
    b = np.random.default_rng(7)
    assert np.array_equal(a.normal(0, 1, 2), b.standard_normal(2))


@pytest.mark.parametrize("alg", ["ula", "pula", "ihpula", "mla", "ula_long"])
def test_gauss_unadjusted_bit_exact(golden_dir, alg):
    (g, tg), _, _ = _targets(golden_dir)
    ref = g[alg]
    th0, xi, _ = O.reference_noise(int(g["seed"]), len(ref))
    name = alg.split("_")[0]
    args = {"ula": (), "pula": (g["M"],), "ihpula": (5,), "mla": (g["beta"],)}[name]
    got = O.run_unadjusted(name, tg, th0, xi, float(g["gamma"]), *args)
    assert np.array_equal(got, ref)


def test_gauss_single_component(golden_dir):
    g = _load(golden_dir, "mix_gauss.npz")
    tg = O.GaussMix([g["mus"][0]], [g["Sigmas"][0]], np.ones(1))
    th0, xi, _ = O.reference_noise(0, 200)
    assert np.array_equal(O.run_unadjusted("ula", tg, th0, xi, 0.1), g["ula_n1"])
    assert np.array_equal(O.run_unadjusted("ihpula", tg, th0, xi, 0.1, 1), g["ihpula_n1"])


@pytest.mark.parametrize("alg", ["ula", "pula", "ihpula", "mla"])
def test_laplace_unadjusted_bit_exact(golden_dir, alg):
    _, (l, tl), _ = _targets(golden_dir)
    ref = l[alg]
    th0, xi, _ = O.reference_noise(int(l["seed"]), len(ref))
    args = {"ula": (), "pula": (l["M"],), "ihpula": (5,), "mla": (l["beta"],)}[alg]
    got = O.run_unadjusted(alg, tl, th0, xi, float(l["gamma"]), *args)
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("alg", ["pgld", "myula", "fbula", "ppula", "lbmumla"])
def test_prox_unadjusted_bit_exact(golden_dir, alg):
    _, _, (p, tp) = _targets(golden_dir)
    ref = p[alg]
    th0, xi, _ = O.reference_noise(int(p["seed"]), len(ref))
    args = {"pgld": (), "myula": (), "fbula": (), "ppula": (p["M"], p["Q"], int(p["t"])),
            "lbmumla": (p["beta"], p["sigma"])}[alg]
    got = O.run_unadjusted(alg, tp, th0, xi, float(p["gamma"]), *args)
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("which", ["gauss", "laplace", "prox"])
def test_adjusted_bit_exact(golden_dir, which):
    (g, tg), (l, tl), (p, tp) = _targets(golden_dir)
    z, tgt, alg = {"gauss": (g, tg, "mala"), "laplace": (l, tl, "mala"),
                   "prox": (p, tp, "mymala")}[which]
    ref, nacc = z[alg], int(z[alg + "_nacc"])
    K = 400
    th0, xi, u = O.reference_noise(int(z["seed"]), K, uniforms=True)
    states, acc, _ = O.run_adjusted(alg, tgt, th0, xi, u, float(z["gamma"]))
    assert acc.sum() == nacc
    assert np.array_equal(states[acc], ref)


@pytest.mark.refonly
def test_live_reference_matches_oracle():
    """Container only: run the real reference with another seed / K / n."""
    from oracle import refshim
    mods = refshim.load_reference()
    mus, Sg, om = refshim.gaussian_mixture_params(4)
    with refshim.quiet():
        s = mods["lmc"].LangevinMonteCarlo(mus, Sg, om, 150, 11)
        ref_ula, ref_ih = s.ula(0.07), s.ihpula(0.07)
        ref_mala, nacc = s.mala(0.07)
    tg = O.GaussMix(mus, Sg, om)
    th0, xi, _ = O.reference_noise(11, 150)
    assert np.array_equal(O.run_unadjusted("ula", tg, th0, xi, 0.07), ref_ula)
    assert np.array_equal(O.run_unadjusted("ihpula", tg, th0, xi, 0.07, 4), ref_ih)
    th0, xi, u = O.reference_noise(11, 150, uniforms=True)
    st, acc, _ = O.run_adjusted("mala", tg, th0, xi, u, 0.07)
    assert acc.sum() == nacc and np.array_equal(st[acc], ref_mala)
