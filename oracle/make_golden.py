"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/mix_*.npz.

Runs the *unmodified* reference sampler classes (imported from /root/reference
through ``oracle/refshim.py``) with the README parameters and stores their
outputs.  The reference tree cannot travel to the GPU box, these vectors can.

    python oracle/make_golden.py            # rewrites tests/golden/mix_*.npz

Configs follow the reference README (README.md:9-24) and the drivers' constants
(lmc.py:278,284; lmc_laplace.py:315,318; prox_lmc.py:374-381).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import refshim  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

M_PRE = np.array([[1.0, 0.1], [0.1, 0.5]])
Q_PRE = np.array([[1.0, 0.1], [0.1, 1.5]])
BETA = np.array([0.7, 0.3])
SIGMA_BM = np.array([0.8, 0.2])


def main(K=1000, K_adj=400, K_long=10000, seed=0):
    mods = refshim.load_reference()
    os.makedirs(OUT, exist_ok=True)

    # ---- config 1: lmc.py, gamma=0.1, n=5 -------------------------------------------
    mus, Sg, om = refshim.gaussian_mixture_params(5)
    with refshim.quiet():
        s = mods["lmc"].LangevinMonteCarlo(mus, Sg, om, K, seed)
        out = {"ula": s.ula(0.1), "pula": s.pula(0.1, M_PRE), "ihpula": s.ihpula(0.1),
               "mla": s.mla(0.1, BETA)}
        s.n = K_adj
        out["mala"], out["mala_nacc"] = s.mala(0.1)
        s.n = K_long
        out["ula_long"] = s.ula(0.1)
        # n=1 exercises the "no eigen-shift" branch of ihpula (lmc.py:164-165)
        m1, S1, o1 = refshim.gaussian_mixture_params(1)
        s1 = mods["lmc"].LangevinMonteCarlo(m1, S1, o1, 200, seed)
        out["ihpula_n1"] = s1.ihpula(0.1)
        out["ula_n1"] = s1.ula(0.1)
    np.savez_compressed(os.path.join(OUT, "mix_gauss.npz"), gamma=0.1, seed=seed,
                        mus=np.array(mus), Sigmas=np.array(Sg), omegas=om,
                        M=M_PRE, beta=BETA, **out)

    # ---- config 2: lmc_laplace.py, gamma=0.05, lamda=0.1, alpha=0.5 -----------------
    mus, al, om = refshim.laplacian_mixture_params(5, 0.5)
    with refshim.quiet():
        s = mods["lmc_laplace"].LangevinMonteCarloLaplacian(mus, al, om, 0.1, K, seed)
        out = {"ula": s.ula(0.05), "pula": s.pula(0.05, M_PRE), "ihpula": s.ihpula(0.05),
               "mla": s.mla(0.05, BETA)}
        s.n = K_adj
        out["mala"], out["mala_nacc"] = s.mala(0.05)
    np.savez_compressed(os.path.join(OUT, "mix_laplace.npz"), gamma=0.05, seed=seed,
                        lamda=0.1, mus=np.array(mus), alphas=al, omegas=om,
                        M=M_PRE, beta=BETA, **out)

    # ---- config 3: prox_lmc.py, gamma=0.05, alpha=0.15, lamda=0.25, t=100 -----------
    mus, Sg, om = refshim.gaussian_mixture_params(5)
    with refshim.quiet():
        s = mods["prox_lmc"].ProximalLangevinMonteCarlo(mus, Sg, om, 0.25, 0.15,
                                                        np.zeros(2), K, seed)
        out = {"pgld": s.pgld(0.05), "myula": s.myula(0.05), "fbula": s.fbula(0.05),
               "ppula": s.ppula(0.05, M_PRE, Q_PRE, 100),
               "lbmumla": s.lbmumla(0.05, BETA, SIGMA_BM)}
        s.n = K_adj
        out["mymala"], out["mymala_nacc"] = s.mymala(0.05)
    np.savez_compressed(os.path.join(OUT, "mix_prox.npz"), gamma=0.05, seed=seed,
                        lamda=0.25, alpha=0.15, mu=np.zeros(2), t=100,
                        mus=np.array(mus), Sigmas=np.array(Sg), omegas=om,
                        M=M_PRE, Q=Q_PRE, beta=BETA, sigma=SIGMA_BM, **out)
    for f in ("mix_gauss", "mix_laplace", "mix_prox"):
        p = os.path.join(OUT, f + ".npz")
        print(p, os.path.getsize(p), "bytes")


if __name__ == "__main__":
    main()
