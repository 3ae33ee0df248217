"""CUDA path against the REFERENCE's own results: the full run of expt_deconv.sh:5-6
(`prox_lmc_deconv.py --image=einstein --sigma=0.75 --tau=0.3 --gamma_mc=15 --gamma_me=15 --N=1000`, seed 0) for both
samplers and all nine models, through the drop-in `prox_lmc_deconv` in fp64 with the reference's noise stream, compared
with the per-sample SNR / PSNR / MSE curves digitised from the reference's committed figures
(tests/golden/deconv_curves.npz, oracle/digitise_figs.py).  See tests/test_oracle_pinned.py for the resolution of
that evidence (2.2e-7 dB per vertex)."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
N = 1000
# Measured on B200 (profiles/r02b_pinned_*_niter_plus_1_passes.json): every vertex of all 54 curves within 1.3e-7 dB,
# i.e. the resolution of the plot, over the whole N = 1000 run.  The same run with `niter` instead of `niter + 1` TV
# passes (LMC_TV_EXTRA_PASS=0, profiles/r02b_pinned_*_niter_passes.json) misses the ME-TV models by 1e-5..3e-5 dB
# and the MYULA L2/MC-TV models by up to 4.6e-7 dB -- the gate below separates the two.
DB_TOL = 3e-7
MODEL_OF = {("l2", 5): 1, ("mc", 5): 2, ("me", 5): 3, ("l2", 6): 4, ("mc", 6): 5, ("me", 6): 6,
            ("l2", 7): 7, ("mc", 7): 8, ("me", 7): 9}      # legend order of prox_lmc_deconv.py:813-821


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


@pytest.fixture(scope="module")
def einstein():
    from PIL import Image
    return np.array(Image.open(os.path.join(GOLDEN, "einstein.png")))


@pytest.fixture(scope="module")
def noise(torch, einstein):
    """default_rng(0).standard_normal(P) per iteration (algs.py:431-433, 561-565), drawn once, kept on the device."""
    P = einstein.size
    rng = np.random.default_rng(0)
    z = torch.empty((N, P), dtype=torch.float64, device="cuda")
    for i in range(0, N, 50):
        z[i:i + 50] = torch.from_numpy(rng.standard_normal((50, P))).cuda()
    return z


def residuals(curves, fig, res, max_idx=None):
    out = {}
    for (term, k), m in MODEL_OF.items():
        r = res["models"][f"{term}_{k}"]
        for name in ("snr", "psnr", "mse"):
            idx, val = curves[f"{fig}_{name}_M{m}_idx"], curves[f"{fig}_{name}_M{m}_val"]
            if max_idx is not None:
                idx, val = idx[idx <= max_idx], val[idx <= max_idx]
            got = np.asarray(r[name])[idx]
            d = np.abs(got - val)
            if name == "mse":      # to dB, so one tolerance serves (the MSE axis resolves 1.1e-4 absolute)
                d = 10 / np.log(10) * np.maximum(d - 2e-4, 0) / val
            out[f"M{m}_{name}"] = float(d.max())
    return out


@pytest.mark.parametrize("alg", ["MYULA", "ULPDA"])
def test_einstein_run_matches_reference_figures(torch, einstein, noise, alg):
    from lmc_atomi_b200.prox_lmc_deconv import prox_lmc_deconv
    curves = np.load(os.path.join(GOLDEN, "deconv_curves.npz"))
    res = prox_lmc_deconv(gamma_mc=15., gamma_me=15., sigma=0.75, tau=0.3, N=N, image=einstein, alg=alg, seed=0,
                          dtype="float64", noise=noise)
    worst = residuals(curves, f"einstein_{alg}", res)
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/pinned_{alg}.json", "w") as f:
        json.dump({"worst_abs_diff_db": worst,
                   "mean_psnr": {k: v["mean_psnr"] for k, v in res["models"].items()}}, f, indent=1)
    bad = {k: v for k, v in worst.items() if not v < DB_TOL}
    assert not bad, bad


def test_einstein_map_matches_reference_figure(torch, einstein):
    """MAP branch (expt_deconv.sh:4: --compute_MAP=True --niter_MAP=1000): adaptive PDHG on the CUDA operators against
    fig_prox_lmc_deconv_einstein_MAP_1000_snr_psnr_mse for all nine models."""
    from lmc_atomi_b200.prox_lmc_deconv import prox_lmc_deconv
    curves = np.load(os.path.join(GOLDEN, "deconv_curves.npz"))
    res = prox_lmc_deconv(gamma_mc=15., gamma_me=15., sigma=0.75, tau=0.3, niter_MAP=N, image=einstein,
                          compute_MAP=True, seed=0, dtype="float64")
    # The driver's step sizes violate the PDHG condition tau mu ||G||^2 < 1 (0.534 x 1 x 8 = 4.3): the
    # iteration settles on a bounded oscillation (residuals p, d ~ 40, never 0) that amplifies rounding differences
    # about 10x per 25 iterations (measured GPU vs CPU oracle, scripts/debug_map.py: 1e-16 -> 1e-9 by iteration 190,
    # identical LSQR iteration counts throughout).  Beyond ~150 iterations no two floating-point implementations
    # (including the reference's own FFT blur) agree to the plot's resolution, so the gate is tight on the early
    # iterates (which pin every rule of the iteration, the step-size adaptations included) and loose on the tail.
    # Measured (profiles/r02f_pinned_MAP.json): first 120 iterations within 1.1e-7 dB for the L2 / ME-TV models and
    # 6e-6 dB for the MC-TV models (anisotropic prior, faster amplification); whole run within 3e-3 dB.
    first50 = residuals(curves, "einstein_MAP", res, max_idx=50)
    first120 = residuals(curves, "einstein_MAP", res, max_idx=120)
    worst = residuals(curves, "einstein_MAP", res)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/pinned_MAP.json", "w") as f:
        json.dump({"worst_abs_diff_db_first_50": first50, "worst_abs_diff_db_first_120": first120,
                   "worst_abs_diff_db": worst,
                   "map_psnr": {k: v["map_psnr"] for k, v in res["models"].items()}}, f, indent=1)
    bad = {k: v for k, v in first50.items() if not v < DB_TOL}
    assert not bad, bad
    bad = {k: v for k, v in first120.items() if not v < 2e-5}
    assert not bad, bad
    bad = {k: v for k, v in worst.items() if not v < 0.02}
    assert not bad, bad
