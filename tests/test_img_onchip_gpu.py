"""GPU parity tests of the on-chip cluster MYULA kernel (csrc/img_myula_onchip.cu): against the
CPU oracle on the reference's noise stream (fp32 tolerance 1e-5 of the dynamic range, north_star)
and against the streaming kernel on the in-kernel Philox stream (same draws, same recipe)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(1.0, np.abs(b).max()))


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


def _problem(dims, kernel=5):
    from oracle import deconv_ref as D
    return D.build_problem(D.synthetic_image(*dims), kernel=kernel)


def test_onchip_support_matrix(torch):
    from lmc_atomi_b200 import _lib, imaging as I
    pb = _problem((64, 64))
    assert I.DeconvEngine((64, 64), pb["H"].h, pb["y"].ravel(), 1.0, dtype="float32").onchip_supported()
    assert not I.DeconvEngine((64, 64), pb["H"].h, pb["y"].ravel(), 1.0, dtype="float64").onchip_supported()
    big = I.DeconvEngine((300, 64), pb["H"].h, np.zeros(300 * 64), 1.0, dtype="float32")
    assert not big.onchip_supported()
    x = torch.zeros((1, 300 * 64), dtype=torch.float32, device="cuda")
    with pytest.raises(_lib.LmcError):
        big.myula(x, 1, 0.1, 0.5, 0.3, engine="onchip")
    assert _lib.load().lmc_img_myula_onchip_clusters() >= 1


# rows in one CTA / spread over several CTAs / ragged / all 8 CTAs, blur 5, 6 (asymmetric), 7
@pytest.mark.parametrize("dims,kernel,N", [((64, 64), 5, 25), ((40, 52), 5, 25), ((33, 95), 6, 20), ((64, 64), 7, 20),
                                           ((100, 70), 5, 12), ((131, 256), 6, 8), ((256, 256), 5, 6)])
def test_onchip_myula_trajectory_vs_oracle(torch, dims, kernel, N):
    from lmc_atomi_b200 import imaging as I
    from oracle import deconv_ref as D
    pb = _problem(dims, kernel)
    P = dims[0] * dims[1]
    noise = D.image_noise(0, N, P)
    tv = pb["tv"]()
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                            niter=N, noise=noise)
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    x = torch.zeros((1, P), dtype=torch.float32, device="cuda")
    samples = torch.empty((1, N, P), dtype=torch.float32, device="cuda")
    mean = torch.zeros((1, P), dtype=torch.float32, device="cuda")
    m2 = torch.zeros((1, P), dtype=torch.float32, device="cuda")
    eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, noise=noise[None], samples=samples,
              mean=mean, m2=m2, engine="onchip")
    got = samples.cpu().numpy()[0]
    errs = [rel(got[n], want[n]) for n in range(N)]
    print(f"on-chip MYULA fp32 {dims} k={kernel}: max rel err {max(errs):.3e}")
    assert max(errs) < 1e-5, errs
    assert rel(x.cpu().numpy()[0], want[-1]) < 1e-5
    assert rel(mean.cpu().numpy()[0], want.mean(axis=0)) < 1e-5
    assert rel(m2.cpu().numpy()[0] / N, want.var(axis=0)) < 1e-4
    # per-chain early exit reproduced: same number of executed TV sweeps as the oracle
    assert eng.read_counters()["tv_sweeps"] == sum(min(u + 1, 10 + D.TV_EXTRA_PASS) for u in tv.updates)


def test_onchip_matches_streaming_on_philox_noise(torch):
    """More chains than co-resident clusters, in-kernel noise, chunked launches with running statistics."""
    from lmc_atomi_b200 import imaging as I
    dims, Cn = (96, 80), 37
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    out = {}
    for engine in ("stream", "onchip"):
        x = torch.zeros((Cn, P), dtype=torch.float32, device="cuda")
        mean = torch.zeros_like(x)
        m2 = torch.zeros_like(x)
        done = 0
        for n in (5, 7):      # two launches: iter_offset / stat_count0 carry the Philox counter and the mean
            eng.myula(x, n, pb["tau_myula"], pb["gamma_myula"], 0.3, 10, 1e-4, seed=11, chain_offset=1000,
                      mean=mean, m2=m2, burnin=2, stat_count0=max(done - 2, 0), iter_offset=done, engine=engine)
            done += n
        torch.cuda.synchronize()
        out[engine] = (x.cpu().numpy(), mean.cpu().numpy(), m2.cpu().numpy())
    for a, b in zip(out["onchip"], out["stream"]):
        assert rel(a, b) < 1e-5
    assert np.abs(out["onchip"][0]).max() > 10      # the chains moved


def test_onchip_tv_niter_edge_cases(torch):
    from lmc_atomi_b200 import imaging as I
    dims, N = (48, 40), 4
    pb = _problem(dims)
    P = dims[0] * dims[1]
    eng = I.DeconvEngine(dims, pb["H"].h, pb["y"].ravel(), 1 / 0.75 ** 2, dtype="float32")
    for niter in (1, 2, 3):
        res = {}
        for engine in ("stream", "onchip"):
            x = torch.zeros((2, P), dtype=torch.float32, device="cuda")
            eng.myula(x, N, pb["tau_myula"], pb["gamma_myula"], 0.3, niter, 1e-4, seed=3, engine=engine)
            res[engine] = x.cpu().numpy()
        assert rel(res["onchip"], res["stream"]) < 1e-5, niter
