"""Image-chain half of the benchmark (BASELINE.json configs[3], configs[4]):

* MYULA deconvolution, 256x256, model M1 (L2(H5) + TV(sigma=0.3, niter=10, rtol=1e-4)),
  tau=0.1125, gamma=0.5625 (prox_lmc_deconv.py:88-94, 465-473), a batch of independent chains
  whose state is far larger than L2;
* ULPDA deconvolution, 512x512, model M1 (L2(H5, niter=50, warm) + L21(0.3) + Gradient),
  tau=0.534375, mu=1, theta=1, gfirst=False (prox_lmc_deconv.py:455-464), 512 chains per GPU (BASELINE.json
  configs[4]: 4096 chains over 8 GPUs).

Synthetic data of the named shapes: a deterministic piecewise-smooth image in [0,255]
(skimage's camera image is unavailable in this image), blurred by the CUDA H5
operator plus N(0, 0.75^2) noise.  Nothing here imports oracle/.
"""
from __future__ import annotations

import numpy as np

from . import imaging, roofline

SIGMA, TAU_REG = 0.75, 0.3


def synthetic_image(ny, nx, seed=1234):
    """Blocks, a disc, a ramp and a little texture, uint8-valued."""
    yy, xx = np.mgrid[0:ny, 0:nx]
    img = 40.0 + 60.0 * (xx / max(nx - 1, 1))
    img[(yy > ny // 8) & (yy < ny // 2) & (xx > nx // 6) & (xx < nx // 2)] = 200.0
    img[(yy - 0.65 * ny) ** 2 + (xx - 0.7 * nx) ** 2 < (0.18 * min(ny, nx)) ** 2] = 20.0
    m = (yy > 0.6 * ny) & (xx < 0.4 * nx)
    img[m] += 50.0 * np.sin(xx[m] * 0.35)
    img += np.random.default_rng(seed).normal(0, 2.0, size=img.shape)
    return np.clip(np.rint(img), 0, 255)


def make_engine(dims, dev, dtype="float32", kernel=5, seed=0):
    """Observation y = H5 img + N(0, sigma^2) (prox_lmc_deconv.py:53-59) and the engine for blur size k."""
    import torch
    img = synthetic_image(*dims)
    h5 = imaging.pad_filter(np.ones((5, 5)) / 25.0, (2, 2))
    x = torch.as_tensor(img.reshape(1, -1)).to(dev, torch.float64)
    y = imaging.blur(imaging.BlurSpec(dims, h5), x)[0].cpu().numpy()
    y = y + np.random.default_rng(seed).normal(0, SIGMA, size=y.shape)
    hk = imaging.pad_filter(np.ones((kernel, kernel)) / kernel ** 2, (kernel // 2, kernel // 2))
    return img, imaging.DeconvEngine(dims, hk, y, 1.0 / SIGMA ** 2, dtype=dtype, device=dev)


def _world():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist, dist.get_world_size(), dist.get_rank()
    return None, 1, 0


def _time_launches(torch, dev, fn, steps, warmup):
    """Seconds per launch, CUDA events on the launching stream, barrier + synchronize on both sides, MAX over
    the ranks when a process group is up (each rank advances its own shard of the chains)."""
    dist, world, _ = _world()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(warmup):
        fn()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    barrier()
    t = torch.tensor([a.elapsed_time(b) * 1e-3 / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _e2e(torch, dev, x, mean, launch, units, reps=2):
    """The same metric through host buffers: per repetition the start states come from pinned host memory and
    the posterior mean goes back to it, both inside the timed region (MAX over ranks)."""
    dist, world, _ = _world()
    h_x = torch.zeros(x.shape, dtype=x.dtype).pin_memory()
    h_mean = torch.empty(mean.shape, dtype=mean.dtype).pin_memory()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        x.copy_(h_x, non_blocking=True)
        launch()
        h_mean.copy_(mean, non_blocking=True)
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) * 1e-3 / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    nbytes = x.numel() * x.element_size()
    return {"value": world * units / float(t.item()), "unit": "chain-iters/s", "h2d_bytes_per_step": nbytes,
            "d2h_bytes_per_step": nbytes}


def _pooled_mean(torch, mean):
    """The only collective of the image path: posterior mean pooled over all chains of all ranks."""
    dist, world, _ = _world()
    s = mean.double().sum(dim=0)
    if world > 1:
        dist.all_reduce(s)
    return s / (mean.shape[0] * world)


# measured once per round with `ncu --set full` (profiles/): dram__bytes_read+write per launch of the
# on-chip kernel divided by the chain-iterations of that launch (256 chains x 10 iterations, 256x256)
ONCHIP_DRAM_BYTES_PER_CHAIN_ITER = (134.6e6 + 86.6e6) / 2560.0


def run_myula(dev, steps=3, warmup=3, chains=1024, dims=(256, 256), iters=20, dtype="float32", engine="auto"):
    import torch
    _, eng = make_engine(dims, dev, dtype)
    L = 1.0 / SIGMA ** 2
    gamma, tau = 1.0 / L, 0.2 / L
    P = dims[0] * dims[1]
    x = torch.zeros((chains, P), dtype=eng.tdt, device=dev)
    mean = torch.zeros_like(x)
    state = {"it": 0, "keep": None}
    _, world, rank = _world()

    def launch():
        state["keep"] = eng.myula(x, iters, tau, gamma, TAU_REG, 10, 1e-4, seed=0, mean=mean,
                                  iter_offset=state["it"], engine=engine, chain_offset=rank * chains)
        state["it"] += iters

    for _ in range(2):      # burn-in so that the TV early exit is in its steady regime
        launch()
    torch.cuda.synchronize(dev)
    eng.counters.zero_()
    it0 = state["it"]
    secs = _time_launches(torch, dev, launch, steps, warmup)
    c = eng.read_counters()
    pooled = _pooled_mean(torch, mean)
    e2e = _e2e(torch, dev, x, mean, launch, chains * iters)
    n_it = state["it"] - it0 - 2 * iters
    j = c["tv_sweeps"] / float(n_it * chains)
    item = 4 if dtype == "float32" else 8
    bpp = roofline.myula_bytes_per_pixel(j, item)
    peaks = roofline.measured_peaks()
    ach = bpp * P * chains * iters / secs / 1e9
    onchip = engine == "onchip" or (engine == "auto" and eng.onchip_supported())
    band = not onchip and (engine == "band" or (engine == "auto" and eng.myula_band_supported()))
    if onchip:
        kernel = "img_myula_onchip_kernel<32,5,%d>" % (1 if dims == (256, 256) else 0)
        policy = ("chain state resident in the shared memory of an 8-CTA cluster for all iterations of a launch; "
                  "%d chains per launch, state of the batch (x + mean) %.0f MB >> 126 MB L2 only for the start/end "
                  "load/store" % (chains, 2 * chains * P * item / 1e6))
        traffic = ONCHIP_DRAM_BYTES_PER_CHAIN_ITER * chains * iters if dims == (256, 256) else None
        note = ("roofline kept as the HBM bound of the streaming formulation ((12+9j) x 4 B per pixel-iteration); "
                "the on-chip kernel removes that traffic (see `traffic`, measured DRAM bytes per launch), so frac can "
                "exceed 1 and the binding resource is instruction issue (smsp__issue_active in profiles/)")
    else:
        kernel = ("img_myula_band_kernel<%s,5>" if band else "img_myula_kernel<%s>") % ("float" if item == 4 else "double")
        policy, traffic, note = "per-chain arrays total far more than the 126 MB L2", None, None
        if band and dims == (512, 512) and item == 4:
            traffic = roofline.MYULA_BAND_DRAM_BYTES_PER_CHAIN_ITER * chains * iters
    out = {"metric": "myula_deconv_chain_iters_per_sec", "value": world * chains * iters / secs,
           "unit": "chain-iters/s", "n_gpus": world, "pooled_mean_finite": bool(torch.isfinite(pooled).all().item()),
           "pixel_iters_per_sec": world * chains * iters * P / secs, "dtype": dtype, "dims": list(dims),
           "chains": chains,
           "iters_per_launch": iters, "ms_per_launch": secs * 1e3, "tv_sweeps_per_iter": j,
           "e2e": e2e, "engine": "onchip" if onchip else ("band" if band else "stream"), "state_bytes": int(10 * chains * P * item), "l2_policy": policy,
           "roofline": {"bound": "hbm", "kernel": kernel,
                        "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                        "bytes_per_pixel_iter": bpp, "traffic": traffic, "peak_source": peaks["source"]}}
    if note:
        out["roofline"]["note"] = note
    return out


def run_ulpda(dev, steps=3, warmup=2, chains=512, dims=(512, 512), iters=2, dtype="float32", engine="auto", group=0):
    import torch
    _, eng = make_engine(dims, dev, dtype)
    L = 1.0 / SIGMA ** 2
    tau, mu = 0.95 / L, 1.0
    P = dims[0] * dims[1]
    z = lambda *s: torch.zeros(s, dtype=eng.tdt, device=dev)
    x, xhat, y, warm, mean = z(chains, P), z(chains, P), z(chains, 2, P), z(chains, P), z(chains, P)
    state = {"it": 0, "keep": None}
    _, world, rank = _world()

    def launch():
        state["keep"] = eng.ulpda(x, xhat, y, warm, iters, tau, mu, 1.0, TAU_REG, True, False, 50, seed=0,
                                  mean=mean, iter_offset=state["it"], engine=engine,
                                  group=group, chain_offset=rank * chains)
        state["it"] += iters

    for _ in range(3):
        launch()
    torch.cuda.synchronize(dev)
    eng.counters.zero_()
    secs = _time_launches(torch, dev, launch, steps, warmup)
    c = eng.read_counters()
    pooled = _pooled_mean(torch, mean)
    e2e = _e2e(torch, dev, x, mean, launch, chains * iters)
    k = c["lsqr_iterations"] / max(1.0, float(c["lsqr_solves"]))
    item = 4 if dtype == "float32" else 8
    bpp = roofline.ulpda_bytes_per_pixel(k, item)
    peaks = roofline.measured_peaks()
    ach = bpp * P * chains * iters / secs / 1e9
    band = engine == "band" or (engine == "auto" and eng.band_supported())
    return {"metric": "ulpda_deconv_chain_iters_per_sec", "value": world * chains * iters / secs,
            "unit": "chain-iters/s", "n_gpus": world, "pooled_mean_finite": bool(torch.isfinite(pooled).all().item()),
            "pixel_iters_per_sec": world * chains * iters * P / secs, "dtype": dtype, "dims": list(dims),
            "chains": chains,
            "iters_per_launch": iters, "ms_per_launch": secs * 1e3, "lsqr_iters_per_solve": k, "e2e": e2e,
            "engine": "band" if (engine == "band" or (engine == "auto" and eng.band_supported())) else "tile",
            "group": group,
            "roofline": {"bound": "hbm", "kernel": ("img_ulpda_band_kernel<%s,5,16>" if (
                engine == "band" or (engine == "auto" and eng.band_supported())) else "img_ulpda_kernel<%s>") % (
                "float" if item == 4 else "double"),
                         "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                         "bytes_per_pixel_iter": bpp,
                         "traffic": (roofline.ULPDA_BAND_DRAM_BYTES_PER_CHAIN_ITER * chains * iters
                                     if (band and dims == (512, 512)) else None),
                         "peak_source": peaks["source"]}}


def run(dev, steps=3, warmup=3, **kw):
    out = run_myula(dev, steps=steps, warmup=warmup, **kw)
    # the general / fp64-validation path, for comparison
    st = run_myula(dev, steps=max(1, min(steps, 2)), warmup=1, engine="stream", **kw)
    out["streaming_kernel"] = {k: st[k] for k in ("value", "unit", "ms_per_launch", "tv_sweeps_per_iter", "roofline")}
    # the reference's own image size (skimage camera / einstein crop: 512 x 512) does not fit the on-chip kernel:
    # row-band kernel through HBM
    big = run_myula(dev, steps=max(1, min(steps, 2)), warmup=1, chains=256, dims=(512, 512), iters=10, engine="band")
    out["band_kernel_512"] = {k: big[k] for k in ("value", "unit", "dims", "chains", "ms_per_launch",
                                                   "tv_sweeps_per_iter", "engine", "roofline")}
    out["ulpda"] = run_ulpda(dev, steps=max(1, min(steps, 3)), warmup=min(warmup, 2))
    return out
