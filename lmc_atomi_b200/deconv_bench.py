"""Image-chain half of the benchmark (BASELINE.json configs[3], configs[4]):

* MYULA deconvolution, 256x256, model M1 (L2(H5) + TV(sigma=0.3, niter=10, rtol=1e-4)),
  tau=0.1125, gamma=0.5625 (prox_lmc_deconv.py:88-94, 465-473), a batch of independent chains
  whose state is far larger than L2;
* ULPDA deconvolution, 512x512, model M1 (L2(H5, niter=50, warm) + L21(0.3) + Gradient),
  tau=0.534375, mu=1, theta=1, gfirst=False (prox_lmc_deconv.py:455-464), 512 chains per GPU (BASELINE.json
  configs[4]: 4096 chains over 8 GPUs).

Input images: crops of the reference's own fig/einstein.png (shipped as tests/golden/einstein.png):
config 4 = the centre 512x512 crop, 2x2 mean-pooled to 256x256; config 5 = rows 77:589, cols 182:694.  When the
fixture is missing a deterministic piecewise-smooth stand-in of the same shape is used and named in `image`.
Blurred by the CUDA H5 operator plus N(0, 0.75^2) noise.  Nothing here imports oracle/.
"""
from __future__ import annotations

import numpy as np

import os

from . import imaging, roofline

SIGMA, TAU_REG = 0.75, 0.3
EINSTEIN = os.path.join(roofline.ROOT, "tests", "golden", "einstein.png")


def bench_image(dims):
    """(image, name): the einstein crops SURVEY 8(d) prescribes for configs 4 and 5, else the synthetic stand-in."""
    if os.path.isfile(EINSTEIN) and tuple(dims) in ((256, 256), (512, 512)):
        from PIL import Image
        full = np.asarray(Image.open(EINSTEIN).convert("L"), dtype=np.float64)      # 667 x 877
        crop = full[77:589, 182:694]
        if tuple(dims) == (256, 256):
            return crop.reshape(256, 2, 256, 2).mean(axis=(1, 3)), "einstein centre crop 512x512, 2x2 mean-pooled"
        return crop, "einstein rows 77:589 cols 182:694"
    return synthetic_image(*dims), "synthetic piecewise-smooth stand-in"


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
    img, _ = bench_image(dims)
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


def _e2e(torch, dev, host_in, host_out, launch_group, n_groups, units, reps=2):
    """The same metric through HOST buffers (pinned, allocated by the caller outside the timed region): per
    repetition every chain's start state comes from host memory and its posterior mean goes back to it, inside the
    timed region.  The chains are processed in `n_groups` groups on two streams, so the H2D copy of group g+1 and
    the D2H copy of group g-1 overlap the kernel of group g.  MAX over ranks.

    launch_group(g, copy_in, copy_out): `copy_in()` enqueues the H2D of group g on the current stream, then the
    kernel, then `copy_out()` the D2H."""
    dist, world, _ = _world()
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    # one untimed repetition: first touch of the pinned pages and of the two streams (without it the figure moved by
    # 10 % between boxes, r02ad 33.1 k vs r02ah 29.5 k for ULPDA at an unchanged device rate)
    cur = torch.cuda.current_stream(dev)
    for st in streams:
        st.wait_stream(cur)
    for g in range(n_groups):
        with torch.cuda.stream(streams[g % 2]):
            launch_group(g)
    for st in streams:
        cur.wait_stream(st)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    cur = torch.cuda.current_stream(dev)
    for st in streams:
        st.wait_stream(cur)
    for _ in range(reps):
        for g in range(n_groups):
            with torch.cuda.stream(streams[g % 2]):
                launch_group(g)
    for st in streams:
        cur.wait_stream(st)
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) * 1e-3 / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    nin = sum(h.numel() * h.element_size() for h in host_in)
    nout = sum(h.numel() * h.element_size() for h in host_out)
    return {"value": world * units / float(t.item()), "unit": "chain-iters/s", "h2d_bytes_per_step": nin,
            "d2h_bytes_per_step": nout, "groups": n_groups,
            "note": "start states from pinned host memory, posterior means back to it, both inside the timed region; "
                    "chain groups double-buffered on two streams"}


def _pooled_mean(torch, mean):
    """The only collective of the image path: posterior mean pooled over all chains of all ranks."""
    dist, world, _ = _world()
    s = mean.double().sum(dim=0)
    if world > 1:
        dist.all_reduce(s)
    return s / (mean.shape[0] * world)


def run_myula(dev, steps=3, warmup=3, chains=1024, dims=(256, 256), iters=250, dtype="float32", engine="auto",
              e2e_groups=4, ncvx=None):
    import torch
    _, eng = make_engine(dims, dev, dtype)
    _, image_name = bench_image(dims)
    L = 1.0 / SIGMA ** 2
    gamma, tau = 1.0 / L, 0.2 / L
    P = dims[0] * dims[1]
    x = torch.zeros((chains, P), dtype=eng.tdt, device=dev)
    mean = torch.zeros_like(x)
    state = {"it": 0, "keep": []}
    _, world, rank = _world()

    def launch(c0=0, c1=chains, advance=True):
        state["keep"].append(eng.myula(x[c0:c1], iters, tau, gamma, TAU_REG, 10, 1e-4, seed=0, mean=mean[c0:c1],
                                       ncvx=ncvx, iter_offset=state["it"],
                                       stat_count0=state["it"] if advance else 0, engine=engine,
                                       chain_offset=rank * chains + c0))
        del state["keep"][:-8]
        if advance:
            state["it"] += iters

    launch()                # burn-in so that the TV early exit is in its steady regime
    torch.cuda.synchronize(dev)
    eng.counters.zero_()
    it0 = state["it"]
    secs = _time_launches(torch, dev, launch, steps, warmup)
    c = eng.read_counters()
    n_it = state["it"] - it0
    pooled = _pooled_mean(torch, mean)
    # e2e: pinned host buffers allocated here, outside the timed region
    h_x = torch.zeros(x.shape, dtype=x.dtype).pin_memory()
    h_mean = torch.empty(mean.shape, dtype=mean.dtype).pin_memory()
    gsz = (chains + e2e_groups - 1) // e2e_groups

    def launch_group(g):
        c0, c1 = g * gsz, min(chains, (g + 1) * gsz)
        x[c0:c1].copy_(h_x[c0:c1], non_blocking=True)
        mean[c0:c1].zero_()
        launch(c0, c1, advance=False)
        h_mean[c0:c1].copy_(mean[c0:c1], non_blocking=True)

    e2e = _e2e(torch, dev, [h_x], [h_mean], launch_group, e2e_groups, chains * iters)
    j = c["tv_sweeps"] / float(n_it * chains)
    item = 4 if dtype == "float32" else 8
    bpp = roofline.myula_bytes_per_pixel(j, item)
    peaks = roofline.measured_peaks()
    rate = chains * iters / secs                      # chain-iters/s on this GPU
    onchip = engine == "onchip" or (engine == "auto" and eng.onchip_supported() and ncvx is None)
    band = not onchip and (engine == "band" or (engine == "auto" and eng.myula_band_supported()))
    if ncvx is not None:     # + the Moreau-envelope gradient: written in phase G, read in phase U (MC-TV)
        bpp += 2 * item
    if onchip:
        kernel = "img_myula_onchip_kernel<32,5,%d>" % (1 if dims == (256, 256) else 0)
        policy = ("chain state resident in the shared memory of an 8-CTA cluster for all iterations of a launch; "
                  "%d chains per launch, state of the batch (x + mean) %.0f MB >> 126 MB L2 only for the start/end "
                  "load/store" % (chains, 2 * chains * P * item / 1e6))
        roof = roofline.onchip_roofline(kernel, rate * P, peaks["sm_max_mhz"])
        roof["hbm_equivalent_of_streaming_formulation_GBs"] = bpp * P * rate / 1e9
    else:
        kernel = ("img_myula_band_kernel<%s,5,2>" if band else "img_myula_kernel<%s>") % ("float" if item == 4 else "double")
        policy = "per-chain arrays total far more than the 126 MB L2"
        roof = roofline.hbm_roofline(kernel, bpp * P, (roofline.myula_band_min_bytes_per_pixel(j, item) +
                                                       (2 * item if ncvx is not None else 0)) * P, rate,
                                     roofline.MYULA_BAND, band and dims == (512, 512) and item == 4 and ncvx is None)
    return {"metric": "myula_deconv_chain_iters_per_sec", "value": world * rate,
            "unit": "chain-iters/s", "n_gpus": world, "pooled_mean_finite": bool(torch.isfinite(pooled).all().item()),
            "pixel_iters_per_sec": world * rate * P, "dtype": dtype, "dims": list(dims), "image": image_name,
            "chains": chains, "iters_per_launch": iters, "ms_per_launch": secs * 1e3, "tv_sweeps_per_iter": j,
            "e2e": e2e, "engine": "onchip" if onchip else ("band" if band else "stream"),
            "state_bytes": int(10 * chains * P * item), "l2_policy": policy, "roofline": roof}


def run_ulpda(dev, steps=3, warmup=2, chains=512, dims=(512, 512), iters=100, dtype="float32", engine="auto", group=0,
              e2e_groups=4, ncvx=None, isotropic=True):
    import torch
    _, eng = make_engine(dims, dev, dtype)
    _, image_name = bench_image(dims)
    L = 1.0 / SIGMA ** 2
    tau, mu = 0.95 / L, 1.0
    P = dims[0] * dims[1]
    z = lambda *s: torch.zeros(s, dtype=eng.tdt, device=dev)
    x, xhat, y, warm, mean = z(chains, P), z(chains, P), z(chains, 2, P), z(chains, P), z(chains, P)
    state = {"it": 0, "keep": []}
    _, world, rank = _world()

    def launch(c0=0, c1=chains, advance=True):
        state["keep"].append(eng.ulpda(x[c0:c1], xhat[c0:c1], y[c0:c1], warm[c0:c1], iters, tau, mu, 1.0, TAU_REG,
                                       isotropic, False, 50, seed=0, mean=mean[c0:c1], ncvx=ncvx,
                                       iter_offset=state["it"],
                                       stat_count0=state["it"] if advance else 0, engine=engine, group=group,
                                       chain_offset=rank * chains + c0))
        del state["keep"][:-8]
        if advance:
            state["it"] += iters

    launch()
    torch.cuda.synchronize(dev)
    eng.counters.zero_()
    secs = _time_launches(torch, dev, launch, steps, warmup)
    c = eng.read_counters()
    pooled = _pooled_mean(torch, mean)
    h_x = torch.zeros(x.shape, dtype=x.dtype).pin_memory()
    h_mean = torch.empty(mean.shape, dtype=mean.dtype).pin_memory()
    gsz = (chains + e2e_groups - 1) // e2e_groups

    def launch_group(g):
        c0, c1 = g * gsz, min(chains, (g + 1) * gsz)
        x[c0:c1].copy_(h_x[c0:c1], non_blocking=True)       # config 5 starts every chain from x0 (host), y = 0
        xhat[c0:c1].copy_(x[c0:c1], non_blocking=True)
        y[c0:c1].zero_()
        warm[c0:c1].zero_()
        mean[c0:c1].zero_()
        launch(c0, c1, advance=False)
        h_mean[c0:c1].copy_(mean[c0:c1], non_blocking=True)

    e2e = _e2e(torch, dev, [h_x], [h_mean], launch_group, e2e_groups, chains * iters, reps=1)
    k = c["lsqr_iterations"] / max(1.0, float(c["lsqr_solves"]))
    item = 4 if dtype == "float32" else 8
    bpp = roofline.ulpda_bytes_per_pixel(k, item)
    if ncvx is not None:     # + arg (write, read) and the pre-stepped argument (write, read)
        bpp += 4 * item
    rate = chains * iters / secs
    band = engine == "band" or (engine == "auto" and eng.band_supported())
    kernel = ("img_ulpda_band_kernel<%s,5,16,2>" if band else "img_ulpda_kernel<%s>") % ("float" if item == 4 else "double")
    roof = roofline.hbm_roofline(kernel, bpp * P, (roofline.ulpda_band_min_bytes_per_pixel(k, item) +
                                                   (4 * item if ncvx is not None else 0)) * P, rate,
                                 roofline.ULPDA_BAND, band and dims == (512, 512) and item == 4 and ncvx is None)
    return {"metric": "ulpda_deconv_chain_iters_per_sec", "value": world * rate,
            "unit": "chain-iters/s", "n_gpus": world, "pooled_mean_finite": bool(torch.isfinite(pooled).all().item()),
            "pixel_iters_per_sec": world * rate * P, "dtype": dtype, "dims": list(dims), "image": image_name,
            "chains": chains, "iters_per_launch": iters, "ms_per_launch": secs * 1e3, "lsqr_iters_per_solve": k,
            "e2e": e2e, "engine": "band" if band else "tile", "group": group, "roofline": roof}


def run(dev, steps=3, warmup=3, **kw):
    out = run_myula(dev, steps=steps, warmup=warmup, **kw)
    # the general / fp64-validation path, for comparison
    st = run_myula(dev, steps=1, warmup=1, engine="stream", iters=10, **kw)
    out["streaming_kernel"] = {k: st[k] for k in ("value", "unit", "ms_per_launch", "tv_sweeps_per_iter", "roofline")}
    # the reference's own image size (skimage camera / einstein crop: 512 x 512) does not fit the on-chip kernel:
    # row-band kernel through HBM
    big = run_myula(dev, steps=max(1, min(steps, 2)), warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
    out["band_kernel_512"] = {k: big[k] for k in ("value", "unit", "dims", "image", "chains", "iters_per_launch",
                                                   "ms_per_launch", "tv_sweeps_per_iter", "engine", "e2e", "roofline")}
    out["ulpda"] = run_ulpda(dev, steps=max(1, min(steps, 2)), warmup=1)
    # non-convex data terms evaluated in-kernel (models M2 / M3 of prox_lmc_deconv.py:106-113, 541-703): MC-TV = a
    # 7-point stencil, ME-TV = an inner TV prox of 50 iterations per sampler iteration
    keys = ("value", "unit", "dims", "image", "chains", "iters_per_launch", "ms_per_launch", "engine", "roofline")
    mc = imaging.MoreauTerm("mc", TAU_REG, 15.0, True)
    me = imaging.MoreauTerm("me", TAU_REG, 15.0, True, 50, 1e-4)
    nc = {}
    r = run_myula(dev, steps=1, warmup=1, chains=256, dims=(512, 512), iters=20, engine="band", ncvx=mc)
    nc["myula_mctv_M2"] = {k: r[k] for k in keys + ("tv_sweeps_per_iter",)}
    r = run_myula(dev, steps=1, warmup=1, chains=64, dims=(512, 512), iters=5, engine="band", ncvx=me)
    nc["myula_metv_M3"] = {k: r[k] for k in keys + ("tv_sweeps_per_iter",)}
    nc["myula_metv_M3"]["roofline"]["note"] = ("the inner TV(dims, 1, 50).prox runs 51 more dual sweeps per iteration; "
                                               "the algorithmic byte count above covers the outer sampler only")
    r = run_ulpda(dev, steps=1, warmup=1, chains=128, dims=(512, 512), iters=10, ncvx=mc, isotropic=False)
    nc["ulpda_mctv_M2"] = {k: r[k] for k in keys + ("lsqr_iters_per_solve",)}
    out["ncvx"] = nc
    return out
