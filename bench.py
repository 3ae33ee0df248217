"""bench.py -- headline benchmark of the sampler hot path (contract: task statement section 4).

Workload (BASELINE.json configs[1]): lmc_laplace.py Laplacian mixture, n=5, lamda=0.1,
alpha=0.5, gamma=0.05, K=50000 steps, 2^20 parallel chains PER GPU (weak scaling),
samplers ULA, MALA, PULA, MLA.  One bench "step" = all four samplers, each one
persistent launch of C chains x K steps with in-kernel Philox noise, posterior moments
and the 100x100 histogram reduced on device, (N>1) one NCCL all-reduce of the statistics.

    python bench.py [--gpus N --steps K --warmup W]         # our arm
    python bench.py --impl reference ...                    # numpy CPU arm (oracle port)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ALGS = ("ula", "mala", "pula", "mla")
GAMMA, LAMDA, ALPHA = 0.05, 0.1, 0.5
M_PRE = np.array([[1.0, 0.1], [0.1, 0.5]])      # lmc_laplace.py:315
BETA = np.array([0.7, 0.3])                      # lmc_laplace.py:318
MUS = [np.array(m) for m in ([0., 0.], [-2., 3.], [2., -3.], [3., 3.], [-2., -2.])]  # lmc_laplace.py:230-245
METRIC = "langevin_chain_steps_per_sec"
UNIT = "chain-steps/s"


def alg_kwargs(alg):
    return {"pula": {"M": M_PRE}, "mla": {"beta": BETA}}.get(alg, {})


# ------------------------------------------------------------------------- CPU arm
def _cpu_worker(args):
    """One host core: the oracle port (numpy restatement of lmc_laplace.py:110-216) on one chain."""
    seed, k_unadj, k_adj = args
    from oracle import mix_ref as O
    tg = O.LaplaceMix(MUS, np.ones(5) * ALPHA, np.ones(5) / 5, LAMDA)
    out = {}
    th0, xi, u = O.reference_noise(seed, max(k_unadj, k_adj), uniforms=True)
    for alg in ALGS:
        t0 = time.perf_counter()
        if alg == "mala":
            O.run_adjusted("mala", tg, th0, xi[:k_adj], u[:k_adj], GAMMA)
            n = k_adj
        else:
            args_ = {"pula": (M_PRE,), "mla": (BETA,)}.get(alg, ())
            O.run_unadjusted(alg, tg, th0, xi[:k_unadj], GAMMA, *args_)
            n = k_unadj
        out[alg] = (n, time.perf_counter() - t0)
    return out


def cpu_sample(k_unadj=3000, k_adj=800, cores=None):
    """Times the CPU port on all host cores (one chain per process).  Returns the rate for
    the benchmark's sampler mix (equal steps per sampler) in chain-steps/s."""
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, [(s, k_unadj, k_adj) for s in range(cores)])
    wall = time.perf_counter() - t0
    per_alg = {}
    for alg in ALGS:
        # per-core seconds per step, averaged over cores
        sps = np.mean([r[alg][1] / r[alg][0] for r in res])
        per_alg[alg] = 1.0 / sps
    mix_rate_core = len(ALGS) / sum(1.0 / per_alg[a] for a in ALGS)
    return {"value": mix_rate_core * cores, "per_core": mix_rate_core, "per_alg_per_core": per_alg,
            "cores": cores, "wall_s": wall,
            "sample": f"{cores} chains (1 per core) x ({k_unadj} steps ULA/PULA/MLA + {k_adj} steps MALA), numpy float64"}


def _cpu_image_worker(args):
    """One host core: the oracle port of the image samplers (numpy + scipy FFT convolution + scipy LSQR,
    oracle/deconv_ref.py restating algs.py:295-587) on one chain."""
    which, seed, dims, iters = args
    from oracle import deconv_ref as D
    from lmc_atomi_b200.deconv_bench import bench_image
    pb = D.build_problem(bench_image(dims)[0], seed=0)      # the image the GPU arm uses (einstein crop)
    t0 = time.perf_counter()
    if which == "myula":
        D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"],
                                         niter=iters, seed=seed)
    else:
        D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1.,
                                       niter=iters, seed=seed, gfirst=False)
    return iters, time.perf_counter() - t0


def cpu_image_sample(which, dims, iters, cores=None):
    """chain-iters/s of the CPU port with one independent chain per host core."""
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.map(_cpu_image_worker, [(which, s, dims, iters) for s in range(cores)])
    per_core = float(np.mean([n / t for n, t in res]))
    return {"value": per_core * cores, "unit": "chain-iters/s", "cores": cores, "kind": "port", "per_core": per_core,
            "sample": f"{cores} chains (1 per core) x {iters} iterations of {dims[0]}x{dims[1]} {which.upper()}, "
                      "numpy float64 + scipy FFT convolution" + (" + scipy LSQR" if which == "ulpda" else "")}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times, vals = [], []
    for i in range(args.warmup + args.steps):
        r = cpu_sample(k_unadj=args.cpu_steps, k_adj=max(args.cpu_steps // 4, 50))
        if i >= args.warmup:
            times.append(r["wall_s"])
            vals.append(r["value"])
    v = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args):
    return {"workload": "lmc_laplace.py Laplacian mixture n=5 lamda=0.1 alpha=0.5 gamma=0.05: "
                        f"ULA+MALA+PULA+MLA, K={args.K} steps x {args.chains} chains per GPU",
            "chains_per_gpu": args.chains, "K": args.K, "samplers": list(ALGS),
            "l2_policy": "no HBM-resident inputs in the timed kernels (state lives in registers); "
                         "final-state/statistics buffers are rewritten every step"}


# ------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                    "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in o.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from lmc_atomi_b200 import roofline
    from lmc_atomi_b200.lmc_laplace import LangevinMonteCarloLaplacian

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    C, K = args.chains, args.K
    sampler = LangevinMonteCarloLaplacian(MUS, np.ones(5) * ALPHA, np.ones(5) / 5, LAMDA, K=K, seed=0)

    ev = {a: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for a in ALGS}

    def one_step(step_idx, timed):
        """device-resident step: 4 persistent launches (+ reductions); returns stats"""
        out = {}
        for alg in ALGS:
            if timed:
                ev[alg][0].record()
            st = sampler.sample_chains(alg, GAMMA, chains=C, K=K, dtype="float32", seed=step_idx,
                                       chain_offset=rank * C, device=dev, _sync=False, **alg_kwargs(alg))
            if timed:
                ev[alg][1].record()
            out[alg] = st
        if world > 1:   # the only collective of the path: posterior statistics
            for alg in ALGS:
                dist.all_reduce(out[alg].extra["moments"])
                dist.all_reduce(out[alg].extra["hist"])
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        one_step(i, False)
    barrier()
    t_alg = {a: 0.0 for a in ALGS}
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        start.record()
        for i in range(args.steps):
            one_step(args.warmup + i, True)
            # per-sampler kernel time (events read after the step; sync is outside the kernels)
            torch.cuda.synchronize()
            for a in ALGS:
                t_alg[a] += ev[a][0].elapsed_time(ev[a][1])
        stop.record()
        barrier()
    ms = start.elapsed_time(stop)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    units = float(world) * C * K * len(ALGS) * args.steps
    value = units / (ms * 1e-3)

    # ---- e2e through the public API with HOST buffers (pinned): start states in, statistics out
    h_theta = torch.randn(C, 2).pin_memory()
    h_out = torch.empty(C, 2).pin_memory()
    e2e_ms = 0.0
    barrier()
    es, ee = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_e2e = max(1, min(args.steps, 2))
    es.record()
    for i in range(n_e2e):
        for alg in ALGS:
            d_theta = h_theta.to(dev, non_blocking=True)
            st = sampler.sample_chains(alg, GAMMA, chains=C, K=K, dtype="float32", seed=100 + i,
                                       chain_offset=rank * C, device=dev, theta0=d_theta, **alg_kwargs(alg))
            h_out.copy_(st.theta, non_blocking=True)
            _ = st.mean  # host read of the reduced statistics (sample_chains already copied them)
    ee.record()
    barrier()
    e2e_ms = es.elapsed_time(ee)
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = float(world) * C * K * len(ALGS) * n_e2e / (float(t.item()) * 1e-3)
    h2d = C * 2 * 4 * len(ALGS)
    d2h = (C * 2 * 4 + 6 * 8 + 100 * 100 * 8 + 4 * 0) * len(ALGS) + C * 4  # + MALA accept counts

    # image half of the metric (configs 3-4): every rank advances its own shard of the image chains
    extra = image_bench(args, dev) if not args.no_image else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = roofline.measured_peaks()
    clocks = clk.summary()
    sm_mhz_peak = peaks["sm_max_mhz"]
    per_alg = {}
    for a in ALGS:
        secs = t_alg[a] * 1e-3 / args.steps
        r = roofline.mix_roofline(a, C * K / secs, sm_mhz_peak)
        per_alg[a] = {"ms_per_launch": secs * 1e3, "chain_steps_per_sec_per_gpu": C * K / secs,
                      "sass_thread_instr_per_step": r["sass_thread_instr_per_step"], "frac": r["frac"],
                      "algorithmic_laneops_per_step": r["algorithmic_laneops_per_step"],
                      "algorithmic_frac": r["algorithmic_frac"],
                      "issue_active_pct_ncu": roofline.MIX_ISSUE_ACTIVE_PCT.get(a)}
    dom = max(ALGS, key=lambda a: t_alg[a])
    secs = t_alg[dom] * 1e-3 / args.steps
    roof = roofline.mix_roofline(dom, C * K / secs, sm_mhz_peak)
    roof.update({"kernel": f"mix_chain_kernel<float,LAPLACE,{dom.upper()},5,PROD>",
                 "traffic": roofline.MIX_DRAM_BYTES_PER_LAUNCH.get(dom),
                 "peak_source": f"148 SMs x 128 lanes x {sm_mhz_peak:.0f} MHz ({peaks['source']} sm_max_mhz)",
                 "note": "register-resident chains: no HBM traffic in steady state, so the bound is instruction "
                         "issue (north_star: FP32 issue rate).  frac = EXECUTED thread-instructions per step (ncu "
                         "capture named in `profile`) x rate / peak = ncu smsp__issue_active; algorithmic_frac uses "
                         "the survey's lane-op count (an FMA counts twice) and is kept for reference only",
                 "per_sampler": per_alg})
    f64 = fp64_line(args, sampler, dev, rank, world, roofline, sm_mhz_peak) if not args.no_f64 else None
    cpu = (cpu_sample(k_unadj=args.cpu_steps, k_adj=max(args.cpu_steps // 4, 50))
           if (world == 1 and not args.no_cpu) else None)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": 2 * len(ALGS) * args.steps, "roofline": roof}
    if f64 is not None:
        line["value_f64"] = f64
    if cpu is not None:
        line["cpu_baseline"] = {"value": cpu["value"], "unit": UNIT, "cores": cpu["cores"], "kind": "port",
                                "sample": cpu["sample"], "per_core": cpu["per_core"]}
    if extra is not None:
        if cpu is not None:   # the CPU port of the image samplers, timed beside them (bounded sample)
            extra["cpu_baseline"] = cpu_image_sample("myula", tuple(extra["dims"]), 8)
            extra["ulpda"]["cpu_baseline"] = cpu_image_sample("ulpda", tuple(extra["ulpda"]["dims"]), 2)
        line["myula_deconv"] = extra
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def fp64_line(args, sampler, dev, rank, world, roofline, sm_mhz):
    """The same workload in the reference's own precision (float64), shorter K (stated): the f32 headline then sits
    beside a same-precision number.  Bound: FP64 issue, 148 SMs x 64 DFMA lanes x clock."""
    import torch
    C, K = args.chains, max(200, args.K // 25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for timed in (False, True):
        if timed:
            torch.cuda.synchronize()
            ev0.record()
        for alg in ALGS:
            sampler.sample_chains(alg, GAMMA, chains=C, K=K, dtype="float64", seed=7, chain_offset=rank * C,
                                  device=dev, _sync=False, **alg_kwargs(alg))
        if timed:
            ev1.record()
    torch.cuda.synchronize()
    secs = ev0.elapsed_time(ev1) * 1e-3
    rate = C * K * len(ALGS) / secs
    peak = roofline.issue_peak_laneops(sm_mhz, lanes=roofline.FP64_LANES_PER_SM)
    algo = sum(roofline.mix_laneops_per_step("laplace", a) for a in ALGS) / len(ALGS)
    return {"value": rate * world, "unit": UNIT, "dtype": "f64", "K": K, "chains_per_gpu": C, "ms": secs * 1e3,
            "note": "per-GPU rate x n_gpus (rank 0 timing); same kernels instantiated for double, IEEE libdevice "
                    "transcendentals; trajectories within 1e-12 of the reference (tests/test_mix_gpu.py)",
            "roofline": {"bound": "fp64-issue", "peak": peak / 1e12, "unit": "Tlane-op/s",
                         "algorithmic_frac": algo * rate / peak, "frac": None,
                         "note": "no ncu capture of the fp64 instantiation: only the algorithmic figure is given"}}


def image_bench(args, dev):
    """Second half of BASELINE.json's metric: MYULA deconvolution iterations/s (config 4)."""
    try:
        from lmc_atomi_b200 import deconv_bench
    except ImportError:
        return None
    return deconv_bench.run(dev, steps=args.steps, warmup=args.warmup)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--chains", type=int, default=1 << 20)
    ap.add_argument("--K", type=int, default=50000)
    ap.add_argument("--cpu-steps", type=int, default=3000)
    ap.add_argument("--no-image", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-f64", action="store_true", help="skip the float64 line (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
