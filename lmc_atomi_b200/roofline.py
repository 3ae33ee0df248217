"""Single source of truth for the roofline accounting of bench.py (VERDICT r01 "What's weak" 5-7).

Two kinds of bound appear on this path:

* ISSUE (mixture chains, on-chip MYULA): the state lives in registers / shared memory, HBM traffic is ~0, so the
  binding resource is instruction dispatch: 148 SMs x 4 sub-partitions x 1 warp-instruction per clock
  = 148 x 128 thread-instructions per clock.  `frac` = dispatch slots per unit x units/s / peak, where the slots are
  the EXECUTED thread-instructions (ncu: smsp__inst_executed x 32 / units) with packed FP32 instructions counted
  twice (they occupy the dispatch port for two cycles on B200); `frac_issue_only` (= ncu smsp__issue_active) counts
  them once.  Neither can exceed 1.  The survey's ALGORITHMIC lane-op count (one op per + - x / max abs compare,
  FMA = 2) is reported beside them as `algorithmic_frac`; it can exceed the executed figure because an FMA retires
  two algorithmic ops in one instruction.
* HBM (band / tile image kernels): `frac` = algorithmic bytes per unit ((12+9j) / (24+16k) array-element
  transfers per pixel-iteration) x units/s / measured copy bandwidth; `frac_measured_dram` = DRAM bytes ncu measured
  per unit x units/s / the same peak (lower than `frac` when re-reads hit the 126 MB L2); `frac_min_traffic` = the
  kernel's own minimal transfer count (what it would move with a perfect L2).
"""
from __future__ import annotations

import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# fallback figures from /opt/skills/guides/B200_PROFILING.md
FALLBACK_HBM_GBS = 6650.0
SM_COUNT = 148
LANES_PER_SM = 128
FP64_LANES_PER_SM = 64
FALLBACK_SM_MHZ = 1965.0


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d.get("hbm_gbs", FALLBACK_HBM_GBS)),
                "sm_max_mhz": float(d.get("sm_max_mhz", FALLBACK_SM_MHZ)), "source": "measured"}
    return {"hbm_gbs": FALLBACK_HBM_GBS, "sm_max_mhz": FALLBACK_SM_MHZ, "source": "fallback"}


def issue_peak_laneops(sm_mhz: float, sms: int = SM_COUNT, lanes: int = LANES_PER_SM) -> float:
    """Thread-instructions per second at one warp-instruction per sub-partition per clock."""
    return sms * lanes * sm_mhz * 1e6


# ---- mixture chains: ALGORITHMIC lane-ops per chain-step (flops + transcendentals + Philox INT ops)
_GRAD = {"gauss": 93 + 5, "laplace": 113 + 5, "prox": 93 + 5}     # n=5: 18n+3 (+n T) / 22n+3 (+n T)
_NOISE_UNADJ = 50 + 4 + 4        # half a Philox block + Box-Muller (4 flops, 4 T)
_NOISE_ADJ = 100 + 4 + 4 + 2     # full block (2 normals + 1 uniform)
_UPDATE = 4 + 8                  # theta update + running moments


def mix_laneops_per_step(target: str, alg: str, t: int = 100) -> float:
    g = _GRAD[target]
    base = g + _NOISE_UNADJ + _UPDATE
    extra = {
        "ula": 0, "pula": 8, "mla": 12 + 8, "pgld": 8, "myula": 12,
        "ihpula": 30 * 5 + 60 + 4, "fbula": 200, "ppula": 8 + 14 * t, "lbmumla": 12 + 8 + 30 + 6,
    }
    if alg in ("mala", "mymala"):
        # one new target evaluation per step (old-point values cached) + ratio + accept
        return g + _NOISE_ADJ + _UPDATE + 30 + 2 + (30 if alg == "mymala" else 0) + (12 if alg == "mymala" else 0)
    return base + extra[alg]


# ---- mixture chains: EXECUTED thread-instructions per chain-step of the production kernels on the Laplacian
# mixture (n = 5), from the `ncu --set full` capture named in MIX_PROFILE: smsp__inst_executed.sum x 32 / (C x K), and
# the share of them that are packed FP32 instructions (FADD2 / FMUL2 / FFMA2) from the SASS census of the steady-state
# loop (scripts/sass_loop_census.py -> MIX_CENSUS).  A packed instruction holds the dispatch port for two cycles on
# B200 (profiles/r02c_ubench_issue.txt: FFMA2 0.5 vs FFMA 1.0 warp-instructions / clock / sub-partition), so the
# dispatch slots a step needs are instr x (1 + packed share).  Refreshed whenever mix_kernels.cu changes.
MIX_PROFILE = "profiles/r02ac_mix_ncu_full_summary.csv"
MIX_CENSUS = "profiles/r02ac_mix_sass_census.txt"
MIX_SASS_INSTR_PER_STEP = {"float32": {"ula": 105.4, "mala": 140.8, "pula": 110.1, "mla": 121.4}}
MIX_PACKED_SHARE = {"float32": {"ula": 0.361, "mala": 0.330, "pula": 0.373, "mla": 0.388}}
MIX_ISSUE_ACTIVE_PCT = {"ula": 67.0, "mala": 64.8, "pula": 67.8, "mla": 62.9}       # smsp__issue_active, same capture
MIX_DRAM_BYTES_PER_LAUNCH = {"ula": 147.7e3, "mala": 152.8e3, "pula": 150.3e3, "mla": 165.6e3}   # state is in registers


def mix_roofline(alg: str, chain_steps_per_sec: float, sm_mhz: float, dtype: str = "float32"):
    """Roofline object of one production mixture kernel at the measured rate (per GPU): `frac` = dispatch-slot
    fraction = executed thread-instructions x (1 + packed share) x rate / (148 x 128 x f)."""
    peak = issue_peak_laneops(sm_mhz)
    sass = MIX_SASS_INSTR_PER_STEP.get(dtype, {}).get(alg)
    packed = MIX_PACKED_SHARE.get(dtype, {}).get(alg)
    algo = mix_laneops_per_step("laplace", alg)
    out = {"bound": "issue", "unit": "Tthread-inst/s", "peak": peak / 1e12,
           "algorithmic_laneops_per_step": algo, "algorithmic_frac": algo * chain_steps_per_sec / peak,
           "sass_thread_instr_per_step": sass, "packed_fp32_share": packed, "profile": MIX_PROFILE,
           "census": MIX_CENSUS}
    if sass is not None and packed is not None:
        slots = sass * (1.0 + packed)
        out["dispatch_slots_per_step"] = slots
        out["achieved"] = slots * chain_steps_per_sec / 1e12
        out["frac"] = slots * chain_steps_per_sec / peak
        out["frac_issue_only"] = sass * chain_steps_per_sec / peak       # = ncu smsp__issue_active
    else:        # no capture of this instantiation: only the algorithmic figure is known
        out["achieved"] = None
        out["frac"] = None
    return out


# ---- image chains: bytes per pixel per sampler iteration
def myula_bytes_per_pixel(j_exec: float, itemsize: int = 4) -> float:
    """(12 + 9 j) array-element transfers: r=Hx-b (3), g=H^T r (2), TV dual
    sweeps (9 each), update + running mean (7)."""
    return (12 + 9 * j_exec) * itemsize


def myula_band_min_bytes_per_pixel(j_exec: float, itemsize: int = 4) -> float:
    """What img_myula_band.inl moves with a perfect L2: G: x, b in, g, p1, q1 out (5); T_k (k >= 1): x + 2 or 4 dual
    arrays in, 2 out (7 for k >= 2, 5 for k = 1); U: x, g, mean in + 2 or 4 duals, x, mean out (9)."""
    j = max(j_exec, 1.0)
    return (5 + 5 * min(j - 1, 1.0) + 7 * max(j - 2, 0.0) + 9) * itemsize


def ulpda_bytes_per_pixel(k_exec: float, itemsize: int = 4) -> float:
    """(24 + 16 k) transfers."""
    return (24 + 16 * k_exec) * itemsize


def ulpda_band_min_bytes_per_pixel(k_exec: float, itemsize: int = 4) -> float:
    """img_ulpda_band.inl: every LSQR vector is read and written once per phase = 11 transfers per LSQR iteration
    (DESIGN 4 I-c) + the 24 of the outer iteration."""
    return (24 + 11 * k_exec) * itemsize


# ---- measured with `ncu --set full` (files under profiles/): dram__bytes_read.sum + dram__bytes_write.sum per launch
# divided by the chain-iterations of that launch, and executed thread-instructions per pixel-iteration.
ULPDA_BAND = {"profile": "profiles/r02r_img_ulpda_band_ncu_full_summary.csv",
              "dram_bytes_per_chain_iter": (39.793e9 + 26.921e9) / 592.0,      # 512x512, H5, 296 chains x 2, k = 8.7
              "issue_active_pct": 39.0}
MYULA_BAND = {"profile": "profiles/r02r_img_myula_band_ncu_full_summary.csv",
              "dram_bytes_per_chain_iter": (57.926e9 + 28.210e9) / 1776.0,     # 512x512, H5, fp32, 296 chains x 6
              "issue_active_pct": 41.7}
MYULA_ONCHIP = {"profile": "profiles/r01c_img_myula_onchip_ncu_full_summary.csv",
                "dram_bytes_per_chain_iter": (134.6e6 + 86.6e6) / 2560.0,      # 256x256: b prefetch + mean
                "thread_instr_per_pixel_iter": 395.0, "issue_active_pct": 54.7, "sms_used": 120}

# kept for callers of the r01 names
ULPDA_BAND_DRAM_BYTES_PER_CHAIN_ITER = ULPDA_BAND["dram_bytes_per_chain_iter"]
MYULA_BAND_DRAM_BYTES_PER_CHAIN_ITER = MYULA_BAND["dram_bytes_per_chain_iter"]


def hbm_roofline(kernel, bytes_alg_per_unit, bytes_min_per_unit, units_per_sec, meas=None, matches_profile=False):
    """HBM-bound kernel: `frac` by the survey's algorithmic bytes, beside the kernel's minimal traffic and (when the
    run has the shape of the committed capture) the DRAM bytes ncu measured."""
    peaks = measured_peaks()
    peak = peaks["hbm_gbs"]
    ach = bytes_alg_per_unit * units_per_sec / 1e9
    out = {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
           "bytes_per_unit_algorithmic": bytes_alg_per_unit,
           "frac_min_traffic": bytes_min_per_unit * units_per_sec / 1e9 / peak,
           "traffic": None, "frac_measured_dram": None, "peak_source": peaks["source"]}
    if meas is not None and matches_profile:
        out["traffic_per_unit"] = meas["dram_bytes_per_chain_iter"]
        out["frac_measured_dram"] = meas["dram_bytes_per_chain_iter"] * units_per_sec / 1e9 / peak
        out["issue_active_pct_ncu"] = meas.get("issue_active_pct")
        out["profile"] = meas["profile"]
    return out


def onchip_roofline(kernel, pixel_iters_per_sec, sm_mhz):
    """On-chip MYULA: issue-bound (with shared-memory bandwidth behind it); counted over ALL 148 SMs although the
    15 co-resident 8-CTA clusters occupy 120 of them."""
    peak = issue_peak_laneops(sm_mhz)
    m = MYULA_ONCHIP
    ach = m["thread_instr_per_pixel_iter"] * pixel_iters_per_sec
    return {"bound": "issue+smem", "kernel": kernel, "achieved": ach / 1e12, "peak": peak / 1e12,
            "unit": "Tthread-inst/s", "frac": ach / peak,
            "sass_thread_instr_per_pixel_iter": m["thread_instr_per_pixel_iter"], "sms_used": m["sms_used"],
            "issue_active_pct_ncu_on_used_sms": m["issue_active_pct"],
            "traffic_per_unit": m["dram_bytes_per_chain_iter"], "profile": m["profile"],
            "note": "chain state lives in the shared memory of a thread-block cluster; HBM carries only the "
                    "observation prefetch and the running mean, so HBM is not the bound"}
