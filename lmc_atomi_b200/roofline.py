"""Single source of truth for the algorithmic work per unit and the peaks.

Mixture chains: zero steady-state HBM traffic -> bound by FP32/INT32 issue
(SMs x 128 lanes x clock) and the MUFU pipe.  Counts are *algorithmic* lane-ops per
chain-step from the reference math with per-component constants precomputed
(one op per + - x / max abs compare; T = one transcendental), not the SASS count.

Image chains: HBM-bound once batched.  Bytes are array-element transfers per pixel
per iteration x sizeof(real), using the inner iterations actually executed.
"""
from __future__ import annotations

import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# fallback figures from /opt/skills/guides/B200_PROFILING.md
FALLBACK_HBM_GBS = 6650.0
SM_COUNT = 148
LANES_PER_SM = 128
FALLBACK_SM_MHZ = 1965.0


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d.get("hbm_gbs", FALLBACK_HBM_GBS)),
                "sm_max_mhz": float(d.get("sm_max_mhz", FALLBACK_SM_MHZ)), "source": "measured"}
    return {"hbm_gbs": FALLBACK_HBM_GBS, "sm_max_mhz": FALLBACK_SM_MHZ, "source": "fallback"}


def issue_peak_laneops(sm_mhz: float, sms: int = SM_COUNT) -> float:
    """FP32/INT32 lane-ops per second at one op per lane per clock."""
    return sms * LANES_PER_SM * sm_mhz * 1e6


# ---- mixture chains: lane-ops per chain-step (flops + transcendentals + Philox INT ops)
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


# ---- image chains: bytes per pixel per sampler iteration
def myula_bytes_per_pixel(j_exec: float, itemsize: int = 4) -> float:
    """(12 + 9 j) array-element transfers: r=Hx-b (3), g=H^T r (2), TV dual
    sweeps (9 each), update + running mean (7)."""
    return (12 + 9 * j_exec) * itemsize


def ulpda_bytes_per_pixel(k_exec: float, itemsize: int = 4) -> float:
    """(24 + 16 k) transfers."""
    return (24 + 16 * k_exec) * itemsize


# ---- measured once per round with `ncu --set full` (profiles/r01e_*): dram__bytes_read.sum + dram__bytes_write.sum
# per launch of the kernel, used for `roofline.traffic`.  Captures are of smaller launches than the bench's
# (ncu replays every kernel ~40 times); the per-unit figures below are what scales.
MIX_DRAM_BYTES_PER_LAUNCH = {"ula": 125.2e3, "mala": 104.7e3, "pula": 126.5e3, "mla": 131.6e3}   # state is in registers
MIX_ISSUE_ACTIVE_PCT = {"ula": 83.0, "mala": 78.3, "pula": 84.4, "mla": 78.3}       # smsp__issue_active, same capture
ULPDA_BAND_DRAM_BYTES_PER_CHAIN_ITER = (34.459e9 + 23.259e9) / 512.0   # 512x512, H5, k = 8: 256 chains x 2 iterations
MYULA_BAND_DRAM_BYTES_PER_CHAIN_ITER = (74.037e9 + 36.859e9) / 2560.0   # 512x512, H5, fp32: 256 chains x 10 iterations (r01g)
