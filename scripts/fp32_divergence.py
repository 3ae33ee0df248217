"""SURVEY 7 hard part 3 / VERDICT r01 weak #3: how long does an fp32 FREE-RUNNING chain stay within the north star's
1e-5 of the reference trajectory?  For every sampler: the fp64 CUDA chain (equal to the reference's numpy chain to
1e-12, tests/test_mix_gpu.py) and the fp32 CUDA chain on the SAME reference noise stream (default_rng(0)), K = 10000
(config 1); reports the first step whose relative deviation max|d| / max(1, |theta|) exceeds 1e-5 / 1e-4 / 1e-2, the
teacher-forced one-step error (fp32 step from the fp64 state) and, for MALA / MYMALA, the first differing decision."""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import ADJ_CASES, CASES, alg_params, make_samplers      # noqa: E402
from lmc_atomi_b200.mixtures import reference_noise                  # noqa: E402

K = 10000
out = {}
S = make_samplers(K=K)
for which, alg in CASES + ADJ_CASES:
    z, s, _ = S[which]
    adj = alg in ("mala", "mymala")
    kw, _ = alg_params(z, alg)
    th0, xi, u = reference_noise(0, K, 2, uniforms=adj)
    g = float(z["gamma"])
    t64, a64 = s.run_with_noise(alg, g, th0[None], xi[None], None if u is None else u[None], dtype="float64", **kw)
    t32, a32 = s.run_with_noise(alg, g, th0[None], xi[None], None if u is None else u[None], dtype="float32", **kw)
    t64, t32 = t64[0], t32[0]
    rel = np.abs(t32 - t64).max(axis=1) / np.maximum(1.0, np.abs(t64).max(axis=1))
    first = lambda tol: int(np.argmax(rel > tol)) if (rel > tol).any() else None
    # teacher-forced: one fp32 step from every fp64 state (chains = steps)
    starts = np.vstack([th0[None], t64[:-1]])
    tf, _ = s.run_with_noise(alg, g, starts, xi[:, None, :], None if u is None else u[:, None], dtype="float32", **kw)
    if adj:       # compare proposals only where both accepted
        ok = a64[0]
        tf_err = np.abs(tf[:, 0][ok] - t64[ok]).max(axis=1) / np.maximum(1.0, np.abs(t64[ok]).max(axis=1))
    else:
        tf_err = np.abs(tf[:, 0] - t64).max(axis=1) / np.maximum(1.0, np.abs(t64).max(axis=1))
    rec = {"first_step_over_1e-5": first(1e-5), "first_step_over_1e-4": first(1e-4), "first_step_over_1e-2": first(1e-2),
           "max_rel_first_200": float(rel[:200].max()), "max_rel_all": float(rel.max()),
           "teacher_forced_median": float(np.median(tf_err)), "teacher_forced_max": float(tf_err.max())}
    if adj:
        diff = np.nonzero(a64[0] != a32[0])[0]
        rec["first_differing_decision"] = int(diff[0]) if len(diff) else None
        rec["accepted_fp64"], rec["accepted_fp32"] = int(a64[0].sum()), int(a32[0].sum())
    out[f"{which}.{alg}"] = rec
    print(which, alg, rec, flush=True)
json.dump(out, open("gpurun_out/fp32_divergence.json", "w"), indent=1)
