"""ctypes binding of the C-ABI declared in include/lmc_b200.h.

The shared library is built in-tree by ``lmc_atomi_b200/build.py`` (nvcc, sm_100a).
There is deliberately no fallback: if the library is missing, cannot be loaded, or a
call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LMC_B200_LIB", os.path.join(HERE, "_lmc_b200.so"))

MAX_COMP = 8
F32, F64 = 0, 1

TARGET_GAUSS, TARGET_LAPLACE, TARGET_GAUSS_L1 = 0, 1, 2
ALGS = {"ula": 0, "mala": 1, "pula": 2, "ihpula": 3, "mla": 4, "pgld": 5, "myula": 6,
        "mymala": 7, "ppula": 8, "fbula": 9, "lbmumla": 10}


class LmcError(RuntimeError):
    pass


class MixTarget(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n", C.c_int32),
                ("mu", (C.c_double * 2) * MAX_COMP),
                ("prec", (C.c_double * 4) * MAX_COMP),
                ("norm", C.c_double * MAX_COMP),
                ("omega", C.c_double * MAX_COMP),
                ("alpha", C.c_double * MAX_COMP),
                ("lamda", C.c_double),
                ("prior_alpha", C.c_double),
                ("prior_mu", C.c_double * 2)]


class MixStep(C.Structure):
    _fields_ = [("alg", C.c_int32), ("t", C.c_int32), ("gamma", C.c_double),
                ("M", C.c_double * 4), ("sqrtM", C.c_double * 4),
                ("Q", C.c_double * 4), ("Qinv", C.c_double * 4),
                ("eta", C.c_double),
                ("beta", C.c_double * 2), ("sigma", C.c_double * 2),
                ("hess_shift", C.c_double)]


class MixIo(C.Structure):
    _fields_ = [("d_theta", C.c_void_p), ("d_noise", C.c_void_p), ("d_uniforms", C.c_void_p),
                ("d_traj", C.c_void_p), ("d_accept", C.c_void_p), ("d_naccept", C.c_void_p),
                ("d_moments", C.c_void_p), ("d_hist", C.c_void_p),
                ("d_workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
                ("init_from_rng", C.c_int32), ("hist_bins", C.c_int32),
                ("hist_lo", C.c_double), ("hist_hi", C.c_double),
                ("hist_every", C.c_int64), ("burnin", C.c_int64)]


class W2Result(C.Structure):
    _fields_ = [("cost", C.c_double), ("w2", C.c_double), ("eps", C.c_double), ("err", C.c_double),
                ("mass", C.c_double), ("iters", C.c_int64), ("converged", C.c_int32)]


IMG_MAX_TAPS = 16


class ImgProblem(C.Structure):
    _fields_ = [("ny", C.c_int32), ("nx", C.c_int32), ("ntap_y", C.c_int32), ("ntap_x", C.c_int32),
                ("lo_y", C.c_int32), ("lo_x", C.c_int32),
                ("wy", C.c_double * IMG_MAX_TAPS), ("wx", C.c_double * IMG_MAX_TAPS),
                ("sigma_l2", C.c_double), ("d_b", C.c_void_p), ("b_chain_stride", C.c_int64)]


class ImgMyulaParams(C.Structure):
    _fields_ = [("tau", C.c_double), ("gamma", C.c_double), ("epsg", C.c_double),
                ("tv_sigma", C.c_double), ("tv_rtol", C.c_double), ("tv_niter", C.c_int32)]


class ImgUlpdaParams(C.Structure):
    _fields_ = [("tau", C.c_double), ("mu", C.c_double), ("theta", C.c_double), ("g_sigma", C.c_double),
                ("g_isotropic", C.c_int32), ("gfirst", C.c_int32), ("lsqr_iter_lim", C.c_int32),
                ("warm", C.c_int32), ("lsqr_atol", C.c_double), ("lsqr_btol", C.c_double),
                ("lsqr_conlim", C.c_double)]


class ImgNcvx(C.Structure):
    _fields_ = [("kind", C.c_int32), ("isotropic", C.c_int32), ("lamda", C.c_double), ("gamma", C.c_double),
                ("niter", C.c_int32), ("rtol", C.c_double)]


class ImgIo(C.Structure):
    _fields_ = [("d_x", C.c_void_p), ("d_y", C.c_void_p), ("d_xhat", C.c_void_p), ("d_warm", C.c_void_p),
                ("d_z", C.c_void_p), ("d_noise", C.c_void_p), ("d_samples", C.c_void_p),
                ("d_ysamples", C.c_void_p), ("d_mean", C.c_void_p), ("d_m2", C.c_void_p),
                ("d_counters", C.c_void_p), ("d_workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
                ("burnin", C.c_int64), ("stat_count0", C.c_int64), ("iter_offset", C.c_uint64),
                ("d_fgrad_extra", C.c_void_p), ("fgrad_scale", C.c_double), ("d_prox_arg", C.c_void_p),
                ("ncvx", C.POINTER(ImgNcvx))]


# name -> (restype, argtypes); must list every symbol include/lmc_b200.h declares
SIGNATURES = {
    "lmc_last_error": (C.c_char_p, []),
    "lmc_version": (C.c_int, []),
    "lmc_device_sm_count": (C.c_int, []),
    "lmc_mix_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "lmc_mix_run": (C.c_int, [C.POINTER(MixTarget), C.POINTER(MixStep), C.c_int, C.c_int64,
                              C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(MixIo), C.c_void_p]),
    "lmc_mix_philox_draws": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_uint64,
                                       C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lmc_img_myula_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int32, C.c_int32, C.c_int64]),
    "lmc_img_ulpda_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int32, C.c_int32, C.c_int64]),
    "lmc_img_ulpda_workspace_bytes_ncvx": (C.c_size_t, [C.c_int, C.c_int32, C.c_int32, C.c_int64, C.c_int32]),
    "lmc_img_myula_run": (C.c_int, [C.POINTER(ImgProblem), C.POINTER(ImgMyulaParams), C.c_int, C.c_int64,
                                    C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(ImgIo), C.c_void_p]),
    "lmc_img_myula_onchip_supported": (C.c_int, [C.POINTER(ImgProblem), C.c_int]),
    "lmc_img_myula_band_supported": (C.c_int, [C.POINTER(ImgProblem), C.c_int]),
    "lmc_img_myula_run_band": (C.c_int, [C.POINTER(ImgProblem), C.POINTER(ImgMyulaParams), C.c_int, C.c_int64,
                                         C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(ImgIo), C.c_void_p]),
    "lmc_img_myula_onchip_clusters": (C.c_int, []),
    "lmc_img_myula_run_onchip": (C.c_int, [C.POINTER(ImgProblem), C.POINTER(ImgMyulaParams), C.c_int, C.c_int64,
                                           C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(ImgIo), C.c_void_p]),
    "lmc_img_ulpda_run": (C.c_int, [C.POINTER(ImgProblem), C.POINTER(ImgUlpdaParams), C.c_int, C.c_int64,
                                    C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(ImgIo), C.c_void_p]),
    "lmc_img_ulpda_band_supported": (C.c_int, [C.POINTER(ImgProblem), C.c_int]),
    "lmc_img_ulpda_run_band": (C.c_int, [C.POINTER(ImgProblem), C.POINTER(ImgUlpdaParams), C.c_int, C.c_int64,
                                         C.c_int64, C.c_uint64, C.c_uint64, C.c_int64, C.POINTER(ImgIo), C.c_void_p]),
    "lmc_img_blur": (C.c_int, [C.POINTER(ImgProblem), C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p,
                               C.c_void_p]),
    "lmc_img_grad": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_int, C.c_int64, C.c_void_p, C.c_void_p,
                               C.c_void_p]),
    "lmc_img_tv_prox": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int32, C.c_double, C.c_int64,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "lmc_img_pdhg_workspace_bytes": (C.c_size_t, []),
    "lmc_img_pdhg_dual": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "lmc_img_pdhg_primal_resid": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_double, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                            C.c_void_p]),
    "lmc_img_proxdual": (C.c_int, [C.c_int, C.c_int, C.c_double, C.c_int64, C.c_void_p, C.c_void_p]),
    "lmc_img_l2_prox": (C.c_int, [C.POINTER(ImgProblem), C.c_int, C.c_double, C.c_int32, C.c_double, C.c_double,
                                  C.c_double, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_size_t, C.c_void_p]),
    "lmc_img_noise": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_uint64, C.c_uint64,
                                C.c_uint64, C.c_void_p, C.c_void_p]),
    "lmc_img_mctv_grad": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_double, C.c_int, C.c_double, C.c_double,
                                    C.c_double, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lmc_img_metv_combine": (C.c_int, [C.c_int, C.c_int64, C.c_double, C.c_double, C.c_double, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "lmc_img_metrics_workspace_bytes": (C.c_size_t, [C.c_int64]),
    "lmc_img_sample_metrics": (C.c_int, [C.POINTER(ImgProblem), C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "lmc_img_sample_metrics_static": (C.c_int, [C.POINTER(ImgProblem), C.c_int, C.c_int, C.c_int64, C.c_int64,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lmc_img_pd_arg": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_double, C.c_int64, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p]),
    "lmc_w2_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "lmc_w2_sinkhorn": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_double, C.c_double,
                                  C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_size_t,
                                  C.POINTER(W2Result), C.c_void_p]),
}

_lib = None


def load():
    """Load the CUDA library (once).  Raises LmcError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise LmcError(
            f"{LIB_PATH} not found: build it with `python -m lmc_atomi_b200.build` "
            "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().lmc_last_error().decode("utf-8", "replace")
        raise LmcError(f"{what} failed ({rc}): {msg}")


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise LmcError("lmc_atomi_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch
