"""lmc_atomi_b200 -- B200-native sampler hot path of 192459/lmc-atomi.

Modules mirror the reference's files for the hot path only:
  lmc.LangevinMonteCarlo, lmc_laplace.LangevinMonteCarloLaplacian,
  prox_lmc.ProximalLangevinMonteCarlo, algs.{MoreauYosidaUnadjustedLangevin,
  UnadjustedLangevinPrimalDual, L2_ncvx_tv}.
All arithmetic runs in hand-written sm_100a CUDA behind the C ABI of
include/lmc_b200.h; there is no CPU fallback.
"""
__version__ = "0.1.0"
