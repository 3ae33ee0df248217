/* lmc_b200.h -- C ABI of the B200-native Langevin sampler hot path.
 *
 * The reference (192459/lmc-atomi) has no FFI: its boundary is the Python call
 * surface of the sampler classes and of algs.py.  Each entry point below replaces
 * the per-iteration Python loop named beside it; the Python side binds them with
 * ctypes (lmc_atomi_b200/_lib.py, stub shown in INTEGRATION.md).
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error; lmc_last_error() then
 *     returns a thread-local description;
 *   - every pointer named d_* is a DEVICE pointer owned by the caller (e.g. a
 *     torch tensor's data_ptr()); the library never allocates or frees user-visible
 *     memory, scratch comes from the caller-supplied workspace;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it
 *     unless stated otherwise;
 *   - dtype: LMC_F32 or LMC_F64 selects the arithmetic type of the state, noise
 *     and trajectory buffers.
 */
#ifndef LMC_B200_H
#define LMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LMC_MAX_COMP 8
#define LMC_F32 0
#define LMC_F64 1

const char* lmc_last_error(void);
int lmc_version(void);
/* number of SMs / name of the current device; <0 when no CUDA device */
int lmc_device_sm_count(void);

/* ------------------------------------------------------------------ mixtures */
enum lmc_target_kind {
  LMC_TARGET_GAUSS = 0,    /* lmc.py:39-78        Gaussian mixture                    */
  LMC_TARGET_LAPLACE = 1,  /* lmc_laplace.py:43-95 Moreau-smoothed Laplacian mixture  */
  LMC_TARGET_GAUSS_L1 = 2  /* prox_lmc.py:42-95    Gaussian mixture x Laplacian prior */
};

enum lmc_alg {
  LMC_ULA = 0,     /* lmc.py:94-104, lmc_laplace.py:110-123 */
  LMC_MALA = 1,    /* lmc.py:108-131, lmc_laplace.py:127-151 */
  LMC_PULA = 2,    /* lmc.py:135-148, lmc_laplace.py:155-170 */
  LMC_IHPULA = 3,  /* lmc.py:152-169, lmc_laplace.py:174-193 */
  LMC_MLA = 4,     /* lmc.py:173-190, lmc_laplace.py:197-216 */
  LMC_PGLD = 5,    /* prox_lmc.py:99-110  */
  LMC_MYULA = 6,   /* prox_lmc.py:114-130 */
  LMC_MYMALA = 7,  /* prox_lmc.py:134-158 */
  LMC_PPULA = 8,   /* prox_lmc.py:162-188 */
  LMC_FBULA = 9,   /* prox_lmc.py:192-208 */
  LMC_LBMUMLA = 10 /* prox_lmc.py:212-255 */
};

/* Target description.  The host fills the per-component constants with the same
 * numpy calls the reference uses (np.linalg.inv / det), so they are bit-identical. */
typedef struct lmc_mix_target {
  int32_t kind;                    /* enum lmc_target_kind */
  int32_t n;                       /* components, 1..LMC_MAX_COMP */
  double mu[LMC_MAX_COMP][2];      /* component locations */
  double prec[LMC_MAX_COMP][4];    /* Sigma_i^-1 row-major (GAUSS, GAUSS_L1) */
  double norm[LMC_MAX_COMP];       /* sqrt((2 pi)^2 |det Sigma_i|) (GAUSS, GAUSS_L1) */
  double omega[LMC_MAX_COMP];      /* mixture weights */
  double alpha[LMC_MAX_COMP];      /* Laplacian scales (LAPLACE) */
  double lamda;                    /* Moreau smoothing (LAPLACE, GAUSS_L1) */
  double prior_alpha;              /* l1 prior scale (GAUSS_L1) */
  double prior_mu[2];              /* l1 prior location, used by the MH ratio only */
} lmc_mix_target;

typedef struct lmc_mix_step {
  int32_t alg;        /* enum lmc_alg */
  int32_t t;          /* PPULA inner dual iterations (prox_lmc.py:165) */
  double gamma;       /* step size */
  double M[4];        /* PULA/PPULA preconditioner, row-major */
  double sqrtM[4];    /* scipy.linalg.sqrtm(M), host-computed */
  double Q[4];        /* PPULA metric */
  double Qinv[4];     /* np.linalg.inv(Q) */
  double eta;         /* PPULA dual step rho - eps (prox_lmc.py:166-168) */
  double beta[2];     /* MLA / LBMUMLA mirror parameter */
  double sigma[2];    /* LBMUMLA: the vector the reference passes in the beta slot
                         of grad_BM_env (prox_lmc.py:251) */
  double hess_shift;  /* IHPULA eigen-shift: 0.05 (lmc.py:162) / 0.02 (lmc_laplace.py:186) */
} lmc_mix_step;

/* Buffers of one run of C chains x K steps.  NULL = not requested. */
typedef struct lmc_mix_io {
  void* d_theta;             /* (C,2) in/out: start state (ignored when init_from_rng) -> final state */
  const void* d_noise;       /* (C,K,2) validation mode: host-supplied N(0,1) draws; NULL = Philox */
  const void* d_uniforms;    /* (C,K) validation mode, adjusted samplers */
  void* d_traj;              /* (C,K,2) state after every step */
  uint8_t* d_accept;         /* (C,K) accept flags (adjusted samplers) */
  uint32_t* d_naccept;       /* (C,) accepted proposals per chain */
  double* d_moments;         /* (6,) count, S(t0), S(t1), S(t0^2), S(t0 t1), S(t1^2) over emitted samples */
  unsigned long long* d_hist;/* (hist_bins,hist_bins) counts, [x-bin][y-bin], ADDED to */
  void* d_workspace;         /* >= lmc_mix_workspace_bytes(C) */
  size_t workspace_bytes;
  int32_t init_from_rng;     /* 1: theta0 ~ N(0,I) from the chain's Philox stream */
  int32_t hist_bins;         /* e.g. 100 (lmc.py:295) */
  double hist_lo, hist_hi;   /* square range, e.g. [-5,5] (lmc.py:199-200) */
  int64_t hist_every;        /* >0: bin every hist_every-th emitted step; 0: final state only */
  int64_t burnin;            /* steps excluded from moments/histogram */
} lmc_mix_io;

size_t lmc_mix_workspace_bytes(int64_t C);

/* Runs C independent chains for K steps in ONE persistent launch (+ a tiny
 * fixed-order reduction of the moment partials).  Chain c uses Philox4x32-10 with
 * key = seed and counter words (step, stream, c + chain_offset), so results do not
 * depend on how chains are sharded over GPUs. */
int lmc_mix_run(const lmc_mix_target* target, const lmc_mix_step* step, int dtype,
                int64_t C, int64_t K, uint64_t seed, uint64_t chain_offset,
                const lmc_mix_io* io, void* stream);

/* Test hook: the N(0,1) pairs (and uniforms) the production path draws:
 * d_normals (C,K,2), d_uniforms (C,K) or NULL, d_theta0 (C,2) or NULL. */
int lmc_mix_philox_draws(int dtype, int adjusted, int64_t C, int64_t K, uint64_t seed,
                         uint64_t chain_offset, void* d_normals, void* d_uniforms,
                         void* d_theta0, void* stream);

/* ------------------------------------------------------------------ image chains */
#define LMC_IMG_MAX_TAPS 16

/* Deconvolution problem: data term sigma_l2/2 ||H x - b||^2 with H a separable 2-D
 * filter, zero boundary -- pylops.signalprocessing.Convolve2D((ny,nx), h, offset)
 * (prox_lmc_deconv.py:55-69) wrapped in pyproximal.L2 (prox_lmc_deconv.py:101-103).
 * Per axis the forward operator is out[i] = sum_t w[t] * in[i + lo + t] with
 * -(ntap-1) <= lo <= 0; the adjoint is derived from the same taps. */
typedef struct lmc_img_problem {
  int32_t ny, nx;
  int32_t ntap_y, ntap_x;
  int32_t lo_y, lo_x;
  double wy[LMC_IMG_MAX_TAPS], wx[LMC_IMG_MAX_TAPS];
  double sigma_l2;          /* pyproximal.L2 sigma = 1 / sigma_noise^2 */
  const void* d_b;          /* observation, (P,) shared by all chains or (C,P) */
  int64_t b_chain_stride;   /* 0 = shared, P = per chain */
} lmc_img_problem;

/* pyproximal.TV.prox semantics, PINNED against the reference's own result figures (oracle/digitise_figs.py,
 * tests/golden/deconv_curves.npz; evidence in DESIGN.md section 2):
 *   - the dual FISTA loop is `iter = 0; while iter <= niter`, i.e. at most niter + 1 passes, each starting with
 *     sol = x - gamma div(rr, ss) and the exit test, so at most niter + 1 evaluations of sol and niter dual updates
 *     ... + 1: LMC_TV_EXTRA_PASS;
 *   - the objective of the exit test is 1/2 ||x - sol||^2 + gamma * TV.__call__(sol) and TV.__call__ carries the
 *     operator's sigma, so the weight of sum |grad sol| is gamma * sigma with gamma = sigma * tau. */
#ifndef LMC_TV_EXTRA_PASS
#define LMC_TV_EXTRA_PASS 1
#endif

/* MoreauYosidaUnadjustedLangevin(proxf=L2, proxg=TV(dims, tv_sigma, tv_niter, tv_rtol),
 * tau, gamma, epsg)  (algs.py:477-587) */
typedef struct lmc_img_myula_params {
  double tau, gamma, epsg;
  double tv_sigma, tv_rtol;
  int32_t tv_niter;
} lmc_img_myula_params;

/* UnadjustedLangevinPrimalDual(proxf=L2(niter=lsqr_iter_lim, warm), proxg=L21|L1(sigma=g_sigma),
 * A=Gradient, tau, mu, theta, gfirst)  (algs.py:295-474) */
typedef struct lmc_img_ulpda_params {
  double tau, mu, theta;
  double g_sigma;           /* pyproximal.L21 / L1 sigma */
  int32_t g_isotropic;      /* 1 = L21(ndim=2) (prox_lmc_deconv.py:116), 0 = L1 (:119) */
  int32_t gfirst;           /* algs.py:435-448 */
  int32_t lsqr_iter_lim;    /* pyproximal.L2 niter (50) */
  int32_t warm;             /* pyproximal.L2 warm start */
  double lsqr_atol, lsqr_btol, lsqr_conlim;   /* scipy lsqr defaults 1e-6, 1e-6, 1e8 */
} lmc_img_ulpda_params;

/* Non-convex data term of the reference, algs.L2_ncvx_tv (algs.py:22-291): f(x) = sigma/2 ||Hx - b||^2 - lamda *
 * env_gamma(TV)(x), evaluated INSIDE the sampler kernels for all N iterations of a launch.
 * kind 1 = MC-TV (Op2 = Gradient): grad_moreau = G^T(w(Gx) Gx), isotropic w = min(1/gamma, 1/|Gx|_2) (|Gx| = 0 read as
 *          1e-9), anisotropic w Gx = (Gx - soft(Gx, gamma)) / gamma   (algs.py:211-217, 271-277);
 * kind 2 = ME-TV (Op2 = None, isotropic): grad_moreau = (x - TV(dims, 1., niter, rtol).prox(x, gamma)) / gamma
 *          (algs.py:169, 221-223, 280-282): the kernel runs the inner TV prox with its own TV machinery.
 * MYULA: grad f(x) = sigma H^T(Hx - b) - lamda grad_moreau(x)  (algs.py:283-291).
 * ULPDA: the argument of the L2 prox gets the in-place pre-step of L2_ncvx_tv.prox, arg += tau lamda grad_moreau(arg)
 *        with the reference's grouping of the scalars (algs.py:208-223); both gfirst orders. */
typedef struct lmc_img_ncvx {
  int32_t kind;             /* 0 none, 1 MC-TV, 2 ME-TV */
  int32_t isotropic;
  double lamda, gamma;
  int32_t niter;            /* ME-TV: inner TV iterations (the driver's niter_l2 = 50) */
  double rtol;              /* ME-TV: inner TV rtol (1e-4) */
} lmc_img_ncvx;

/* Buffers of one run of C image chains x N sampler iterations.  NULL = not requested. */
typedef struct lmc_img_io {
  void* d_x;                /* (C,P) in/out sampler state */
  void* d_y;                /* (C,2,P) in/out ULPDA dual state (ignored by MYULA) */
  void* d_xhat;             /* (C,P) in/out ULPDA extrapolated point */
  void* d_warm;             /* (C,P) in/out LSQR warm start (pyproximal.L2.x0) */
  const void* d_z;          /* (P,) optional ULPDA linear term z (algs.py:437-446), shared */
  const void* d_noise;      /* (C,N,P) validation mode: host-supplied N(0,1); NULL = Philox */
  void* d_samples;          /* (C,N,P) every sample (the reference's return value) */
  void* d_ysamples;         /* (C,N,2,P) ULPDA returny */
  void* d_mean;             /* (C,P) running posterior mean (Welford), updated in place */
  void* d_m2;               /* (C,P) running sum of squared deviations */
  unsigned long long* d_counters; /* (8,) added to: [0] TV sweeps, [1] LSQR iterations, [2] LSQR solves */
  void* d_workspace;
  size_t workspace_bytes;
  int64_t burnin;           /* global iteration index of the first sample folded into mean/m2 */
  int64_t stat_count0;      /* samples already folded into mean/m2 BEFORE this call (chunked runs): the sample of
                               global iteration it >= max(burnin, iter_offset) is folded in with count
                               stat_count0 + (it - max(burnin, iter_offset)) + 1 */
  uint64_t iter_offset;     /* global index of this call's first iteration (Philox counter) */
  /* Non-convex data terms (algs.L2_ncvx_tv, algs.py:22-291): the Moreau-envelope term is evaluated by the
   * unit entry points below between launches, so both need N == 1 (and gfirst == 0 for ULPDA).
   * Streaming / band kernels only (not lmc_img_myula_run_onchip). */
  const void* d_fgrad_extra; /* (C,P) MYULA: grad f(x) = sigma H^T(Hx - b) + fgrad_scale * this; read every iteration of the
                                launch: a per-iteration term (the Moreau gradient, algs.py:283-291, fgrad_scale = -lamda)
                                needs N == 1, a constant one (alpha q of L2 / L2_ncvx_tv, algs.py:288-289) any N */
  double fgrad_scale;
  const void* d_prox_arg;    /* (C,P) ULPDA: replaces x - tau (G^T y + z) as the argument of the L2 prox (algs.py:224-256) */
  const lmc_img_ncvx* ncvx;  /* HOST pointer or NULL: non-convex term evaluated in-kernel, any N (streaming / band kernels
                                of both samplers; ULPDA needs lmc_img_ulpda_workspace_bytes_ncvx); excludes
                                d_fgrad_extra / d_prox_arg */
} lmc_img_io;

size_t lmc_img_myula_workspace_bytes(int dtype, int32_t ny, int32_t nx, int64_t C);
size_t lmc_img_ulpda_workspace_bytes(int dtype, int32_t ny, int32_t nx, int64_t C);
/* ... with a non-convex term evaluated in-kernel (io->ncvx, kind as in lmc_img_ncvx): adds the prox-argument buffers
 * and, for ME-TV, the state of the inner TV pipeline */
size_t lmc_img_ulpda_workspace_bytes_ncvx(int dtype, int32_t ny, int32_t nx, int64_t C, int32_t kind);

/* C chains x N iterations in ONE persistent cooperative launch.  Chain c draws its
 * noise from Philox4x32-10 keyed by seed with counter (pixel-quad, iteration,
 * c + chain_offset), so results do not depend on the sharding over GPUs. */
int lmc_img_myula_run(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype, int64_t C,
                      int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io, void* stream);
int lmc_img_ulpda_run(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, int dtype, int64_t C,
                      int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io, void* stream);

/* ULPDA, row-band formulation (csrc/img_ulpda_band.inl): full-width bands of 16 rows staged in shared
 * memory, fused two-pass H^T H with compile-time taps, chains advanced in groups of `group` chains
 * (0 = all chains in lockstep; a smaller group keeps the LSQR vectors L2 resident), each group
 * running all N iterations before the next.  Same arguments, workspace, noise stream and results (same arithmetic per pixel) as
 * lmc_img_ulpda_run; requires 5/6/7 blur taps on both axes (lmc_img_ulpda_band_supported() != 0). */
int lmc_img_ulpda_band_supported(const lmc_img_problem* pb, int dtype);
int lmc_img_ulpda_run_band(const lmc_img_problem* pb, const lmc_img_ulpda_params* pr, int dtype, int64_t C,
                           int64_t N, uint64_t seed, uint64_t chain_offset, int64_t group, const lmc_img_io* io,
                           void* stream);

/* MYULA, row-band formulation (csrc/img_myula_band.inl): bands of 16 full-width rows, fused two-pass H^T H with
 * compile-time taps, TV sweeps on 8-row sub-bands, two columns per thread, work items drawn from a global queue.
 * Same arguments, workspace and results as lmc_img_myula_run; requires 5/6/7 blur taps on both axes, an even
 * width and buffers aligned to two elements (lmc_img_myula_band_supported() != 0). */
int lmc_img_myula_band_supported(const lmc_img_problem* pb, int dtype);
int lmc_img_myula_run_band(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype, int64_t C,
                           int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io, void* stream);

/* MYULA with the whole chain state ON CHIP (csrc/img_myula_onchip.cu): one thread-block
 * cluster of 8 CTAs per chain, x and the TV dual field in (distributed) shared memory for
 * all N iterations.  Same arguments, noise stream and results (to fp32 rounding) as
 * lmc_img_myula_run; requires LMC_F32, ny <= 256, nx <= 256, blur taps <= 7 per axis
 * (lmc_img_myula_onchip_supported() != 0); needs no workspace but io->d_counters.
 * lmc_img_myula_onchip_clusters(): co-resident clusters on the current device (<0 on error). */
int lmc_img_myula_onchip_supported(const lmc_img_problem* pb, int dtype);
int lmc_img_myula_onchip_clusters(void);
int lmc_img_myula_run_onchip(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype, int64_t C,
                             int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io, void* stream);

/* Unit entry points (operators of SURVEY 8a rows I1-I5), batched over C images.
 * d_in/d_out are (C,P) unless stated; workspace as for the sampler of the same family. */
int lmc_img_blur(const lmc_img_problem* pb, int dtype, int adjoint, int64_t C, const void* d_in, void* d_out,
                 void* stream);                                   /* Convolve2D matvec / rmatvec */
int lmc_img_grad(int dtype, int32_t ny, int32_t nx, int adjoint, int64_t C, const void* d_in, void* d_out,
                 void* stream);                                   /* Gradient: (C,P)->(C,2,P) / adjoint */
int lmc_img_tv_prox(int dtype, int32_t ny, int32_t nx, double gamma, double sigma, int32_t niter, double rtol,
                    int64_t C, const void* d_in, void* d_out, unsigned long long* d_counters, void* d_workspace,
                    size_t workspace_bytes, void* stream);        /* pyproximal.TV(sigma).prox with sigma*tau = gamma */
int lmc_img_proxdual(int dtype, int isotropic, double sigma, int64_t n_pix, void* d_v, void* stream);
                                                                  /* L21/L1 proxdual, in place on (2,n_pix) */
int lmc_img_l2_prox(const lmc_img_problem* pb, int dtype, double tau, int32_t iter_lim, double atol, double btol,
                    double conlim, int64_t C, const void* d_in, void* d_warm, void* d_out,
                    unsigned long long* d_counters, void* d_workspace, size_t workspace_bytes,
                    void* stream);                                /* pyproximal.L2.prox (LSQR, warm start) */
int lmc_img_noise(int dtype, int32_t ny, int32_t nx, int64_t C, int64_t N, uint64_t seed, uint64_t chain_offset,
                  uint64_t iter_offset, void* d_out, void* stream); /* test hook: the Philox normals (C,N,P) */

/* Pieces of algs.L2_ncvx_tv, batched over C images, d_in != d_out:
 * lmc_img_mctv_grad   MC-TV (Op2 = Gradient): m = G^T(w(Gx) Gx), isotropic: w = min(1/gamma, 1/|Gx|_2) per pixel
 *                     (|Gx| = 0 read as 1e-9, algs.py:211-213,273-275); anisotropic: w Gx = Gx - soft(Gx, gamma)
 *                     (algs.py:215-217,277).  out = base * x + scale * (m / post_div).
 * lmc_img_metv_combine ME-TV (Op2 = None): d = x - p with p = TV_1.prox(x, gamma) from lmc_img_tv_prox;
 *                     out = base * x + (den != 0 ? d / den : scale * d)   (algs.py:221-223, 280-282).
 * lmc_img_pd_arg      out = x - tau * (G^T y + z): the primal argument of ULPDA (algs.py:436-446); d_z may be NULL. */
int lmc_img_mctv_grad(int dtype, int32_t ny, int32_t nx, double gamma, int isotropic, double base, double scale,
                      double post_div, int64_t C, const void* d_x, void* d_out, void* stream);
int lmc_img_metv_combine(int dtype, int64_t n, double base, double scale, double den, const void* d_x,
                         const void* d_p, void* d_out, void* stream);
int lmc_img_pd_arg(int dtype, int32_t ny, int32_t nx, double tau, int64_t C, const void* d_x, const void* d_y,
                   const void* d_z, void* d_out, void* stream);

/* MAP branch of the driver (prox_lmc_deconv.py:138-445): pyproximal.optimization.primaldual.AdaptivePrimalDual with
 * A = Gradient.  One image (C = 1).  The host loop (lmc_atomi_b200/primaldual.py) calls lmc_img_pd_arg and
 * lmc_img_l2_prox for the primal step, then
 * lmc_img_pdhg_dual:         Ax <- G x (in place over the previous Ax, (2,P)); y <- proxdual_{L21|L1}(y + mu (2 Ax -
 *                            Ax_old), mu) in place ((2,P)); *d_resid = ||(y_old - y)/mu - (Ax_old - Ax)||_2
 * lmc_img_pdhg_primal_resid: ATy <- G^T y (in place over the previous ATy, (P,));
 *                            *d_resid = ||(x_old - x)/tau - (ATy_old - ATy)||_2
 * (fp64 sums in a fixed order; d_resid is a device double). */
size_t lmc_img_pdhg_workspace_bytes(void);
int lmc_img_pdhg_dual(int dtype, int32_t ny, int32_t nx, double mu, double g_sigma, int isotropic, const void* d_x,
                      void* d_Ax, void* d_y, double* d_resid, void* d_workspace, size_t workspace_bytes, void* stream);
int lmc_img_pdhg_primal_resid(int dtype, int32_t ny, int32_t nx, double tau, const void* d_xold, const void* d_x,
                              const void* d_y, void* d_ATy, double* d_resid, void* d_workspace,
                              size_t workspace_bytes, void* stream);

/* Per-sample metrics on the device, replacing the reference's per-iteration host callback
 * (prox_lmc_deconv.py:35-36,128-133): for each of the M images of d_x (M,P) -- e.g. the (C,N,P) samples of a
 * run, images_per_chain = N -- d_out (M,4) doubles receives
 *   [0] ||H x - b||^2, [1] sum_pixels |G x| (isotropic l2 / anisotropic l1), [2] ||x - x*||^2, [3] ||x*||^2
 * (x* = d_xtrue (P,), may be NULL), from which cost = sigma_l2/2 [0] + g_sigma [1], err = sqrt([2]),
 * SNR = 10 log10([3]/[2]), MSE = [2]/P, PSNR = 10 log10(255^2 / MSE).  fp64 sums, fixed order. */
size_t lmc_img_metrics_workspace_bytes(int64_t M);
int lmc_img_sample_metrics(const lmc_img_problem* pb, int dtype, int isotropic, int64_t M, int64_t images_per_chain,
                           const void* d_x, const void* d_xtrue, double* d_out, void* d_workspace,
                           size_t workspace_bytes, void* stream);

/* lmc_img_sample_metrics without a workspace argument and without any allocation (no cudaMalloc / cudaFree or
 * allocating library call): the partial sums live in a static device array inside the library, 512 KiB per GPU, held
 * until exit.  NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * allocate (CUDA graph capture, serving loops).  Same arguments, kernels and results (bit for bit) as
 * lmc_img_sample_metrics; M > LMC_METRICS_STATIC_MAX_IMAGES is rejected (-1, lmc_last_error()).  All calls on a
 * device share the one array: calls that overlap on the GPU (different streams, no dependency between them) race --
 * serialise them.  Nothing has to be cleared between calls (every partial is written before it is read). */
#define LMC_METRICS_STATIC_MAX_IMAGES 1024
int lmc_img_sample_metrics_static(const lmc_img_problem* pb, int dtype, int isotropic, int64_t M,
                                  int64_t images_per_chain, const void* d_x, const void* d_xtrue, double* d_out,
                                  void* stream);

/* 2-Wasserstein distance between two uniformly weighted point sets on the device -- the evaluation the
 * reference's mixture drivers run on the CPU with POT, ot.emd2(a, b, ot.dist(Z1, Z[:k+1]))**.5
 * (lmc.py:386-416, lmc_laplace.py:426-451).  d_x (n,dim), d_y (m,dim) row-major fp64, dim <= 4.  Log-domain
 * Sinkhorn with epsilon scaling (eps: eps_start -> eps_final by factors eps_scale; eps_start <= 0 = squared
 * diagonal of the joint bounding box), at most max_iters iterations per eps, stopping when the potentials move
 * by less than tol*eps (checked every check_every iterations); omega in [1,2) over-relaxes both updates.  out->cost is the primal value <C, P_eps>,
 * out->w2 its square root: an approximation of the exact EMD that tightens as eps_final -> 0.
 * warm_start != 0 keeps the x-side potentials (first n doubles of the workspace) of the previous call.
 * Unlike the sampler entry points this call SYNCHRONISES the stream (it is an evaluation, with a host loop). */
typedef struct lmc_w2_result {
  double cost, w2, eps, err, mass;
  int64_t iters;
  int32_t converged;
} lmc_w2_result;
size_t lmc_w2_workspace_bytes(int64_t n, int64_t m);
int lmc_w2_sinkhorn(const double* d_x, int64_t n, const double* d_y, int64_t m, int32_t dim, double eps_final,
                    double eps_start, double eps_scale, double tol, double omega, int32_t max_iters,
                    int32_t check_every, int32_t warm_start, void* d_workspace, size_t workspace_bytes, lmc_w2_result* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LMC_B200_H */
