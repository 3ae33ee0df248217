// Host side of img_myula.cu: workspace layout, argument marshalling, C entry points.
#include "img_host_common.inl"

namespace {

template <typename real>
size_t myula_ws_bytes(int ny, int nx, long long C) {
  const size_t P = (size_t)ny * nx;
  size_t tiles = (size_t)((ny + TH - 1) / TH) * ((nx + TW - 1) / TW);
  const size_t bands = (size_t)(ny + MB_RB - 1) / MB_RB;   // the band kernel keeps one partial per band
  if (bands > tiles) tiles = bands;
  return 256 /* n_active + counters */ + align_up(C * sizeof(TvState)) + align_up(C * tiles * sizeof(double)) +
         align_up((size_t)C * P * sizeof(real)) /* grad */ + align_up((size_t)6 * C * P * sizeof(real)) /* duals */ +
         align_up((size_t)C * P * sizeof(real)) /* Moreau-envelope gradient of an in-kernel non-convex term */;
}

// the band kernel covers the H5 / H6 / H7 blurs of the driver; two image columns per thread on even widths (the
// gradient phase when the tap windows also start on the driver's columns), one column per thread on odd widths
// (the reference's own einstein.png is 667 x 877)
template <typename real>
bool myula_band_ok(const lmc_img_problem* pb) {
  if (pb->ntap_y != pb->ntap_x || pb->ntap_y < 5 || pb->ntap_y > 7) return false;
  return myula_band_smem_bytes<real, 7>(band_pitch(pb->nx)) <= 200 * 1024;
}

// The band kernels are instantiated per tap count in their own translation units (img_myula_band{5,6,7}.cu, compiled
// in parallel): (dtype, GW, TW, NC) are resolved there.  A / B are MyulaArgs<float|double> of this layout.
}  // namespace
int lmc_myula_band_launch_nt5(int dtype, const void* A, const void* B, int gw, int tw, cudaStream_t st);
int lmc_myula_band_launch_nt6(int dtype, const void* A, const void* B, int gw, int tw, cudaStream_t st);
int lmc_myula_band_launch_nt7(int dtype, const void* A, const void* B, int gw, int tw, cudaStream_t st);
namespace {

template <typename real>
int launch_myula_band(int nt, const MyulaArgs<real>& A, const MyulaArgs<real>& B, int gw, int tw, cudaStream_t st) {
  const int dtype = sizeof(real) == 8 ? LMC_F64 : LMC_F32;
  switch (nt) {
    case 5: return lmc_myula_band_launch_nt5(dtype, &A, &B, gw, tw, st);
    case 6: return lmc_myula_band_launch_nt6(dtype, &A, &B, gw, tw, st);
    default: return lmc_myula_band_launch_nt7(dtype, &A, &B, gw, tw, st);
  }
}

template <typename real>
int run_myula(const lmc_img_problem* pb, const lmc_img_myula_params* pr, long long C, long long N, uint64_t seed,
              uint64_t chain_offset, const lmc_img_io* io, int mode, const void* tv_in, void* tv_out,
              cudaStream_t st, bool band = false) {
  MyulaArgs<real> A, B;
  memset(&A, 0, sizeof(A));
  memset(&B, 0, sizeof(B));
  B.mode = -1;
  fill_geom(pb, A.g);
  if (band) {   // one objective partial per band of MB_RB rows
    A.g.tiles_x = 1;
    A.g.tiles_y = (pb->ny + MB_RB - 1) / MB_RB;
    A.g.tiles = A.g.tiles_y;
    A.pitch = band_pitch(pb->nx);
  }
  const size_t need = myula_ws_bytes<real>(pb->ny, pb->nx, C);
  LMC_REQUIRE(io->d_workspace != nullptr && io->workspace_bytes >= need, "image workspace too small (%zu < %zu)",
              io->workspace_bytes, need);
  char* ws = (char*)io->d_workspace;
  A.n_active = (int*)ws;
  A.queue = (int*)(ws + 128);
  unsigned long long* local_counters = (unsigned long long*)(ws + 64);
  ws += 256;
  A.st = (TvState*)ws;
  ws += align_up(C * sizeof(TvState));
  A.partial = (double*)ws;
  ws += align_up((size_t)C * A.g.tiles * sizeof(double));
  A.grad = (real*)ws;
  ws += align_up((size_t)C * A.g.P * sizeof(real));
  A.D = (real*)ws;
  ws += align_up((size_t)6 * C * A.g.P * sizeof(real));
  A.gextra_w = (real*)ws;
  A.counters = io->d_counters ? io->d_counters : local_counters;
  A.C = C;
  A.N = N;
  A.mode = mode;
  A.key = img_key(seed);
  A.chain_offset = chain_offset;
  A.iter_offset = io->iter_offset;
  A.burnin = io->burnin;
  // samples folded in by this call before global iteration `it` = it - max(burnin, iter_offset); the kernels evaluate
  // count = stat_base + (it - burnin) + 1, so stat_base absorbs the part of iter_offset beyond the burn-in
  A.stat_count0 = io->stat_count0 - ((long long)io->iter_offset > io->burnin ? (long long)io->iter_offset - io->burnin : 0);
  A.tv_passes = pr->tv_niter + LMC_TV_EXTRA_PASS;
  A.tv_rtol = pr->tv_rtol;
  if (mode == 0) {
    A.x = (real*)io->d_x;
    A.b = (const real*)pb->d_b;
    A.b_stride = pb->b_chain_stride;
    A.noise = (const real*)io->d_noise;
    A.samples = (real*)io->d_samples;
    A.mean = (real*)io->d_mean;
    A.m2 = (real*)io->d_m2;
    A.gextra = (const real*)io->d_fgrad_extra;
    A.gscale = (real)io->fgrad_scale;
    A.sigma_l2 = (real)pb->sigma_l2;
    A.tau = (real)pr->tau;
    A.a_x = (real)(1 - pr->tau / pr->gamma);
    A.tog = (real)(pr->tau / pr->gamma);
    A.s2t = (real)sqrt(2 * pr->tau);
    A.tv_gamma = (real)(pr->tv_sigma * (pr->epsg * pr->gamma));   // TV.prox(x, epsg*gamma): sigma * tau
    A.tv_objw = (double)A.tv_gamma * pr->tv_sigma;                 // obj = .5|x-sol|^2 + gamma * TV.__call__(sol)
    // non-convex data term evaluated in-kernel (algs.L2_ncvx_tv, algs.py:270-291): grad f = sigma H^T(Hx - b)
    // - lamda * grad_moreau(x), for all N iterations of the launch
    if (io->ncvx != nullptr && io->ncvx->kind != 0) {
      const lmc_img_ncvx& nc = *io->ncvx;
      A.gextra = A.gextra_w;
      A.gscale = (real)(-nc.lamda);
      if (nc.kind == 1) {          // MC-TV: 7-point stencil in phase G
        A.nc.kind = 1;
        A.nc.W.inv_gamma = (real)(1.0 / nc.gamma);
        A.nc.W.gamma = (real)nc.gamma;
        A.nc.W.iso = nc.isotropic;
        A.nc.post_div = (real)(nc.isotropic ? 1.0 : nc.gamma);
      } else {                     // ME-TV: prelude B = TV(dims, 1., niter, rtol).prox(x, gamma), out = (x - prox)/gamma
        B = A;
        B.mode = 2;
        B.nc.kind = 0;
        B.gextra = nullptr;
        B.out = A.gextra_w;
        B.me_den = (real)nc.gamma;
        B.tv_gamma = (real)nc.gamma;                  // sigma (1) * tau (gamma)
        B.tv_objw = (double)B.tv_gamma * 1.0;
        B.tv_passes = nc.niter + LMC_TV_EXTRA_PASS;
        B.tv_rtol = nc.rtol;
        B.counters = local_counters + 4;              // inner sweeps are not the sampler's TV sweeps
        B.noise = nullptr; B.samples = nullptr; B.mean = nullptr; B.m2 = nullptr;
      }
    }
  } else {
    A.x = (real*)const_cast<void*>(tv_in);
    A.out = (real*)tv_out;
    A.tv_gamma = (real)pr->tv_sigma;   // lmc_img_tv_prox passes gamma = sigma * tau here ...
    A.tv_objw = (double)A.tv_gamma * pr->epsg;   // ... and the operator's sigma here
  }
  LMC_CUDA_OK(cudaMemsetAsync(io->d_workspace, 0, 256, st));
  if (band) {
    if (pb->nx & 1)   // odd width: rows are not pair-aligned, one column per thread in every phase
      return launch_myula_band<real>(pb->ntap_y, A, B, 1, 1, st);
    const void* ptrs[] = {A.x, A.b, A.grad, A.D};
    for (const void* p : ptrs)
      LMC_REQUIRE(reinterpret_cast<size_t>(p) % (2 * sizeof(real)) == 0,
                  "MYULA band kernel: buffers must be aligned to two elements");
    // two columns per thread in the gradient phase when the blur window is the driver's (BandWin)
    const int lo = pb->ntap_y == 7 ? BandWin<7>::LO : BandWin<5>::LO;
    return launch_myula_band<real>(pb->ntap_y, A, B, pb->lo_x == lo ? 2 : 1, 2, st);
  }
  return launch_persistent(img_myula_kernel<real>, A, C * A.g.tiles, (size_t)3 * RS * RS * sizeof(real), st, &B);
}

}  // namespace

extern "C" size_t lmc_img_myula_workspace_bytes(int dtype, int32_t ny, int32_t nx, int64_t C) {
  if (ny < 1 || nx < 1 || C < 1) return 256;
  return dtype == LMC_F64 ? myula_ws_bytes<double>(ny, nx, C) : myula_ws_bytes<float>(ny, nx, C);
}

extern "C" int lmc_img_myula_run(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype, int64_t C,
                                 int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io,
                                 void* stream) {
  if (check_problem(pb, "lmc_img_myula_run") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_myula_run: null descriptor");
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_myula_run: dtype must be LMC_F32 or LMC_F64");
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_myula_run: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->gamma > 0, "lmc_img_myula_run: tau and gamma must be positive");
  LMC_REQUIRE(pr->tv_niter >= 0 && pr->tv_sigma * pr->epsg * pr->gamma > 0,
              "lmc_img_myula_run: TV needs niter >= 0 and sigma*epsg*gamma > 0");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x != nullptr && pb->d_b != nullptr, "lmc_img_myula_run: d_x and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_myula_run: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(io->ncvx == nullptr || io->ncvx->kind == 0 || io->d_fgrad_extra == nullptr,
              "lmc_img_myula_run: ncvx and d_fgrad_extra are two ways to pass the same term");
  LMC_REQUIRE(io->ncvx == nullptr || (io->ncvx->kind >= 0 && io->ncvx->kind <= 2 && (io->ncvx->kind == 0 ||
              (io->ncvx->gamma > 0 && (io->ncvx->kind == 1 || (io->ncvx->isotropic && io->ncvx->niter >= 0))))),
              "lmc_img_myula_run: bad ncvx descriptor (ME-TV needs isotropic = 1)");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_myula<double>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st)
                          : run_myula<float>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st);
}

extern "C" int lmc_img_myula_band_supported(const lmc_img_problem* pb, int dtype) {
  if (pb == nullptr || pb->ny < 1 || pb->nx < 1) return 0;
  if (dtype == LMC_F32) return myula_band_ok<float>(pb) ? 1 : 0;
  if (dtype == LMC_F64) return myula_band_ok<double>(pb) ? 1 : 0;
  return 0;
}

extern "C" int lmc_img_myula_run_band(const lmc_img_problem* pb, const lmc_img_myula_params* pr, int dtype, int64_t C,
                                 int64_t N, uint64_t seed, uint64_t chain_offset, const lmc_img_io* io,
                                 void* stream) {
  if (check_problem(pb, "lmc_img_myula_run_band") != 0) return -1;
  LMC_REQUIRE(pr != nullptr && io != nullptr, "lmc_img_myula_run_band: null descriptor");
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_myula_run_band: dtype must be LMC_F32 or LMC_F64");
  LMC_REQUIRE(C >= 0 && N >= 0, "lmc_img_myula_run_band: negative C or N");
  LMC_REQUIRE(pr->tau > 0 && pr->gamma > 0, "lmc_img_myula_run_band: tau and gamma must be positive");
  LMC_REQUIRE(pr->tv_niter >= 0 && pr->tv_sigma * pr->epsg * pr->gamma > 0,
              "lmc_img_myula_run_band: TV needs niter >= 0 and sigma*epsg*gamma > 0");
  LMC_REQUIRE(lmc_img_myula_band_supported(pb, dtype),
              "lmc_img_myula_run_band: needs 5/6/7 blur taps on both axes and a row that fits in shared memory "
              "(use lmc_img_myula_run)");
  if (C == 0 || N == 0) return 0;
  LMC_REQUIRE(io->d_x != nullptr && pb->d_b != nullptr, "lmc_img_myula_run_band: d_x and d_b are required");
  LMC_REQUIRE(pb->b_chain_stride == 0 || pb->b_chain_stride == (int64_t)pb->ny * pb->nx,
              "lmc_img_myula_run_band: b_chain_stride must be 0 or ny*nx");
  LMC_REQUIRE(io->ncvx == nullptr || io->ncvx->kind == 0 || io->d_fgrad_extra == nullptr,
              "lmc_img_myula_run_band: ncvx and d_fgrad_extra are two ways to pass the same term");
  LMC_REQUIRE(io->ncvx == nullptr || (io->ncvx->kind >= 0 && io->ncvx->kind <= 2 && (io->ncvx->kind == 0 ||
              (io->ncvx->gamma > 0 && (io->ncvx->kind == 1 || (io->ncvx->isotropic && io->ncvx->niter >= 0))))),
              "lmc_img_myula_run_band: bad ncvx descriptor (ME-TV needs isotropic = 1)");
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_myula<double>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st, true)
                          : run_myula<float>(pb, pr, C, N, seed, chain_offset, io, 0, nullptr, nullptr, st, true);
}

extern "C" int lmc_img_tv_prox(int dtype, int32_t ny, int32_t nx, double gamma, double sigma, int32_t niter,
                               double rtol, int64_t C, const void* d_in, void* d_out, unsigned long long* d_counters,
                               void* d_workspace, size_t workspace_bytes, void* stream) {
  LMC_REQUIRE(dtype == LMC_F32 || dtype == LMC_F64, "lmc_img_tv_prox: bad dtype");
  LMC_REQUIRE(ny >= 1 && nx >= 1 && C >= 0 && niter >= 0 && gamma > 0 && sigma > 0, "lmc_img_tv_prox: bad arguments");
  if (C == 0) return 0;
  LMC_REQUIRE(d_in != nullptr && d_out != nullptr, "lmc_img_tv_prox: null buffer");
  lmc_img_problem pb;
  identity_problem(pb, ny, nx);
  lmc_img_myula_params pr;
  memset(&pr, 0, sizeof(pr));
  pr.tv_sigma = gamma;
  pr.epsg = sigma;
  pr.tv_niter = niter;
  pr.tv_rtol = rtol;
  lmc_img_io io;
  memset(&io, 0, sizeof(io));
  io.d_workspace = d_workspace;
  io.workspace_bytes = workspace_bytes;
  io.d_counters = d_counters;
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == LMC_F64 ? run_myula<double>(&pb, &pr, C, 1, 0, 0, &io, 1, d_in, d_out, st)
                          : run_myula<float>(&pb, &pr, C, 1, 0, 0, &io, 1, d_in, d_out, st);
}
