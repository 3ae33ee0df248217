// Host-side helpers shared by the image-chain translation units (included, not compiled alone).
namespace {

inline uint2 img_key(uint64_t seed) { return make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)); }

inline int check_problem(const lmc_img_problem* pb, const char* who) {
  LMC_REQUIRE(pb != nullptr, "%s: null problem descriptor", who);
  LMC_REQUIRE(pb->ny >= 1 && pb->nx >= 1, "%s: bad image size %d x %d", who, pb->ny, pb->nx);
  LMC_REQUIRE(pb->ntap_y >= 1 && pb->ntap_y <= lmcimg::HMAX + 1 && pb->ntap_x >= 1 && pb->ntap_x <= lmcimg::HMAX + 1,
              "%s: blur taps per axis must be in 1..%d (got %d, %d)", who, lmcimg::HMAX + 1, pb->ntap_y, pb->ntap_x);
  LMC_REQUIRE(pb->lo_y <= 0 && pb->lo_y >= -(pb->ntap_y - 1) && pb->lo_x <= 0 && pb->lo_x >= -(pb->ntap_x - 1),
              "%s: blur window must contain the output pixel (lo_y=%d, lo_x=%d)", who, pb->lo_y, pb->lo_x);
  return 0;
}

template <typename real>
void fill_geom(const lmc_img_problem* pb, lmcimg::Geom<real>& g) {
  memset(&g, 0, sizeof(g));
  g.ny = pb->ny;
  g.nx = pb->nx;
  g.tiles_y = (pb->ny + lmcimg::TH - 1) / lmcimg::TH;
  g.tiles_x = (pb->nx + lmcimg::TW - 1) / lmcimg::TW;
  g.tiles = g.tiles_y * g.tiles_x;
  g.P = (long long)pb->ny * pb->nx;
  g.ntap_y = pb->ntap_y;
  g.ntap_x = pb->ntap_x;
  g.lo_y = pb->lo_y;
  g.lo_x = pb->lo_x;
  g.e = (pb->ntap_y > pb->ntap_x ? pb->ntap_y : pb->ntap_x) - 1;
  for (int t = 0; t < pb->ntap_y; ++t) g.wy[t] = (real)pb->wy[t];
  for (int t = 0; t < pb->ntap_x; ++t) g.wx[t] = (real)pb->wx[t];
}

inline void identity_problem(lmc_img_problem& pb, int ny, int nx) {
  memset(&pb, 0, sizeof(pb));
  pb.ny = ny;
  pb.nx = nx;
  pb.ntap_y = pb.ntap_x = 1;
  pb.wy[0] = pb.wx[0] = 1.0;
}

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// cooperative launch: one CTA per SM x occupancy, capped by the work items; `args2` = optional second kernel argument
template <typename Kern, typename Args>
int launch_persistent(Kern kern, const Args& args, long long items, size_t smem, cudaStream_t st,
                      const Args* args2 = nullptr) {
  int dev = 0, sms = 0, occ = 0, coop = 0;
  LMC_CUDA_OK(cudaGetDevice(&dev));
  LMC_CUDA_OK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  LMC_REQUIRE(coop != 0, "device does not support cooperative launches");
  LMC_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  LMC_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, lmcimg::NTHREADS, smem));
  LMC_REQUIRE(occ >= 1, "image kernel does not fit on an SM (smem %zu)", smem);
  long long cap = (long long)sms * occ;
  int grid = (int)(items < cap ? items : cap);
  if (grid < 1) grid = 1;
  void* kargs[] = {(void*)&args, (void*)args2};
  LMC_CUDA_OK(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(lmcimg::NTHREADS), kargs, smem, st));
  return 0;
}

}  // namespace
