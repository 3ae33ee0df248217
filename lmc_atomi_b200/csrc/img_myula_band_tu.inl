// Included by img_myula_band{5,6,7}.cu with LMC_BAND_NT defined: the band-kernel instantiations of one tap count.
#include "img_myula_tile.cuh"
#include "img_myula_band.inl"

namespace {

template <typename real, int NT, int GW, int TW>
int launch_myula_band(const MyulaArgs<real>& A, const MyulaArgs<real>& B, cudaStream_t st) {
  const size_t smem = myula_band_smem_bytes<real, NT>(A.pitch);
  const bool nc = A.nc.kind != 0 || B.mode == 2;
  auto kern = nc ? img_myula_band_kernel<real, NT, GW, TW, true> : img_myula_band_kernel<real, NT, GW, TW, false>;
  int dev = 0, sms = 0, occ = 0;
  LMC_CUDA_OK(cudaGetDevice(&dev));
  LMC_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  LMC_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, BAND_BT, smem));
  LMC_REQUIRE(occ >= 1, "MYULA band kernel does not fit on an SM (smem %zu)", smem);
  const long long cap = (long long)sms * occ, items = A.C * A.g.tiles;
  LMC_REQUIRE(items < (1ll << 31) && A.g.P < (1ll << 31), "MYULA band kernel: 32-bit item / pixel indices exceeded");
  int grid = (int)(items < cap ? items : cap);
  if (grid < 1) grid = 1;
  void* kargs[] = {(void*)&A, (void*)&B};
  LMC_CUDA_OK(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(BAND_BT), kargs, smem, st));
  return 0;
}


template <typename real>
int dispatch_band(const void* Av, const void* Bv, int gw, int tw, cudaStream_t st) {
  constexpr int NT = LMC_BAND_NT;
  const MyulaArgs<real>& A = *static_cast<const MyulaArgs<real>*>(Av);
  const MyulaArgs<real>& B = *static_cast<const MyulaArgs<real>*>(Bv);
  if (tw == 1) return launch_myula_band<real, NT, 1, 1>(A, B, st);
  return gw == 2 ? launch_myula_band<real, NT, 2, 2>(A, B, st) : launch_myula_band<real, NT, 1, 2>(A, B, st);
}

}  // namespace

#define LMC_CAT2(a, b) a##b
#define LMC_CAT(a, b) LMC_CAT2(a, b)
int LMC_CAT(lmc_myula_band_launch_nt, LMC_BAND_NT)(int dtype, const void* A, const void* B, int gw, int tw,
                                                   cudaStream_t st) {
  return dtype == LMC_F64 ? dispatch_band<double>(A, B, gw, tw, st) : dispatch_band<float>(A, B, gw, tw, st);
}
