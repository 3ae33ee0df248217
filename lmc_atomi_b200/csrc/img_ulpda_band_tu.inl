// Included by img_ulpda_band{5,6,7}.cu with LMC_BAND_NT defined: the band-kernel instantiations of one tap count.
#include "img_ulpda_tile.cuh"
#include "img_ulpda_band.inl"
#include "img_host_common.inl"

namespace {

#ifndef LMC_ULPDA_BAND_RB
#define LMC_ULPDA_BAND_RB 16
#endif
constexpr int BAND_RB = LMC_ULPDA_BAND_RB;

template <typename real, int NT, int W>
int launch_band(UlpdaArgs<real>& A, const MyulaArgs<real>& B, cudaStream_t st) {
  size_t smem = Band<real, NT, BAND_RB>::smem_bytes(A.pitch);
  if (B.mode == 3 && smem < (size_t)4 * RS * RS * sizeof(real)) smem = (size_t)4 * RS * RS * sizeof(real);   // tile TV phases
  auto kern = img_ulpda_band_kernel<real, NT, BAND_RB, W>;
  int dev = 0, sms = 0, occ = 0;
  LMC_CUDA_OK(cudaGetDevice(&dev));
  LMC_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LMC_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  LMC_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, BAND_BT, smem));
  LMC_REQUIRE(occ >= 1, "ULPDA band kernel does not fit on an SM (smem %zu)", smem);
  const long long cap = (long long)sms * occ;
  const long long g = A.group < A.C ? A.group : A.C;
  const long long items = g * A.g.tiles;
  LMC_REQUIRE(items < (1ll << 31) && A.g.P < (1ll << 31), "ULPDA band kernel: 32-bit item / pixel indices exceeded");
  int grid = (int)(items < cap ? items : cap);
  if (grid < 1) grid = 1;
  void* kargs[] = {(void*)&A, (void*)&B};
  LMC_CUDA_OK(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(BAND_BT), kargs, smem, st));
  return 0;
}

template <typename real>
int dispatch_band(void* Av, const void* Bv, int w, cudaStream_t st) {
  constexpr int NT = LMC_BAND_NT;
  UlpdaArgs<real>& A = *static_cast<UlpdaArgs<real>*>(Av);
  const MyulaArgs<real>& B = *static_cast<const MyulaArgs<real>*>(Bv);
  return w == 2 ? launch_band<real, NT, 2>(A, B, st) : launch_band<real, NT, 1>(A, B, st);
}

}  // namespace

#define LMC_CAT2(a, b) a##b
#define LMC_CAT(a, b) LMC_CAT2(a, b)
int LMC_CAT(lmc_ulpda_band_launch_nt, LMC_BAND_NT)(int dtype, void* A, const void* B, int w, cudaStream_t st) {
  return dtype == LMC_F64 ? dispatch_band<double>(A, B, w, st) : dispatch_band<float>(A, B, w, st);
}
