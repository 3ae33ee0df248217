// Moreau-envelope gradient of the reference's non-convex data term algs.L2_ncvx_tv (algs.py:201-223, 270-291),
// MC-TV case (Op2 = Gradient): m(x) = G^T( w(Gx) Gx ) at one pixel, as a 7-point stencil.  Shared by the unit
// kernel (img_ncvx.cu) and the sampler kernels that evaluate the term in-kernel (img_myula*.{cu,inl}, img_ulpda*).
#pragma once
#include "lmc_common.cuh"

namespace lmcimg {

template <typename real>
struct MctvW {
  real inv_gamma, gamma;
  int iso;
  // weighted forward differences (v0, v1) of the pixel whose differences are (d0, d1)
  __device__ __forceinline__ void operator()(real d0, real d1, real& v0, real& v1) const {
    if (iso) {
      real n = Math<real>::sqrt(d0 * d0 + d1 * d1);
      n = (n != real(0)) ? n : real(1e-9);                       // algs.py:212,274
      const real w = Math<real>::min(inv_gamma, real(1) / n);    // np.minimum(1/gamma, 1/|Gx|)
      v0 = w * d0;
      v1 = w * d1;
    } else {
      // Gx - L1(1).prox(Gx, gamma): the residual of the soft threshold
      const real s0 = Math<real>::max(Math<real>::abs(d0) - gamma, real(0));
      const real s1 = Math<real>::max(Math<real>::abs(d1) - gamma, real(0));
      v0 = d0 - (d0 < real(0) ? -s0 : s0);
      v1 = d1 - (d1 < real(0) ? -s1 : s1);
    }
  }
};

// G^T(w(Gx) Gx) at pixel (i, j).  at(di, dj) returns x(i + di, j + dj); it is only called for in-image neighbours.
// G = forward differences, last row / column zero; accumulation order of pylops.Gradient.rmatvec
// (axis 0, then axis 1).
template <typename real, class At>
__device__ __forceinline__ real mctv_at(const MctvW<real>& W, int i, int j, int ny, int nx, At&& at) {
  const real xc = at(0, 0);
  const bool hd = i < ny - 1, hr = j < nx - 1, hu = i >= 1, hl = j >= 1;
  real v0c = real(0), v1c = real(0), v0u = real(0), v1l = real(0), dmy;
  W(hd ? at(1, 0) - xc : real(0), hr ? at(0, 1) - xc : real(0), v0c, v1c);      // differences at (i, j)
  if (hu) {   // axis-0 component at (i-1, j): needs both differences of that pixel
    const real xu = at(-1, 0);
    W(xc - xu, hr ? at(-1, 1) - xu : real(0), v0u, dmy);
  }
  if (hl) {   // axis-1 component at (i, j-1)
    const real xl = at(0, -1);
    W(hd ? at(1, -1) - xl : real(0), xc - xl, dmy, v1l);
  }
  real acc = real(0);
  if (hd) acc -= v0c;
  if (hu) acc += v0u;
  if (hr) acc -= v1c;
  if (hl) acc += v1l;
  return acc;
}

// what the samplers need of an L2_ncvx_tv term (lmc_img_ncvx of the C ABI, resolved for the kernel's real type)
template <typename real>
struct NcvxK {
  int kind;            // 0 none, 1 MC-TV, 2 ME-TV
  MctvW<real> W;       // MC-TV weights
  real post_div;       // MC-TV: m / post_div  (1 isotropic, gamma anisotropic; algs.py:277)
};

}  // namespace lmcimg
