// ULPDA row-band kernels for 6-tap blurs (see img_ulpda_band_tu.inl).
#define LMC_BAND_NT 6
#include "img_ulpda_band_tu.inl"
