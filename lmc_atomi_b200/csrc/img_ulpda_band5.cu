// ULPDA row-band kernels for 5-tap blurs (see img_ulpda_band_tu.inl).
#define LMC_BAND_NT 5
#include "img_ulpda_band_tu.inl"
