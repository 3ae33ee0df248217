// ULPDA row-band kernels for 7-tap blurs (see img_ulpda_band_tu.inl).
#define LMC_BAND_NT 7
#include "img_ulpda_band_tu.inl"
