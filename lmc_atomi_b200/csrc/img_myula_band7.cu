// MYULA row-band kernels for 7-tap blurs (see img_myula_band_tu.inl).
#define LMC_BAND_NT 7
#include "img_myula_band_tu.inl"
