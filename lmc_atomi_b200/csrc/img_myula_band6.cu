// MYULA row-band kernels for 6-tap blurs (see img_myula_band_tu.inl).
#define LMC_BAND_NT 6
#include "img_myula_band_tu.inl"
