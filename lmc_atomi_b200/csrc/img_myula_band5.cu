// MYULA row-band kernels for 5-tap blurs (see img_myula_band_tu.inl).
#define LMC_BAND_NT 5
#include "img_myula_band_tu.inl"
