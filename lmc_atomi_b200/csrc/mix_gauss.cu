// Translation unit 1 of the mixture-chain kernels: the instantiations of one target kind (see mix_kernels.cu).
#define LMC_MIX_TU 1
#include "mix_kernels.cu"
