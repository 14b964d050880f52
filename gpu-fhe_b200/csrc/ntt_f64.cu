// Kernels of the FP64 arithmetic policy: moduli below 2^50 (error-free products on the FP64 pipe).
#include "ntt_kernels.inl"
namespace gfhe { template struct Launcher<Fp50>; }
