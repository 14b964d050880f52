// Kernels of the exact (Harvey) arithmetic policy: any modulus below 2^62.
#include "ntt_kernels.inl"
namespace gfhe { template struct Launcher<Harvey>; }
