// Kernels of the 32-bit arithmetic policy: moduli below 2^30 (residues still stored as 64-bit words).
#include "ntt_kernels.inl"
namespace gfhe { template struct Launcher<Shoup32>; }
