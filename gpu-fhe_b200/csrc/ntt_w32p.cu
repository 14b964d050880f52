// Kernels of the 32-bit arithmetic policy on the packed layout: moduli below 2^30, residues stored as 32-bit words.
#include "ntt_kernels.inl"
namespace gfhe { template struct Launcher<Shoup32P>; }
