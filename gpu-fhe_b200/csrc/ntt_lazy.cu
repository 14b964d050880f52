// Kernels of the lazy arithmetic policy: moduli in [2^19, 2^60) (four spare bits per residue).
#include "ntt_kernels.inl"
namespace gfhe { template struct Launcher<Lazy60>; }
