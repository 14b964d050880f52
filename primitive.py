"""Drop-in for /root/reference/primitive.py -- re-exports arithmetic (primitive.py:1), keeps the
placeholder XXX() (primitive.py:3-4) verbatim, and adds what the module was meant to hold: the
primitive roots of unity and the RNS ring they define.
"""
from arithmetic import *  # noqa: F401,F403
from arithmetic import _pkg

_r = _pkg.ring

__all__ = ["np", "vec_add", "vec_sub", "vec_mul", "NTT", "iNTT", "XXX",
           "find_psi", "ntt_primes", "is_probable_prime", "set_ring", "get_ring", "ring", "make_plan", "clear_plans",
           "save_plan", "load_plan"]

find_psi = _r.find_psi                    # primitive 2N-th root of unity mod q
ntt_primes = _r.ntt_primes                # largest `bits`-bit primes = 1 mod 2N
is_probable_prime = _r.is_probable_prime
set_ring = _r.set_ring                    # moduli used by NTT(x)/iNTT(x) when MOD= is not given
get_ring = _r.get_ring
ring = _r.ring                            # context manager form of set_ring
clear_plans = _r.clear_plans
save_plan = _r.save_plan                  # serialisable table cache: plan image to a file ...
load_plan = _r.load_plan                  # ... and back into HBM without root search or table generation


def XXX():
    print("XXX")


def make_plan(N, moduli, device=None):
    """Generate the twiddle + Shoup tables for (N, moduli) now (otherwise built on first use) and
    keep them resident in the device's HBM."""
    import torch
    dev = torch.cuda.current_device() if device is None else int(device)
    return _r.get_plan(dev, N, _r._norm_moduli(moduli))
