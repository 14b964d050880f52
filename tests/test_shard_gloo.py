"""CPU suite, part 4: the N>1 path on world_size-2 (and 4) gloo -- partition, per-rank compute
stand-in (the C oracle, as checker only), final all-gather back to [L, B, N]."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, L, B, N, q_bits, result_dir, cyclic=False):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import shard
    from oracle import cref
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    moduli = O.ntt_primes(q_bits, L)
    rng = np.random.default_rng(1234)          # same full input on every rank
    a = np.stack([rng.integers(0, q, (B, N), dtype=np.uint64) for q in moduli])
    sh = shard.partition(L, B, world, rank, cyclic=cyclic)
    mine = shard.take(torch.from_numpy(a.view(np.int64)), sh)
    my_moduli = [moduli[l] for l in sh.limb_ids]
    R = cref.Ring(N, my_moduli, [O.find_psi(q, N) for q in my_moduli])
    local = torch.from_numpy(R.ntt(mine.numpy().view(np.uint64)).view(np.int64))
    full = shard.all_gather_limbs(local, sh)
    Rf = cref.Ring(N, moduli, [O.find_psi(q, N) for q in moduli])
    ok = np.array_equal(full.numpy().view(np.uint64), Rf.ntt(a))
    open(os.path.join(result_dir, "rank%d" % rank), "w").write("ok" if ok else "bad")
    dist.destroy_process_group()


@pytest.mark.parametrize("cyclic", [False, True])
@pytest.mark.parametrize("world,L,B", [(2, 4, 3), (2, 1, 4), (4, 2, 4)])
def test_sharded_ntt_allgather_gloo(tmp_path, world, L, B, cyclic):
    """block map: one gather of the whole shard; cyclic map: one (asynchronous) gather per local limb, each landing in its
    contiguous slab of the result"""
    mp.spawn(_worker, args=(world, _free_port(), L, B, 64, 30, str(tmp_path), cyclic), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / ("rank%d" % r)).read() == "ok"
