#!/usr/bin/env python
"""Under torchrun: every rank asks shard.choose_local_devices for its GPU, the ranks compare answers (distinct devices,
one aligned block) and run one NCCL all-reduce on them.
    GFHE_DEVICE_MAP=probe python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 tools/devmap_check.py
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import arithmetic  # noqa: E402,F401
from gpu_fhe_b200 import shard  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lw, lr = int(os.environ.get("LOCAL_WORLD_SIZE", world)), int(os.environ["LOCAL_RANK"])
t0 = time.time()
device, probe = shard.choose_local_devices(lw, lr)
dt = time.time() - t0
torch.cuda.set_device(device)
dist.init_process_group("nccl", device_id=torch.device("cuda", device))
mine = torch.zeros(world, dtype=torch.int64, device="cuda")
mine[rank] = device + 1
dist.all_reduce(mine)
devs = [int(x) - 1 for x in mine.tolist()]
ok = len(set(devs)) == world and devs == list(range(devs[0], devs[0] + world)) and devs[0] % world == 0
if rank == 0:
    print("devices per rank:", devs, "| probe:", probe, "| rank 0 waited %.1f s |" % dt, "OK" if ok else "BAD", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
