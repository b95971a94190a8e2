"""Mode G device loop with a prefetch group (torchrun, one rank per GPU): per iteration ONE roots exchange over
NVLink, G global stratified draws in one launch (shares packed), one gather and one write-back of the packed
share.  CASES = comma-separated global_batch:group.  Atari-u8 rows, capacity CAP per shard."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from reth_b200 import ReplayShard
from reth_b200.sharded import ShardGroup

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cap = int(os.environ.get("CAP", 1 << 18)); iters = int(os.environ.get("ITERS", 200))
CASES = [tuple(int(x) for x in c.split(":")) for c in os.environ.get("CASES", "4096:1,4096:8").split(",")]
for GB, G in CASES:
    sh = ReplayShard(cap, ATARI, max_batch=GB, ring_depth=max(4, int(os.environ.get('DEPTH', 2)) * G), device=local)
    g = torch.Generator(device="cuda").manual_seed(rank)
    for st in range(0, cap, 1 << 14):
        n = min(1 << 14, cap - st)
        s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
        z = torch.zeros(n, device="cuda")
        sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
    sg = ShardGroup(sh)
    sg.connect_board()
    steps = iters * G
    prio = torch.rand(steps * GB, device="cuda", generator=g) + 0.01
    sh.step_loop(8 * G, GB, prio, seed=7, alpha=0.6, beta=0.4, global_mode=True, group=G)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sh.step_loop(steps, GB, prio, seed=7, alpha=0.6, beta=0.4, global_mode=True, group=G)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        us = float(ms.item()) * 1e3 / steps
        print(json.dumps({"n_gpus": world, "mode": "loop", "global_batch": GB, "prefetch_group": G, "capacity_per_shard": cap,
                          "us_per_global_batch": us, "transitions_per_s": GB / us * 1e6,
                          "note": "device loop; per iteration: roots exchange kernel (NVLink peer stores), G global draws "
                                  "(shares packed), gather || write-back of the packed share"}), flush=True)
    dist.barrier()
    sh.close()
    del prio
dist.destroy_process_group()
