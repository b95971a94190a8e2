"""BASELINE config 5 at N GPUs in Mode G (the multi-GPU priority-update path that couples the shards): tree-only
shards, strict alternation, per iteration  roots exchange (board) -> global stratified draw of this shard's share ->
write-back of the share.  Global batch 64..8192 x capacity per shard, beside Mode R (independent shards, each drawing
and updating batch / N per iteration).  torchrun, one rank per GPU; rank 0 appends JSON lines to OUT."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from reth_b200 import ReplayShard
from reth_b200.sharded import ShardGroup

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = os.environ.get("OUT", "gpurun_out/sweep_config5_modeg.jsonl")
caps = [1 << 16, 1 << 20, (1 << 24) // world]
batches = [64, 512, 2048, 8192]
steps = int(os.environ.get("STEPS", 300))


def timed(fn):
    fn(20)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(steps); e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()) * 1e3 / steps


for cap in caps:
    g = torch.Generator(device="cuda").manual_seed(rank)
    w = torch.randn(cap, device="cuda", generator=g).abs() * (1 + 9 * (torch.rand(cap, device="cuda", generator=g) < 0.05)) + 1e-3
    for GB in batches:
        sh = ReplayShard(cap, [], max_batch=GB, store_rows=False, ring_depth=4, device=local)
        for st in range(0, cap, 16384):
            sh.append([], w[st:st + 16384].float(), alpha=0.6)
        sg = ShardGroup(sh)
        sg.connect_board()
        prio = torch.rand(steps * GB, device="cuda", generator=g) + 0.01
        us_g = timed(lambda n: sh.step_loop(n, GB, prio, seed=3, global_mode=True, gather=False))
        B = max(1, GB // world)
        us_r = timed(lambda n: sh.step_loop(n, B, prio, seed=3, gather=False))
        if rank == 0:
            row = {"n_gpus": world, "capacity_per_shard": cap, "global_batch": GB, "mode_g_us": us_g, "mode_g_Midx_per_s": GB / us_g,
                   "mode_r_us_per_shard_batch": us_r, "mode_r_batch_per_shard": B, "mode_r_Midx_per_s": world * B / us_r}
            print(row, flush=True)
            with open(out, "a") as f:
                f.write(json.dumps(row) + "\n")
        dist.barrier()
        sh.close()
dist.destroy_process_group()
