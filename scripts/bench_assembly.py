"""Mode G with cross-shard batch assembly over NVLink, one rank per GPU (torchrun).
Per step: all-gather of shard roots -> sample_global -> gather_to(learner rank 0) on every rank
-> all-gather (orders the learner's reads).  Reports global batches/s and the NVLink payload rate
into the learner GPU.  Rows: Atari-u8 (56 464 B); capacity per shard CAP; global batch GB."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from reth_b200 import ReplayShard
from reth_b200.sharded import ShardGroup

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
R = 56464
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cap = int(os.environ.get("CAP", 1 << 18)); steps = int(os.environ.get("STEPS", 200))
# CASES: comma-separated mode:global_batch.  assemble = single learner on rank 0 (slices travel over
# NVLink); local = data-parallel learners, each rank consumes its own slice (gather + write-back stay local);
# loop = the data-parallel step as ONE device loop (reth_replay_step_loop, Mode G flag): roots exchange by peer
# stores into every rank's roots board -> global draw -> gather || write-back, no host round trip per step.
# BOARD=1: local / assemble use the roots board instead of the NCCL all-gather (loop always does).
CASES = [(c.split(":")[0], int(c.split(":")[1])) for c in os.environ.get("CASES", "assemble:4096").split(",")]
GBMAX = max(gb for _, gb in CASES)
sh = ReplayShard(cap, ATARI, max_batch=GBMAX, ring_depth=4, device=local)
g = torch.Generator(device="cuda").manual_seed(rank)
for st in range(0, cap, 1 << 14):
    n = min(1 << 14, cap - st)
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01)
torch.cuda.synchronize()
sg = ShardGroup(sh)
sg.connect_learner(0)
BOARD = bool(int(os.environ.get("BOARD", "0")))
if BOARD or any(m == "loop" for m, _ in CASES):
    nccl_roots = ShardGroup(sh)                 # keeps the NCCL exchange for cases that do not use the board
    sg.connect_board()
    if not BOARD:
        sg.board = False
        sg.all_roots = nccl_roots.all_roots
prio_loop = torch.rand(steps * GBMAX, device="cuda", generator=g) + 0.01
newp = torch.rand(GBMAX, device="cuda", generator=g) + 0.01

for mode, GB in CASES:
    def step(k):
        sg.exchange_roots()
        sh.sample_global(GB, sg.all_roots, world, rank, beta=0.4, seed=7, slot=k % 2)
        if mode == "assemble":
            sg.assemble(slot=k % 2, dst_slot=2 + k % 2)
        else:
            sh.gather(k % 2)
            sh.update_slot(k % 2, newp[:GB], alpha=0.6)

    if mode == "loop":
        sh.step_loop(12, GB, prio_loop, seed=7, alpha=0.6, beta=0.4, global_mode=True)
    else:
        for k in range(10):
            step(k)
        sg.exchange_roots()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if mode == "loop":
        sh.step_loop(steps, GB, prio_loop, seed=7, alpha=0.6, beta=0.4, global_mode=True)
    else:
        for k in range(steps):
            step(k)
        sg.exchange_roots()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        us = float(ms.item()) * 1e3 / steps
        out = {"n_gpus": world, "mode": mode, "global_batch": GB, "capacity_per_shard": cap, "us_per_global_batch": us,
               "transitions_per_s": GB / us * 1e6}
        if mode == "assemble":
            out["nvlink_payload_GBps_into_learner"] = GB * (world - 1) / world * R / us / 1e3
            out["note"] = "every rank gathers its slice straight into rank 0's batch ring (peer stores from the gather kernel)"
        elif mode == "loop":
            out["note"] = ("device loop, per step: roots exchange kernel (peer stores into every rank's roots board over NVLink + "
                           "wait on the own board), draw of this shard's share, local gather || priority write-back of the share")
        else:
            out["roots"] = "board" if sg.board else "nccl"
            out["note"] = ("per step: all-gather of shard roots, draw of this shard's share of the global stratified batch, "
                           "local gather, priority write-back of the share")
        print(json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
