"""Where does the time of the config-4 loop go (reth_replay_step_loop_ingest, 64 rows appended per learner step)?
One process, the loop timed with parts of the ingest switched off through RETH_C4_PROBE (replay.cu, timing only:
bits 0/1 drop the ordering between the row copy and the gathers, which is unsafe outside a timing run).
Prints one JSON line per variant."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as entry
entry.build()
from reth_b200 import ReplayShard

ATARI = [(np.uint8, (4, 84, 84)), (np.int64, ()), (np.float32, ()), (np.uint8, (4, 84, 84)), (np.float32, ())]
cap, B, G, chunk = 1 << 19, 512, 8, 8
sh = ReplayShard(cap, ATARI, max_batch=B, ring_depth=3 * G)
g = torch.Generator(device="cuda").manual_seed(0)
for st in range(0, cap, 1 << 14):
    n = 1 << 14
    s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
    z = torch.zeros(n, device="cuda")
    sh.append([s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g) + 0.01, alpha=0.5)
prio = (torch.rand(G * B, device="cuda") + 0.01).repeat(chunk)
n = 64 * G
s0 = torch.randint(0, 256, (n, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=g)
z = torch.zeros(n, device="cuda")
cols, w = [s0, z.long(), z, s0, z], torch.rand(n, device="cuda", generator=g)


def timed(label, probe=None, ingest=True, groups=400, **kw):
    if probe is None:
        os.environ.pop("RETH_C4_PROBE", None)
    else:
        os.environ["RETH_C4_PROBE"] = str(probe)

    def loop(k):
        for _ in range(0, k, chunk):
            sh.step_loop(G * chunk, B, prio, alpha=0.5, beta=0.4, group=G, ingest=(cols, w) if ingest else None, **kw)
    loop(chunk * 2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(groups)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / groups
    print(json.dumps({"variant": label, "probe": probe, "us_per_group": round(us, 2),
                      "sampled_M_per_s": round(G * B / us, 2)}), flush=True)


if os.environ.get("C4_ONLY"):                 # shape sweeps (RETH_UPDATE_CLUSTER=...): the two lines that matter
    timed("plain loop, no ingest", ingest=False)
    timed("ingest (product path)")
    timed("tree stream alone with ingest leaves folded (no gathers, no copy)", probe=8, gather=False)
    sys.exit(0)
timed("plain loop, no ingest", ingest=False)
timed("ingest (product path: leaves in the write-back's launch)")
timed("ingest, leaves in a launch of their own", probe=16)
timed("ingest, copy/gather ordering dropped", probe=3)
timed("ingest, no leaf update", probe=4)
timed("ingest, no leaf update, ordering dropped", probe=7)
timed("ingest, no row copy (leaves + waits only)", probe=8)
timed("ingest, leaves only (no copy, no waits)", probe=11)
timed("tree stream alone: no gathers, no ingest", ingest=False, gather=False)
timed("tree stream alone with ingest leaves folded (no gathers, no copy)", probe=8, gather=False)
timed("tree stream alone with ingest leaves in their own launch", probe=24, gather=False)
timed("ingest, one gather stream", one_gather_stream=True)
timed("plain loop, one gather stream", ingest=False, one_gather_stream=True)
os.environ.pop("RETH_C4_PROBE", None)
