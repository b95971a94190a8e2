import os, sys, threading
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
os.environ.setdefault("RETH_BOARD_TIMEOUT_MS", "3000")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from reth_b200 import ReplayShard as Shard, _cabi
from reth_b200.shard import _DevArray
import test_board_gpu as T
G = int(os.environ.get("G", 2))
cap, GB, n_shards, steps, seed = 4096, 256, 2, int(os.environ.get("STEPS", 8)), 1234
struct = [(np.uint8, (64,)), (np.int64, ())]
rng = np.random.default_rng(3)
shards = []
for g in range(n_shards):
    sh = Shard(cap, struct, max_batch=GB, ring_depth=4)
    sh.append([rng.integers(0, 256, (cap, 64), dtype=np.uint8), np.arange(cap)], (rng.random(cap) + 0.01).astype(np.float32))
    shards.append(sh)
T._connect(shards)
prio = [torch.rand(steps * GB, device="cuda") + 0.01 for g in range(n_shards)]
# warm every kernel
for sh in shards:
    sh.sample(GB, slot=0); sh.gather(0); sh.update_slot(0, prio[0][:GB])
torch.cuda.synchronize()
streams = [torch.cuda.Stream() for _ in range(n_shards)]
errs = []
def run(g):
    try:
        shards[g].step_loop(steps, GB, prio[g], seed=seed, global_mode=True, group=G, stream=streams[g])
    except Exception as e:
        errs.append(e)
th = [threading.Thread(target=run, args=(g,)) for g in range(n_shards)]
[t.start() for t in th]; [t.join() for t in th]
torch.cuda.synchronize()
print("errs", errs)
for g, sh in enumerate(shards):
    ex = _cabi.BoardExport.from_buffer_copy(sh.board_export())
    b = torch.as_tensor(_DevArray(ex.local_ptr, (2, 64, 4), np.float64, sh), device="cuda").cpu().numpy()
    seqs = b.view(np.uint64)[:, :n_shards, 2]
    print("shard", g, "board seq [parity][src]:", seqs.tolist())
    try:
        sh.tree_arrays(); print("shard", g, "ok")
    except Exception as e:
        print("shard", g, "ERR", e)
