"""End to end through the SERVER-PROCESS layout (the reference's own: trainer.py:49-66): server in
its own process, this process holds Client + TorchCudaLoader (CUDA IPC batch ring).  Atari-u8 rows,
batch 512; capacity bounded (rows travel over ZeroMQ to get in)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from reth_b200 import service_mp

if __name__ == "__main__":
    cap, B = int(os.environ.get("CAP", 1 << 16)), 512
    G, ahead = int(os.environ.get("GROUP", 8)), int(os.environ.get("AHEAD", 2))
    defaults = bool(os.environ.get("DEFAULTS"))             # what a caller of the reference API gets untouched
    if defaults:
        proc, addr = service_mp.start_per(cap, B, alpha=0.6, beta=0.4, sample_start=B)
    else:
        proc, addr = service_mp.start_per(cap, B, alpha=0.6, beta=0.4, sample_start=B, max_ring=max(ahead + 1, 3) * G)
    try:
        client = service_mp.Client(addr, wire=os.environ.get("WIRE", "auto"))
        rng = np.random.default_rng(0)
        n = 2048
        s0 = rng.integers(0, 256, (n, 4, 84, 84), dtype=np.uint8)
        cols = [s0, rng.integers(0, 18, n), np.zeros(n, np.float32), s0, np.zeros(n, np.float32)]
        if defaults:
            loader = service_mp.TorchCudaLoader(addr)
        else:
            loader = service_mp.TorchCudaLoader(addr, buffer_size=(ahead + 1) * G, prefetch=G, ahead=ahead)
        wts = np.abs(rng.standard_normal(n)).astype(np.float32)
        client.append(cols, wts)
        loader.drain()                                            # shard created, ring attached: steady state from here
        t0 = time.perf_counter()
        for st in range(n, cap, n):
            client.append(cols, wts)
        loader.drain()
        t_fill = time.perf_counter() - t0
        data, idx, w = loader.sample()
        newp = np.abs(rng.standard_normal((64, B))).astype(np.float32)
        for i in range(64):
            data, idx, w = loader.sample(); client.update_priorities(idx, newp[i % 64])
        loader.drain()
        iters = int(os.environ.get("ITERS", 2000))
        pr = None
        if os.environ.get("PROFILE"):
            import cProfile
            pr = cProfile.Profile()
            pr.enable()
        t0 = time.perf_counter()
        for i in range(iters):
            data, idx, w = loader.sample()
            client.update_priorities(idx, newp[i % 64])
        loader.drain()
        dt = time.perf_counter() - t0
        if pr is not None:
            import pstats
            pr.disable()
            pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(18)
        print(json.dumps({"layout": "server process + CUDA IPC loader", "capacity": cap, "batch": B, "prefetch_group": loader.group,
                          "ahead": loader.ahead, "defaults": defaults, "transitions_per_s": iters * B / dt, "us_per_iteration": dt / iters * 1e6,
                          "wire": client.wire, "ingest_rows_per_s": (cap - n) / t_fill, "ingest_GBps": (cap - n) * 56464 / t_fill / 1e9}))
        loader.close()
    finally:
        proc.terminate(); proc.join()
