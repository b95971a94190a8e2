import os, sys, traceback
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
os.environ.setdefault("RETH_BOARD_TIMEOUT_MS", "3000")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_board_gpu as T
for G in [int(x) for x in os.environ.get("GS", "2").split(",")]:
    try:
        T._loop_body(G); print("body", G, "ok", flush=True)
    except BaseException:
        print("body", G, "FAILED", traceback.format_exc()[-600:], flush=True)
